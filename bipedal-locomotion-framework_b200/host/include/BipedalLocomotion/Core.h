/**
 * Core.h -- logging, small math types and the parameter handler of the host layer.
 *
 * The host layer re-provides, with a batch dimension, the C++ surface of the reference's IK tick so that
 * code written against bipedal-locomotion-framework keeps its shape.  The reference's types
 * come from Eigen / manif / iDynTree / spdlog, none of which exists in this image; the stand-ins below keep
 * the coefficient orders those libraries use:
 *   - Quaterniond coefficients (x, y, z, w)                       (Eigen / manif storage)
 *   - SE3Tangentd::coeffs() = [lin(3); ang(3)]                    (manif; SE3Task.cpp:323-336)
 *   - log()->error("[Class::method] ...") + return false          (src/TextLogging, every fallible call)
 *   - IParametersHandler::getParameter / setParameter / getGroup / setGroup / clone
 *     (src/ParametersHandler/include/BipedalLocomotion/ParametersHandler/IParametersHandler.h:42-260)
 *     with StdImplementation as the only back-end (TOML / YARP files are out of scope).
 */
#ifndef BLF_B200_CORE_H
#define BLF_B200_CORE_H

#include <array>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <variant>
#include <vector>

namespace BipedalLocomotion
{

// ------------------------------------------------------------------ TextLogging
class Logger
{
public:
    enum class Level { debug, info, warn, error, off };
    void setLevel(Level l) { m_level = l; }
    template <typename... Args> void error(const std::string& fmt, const Args&... a) { emit(Level::error, "error", fmt, a...); }
    template <typename... Args> void warn(const std::string& fmt, const Args&... a) { emit(Level::warn, "warning", fmt, a...); }
    template <typename... Args> void info(const std::string& fmt, const Args&... a) { emit(Level::info, "info", fmt, a...); }
    template <typename... Args> void debug(const std::string& fmt, const Args&... a) { emit(Level::debug, "debug", fmt, a...); }
    const std::string& lastError() const { return m_lastError; }

private:
    Level m_level{Level::warn};
    std::string m_lastError;
    static void subst(std::ostringstream& os, const std::string& fmt, size_t pos) { os << fmt.substr(pos); }
    template <typename T, typename... Rest>
    static void subst(std::ostringstream& os, const std::string& fmt, size_t pos, const T& v, const Rest&... rest)
    {
        const size_t p = fmt.find("{}", pos);
        if (p == std::string::npos) {
            os << fmt.substr(pos);
            return;
        }
        os << fmt.substr(pos, p - pos) << v;
        subst(os, fmt, p + 2, rest...);
    }
    template <typename... Args> void emit(Level l, const char* tag, const std::string& fmt, const Args&... a)
    {
        std::ostringstream os;
        subst(os, fmt, 0, a...);
        if (l == Level::error) m_lastError = os.str();
        if (l >= m_level) std::fprintf(stderr, "[blf][%s] %s\n", tag, os.str().c_str());
    }
};

inline Logger* log()
{
    static Logger logger;
    return &logger;
}

// ------------------------------------------------------------------ math stand-ins
using Vector3d = std::array<double, 3>;
using VectorXd = std::vector<double>;

struct Quaterniond {
    double x{0}, y{0}, z{0}, w{1};
    std::array<double, 4> coeffs() const { return {x, y, z, w}; }
    void normalize()
    {
        const double n = std::sqrt(x * x + y * y + z * z + w * w);
        x /= n, y /= n, z /= n, w /= n;
    }
};

/** manif::SO3d stand-in: a unit quaternion (ManifConversions.h:44-50 normalises before constructing). */
struct SO3d {
    Quaterniond q;
    SO3d() = default;
    explicit SO3d(const Quaterniond& quat) : q(quat) { q.normalize(); }
    const Quaterniond& quat() const { return q; }
    static SO3d Identity() { return SO3d(); }
    /** manif SO3Tangent::exp of a rotation vector */
    static SO3d Exp(const Vector3d& phi)
    {
        const double th = std::sqrt(phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2]);
        const double k = th > 1e-8 ? std::sin(0.5 * th) / th : 0.5 - th * th / 48.0;
        return SO3d(Quaterniond{k * phi[0], k * phi[1], k * phi[2], std::cos(0.5 * th)});
    }
    SO3d inverse() const { return SO3d(Quaterniond{-q.x, -q.y, -q.z, q.w}); }
    /** composition: (*this) * o */
    SO3d operator*(const SO3d& o) const
    {
        const Quaterniond &a = q, &b = o.q;
        return SO3d(Quaterniond{a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y, a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x,
                                a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w, a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z});
    }
    /** manif SO3::log as a rotation vector */
    Vector3d Log() const
    {
        const double s = q.w < 0 ? -1.0 : 1.0;
        const double x = s * q.x, y = s * q.y, z = s * q.z, w = s * q.w;
        const double n = std::sqrt(x * x + y * y + z * z);
        const double k = n < 1e-10 ? 2.0 / w * (1.0 - n * n / (3.0 * w * w)) : 2.0 * std::atan2(n, w) / n;
        return {k * x, k * y, k * z};
    }
    std::array<double, 9> rotation() const
    {
        const double x = q.x, y = q.y, z = q.z, w = q.w;
        return {1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w), 2 * (x * y + z * w), 1 - 2 * (x * x + z * z),
                2 * (y * z - x * w),     2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)};
    }
};

struct SO3Tangentd {
    Vector3d w{0, 0, 0};
    static SO3Tangentd Zero() { return {}; }
    const Vector3d& coeffs() const { return w; }
};

/** manif::SE3d::Tangent stand-in; coeffs() = [lin; ang]. */
struct SE3Tangentd {
    Vector3d linear{0, 0, 0}, angular{0, 0, 0};
    static SE3Tangentd Zero() { return {}; }
    const Vector3d& lin() const { return linear; }
    const Vector3d& ang() const { return angular; }
    std::array<double, 6> coeffs() const { return {linear[0], linear[1], linear[2], angular[0], angular[1], angular[2]}; }
};

/** manif::SE3d stand-in. */
struct SE3d {
    Vector3d p{0, 0, 0};
    SO3d R;
    SE3d() = default;
    SE3d(const Vector3d& position, const SO3d& rot) : p(position), R(rot) {}
    const Vector3d& translation() const { return p; }
    const Quaterniond& quat() const { return R.quat(); }
    using Tangent = SE3Tangentd;
};

// ------------------------------------------------------------------ ParametersHandler
namespace ParametersHandler
{

class IParametersHandler
{
public:
    using shared_ptr = std::shared_ptr<IParametersHandler>;
    using weak_ptr = std::weak_ptr<IParametersHandler>;
    virtual ~IParametersHandler() = default;

    virtual bool getParameter(const std::string& name, int& v) const = 0;
    virtual bool getParameter(const std::string& name, double& v) const = 0;
    virtual bool getParameter(const std::string& name, std::string& v) const = 0;
    virtual bool getParameter(const std::string& name, bool& v) const = 0;
    virtual bool getParameter(const std::string& name, std::chrono::nanoseconds& v) const = 0;
    virtual bool getParameter(const std::string& name, std::vector<int>& v) const = 0;
    virtual bool getParameter(const std::string& name, std::vector<double>& v) const = 0;
    virtual bool getParameter(const std::string& name, std::vector<std::string>& v) const = 0;
    virtual bool getParameter(const std::string& name, std::vector<bool>& v) const = 0;

    virtual void setParameter(const std::string& name, const int& v) = 0;
    virtual void setParameter(const std::string& name, const double& v) = 0;
    virtual void setParameter(const std::string& name, const std::string& v) = 0;
    virtual void setParameter(const std::string& name, const char* v) = 0;
    virtual void setParameter(const std::string& name, const bool& v) = 0;
    virtual void setParameter(const std::string& name, const std::chrono::nanoseconds& v) = 0;
    virtual void setParameter(const std::string& name, const std::vector<int>& v) = 0;
    virtual void setParameter(const std::string& name, const std::vector<double>& v) = 0;
    virtual void setParameter(const std::string& name, const std::vector<std::string>& v) = 0;
    virtual void setParameter(const std::string& name, const std::vector<bool>& v) = 0;

    virtual weak_ptr getGroup(const std::string& name) const = 0;
    virtual bool setGroup(const std::string& name, shared_ptr group) = 0;
    virtual std::string toString() const = 0;
    virtual bool isEmpty() const = 0;
    virtual void clear() = 0;
    virtual shared_ptr clone() const = 0;
};

/** In-memory handler (src/ParametersHandler/.../StdImplementation.h): name -> typed value, nested groups. */
class StdImplementation : public IParametersHandler
{
    using Value = std::variant<int, double, std::string, bool, std::chrono::nanoseconds, std::vector<int>, std::vector<double>,
                               std::vector<std::string>, std::vector<bool>>;
    std::map<std::string, Value> m_values;
    std::map<std::string, shared_ptr> m_groups;

    template <typename T> bool get(const std::string& name, T& v) const
    {
        auto it = m_values.find(name);
        if (it == m_values.end()) return false;
        if (auto p = std::get_if<T>(&it->second)) {
            v = *p;
            return true;
        }
        return false;
    }

public:
    bool getParameter(const std::string& name, int& v) const override { return get(name, v); }
    bool getParameter(const std::string& name, double& v) const override
    {
        if (get(name, v)) return true;
        int i;
        if (get(name, i)) {  // an integer literal is a valid double, as in the file back-ends
            v = i;
            return true;
        }
        return false;
    }
    bool getParameter(const std::string& name, std::string& v) const override { return get(name, v); }
    bool getParameter(const std::string& name, bool& v) const override { return get(name, v); }
    bool getParameter(const std::string& name, std::chrono::nanoseconds& v) const override
    {
        if (get(name, v)) return true;
        double s;  // seconds as a floating-point number (the way the ini / toml files carry sampling_time)
        if (getParameter(name, s)) {
            v = std::chrono::nanoseconds(static_cast<long long>(std::llround(s * 1e9)));
            return true;
        }
        return false;
    }
    bool getParameter(const std::string& name, std::vector<int>& v) const override { return get(name, v); }
    bool getParameter(const std::string& name, std::vector<double>& v) const override { return get(name, v); }
    bool getParameter(const std::string& name, std::vector<std::string>& v) const override { return get(name, v); }
    bool getParameter(const std::string& name, std::vector<bool>& v) const override { return get(name, v); }

    void setParameter(const std::string& name, const int& v) override { m_values[name] = v; }
    void setParameter(const std::string& name, const double& v) override { m_values[name] = v; }
    void setParameter(const std::string& name, const std::string& v) override { m_values[name] = v; }
    void setParameter(const std::string& name, const char* v) override { m_values[name] = std::string(v); }
    void setParameter(const std::string& name, const bool& v) override { m_values[name] = v; }
    void setParameter(const std::string& name, const std::chrono::nanoseconds& v) override { m_values[name] = v; }
    void setParameter(const std::string& name, const std::vector<int>& v) override { m_values[name] = v; }
    void setParameter(const std::string& name, const std::vector<double>& v) override { m_values[name] = v; }
    void setParameter(const std::string& name, const std::vector<std::string>& v) override { m_values[name] = v; }
    void setParameter(const std::string& name, const std::vector<bool>& v) override { m_values[name] = v; }

    weak_ptr getGroup(const std::string& name) const override
    {
        auto it = m_groups.find(name);
        return it == m_groups.end() ? weak_ptr() : weak_ptr(it->second);
    }
    bool setGroup(const std::string& name, shared_ptr group) override
    {
        if (!group) return false;
        m_groups[name] = group->clone();  // the reference stores a copy of the group
        return true;
    }
    std::string toString() const override
    {
        std::string s;
        for (auto& kv : m_values) s += kv.first + " ";
        for (auto& kv : m_groups) s += "[" + kv.first + "] ";
        return s;
    }
    bool isEmpty() const override { return m_values.empty() && m_groups.empty(); }
    void clear() override { m_values.clear(), m_groups.clear(); }
    shared_ptr clone() const override
    {
        auto c = std::make_shared<StdImplementation>();
        c->m_values = m_values;
        for (auto& kv : m_groups) c->m_groups[kv.first] = kv.second->clone();
        return c;
    }
};

}  // namespace ParametersHandler
}  // namespace BipedalLocomotion

#endif
