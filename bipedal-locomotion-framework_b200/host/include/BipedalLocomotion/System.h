/**
 * System.h -- the System:: interface types the IK tick is built on, kept with the reference's names:
 * VariablesHandler (src/System/src/VariablesHandler.cpp:58-120), WeightProvider / ConstantWeightProvider
 * (src/System/src/ConstantWeightProvider.cpp:14-52), LinearTask (src/System/include/BipedalLocomotion/System/
 * LinearTask.h:33-135), Source / Advanceable (Advanceable.h:35-69), ILinearTaskSolver (ILinearTaskSolver.h:38-147)
 * and the string -> builder Factory (Factory.h:25-80).
 */
#ifndef BLF_B200_SYSTEM_H
#define BLF_B200_SYSTEM_H

#include <BipedalLocomotion/Core.h>

#include <functional>
#include <unordered_map>

namespace BipedalLocomotion
{
namespace System
{

class VariablesHandler
{
public:
    class VariableDescription
    {
    public:
        static constexpr std::ptrdiff_t InvalidIndex{-1};
        std::ptrdiff_t offset{InvalidIndex};
        std::ptrdiff_t size{InvalidIndex};
        std::string name;
        bool isValid() const { return offset >= 0 && size >= 0; }
        static VariableDescription InvalidVariable() { return VariableDescription(); }
    };

    /** keys: variables_name (vector<string>), variables_size (vector<int>); erases the previous content */
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> handler) noexcept
    {
        constexpr auto logPrefix = "[VariablesHandler::initialize]";
        auto ptr = handler.lock();
        if (!ptr) {
            log()->error("{} The handler is not valid.", logPrefix);
            return false;
        }
        std::vector<std::string> names;
        std::vector<int> sizes;
        if (!ptr->getParameter("variables_name", names)) {
            log()->error("{} Unable to get the name of the variables.", logPrefix);
            return false;
        }
        if (!ptr->getParameter("variables_size", sizes)) {
            log()->error("{} Unable to get the size of the variables.", logPrefix);
            return false;
        }
        if (names.size() != sizes.size()) {
            log()->error("{} The size of the list containing the names is different from the size of the list containing the sizes.", logPrefix);
            return false;
        }
        clear();
        for (size_t i = 0; i < names.size(); i++) {
            if (sizes[i] <= 0) {
                log()->error("{} The size of the variable {} must be strictly positive.", logPrefix, names[i]);
                return false;
            }
            if (!addVariable(names[i], static_cast<std::size_t>(sizes[i]))) return false;
        }
        return true;
    }

    bool addVariable(const std::string& name, const std::size_t& size) noexcept
    {
        if (m_variables.count(name)) {
            log()->error("[VariablesHandler::addVariable] The variable named {} already exists.", name);
            return false;
        }
        VariableDescription d;
        d.offset = static_cast<std::ptrdiff_t>(m_numberOfVariables);
        d.size = static_cast<std::ptrdiff_t>(size);
        d.name = name;
        m_variables[name] = d;
        m_numberOfVariables += size;
        return true;
    }
    const VariableDescription& getVariable(const std::string& name) const noexcept
    {
        auto it = m_variables.find(name);
        return it == m_variables.end() ? m_invalid : it->second;
    }
    bool getVariable(const std::string& name, VariableDescription& description) const noexcept
    {
        description = getVariable(name);
        return description.isValid();
    }
    const std::size_t& getNumberOfVariables() const noexcept { return m_numberOfVariables; }
    void clear() noexcept
    {
        m_variables.clear();
        m_numberOfVariables = 0;
    }

private:
    std::unordered_map<std::string, VariableDescription> m_variables;
    std::size_t m_numberOfVariables{0};
    VariableDescription m_invalid{VariableDescription::InvalidVariable()};
};

/** OutputPort<VectorXd>: what QPInverseKinematics::advance reads every tick (QPInverseKinematics.cpp:450-451). */
struct WeightProvider {
    virtual ~WeightProvider() = default;
    virtual const VectorXd& getOutput() const = 0;
    virtual bool isOutputValid() const = 0;
};
using WeightProviderPort = WeightProvider;

struct ConstantWeightProvider : public WeightProvider {
    ConstantWeightProvider() = default;
    explicit ConstantWeightProvider(const VectorXd& w) : m_weight(w) {}
    /** key: weight */
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> handler)
    {
        constexpr auto logPrefix = "[ConstantWeightProvider::initialize]";
        auto ptr = handler.lock();
        if (!ptr) {
            log()->error("{} Invalid parameter handler.", logPrefix);
            return false;
        }
        if (!ptr->getParameter("weight", m_weight)) {
            log()->error("{} Unable to get the parameter named 'weight'.", logPrefix);
            return false;
        }
        return true;
    }
    const VectorXd& getOutput() const override { return m_weight; }
    bool isOutputValid() const override { return !m_weight.empty(); }

private:
    VectorXd m_weight;
};

/** Dense row-major matrix returned by getA() (stands in for Eigen::MatrixXd). */
struct MatrixXd {
    std::size_t rows{0}, cols{0};
    std::vector<double> data;
    void resize(std::size_t r, std::size_t c)
    {
        rows = r, cols = c;
        data.assign(r * c, 0.0);
    }
    double& operator()(std::size_t r, std::size_t c) { return data[r * cols + c]; }
    double operator()(std::size_t r, std::size_t c) const { return data[r * cols + c]; }
};

class LinearTask
{
protected:
    MatrixXd m_A;
    VectorXd m_b;
    std::string m_description{"Generic Linear Task Element"};

public:
    enum class Type { equality, inequality };
    virtual bool setVariablesHandler(const VariablesHandler&) { return true; }
    virtual bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler>) { return true; }
    virtual bool update() { return true; }
    const MatrixXd& getA() const { return m_A; }
    const VectorXd& getB() const { return m_b; }
    const std::string& getDescription() const { return m_description; }
    virtual std::size_t size() const = 0;
    virtual Type type() const = 0;
    virtual bool isValid() const = 0;
    /** A x - b  (LinearTask.cpp:61-66) */
    VectorXd getResidual(const VectorXd& solution) const
    {
        VectorXd r(m_b.size(), 0.0);
        for (std::size_t i = 0; i < m_A.rows; i++) {
            double v = -m_b[i];
            for (std::size_t j = 0; j < m_A.cols && j < solution.size(); j++) v += m_A(i, j) * solution[j];
            r[i] = v;
        }
        return r;
    }
    virtual ~LinearTask() = default;
};

/** Advanceable<EmptySignal, Output> = Source<Output> (Source.h:22, Advanceable.h:35-69). */
template <class Output> class Source
{
public:
    virtual ~Source() = default;
    virtual bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler>) { return true; }
    virtual bool advance() = 0;
    virtual const Output& getOutput() const = 0;
    virtual bool isOutputValid() const = 0;
};

template <class _Task, class _State> class ILinearTaskSolver : public Source<_State>
{
public:
    using Task = _Task;
    using State = _State;
    virtual bool addTask(std::shared_ptr<Task> task, const std::string& taskName, std::size_t priority,
                         std::shared_ptr<const WeightProviderPort> weightProvider = nullptr) = 0;
    virtual bool addTask(std::shared_ptr<Task> task, const std::string& taskName, std::size_t priority, const VectorXd& weight) = 0;
    virtual bool setTaskWeight(const std::string& taskName, std::shared_ptr<const WeightProviderPort> weightProvider) = 0;
    virtual bool setTaskWeight(const std::string& taskName, const VectorXd& weight) = 0;
    virtual std::weak_ptr<const WeightProviderPort> getTaskWeightProvider(const std::string& taskName) const = 0;
    virtual std::vector<std::string> getTaskNames() const = 0;
    virtual bool finalize(const VariablesHandler& handler) = 0;
    virtual std::weak_ptr<Task> getTask(const std::string& name) const = 0;
    virtual std::string toString() const = 0;
    virtual const VectorXd& getRawSolution() const = 0;
};

/** string -> builder map (Factory.h:25-80); registration through BLF_REGISTER_IK_TASK-style statics. */
template <class _Type> class Factory
{
public:
    using Builder = std::function<std::shared_ptr<_Type>()>;
    static std::unordered_map<std::string, Builder>& getMapFactory()
    {
        static std::unordered_map<std::string, Builder> map;
        return map;
    }
    static std::string registerBuilder(const std::string& idKey, Builder classBuilder)
    {
        getMapFactory().insert({idKey, std::move(classBuilder)});
        return idKey;
    }
    static std::shared_ptr<_Type> createInstance(const std::string& idKey)
    {
        auto it = getMapFactory().find(idKey);
        if (it == getMapFactory().end()) return nullptr;
        return it->second();
    }
};

}  // namespace System
}  // namespace BipedalLocomotion

#endif
