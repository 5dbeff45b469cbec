/**
 * Contacts.h -- Contacts::PlannedContact and Contacts::ContactList, as far as Planners::SwingFootPlanner needs them.
 *
 * Same names and semantics as
 *   src/Contacts/include/BipedalLocomotion/Contacts/Contact.h       (PlannedContact: pose, activationTime, deactivationTime, name, index)
 *   src/Contacts/include/BipedalLocomotion/Contacts/ContactList.h   (addContact, getActiveContact, getNextContact, getPresentContact, ...)
 *   src/Contacts/src/ContactList.cpp:52-96, 215-265
 * with SE3d of Core.h in place of manif::SE3d.  Contacts are kept sorted by activation time and may not overlap.
 */
#ifndef BLF_B200_CONTACTS_H
#define BLF_B200_CONTACTS_H

#include <algorithm>
#include <chrono>
#include <string>
#include <vector>

#include <BipedalLocomotion/Core.h>

namespace BipedalLocomotion
{
namespace Contacts
{

struct PlannedContact {
    SE3d pose;
    std::chrono::nanoseconds activationTime{0}, deactivationTime{std::chrono::nanoseconds::max()};
    std::string name{"Contact"};
    int index{-1};
    bool isContactActive(const std::chrono::nanoseconds& t) const { return activationTime <= t && t < deactivationTime; }
};

class ContactList
{
    std::vector<PlannedContact> m_contacts;
    std::string m_defaultName{"ContactList"};
    int m_defaultIndex{-1};

public:
    using const_iterator = std::vector<PlannedContact>::const_iterator;

    void setDefaultName(const std::string& name) { m_defaultName = name; }
    const std::string& defaultName() const { return m_defaultName; }
    void setDefaultIndex(int index) { m_defaultIndex = index; }

    /** ContactList.cpp:52-81: activation <= deactivation and no overlap with a contact already in the list */
    bool addContact(const PlannedContact& c)
    {
        constexpr auto errorPrefix = "[ContactList::addContact]";
        if (c.activationTime > c.deactivationTime) {
            log()->error("{} The activation time cannot be greater than the deactivation time.", errorPrefix);
            return false;
        }
        for (const auto& o : m_contacts)
            if (!(c.deactivationTime <= o.activationTime || o.deactivationTime <= c.activationTime)) {
                log()->error("{} Failed to insert new element.", errorPrefix);
                return false;
            }
        const auto at = std::upper_bound(m_contacts.begin(), m_contacts.end(), c,
                                         [](const PlannedContact& a, const PlannedContact& b) { return a.activationTime < b.activationTime; });
        m_contacts.insert(at, c);
        return true;
    }
    bool addContact(const SE3d& transform, const std::chrono::nanoseconds& activationTime, const std::chrono::nanoseconds& deactivationTime)
    {
        PlannedContact c;
        c.pose = transform, c.activationTime = activationTime, c.deactivationTime = deactivationTime, c.name = m_defaultName, c.index = m_defaultIndex;
        return addContact(c);
    }
    std::size_t size() const { return m_contacts.size(); }
    void clear() { m_contacts.clear(); }
    const_iterator begin() const { return m_contacts.begin(); }
    const_iterator end() const { return m_contacts.end(); }
    const_iterator cbegin() const { return m_contacts.cbegin(); }
    const_iterator cend() const { return m_contacts.cend(); }
    const_iterator firstContact() const { return m_contacts.cbegin(); }
    const_iterator lastContact() const { return m_contacts.empty() ? m_contacts.cend() : m_contacts.cend() - 1; }
    const PlannedContact& operator[](std::size_t i) const { return m_contacts[i]; }

    /** the last contact whose activation time is not after `time` (:215-233) */
    const_iterator getPresentContact(const std::chrono::nanoseconds& time) const
    {
        for (auto it = m_contacts.crbegin(); it != m_contacts.crend(); ++it)
            if (it->activationTime <= time) return (it + 1).base();
        return m_contacts.cend();
    }
    /** the contact with activationTime <= time < deactivationTime (:236-254) */
    const_iterator getActiveContact(const std::chrono::nanoseconds& time) const
    {
        for (auto it = m_contacts.crbegin(); it != m_contacts.crend(); ++it)
            if (it->activationTime <= time && it->deactivationTime > time) return (it + 1).base();
        return m_contacts.cend();
    }
    /** the first contact that becomes active after `time` (:256-265) */
    const_iterator getNextContact(const std::chrono::nanoseconds& time) const
    {
        return std::find_if(m_contacts.cbegin(), m_contacts.cend(), [time](const PlannedContact& a) { return a.activationTime > time; });
    }
};

}  // namespace Contacts
}  // namespace BipedalLocomotion

#endif
