// TEST INFRASTRUCTURE (oracle/): stand-in for <mitsuba/core/object.h>: a reference-counted
// base, ref<T>, TraversalCallback and ParamFlags with the shapes src/emitters/sunsky.cpp
// uses (reference: include/mitsuba/core/object.h).
#pragma once
#include <mitsuba/core/class.h>
#include <atomic>
#include <string>
#include <vector>

NAMESPACE_BEGIN(mitsuba)

class Object {
public:
    Object() { }
    Object(const Object &) { }
    virtual ~Object() { }
    void inc_ref() const { ++m_rc; }
    void dec_ref() const { if (--m_rc == 0) delete this; }
private:
    mutable std::atomic<int> m_rc { 0 };
};

template <typename T> class ref {
public:
    ref() { }
    ref(T *p) : m_p(p) { if (m_p) ((Object *) m_p)->inc_ref(); }
    ref(const ref &r) : m_p(r.m_p) { if (m_p) ((Object *) m_p)->inc_ref(); }
    ref(ref &&r) noexcept : m_p(r.m_p) { r.m_p = nullptr; }
    ~ref() { if (m_p) ((Object *) m_p)->dec_ref(); }
    ref &operator=(const ref &r) {
        if (r.m_p) ((Object *) r.m_p)->inc_ref();
        if (m_p) ((Object *) m_p)->dec_ref();
        m_p = r.m_p;
        return *this;
    }
    T *operator->() const { return m_p; }
    T &operator*() const { return *m_p; }
    T *get() const { return m_p; }
    operator T *() const { return m_p; }
    explicit operator bool() const { return m_p != nullptr; }
private:
    T *m_p = nullptr;
};

enum class ParamFlags : uint32_t { Differentiable = 0x0, NonDifferentiable = 0x1, Discontinuous = 0x2 };
constexpr uint32_t operator+(ParamFlags f) { return (uint32_t) f; }

/// Records the names handed to traverse() so the harness can check the property surface
class TraversalCallback {
public:
    template <typename T> void put_parameter(const std::string &name, T &, uint32_t) { names.push_back(name); }
    void put_object(const std::string &name, Object *, uint32_t) { names.push_back(name); }
    std::vector<std::string> names;
};

NAMESPACE_END(mitsuba)
