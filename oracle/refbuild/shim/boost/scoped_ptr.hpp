#pragma once
#include <memory>
namespace boost {
template <class T> class scoped_ptr {
    T *p;
    scoped_ptr(const scoped_ptr &); scoped_ptr &operator=(const scoped_ptr &);
public:
    explicit scoped_ptr(T *q = nullptr) : p(q) {}
    ~scoped_ptr() { delete p; }
    void reset(T *q = nullptr) { if (q != p) { delete p; p = q; } }
    T &operator*() const { return *p; }
    T *operator->() const { return p; }
    T *get() const { return p; }
    explicit operator bool() const { return p != nullptr; }
    bool operator!() const { return p == nullptr; }
};
}
