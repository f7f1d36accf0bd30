#pragma once
#include <cstddef>
namespace boost {
template <class T> class scoped_array {
    T *p;
    scoped_array(const scoped_array &); scoped_array &operator=(const scoped_array &);
public:
    explicit scoped_array(T *q = nullptr) : p(q) {}
    ~scoped_array() { delete[] p; }
    void reset(T *q = nullptr) { if (q != p) { delete[] p; p = q; } }
    T &operator[](std::ptrdiff_t i) const { return p[i]; }
    T *get() const { return p; }
};
}
