#pragma once
#include <thread>
#include <chrono>
#include <stdexcept>
#include "mutex.hpp"
namespace boost {
struct thread_resource_error : std::runtime_error { thread_resource_error() : std::runtime_error("thread resource error") {} };
struct thread_interrupted {};
class thread : public std::thread {
public:
    using std::thread::thread;
    static unsigned hardware_concurrency() { return std::thread::hardware_concurrency(); }
};
namespace this_thread { using std::this_thread::get_id; using std::this_thread::yield; template <class D> void sleep(const D &d) { std::this_thread::sleep_for(d); } }
template <class T> class thread_specific_ptr {
    static thread_local T *p;   // one per instantiating type is enough for the two uses in thread.cpp
public:
    thread_specific_ptr() {}
    explicit thread_specific_ptr(void (*)(T *)) {}
    T *get() const { return p; }
    T *operator->() const { return p; }
    T &operator*() const { return *p; }
    void reset(T *q = nullptr) { p = q; }
};
template <class T> thread_local T *thread_specific_ptr<T>::p = nullptr;
}
