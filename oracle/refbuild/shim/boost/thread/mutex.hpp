#pragma once
#include <mutex>
#include <chrono>
#include "thread_time.hpp"
namespace boost {
using std::mutex;
template <class M> using lock_guard = std::lock_guard<M>;
template <class M> class unique_lock : public std::unique_lock<M> {
public:
    using std::unique_lock<M>::unique_lock;
    unique_lock(M &m, const system_time &t) : std::unique_lock<M>(m, t) {}
    bool timed_lock(const system_time &t) { return this->try_lock_until(t); }
};
class timed_mutex : public std::timed_mutex { public: typedef boost::unique_lock<timed_mutex> scoped_lock; bool timed_lock(const system_time &t) { return try_lock_until(t); } };
}
