#pragma once
#include <condition_variable>
#include "mutex.hpp"
#include "thread_time.hpp"
namespace boost {
class condition_variable_any2 : public std::condition_variable_any {
public:
    template <class L> bool timed_wait(L &l, const system_time &t) { return wait_until(l, t) == std::cv_status::no_timeout; }
};
typedef condition_variable_any2 condition_variable_any;
class condition_variable2 : public std::condition_variable {
public:
    template <class L> bool timed_wait(L &l, const system_time &t) { return wait_until(l, t) == std::cv_status::no_timeout; }
};
typedef condition_variable2 condition_variable;
}
