#pragma once
#include "mutex.hpp"
namespace boost {
using std::recursive_mutex;
class recursive_timed_mutex : public std::recursive_timed_mutex { public: typedef boost::unique_lock<recursive_timed_mutex> scoped_lock; bool timed_lock(const system_time &t) { return try_lock_until(t); } };
}
