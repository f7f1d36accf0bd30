#pragma once
#include <chrono>
namespace boost {
typedef std::chrono::system_clock::time_point system_time;
inline system_time get_system_time() { return std::chrono::system_clock::now(); }
namespace posix_time {
typedef boost::system_time ptime;
inline std::chrono::milliseconds milliseconds(long long ms) { return std::chrono::milliseconds(ms); }
}
}
