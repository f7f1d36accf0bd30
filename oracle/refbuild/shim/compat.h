// Force-included into every reference TU: headers Boost used to drag in transitively.
#pragma once
#include <unistd.h>
#include <cassert>
#include <functional>
#include <type_traits>
#include <cstring>
#include <cstdlib>
#include <memory>
namespace boost { using std::is_same; template <class T> inline void hash_combine(std::size_t &seed, const T &v) { seed ^= std::hash<T>()(v) + 0x9e3779b9 + (seed << 6) + (seed >> 2); } }
