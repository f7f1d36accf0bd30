#pragma once
#include <fstream>
#include "../filesystem.hpp"
namespace boost { namespace filesystem { using std::ifstream; using std::ofstream; using std::fstream; } }
