#pragma once
#include <system_error>
namespace boost { namespace system { using std::error_code; } }
