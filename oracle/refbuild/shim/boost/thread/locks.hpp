#pragma once
#include "mutex.hpp"
