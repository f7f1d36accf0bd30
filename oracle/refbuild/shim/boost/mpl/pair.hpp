#pragma once
#include "vector.hpp"
