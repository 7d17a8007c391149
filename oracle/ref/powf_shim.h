// TEST INFRASTRUCTURE (oracle). libstdc++ 13 has no std::powf; the reference calls it
// (src/shader/shader.cpp:41-43). Force-included when compiling the reference sources in place.
#pragma once
#include <cmath>
namespace std { using ::powf; }
