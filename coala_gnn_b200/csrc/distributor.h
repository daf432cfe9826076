// internal C++ view of the distributor handle (the C ABI is in include/coala_b200.h)
#pragma once
#include <cstdint>
#include "../../include/coala_b200.h"

// device int32 copy of color[N] (NULL when the distributor carries no colour information)
const int32_t* coala_nd_color_dev_i32(coala_nd* h);
