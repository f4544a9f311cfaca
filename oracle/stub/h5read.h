/* Stub standing in for the reference's h5read/include/h5read.h when compiling
 * baseline/spotfinder/standalone.cc in place (it includes <h5read.h> but uses
 * nothing from it).  Written for this repo; contains no reference code. */
#pragma once
#include <cstddef>
#include <cstdint>
