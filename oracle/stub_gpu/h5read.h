/* Stand-in for h5read/include/h5read.h: the GPU sources only need the pixel type. */
#pragma once
#include <cstddef>
#include <cstdint>
#ifdef PIXEL_DATA_32BIT
using pixel_t = uint32_t;
#else
using pixel_t = uint16_t;
#endif
