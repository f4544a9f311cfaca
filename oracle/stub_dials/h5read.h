/* Stand-in for h5read/include/h5read.h: the pixel type and the image_t view of the baseline C ABI. */
#pragma once
#include <cstddef>
#include <cstdint>
#ifdef PIXEL_DATA_32BIT
typedef uint32_t image_t_type;
#else
typedef uint16_t image_t_type;
#endif
typedef struct image_t {
    image_t_type* data;
    uint8_t* mask;
    size_t slow;
    size_t fast;
} image_t;
