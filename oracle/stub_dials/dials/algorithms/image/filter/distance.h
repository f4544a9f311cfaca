// Stand-in for dials/algorithms/image/filter/distance.h (see ../../../../README.md).  Not DIALS
// code: the two-pass chessboard (Chebyshev) distance transform, written from its textbook
// definition.  dst(j, i) = max(|dj|, |di|) to the nearest pixel with src == value; every
// distance is at least height + width where there is no such pixel.
#pragma once
#include <algorithm>
#include <cstddef>

#include <scitbx/array_family/accessors/c_grid.h>
#include <scitbx/array_family/tiny_types.h>

namespace dials { namespace algorithms {
namespace af = scitbx::af;
using scitbx::af::int2;

template <typename InputType, typename OutputType>
void chebyshev_distance(const af::const_ref<InputType, af::c_grid<2>>& src, InputType value,
                        af::ref<OutputType, af::c_grid<2>> dst) {
    const std::size_t height = src.accessor()[0], width = src.accessor()[1];
    const OutputType far = (OutputType)(height + width);
    for (std::size_t j = 0; j < height; ++j)      // forward: north-west, north, north-east, west
        for (std::size_t i = 0; i < width; ++i) {
            if (src(j, i) == value) {
                dst(j, i) = 0;
                continue;
            }
            OutputType best = far;
            if (j > 0) {
                best = std::min(best, dst(j - 1, i));
                if (i > 0) best = std::min(best, dst(j - 1, i - 1));
                if (i + 1 < width) best = std::min(best, dst(j - 1, i + 1));
            }
            if (i > 0) best = std::min(best, dst(j, i - 1));
            dst(j, i) = best + 1;
        }
    for (std::size_t j = height; j-- > 0;)        // backward: south-east, south, south-west, east
        for (std::size_t i = width; i-- > 0;) {
            OutputType best = dst(j, i);
            if (j + 1 < height) {
                best = std::min(best, (OutputType)(dst(j + 1, i) + 1));
                if (i > 0) best = std::min(best, (OutputType)(dst(j + 1, i - 1) + 1));
                if (i + 1 < width) best = std::min(best, (OutputType)(dst(j + 1, i + 1) + 1));
            }
            if (i + 1 < width) best = std::min(best, (OutputType)(dst(j, i + 1) + 1));
            dst(j, i) = best;
        }
}
}}  // namespace dials::algorithms
