// trusted.cuh -- the trusted range (reader.get_trusted_range()[1]; spotfinder/spotfinder.cc:482,
// 868,879 hands it to the kernels, spotfinder/kernels/thresholding.cu:208-209 rejects a centre
// pixel above it).  Here a mask-valid pixel above trusted_max is treated exactly like a masked
// one, for the centre AND for every window it lies in (identical to running the
// CPU algorithm with mask' = mask && v <= trusted_max).
//
// The exact per-pixel routines apply that rule themselves (ThreshArgs::tmax / vlim).  The
// 16-bit streaming loop, however, sums such a pixel clamped instead of leaving it out, so its
// pruning bounds are not bounds for the windows around it.  Those pixels are rare; whoever meets
// one as a centre pixel lists it (note_bad_pixel) and the kernels below repair the result:
//   * trusted_fixup_kernel   (dispersion): every pixel of the 7x7 neighbourhood of a listed pixel
//     is decided again by the exact routine and its strong bit set or cleared accordingly
//     (idempotent: the counters follow the bit transitions, so overlapping neighbourhoods and
//     pixels the main kernel already decided correctly are harmless);
//   * trusted_clear_kernel   (dispersion_extended): a listed pixel is not a valid background
//     pixel of the final pass -- its bit in the eroded plane is cleared (baseline.cpp:568
//     `dst = mask ? ... : false` with the per-frame mask').
// If a list overflows, the whole frame is walked instead (exact, slow).
#pragma once
#include "threshold.cuh"

namespace ffsb {

// grid (blocks, n_frames)
template <typename PixelT, int KH>
__global__ void __launch_bounds__(256) trusted_fixup_kernel(const ThreshArgs a) {
    const int z = a.slot0 + blockIdx.y;
    const unsigned n_bad = a.bad_count[(size_t)z * a.zstride_words];
    if (n_bad == 0) return;
    constexpr int KS = 2 * KH + 1;
    const uint8_t* frame = a.image + (size_t)z * a.frame_stride;
    const uint8_t* valid = a.valid_bits + (size_t)z * a.valid_frame_stride;
    const bool whole = n_bad > a.bad_cap;
    const long long total = whole ? (long long)a.width * a.height : (long long)n_bad * KS * KS;
    const uint32_t* list = a.bad + (size_t)z * a.bad_cap;
    int delta = 0;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        int col, row;
        if (whole) {
            row = (int)(t / a.width);
            col = (int)(t % a.width);
        } else {
            const uint32_t e = list[t / (KS * KS)];
            const int k = (int)(t % (KS * KS));
            col = (int)(e & 0x1fff) + k % KS - KH;
            row = (int)(e >> 13) + k / KS - KH;
            if (col < 0 || row < 0 || col >= a.width || row >= a.height) continue;
        }
        const bool strong = resolve_pixel<PixelT, KH, false>(a, frame, valid, col, row);
        const size_t word = (size_t)row * (a.strong_pitch >> 2) + (col >> 5);
        uint32_t* w = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(a.strong_bits) +
                                                  (size_t)z * a.strong_frame_stride) + word;
        const uint32_t bit = 1u << (col & 31);
        if (strong) {
            if (!(atomicOr(w, bit) & bit)) {
                if (a.chunk_count) atomicAdd(a.chunk_count + (size_t)z * a.zstride_words + (word >> 10), 1u);
                ++delta;
            }
        } else if (*w & bit) {
            if (atomicAnd(w, ~bit) & bit) {
                if (a.chunk_count) atomicSub(a.chunk_count + (size_t)z * a.zstride_words + (word >> 10), 1u);
                --delta;
            }
        }
    }
    if (delta) atomicAdd(a.strong_count + (size_t)z * a.zstride_words, (unsigned)delta);
}

// grid (blocks, n_frames).  b_bits: the eroded planes (ext2), one per slot.
template <typename PixelT>
__global__ void __launch_bounds__(256) trusted_clear_kernel(const ThreshArgs a, uint32_t* b_bits, size_t bits_stride) {
    const int z = a.slot0 + blockIdx.y;
    const unsigned n_bad = a.bad_count[(size_t)z * a.zstride_words];
    if (n_bad == 0) return;
    uint32_t* B = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(b_bits) + (size_t)z * bits_stride);
    const int wpr = (int)(a.strong_pitch >> 2);
    if (n_bad <= a.bad_cap) {
        const uint32_t* list = a.bad + (size_t)z * a.bad_cap;
        for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < n_bad; t += gridDim.x * blockDim.x) {
            const int col = list[t] & 0x1fff, row = list[t] >> 13;
            atomicAnd(B + (size_t)row * wpr + (col >> 5), ~(1u << (col & 31)));
        }
        return;
    }
    const uint8_t* frame = a.image + (size_t)z * a.frame_stride;
    const long long total = (long long)a.width * a.height;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        const int row = (int)(t / a.width), col = (int)(t % a.width);
        const uint32_t v = reinterpret_cast<const PixelT*>(frame + (size_t)row * a.pitch)[col];
        if (v > a.tmax) atomicAnd(B + (size_t)row * wpr + (col >> 5), ~(1u << (col & 31)));
    }
}

}  // namespace ffsb
