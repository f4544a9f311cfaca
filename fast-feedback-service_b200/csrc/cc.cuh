// cc.cuh -- connected components and per-spot reduction on the device.
//
// Reference: spotfinder/connected_components/connected_components.cc:17-139 (graph: one
// vertex per strong pixel in ascending linear index, edges to index+1 and index+width --
// the index+1 edge has NO row-wrap check, :62-70), :238-266 (2D spots), :270-470 (3D: plus
// an edge between identical linear indices of consecutive slices), and
// connected_components.hpp:31-268 (Reflection3D: bbox, centre of mass, peak).
//
// Design: strong pixels are sparse (1e3..1e4 of 1.8e7), so the strong bit-plane is first
// compacted into a sorted list of linear indices (ordered stream compaction: popcount per
// chunk -> scan -> extract).  Neighbours are then found in that sorted list (index+1 is the
// next entry; index+width / same index in the next slice by binary search), joined with a
// lock-free union-find whose root is always the minimum compact id -- i.e. the minimum
// (z, linear index) of the component, which is the canonical label and gives the same
// component order as boost::connected_components.  Per-spot sums are exact integers
// (atomics on the root's accumulator, warp-aggregated over runs of equal roots); the centre
// of mass is one fp64 divide per axis, stored as float like the reference does.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ffs_b200.h"

namespace ffsb {

struct SpotAcc {
    unsigned long long sum_i;    // sum I
    unsigned long long sum_xi;   // sum (2x+1) I
    unsigned long long sum_yi;   // sum (2y+1) I
    unsigned long long sum_zi;   // sum (2z+1) I
    unsigned long long peak;     // I << 32 | ~compact_id  (max: brightest, ties -> smallest id)
    uint32_t count;
    uint32_t x_min, x_max, y_min, y_max, z_min, z_max;
    uint32_t pad;
};
static_assert(sizeof(SpotAcc) == 72, "SpotAcc layout");

enum { CNT_STRONG = 0, CNT_STRONG_RAW, CNT_COMPONENTS, CNT_SPOTS_TOTAL, CNT_STRONG_FILTERED,
       CNT_SPOTS_FILTERED, CNT_STATUS, CNT_PAD, CNT_N };

constexpr int kChunkWords = 1024;  // strong-plane words per compaction block
static_assert(kChunkWords == 1 << 10, "set_strong (threshold.cuh) derives the chunk with a shift by 10");

struct CCArgs {
    int width, height, n_frames, slot0;
    const uint32_t* strong_bits;  // slot 0 plane
    size_t bits_pitch, bits_stride;
    const uint8_t* image;         // slot 0 frame
    size_t pitch, frame_stride;
    int bpp;
    int words_per_row, n_words, n_chunks;
    uint32_t cap;                 // compact-list capacity per slot
    uint32_t* px_index;           // [slot][cap] linear index, ascending
    uint32_t* px_value;           // [slot][cap]
    uint32_t* parent;             // [slot][cap]
    SpotAcc* acc;                 // [slot][cap]
    uint32_t* chunk_count;        // slot z at + z * chunk_stride: counts, then exclusive offsets
    size_t chunk_stride;          // uint32 units (the per-slot zero region, context.cuh d_zero)
    uint32_t* counts;             // [slot][CNT_N]
    ffs_spot* spots;              // [slot][cap_spots]
    uint32_t cap_spots;
    int row_wrap;
    uint32_t min_spot_size;
    float max_sep;
};

// ---- union-find (root = minimum id) ----------------------------------------------------
__device__ __forceinline__ uint32_t uf_find(const uint32_t* parent, uint32_t x) {
    uint32_t p = parent[x];
    while (p != x) {
        x = p;
        p = parent[x];
    }
    return x;
}

__device__ __forceinline__ void uf_union(uint32_t* parent, uint32_t a, uint32_t b) {
    for (;;) {
        a = uf_find(parent, a);
        b = uf_find(parent, b);
        if (a == b) return;
        if (a < b) {
            const uint32_t t = a;
            a = b;
            b = t;
        }  // a > b: hook a under b
        const uint32_t old = atomicMin(parent + a, b);
        if (old == a) return;
        a = old;  // someone hooked a meanwhile: merge that root with b
    }
}

// first position in [lo, hi) with list[pos] >= key
__device__ __forceinline__ uint32_t lower_bound(const uint32_t* list, uint32_t lo, uint32_t hi,
                                                uint32_t key) {
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (list[mid] < key) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ void acc_init(SpotAcc* a) {
    a->sum_i = a->sum_xi = a->sum_yi = a->sum_zi = 0ull;
    a->peak = 0ull;
    a->count = 0;
    a->x_min = a->y_min = a->z_min = 0xffffffffu;
    a->x_max = a->y_max = a->z_max = 0;
    a->pad = 0;
}

// Reflection3D::center_of_mass (hpp:71-103) + peak_centroid_distance (hpp:111-206).
__device__ __forceinline__ void acc_to_spot(const SpotAcc& a, uint64_t label, uint32_t peak_x,
                                            uint32_t peak_y, int32_t peak_z, ffs_spot* s) {
    s->label = label;
    s->x_min = a.x_min;
    s->x_max = a.x_max;
    s->y_min = a.y_min;
    s->y_max = a.y_max;
    s->z_min = (int32_t)a.z_min;
    s->z_max = (int32_t)a.z_max;
    s->num_pixels = a.count;
    s->peak_x = peak_x;
    s->peak_y = peak_y;
    s->peak_z = peak_z;
    s->peak_intensity = a.peak >> 32;
    s->sum_intensity = a.sum_i;
    if (a.sum_i == 0) {  // the reference throws here; cannot happen for threshold output
        s->com_x = s->com_y = s->com_z = s->peak_centroid_distance = __int_as_float(0x7fc00000);
        return;
    }
    const double wt = (double)a.sum_i;
    s->com_x = (float)(((double)a.sum_xi * 0.5) / wt);
    s->com_y = (float)(((double)a.sum_yi * 0.5) / wt);
    s->com_z = (float)(((double)a.sum_zi * 0.5) / wt);
    const float dx = __fsub_rn(__fadd_rn((float)peak_x, 0.5f), s->com_x);
    const float dy = __fsub_rn(__fadd_rn((float)peak_y, 0.5f), s->com_y);
    const float dz = __fsub_rn(__fadd_rn((float)peak_z, 0.5f), s->com_z);
    s->peak_centroid_distance =
      __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
}

// block-wide exclusive scan of one value per thread (blockDim.x == 256); returns the
// exclusive prefix, *total receives the block sum.
__device__ __forceinline__ uint32_t block_exclusive_scan_256(uint32_t v, uint32_t* total) {
    __shared__ uint32_t warp_sums[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, off);
        if (lane >= off) inc += t;
    }
    __syncthreads();  // protect warp_sums reuse across calls
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
        const uint32_t s = warp_sums[w];
        if (w < warp) base += s;
        tot += s;
    }
    *total = tot;
    return base + inc - v;
}

}  // namespace ffsb

#include "cc_kernels.cuh"
