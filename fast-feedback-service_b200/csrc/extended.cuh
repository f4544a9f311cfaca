// extended.cuh -- DispersionExtendedThreshold (baseline/spotfinder/baseline.cpp:325-813).
//
//   pass 1   compute_dispersion_threshold (:415-475): D = mask && m >= min_count && a > c
//            -> resolve_kernel<.., EXT_PASS1 = true> over every pixel (exact integer window
//               gather + fp64 test), strong plane = the ext1 bit-plane.
//   erosion  erode_dispersion_mask (:552-571): B = mask ? !(D && chebyshev_distance >= 3) : 0.
//            distance >= 3 to the nearest false pixel == "all of the in-image 5x5
//            neighbourhood is true", so this is pure bit-plane logic (erode_kernel).  Masked
//            pixels have D = 0 and therefore erode their neighbours, as in the reference CPU
//            code (the reference GPU kernel skips them: spotfinder/kernels/erosion.cu:101-105).
//   pass 2   compute_sat(table, src, B) + compute_final_threshold (:580-645): 11x11 window
//            over B-valid pixels; strong iff the pixel is in the eroded core (!B), v > threshold
//            and v >= mean + nsig_s*sqrt(mean).  Only core pixels can be strong and they are
//            sparse, so erode_kernel emits them as a candidate list and final_kernel gathers
//            the 11x11 window for those alone -- no second box filter over the frame.
#pragma once
#include "threshold.cuh"

namespace ffsb {

struct ErodeArgs {
    int width, height, n_frames, slot0;
    int words_per_row;
    uint32_t inv_wpr;            // floor(2^32 / words_per_row) + 1, or 0: divide
    const uint32_t* mask_bits;   // validity plane of the dataset
    const uint32_t* d_bits;      // pass-1 plane(s) (ext1)
    uint32_t* b_bits;            // output plane(s) (ext2): valid and not in the core
    size_t bits_stride;          // bytes per plane
    uint32_t* cand;              // per slot: core pixels  col | row << 13
    unsigned int* cand_count;
    unsigned int cand_cap;
    size_t zstride_words;        // counters of consecutive slots are this many uint32 apart
};

// One thread per 32-bit word of the plane.  grid (blocks, n_frames).
__global__ void __launch_bounds__(256) erode_kernel(const ErodeArgs a) {
    const int z = a.slot0 + blockIdx.y;
    const uint32_t* D = reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(a.d_bits) +
                                                          (size_t)z * a.bits_stride);
    uint32_t* B = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(a.b_bits) + (size_t)z * a.bits_stride);
    const int wpr = a.words_per_row;
    const int n_words = wpr * a.height;
    const int last_word = (a.width - 1) >> 5;
    const uint32_t last_valid = (a.width & 31) ? ((1u << (a.width & 31)) - 1u) : 0xffffffffu;
    const int lane = threadIdx.x & 31;
    // warp-uniform trip count: the candidate push below is a warp-aggregated append (one
    // atomic per warp and word row instead of one per core pixel on a single address)
    for (int t0 = blockIdx.x * blockDim.x + (threadIdx.x & ~31); t0 < n_words; t0 += gridDim.x * blockDim.x) {
        const int t = t0 + lane;
        const int row = a.inv_wpr ? (int)__umulhi((uint32_t)t, a.inv_wpr) : t / wpr, wi = t - row * wpr;
        const bool in_plane = t < n_words;
        uint32_t core = 0;
        if (in_plane && wi > last_word) {
            B[t] = 0;
        } else if (in_plane && (D[t] & (wi == last_word ? last_valid : 0xffffffffu)) == 0u) {
            // no D pixel in this word, so no core pixel either (the core is a subset of D): the 5x5
            // neighbourhood is not read.  D is sparse -- on background nearly every word ends here.
            B[t] = a.mask_bits[t] & (wi == last_word ? last_valid : 0xffffffffu);
        } else if (in_plane) {
            core = 0xffffffffu;
            // out-of-image pixels are ignored by the distance transform: they read as "true"
            auto word = [&](int r, int w) -> uint32_t {
                if (r < 0 || r >= a.height || w < 0 || w > last_word) return 0xffffffffu;
                uint32_t v = D[(size_t)r * wpr + w];
                if (w == last_word) v |= ~last_valid;
                return v;
            };
#pragma unroll
            for (int dr = -2; dr <= 2; ++dr) {
                const int r = row + dr;
                const uint32_t c = word(r, wi), l = word(r, wi - 1), rt = word(r, wi + 1);
                // bit i of the result = AND of bits i-2 .. i+2 of the row
                uint32_t h = c;
                h &= (c << 1) | (l >> 31);
                h &= (c << 2) | (l >> 30);
                h &= (c >> 1) | (rt << 31);
                h &= (c >> 2) | (rt << 30);
                core &= h;
            }
            const uint32_t m = a.mask_bits[t];
            const uint32_t valid_cols = wi == last_word ? last_valid : 0xffffffffu;
            core &= m & valid_cols;  // the centre itself is in D (dr = 0), D implies mask
            B[t] = m & ~core & valid_cols;
        }
        // core pixels are the only possible strong pixels of the final pass
        if (__any_sync(0xffffffffu, core != 0)) {
            const uint32_t cnt = __popc(core);
            uint32_t incl = cnt;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const uint32_t v = __shfl_up_sync(0xffffffffu, incl, off);
                if (lane >= off) incl += v;
            }
            uint32_t base = 0;
            if (lane == 31) base = atomicAdd(a.cand_count + (size_t)z * a.zstride_words, incl);
            base = __shfl_sync(0xffffffffu, base, 31);
            uint32_t slot = base + incl - cnt;
            for (uint32_t bits = core; bits; bits &= bits - 1, ++slot) {
                const int b = __ffs(bits) - 1;
                if (slot < a.cand_cap)
                    a.cand[(size_t)z * a.cand_cap + slot] = (uint32_t)(wi * 32 + b) | ((uint32_t)row << 13);
            }
        }
    }
}

// compute_final_threshold for one pixel (baseline.cpp:630-643).  valid = the B plane of this
// frame (validity of the second summed-area table), mask = the dataset's validity plane.
template <typename PixelT, int KH>
__device__ __forceinline__ bool resolve_final_pixel(const ThreshArgs& a, const uint8_t* frame,
                                                    const uint8_t* bplane, const uint8_t* mask, int col,
                                                    int row) {
    const size_t boff = (size_t)row * a.valid_pitch + (col >> 3);
    if (!((mask[boff] >> (col & 7)) & 1u)) return false;    // mask[k]
    if ((bplane[boff] >> (col & 7)) & 1u) return false;     // dispersion_mask = !dst[k]
    const int x0 = max(col - KH, 0), x1 = min(col + KH, a.width - 1);
    const int y0 = max(row - KH, 0), y1 = min(row + KH, a.height - 1);
    uint32_t m = 0;
    unsigned long long x = 0;
    if (x1 - x0 == 2 * KH && y1 - y0 == 2 * KH) {
        // window inside the image (nearly every core pixel): no per-pixel bound test, 32-bit sum
        // ((2 KH + 1)^2 counted pixels of at most 2^24 - 1 stay below 2^32), count by popc where no
        // pixel can be left out (16-bit data without a trusted range)
        constexpr int W = 2 * KH + 1;
        static_assert(W * W < 256, "32-bit window sum");
        const bool all_count = sizeof(PixelT) == 2 && a.vlim >= 0xffffu;
        uint32_t xs = 0;
        for (int yy = y0; yy <= y1; ++yy) {
            const PixelT* prow = reinterpret_cast<const PixelT*>(frame + (size_t)yy * a.pitch) + x0;
            const uint8_t* vrow = bplane + (size_t)yy * a.valid_pitch + (x0 >> 3);
            uint32_t bits = (uint32_t)vrow[0] | (uint32_t)vrow[1] << 8;
            if ((x0 & 7) + W > 16) bits |= (uint32_t)vrow[2] << 16;
            bits = (bits >> (x0 & 7)) & ((1u << W) - 1u);
            if (all_count) {
                m += __popc(bits);
#pragma unroll
                for (int k = 0; k < W; ++k) xs += ((bits >> k) & 1u) ? (uint32_t)prow[k] : 0u;
            } else {
#pragma unroll
                for (int k = 0; k < W; ++k) {
                    const uint32_t pv = prow[k];
                    const bool take = ((bits >> k) & 1u) && pv <= a.vlim;
                    m += take ? 1u : 0u;
                    xs += take ? pv : 0u;
                }
            }
        }
        x = xs;
    } else {
        for (int yy = y0; yy <= y1; ++yy) {
            const PixelT* prow = reinterpret_cast<const PixelT*>(frame + (size_t)yy * a.pitch);
            const uint8_t* vrow = bplane + (size_t)yy * a.valid_pitch;
            const int byte0 = x0 >> 3;
            uint32_t bits = vrow[byte0];
            if ((x1 >> 3) > byte0) bits |= (uint32_t)vrow[byte0 + 1] << 8;
            if ((x1 >> 3) > byte0 + 1) bits |= (uint32_t)vrow[byte0 + 2] << 16;
            bits >>= (x0 & 7);
#pragma unroll
            for (int k = 0; k < 2 * KH + 1; ++k) {
                const int xx = x0 + k;
                if (xx <= x1) {
                    const uint32_t pv = prow[xx];
                    if (((bits >> k) & 1u) && pv <= a.vlim) {
                        m += 1;
                        x += pv;
                    }
                }
            }
        }
    }
    const uint32_t vc = reinterpret_cast<const PixelT*>(frame + (size_t)row * a.pitch)[col];
    if (vc > a.tmax) return false;                          // above the trusted range: mask'[k] is false
    const double vd = (double)vc;
    if (!(vd > a.threshold)) return false;                  // global_mask
    // double mean = (m >= 2 ? (x / m) : 0);  local_mask = src[k] >= mean + nsig_s_ * sqrt(mean)
    const double mean = m >= 2 ? __ddiv_rn((double)x, (double)m) : 0.0;
    // with a gain map (baseline.cpp:706-707): src[k] >= mean + nsig_s_ * std::sqrt(gain[k] * mean)
    const double var = a.gain ? __dmul_rn(a.gain[(size_t)row * a.width + col], mean) : mean;
    return vd >= __dadd_rn(mean, __dmul_rn(a.nsig_s, __dsqrt_rn(var)));
}

// grid (blocks, n_frames): one thread per core pixel (or per pixel if the list overflowed)
template <typename PixelT, int KH>
__global__ void __launch_bounds__(256) final_kernel(const ThreshArgs a, const uint8_t* mask_bits) {
    const int z = a.slot0 + blockIdx.y;
    const uint8_t* frame = a.image + (size_t)z * a.frame_stride;
    const uint8_t* bplane = a.valid_bits + (size_t)z * a.valid_frame_stride;
    const unsigned n_cand = a.cand_count[(size_t)z * a.zstride_words];
    const bool brute = a.force_brute || n_cand > a.cand_cap;
    const long long total = brute ? (long long)a.width * a.height : (long long)n_cand;
    const uint32_t* list = a.cand + (size_t)z * a.cand_cap;
    unsigned found = 0;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        int col, row;
        if (brute) {
            row = (int)(t / a.width);
            col = (int)(t % a.width);
        } else {
            const uint32_t e = list[t];
            col = e & 0x1fff;
            row = e >> 13;
        }
        if (resolve_final_pixel<PixelT, KH>(a, frame, bplane, mask_bits, col, row)) {
            set_strong(a, z, col, row);
            ++found;
        }
    }
    for (int off = 16; off > 0; off >>= 1) found += __shfl_down_sync(0xffffffffu, found, off);
    if ((threadIdx.x & 31) == 0 && found) atomicAdd(a.strong_count + (size_t)z * a.zstride_words, found);
}

}  // namespace ffsb
