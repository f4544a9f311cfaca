// threshold.cuh -- DIALS dispersion threshold (DispersionThreshold,
// baseline/spotfinder/standalone.cc:27-206) as hand-written sm_100a kernels.
//
// Two kernels per batch of frames:
//
//  1. prefilter_kernel  (HBM-streaming; this is the kernel the roofline is quoted on)
//     A *warp task* is a strip of OUT_COLS output columns x tile_rows output rows of one
//     frame.  Every warp owns its task end to end: its own TMA pipeline (STAGES stages of
//     SR rows: image tile + validity-bit tile, zero-filled out of bounds, which is exactly
//     the reference's window clipping), its own mbarriers, its own ring of the last KS
//     rows of packed words in shared memory and its own candidate queue; warps never
//     synchronise with each other.  Lane l owns PPL adjacent columns; the first and last
//     lane are halo lanes.
//     Window sums are exact integers.  For 16-bit pixels count and sum share one 32-bit
//     word  w = valid ? (1 << 24) | v : 0  (49 * 65535 < 2^24), so ONE separable box sum
//     (vertical running sum, horizontal sum via warp shuffles) gives both m = H >> 24 and
//     x = H & 0xffffff.  32-bit pixels use a sum plane and a byte-packed count plane
//     (dp4a horizontal sums).  Each pixel is then put through a cheap NECESSARY condition
//     for the signal test  b > d  (b = m*v - x, d = nsig_s*sqrt(x*m)); it is conservative,
//     never exact.  Survivors (~0.5 % of background pixels) are appended to a per-frame
//     candidate list in global memory (32 at a time per warp).
//
//  2. resolve_kernel  (one thread per candidate)
//     Re-gathers the clipped window from the L2-resident frame as exact integers
//     (m, x as uint32, y = sum of squares as uint64) and evaluates the reference's test in
//     fp64 with round-to-nearest intrinsics in the reference's operation order
//     (standalone.cc:164-171).  Strong pixels are OR-ed into a pre-zeroed 1-bit plane.
//     If a frame's candidate list overflowed, every pixel of that frame is resolved this
//     way instead (slow, exact) -- there is no approximate path.
#pragma once
#include "ptx.cuh"

namespace ffsb {

struct ThreshArgs {
    int width, height;
    int n_frames;
    int slot0;        // first device slot of the batch (TMA z coordinate offset)
    int tile_rows;    // output rows per warp task
    int n_strips;     // ceil(width / OUT_COLS)
    int n_row_tiles;  // ceil(height / tile_rows)
    const uint8_t* image;  // slot 0 of the device frame ring, pitched
    size_t pitch;          // bytes per image row
    size_t frame_stride;   // bytes per frame
    const uint8_t* valid_bits;  // 1 bit per pixel, little-endian within a byte
    size_t valid_pitch;         // bytes per row
    size_t valid_frame_stride;  // 0: one plane shared by all frames (the mask)
    uint32_t* strong_bits;      // output plane(s), pre-zeroed
    size_t strong_pitch;        // bytes per row (multiple of 4)
    size_t strong_frame_stride; // bytes per frame plane
    unsigned int* strong_count; // per slot (slot z at + z * zstride_words), pre-zeroed
    uint32_t* cand;             // per slot: cand_cap entries  col | row << 13
    unsigned int* cand_count;   // per slot, pre-zeroed (may run past cand_cap)
    unsigned int cand_cap;
    int force_brute;            // debug: resolve every pixel exactly
    double nsig_b, nsig_s, threshold;
    int min_count;
    float prefilter_k;  // nsig_s^2 * (1 - 1e-4): conservative
    float sig_k;        // nsig_s * (1 - 1e-4)            (fused kernel)
    int thr_floor;      // floor(threshold): v > threshold <=> v > thr_floor for integer v
    uint32_t* chunk_count;  // per slot: strong pixels per 1024-word chunk of the strong plane (the
    int n_chunks;           // connected-components compaction's first pass); NULL = do not count
    const double* gain;     // optional per-pixel gain map, width*height row-major (threshold_w_gain,
                            // baseline.cpp:187-250); NULL = the plain test
    float g_lo, g_hi;       // extended first pass: bracket of the gain over the valid pixels (1, 1 without a map)
    const uint8_t* mcount;  // per pixel: mask-valid pixels in its image-clipped 7x7 window
    size_t m_pitch;         // (group_info_kernel); 49 <=> window inside the image and fully valid
    // Trusted range (reader.get_trusted_range()[1], spotfinder.cc:482,868,879): a pixel
    // above trusted_max is treated exactly like a masked one -- never strong, never in a window sum.
    uint32_t tmax;          // a centre pixel above this is invalid (0xffffffff = no trusted range)
    uint32_t vlim;          // a window pixel is counted iff v <= vlim = min(tmax, 2^24 - 1)  (standalone.cc:90)
    uint32_t* bad;          // per slot: bad_cap entries col | row << 13 of mask-valid pixels above tmax,
    unsigned int* bad_count;  // appended by whoever meets one as a centre pixel (may run past bad_cap)
    unsigned int bad_cap;
    // strong plane, chunk counts and the four counters of a slot live in ONE per-slot region that a
    // single memset clears (context.cuh d_zero): consecutive slots are zstride_words uint32 apart
    size_t zstride_words;
};

// A mask-valid centre pixel above the trusted maximum: remember it for trusted_fixup_kernel.
__device__ __forceinline__ void note_bad_pixel(const ThreshArgs& a, int z, int col, int row) {
    if (!a.bad_count) return;
    const unsigned k = atomicAdd(a.bad_count + (size_t)z * a.zstride_words, 1u);
    if (k < a.bad_cap) a.bad[(size_t)z * a.bad_cap + k] = (uint32_t)col | ((uint32_t)row << 13);
}

template <typename PixelT, int KH_, int PPL_, int SR_>
struct ThreshGeo {
    static constexpr int KH = KH_;
    static constexpr int PPL = PPL_;
    static constexpr int SR = SR_;                        // rows per TMA stage
    static constexpr int KS = 2 * KH + 1;                 // ring depth
    static constexpr int HL = (KH + PPL - 1) / PPL;       // halo lanes per side
    static constexpr int LOAD_COLS = 32 * PPL;
    static constexpr int OUT_COLS = (32 - 2 * HL) * PPL;
    static constexpr int VPAD = (8 - (HL * PPL) % 8) % 8; // validity tile starts VPAD px earlier
    // TMA box starts must be 16-byte aligned in global memory (measured: a misaligned start
    // raises 'illegal instruction'), so the validity box starts up to 15 bytes early.
    static constexpr int VROW_BYTES = ((VPAD + LOAD_COLS + 7) / 8 + 15 + 15) / 16 * 16;
    static constexpr int IMG_STAGE_BYTES = SR * LOAD_COLS * (int)sizeof(PixelT);
    static constexpr int VALID_STAGE_BYTES = (SR * VROW_BYTES + 127) / 128 * 128;
    static constexpr int STAGE_BYTES = IMG_STAGE_BYTES + VALID_STAGE_BYTES;
    static constexpr int TX_BYTES = IMG_STAGE_BYTES + SR * VROW_BYTES;
    static constexpr int RING_PLANES = PPL / 4;           // 16-byte planes per ring row
    static constexpr int RING_ROW_BYTES = RING_PLANES * 512;
    // 32-bit pixels keep one extra plane: byte-packed counts + flags
    static constexpr int RING_BYTES = KS * RING_ROW_BYTES + (sizeof(PixelT) == 4 ? KS * 256 : 0);
    static constexpr int QUEUE_CAP = 32 + OUT_COLS;       // entries; a row adds <= OUT_COLS
    static constexpr int QUEUE_BYTES = (QUEUE_CAP * 4 + 127) / 128 * 128;
    static_assert(HL == 1, "one halo lane per side");
    static_assert(OUT_COLS % 8 == 0, "strips must start on a validity byte");
    static_assert(IMG_STAGE_BYTES % 128 == 0, "TMA destination alignment");
    static_assert(LOAD_COLS <= 256, "TMA box limit");
    __host__ __device__ static constexpr int smem_per_warp(int stages) {
        return stages * STAGE_BYTES + RING_BYTES + QUEUE_BYTES + 128 /* mbarriers, counter */;
    }
};

// ---------------------------------------------------------------------------------------
// Exact evaluation of one pixel: integer window gather + fp64 test in the reference's
// operation order.  extended_pass1 = true gives DispersionExtendedThreshold's first pass
// (baseline.cpp:465-469: "a > c" only, no global threshold, no signal test).
// The fp64 decision from exact integer window sums, in the reference's operation order
// (standalone.cc:164-170; with a gain map baseline.cpp:240-246).  min_count is checked by the
// caller's common prologue here as well so that every resolve routine shares one copy.
__device__ __forceinline__ bool dispersion_decide(const ThreshArgs& a, uint32_t m, uint32_t x,
                                                  unsigned long long y, uint32_t v, int col, int row) {
    const double vd = (double)v, md = (double)m, xd = (double)x, yd = (double)y;
    if (v > a.tmax) return false;  // above the trusted range: an invalid pixel
    if (!((int)m >= a.min_count)) return false;
    if (!(vd > a.threshold)) return false;
    const double b = __dsub_rn(__dmul_rn(md, vd), xd);
    if (a.gain) {
        // double a = m * y - x * x;  double c = gain[k] * x * (m - 1 + nsig_b_ * std::sqrt(2 * (m - 1)));
        // double d = nsig_s_ * std::sqrt(gain[k] * x * m);
        const double g = a.gain[(size_t)row * a.width + col];
        const double gx = __dmul_rn(g, xd);
        const double d = __dmul_rn(a.nsig_s, __dsqrt_rn(__dmul_rn(gx, md)));
        if (!(b > d)) return false;
        const double aa = __dsub_rn(__dmul_rn(md, yd), __dmul_rn(xd, xd));
        const double m1 = __dsub_rn(md, 1.0);
        const double cc = __dmul_rn(gx, __dadd_rn(m1, __dmul_rn(a.nsig_b, __dsqrt_rn(__dmul_rn(2.0, m1)))));
        return aa > cc;
    }
    const double d = __dmul_rn(a.nsig_s, __dsqrt_rn(__dmul_rn(xd, md)));
    if (!(b > d)) return false;
    const double aa = __dsub_rn(__dsub_rn(__dmul_rn(md, yd), __dmul_rn(xd, xd)),
                                __dmul_rn(xd, __dsub_rn(md, 1.0)));
    const double cc =
      __dmul_rn(__dmul_rn(xd, a.nsig_b), __dsqrt_rn(__dmul_rn(2.0, __dsub_rn(md, 1.0))));
    return aa > cc;
}

template <typename PixelT, int KH, bool EXT_PASS1>
__device__ __forceinline__ bool resolve_pixel(const ThreshArgs& a, const uint8_t* frame,
                                              const uint8_t* valid, int col, int row) {
    const int x0 = max(col - KH, 0), x1 = min(col + KH, a.width - 1);
    const int y0 = max(row - KH, 0), y1 = min(row + KH, a.height - 1);
    if (!((valid[(size_t)row * a.valid_pitch + (col >> 3)] >> (col & 7)) & 1u)) return false;
    uint32_t m = 0, x = 0;
    unsigned long long y = 0;
    for (int yy = y0; yy <= y1; ++yy) {
        const PixelT* prow = reinterpret_cast<const PixelT*>(frame + (size_t)yy * a.pitch);
        const uint8_t* vrow = valid + (size_t)yy * a.valid_pitch;
        // 2KH+1 <= 16 columns span at most 3 bytes
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
                // int mm = (mask[k] && src[k] < BIG) ? 1 : 0;   standalone.cc:90
                if (((bits >> k) & 1u) && pv <= a.vlim) {
                    m += 1;
                    x += pv;
                    y += (unsigned long long)pv * pv;
                }
            }
        }
    }
    if (!EXT_PASS1)
        return dispersion_decide(a, m, x, y, (uint32_t) reinterpret_cast<const PixelT*>(frame + (size_t)row * a.pitch)[col],
                                 col, row);
    const uint32_t vc = reinterpret_cast<const PixelT*>(frame + (size_t)row * a.pitch)[col];
    if (vc > a.tmax) return false;  // above the trusted range: an invalid pixel (the caller notes it)
    const double vd = (double)vc;
    const double md = (double)m, xd = (double)x, yd = (double)y;
    // if (mask[k] && m >= min_count_ && x >= 0 && src[k] > threshold_)
    if (!((int)m >= a.min_count)) return false;
    if (!EXT_PASS1) {
        if (!(vd > a.threshold)) return false;
        // double b = m * src[k] - x;   double d = nsig_s_ * std::sqrt(x * m);
        const double b = __dsub_rn(__dmul_rn(md, vd), xd);
        const double d = __dmul_rn(a.nsig_s, __dsqrt_rn(__dmul_rn(xd, md)));
        if (!(b > d)) return false;
    }
    if (EXT_PASS1 && a.gain) {
        // baseline.cpp:535-537: double a = m * y - x * x;
        //                       double c = gain[k] * x * (m - 1 + nsig_b_ * std::sqrt(2 * (m - 1)));
        const double g = a.gain[(size_t)row * a.width + col];
        const double m1 = __dsub_rn(md, 1.0);
        const double aa = __dsub_rn(__dmul_rn(md, yd), __dmul_rn(xd, xd));
        const double cc = __dmul_rn(__dmul_rn(g, xd),
                                    __dadd_rn(m1, __dmul_rn(a.nsig_b, __dsqrt_rn(__dmul_rn(2.0, m1)))));
        return aa > cc;
    }
    // double a = m * y - x * x - x * (m - 1);  double c = x * nsig_b_ * std::sqrt(2 * (m - 1));
    const double aa = __dsub_rn(__dsub_rn(__dmul_rn(md, yd), __dmul_rn(xd, xd)),
                                __dmul_rn(xd, __dsub_rn(md, 1.0)));
    const double cc =
      __dmul_rn(__dmul_rn(xd, a.nsig_b), __dsqrt_rn(__dmul_rn(2.0, __dsub_rn(md, 1.0))));
    return aa > cc;
}

// Same decision as resolve_pixel (standard dispersion), gathering each window row with 16-byte
// loads: a scattered gather is bound by L1 wavefronts (one per distinct line touched per
// lane), so 2-3 wide loads + 2 validity bytes per row replace 7 + 3 narrow ones.
template <typename PixelT, int KH>
__device__ __forceinline__ bool resolve_pixel_wide(const ThreshArgs& a, const uint8_t* frame,
                                                   const uint8_t* valid, int col, int row) {
    constexpr int PC = 16 / (int)sizeof(PixelT);           // pixels per 16-byte chunk
    constexpr int NCH = sizeof(PixelT) == 2 ? 2 : 3;       // chunks covering any 7-pixel window
    // the caller has already established mask[k] for the centre pixel
    const int y0 = max(row - KH, 0), y1 = min(row + KH, a.height - 1);
    const int xs = col - KH;                               // may be negative
    const int ch0 = xs >> (sizeof(PixelT) == 2 ? 3 : 2);   // floor(xs / PC)
    const int base_px = ch0 * PC;
    const int n_chunks = (int)(a.pitch >> 4);
    const int vb0 = base_px >> 3;                          // floor; base_px is a multiple of 4
    const uint32_t win = 0x7fu << (xs - base_px);
    uint32_t m = 0, x = 0;
    unsigned long long y = 0;
    // all rows unrolled so that the loads of the whole window are in flight together
#pragma unroll
    for (int dy = -KH; dy <= KH; ++dy) {
        const int yy = row + dy;
        const bool in_rows = yy >= y0 && yy <= y1;
        const int yc = min(max(yy, y0), y1);
        const uint4* prow = reinterpret_cast<const uint4*>(frame + (size_t)yc * a.pitch);
        const uint8_t* vrow = valid + (size_t)yc * a.valid_pitch;
        uint32_t bits = 0;
        if (in_rows && vb0 >= 0 && vb0 < (int)a.valid_pitch) bits = vrow[vb0];
        if (in_rows && vb0 + 1 >= 0 && vb0 + 1 < (int)a.valid_pitch) bits |= (uint32_t)vrow[vb0 + 1] << 8;
        if (sizeof(PixelT) == 4) bits >>= (base_px & 4);
        const uint32_t sel = bits & win;
        uint4 ch[NCH];
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
            const int ci = ch0 + k;
            ch[k] = (ci >= 0 && ci < n_chunks) ? prow[ci] : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
            const uint32_t wv[4] = {ch[k].x, ch[k].y, ch[k].z, ch[k].w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if constexpr (sizeof(PixelT) == 2) {
                    // (16-bit data with a trusted range goes through resolve_pixel instead: see resolve_any)
                    const uint32_t lo = (sel >> (k * 8 + 2 * q)) & 1u ? (wv[q] & 0xffffu) : 0u;
                    const uint32_t hi = (sel >> (k * 8 + 2 * q + 1)) & 1u ? (wv[q] >> 16) : 0u;
                    x += lo + hi;
                    y += (unsigned long long)lo * lo;
                    y += (unsigned long long)hi * hi;
                } else {
                    // int mm = (mask[k] && src[k] < BIG) ? 1 : 0;   standalone.cc:90
                    const bool counted = ((sel >> (k * 4 + q)) & 1u) && wv[q] <= a.vlim;
                    const uint32_t pv = counted ? wv[q] : 0u;
                    m += counted ? 1u : 0u;
                    x += pv;
                    y += (unsigned long long)pv * pv;
                }
            }
        }
        if constexpr (sizeof(PixelT) == 2) m += __popc(sel);
    }
    const uint32_t vc = reinterpret_cast<const PixelT*>(frame + (size_t)row * a.pitch)[col];
    return dispersion_decide(a, m, x, y, vc, col, row);
}

// Exact test for a pixel whose 7x7 window lies inside the image and is fully mask-valid
// (mcount == 49): no validity loads, no clipping; the window position only selects which
// 16-bit halves of the two 16-byte chunks per row take part, and that selection is turned
// into AND masks once per pixel.  x via IDP.2A, y via 64-bit IMAD.WIDE accumulation.
// Same fp64 comparison, same operation order as resolve_pixel (standalone.cc:113-174).
template <typename PixelT, int KH>
__device__ __forceinline__ bool resolve_pixel_full(const ThreshArgs& a, const uint8_t* frame, int col, int row) {
    const int xs = col - KH;  // >= 0 here
    const int n_chunks = (int)(a.pitch >> 4);
    uint32_t m = (2 * KH + 1) * (2 * KH + 1), x = 0;
    unsigned long long y = 0;
    // the centre pixel: its line is loaded by the window gather below as well (L1 hit)
    const uint32_t vc = reinterpret_cast<const PixelT*>(frame + (size_t)row * a.pitch)[col];
    if constexpr (sizeof(PixelT) == 2) {
        const int ch0 = xs >> 3, o = xs & 7;
        const int ch1 = min(ch0 + 1, n_chunks - 1);  // only read when the window reaches it
        const uint32_t win = 0x7fu << o;             // halves 0..13 of the two chunks
        uint32_t M[7];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t zz = ((win >> (4 * k)) & 0xfu) * 0x10204080u;
            M[2 * k] = ptx::prmt(zz, 0u, 0x9988u);
            if (2 * k + 1 < 7) M[2 * k + 1] = ptx::prmt(zz, 0u, 0xbbaau);
        }
#pragma unroll
        for (int dy = -KH; dy <= KH; ++dy) {
            const uint4* prow = reinterpret_cast<const uint4*>(frame + (size_t)(row + dy) * a.pitch);
            // the gather is bound by L1 wavefronts (one per lane and load): skip the second
            // chunk when the 7 pixels fit into the first (window offset 0 or 1)
            const uint4 A = prow[ch0];
            const uint4 B = o > 1 ? prow[ch1] : make_uint4(0, 0, 0, 0);
            const uint32_t w[7] = {A.x, A.y, A.z, A.w, B.x, B.y, B.z};
#pragma unroll
            for (int q = 0; q < 7; ++q) {
                const uint32_t t = w[q] & M[q];
                x = __dp2a_lo(t, 0x0101u, x);
                const uint32_t lo = t & 0xffffu, hi = t >> 16;
                y += (unsigned long long)lo * lo;
                y += (unsigned long long)hi * hi;
            }
        }
    } else {
        const int ch0 = xs >> 2, o = xs & 3;
        const int ch2 = min(ch0 + 2, n_chunks - 1);
        const uint32_t win = 0x7fu << o;  // pixels 0..9 of three chunks
#pragma unroll
        for (int dy = -KH; dy <= KH; ++dy) {
            const uint4* prow = reinterpret_cast<const uint4*>(frame + (size_t)(row + dy) * a.pitch);
            const uint4 A = prow[ch0], B = prow[ch0 + 1], C = prow[ch2];
            const uint32_t w[10] = {A.x, A.y, A.z, A.w, B.x, B.y, B.z, B.w, C.x, C.y};
#pragma unroll
            for (int q = 0; q < 10; ++q) {
                // int mm = (mask[k] && src[k] < BIG) ? 1 : 0;   standalone.cc:90
                const bool in = (win >> q) & 1u;
                const bool counted = in && w[q] <= a.vlim;
                const uint32_t pv = counted ? w[q] : 0u;
                m -= (in && !counted) ? 1u : 0u;
                x += pv;
                y += (unsigned long long)pv * pv;
            }
        }
    }
    return dispersion_decide(a, m, x, y, vc, col, row);
}

// resolve_pixel_full for 16-bit frames whose window holds only pixels < 256 (the common case on
// photon-counting data).  Returns 0 = not strong, 1 = strong, 2 = "a window pixel is >= 256":
// the caller sends that pixel through the general exact routine.
//   * x and y = sum v^2 both by IDP.2A: with v < 256 the low bytes of a word's two halves are
//     the pixels themselves, so d = c + lo*lo + hi*hi is one instruction; 49 * 255^2 fits 32 bits.
//   * m = 49, so a = 49 y - x^2 - 48 x and b = 49 v - x are exact 32-bit integers; they are
//     compared with fp32 evaluations of c = x nsig_b sqrt(96) and d = nsig_s sqrt(49 x) first and
//     only a pixel within 1e-4 (relative) of either boundary runs the fp64 sequence of the
//     reference (dispersion_decide).  fp32 rounding and MUFU.SQRT are below 1e-6, so a decision
//     taken outside that band equals the fp64 one.
__device__ __forceinline__ int full_small_decide(const ThreshArgs& a, uint32_t x, uint32_t y, uint32_t any,
                                                 uint32_t vc, int col, int row);

template <int KH>
__device__ __forceinline__ int resolve_pixel_full_small(const ThreshArgs& a, const uint8_t* frame, int col, int row) {
    static_assert(KH == 3, "constants below are for the 7x7 window");
    const int xs = col - KH;  // >= 0 here
    const int n_chunks = (int)(a.pitch >> 4);
    const int ch0 = xs >> 3, o = xs & 7;
    const int ch1 = min(ch0 + 1, n_chunks - 1);  // only read when the window reaches it
    const uint32_t win = 0x7fu << o;             // halves 0..13 of the two chunks
    uint32_t M[7];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t zz = ((win >> (4 * k)) & 0xfu) * 0x10204080u;
        M[2 * k] = ptx::prmt(zz, 0u, 0x9988u);
        if (2 * k + 1 < 7) M[2 * k + 1] = ptx::prmt(zz, 0u, 0xbbaau);
    }
    uint32_t x = 0, y = 0, any = 0;
#pragma unroll
    for (int dy = -KH; dy <= KH; ++dy) {
#if defined(FUSED_EXP_L1)  // timing experiment (wrong results): every gather hits the same few L1 lines
        const uint4* prow = reinterpret_cast<const uint4*>(frame + (size_t)(dy + KH) * a.pitch);
        const uint4 A = prow[ch0 & 1];
        const uint4 B = o > 1 ? prow[ch1 & 1] : make_uint4(0, 0, 0, 0);
#else
        const uint4* prow = reinterpret_cast<const uint4*>(frame + (size_t)(row + dy) * a.pitch);
        const uint4 A = prow[ch0];
        const uint4 B = o > 1 ? prow[ch1] : make_uint4(0, 0, 0, 0);
#endif
        const uint32_t w[7] = {A.x, A.y, A.z, A.w, B.x, B.y, B.z};
#pragma unroll
        for (int q = 0; q < 7; ++q) {
            const uint32_t t = w[q] & M[q];
            any |= t;
            x = __dp2a_lo(t, 0x0101u, x);
            y = __dp2a_lo(t, ptx::prmt(t, 0u, 0x4420u), y);  // bytes {t.b0, t.b2}: the two pixels if < 256
        }
    }
    const uint32_t vc = reinterpret_cast<const uint16_t*>(frame + (size_t)row * a.pitch)[col];
    return full_small_decide(a, x, y, any, vc, col, row);
}

// The decision of resolve_pixel_full_small from the window's sums (m = 49).
__device__ __forceinline__ int full_small_decide(const ThreshArgs& a, uint32_t x, uint32_t y, uint32_t any,
                                                 uint32_t vc, int col, int row) {
    if (any & 0xff00ff00u) return 2;
    if (a.gain) return dispersion_decide(a, 49u, x, (unsigned long long)y, vc, col, row) ? 1 : 0;
    if (!(49 >= a.min_count) || !((int)vc > a.thr_floor)) return 0;
    const int ai = (int)(49u * y - x * x - 48u * x);   // |.| < 2^28
    const int bi = (int)(49u * vc - x);
    const float xf = (float)x;
    const float cf = xf * ((float)a.nsig_b * 9.797959f);            // x nsig_b sqrt(2 * 48)
    float sq;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(49.0f * xf));
    const float df = (float)a.nsig_s * sq;                           // nsig_s sqrt(x m)
    const float af = (float)ai, bf = (float)bi;
    constexpr float LO = 1.0f - 1e-4f, HI = 1.0f + 1e-4f;
    if (af < cf * LO || bf < df * LO) return 0;          // clearly fails a > c or b > d
    if (af > cf * HI && bf > df * HI) return 1;          // clearly passes both
    return dispersion_decide(a, 49u, x, (unsigned long long)y, vc, col, row) ? 1 : 0;
}

__device__ __forceinline__ void set_strong(const ThreshArgs& a, int z, int col, int row) {
    const size_t word = (size_t)row * (a.strong_pitch >> 2) + (col >> 5);
    uint32_t* w = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(a.strong_bits) +
                                              (size_t)z * a.strong_frame_stride) + word;
    const uint32_t bit = 1u << (col & 31);
    const uint32_t before = atomicOr(w, bit);
    // strong pixels are sparse, so counting them per compaction chunk here (one more atomic
    // per strong pixel) replaces a kernel that re-reads every strong plane (cc_count_kernel)
    if (a.chunk_count && !(before & bit)) atomicAdd(a.chunk_count + (size_t)z * a.zstride_words + (word >> 10), 1u);
}

// grid: (blocks, n_frames).  One thread per candidate of frame blockIdx.y (grid-stride).
template <typename PixelT, int KH, bool EXT_PASS1>
__global__ void __launch_bounds__(256) resolve_kernel(const ThreshArgs a) {
    const int z = a.slot0 + blockIdx.y;
    const uint8_t* frame = a.image + (size_t)z * a.frame_stride;
    const uint8_t* valid = a.valid_bits + (size_t)z * a.valid_frame_stride;
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
        if (EXT_PASS1 && a.tmax != 0xffffffffu) {  // extended: trusted_clear_kernel needs the list (trusted.cuh)
            const uint32_t vc = reinterpret_cast<const PixelT*>(frame + (size_t)row * a.pitch)[col];
            if (vc > a.tmax && ((valid[(size_t)row * a.valid_pitch + (col >> 3)] >> (col & 7)) & 1u))
                note_bad_pixel(a, z, col, row);
        }
        if (resolve_pixel<PixelT, KH, EXT_PASS1>(a, frame, valid, col, row)) {
            set_strong(a, z, col, row);
            ++found;
        }
    }
    for (int off = 16; off > 0; off >>= 1) found += __shfl_down_sync(0xffffffffu, found, off);
    if ((threadIdx.x & 31) == 0 && found) atomicAdd(a.strong_count + (size_t)z * a.zstride_words, found);
}

}  // namespace ffsb

#include "threshold_prefilter.cuh"
