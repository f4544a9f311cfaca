// dispersion_fused.cuh -- DIALS dispersion threshold (DispersionThreshold,
// baseline/spotfinder/standalone.cc:27-206) as ONE hand-written sm_100a kernel: the frame is
// read from HBM exactly once and the strong-pixel bit-plane is the only output.
//
// A *warp task* is a strip of 240 output columns x tile_rows output rows of one frame.  Every
// warp owns its task end to end -- its own TMA pipeline (STAGES stages of 4 rows: image tile
// + group-info tile, zero-filled out of bounds, which is the reference's window clipping),
// its own mbarriers, its own shared-memory ring of the last 8 rows and its own work queues --
// so warps never synchronise with each other after the first instruction.
// Lane l owns the 8 adjacent columns of one *pixel group*; the first and last lane of the
// warp are halo lanes.
//
// The exact test (integer window sums m, x, y and the fp64 comparison in the reference's
// operation order, resolve_pixel() in threshold.cuh) costs ~250 instructions per pixel, so
// the streaming loop only has to decide, for each group of 8 pixels, whether any of them can
// possibly pass.  A strong pixel satisfies (standalone.cc:164-170)
//      b > d   <=>   v > mu + nsig_s * sqrt(mu),   mu = x / m   (window mean),
// whose right-hand side increases with mu.  The loop therefore forms a LOWER bound of x and
// uses an UPPER bound of m:
//   * pixels are masked, clamped to 9362 (7 * 9362 < 2^16) and packed two per 32-bit word, so
//     the vertical 7-row running sum is 4 IADD3 per 8 pixels; the horizontal 7-tap sums are
//     formed in 32 bits from 4 warp shuffles; x_lb = min over the group's 8 windows;
//   * m_ub = number of mask-valid pixels in the window, maximised over the group -- a
//     property of the mask alone, precomputed once per mask (group_info_kernel) and delivered
//     by TMA next to the validity bits (2 bytes per group);
//   * the group is a candidate iff  max(masked v) > floor((mu_lb + nsig_s' sqrt(mu_lb)))
//     with mu_lb = x_lb / m_ub and safety margins far above the float rounding error.
// Candidate groups (~10 % on Poisson background) are queued in shared memory; whenever 32
// are pending the warp expands them to candidate pixels (v > T of the group) and resolves 32
// pixels at a time, one per lane, with the exact test, gathering the 7x7 window from the
// frame rows it has just streamed (L2 hits).  Nothing is approximate: the bound only prunes.
#pragma once
#include "threshold.cuh"

namespace ffsb {

#ifndef FUSED_DEBUG_MODES
#define FUSED_DEBUG_MODES 0  // 1: FFS_FUSED_DEBUG=5 (TMA only) is honoured by the streaming loop
#endif
#ifndef FUSED_ROWS_PER_BATCH
#define FUSED_ROWS_PER_BATCH 2  // measured: 4 is no faster at equal occupancy and costs 18 registers
#endif
#ifndef FUSED_DP2A
#define FUSED_DP2A 1
#endif
#ifndef FUSED_WARPS_PER_SM_16BIT
#define FUSED_WARPS_PER_SM_16BIT 16  // register cap of the 16-bit kernel = 65536 / (32 * this); 12 for A/B builds with more rows per batch
#endif
#ifndef FUSED_LUT_MASK
#define FUSED_LUT_MASK 1  // 16-bit front end: validity nibble -> {mask & clamp} words from a 16-entry shared table
#endif

template <typename PixelT>
struct FusedGeo {
    static constexpr int KH = 3;
    static constexpr int SR = 4;            // rows per TMA stage (== half of the ring)
    static constexpr int PPL = 8;           // pixels per lane = one pixel group
    static constexpr int LOAD_COLS = 32 * PPL;
    static constexpr int OUT_LANES = 30;
    static constexpr int OUT_COLS = OUT_LANES * PPL;
    static constexpr unsigned CLAMP = 9362u;  // 7 * 9362 = 65534: a column sum fits 16 bits
    static constexpr int IMG_ROW_BYTES = LOAD_COLS * (int)sizeof(PixelT);
    static constexpr int INFO_ROW_BYTES = 80;  // 32 groups * 2 B + <= 14 B alignment slack
    static constexpr int IMG_STAGE_BYTES = SR * IMG_ROW_BYTES;
    static constexpr int INFO_STAGE_BYTES = (SR * INFO_ROW_BYTES + 127) / 128 * 128;
    static constexpr int STAGE_BYTES = IMG_STAGE_BYTES + INFO_STAGE_BYTES;
    static constexpr int TX_BYTES = IMG_STAGE_BYTES + SR * INFO_ROW_BYTES;
    static constexpr int RING_ROWS = 2 * SR;
    static constexpr int RING_ROW_BYTES = 32 * 16;
    static constexpr int RING_BYTES = RING_ROWS * RING_ROW_BYTES;
    // resident warps per SM the kernel is compiled for (register cap = 65536 / (32 * this)):
    // 16-bit: 4 CTAs of 4 warps at 128 registers and 13.8 KB of shared memory per warp; the
    // 32-bit front end needs more registers
    static constexpr int MIN_WARPS_PER_SM = sizeof(PixelT) == 2 ? FUSED_WARPS_PER_SM_16BIT : 12;
    static constexpr int GQ_CAP = 96;        // pending candidate groups (< 32 + 2 rows * 30)
    static constexpr int PQ_CAP = 320;       // pending candidate pixels, both kinds (< 32 + 32 + 32 * 8)
    // per queued group: 4 B meta + 32 B window sums (the centre-row pixels are re-read from the
    // frame in stage 2: carrying them cost 1.5 KB of shared memory per warp, i.e. occupancy)
    // plus one double-ended pixel queue: pixels whose window is fully valid (fast exact test)
    // grow from the bottom, all others from the top
    static constexpr int QUEUE_BYTES = GQ_CAP * (4 + 32) + PQ_CAP * 4;
    static_assert(RING_ROWS > 2 * KH, "ring must hold the 7-row window");
    static_assert(IMG_STAGE_BYTES % 128 == 0 && QUEUE_BYTES % 128 == 0, "TMA destination alignment");
    __host__ __device__ static constexpr int smem_per_warp(int stages) {
        return stages * STAGE_BYTES + RING_BYTES + QUEUE_BYTES + 128 /* mbarriers */;
    }
};

// group info plane: per row, per group of 8 pixels, 2 bytes {validity bits, m_ub | full << 6}.
// m_ub = max over the group's mask-valid pixels of the number of valid pixels in the
// image-clipped 7x7 window (standalone.cc:126-130 clipping).
__global__ void group_info_kernel(const uint8_t* __restrict__ valid_bits, size_t bits_pitch, int width,
                                  int height, uint8_t* __restrict__ info, size_t info_pitch,
                                  uint8_t* __restrict__ mplane, size_t m_pitch,
                                  uint8_t* __restrict__ info4, size_t info4_pitch) {
    const int groups = (width + 7) / 8;
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (long long)groups * height) return;
    const int row = (int)(t / groups), gi = (int)(t % groups);
    const int bytes_per_row = (width + 7) / 8;
    uint32_t m[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int dy = -3; dy <= 3; ++dy) {
        const int r = row + dy;
        if (r < 0 || r >= height) continue;
        const uint8_t* vrow = valid_bits + (size_t)r * bits_pitch;
        // 24 validity bits: columns 8*gi - 8 .. 8*gi + 15 (pack_mask_kernel clears bits >= width)
        uint32_t b = (uint32_t)vrow[gi] << 8;
        if (gi > 0) b |= vrow[gi - 1];
        if (gi + 1 < bytes_per_row) b |= (uint32_t)vrow[gi + 1] << 16;
#pragma unroll
        for (int p = 0; p < 8; ++p) m[p] += __popc((b >> (8 + p - 3)) & 0x7fu);
    }
    const uint32_t vb = valid_bits[(size_t)row * bits_pitch + gi];
    uint32_t mub = 0, mlb = 49;
#pragma unroll
    for (int p = 0; p < 8; ++p)
        if ((vb >> p) & 1u) {
            mub = max(mub, m[p]);
            mlb = min(mlb, m[p]);
        }
    uint8_t* out = info + (size_t)row * info_pitch + 2 * gi;
    out[0] = (uint8_t)vb;
    // bit 6: every valid pixel of the group has a full window (inside the image, all 49 valid)
    out[1] = (uint8_t)(mub | (mlb == 49u ? 0x40u : 0u));
    // 4-byte variant for the extended first pass (extended_stream.cuh): + the smallest count
    *reinterpret_cast<uint32_t*>(info4 + (size_t)row * info4_pitch + 4 * gi) =
      vb | ((mub | (mlb == 49u ? 0x40u : 0u)) << 8) | ((mlb & 63u) << 16);
    // per-pixel window counts (extended_stream.cuh): 8 bytes per group, 0 beyond the width
    uint2 mm;
    mm.x = m[0] | (m[1] << 8) | (m[2] << 16) | (m[3] << 24);
    mm.y = m[4] | (m[5] << 8) | (m[6] << 16) | (m[7] << 24);
    const int left = width - 8 * gi;  // pixels of this group inside the image
    if (left < 8) {
        const unsigned long long keep = (1ull << (8 * left)) - 1ull;
        const unsigned long long v = ((unsigned long long)mm.y << 32 | mm.x) & keep;
        mm.x = (uint32_t)v;
        mm.y = (uint32_t)(v >> 32);
    }
    *reinterpret_cast<uint2*>(mplane + (size_t)row * m_pitch + 8 * gi) = mm;
}

// ---- row front end: masked + clamped packed words, max of the masked raw values -----------
// lut_s (16-bit pixels, FUSED_LUT_MASK): shared address of a 16-entry table, entry n = the two
// words {0xffff per set bit of the nibble n} & {CLAMP, CLAMP}; an unsigned min against it masks
// and clamps in one instruction.  vmax is then the max of the CLAMPED values: the caller
// compares it with min(T, CLAMP - 1), so a group holding a pixel >= CLAMP is always a candidate
// and a group of smaller pixels is tested as before.
template <typename PixelT>
__device__ __forceinline__ void fused_load_row(const uint8_t* pimg, const uint8_t* pinfo, uint32_t lut_s,
                                               uint32_t vlim, uint32_t (&c)[4], uint32_t& vmax, uint32_t& info) {
    constexpr uint32_t CL = FusedGeo<PixelT>::CLAMP;
    info = *reinterpret_cast<const uint16_t*>(pinfo);
    if constexpr (sizeof(PixelT) == 2 && FUSED_LUT_MASK) {
        const uint4 raw = ptx::lds128(pimg);
        // the table is 128-byte aligned: (index & 0x78) | base is one LOP3
        const uint2 ml = ptx::lds64_s(((info << 3) & 0x78u) | lut_s);
        const uint2 mh = ptx::lds64_s(((info >> 1) & 0x78u) | lut_s);
        c[0] = __vminu2(raw.x, ml.x);
        c[1] = __vminu2(raw.y, ml.y);
        c[2] = __vminu2(raw.z, mh.x);
        c[3] = __vminu2(raw.w, mh.y);
        const uint32_t m2 = __vmaxu2(__vimax3_u16x2(c[0], c[1], c[2]), c[3]);
        vmax = max(m2 & 0xffffu, m2 >> 16);
    } else if constexpr (sizeof(PixelT) == 2) {
        const uint4 raw = ptx::lds128(pimg);
        // bit k of the validity byte -> sign bit of byte k; prmt replicates sign bits to bytes
        const uint32_t zlo = (info & 0xfu) * 0x10204080u;
        const uint32_t zhi = ((info >> 4) & 0xfu) * 0x10204080u;
        const uint32_t w0 = raw.x & ptx::prmt(zlo, 0u, 0x9988u);
        const uint32_t w1 = raw.y & ptx::prmt(zlo, 0u, 0xbbaau);
        const uint32_t w2 = raw.z & ptx::prmt(zhi, 0u, 0x9988u);
        const uint32_t w3 = raw.w & ptx::prmt(zhi, 0u, 0xbbaau);
        const uint32_t m2 = __vmaxu2(__vimax3_u16x2(w0, w1, w2), w3);
        vmax = max(m2 & 0xffffu, m2 >> 16);
        constexpr uint32_t CL2 = CL | (CL << 16);
        c[0] = __vminu2(w0, CL2);
        c[1] = __vminu2(w1, CL2);
        c[2] = __vminu2(w2, CL2);
        c[3] = __vminu2(w3, CL2);
    } else {
        const uint4 ra = ptx::lds128(pimg);
        const uint4 rb = ptx::lds128(pimg + 16);
        // lut_s (32-bit pixels): 16-entry table, entry n = four AND masks {~0 per set bit of the nibble n};
        // 256-byte aligned, so (index << 4) | base is one LOP3
        const uint4 mlo = ptx::lds128_s(((info << 4) & 0xf0u) | lut_s);
        const uint4 mhi = ptx::lds128_s((info & 0xf0u) | lut_s);
        const uint32_t v[8] = {ra.x & mlo.x, ra.y & mlo.y, ra.z & mlo.z, ra.w & mlo.w,
                               rb.x & mhi.x, rb.y & mhi.y, rb.z & mhi.z, rb.w & mhi.w};
        uint32_t cl[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) cl[j] = min(v[j], CL);
        vmax = max(max(max(v[0], v[1]), max(v[2], v[3])), max(max(v[4], v[5]), max(v[6], v[7])));
        // pixels >= 2^24 are not counted by the reference (standalone.cc:90), nor are pixels
        // above the trusted range: they add 0  (vlim = min(trusted_max, 2^24 - 1)).  Rare: only a
        // group that holds one takes this branch.
        if (vmax > vlim) {
#pragma unroll
            for (int j = 0; j < 8; ++j) cl[j] = v[j] <= vlim ? cl[j] : 0u;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) c[k] = cl[2 * k] | (cl[2 * k + 1] << 16);
    }
}

// ---- rare path, deliberately NOT inlined (one copy of ~2000 instructions instead of five) ----
// Expands the last `n` queued groups to candidate pixels with each pixel's own window sum,
// then resolves the queued pixels 32 at a time (all of them if flush_all) with the exact
// test.  state = qn | pqn << 8 | pgn << 20 (all warp-uniform); the new state is returned.
template <typename PixelT>
__device__ __noinline__ uint32_t fused_resolve(const ThreshArgs* __restrict__ ap, uint8_t* qbase,
                                               const float* rcp_tab, int x0, int r0, int z, uint32_t state,
                                               int n, bool flush_all) {
    using G = FusedGeo<PixelT>;
    constexpr int PPL = G::PPL, KH = G::KH;
    const ThreshArgs& a = *ap;
    const int lane = threadIdx.x & 31;
    uint32_t* gq = reinterpret_cast<uint32_t*>(qbase);
    uint4* hq = reinterpret_cast<uint4*>(gq + G::GQ_CAP);
    uint32_t* pq = reinterpret_cast<uint32_t*>(hq + 2 * G::GQ_CAP);  // [0, pqn): pixels with a full 7x7 window
    uint32_t* pg = pq + G::PQ_CAP - 1;                           // pg[-k], k < pgn: pixels near the mask / the edge
    const uint8_t* frame = a.image + (size_t)z * a.frame_stride;
    const uint8_t* valid = a.valid_bits;
    const float sk = a.sig_k;
    const int thr_floor = a.thr_floor;
    int qn = state & 0xff, pqn = (state >> 8) & 0xfff, pgn = state >> 20;
    unsigned found = 0;
    const uint32_t lt_mask = (1u << lane) - 1u;
    // resolves one full-window pixel per active lane; 16-bit pixels whose window holds a value
    // >= 256 are handed to the general queue (pg) -- called by converged warps only
    auto resolve_full = [&](bool active, uint32_t e) {
        const int col = e & 0x1fff, row = e >> 13;
        int r = 0;
        if (active && a.force_brute != 2) {
            if constexpr (sizeof(PixelT) == 2) r = resolve_pixel_full_small<KH>(a, frame, col, row);
            else r = resolve_pixel_full<PixelT, KH>(a, frame, col, row) ? 1 : 0;
        }
        if (r == 1) {
            set_strong(a, z, col, row);
            ++found;
        }
        if constexpr (sizeof(PixelT) == 2) {
            const unsigned again = __ballot_sync(0xffffffffu, r == 2);
            if (again) {
                if (r == 2) pg[-(pgn + __popc(again & lt_mask))] = e;
                pgn += __popc(again);
            }
        }
    };
    auto resolve_any = [&](uint32_t e) {
        const int col = e & 0x1fff, row = e >> 13;
        if (a.force_brute == 2) return;
        // 16-bit frames with a trusted range: the general routine leaves pixels above it out of
        // the sums (the wide gather counts by mask bits alone)
        const bool strong = (sizeof(PixelT) == 2 && a.vlim < 0xffffu)
                              ? resolve_pixel<PixelT, KH, false>(a, frame, valid, col, row)
                              : resolve_pixel_wide<PixelT, KH>(a, frame, valid, col, row);
        if (strong) {
            set_strong(a, z, col, row);
            ++found;
        }
    };
    if (n > 0) {
        uint32_t bits = 0, full = 0, badb = 0;
        int col0 = 0, row = 0;
        const uint32_t tmax = a.tmax;
        if (lane < n) {
            const int qi = qn - n + lane;
            const uint32_t e = gq[qi];
            const uint4 ha = hq[2 * qi], hb = hq[2 * qi + 1];
            const uint32_t hx[8] = {ha.x, ha.y, ha.z, ha.w, hb.x, hb.y, hb.z, hb.w};
            const float rc = rcp_tab[(e >> 14) & 0x7fu];
            full = ((e >> 20) & 1u) ? 0xffu : 0u;  // group flag: all windows full (m == 49)
            col0 = x0 + (int)(e & 31u) * PPL;
            row = r0 + (int)((e >> 5) & 0x1ffu);
            uint32_t v[8], vb = 0xffu;
            {
                // the centre row's raw pixels and validity bits, from the frame (L2 hits)
                vb = valid[(size_t)row * a.valid_pitch + (col0 >> 3)];
                const PixelT* p = reinterpret_cast<const PixelT*>(frame + (size_t)row * a.pitch) + col0;
                if constexpr (sizeof(PixelT) == 2) {
                    const uint4 cw = reinterpret_cast<const uint4*>(p)[0];
                    v[0] = cw.x & 0xffffu; v[1] = cw.x >> 16; v[2] = cw.y & 0xffffu; v[3] = cw.y >> 16;
                    v[4] = cw.z & 0xffffu; v[5] = cw.z >> 16; v[6] = cw.w & 0xffffu; v[7] = cw.w >> 16;
                } else {
                    const uint4 ra = reinterpret_cast<const uint4*>(p)[0], rb = reinterpret_cast<const uint4*>(p)[1];
                    v[0] = ra.x; v[1] = ra.y; v[2] = ra.z; v[3] = ra.w;
                    v[4] = rb.x; v[5] = rb.y; v[6] = rb.z; v[7] = rb.w;
                }
            }
            // per-pixel bound: the pixel's own window sum (still a lower bound of x) over m_ub
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float mu = (float)hx[j] * rc;
                float sq;
                asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(mu));
                uint32_t T = (uint32_t)max(__float2int_rz(fmaf(sk, sq, mu)), thr_floor);
                bits |= (((vb >> j) & 1u) && v[j] > T) ? (1u << j) : 0u;
                badb |= (((vb >> j) & 1u) && v[j] > tmax) ? (1u << j) : 0u;
            }
        }
        // Mask-valid pixels above the trusted range.  16-bit pixels reach the streaming sums
        // clamped, not zeroed, so the bounds of the windows around such a pixel are not bounds:
        // the pixel is listed and trusted_fixup_kernel decides its 7x7 neighbourhood exactly.
        // (Its group always gets here: its clamped value exceeds every group threshold.)
        for (; badb; badb &= badb - 1) note_bad_pixel(a, z, col0 + __ffs(badb) - 1, row);
        qn -= n;
        const uint32_t bf = bits & full, bg = bits & ~full;
        // one scan for both queues: fast count in the low half, general count in the high half
        const uint32_t cnt = (uint32_t)__popc(bf) | ((uint32_t)__popc(bg) << 16);
        uint32_t incl = cnt;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off) incl += t;
        }
        const uint32_t excl = incl - cnt;
        int pos = pqn + (int)(excl & 0xffffu);
        for (uint32_t b = bf; b;) {
            const int j = __ffs(b) - 1;
            b &= b - 1;
            pq[pos++] = (uint32_t)(col0 + j) | ((uint32_t)row << 13);
        }
        pos = pgn + (int)(excl >> 16);
        for (uint32_t b = bg; b;) {
            const int j = __ffs(b) - 1;
            b &= b - 1;
            pg[-(pos++)] = (uint32_t)(col0 + j) | ((uint32_t)row << 13);
        }
        const uint32_t tot = __shfl_sync(0xffffffffu, incl, 31);
        pqn += (int)(tot & 0xffffu);
        pgn += (int)(tot >> 16);
        __syncwarp();
        while (pqn >= 32) {
            pqn -= 32;
            resolve_full(true, pq[pqn + lane]);
        }
        __syncwarp();  // pixels handed over to pg above
        while (pgn >= 32) {
            pgn -= 32;
            resolve_any(pg[-(pgn + lane)]);
        }
        __syncwarp();
    }
    if (flush_all) {
        resolve_full(lane < pqn, lane < pqn ? pq[lane] : 0u);
        __syncwarp();
        while (pgn > 0) {  // < 64: the residue plus what the line above handed over
            const int k = min(pgn, 32);
            pgn -= k;
            if (lane < k) resolve_any(pg[-(pgn + lane)]);
        }
        pqn = pgn = 0;
    }
    for (int off = 16; off > 0; off >>= 1) found += __shfl_down_sync(0xffffffffu, found, off);
    if (lane == 0 && found) atomicAdd(a.strong_count + (size_t)z * a.zstride_words, found);
    return (uint32_t)qn | ((uint32_t)pqn << 8) | ((uint32_t)pgn << 20);
}

template <typename PixelT, int STAGES, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, FusedGeo<PixelT>::MIN_WARPS_PER_SM / WARPS)
  fused_dispersion_kernel(const __grid_constant__ CUtensorMap tm_img,
                          const __grid_constant__ CUtensorMap tm_info,
                          const __grid_constant__ ThreshArgs a) {
    using G = FusedGeo<PixelT>;
    constexpr int SR = G::SR, KH = G::KH, PPL = G::PPL, RB = FUSED_ROWS_PER_BATCH;
    static_assert(SR % RB == 0 && RB % 2 == 0, "the queue is checked every two rows");
    static_assert(SR == 4, "row loop is written for 4-row stages and an 8-row ring");

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ float rcp_tab[128];  // 0.99999 / m_ub, indexed by the info byte (m_ub | full << 6)
    // The out-of-line resolve path takes the arguments by pointer; a pointer to the kernel
    // parameter turns every field access into a generic load from the parameter window
    // (long-scoreboard latency on each call), a shared-memory copy is one LDS away.
    __shared__ ThreshArgs sa;
    // 16-bit: entry n = two words {0xffff per set bit of the nibble n} & {CLAMP, CLAMP} (128 B);
    // 32-bit: four AND masks (256 B).  Sized per pixel type: the 16-bit kernel's 4 CTAs per SM leave
    // no room for another kilobyte of static shared memory.
    constexpr int kLutBytes = sizeof(PixelT) == 2 ? 128 : 256;
    __shared__ __align__(kLutBytes) uint8_t mask_lut[kLutBytes];
    if (threadIdx.x < 16) {
        constexpr uint32_t CL = G::CLAMP;
        const uint32_t n = threadIdx.x;
        if constexpr (sizeof(PixelT) == 2)
            reinterpret_cast<uint2*>(mask_lut)[n] = make_uint2(((n & 1u) ? CL : 0u) | ((n & 2u) ? CL << 16 : 0u),
                                                               ((n & 4u) ? CL : 0u) | ((n & 8u) ? CL << 16 : 0u));
        else
            reinterpret_cast<uint4*>(mask_lut)[n] =
              make_uint4((n & 1u) ? ~0u : 0u, (n & 2u) ? ~0u : 0u, (n & 4u) ? ~0u : 0u, (n & 8u) ? ~0u : 0u);
    }
    uint32_t lut_s;  // opaque move: keeps the address in a register instead of re-deriving it per stage
    asm volatile("mov.u32 %0, %1;" : "=r"(lut_s) : "r"(ptx::smem_u32(mask_lut)));
    for (int t = threadIdx.x; t < 128; t += WARPS * 32) rcp_tab[t] = 0.99999f / (float)max(t & 63, 1);
    for (int t = threadIdx.x; t < (int)(sizeof(ThreshArgs) / 4); t += WARPS * 32)
        reinterpret_cast<uint32_t*>(&sa)[t] = reinterpret_cast<const uint32_t*>(&a)[t];
    __syncthreads();  // the only block-wide barrier

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const long long task = (long long)blockIdx.x * WARPS + warp;
    const long long n_tasks = (long long)a.n_frames * a.n_row_tiles * a.n_strips;
    if (task >= n_tasks) return;
    const int strip = (int)(task % a.n_strips);
    const int rt = (int)((task / a.n_strips) % a.n_row_tiles);
    const int z = a.slot0 + (int)(task / ((long long)a.n_strips * a.n_row_tiles));

    uint8_t* smem_al = smem_raw + ((128u - (ptx::smem_u32(smem_raw) & 127u)) & 127u);
    uint8_t* wsm = smem_al + (size_t)warp * G::smem_per_warp(STAGES);
    uint8_t* stage_base = wsm;
    uint8_t* ring = wsm + STAGES * G::STAGE_BYTES;
    uint32_t* gq = reinterpret_cast<uint32_t*>(ring + G::RING_BYTES);
    uint4* hq = reinterpret_cast<uint4*>(gq + G::GQ_CAP);   // 2 x uint4 per queued group: its 8 window sums
    uint64_t* bars = reinterpret_cast<uint64_t*>(ring + G::RING_BYTES + G::QUEUE_BYTES);
    static_assert(G::QUEUE_BYTES % 128 == 0, "mbarrier alignment");

    const int r0 = rt * a.tile_rows;
    const int row_end = min(r0 + a.tile_rows, a.height);
    const int n_load_rows = (row_end - r0) + 2 * KH;
    const int n_groups = (n_load_rows + SR - 1) / SR;
    const int x0 = strip * G::OUT_COLS - PPL;   // first loaded column (halo lane 0)
    const int y_first = r0 - KH;                // first loaded row
    const int info_b = 2 * (strip * G::OUT_LANES - 1);  // byte offset of lane 0's group in an info row
    const int info_b0 = info_b & ~15;           // 16-byte aligned TMA start (may be -16)

    if (lane == 0) {
        for (int s = 0; s < STAGES; ++s) ptx::mbar_init(bars + s, 1);
        ptx::fence_mbar_init();
    }
    {
        const uint4 zero = make_uint4(0, 0, 0, 0);
        for (int k = lane; k < G::RING_BYTES / 16; k += 32) reinterpret_cast<uint4*>(ring)[k] = zero;
    }
    __syncwarp();

    auto issue = [&](int g, int s) {
        uint8_t* dst = stage_base + s * G::STAGE_BYTES;
        ptx::mbar_arrive_expect_tx(bars + s, G::TX_BYTES);
        ptx::tma_load_3d(dst, &tm_img, bars + s, x0, y_first + g * SR, z);
        ptx::tma_load_2d(dst + G::IMG_STAGE_BYTES, &tm_info, bars + s, info_b0, y_first + g * SR);
    };
    if (lane == 0) {
        for (int g = 0; g < STAGES && g < n_groups; ++g) issue(g, g);
    }

    const bool out_lane = lane >= 1 && lane <= G::OUT_LANES;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const float sk = a.sig_k;             // nsig_s * (1 - 1e-4)
    const int thr_floor = a.thr_floor;
    const uint32_t vlim = sizeof(PixelT) == 4 ? a.vlim : 0u;  // 16-bit front end does not use it (registers)
    int qn = 0;              // warp-uniform fill level of the group queue
    uint32_t pstate = 0;     // fill levels of the two pixel queues (<< 8, << 20)
    uint8_t* qbase = reinterpret_cast<uint8_t*>(gq);

    uint32_t V[4] = {0, 0, 0, 0};   // packed column sums of the last 7 rows
    uint32_t hv1 = 0, hv2 = 0, hv3 = 0;  // max masked value of rows i-1, i-2, i-3
    uint32_t hi1 = 0, hi2 = 0, hi3 = 0;  // group info of rows i-1, i-2, i-3
    uint8_t* ring_lane = ring + lane * 16;
    const int pimg_off = lane * PPL * (int)sizeof(PixelT);
    const int pinfo_off = G::IMG_STAGE_BYTES + (info_b - info_b0) + 2 * lane;

    int s = 0;
    uint32_t phase = 0;
    bool ready = false;  // the next stage's barrier was already seen complete (polled early)
    for (int g = 0; g < n_groups; ++g) {
        if (!ready) ptx::mbar_wait(bars + s, phase);
        const uint8_t* st = stage_base + s * G::STAGE_BYTES;
        uint8_t* rcur = ring_lane + ((g & 1) ? SR * G::RING_ROW_BYTES : 0);
        uint8_t* roth = ring_lane + ((g & 1) ? 0 : SR * G::RING_ROW_BYTES);
#if FUSED_DEBUG_MODES
        if (a.force_brute != 5)   // 5: timing experiment, TMA traffic only
#endif
        {
        // The stage body is written in three straight-line phases so that the four rows'
        // dependent chains (LDS -> sums -> SHFL -> sums -> MUFU -> VOTE) interleave.
        // FUSED_ROWS_PER_BATCH rows at a time (4: one batch per stage; 2: two, fewer live registers)
#pragma unroll
        for (int hb = 0; hb < SR / RB; ++hb) {
        // ---- phase A: load, mask, clamp, vertical running sums, ring update ------------------
        uint32_t Vs[RB][4], vmc[RB], mub[RB];
#pragma unroll
        for (int uu = 0; uu < RB; ++uu) {
            const int u = hb * RB + uu;  // row of the stage
            uint32_t c[4], vm, info;
            fused_load_row<PixelT>(st + u * G::IMG_ROW_BYTES + pimg_off, st + pinfo_off + u * G::INFO_ROW_BYTES,
                                   lut_s, vlim, c, vm, info);
            // V += row(i) - row(i-7); ring slot of row i = i & 7
            uint8_t* oldp = (u + 1 < SR) ? rcur + (u + 1) * G::RING_ROW_BYTES : roth;
            const uint4 old = ptx::lds128(oldp);
            V[0] = V[0] + c[0] - old.x;
            V[1] = V[1] + c[1] - old.y;
            V[2] = V[2] + c[2] - old.z;
            V[3] = V[3] + c[3] - old.w;
            *reinterpret_cast<uint4*>(rcur + u * G::RING_ROW_BYTES) = make_uint4(c[0], c[1], c[2], c[3]);
#pragma unroll
            for (int k = 0; k < 4; ++k) Vs[uu][k] = V[k];
            vmc[uu] = hv3;          // centre row (i - 3): max masked value and m_ub
            mub[uu] = hi3 >> 8;
            hv3 = hv2; hv2 = hv1; hv1 = vm;
            hi3 = hi2; hi2 = hi1; hi1 = info;
        }
        // ---- phase B: horizontal sums, group threshold, pass flag (no branches) --------------
        uint32_t Hx[RB][8];
        unsigned ball[RB];
#pragma unroll
        for (int uu = 0; uu < RB; ++uu) {
            const int u = hb * RB + uu;  // row of the stage
            const int i = g * SR + u;  // load-row index within the task
            const int out_row = r0 + i - 2 * KH;
            const uint32_t l2 = __shfl_up_sync(0xffffffffu, Vs[uu][2], 1);
            const uint32_t l3 = __shfl_up_sync(0xffffffffu, Vs[uu][3], 1);
            const uint32_t rr0 = __shfl_down_sync(0xffffffffu, Vs[uu][0], 1);
            const uint32_t rr1 = __shfl_down_sync(0xffffffffu, Vs[uu][1], 1);
#if FUSED_DP2A
            // 7-tap sums of 16-bit halves with IDP.2A (d = c + a.lo * b0 + a.hi * b1): the
            // integer dot product issues on the FMA pipe, which this loop otherwise leaves idle,
            // while every unpack / add it replaces would queue on the ALU pipe.
            // Words: W0 = l2, W1 = l3, W2..W5 = own V[0..3], W6 = rr0, W7 = rr1; window j = halves j .. j+6
            // of the 16-half sequence, starting at W0.hi.
            constexpr uint32_t LO = 0x0001u, HI = 0x0100u, BOTH = 0x0101u;
            const uint32_t* Wv = Vs[uu];
            const uint32_t a23 = __dp2a_lo(Wv[0], BOTH, __dp2a_lo(Wv[1], BOTH, 0u));
            const uint32_t a45 = __dp2a_lo(Wv[2], BOTH, __dp2a_lo(Wv[3], BOTH, 0u));
            const uint32_t t1 = __dp2a_lo(l3, BOTH, a23);      // W1 + W2 + W3
            const uint32_t t2 = __dp2a_lo(Wv[2], BOTH, a23);   // W2 + W3 + W4
            const uint32_t t3 = __dp2a_lo(Wv[1], BOTH, a45);   // W3 + W4 + W5
            const uint32_t t4 = __dp2a_lo(rr0, BOTH, a45);     // W4 + W5 + W6
            Hx[uu][0] = __dp2a_lo(l2, HI, t1);
            Hx[uu][1] = __dp2a_lo(Wv[2], LO, t1);
            Hx[uu][2] = __dp2a_lo(l3, HI, t2);
            Hx[uu][3] = __dp2a_lo(Wv[3], LO, t2);
            Hx[uu][4] = __dp2a_lo(Wv[0], HI, t3);
            Hx[uu][5] = __dp2a_lo(rr0, LO, t3);
            Hx[uu][6] = __dp2a_lo(Wv[1], HI, t4);
            Hx[uu][7] = __dp2a_lo(rr1, LO, t4);
            const uint32_t xmin = min(min(min(Hx[uu][0], Hx[uu][1]), min(Hx[uu][2], Hx[uu][3])),
                                      min(min(Hx[uu][4], Hx[uu][5]), min(Hx[uu][6], Hx[uu][7])));
#else
            uint32_t E[14];
            E[0] = l2 >> 16;
            E[1] = l3 & 0xffffu;
            E[2] = l3 >> 16;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                E[3 + 2 * k] = Vs[uu][k] & 0xffffu;
                E[4 + 2 * k] = Vs[uu][k] >> 16;
            }
            E[11] = rr0 & 0xffffu;
            E[12] = rr0 >> 16;
            E[13] = rr1 & 0xffffu;
            Hx[uu][0] = E[0] + E[1] + E[2] + E[3] + E[4] + E[5] + E[6];
            uint32_t xmin = Hx[uu][0];
#pragma unroll
            for (int j = 1; j < 8; ++j) {
                Hx[uu][j] = Hx[uu][j - 1] + E[j + 6] - E[j - 1];
                xmin = min(xmin, Hx[uu][j]);
            }
#endif
            // group threshold on the raw value: T = floor(mu + nsig_s' sqrt(mu)), mu = x_lb / m_ub
            const float mu = (float)xmin * rcp_tab[mub[uu]];
            float sq;
            asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(mu));
            int T = max(__float2int_rz(fmaf(sk, sq, mu)), thr_floor);
            if constexpr (sizeof(PixelT) == 2 && FUSED_LUT_MASK) T = min(T, (int)G::CLAMP - 1);  // vmc is clamped
            ball[uu] = __ballot_sync(0xffffffffu,
                                    out_lane && vmc[uu] > (uint32_t)T && i >= 2 * KH && out_row < row_end);
        }
        // poll the next stage's barrier now; the answer arrives while phase C runs
        if (hb == SR / RB - 1) {
            const int sn = (s + 1 == STAGES) ? 0 : s + 1;
            ready = g + 1 < n_groups && ptx::mbar_try_wait(bars + sn, (s + 1 == STAGES) ? phase ^ 1u : phase);
        }
        // ---- phase C: queue the passing groups (predicated stores, no divergence); resolve
        // when 32 are pending -- checked every two rows, which GQ_CAP allows for ----------------
        const uint32_t gq_s = ptx::smem_u32(gq), hq_s = ptx::smem_u32(hq);
#pragma unroll
        for (int uu = 0; uu < RB; ++uu) {
            const int u = hb * RB + uu;  // row of the stage
            const int qi = qn + __popc(ball[uu] & lt_mask);
            ptx::push_group_if(ball[uu] & (1u << lane), gq_s + 4 * qi,
                               (uint32_t)lane | ((uint32_t)(g * SR + u - 2 * KH) << 5) | (mub[uu] << 14),
                               hq_s + 32 * qi, Hx[uu]);
            qn += __popc(ball[uu]);
#if FUSED_DEBUG_MODES
            if (a.force_brute == 3) qn = 0;  // timing experiment: streaming loop + queue pushes only
#endif
#if FUSED_DEBUG_MODES
            if (a.force_brute == 3) qn = 0;  // timing experiment: streaming loop + queue pushes only
#endif
            if ((u & 1) && qn >= 32) {
                __syncwarp();
                do {
                    const uint32_t st2 = fused_resolve<PixelT>(&sa, qbase, rcp_tab, x0, r0, z,
                                                               (uint32_t)qn | pstate, 32, false);
                    qn = st2 & 0xff;
                    pstate = st2 & ~0xffu;
                } while (qn >= 32);
            }
        }
                }
        }
        // refill the stage we just consumed
        // Every lane's reads of this stage have completed (their values were consumed above) and
        // __syncwarp orders them before the elected lane's TMA write, as an mbarrier release /
        // acquire pair would in a producer-consumer pipeline; no cross-proxy fence is needed for
        // this read-then-async-write reuse.
        __syncwarp();
        if (g + STAGES < n_groups && ptx::elect_one()) issue(g + STAGES, s);
        if (++s == STAGES) {
            s = 0;
            phase ^= 1u;
        }
    }
    __syncwarp();
    if (qn > 0 || pstate)
        fused_resolve<PixelT>(&sa, qbase, rcp_tab, x0, r0, z, (uint32_t)qn | pstate, qn, true);
}

}  // namespace ffsb
