// extended_stream.cuh -- first pass of DispersionExtendedThreshold as an HBM-streaming kernel.
//
//   compute_dispersion_threshold (baseline/spotfinder/baseline.cpp:415-475):
//       D[k] = mask[k] && m >= min_count && x >= 0 && a > c
//       a = m*y - x*x - x*(m-1),  c = x * nsig_b * sqrt(2*(m-1))       (7x7 window sums m, x, y)
//
// Unlike the standard algorithm there is no sparse signal test to prune with: every pixel
// needs its own window sums.  The kernel shares the fused dispersion kernel's skeleton (one
// warp per 240-column strip, TMA-fed 4-row stages, 8-row shared ring, no block barriers) and
// forms EXACT separable box sums for all pixels whose window holds no pixel above 1169:
//   x : values packed two per word, vertical running sum in packed form, horizontal 7-tap
//       sums from warp shuffles (as in dispersion_fused.cuh);
//   y : squares (<= 1169^2, so a 49-pixel sum stays below 2^26) as 32-bit running column sums
//       -- the outgoing row's squares are recomputed from the ring's packed values -- and
//       7-tap sums from 6 shuffles;
//   m : the number of mask-valid pixels in the clipped window is a property of the mask,
//       precomputed once (group_info_kernel): the largest and smallest count over each group's
//       valid pixels travel in the 4-byte group info the TMA delivers beside the frame; the
//       per-pixel counts are read from the (L2-resident) plane by the rows that need them.
// The comparison  m*y - x^2 > x * K(m),  K(m) = m-1 + nsig_b*sqrt(2(m-1)),  is made in fp32 with
// an explicit error bound; only pixels inside that bound, and windows that contain a pixel
// above 1169 (or, for 32-bit data, an uncounted pixel >= 2^24), are queued and decided by the
// exact integer/fp64 evaluation (resolve_pixel<.., EXT_PASS1 = true>).  The D plane is written
// as one byte per pixel group; queued pixels are OR-ed in afterwards.
#pragma once
#include "dispersion_fused.cuh"

#ifndef EXT_MIN_CTAS
#define EXT_MIN_CTAS 5  // 3-warp CTAs per SM the first pass is compiled for; 6 (<= 112 registers) spills: 77.9 vs 55.2 us/frame
#endif
#ifndef EXT_ROW_UNROLL
#define EXT_ROW_UNROLL 2  // rows of a stage unrolled together (4 = whole stage: more code, no faster)
#endif

namespace ffsb {

template <typename PixelT>
struct ExtGeo {
    using F = FusedGeo<PixelT>;
    static constexpr int SR = 4;
    static constexpr unsigned CLAMP = 1169u;     // 49 * 1169^2 < 2^26
    // group info, 4 bytes per group {validity, m_ub | full << 6, m_lb, 0}: 32 groups + <= 12 B alignment slack
    static constexpr int INFO4_ROW_BYTES = 144;
    static constexpr int INFO4_STAGE_BYTES = (SR * INFO4_ROW_BYTES + 127) / 128 * 128;
    static constexpr int STAGE_BYTES = F::IMG_STAGE_BYTES + INFO4_STAGE_BYTES;
    static constexpr int TX_BYTES = F::IMG_STAGE_BYTES + SR * INFO4_ROW_BYTES;
    static constexpr int GQ_CAP = 64, PQ_CAP = 288;
    static constexpr int QUEUE_BYTES = (GQ_CAP + PQ_CAP) * 4;
    static_assert(QUEUE_BYTES % 128 == 0, "alignment");
    __host__ __device__ static constexpr int smem_per_warp(int stages) {
        return stages * STAGE_BYTES + F::RING_BYTES + QUEUE_BYTES + 128;
    }
};

// Exact decision for queued pixels (out of line: rare path).  state = qn | pqn << 8.
template <typename PixelT>
__device__ __noinline__ uint32_t ext_resolve(const ThreshArgs* __restrict__ ap, uint32_t* gq, int x0, int r0,
                                             int z, uint32_t state, int n, bool flush_all) {
    using E = ExtGeo<PixelT>;
    const ThreshArgs& a = *ap;
    const int lane = threadIdx.x & 31;
    uint32_t* pq = gq + E::GQ_CAP;
    const uint8_t* frame = a.image + (size_t)z * a.frame_stride;
    int qn = state & 0xff, pqn = state >> 8;
    unsigned found = 0;
    auto resolve_entry = [&](uint32_t e) {
        const int col = e & 0x1fff, row = e >> 13;
        if (a.tmax != 0xffffffffu) {  // a mask-valid centre above the trusted range (trusted.cuh)
            const uint32_t vc = reinterpret_cast<const PixelT*>(frame + (size_t)row * a.pitch)[col];
            if (vc > a.tmax && ((a.valid_bits[(size_t)row * a.valid_pitch + (col >> 3)] >> (col & 7)) & 1u))
                note_bad_pixel(a, z, col, row);
        }
        if (resolve_pixel<PixelT, 3, true>(a, frame, a.valid_bits, col, row)) {
            set_strong(a, z, col, row);
            ++found;
        }
    };
    if (n > 0) {
        uint32_t bits = 0;
        int col0 = 0, row = 0;
        if (lane < n) {
            const uint32_t e = gq[qn - n + lane];
            bits = e >> 14;
            col0 = x0 + (int)(e & 31u) * 8;
            row = r0 + (int)((e >> 5) & 0x1ffu);
        }
        qn -= n;
        int cnt = __popc(bits), incl = cnt;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off) incl += t;
        }
        int pos = pqn + incl - cnt;
        while (bits) {
            const int j = __ffs(bits) - 1;
            bits &= bits - 1;
            pq[pos++] = (uint32_t)(col0 + j) | ((uint32_t)row << 13);
        }
        pqn += __shfl_sync(0xffffffffu, incl, 31);
        __syncwarp();
        while (pqn >= 32) {
            pqn -= 32;
            resolve_entry(pq[pqn + lane]);
        }
        __syncwarp();
    }
    if (flush_all && pqn > 0) {
        if (lane < pqn) resolve_entry(pq[lane]);
        pqn = 0;
    }
    for (int off = 16; off > 0; off >>= 1) found += __shfl_down_sync(0xffffffffu, found, off);
    if (lane == 0 && found) atomicAdd(a.strong_count + (size_t)z * a.zstride_words, found);
    return (uint32_t)qn | ((uint32_t)pqn << 8);
}

// A row that survived the integer prune (some lane holds a pixel whose test is not clearly false):
// per-pixel fp32 comparison with an explicit error bound; pixels inside the bound, and windows
// whose sums are not exact (flagged), go to the exact routine.  Out of line on purpose: inlined
// into the four-row loop body it pushed the hot loop out of the instruction cache
// (stall_no_instruction was the top stall reason, profiles/r2_ext_pass1_ncu_full.json).
// state = qn | pqn << 8; returns {new state, D bits set by this lane}.
template <typename PixelT>
__device__ __noinline__ uint2 ext_row_slow(const ThreshArgs* __restrict__ ap, const float4* __restrict__ ktab,
                                           uint32_t* gq, uint32_t state, uint4 hxa, uint4 hxb, uint4 hya, uint4 hyb,
                                           uint2 mw, uint32_t vb, uint32_t flags, int out_row, int group, int x0,
                                           int r0, int z) {
    const ThreshArgs& a = *ap;  // mw: the per-pixel window counts of this group's 8 pixels (prefetched by the caller)
    const int lane = threadIdx.x & 31;
    const bool flagged = flags & 1u, out_lane = flags & 2u, store_lane = flags & 4u;
    const uint32_t Hx[8] = {hxa.x, hxa.y, hxa.z, hxa.w, hxb.x, hxb.y, hxb.z, hxb.w};
    const uint32_t Hy[8] = {hya.x, hya.y, hya.z, hya.w, hyb.x, hyb.y, hyb.z, hyb.w};
    int qn = state & 0xff, pqn = state >> 8;
    uint32_t dbits = 0, border = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const uint32_t m = ((j < 4 ? mw.x : mw.y) >> (8 * (j & 3))) & 0xffu;
        const float4 mk = ktab[m];
        const float xf = (float)Hx[j], yf = (float)Hy[j];
        const float t = mk.x * yf, uu = xf * xf, rhs_lo = xf * mk.y, rhs_hi = xf * mk.z;
        const float lhs = t - uu;
        const float sl = 4e-7f * (t + uu + rhs_hi);
        const bool ok = ((vb >> j) & 1u) && (int)m >= a.min_count;
        const bool hi = (lhs - sl) > rhs_hi, lo = (lhs + sl) > rhs_lo;
        dbits |= (ok && hi) ? (1u << j) : 0u;
        border |= (ok && lo && !hi) ? (1u << j) : 0u;
    }
    if (flagged) {  // sums are not exact here: decide every valid pixel exactly
        dbits = 0;
        border = vb;
    }
    uint8_t* dplane = reinterpret_cast<uint8_t*>(a.strong_bits) + (size_t)z * a.strong_frame_stride;
    if (store_lane) dplane[(size_t)out_row * a.strong_pitch + group] = (uint8_t)dbits;
    const bool push = out_lane && border != 0;
    const unsigned ball = __ballot_sync(0xffffffffu, push);
    if (ball) {
        if (push)
            gq[qn + __popc(ball & ((1u << lane) - 1u))] =
              (uint32_t)lane | ((uint32_t)(out_row - r0) << 5) | (border << 14);
        qn += __popc(ball);
        __syncwarp();
        if (qn >= 32) {
            const uint32_t st2 = ext_resolve<PixelT>(ap, gq, x0, r0, z, (uint32_t)qn | ((uint32_t)pqn << 8), 32, false);
            qn = st2 & 0xff;
            pqn = st2 >> 8;
        }
    }
    return make_uint2((uint32_t)qn | ((uint32_t)pqn << 8), out_lane ? (uint32_t)__popc(dbits) : 0u);
}

template <typename PixelT, int STAGES, int WARPS>
#ifdef EXT_MAXNREG  // A/B builds: an explicit register cap instead of the resident-CTA hint
__global__ void __maxnreg__(EXT_MAXNREG)
#else
__global__ void __launch_bounds__(WARPS * 32, WARPS == 3 ? EXT_MIN_CTAS : 1)
#endif
  ext_pass1_kernel(const __grid_constant__ CUtensorMap tm_img, const __grid_constant__ CUtensorMap tm_info4,
                   const __grid_constant__ ThreshArgs a) {
    using G = FusedGeo<PixelT>;
    using E = ExtGeo<PixelT>;
    constexpr int SR = 4, KH = 3, PPL = 8;
    constexpr uint32_t CL = E::CLAMP;
    constexpr int kRowUnroll = EXT_ROW_UNROLL;

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // {m, K(m)*g_lo, K(m)*g_hi},  K(m) = m-1 + nsig_b*sqrt(2(m-1)).  Without a gain map g_lo = g_hi = 1;
    // with one (baseline.cpp:536: c = gain*x*K(m)) the smallest / largest gain bracket the
    // right-hand side and the pixels in between are decided exactly.
    __shared__ float4 ktab[64];
    // floor(K(m) * g_lo), rounded down with a margin: m*y - x^2 <= x * ktabi[m'] for some m' <= m
    // proves a <= c, i.e. D = 0, in exact integer arithmetic (see the row prune below)
    __shared__ uint32_t ktabi[64];
    __shared__ ThreshArgs sa;  // the out-of-line paths read their arguments from here (see dispersion_fused.cuh)
    // 32-bit front end: validity nibble -> four AND masks (one LDS.128 instead of a test + select per pixel)
    __shared__ __align__(16) uint4 vmask_lut[16];
    if (threadIdx.x < 16) {
        const uint32_t n = threadIdx.x;
        vmask_lut[n] = make_uint4((n & 1u) ? ~0u : 0u, (n & 2u) ? ~0u : 0u, (n & 4u) ? ~0u : 0u, (n & 8u) ? ~0u : 0u);
    }
    for (int t = threadIdx.x; t < (int)(sizeof(ThreshArgs) / 4); t += WARPS * 32)
        reinterpret_cast<uint32_t*>(&sa)[t] = reinterpret_cast<const uint32_t*>(&a)[t];
    for (int t = threadIdx.x; t < 64; t += WARPS * 32) {
        const float m1 = (float)max(t - 1, 0);
        const float k = m1 + (float)a.nsig_b * sqrtf(2.0f * m1);
        ktab[t] = make_float4((float)t, k * a.g_lo, k * a.g_hi, 0.f);
        const double kd = ((double)max(t - 1, 0) + a.nsig_b * sqrt(2.0 * (double)max(t - 1, 0))) * (double)a.g_lo;
        ktabi[t] = (uint32_t)floor(fmax(kd * (1.0 - 1e-6), 0.0));
    }
    __syncthreads();

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const long long task = (long long)blockIdx.x * WARPS + warp;
    const long long n_tasks = (long long)a.n_frames * a.n_row_tiles * a.n_strips;
    if (task >= n_tasks) return;
    const int strip = (int)(task % a.n_strips);
    const int rt = (int)((task / a.n_strips) % a.n_row_tiles);
    const int z = a.slot0 + (int)(task / ((long long)a.n_strips * a.n_row_tiles));

    uint8_t* smem_al = smem_raw + ((128u - (ptx::smem_u32(smem_raw) & 127u)) & 127u);
    uint8_t* wsm = smem_al + (size_t)warp * E::smem_per_warp(STAGES);
    uint8_t* stage_base = wsm;
    uint8_t* ring = wsm + STAGES * E::STAGE_BYTES;
    uint32_t* gq = reinterpret_cast<uint32_t*>(ring + G::RING_BYTES);
    uint64_t* bars = reinterpret_cast<uint64_t*>(ring + G::RING_BYTES + E::QUEUE_BYTES);

    const int r0 = rt * a.tile_rows;
    const int row_end = min(r0 + a.tile_rows, a.height);
    const int n_load_rows = (row_end - r0) + 2 * KH;
    const int n_groups = (n_load_rows + SR - 1) / SR;
    const int x0 = strip * G::OUT_COLS - PPL;
    const int y_first = r0 - KH;
    const int info_b = 4 * (strip * G::OUT_LANES - 1);  // byte offset of lane 0's group in an info4 row
    const int info_b0 = info_b & ~15;

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
        uint8_t* dst = stage_base + s * E::STAGE_BYTES;
        ptx::mbar_arrive_expect_tx(bars + s, E::TX_BYTES);
        ptx::tma_load_3d(dst, &tm_img, bars + s, x0, y_first + g * SR, z);
        ptx::tma_load_2d(dst + G::IMG_STAGE_BYTES, &tm_info4, bars + s, info_b0, y_first + g * SR);
    };
    if (lane == 0) {
        for (int g = 0; g < STAGES && g < n_groups; ++g) issue(g, g);
    }

    const bool out_lane = lane >= 1 && lane <= G::OUT_LANES;
    const int group = strip * G::OUT_LANES + lane - 1;            // pixel group of this lane
    const bool store_lane = out_lane && group < (int)a.strong_pitch;
    int qn = 0, pqn = 0;
    unsigned n_set = 0;

    uint32_t V[4] = {0, 0, 0, 0};                  // packed column sums of x over the last 7 rows
    uint32_t Vy[8] = {0, 0, 0, 0, 0, 0, 0, 0};     // column sums of the squares
    uint32_t big7 = 0;                             // bit r: row i-r of this lane holds a pixel > CLAMP
    uint32_t hi1 = 0, hi2 = 0, hi3 = 0;            // group info of rows i-1, i-2, i-3
    // per-pixel window counts (one byte each) of this lane's group in rows i-1, i-2, i-3.  A group whose
    // valid pixels all have a full window (info bit 14) is 49 everywhere; the others -- next to a module
    // gap, the image edge or a dead pixel -- read theirs from the mask's count plane (L2-resident) when
    // the row is LOADED, three row iterations before it is the centre row: the latency never shows.
    constexpr uint32_t kFull4 = 0x31313131u;  // 49, 49, 49, 49
    uint2 mq1 = make_uint2(kFull4, kFull4), mq2 = mq1, mq3 = mq1;
    const uint32_t kf49 = ktabi[49];
    uint8_t* ring_lane = ring + lane * 16;
    const int pimg_off = lane * PPL * (int)sizeof(PixelT);
    const int pinfo_off = G::IMG_STAGE_BYTES + (info_b - info_b0) + 4 * lane;

    // this lane's 8 window counts in load row i (read only where the group info says "not full")
    const uint8_t* mrow = a.mcount + (ptrdiff_t)y_first * (ptrdiff_t)a.m_pitch + 8 * (ptrdiff_t)group;
    int s = 0;
    uint32_t phase = 0;
    for (int g = 0; g < n_groups; ++g) {
        ptx::mbar_wait(bars + s, phase);
        const uint8_t* st = stage_base + s * E::STAGE_BYTES;
        uint8_t* rcur = ring_lane + ((g & 1) ? SR * G::RING_ROW_BYTES : 0);
        uint8_t* roth = ring_lane + ((g & 1) ? 0 : SR * G::RING_ROW_BYTES);
#pragma unroll kRowUnroll
        for (int u = 0; u < SR; ++u) {
            const int i = g * SR + u;
            // ---- load, mask, clamp ------------------------------------------------------------
            const uint32_t info = *reinterpret_cast<const uint32_t*>(st + pinfo_off + u * E::INFO4_ROW_BYTES);
            uint2 mq0 = make_uint2(kFull4, kFull4);
            if ((info & 0xffu) && !(info & 0x4000u))  // (rows / groups outside the image are zero-filled: info == 0)
                mq0 = *reinterpret_cast<const uint2*>(mrow);
            mrow += a.m_pitch;
            uint32_t c[4];
            bool big;
            if constexpr (sizeof(PixelT) == 2) {
                const uint4 raw = ptx::lds128(st + u * G::IMG_ROW_BYTES + pimg_off);
                const uint32_t zlo = (info & 0xfu) * 0x10204080u;
                const uint32_t zhi = ((info >> 4) & 0xfu) * 0x10204080u;
                const uint32_t w0 = raw.x & ptx::prmt(zlo, 0u, 0x9988u);
                const uint32_t w1 = raw.y & ptx::prmt(zlo, 0u, 0xbbaau);
                const uint32_t w2 = raw.z & ptx::prmt(zhi, 0u, 0x9988u);
                const uint32_t w3 = raw.w & ptx::prmt(zhi, 0u, 0xbbaau);
                constexpr uint32_t CL2 = CL | (CL << 16);
                c[0] = __vminu2(w0, CL2);
                c[1] = __vminu2(w1, CL2);
                c[2] = __vminu2(w2, CL2);
                c[3] = __vminu2(w3, CL2);
                big = ((w0 ^ c[0]) | (w1 ^ c[1]) | (w2 ^ c[2]) | (w3 ^ c[3])) != 0;
            } else {
                const uint4 ra = ptx::lds128(st + u * G::IMG_ROW_BYTES + pimg_off);
                const uint4 rb = ptx::lds128(st + u * G::IMG_ROW_BYTES + pimg_off + 16);
                const uint4 mlo = vmask_lut[info & 0xfu], mhi = vmask_lut[(info >> 4) & 0xfu];
                const uint32_t v[8] = {ra.x & mlo.x, ra.y & mlo.y, ra.z & mlo.z, ra.w & mlo.w,
                                       rb.x & mhi.x, rb.y & mhi.y, rb.z & mhi.z, rb.w & mhi.w};
                uint32_t cl[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) cl[j] = min(v[j], CL);
                const uint32_t vmax = max(max(max(v[0], v[1]), max(v[2], v[3])), max(max(v[4], v[5]), max(v[6], v[7])));
                big = vmax > CL;  // includes uncounted pixels >= 2^24
#pragma unroll
                for (int k = 0; k < 4; ++k) c[k] = cl[2 * k] | (cl[2 * k + 1] << 16);
            }
            // ---- vertical running sums; the outgoing row's squares come from the ring ------------
            uint8_t* oldp = (u + 1 < SR) ? rcur + (u + 1) * G::RING_ROW_BYTES : roth;
            const uint4 old = ptx::lds128(oldp);
            const uint32_t o[4] = {old.x, old.y, old.z, old.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                V[k] = V[k] + c[k] - o[k];
                const uint32_t nl = c[k] & 0xffffu, nh = c[k] >> 16, ol = o[k] & 0xffffu, oh = o[k] >> 16;
                Vy[2 * k] = Vy[2 * k] + nl * nl - ol * ol;
                Vy[2 * k + 1] = Vy[2 * k + 1] + nh * nh - oh * oh;
            }
            *reinterpret_cast<uint4*>(rcur + u * G::RING_ROW_BYTES) = make_uint4(c[0], c[1], c[2], c[3]);
            big7 = ((big7 << 1) | (big ? 1u : 0u)) & 0x7fu;

            const int out_row = r0 + i - 2 * KH;
            if (i >= 2 * KH && out_row < row_end) {  // warp-uniform
                // windows that hold a clamped pixel: this lane's or a neighbour's last 7 rows
                const unsigned bm = __ballot_sync(0xffffffffu, big7 != 0);
                const bool flagged = ((bm >> (lane > 0 ? lane - 1 : 0)) & (lane > 0 ? 7u : 3u)) != 0;
                // ---- horizontal sums ------------------------------------------------------------
                const uint32_t l2 = __shfl_up_sync(0xffffffffu, V[2], 1);
                const uint32_t l3 = __shfl_up_sync(0xffffffffu, V[3], 1);
                const uint32_t rr0 = __shfl_down_sync(0xffffffffu, V[0], 1);
                const uint32_t rr1 = __shfl_down_sync(0xffffffffu, V[1], 1);
                uint32_t Ey[14];
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    Ey[k] = __shfl_up_sync(0xffffffffu, Vy[5 + k], 1);
                    Ey[11 + k] = __shfl_down_sync(0xffffffffu, Vy[k], 1);
                }
#pragma unroll
                for (int k = 0; k < 8; ++k) Ey[3 + k] = Vy[k];
                uint32_t Hx[8], Hy[8];
                {   // 7-tap sums of the packed 16-bit column sums with IDP.2A (see dispersion_fused.cuh)
                    constexpr uint32_t LO = 0x0001u, HI = 0x0100u, BOTH = 0x0101u;
                    const uint32_t a23 = __dp2a_lo(V[0], BOTH, __dp2a_lo(V[1], BOTH, 0u));
                    const uint32_t a45 = __dp2a_lo(V[2], BOTH, __dp2a_lo(V[3], BOTH, 0u));
                    const uint32_t t1 = __dp2a_lo(l3, BOTH, a23), t2 = __dp2a_lo(V[2], BOTH, a23);
                    const uint32_t t3 = __dp2a_lo(V[1], BOTH, a45), t4 = __dp2a_lo(rr0, BOTH, a45);
                    Hx[0] = __dp2a_lo(l2, HI, t1);
                    Hx[1] = __dp2a_lo(V[2], LO, t1);
                    Hx[2] = __dp2a_lo(l3, HI, t2);
                    Hx[3] = __dp2a_lo(V[3], LO, t2);
                    Hx[4] = __dp2a_lo(V[0], HI, t3);
                    Hx[5] = __dp2a_lo(rr0, LO, t3);
                    Hx[6] = __dp2a_lo(V[1], HI, t4);
                    Hx[7] = __dp2a_lo(rr1, LO, t4);
                }
                Hy[0] = Ey[0] + Ey[1] + Ey[2] + Ey[3] + Ey[4] + Ey[5] + Ey[6];
#pragma unroll
                for (int j = 1; j < 8; ++j) Hy[j] = Hy[j - 1] + Ey[j + 6] - Ey[j - 1];
                // ---- row prune: on background the test fails by a wide margin for every pixel:
                //     m * y <= x * (x + floor(K(m) g_lo))   ==>   m*y - x^2 <= x K(m) g   ==>   D = 0
                // in exact 32-bit integers (m*y < 49 * 2^26, x (x + K) < 3.3e9) with the pixel's own m.
                // Only when some lane of the warp has a pixel that survives (5 % of the rows on the
                // synthetic frames) does the row take the per-pixel fp32 path; the D plane is pre-zeroed,
                // so a pruned row stores nothing.  (A bound with one m per group, m_hi * y <= x (x + K(m_lo)),
                // left 65 % of the rows unpruned: every strip that holds a module-gap or image edge.)
                const uint32_t vb = hi3 & 0xffu;  // validity of the centre row's pixels
                {
                    // 4 instructions per pixel (IMAD, IADD, IMAD, ISETP.OR): validity is not tested per
                    // pixel.  Full groups: a group without a valid pixel is skipped as a whole, and a
                    // (rare) partly valid one can only send the row to the per-pixel path, which tests
                    // validity itself.  Other groups: an invalid pixel's count is masked to m = 0 first,
                    // so its test reads 0 > x (x + K(0)) and never holds.
                    bool any = false;
                    // warp-uniform choice: the general path is right for full groups too (m = 49), so a
                    // row with one group next to a gap runs it alone instead of both paths diverged
                    if (__all_sync(0xffffffffu, mq3.x == kFull4 && mq3.y == kFull4)) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) any |= 49u * Hy[j] > Hx[j] * (Hx[j] + kf49);
                        any = any && vb != 0u;
                    } else {
                        const uint32_t zlo = (vb & 0xfu) * 0x10204080u, zhi = (vb >> 4) * 0x10204080u;
                        const uint32_t ma = mq3.x & ptx::prmt(zlo, 0u, 0xba98u) & 0x3f3f3f3fu;
                        const uint32_t mb = mq3.y & ptx::prmt(zhi, 0u, 0xba98u) & 0x3f3f3f3fu;
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const uint32_t mj = ((j < 4 ? ma : mb) >> (8 * (j & 3))) & 0xffu;
                            any |= mj * Hy[j] > Hx[j] * (Hx[j] + ktabi[mj]);
                        }
                    }
                    const bool maybe = flagged || any;
                    if (__ballot_sync(0xffffffffu, out_lane && maybe) == 0u) {
                        hi3 = hi2; hi2 = hi1; hi1 = info;
                        mq3 = mq2; mq2 = mq1; mq1 = mq0;
                        continue;
                    }
                }
                // ---- the row holds a candidate: per-pixel test, store, queue (out of line) -------------
                const uint2 r2 = ext_row_slow<PixelT>(
                  &sa, ktab, gq, (uint32_t)qn | ((uint32_t)pqn << 8), make_uint4(Hx[0], Hx[1], Hx[2], Hx[3]),
                  make_uint4(Hx[4], Hx[5], Hx[6], Hx[7]), make_uint4(Hy[0], Hy[1], Hy[2], Hy[3]),
                  make_uint4(Hy[4], Hy[5], Hy[6], Hy[7]), mq3, vb,
                  (flagged ? 1u : 0u) | (out_lane ? 2u : 0u) | (store_lane ? 4u : 0u), out_row, group, x0, r0, z);
                qn = r2.x & 0xff;
                pqn = r2.x >> 8;
                n_set += r2.y;  // one atomic per task, at the end
            }
            hi3 = hi2; hi2 = hi1; hi1 = info;
            mq3 = mq2; mq2 = mq1; mq1 = mq0;
        }
        __syncwarp();  // see dispersion_fused.cuh on this read-then-async-write reuse
        if (g + STAGES < n_groups && ptx::elect_one()) issue(g + STAGES, s);
        if (++s == STAGES) {
            s = 0;
            phase ^= 1u;
        }
    }
    __syncwarp();
    for (int off = 16; off > 0; off >>= 1) n_set += __shfl_down_sync(0xffffffffu, n_set, off);
    if (lane == 0 && n_set) atomicAdd(a.strong_count + (size_t)z * a.zstride_words, n_set);
    if (qn > 0 || pqn > 0)
        ext_resolve<PixelT>(&sa, gq, x0, r0, z, (uint32_t)qn | ((uint32_t)pqn << 8), qn, true);
}

}  // namespace ffsb
