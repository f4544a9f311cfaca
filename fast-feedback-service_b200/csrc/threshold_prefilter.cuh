// threshold_prefilter.cuh -- kernel 1 of threshold.cuh (see the header comment there).
#pragma once

namespace ffsb {

// Flush `n` (<= 32) queue entries starting at q[first] to the frame's global candidate list.
__device__ __forceinline__ void flush_candidates(const ThreshArgs& a, int z, const uint32_t* q,
                                                 int first, int n, int lane) {
    unsigned base = 0;
    if (lane == 0) base = atomicAdd(a.cand_count + (size_t)z * a.zstride_words, (unsigned)n);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (lane < n && base + lane < a.cand_cap)
        a.cand[(size_t)z * a.cand_cap + base + lane] = q[first + lane];
    // an overflowing list is detected by resolve_kernel from cand_count > cand_cap
}

template <typename PixelT, int KH, int PPL, int SR, int STAGES, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
  prefilter_kernel(const __grid_constant__ CUtensorMap tm_img,
                   const __grid_constant__ CUtensorMap tm_valid, const ThreshArgs a) {
    using G = ThreshGeo<PixelT, KH, PPL, SR>;
    constexpr int KS = G::KS;
    constexpr bool P16 = sizeof(PixelT) == 2;
    static_assert(P16 ? PPL == 8 : PPL == 4, "one 16-byte shared load per lane per row");

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const long long task = (long long)blockIdx.x * WARPS + warp;
    const long long n_tasks = (long long)a.n_frames * a.n_row_tiles * a.n_strips;
    if (task >= n_tasks) return;  // whole warp leaves; warps never synchronise with each other
    const int strip = (int)(task % a.n_strips);
    const int rt = (int)((task / a.n_strips) % a.n_row_tiles);
    const int z = a.slot0 + (int)(task / ((long long)a.n_strips * a.n_row_tiles));

    // TMA destinations need 128-byte alignment: align the dynamic window ourselves
    uint8_t* smem_al = smem_raw + ((128u - (ptx::smem_u32(smem_raw) & 127u)) & 127u);
    uint8_t* wsm = smem_al + (size_t)warp * G::smem_per_warp(STAGES);
    uint8_t* stage_base = wsm;
    uint8_t* ring = wsm + STAGES * G::STAGE_BYTES;
    uint32_t* queue = reinterpret_cast<uint32_t*>(ring + G::RING_BYTES);
    uint64_t* bars = reinterpret_cast<uint64_t*>(ring + G::RING_BYTES + G::QUEUE_BYTES);
    int* qcount = reinterpret_cast<int*>(bars + STAGES);

    const int r0 = rt * a.tile_rows;                      // first output row
    const int row_end = min(r0 + a.tile_rows, a.height);  // one past last output row
    const int n_load_rows = (row_end - r0) + 2 * KH;
    const int n_groups = (n_load_rows + SR - 1) / SR;
    const int x0 = strip * G::OUT_COLS - G::HL * PPL;      // first loaded column (may be < 0)
    const int y_first = r0 - KH;                           // first loaded row   (may be < 0)
    const int vbyte_first = (x0 - G::VPAD) >> 3;           // x0 - VPAD is a multiple of 8
    const int vbyte0 = vbyte_first & ~15;                  // 16-byte aligned TMA start (may be < 0)
    const int voff = vbyte_first - vbyte0;

    if (lane == 0) {
        for (int s = 0; s < STAGES; ++s) ptx::mbar_init(bars + s, 1);
        *qcount = 0;
        ptx::fence_mbar_init();
    }
    // zero this warp's ring (lane-private addresses from here on)
    {
        uint4 zero = make_uint4(0, 0, 0, 0);
        for (int k = lane; k < G::RING_BYTES / 16; k += 32) reinterpret_cast<uint4*>(ring)[k] = zero;
    }
    __syncwarp();

    auto issue = [&](int g) {
        const int s = g % STAGES;
        uint8_t* dst = stage_base + s * G::STAGE_BYTES;
        ptx::mbar_arrive_expect_tx(bars + s, G::TX_BYTES);
        ptx::tma_load_3d(dst, &tm_img, bars + s, x0, y_first + g * SR, z);
        ptx::tma_load_3d(dst + G::IMG_STAGE_BYTES, &tm_valid, bars + s, vbyte0, y_first + g * SR,
                         a.valid_frame_stride ? z : 0);
    };
    if (lane == 0) {
        for (int g = 0; g < STAGES && g < n_groups; ++g) issue(g);
    }

    uint32_t V[PPL];
#pragma unroll
    for (int j = 0; j < PPL; ++j) V[j] = 0;
    uint32_t Vc = 0;  // 32-bit pixels: byte-packed per-column counts

    const bool out_lane = lane >= G::HL && lane < 32 - G::HL;
    const int col0 = x0 + lane * PPL;  // this lane's first column
    const int bitoff = G::VPAD + lane * PPL;
    const int vbyte_lane = voff + (bitoff >> 3);
    const int vshift = bitoff & 7;
    uint8_t* ring_lane = ring + lane * 16;
    uint8_t* ringc_lane = ring + KS * G::RING_ROW_BYTES + lane * 8;  // 32-bit pixels only

    int slot = 0;             // ring slot of the incoming row  (i % KS)
    int cslot = KS - KH;      // ring slot of the centre row    ((i - KH) % KS)
    int i = 0;                // load-row index within the task
    for (int g = 0; g < n_groups; ++g) {
        const int s = g % STAGES;
        ptx::mbar_wait(bars + s, (uint32_t)((g / STAGES) & 1));
        const uint8_t* simg = stage_base + s * G::STAGE_BYTES;
        const uint8_t* sval = simg + G::IMG_STAGE_BYTES;

#pragma unroll 1
        for (int u = 0; u < SR; ++u, ++i) {
            // ---- load + unpack one row ------------------------------------------------
            const uint4 raw = ptx::lds128(simg + (u * G::LOAD_COLS + lane * PPL) * sizeof(PixelT));
            const uint32_t mb = ((uint32_t)sval[u * G::VROW_BYTES + vbyte_lane] >> vshift) & ((1u << PPL) - 1u);
            const uint32_t r[4] = {raw.x, raw.y, raw.z, raw.w};
            uint32_t w[PPL];
            uint32_t flags = 0;
            if constexpr (P16) {
#pragma unroll
                for (int j = 0; j < PPL; ++j) {
                    // bytes {v.lo, v.hi, 0x00, 0x01}: value in bits 0..15, count in bit 24
                    const uint32_t t = __byte_perm(r[j >> 1], 0x00000100u, (j & 1) ? 0x5432 : 0x5410);
                    w[j] = (mb & (1u << j)) ? t : 0u;
                }
            } else {
                uint32_t cbits = 0;
#pragma unroll
                for (int j = 0; j < PPL; ++j) {
                    const bool counted = (mb & (1u << j)) && r[j] < (1u << 24);
                    w[j] = counted ? r[j] : 0u;
                    cbits |= counted ? (1u << j) : 0u;
                }
                flags = mb | (cbits << 8);
            }
            // ---- vertical running sum: V += new - old; ring[slot] = new -----------------
            uint8_t* rrow = ring_lane + slot * G::RING_ROW_BYTES;
#pragma unroll
            for (int h = 0; h < G::RING_PLANES; ++h) {
                const uint4 old = ptx::lds128(rrow + h * 512);
                V[4 * h + 0] = V[4 * h + 0] + w[4 * h + 0] - old.x;
                V[4 * h + 1] = V[4 * h + 1] + w[4 * h + 1] - old.y;
                V[4 * h + 2] = V[4 * h + 2] + w[4 * h + 2] - old.z;
                V[4 * h + 3] = V[4 * h + 3] + w[4 * h + 3] - old.w;
                *reinterpret_cast<uint4*>(rrow + h * 512) =
                  make_uint4(w[4 * h + 0], w[4 * h + 1], w[4 * h + 2], w[4 * h + 3]);
            }
            if constexpr (!P16) {
                // spread 4 count bits to 4 bytes (bit k -> byte k) and keep the row's flags
                const uint32_t cspread = (((flags >> 8) & 0xfu) * 0x00204081u) & 0x01010101u;
                uint2* rc = reinterpret_cast<uint2*>(ringc_lane + slot * 256);
                const uint2 oldc = *rc;
                Vc = Vc + cspread - oldc.x;
                *rc = make_uint2(cspread, flags);
            }

            // ---- test the centre row of the window that is now complete -----------------
            const int out_row = r0 + i - 2 * KH;
            if (i >= 2 * KH && out_row < row_end) {  // warp-uniform
                uint32_t cb = 0;                     // candidate bits of this lane
                uint32_t Hx[PPL];
                {
                    uint32_t E[PPL + 2 * KH];
#pragma unroll
                    for (int k = 1; k <= KH; ++k) {
                        E[KH - k] = __shfl_up_sync(0xffffffffu, V[PPL - k], 1);
                        E[KH + PPL - 1 + k] = __shfl_down_sync(0xffffffffu, V[k - 1], 1);
                    }
#pragma unroll
                    for (int j = 0; j < PPL; ++j) E[KH + j] = V[j];
                    uint32_t h = 0;
#pragma unroll
                    for (int t = 0; t < KS; ++t) h += E[t];
                    Hx[0] = h;
#pragma unroll
                    for (int j = 1; j < PPL; ++j) {
                        h = h + E[j + 2 * KH] - E[j - 1];
                        Hx[j] = h;
                    }
                }
                const uint8_t* crow = ring_lane + cslot * G::RING_ROW_BYTES;
                if constexpr (P16) {
                    // H = m << 24 | x ;  b = m*v - x = m*w_c - H  (w_c = 1<<24 | v, or 0)
                    uint32_t wc[PPL];
#pragma unroll
                    for (int h = 0; h < G::RING_PLANES; ++h) {
                        const uint4 c4 = ptx::lds128(crow + h * 512);
                        wc[4 * h + 0] = c4.x;
                        wc[4 * h + 1] = c4.y;
                        wc[4 * h + 2] = c4.z;
                        wc[4 * h + 3] = c4.w;
                    }
                    int b[PPL];
                    uint32_t pmin = 0xffffffffu;
#pragma unroll
                    for (int j = 0; j < PPL; ++j) {
                        const uint32_t m = Hx[j] >> 24;
                        b[j] = (int)(m * wc[j] - Hx[j]);
                        pmin = min(pmin, (Hx[j] & 0xffffffu) * m);
                    }
                    // b > d  =>  b > sqrt(k * x*m) for every k < nsig_s^2, x*m >= pmin
                    const int T = (int)(sqrtf(a.prefilter_k * (float)pmin) * 0.99999f);
#pragma unroll
                    for (int j = 0; j < PPL; ++j) cb |= (b[j] > T) ? (1u << j) : 0u;
                } else {
                    const uint4 c4 = ptx::lds128(crow);
                    const uint32_t wc[4] = {c4.x, c4.y, c4.z, c4.w};
                    const uint32_t cflags = reinterpret_cast<const uint2*>(ringc_lane + cslot * 256)->y;
                    const uint32_t cl = __shfl_up_sync(0xffffffffu, Vc, 1);
                    const uint32_t cr = __shfl_down_sync(0xffffffffu, Vc, 1);
                    const uint32_t call = __dp4a(Vc, 0x01010101u, 0u);
                    uint32_t Hm[4];
                    // columns -3..-1 are bytes 1..3 of the left lane, +1..+3 bytes 0..2 of the right
                    Hm[0] = __dp4a(cl, 0x01010100u, call);
                    Hm[1] = __dp4a(cl, 0x01010000u, __dp4a(cr, 0x00000001u, call));
                    Hm[2] = __dp4a(cl, 0x01000000u, __dp4a(cr, 0x00000101u, call));
                    Hm[3] = __dp4a(cr, 0x00010101u, call);
                    float pmin = 3.0e38f;
                    float bf[PPL], sf[PPL];
#pragma unroll
                    for (int j = 0; j < PPL; ++j) {
                        const float mf = (float)Hm[j], xf = (float)Hx[j];
                        const float mv = mf * (float)wc[j];
                        bf[j] = mv - xf;
                        sf[j] = mv + xf;
                        pmin = fminf(pmin, xf * mf);
                    }
                    const float tf = sqrtf(a.prefilter_k * pmin) * 0.9999f;
#pragma unroll
                    for (int j = 0; j < PPL; ++j) {
                        const bool mask_ok = cflags & (1u << j);
                        const bool counted = cflags & (0x100u << j);
                        // float error of bf is below 4e-7 * (m*v + x); a mask-valid pixel
                        // >= 2^24 (not counted) is always a candidate
                        const bool c = mask_ok && (!counted || bf[j] + 1e-6f * sf[j] > tf);
                        cb |= c ? (1u << j) : 0u;
                    }
                }
                if (!out_lane) cb = 0;
                if (__any_sync(0xffffffffu, cb != 0)) {
                    while (cb) {
                        const int j = __ffs(cb) - 1;
                        cb &= cb - 1;
                        const int qs = atomicAdd(qcount, 1);
                        queue[qs] = (uint32_t)(col0 + j) | ((uint32_t)out_row << 13);
                    }
                    __syncwarp();
                    int qn = *qcount;
                    if (qn >= 32) {
                        int first = 0;
                        for (; qn - first >= 32; first += 32) flush_candidates(a, z, queue, first, 32, lane);
                        const int rem = qn - first;  // < 32: move the tail to the front
                        const uint32_t t0 = lane < rem ? queue[first + lane] : 0u;
                        __syncwarp();
                        if (lane < rem) queue[lane] = t0;
                        if (lane == 0) *qcount = rem;
                        __syncwarp();
                    }
                }
            }
            slot = (slot + 1 == KS) ? 0 : slot + 1;
            cslot = (cslot + 1 == KS) ? 0 : cslot + 1;
        }
        // ---- refill the stage we just consumed ---------------------------------------------
        __syncwarp();
        if (lane == 0 && g + STAGES < n_groups) {
            ptx::fence_proxy_async();
            issue(g + STAGES);
        }
    }
    __syncwarp();
    const int qn = *qcount;
    if (qn > 0) flush_candidates(a, z, queue, 0, qn, lane);
}

}  // namespace ffsb
