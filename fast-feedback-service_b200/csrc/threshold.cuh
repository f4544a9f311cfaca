// threshold.cuh -- DIALS dispersion threshold (DispersionThreshold,
// baseline/spotfinder/standalone.cc:27-206) as a hand-written sm_100a kernel.
//
// Work decomposition
//   A *warp task* is a strip of OUT_COLS output columns x tile_rows output rows of one
//   frame.  Every warp owns its task end to end: its own TMA pipeline (STAGES stages of
//   KS = 2*KH+1 rows, image tile + validity-bit tile, zero-filled out of bounds, which is
//   exactly the reference's window clipping), its own mbarriers and its own candidate
//   queue, so there is no block-level synchronisation anywhere.
//   Lane l owns PPL adjacent columns; the first/last HL lanes are halo lanes.
//
// Arithmetic (bit-exact with the CPU oracle by construction)
//   * window sums are exact integers.  For 16-bit pixels count and sum share one 32-bit
//     word  w = valid ? (1 << 24) | v : 0  (49 * 65535 < 2^24), so ONE separable box sum
//     (vertical running sum in registers, horizontal sum via warp shuffles) gives both
//     m = H >> 24 and x = H & 0xffffff.  32-bit pixels use a sum plane and a byte-packed
//     count plane (dp4a horizontal sums).
//   * every pixel goes through an integer/float *necessary* condition for the
//     signal test  b > d  (b = m*v - x, d = nsig_s*sqrt(x*m)); it is conservative, never
//     exact.  Survivors (~0.5 % of background pixels) are queued per warp and resolved 32
//     at a time in fp64 with round-to-nearest intrinsics in the oracle's operation order
//     (standalone.cc:164-171): the sum of squares y is only ever needed for those, and is
//     gathered there as an exact uint64 from the (L1/L2-resident) tile.
//   * strong pixels are OR-ed into a pre-zeroed 1-bit plane.
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
    uint32_t* strong_bits;      // output plane(s), pre-zeroed
    size_t strong_pitch;        // bytes per row (multiple of 4)
    size_t strong_frame_stride; // bytes per frame plane
    unsigned int* strong_count; // per frame, pre-zeroed
    double nsig_b, nsig_s, threshold;
    int min_count;
    float prefilter_k;  // nsig_s^2 * (1 - 1e-4): conservative
};

template <typename PixelT, int KH_, int PPL_>
struct ThreshGeo {
    static constexpr int KH = KH_;
    static constexpr int PPL = PPL_;
    static constexpr int KS = 2 * KH + 1;                 // rows per stage == ring depth
    static constexpr int HL = (KH + PPL - 1) / PPL;       // halo lanes per side
    static constexpr int LOAD_COLS = 32 * PPL;
    static constexpr int OUT_COLS = (32 - 2 * HL) * PPL;
    static constexpr int VPAD = (8 - (HL * PPL) % 8) % 8; // validity tile starts VPAD px earlier
    // TMA box starts must be 16-byte aligned in global memory (measured: a misaligned start
    // raises 'illegal instruction'), so the validity box starts up to 15 bytes early.
    static constexpr int VROW_BYTES = ((VPAD + LOAD_COLS + 7) / 8 + 15 + 15) / 16 * 16;
    static constexpr int IMG_STAGE_BYTES = KS * LOAD_COLS * (int)sizeof(PixelT);
    static constexpr int VALID_STAGE_BYTES = (KS * VROW_BYTES + 127) / 128 * 128;
    static constexpr int STAGE_BYTES = IMG_STAGE_BYTES + VALID_STAGE_BYTES;
    static constexpr int TX_BYTES = IMG_STAGE_BYTES + KS * VROW_BYTES;
    static constexpr int QUEUE_CAP = 32 + OUT_COLS;       // entries; a row adds <= OUT_COLS
    static constexpr int QUEUE_BYTES = (QUEUE_CAP * 12 + 127) / 128 * 128;
    static_assert(HL == 1, "one halo lane per side");
    static_assert(OUT_COLS % 8 == 0, "strips must start on a validity byte");
    static_assert(IMG_STAGE_BYTES % 128 == 0, "TMA destination alignment");
    static_assert(LOAD_COLS <= 256, "TMA box limit");
    __host__ __device__ static constexpr int smem_per_warp(int stages) {
        return stages * STAGE_BYTES + QUEUE_BYTES + 128 /* mbarriers + queue counter */;
    }
};

// ---------------------------------------------------------------------------------------
// Exact resolution of one candidate, fp64, oracle operation order (standalone.cc:164-171).
// m, x: exact window count / sum.  v: raw centre pixel (mask-valid).
template <typename PixelT, int KH>
__device__ __forceinline__ bool resolve_dispersion(const ThreshArgs& a, const uint8_t* frame,
                                                   int col, int row, uint32_t m, uint32_t x,
                                                   uint32_t v) {
    if (sizeof(PixelT) == 4)  // the queue drops values >= 2^24: take the raw centre pixel
        v = reinterpret_cast<const PixelT*>(frame + (size_t)row * a.pitch)[col];
    const double md = (double)m, xd = (double)x, vd = (double)v;
    if (!((int)m >= a.min_count && vd > a.threshold)) return false;
    // double b = m * src[k] - x;   double d = nsig_s_ * std::sqrt(x * m);
    const double b = __dsub_rn(__dmul_rn(md, vd), xd);
    const double d = __dmul_rn(a.nsig_s, __dsqrt_rn(__dmul_rn(xd, md)));
    if (!(b > d)) return false;
    // y = sum of squares over valid (mask && v < 2^24) pixels of the clipped window
    const int x0 = max(col - KH, 0), x1 = min(col + KH, a.width - 1);
    const int y0 = max(row - KH, 0), y1 = min(row + KH, a.height - 1);
    unsigned long long y = 0;
    for (int yy = y0; yy <= y1; ++yy) {
        const PixelT* prow = reinterpret_cast<const PixelT*>(frame + (size_t)yy * a.pitch);
        const uint8_t* vrow = a.valid_bits + (size_t)yy * a.valid_pitch;
        // 2KH+1 <= 16 columns span at most 3 bytes
        const int byte0 = x0 >> 3;
        uint32_t bits = vrow[byte0];
        if ((x1 >> 3) > byte0) bits |= (uint32_t)vrow[byte0 + 1] << 8;
        if ((x1 >> 3) > byte0 + 1) bits |= (uint32_t)vrow[byte0 + 2] << 16;
        bits >>= (x0 & 7);
        for (int xx = x0; xx <= x1; ++xx, bits >>= 1) {
            const unsigned long long pv = prow[xx];
            if ((bits & 1u) && pv < (1ull << 24)) y += pv * pv;
        }
    }
    const double yd = (double)y;
    // double a = m * y - x * x - x * (m - 1);  double c = x * nsig_b_ * std::sqrt(2 * (m - 1));
    const double aa = __dsub_rn(__dsub_rn(__dmul_rn(md, yd), __dmul_rn(xd, xd)),
                                __dmul_rn(xd, __dsub_rn(md, 1.0)));
    const double cc =
      __dmul_rn(__dmul_rn(xd, a.nsig_b), __dsqrt_rn(__dmul_rn(2.0, __dsub_rn(md, 1.0))));
    return aa > cc;
}

// Resolve up to 32 queued candidates (one per lane); entries are {col | row<<13 | m<<26, x, v}.
template <typename PixelT, int KH>
__device__ __forceinline__ void drain32(const ThreshArgs& a, int z, const uint32_t* q, int first,
                                        int n, int lane) {
    bool strong = false;
    int col = 0, row = 0;
    if (lane < n) {
        const uint32_t* e = q + 3 * (first + lane);
        const uint32_t e0 = e[0];
        col = e0 & 0x1fff;
        row = (e0 >> 13) & 0x1fff;
        strong = resolve_dispersion<PixelT, KH>(a, a.image + (size_t)z * a.frame_stride, col, row,
                                                e0 >> 26, e[1], e[2]);
    }
    if (strong) {
        uint32_t* w = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(a.strong_bits) +
                                                  (size_t)z * a.strong_frame_stride +
                                                  (size_t)row * a.strong_pitch) +
                      (col >> 5);
        atomicOr(w, 1u << (col & 31));
    }
    const unsigned ballot = __ballot_sync(0xffffffffu, strong);
    if (lane == 0 && ballot) atomicAdd(a.strong_count + z, __popc(ballot));
}

// ---------------------------------------------------------------------------------------
template <typename PixelT, int KH, int PPL, int STAGES, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
  dispersion_kernel(const __grid_constant__ CUtensorMap tm_img,
                    const __grid_constant__ CUtensorMap tm_valid, const ThreshArgs a) {
    using G = ThreshGeo<PixelT, KH, PPL>;
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
    uint32_t* queue = reinterpret_cast<uint32_t*>(wsm + STAGES * G::STAGE_BYTES);
    uint64_t* bars = reinterpret_cast<uint64_t*>(wsm + STAGES * G::STAGE_BYTES + G::QUEUE_BYTES);
    int* qcount = reinterpret_cast<int*>(bars + STAGES);

    const int r0 = rt * a.tile_rows;                      // first output row
    const int row_end = min(r0 + a.tile_rows, a.height);  // one past last output row
    const int n_load_rows = (row_end - r0) + 2 * KH;
    const int n_groups = (n_load_rows + KS - 1) / KS;
    const int x0 = strip * G::OUT_COLS - G::HL * PPL;      // first loaded column (may be < 0)
    const int y_first = r0 - KH;                           // first loaded row   (may be < 0)
    const int vbyte_first = (x0 - G::VPAD) >> 3;           // arithmetic shift: x0-VPAD is a multiple of 8
    const int vbyte0 = vbyte_first & ~15;                  // 16-byte aligned TMA start (may be < 0)
    const int voff = vbyte_first - vbyte0;

    if (lane == 0) {
        for (int s = 0; s < STAGES; ++s) ptx::mbar_init(bars + s, 1);
        *qcount = 0;
        ptx::fence_mbar_init();
    }
    __syncwarp();

    auto issue = [&](int g) {
        const int s = g % STAGES;
        uint8_t* dst = stage_base + s * G::STAGE_BYTES;
        ptx::mbar_arrive_expect_tx(bars + s, G::TX_BYTES);
        ptx::tma_load_3d(dst, &tm_img, bars + s, x0, y_first + g * KS, z);
        ptx::tma_load_2d(dst + G::IMG_STAGE_BYTES, &tm_valid, bars + s, vbyte0, y_first + g * KS);
    };
    if (lane == 0) {
        for (int g = 0; g < STAGES && g < n_groups; ++g) issue(g);
    }

    // vertical running sums and the ring of the last KS rows of packed words
    uint32_t ring[KS][PPL];
    uint32_t V[PPL];
    // 32-bit pixels: byte-packed per-column counts (4 columns in one register)
    uint32_t ringc[P16 ? 1 : KS];
    uint32_t ringf[P16 ? 1 : KS];  // mask bits of the row (centre-pixel validity)
    uint32_t Vc = 0;
#pragma unroll
    for (int u = 0; u < KS; ++u) {
#pragma unroll
        for (int j = 0; j < PPL; ++j) ring[u][j] = 0;
        if (!P16) {
            ringc[u] = 0;
            ringf[u] = 0;
        }
    }
#pragma unroll
    for (int j = 0; j < PPL; ++j) V[j] = 0;

    const bool out_lane = lane >= G::HL && lane < 32 - G::HL;
    const int col0 = x0 + lane * PPL;  // this lane's first column

    for (int g = 0; g < n_groups; ++g) {
        const int s = g % STAGES;
        ptx::mbar_wait(bars + s, (uint32_t)((g / STAGES) & 1));
        const uint8_t* simg = stage_base + s * G::STAGE_BYTES;
        const uint8_t* sval = simg + G::IMG_STAGE_BYTES;

#pragma unroll
        for (int u = 0; u < KS; ++u) {
            const int i = g * KS + u;  // index of this load row within the task
            // ---- load + unpack one row ------------------------------------------------
            const uint4 raw = ptx::lds128(simg + (u * G::LOAD_COLS + lane * PPL) * sizeof(PixelT));
            uint32_t mb;
            {
                const int bitoff = G::VPAD + lane * PPL;
                mb = sval[u * G::VROW_BYTES + voff + (bitoff >> 3)];
                mb = (mb >> (bitoff & 7)) & ((1u << PPL) - 1u);
            }
            uint32_t w[PPL];
            if constexpr (P16) {
                const uint32_t r[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
                for (int j = 0; j < PPL; ++j) {
                    // bytes {v.lo, v.hi, 0x00, 0x01}: value in bits 0..15, count in bit 24
                    const uint32_t t = __byte_perm(r[j >> 1], 0x00000100u, (j & 1) ? 0x5432 : 0x5410);
                    w[j] = (mb & (1u << j)) ? t : 0u;
                }
#pragma unroll
                for (int j = 0; j < PPL; ++j) {
                    V[j] = V[j] + w[j] - ring[u][j];
                    ring[u][j] = w[j];
                }
            } else {
                const uint32_t r[4] = {raw.x, raw.y, raw.z, raw.w};
                uint32_t cbits = 0;
#pragma unroll
                for (int j = 0; j < PPL; ++j) {
                    const bool counted = (mb & (1u << j)) && r[j] < (1u << 24);
                    w[j] = counted ? r[j] : 0u;
                    cbits |= counted ? (1u << j) : 0u;
                }
                // spread 4 bits to 4 bytes (bit k -> byte k)
                const uint32_t cspread = (cbits * 0x00204081u) & 0x01010101u;
#pragma unroll
                for (int j = 0; j < PPL; ++j) {
                    V[j] = V[j] + w[j] - ring[u][j];
                    // keep the RAW value in the ring when the pixel is mask-valid but
                    // excluded from the sums (>= 2^24): the centre test needs it
                    ring[u][j] = w[j];
                }
                Vc = Vc + cspread - ringc[u];
                ringc[u] = cspread;
                ringf[u] = mb | (cbits << 8);
            }

            // ---- test the centre row of the window that is now complete -----------------
            const int out_row = r0 + i - 2 * KH;
            if (i >= 2 * KH && out_row < row_end) {  // warp-uniform
                const int cs = (u + KS - KH) % KS;  // ring slot of the centre row
                uint32_t cb = 0;                        // candidate bits of this lane
                uint32_t Hx[PPL];
                uint32_t Hm[PPL];
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
                if constexpr (P16) {
                    // H = m << 24 | x ;  b = m*v - x = m*w_c - H  (w_c = 1<<24 | v, or 0)
                    int b[PPL];
                    uint32_t pmin = 0xffffffffu;
#pragma unroll
                    for (int j = 0; j < PPL; ++j) {
                        const uint32_t m = Hx[j] >> 24;
                        b[j] = (int)(m * ring[cs][j] - Hx[j]);
                        pmin = min(pmin, (Hx[j] & 0xffffffu) * m);
                        Hm[j] = m;
                    }
                    // b > d  =>  b > sqrt(k * x*m) for every k < nsig_s^2, x*m >= pmin
                    const float tf = sqrtf(a.prefilter_k * (float)pmin) * 0.99999f;
                    const int T = (int)tf;
#pragma unroll
                    for (int j = 0; j < PPL; ++j) cb |= (b[j] > T) ? (1u << j) : 0u;
                } else {
                    const uint32_t cl = __shfl_up_sync(0xffffffffu, Vc, 1);
                    const uint32_t cr = __shfl_down_sync(0xffffffffu, Vc, 1);
                    const uint32_t call = __dp4a(Vc, 0x01010101u, 0u);
                    // columns -3..-1 live in bytes 1..3 of the left lane, +1..+3 in bytes 0..2 of the right
                    Hm[0] = __dp4a(cl, 0x01010100u, call);
                    Hm[1] = __dp4a(cl, 0x01010000u, __dp4a(cr, 0x00000001u, call));
                    Hm[2] = __dp4a(cl, 0x01000000u, __dp4a(cr, 0x00000101u, call));
                    Hm[3] = __dp4a(cr, 0x00010101u, call);
                    const uint32_t flags = ringf[cs];
                    float pmin = 3.0e38f;
                    float bf[PPL], sf[PPL];
#pragma unroll
                    for (int j = 0; j < PPL; ++j) {
                        const float mf = (float)Hm[j], xf = (float)Hx[j];
                        const float mv = mf * (float)ring[cs][j];
                        bf[j] = mv - xf;
                        sf[j] = mv + xf;
                        pmin = fminf(pmin, xf * mf);
                    }
                    const float tf = sqrtf(a.prefilter_k * pmin) * 0.9999f;
#pragma unroll
                    for (int j = 0; j < PPL; ++j) {
                        const bool mask_ok = flags & (1u << j);
                        const bool counted = flags & (0x100u << j);
                        // float error of bf is below 4e-7 * (m*v + x)
                        const bool cand = mask_ok && (!counted || bf[j] + 1e-6f * sf[j] > tf);
                        cb |= cand ? (1u << j) : 0u;
                    }
                }
                if (!out_lane) cb = 0;
                if (__any_sync(0xffffffffu, cb != 0)) {
#pragma unroll
                    for (int j = 0; j < PPL; ++j) {
                        if (cb & (1u << j)) {
                            const int slot = atomicAdd(qcount, 1);
                            uint32_t* e = queue + 3 * slot;
                            uint32_t m, x, v;
                            if constexpr (P16) {
                                m = Hm[j];
                                x = Hx[j] & 0xffffffu;
                                v = ring[cs][j] & 0xffffu;
                            } else {
                                m = Hm[j];
                                x = Hx[j];
                                v = ring[cs][j];  // 0 when the pixel is >= 2^24: re-read below
                            }
                            e[0] = (uint32_t)(col0 + j) | ((uint32_t)out_row << 13) | (m << 26);
                            e[1] = x;
                            e[2] = v;
                        }
                    }
                    __syncwarp();
                    int qn = *qcount;
                    if (qn >= 32) {
                        int first = 0;
                        while (qn - first >= 32) {
                            drain32<PixelT, KH>(a, z, queue, first, 32, lane);
                            first += 32;
                        }
                        // move the tail to the front
                        const int rem = qn - first;
                        uint32_t t0 = 0, t1 = 0, t2 = 0;
                        if (lane < rem) {
                            t0 = queue[3 * (first + lane)];
                            t1 = queue[3 * (first + lane) + 1];
                            t2 = queue[3 * (first + lane) + 2];
                        }
                        __syncwarp();
                        if (lane < rem) {
                            queue[3 * lane] = t0;
                            queue[3 * lane + 1] = t1;
                            queue[3 * lane + 2] = t2;
                        }
                        if (lane == 0) *qcount = rem;
                        __syncwarp();
                    }
                }
            }
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
    if (qn > 0) drain32<PixelT, KH>(a, z, queue, 0, qn, lane);
}

}  // namespace ffsb
