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
    static constexpr int GQ_CAP = 64;        // pending candidate groups (< 32 + 30)
    static constexpr int PQ_CAP = 288;       // pending candidate pixels (< 32 + 32 * 8)
    static constexpr int QUEUE_BYTES = (GQ_CAP + PQ_CAP) * 4;
    static_assert(RING_ROWS > 2 * KH, "ring must hold the 7-row window");
    static_assert(IMG_STAGE_BYTES % 128 == 0 && QUEUE_BYTES % 128 == 0, "TMA destination alignment");
    __host__ __device__ static constexpr int smem_per_warp(int stages) {
        return stages * STAGE_BYTES + RING_BYTES + QUEUE_BYTES + 128 /* mbarriers */;
    }
};

// group info plane: per row, per group of 8 pixels, 2 bytes {validity bits, m_ub}.
// m_ub = max over the group's mask-valid pixels of the number of valid pixels in the
// image-clipped 7x7 window (standalone.cc:126-130 clipping).
__global__ void group_info_kernel(const uint8_t* __restrict__ valid_bits, size_t bits_pitch, int width,
                                  int height, uint8_t* __restrict__ info, size_t info_pitch) {
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
    uint32_t mub = 0;
#pragma unroll
    for (int p = 0; p < 8; ++p)
        if ((vb >> p) & 1u) mub = max(mub, m[p]);
    uint8_t* out = info + (size_t)row * info_pitch + 2 * gi;
    out[0] = (uint8_t)vb;
    out[1] = (uint8_t)mub;
}

// ---- row front end: masked + clamped packed words, max of the masked raw values -----------
template <typename PixelT>
__device__ __forceinline__ void fused_load_row(const uint8_t* pimg, const uint8_t* pinfo, uint32_t (&c)[4],
                                               uint32_t& vmax, uint32_t& info) {
    constexpr uint32_t CL = FusedGeo<PixelT>::CLAMP;
    info = *reinterpret_cast<const uint16_t*>(pinfo);
    if constexpr (sizeof(PixelT) == 2) {
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
        const uint32_t r[8] = {ra.x, ra.y, ra.z, ra.w, rb.x, rb.y, rb.z, rb.w};
        uint32_t cl[8];
        vmax = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint32_t v = (info & (1u << j)) ? r[j] : 0u;
            vmax = max(vmax, v);
            // pixels >= 2^24 are not counted by the reference (standalone.cc:90): they add 0
            cl[j] = v < (1u << 24) ? min(v, CL) : 0u;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) c[k] = cl[2 * k] | (cl[2 * k + 1] << 16);
    }
}

template <typename PixelT, int STAGES, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
  fused_dispersion_kernel(const __grid_constant__ CUtensorMap tm_img,
                          const __grid_constant__ CUtensorMap tm_info, const ThreshArgs a) {
    using G = FusedGeo<PixelT>;
    constexpr int SR = G::SR, KH = G::KH, PPL = G::PPL;
    static_assert(SR == 4, "row loop is written for 4-row stages and an 8-row ring");

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ float rcp_tab[64];  // 0.99999 / m_ub
    if (threadIdx.x < 64) rcp_tab[threadIdx.x] = 0.99999f / (float)max((int)threadIdx.x, 1);
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
    uint32_t* pq = gq + G::GQ_CAP;
    uint64_t* bars = reinterpret_cast<uint64_t*>(ring + G::RING_BYTES + G::QUEUE_BYTES);

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

    const uint8_t* frame = a.image + (size_t)z * a.frame_stride;
    const uint8_t* valid = a.valid_bits;
    const bool out_lane = lane >= 1 && lane <= G::OUT_LANES;
    const uint32_t lt_mask = (1u << lane) - 1u;
    int qn = 0, pqn = 0;     // warp-uniform queue fill levels
    unsigned found = 0;

    // exact test of one queued pixel
    auto resolve_entry = [&](uint32_t e) {
        const int col = e & 0x1fff, row = e >> 13;
        if (a.force_brute == 2) return;
        if (resolve_pixel<PixelT, KH, false>(a, frame, valid, col, row)) {
            set_strong(a, z, col, row);
            ++found;
        }
    };
    // expand `n` queued groups (the last n of gq) to candidate pixels, then resolve full rounds
    auto resolve_groups = [&](int n) {
        uint32_t bits = 0, e = 0;
        int col0 = 0, row = 0;
        if (lane < n) {
            e = gq[qn - n + lane];
            const uint32_t T = e >> 14;
            col0 = x0 + (int)(e & 31u) * PPL;
            row = r0 + (int)((e >> 5) & 0x1ffu);
            const uint32_t vb = valid[(size_t)row * a.valid_pitch + (col0 >> 3)];
            const PixelT* p = reinterpret_cast<const PixelT*>(frame + (size_t)row * a.pitch) + col0;
            uint32_t v[8];
            if constexpr (sizeof(PixelT) == 2) {
                const uint4 raw = *reinterpret_cast<const uint4*>(p);
                v[0] = raw.x & 0xffffu; v[1] = raw.x >> 16; v[2] = raw.y & 0xffffu; v[3] = raw.y >> 16;
                v[4] = raw.z & 0xffffu; v[5] = raw.z >> 16; v[6] = raw.w & 0xffffu; v[7] = raw.w >> 16;
            } else {
                const uint4 ra = reinterpret_cast<const uint4*>(p)[0], rb = reinterpret_cast<const uint4*>(p)[1];
                v[0] = ra.x; v[1] = ra.y; v[2] = ra.z; v[3] = ra.w;
                v[4] = rb.x; v[5] = rb.y; v[6] = rb.z; v[7] = rb.w;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) bits |= (((vb >> j) & 1u) && v[j] > T) ? (1u << j) : 0u;
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
    };

    uint32_t V[4] = {0, 0, 0, 0};   // packed column sums of the last 7 rows
    uint32_t hv1 = 0, hv2 = 0, hv3 = 0;  // max masked value of rows i-1, i-2, i-3
    uint32_t hi1 = 0, hi2 = 0, hi3 = 0;  // group info of rows i-1, i-2, i-3
    uint8_t* ring_lane = ring + lane * 16;
    const int pimg_off = lane * PPL * (int)sizeof(PixelT);
    const int pinfo_off = G::IMG_STAGE_BYTES + (info_b - info_b0) + 2 * lane;
    const float sk = a.sig_k;             // nsig_s * (1 - 1e-4)
    const int thr_floor = a.thr_floor;

    int s = 0;
    uint32_t phase = 0;
    for (int g = 0; g < n_groups; ++g) {
        ptx::mbar_wait(bars + s, phase);
        const uint8_t* st = stage_base + s * G::STAGE_BYTES;
        uint8_t* rcur = ring_lane + ((g & 1) ? SR * G::RING_ROW_BYTES : 0);
        uint8_t* roth = ring_lane + ((g & 1) ? 0 : SR * G::RING_ROW_BYTES);
#pragma unroll
        for (int u = 0; u < SR; ++u) {
            const int i = g * SR + u;  // load-row index within the task
            uint32_t c[4], vm, info;
            fused_load_row<PixelT>(st + u * G::IMG_ROW_BYTES + pimg_off, st + pinfo_off + u * G::INFO_ROW_BYTES,
                                   c, vm, info);
            // vertical running sum over 7 rows: V += row(i) - row(i-7); ring slot of row i = i & 7
            uint8_t* oldp = (u + 1 < SR) ? rcur + (u + 1) * G::RING_ROW_BYTES : roth;
            const uint4 old = ptx::lds128(oldp);
            V[0] = V[0] + c[0] - old.x;
            V[1] = V[1] + c[1] - old.y;
            V[2] = V[2] + c[2] - old.z;
            V[3] = V[3] + c[3] - old.w;
            *reinterpret_cast<uint4*>(rcur + u * G::RING_ROW_BYTES) = make_uint4(c[0], c[1], c[2], c[3]);

            const int out_row = r0 + i - 2 * KH;
            if (i >= 2 * KH && out_row < row_end) {  // warp-uniform
                // horizontal 7-tap sums of the column sums, 32-bit
                const uint32_t l2 = __shfl_up_sync(0xffffffffu, V[2], 1);
                const uint32_t l3 = __shfl_up_sync(0xffffffffu, V[3], 1);
                const uint32_t rr0 = __shfl_down_sync(0xffffffffu, V[0], 1);
                const uint32_t rr1 = __shfl_down_sync(0xffffffffu, V[1], 1);
                uint32_t E[14];
                E[0] = l2 >> 16;
                E[1] = l3 & 0xffffu;
                E[2] = l3 >> 16;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    E[3 + 2 * k] = V[k] & 0xffffu;
                    E[4 + 2 * k] = V[k] >> 16;
                }
                E[11] = rr0 & 0xffffu;
                E[12] = rr0 >> 16;
                E[13] = rr1 & 0xffffu;
                uint32_t h = E[0] + E[1] + E[2] + E[3] + E[4] + E[5] + E[6];
                uint32_t xmin = h;
#pragma unroll
                for (int j = 1; j < 8; ++j) {
                    h = h + E[j + 6] - E[j - 1];
                    xmin = min(xmin, h);
                }
                // group threshold on the raw value: T = floor(mu + nsig_s' sqrt(mu)), mu = x_lb / m_ub
                const float mu = (float)xmin * rcp_tab[hi3 >> 8];
                float sq;
                asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(mu));
                int T = max(__float2int_rz(fmaf(sk, sq, mu)), thr_floor);
                if (a.force_brute == 3) T = thr_floor;  // debug: no pruning
                const bool pass = out_lane && hv3 > (uint32_t)T;
                const unsigned ball = __ballot_sync(0xffffffffu, pass);
                if (ball) {
                    if (pass)
                        gq[qn + __popc(ball & lt_mask)] =
                          (uint32_t)lane | ((uint32_t)(out_row - r0) << 5) | ((uint32_t)min(T, 0x3ffff) << 14);
                    qn += __popc(ball);
                    __syncwarp();
                    if (qn >= 32) resolve_groups(32);
                }
            }
            hv3 = hv2; hv2 = hv1; hv1 = vm;
            hi3 = hi2; hi2 = hi1; hi1 = info;
        }
        // refill the stage we just consumed
        __syncwarp();
        if (lane == 0 && g + STAGES < n_groups) {
            ptx::fence_proxy_async();
            issue(g + STAGES, s);
        }
        if (++s == STAGES) {
            s = 0;
            phase ^= 1u;
        }
    }
    __syncwarp();
    if (qn > 0) resolve_groups(qn);
    if (lane < pqn) resolve_entry(pq[lane]);
    for (int off = 16; off > 0; off >>= 1) found += __shfl_down_sync(0xffffffffu, found, off);
    if (lane == 0 && found) atomicAdd(a.strong_count + z, found);
}

}  // namespace ffsb
