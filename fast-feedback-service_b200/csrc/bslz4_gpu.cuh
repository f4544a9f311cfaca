// bslz4_gpu.cuh -- bitshuffle-LZ4 chunk decode ON THE DEVICE.
//
// The reference decompresses every frame on the host before the upload
//   bshuf_decompress_lz4(buffer + 12, host_image, px, sizeof(pixel_t), 0)
// (spotfinder/spotfinder.cc:823-841, h5read/src/read_chunks.cc:21-25) and notes that the
// decompression should eventually be offloaded.  Here the COMPRESSED chunk crosses PCIe
// (4-10x fewer bytes than the frame, and the H2D link is the end-to-end roofline) and one
// warp decodes one bitshuffle block:
//   1. (the compressed payload is read in place: global memory through L1)
//   2. the raw LZ4 block is decoded into a shared buffer: the token / length parse is
//      warp-uniform, literal and match copies are warp-parallel (word copies through a
//      funnel shift; run-length matches with offset 1/2/4 become pattern stores; short
//      overlapping matches are copied as the periodic sequence they denote);
//   3. the block -- elem_size*8 bit planes of n/8 bytes -- is transposed back to pixels
//      (8x8 bit-matrix transposes in registers) and written to the pitched device frame the
//      threshold kernel reads.
// Stream format: see csrc/cli/bslz4.hpp (host decoder, same restated format).  The host
// only walks the 4-byte block headers to build the offset table (ffs_submit_compressed).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ffsb {

constexpr int kBslz4MaxBlockBytes = 8192;                       // bitshuffle's default target block
constexpr int kBslz4InBytes = kBslz4MaxBlockBytes + 8192 / 255 + 64 + 32;  // LZ4 worst case + slack
constexpr int kBslz4SmemPerWarp = kBslz4MaxBlockBytes + 16;  // the decoded (bit-transposed) block

// Per-frame values live in device memory (uploaded with the block table), so the kernel
// arguments of a ring slot never change and the slot's work can be replayed as a CUDA graph.
struct Bslz4Dyn {
    uint32_t n_blocks;
    uint32_t block_elems;       // elements per full block (multiple of 8)
    uint32_t last_elems;        // elements of the last block
    uint32_t tail_off;          // offset of the raw trailing (n % 8) elements
    uint32_t tail_elems;
    uint32_t tail_elem0;        // index of the first trailing element
    uint32_t pad[2];
};
static_assert(sizeof(Bslz4Dyn) == 32, "header sits between the uint2 block table and the chunk");

struct Bslz4Args {
    const uint8_t* chunk;       // device copy of the whole chunk (header + body), 16-byte aligned
    const Bslz4Dyn* dyn;        // = chunk - 32; PRECEDED by per block: uint2 {payload offset in chunk, payload
                                // bytes}, and before that uint32 work order[n_blocks]: warp w decodes block
                                // order[w]  (written backwards from the header so one copy uploads table + chunk)
    uint8_t* frame;             // pitched destination
    size_t pitch;
    uint32_t width;
    uint32_t elem_size;         // 2 or 4
    uint32_t inv_width;         // floor(2^32 / width) + 1: n / width == umulhi(n, inv_width) for n*width < 2^32
    unsigned int* status;       // set to nonzero on malformed input; status[1] = work counter of the
                                // persistent variant (BSLZ4_PERSISTENT), zeroed with it before every launch
};

// 8x8 bit-matrix transpose of the 64-bit value hi:lo (byte r = row r).
__device__ __forceinline__ void transpose8(uint32_t& lo, uint32_t& hi) {
    uint32_t t;
    t = (lo ^ (lo >> 7)) & 0x00AA00AAu;  lo = lo ^ t ^ (t << 7);
    t = (hi ^ (hi >> 7)) & 0x00AA00AAu;  hi = hi ^ t ^ (t << 7);
    t = (lo ^ (lo >> 14)) & 0x0000CCCCu; lo = lo ^ t ^ (t << 14);
    t = (hi ^ (hi >> 14)) & 0x0000CCCCu; hi = hi ^ t ^ (t << 14);
    // x ^ (x >> 28) masked with 0x00000000F0F0F0F0: only the low word of t is non-zero
    t = (lo ^ ((lo >> 28) | (hi << 4))) & 0xF0F0F0F0u;
    lo = lo ^ t ^ (t << 28);
    hi = hi ^ (t >> 4);
}

// Warp-parallel copy of len bytes to shared memory from shared or global memory.  Word
// stores; the source is realigned with a funnel shift (it may be read up to 4 bytes past its
// end: every buffer is padded).  SYNC_EACH_ROUND: the source may have been written by the
// previous 128-byte round (overlapping match with offset >= 132).
template <bool SYNC_EACH_ROUND>
__device__ __forceinline__ void warp_copy(uint8_t* dst, const uint8_t* src, int len, int lane) {
    if (len <= 32) {  // the common case: one masked byte copy
        if (lane < len) dst[lane] = src[lane];
        return;
    }
    const int head = (int)((4u - ((uint32_t)reinterpret_cast<uintptr_t>(dst) & 3u)) & 3u);
    if (lane < head) dst[lane] = src[lane];
    const int nwords = (len - head) >> 2;
    const uint8_t* s = src + head;
    const uint32_t sh = ((uint32_t)reinterpret_cast<uintptr_t>(s) & 3u) * 8u;
    const uint32_t* sa = reinterpret_cast<const uint32_t*>(s - (sh >> 3));
    uint32_t* d = reinterpret_cast<uint32_t*>(dst + head);
    if (SYNC_EACH_ROUND) __syncwarp();
    for (int base = 0; base < nwords; base += 32) {  // uniform trip count
        const int w = base + lane;
        if (w < nwords) d[w] = __funnelshift_r(sa[w], sa[w + 1], sh);
        if (SYNC_EACH_ROUND) __syncwarp();
    }
    const int done = head + (nwords << 2);
    if (lane < len - done) dst[done + lane] = src[done + lane];
}

// Literal runs longer than the parse window: global -> shared, four rounds of loads in flight
// (the source is read-only global memory, so no round waits for the previous round's stores).
__device__ __forceinline__ void warp_copy_in(uint8_t* dst, const uint8_t* __restrict__ src, int len, int lane) {
    if (len <= 32) {
        if (lane < len) dst[lane] = src[lane];
        return;
    }
    const int head = (int)((4u - ((uint32_t)reinterpret_cast<uintptr_t>(dst) & 3u)) & 3u);
    if (lane < head) dst[lane] = src[lane];
    const int nwords = (len - head) >> 2;
    const uint8_t* s = src + head;
    const uint32_t sh = ((uint32_t)reinterpret_cast<uintptr_t>(s) & 3u) * 8u;
    const uint32_t* sa = reinterpret_cast<const uint32_t*>(s - (sh >> 3));
    uint32_t* d = reinterpret_cast<uint32_t*>(dst + head);
    for (int base = 0; base < nwords; base += 128) {  // uniform trip count
        uint32_t lo[4], hi[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int w = base + 32 * u + lane;
            lo[u] = hi[u] = 0;
            if (w < nwords) {
                lo[u] = __ldg(sa + w);
                hi[u] = __ldg(sa + w + 1);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int w = base + 32 * u + lane;
            if (w < nwords) d[w] = __funnelshift_r(lo[u], hi[u], sh);
        }
    }
    const int done = head + (nwords << 2);
    if (lane < len - done) dst[done + lane] = src[done + lane];
}

// Decode one raw LZ4 block from in[0..in_len) (global memory, read through L1) to
// out[0..out_len) (shared).  Returns false on malformed input.  All lanes take the same path.
//
// The parse is a serial chain (token -> literal length -> offset -> match length -> next token),
// so what matters is the latency and the instruction count of one sequence:
//   * the 32 bytes that follow the parse position sit in a register window, one byte per lane
//     (ONE coalesced load per sequence, issued as soon as the next position is known, i.e.
//     before the match copy runs); token, offset and the first length-extension byte come out of
//     it by warp shuffles, and literal runs of up to 14 bytes are stored straight from it;
//   * matches of up to 128 bytes (the bulk) are copied as at most 4 bytes per lane, no
//     alignment prologue; longer ones take the word / pattern copies below.
__device__ __forceinline__ bool lz4_warp_decode(const uint8_t* __restrict__ in, int in_len, uint8_t* out,
                                                int out_len, int lane) {
    constexpr unsigned FULL = 0xffffffffu;
    int ip = 0, op = 0;
    // length extension bytes from ip on: 255, 255, ..., b < 255  (long lengths only)
    auto extend = [&](int& len) -> bool {
        for (;;) {
            const int idx = ip + lane;
            const uint32_t b = idx < in_len ? in[idx] : 0u;
            const unsigned nz = __ballot_sync(FULL, b != 255u);
            if (nz) {
                const int first = __ffs(nz) - 1;
                if (ip + first >= in_len) return false;
                len += 255 * first + (int)__shfl_sync(FULL, b, first);
                ip += first + 1;
                return true;
            }
            len += 255 * 32;
            ip += 32;
            if (ip >= in_len) return false;
        }
    };
    auto load_win = [&](int p) -> uint32_t {
        const int i = p + lane;
        return i < in_len ? (uint32_t)in[i] : 0u;
    };
    // first touch of the payload: all of its lines at once instead of one miss after the other
#pragma nv_diag_suppress 550  // the loaded value is deliberately never read (nothing waits for it)
    for (int o = lane * 128; o < in_len; o += 32 * 128) {
        uint32_t sink;
        asm volatile("ld.global.ca.u8 %0, [%1];" : "=r"(sink) : "l"(in + o));
    }
#pragma nv_diag_default 550
    uint32_t win = load_win(0);  // byte ip + lane of the input, 0 beyond its end
    while (ip < in_len) {
        const uint32_t token = __shfl_sync(FULL, win, 0);
        int lit = (int)(token >> 4);
        int mlen = (int)(token & 15u);
        const bool in_win = lit < 15;  // token, literals, offset and one extension byte: lanes 0 .. lit + 3 <= 17
        // issued early, used after the literals (lanes beyond the input hold 0)
        const uint32_t o0 = __shfl_sync(FULL, win, (lit + 1) & 31);
        const uint32_t o1 = __shfl_sync(FULL, win, (lit + 2) & 31);
        const uint32_t e0 = __shfl_sync(FULL, win, (lit + 3) & 31);
        ++ip;
        if (in_win) {
            if (lit > in_len - ip || lit > out_len - op) return false;
            if (lane >= 1 && lane <= lit) out[op + lane - 1] = (uint8_t)win;
        } else {
            if (!extend(lit)) return false;
            if (lit > in_len - ip || lit > out_len - op) return false;
            warp_copy_in(out + op, in + ip, lit, lane);
        }
        ip += lit;
        op += lit;
        if (ip >= in_len) break;  // the last sequence has no match
        if (in_len - ip < 2) return false;
        int offset;
        if (in_win) offset = (int)(o0 | (o1 << 8));
        else offset = in[ip] | (in[ip + 1] << 8);
        ip += 2;
        if (mlen == 15) {
            if (in_win && e0 != 255u) {
                if (ip >= in_len) return false;
                mlen += (int)e0;
                ++ip;
            } else if (!extend(mlen)) {
                return false;
            }
        }
        mlen += 4;
        if (offset == 0 || offset > op || mlen > out_len - op) return false;
        win = load_win(ip);  // the next sequence's window: in flight while the match is copied
        if (lit) __syncwarp();  // literals just written may be the match source
        uint8_t* dst = out + op;
        if (mlen <= 128) {
            if (offset >= mlen) {  // no overlap: <= 4 bytes per lane, loads before stores
                const uint8_t* s = dst - offset;
                uint32_t b[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) b[u] = (lane + 32 * u < mlen) ? (uint32_t)s[lane + 32 * u] : 0u;
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (lane + 32 * u < mlen) dst[lane + 32 * u] = (uint8_t)b[u];
            } else if ((offset & (offset - 1)) == 0 && offset <= 32) {
                // run with a power-of-two period: every byte a lane writes is the same source byte
                const uint8_t v = dst[(lane & (offset - 1)) - offset];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (lane + 32 * u < mlen) dst[lane + 32 * u] = v;
            } else if (offset > 32) {
                // overlapping, period longer than a round: a round reads what earlier rounds wrote
                for (int base = 0; base < mlen; base += 32) {  // uniform trip count
                    const int k = base + lane;
                    if (k < mlen) dst[k] = dst[k - offset];
                    __syncwarp();
                }
            } else {
                // short period: the periodic sequence it denotes
                const int step = 32 % offset;
                int km = lane % offset;
                for (int k = lane; k < mlen; k += 32) {
                    dst[k] = dst[km - offset];
                    km += step;
                    if (km >= offset) km -= offset;
                }
            }
        } else if (offset >= mlen) {
            warp_copy<false>(dst, dst - offset, mlen, lane);
        } else if (offset == 1 || offset == 2 || offset == 4) {
            // run: the 4-byte pattern seen from an aligned word is the same for every word
            const uint32_t a0 = (uint32_t)reinterpret_cast<uintptr_t>(out);  // 16-byte aligned base
            uint32_t pat = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int p = (int)(((uint32_t)j - (uint32_t)op - a0) & (uint32_t)(offset - 1));
                pat |= (uint32_t)dst[p - offset] << (8 * j);
            }
            const int head = min((int)((16u - ((a0 + (uint32_t)op) & 15u)) & 15u), mlen);
            if (lane < head) dst[lane] = (uint8_t)(pat >> (8 * ((a0 + op + lane) & 3u)));
            const int nq = (mlen - head) >> 4;
            uint4* d = reinterpret_cast<uint4*>(dst + head);
            const uint4 p4 = make_uint4(pat, pat, pat, pat);
            for (int w = lane; w < nq; w += 32) d[w] = p4;
            const int done = head + (nq << 4);
            if (lane < mlen - done) dst[done + lane] = (uint8_t)(pat >> (8 * ((a0 + op + done + lane) & 3u)));
        } else if (offset >= 132) {
            warp_copy<true>(dst, dst - offset, mlen, lane);
        } else {
            // short overlapping match = periodic sequence with period `offset`
            const int step = 32 % offset;
            int km = lane % offset;
            for (int k = lane; k < mlen; k += 32) {
                dst[k] = dst[km - offset];
                km += step;
                if (km >= offset) km -= offset;
            }
        }
        op += mlen;
        __syncwarp();
    }
    return op == out_len;
}

#ifdef BSLZ4_TIMELINE  // developer build (make variant EXTRA=-DBSLZ4_TIMELINE): per-warp {start, LZ4 done, end, block | smid << 16}
__device__ uint4 g_bslz4_timeline[16384];
__device__ __forceinline__ uint32_t tl_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return (uint32_t)t;
}
extern "C" int ffs_debug_bslz4_timeline(void* out, size_t bytes) {
    return (int)cudaMemcpyFromSymbol(out, g_bslz4_timeline, bytes < sizeof(g_bslz4_timeline) ? bytes : sizeof(g_bslz4_timeline));
}
#endif

#ifndef BSLZ4_MIN_CTAS
#define BSLZ4_MIN_CTAS 13  // resident CTAs per SM the kernel is compiled for: shared memory allows 13 of 2 warps (<= 72 registers; left
                           // to itself ptxas takes 82-96 and 10 CTAs fit: same speed, 61.5 vs 62.7 us u16, 93.1 vs 92.8 u32)
#endif
// grid: ceil(n_blocks / WARPS) CTAs of WARPS warps; dynamic smem WARPS * kBslz4SmemPerWarp.
template <typename PixelT, int WARPS>
#ifdef BSLZ4_PERSISTENT
__global__ void __launch_bounds__(WARPS * 32, 1) bslz4_decode_kernel(const Bslz4Args a) {
#else
__global__ void __launch_bounds__(WARPS * 32, BSLZ4_MIN_CTAS) bslz4_decode_kernel(const Bslz4Args a) {
#endif
    extern __shared__ __align__(16) uint8_t bsmem[];
    const Bslz4Dyn dyn = *a.dyn;
    const uint2* blocks = reinterpret_cast<const uint2*>(a.dyn) - dyn.n_blocks;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#ifndef BSLZ4_PERSISTENT
    const uint32_t w_id = blockIdx.x * WARPS + warp;
#endif
    constexpr int ES = (int)sizeof(PixelT);
    uint8_t* out = bsmem + (size_t)warp * kBslz4SmemPerWarp;

    // raw trailing elements (n % 8), copied by the first thread block
    if (blockIdx.x == 0 && threadIdx.x < dyn.tail_elems) {
        const uint32_t e = dyn.tail_elem0 + threadIdx.x;
        const uint8_t* p = a.chunk + dyn.tail_off + (size_t)threadIdx.x * ES;
        uint32_t val = 0;
        for (int k = 0; k < ES; ++k) val |= (uint32_t)p[k] << (8 * k);
        const PixelT v = (PixelT)val;
        const uint32_t row = e / a.width, col = e % a.width;
        reinterpret_cast<PixelT*>(a.frame + (size_t)row * a.pitch)[col] = v;
    }
#ifdef BSLZ4_PERSISTENT
    // one CTA per SM with as many warps as its shared memory holds; every warp pulls the next block of
    // the work order until none is left (no second wave of CTAs waiting for a whole CTA to retire)
#define BSLZ4_NEXT continue
    for (;; __syncwarp()) {
    uint32_t w_id = 0;
    if (lane == 0) w_id = atomicAdd(a.status + 1, 1u);
    w_id = __shfl_sync(0xffffffffu, w_id, 0);
    if (w_id >= dyn.n_blocks) break;
#else
#define BSLZ4_NEXT return
    if (w_id >= dyn.n_blocks) return;
    {
#endif
    const uint32_t b = (reinterpret_cast<const uint32_t*>(blocks) - dyn.n_blocks)[w_id];  // longest payload first
    if (b >= dyn.n_blocks) {
        if (lane == 0) atomicExch(a.status, 3u);
        BSLZ4_NEXT;
    }
    const uint2 blk = blocks[b];
    const int n_el = (b + 1 == dyn.n_blocks) ? (int)dyn.last_elems : (int)dyn.block_elems;
    const int out_len = n_el * ES;
    const int in_len = (int)blk.y;
    if (in_len > kBslz4InBytes - 32 || out_len > kBslz4MaxBlockBytes) {
        if (lane == 0) atomicExch(a.status, 1u);
        BSLZ4_NEXT;
    }
    const uint8_t* in = a.chunk + blk.x;
#ifdef BSLZ4_TIMELINE
    const uint32_t tl0 = tl_now();
#endif
    // 2. LZ4
    if (!lz4_warp_decode(in, in_len, out, out_len, lane)) {
        if (lane == 0) atomicExch(a.status, 2u);
        BSLZ4_NEXT;
    }
    __syncwarp();
#ifdef BSLZ4_TIMELINE
    const uint32_t tl1 = tl_now();
#endif
    // 3. bit planes -> pixels.  Plane (byte j, bit k) is row 8j+k, n_el/8 bytes per row;
    //    column c holds bit (e & 7) of elements e = 8c .. 8c+7.
    const int row_bytes = n_el >> 3;
    // 32-bit index arithmetic (frames have < 2^32 pixels); one division per warp
    const uint32_t e0 = b * dyn.block_elems;
    const uint32_t row0 = e0 / a.width, col0 = e0 - row0 * a.width;
    const bool aligned4 = (a.width & 3u) == 0;
    const uint32_t pitch32 = (uint32_t)a.pitch;  // a device frame is far below 4 GB
    // A lane takes 4 adjacent columns (32 pixels) per round when the plane rows are word
    // aligned (16 LDS.32 instead of 64 LDS.U8); ragged last blocks take one column per round.
    const bool quad = (row_bytes & 3) == 0;
    const int n_units = quad ? row_bytes >> 2 : row_bytes;
    for (int unit = lane; unit < n_units; unit += 32) {
      uint32_t pl[8 * ES];  // plane words: byte q = column 4*unit + q (quad) / byte 0 = column (single)
#pragma unroll
      for (int r = 0; r < 8 * ES; ++r)
          pl[r] = quad ? *reinterpret_cast<const uint32_t*>(out + r * row_bytes + 4 * unit)
                       : (uint32_t)out[r * row_bytes + unit];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (q > 0 && !quad) break;
        const int c = quad ? 4 * unit + q : unit;
        uint32_t w[2 * ES];  // per byte j of the element: 8 output bytes (element i -> byte i)
#pragma unroll
        for (int j = 0; j < ES; ++j) {
            // gather byte q of planes 8j .. 8j+7
            const uint32_t sel = 0x4040u + 0x11u * (uint32_t)q;   // {x.bq, y.bq} in the low half
            const uint32_t a01 = __byte_perm(pl[8 * j + 0], pl[8 * j + 1], sel);
            const uint32_t a23 = __byte_perm(pl[8 * j + 2], pl[8 * j + 3], sel);
            const uint32_t a45 = __byte_perm(pl[8 * j + 4], pl[8 * j + 5], sel);
            const uint32_t a67 = __byte_perm(pl[8 * j + 6], pl[8 * j + 7], sel);
            uint32_t lo = __byte_perm(a01, a23, 0x5410);
            uint32_t hi = __byte_perm(a45, a67, 0x5410);
            if (lo | hi) transpose8(lo, hi);
            w[2 * j] = lo;
            w[2 * j + 1] = hi;
        }
        // interleave: pixel i = bytes (w[0..1] byte i, w[2..3] byte i, ...)
        uint32_t px[2 * ES];  // 8 pixels * ES bytes
        if constexpr (ES == 2) {
            px[0] = __byte_perm(w[0], w[2], 0x5140);  // pixels 0,1
            px[1] = __byte_perm(w[0], w[2], 0x7362);  // pixels 2,3
            px[2] = __byte_perm(w[1], w[3], 0x5140);  // pixels 4,5
            px[3] = __byte_perm(w[1], w[3], 0x7362);  // pixels 6,7
        } else {
            // pixel i = b0[i] | b1[i] << 8 | b2[i] << 16 | b3[i] << 24
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const uint32_t t01a = __byte_perm(w[0 + h], w[2 + h], 0x5140);  // b0[0] b1[0] b0[1] b1[1]
                const uint32_t t01b = __byte_perm(w[0 + h], w[2 + h], 0x7362);  // b0[2] b1[2] b0[3] b1[3]
                const uint32_t t23a = __byte_perm(w[4 + h], w[6 + h], 0x5140);
                const uint32_t t23b = __byte_perm(w[4 + h], w[6 + h], 0x7362);
                px[4 * h + 0] = __byte_perm(t01a, t23a, 0x5410);
                px[4 * h + 1] = __byte_perm(t01a, t23a, 0x7632);
                px[4 * h + 2] = __byte_perm(t01b, t23b, 0x5410);
                px[4 * h + 3] = __byte_perm(t01b, t23b, 0x7632);
            }
        }
        const uint32_t eo = col0 + 8u * (uint32_t)c;  // offset from the start of row0
        if (aligned4) {
            // width % 4 == 0: each half (4 pixels) lies in one row and is 4*ES-byte aligned
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const uint32_t eh = eo + 4 * h;
                const uint32_t dr = __umulhi(eh, a.inv_width);  // eh < width + 8192
                const uint32_t row = row0 + dr, col = eh - dr * a.width;
                uint8_t* dstp = a.frame + (row * pitch32 + col * ES);
                if constexpr (ES == 2)
                    *reinterpret_cast<uint2*>(dstp) = make_uint2(px[2 * h], px[2 * h + 1]);
                else
                    *reinterpret_cast<uint4*>(dstp) = make_uint4(px[4 * h], px[4 * h + 1], px[4 * h + 2], px[4 * h + 3]);
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const uint32_t ei = eo + i;
                const uint32_t dr = __umulhi(ei, a.inv_width);
                const uint32_t row = row0 + dr, col = ei - dr * a.width;
                PixelT v;
                if constexpr (ES == 2) v = (PixelT)(px[i >> 1] >> (16 * (i & 1)));
                else v = (PixelT)px[i];
                reinterpret_cast<PixelT*>(a.frame + (size_t)row * a.pitch)[col] = v;
            }
        }
      }
    }
#ifdef BSLZ4_TIMELINE
    if (lane == 0 && w_id < 16384) {
        uint32_t smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        g_bslz4_timeline[w_id] = make_uint4(tl0, tl1, tl_now(), b | (smid << 16));
    }
#endif
    }  // one block
#undef BSLZ4_NEXT
}

}  // namespace ffsb
