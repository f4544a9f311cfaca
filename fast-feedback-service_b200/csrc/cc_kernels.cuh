// cc_kernels.cuh -- the 2D (per-frame, batched) connected-components kernels.
#pragma once

namespace ffsb {

// K1: popcount per chunk.  grid (n_chunks, n_frames), 256 threads x 4 words.
__global__ void __launch_bounds__(256) cc_count_kernel(const CCArgs a) {
    const int z = a.slot0 + blockIdx.y;
    const uint32_t* plane = reinterpret_cast<const uint32_t*>(
      reinterpret_cast<const uint8_t*>(a.strong_bits) + (size_t)z * a.bits_stride);
    const int w0 = blockIdx.x * kChunkWords + threadIdx.x * 4;
    uint32_t c = 0;
    if (w0 + 3 < a.n_words) {
        const uint4 v = *reinterpret_cast<const uint4*>(plane + w0);
        c = __popc(v.x) + __popc(v.y) + __popc(v.z) + __popc(v.w);
    } else {
        for (int q = 0; q < 4; ++q)
            if (w0 + q < a.n_words) c += __popc(plane[w0 + q]);
    }
    for (int off = 16; off > 0; off >>= 1) c += __shfl_down_sync(0xffffffffu, c, off);
    __shared__ uint32_t ws[8];
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int w = 0; w < 8; ++w) t += ws[w];
        a.chunk_count[(size_t)z * a.chunk_stride + blockIdx.x] = t;
    }
}

// K2: exclusive scan of the chunk counts of one frame.  grid (n_frames), 256 threads.
__global__ void __launch_bounds__(256) cc_scan_kernel(const CCArgs a) {
    const int z = a.slot0 + blockIdx.x;
    uint32_t* cc = a.chunk_count + (size_t)z * a.chunk_stride;
    uint32_t running = 0;
    for (int base = 0; base < a.n_chunks; base += 256) {
        const int k = base + threadIdx.x;
        const uint32_t v = k < a.n_chunks ? cc[k] : 0u;
        uint32_t tot;
        const uint32_t ex = block_exclusive_scan_256(v, &tot);
        if (k < a.n_chunks) cc[k] = running + ex;
        running += tot;
    }
    if (threadIdx.x == 0) {
        uint32_t* cnt = a.counts + (size_t)z * CNT_N;
        cnt[CNT_STRONG_RAW] = running;
        cnt[CNT_STRONG] = min(running, a.cap);
        cnt[CNT_STATUS] = running > a.cap ? (uint32_t)FFS_ERR_CAPACITY : 0u;
        cnt[CNT_COMPONENTS] = cnt[CNT_SPOTS_TOTAL] = cnt[CNT_STRONG_FILTERED] = cnt[CNT_SPOTS_FILTERED] = 0;
    }
}

// K3: ordered extraction.  grid (n_chunks, n_frames), 256 threads x 4 words.
__global__ void __launch_bounds__(256) cc_extract_kernel(const CCArgs a) {
    const int z = a.slot0 + blockIdx.y;
    const uint32_t* plane = reinterpret_cast<const uint32_t*>(
      reinterpret_cast<const uint8_t*>(a.strong_bits) + (size_t)z * a.bits_stride);
    const uint8_t* frame = a.image + (size_t)z * a.frame_stride;
    const int w0 = blockIdx.x * kChunkWords + threadIdx.x * 4;
    uint32_t words[4];
    uint32_t c = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        words[q] = (w0 + q < a.n_words) ? plane[w0 + q] : 0u;
        c += __popc(words[q]);
    }
    uint32_t tot;
    uint32_t off = block_exclusive_scan_256(c, &tot);
    if (tot == 0) return;
    off += a.chunk_count[(size_t)z * a.chunk_stride + blockIdx.x];
    uint32_t* idx = a.px_index + (size_t)z * a.cap;
    uint32_t* val = a.px_value + (size_t)z * a.cap;
    uint32_t* par = a.parent + (size_t)z * a.cap;
    SpotAcc* acc = a.acc + (size_t)z * a.cap;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        uint32_t w = words[q];
        if (!w) continue;
        const int p = w0 + q;
        const int row = p / a.words_per_row;
        const int cbase = (p - row * a.words_per_row) * 32;
        while (w) {
            const int b = __ffs(w) - 1;
            w &= w - 1;
            if (off < a.cap) {
                const int col = cbase + b;
                idx[off] = (uint32_t)row * (uint32_t)a.width + (uint32_t)col;
                val[off] = a.bpp == 2
                             ? (uint32_t) reinterpret_cast<const uint16_t*>(frame + (size_t)row * a.pitch)[col]
                             : reinterpret_cast<const uint32_t*>(frame + (size_t)row * a.pitch)[col];
                par[off] = off;
                acc_init(acc + off);
            }
            ++off;
        }
    }
}

// K4: link each strong pixel to its right and lower neighbour.  grid (blocks, n_frames).
__global__ void __launch_bounds__(256) cc_link_kernel(const CCArgs a) {
    const int z = a.slot0 + blockIdx.y;
    const uint32_t S = a.counts[(size_t)z * CNT_N + CNT_STRONG];
    const uint32_t* idx = a.px_index + (size_t)z * a.cap;
    uint32_t* par = a.parent + (size_t)z * a.cap;
    const uint32_t W = (uint32_t)a.width;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < S; i += gridDim.x * blockDim.x) {
        const uint32_t k = idx[i];
        // edge to k+1 (connected_components.cc:62-66; row_wrap = reference behaviour)
        if (i + 1 < S && idx[i + 1] == k + 1 && (a.row_wrap || (k + 1) % W != 0)) uf_union(par, i, i + 1);
        // edge to k+width (:67-70): search the next `width` indices
        const uint32_t hi = min(S, i + 1 + W);
        const uint32_t j = lower_bound(idx, i + 1, hi, k + W);
        if (j < hi && idx[j] == k + W) uf_union(par, i, j);
    }
}

// Add one pixel (or a warp-aggregated run) into its root's accumulator.
__device__ __forceinline__ void acc_add(SpotAcc* r, uint32_t cnt, unsigned long long si,
                                        unsigned long long sx, unsigned long long sy,
                                        unsigned long long sz, unsigned long long peak, uint32_t xmin,
                                        uint32_t xmax, uint32_t ymin, uint32_t ymax, uint32_t zmin,
                                        uint32_t zmax) {
    atomicAdd(&r->count, cnt);
    atomicAdd(&r->sum_i, si);
    atomicAdd(&r->sum_xi, sx);
    atomicAdd(&r->sum_yi, sy);
    if (sz) atomicAdd(&r->sum_zi, sz);
    atomicMax(&r->peak, peak);
    atomicMin(&r->x_min, xmin);
    atomicMax(&r->x_max, xmax);
    atomicMin(&r->y_min, ymin);
    atomicMax(&r->y_max, ymax);
    atomicMin(&r->z_min, zmin);
    atomicMax(&r->z_max, zmax);
}

// Warp-level segmented reduction over runs of equal `root` in consecutive lanes, then one
// set of atomics per run (pixels of one spot in one row sit in consecutive list entries).
__device__ __forceinline__ void warp_segmented_acc(SpotAcc* accbase, bool active, uint32_t root,
                                                   uint32_t x, uint32_t y, uint32_t zpos, uint32_t inten,
                                                   uint32_t id) {
    const int lane = threadIdx.x & 31;
    uint32_t cnt = active ? 1u : 0u;
    unsigned long long si = active ? inten : 0ull;
    unsigned long long sx = si * (2ull * x + 1ull), sy = si * (2ull * y + 1ull);
    unsigned long long sz = si * (2ull * zpos + 1ull);
    unsigned long long peak = active ? (((unsigned long long)inten << 32) | (uint32_t)~id) : 0ull;
    uint32_t xmin = x, xmax = x, ymin = y, ymax = y, zmin = zpos, zmax = zpos;
    const uint32_t key = active ? root : 0xffffffffu - lane;  // inactive lanes never match
    const uint32_t prev = __shfl_up_sync(0xffffffffu, key, 1);
    const bool head = lane == 0 || prev != key;
    // a run ends where the next head starts (equal roots may reappear later in the warp,
    // e.g. the next row of the same spot, so key equality alone is not enough)
    const unsigned heads = __ballot_sync(0xffffffffu, head);
    const unsigned later = lane == 31 ? 0u : heads >> (lane + 1);
    const int run_end = later ? lane + __ffs(later) : 32;  // one past the last lane of my run
    // segmented inclusive suffix reduce: fold lane+off into lane while still in the same run
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t k2 = __shfl_down_sync(0xffffffffu, key, off);
        const uint32_t c2 = __shfl_down_sync(0xffffffffu, cnt, off);
        const unsigned long long si2 = __shfl_down_sync(0xffffffffu, si, off);
        const unsigned long long sx2 = __shfl_down_sync(0xffffffffu, sx, off);
        const unsigned long long sy2 = __shfl_down_sync(0xffffffffu, sy, off);
        const unsigned long long sz2 = __shfl_down_sync(0xffffffffu, sz, off);
        const unsigned long long pk2 = __shfl_down_sync(0xffffffffu, peak, off);
        const uint32_t xn2 = __shfl_down_sync(0xffffffffu, xmin, off);
        const uint32_t xx2 = __shfl_down_sync(0xffffffffu, xmax, off);
        const uint32_t yn2 = __shfl_down_sync(0xffffffffu, ymin, off);
        const uint32_t yx2 = __shfl_down_sync(0xffffffffu, ymax, off);
        const uint32_t zn2 = __shfl_down_sync(0xffffffffu, zmin, off);
        const uint32_t zx2 = __shfl_down_sync(0xffffffffu, zmax, off);
        if (lane + off < run_end) {
            (void)k2;
            cnt += c2;
            si += si2;
            sx += sx2;
            sy += sy2;
            sz += sz2;
            peak = max(peak, pk2);
            xmin = min(xmin, xn2);
            xmax = max(xmax, xx2);
            ymin = min(ymin, yn2);
            ymax = max(ymax, yx2);
            zmin = min(zmin, zn2);
            zmax = max(zmax, zx2);
        }
    }
    if (active && head) acc_add(accbase + root, cnt, si, sx, sy, sz, peak, xmin, xmax, ymin, ymax, zmin, zmax);
}

// K5: flatten + per-spot reduction.  grid (blocks, n_frames).
__global__ void __launch_bounds__(256) cc_reduce_kernel(const CCArgs a) {
    const int z = a.slot0 + blockIdx.y;
    const uint32_t S = a.counts[(size_t)z * CNT_N + CNT_STRONG];
    const uint32_t* idx = a.px_index + (size_t)z * a.cap;
    const uint32_t* val = a.px_value + (size_t)z * a.cap;
    uint32_t* par = a.parent + (size_t)z * a.cap;
    SpotAcc* acc = a.acc + (size_t)z * a.cap;
    const uint32_t W = (uint32_t)a.width;
    const uint32_t step = gridDim.x * blockDim.x;
    const uint32_t S_round = (S + 31u) & ~31u;  // keep warps converged for the shuffles
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < S_round; i += step) {
        const bool active = i < S;
        uint32_t root = 0, x = 0, y = 0, v = 0;
        if (active) {
            root = uf_find(par, i);
            const uint32_t k = idx[i];
            y = k / W;
            x = k - y * W;
            v = val[i];
        }
        warp_segmented_acc(acc, active, root, x, y, 0u, v, i);
        // roots are final here (K4 finished), so storing the root is a pure compression
        if (active) par[i] = root;
    }
}

// K6: ordered compaction of roots into spot records + per-frame counters.
// grid (n_frames), 1024 threads; loops over the frame's compact list 1024 entries at a time.
// A root's rank is the number of roots before it in list order: ballot + popc inside the warp,
// one shared-memory pass over the 32 warp totals across warps.
constexpr int kFinalizeThreads = 1024;
__global__ void __launch_bounds__(kFinalizeThreads) cc_finalize_kernel(const CCArgs a) {
    const int z = a.slot0 + blockIdx.x;
    uint32_t* cnt = a.counts + (size_t)z * CNT_N;
    const uint32_t S = cnt[CNT_STRONG];
    const uint32_t* idx = a.px_index + (size_t)z * a.cap;
    const uint32_t* par = a.parent + (size_t)z * a.cap;
    const SpotAcc* acc = a.acc + (size_t)z * a.cap;
    ffs_spot* spots = a.spots + (size_t)z * a.cap_spots;
    const uint32_t W = (uint32_t)a.width;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __shared__ uint32_t warp_roots[2][32];  // double-buffered: one barrier per round
    __shared__ uint32_t totals[3];
    if (threadIdx.x < 3) totals[threadIdx.x] = 0;
    uint32_t running = 0, n_total = 0, n_strong = 0, n_filtered = 0;
    int buf = 0;
    for (uint32_t base = 0; base < S; base += kFinalizeThreads, buf ^= 1) {
        const uint32_t i = base + threadIdx.x;
        const bool is_root = i < S && par[i] == i;
        const unsigned ball = __ballot_sync(0xffffffffu, is_root);
        if (lane == 0) warp_roots[buf][warp] = __popc(ball);
        // start the root's loads before the barrier
        SpotAcc r;
        uint32_t first = 0;
        if (is_root) {
            r = acc[i];
            first = idx[i];
        }
        __syncthreads();
        const uint32_t wr = warp_roots[buf][lane];
        uint32_t before = lane < warp ? wr : 0u, all = wr;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            before += __shfl_xor_sync(0xffffffffu, before, off);
            all += __shfl_xor_sync(0xffffffffu, all, off);
        }
        if (is_root) {
            const uint32_t rank = running + before + __popc(ball & ((1u << lane) - 1u));
            const uint32_t pid = ~(uint32_t)(r.peak & 0xffffffffu);
            const uint32_t pk = idx[pid];
            ffs_spot s;
            acc_to_spot(r, (uint64_t)first, pk % W, pk / W, 0, &s);
            if (rank < a.cap_spots) spots[rank] = s;
            // get_boxes keeps num_pixels >= min_spot_size (connected_components.cc:121-131);
            // filter_reflections then drops distance > max_sep (:207-236)
            if (!(a.min_spot_size > 0 && s.num_pixels < a.min_spot_size)) {
                ++n_total;
                n_strong += s.num_pixels;
                if (!(a.max_sep > 0 && s.peak_centroid_distance > a.max_sep)) ++n_filtered;
            }
        }
        running += all;
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        n_total += __shfl_xor_sync(0xffffffffu, n_total, off);
        n_strong += __shfl_xor_sync(0xffffffffu, n_strong, off);
        n_filtered += __shfl_xor_sync(0xffffffffu, n_filtered, off);
    }
    __syncthreads();  // totals[] initialised
    if (lane == 0 && (n_total | n_strong | n_filtered)) {
        atomicAdd(&totals[0], n_total);
        atomicAdd(&totals[1], n_strong);
        atomicAdd(&totals[2], n_filtered);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        cnt[CNT_SPOTS_TOTAL] = totals[0];
        cnt[CNT_STRONG_FILTERED] = totals[1];
        cnt[CNT_SPOTS_FILTERED] = totals[2];
        cnt[CNT_COMPONENTS] = running;
        if (running > a.cap_spots) cnt[CNT_STATUS] = (uint32_t)FFS_ERR_CAPACITY;
    }
}

}  // namespace ffsb
