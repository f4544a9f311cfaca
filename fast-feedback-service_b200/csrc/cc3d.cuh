// cc3d.cuh -- 3D connected components over a stack of per-frame strong-pixel lists
// (ConnectedComponents::find_3d_components, connected_components.cc:270-470).
// The lists of all frames are concatenated in ascending (z, linear index) order, so a
// compact id orders pixels exactly as the reference's global vertex ids do (:292-309).
#pragma once
#include "cc.cuh"

namespace ffsb {

struct CC3Args {
    int width, height;
    uint32_t n_slices;
    uint32_t total;            // pixels in the stack
    const uint32_t* fo;        // [n_slices + 1] offsets of each slice's list
    const uint32_t* idx;       // [total] linear index within the slice
    const uint32_t* val;       // [total]
    const uint32_t* zarr;      // [total] slice position
    uint32_t* parent;          // [total]
    uint32_t* rank;            // [total] output order of a root's spot
    SpotAcc* acc;              // [total]
    ffs_spot* spots;           // [cap_spots]
    unsigned long long* sums;  // [cap_spots][4] exact sum I, sum (2x+1)I, sum (2y+1)I, sum (2z+1)I
    uint32_t cap_spots;
    uint32_t* n_components;    // device scalar
    int row_wrap;
};

__global__ void __launch_bounds__(256) cc3_init_kernel(const CC3Args a) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < a.total; i += gridDim.x * blockDim.x) {
        a.parent[i] = i;
        acc_init(a.acc + i);
    }
}

__global__ void __launch_bounds__(256) cc3_link_kernel(const CC3Args a) {
    const uint32_t W = (uint32_t)a.width;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < a.total; i += gridDim.x * blockDim.x) {
        const uint32_t z = a.zarr[i];
        const uint32_t end = a.fo[z + 1];
        const uint32_t k = a.idx[i];
        // the 2D edges of the slice (connected_components.cc:316-343)
        if (i + 1 < end && a.idx[i + 1] == k + 1 && (a.row_wrap || (k + 1) % W != 0)) uf_union(a.parent, i, i + 1);
        const uint32_t hi = min(end, i + 1 + W);
        const uint32_t j = lower_bound(a.idx, i + 1, hi, k + W);
        if (j < hi && a.idx[j] == k + W) uf_union(a.parent, i, j);
        // same linear index in the next slice (:352-370)
        if (z + 1 < a.n_slices) {
            const uint32_t lo2 = end, hi2 = a.fo[z + 2];
            const uint32_t j2 = lower_bound(a.idx, lo2, hi2, k);
            if (j2 < hi2 && a.idx[j2] == k) uf_union(a.parent, i, j2);
        }
    }
}

__global__ void __launch_bounds__(256) cc3_reduce_kernel(const CC3Args a) {
    const uint32_t W = (uint32_t)a.width;
    const uint32_t step = gridDim.x * blockDim.x;
    const uint32_t total_round = (a.total + 31u) & ~31u;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total_round; i += step) {
        const bool active = i < a.total;
        uint32_t root = 0, x = 0, y = 0, z = 0, v = 0;
        if (active) {
            root = uf_find(a.parent, i);
            const uint32_t k = a.idx[i];
            y = k / W;
            x = k - y * W;
            z = a.zarr[i];
            v = a.val[i];
        }
        warp_segmented_acc(a.acc, active, root, x, y, z, v, i);
        if (active) a.parent[i] = root;
    }
}

// one block: ordered compaction of roots -> spots
__global__ void __launch_bounds__(256) cc3_finalize_kernel(const CC3Args a) {
    const uint32_t W = (uint32_t)a.width;
    const uint64_t px = (uint64_t)a.width * a.height;
    uint32_t running = 0;
    for (uint32_t base = 0; base < a.total; base += 256) {
        const uint32_t i = base + threadIdx.x;
        const bool is_root = i < a.total && a.parent[i] == i;
        uint32_t tot;
        const uint32_t rank = running + block_exclusive_scan_256(is_root ? 1u : 0u, &tot);
        running += tot;
        if (is_root) {
            a.rank[i] = rank;
            if (rank < a.cap_spots) {
                const SpotAcc r = a.acc[i];
                const uint32_t pid = ~(uint32_t)(r.peak & 0xffffffffu);
                const uint32_t pk = a.idx[pid];
                acc_to_spot(r, (uint64_t)a.zarr[i] * px + a.idx[i], pk % W, pk / W, (int32_t)a.zarr[pid],
                            a.spots + rank);
                a.sums[4ull * rank + 0] = r.sum_i;
                a.sums[4ull * rank + 1] = r.sum_xi;
                a.sums[4ull * rank + 2] = r.sum_yi;
                a.sums[4ull * rank + 3] = r.sum_zi;
            }
        }
    }
    if (threadIdx.x == 0) *a.n_components = running;
}

}  // namespace ffsb
