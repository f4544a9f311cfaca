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
    unsigned long long* moments;  // [cap_spots][4] sum I u^2, I v^2, I u v, I w^2 (bbox-relative), pre-zeroed
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

// one block of 1024 threads: ordered compaction of roots -> spots (ranks as in cc_finalize_kernel:
// ballot + popc inside a warp, one pass over the 32 warp totals, one barrier per 1024 entries)
__global__ void __launch_bounds__(kFinalizeThreads) cc3_finalize_kernel(const CC3Args a) {
    const uint32_t W = (uint32_t)a.width;
    const uint64_t px = (uint64_t)a.width * a.height;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __shared__ uint32_t warp_roots[2][32];
    uint32_t running = 0;
    int buf = 0;
    for (uint32_t base = 0; base < a.total; base += kFinalizeThreads, buf ^= 1) {
        const uint32_t i = base + threadIdx.x;
        const bool is_root = i < a.total && a.parent[i] == i;
        const unsigned ball = __ballot_sync(0xffffffffu, is_root);
        if (lane == 0) warp_roots[buf][warp] = __popc(ball);
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
        running += all;
    }
    if (threadIdx.x == 0) *a.n_components = running;
}

// Second moments for the Kabsch-space variances, relative to each spot's bounding-box minimum
// (exact in 64 bits: u, v, w are at most a spot's extent).  Runs after cc3_finalize_kernel:
// parent[] holds roots, rank[] their output positions, spots[] the bounding boxes.
__global__ void __launch_bounds__(256) cc3_moments_kernel(const CC3Args a) {
    const uint32_t W = (uint32_t)a.width;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < a.total; i += gridDim.x * blockDim.x) {
        const uint32_t r = a.rank[a.parent[i]];
        if (r >= a.cap_spots) continue;
        const ffs_spot* s = a.spots + r;
        const uint32_t k = a.idx[i];
        const uint32_t y = k / W, x = k - y * W;
        const unsigned long long u = x - s->x_min, v = y - s->y_min, w = a.zarr[i] - (uint32_t)s->z_min;
        const unsigned long long I = a.val[i];
        if (I == 0) continue;
        unsigned long long* m = a.moments + 4ull * r;
        if (u) atomicAdd(m + 0, I * u * u);
        if (v) atomicAdd(m + 1, I * v * v);
        if (u && v) atomicAdd(m + 2, I * u * v);
        if (w) atomicAdd(m + 3, I * w * w);
    }
}

struct KabschArgs {
    ffs_kabsch_geometry g;
    const ffs_spot* spots;
    const unsigned long long* sums;
    const unsigned long long* moments;
    uint32_t n;
    double* sigma_b;
    double* sigma_m;
    int32_t* depth;
};

// One thread per spot: the reference's per-pixel loop (connected_components.cc:181-197)
//     eps1 = e1 . (s1' - s1) / |s1|,  eps2 = e2 . (s1' - s1) / |s1|,  eps3 = (phi' - phi) * zeta
// is linear in the pixel coordinates, so its intensity-weighted squares are quadratic forms of
// the spot's second moments about the centre of mass.
__global__ void __launch_bounds__(128) kabsch_kernel(const KabschArgs a) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= a.n) return;
    const ffs_spot s = a.spots[t];
    const unsigned long long S0 = a.sums[4ull * t], Sx = a.sums[4ull * t + 1], Sy = a.sums[4ull * t + 2],
                             Sz = a.sums[4ull * t + 3];
    const unsigned long long* mo = a.moments + 4ull * t;
    const double cx = (double)s.com_x, cy = (double)s.com_y, cz = (double)s.com_z;
    // Panel(distance, beam centre, pixel size): fast = +x, slow = -y, origin = (0, 0, -distance)
    // - bcx * psx * fast - bcy * psy * slow   (dx2 detector.cxx:256-267)
    const double psx = a.g.pixel_size_x_mm, psy = a.g.pixel_size_y_mm;
    const double fast[3] = {1.0, 0.0, 0.0}, slow[3] = {0.0, -1.0, 0.0};
    double origin[3] = {0.0, 0.0, -1.0 * a.g.distance_mm};
    for (int k = 0; k < 3; ++k) {
        origin[k] -= a.g.beam_center_x_px * psx * fast[k];
        origin[k] -= a.g.beam_center_y_px * psy * slow[k];
    }
    const double xmm = cx * psx, ymm = cy * psy;  // px_to_mm, no parallax correction
    double s1[3];
    for (int k = 0; k < 3; ++k) s1[k] = fast[k] * xmm + slow[k] * ymm + origin[k];  // get_lab_coord
    const double s0[3] = {0.0, 0.0, -1.0 / a.g.wavelength};
    auto cross = [](const double* p, const double* q, double* r) {
        r[0] = p[1] * q[2] - p[2] * q[1];
        r[1] = p[2] * q[0] - p[0] * q[2];
        r[2] = p[0] * q[1] - p[1] * q[0];
    };
    auto normalize = [](double* p) {
        const double nrm = sqrt(p[0] * p[0] + p[1] * p[1] + p[2] * p[2]);
        p[0] /= nrm; p[1] /= nrm; p[2] /= nrm;
    };
    double e1[3], e2[3];
    cross(s1, s0, e1);
    normalize(e1);
    cross(s1, e1, e2);
    normalize(e2);
    const double mags1 = sqrt(s1[0] * s1[0] + s1[1] * s1[1] + s1[2] * s1[2]);
    const double zeta = e1[0];  // m2 = (1, 0, 0)
    // pixel centre minus centre of mass, per axis: (u + off), off = min + 0.5 - com
    const double ox = (double)s.x_min + 0.5 - cx, oy = (double)s.y_min + 0.5 - cy,
                 oz = (double)s.z_min + 0.5 - cz;
    const double I0 = (double)S0;
    // first moments about the box minimum, exact integers: sum I u = (S(2x+1)I - S I)/2 - x_min S I
    const double Iu = (double)((Sx - S0) / 2ull - (unsigned long long)s.x_min * S0);
    const double Iv = (double)((Sy - S0) / 2ull - (unsigned long long)s.y_min * S0);
    const double Iw = (double)((Sz - S0) / 2ull - (unsigned long long)(uint32_t)s.z_min * S0);
    const double Ixx = (double)mo[0] + 2.0 * ox * Iu + ox * ox * I0;
    const double Iyy = (double)mo[1] + 2.0 * oy * Iv + oy * oy * I0;
    const double Ixy = (double)mo[2] + ox * Iv + oy * Iu + ox * oy * I0;
    const double Izz = (double)mo[3] + 2.0 * oz * Iw + oz * oz * I0;
    // delta_s1 = dx * psx * fast + dy * psy * slow
    double A1 = 0, B1 = 0, A2 = 0, B2 = 0;
    for (int k = 0; k < 3; ++k) {
        A1 += e1[k] * fast[k];
        B1 += e1[k] * slow[k];
        A2 += e2[k] * fast[k];
        B2 += e2[k] * slow[k];
    }
    A1 *= psx / mags1; B1 *= psy / mags1; A2 *= psx / mags1; B2 *= psy / mags1;
    const double varx = (A1 * A1 * Ixx + 2.0 * A1 * B1 * Ixy + B1 * B1 * Iyy) / I0;
    const double vary = (A2 * A2 * Ixx + 2.0 * A2 * B2 * Ixy + B2 * B2 * Iyy) / I0;
    const double kz = a.g.oscillation_width_deg * (3.14159265358979323846 / 180.0) * zeta;
    const double varz = kz * kz * Izz / I0;
    if (a.sigma_b) a.sigma_b[t] = (varx + vary) / 2.0;
    if (a.sigma_m) a.sigma_m[t] = varz;
    if (a.depth) a.depth[t] = s.z_max - s.z_min + 1;
}

}  // namespace ffsb
