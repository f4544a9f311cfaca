// api_sweep.inl -- rotation sweeps: record per-frame strong-pixel lists, then label the whole
// stack on the device (cc3d.cuh).  Replaces rotation_slices + find_3d_components
// (spotfinder.cc:913-918,1099-1112; connected_components.cc:270-470).

extern "C" int ffs_sweep_begin(ffs_ctx* c, uint64_t max_frames) {
    if (!c) return FFS_ERR_ARG;
    std::lock_guard<std::mutex> lk(c->mu);
    c->sweep_on = true;
    c->sweep_ids.clear();
    c->sweep_index.clear();
    c->sweep_value.clear();
    c->sweep_ids.reserve(max_frames);
    c->sweep_index.reserve(max_frames);
    c->sweep_value.reserve(max_frames);
    c->sweep_plane[0].clear();
    c->sweep_plane[1].clear();
    c->sweep_sums.clear();
    c->sweep_moments.clear();
    c->sweep_overflow = false;
    return FFS_OK;
}

// called from ffs_wait / ffs_sweep_record_slot once the slot's results are complete
static int sweep_record(ffs_ctx* c, int slot, uint64_t frame_id, uint32_t n_strong) {
    // a frame whose strong-pixel list hit its capacity would contribute a truncated plane and the
    // 3D components would come out wrong without notice: remember it, ffs_sweep_finish refuses
    if ((int32_t)c->h_frame_counts[(size_t)slot * CNT_N + CNT_STATUS] != 0 && c->slots[slot].state == 2)
        c->sweep_overflow = true;
    std::vector<uint32_t> idx(n_strong), val(n_strong);
    if (n_strong) {
        FFS_CK(c, cudaMemcpy(idx.data(), c->d_px_index + (size_t)slot * c->cap_strong, n_strong * 4ull,
                             cudaMemcpyDeviceToHost));
        FFS_CK(c, cudaMemcpy(val.data(), c->d_px_value + (size_t)slot * c->cap_strong, n_strong * 4ull,
                             cudaMemcpyDeviceToHost));
    }
    std::lock_guard<std::mutex> lk(c->mu);
    c->sweep_ids.push_back(frame_id);
    c->sweep_index.push_back(std::move(idx));
    c->sweep_value.push_back(std::move(val));
    return FFS_OK;
}

// Slot-API counterpart of the recording done by ffs_wait: record the last ffs_process
// result of `slot` as frame `frame_id` of the sweep.
extern "C" int ffs_sweep_record_slot(ffs_ctx* c, int slot, uint64_t frame_id) {
    if (!slots_ok(c, slot, 1) || !c->sweep_on) return FFS_ERR_ARG;
    FFS_CK(c, cudaSetDevice(c->p.device));
    FFS_CK(c, cudaDeviceSynchronize());
    uint32_t cnt[CNT_N];
    FFS_CK(c, cudaMemcpy(cnt, c->d_frame_counts + (size_t)slot * CNT_N, sizeof(cnt), cudaMemcpyDeviceToHost));
    if ((int32_t)cnt[CNT_STATUS] != 0) return FFS_ERR_CAPACITY;
    return sweep_record(c, slot, frame_id, cnt[CNT_STRONG]);
}

extern "C" int ffs_sweep_finish(ffs_ctx* c, ffs_spot** spots_out, size_t* n_out) {
    if (!c || !spots_out || !n_out || !c->sweep_on) return FFS_ERR_ARG;
    *spots_out = nullptr;
    *n_out = 0;
    FFS_CK(c, cudaSetDevice(c->p.device));
    // slices ordered by frame id (std::map order, spotfinder.cc:1105-1108); z = position
    const size_t nz = c->sweep_ids.size();
    std::vector<size_t> order(nz);
    for (size_t i = 0; i < nz; ++i) order[i] = i;
    std::sort(order.begin(), order.end(), [&](size_t x, size_t y) { return c->sweep_ids[x] < c->sweep_ids[y]; });
    c->sweep_on = false;
    if (c->sweep_overflow) {
        c->err = "a frame of the sweep exceeded max_strong_per_frame: 3D components would be wrong";
        return FFS_ERR_CAPACITY;
    }
    std::vector<uint32_t> fo(nz + 1, 0);
    uint64_t total64 = 0;
    for (size_t z = 0; z < nz; ++z) {
        total64 += c->sweep_index[order[z]].size();
        fo[z + 1] = (uint32_t)total64;
    }
    if (total64 >= 0xffffffffull) {  // compact ids are 32-bit
        c->err = "sweep holds more than 2^32 strong pixels";
        return FFS_ERR_CAPACITY;
    }
    const size_t total = (size_t)total64;
    if (total == 0) return FFS_OK;
    std::vector<uint32_t> idx(total), val(total), zarr(total);
    for (size_t z = 0; z < nz; ++z) {
        const auto& si = c->sweep_index[order[z]];
        const auto& sv = c->sweep_value[order[z]];
        std::copy(si.begin(), si.end(), idx.begin() + fo[z]);
        std::copy(sv.begin(), sv.end(), val.begin() + fo[z]);
        std::fill(zarr.begin() + fo[z], zarr.begin() + fo[z + 1], (uint32_t)z);
    }
    uint32_t *d_fo = nullptr, *d_idx = nullptr, *d_val = nullptr, *d_z = nullptr, *d_par = nullptr, *d_rank = nullptr,
             *d_n = nullptr;
    SpotAcc* d_acc = nullptr;
    ffs_spot* d_spots = nullptr;
    unsigned long long *d_sums = nullptr, *d_mom = nullptr;
    int rc = FFS_OK;
    ffs_spot* out = nullptr;
    bool keep_out = false;
    auto cleanup = [&]() {
        if (!keep_out) free(out);
        cudaFree(d_fo); cudaFree(d_idx); cudaFree(d_val); cudaFree(d_z); cudaFree(d_par);
        cudaFree(d_rank); cudaFree(d_n); cudaFree(d_acc); cudaFree(d_spots); cudaFree(d_sums); cudaFree(d_mom);
    };
#define SWEEP_CK(call)                                                        \
    do {                                                                      \
        cudaError_t e__ = (call);                                             \
        if (e__ != cudaSuccess) {                                             \
            c->err = std::string(#call) + ": " + cudaGetErrorString(e__);     \
            cleanup();                                                        \
            return FFS_ERR_CUDA;                                              \
        }                                                                     \
    } while (0)
    SWEEP_CK(cudaMalloc(&d_fo, (nz + 1) * 4));
    SWEEP_CK(cudaMalloc(&d_idx, total * 4));
    SWEEP_CK(cudaMalloc(&d_val, total * 4));
    SWEEP_CK(cudaMalloc(&d_z, total * 4));
    SWEEP_CK(cudaMalloc(&d_par, total * 4));
    SWEEP_CK(cudaMalloc(&d_rank, total * 4));
    SWEEP_CK(cudaMalloc(&d_n, 4));
    SWEEP_CK(cudaMalloc(&d_acc, total * sizeof(SpotAcc)));
    SWEEP_CK(cudaMalloc(&d_spots, total * sizeof(ffs_spot)));
    SWEEP_CK(cudaMalloc(&d_sums, total * 4 * sizeof(unsigned long long)));
    SWEEP_CK(cudaMalloc(&d_mom, total * 4 * sizeof(unsigned long long)));
    SWEEP_CK(cudaMemsetAsync(d_mom, 0, total * 4 * sizeof(unsigned long long), c->stream));
    SWEEP_CK(cudaMemcpy(d_fo, fo.data(), (nz + 1) * 4, cudaMemcpyHostToDevice));
    SWEEP_CK(cudaMemcpy(d_idx, idx.data(), total * 4, cudaMemcpyHostToDevice));
    SWEEP_CK(cudaMemcpy(d_val, val.data(), total * 4, cudaMemcpyHostToDevice));
    SWEEP_CK(cudaMemcpy(d_z, zarr.data(), total * 4, cudaMemcpyHostToDevice));
    CC3Args a{};
    a.width = (int)c->p.width;
    a.height = (int)c->p.height;
    a.n_slices = (uint32_t)nz;
    a.total = (uint32_t)total;
    a.fo = d_fo; a.idx = d_idx; a.val = d_val; a.zarr = d_z;
    a.parent = d_par; a.rank = d_rank; a.acc = d_acc; a.spots = d_spots; a.sums = d_sums; a.moments = d_mom;
    a.cap_spots = (uint32_t)total;
    a.n_components = d_n;
    a.row_wrap = c->p.cc_row_wrap;
    const unsigned grid = (unsigned)std::min<size_t>((total + 255) / 256, (size_t)c->sm_count * 8);
    cc3_init_kernel<<<grid, 256, 0, c->stream>>>(a);
    cc3_link_kernel<<<grid, 256, 0, c->stream>>>(a);
    cc3_reduce_kernel<<<grid, 256, 0, c->stream>>>(a);
    cc3_finalize_kernel<<<1, kFinalizeThreads, 0, c->stream>>>(a);
    cc3_moments_kernel<<<grid, 256, 0, c->stream>>>(a);
    c->launches += 5;
    SWEEP_CK(cudaGetLastError());
    uint32_t n = 0;
    SWEEP_CK(cudaMemcpyAsync(&n, d_n, 4, cudaMemcpyDeviceToHost, c->stream));
    SWEEP_CK(cudaStreamSynchronize(c->stream));
    out = n ? (ffs_spot*)malloc(sizeof(ffs_spot) * n) : nullptr;
    if (n && !out) {
        cleanup();
        return FFS_ERR_NOMEM;
    }
    if (n) SWEEP_CK(cudaMemcpy(out, d_spots, sizeof(ffs_spot) * n, cudaMemcpyDeviceToHost));
    c->sweep_sums.assign((size_t)n * 4, 0);
    if (n) SWEEP_CK(cudaMemcpy(c->sweep_sums.data(), d_sums, sizeof(uint64_t) * 4 * n, cudaMemcpyDeviceToHost));
    c->sweep_moments.assign((size_t)n * 4, 0);
    if (n) SWEEP_CK(cudaMemcpy(c->sweep_moments.data(), d_mom, sizeof(uint64_t) * 4 * n, cudaMemcpyDeviceToHost));
    // boundary planes for slab-sharded sweeps: {linear index, component rank} of the first
    // and last slice
    {
        std::vector<uint32_t> par(total), rank(total);
        SWEEP_CK(cudaMemcpy(par.data(), d_par, total * 4, cudaMemcpyDeviceToHost));
        SWEEP_CK(cudaMemcpy(rank.data(), d_rank, total * 4, cudaMemcpyDeviceToHost));
        for (int which = 0; which < 2; ++which) {
            const size_t z = which == 0 ? 0 : nz - 1;
            auto& pl = c->sweep_plane[which];
            pl.clear();
            for (uint32_t i = fo[z]; i < fo[z + 1]; ++i) {
                pl.push_back(idx[i]);
                pl.push_back(rank[par[i]]);
            }
        }
    }
#undef SWEEP_CK
    keep_out = true;
    cleanup();
    *spots_out = out;
    *n_out = n;
    return rc;
}

extern "C" int ffs_sweep_export_plane(ffs_ctx* c, int which, uint32_t** pairs, size_t* n_pixels) {
    if (!c || !pairs || !n_pixels || which < 0 || which > 1) return FFS_ERR_ARG;
    const auto& pl = c->sweep_plane[which];
    *n_pixels = pl.size() / 2;
    *pairs = nullptr;
    if (pl.empty()) return FFS_OK;
    *pairs = (uint32_t*)malloc(pl.size() * 4);
    if (!*pairs) return FFS_ERR_NOMEM;
    memcpy(*pairs, pl.data(), pl.size() * 4);
    return FFS_OK;
}

extern "C" int ffs_sweep_export_sums(ffs_ctx* c, uint64_t** sums, size_t* n_spots) {
    if (!c || !sums || !n_spots) return FFS_ERR_ARG;
    *n_spots = c->sweep_sums.size() / 4;
    *sums = nullptr;
    if (c->sweep_sums.empty()) return FFS_OK;
    *sums = (uint64_t*)malloc(c->sweep_sums.size() * sizeof(uint64_t));
    if (!*sums) return FFS_ERR_NOMEM;
    memcpy(*sums, c->sweep_sums.data(), c->sweep_sums.size() * sizeof(uint64_t));
    return FFS_OK;
}

extern "C" int ffs_sweep_export_moments(ffs_ctx* c, uint64_t** moments, size_t* n_spots) {
    if (!c || !moments || !n_spots) return FFS_ERR_ARG;
    *n_spots = c->sweep_moments.size() / 4;
    *moments = nullptr;
    if (c->sweep_moments.empty()) return FFS_OK;
    *moments = (uint64_t*)malloc(c->sweep_moments.size() * sizeof(uint64_t));
    if (!*moments) return FFS_ERR_NOMEM;
    memcpy(*moments, c->sweep_moments.data(), c->sweep_moments.size() * sizeof(uint64_t));
    return FFS_OK;
}

extern "C" int ffs_kabsch_variances(ffs_ctx* c, const ffs_kabsch_geometry* g, const ffs_spot* spots,
                                    const uint64_t* sums, const uint64_t* moments, size_t n,
                                    double* sigma_b, double* sigma_m, int32_t* depth) {
    if (!c || !g || (n && (!spots || !sums || !moments))) return FFS_ERR_ARG;
    if (n == 0) return FFS_OK;
    FFS_CK(c, cudaSetDevice(c->p.device));
    ffs_spot* d_spots = nullptr;
    unsigned long long *d_sums = nullptr, *d_mom = nullptr;
    double *d_b = nullptr, *d_m = nullptr;
    int32_t* d_d = nullptr;
    auto cleanup = [&]() {
        cudaFree(d_spots); cudaFree(d_sums); cudaFree(d_mom); cudaFree(d_b); cudaFree(d_m); cudaFree(d_d);
    };
#define KAB_CK(call)                                                          \
    do {                                                                      \
        cudaError_t e__ = (call);                                             \
        if (e__ != cudaSuccess) {                                             \
            c->err = std::string(#call) + ": " + cudaGetErrorString(e__);     \
            cleanup();                                                        \
            return FFS_ERR_CUDA;                                              \
        }                                                                     \
    } while (0)
    KAB_CK(cudaMalloc(&d_spots, n * sizeof(ffs_spot)));
    KAB_CK(cudaMalloc(&d_sums, n * 4 * sizeof(uint64_t)));
    KAB_CK(cudaMalloc(&d_mom, n * 4 * sizeof(uint64_t)));
    KAB_CK(cudaMalloc(&d_b, n * sizeof(double)));
    KAB_CK(cudaMalloc(&d_m, n * sizeof(double)));
    KAB_CK(cudaMalloc(&d_d, n * sizeof(int32_t)));
    KAB_CK(cudaMemcpyAsync(d_spots, spots, n * sizeof(ffs_spot), cudaMemcpyHostToDevice, c->stream));
    KAB_CK(cudaMemcpyAsync(d_sums, sums, n * 4 * sizeof(uint64_t), cudaMemcpyHostToDevice, c->stream));
    KAB_CK(cudaMemcpyAsync(d_mom, moments, n * 4 * sizeof(uint64_t), cudaMemcpyHostToDevice, c->stream));
    KabschArgs a{};
    a.g = *g;
    a.spots = d_spots;
    a.sums = d_sums;
    a.moments = d_mom;
    a.n = (uint32_t)n;
    a.sigma_b = d_b;
    a.sigma_m = d_m;
    a.depth = d_d;
    kabsch_kernel<<<(unsigned)((n + 127) / 128), 128, 0, c->stream>>>(a);
    c->launches++;
    KAB_CK(cudaGetLastError());
    if (sigma_b) KAB_CK(cudaMemcpyAsync(sigma_b, d_b, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (sigma_m) KAB_CK(cudaMemcpyAsync(sigma_m, d_m, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (depth) KAB_CK(cudaMemcpyAsync(depth, d_d, n * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    KAB_CK(cudaStreamSynchronize(c->stream));
#undef KAB_CK
    cleanup();
    return FFS_OK;
}
