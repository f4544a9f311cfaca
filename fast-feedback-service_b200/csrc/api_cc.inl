// api_cc.inl -- connected components / spot reduction: setup, ffs_process, ffs_fetch, labels.

static int cc_setup(ffs_ctx* c) {
    const ffs_params& p = c->p;
    const size_t px = (size_t)p.width * p.height;
    const size_t S = p.n_slots;
    c->cap_strong = p.max_strong_per_frame
                      ? p.max_strong_per_frame
                      : (uint32_t)std::min<size_t>(std::max<size_t>(px / 16, 65536), 1u << 20);
    c->cap_spots = p.max_spots_per_frame ? p.max_spots_per_frame : std::max<uint32_t>(c->cap_strong / 4, 1024);
    FFS_CK(c, cudaMalloc(&c->d_px_index, sizeof(uint32_t) * c->cap_strong * S));
    FFS_CK(c, cudaMalloc(&c->d_px_value, sizeof(uint32_t) * c->cap_strong * S));
    FFS_CK(c, cudaMalloc(&c->d_px_parent, sizeof(uint32_t) * c->cap_strong * S));
    FFS_CK(c, cudaMalloc(&c->d_acc, sizeof(SpotAcc) * c->cap_strong * S));
    FFS_CK(c, cudaMalloc(&c->d_spots, sizeof(ffs_spot) * c->cap_spots * S));
    FFS_CK(c, cudaMalloc(&c->d_frame_counts, sizeof(uint32_t) * CNT_N * S));
    FFS_CK(c, cudaMemset(c->d_frame_counts, 0, sizeof(uint32_t) * CNT_N * S));
    return FFS_OK;
}

static void cc_teardown(ffs_ctx* c) {
    cudaFree(c->d_px_index);
    cudaFree(c->d_px_value);
    cudaFree(c->d_px_parent);
    cudaFree(c->d_acc);
    cudaFree(c->d_spots);
    cudaFree(c->d_frame_counts);
}

static void fill_cc_args(ffs_ctx* c, CCArgs& a, int first_slot, int n) {
    a.width = (int)c->p.width;
    a.height = (int)c->p.height;
    a.n_frames = n;
    a.slot0 = first_slot;
    a.strong_bits = reinterpret_cast<const uint32_t*>(c->d_strong);
    a.bits_pitch = c->bits_pitch;
    a.bits_stride = c->zstride;  // strong planes live in the per-slot zero region
    a.image = c->d_frames;
    a.pitch = c->pitch;
    a.frame_stride = c->frame_stride;
    a.bpp = (int)c->p.bytes_per_pixel;
    a.words_per_row = (int)(c->bits_pitch / 4);
    a.n_words = a.words_per_row * a.height;
    a.n_chunks = c->n_chunks;
    a.cap = c->cap_strong;
    a.px_index = c->d_px_index;
    a.px_value = c->d_px_value;
    a.parent = c->d_px_parent;
    a.acc = c->d_acc;
    a.chunk_count = c->d_chunk_count;
    a.chunk_stride = c->zstride / 4;
    a.counts = c->d_frame_counts;
    a.spots = c->d_spots;
    a.cap_spots = c->cap_spots;
    a.row_wrap = c->p.cc_row_wrap;
    a.min_spot_size = c->p.min_spot_size;
    a.max_sep = c->p.max_peak_centroid_separation;
}

// strong bit-planes of slots [first, first+n) -> spot lists + counters (all on `st`)
static int cc_impl(ffs_ctx* c, int first_slot, int n, cudaStream_t st) {
    CCArgs a{};
    fill_cc_args(c, a, first_slot, n);
    const dim3 gchunks((unsigned)a.n_chunks, (unsigned)n);
    const dim3 gpix((unsigned)std::max(1, c->sm_count / std::max(1, std::min(n, 8))), (unsigned)n);
    // chunk_count was filled by the threshold kernels as they set the strong bits (set_strong);
    // cc_count_kernel is kept for planes that come from elsewhere
    static const int recount = env_int("FFS_CC_RECOUNT", 0);
    if (recount) {
        cc_count_kernel<<<gchunks, 256, 0, st>>>(a);
        c->launches++;
    }
    cc_scan_kernel<<<(unsigned)n, 256, 0, st>>>(a);
    cc_extract_kernel<<<gchunks, 256, 0, st>>>(a);
    cc_link_kernel<<<gpix, 256, 0, st>>>(a);
    cc_reduce_kernel<<<gpix, 256, 0, st>>>(a);
    cc_finalize_kernel<<<(unsigned)n, kFinalizeThreads, 0, st>>>(a);
    c->launches += 5;
    FFS_CK(c, cudaGetLastError());
    return FFS_OK;
}

extern "C" int ffs_process(ffs_ctx* c, int first_slot, int n, void* stream) {
    if (!slots_ok(c, first_slot, n)) return FFS_ERR_ARG;
    FFS_CK(c, cudaSetDevice(c->p.device));
    cudaStream_t st = pick_stream(c, stream);
    int rc = threshold_impl(c, first_slot, n, st);
    if (rc) return rc;
    return cc_impl(c, first_slot, n, st);
}

static void counts_to_result(const uint32_t* cnt, uint64_t frame_id, ffs_frame_result* r) {
    r->frame_id = frame_id;
    r->num_strong_pixels = cnt[CNT_STRONG_RAW];
    r->n_components = cnt[CNT_COMPONENTS];
    r->n_spots_total = cnt[CNT_SPOTS_TOTAL];
    r->num_strong_pixels_filtered = cnt[CNT_STRONG_FILTERED];
    r->n_spots_filtered = cnt[CNT_SPOTS_FILTERED];
    r->status = (int32_t)cnt[CNT_STATUS];
}

extern "C" int ffs_fetch(ffs_ctx* c, int slot, ffs_frame_result* result, ffs_spot* spots,
                         size_t max_spots, size_t* n_spots) {
    if (!slots_ok(c, slot, 1)) return FFS_ERR_ARG;
    FFS_CK(c, cudaSetDevice(c->p.device));
    FFS_CK(c, cudaDeviceSynchronize());
    uint32_t cnt[CNT_N];
    FFS_CK(c, cudaMemcpy(cnt, c->d_frame_counts + (size_t)slot * CNT_N, sizeof(cnt), cudaMemcpyDeviceToHost));
    if (result) counts_to_result(cnt, (uint64_t)slot, result);
    const size_t have = std::min<size_t>(cnt[CNT_COMPONENTS], c->cap_spots);
    if (n_spots) *n_spots = have;
    if (spots && max_spots) {
        const size_t k = std::min(have, max_spots);
        if (k) FFS_CK(c, cudaMemcpy(spots, c->d_spots + (size_t)slot * c->cap_spots, k * sizeof(ffs_spot),
                                    cudaMemcpyDeviceToHost));
    }
    return FFS_OK;
}

extern "C" int ffs_get_labels(ffs_ctx* c, int slot, int64_t* labels) {
    if (!slots_ok(c, slot, 1) || !labels) return FFS_ERR_ARG;
    FFS_CK(c, cudaSetDevice(c->p.device));
    FFS_CK(c, cudaDeviceSynchronize());
    uint32_t cnt[CNT_N];
    FFS_CK(c, cudaMemcpy(cnt, c->d_frame_counts + (size_t)slot * CNT_N, sizeof(cnt), cudaMemcpyDeviceToHost));
    const size_t S = cnt[CNT_STRONG];
    std::vector<uint32_t> idx(S), par(S);
    if (S) {
        FFS_CK(c, cudaMemcpy(idx.data(), c->d_px_index + (size_t)slot * c->cap_strong, S * 4, cudaMemcpyDeviceToHost));
        FFS_CK(c, cudaMemcpy(par.data(), c->d_px_parent + (size_t)slot * c->cap_strong, S * 4, cudaMemcpyDeviceToHost));
    }
    const size_t px = (size_t)c->p.width * c->p.height;
    for (size_t k = 0; k < px; ++k) labels[k] = -1;
    for (size_t i = 0; i < S; ++i) labels[idx[i]] = (int64_t)idx[par[i]];  // parents are flattened roots
    return FFS_OK;
}

// filter_reflections, connected_components.cc:207-236 (host-side list utility)
extern "C" size_t ffs_filter_spots(ffs_spot* spots, size_t n, uint32_t min_spot_size, float max_sep,
                                   size_t* n_removed_size, size_t* n_removed_sep) {
    size_t kept = n, rs = 0, rp = 0;
    if (spots && min_spot_size > 0) {
        size_t o = 0;
        for (size_t i = 0; i < kept; ++i)
            if (!(spots[i].num_pixels < min_spot_size)) spots[o++] = spots[i];
        rs = kept - o;
        kept = o;
    }
    if (spots && max_sep > 0) {
        size_t o = 0;
        for (size_t i = 0; i < kept; ++i)
            if (!(spots[i].peak_centroid_distance > max_sep)) spots[o++] = spots[i];
        rp = kept - o;
        kept = o;
    }
    if (n_removed_size) *n_removed_size = rs;
    if (n_removed_sep) *n_removed_sep = rp;
    return kept;
}
