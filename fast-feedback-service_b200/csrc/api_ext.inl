// api_ext.inl -- dispersion_extended launch sequence (see extended.cuh).
static int extended_setup(ffs_ctx*) { return FFS_OK; }

template <typename PixelT, int kWarps>
static int launch_ext_pass1(ffs_ctx* c, ThreshArgs& a, int n, cudaStream_t st) {
    using G = FusedGeo<PixelT>;
    using E = ExtGeo<PixelT>;
    constexpr int kStages = 2;
    auto kern = ext_pass1_kernel<PixelT, kStages, kWarps>;
    const int smem = kWarps * E::smem_per_warp(kStages) + 128;
    static bool attr_done[64] = {};
    if (!attr_done[c->p.device & 63]) {
        FFS_CK(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        FFS_CK(c, cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout,
                                       (int)cudaSharedmemCarveoutMaxShared));
        attr_done[c->p.device & 63] = true;
    }
    a.force_brute = 0;
    a.n_strips = (a.width + G::OUT_COLS - 1) / G::OUT_COLS;
    static int ctas_per_sm[64] = {};
    if (!ctas_per_sm[c->p.device & 63]) {
        int nb = 0;
        FFS_CK(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, kWarps * 32, smem));
        ctas_per_sm[c->p.device & 63] = std::max(1, nb);
    }
    a.tile_rows = pick_wave_tile_rows(c, n, a.n_strips, a.height, kWarps, ctas_per_sm[c->p.device & 63],
                                      std::max(1, (int)(200 * 1024 / E::smem_per_warp(kStages))));
    a.n_row_tiles = (a.height + a.tile_rows - 1) / a.tile_rows;
    const long long tasks = (long long)n * a.n_strips * a.n_row_tiles;
    kern<<<(unsigned)((tasks + kWarps - 1) / kWarps), kWarps * 32, smem, st>>>(c->tm_img8, c->tm_info4, a);
    FFS_CK(c, cudaGetLastError());
    return FFS_OK;
}

template <typename PixelT>
static int launch_extended(ffs_ctx* c, int first_slot, int n, cudaStream_t st) {
    // threshold_impl has already zeroed the strong planes, strong counters and candidate counters
    FFS_CK(c, cudaMemsetAsync(c->d_ext1 + (size_t)first_slot * c->bits_stride, 0, c->bits_stride * n, st));
    const dim3 grid((unsigned)(c->sm_count * 8), (unsigned)n);
    // pass 1: D = mask && m >= min_count && a > c, every pixel resolved exactly
    ThreshArgs a{};
    fill_thresh_args(c, a, first_slot, n);
    a.strong_bits = reinterpret_cast<uint32_t*>(c->d_ext1);
    a.strong_frame_stride = c->bits_stride;  // the D planes are a plain array of bit-planes
    a.strong_count = c->d_ext_count;
    a.chunk_count = nullptr;  // the D plane is not the connected-components input
    static const int brute = env_int("FFS_EXT_BRUTE", 0);  // A/B: resolve every pixel exactly
    cudaEvent_t e0 = nullptr, e1 = nullptr;  // ffs_profile_*: first pass + erosion + final pass together
    prof_pair(c, e0, e1);
    if (e0) cudaEventRecord(e0, st);
    if (brute || c->gain_brute || trusted_needs_brute(c)) {  // a gain map with non-positive entries cannot be bracketed
        a.force_brute = 1;
        resolve_kernel<PixelT, 3, true><<<grid, 256, 0, st>>>(a);
    } else {
        // streaming first pass (extended_stream.cuh).  Warps are independent tasks; their number per CTA
        // only decides how the per-warp shared memory (11 KB) packs into an SM: 3 per CTA -> 6 CTAs =
        // 18 resident warps.  Measured (Eiger2 16M, 32 frames): 3 -> 55.9 us/frame, 4 (16 warps/SM) -> 63.4,
        // 10 (20 warps/SM) -> 63.1: a CTA's shared memory is released only when its slowest warp ends,
        // and the kernel is bound by ALU-pipe issue, not by latency, so more resident warps do not pay.
        static const int ext_warps = env_int("FFS_EXT_WARPS", 3);
        int rc = ext_warps == 3   ? launch_ext_pass1<PixelT, 3>(c, a, n, st)
                 : ext_warps == 4 ? launch_ext_pass1<PixelT, 4>(c, a, n, st)
                                  : launch_ext_pass1<PixelT, 10>(c, a, n, st);
        if (rc) return rc;
    }
    // erosion -> B plane (validity of the second pass) + list of core pixels
    ErodeArgs e{};
    e.width = (int)c->p.width;
    e.height = (int)c->p.height;
    e.n_frames = n;
    e.slot0 = first_slot;
    e.words_per_row = (int)(c->bits_pitch / 4);
    // row = t / words_per_row as one multiply-high; exact while t * words_per_row < 2^32 (0: divide)
    e.inv_wpr = ((uint64_t)e.words_per_row * e.words_per_row * (uint64_t)c->p.height < (1ull << 32))
                  ? (uint32_t)((1ull << 32) / (uint32_t)e.words_per_row) + 1u : 0u;
    e.mask_bits = reinterpret_cast<const uint32_t*>(c->d_valid_bits);
    e.d_bits = reinterpret_cast<const uint32_t*>(c->d_ext1);
    e.b_bits = reinterpret_cast<uint32_t*>(c->d_ext2);
    e.bits_stride = c->bits_stride;
    e.cand = c->d_cand;
    e.cand_count = c->d_cand_count;
    e.cand_cap = c->cand_cap;
    e.zstride_words = c->zstride / 4;
    erode_kernel<<<dim3((unsigned)(c->sm_count * 2), (unsigned)n), 256, 0, st>>>(e);
    // final pass over the core pixels: 11x11 window over B-valid pixels
    ThreshArgs f{};
    fill_thresh_args(c, f, first_slot, n);
    if (c->trusted_on) {  // pixels above the trusted range are not valid background (trusted.cuh)
        trusted_clear_kernel<PixelT><<<dim3(8, (unsigned)n), 256, 0, st>>>(f, reinterpret_cast<uint32_t*>(c->d_ext2),
                                                                           c->bits_stride);
        c->launches++;
    }
    f.valid_bits = c->d_ext2;
    f.valid_frame_stride = c->bits_stride;
    final_kernel<PixelT, 5><<<dim3((unsigned)(c->sm_count * 2), (unsigned)n), 256, 0, st>>>(f, c->d_valid_bits);
    if (e1) cudaEventRecord(e1, st);
    c->launches += 3;
    FFS_CK(c, cudaGetLastError());
    return FFS_OK;
}

static int extended_impl(ffs_ctx* c, int first_slot, int n, cudaStream_t st) {
    if (c->p.bytes_per_pixel == 2) return launch_extended<uint16_t>(c, first_slot, n, st);
    return launch_extended<uint32_t>(c, first_slot, n, st);
}
