// api_ext.inl -- dispersion_extended launch sequence (see extended.cuh).
static int extended_setup(ffs_ctx*) { return FFS_OK; }

template <typename PixelT>
static int launch_extended(ffs_ctx* c, int first_slot, int n, cudaStream_t st) {
    // threshold_impl has already zeroed the strong planes, strong counters and candidate counters
    FFS_CK(c, cudaMemsetAsync(c->d_ext1 + (size_t)first_slot * c->bits_stride, 0, c->bits_stride * n, st));
    const dim3 grid((unsigned)(c->sm_count * 8), (unsigned)n);
    // pass 1: D = mask && m >= min_count && a > c, every pixel resolved exactly
    ThreshArgs a{};
    fill_thresh_args(c, a, first_slot, n);
    a.strong_bits = reinterpret_cast<uint32_t*>(c->d_ext1);
    a.strong_count = c->d_ext_count;
    a.force_brute = 1;
    resolve_kernel<PixelT, 3, true><<<grid, 256, 0, st>>>(a);
    // erosion -> B plane (validity of the second pass) + list of core pixels
    ErodeArgs e{};
    e.width = (int)c->p.width;
    e.height = (int)c->p.height;
    e.n_frames = n;
    e.slot0 = first_slot;
    e.words_per_row = (int)(c->bits_pitch / 4);
    e.mask_bits = reinterpret_cast<const uint32_t*>(c->d_valid_bits);
    e.d_bits = reinterpret_cast<const uint32_t*>(c->d_ext1);
    e.b_bits = reinterpret_cast<uint32_t*>(c->d_ext2);
    e.bits_stride = c->bits_stride;
    e.cand = c->d_cand;
    e.cand_count = c->d_cand_count;
    e.cand_cap = c->cand_cap;
    erode_kernel<<<dim3((unsigned)(c->sm_count * 2), (unsigned)n), 256, 0, st>>>(e);
    // final pass over the core pixels: 11x11 window over B-valid pixels
    ThreshArgs f{};
    fill_thresh_args(c, f, first_slot, n);
    f.valid_bits = c->d_ext2;
    f.valid_frame_stride = c->bits_stride;
    final_kernel<PixelT, 5><<<dim3((unsigned)(c->sm_count * 2), (unsigned)n), 256, 0, st>>>(f, c->d_valid_bits);
    c->launches += 3;
    FFS_CK(c, cudaGetLastError());
    return FFS_OK;
}

static int extended_impl(ffs_ctx* c, int first_slot, int n, cudaStream_t st) {
    if (c->p.bytes_per_pixel == 2) return launch_extended<uint16_t>(c, first_slot, n, st);
    return launch_extended<uint32_t>(c, first_slot, n, st);
}
