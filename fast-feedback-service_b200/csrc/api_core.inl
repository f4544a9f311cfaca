// api_core.inl -- lifetime, mask, slot upload, threshold launch, parity hooks.

static constexpr int kThreshStages = 3;  // TMA stages per warp
static constexpr int kThreshSR = 4;      // rows per stage
static constexpr int kThreshWarps = 4;   // warps (independent tasks) per CTA

extern "C" void ffs_default_params(ffs_params* p) {
    if (!p) return;
    memset(p, 0, sizeof(*p));
    p->bytes_per_pixel = 2;
    p->algorithm = FFS_ALGO_DISPERSION;
    p->kernel_half = 3;
    p->min_count = 2;
    p->nsig_b = 6.0;
    p->nsig_s = 3.0;
    p->threshold = 0.0;
    p->min_spot_size = 3;
    p->min_spot_size_3d = 3;
    p->max_peak_centroid_separation = 2.0f;
    p->cc_row_wrap = 1;
    p->device = 0;
    p->n_slots = 4;
}

extern "C" const char* ffs_version(void) { return "ffs_b200 0.1.0 (sm_100a)"; }

extern "C" const char* ffs_last_error(const ffs_ctx* ctx) {
    return ctx ? ctx->err.c_str() : g_create_error.c_str();
}

extern "C" int ffs_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

extern "C" int ffs_device_description(int device, char* buf, size_t buflen) {
    cudaDeviceProp prop;
    if (!buf || cudaGetDeviceProperties(&prop, device) != cudaSuccess) return FFS_ERR_ARG;
    snprintf(buf, buflen, "%s (CUDA %d.%d)", prop.name, prop.major, prop.minor);
    return FFS_OK;
}

static int repack_mask(ffs_ctx* c) {
    const long long words = (long long)(c->bits_pitch / 4) * c->p.height;
    FFS_CK(c, cudaMemsetAsync(c->d_valid_count, 0, sizeof(unsigned long long), c->stream));
    pack_mask_kernel<<<(unsigned)((words + 255) / 256), 256, 0, c->stream>>>(
      c->d_mask, (int)c->p.width, (int)c->p.height, reinterpret_cast<uint32_t*>(c->d_valid_bits),
      c->bits_pitch, c->d_valid_count);
    c->launches++;
    {
        const long long groups = (long long)((c->p.width + 7) / 8) * c->p.height;
        group_info_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, c->stream>>>(
          c->d_valid_bits, c->bits_pitch, (int)c->p.width, (int)c->p.height, c->d_info, c->info_pitch,
          c->d_mcount, c->m_pitch, c->d_info4, c->info4_pitch);
        c->launches++;
    }
    unsigned long long n = 0;
    FFS_CK(c, cudaMemcpyAsync(&n, c->d_valid_count, sizeof(n), cudaMemcpyDeviceToHost, c->stream));
    FFS_CK(c, cudaStreamSynchronize(c->stream));
    c->valid_count = n;
    return FFS_OK;
}

static int create_impl(ffs_ctx* c, const uint8_t* mask) {
    const ffs_params& p = c->p;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        c->err = "no CUDA device: libffs_b200 has no CPU fallback";
        return FFS_ERR_NO_DEVICE;
    }
    FFS_CK(c, cudaSetDevice(p.device));
    cudaDeviceProp prop;
    FFS_CK(c, cudaGetDeviceProperties(&prop, p.device));
    if (prop.major != 10) {
        c->err = std::string("device ") + prop.name + " is not sm_100: kernels are built for sm_100a only";
        return FFS_ERR_NO_DEVICE;
    }
    c->sm_count = prop.multiProcessorCount;
    const size_t px = (size_t)p.width * p.height;
    // TMA needs 16-byte-aligned row strides.  A row that already is (32-bit Eiger2 16M, Jungfrau)
    // keeps the dense layout, so a frame arrives with one contiguous copy and nothing else;
    // otherwise rows are padded (to 128 bytes) and the streaming path re-pitches on the device.
    {
        const size_t row = (size_t)p.width * p.bytes_per_pixel;
        c->pitch = (row % 16 == 0 && !env_int("FFS_PITCH128", 0)) ? row : round_up(row, 128);
    }
    c->frame_stride = c->pitch * p.height;
    c->bits_pitch = round_up((p.width + 7) / 8, 16);
    c->bits_stride = round_up(c->bits_pitch * p.height, 256);
    const int window = (2 * p.kernel_half + 1) * (2 * p.kernel_half + 1);
    c->eff_min_count = p.min_count <= 0 ? window : p.min_count;  // standalone.cc:60-64

    FFS_CK(c, cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    const size_t S = p.n_slots;
    FFS_CK(c, cudaMalloc(&c->d_mask, px));
    FFS_CK(c, cudaMalloc(&c->d_valid_bits, c->bits_stride));
    c->info_pitch = round_up(2 * (size_t)((p.width + 7) / 8), 16);
    FFS_CK(c, cudaMalloc(&c->d_info, c->info_pitch * p.height));
    FFS_CK(c, cudaMemset(c->d_info, 0, c->info_pitch * p.height));
    c->info4_pitch = round_up(4 * (size_t)((p.width + 7) / 8), 16);
    FFS_CK(c, cudaMalloc(&c->d_info4, c->info4_pitch * p.height));
    FFS_CK(c, cudaMemset(c->d_info4, 0, c->info4_pitch * p.height));
    c->m_pitch = round_up(8 * (size_t)((p.width + 7) / 8), 16);
    FFS_CK(c, cudaMalloc(&c->d_mcount, c->m_pitch * p.height));
    FFS_CK(c, cudaMemset(c->d_mcount, 0, c->m_pitch * p.height));
    FFS_CK(c, cudaMalloc(&c->d_frames, c->frame_stride * S));
    {
        const int n_words = (int)(c->bits_pitch / 4) * (int)p.height;
        c->n_chunks = (n_words + kChunkWords - 1) / kChunkWords;
        const size_t chunk_bytes = round_up(sizeof(uint32_t) * c->n_chunks, 16);
        c->zstride = round_up(c->bits_stride + chunk_bytes + 4 * sizeof(unsigned), 256);
        FFS_CK(c, cudaMalloc(&c->d_zero, c->zstride * S));
        FFS_CK(c, cudaMemset(c->d_zero, 0, c->zstride * S));
        c->d_strong = c->d_zero;
        c->d_chunk_count = reinterpret_cast<uint32_t*>(c->d_zero + c->bits_stride);
        c->d_strong_count = reinterpret_cast<unsigned*>(c->d_zero + c->bits_stride + chunk_bytes);
        c->d_cand_count = c->d_strong_count + 1;
        c->d_ext_count = c->d_strong_count + 2;
        c->d_bad_count = c->d_strong_count + 3;
    }
    if (p.trusted_max > 0 && (uint64_t)p.trusted_max < (p.bytes_per_pixel == 2 ? 0xffffull : 0xffffffffull)) {
        c->trusted_on = true;
        c->tmax = (uint32_t)p.trusted_max;
        c->vlim = std::min<uint32_t>(c->tmax, (1u << 24) - 1u);
        FFS_CK(c, cudaMalloc(&c->d_bad, sizeof(uint32_t) * (size_t)c->bad_cap * S));
    }
    c->cand_cap = env_int("FFS_CAND_CAP", (int)std::max<size_t>(px / 8, 4096));
    FFS_CK(c, cudaMalloc(&c->d_cand, sizeof(uint32_t) * (size_t)c->cand_cap * S));
    FFS_CK(c, cudaMalloc(&c->d_valid_count, sizeof(unsigned long long)));
    FFS_CK(c, cudaMemset(c->d_frames, 0, c->frame_stride * S));
    if (p.algorithm == FFS_ALGO_DISPERSION_EXTENDED) {
        FFS_CK(c, cudaMalloc(&c->d_ext1, c->bits_stride * S));
        FFS_CK(c, cudaMalloc(&c->d_ext2, c->bits_stride * S));
        FFS_CK(c, cudaMemset(c->d_ext1, 0, c->bits_stride * S));
        FFS_CK(c, cudaMemset(c->d_ext2, 0, c->bits_stride * S));
    }
    if (mask)
        FFS_CK(c, cudaMemcpy(c->d_mask, mask, px, cudaMemcpyHostToDevice));
    else
        FFS_CK(c, cudaMemset(c->d_mask, 1, px));
    int rc = repack_mask(c);
    if (rc) return rc;

    // TMA descriptors: image ring (x, y, slot) and validity bytes (x, y)
    {
        const bool p16 = p.bytes_per_pixel == 2;
        cuuint64_t dims[3] = {p.width, p.height, (cuuint64_t)S};
        cuuint64_t strides[2] = {c->pitch, c->frame_stride};
        cuuint32_t box[3] = {p16 ? 256u : 128u, (cuuint32_t)kThreshSR, 1};
        rc = make_tensor_map(c, &c->tm_img, p16 ? CU_TENSOR_MAP_DATA_TYPE_UINT16 : CU_TENSOR_MAP_DATA_TYPE_UINT32,
                             3, c->d_frames, dims, strides, box);
        if (rc) return rc;
        cuuint64_t vdims[3] = {(cuuint64_t)(p.width + 7) / 8, p.height, 1};
        cuuint64_t vstrides[2] = {c->bits_pitch, c->bits_stride};
        cuuint32_t vbox[3] = {p16 ? (cuuint32_t)ThreshGeo<uint16_t, 3, 8, kThreshSR>::VROW_BYTES
                                  : (cuuint32_t)ThreshGeo<uint32_t, 3, 4, kThreshSR>::VROW_BYTES,
                              (cuuint32_t)kThreshSR, 1};
        rc = make_tensor_map(c, &c->tm_valid, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, c->d_valid_bits,
                             vdims, vstrides, vbox);
        if (rc) return rc;
        cuuint32_t box8[3] = {256u, (cuuint32_t)FusedGeo<uint16_t>::SR, 1};
        rc = make_tensor_map(c, &c->tm_img8, p16 ? CU_TENSOR_MAP_DATA_TYPE_UINT16 : CU_TENSOR_MAP_DATA_TYPE_UINT32,
                             3, c->d_frames, dims, strides, box8);
        if (rc) return rc;
        cuuint64_t idims[2] = {2 * (cuuint64_t)((p.width + 7) / 8), p.height};
        cuuint64_t istrides[1] = {c->info_pitch};
        cuuint32_t ibox[2] = {(cuuint32_t)FusedGeo<uint16_t>::INFO_ROW_BYTES, (cuuint32_t)FusedGeo<uint16_t>::SR};
        rc = make_tensor_map(c, &c->tm_info, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, c->d_info, idims, istrides, ibox);
        if (rc) return rc;
        cuuint64_t i4dims[2] = {4 * (cuuint64_t)((p.width + 7) / 8), p.height};
        cuuint64_t i4strides[1] = {c->info4_pitch};
        cuuint32_t i4box[2] = {(cuuint32_t)ExtGeo<uint16_t>::INFO4_ROW_BYTES, 4};
        rc = make_tensor_map(c, &c->tm_info4, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, c->d_info4, i4dims, i4strides, i4box);
        if (rc) return rc;
        c->have_tm = true;
    }
    rc = extended_setup(c);
    if (rc) return rc;
    rc = cc_setup(c);
    if (rc) return rc;
    return stream_setup(c);
}

extern "C" int ffs_create(const ffs_params* p, const uint8_t* mask, ffs_ctx** out) {
    if (!p || !out) return FFS_ERR_ARG;
    *out = nullptr;
    if (p->width == 0 || p->height == 0 || p->width > (uint32_t)kMaxDim ||
        p->height > (uint32_t)kMaxDim || (p->bytes_per_pixel != 2 && p->bytes_per_pixel != 4) ||
        p->kernel_half != 3 || p->n_slots < 1 || p->nsig_b < 0 || p->nsig_s < 0 ||
        p->threshold < 0 ||
        (p->algorithm != FFS_ALGO_DISPERSION && p->algorithm != FFS_ALGO_DISPERSION_EXTENDED)) {
        g_create_error = "ffs_create: invalid parameters (kernel_half must be 3, dimensions <= 8192)";
        return FFS_ERR_ARG;
    }
    ffs_ctx* c = new (std::nothrow) ffs_ctx();
    if (!c) return FFS_ERR_NOMEM;
    c->p = *p;
    int rc = create_impl(c, mask);
    if (rc) {
        g_create_error = c->err;
        ffs_destroy(c);
        return rc;
    }
    *out = c;
    return FFS_OK;
}

extern "C" void ffs_destroy(ffs_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->p.device);
    cudaDeviceSynchronize();
    stream_teardown(c);
    bslz4_teardown(c);
    cc_teardown(c);
    cudaFree(c->d_mask);
    cudaFree(c->d_valid_bits);
    cudaFree(c->d_info);
    cudaFree(c->d_mcount);
    cudaFree(c->d_info4);
    cudaFree(c->d_gain);
    cudaFree(c->d_frames);
    cudaFree(c->d_zero);
    cudaFree(c->d_check);
    if (c->check_stream) cudaStreamDestroy(c->check_stream);
    cudaFree(c->d_ext1);
    cudaFree(c->d_ext2);
    cudaFree(c->d_cand);
    cudaFree(c->d_bad);
    cudaFree(c->d_valid_count);
    for (auto& ev : c->prof_events) {
        cudaEventDestroy(ev.first);
        cudaEventDestroy(ev.second);
    }
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

extern "C" int ffs_profile_enable(ffs_ctx* c, int on) {
    if (!c) return FFS_ERR_ARG;
    std::lock_guard<std::mutex> lk(c->mu);
    c->prof_on = on != 0;
    return FFS_OK;
}

extern "C" int ffs_profile_read(ffs_ctx* c, double* total_ms, uint64_t* n_launches) {
    if (!c) return FFS_ERR_ARG;
    std::lock_guard<std::mutex> lk(c->mu);
    double tot = 0;
    for (size_t i = 0; i < c->prof_used; ++i) {
        float ms = 0;
        FFS_CK(c, cudaEventSynchronize(c->prof_events[i].second));
        FFS_CK(c, cudaEventElapsedTime(&ms, c->prof_events[i].first, c->prof_events[i].second));
        tot += ms;
    }
    if (total_ms) *total_ms = tot;
    if (n_launches) *n_launches = c->prof_used;
    c->prof_used = 0;
    return FFS_OK;
}

extern "C" uint64_t ffs_kernel_launches(const ffs_ctx* c) { return c ? c->launches.load() : 0; }
extern "C" size_t ffs_frame_pitch(const ffs_ctx* c) { return c ? c->pitch : 0; }
extern "C" size_t ffs_frame_bytes_device(const ffs_ctx* c) { return c ? c->frame_stride : 0; }
extern "C" uint64_t ffs_mask_valid_count(ffs_ctx* c) { return c ? c->valid_count : 0; }
extern "C" void ffs_free(void* p) { free(p); }

// threshold_w_gain: a per-pixel gain map for the standard dispersion test (baseline.cpp:289-310)
// and for dispersion_extended (baseline.cpp:772-805).
extern "C" int ffs_set_gain(ffs_ctx* c, const double* gain) {
    if (!c) return FFS_ERR_ARG;
    FFS_CK(c, cudaSetDevice(c->p.device));
    FFS_CK(c, cudaDeviceSynchronize());
    std::lock_guard<std::mutex> lk(c->mu);
    for (auto& s : c->slots)  // the slot graphs captured the old launch parameters
        for (auto& g : s.graph)
            if (g) {
                cudaGraphExecDestroy(g);
                g = nullptr;
            }
    if (!gain) {
        cudaFree(c->d_gain);
        c->d_gain = nullptr;
        c->gain_min = c->gain_max = 1.0;
        c->gain_brute = false;
        return FFS_OK;
    }
    const size_t n = (size_t)c->p.width * c->p.height;
    std::vector<uint8_t> mask(n);
    FFS_CK(c, cudaMemcpy(mask.data(), c->d_mask, n, cudaMemcpyDeviceToHost));
    double gmin = INFINITY, gmax = 0.0;
    bool bad = false;
    for (size_t k = 0; k < n; ++k) {
        if (!mask[k]) continue;
        const double g = gain[k];
        if (!(g > 0.0) || !std::isfinite(g)) bad = true;
        else {
            gmin = std::min(gmin, g);
            gmax = std::max(gmax, g);
        }
    }
    if (!c->d_gain) FFS_CK(c, cudaMalloc(&c->d_gain, n * sizeof(double)));
    FFS_CK(c, cudaMemcpy(c->d_gain, gain, n * sizeof(double), cudaMemcpyHostToDevice));
    c->gain_min = std::isfinite(gmin) ? gmin : 1.0;
    c->gain_max = std::isfinite(gmin) ? gmax : 1.0;
    c->gain_brute = bad;  // sqrt of a non-positive gain: the reference yields NaN there; decide every pixel exactly
    return FFS_OK;
}

extern "C" int ffs_apply_resolution_mask(ffs_ctx* c, float wavelength, float distance_m, float bcx,
                                         float bcy, float psx, float psy, float dmin, float dmax) {
    if (!c) return FFS_ERR_ARG;
    FFS_CK(c, cudaSetDevice(c->p.device));
    dim3 block(32, 32), grid((c->p.width + 31) / 32, (c->p.height + 31) / 32);  // spotfinder.cc:119-121
    resolution_mask_kernel<<<grid, block, 0, c->stream>>>(c->d_mask, (int)c->p.width, (int)c->p.height,
                                                          wavelength, distance_m, bcx, bcy, psx, psy,
                                                          dmin, dmax);
    c->launches++;
    FFS_CK(c, cudaGetLastError());
    return repack_mask(c);
}

extern "C" int ffs_get_mask(ffs_ctx* c, uint8_t* out) {
    if (!c || !out) return FFS_ERR_ARG;
    FFS_CK(c, cudaSetDevice(c->p.device));
    FFS_CK(c, cudaMemcpy(out, c->d_mask, (size_t)c->p.width * c->p.height, cudaMemcpyDeviceToHost));
    return FFS_OK;
}

static bool slots_ok(const ffs_ctx* c, int first, int n) {
    return c && first >= 0 && n >= 1 && first + n <= c->p.n_slots;
}

extern "C" int ffs_upload(ffs_ctx* c, int first_slot, int n, const void* host_frames) {
    if (!slots_ok(c, first_slot, n) || !host_frames) return FFS_ERR_ARG;
    FFS_CK(c, cudaSetDevice(c->p.device));
    const size_t row = (size_t)c->p.width * c->p.bytes_per_pixel;
    for (int i = 0; i < n; ++i) {
        FFS_CK(c, cudaMemcpy2DAsync(c->d_frames + (size_t)(first_slot + i) * c->frame_stride, c->pitch,
                                    (const uint8_t*)host_frames + (size_t)i * row * c->p.height, row,
                                    row, c->p.height, cudaMemcpyHostToDevice, c->stream));
    }
    FFS_CK(c, cudaStreamSynchronize(c->stream));
    return FFS_OK;
}

// Pick output rows per warp task: R + 2*KH a multiple of the stage depth (7), as large as
// possible (less halo re-reading) while still giving every SM's warp slots ~2 tasks.
static int pick_tile_rows(const ffs_ctx* c, int n_frames, int out_cols, int ks, int kh) {
    const int forced = env_int("FFS_TILE_ROWS", 0);
    if (forced > 0) return forced;
    const long long want = 2LL * c->sm_count * 12;
    const int n_strips = ((int)c->p.width + out_cols - 1) / out_cols;
    int best = ks * 8 - 2 * kh;
    for (int groups = 34; groups >= 8; groups -= 2) {
        const int R = groups * ks - 2 * kh;  // R + 2*kh a multiple of the stage depth
        const long long tasks = (long long)n_frames * n_strips * (((int)c->p.height + R - 1) / R);
        best = R;
        if (tasks >= want) break;
    }
    return best;
}

static void fill_thresh_args(ffs_ctx* c, ThreshArgs& a, int first_slot, int n) {
    a.width = (int)c->p.width;
    a.height = (int)c->p.height;
    a.n_frames = n;
    a.slot0 = first_slot;
    a.image = c->d_frames;
    a.pitch = c->pitch;
    a.frame_stride = c->frame_stride;
    a.valid_bits = c->d_valid_bits;
    a.valid_pitch = c->bits_pitch;
    a.valid_frame_stride = 0;
    a.strong_bits = reinterpret_cast<uint32_t*>(c->d_strong);
    a.strong_pitch = c->bits_pitch;
    a.strong_frame_stride = c->zstride;
    a.zstride_words = c->zstride / 4;
    a.strong_count = c->d_strong_count;
    a.cand = c->d_cand;
    a.cand_count = c->d_cand_count;
    a.cand_cap = c->cand_cap;
    a.force_brute = env_int("FFS_FORCE_BRUTE", 0) || c->gain_brute || (c->trusted_on && c->tmax < 16384u);
    a.nsig_b = c->p.nsig_b;
    a.nsig_s = c->p.nsig_s;
    a.threshold = c->p.threshold;
    a.min_count = c->eff_min_count;
    a.prefilter_k = (float)(c->p.nsig_s * c->p.nsig_s * (1.0 - 1e-4));
    // with a gain map b > d reads v > mu + nsig_s sqrt(g) sqrt(mu): the smallest gain keeps the
    // pruning bound a lower bound
    a.sig_k = (float)(c->p.nsig_s * std::sqrt(c->d_gain ? c->gain_min : 1.0) * (1.0 - 1e-4));
    a.gain = c->d_gain;
    // float brackets rounded outwards (the kernel's own error bound covers the table rounding)
    a.g_lo = c->d_gain ? (float)(c->gain_min * (1.0 - 1e-6)) : 1.0f;
    a.g_hi = c->d_gain ? (float)(c->gain_max * (1.0 + 1e-6)) : 1.0f;
    a.thr_floor = (int)std::min(std::floor(c->p.threshold), 1073741824.0);
    a.mcount = c->d_mcount;
    a.m_pitch = c->m_pitch;
    a.chunk_count = c->d_chunk_count;  // zeroed by threshold_impl; consumed by cc_scan_kernel
    a.n_chunks = c->n_chunks;
    a.tmax = c->tmax;
    a.vlim = c->vlim;
    a.bad = c->d_bad;
    a.bad_count = c->trusted_on ? c->d_bad_count : nullptr;
    a.bad_cap = c->bad_cap;
}

// The streaming kernels list a pixel above the trusted range because its clamped value makes
// its group / window a candidate; that needs trusted_max to sit above every clamp and group
// threshold (9362 + 3 sqrt(9362)).  A smaller trusted_max (no real detector has one) takes the
// exact per-pixel path.
static bool trusted_needs_brute(const ffs_ctx* c) { return c->trusted_on && c->tmax < 16384u; }

template <typename PixelT>
static int launch_trusted_fixup(ffs_ctx* c, int first_slot, int n, cudaStream_t st) {
    ThreshArgs a{};
    fill_thresh_args(c, a, first_slot, n);
    trusted_fixup_kernel<PixelT, 3><<<dim3(32, (unsigned)n), 256, 0, st>>>(a);
    c->launches++;
    FFS_CK(c, cudaGetLastError());
    return FFS_OK;
}

// ffs_profile_*: the next event pair that brackets the threshold kernel(s) of one launch
static void prof_pair(ffs_ctx* c, cudaEvent_t& e0, cudaEvent_t& e1) {
    if (!c->prof_on) return;
    std::lock_guard<std::mutex> lk(c->mu);
    if (c->prof_used == c->prof_events.size()) {
        std::pair<cudaEvent_t, cudaEvent_t> ev;
        cudaEventCreate(&ev.first);
        cudaEventCreate(&ev.second);
        c->prof_events.push_back(ev);
    }
    e0 = c->prof_events[c->prof_used].first;
    e1 = c->prof_events[c->prof_used].second;
    c->prof_used++;
}

template <typename PixelT, int PPL>
static int launch_dispersion(ffs_ctx* c, int first_slot, int n, cudaStream_t st) {
    using G = ThreshGeo<PixelT, 3, PPL, kThreshSR>;
    auto kern = prefilter_kernel<PixelT, 3, PPL, kThreshSR, kThreshStages, kThreshWarps>;
    const int smem = kThreshWarps * G::smem_per_warp(kThreshStages) + 128;
    static bool attr_done[64] = {};
    if (!attr_done[c->p.device & 63]) {
        FFS_CK(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_done[c->p.device & 63] = true;
    }
    ThreshArgs a{};
    fill_thresh_args(c, a, first_slot, n);
    a.tile_rows = pick_tile_rows(c, n, G::OUT_COLS, kThreshSR, 3);
    a.n_strips = (a.width + G::OUT_COLS - 1) / G::OUT_COLS;
    a.n_row_tiles = (a.height + a.tile_rows - 1) / a.tile_rows;
    if (c->trusted_on) a.force_brute = 1;  // the round-1 prefilter knows no trusted range
    if (!a.force_brute) {
        const long long tasks = (long long)n * a.n_strips * a.n_row_tiles;
        const unsigned grid = (unsigned)((tasks + kThreshWarps - 1) / kThreshWarps);
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        prof_pair(c, e0, e1);
        if (e0) cudaEventRecord(e0, st);
        kern<<<grid, kThreshWarps * 32, smem, st>>>(c->tm_img, c->tm_valid, a);
        if (e1) cudaEventRecord(e1, st);
        c->launches++;
        FFS_CK(c, cudaGetLastError());
    }
    dim3 rgrid((unsigned)(c->sm_count * 4), (unsigned)n);
    resolve_kernel<PixelT, 3, false><<<rgrid, 256, 0, st>>>(a);
    c->launches++;
    FFS_CK(c, cudaGetLastError());
    return FFS_OK;
}

// Rows per warp task of the streaming kernels (R + 6 a multiple of the 4-row stage depth).
// Start from "as large as possible while every resident warp still gets ~2 tasks"; then, because
// the last, partly filled wave of CTAs is idle time, prefer the tile height whose wave count sits
// just below an integer, weighted by the 6 halo rows a tile re-reads (measured on a 32-frame
// Eiger2 16M batch, tools/tile_sweep.sh: 12 row tiles = 2.92 waves 23.4 us/frame, 18 tiles = 4.38
// waves 24.2).  A single frame -- the streaming path -- was ~1.3 waves under the first rule and
// its second, nearly empty wave doubled the kernel time; one full wave of taller tiles does not.
static int pick_wave_tile_rows(const ffs_ctx* c, int n, int n_strips, int height, int warps_per_cta,
                               int ctas_per_sm, int resident_warps_per_sm) {
    const int forced = env_int("FFS_TILE_ROWS", 0);
    if (forced > 0) return std::min(forced, 500);
    const long long want = 2LL * c->sm_count * resident_warps_per_sm;
    int best = 4 * 8 - 6;
    for (int groups = 64; groups >= 8; groups -= 2) {
        best = groups * 4 - 6;
        const long long tasks = (long long)n * n_strips * ((height + best - 1) / best);
        if (tasks >= want) break;
    }
    const double slots = (double)c->sm_count * std::max(1, ctas_per_sm);
    auto waves_of = [&](int rows) {
        const long long t = (long long)n * n_strips * ((height + rows - 1) / rows);
        return (double)((t + warps_per_cta - 1) / warps_per_cta) / slots;
    };
    static const int one_wave = env_int("FFS_ONE_WAVE", 1);
    if (waves_of(best) > 1.5 || one_wave) {
        double best_score = 0.0;
        for (int groups = 129; groups >= 6; groups -= 1) {  // rows <= 510: queue entries hold 9-bit row offsets
            const int rows = groups * 4 - 6;
            const double w = waves_of(rows);
            if (w < (one_wave ? 0.85 : 1.5)) continue;
            const double score = w / std::ceil(w) * rows / (rows + 6.0);
            if (score > best_score + 1e-9) {
                best_score = score;
                best = rows;
            }
        }
        // same number of row tiles, split evenly (R + 6 still a multiple of the stage depth)
        const int tiles = (height + best - 1) / best;
        best = ((height + tiles - 1) / tiles + 6 + 3) / 4 * 4 - 6;
    }
    return best;
}

// The one-kernel path (dispersion_fused.cuh): TMA streaming + in-kernel exact resolution.
template <typename PixelT, int kWarps, int kStages>
static int launch_fused_w(ffs_ctx* c, int first_slot, int n, cudaStream_t st) {
    using G = FusedGeo<PixelT>;
    auto kern = fused_dispersion_kernel<PixelT, kStages, kWarps>;
    const int smem = kWarps * G::smem_per_warp(kStages) + 128;
    static bool attr_done[64] = {};
    if (!attr_done[c->p.device & 63]) {
        FFS_CK(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        // occupancy is bounded by shared memory: ask for the whole carve-out
        // (FFS_FUSED_CARVEOUT=<percent>: developer A/B -- fewer resident warps, more L1 for stage 3's gathers)
        static const int carve = env_int("FFS_FUSED_CARVEOUT", 0);
        FFS_CK(c, cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout,
                                       carve > 0 ? carve : (int)cudaSharedmemCarveoutMaxShared));
        attr_done[c->p.device & 63] = true;
    }
    ThreshArgs a{};
    fill_thresh_args(c, a, first_slot, n);
    a.n_strips = (a.width + G::OUT_COLS - 1) / G::OUT_COLS;
    static const int dbg = env_int("FFS_FUSED_DEBUG", 0);
    a.force_brute = dbg;  // 2: queue but do not resolve (timing experiments only)
    a.tile_rows = pick_wave_tile_rows(c, n, a.n_strips, a.height, kWarps, G::MIN_WARPS_PER_SM / kWarps,
                                      std::max(1, (int)(200 * 1024 / G::smem_per_warp(kStages))));
    a.n_row_tiles = (a.height + a.tile_rows - 1) / a.tile_rows;
    const long long tasks = (long long)n * a.n_strips * a.n_row_tiles;
    const unsigned grid = (unsigned)((tasks + kWarps - 1) / kWarps);
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    prof_pair(c, e0, e1);
    if (e0) cudaEventRecord(e0, st);
    kern<<<grid, kWarps * 32, smem, st>>>(c->tm_img8, c->tm_info, a);
    if (e1) cudaEventRecord(e1, st);
    c->launches++;
    FFS_CK(c, cudaGetLastError());
    return FFS_OK;
}

template <typename PixelT>
static int launch_fused(ffs_ctx* c, int first_slot, int n, cudaStream_t st) {
    static const int w = env_int("FFS_FUSED_WARPS", sizeof(PixelT) == 2 ? 4 : 3);  // warps (independent tasks) per CTA
    static const int stg = env_int("FFS_FUSED_STAGES", 2);
    if (stg == 2) {
        switch (w) {
            case 1: return launch_fused_w<PixelT, 1, 2>(c, first_slot, n, st);
            case 2: return launch_fused_w<PixelT, 2, 2>(c, first_slot, n, st);
            case 4: return launch_fused_w<PixelT, 4, 2>(c, first_slot, n, st);
            case 5: return launch_fused_w<PixelT, 5, 2>(c, first_slot, n, st);
            case 3: return launch_fused_w<PixelT, 3, 2>(c, first_slot, n, st);
            default: return launch_fused_w<PixelT, 2, 2>(c, first_slot, n, st);
        }
    }
    switch (w) {
        case 1: return launch_fused_w<PixelT, 1, 3>(c, first_slot, n, st);
        case 2: return launch_fused_w<PixelT, 2, 3>(c, first_slot, n, st);
        case 4: return launch_fused_w<PixelT, 4, 3>(c, first_slot, n, st);
        case 3: return launch_fused_w<PixelT, 3, 3>(c, first_slot, n, st);
        default: return launch_fused_w<PixelT, 2, 3>(c, first_slot, n, st);
    }
}

static int threshold_impl(ffs_ctx* c, int first_slot, int n, cudaStream_t st) {
    // strong planes, chunk counts and counters of the batch: one memset (they share a per-slot region)
    FFS_CK(c, cudaMemsetAsync(c->d_zero + (size_t)first_slot * c->zstride, 0, c->zstride * n, st));
    if (c->p.algorithm == FFS_ALGO_DISPERSION_EXTENDED) return extended_impl(c, first_slot, n, st);
    static const int two_kernel = env_int("FFS_TWO_KERNEL", 0);  // A/B: the round-1 prefilter + resolve pair
    if (two_kernel || env_int("FFS_FORCE_BRUTE", 0) || c->gain_brute || trusted_needs_brute(c)) {
        if (c->p.bytes_per_pixel == 2) return launch_dispersion<uint16_t, 8>(c, first_slot, n, st);
        return launch_dispersion<uint32_t, 4>(c, first_slot, n, st);
    }
    int rc = c->p.bytes_per_pixel == 2 ? launch_fused<uint16_t>(c, first_slot, n, st)
                                       : launch_fused<uint32_t>(c, first_slot, n, st);
    // 16-bit pixels above the trusted range entered the streaming sums clamped: repair their
    // neighbourhoods (32-bit pixels are left out of the sums by the front end itself)
    if (rc == FFS_OK && c->trusted_on && c->p.bytes_per_pixel == 2) rc = launch_trusted_fixup<uint16_t>(c, first_slot, n, st);
    return rc;
}

extern "C" int ffs_threshold(ffs_ctx* c, int first_slot, int n, void* stream) {
    if (!slots_ok(c, first_slot, n)) return FFS_ERR_ARG;
    FFS_CK(c, cudaSetDevice(c->p.device));
    return threshold_impl(c, first_slot, n, pick_stream(c, stream));
}

extern "C" int ffs_sync(ffs_ctx* c, void* stream) {
    if (!c) return FFS_ERR_ARG;
    FFS_CK(c, cudaStreamSynchronize(pick_stream(c, stream)));
    return FFS_OK;
}

static int get_bits_plane(ffs_ctx* c, const uint8_t* d_plane, uint8_t* out) {
    const size_t row = (c->p.width + 7) / 8;
    FFS_CK(c, cudaSetDevice(c->p.device));
    FFS_CK(c, cudaDeviceSynchronize());
    FFS_CK(c, cudaMemcpy2D(out, row, d_plane, c->bits_pitch, row, c->p.height, cudaMemcpyDeviceToHost));
    return FFS_OK;
}

static int get_u8_plane(ffs_ctx* c, const uint8_t* d_plane, uint8_t* out) {
    const size_t px = (size_t)c->p.width * c->p.height;
    FFS_CK(c, cudaSetDevice(c->p.device));
    FFS_CK(c, cudaDeviceSynchronize());
    uint8_t* tmp = nullptr;
    FFS_CK(c, cudaMalloc(&tmp, px));
    unpack_bits_kernel<<<(unsigned)((px + 255) / 256), 256, 0, c->stream>>>(
      reinterpret_cast<const uint32_t*>(d_plane), c->bits_pitch, (int)c->p.width, (int)c->p.height, tmp);
    c->launches++;
    cudaError_t e = cudaMemcpyAsync(out, tmp, px, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(tmp);
    FFS_CK(c, e);
    return FFS_OK;
}

extern "C" int ffs_get_strong_bits(ffs_ctx* c, int slot, uint8_t* out) {
    if (!slots_ok(c, slot, 1) || !out) return FFS_ERR_ARG;
    return get_bits_plane(c, c->d_strong + (size_t)slot * c->zstride, out);
}

extern "C" int ffs_get_strong_u8(ffs_ctx* c, int slot, uint8_t* out) {
    if (!slots_ok(c, slot, 1) || !out) return FFS_ERR_ARG;
    return get_u8_plane(c, c->d_strong + (size_t)slot * c->zstride, out);
}

extern "C" int ffs_get_extended_stage_u8(ffs_ctx* c, int slot, int stage, uint8_t* out) {
    if (!slots_ok(c, slot, 1) || !out || (stage != 1 && stage != 2) || !c->d_ext1) return FFS_ERR_ARG;
    return get_u8_plane(c, (stage == 1 ? c->d_ext1 : c->d_ext2) + (size_t)slot * c->bits_stride, out);
}
