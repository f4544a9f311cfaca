// api_selfcheck.inl -- the executable's --validate (spotfinder.cc:1010-1052).
//
// The reference re-runs its CPU DIALS code on the frame and compares the strong pixels.  This
// library holds no CPU implementation (the product must fail rather than fall back), so the
// second opinion comes from the device: every pixel of the frame is decided again by the exact
// per-pixel routine (resolve_kernel in brute mode: direct clipped window sums, fp64 test in the
// reference's operation order -- no streaming sums, no bounds, no pruning, no candidate queues),
// into a plane of its own, and the two bit-planes are compared.  It checks the machinery of the
// fused kernel, not the arithmetic of the decision (that is what tests/ and the oracle are for).
// Standard dispersion only, like the reference's --validate.

static __global__ void plane_diff_kernel(const uint32_t* __restrict__ a, const uint32_t* __restrict__ b, size_t n_words,
                                         unsigned* __restrict__ diff) {
    unsigned d = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += (size_t)gridDim.x * blockDim.x)
        d += __popc(a[i] ^ b[i]);
    for (int off = 16; off > 0; off >>= 1) d += __shfl_down_sync(0xffffffffu, d, off);
    if ((threadIdx.x & 31) == 0 && d) atomicAdd(diff, d);
}

extern "C" int ffs_self_check(ffs_ctx* c, int on) {
    if (!c) return FFS_ERR_ARG;
    if (on && c->p.algorithm != FFS_ALGO_DISPERSION) {
        c->err = "ffs_self_check: standard dispersion only";
        return FFS_ERR_ARG;
    }
    FFS_CK(c, cudaSetDevice(c->p.device));
    if (on && !c->d_check) {
        c->check_bytes = round_up(c->bits_stride + 4 * sizeof(unsigned), 256);
        FFS_CK(c, cudaMalloc(&c->d_check, c->check_bytes));
        FFS_CK(c, cudaStreamCreateWithFlags(&c->check_stream, cudaStreamNonBlocking));
    }
    c->self_check = on != 0;
    return FFS_OK;
}

extern "C" int ffs_last_self_check(ffs_ctx* c, uint32_t* n_strong_exact, uint32_t* n_mismatch) {
    if (!c || !c->self_check) return FFS_ERR_ARG;
    if (n_strong_exact) *n_strong_exact = c->last_check[0];
    if (n_mismatch) *n_mismatch = c->last_check[1];
    return FFS_OK;
}

// Called by ffs_wait while it still owns the slot (the frame and its strong plane are in place).
template <typename PixelT>
static int self_check_slot_t(ffs_ctx* c, int slot) {
    cudaStream_t st = c->check_stream;
    FFS_CK(c, cudaMemsetAsync(c->d_check, 0, c->check_bytes, st));
    unsigned* counters = reinterpret_cast<unsigned*>(c->d_check + c->bits_stride);  // {strong, candidates (0), diff}
    ThreshArgs a{};
    fill_thresh_args(c, a, slot, 1);
    a.force_brute = 1;
    a.strong_bits = reinterpret_cast<uint32_t*>(c->d_check);
    a.strong_frame_stride = 0;  // one plane, whatever the slot number
    a.zstride_words = 0;
    a.strong_count = counters;
    a.cand_count = counters + 1;
    a.chunk_count = nullptr;
    a.bad_count = nullptr;
    resolve_kernel<PixelT, 3, false><<<dim3((unsigned)(c->sm_count * 4), 1u), 256, 0, st>>>(a);
    c->launches++;
    FFS_CK(c, cudaGetLastError());
    const size_t n_words = c->bits_stride / 4;
    plane_diff_kernel<<<(unsigned)c->sm_count * 2, 256, 0, st>>>(
        reinterpret_cast<const uint32_t*>(c->d_check),
        reinterpret_cast<const uint32_t*>(c->d_strong + (size_t)slot * c->zstride), n_words, counters + 2);
    c->launches++;
    FFS_CK(c, cudaGetLastError());
    unsigned h[3] = {0, 0, 0};
    FFS_CK(c, cudaMemcpyAsync(h, counters, sizeof h, cudaMemcpyDeviceToHost, st));
    FFS_CK(c, cudaStreamSynchronize(st));
    c->last_check[0] = h[0];
    c->last_check[1] = h[2];
    return FFS_OK;
}

static int self_check_slot(ffs_ctx* c, int slot) {
    return c->p.bytes_per_pixel == 2 ? self_check_slot_t<uint16_t>(c, slot) : self_check_slot_t<uint32_t>(c, slot);
}
