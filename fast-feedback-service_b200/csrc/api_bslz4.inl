// api_bslz4.inl -- compressed ingestion: upload the raw bitshuffle-LZ4 chunk, decode on the
// device (bslz4_gpu.cuh), continue with the ordinary per-frame pipeline.

static inline uint32_t be32h(const uint8_t* p) {
    return (uint32_t)p[0] << 24 | (uint32_t)p[1] << 16 | (uint32_t)p[2] << 8 | p[3];
}

// the slot's chunk in the pinned host buffer / in device memory (its table ends right in front of it)
static inline uint8_t* bslz4_h_chunk(ffs_ctx* c, int slot) { return c->h_chunks + (size_t)slot * c->chunk_stride + c->table_cap; }
static inline uint8_t* bslz4_d_chunk(ffs_ctx* c, int slot) { return c->d_chunks + (size_t)slot * c->chunk_stride + c->table_cap; }
// offset, from the chunk's start backwards, at which the upload begins (16-byte aligned)
static inline size_t bslz4_table_bytes(uint32_t n_blocks) {
    return round_up(sizeof(Bslz4Dyn) + (sizeof(uint2) + sizeof(uint32_t)) * (size_t)n_blocks, 16);
}
// one copy: block table + chunk
static inline cudaError_t bslz4_upload(ffs_ctx* c, int slot, uint32_t n_blocks, size_t chunk_bytes, cudaStream_t st) {
    const size_t tb = bslz4_table_bytes(n_blocks);
    return cudaMemcpyAsync(bslz4_d_chunk(c, slot) - tb, bslz4_h_chunk(c, slot) - tb, tb + chunk_bytes,
                           cudaMemcpyHostToDevice, st);
}

static int bslz4_setup(ffs_ctx* c) {
    if (c->h_chunks) return FFS_OK;
    const size_t S = c->p.n_slots;
    const size_t es = c->p.bytes_per_pixel;
    const size_t n = (size_t)c->p.width * c->p.height;
    const size_t min_block = 128;  // elements; the smallest block bitshuffle ever chooses
    c->max_blocks = (uint32_t)(n / min_block + 2);
    // worst case: incompressible data in the smallest blocks -- per block a 4-byte header and
    // the LZ4 bound (len + len/255 + 16) -- plus the filter header and slack
    const size_t frame_bytes = n * es;
    c->chunk_cap = round_up(frame_bytes + frame_bytes / 255 + 20 * (size_t)c->max_blocks + 12 + 64, 256);
    FFS_CK(c, cudaSetDevice(c->p.device));
    // in front of every chunk: uint32 work order (longest payload first) | uint2 {offset, length} per
    // block | Bslz4Dyn, written backwards from the chunk's start (context.cuh)
    c->table_cap = round_up(sizeof(Bslz4Dyn) + (sizeof(uint2) + sizeof(uint32_t)) * (size_t)c->max_blocks, 256);
    c->chunk_stride = c->table_cap + c->chunk_cap;
    FFS_CK(c, cudaHostAlloc(&c->h_chunks, c->chunk_stride * S, cudaHostAllocDefault));
    FFS_CK(c, cudaMalloc(&c->d_chunks, c->chunk_stride * S + 64));
    FFS_CK(c, cudaMalloc(&c->d_dec_status, 2 * sizeof(unsigned) * S));  // per slot {status, work counter}
    FFS_CK(c, cudaHostAlloc(&c->h_dec_status, sizeof(unsigned) * S, cudaHostAllocDefault));
    memset(c->h_dec_status, 0, sizeof(unsigned) * S);
    return FFS_OK;
}

static void bslz4_teardown(ffs_ctx* c) {
    cudaFreeHost(c->h_chunks);
    cudaFree(c->d_chunks);
    cudaFree(c->d_dec_status);
    cudaFreeHost(c->h_dec_status);
}

// Walk the block headers of `chunk` (host memory) and fill the slot's block table.
static int bslz4_index(ffs_ctx* c, int slot, const uint8_t* chunk, size_t len, Bslz4Dyn& a) {
    const size_t es = c->p.bytes_per_pixel;
    const size_t n = (size_t)c->p.width * c->p.height;
    if (len < 12 || len > c->chunk_cap) return FFS_ERR_FORMAT;
    const uint64_t total = (uint64_t)be32h(chunk) << 32 | be32h(chunk + 4);
    if (total != (uint64_t)n * es) return FFS_ERR_FORMAT;
    size_t block = be32h(chunk + 8) / es;
    if (block == 0) block = std::max<size_t>((8192 / es) / 8 * 8, 128);  // bshuf_default_block_size
    if (block % 8 || block * es > (size_t)kBslz4MaxBlockBytes) return FFS_ERR_FORMAT;
    const size_t full = n / block;
    size_t last = n % block;
    last -= last % 8;
    const size_t n_blocks = full + (last ? 1 : 0);
    if (n_blocks > c->max_blocks) return FFS_ERR_FORMAT;
    Bslz4Dyn* hdr = reinterpret_cast<Bslz4Dyn*>(bslz4_h_chunk(c, slot)) - 1;
    uint2* tab = reinterpret_cast<uint2*>(hdr) - n_blocks;
    size_t pos = 12;
    for (size_t b = 0; b < n_blocks; ++b) {
        if (len - pos < 4) return FFS_ERR_FORMAT;
        const uint32_t clen = be32h(chunk + pos);
        pos += 4;
        if (clen > len - pos || clen > (uint32_t)(kBslz4InBytes - 32)) return FFS_ERR_FORMAT;
        tab[b] = make_uint2((uint32_t)pos, clen);
        pos += clen;
    }
    const size_t tail = n % 8;
    if (len - pos < tail * es) return FFS_ERR_FORMAT;
    a.n_blocks = (uint32_t)n_blocks;
    a.block_elems = (uint32_t)block;
    a.last_elems = (uint32_t)(last ? last : block);
    a.tail_off = (uint32_t)pos;
    a.tail_elems = (uint32_t)tail;
    a.tail_elem0 = (uint32_t)(n - tail);
    a.pad[0] = a.pad[1] = 0;
    *hdr = a;
    // Work order: longest compressed payload first.  A block's decode time grows with its payload and
    // the kernel is ~1.15 waves of warps, so the blocks that start last decide when it ends: make those
    // the cheapest ones (module-gap blocks compress to a few bytes).  Counting sort, O(n_blocks).
    {
        uint32_t* order = reinterpret_cast<uint32_t*>(tab) - n_blocks;
        constexpr int kBuckets = 64;
        uint32_t start[kBuckets + 1] = {};
        auto bucket = [](uint32_t clen) { return kBuckets - 1 - (int)std::min<uint32_t>(clen >> 7, kBuckets - 1); };
        for (size_t b = 0; b < n_blocks; ++b) start[bucket(tab[b].y) + 1]++;
        for (int k = 0; k < kBuckets; ++k) start[k + 1] += start[k];
        for (size_t b = 0; b < n_blocks; ++b) order[start[bucket(tab[b].y)]++] = (uint32_t)b;
    }
    return FFS_OK;
}

static int bslz4_launch(ffs_ctx* c, int slot, uint32_t n_blocks, cudaStream_t st) {
#ifdef BSLZ4_PERSISTENT
    constexpr int kWarps = 28;  // 28 x 8208 B of shared memory: one CTA per SM
#else
    constexpr int kWarps = 2;
#endif
    Bslz4Args a{};
    a.chunk = bslz4_d_chunk(c, slot);
    a.dyn = reinterpret_cast<const Bslz4Dyn*>(a.chunk) - 1;
    a.frame = c->d_frames + (size_t)slot * c->frame_stride;
    a.pitch = c->pitch;
    a.width = c->p.width;
    a.elem_size = c->p.bytes_per_pixel;
    a.inv_width = (uint32_t)((1ull << 32) / c->p.width) + 1u;
    a.status = c->d_dec_status + 2 * slot;
    const int smem = kWarps * kBslz4SmemPerWarp;
#ifdef BSLZ4_PERSISTENT
    int n_sm = 0;
    FFS_CK(c, cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, c->p.device));
    const unsigned grid = std::max(1u, std::min((unsigned)n_sm, (n_blocks + kWarps - 1) / kWarps));
#else
    const unsigned grid = std::max(1u, (n_blocks + kWarps - 1) / kWarps);
#endif
    static bool attr_done[64][2] = {};
    const int pi = c->p.bytes_per_pixel == 2 ? 0 : 1;
    if (!attr_done[c->p.device & 63][pi]) {
        if (pi == 0)
            FFS_CK(c, cudaFuncSetAttribute(bslz4_decode_kernel<uint16_t, kWarps>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        else
            FFS_CK(c, cudaFuncSetAttribute(bslz4_decode_kernel<uint32_t, kWarps>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_done[c->p.device & 63][pi] = true;
    }
    FFS_CK(c, cudaMemsetAsync(c->d_dec_status + 2 * slot, 0, 2 * sizeof(unsigned), st));
    if (pi == 0)
        bslz4_decode_kernel<uint16_t, kWarps><<<grid, kWarps * 32, smem, st>>>(a);
    else
        bslz4_decode_kernel<uint32_t, kWarps><<<grid, kWarps * 32, smem, st>>>(a);
    c->launches++;
    FFS_CK(c, cudaGetLastError());
    return FFS_OK;
}

extern "C" int ffs_acquire_compressed(ffs_ctx* c, int* slot, void** pinned_chunk, size_t* capacity) {
    if (!c || !slot || !pinned_chunk) return FFS_ERR_ARG;
    {
        std::lock_guard<std::mutex> lk(c->mu);
        int rc = bslz4_setup(c);
        if (rc) return rc;
    }
    void* frame = nullptr;
    int rc = ffs_acquire(c, slot, &frame);
    if (rc) return rc;
    *pinned_chunk = bslz4_h_chunk(c, *slot);
    if (capacity) *capacity = c->chunk_cap;
    return FFS_OK;
}

extern "C" int ffs_slot_buffers(ffs_ctx* c, int slot, void** pinned_pixels, void** pinned_chunk,
                                size_t* chunk_capacity) {
    if (!slots_ok(c, slot, 1)) return FFS_ERR_ARG;
    std::lock_guard<std::mutex> lk(c->mu);
    if (pinned_chunk || chunk_capacity) {
        int rc = bslz4_setup(c);
        if (rc) return rc;
    }
    if (pinned_pixels)
        *pinned_pixels = c->h_frames + (size_t)slot * ((size_t)c->p.width * c->p.height * c->p.bytes_per_pixel);
    if (pinned_chunk) *pinned_chunk = bslz4_h_chunk(c, slot);
    if (chunk_capacity) *chunk_capacity = c->chunk_cap;
    return FFS_OK;
}

extern "C" int ffs_submit_compressed(ffs_ctx* c, int slot, uint64_t frame_id, size_t chunk_bytes) {
    if (!slots_ok(c, slot, 1) || !c->h_chunks) return FFS_ERR_ARG;
    // The header walk (one dependent, usually cache-missing read per block: ~75 ns each when
    // the chunk was DMA-ed into the buffer rather than written by this core) runs outside the
    // context lock: the slot belongs to the caller until it is submitted, so producer threads
    // index their chunks in parallel.
    const uint8_t* chunk = bslz4_h_chunk(c, slot);
    Bslz4Dyn dyn{};
    int rc = bslz4_index(c, slot, chunk, chunk_bytes, dyn);
    std::unique_lock<std::mutex> lk(c->mu);
    StreamSlot& s = c->slots[slot];
    if (s.state != 1) return FFS_ERR_STATE;
    if (rc) {
        c->err = "ffs_submit_compressed: malformed or unsupported bitshuffle-LZ4 chunk";
        return rc;
    }
    FFS_CK(c, cudaSetDevice(c->p.device));
    cudaStream_t st = c->streams[s.stream];
    if (c->frame_timing) {  // "Copy" of the timing report starts with the chunk's upload
        if (!s.t_ev[0]) FFS_CK(c, cudaEventCreate(&s.t_ev[0]));
        FFS_CK(c, cudaEventRecord(s.t_ev[0], st));
    }
    // the upload, whose size changes from frame to frame, stays outside the slot's graph
    FFS_CK(c, bslz4_upload(c, slot, dyn.n_blocks, chunk_bytes, st));
    s.compressed = true;
    return submit_tail(c, slot, 1, dyn.n_blocks, frame_id, st);
}

extern "C" int ffs_decode_chunk_device(ffs_ctx* c, int slot, const void* chunk, size_t chunk_bytes,
                                       void* frame_out) {
    if (!slots_ok(c, slot, 1) || !chunk) return FFS_ERR_ARG;
    std::unique_lock<std::mutex> lk(c->mu);
    int rc = bslz4_setup(c);
    if (rc) return rc;
    if (chunk_bytes > c->chunk_cap) return FFS_ERR_FORMAT;
    uint8_t* h = bslz4_h_chunk(c, slot);
    memcpy(h, chunk, chunk_bytes);
    Bslz4Dyn dyn{};
    rc = bslz4_index(c, slot, h, chunk_bytes, dyn);
    if (rc) return rc;
    FFS_CK(c, cudaSetDevice(c->p.device));
    cudaStream_t st = c->stream;
    FFS_CK(c, bslz4_upload(c, slot, dyn.n_blocks, chunk_bytes, st));
    rc = bslz4_launch(c, slot, dyn.n_blocks, st);
    if (rc) return rc;
    unsigned status = 0;
    FFS_CK(c, cudaMemcpyAsync(&status, c->d_dec_status + 2 * slot, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
    FFS_CK(c, cudaStreamSynchronize(st));
    if (status) return FFS_ERR_FORMAT;
    if (frame_out) {
        const size_t row = (size_t)c->p.width * c->p.bytes_per_pixel;
        FFS_CK(c, cudaMemcpy2D(frame_out, row, c->d_frames + (size_t)slot * c->frame_stride, c->pitch, row,
                               c->p.height, cudaMemcpyDeviceToHost));
    }
    return FFS_OK;
}
