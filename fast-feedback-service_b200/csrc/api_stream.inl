// api_stream.inl -- streaming API: pinned host ring -> multi-stream H2D -> kernels -> small
// result D2H.  Replaces the reference's per-thread frame loop (spotfinder.cc:725-1092): one
// pinned frame + one stream + one device frame per ring slot, no full-mask D2H, no host CC.

static constexpr uint32_t kAsyncSpots = 4096;  // spots copied back asynchronously per frame

static int stream_setup(ffs_ctx* c) {
    const size_t S = c->p.n_slots;
    const size_t frame_bytes = (size_t)c->p.width * c->p.height * c->p.bytes_per_pixel;
    const int n_streams = std::max(1, std::min<int>((int)S, env_int("FFS_STREAMS", 4)));
    FFS_CK(c, cudaHostAlloc(&c->h_frames, frame_bytes * S, cudaHostAllocDefault));
    if (c->pitch != (size_t)c->p.width * c->p.bytes_per_pixel) FFS_CK(c, cudaMalloc(&c->d_stage, frame_bytes * S));
    FFS_CK(c, cudaHostAlloc(&c->h_frame_counts, sizeof(uint32_t) * CNT_N * S, cudaHostAllocDefault));
    FFS_CK(c, cudaHostAlloc(&c->h_spots, sizeof(ffs_spot) * kAsyncSpots * S, cudaHostAllocDefault));
    c->streams.resize(n_streams);
    for (auto& s : c->streams) FFS_CK(c, cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    c->slots.resize(S);
    for (size_t i = 0; i < S; ++i) {
        FFS_CK(c, cudaEventCreateWithFlags(&c->slots[i].done, cudaEventDisableTiming));
        c->slots[i].stream = (int)(i % n_streams);
    }
    return FFS_OK;
}

static void stream_teardown(ffs_ctx* c) {
    for (auto& s : c->slots) {
        if (s.done) cudaEventDestroy(s.done);
        for (auto& e : s.t_ev)
            if (e) cudaEventDestroy(e);
        for (auto& g : s.graph)
            if (g) cudaGraphExecDestroy(g);
    }
    for (auto& s : c->streams)
        if (s) cudaStreamDestroy(s);
    cudaFreeHost(c->h_frames);
    cudaFree(c->d_stage);
    cudaFreeHost(c->h_frame_counts);
    cudaFreeHost(c->h_spots);
}

extern "C" int ffs_acquire(ffs_ctx* c, int* slot, void** pinned_pixels) {
    if (!c || !slot || !pinned_pixels) return FFS_ERR_ARG;
    std::unique_lock<std::mutex> lk(c->mu);
    for (;;) {
        for (size_t i = 0; i < c->slots.size(); ++i) {
            if (c->slots[i].state == 0) {
                c->slots[i].state = 1;
                *slot = (int)i;
                *pinned_pixels = c->h_frames + i * ((size_t)c->p.width * c->p.height * c->p.bytes_per_pixel);
                return FFS_OK;
            }
        }
        // every slot is being filled or in flight: wait until ffs_wait frees one
        bool any_inflight = !c->inflight.empty();
        if (!any_inflight) {
            bool any_filling = false;
            for (auto& s : c->slots) any_filling |= s.state == 1;
            if (!any_filling) return FFS_ERR_STATE;
        }
        c->cv.wait(lk);
    }
}

extern "C" int ffs_release(ffs_ctx* c, int slot) {
    if (!slots_ok(c, slot, 1)) return FFS_ERR_ARG;
    {
        std::lock_guard<std::mutex> lk(c->mu);
        if (c->slots[slot].state != 1) return FFS_ERR_STATE;
        c->slots[slot].state = 0;
    }
    c->cv.notify_all();
    return FFS_OK;
}

static int bslz4_launch(ffs_ctx* c, int slot, uint32_t n_blocks, cudaStream_t st);

// The device work of one frame in ring slot `slot`, queued on st.  mode 0: H2D of the dense
// pinned frame; mode 1: decode of the (already uploaded) compressed chunk.  Then threshold,
// connected components and the small result D2H.
static int slot_work(ffs_ctx* c, int slot, int mode, uint32_t n_blocks, cudaStream_t st) {
    cudaEvent_t* tev = c->frame_timing ? c->slots[slot].t_ev : nullptr;
    if (tev) {
        for (int k = 0; k < 4; ++k)
            if (!tev[k]) FFS_CK(c, cudaEventCreate(&tev[k]));
        if (mode == 0) FFS_CK(c, cudaEventRecord(tev[0], st));  // (compressed: recorded before the chunk's H2D)
    }
    if (mode == 0) {
        const size_t row = (size_t)c->p.width * c->p.bytes_per_pixel, dense = row * c->p.height;
        uint8_t* dst = c->d_frames + (size_t)slot * c->frame_stride;
        static const int copy2d = env_int("FFS_H2D_2D", 0);  // A/B: the pitched copy of round 1
        if (c->pitch == row) {
            FFS_CK(c, cudaMemcpyAsync(dst, c->h_frames + (size_t)slot * dense, dense, cudaMemcpyHostToDevice, st));
        } else if (copy2d) {
            FFS_CK(c, cudaMemcpy2DAsync(dst, c->pitch, c->h_frames + (size_t)slot * dense, row, row, c->p.height,
                                        cudaMemcpyHostToDevice, st));
        } else {
            // one contiguous copy (the link's best case), then the row pitch on the device
            uint8_t* stage = c->d_stage + (size_t)slot * dense;
            FFS_CK(c, cudaMemcpyAsync(stage, c->h_frames + (size_t)slot * dense, dense, cudaMemcpyHostToDevice, st));
            const int vb = row % 8 == 0 ? 8 : row % 4 == 0 ? 4 : 2;
            const dim3 grid((unsigned)((row / vb + 255) / 256), c->p.height);
            if (vb == 8) repitch_kernel<uint2><<<grid, 256, 0, st>>>(stage, row, dst, c->pitch);
            else if (vb == 4) repitch_kernel<uint32_t><<<grid, 256, 0, st>>>(stage, row, dst, c->pitch);
            else repitch_kernel<uint16_t><<<grid, 256, 0, st>>>(stage, row, dst, c->pitch);
            c->launches++;
            FFS_CK(c, cudaGetLastError());
        }
    } else {
        int rc = bslz4_launch(c, slot, n_blocks, st);
        if (rc) return rc;
        FFS_CK(c, cudaMemcpyAsync(c->h_dec_status + slot, c->d_dec_status + 2 * slot, sizeof(unsigned),
                                  cudaMemcpyDeviceToHost, st));
    }
    if (tev) FFS_CK(c, cudaEventRecord(tev[1], st));
    int rc = threshold_impl(c, slot, 1, st);
    if (rc == FFS_OK) rc = cc_impl(c, slot, 1, st);
    if (rc) return rc;
    if (tev) FFS_CK(c, cudaEventRecord(tev[2], st));
    FFS_CK(c, cudaMemcpyAsync(c->h_frame_counts + (size_t)slot * CNT_N, c->d_frame_counts + (size_t)slot * CNT_N,
                              sizeof(uint32_t) * CNT_N, cudaMemcpyDeviceToHost, st));
    FFS_CK(c, cudaMemcpyAsync(c->h_spots + (size_t)slot * kAsyncSpots, c->d_spots + (size_t)slot * c->cap_spots,
                              sizeof(ffs_spot) * std::min(kAsyncSpots, c->cap_spots), cudaMemcpyDeviceToHost, st));
    if (tev) FFS_CK(c, cudaEventRecord(tev[3], st));
    return FFS_OK;
}

// Queue one frame.  The launch sequence of a slot never changes (same buffers, same launch
// parameters), so from its second use on it is captured into a CUDA graph and replayed with
// one cudaGraphLaunch instead of ~20 runtime calls: the per-frame host cost is what bounds
// small-frame / compressed-ingest rates (FFS_GRAPH=0 disables).  Called with c->mu held.
static int submit_tail(ffs_ctx* c, int slot, int mode, uint32_t n_blocks, uint64_t frame_id, cudaStream_t st) {
    StreamSlot& s = c->slots[slot];
    static const int use_graph = env_int("FFS_GRAPH", 1);
    const bool graph_ok = use_graph && !c->prof_on && !c->frame_timing && s.uses[mode] >= 1;
    if (graph_ok && s.graph[mode] && mode == 1 && s.graph_blocks != n_blocks) {
        cudaGraphExecDestroy(s.graph[mode]);  // block size changed: rebuild
        s.graph[mode] = nullptr;
    }
    if (graph_ok && !s.graph[mode]) {
        const uint64_t before = c->launches.load();
        cudaGraph_t g = nullptr;
        FFS_CK(c, cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        int rc = slot_work(c, slot, mode, n_blocks, st);
        cudaError_t e = cudaStreamEndCapture(st, &g);
        const uint64_t after = c->launches.load();
        c->launches -= (after - before);  // nothing ran during capture
        if (rc) {
            if (g) cudaGraphDestroy(g);
            return rc;
        }
        FFS_CK(c, e);
        e = cudaGraphInstantiate(&s.graph[mode], g, 0);
        cudaGraphDestroy(g);
        FFS_CK(c, e);
        s.graph_kernels[mode] = (uint32_t)(after - before);
        s.graph_blocks = n_blocks;
    }
    if (graph_ok) {
        FFS_CK(c, cudaGraphLaunch(s.graph[mode], st));
        c->launches += s.graph_kernels[mode];
    } else {
        int rc = slot_work(c, slot, mode, n_blocks, st);
        if (rc) return rc;
    }
    s.uses[mode]++;
    FFS_CK(c, cudaEventRecord(s.done, st));
    s.frame_id = frame_id;
    s.state = 2;
    c->inflight.push_back(slot);
    c->cv.notify_all();
    return FFS_OK;
}

extern "C" int ffs_submit(ffs_ctx* c, int slot, uint64_t frame_id) {
    if (!slots_ok(c, slot, 1)) return FFS_ERR_ARG;
    std::unique_lock<std::mutex> lk(c->mu);
    StreamSlot& s = c->slots[slot];
    if (s.state != 1) return FFS_ERR_STATE;
    FFS_CK(c, cudaSetDevice(c->p.device));
    cudaStream_t st = c->streams[s.stream];
    s.compressed = false;
    return submit_tail(c, slot, 0, 0, frame_id, st);
}

extern "C" int ffs_submit_copy(ffs_ctx* c, uint64_t frame_id, const void* pixels) {
    if (!c || !pixels) return FFS_ERR_ARG;
    int slot = -1;
    void* dst = nullptr;
    int rc = ffs_acquire(c, &slot, &dst);
    if (rc) return rc;
    memcpy(dst, pixels, (size_t)c->p.width * c->p.height * c->p.bytes_per_pixel);
    return ffs_submit(c, slot, frame_id);
}

extern "C" int ffs_pending(ffs_ctx* c) {
    if (!c) return 0;
    std::lock_guard<std::mutex> lk(c->mu);
    return (int)c->inflight.size();
}

extern "C" int ffs_wait(ffs_ctx* c, ffs_frame_result* result, ffs_spot* spots, size_t max_spots,
                        size_t* n_spots) {
    if (!c) return FFS_ERR_ARG;
    int slot;
    {
        std::lock_guard<std::mutex> lk(c->mu);
        if (c->inflight.empty()) return FFS_ERR_STATE;
        slot = c->inflight.front();
        c->inflight.pop_front();
    }
    StreamSlot& s = c->slots[slot];
    FFS_CK(c, cudaSetDevice(c->p.device));
    FFS_CK(c, cudaEventSynchronize(s.done));
    const uint32_t* cnt = c->h_frame_counts + (size_t)slot * CNT_N;
    if (result) counts_to_result(cnt, s.frame_id, result);
    if (c->frame_timing && s.t_ev[3]) {
        for (int k = 0; k < 3; ++k)
            if (cudaEventElapsedTime(&c->last_timing[k], s.t_ev[k], s.t_ev[k + 1]) != cudaSuccess) c->last_timing[k] = 0;
        cudaGetLastError();
    }
    if (s.compressed && c->h_dec_status && c->h_dec_status[slot] != 0) {
        // malformed bitshuffle-LZ4 input: the frame was not decoded, its result is void
        if (result) result->status = FFS_ERR_FORMAT;
    }
    const size_t have = std::min<size_t>(cnt[CNT_COMPONENTS], c->cap_spots);
    if (n_spots) *n_spots = have;
    int rc = FFS_OK;
    if (spots && max_spots) {
        const size_t k = std::min(have, max_spots);
        const size_t k_async = std::min<size_t>(k, kAsyncSpots);
        memcpy(spots, c->h_spots + (size_t)slot * kAsyncSpots, k_async * sizeof(ffs_spot));
        if (k > k_async) {  // rare: more spots than the async window
            cudaError_t e = cudaMemcpy(spots + k_async, c->d_spots + (size_t)slot * c->cap_spots + k_async,
                                       (k - k_async) * sizeof(ffs_spot), cudaMemcpyDeviceToHost);
            if (e != cudaSuccess) {
                c->err = cudaGetErrorString(e);
                rc = FFS_ERR_CUDA;
            }
        }
    }
    if (rc == FFS_OK && c->self_check) rc = self_check_slot(c, slot);
    if (rc == FFS_OK && c->sweep_on) rc = sweep_record(c, slot, s.frame_id, cnt[CNT_STRONG]);
    if (rc == FFS_OK && c->keep_pixels) {  // --writeout: the frame's strong-pixel list, before the slot is reused
        c->last_pixels.resize(cnt[CNT_STRONG]);
        if (cnt[CNT_STRONG]) {
            cudaError_t e = cudaMemcpy(c->last_pixels.data(), c->d_px_index + (size_t)slot * c->cap_strong,
                                       cnt[CNT_STRONG] * 4ull, cudaMemcpyDeviceToHost);
            if (e != cudaSuccess) {
                c->err = cudaGetErrorString(e);
                rc = FFS_ERR_CUDA;
            }
        }
    }
    if (rc == FFS_OK && c->keep_pixels && c->keep_frame) {  // the dense frame as the device saw it
        const size_t row = (size_t)c->p.width * c->p.bytes_per_pixel;
        c->last_frame.resize(row * c->p.height);
        cudaError_t e = cudaMemcpy2D(c->last_frame.data(), row, c->d_frames + (size_t)slot * c->frame_stride, c->pitch,
                                     row, c->p.height, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) {
            c->err = cudaGetErrorString(e);
            rc = FFS_ERR_CUDA;
        }
    }
    {
        std::lock_guard<std::mutex> lk(c->mu);
        s.state = 0;
    }
    c->cv.notify_all();
    return rc;
}

// Per-image timing of the reference's `--threads 1` report (spotfinder.cc:1055-1076).
extern "C" int ffs_frame_timing(ffs_ctx* c, int on) {
    if (!c) return FFS_ERR_ARG;
    std::lock_guard<std::mutex> lk(c->mu);
    c->frame_timing = on != 0;
    return FFS_OK;
}

extern "C" int ffs_last_frame_timing(ffs_ctx* c, float* copy_ms, float* kernel_ms, float* post_copy_ms) {
    if (!c || !c->frame_timing) return FFS_ERR_ARG;
    if (copy_ms) *copy_ms = c->last_timing[0];
    if (kernel_ms) *kernel_ms = c->last_timing[1];
    if (post_copy_ms) *post_copy_ms = c->last_timing[2];
    return FFS_OK;
}

extern "C" int ffs_keep_pixel_lists(ffs_ctx* c, int on) {
    if (!c) return FFS_ERR_ARG;
    c->keep_pixels = on != 0;
    c->keep_frame = on == 2;
    return FFS_OK;
}

extern "C" int ffs_last_frame(ffs_ctx* c, const void** pixels, size_t* bytes) {
    if (!c || !pixels || !c->keep_frame) return FFS_ERR_ARG;
    *pixels = c->last_frame.data();
    if (bytes) *bytes = c->last_frame.size();
    return FFS_OK;
}

extern "C" int ffs_last_pixel_list(ffs_ctx* c, const uint32_t** index, size_t* n) {
    if (!c || !index || !n || !c->keep_pixels) return FFS_ERR_ARG;
    *index = c->last_pixels.data();
    *n = c->last_pixels.size();
    return FFS_OK;
}
