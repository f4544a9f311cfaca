    // ---- continuation of main(): context, frame loop, results (spotfinder.cc:606-1331) ----------
    ffs_params p;
    ffs_default_params(&p);
    p.width = width;
    p.height = height;
    p.bytes_per_pixel = sizeof(pixel_t);
    p.algorithm = algo == "dispersion" ? FFS_ALGO_DISPERSION : FFS_ALGO_DISPERSION_EXTENDED;
    p.min_spot_size = args.min_spot_size;
    p.min_spot_size_3d = args.min_spot_size_3d;
    p.max_peak_centroid_separation = args.max_peak_centroid_separation;
    // max_valid_pixel_value of the reference's kernels (spotfinder.cc:482,868,879): the int64 cutoff
    // is narrowed to pixel_t at the call (spotfinder.cuh:16), reproduced here
    p.trusted_max = (int64_t)(pixel_t)reader.get_trusted_range()[1];
    p.device = args.device;
    p.n_slots = (int32_t)std::min<uint32_t>(std::max<uint32_t>(args.threads + 2, 6), 24);
    // One context (buffers, streams, pinned ring) per GPU.  `--devices` shards the stream frame by
    // frame over several GPUs inside this process, as the reference shards it over host threads
    // (shared atomic frame counter, spotfinder.cc:686,752); -d/--device alone is the reference's CLI.
    std::vector<int> devices = args.devices.empty() ? std::vector<int>{args.device} : args.devices;
    const size_t n_dev = devices.size();
    if (n_dev > 1) {
        printf("Sharding frames over %zu GPUs:", n_dev);
        for (int d : devices) printf(" %d", d);
        printf("\n");
    }
    std::vector<ffs_ctx*> ctxs(n_dev, nullptr);
    auto destroy_all = [&]() {
        for (ffs_ctx* c : ctxs) ffs_destroy(c);
    };
    for (size_t d = 0; d < n_dev; ++d) {
        p.device = devices[d];
        if (int rc = ffs_create(&p, reader.get_mask().data(), &ctxs[d])) {
            printf("Error: could not initialise the GPU spotfinder on device %d (%d): %s\n", devices[d], rc,
                   ffs_last_error(nullptr));
            destroy_all();
            return 1;
        }
    }
    ffs_ctx* ctx = ctxs[0];
    printf("Mask has %llu valid pixels of %u\n", (unsigned long long)ffs_mask_valid_count(ctx), width * height);
    // --writeout (spotfinder.cc:626-646,653-679): white = valid, red = masked
    auto write_mask_png = [&](const char* name, const uint8_t* m) {
        std::vector<uint8_t> rgb((size_t)width * height * 3, 255);
        for (size_t k = 0; k < (size_t)width * height; ++k)
            if (!m[k]) rgb[3 * k + 1] = rgb[3 * k + 2] = 0;
        if (!pngw::write_rgb(name, rgb.data(), width, height)) printf("Warning: could not write %s\n", name);
    };
    if (args.writeout) {
        write_mask_png("mask_source.png", reader.get_mask().data());
        for (ffs_ctx* c : ctxs) ffs_keep_pixel_lists(c, 2);
    }
    if (args.dmin > 0 || args.dmax > 0) {  // apply_resolution_filtering, spotfinder.cc:110-137,650-652
        for (ffs_ctx* c : ctxs)
            if (ffs_apply_resolution_mask(c, wavelength, det.distance, det.beam_center_x, det.beam_center_y,
                                          det.pixel_size_x, det.pixel_size_y, args.dmin, args.dmax)) {
                printf("Error: resolution mask failed: %s\n", ffs_last_error(c));
                destroy_all();
                return 1;
            }
        printf("Resolution mask applied: %llu valid pixels remain\n", (unsigned long long)ffs_mask_valid_count(ctx));
        if (args.writeout) {
            std::vector<uint8_t> calc((size_t)width * height);
            if (ffs_get_mask(ctx, calc.data()) == FFS_OK) write_mask_png("mask_calculated.png", calc.data());
        }
    }
    // --threads 1: the per-image Copy / Kernel / Post Copy report (spotfinder.cc:1055-1076)
    const bool timing_report = args.threads == 1;
    if (timing_report)
        for (ffs_ctx* c : ctxs) ffs_frame_timing(c, 1);
    // --validate (spotfinder.cc:1010-1052): the reference re-runs its CPU DIALS code per image.  No CPU
    // implementation lives in this build; the second opinion is the device's exact per-pixel routine
    // (ffs_self_check, csrc/api_selfcheck.inl).  tests/validate_stream.py is the check against the CPU oracle.
    bool validating = false;
    if (args.validate) {
        validating = true;
        for (ffs_ctx* c : ctxs)
            if (ffs_self_check(c, 1) != FFS_OK) validating = false;
        if (validating)
            printf("Validation: every image is decided a second time by the exact per-pixel kernel (direct window sums, "
                   "no pruning) and the strong pixels are compared\n");
        else
            printf("Note: --validate covers standard dispersion only; not validating\n");
    }
    // --save-h5 on stills: kept centroids per image, keyed by image number (spotfinder.cc:706-721,928-932)
    std::map<int, std::vector<float>> reflection_centers_2d;
    std::unique_ptr<PipeHandler> pipe;
    if (args.pipe_fd != -1) pipe = std::make_unique<PipeHandler>(args.pipe_fd);
    // Rotation data: every GPU owns a CONTIGUOUS slab of frames (3D components couple consecutive
    // frames only), labels it, and the slabs are merged at the end (ffs_sweep_merge_host).
    auto device_of = [&](uint32_t image_num) -> size_t {
        if (n_dev == 1) return 0;
        return rotation ? std::min<size_t>(n_dev - 1, (size_t)image_num * n_dev / num_images) : image_num % n_dev;
    };
    auto slab_begin = [&](size_t d) -> uint32_t {  // first frame with device_of(frame) == d
        return (uint32_t)(((uint64_t)d * num_images + n_dev - 1) / n_dev);
    };
    if (rotation)
        for (size_t d = 0; d < n_dev; ++d) ffs_sweep_begin(ctxs[d], num_images / n_dev + 1);

    const size_t px = (size_t)width * height;
    std::atomic<uint32_t> next_image{0}, completed{0};
    std::vector<std::atomic<uint32_t>> submitted(n_dev), collected(n_dev);
    for (size_t d = 0; d < n_dev; ++d) submitted[d] = collected[d] = 0;
    std::mutex print_mutex;
    std::atomic<int> workers_alive{(int)args.threads};
    std::atomic<bool> had_error{false};
    std::atomic<uint32_t> failed_images{0};
    std::atomic<uint32_t> validation_failures{0};
    std::vector<int> thread_of(num_images, 0);
    std::mutex wait_mutex;
    double time_waiting_for_images = 0;
    auto last_image_received = clk::now();
    const auto all_images_start = clk::now();
    printf("Finding spots in %u images (%u x %u) with %u threads\n", num_images, width, height, args.threads);

    // developer knob (honoured only with FFS_DEV_KNOBS=1, like the library's): decode chunks on the host
    const bool host_decode = getenv("FFS_DEV_KNOBS") && atoi(getenv("FFS_DEV_KNOBS")) != 0 &&
                             getenv("FFS_HOST_DECODE") && atoi(getenv("FFS_HOST_DECODE")) != 0;
    // developer knob: where the worker threads' time goes (summed over threads, printed at the end)
    const bool cli_timing = getenv("FFS_DEV_KNOBS") && atoi(getenv("FFS_DEV_KNOBS")) != 0 && getenv("FFS_CLI_TIMING");
    std::atomic<long long> ns_avail{0}, ns_acquire{0}, ns_read{0}, ns_submit{0};
    auto ns_since = [](clk::time_point t) { return (long long)std::chrono::duration_cast<std::chrono::nanoseconds>(clk::now() - t).count(); };
    auto worker = [&](int thread_id) {
        std::vector<uint8_t> chunk, scratch;
        while (!g_stop) {
            const uint32_t image_num = next_image.fetch_add(1);
            if (image_num >= num_images) break;
            const size_t offset_image_num = image_num + args.start_index;
            const size_t dev = device_of(image_num);
            ffs_ctx* ctx = ctxs[dev];
            {   // availability poll every 100 ms, abort after `timeout` s without a new image (:770-787)
                const auto wait_start = clk::now();
                while (!reader.is_image_available(offset_image_num) && !g_stop) {
                    std::lock_guard<std::mutex> lk(wait_mutex);
                    if (seconds_since(last_image_received) > args.timeout) {
                        printf("Timeout waiting for image %zu\n", offset_image_num);
                        g_stop = true;  // graceful stop and exit 0, as the reference (spotfinder.cc:778-782)
                        break;
                    }
                    std::this_thread::sleep_for(100ms);
                }
                if (g_stop) break;
                std::lock_guard<std::mutex> lk(wait_mutex);
                last_image_received = clk::now();
                time_waiting_for_images += seconds_since(wait_start);
                if (cli_timing) ns_avail += ns_since(wait_start);
            }
            int slot = -1;
            void* pinned = nullptr;
            auto t_phase = clk::now();
            if (ffs_acquire(ctx, &slot, &pinned) != FFS_OK) break;
            if (cli_timing) ns_acquire += ns_since(t_phase);
            thread_of[image_num] = thread_id;
            // The compressed chunk is read straight into the slot's pinned upload buffer, uploaded
            // and decoded on the device (bslz4_gpu.cuh).  The reference decompresses here on the
            // host (spotfinder.cc:823-841); that path is kept for chunks the device decoder does
            // not take (FFS_HOST_DECODE=1 forces it).
            if (reader.chunks_are_raw_pixels()) {
                // a source stored without a compression filter (reader_h5.hpp): pixels go straight
                // into the pinned ring slot
                if (!reader.read_pixels(offset_image_num, pinned, px * sizeof(pixel_t))) {
                    printf("Error: could not read image %zu\n", offset_image_num);
                    g_stop = true;
                    had_error = true;
                    ffs_release(ctx, slot);
                    break;
                }
                if (ffs_submit(ctx, slot, image_num) != FFS_OK) {
                    printf("Error: submit failed: %s\n", ffs_last_error(ctx));
                    g_stop = true;
                    had_error = true;
                    break;
                }
                submitted[dev]++;
                continue;
            }
            void* pchunk = nullptr;
            size_t cap = 0;
            const bool direct = !host_decode && ffs_slot_buffers(ctx, slot, nullptr, &pchunk, &cap) == FFS_OK;
            size_t chunk_bytes = 0;
            const uint8_t* chunk_data = nullptr;
            t_phase = clk::now();
            for (;;) {  // zero-length chunk = writer not finished: sleep and retry (:805-821)
                size_t got = direct ? reader.read_raw_chunk(offset_image_num, static_cast<uint8_t*>(pchunk), cap) : SIZE_MAX;
                if (got == SIZE_MAX) {  // host path, or a chunk larger than the upload buffer
                    got = reader.get_raw_chunk(offset_image_num, chunk) ? chunk.size() : 0;
                    chunk_data = chunk.data();
                } else {
                    chunk_data = static_cast<const uint8_t*>(pchunk);
                }
                if (got == 0) {
                    printf("Race Condition?!?? Got buffer size 0 for image %u. Sleeping.\n", image_num);
                    std::this_thread::sleep_for(100ms);
                    if (g_stop) break;
                    continue;
                }
                chunk_bytes = got;
                break;
            }
            if (cli_timing) ns_read += ns_since(t_phase);
            t_phase = clk::now();
            int rc = FFS_ERR_FORMAT;
            if (chunk_bytes == 0) {  // stopped while waiting for the writer
                ffs_release(ctx, slot);
                break;
            }
            if (direct && chunk_data == pchunk) rc = ffs_submit_compressed(ctx, slot, image_num, chunk_bytes);
            if (rc == FFS_ERR_FORMAT) {
                // decompress straight into the pinned ring slot (no intermediate host copy)
                if (bslz4::decompress_chunk(chunk_data, chunk_bytes, pinned, px, sizeof(pixel_t), scratch) != 0) {
                    printf("Error: could not decompress image %zu\n", offset_image_num);
                    g_stop = true;
                    had_error = true;
                    memset(pinned, 0, px * sizeof(pixel_t));
                }
                rc = ffs_submit(ctx, slot, image_num);
            }
            if (rc != FFS_OK) {
                printf("Error: submit failed: %s\n", ffs_last_error(ctx));
                g_stop = true;
                had_error = true;
                break;
            }
            if (cli_timing) ns_submit += ns_since(t_phase);
            submitted[dev]++;
        }
        workers_alive--;
    };

    auto collector = [&](size_t dev) {
        ffs_ctx* ctx = ctxs[dev];
        std::vector<ffs_spot> spots(65536);
        for (;;) {
            ffs_frame_result r;
            size_t n = 0;
            const int rc = ffs_wait(ctx, &r, spots.data(), spots.size(), &n);
            if (rc == FFS_ERR_STATE) {
                if (workers_alive == 0 && collected[dev] >= submitted[dev]) break;
                std::this_thread::sleep_for(200us);
                continue;
            }
            if (rc != FFS_OK) {
                // ffs_wait has already popped the slot: count it, or the exit test above never holds
                printf("Error: %s\n", ffs_last_error(ctx));
                g_stop = true;
                had_error = true;
                collected[dev]++;
                continue;
            }
            if (r.status == FFS_ERR_FORMAT) {
                // the device decoder rejected the chunk after the header walk accepted it: the ring
                // slot holds a partial frame, so nothing of this image is reported
                printf("Error: could not decompress image %llu\n", (unsigned long long)r.frame_id);
                had_error = true;
                failed_images++;
                collected[dev]++;
                continue;
            }
            std::lock_guard<std::mutex> plk(print_mutex);  // one image's lines stay together
            if (r.status != 0)
                printf("Warning: image %llu exceeded a spot-list capacity; results truncated\n", (unsigned long long)r.frame_id);
            if (n > spots.size()) n = spots.size();
            const uint32_t image_num = (uint32_t)r.frame_id;
            // ConnectedComponents::generate_boxes log lines (connected_components.cc:119,133)
            printf("Extracted %u spots\n", r.n_components);
            if (args.min_spot_size > 0)
                printf("Removed %u spots with size < %u pixels\n", r.n_components - r.n_spots_total, args.min_spot_size);
            std::string centers;
            if (!rotation && (args.save_h5 || args.output_for_index)) {  // find_2d_components (:238-266)
                const size_t kept = ffs_filter_spots(spots.data(), n, args.min_spot_size, args.max_peak_centroid_separation,
                                                     nullptr, nullptr);
                if (r.n_spots_total > kept)
                    printf("Filtered %zu spots with peak-centroid distance > %s\n", (size_t)r.n_spots_total - kept,
                           shortest(args.max_peak_centroid_separation).c_str());
                for (size_t i = 0; i < kept; ++i) {
                    if (i) centers += ",";
                    centers += jsonmini::number(spots[i].com_x) + "," + jsonmini::number(spots[i].com_y) + "," +
                               jsonmini::number(spots[i].com_z);
                }
                if (args.save_h5) {  // print_mutex is held
                    auto& v = reflection_centers_2d[(int)(image_num + args.start_index)];
                    v.clear();
                    for (size_t i = 0; i < kept; ++i) v.insert(v.end(), {spots[i].com_x, spots[i].com_y, spots[i].com_z});
                }
            }
            if (pipe) {  // keys in sorted order, as nlohmann::json dumps them (spotfinder.cc:997-1008)
                std::string js = "{\"file\":" + jsonmini::escape(file) + ",\"file-number\":" + std::to_string(image_num) +
                                 ",\"n_spots_total\":" + std::to_string(r.n_spots_total) +
                                 ",\"num_strong_pixels\":" + std::to_string(r.num_strong_pixels);
                if (args.output_for_index) js += ",\"spot_centers\":[" + centers + "]";
                pipe->send(js + "}");
            }
            if (args.writeout) {  // image_NNNNN.png + pixels_NNNNN.txt (spotfinder.cc:939-994)
                const uint32_t* pidx = nullptr;
                size_t np = 0;
                const void* fr = nullptr;
                if (ffs_last_pixel_list(ctx, &pidx, &np) == FFS_OK && ffs_last_frame(ctx, &fr, nullptr) == FFS_OK) {
                    const pixel_t* img = static_cast<const pixel_t*>(fr);
                    std::vector<uint8_t> rgb(px * 3);
                    for (size_t k = 0; k < px; ++k) {
                        const uint8_t gray = (uint8_t)std::max(0.0f, 255.99f - static_cast<float>(img[k]) * 10);
                        rgb[3 * k] = rgb[3 * k + 1] = rgb[3 * k + 2] = gray;
                    }
                    auto blue = [&](long x, long y) {  // (the reference does not clip the frame: we do)
                        if (x < 0 || y < 0 || x >= (long)width || y >= (long)height) return;
                        uint8_t* q = &rgb[3 * ((size_t)y * width + x)];
                        q[0] = 0; q[1] = 0; q[2] = 255;
                    };
                    for (size_t i = 0; i < n; ++i) {  // get_boxes(): components of at least min_spot_size pixels
                        const ffs_spot& b = spots[i];
                        if (args.min_spot_size > 0 && b.num_pixels < args.min_spot_size) continue;
                        for (long edge = 5; edge <= 7; ++edge) {
                            for (long x = (long)b.x_min - edge; x <= (long)b.x_max + edge; ++x) {
                                blue(x, (long)b.y_min - edge);
                                blue(x, (long)b.y_max + edge);
                            }
                            for (long y = (long)b.y_min - edge; y <= (long)b.y_max + edge; ++y) {
                                blue((long)b.x_min - edge, y);
                                blue((long)b.x_max + edge, y);
                            }
                        }
                    }
                    char name[64];
                    snprintf(name, sizeof name, "pixels_%05u.txt", image_num);
                    FILE* pf = fopen(name, "w");
                    for (size_t i = 0; i < np; ++i) {
                        uint8_t* q = &rgb[3 * (size_t)pidx[i]];
                        q[0] = 255; q[1] = 0; q[2] = 0;
                        if (pf) fprintf(pf, "%4d, %4d\n", (int)(pidx[i] % width), (int)(pidx[i] / width));
                    }
                    if (pf) fclose(pf);
                    snprintf(name, sizeof name, "image_%05u.png", image_num);
                    if (!pngw::write_rgb(name, rgb.data(), width, height)) printf("Warning: could not write %s\n", name);
                }
            }
            uint32_t exact_strong = 0, mismatch = 0;
            if (validating && ffs_last_self_check(ctx, &exact_strong, &mismatch) == FFS_OK) {
                // the reference prints this line instead of the per-image summary (spotfinder.cc:1036-1051)
                if (mismatch == 0)
                    printf("Thread %2d, Image %4u: Compared: \033[32mMatch %u px\033[0m\n", thread_of[image_num], image_num,
                           r.num_strong_pixels);
                else {
                    printf("Thread %2d, Image %4u: Compared: \033[1;31mMismatch (%u px from kernel)\033[0m\n",
                           thread_of[image_num], image_num, r.num_strong_pixels);
                    validation_failures++;
                    had_error = true;  // unlike the reference, a failed validation is not a clean exit
                }
                completed++;
                collected[dev]++;
                continue;
            }
            float t_copy = 0, t_kernel = 0, t_post = 0;
            if (timing_report && ffs_last_frame_timing(ctx, &t_copy, &t_kernel, &t_post) == FFS_OK) {
                // the reference's single-thread block (spotfinder.cc:1055-1076); "Post" is its host-side
                // connected components, which run inside "Kernel" here
                const float total = t_copy + t_kernel + t_post;
                printf("Thread %2d finished image %4u\n"
                       "       Copy: %5.1f ms\n"
                       "     Kernel: %5.1f ms\n"
                       "  Post Copy: %5.1f ms\n"
                       "       Post: %5.1f ms\n"
                       "             \xE2\x95\x90\xE2\x95\x90\xE2\x95\x90\xE2\x95\x90\xE2\x95\x90\xE2\x95\x90\xE2\x95\x90\xE2\x95\x90\n"
                       "     Total:  %5.1f ms (%.1f GBps)\n"
                       "    \033[1m%u\033[0m strong pixels\n"
                       "    \033[1m%u\033[0m filtered reflections (\033[1m%u\033[0m pixels)\n",
                       thread_of[image_num], image_num, t_copy, t_kernel, t_post, 0.0f, total,
                       total > 0 ? (float)((double)px * sizeof(pixel_t) / (total * 1e-3) / 1e9) : 0.0f,
                       r.num_strong_pixels, r.n_spots_total, r.num_strong_pixels_filtered);
            } else {
                printf("Thread %2d finished image %4u with %5u strong pixels, %4u filtered reflections (%u pixels)\n",
                       thread_of[image_num], image_num, r.num_strong_pixels, r.n_spots_total,
                       r.num_strong_pixels_filtered);
            }
            completed++;
            collected[dev]++;
        }
    };

    {
        std::vector<std::thread> threads;
        for (uint32_t t = 0; t < args.threads; ++t) threads.emplace_back(worker, (int)t);
        std::vector<std::thread> colls;
        for (size_t d = 0; d < n_dev; ++d) colls.emplace_back(collector, d);
        for (auto& t : threads) t.join();
        for (auto& t : colls) t.join();
    }

    if (rotation) {  // spotfinder.cc:1099-1150 + find_3d_components (connected_components.cc:448-470)
        printf("Processing 3D spots\n");
        ffs_spot* s3 = nullptr;
        size_t n3 = 0;
        uint64_t *msums = nullptr, *mmoments = nullptr;  // merged sums / moments (several GPUs)
        if (n_dev == 1) {
            if (ffs_sweep_finish(ctx, &s3, &n3) != FFS_OK) {
                printf("Error: 3D connected components failed: %s\n", ffs_last_error(ctx));
                return 1;
            }
        } else {
            // every GPU labels its slab; slabs are merged through their boundary planes (csrc/api_multi.inl)
            std::vector<ffs_spot*> sp(n_dev, nullptr);
            std::vector<size_t> ns(n_dev, 0), nfirst(n_dev, 0), nlast(n_dev, 0);
            std::vector<uint64_t*> sums(n_dev, nullptr), moms(n_dev, nullptr);
            std::vector<uint32_t*> first(n_dev, nullptr), last(n_dev, nullptr);
            std::vector<uint64_t> z0(n_dev), nf(n_dev);
            bool ok = true;
            for (size_t d = 0; d < n_dev && ok; ++d) {
                size_t n_s = 0, n_m = 0;
                ok = ffs_sweep_finish(ctxs[d], &sp[d], &ns[d]) == FFS_OK &&
                     ffs_sweep_export_sums(ctxs[d], &sums[d], &n_s) == FFS_OK &&
                     ffs_sweep_export_moments(ctxs[d], &moms[d], &n_m) == FFS_OK &&
                     ffs_sweep_export_plane(ctxs[d], 0, &first[d], &nfirst[d]) == FFS_OK &&
                     ffs_sweep_export_plane(ctxs[d], 1, &last[d], &nlast[d]) == FFS_OK;
                if (!ok) printf("Error: 3D connected components failed on device %d: %s\n", devices[d], ffs_last_error(ctxs[d]));
                z0[d] = slab_begin(d);
                nf[d] = (d + 1 < n_dev ? slab_begin(d + 1) : num_images) - slab_begin(d);
            }
            if (ok && ffs_sweep_merge_host((int)n_dev, width, height, z0.data(), nf.data(), sp.data(), ns.data(),
                                           sums.data(), moms.data(), first.data(), nfirst.data(), last.data(),
                                           nlast.data(), &s3, &msums, &mmoments, &n3) != FFS_OK) {
                printf("Error: slab merge failed\n");
                ok = false;
            }
            for (size_t d = 0; d < n_dev; ++d) {
                ffs_free(sp[d]); ffs_free(sums[d]); ffs_free(moms[d]); ffs_free(first[d]); ffs_free(last[d]);
            }
            if (!ok) return 1;
        }
        printf("Calculated %zu spots\n", n3);
        // Kabsch-space variances of every component (spotfinder.cc:1152-1196), evaluated before
        // the filter compacts the array; the kept spots are picked out by label afterwards
        std::vector<double> var_b(n3), var_m(n3);
        std::vector<int32_t> depth(n3);
        std::vector<uint64_t> labels3(n3);
        for (size_t i = 0; i < n3; ++i) labels3[i] = s3[i].label;
        if (n3) {
            uint64_t *sums = nullptr, *moments = nullptr;
            size_t ns = 0, nm = 0;
            ffs_kabsch_geometry kg{};
            kg.distance_mm = det.distance * 1000;
            kg.beam_center_x_px = det.beam_center_x;
            kg.beam_center_y_px = det.beam_center_y;
            kg.pixel_size_x_mm = det.pixel_size_x * 1000;
            kg.pixel_size_y_mm = det.pixel_size_y * 1000;
            kg.wavelength = wavelength;
            kg.oscillation_width_deg = oscillation_width;
            if (n_dev > 1) {
                sums = msums;
                moments = mmoments;
                ns = nm = n3;
            }
            if ((n_dev == 1 && (ffs_sweep_export_sums(ctx, &sums, &ns) != FFS_OK ||
                                ffs_sweep_export_moments(ctx, &moments, &nm) != FFS_OK)) ||
                ns != n3 || nm != n3 ||
                ffs_kabsch_variances(ctx, &kg, s3, sums, moments, n3, var_b.data(), var_m.data(), depth.data()) != FFS_OK) {
                printf("Error: spot variance calculation failed: %s\n", ffs_last_error(ctx));
                return 1;
            }
            ffs_free(sums);
            ffs_free(moments);
        }
        size_t rs = 0, rp = 0;
        const size_t kept = ffs_filter_spots(s3, n3, args.min_spot_size_3d, args.max_peak_centroid_separation, &rs, &rp);
        if (rs > 0) printf("Filtered %zu spots with size < %u pixels\n", rs, args.min_spot_size_3d);
        if (rp > 0) printf("Filtered %zu spots with peak-centroid distance > %s\n", rp, shortest(args.max_peak_centroid_separation).c_str());
        printf("Found %zu spots\n", kept);
        std::vector<double> kept_var_b, kept_var_m;
        std::vector<int> kept_depth;
        {   // estimated global values over the kept reflections (spotfinder.cc:1180-1212)
            double sum_b = 0, sum_m = 0;
            size_t n_m = 0, j = 0;
            for (size_t i = 0; i < kept; ++i) {
                while (j < n3 && labels3[j] != s3[i].label) ++j;  // the filter is stable
                if (j == n3) break;
                kept_var_b.push_back(var_b[j]);
                kept_var_m.push_back(var_m[j]);
                kept_depth.push_back(depth[j]);
                sum_b += var_b[j];
                if (depth[j] >= 5) {  // min_bbox_depth
                    sum_m += var_m[j];
                    ++n_m;
                }
            }
            const double rad_to_deg = 180.0 / M_PI;
            if (kept) printf("Estimated sigma_b (degrees): %.6f\n", std::sqrt(sum_b / kept) * rad_to_deg);
            if (n_m) printf("Estimated sigma_m (degrees): %.6f, calculated on %zu spots\n", std::sqrt(sum_m / n_m) * rad_to_deg, n_m);
        }
        if (args.writeout) {
            std::ofstream out("3d_reflections.txt");
            for (size_t i = 0; i < kept; ++i)
                out << "X: [" << s3[i].x_min << ", " << s3[i].x_max << "] Y: [" << s3[i].y_min << ", " << s3[i].y_max
                    << "] Z: [" << s3[i].z_min << ", " << s3[i].z_max << "] COM: (" << s3[i].com_x << ", " << s3[i].com_y
                    << ", " << s3[i].com_z << ")\n";
        }
        if (args.save_h5) {  // spotfinder.cc:1219-1257
            try {
                std::vector<double> flat_coms;
                flat_coms.reserve(kept * 3);
                for (size_t i = 0; i < kept; ++i) flat_coms.insert(flat_coms.end(), {s3[i].com_x, s3[i].com_y, s3[i].com_z});
                h5results::Table table;
                table.add_column("xyzobs.px.value", kept, 3, flat_coms);
                std::vector<int> id(kept, (int)table.get_experiment_ids()[0]);
                table.add_column("id", kept, 1, id);
                table.add_column("sigma_b_variance", kept_var_b.size(), 1, kept_var_b);
                table.add_column("sigma_m_variance", kept_var_m.size(), 1, kept_var_m);
                table.add_column("spot_extent_z", kept_depth.size(), 1, kept_depth);
                table.write("results_ffs.h5", "dials/processing/group_0");
                printf("Successfully wrote 3D reflections to HDF5 file\n");
            } catch (const std::exception& e) {
                printf("Error writing data to HDF5 file: %s\n", e.what());
            }
        }
        ffs_free(s3);
        printf("3D spot analysis complete\n");
    } else if (args.save_h5) {  // stills: spotfinder.cc:1258-1299
        printf("Processing 2D spots\n");
        try {
            std::vector<double> flat_coms;
            std::vector<int> ids;
            int id = 0;
            for (const auto& kv : reflection_centers_2d) {  // std::map: ascending image number
                for (float c : kv.second) flat_coms.push_back((double)c);
                ids.insert(ids.end(), kv.second.size() / 3, id);
                ++id;
            }
            h5results::Table table;  // one experiment per image
            for (int i = 0; i < id - 1; ++i) table.generate_new_attributes();
            table.add_column("xyzobs.px.value", flat_coms.size() / 3, 3, flat_coms);
            table.add_column("id", ids.size(), 1, ids);
            table.write("results_ffs.h5", "dials/processing/group_0");
            printf("Successfully wrote %zu 2D reflections to HDF5 file\n", ids.size());
        } catch (const std::exception& e) {
            printf("Error writing data to HDF5 file: %s\n", e.what());
        }
        printf("2D spot analysis complete\n");
    }

    if (cli_timing)
        printf("worker time, summed over threads (s): availability %.3f, slot acquire %.3f, chunk read %.3f, submit %.3f\n",
               ns_avail * 1e-9, ns_acquire * 1e-9, ns_read * 1e-9, ns_submit * 1e-9);
    const double total_time = seconds_since(all_images_start);
    const uint32_t done = completed;
    printf("\n%u images in %.2f s (%.2f GBps) (%.1f fps)\n", done, total_time,
           (double)px * sizeof(pixel_t) * done / total_time / 1e9, done / total_time);
    if (time_waiting_for_images < 10)
        printf("Total time waiting for images to appear: %.0f ms\n", time_waiting_for_images * 1000);
    else
        printf("Total time waiting for images to appear: %.2f s\n", time_waiting_for_images);
    pipe.reset();
    destroy_all();
    // a run that stopped on an error must not look like a clean one to the calling service
    return had_error ? 1 : 0;
