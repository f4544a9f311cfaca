// h2d.cu -- the host->device ceiling the streaming path is measured against (SURVEY.md 8(d):
// "practical peak must be measured on the box with a pinned 1 GiB cudaMemcpyAsync loop per GPU
// and with all 8 concurrently").  One host thread per GPU, all started together; every variant
// is one plain cudaMemcpyAsync / cudaMemcpy2DAsync per copy.
//
//   1d_1gib      : 1 GiB contiguous pinned -> device, one stream                (the link's best case)
//   1d_frame     : frame-sized (W*H*bpp) contiguous copies, one stream
//   2d_frame     : frame-sized cudaMemcpy2DAsync, dense rows -> 128-byte-aligned pitch
//                  (what ffs_submit queues, csrc/api_stream.inl)
//   1d_frame_2s  : frame-sized contiguous copies alternating over two streams
//   2d_frame_2s  : the pitched copy alternating over two streams
//   *_w          : the same while one host thread per GPU keeps writing a ring of pinned frames
//                  (the producer of the real frame loop, spotfinder.cc:828-834)
//
// Output: one JSON object on stdout: per variant the per-GPU GB/s (CUDA events on the copy
// stream) and the aggregate GB/s over the common wall-clock window.
//
// build: nvcc -O2 -std=c++17 -cudart shared -o tools/ubench/h2d tools/ubench/h2d.cu -lpthread
// run  : tools/ubench/h2d --gpus 8 --width 4148 --height 4362 --bpp 2 --seconds 1.0
#include <cuda_runtime.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e_ = (x);                                                                   \
        if (e_ != cudaSuccess) {                                                                \
            fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_)); \
            exit(2);                                                                            \
        }                                                                                       \
    } while (0)

using clk = std::chrono::steady_clock;

struct Barrier {
    std::atomic<int> count{0};
    std::atomic<int> gen{0};
    int n;
    explicit Barrier(int n) : n(n) {}
    void wait() {
        const int g = gen.load();
        if (count.fetch_add(1) + 1 == n) {
            count = 0;
            gen++;
        } else {
            while (gen.load() == g) std::this_thread::yield();
        }
    }
};

struct Variant {
    const char* name;
    int kind;     // 0 1d_1gib, 1 1d_frame, 2 2d_frame
    int streams;  // 1 or 2
    bool writer;
};

int main(int argc, char** argv) {
    int gpus = 1, width = 4148, height = 4362, bpp = 2;
    double seconds = 1.0;
    for (int i = 1; i + 1 < argc; i += 2) {
        std::string a = argv[i];
        if (a == "--gpus") gpus = atoi(argv[i + 1]);
        else if (a == "--width") width = atoi(argv[i + 1]);
        else if (a == "--height") height = atoi(argv[i + 1]);
        else if (a == "--bpp") bpp = atoi(argv[i + 1]);
        else if (a == "--seconds") seconds = atof(argv[i + 1]);
    }
    int have = 0;
    CK(cudaGetDeviceCount(&have));
    if (gpus > have) gpus = have;
    const size_t row = (size_t)width * bpp, pitch = (row + 127) / 128 * 128;
    const size_t frame = row * height, dframe = pitch * height;
    const size_t gib = 1ull << 30;
    const int ring = 8;  // pinned frames per GPU (the product's ring depth)

    const Variant variants[] = {
        {"1d_1gib", 0, 1, false},     {"1d_frame", 1, 1, false},      {"2d_frame", 2, 1, false},
        {"1d_frame_2s", 1, 2, false}, {"2d_frame_2s", 2, 2, false},   {"1d_frame_w", 1, 1, true},
        {"2d_frame_w", 2, 1, true},   {"2d_frame_2s_w", 2, 2, true},
    };
    const int nv = (int)(sizeof variants / sizeof variants[0]);

    struct PerGpu {
        uint8_t *h_big = nullptr, *h_ring = nullptr, *h_wr = nullptr, *d_big = nullptr, *d_frames = nullptr;
        cudaStream_t st[2];
        cudaEvent_t e0, e1;
    };
    std::vector<PerGpu> G(gpus);
    for (int g = 0; g < gpus; ++g) {
        CK(cudaSetDevice(g));
        CK(cudaHostAlloc(&G[g].h_big, gib, cudaHostAllocDefault));
        CK(cudaHostAlloc(&G[g].h_ring, frame * ring, cudaHostAllocDefault));
        CK(cudaHostAlloc(&G[g].h_wr, frame * ring, cudaHostAllocDefault));
        memset(G[g].h_big, 1, gib);
        memset(G[g].h_ring, 2, frame * ring);
        memset(G[g].h_wr, 3, frame * ring);
        CK(cudaMalloc(&G[g].d_big, gib));
        CK(cudaMalloc(&G[g].d_frames, dframe * ring));
        for (auto& s : G[g].st) CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
        CK(cudaEventCreate(&G[g].e0));
        CK(cudaEventCreate(&G[g].e1));
    }

    printf("{\"gpus\": %d, \"width\": %d, \"height\": %d, \"bytes_per_pixel\": %d, \"frame_bytes\": %zu, "
           "\"device_pitch\": %zu, \"seconds_per_variant\": %.2f, \"variants\": {",
           gpus, width, height, bpp, frame, pitch, seconds);
    for (int v = 0; v < nv; ++v) {
        const Variant& V = variants[v];
        std::vector<double> gbs(gpus, 0.0), bytes(gpus, 0.0);
        std::vector<double> t_begin(gpus), t_end(gpus);
        Barrier bar(gpus);
        std::atomic<bool> stop_writers{false};
        std::vector<std::thread> writers;
        std::atomic<unsigned long long> written{0};
        if (V.writer)
            for (int g = 0; g < gpus; ++g)
                writers.emplace_back([&, g] {
                    int k = 0;
                    while (!stop_writers) {
                        memset(G[g].h_wr + (size_t)(k % ring) * frame, k & 0xff, frame);
                        written += frame;
                        ++k;
                    }
                });
        const auto origin = clk::now();
        std::vector<std::thread> th;
        for (int g = 0; g < gpus; ++g)
            th.emplace_back([&, g] {
                CK(cudaSetDevice(g));
                PerGpu& P = G[g];
                auto copy = [&](int k) {
                    cudaStream_t s = P.st[V.streams == 2 ? (k & 1) : 0];
                    if (V.kind == 0) CK(cudaMemcpyAsync(P.d_big, P.h_big, gib, cudaMemcpyHostToDevice, s));
                    else if (V.kind == 1)
                        CK(cudaMemcpyAsync(P.d_frames + (size_t)(k % ring) * dframe, P.h_ring + (size_t)(k % ring) * frame,
                                           frame, cudaMemcpyHostToDevice, s));
                    else
                        CK(cudaMemcpy2DAsync(P.d_frames + (size_t)(k % ring) * dframe, pitch,
                                             P.h_ring + (size_t)(k % ring) * frame, row, row, height,
                                             cudaMemcpyHostToDevice, s));
                };
                const size_t per = V.kind == 0 ? gib : frame;
                // warm-up + calibrate the copy count for ~`seconds`
                for (int k = 0; k < 4; ++k) copy(k);
                for (auto& s : P.st) CK(cudaStreamSynchronize(s));
                const int n = std::max(4, (int)(seconds * 40e9 / (double)per));
                bar.wait();
                t_begin[g] = std::chrono::duration<double>(clk::now() - origin).count();
                // events on stream 0; with two streams, stream 1 is fenced to it at both ends
                CK(cudaEventRecord(P.e0, P.st[0]));
                if (V.streams == 2) CK(cudaStreamWaitEvent(P.st[1], P.e0, 0));
                for (int k = 0; k < n; ++k) copy(k);
                if (V.streams == 2) {
                    cudaEvent_t j;
                    CK(cudaEventCreateWithFlags(&j, cudaEventDisableTiming));
                    CK(cudaEventRecord(j, P.st[1]));
                    CK(cudaStreamWaitEvent(P.st[0], j, 0));
                    CK(cudaEventRecord(P.e1, P.st[0]));
                    CK(cudaEventSynchronize(P.e1));
                    CK(cudaEventDestroy(j));
                } else {
                    CK(cudaEventRecord(P.e1, P.st[0]));
                    CK(cudaEventSynchronize(P.e1));
                }
                t_end[g] = std::chrono::duration<double>(clk::now() - origin).count();
                float ms = 0;
                CK(cudaEventElapsedTime(&ms, P.e0, P.e1));
                bytes[g] = (double)per * n;
                gbs[g] = bytes[g] / (ms * 1e-3) / 1e9;
            });
        for (auto& t : th) t.join();
        const unsigned long long wr = written.load();
        stop_writers = true;
        for (auto& t : writers) t.join();
        double b0 = 1e30, e1 = 0, total = 0;
        for (int g = 0; g < gpus; ++g) {
            b0 = std::min(b0, t_begin[g]);
            e1 = std::max(e1, t_end[g]);
            total += bytes[g];
        }
        printf("%s\"%s\": {\"per_gpu_gbs\": [", v ? ", " : "", V.name);
        for (int g = 0; g < gpus; ++g) printf("%s%.2f", g ? ", " : "", gbs[g]);
        printf("], \"aggregate_gbs\": %.2f", total / (e1 - b0) / 1e9);
        if (V.writer) printf(", \"host_writer_gbs\": %.2f", (double)wr / (e1 - b0) / 1e9);
        printf("}");
        fflush(stdout);
    }
    printf("}}\n");
    return 0;
}
