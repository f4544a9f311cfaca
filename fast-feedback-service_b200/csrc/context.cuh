// context.cuh -- the ffs_ctx object behind the C ABI: device buffers, TMA descriptors,
// streams, the pinned ring of the streaming API.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/ffs_b200.h"

namespace ffsb {

constexpr int kMaxDim = 8192;  // queue entries pack col/row into 13 bits each

inline size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Per-spot accumulator written by the reduction kernel (see cc.cuh).
struct SpotAcc;

struct StreamSlot {
    cudaEvent_t done = nullptr;  // recorded after the slot's result D2H
    uint64_t frame_id = 0;
    int state = 0;               // 0 free, 1 acquired (host filling), 2 in flight
    int stream = 0;
    bool compressed = false;     // submitted through ffs_submit_compressed
    // The slot's whole per-frame device work, captured once and replayed (all pointers and
    // launch parameters of a ring slot are fixed): [0] dense-frame path, [1] compressed path.
    cudaGraphExec_t graph[2] = {nullptr, nullptr};
    uint32_t graph_kernels[2] = {0, 0};  // kernels inside each graph (launch accounting)
    uint32_t graph_blocks = 0;           // bitshuffle blocks the compressed graph was built for
    uint32_t uses[2] = {0, 0};
    cudaEvent_t t_ev[4] = {nullptr, nullptr, nullptr, nullptr};  // ffs_frame_timing: start, copied, kernels, done
};

}  // namespace ffsb

struct ffs_ctx {
    ffs_params p{};
    int sm_count = 0;
    std::string err;
    std::atomic<uint64_t> launches{0};

    // geometry
    size_t pitch = 0;          // bytes per device image row (multiple of 128)
    size_t frame_stride = 0;   // bytes per device frame
    size_t bits_pitch = 0;     // bytes per row of a 1-bit plane (multiple of 16)
    size_t bits_stride = 0;    // bytes per 1-bit plane
    int eff_min_count = 2;

    // device memory
    uint8_t* d_mask = nullptr;        // dense u8 mask
    uint8_t* d_valid_bits = nullptr;  // 1-bit validity plane
    uint8_t* d_info = nullptr;        // per group of 8 pixels: {validity byte, m_ub} (dispersion_fused.cuh)
    size_t info_pitch = 0;            // bytes per row of d_info (multiple of 16)
    double* d_gain = nullptr;         // optional gain map (ffs_set_gain), width*height
    double gain_min = 1.0;            // over the mask-valid pixels: scales the pruning bound
    double gain_max = 1.0;            // (extended first pass: bracket of the right-hand side)
    bool gain_brute = false;          // gain map with non-positive / non-finite entries: no pruning
    uint8_t* d_info4 = nullptr;       // per group of 8 pixels: {validity byte, m_ub | full << 6, m_lb, 0} (extended_stream.cuh)
    size_t info4_pitch = 0;
    uint8_t* d_mcount = nullptr;      // per pixel: valid pixels in its clipped 7x7 window (extended_stream.cuh)
    size_t m_pitch = 0;
    uint8_t* d_frames = nullptr;      // n_slots pitched frames
    // Everything a frame's threshold must find zeroed sits in ONE region per slot, cleared by one
    // memset per batch: [strong bit-plane | chunk counts | strong, cand, ext, bad counters]
    uint8_t* d_zero = nullptr;        // n_slots * zstride bytes
    size_t zstride = 0;               // bytes per slot region (multiple of 256)
    uint8_t* d_strong = nullptr;      // = d_zero: slot z's strong bit-plane at + z * zstride
    uint8_t* d_ext1 = nullptr;        // extended: first-pass dispersion bit-planes
    uint8_t* d_ext2 = nullptr;        // extended: eroded validity bit-planes
    unsigned int* d_strong_count = nullptr;  // n_slots (+ n_slots candidate counters behind)
    unsigned int* d_cand_count = nullptr;    // alias into d_strong_count
    unsigned int* d_ext_count = nullptr;     // alias: scratch counter of the extended first pass
    uint32_t* d_cand = nullptr;              // n_slots * cand_cap candidate pixels
    unsigned int cand_cap = 0;
    // trusted range (trusted.cuh): mask-valid pixels above trusted_max met by the kernels
    bool trusted_on = false;
    uint32_t tmax = 0xffffffffu, vlim = (1u << 24) - 1u;
    uint32_t* d_bad = nullptr;               // n_slots * bad_cap
    unsigned int* d_bad_count = nullptr;     // n_slots
    unsigned int bad_cap = 4096;
    unsigned long long* d_valid_count = nullptr;
    uint64_t valid_count = 0;

    // connected components / spots (cc.cuh), per slot
    uint32_t cap_strong = 0, cap_spots = 0;
    uint32_t* d_px_index = nullptr;   // n_slots * cap_strong : linear index of strong pixel
    uint32_t* d_px_value = nullptr;   // n_slots * cap_strong : raw pixel value
    uint32_t* d_px_parent = nullptr;  // n_slots * cap_strong : union-find parent (compact id)
    uint32_t* d_chunk_count = nullptr;  // n_slots * n_chunks
    int n_chunks = 0;
    ffsb::SpotAcc* d_acc = nullptr;    // n_slots * cap_strong accumulators (indexed by root id)
    ffs_spot* d_spots = nullptr;      // n_slots * cap_spots
    uint32_t* d_frame_counts = nullptr;  // n_slots * 8 counters (see cc.cuh)
    // pinned mirrors for the streaming API
    uint32_t* h_frame_counts = nullptr;  // n_slots * 8
    ffs_spot* h_spots = nullptr;         // n_slots * cap_spots
    uint8_t* h_frames = nullptr;         // n_slots dense pinned frames
    uint8_t* d_stage = nullptr;          // n_slots dense device frames (H2D target when pitch != row bytes)

    // compressed ingestion (bslz4_gpu.cuh), allocated on first use
    // Per slot, in the pinned host buffer and in the device buffer alike:
    //   [ ... unused ... | work order u32[n] | block table uint2[n] | Bslz4Dyn | chunk bytes ... ]
    //                                                                ^ table_cap - 32   ^ table_cap
    // The table is written BACKWARDS from the fixed position of its header, so that it ends where the
    // chunk starts and ONE copy uploads both (a second, 50 KB copy per frame cost 8.6 us of a 146 us
    // frame period), while the kernel finds the header -- and from n_blocks everything else -- at an
    // address that never changes (the slot's CUDA graph stays valid).
    size_t chunk_cap = 0;             // bytes per slot available to the chunk itself
    size_t table_cap = 0;             // bytes per slot in front of the chunk (header + table for max_blocks)
    size_t chunk_stride = 0;          // table_cap + chunk_cap
    uint32_t max_blocks = 0;
    uint8_t* h_chunks = nullptr;      // pinned, n_slots * chunk_stride
    uint8_t* d_chunks = nullptr;
    unsigned int* d_dec_status = nullptr;  // n_slots x {status, work counter}
    unsigned int* h_dec_status = nullptr;  // pinned

    // TMA descriptors
    CUtensorMap tm_img{};     // 3D: x, y, slot
    CUtensorMap tm_valid{};   // 2D validity bytes
    CUtensorMap tm_info{};    // 2D group info bytes
    CUtensorMap tm_info4{};   // 2D 4-byte group info (extended first pass)
    CUtensorMap tm_img8{};    // 3D image, 256-pixel box (fused kernel; == tm_img for 16-bit pixels)
    CUtensorMap tm_valid2{};  // 3D eroded validity (extended second pass): x, y, slot
    bool have_tm = false;

    // streams
    cudaStream_t stream = nullptr;  // default work stream of the slot API
    std::vector<cudaStream_t> streams;  // streaming API
    std::vector<ffsb::StreamSlot> slots;
    std::deque<int> inflight;  // submitted slots, oldest first
    std::mutex mu;
    std::condition_variable cv;
    int next_stream = 0;

    // per-kernel profiling (ffs_profile_*)
    bool prof_on = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events;
    size_t prof_used = 0;

    // sweep (3D) recording
    bool sweep_on = false;
    bool sweep_overflow = false;  // a recorded frame hit a capacity: ffs_sweep_finish refuses
    std::vector<uint64_t> sweep_ids;
    std::vector<std::vector<uint32_t>> sweep_index;  // per frame strong-pixel linear indices
    std::vector<std::vector<uint32_t>> sweep_value;
    std::vector<uint64_t> sweep_sums;      // per 3D component: exact sum I, (2x+1)I, (2y+1)I, (2z+1)I
    bool frame_timing = false;             // ffs_frame_timing: per-frame copy / kernel / post-copy times
    float last_timing[3] = {0, 0, 0};      // of the last ffs_wait frame, ms
    bool self_check = false;               // ffs_self_check: ffs_wait decides every frame a second time (api_selfcheck.inl)
    uint8_t* d_check = nullptr;            // [bit-plane | strong count, candidate count (0), differing bits]
    size_t check_bytes = 0;
    cudaStream_t check_stream = nullptr;
    uint32_t last_check[2] = {0, 0};       // {strong pixels of the exact pass, differing pixels} of the last ffs_wait frame
    bool keep_pixels = false;              // ffs_keep_pixel_lists
    bool keep_frame = false;
    std::vector<uint8_t> last_frame;       // dense pixels of the last ffs_wait frame (keep mode 2)
    std::vector<uint32_t> last_pixels;     // strong-pixel linear indices of the last ffs_wait frame
    std::vector<uint64_t> sweep_moments;   // per 3D component: bbox-relative sum I u^2, I v^2, I u v, I w^2
    std::vector<uint32_t> sweep_plane[2];  // {index, component rank} of the first / last slice
};
