// mask.cuh -- once-per-dataset mask kernels: u8 mask <-> 1-bit validity plane, and the
// resolution mask (spotfinder/kernels/masking.cu:37-147).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ffsb {

// One thread packs 32 pixels of one row into one 32-bit word (bit x&31 of word x>>5).
// mask: dense width*height bytes.  bits: rows of `pitch_bytes` bytes.
__global__ void pack_mask_kernel(const uint8_t* __restrict__ mask, int width, int height,
                                 uint32_t* __restrict__ bits, size_t pitch_bytes,
                                 unsigned long long* valid_count) {
    const int words_per_row = (int)(pitch_bytes / 4);
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long total = (long long)words_per_row * height;
    unsigned cnt = 0;
    if (t < total) {
        const int row = (int)(t / words_per_row);
        const int wx = (int)(t % words_per_row);
        const uint8_t* src = mask + (size_t)row * width;
        uint32_t word = 0;
        const int x0 = wx * 32;
        for (int k = 0; k < 32; ++k) {
            const int x = x0 + k;
            if (x < width && src[x] != 0) word |= 1u << k;
        }
        bits[(size_t)row * words_per_row + wx] = word;
        cnt = __popc(word);
    }
    // block reduce the count
    for (int off = 16; off > 0; off >>= 1) cnt += __shfl_down_sync(0xffffffffu, cnt, off);
    if ((threadIdx.x & 31) == 0 && cnt && valid_count) atomicAdd(valid_count, (unsigned long long)cnt);
}

// bits (rows of pitch_bytes) -> dense u8 plane of 0/1.
__global__ void unpack_bits_kernel(const uint32_t* __restrict__ bits, size_t pitch_bytes,
                                   int width, int height, uint8_t* __restrict__ out) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long total = (long long)width * height;
    if (t >= total) return;
    const int row = (int)(t / width), x = (int)(t % width);
    const uint32_t w = bits[(size_t)row * (pitch_bytes / 4) + (x >> 5)];
    out[t] = (w >> (x & 31)) & 1u;
}

// Dense frame (rows of row_bytes) -> pitched frame (rows of pitch bytes), V = vector type of the
// widest width that divides row_bytes.  grid (ceil(row_vecs / 256), height).  The host -> device
// copy of a frame is ONE contiguous cudaMemcpyAsync (measured 12 % faster over PCIe than the
// pitched cudaMemcpy2DAsync, tools/ubench/h2d.cu); this pass restores the 16-byte-aligned row
// pitch TMA needs at HBM speed (2 x 36 MB per Eiger2 16M frame).
template <typename V>
__global__ void __launch_bounds__(256) repitch_kernel(const uint8_t* __restrict__ dense, size_t row_bytes,
                                                      uint8_t* __restrict__ pitched, size_t pitch) {
    const unsigned row_vecs = (unsigned)(row_bytes / sizeof(V));
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= row_vecs) return;
    const size_t y = blockIdx.y;
    reinterpret_cast<V*>(pitched + y * pitch)[i] = reinterpret_cast<const V*>(dense + y * row_bytes)[i];
}

// apply_resolution_mask, kernels/masking.cu:99-147 with get_distance_from_center (:37-53)
// and get_resolution (:64-73): float math; "0.5 * atanf()" is a double product narrowed
// back to float exactly as the reference writes it.
__global__ void resolution_mask_kernel(uint8_t* mask, int width, int height, float wavelength,
                                       float distance, float bcx, float bcy, float psx, float psy,
                                       float dmin, float dmax) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= width || y >= height) return;
    const size_t k = (size_t)y * width + x;
    if (mask[k] == 0) return;
    const float dx = (((float)x + 0.5f) - bcx) * psx;
    const float dy = (((float)y + 0.5f) - bcy) * psy;
    const float r = sqrtf(dx * dx + dy * dy);
    const float theta = (float)(0.5 * atanf(r / distance));
    const float d = wavelength / (2 * sinf(theta));
    uint8_t out = 1;
    if (dmin > 0 && d < dmin) out = 0;
    else if (dmax > 0 && d > dmax) out = 0;
    mask[k] = out;
}

}  // namespace ffsb
