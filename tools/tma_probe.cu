// tma_probe.cu -- minimal TMA experiments (developer tool, not part of the product).
#include "../fast-feedback-service_b200/csrc/ptx.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>
using namespace ffsb;
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
__global__ void probe(const __grid_constant__ CUtensorMap tm_img, const __grid_constant__ CUtensorMap tm_val, int mode, int cx, int cy, int vx, uint32_t* out) {
    extern __shared__ __align__(1024) uint8_t sm[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(sm + 8192);
    int lane = threadIdx.x;
    if (lane == 0) { ptx::mbar_init(bar, 1); ptx::fence_mbar_init(); }
    __syncwarp();
    if (lane == 0) {
        uint32_t tx = 0;
        if (mode & 1) tx += 3584;
        if (mode & 2) tx += 224;
        ptx::mbar_arrive_expect_tx(bar, tx);
        if (mode & 1) ptx::tma_load_3d(sm, &tm_img, bar, cx, cy, 0);
        if (mode & 2) ptx::tma_load_2d(sm + 3584, &tm_val, bar, vx, cy);
    }
    ptx::mbar_wait(bar, 0);
    uint32_t s = 0;
    for (int i = lane; i < (3584 + 256) / 4; i += 32) s += reinterpret_cast<uint32_t*>(sm)[i];
    out[lane] = s;
}
int main(int argc, char** argv) {
    int mode = atoi(argv[1]), cx = atoi(argv[2]), cy = atoi(argv[3]), vx = atoi(argv[4]);
    int W = argc > 5 ? atoi(argv[5]) : 64, H = argc > 6 ? atoi(argv[6]) : 64;
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    EncodeTiledFn fn = (EncodeTiledFn)p;
    size_t pitch = (W * 2 + 127) / 128 * 128, vp = ((W + 7) / 8 + 15) / 16 * 16;
    uint8_t *dimg, *dval; uint32_t* dout;
    cudaMalloc(&dimg, pitch * H); cudaMalloc(&dval, vp * H); cudaMalloc(&dout, 128);
    cudaMemset(dimg, 1, pitch * H); cudaMemset(dval, 0xff, vp * H);
    CUtensorMap ti, tv; cuuint32_t es[3] = {1, 1, 1};
    cuuint64_t d3[3] = {(cuuint64_t)W, (cuuint64_t)H, 1}, s3[2] = {pitch, pitch * H}; cuuint32_t b3[3] = {256, 7, 1};
    CUresult r1 = fn(&ti, CU_TENSOR_MAP_DATA_TYPE_UINT16, 3, dimg, d3, s3, b3, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    cuuint64_t d2[2] = {(cuuint64_t)(W + 7) / 8, (cuuint64_t)H}, s2[1] = {vp}; cuuint32_t b2[2] = {32, 7};
    CUresult r2 = fn(&tv, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, dval, d2, s2, b2, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode %d %d | ", (int)r1, (int)r2);
    probe<<<1, 32, 16384>>>(ti, tv, mode, cx, cy, vx, dout);
    cudaError_t e = cudaDeviceSynchronize();
    uint32_t h[32] = {}; cudaMemcpy(h, dout, 128, cudaMemcpyDeviceToHost);
    uint64_t tot = 0; for (int i = 0; i < 32; ++i) tot += h[i];
    printf("mode %d c(%d,%d) vx %d W %d H %d -> %s sum %llu\n", mode, cx, cy, vx, W, H, cudaGetErrorString(e), (unsigned long long)tot);
    return 0;
}
