// ffs_b200.cu -- single translation unit of libffs_b200.so (C ABI in include/ffs_b200.h).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared ...
#include "context.cuh"
#include "mask.cuh"
#include "threshold.cuh"
#include "dispersion_fused.cuh"
#include "extended_stream.cuh"
#include "extended.cuh"
#include "trusted.cuh"
#include "cc.cuh"
#include "cc3d.cuh"
#include "bslz4_gpu.cuh"

#include <algorithm>
#include <cmath>
#include <tuple>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace ffsb {

static thread_local std::string g_create_error;

#define FFS_CK(ctx, call)                                                               \
    do {                                                                                \
        cudaError_t e__ = (call);                                                       \
        if (e__ != cudaSuccess) {                                                       \
            (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e__);           \
            return FFS_ERR_CUDA;                                                        \
        }                                                                               \
    } while (0)

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
              cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// rank-2 or rank-3 tiled descriptor, no swizzle, zero fill out of bounds
static int make_tensor_map(ffs_ctx* c, CUtensorMap* map, CUtensorMapDataType dt, int rank,
                           void* base, const cuuint64_t* dims, const cuuint64_t* strides_bytes,
                           const cuuint32_t* box) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) {
        c->err = "cuTensorMapEncodeTiled not available from the driver";
        return FFS_ERR_CUDA;
    }
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(map, dt, (cuuint32_t)rank, base, dims, strides_bytes, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        c->err = "cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r);
        return FFS_ERR_CUDA;
    }
    return FFS_OK;
}

static inline cudaStream_t pick_stream(ffs_ctx* c, void* s) {
    return s ? reinterpret_cast<cudaStream_t>(s) : c->stream;
}

// Developer A/B knobs (FFS_TWO_KERNEL, FFS_FORCE_BRUTE, FFS_TILE_ROWS, ... -- the table is in
// DESIGN.md section 9) are read ONLY when FFS_DEV_KNOBS=1 is set as well, so a stray variable in a
// service environment cannot silently change (or slow down 20x) the product path.
static int env_int(const char* name, int dflt) {
    static const bool dev = [] {
        const char* d = getenv("FFS_DEV_KNOBS");
        return d && atoi(d) != 0;
    }();
    if (!dev) return dflt;
    const char* v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

}  // namespace ffsb

using namespace ffsb;

static int extended_setup(ffs_ctx* c);
static int extended_impl(ffs_ctx* c, int first_slot, int n, cudaStream_t st);
static int cc_setup(ffs_ctx* c);
static void cc_teardown(ffs_ctx* c);
static int stream_setup(ffs_ctx* c);
static void stream_teardown(ffs_ctx* c);
static void bslz4_teardown(ffs_ctx* c);
static int sweep_record(ffs_ctx* c, int slot, uint64_t frame_id, uint32_t n_strong);

#include "api_core.inl"
#include "api_ext.inl"
#include "api_cc.inl"
#include "api_selfcheck.inl"
#include "api_stream.inl"
#include "api_bslz4.inl"
#include "api_sweep.inl"
#include "api_multi.inl"
#include "api_shim.inl"
#include "api_noalloc.inl"
