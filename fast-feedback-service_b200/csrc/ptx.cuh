// ptx.cuh -- thin inline-PTX wrappers for the sm_100a features the kernels use:
// mbarrier, TMA tiled loads (cp.async.bulk.tensor), proxy fences.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace ffsb {
namespace ptx {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}

// Make mbarrier initialisation visible to the async (TMA) proxy.
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

// Order prior generic-proxy accesses to shared memory before later async-proxy ones
// (a stage that was just read with ld.shared is about to be refilled by TMA).
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
    return ok != 0;
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    // a transaction-count mismatch would otherwise hang the GPU: trap after ~seconds instead
    unsigned spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins > (1u << 24)) __trap();
    }
}

// 3D tiled TMA load global -> shared, completion signalled on an mbarrier.
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar,
                                            int c0, int c1, int c2) {
    asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4, %5}], [%2];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}

__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar,
                                            int c0, int c1) {
    asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}

__device__ __forceinline__ void prefetch_tensormap(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}

// prmt.b32 in its default mode: selector nibble bit 3 replicates the sign bit of the selected
// byte over the destination byte (__byte_perm masks that bit off).
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
    return d;
}

__device__ __forceinline__ uint4 lds128(const void* p) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "r"(smem_u32(p))
                 : "memory");
    return v;
}

// 8-byte shared-memory load from a 32-bit shared address
__device__ __forceinline__ uint4 lds128_s(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}

__device__ __forceinline__ uint2 lds64_s(uint32_t addr) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}

// One lane of a converged warp (elect.sync): ptxas recognises the single-thread region and
// feeds UTMALDG's uniform operands directly, where `lane == 0` costs a lane-serialising loop.
__device__ __forceinline__ bool elect_one() {
    uint32_t p;
    asm volatile(
      "{\n\t"
      ".reg .pred q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, q;\n\t"
      "}"
      : "=r"(p));
    return p != 0;
}

// Predicated shared-memory accesses: the compiler turns `if (p) { several stores }` into a
// divergent branch with a reconvergence barrier; a queue push is cheaper as predicated
// instructions on the straight-line path.
__device__ __forceinline__ void sts128_if(bool p, uint32_t addr, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
    asm volatile(
      "{\n\t"
      ".reg .pred q;\n\t"
      "setp.ne.u32 q, %0, 0;\n\t"
      "@q st.shared.v4.u32 [%1], {%2, %3, %4, %5};\n\t"
      "}" ::"r"((uint32_t)p),
      "r"(addr), "r"(x), "r"(y), "r"(z), "r"(w)
      : "memory");
}

// One queue record under one predicate: 4-byte meta word + 8 window sums (32 bytes).
__device__ __forceinline__ void push_group_if(uint32_t pbit, uint32_t a_meta, uint32_t meta, uint32_t a_sums,
                                              const uint32_t (&h)[8]) {
    asm volatile(
      "{\n\t"
      ".reg .pred q;\n\t"
      "setp.ne.u32 q, %0, 0;\n\t"
      "@q st.shared.u32 [%1], %2;\n\t"
      "@q st.shared.v4.u32 [%3], {%4, %5, %6, %7};\n\t"
      "@q st.shared.v4.u32 [%3+16], {%8, %9, %10, %11};\n\t"
      "}" ::"r"(pbit),
      "r"(a_meta), "r"(meta), "r"(a_sums), "r"(h[0]), "r"(h[1]), "r"(h[2]), "r"(h[3]), "r"(h[4]), "r"(h[5]),
      "r"(h[6]), "r"(h[7])
      : "memory");
}

__device__ __forceinline__ void sts32_if(bool p, uint32_t addr, uint32_t x) {
    asm volatile(
      "{\n\t"
      ".reg .pred q;\n\t"
      "setp.ne.u32 q, %0, 0;\n\t"
      "@q st.shared.u32 [%1], %2;\n\t"
      "}" ::"r"((uint32_t)p),
      "r"(addr), "r"(x)
      : "memory");
}

}  // namespace ptx
}  // namespace ffsb
