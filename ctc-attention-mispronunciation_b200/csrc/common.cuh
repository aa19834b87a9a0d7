// common.cuh -- shared helpers for the sm_100a kernels of libmddctc.so
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#include <algorithm>

#include "../../include/mddctc.h"

namespace mdd {

extern int64_t g_launch_count;   // defined in capi.cu

#define MDD_CHECK_LAUNCH()                                                            \
    do {                                                                              \
        ++mdd::g_launch_count;                                                        \
        if (cudaPeekAtLastError() != cudaSuccess) { cudaGetLastError(); return MDD_E_CUDA; } \
    } while (0)

constexpr int kNumSM = 148;   // B200 (grid sizing of the streaming kernels)
int sm_count();               // SMs of the current device (capi.cu)

__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

__device__ __forceinline__ float neg_inf() { return -CUDART_INF_F; }

// log(exp(a)+exp(b)) with the aten convention: max-shift, shift 0 when both are -inf.
__device__ __forceinline__ float lse2(float a, float b) {
    float m = fmaxf(a, b);
    if (m == -CUDART_INF_F) return -CUDART_INF_F;
    return m + __logf(__expf(a - m) + __expf(b - m));
}
__device__ __forceinline__ float lse3(float a, float b, float c) {
    float m = fmaxf(fmaxf(a, b), c);
    if (m == -CUDART_INF_F) return -CUDART_INF_F;
    return m + __logf(__expf(a - m) + __expf(b - m) + __expf(c - m));
}
// Accurate variants (libdevice expf/logf, ~1-2 ulp) for the log-space reference path.
__device__ __forceinline__ float lse2_acc(float a, float b) {
    float m = fmaxf(a, b);
    if (m == -CUDART_INF_F) return -CUDART_INF_F;
    return m + logf(expf(a - m) + expf(b - m));
}
__device__ __forceinline__ float lse3_acc(float a, float b, float c) {
    float m = fmaxf(fmaxf(a, b), c);
    if (m == -CUDART_INF_F) return -CUDART_INF_F;
    return m + logf(expf(a - m) + expf(b - m) + expf(c - m));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// ---------------------------------------------------------------------------------------------------------------
// mbarrier + 1-D bulk async copy (TMA engine, SASS UBLKCP): global -> shared, completion counted in bytes on an
// mbarrier.  Source, destination and size must be multiples of 16 B.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}
// The same wait for a warp that is AHEAD of its partner (helpers waiting for a buffer to come free): the thread is
// suspended by the hardware for up to `ns` at a time and woken by the phase completing, instead of polling -- a polling
// warp takes issue slots from the warp it is waiting for when the two share a scheduler.
__device__ __forceinline__ void mbar_wait_idle(uint64_t* bar, uint32_t parity, uint32_t ns = 4000) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAITI_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
        "@p bra DONEI_%=;\n"
        "bra WAITI_%=;\n"
        "DONEI_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity), "r"(ns)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// shared -> global bulk store (bulk_group completion)
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }
// ---------------------------------------------------------------------------------------------------------------
// Programmatic dependent launch: a kernel launched with launch_dependent() may be scheduled while its predecessor in
// the stream is still running (once every CTA of the predecessor has executed pdl_launch_dependents() or left); it must
// not touch anything the predecessor produces before pdl_wait(), which returns when the predecessor has completed and
// its writes are visible.  Both instructions are no-ops in a kernel launched the ordinary way.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool pdl_enabled();           // MDD_PDL != 0 (capi.cu)

template <typename... KArgs, typename... Args>
inline cudaError_t launch_dependent(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

// make generic-proxy shared-memory writes visible to the async proxy (before bulk_s2g)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

}  // namespace mdd
