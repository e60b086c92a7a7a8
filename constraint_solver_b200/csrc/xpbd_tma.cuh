// 1-D TMA bulk copy (cp.async.bulk, SASS: UBLKCP) + mbarrier helpers for staging the hull
// table into shared memory.  sm_90+ PTX; this library only targets sm_100a.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace xpbd {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(arrivals) : "memory");
}
// Makes the initialised barrier visible to the async proxy before a bulk copy signals it.
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// global -> shared bulk copy; src, dst 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void bulk_copy_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: returns false if the copy has not completed after ~2^20 probes (never observed;
// the caller then falls back to plain loads instead of hanging the GPU).
__device__ __forceinline__ bool mbar_wait_bounded(uint64_t* bar, uint32_t parity) {
    for (int i = 0; i < (1 << 20); ++i)
        if (mbar_try_wait(bar, parity)) return true;
    return false;
}

// Block-cooperative staging of `bytes` (multiple of 16) from global to shared memory with one
// bulk copy issued by thread 0.  `bar` is a shared-memory mbarrier owned by the caller.
// All threads of the block must call this; returns after the data is visible to every thread.
__device__ __forceinline__ void stage_bulk(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar,
                                           unsigned int* fallback_flag) {
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        fence_mbar_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(bar, bytes);
        bulk_copy_g2s(dst_smem, src_gmem, bytes, bar);
    }
    bool ok = mbar_wait_bounded(bar, 0);
    if (!__syncthreads_and(ok)) {
        // Fallback (diagnostic path): cooperative 16-byte loads. Flagged so tests can see it happened.
        const uint4* s = reinterpret_cast<const uint4*>(src_gmem);
        uint4* d = reinterpret_cast<uint4*>(dst_smem);
        for (uint32_t i = threadIdx.x; i < bytes / 16; i += blockDim.x) d[i] = s[i];
        if (threadIdx.x == 0 && fallback_flag) atomicAdd(fallback_flag, 1u);
        __syncthreads();
    }
}

}  // namespace xpbd
