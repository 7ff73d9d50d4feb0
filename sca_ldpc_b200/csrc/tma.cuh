// 1-D bulk-async copy (TMA, cp.async.bulk; SASS UBLKCP) + mbarrier helpers shared by the persistent check-node
// kernels of both decoders.  Included INSIDE namespace scaldpc.  Device only (the host model stages synchronously).
#pragma once
#ifndef SCALDPC_HOST_MODEL
__device__ __forceinline__ uint32_t sc_smem_u32(const void *ptr) { return (uint32_t)__cvta_generic_to_shared(ptr); }
__device__ __forceinline__ void sc_mbar_init(uint64_t *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(sc_smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void sc_mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sc_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void sc_mbar_wait(uint64_t *bar, uint32_t phase)
{
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(sc_smem_u32(bar)), "r"(phase)
            : "memory");
    } while (!ok);
}
// 1-D bulk async copy global -> shared, completion counted in bytes on `bar`.  16-byte aligned, size % 16 == 0.
__device__ __forceinline__ void sc_bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     sc_smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(sc_smem_u32(bar))
                 : "memory");
}
// Orders earlier generic-proxy accesses of shared memory before later async-proxy (TMA) writes to it.
__device__ __forceinline__ void sc_fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
#endif
