// Bulk asynchronous copies (the TMA engine: cp.async.bulk, SASS UBLKCP) tracked by shared-memory mbarriers, and the
// FP64 tensor-pipe tile (mma.sync.m8n8k4.f64, SASS DMMA.8x8x4).  tcgen05 / TMEM have no FP64 kind, so these two are what
// Blackwell offers a double-precision contraction: the TMA engine moves the operand rows, DMMA retires 256 FMAs per issue.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace msckf {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
// makes the barrier initialisation (generic proxy) visible to the async proxy that will complete transactions on it
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// orders generic-proxy accesses (ordinary loads / stores) against async-proxy accesses (bulk copies) of the same locations
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

// one arrival that also announces `bytes` of pending bulk-copy traffic
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// Waits for the phase of the given parity.  try_wait suspends the thread for a hardware-defined time slice per attempt;
// a barrier that never completes (wrong byte count, lost copy) traps after a few seconds instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  unsigned done = 0;
  int spins = 0;
  while (true) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n"
        "selp.u32 %0, 1, 0, P1;\n"
        "}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (done) break;
    if (++spins > (1 << 22)) __trap();
  }
}
// global -> shared bulk copy: 16-byte aligned source and destination, size a multiple of 16 bytes
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src),
               "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

}  // namespace msckf
