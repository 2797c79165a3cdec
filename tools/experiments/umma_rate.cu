// How many cycles does one tcgen05.mma kind::tf32 (K = 8) take, as a function of M, N and of how many independent
// accumulators the instruction stream rotates over?  One CTA, one issuing thread, 512 MMAs back to back, timed with
// clock64 from the first issue to the arrival of the final tcgen05.commit.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o umma_rate umma_rate.cu && ./umma_rate
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void rate(int M, int N, int nacc, int a_major, int mode, int step, int kind, int nw, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tslot;
  const uint32_t base = (smem_u32(smem) + 1023u) & ~1023u;
  for (int w = threadIdx.x; w < 32768; w += blockDim.x) reinterpret_cast<float*>(smem + (base - smem_u32(smem)))[w] = 1.f;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(nw));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tslot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tmem = tslot;
  if ((threadIdx.x & 31) == 0 && (int)(threadIdx.x >> 5) < nw) {
    const uint32_t wid = threadIdx.x >> 5;
    const uint32_t layout = a_major ? 1u : 0u;
    // kind 0: K-major no swizzle, canonical (LBO 128, SBO 256); kind 1: the stem forward's overlapping form (LBO 16, SBO 128,
    // start 48 B off alignment); kind 2: K-major 128-byte swizzle (SBO 1024)
    const uint32_t lbo = a_major ? 4096u : (kind == 1 ? 16u : 128u);
    const uint32_t sbo = a_major ? 512u : (kind == 1 ? 128u : (kind == 2 ? 1024u : 256u));
    const uint32_t lay = a_major ? 1u : (kind == 2 ? 2u : 0u);
    const uint32_t a_lo0 = (((base + (kind == 1 ? 48u : 0u)) & 0x3FFFFu) >> 4) | ((lbo >> 4) << 16);
    const uint32_t b_lo = (((base + 65536u) & 0x3FFFFu) >> 4) | ((a_major ? 4096u >> 4 : 128u >> 4) << 16);
    const uint32_t hi = (sbo >> 4) | (1u << 14) | (lay << 29);
    const uint32_t hib = ((a_major ? 512u : 256u) >> 4) | (1u << 14) | (layout << 29);
    uint32_t a_lo = a_lo0;
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_major << 15) | ((uint32_t)a_major << 16) |
                           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    const long long t0 = clock64();
    uint32_t elected = 1;
    if (mode == 1) asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(elected));
#pragma unroll 8
    for (int i = 0; i < 512; ++i) {
      const uint32_t d = tmem + (wid * (uint32_t)nacc + (uint32_t)(i & (nacc - 1))) * (uint32_t)N;
      if (elected) asm volatile(
          "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tmov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %6};\n\tsetp.ne.b32 p, %5, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::tf32 [%0], da, db, %4, p;\n\t}" ::"r"(d), "r"(a_lo), "r"(hi), "r"(b_lo), "r"(idesc), "r"(1u), "r"(hib));
      a_lo = a_lo0 + ((uint32_t)((i + 1) * step) & 1023u);
    }
    const long long t1 = clock64();
    if (elected) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)));
    uint32_t ok = 0;
    while (!ok)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(ok) : "r"(smem_u32(&bar)));
    const long long t2 = clock64();
    if (threadIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

int main() {
  long long *d, h[2];
  cudaMalloc(&d, 16);
  cudaFuncSetAttribute(rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  printf("512 MMAs (kind::tf32, K = 8) per issuing warp, each warp into its own TMEM accumulator; clock64 of warp 0 from its first issue to the arrival of every warp's commit\n");
  const int Ns[] = {8, 16, 48, 64, 128, 256};
  const int mode = 0;
  for (int cfg = 0; cfg < 5; cfg += 1)
    for (int M = 64; M <= 128; M += 64)
      for (int ni = 0; ni < 6; ++ni)
        for (int nw = 1; nw <= 4; nw *= 2) {
          const int nacc = 1;
          if (Ns[ni] * nw > 512) continue;
          const int major = cfg == 4, kind = cfg < 3 ? cfg : (cfg == 3 ? 1 : 0), step = cfg == 0 ? 0 : (cfg == 3 ? 3 : 64);
          const int N = Ns[ni];

          for (int rep = 0; rep < 2; ++rep) {
            rate<<<1, 128, 200 * 1024>>>(M, N, nacc, major, mode, step, kind, nw, d);
            if (cudaDeviceSynchronize() != cudaSuccess) { printf("error\n"); return 1; }
          }
          cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
          printf("cfg %d (major %s kind %d step %d) M %3d N %3d issuing warps %d : %6.1f cycles per MMA per warp, %6.1f per MMA overall\n", cfg, major ? "MN" : "K", kind, step * 16, M, N, nw, h[1] / 512.0, h[1] / 512.0 / nw);
        }
  return 0;
}
