// Which shared-memory words does tcgen05.mma read for operand A under a given descriptor?  One CTA, one MMA
// (kind::tf32, K = 8) with B = identity (K-major, no swizzle, N = 8), so D[m][n] = A(m, k = n) = the value stored at
// the address the hardware computed for A(m, k).  Shared memory word w holds the float (w & 2047) in pass 0 and
// (w >> 11) in pass 1 (both exact in tf32), so the two passes give the word index.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -o umma_probe umma_probe.cu
//   ./umma_probe <M 64|128> <a_major 0=K 1=MN> <layout 0=none 2=sw128> <lbo bytes> <sbo bytes> [start byte offset]
//   prints, for every TMEM lane that is not all zero, the 8 word indices read for k = 0..7.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void probe(int M, int a_major, int layout, int lbo, int sbo, int start, int pass, float* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tslot;
  const uint32_t base = (smem_u32(smem) + 1023u) & ~1023u;
  float* A = reinterpret_cast<float*>(smem + (base - smem_u32(smem)));        // 64 KB of "address" values
  float* B = A + 16384;                                                       // identity, K-major, no swizzle
  for (int w = threadIdx.x; w < 16384; w += blockDim.x) A[w] = (float)(pass ? (w >> 11) : (w & 2047));
  for (int i = threadIdx.x; i < 64; i += blockDim.x) {                        // B[n][k]: n row, k: half h = k/4
    const int n = i >> 3, k = i & 7;
    B[(k >> 2) * 32 + n * 4 + (k & 3)] = n == k ? 1.f : 0.f;                  // core matrix h at 128 B * h
  }
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 32;" ::"r"(smem_u32(&tslot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;");     // generic-proxy writes above -> async-proxy reads by the MMA
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tmem = tslot;
  if (threadIdx.x == 0) {
    const uint32_t aaddr = base + start, baddr = base + 65536;
    const uint64_t da = (uint64_t)((aaddr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) |
                        (1ull << 46) | ((uint64_t)layout << 61);
    const uint64_t db = (uint64_t)((baddr & 0x3FFFFu) >> 4) | ((uint64_t)(128 >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) |
                        (1ull << 46);
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_major << 15) | ((8u >> 3) << 17) |
                           ((uint32_t)(M >> 4) << 24);
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(0u));
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)));
  }
  uint32_t ok = 0;
  for (int spin = 0; spin < (1 << 24) && !ok; ++spin)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(&bar)));
  asm volatile("tcgen05.fence::after_thread_sync;");
  uint32_t v[8];
  const uint32_t ta = tmem + ((uint32_t)((threadIdx.x >> 5) * 32) << 16);
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(ta));
  asm volatile("tcgen05.wait::ld.sync.aligned;");
  for (int j = 0; j < 8; ++j) out[threadIdx.x * 8 + j] = ok ? __uint_as_float(v[j]) : -1.f;
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;" ::"r"(tmem));
}

int main(int argc, char** argv) {
  if (argc < 6) { printf("usage: M a_major layout lbo sbo [start]\n"); return 1; }
  const int M = atoi(argv[1]), a_major = atoi(argv[2]), layout = atoi(argv[3]), lbo = atoi(argv[4]), sbo = atoi(argv[5]);
  const int start = argc > 6 ? atoi(argv[6]) : 0;
  float *d, h[2][1024];
  cudaMalloc(&d, sizeof(float) * 1024);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 70 * 1024);
  for (int pass = 0; pass < 2; ++pass) {
    cudaMemset(d, 0, sizeof(float) * 1024);
    probe<<<1, 128, 70 * 1024>>>(M, a_major, layout, lbo, sbo, start, pass, d);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 2; }
    cudaMemcpy(h[pass], d, sizeof(float) * 1024, cudaMemcpyDeviceToHost);
  }
  printf("M=%d a_major=%d layout=%d lbo=%d sbo=%d start=%d: byte offsets of A(m, k=0..7) per TMEM lane\n", M, a_major, layout, lbo, sbo, start);
  for (int lane = 0; lane < 128; ++lane) {
    bool any = false;
    for (int j = 0; j < 8; ++j) any |= h[0][lane * 8 + j] != 0.f || h[1][lane * 8 + j] != 0.f;
    if (!any && lane != 0) continue;
    printf("lane %3d:", lane);
    for (int j = 0; j < 8; ++j) printf(" %6d", 4 * ((int)h[0][lane * 8 + j] + 2048 * (int)h[1][lane * 8 + j]));
    printf("\n");
  }
  return 0;
}
