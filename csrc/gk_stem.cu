// ResNet stem convolution (7x7 / stride 2 / padding 3 on 3 channels) as an implicit GEMM on the sm_100a tensor cores.
//
// The reference runs torchvision's resnet50 (dl_trainer.py:187-196 picks the model, the stem is its conv1) through
// cuDNN.  Here the stem is computed on the 2x2 space-to-depth image X [N, Hp, Wp, 12] (Hp = H/2 + 3, Wp = W/2 + 3,
// two zero rows / columns in front and one behind; see models/fuse.py), where it is a 4x4 / stride 1 convolution:
//
//     y[n, oh, ow, o] = sum_{a<4} sum_{k<48} X_flat[(p + a*Wp) * 12 + k] * Wk[o][a*48 + k],   p = (n*Hp + oh)*Wp + ow
//
// i.e. for every kernel row a the operand rows are 48 consecutive floats starting 12 floats apart: a matrix with
// OVERLAPPING rows (row pitch 48 B, row length 192 B).  A TMA tensor map describes exactly that (dims {48, P, 4},
// strides {4, 48, 48*Wp} bytes), so the im2col expansion happens in the copy engine: no gather kernel, no expanded
// copy in HBM.  The flat index p runs over the padded grid, so a 128-row tile is 128 consecutive p; the 3 of every
// Wp positions (and 3 of every Hp rows) that are not output pixels are computed and dropped by the epilogue.
//
// One persistent CTA per SM, warp-specialised:
//   warp 0   TMA producer: per kernel row two 128 x 32-float boxes (128-byte swizzle; the second box's upper half is
//            out of bounds in k and arrives as zeros) into a 4-deep ring; the 64 x 192 weight matrix once.
//   warp 1   one lane issues tcgen05.mma kind::tf32, M = 128, N = 64, K = 8, six per kernel row, accumulating the
//            four kernel rows in TMEM; tcgen05.commit releases ring slots and publishes the accumulator.
//   warp 2   allocates / frees 128 TMEM columns (two accumulators: the MMAs of tile t+1 overlap the epilogue of t).
//   warps 4-7  epilogue: tcgen05.ld the 128 x 64 fp32 tile (warp w owns lanes 32(w%4)..), transpose through padded
//            shared memory, store whole 128-byte lines of the NHWC output.
// Every mbarrier wait is bounded (2 s on %globaltimer): on expiry the kernel records a code and traps rather than
// spinning forever.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../include/gk_b200.h"

namespace {

constexpr int ST_M = 128;                          // flat padded positions per tile
constexpr int ST_N = 64;                           // output channels
constexpr int ST_TAPS = 4;                         // kernel rows of the 4x4 form
constexpr int ST_KROW = 48;                        // 4 positions x 12 channels per kernel row
constexpr int ST_STAGES = 4;
constexpr int ST_SLAB_A = ST_M * 128;              // 128 rows x 128 B
constexpr int ST_STAGE_A = 2 * ST_SLAB_A;          // k 0..31 and k 32..63 (48..63 zero)
constexpr int ST_SLAB_B = ST_N * 128;
constexpr int ST_B_BYTES = ST_TAPS * 2 * ST_SLAB_B;
constexpr int ST_EPI_PITCH = 36;                   // floats; 32 + 4 keeps the float4 phases conflict-free
constexpr int ST_EPI_BYTES = 4 * 32 * ST_EPI_PITCH * 4;
constexpr int ST_THREADS = 256;
constexpr int ST_TMEM_COLS = 128;
constexpr int ST_SMEM = ST_STAGES * ST_STAGE_A + ST_B_BYTES + ST_EPI_BYTES + 256 + 1024;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ uint32_t mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok;
}
__device__ __forceinline__ uint64_t gtime_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// Bounded wait: a protocol error must end the kernel, not hang the device.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, unsigned int* err, uint32_t code) {
  if (mbar_try_wait(bar, parity)) return;
  const uint64_t t0 = gtime_ns();
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 1023u) == 0u && gtime_ns() - t0 > 2000000000ull) {
      if (err) atomicExch(err, code);
      __threadfence_system();
      __trap();
    }
  }
}

__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(m) : "memory");
}
// Loads from this GPU's memory only, no multicast.
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(dst), "l"(m), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(m), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// K-major operand, 128-byte swizzle: rows 128 B apart, 8-row groups 1024 B apart (SBO), LBO unused (1), version 1.
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// D fp32 (1 << 4), A and B tf32 (2 << 7, 2 << 10), both K-major, N >> 3 at bit 17, M >> 4 at bit 24.
constexpr uint32_t ST_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(ST_N >> 3) << 17) | ((uint32_t)(ST_M >> 4) << 24);

__device__ __forceinline__ void mma_tf32(uint32_t d_tmem, uint64_t da, uint64_t db, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(da), "l"(db), "r"(ST_IDESC), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

struct StemGeom {
  int nimg, Hp, Wp, OH, OW, ntiles;
};

__global__ void __launch_bounds__(ST_THREADS, 1)
k_stem_fwd_tc(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
              float* __restrict__ out, const StemGeom g, unsigned int* __restrict__ err) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;     // the swizzle atoms repeat every 1024 B
  const uint32_t sA = base;
  const uint32_t sB = sA + ST_STAGES * ST_STAGE_A;
  const uint32_t sE = sB + ST_B_BYTES;
  const uint32_t sBar = sE + ST_EPI_BYTES;
  auto full = [&](uint32_t s) { return sBar + 8u * s; };
  auto empty = [&](uint32_t s) { return sBar + 8u * (ST_STAGES + s); };
  const uint32_t bfull = sBar + 8u * (2 * ST_STAGES);
  auto tfull = [&](uint32_t s) { return bfull + 8u + 8u * s; };
  auto tempty = [&](uint32_t s) { return bfull + 24u + 8u * s; };
  const uint32_t tslot = bfull + 40u;
  float* const epi = reinterpret_cast<float*>(smem_raw + (sE - raw));
  volatile uint32_t* const tslot_p = reinterpret_cast<volatile uint32_t*>(smem_raw + (tslot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&mapA);
    tma_prefetch_desc(&mapB);
  }
  if (warp == 1 && lane == 0) {
    for (uint32_t s = 0; s < ST_STAGES; ++s) {
      mbar_init(full(s), 1);
      mbar_init(empty(s), 1);
    }
    mbar_init(bfull, 1);
    for (uint32_t s = 0; s < 2; ++s) {
      mbar_init(tfull(s), 1);
      mbar_init(tempty(s), 4);                      // one arrival per epilogue warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tslot), "r"(ST_TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot_p;

  if (warp == 0) {
    if (lane == 0) {                                // ---- TMA producer
      mbar_expect_tx(bfull, ST_B_BYTES);
      for (int a = 0; a < ST_TAPS; ++a)
        for (int h = 0; h < 2; ++h)
          tma_load_2d(sB + (a * 2 + h) * ST_SLAB_B, &mapB, bfull, a * ST_KROW + h * 32, 0);
      uint32_t stage = 0, phase = 0;
      for (int tile = blockIdx.x; tile < g.ntiles; tile += gridDim.x) {
        const int p0 = tile * ST_M;
        for (int a = 0; a < ST_TAPS; ++a) {
          mbar_wait(empty(stage), phase ^ 1u, err, 1);
          mbar_expect_tx(full(stage), ST_STAGE_A);
          const uint32_t dst = sA + stage * ST_STAGE_A;
          tma_load_3d(dst, &mapA, full(stage), 0, p0, a);
          tma_load_3d(dst + ST_SLAB_A, &mapA, full(stage), 32, p0, a);
          if (++stage == ST_STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {                                // ---- MMA issuer
      mbar_wait(bfull, 0, err, 2);
      uint32_t stage = 0, phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < g.ntiles; tile += gridDim.x, ++it) {
        const uint32_t as = it & 1, aph = (it >> 1) & 1;
        mbar_wait(tempty(as), aph ^ 1u, err, 3);    // the epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d = tmem + as * ST_N;
        for (int a = 0; a < ST_TAPS; ++a) {
          mbar_wait(full(stage), phase, err, 4);
          tc_fence_after();
          const uint32_t a0 = sA + stage * ST_STAGE_A, b0 = sB + a * 2 * ST_SLAB_B;
#pragma unroll
          for (int kk = 0; kk < ST_KROW / 8; ++kk) {  // k 0..31 from the first slab, 32..47 from the second
            const uint32_t h = kk >> 2, ko = (kk & 3) * 32;
            mma_tf32(d, smem_desc_sw128(a0 + h * ST_SLAB_A + ko), smem_desc_sw128(b0 + h * ST_SLAB_B + ko),
                     (a | kk) != 0 ? 1u : 0u);
          }
          mma_commit(empty(stage));                 // slot free once these MMAs have read it
          if (++stage == ST_STAGES) { stage = 0; phase ^= 1u; }
        }
        mma_commit(tfull(as));
      }
    }
  } else if (warp >= 4) {                           // ---- epilogue
    const int q = warp & 3;                         // TMEM lanes 32q .. 32q+31
    float* const stg = epi + q * 32 * ST_EPI_PITCH;
    const long long plane = (long long)g.Hp * g.Wp;
    int it = 0;
    for (int tile = blockIdx.x; tile < g.ntiles; tile += gridDim.x, ++it) {
      const uint32_t as = it & 1, aph = (it >> 1) & 1;
      // where this lane's row goes: p -> (n, oh, ow); -1 for the padded positions
      const long long p = (long long)tile * ST_M + q * 32 + lane;
      const int n = (int)(p / plane);
      const int rem = (int)(p - (long long)n * plane);
      const int oh = rem / g.Wp, ow = rem - oh * g.Wp;
      const long long off = (n < g.nimg && oh < g.OH && ow < g.OW)
                                ? (((long long)n * g.OH + oh) * g.OW + ow) * ST_N : -1ll;
      mbar_wait(tfull(as), aph, err, 5);
      tc_fence_after();
      uint32_t v[64];
      const uint32_t ta = tmem + ((uint32_t)(q * 32) << 16) + as * ST_N;
      tmem_ld32(ta, v);
      tmem_ld32(ta + 32, v + 32);
      tmem_wait_ld();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty(as));       // the accumulator is in registers: MMAs may overwrite it
#pragma unroll
      for (int half = 0; half < 2; ++half) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(stg + lane * ST_EPI_PITCH + j * 4) =
              make_uint4(v[half * 32 + 4 * j], v[half * 32 + 4 * j + 1], v[half * 32 + 4 * j + 2], v[half * 32 + 4 * j + 3]);
        __syncwarp();
#pragma unroll
        for (int r = 0; r < 8; ++r) {               // 4 rows x 128 B per instruction
          const int row = r * 4 + (lane >> 3), ch = lane & 7;
          const uint4 x = *reinterpret_cast<const uint4*>(stg + row * ST_EPI_PITCH + ch * 4);
          const long long o = __shfl_sync(0xffffffffu, off, row);
          if (o >= 0) *reinterpret_cast<uint4*>(out + o + half * 32 + ch * 4) = x;
        }
        __syncwarp();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(ST_TMEM_COLS) : "memory");
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

int stem_sm_count() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0, n = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    sms = n > 0 ? n : 148;
  }
  return sms;
}

}  // namespace

extern "C" {

// x: the padded space-to-depth image [nimg, Hp, Wp, 12] fp32 (Hp = OH + 3, Wp = OW + 3), wk: [64, 192] fp32 with
// wk[o][a*48 + b*12 + c] the weight of kernel row a, column b, channel c; y: [nimg, OH, OW, 64] fp32 (NHWC).
// err (optional, device uint32) receives the code of a wait that timed out.  Returns a cudaError_t, or -2 when the
// shape is not covered, -3 when the driver cannot encode the tensor maps.
int gk_stem_fwd(const void* x, const void* wk, void* y, int nimg, int OH, int OW, void* err, void* stream) {
  if (nimg <= 0 || OH <= 0 || OW <= 0) return -2;
  const int Hp = OH + 3, Wp = OW + 3;
  const long long P = (long long)nimg * Hp * Wp;
  const long long rows = P - 3LL * Wp - 3;           // rows whose 4 x 48 floats lie inside x
  if (rows <= 0 || P > (1LL << 31) - ST_M) return -2;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(wk) | reinterpret_cast<uintptr_t>(y)) & 15) return -2;
  EncodeTiledFn enc = encode_fn();
  if (!enc) return -3;
  CUtensorMap mapA, mapB;
  {
    const cuuint64_t dims[3] = {(cuuint64_t)ST_KROW, (cuuint64_t)rows, (cuuint64_t)ST_TAPS};
    const cuuint64_t strides[2] = {12 * sizeof(float), (cuuint64_t)Wp * 12 * sizeof(float)};
    const cuuint32_t box[3] = {32, ST_M, 1};
    const cuuint32_t es[3] = {1, 1, 1};
    if (enc(&mapA, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(x), dims, strides, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return -3;
  }
  {
    const cuuint64_t dims[2] = {(cuuint64_t)(ST_TAPS * ST_KROW), (cuuint64_t)ST_N};
    const cuuint64_t strides[1] = {ST_TAPS * ST_KROW * sizeof(float)};
    const cuuint32_t box[2] = {32, ST_N};
    const cuuint32_t es[2] = {1, 1};
    if (enc(&mapB, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(wk), dims, strides, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return -3;
  }
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(k_stem_fwd_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, ST_SMEM);
    if (e != cudaSuccess) return (int)e;
    attr_set = true;
  }
  StemGeom g;
  g.nimg = nimg; g.Hp = Hp; g.Wp = Wp; g.OH = OH; g.OW = OW;
  g.ntiles = (int)((rows + ST_M - 1) / ST_M);
  const int sms = stem_sm_count();
  const int grid = g.ntiles < sms ? g.ntiles : sms;
  k_stem_fwd_tc<<<grid, ST_THREADS, ST_SMEM, static_cast<cudaStream_t>(stream)>>>(
      mapA, mapB, static_cast<float*>(y), g, static_cast<unsigned int*>(err));
  return (int)cudaGetLastError();
}

}  // extern "C"
