// ResNet stem convolution (7x7 / stride 2 / padding 3 on 3 channels) as an implicit GEMM on the sm_100a tensor cores.
//
// The reference runs torchvision's resnet50 (dl_trainer.py:187-196 picks the model, the stem is its conv1) through
// cuDNN.  Here the stem is computed on the 2x2 space-to-depth image X [N, Hp, Wp, 12] (Hp = H/2 + 3, Wp = W/2 + 3,
// two zero rows / columns in front and one behind; k_stem_s2d below writes it), where it is a 4x4 / stride 1
// convolution.  With q = (n*Hp + i)*Wp + j the flat position in the padded grid,
//
//     y[q, o] = sum_{a<4} sum_{b<4} sum_{c<12} X[q + a*Wp + b, c] * W4[o, c, a, b]
//
// for every q that is an output pixel (i < OH, j < OW); the other 3 of every Wp positions and 3 of every Hp rows
// are computed and dropped.  A tile is 128 consecutive q.
//
// No im2col, not even in shared memory.  X is kept in shared memory ONCE, as three planes of 16-byte quarters
// (plane t holds channels 4t..4t+3 of every position, 16 B apart).  In the un-swizzled K-major operand layout of
// tcgen05.mma a "core matrix" is 8 rows x 16 B stored contiguously, 8-row groups SBO apart, the two 16-byte halves
// of a K = 8 tf32 step LBO apart.  With SBO = 128 B and LBO = 16 B a descriptor that starts at position q + a*Wp + b
// of plane t therefore reads, for row r, channels 4t..4t+3 of positions q+r+a*Wp+b and q+r+a*Wp+b+1: the operand
// rows of consecutive output pixels overlap in shared memory exactly as the convolution windows overlap in the
// image.  One tile is 4 (a) x 3 (t) x 2 (b pairs) = 24 MMAs of M = 128, N = 64, K = 8; the weights are packed on the
// host side in that order.  (The first version staged each kernel row through 128-byte-swizzled TMA boxes over an
// overlapping-row tensor map: correct, 59 us, bound by the 4x-expanded L2 -> shared memory traffic; this one reads
// X from L2 1.7 times.)
//
// One persistent CTA per SM owns a contiguous range of tiles and walks it in chunks of 4 tiles (512 + 3*Wp + 3
// positions of X per chunk, double-buffered).  Warp-specialised:
//   warp 0   producer: 12 TMA boxes {4 floats, boxP positions} per chunk (tensor map dims {4, P, 3}, strides
//            {4, 48, 16} B: the plane split happens in the copy engine); one bulk copy of the packed weights.
//   warps 1-4  MMA issuers: one lane each issues the tcgen05.mma kind::tf32 instructions of every fourth tile into
//            its own TMEM accumulator (one thread can issue an MMA of this size only every ~70-130 cycles, see
//            tools/experiments/umma_rate.cu); tcgen05.commit publishes the accumulator and releases the chunk buffer.
//   warp 5   allocates / frees 256 TMEM columns (four 64-column accumulators).
//   warps 6-13  two groups of 4 epilogue warps (alternate tiles), in a group one warp per TMEM lane quarter: tcgen05.ld the 128 x 64 fp32 tile, transpose through padded shared
//            memory, store whole 128-byte lines of the NHWC output (the valid rows of a tile are contiguous in y);
//            optionally accumulate the per-channel sums of y for the BatchNorm that follows.
// Every mbarrier wait is bounded (2 s on %globaltimer): on expiry the kernel records a code and traps.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "../include/gk_b200.h"

namespace {

constexpr int ST_M = 128;                          // flat padded positions per tile
constexpr int ST_N = 64;                           // output channels
constexpr int ST_TAPS = 4;                         // kernel rows (and columns) of the 4x4 form
constexpr int ST_CHUNK = 4;                        // tiles per chunk of X
constexpr int ST_NMMA = ST_TAPS * 3 * 2;           // K = 8 steps per tile
constexpr int ST_B_MMA = 2 * ST_N * 16;            // weight bytes per step: 2 halves x 64 rows x 16 B
constexpr int ST_B_BYTES = ST_NMMA * ST_B_MMA;     // 48 KB
constexpr int ST_EPI_PITCH = 36;                   // floats; 32 + 4 keeps the float4 phases conflict-free
constexpr int ST_EPI_GROUPS = 2;                   // epilogue warp groups working on alternate tiles (a tile's epilogue
                                                   // is ~3000 cycles of latency for a 4-warp group)
constexpr int ST_EPI_WARPS = 4 * ST_EPI_GROUPS;    // per group one warp per TMEM lane quarter
constexpr int ST_EPI_BYTES = ST_EPI_WARPS * 32 * ST_EPI_PITCH * 4;
constexpr int ST_NI = 4;                           // MMA-issuing warps = TMEM accumulators
constexpr int ST_W_ALLOC = 1 + ST_NI;              // warp 0 producer, 1..ST_NI issuers, then TMEM allocator, then 4 epilogue
constexpr int ST_W_EPI = 2 + ST_NI;
constexpr int ST_THREADS = 32 * (ST_W_EPI + ST_EPI_WARPS);
constexpr int WG_THREADS = 32 * (ST_W_EPI + 4);    // the weight-gradient kernel has one light epilogue: 4 warps
constexpr int ST_TMEM_COLS = 256;
static_assert(ST_NI % ST_EPI_GROUPS == 0, "an accumulator must always be drained by the same epilogue group");
constexpr int ST_FIXED = ST_B_BYTES + ST_EPI_BYTES + 1024;  // barriers + statistics; + 2 chunk buffers of 3 planes

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

// D fp32 (1 << 4), A and B tf32 (2 << 7, 2 << 10), both K-major, N >> 3 at bit 17, M >> 4 at bit 24.
constexpr uint32_t ST_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(ST_N >> 3) << 17) | ((uint32_t)(ST_M >> 4) << 24);

// tcgen05.mma kind::tf32 with the shared-memory descriptors given as (lo, hi) words: lo = start address >> 4 |
// LBO >> 4 << 16, hi = SBO >> 4 | version 1 | layout type << 29.  Stepping an operand is then one 32-bit add on lo
// (an issuing thread's instruction stream is one dependent chain: what stands between two MMAs costs cycles).
__device__ __forceinline__ void mma_tf32_w(uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                           uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "mov.b64 da, {%1, %2};\n\t"
      "mov.b64 db, {%3, %4};\n\t"
      "setp.ne.b32 p, %6, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], da, db, %5, p;\n\t}"
      ::"r"(d_tmem), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ uint32_t desc_lo(uint32_t saddr, uint32_t lbo) { return ((saddr & 0x3FFFFu) >> 4) | ((lbo >> 4) << 16); }
__device__ __forceinline__ uint32_t desc_hi(uint32_t sbo, uint32_t layout) { return (sbo >> 4) | (1u << 14) | (layout << 29); }

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

__device__ __forceinline__ void bulk_load_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

struct StemGeom {
  int nimg, Hp, Wp, OH, OW, ntiles;
  int boxP;                                          // positions per TMA box; a chunk buffer plane holds 4 boxes
};

template <bool STATS>
__global__ void __launch_bounds__(ST_THREADS, 1)
k_stem_fwd_tc(const __grid_constant__ CUtensorMap mapX, const float* __restrict__ wpk, float* __restrict__ out,
              float* __restrict__ stats, const StemGeom g, unsigned int* __restrict__ err) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t plane = (uint32_t)g.boxP * 4u * 16u;            // bytes per plane (boxP is a multiple of 8)
  const uint32_t sB = base;
  const uint32_t sE = sB + ST_B_BYTES;
  const uint32_t sBar = sE + ST_EPI_BYTES;
  const uint32_t sX = sBar + 1024u;                              // 2 buffers x 3 planes
  float* const sstat = reinterpret_cast<float*>(smem_raw + (sBar + 256u - raw));   // 64 sums + 64 sums of squares
  auto full = [&](uint32_t s) { return sBar + 8u * s; };
  auto empty = [&](uint32_t s) { return sBar + 16u + 8u * s; };
  const uint32_t bfull = sBar + 32u;
  auto tfull = [&](uint32_t s) { return sBar + 40u + 8u * s; };
  auto tempty = [&](uint32_t s) { return sBar + 40u + 8u * ST_NI + 8u * s; };
  const uint32_t tslot = sBar + 40u + 16u * ST_NI;
  float* const epi = reinterpret_cast<float*>(smem_raw + (sE - raw));
  volatile uint32_t* const tslot_p = reinterpret_cast<volatile uint32_t*>(smem_raw + (tslot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // this CTA's contiguous range of tiles
  const int per = g.ntiles / (int)gridDim.x, extra = g.ntiles % (int)gridDim.x;
  const int bid = (int)blockIdx.x;
  const int t0 = bid * per + (bid < extra ? bid : extra);
  const int nt = per + (bid < extra ? 1 : 0);
  const int nchunk = (nt + ST_CHUNK - 1) / ST_CHUNK;

  if (warp == 0 && lane == 0) tma_prefetch_desc(&mapX);
  if (threadIdx.x < 2 * ST_N) sstat[threadIdx.x] = 0.f;
  if (warp == 1 && lane == 0) {
    for (uint32_t s = 0; s < 2; ++s) {
      mbar_init(full(s), 1);
      mbar_init(empty(s), ST_NI);                   // every issuer releases the chunk buffer
    }
    for (uint32_t s = 0; s < ST_NI; ++s) {
      mbar_init(tfull(s), 1);
      mbar_init(tempty(s), 4);                      // one arrival per warp of the epilogue group that owns the tile
    }
    mbar_init(bfull, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == ST_W_ALLOC) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tslot), "r"(ST_TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot_p;

  if (warp == 0) {
    if (lane == 0) {                                // ---- producer
      mbar_expect_tx(bfull, ST_B_BYTES);
      bulk_load_1d(sB, wpk, ST_B_BYTES, bfull);
      for (int c = 0; c < nchunk; ++c) {
        const uint32_t buf = c & 1, ph = (c >> 1) & 1;
        mbar_wait(empty(buf), ph ^ 1u, err, 1);
        mbar_expect_tx(full(buf), 3u * plane);
        const int q0 = (t0 + c * ST_CHUNK) * ST_M;
        const uint32_t dst = sX + buf * 3u * plane;
        for (int t = 0; t < 3; ++t)
          for (int k = 0; k < 4; ++k)
            tma_load_3d(dst + t * plane + (uint32_t)(k * g.boxP) * 16u, &mapX, full(buf), 0, q0 + k * g.boxP, t);
      }
    }
  } else if (warp <= ST_NI) {
    if (lane == 0) {                                // ---- MMA issuers: issuer `me` owns accumulator `me` and every ST_NI-th tile
      // (one thread can issue an MMA of this size only every ~100 cycles - the descriptor arithmetic and the issue
      // itself are a dependent instruction chain - so several warps, each on its own scheduler, issue in parallel)
      const uint32_t me = (uint32_t)warp - 1u;
      mbar_wait(bfull, 0, err, 2);
      const uint32_t hi_plain = desc_hi(128u, 0u);  // both operands: K-major, no swizzle, 8-row groups 128 B apart
      const uint32_t d = tmem + me * ST_N;
      uint32_t use = 0;
      int it = 0;
      for (int c = 0; c < nchunk; ++c) {
        const uint32_t buf = c & 1, ph = (c >> 1) & 1;
        mbar_wait(full(buf), ph, err, 4);
        tc_fence_after();
        const uint32_t xb = sX + buf * 3u * plane;
        const int ntc = nt - c * ST_CHUNK < ST_CHUNK ? nt - c * ST_CHUNK : ST_CHUNK;
        for (int tl = 0; tl < ntc; ++tl, ++it) {
          if ((uint32_t)it % ST_NI != me) continue;
          mbar_wait(tempty(me), (use & 1u) ^ 1u, err, 3);  // the epilogue has drained this accumulator
          ++use;
          tc_fence_after();
          uint32_t b_lo = desc_lo(sB, ST_N * 16u);
#pragma unroll 1
          for (int a = 0; a < ST_TAPS; ++a) {
            const uint32_t a_lo = desc_lo(xb + (uint32_t)(tl * ST_M + a * g.Wp) * 16u, 16u);
#pragma unroll
            for (int t = 0; t < 3; ++t) {
#pragma unroll
              for (int bp = 0; bp < 2; ++bp) {
                mma_tf32_w(d, a_lo + (uint32_t)t * (plane >> 4) + (uint32_t)bp * 2u, hi_plain, b_lo, hi_plain, ST_IDESC,
                           (a | t | bp) != 0 ? 1u : 0u);
                b_lo += ST_B_MMA >> 4;
              }
            }
          }
          mma_commit(tfull(me));
        }
        mma_commit(empty(buf));                     // this issuer's MMAs on the chunk buffer are done (or it had none)
      }
    }
  } else if (warp >= ST_W_EPI) {                    // ---- epilogue: warp = TMEM lane quarter q (lanes 32q .. 32q+31)
    // These four warps are a serial instruction stream each and the kernel is sensitive to every instruction in
    // their loop: 32-bit index arithmetic (the host checks the ranges), one shuffle per row.
    const int q = warp & 3;                         // hardware rule: a warp reads TMEM lanes 32 (warp % 4) .. + 31
    const int grp = (warp - ST_W_EPI) >> 2;         // tiles it = grp, grp + ST_EPI_GROUPS, ...
    float* const stg = epi + (warp - ST_W_EPI) * 32 * ST_EPI_PITCH;
    const uint32_t pl = (uint32_t)(g.Hp * g.Wp), Wp = (uint32_t)g.Wp;
    unsigned long long outb_u = reinterpret_cast<unsigned long long>(out) + (unsigned)(lane & 7) * 16u;
    asm volatile("" : "+l"(outb_u));                // keep it in registers (else re-derived from the parameter bank per store)
    char* const outb = reinterpret_cast<char*>(outb_u);                    // this lane's 16 bytes of a 128-byte line
    float cs[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}}, cq[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
    for (int it = grp; it < nt; it += ST_EPI_GROUPS) {
      const uint32_t as = (uint32_t)it % ST_NI, aph = ((uint32_t)it / ST_NI) & 1u;
      // where this lane's row goes: position -> (n, i, j); 0xffffffff for the positions that are not output pixels
      const uint32_t p = (uint32_t)(t0 + it) * ST_M + (uint32_t)(q * 32 + lane);
      const uint32_t n = p / pl, rem = p - n * pl;
      const uint32_t oh = rem / Wp, ow = rem - oh * Wp;
      const uint32_t off = (n < (uint32_t)g.nimg && oh < (uint32_t)g.OH && ow < (uint32_t)g.OW)     // byte offset of the row
                               ? ((n * (uint32_t)g.OH + oh) * (uint32_t)g.OW + ow) * (ST_N * 4u) : 0xffffffffu;
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
          const uint32_t o = __shfl_sync(0xffffffffu, off, row);
          if (o != 0xffffffffu) {
            *reinterpret_cast<uint4*>(outb + o + half * 128) = x;
            if (STATS) {  // per-channel sum and sum of squares of the output pixels, for the BatchNorm that follows
              const float f0 = __uint_as_float(x.x), f1 = __uint_as_float(x.y), f2 = __uint_as_float(x.z), f3 = __uint_as_float(x.w);
              cs[half][0] += f0; cs[half][1] += f1; cs[half][2] += f2; cs[half][3] += f3;
              cq[half][0] = fmaf(f0, f0, cq[half][0]); cq[half][1] = fmaf(f1, f1, cq[half][1]);
              cq[half][2] = fmaf(f2, f2, cq[half][2]); cq[half][3] = fmaf(f3, f3, cq[half][3]);
            }
          }
        }
        __syncwarp();
      }
    }
    if (STATS) {                                    // lanes l, l+8, l+16, l+24 hold the same four channels
#pragma unroll
      for (int half = 0; half < 2; ++half)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float a = cs[half][j], b = cq[half][j];
          a += __shfl_xor_sync(0xffffffffu, a, 8);  b += __shfl_xor_sync(0xffffffffu, b, 8);
          a += __shfl_xor_sync(0xffffffffu, a, 16); b += __shfl_xor_sync(0xffffffffu, b, 16);
          if (lane < 8) {
            atomicAdd(sstat + half * 32 + lane * 4 + j, a);
            atomicAdd(sstat + ST_N + half * 32 + lane * 4 + j, b);
          }
        }
      asm volatile("bar.sync 1, %0;" ::"n"(ST_EPI_WARPS * 32) : "memory");      // the epilogue warps only
      const int et = (int)threadIdx.x - ST_W_EPI * 32;
      if (et < 2 * ST_N) stats[(size_t)bid * 2 * ST_N + et] = sstat[et];
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == ST_W_ALLOC) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(ST_TMEM_COLS) : "memory");
  }
}

__device__ __forceinline__ float to_tf32(float v) {   // round to nearest: the MMA itself would truncate
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  return __uint_as_float(r);
}

// x [N, H, W, 3] (NHWC) -> X [N, Hp, Wp, 12], X[n, i, j, (dy*2+dx)*3 + c] = x[n, 2(i-2)+dy, 2(j-2)+dx, c], zero in the
// two leading and one trailing rows / columns.  One thread per (position, dy): six consecutive floats in, six out,
// both 8-byte aligned.
// NCHW = true reads the planar [N, 3, H, W] batch as a data loader delivers it (per (c, dy) the two dx values are one
// aligned float2; no separate NCHW -> NHWC conversion of the input is needed then).
template <bool NCHW>
__global__ void __launch_bounds__(256) k_stem_s2d(const float* __restrict__ x, float* __restrict__ X, int nimg, int H, int W,
                                                  long long nhalf) {
  const uint32_t Hp = H / 2 + 3, Wp = W / 2 + 3;   // positions < 2^31 (host check): 32-bit index arithmetic
  for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < (uint32_t)nhalf; g += gridDim.x * blockDim.x) {
    const int dy = (int)(g & 1u);
    const uint32_t pos = g >> 1;
    const uint32_t ni = pos / Wp;
    const int j = (int)(pos - ni * Wp);
    const uint32_t nn = ni / Hp;
    const int i = (int)(ni - nn * Hp), n = (int)nn;
    float2 v0 = make_float2(0.f, 0.f), v1 = v0, v2 = v0;
    if (i >= 2 && i < Hp - 1 && j >= 2 && j < Wp - 1) {
      if (NCHW) {
        const size_t plane = (size_t)H * W;
        const float* r = x + (size_t)n * 3 * plane + (size_t)(2 * (i - 2) + dy) * W + 2 * (j - 2);
        const float2 c0 = __ldg(reinterpret_cast<const float2*>(r)), c1 = __ldg(reinterpret_cast<const float2*>(r + plane)),
                     c2 = __ldg(reinterpret_cast<const float2*>(r + 2 * plane));
        v0 = make_float2(c0.x, c1.x);               // dx = 0: c0 c1 c2, dx = 1: c0 c1 c2
        v1 = make_float2(c2.x, c0.y);
        v2 = make_float2(c1.y, c2.y);
      } else {
        const float2* r = reinterpret_cast<const float2*>(x + (((long long)n * H + 2 * (i - 2) + dy) * W + 2 * (j - 2)) * 3);
        v0 = __ldg(r);
        v1 = __ldg(r + 1);
        v2 = __ldg(r + 2);
      }
      v0.x = to_tf32(v0.x); v0.y = to_tf32(v0.y);
      v1.x = to_tf32(v1.x); v1.y = to_tf32(v1.y);
      v2.x = to_tf32(v2.x); v2.y = to_tf32(v2.y);
    }
    float2* o = reinterpret_cast<float2*>(X + (size_t)g * 6);
    o[0] = v0;
    o[1] = v1;
    o[2] = v2;
  }
}

// weight [64, 3, 7, 7] -> the 24 K = 8 steps of the kernel above, each [half 2][row group 8][row 8][4 floats]:
// step (a, t, bp), half kc, row o = 8*og + r, float e  <-  W4[o, 4t + e, a, 2bp + kc], where W4 is the 4x4 form
// W4[o, (dy*2+dx)*3 + ch, a, b] = w[o, ch, 2a+dy-1, 2b+dx-1] (zero where an index is -1).
__global__ void k_stem_pack_w(const float* __restrict__ w, float* __restrict__ wpk) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= ST_NMMA * 2 * ST_N * 4) return;
  const int e = g & 3, r = (g >> 2) & 7, og = (g >> 5) & 7, kc = (g >> 8) & 1, m = g >> 9;
  const int bp = m & 1, t = (m >> 1) % 3, a = m / 6;
  const int o = og * 8 + r, c12 = 4 * t + e, s = c12 / 3, ch = c12 % 3, dy = s >> 1, dx = s & 1;
  const int u = 2 * a + dy - 1, v = 2 * (2 * bp + kc) + dx - 1;
  wpk[g] = (u >= 0 && v >= 0) ? to_tf32(w[((o * 3 + ch) * 7 + u) * 7 + v]) : 0.f;
}

// ------------------------------------------------------------------------------------------------ weight gradient
// gW4[o, c, a, b] = sum over output pixels q of gY[q, o] * X[q + a*Wp + b, c]: a GEMM whose reduction index is the
// pixel, so both operands are "MN-major" (the pixel index is the slow one in memory).  Measured on a B200
// (tools/experiments/umma_probe.cu): with kind::tf32 an MN-major operand is read only under layout type 1, the
// 128-byte swizzle with 32-byte atoms (every other layout type yields zeros): 4 K rows x 128 B per atom, the 32-byte
// chunk index XORed with the K row, MN atoms LBO apart, K atoms SBO apart.  TMA's SWIZZLE_128B_ATOM_32B writes that.
//   A (M = 64 channels o, K = 8 pixels): a row of gY, [OW pixels][64 floats], as two column halves [OW][32 floats];
//     a K step is 8 rows = 1024 B, LBO = one half, SBO = 512.
//   B (N = 48 = 4 column shifts b x 12 channels, K = 8 pixels): row j of the operand is the 48 floats of X that start
//     at position (input row, j) - consecutive rows overlap in memory (48 B apart), which an overlapping-row tensor map
//     expresses; they land as two boxes [OW][32 floats] (the upper half of the second is out of bounds = zeros).
// Per K step 4 MMAs (one per kernel row a) of M = 64, N = 48 accumulate into 4 x 48 = 192 TMEM columns; row m of an
// M = 64 accumulator lives in lane (m % 16) + 32 (m / 16).  Every CTA owns a contiguous range of output rows (n, oh),
// keeps the expanded input rows it needs in a 5-slot ring (each is loaded once and used by up to four output rows),
// double-buffers the gY rows, and finally writes its 64 x 192 partial sums; k_stem_wgrad_fold adds the partials and
// folds the 4x4 form back to [64, 3, 7, 7].
constexpr int WG_SLOTS = 5;
constexpr int WG_GST = 2;
constexpr int WG_COLS = 192;
constexpr int WG_TMEM_COLS = 256;
constexpr uint32_t WG_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((48u >> 3) << 17) | ((64u >> 4) << 24);

struct WgGeom {
  int nimg, Hp, Wp, OH, OW, nrows;
  int dbg;                                           // 1: dump all 128 TMEM lanes (part has 128 rows per CTA)
};

__global__ void __launch_bounds__(WG_THREADS, 1)
k_stem_wgrad_tc(const __grid_constant__ CUtensorMap mapG, const __grid_constant__ CUtensorMap mapX,
                float* __restrict__ part, const WgGeom g, unsigned int* __restrict__ err) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t half = (uint32_t)g.OW * 128u;                   // one [OW][32 floats] box (OW % 8 == 0: 1024-multiple)
  const uint32_t sG = base;
  const uint32_t sX = sG + WG_GST * 2u * half;
  const uint32_t sBar = sX + WG_SLOTS * 2u * half;
  auto gfull = [&](uint32_t s) { return sBar + 8u * s; };
  auto gempty = [&](uint32_t s) { return sBar + 16u + 8u * s; };
  auto xfull = [&](uint32_t s) { return sBar + 32u + 8u * s; };
  auto xempty = [&](uint32_t s) { return sBar + 72u + 8u * s; };
  const uint32_t tfull = sBar + 112u;
  const uint32_t tslot = sBar + 120u;
  volatile uint32_t* const tslot_p = reinterpret_cast<volatile uint32_t*>(smem_raw + (tslot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int per = g.nrows / (int)gridDim.x, extra = g.nrows % (int)gridDim.x;
  const int bid = (int)blockIdx.x;
  const int r0 = bid * per + (bid < extra ? bid : extra);
  const int r1 = r0 + per + (bid < extra ? 1 : 0);
  const int nks = g.OW >> 3;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&mapG);
    tma_prefetch_desc(&mapX);
  }
  if (warp == 1 && lane == 0) {
    for (uint32_t s = 0; s < WG_GST; ++s) {
      mbar_init(gfull(s), 1);
      mbar_init(gempty(s), ST_TAPS);                // every issuer releases the gY row
    }
    for (uint32_t s = 0; s < WG_SLOTS; ++s) {
      mbar_init(xfull(s), 1);
      mbar_init(xempty(s), 1);
    }
    mbar_init(tfull, ST_TAPS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == ST_W_ALLOC) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tslot), "r"(WG_TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot_p;

  if (warp == 0) {
    if (lane == 0) {                                // ---- producer
      uint32_t seq = 0;
      long long lastI = -1;
      int rc = 0;
      for (int R = r0; R < r1; ++R, ++rc) {
        const int n = R / g.OH, oh = R - n * g.OH;
        const long long I0 = (long long)n * g.Hp + oh;
        const uint32_t gs = rc % WG_GST;            // gY first: the issuers release X slots only after they see it
        mbar_wait(gempty(gs), ((rc / WG_GST) & 1u) ^ 1u, err, 12);
        mbar_expect_tx(gfull(gs), 2u * half);
        tma_load_2d(sG + gs * 2u * half, &mapG, gfull(gs), 0, R * g.OW);
        tma_load_2d(sG + gs * 2u * half + half, &mapG, gfull(gs), 32, R * g.OW);
        for (long long I = (I0 > lastI + 1 ? I0 : lastI + 1); I <= I0 + 3; ++I, ++seq) {
          const uint32_t s = seq % WG_SLOTS;
          mbar_wait(xempty(s), ((seq / WG_SLOTS) & 1u) ^ 1u, err, 11);
          mbar_expect_tx(xfull(s), 2u * half);
          tma_load_2d(sX + s * 2u * half, &mapX, xfull(s), 0, (int)(I * g.Wp));
          tma_load_2d(sX + s * 2u * half + half, &mapX, xfull(s), 32, (int)(I * g.Wp));
        }
        lastI = I0 + 3;
      }
    }
  } else if (warp <= ST_TAPS) {
    if (lane == 0) {                                // ---- MMA issuers: issuer a owns kernel row a and its 48 TMEM columns
      const uint32_t a = (uint32_t)warp - 1u;
      const uint32_t hi = desc_hi(512u, 1u);
      const uint32_t d = tmem + a * 48u;
      uint32_t seq = 0, pending = 0xffffffffu;
      long long lastI = -1;
      int rc = 0;
      for (int R = r0; R < r1; ++R, ++rc) {
        const int n = R / g.OH, oh = R - n * g.OH;
        const long long I0 = (long long)n * g.Hp + oh;
        const long long first = I0 > lastI + 1 ? I0 : lastI + 1;
        const uint32_t cnt = (uint32_t)(I0 + 4 - first);
        const uint32_t mine = seq + cnt - 4u + a;             // the load that brought input row I0 + a
        seq += cnt;
        lastI = I0 + 3;
        const uint32_t gs = rc % WG_GST;
        mbar_wait(gfull(gs), (rc / WG_GST) & 1u, err, 14);
        // gY row R is only loaded once every issuer has finished output row R - 2, and the other users of the slot
        // freed here read it at output rows <= R - 2: now it is safe to hand it back (after this thread's own MMAs).
        // This must precede the wait for this row's X data: at an image boundary the producer needs these slots.
        if (pending != 0xffffffffu) mma_commit(xempty(pending));
        mbar_wait(xfull(mine % WG_SLOTS), (mine / WG_SLOTS) & 1u, err, 13);
        tc_fence_after();
        uint32_t a_lo = desc_lo(sG + gs * 2u * half, half);
        uint32_t x_lo = desc_lo(sX + (mine % WG_SLOTS) * 2u * half, half);
#pragma unroll 2
        for (int ks = 0; ks < nks; ++ks) {
          mma_tf32_w(d, a_lo, hi, x_lo, hi, WG_IDESC, (rc | ks) != 0 ? 1u : 0u);
          a_lo += 1024u >> 4;
          x_lo += 1024u >> 4;
        }
        mma_commit(gempty(gs));
        // input row I0 is not needed after this output row; at the end of an image neither are the other three.  The
        // slot is handed back at the start of the next output row (see above); after the last row nothing is reloaded.
        const bool slide = R + 1 < r1 && (R + 1) / g.OH == n;
        pending = (a == 0 || !slide) ? mine % WG_SLOTS : 0xffffffffu;
      }
      mma_commit(tfull);
    }
  } else if (warp >= ST_W_EPI) {                    // ---- epilogue: the partial sums of this CTA
    const int q = warp & 3;
    mbar_wait(tfull, 0, err, 15);
    tc_fence_after();
    float* dst = g.dbg ? part + ((size_t)bid * 128 + 32 * q + lane) * WG_COLS
                       : part + ((size_t)bid * ST_N + 16 * q + (lane & 15)) * WG_COLS;
#pragma unroll 1
    for (int c = 0; c < WG_COLS / 32; ++c) {
      uint32_t v[32];
      tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)c * 32u, v);
      tmem_wait_ld();
      if (lane < 16 || g.dbg) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(dst + c * 32 + j * 4) = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      }
    }
    tc_fence_before();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == ST_W_ALLOC) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(WG_TMEM_COLS) : "memory");
  }
}

// gw[o, ch, u, v] = sum over CTAs of part[cta][o][a*48 + b*12 + c], 2a + dy - 1 = u, 2b + dx - 1 = v,
// c = (dy*2 + dx)*3 + ch.  One block per output channel o: thread (column, quarter j) adds every fourth partial
// (coalesced 768-byte rows, fixed order: deterministic), the quarters meet in shared memory, then 147 threads pick
// their column.  (One thread per output element walking all partials serially took 15 us.)
constexpr int WG_FOLD_SPLIT = 4;
__global__ void __launch_bounds__(WG_COLS * WG_FOLD_SPLIT) k_stem_wgrad_fold(const float* __restrict__ part, int nparts,
                                                                            float* __restrict__ gw) {
  __shared__ float red[WG_FOLD_SPLIT][WG_COLS];
  const int o = blockIdx.x, col = threadIdx.x % WG_COLS, j = threadIdx.x / WG_COLS;
  const float* p = part + (size_t)o * WG_COLS + col;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  int k = j;
  for (; k + 3 * WG_FOLD_SPLIT < nparts; k += 4 * WG_FOLD_SPLIT) {
    s0 += __ldg(p + (size_t)(k) * ST_N * WG_COLS);
    s1 += __ldg(p + (size_t)(k + WG_FOLD_SPLIT) * ST_N * WG_COLS);
    s2 += __ldg(p + (size_t)(k + 2 * WG_FOLD_SPLIT) * ST_N * WG_COLS);
    s3 += __ldg(p + (size_t)(k + 3 * WG_FOLD_SPLIT) * ST_N * WG_COLS);
  }
  for (; k < nparts; k += WG_FOLD_SPLIT) s0 += __ldg(p + (size_t)k * ST_N * WG_COLS);
  red[j][col] = (s0 + s1) + (s2 + s3);
  __syncthreads();
  const int i = threadIdx.x;
  if (i < 147) {
    const int v = i % 7, u = (i / 7) % 7, ch = i / 49;
    const int dy = (u + 1) & 1, a = (u + 1) >> 1, dx = (v + 1) & 1, b = (v + 1) >> 1;
    const int c = a * 48 + b * 12 + (dy * 2 + dx) * 3 + ch;
    gw[o * 147 + i] = (red[0][c] + red[1][c]) + (red[2][c] + red[3][c]);
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

int stem_box(int Wp) { return ((ST_CHUNK * ST_M + 3 * Wp + 3 + 3) / 4 + 7) & ~7; }

}  // namespace

extern "C" {

// 1 when this driver can encode the tensor maps the kernels use (cuTensorMapEncodeTiled is exported and accepts the
// un-swizzled and the 128-byte / 32-byte-atom swizzled forms); nothing is launched, `probe` is any 16-byte aligned
// device pointer (it is not dereferenced).
int gk_stem_driver_ok(const void* probe) {
  EncodeTiledFn enc = encode_fn();
  if (!enc || probe == nullptr || (reinterpret_cast<uintptr_t>(probe) & 15)) return 0;
  CUtensorMap m;
  const cuuint32_t es[3] = {1, 1, 1};
  {
    const cuuint64_t dims[3] = {4, 1024, 3};
    const cuuint64_t strides[2] = {48, 16};
    const cuuint32_t box[3] = {4, 256, 1};
    if (enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(probe), dims, strides, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return 0;
  }
  {
    const cuuint64_t dims[2] = {48, 1024};
    const cuuint64_t strides[1] = {48};
    const cuuint32_t box[2] = {32, 112};
    if (enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(probe), dims, strides, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return 0;
  }
  return 1;
}

// 1 when gk_stem_fwd covers an H x W image (even sizes; the chunk of X must fit four 256-position TMA boxes).
int gk_stem_supported(int H, int W) {
  return H >= 2 && W >= 2 && H % 2 == 0 && W % 2 == 0 && stem_box(W / 2 + 3) <= 256;
}

// x [nimg, H, W, 3] fp32 NHWC (nchw = 0) or [nimg, 3, H, W] (nchw = 1)  ->  X [nimg, H/2+3, W/2+3, 12] fp32 (values
// rounded to tf32).
int gk_stem_s2d_layout(const void* x, void* X, int nimg, int H, int W, int nchw, void* stream);
int gk_stem_s2d(const void* x, void* X, int nimg, int H, int W, void* stream) {
  return gk_stem_s2d_layout(x, X, nimg, H, W, 0, stream);
}
int gk_stem_s2d_layout(const void* x, void* X, int nimg, int H, int W, int nchw, void* stream) {
  if (nimg <= 0 || H % 2 || W % 2) return -2;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(X)) & 7) return -2;
  const long long nhalf = (long long)nimg * (H / 2 + 3) * (W / 2 + 3) * 2;
  if (nhalf >= (1LL << 32) - (1LL << 24)) return -2;
  const int threads = 256;
  long long blocks = (nhalf + threads - 1) / threads;
  if (blocks > 148 * 32) blocks = 148 * 32;
  if (nchw)
    k_stem_s2d<true><<<(int)blocks, threads, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const float*>(x), static_cast<float*>(X), nimg, H, W, nhalf);
  else
    k_stem_s2d<false><<<(int)blocks, threads, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const float*>(x), static_cast<float*>(X), nimg, H, W, nhalf);
  return (int)cudaGetLastError();
}

// w [64, 3, 7, 7] fp32  ->  wpk [12288] fp32 in the kernel's operand order (values rounded to tf32).
int gk_stem_pack_w(const void* w, void* wpk, void* stream) {
  k_stem_pack_w<<<(ST_NMMA * 2 * ST_N * 4 + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const float*>(w), static_cast<float*>(wpk));
  return (int)cudaGetLastError();
}

// X: the padded space-to-depth image [nimg, OH+3, OW+3, 12] fp32, wpk: gk_stem_pack_w's output,
// y: [nimg, OH, OW, 64] fp32 (NHWC).  err (optional, device uint32) receives the code of a wait that timed out.
// Returns a cudaError_t, or -2 when the shape is not covered, -3 when the driver cannot encode the tensor map.
int gk_stem_fwd_stats(const void* X, const void* wpk, void* y, void* stats, int nimg, int OH, int OW, void* err,
                      void* stream);

int gk_stem_fwd(const void* X, const void* wpk, void* y, int nimg, int OH, int OW, void* err, void* stream) {
  return gk_stem_fwd_stats(X, wpk, y, nullptr, nimg, OH, OW, err, stream);
}

// Floats of `stats` gk_stem_fwd_stats fills: one row of 64 sums + 64 sums of squares of y per CTA (rows of CTAs
// that had no tile are zero); gk_stem_stats_rows() rows.
int gk_stem_stats_rows(void) { return stem_sm_count(); }

// gk_stem_fwd that also leaves the per-channel sum and sum of squares of y over all output pixels in `stats`
// ([gk_stem_stats_rows()][128] floats, partial per CTA): the batch statistics of the BatchNorm that follows, for
// free (gk_bn_pool_fwd_sums consumes them).
int gk_stem_fwd_stats(const void* X, const void* wpk, void* y, void* stats, int nimg, int OH, int OW, void* err,
                      void* stream) {
  if (nimg <= 0 || OH <= 0 || OW <= 0) return -2;
  const int Hp = OH + 3, Wp = OW + 3;
  const long long P = (long long)nimg * Hp * Wp;
  const long long rows = P - 3LL * Wp - 3;           // one past the last position that is an output pixel
  const int boxP = stem_box(Wp);
  if (rows <= 0 || P > (1LL << 31) - 2048 || boxP > 256) return -2;
  if ((long long)nimg * OH * OW * ST_N * 4 >= (1LL << 32) - 1) return -2;  // 32-bit byte offsets in the epilogue
  if ((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(wpk) | reinterpret_cast<uintptr_t>(y)) & 15) return -2;
  EncodeTiledFn enc = encode_fn();
  if (!enc) return -3;
  CUtensorMap mapX;
  {
    const cuuint64_t dims[3] = {4, (cuuint64_t)P, 3};                       // {float in quarter, position, quarter}
    const cuuint64_t strides[2] = {12 * sizeof(float), 4 * sizeof(float)};
    const cuuint32_t box[3] = {4, (cuuint32_t)boxP, 1};
    const cuuint32_t es[3] = {1, 1, 1};
    if (enc(&mapX, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(X), dims, strides, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return -3;
  }
  const int smem = ST_FIXED + 2 * 3 * boxP * 4 * 16 + 1024;
  {
    cudaError_t e = stats != nullptr
                        ? cudaFuncSetAttribute(k_stem_fwd_tc<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)
                        : cudaFuncSetAttribute(k_stem_fwd_tc<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
  }
  StemGeom g;
  g.nimg = nimg; g.Hp = Hp; g.Wp = Wp; g.OH = OH; g.OW = OW; g.boxP = boxP;
  g.ntiles = (int)((rows + ST_M - 1) / ST_M);
  const int sms = stem_sm_count();
  const int grid = g.ntiles < sms ? g.ntiles : sms;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (stats != nullptr && grid < sms) {              // rows of the CTAs that do not exist
    cudaError_t e = cudaMemsetAsync(static_cast<float*>(stats) + (size_t)grid * 2 * ST_N, 0,
                                    (size_t)(sms - grid) * 2 * ST_N * sizeof(float), s);
    if (e != cudaSuccess) return (int)e;
  }
  if (stats != nullptr)
    k_stem_fwd_tc<true><<<grid, ST_THREADS, smem, s>>>(mapX, static_cast<const float*>(wpk), static_cast<float*>(y),
                                                       static_cast<float*>(stats), g, static_cast<unsigned int*>(err));
  else
    k_stem_fwd_tc<false><<<grid, ST_THREADS, smem, s>>>(mapX, static_cast<const float*>(wpk), static_cast<float*>(y),
                                                        nullptr, g, static_cast<unsigned int*>(err));
  return (int)cudaGetLastError();
}

static int wgrad_smem(int OW) { return (WG_GST + WG_SLOTS) * 2 * OW * 128 + 256 + 1024; }

// 1 when gk_stem_wgrad covers an output of OH x OW pixels.
int gk_stem_wgrad_supported(int OH, int OW) {
  return OH > 0 && OW >= 8 && OW % 8 == 0 && OW <= 256 && wgrad_smem(OW) <= 227 * 1024;
}

// Floats of workspace gk_stem_wgrad needs (one 64 x 192 partial per CTA).
long long gk_stem_wgrad_workspace(void) { return (long long)stem_sm_count() * ST_N * WG_COLS; }

// gy [nimg, OH, OW, 64] fp32 (NHWC), X: the padded space-to-depth image of the forward pass  ->  gw [64, 3, 7, 7].
int gk_stem_wgrad(const void* gy, const void* X, void* gw, void* work, int nimg, int OH, int OW, void* err, void* stream) {
  if (nimg <= 0 || !gk_stem_wgrad_supported(OH, OW)) return -2;
  const int Hp = OH + 3, Wp = OW + 3;
  const long long P = (long long)nimg * Hp * Wp;
  const long long nrows = (long long)nimg * OH;
  if (P > (1LL << 31) - 2048 || nrows * OW > (1LL << 31) - 1) return -2;
  if ((reinterpret_cast<uintptr_t>(gy) | reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(work)) & 15) return -2;
  EncodeTiledFn enc = encode_fn();
  if (!enc) return -3;
  WgGeom g;
  g.nimg = nimg; g.Hp = Hp; g.Wp = Wp; g.OH = OH; g.OW = OW; g.nrows = (int)nrows;
  g.dbg = getenv("GK_STEM_WG_DEBUG") != nullptr;
  CUtensorMap mapG, mapX;
  {
    const cuuint64_t dims[2] = {(cuuint64_t)ST_N, (cuuint64_t)(nrows * OW)};
    const cuuint64_t strides[1] = {ST_N * sizeof(float)};
    const cuuint32_t box[2] = {32, (cuuint32_t)OW};
    const cuuint32_t es[2] = {1, 1};
    if (enc(&mapG, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(gy), dims, strides, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return -3;
  }
  {
    // rows = positions, 48 floats each, 48 B apart: consecutive rows overlap (the window of a position is the next
    // 4 positions' channels); the last rows' windows would run past the image: they are cut off (zeros)
    const cuuint64_t dims[2] = {48, (cuuint64_t)(P - 3)};
    const cuuint64_t strides[1] = {12 * sizeof(float)};
    const cuuint32_t box[2] = {32, (cuuint32_t)OW};
    const cuuint32_t es[2] = {1, 1};
    if (enc(&mapX, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(X), dims, strides, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return -3;
  }
  const int smem = wgrad_smem(OW);
  {
    cudaError_t e = cudaFuncSetAttribute(k_stem_wgrad_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
  }
  const int sms = stem_sm_count();
  const int grid = nrows < sms ? (int)nrows : sms;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  k_stem_wgrad_tc<<<grid, WG_THREADS, smem, s>>>(mapG, mapX, static_cast<float*>(work), g, static_cast<unsigned int*>(err));
  k_stem_wgrad_fold<<<ST_N, WG_COLS * WG_FOLD_SPLIT, 0, s>>>(static_cast<const float*>(work), grid, static_cast<float*>(gw));
  return (int)cudaGetLastError();
}

}  // extern "C"
