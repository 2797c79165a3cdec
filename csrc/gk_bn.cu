// gk_bn.cu — fused training-mode batch norm for NHWC fp32 activations (sm_100a).
//
// Why it exists: on a B200 the ResNet-50 step of the benchmark spends 29 % of its kernel time in cuDNN's
// NHWC batch norm and another 9 % in the stand-alone ReLU / residual-add kernels around it
// (profiles/step_breakdown_resnet50_v1.txt) — all of it HBM-bound elementwise work.  The reference gets
// these ops from torch.nn; here y = relu(bn(x) [+ residual]) is ONE pass after the
// statistics pass, and the backward recomputes the ReLU mask instead of reading a saved activation:
//   forward : stats pass reads x (4 B/elem), apply pass reads x (+ residual) and writes y (8-12 B/elem)
//   backward: reduce pass reads dy, x (8 B/elem), dx pass reads dy, x and writes dx (12 B/elem)
// versus 20-32 B/elem forward and 32+ B/elem backward for the unfused sequence.
//
// Layout: x is [M = N*H*W rows][C channels], C innermost, C % 4 == 0.  A thread owns one quad of channels
// (float4) and walks rows; 16 or 32 consecutive threads cover consecutive quads (coalesced 256-512 B).
#include <cuda_runtime.h>
#include <stdint.h>

#include "../include/gk_b200.h"

#define BN_T 256

__device__ __forceinline__ float4 bn_ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void bn_st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ float4 f4_add(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float4 f4_fma(float4 a, float4 b, float4 c) {
  return make_float4(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y), fmaf(a.z, b.z, c.z), fmaf(a.w, b.w, c.w));
}

// What decides the ReLU mask of a saved forward output: the output itself or, with flag bit 2 (value 4), one
// byte per channel quad written by the forward kernel (bit j = output j of the quad was positive) — a sixteenth
// of the bytes.
__device__ __forceinline__ float4 bn_mask_src(const float* y, int64_t o, int64_t r, int C, int q, bool bits) {
  if (bits) {
    const unsigned b = reinterpret_cast<const unsigned char*>(y)[r * (C >> 2) + q];
    return make_float4((b & 1u) ? 1.f : 0.f, (b & 2u) ? 1.f : 0.f, (b & 4u) ? 1.f : 0.f, (b & 8u) ? 1.f : 0.f);
  }
  return bn_ld4(y + o);
}
__device__ __forceinline__ unsigned char bn_mask_bits(float4 v) {
  return (unsigned char)((v.x > 0.f ? 1u : 0u) | (v.y > 0.f ? 2u : 0u) | (v.z > 0.f ? 4u : 0u) | (v.w > 0.f ? 8u : 0u));
}

// true in the last block to arrive (self-resetting ticket), same pattern as gk_cuda.cu
__device__ __forceinline__ bool bn_last_block(uint32_t* ticket, uint32_t nblocks) {
  __shared__ int s_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const uint32_t t = atomicAdd(ticket, 1u);
    s_last = (t == nblocks - 1);
    if (s_last) *ticket = 0;
  }
  __syncthreads();
  if (s_last) __threadfence();
  return s_last != 0;
}

// Two per-channel sums over this block's rows, added into accum[0/1][C] (fp64, zero between uses).
// MODE 0 (forward): (sum x, sum x^2).
// MODE 1 (backward): (sum g, sum g*xhat) with g = dy masked by the ReLU of the forward output.
template <int MODE>
__device__ __forceinline__ void bn_block_sums(const float* __restrict__ x, const float* __restrict__ dy,
                                              const float* __restrict__ y, int64_t M, int C, int rows_per_block,
                                              int qpb, float4 mean, float4 istd, float4 scale, float4 shift,
                                              int relu, double* __restrict__ partial,
                                              float* __restrict__ gout) {
  __shared__ float4 sm[2][BN_T];
  const int tid = threadIdx.x;
  const int ql = tid % qpb, rl = tid / qpb, RL = BN_T / qpb;   // threads with rl >= RL idle (qpb need not divide BN_T)
  const int q = blockIdx.y * qpb + ql;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = (r0 + rows_per_block < M) ? r0 + rows_per_block : M;
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
#pragma unroll 4
  for (int64_t r = (rl < RL) ? r0 + rl : r1; r < r1; r += RL) {
    const float4 xv = bn_ld4(x + r * C + 4 * q);
    if (MODE == 0) {
      a = f4_add(a, xv);
      b = f4_fma(xv, xv, b);
    } else {
      float4 g = bn_ld4(dy + r * C + 4 * q);
      if (relu) {
        if (y != nullptr) {                       // residual case: the mask comes from the saved output
          const float4 yv = bn_ld4(y + r * C + 4 * q);
          g.x = yv.x > 0.f ? g.x : 0.f; g.y = yv.y > 0.f ? g.y : 0.f;
          g.z = yv.z > 0.f ? g.z : 0.f; g.w = yv.w > 0.f ? g.w : 0.f;
        } else {                                  // recompute z = x*scale + shift instead of reading y
          g.x = fmaf(xv.x, scale.x, shift.x) > 0.f ? g.x : 0.f; g.y = fmaf(xv.y, scale.y, shift.y) > 0.f ? g.y : 0.f;
          g.z = fmaf(xv.z, scale.z, shift.z) > 0.f ? g.z : 0.f; g.w = fmaf(xv.w, scale.w, shift.w) > 0.f ? g.w : 0.f;
        }
      }
      if (gout != nullptr) bn_st4(gout + r * C + 4 * q, g);   // masked gradient = gradient of the residual branch
      const float4 xh = make_float4((xv.x - mean.x) * istd.x, (xv.y - mean.y) * istd.y, (xv.z - mean.z) * istd.z,
                                    (xv.w - mean.w) * istd.w);
      a = f4_add(a, g);
      b = f4_fma(g, xh, b);
    }
  }
  sm[0][tid] = a;
  sm[1][tid] = b;
  __syncthreads();
  if (rl == 0) {
    for (int i = 1; i < RL; ++i) {
      a = f4_add(a, sm[0][i * qpb + ql]);
      b = f4_add(b, sm[1][i * qpb + ql]);
    }
    // fp64 atomics into accum[0..C) / accum[C..2C): one pair per channel and block.  The order of the adds
    // varies, but at 2^-53 relative the sums round to the same fp32 statistics in practice, and the last
    // block needs O(1) work instead of a serial sweep over per-block partials.
    double* acc = partial;
    atomicAdd(acc + 4 * q + 0, (double)a.x); atomicAdd(acc + 4 * q + 1, (double)a.y);
    atomicAdd(acc + 4 * q + 2, (double)a.z); atomicAdd(acc + 4 * q + 3, (double)a.w);
    atomicAdd(acc + C + 4 * q + 0, (double)b.x); atomicAdd(acc + C + 4 * q + 1, (double)b.y);
    atomicAdd(acc + C + 4 * q + 2, (double)b.z); atomicAdd(acc + C + 4 * q + 3, (double)b.w);
  }
}

// ------------------------------------------------------------------------------- forward
// Pass 1: per-channel sum / sum of squares; the last block turns them into mean, 1/std, updates the running
// statistics (torch semantics: biased variance for normalisation, unbiased for running_var) and bumps
// num_batches_tracked.
__global__ void __launch_bounds__(BN_T) k_bn_stats(const float* __restrict__ x, int64_t M, int C, int rows_per_block,
                                                    int qpb, double* __restrict__ partial, uint32_t* ticket,
                                                    float* __restrict__ save_mean, float* __restrict__ save_istd,
                                                    float* running_mean, float* running_var, int64_t* num_batches,
                                                    float eps, float momentum) {
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  bn_block_sums<0>(x, nullptr, nullptr, M, C, rows_per_block, qpb, z, z, z, z, 0, partial, nullptr);
  if (bn_last_block(ticket, gridDim.x * gridDim.y)) {
    for (int c = threadIdx.x; c < C; c += BN_T) {
      const double s = __ldcg(partial + c), ss = __ldcg(partial + C + c);
      partial[c] = 0;                 // ready for the next use
      partial[C + c] = 0;
      const double mean = s / (double)M;
      double var = ss / (double)M - mean * mean;
      if (var < 0) var = 0;
      save_mean[c] = (float)mean;
      save_istd[c] = (float)(1.0 / sqrt(var + (double)eps));
      if (running_mean != nullptr) {
        const double unbiased = M > 1 ? var * (double)M / (double)(M - 1) : var;
        running_mean[c] = (float)((1.0 - momentum) * (double)running_mean[c] + momentum * mean);
        running_var[c] = (float)((1.0 - momentum) * (double)running_var[c] + momentum * unbiased);
      }
    }
    if (threadIdx.x == 0 && num_batches != nullptr) *num_batches += 1;
  }
}

// Pass 2: y = [relu]( (x - mean) * istd * gamma + beta [+ residual] )
__global__ void __launch_bounds__(BN_T) k_bn_apply(const float* __restrict__ x, const float* __restrict__ res,
                                                    float* __restrict__ y, const float* __restrict__ gamma,
                                                    const float* __restrict__ beta, const float* __restrict__ save_mean,
                                                    const float* __restrict__ save_istd, int64_t M, int C,
                                                    int rows_per_block, int qpb, int relu) {
  const int tid = threadIdx.x;
  const int ql = tid % qpb, rl = tid / qpb, RL = BN_T / qpb;
  const int q = blockIdx.y * qpb + ql;
  const float4 g = bn_ld4(gamma + 4 * q), bt = bn_ld4(beta + 4 * q);
  const float4 mu = bn_ld4(save_mean + 4 * q), is = bn_ld4(save_istd + 4 * q);
  const float4 sc = make_float4(g.x * is.x, g.y * is.y, g.z * is.z, g.w * is.w);
  const float4 sh = make_float4(fmaf(-mu.x, sc.x, bt.x), fmaf(-mu.y, sc.y, bt.y), fmaf(-mu.z, sc.z, bt.z),
                                fmaf(-mu.w, sc.w, bt.w));
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = (r0 + rows_per_block < M) ? r0 + rows_per_block : M;
#pragma unroll 4
  for (int64_t r = (rl < RL) ? r0 + rl : r1; r < r1; r += RL) {
    const float4 xv = bn_ld4(x + r * C + 4 * q);
    float4 o = make_float4(fmaf(xv.x, sc.x, sh.x), fmaf(xv.y, sc.y, sh.y), fmaf(xv.z, sc.z, sh.z), fmaf(xv.w, sc.w, sh.w));
    if (res != nullptr) o = f4_add(o, bn_ld4(res + r * C + 4 * q));
    if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
    bn_st4(y + r * C + 4 * q, o);
  }
}

// ------------------------------------------------------------------------------ backward
__device__ __forceinline__ void bn_load_channel_consts(const float* gamma, const float* beta, const float* save_mean,
                                                       const float* save_istd, int q, float4& mu, float4& is,
                                                       float4& sc, float4& sh) {
  const float4 g = bn_ld4(gamma + 4 * q), bt = bn_ld4(beta + 4 * q);
  mu = bn_ld4(save_mean + 4 * q);
  is = bn_ld4(save_istd + 4 * q);
  sc = make_float4(g.x * is.x, g.y * is.y, g.z * is.z, g.w * is.w);      // same expressions as the forward,
  sh = make_float4(fmaf(-mu.x, sc.x, bt.x), fmaf(-mu.y, sc.y, bt.y),      // so the recomputed ReLU mask is
                   fmaf(-mu.z, sc.z, bt.z), fmaf(-mu.w, sc.w, bt.w));     // exactly the forward's
}

// Pass 1: per-channel sum g and sum g*xhat; the last block writes dgamma, dbeta and the two means pass 2 needs.
__global__ void __launch_bounds__(BN_T) k_bn_bwd_reduce(const float* __restrict__ dy, const float* __restrict__ x,
                                                         const float* __restrict__ y, const float* __restrict__ gamma,
                                                         const float* __restrict__ beta,
                                                         const float* __restrict__ save_mean,
                                                         const float* __restrict__ save_istd, int64_t M, int C,
                                                         int rows_per_block, int qpb, int relu,
                                                         double* __restrict__ accum, uint32_t* ticket,
                                                         float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                         float* __restrict__ coef /* [2][C]: mean g, mean g*xhat */,
                                                         float* __restrict__ gout /* masked g or null */) {
  float4 mu, is, sc, sh;
  bn_load_channel_consts(gamma, beta, save_mean, save_istd, blockIdx.y * qpb + threadIdx.x % qpb, mu, is, sc, sh);
  bn_block_sums<1>(x, dy, y, M, C, rows_per_block, qpb, mu, is, sc, sh, relu, accum, gout);
  if (bn_last_block(ticket, gridDim.x * gridDim.y)) {
    for (int c = threadIdx.x; c < C; c += BN_T) {
      const double sg = __ldcg(accum + c), sgx = __ldcg(accum + C + c);
      accum[c] = 0;
      accum[C + c] = 0;
      dbeta[c] = (float)sg;
      dgamma[c] = (float)sgx;
      coef[c] = (float)(sg / (double)M);
      coef[C + c] = (float)(sgx / (double)M);
    }
  }
}

// Pass 2: dx = gamma*istd * (g - mean(g) - xhat * mean(g*xhat)).  With a residual branch pass 1 has already
// written the masked gradient (= dres); the host then passes it here as dy with relu = 0.
__global__ void __launch_bounds__(BN_T) k_bn_bwd_dx(const float* __restrict__ dy, const float* __restrict__ x,
                                                     const float* __restrict__ y, const float* __restrict__ gamma,
                                                     const float* __restrict__ beta, const float* __restrict__ save_mean,
                                                     const float* __restrict__ save_istd, const float* __restrict__ coef,
                                                     float* __restrict__ dx, int64_t M, int C, int rows_per_block,
                                                     int qpb, int relu) {
  const int tid = threadIdx.x;
  const int ql = tid % qpb, rl = tid / qpb, RL = BN_T / qpb;
  const int q = blockIdx.y * qpb + ql;
  float4 mu, is, sc, sh;
  bn_load_channel_consts(gamma, beta, save_mean, save_istd, q, mu, is, sc, sh);
  const float4 mg = bn_ld4(coef + 4 * q), mgx = bn_ld4(coef + C + 4 * q);
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = (r0 + rows_per_block < M) ? r0 + rows_per_block : M;
#pragma unroll 4
  for (int64_t r = (rl < RL) ? r0 + rl : r1; r < r1; r += RL) {
    const float4 xv = bn_ld4(x + r * C + 4 * q);
    float4 g = bn_ld4(dy + r * C + 4 * q);
    if (relu) {
      if (y != nullptr) {
        const float4 yv = bn_ld4(y + r * C + 4 * q);
        g.x = yv.x > 0.f ? g.x : 0.f; g.y = yv.y > 0.f ? g.y : 0.f;
        g.z = yv.z > 0.f ? g.z : 0.f; g.w = yv.w > 0.f ? g.w : 0.f;
      } else {
        g.x = fmaf(xv.x, sc.x, sh.x) > 0.f ? g.x : 0.f; g.y = fmaf(xv.y, sc.y, sh.y) > 0.f ? g.y : 0.f;
        g.z = fmaf(xv.z, sc.z, sh.z) > 0.f ? g.z : 0.f; g.w = fmaf(xv.w, sc.w, sh.w) > 0.f ? g.w : 0.f;
      }
    }
    float4 o;
    o.x = sc.x * (g.x - mg.x - (xv.x - mu.x) * is.x * mgx.x);
    o.y = sc.y * (g.y - mg.y - (xv.y - mu.y) * is.y * mgx.y);
    o.z = sc.z * (g.z - mg.z - (xv.z - mu.z) * is.z * mgx.z);
    o.w = sc.w * (g.w - mg.w - (xv.w - mu.w) * is.w * mgx.w);
    bn_st4(dx + r * C + 4 * q, o);
  }
}

// --------------------------------------------------------------------------------- host
namespace {
struct BnGeom {
  int qpb, rpb;
  dim3 grid;
};

int bn_sm_count() {
  static int nsm[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (nsm[dev] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    nsm[dev] = v;
  }
  return nsm[dev];
}

// qpb = the largest divisor of C/4 that is <= 32 (consecutive threads then read consecutive quads of one row);
// about eight resident blocks per SM, but never fewer than four rows per thread.
bool bn_geometry(int64_t M, int C, BnGeom* g) {
  if (M <= 0 || C <= 0 || (C & 3)) return false;
  const int quads = C / 4;
  int qpb = quads < 32 ? quads : 32;
  while (quads % qpb) --qpb;
  const int RL = BN_T / qpb;
  const int gy = quads / qpb;
  int64_t gx_target = ((int64_t)bn_sm_count() * 8 + gy - 1) / gy;
  if (gx_target < 1) gx_target = 1;
  int64_t rpb = (M + gx_target - 1) / gx_target;
  if (rpb < 4 * RL) rpb = 4 * RL;
  if (rpb > (int64_t)1 << 30) return false;
  const int64_t gx = (M + rpb - 1) / rpb;
  g->qpb = qpb;
  g->rpb = (int)rpb;
  g->grid = dim3((unsigned)gx, (unsigned)gy, 1);
  return true;
}

inline bool bn_aligned(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
}  // namespace

extern "C" int gk_bn_fwd(const float* x, const float* res, float* y, const float* gamma, const float* beta,
                         float* running_mean, float* running_var, int64_t* num_batches, float* save_mean,
                         float* save_istd, double* accum, uint32_t* ticket, int64_t M, int C, float eps,
                         float momentum, int relu, void* stream) {
  BnGeom g;
  if (!bn_geometry(M, C, &g)) return (int)cudaErrorInvalidValue;
  if (!bn_aligned(x) || !bn_aligned(y) || !bn_aligned(res) || !bn_aligned(gamma) || !bn_aligned(beta) ||
      !bn_aligned(save_mean) || !bn_aligned(save_istd))
    return (int)cudaErrorMisalignedAddress;
  cudaStream_t s = (cudaStream_t)stream;
  k_bn_stats<<<g.grid, BN_T, 0, s>>>(x, M, C, g.rpb, g.qpb, accum, ticket, save_mean, save_istd, running_mean,
                                     running_var, num_batches, eps, momentum);
  k_bn_apply<<<g.grid, BN_T, 0, s>>>(x, res, y, gamma, beta, save_mean, save_istd, M, C, g.rpb, g.qpb, relu);
  return (int)cudaGetLastError();
}

extern "C" int gk_bn_bwd(const float* dy, const float* x, const float* y, const float* gamma, const float* beta,
                         const float* save_mean, const float* save_istd, float* dx, float* dres, float* dgamma,
                         float* dbeta, float* coef, double* accum, uint32_t* ticket, int64_t M, int C, int relu,
                         void* stream) {
  BnGeom g;
  if (!bn_geometry(M, C, &g)) return (int)cudaErrorInvalidValue;
  if (!bn_aligned(dy) || !bn_aligned(x) || !bn_aligned(y) || !bn_aligned(dx) || !bn_aligned(dres) ||
      !bn_aligned(gamma) || !bn_aligned(beta) || !bn_aligned(save_mean) || !bn_aligned(save_istd) ||
      !bn_aligned(coef))
    return (int)cudaErrorMisalignedAddress;
  cudaStream_t s = (cudaStream_t)stream;
  // y != null: mask from the saved output (needed when a residual was added); dres != null: also emit masked g
  k_bn_bwd_reduce<<<g.grid, BN_T, 0, s>>>(dy, x, y, gamma, beta, save_mean, save_istd, M, C, g.rpb, g.qpb, relu,
                                          accum, ticket, dgamma, dbeta, coef, dres);
  if (dres != nullptr)
    k_bn_bwd_dx<<<g.grid, BN_T, 0, s>>>(dres, x, nullptr, gamma, beta, save_mean, save_istd, coef, dx, M, C, g.rpb,
                                        g.qpb, 0);
  else
    k_bn_bwd_dx<<<g.grid, BN_T, 0, s>>>(dy, x, y, gamma, beta, save_mean, save_istd, coef, dx, M, C, g.rpb, g.qpb,
                                        relu);
  return (int)cudaGetLastError();
}

// ======================================================================= single-kernel variants
// One cooperative kernel per direction: every block keeps its slab of the activation in shared memory,
// the per-channel sums cross blocks through fp64 atomics and a per-channel-slab barrier, and the second
// phase runs out of shared memory.  HBM traffic: forward read x (+res) once, write y; backward read dy, x
// (+y with a residual) once, write dx (+dres).  The grid is (gx row slabs) x (gy channel slabs) with
// gx*gy <= number of SMs, one block per SM, launched with cudaLaunchCooperativeKernel so that all blocks
// are resident while they wait for each other.  Rows that do not fit in shared memory (the four largest
// ResNet-50 shapes at batch 32) are re-read from L2/HBM in the second phase.
//
// Scratch words (uint64) of one layer: bar[2y] arrivals and bar[2y+1] generation of channel slab y
// (y < BNF_MAX_COLUMNS), bar[96..111] optional phase time stamps, bar[127] error flag.
// threads per block of the cooperative kernels: template parameter T_ (512 or 1024, chosen per layer by
// bnf_geometry)
// unroll depths of the streaming loops = loads in flight per thread (tuned on a B200, profiles/bn_unroll_sweep.txt)
#ifndef BNF_U_F1
#define BNF_U_F1 8
#endif
#ifndef BNF_U_F2
#define BNF_U_F2 4
#endif
#ifndef BNF_U_B1
#define BNF_U_B1 4
#endif
#ifndef BNF_U_B2
#define BNF_U_B2 4
#endif
// Row assignment.  Interleaved (default): in iteration i block b handles rows (i*gridDim.x + b)*RL .. +RL-1, so
// at any moment the blocks of a channel slab read one contiguous region (the access pattern of a streaming
// copy: all DRAM channels busy, blocks finish together) and the rows read last — the ones phase 2 starts with —
// are the tail of the tensor for every block.  BNF_INTERLEAVE=0: each block owns a contiguous slab of rows.
#ifndef BNF_INTERLEAVE
#define BNF_INTERLEAVE 1
#endif
#if BNF_INTERLEAVE
#define BNF_ROWS_DECL                                                                     \
  const int64_t rstep = (int64_t)gridDim.x * RL, rbase = (int64_t)blockIdx.x * RL + rl;   \
  const int n_it = rbase < M ? (int)((M - rbase + rstep - 1) / rstep) : 0;                \
  (void)rows_per_block;
#else
#define BNF_ROWS_DECL                                                                     \
  const int64_t rstep = RL, r0_ = (int64_t)blockIdx.x * rows_per_block;                   \
  const int64_t r1_ = (r0_ + rows_per_block < M) ? r0_ + rows_per_block : M;              \
  const int64_t rbase = r0_ + rl;                                                         \
  const int n_it = rbase < r1_ ? (int)((r1_ - rbase + RL - 1) / RL) : 0;
#endif
// the 1024-thread instantiations have 64 registers per thread: their own (smaller) depths
#ifndef BNF_U_F1_BIG
#define BNF_U_F1_BIG 4
#endif
#ifndef BNF_U_F2_BIG
#define BNF_U_F2_BIG 4
#endif
#ifndef BNF_U_B1_BIG
#define BNF_U_B1_BIG 2
#endif
#ifndef BNF_U_B2_BIG
#define BNF_U_B2_BIG 4
#endif
template <int T_> struct BnfUnroll {
  static constexpr int F1 = T_ >= 1024 ? BNF_U_F1_BIG : BNF_U_F1, F2 = T_ >= 1024 ? BNF_U_F2_BIG : BNF_U_F2,
                       B1 = T_ >= 1024 ? BNF_U_B1_BIG : BNF_U_B1, B2 = T_ >= 1024 ? BNF_U_B2_BIG : BNF_U_B2;
};
#define BNF_SPIN_LIMIT (1 << 22)
#define BNF_MAX_COLUMNS 48
#define BNF_STAMP_WORD 96
#define BNF_ERR_WORD 127

__device__ __forceinline__ unsigned long long bnf_ld_relaxed(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void bnf_st_release(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long bnf_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Sense-reversing barrier among the gridDim.x blocks that share blockIdx.y.  `gen` is the generation word
// as read when the block started (it cannot advance before this block arrives).  The last block to arrive
// resets the arrival count and publishes gen+1; the others poll the generation word.  The wait is bounded:
// on expiry the error word is set and the block carries on (wrong numbers, no hang) — that can only happen
// if the blocks were not co-resident, which the cooperative launch rules out.
__device__ __forceinline__ void bnf_column_barrier(unsigned long long* bar, unsigned long long gen) {
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long* arrive = bar + 2 * blockIdx.y;
    unsigned long long* genp = arrive + 1;
    __threadfence();
    const unsigned long long old = atomicAdd(arrive, 1ULL);
    if (old == (unsigned long long)gridDim.x - 1ULL) {
      *arrive = 0;
      bnf_st_release(genp, gen + 1ULL);
    } else {
      int spins = 0;
      while (bnf_ld_relaxed(genp) == gen) {
        if (++spins > BNF_SPIN_LIMIT) {
          *reinterpret_cast<int*>(bar + BNF_ERR_WORD) = 1;
          break;
        }
      }
    }
    __threadfence();
  }
  __syncthreads();
}

// block-level reduction of (a, b) over the row lanes that share a channel quad; result valid where rl == 0
template <int T_>
__device__ __forceinline__ void bnf_reduce_rows(float4 (*red)[T_], float4& a, float4& b, int tid, int rl, int RL,
                                                int qpb) {
  red[0][tid] = a;
  red[1][tid] = b;
  __syncthreads();
  for (int s = RL >> 1; s > 0; s >>= 1) {
    if (rl < s) {
      a = f4_add(a, red[0][tid + s * qpb]);
      b = f4_add(b, red[1][tid + s * qpb]);
      red[0][tid] = a;
      red[1][tid] = b;
    }
    __syncthreads();
  }
}

__device__ __forceinline__ void bnf_atomic_add8(double* acc, int C, int q, float4 a, float4 b) {
  atomicAdd(acc + 4 * q + 0, (double)a.x); atomicAdd(acc + 4 * q + 1, (double)a.y);
  atomicAdd(acc + 4 * q + 2, (double)a.z); atomicAdd(acc + 4 * q + 3, (double)a.w);
  atomicAdd(acc + C + 4 * q + 0, (double)b.x); atomicAdd(acc + C + 4 * q + 1, (double)b.y);
  atomicAdd(acc + C + 4 * q + 2, (double)b.z); atomicAdd(acc + C + 4 * q + 3, (double)b.w);
}

// the eight totals of one channel quad, fetched from L2 in one batch
__device__ __forceinline__ void bnf_load_totals(const double* acc, int C, int q, double* s, double* ss) {
  const double2 s01 = __ldcg(reinterpret_cast<const double2*>(acc + 4 * q));
  const double2 s23 = __ldcg(reinterpret_cast<const double2*>(acc + 4 * q + 2));
  const double2 t01 = __ldcg(reinterpret_cast<const double2*>(acc + C + 4 * q));
  const double2 t23 = __ldcg(reinterpret_cast<const double2*>(acc + C + 4 * q + 2));
  s[0] = s01.x; s[1] = s01.y; s[2] = s23.x; s[3] = s23.y;
  ss[0] = t01.x; ss[1] = t01.y; ss[2] = t23.x; ss[3] = t23.y;
}

__device__ __forceinline__ void bnf_cp_async16(void* smem, const void* gmem) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void bnf_cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}

// ---- optional: TMA bulk copies (cp.async.bulk, SASS UBLKCP) into the shared-memory slab, completion on an
// mbarrier.  One thread per row lane issues one copy per cached row (the row's channel slab, 32-512 B
// contiguous) instead of every thread issuing a 16-byte cp.async.  Built with -DBNF_TMA=1; measured on a
// B200 it LOSES against cp.async (sum over the 53 ResNet-50 layers 3.65 ms vs 3.45 ms, the 7x7 layers with
// their 32-64 B rows 35 % slower, the 100 MB layers equal; profiles/bn_unroll_sweep.txt), because the rows of
// an NHWC channel slab are far below the size where one bulk copy beats sixteen vector copies — so the
// default stays cp.async.
#ifndef BNF_TMA
#define BNF_TMA 0
#endif
__device__ __forceinline__ unsigned bnf_saddr(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bnf_mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bnf_saddr(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void bnf_mbar_expect(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bnf_saddr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bnf_bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(bnf_saddr(dst)), "l"(src), "r"(bytes), "r"(bnf_saddr(bar)) : "memory");
}
__device__ __forceinline__ void bnf_mbar_wait(unsigned long long* bar, unsigned parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "BNF_WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra BNF_DONE_%=;\n\t"
      "bra BNF_WAIT_%=;\n\t"
      "BNF_DONE_%=:\n\t}"
      ::"r"(bnf_saddr(bar)), "r"(parity) : "memory");
}

__device__ __forceinline__ float bnf_rsqrt(float a) {   // rsqrt.approx plus one Newton step: < 1 ulp
  const float y = rsqrtf(a);
  return y * fmaf(-0.5f * a * y, y, 1.5f);
}

// flags: bit 0 = ReLU, bit 1 = write phase time stamps (block 0,0 only; profiling aid)
template <int T_>
__global__ void __launch_bounds__(T_, 1)
k_bn_fwd_fused(const float* __restrict__ x, const float* __restrict__ res, float* __restrict__ y,
               const float* __restrict__ gamma, const float* __restrict__ beta, float* __restrict__ save_mean,
               float* __restrict__ save_istd, float* running_mean, float* running_var, int64_t* num_batches,
               double* accum /* [2 parities][2][C] */, unsigned long long* bar, int64_t M, int C,
               int rows_per_block, int qshift, int cached, float eps, float momentum, double inv_m, double unbias,
               int flags, unsigned char* __restrict__ mask /* optional: ReLU bits, one byte per channel quad */) {
  extern __shared__ float4 tile[];          // [cached][T_]
  __shared__ float4 red[2][T_];
  __shared__ unsigned long long s_gen;
#if T_MA
  __shared__ __align__(8) unsigned long long s_mbar;
#endif
  const int tid = threadIdx.x;
  const int qpb = 1 << qshift, ql = tid & (qpb - 1), rl = tid >> qshift, RL = T_ >> qshift;
  const int q = blockIdx.y * qpb + ql;
  const bool stamp = (flags & 2) && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0;
  unsigned long long* ts = bar + BNF_STAMP_WORD;
  if (stamp) ts[0] = bnf_now();
  if (tid == 0 && gridDim.x > 1) s_gen = bnf_ld_relaxed(bar + 2 * blockIdx.y + 1);
#if T_MA
  if (tid == 0) bnf_mbar_init(&s_mbar, (unsigned)RL);      // one arrival per row lane
  __syncthreads();
#endif
  const int relu = flags & 1;
  BNF_ROWS_DECL
  const float* xp = x + 4 * q;

  // what the second phase needs besides the sums is fetched now, so its latency hides behind phase 1
  const float4 gm = bn_ld4(gamma + 4 * q), bt = bn_ld4(beta + 4 * q);
  const bool keeper = blockIdx.x == 0 && rl == 0;      // writes the per-channel results of its quad
  float4 rm = make_float4(0.f, 0.f, 0.f, 0.f), rv = rm;
  if (keeper && running_mean != nullptr) {
    rm = bn_ld4(running_mean + 4 * q);
    rv = bn_ld4(running_var + 4 * q);
  }

  // phase 1: one read of x.  The rows that fit in shared memory are fetched with cp.async (all in flight
  // at once, no register staging); the rest streams through registers meanwhile.
  const int nc = n_it < cached ? n_it : cached;
#if T_MA
  if (ql == 0) {
    bnf_mbar_expect(&s_mbar, (unsigned)(nc * qpb * 16));
    for (int i = 0; i < nc; ++i)
      bnf_bulk_g2s(&tile[i * T_ + tid], xp + (rbase + (int64_t)i * rstep) * C, (unsigned)(qpb * 16), &s_mbar);
  }
#else
  for (int i = 0; i < nc; ++i) bnf_cp_async16(&tile[i * T_ + tid], xp + (rbase + (int64_t)i * rstep) * C);
#endif
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
#pragma unroll BnfUnroll<T_>::F1
  for (int i = nc; i < n_it; ++i) {
    const float4 xv = bn_ld4(xp + (rbase + (int64_t)i * rstep) * C);
    a = f4_add(a, xv);
    b = f4_fma(xv, xv, b);
  }
#if T_MA
  bnf_mbar_wait(&s_mbar, 0);
#else
  bnf_cp_async_wait_all();
#endif
#pragma unroll 8
  for (int i = 0; i < nc; ++i) {
    const float4 xv = tile[i * T_ + tid];
    a = f4_add(a, xv);
    b = f4_fma(xv, xv, b);
  }
  if (stamp) ts[1] = bnf_now();
  bnf_reduce_rows(red, a, b, tid, rl, RL, qpb);
  const bool alone = gridDim.x == 1;        // this block owns its channels: no atomics, no barrier
  const unsigned long long gen = alone ? 0ULL : s_gen;
  double* acc = accum + (gen & 1ULL) * 2 * (size_t)C;
  double s[4], ss[4];
  if (alone) {
    const float4 ta = red[0][ql], tb = red[1][ql];
    s[0] = ta.x; s[1] = ta.y; s[2] = ta.z; s[3] = ta.w;
    ss[0] = tb.x; ss[1] = tb.y; ss[2] = tb.z; ss[3] = tb.w;
    if (stamp) ts[2] = ts[3] = bnf_now();
  } else {
    if (rl == 0) bnf_atomic_add8(acc, C, q, a, b);
    if (stamp) ts[2] = bnf_now();
    bnf_column_barrier(bar, gen);
    if (stamp) ts[3] = bnf_now();
    bnf_load_totals(acc, C, q, s, ss);
  }

  // per-channel constants, recomputed by every thread for its own quad (no divisions, no fp64 sqrt)
  float mu[4], is[4];
  double varv[4], meanv[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const double mean = s[j] * inv_m;
    double var = fma(ss[j], inv_m, -mean * mean);
    if (var < 0) var = 0;
    meanv[j] = mean;
    varv[j] = var;
    mu[j] = (float)mean;
    is[j] = bnf_rsqrt((float)var + eps);
  }
  const float4 sc = make_float4(gm.x * is[0], gm.y * is[1], gm.z * is[2], gm.w * is[3]);
  const float4 sh = make_float4(fmaf(-mu[0], sc.x, bt.x), fmaf(-mu[1], sc.y, bt.y), fmaf(-mu[2], sc.z, bt.z),
                                fmaf(-mu[3], sc.w, bt.w));
  if (stamp) ts[4] = bnf_now();

  // y == NULL: statistics only (the caller applies them itself, e.g. fused with a max pool)

  // phase 2: normalise out of shared memory (rows beyond the cached part come from L2/HBM again)
  // (last rows first: they are the ones most likely still in L2)
  if (y != nullptr) {
#pragma unroll BnfUnroll<T_>::F2
    for (int i = n_it - 1; i >= 0; --i) {
      const int64_t r = rbase + (int64_t)i * rstep;
      const float4 xv = (i < nc) ? tile[i * T_ + tid] : bn_ld4(xp + r * C);
      float4 o = f4_fma(xv, sc, sh);
      if (res != nullptr) o = f4_add(o, bn_ld4(res + r * C + 4 * q));
      if (mask != nullptr) mask[r * (C >> 2) + q] = bn_mask_bits(o);
      if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
      bn_st4(y + r * C + 4 * q, o);
    }
  }
  // per-channel results last: their latency (running statistics are read-modify-write) is off the critical path
  if (keeper) {
    double* other = accum + ((gen + 1ULL) & 1ULL) * 2 * (size_t)C;   // dirty from the previous launch
    if (alone) other = acc;                                           // (nothing is dirty: keep the stores valid)
    const float rmv[4] = {rm.x, rm.y, rm.z, rm.w}, rvv[4] = {rv.x, rv.y, rv.z, rv.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = 4 * q + j;
      other[c] = 0;
      other[C + c] = 0;
      save_mean[c] = mu[j];
      save_istd[c] = is[j];
      if (running_mean != nullptr) {
        running_mean[c] = (float)((1.0 - momentum) * (double)rmv[j] + momentum * meanv[j]);
        running_var[c] = (float)((1.0 - momentum) * (double)rvv[j] + momentum * (varv[j] * unbias));
      }
    }
    if (blockIdx.y == 0 && tid == 0 && num_batches != nullptr) *num_batches += 1;
  }
  if (stamp) ts[5] = bnf_now();
}

__device__ __forceinline__ float4 bn_mask_pos(float4 g, float4 z) {
  g.x = z.x > 0.f ? g.x : 0.f; g.y = z.y > 0.f ? g.y : 0.f;
  g.z = z.z > 0.f ? g.z : 0.f; g.w = z.w > 0.f ? g.w : 0.f;
  return g;
}

// Gradient arriving through a 3x3 / stride 2 / padding 1 max pool that followed the BN+ReLU (ResNet stem):
// the pooled gradient dy[N][OH][OW][C] and the window position (0..8, kh*3+kw) of each maximum.
struct BnPool {
  const float* dy;
  const unsigned char* idx;
  int H, W, OH, OW;
};

// gradient of one input pixel's channel quad = sum over the (1, 2 or 4) windows that contain it and whose
// maximum sits exactly there
__device__ __forceinline__ float4 bn_pool_gather(const BnPool& P, int64_t r, int C, int q) {
  const int hw = P.H * P.W;
  const int n = (int)(r / hw), rem = (int)(r - (int64_t)n * hw);
  const int h = rem / P.W, w = rem - h * P.W;
  float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
  const int oh0 = h >> 1, ow0 = w >> 1;                 // window oh covers rows 2*oh-1 .. 2*oh+1
#pragma unroll
  for (int dh = 0; dh < 2; ++dh) {
    const int oh = oh0 + dh;
    if (oh >= P.OH || (dh == 1 && !(h & 1))) continue;  // even rows belong to one window only
#pragma unroll
    for (int dw = 0; dw < 2; ++dw) {
      const int ow = ow0 + dw;
      if (ow >= P.OW || (dw == 1 && !(w & 1))) continue;
      const unsigned code = (unsigned)((h - (2 * oh - 1)) * 3 + (w - (2 * ow - 1)));
      const int64_t o = (((int64_t)n * P.OH + oh) * P.OW + ow) * C + 4 * q;
      const uchar4 id = *reinterpret_cast<const uchar4*>(P.idx + o);
      const float4 d = bn_ld4(P.dy + o);
      g.x += id.x == code ? d.x : 0.f; g.y += id.y == code ? d.y : 0.f;
      g.z += id.z == code ? d.z : 0.f; g.w += id.w == code ? d.w : 0.f;
    }
  }
  return g;
}

// dy2: optional second gradient of the same output (the output was handed out twice, see ops/bn.py), added
// on the fly.  POOL: the gradient comes through the max pool described by P instead of dy.
template <int T_, bool POOL, bool DUAL>
__global__ void __launch_bounds__(T_, 1)
k_bn_bwd_fused(const float* __restrict__ dy, const float* __restrict__ dy2, const float* __restrict__ x,
               const float* __restrict__ y, const float* __restrict__ gamma, const float* __restrict__ beta,
               const float* __restrict__ save_mean, const float* __restrict__ save_istd, float* __restrict__ dx,
               float* __restrict__ dres, float* __restrict__ dgamma, float* __restrict__ dbeta, double* accum,
               unsigned long long* bar, int64_t M, int C, int rows_per_block, int qshift, int cached,
               double inv_m, int flags, BnPool P) {
  extern __shared__ float4 tile[];          // [cached][2][T_]: masked gradient, x
  __shared__ float4 red[2][T_];
  __shared__ unsigned long long s_gen;
#if T_MA
  __shared__ __align__(8) unsigned long long s_mbar;
#endif
  const int tid = threadIdx.x;
  const int qpb = 1 << qshift, ql = tid & (qpb - 1), rl = tid >> qshift, RL = T_ >> qshift;
  const int q = blockIdx.y * qpb + ql;
  const bool stamp = (flags & 2) && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0;
  unsigned long long* ts = bar + BNF_STAMP_WORD + 8;
  if (stamp) ts[0] = bnf_now();
  if (tid == 0 && gridDim.x > 1) s_gen = bnf_ld_relaxed(bar + 2 * blockIdx.y + 1);
#if T_MA
  if (tid == 0) bnf_mbar_init(&s_mbar, (unsigned)RL);
  __syncthreads();
#endif
  const int relu = flags & 1;
  const bool mbits = (flags & 4) != 0;
  BNF_ROWS_DECL
  float4 mu, is, sc, sh;
  bn_load_channel_consts(gamma, beta, save_mean, save_istd, q, mu, is, sc, sh);

  // phase 1: dy and x of the rows that fit go to shared memory with cp.async; the rest streams through
  // registers meanwhile.  g = dy masked by the forward ReLU replaces dy in shared memory.
  const int nc = n_it < cached ? n_it : cached;
  auto load_g = [&](int64_t r, int64_t o) -> float4 {
    if (POOL) return bn_pool_gather(P, r, C, q);
    float4 g = bn_ld4(dy + o);
    if (DUAL) g = f4_add(g, bn_ld4(dy2 + o));
    return g;
  };
#if T_MA
  if (ql == 0) {
    bnf_mbar_expect(&s_mbar, (unsigned)(nc * qpb * 16 * (POOL ? 1 : 2)));
    for (int i = 0; i < nc; ++i) {
      const int64_t o = (rbase + (int64_t)i * rstep) * C + 4 * q;
      if (!POOL) bnf_bulk_g2s(&tile[(2 * i) * T_ + tid], dy + o, (unsigned)(qpb * 16), &s_mbar);
      bnf_bulk_g2s(&tile[(2 * i + 1) * T_ + tid], x + o, (unsigned)(qpb * 16), &s_mbar);
    }
  }
#else
  for (int i = 0; i < nc; ++i) {
    const int64_t o = (rbase + (int64_t)i * rstep) * C + 4 * q;
    if (!POOL) bnf_cp_async16(&tile[(2 * i) * T_ + tid], dy + o);
    bnf_cp_async16(&tile[(2 * i + 1) * T_ + tid], x + o);
  }
#endif
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
#pragma unroll BnfUnroll<T_>::B1
  for (int i = nc; i < n_it; ++i) {
    const int64_t r = rbase + (int64_t)i * rstep;
    const int64_t o = r * C + 4 * q;
    const float4 xv = bn_ld4(x + o);
    float4 g = load_g(r, o);
    if (relu) g = bn_mask_pos(g, y != nullptr ? bn_mask_src(y, o, r, C, q, mbits) : f4_fma(xv, sc, sh));
    if (dres != nullptr) bn_st4(dres + o, g);
    const float4 xh = make_float4((xv.x - mu.x) * is.x, (xv.y - mu.y) * is.y, (xv.z - mu.z) * is.z,
                                  (xv.w - mu.w) * is.w);
    a = f4_add(a, g);
    b = f4_fma(g, xh, b);
  }
#if T_MA
  bnf_mbar_wait(&s_mbar, 0);
#else
  bnf_cp_async_wait_all();
#endif
#pragma unroll BnfUnroll<T_>::B1
  for (int i = 0; i < nc; ++i) {
    const int64_t r = rbase + (int64_t)i * rstep;
    const int64_t o = r * C + 4 * q;
    const float4 xv = tile[(2 * i + 1) * T_ + tid];
    float4 g;
    if (POOL) {
      g = bn_pool_gather(P, r, C, q);
    } else {
      g = tile[(2 * i) * T_ + tid];
      if (DUAL) g = f4_add(g, bn_ld4(dy2 + o));
    }
    if (relu) g = bn_mask_pos(g, y != nullptr ? bn_mask_src(y, o, r, C, q, mbits) : f4_fma(xv, sc, sh));
    if (POOL || DUAL || relu) tile[(2 * i) * T_ + tid] = g;
    if (dres != nullptr) bn_st4(dres + o, g);
    const float4 xh = make_float4((xv.x - mu.x) * is.x, (xv.y - mu.y) * is.y, (xv.z - mu.z) * is.z,
                                  (xv.w - mu.w) * is.w);
    a = f4_add(a, g);
    b = f4_fma(g, xh, b);
  }
  if (stamp) ts[1] = bnf_now();
  bnf_reduce_rows(red, a, b, tid, rl, RL, qpb);
  const bool alone = gridDim.x == 1;
  const unsigned long long gen = alone ? 0ULL : s_gen;
  double* acc = accum + (gen & 1ULL) * 2 * (size_t)C;
  double sg[4], sgx[4];
  if (alone) {
    const float4 ta = red[0][ql], tb = red[1][ql];
    sg[0] = ta.x; sg[1] = ta.y; sg[2] = ta.z; sg[3] = ta.w;
    sgx[0] = tb.x; sgx[1] = tb.y; sgx[2] = tb.z; sgx[3] = tb.w;
    if (stamp) ts[2] = ts[3] = bnf_now();
  } else {
    if (rl == 0) bnf_atomic_add8(acc, C, q, a, b);
    if (stamp) ts[2] = bnf_now();
    bnf_column_barrier(bar, gen);
    if (stamp) ts[3] = bnf_now();
    bnf_load_totals(acc, C, q, sg, sgx);
  }
  float mg[4], mgx[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    mg[j] = (float)(sg[j] * inv_m);
    mgx[j] = (float)(sgx[j] * inv_m);
  }
  if (blockIdx.x == 0 && rl == 0) {
    double* other = alone ? acc : accum + ((gen + 1ULL) & 1ULL) * 2 * (size_t)C;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      other[4 * q + j] = 0;
      other[C + 4 * q + j] = 0;
      dbeta[4 * q + j] = (float)sg[j];
      dgamma[4 * q + j] = (float)sgx[j];
    }
  }
  if (stamp) ts[4] = bnf_now();
  {
#pragma unroll BnfUnroll<T_>::B2
    for (int i = n_it - 1; i >= 0; --i) {      // last rows first: most likely still in L2
      const int64_t r = rbase + (int64_t)i * rstep;
      const int64_t o = r * C + 4 * q;
      float4 g, xv;
      if (i < nc) {
        g = tile[(2 * i) * T_ + tid];
        xv = tile[(2 * i + 1) * T_ + tid];
      } else {
        xv = bn_ld4(x + o);
        if (dres != nullptr) {
          g = bn_ld4(dres + o);            // already masked by phase 1 (written by this very thread)
        } else {
          g = load_g(r, o);
          if (relu) g = bn_mask_pos(g, y != nullptr ? bn_mask_src(y, o, r, C, q, mbits) : f4_fma(xv, sc, sh));
        }
      }
      float4 d;
      d.x = sc.x * (g.x - mg[0] - (xv.x - mu.x) * is.x * mgx[0]);
      d.y = sc.y * (g.y - mg[1] - (xv.y - mu.y) * is.y * mgx[1]);
      d.z = sc.z * (g.z - mg[2] - (xv.z - mu.z) * is.z * mgx[2]);
      d.w = sc.w * (g.w - mg[3] - (xv.w - mu.w) * is.w * mgx[3]);
      bn_st4(dx + o, d);
    }
  }
  if (stamp) ts[5] = bnf_now();
}

namespace {
// cudaLaunchCooperativeKernel reports e.g. cudaErrorCooperativeLaunchTooLarge synchronously; the error is not
// sticky, but it stays in the runtime's "last error" slot where the next launch check of the host framework
// would find it.  Take it out so that the caller can fall back to the two-kernel path.
inline int bn_coop_result(cudaError_t e) {
  if (e != cudaSuccess) (void)cudaGetLastError();
  return (int)e;
}

int g_qcap = 9;           // upper bound on log2 of the channel slab width (in quads), cooperative path
int g_private_kb = 128;   // largest per-block slab of the barrier-free path

struct BnfGeom {
  int qshift, rpb, iters, threads;
  dim3 grid;
};

int g_big_mb = 40;        // activations of at least this many MB run 1024 threads per block (else 512)
int g_pair_big_mb = 40;   // same for the two-BN kernels (per input tensor)
inline size_t bnf_smem_budget(int threads) { return (size_t)216 * 1024 - (size_t)2 * threads * 16; }

bool bnf_geometry(int64_t M, int C, BnfGeom* g, bool pair = false) {
  if (M < 2 || C <= 0 || (C & 3)) return false;
  const int quads = C / 4;
  // 1024 threads per block hide the load latency of the big streaming layers better (-15 % on the 100 MB
  // shapes although the kernels then spill a few registers), 512 win on the small ones
  // (profiles/bn_unroll_sweep.txt)
  const int T = (M * (int64_t)C * 4 >= (int64_t)(pair ? g_pair_big_mb : g_big_mb) << 20) ? 1024 : 512;
  g->threads = T;
  int qmax = 9;                             // widest power-of-two quad slab that divides the row
  while (quads % (1 << qmax)) --qmax;
  const int nsm = bn_sm_count();
  // (a) small activations with many channels: one block owns a channel slab over ALL rows, so there is
  //     nothing to exchange between blocks (>= 32 blocks of <= g_private_kb KB each, full 32 B sectors).
  //     Only worth it while the slabs are small: a strided 32 B column is the worst DRAM access pattern.
  int lg = 0;
  while ((1 << (lg + 1)) <= quads) ++lg;
  int qs = lg - 7 < 1 ? 1 : lg - 7;
  while ((quads >> qs) > nsm && qs < qmax) ++qs;
  if (qs <= qmax && (quads >> qs) >= 32 && (quads >> qs) <= nsm &&
      M * (int64_t)(16 << qs) <= (int64_t)g_private_kb * 1024) {
    const int RL = T >> qs;
    int64_t rpb = (M + RL - 1) / RL * RL;
    g->qshift = qs;
    g->rpb = (int)rpb;
    g->iters = (int)(rpb / RL);
    g->grid = dim3(1, (unsigned)(quads >> qs), 1);
    return true;
  }
  // (b) everything else: one block per SM, channel slabs chosen so that about eight (four for activations
  //     beyond 64 MB, where longer contiguous runs pay) slabs share the SMs: fewer blocks per barrier and
  //     per accumulator address than whole rows would give (measured: profiles/bn_geometry_sweep.txt)
  const int target = M * (int64_t)C * 4 > ((int64_t)64 << 20) ? 2 : 3;
  int qshift = lg - target < 3 ? 3 : lg - target;
  if (qshift > qmax) qshift = qmax;
  if (qshift > g_qcap) qshift = g_qcap;
  if (qshift < 3) return false;            // < 128 B contiguous per row: leave it to the two-kernel path
  const int qpb = 1 << qshift;
  const int gy = quads / qpb;
  if (gy > nsm || gy > BNF_MAX_COLUMNS) return false;
  const int RL = T / qpb;
  int64_t gx = nsm / gy;
  const int64_t useful = (M + RL - 1) / RL;
  if (gx > useful) gx = useful;
  int64_t rpb = (M + gx - 1) / gx;
  rpb = (rpb + RL - 1) / RL * RL;
  if (rpb > (int64_t)1 << 30) return false;
  gx = (M + rpb - 1) / rpb;
  g->qshift = qshift;
  g->rpb = (int)rpb;
  g->iters = (int)(rpb / RL);
  g->grid = dim3((unsigned)gx, (unsigned)gy, 1);
  return true;
}

int bnf_prepare() {
  static bool ready[64] = {false};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= 64) return (int)cudaErrorInvalidDevice;
  if (!ready[dev]) {
#define BNF_SET_ATTR(fn, threads)                                                                                  \
  e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bnf_smem_budget(threads));      \
  if (e != cudaSuccess) return (int)e;
    BNF_SET_ATTR(k_bn_fwd_fused<512>, 512)
    BNF_SET_ATTR(k_bn_fwd_fused<1024>, 1024)
    BNF_SET_ATTR((k_bn_bwd_fused<512, false, false>), 512)
    BNF_SET_ATTR((k_bn_bwd_fused<512, false, true>), 512)
    BNF_SET_ATTR((k_bn_bwd_fused<512, true, false>), 512)
    BNF_SET_ATTR((k_bn_bwd_fused<1024, false, false>), 1024)
    BNF_SET_ATTR((k_bn_bwd_fused<1024, false, true>), 1024)
    BNF_SET_ATTR((k_bn_bwd_fused<1024, true, false>), 1024)
#undef BNF_SET_ATTR
    ready[dev] = true;
  }
  return 0;
}
}  // namespace

// Tuning knobs of the geometry heuristic (benchmarks only; not thread-safe, set before first use).
extern "C" int gk_bn_set_tuning(int qcap_log2, int private_kb) {
  if (qcap_log2 < 3 || qcap_log2 > 9 || private_kb < 0) return (int)cudaErrorInvalidValue;
  g_qcap = qcap_log2;
  g_private_kb = private_kb;
  return 0;
}

extern "C" int gk_bn_set_big_mb(int mb) {      // activations of at least mb MB run 1024 threads per block
  if (mb < 0) return (int)cudaErrorInvalidValue;
  g_big_mb = mb;
  return 0;
}

extern "C" int gk_bn_set_pair_big_mb(int mb) {  // the same threshold for gk_bn2_fwd_fused / gk_bn2_bwd_fused
  if (mb < 0) return (int)cudaErrorInvalidValue;
  g_pair_big_mb = mb;
  return 0;
}

extern "C" int gk_bn_fused_supported(int64_t M, int C) {
  BnfGeom g;
  return bnf_geometry(M, C, &g) ? 1 : 0;
}

extern "C" int gk_bn_fwd_fused_m(const float* x, const float* res, float* y, const float* gamma, const float* beta,
                                 float* running_mean, float* running_var, int64_t* num_batches, float* save_mean,
                                 float* save_istd, double* accum, unsigned long long* bar, int64_t M, int C,
                                 float eps, float momentum, int flags, unsigned char* mask, void* stream);

extern "C" int gk_bn_fwd_fused(const float* x, const float* res, float* y, const float* gamma, const float* beta,
                               float* running_mean, float* running_var, int64_t* num_batches, float* save_mean,
                               float* save_istd, double* accum, unsigned long long* bar, int64_t M, int C,
                               float eps, float momentum, int flags, void* stream) {
  return gk_bn_fwd_fused_m(x, res, y, gamma, beta, running_mean, running_var, num_batches, save_mean, save_istd, accum,
                           bar, M, C, eps, momentum, flags, nullptr, stream);
}

extern "C" int gk_bn_fwd_fused_m(const float* x, const float* res, float* y, const float* gamma, const float* beta,
                                 float* running_mean, float* running_var, int64_t* num_batches, float* save_mean,
                                 float* save_istd, double* accum, unsigned long long* bar, int64_t M, int C,
                                 float eps, float momentum, int flags, unsigned char* mask, void* stream) {
  BnfGeom g;
  if (!bnf_geometry(M, C, &g)) return (int)cudaErrorInvalidValue;
  if (!bn_aligned(x) || !bn_aligned(y) || !bn_aligned(res) || !bn_aligned(gamma) || !bn_aligned(beta) ||
      !bn_aligned(save_mean) || !bn_aligned(save_istd) || !bn_aligned(accum))
    return (int)cudaErrorMisalignedAddress;
  int rc = bnf_prepare();
  if (rc) return rc;
  const int T = g.threads;
  int cached = (int)(bnf_smem_budget(T) / ((size_t)T * 16));
  if (g.iters < cached) cached = g.iters;
  size_t smem = (size_t)cached * T * 16;
  double inv_m = 1.0 / (double)M, unbias = M > 1 ? (double)M / (double)(M - 1) : 1.0;
  void* args[] = {&x, &res, &y, &gamma, &beta, &save_mean, &save_istd, &running_mean, &running_var, &num_batches,
                  &accum, &bar, &M, &C, &g.rpb, &g.qshift, &cached, &eps, &momentum, &inv_m, &unbias, &flags, &mask};
  void* fn = T == 1024 ? (void*)k_bn_fwd_fused<1024> : (void*)k_bn_fwd_fused<512>;
  return bn_coop_result(cudaLaunchCooperativeKernel(fn, g.grid, dim3(T), args, smem,
                                          (cudaStream_t)stream));
}

namespace {
int bn_bwd_fused_launch(const float* dy, const float* dy2, const float* x, const float* y, const float* gamma,
                        const float* beta, const float* save_mean, const float* save_istd, float* dx, float* dres,
                        float* dgamma, float* dbeta, double* accum, unsigned long long* bar, int64_t M, int C,
                        int flags, BnPool P, bool pool, void* stream) {
  BnfGeom g;
  if (!bnf_geometry(M, C, &g)) return (int)cudaErrorInvalidValue;
  if (!bn_aligned(dy) || !bn_aligned(dy2) || !bn_aligned(x) || !bn_aligned(y) || !bn_aligned(dx) ||
      !bn_aligned(dres) || !bn_aligned(gamma) || !bn_aligned(beta) || !bn_aligned(save_mean) ||
      !bn_aligned(save_istd) || !bn_aligned(accum) || !bn_aligned(P.dy))
    return (int)cudaErrorMisalignedAddress;
  if (dres != nullptr && (flags & 1) && y == nullptr) return (int)cudaErrorInvalidValue;
  int rc = bnf_prepare();
  if (rc) return rc;
  const int T = g.threads;
  int cached = (int)(bnf_smem_budget(T) / ((size_t)T * 32));
  if (g.iters < cached) cached = g.iters;
  size_t smem = (size_t)cached * T * 32;
  double inv_m = 1.0 / (double)M;
  void* args[] = {&dy, &dy2, &x, &y, &gamma, &beta, &save_mean, &save_istd, &dx, &dres, &dgamma, &dbeta, &accum,
                  &bar, &M, &C, &g.rpb, &g.qshift, &cached, &inv_m, &flags, &P};
  void* fn;
  if (T == 1024)
    fn = pool ? (void*)k_bn_bwd_fused<1024, true, false>
              : (dy2 != nullptr ? (void*)k_bn_bwd_fused<1024, false, true> : (void*)k_bn_bwd_fused<1024, false, false>);
  else
    fn = pool ? (void*)k_bn_bwd_fused<512, true, false>
              : (dy2 != nullptr ? (void*)k_bn_bwd_fused<512, false, true> : (void*)k_bn_bwd_fused<512, false, false>);
  return bn_coop_result(cudaLaunchCooperativeKernel(fn, g.grid, dim3(T), args, smem, (cudaStream_t)stream));
}
}  // namespace

extern "C" int gk_bn_bwd_fused(const float* dy, const float* x, const float* y, const float* gamma,
                               const float* beta, const float* save_mean, const float* save_istd, float* dx,
                               float* dres, float* dgamma, float* dbeta, double* accum, unsigned long long* bar,
                               int64_t M, int C, int flags, void* stream) {
  BnPool none = {nullptr, nullptr, 0, 0, 0, 0};
  return bn_bwd_fused_launch(dy, nullptr, x, y, gamma, beta, save_mean, save_istd, dx, dres, dgamma, dbeta, accum,
                             bar, M, C, flags, none, false, stream);
}

extern "C" int gk_bn_bwd_fused2(const float* dy, const float* dy2, const float* x, const float* y,
                                const float* gamma, const float* beta, const float* save_mean,
                                const float* save_istd, float* dx, float* dres, float* dgamma, float* dbeta,
                                double* accum, unsigned long long* bar, int64_t M, int C, int flags, void* stream) {
  BnPool none = {nullptr, nullptr, 0, 0, 0, 0};
  return bn_bwd_fused_launch(dy, dy2, x, y, gamma, beta, save_mean, save_istd, dx, dres, dgamma, dbeta, accum, bar,
                             M, C, flags, none, false, stream);
}

// ------------------------------------------------------------- BN + ReLU + 3x3/2 max pool (ResNet stem)
// Forward: statistics with k_bn_stats, then one pass that normalises the nine inputs of every pooled pixel
// on the fly (they are still in L2 from the statistics pass), writes the pooled activation and the window
// position of each maximum (first maximum in scan order, like at::max_pool2d).  The full-resolution
// activation is never written.  Backward: k_bn_bwd_fused<true> gathers the pooled gradient back.
// Grid: x = (pooled column, channel quad) in blocks of 256 threads, y = (image, pooled row): no 64-bit divisions
// on the way to the nine taps (the first version decomposed one flat 64-bit index per thread and was
// instruction-bound: 48 us for 103 MB).
__global__ void __launch_bounds__(256) k_bn_pool_apply(const float* __restrict__ x, float* __restrict__ yp,
                                                       unsigned char* __restrict__ pidx,
                                                       const float* __restrict__ gamma, const float* __restrict__ beta,
                                                       const float* __restrict__ save_mean,
                                                       const float* __restrict__ save_istd, int N, int H, int W, int C,
                                                       int OH, int OW) {
  const unsigned quads = (unsigned)C >> 2;
  const unsigned t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (unsigned)OW * quads) return;
  const unsigned ow = t / quads, q = t - ow * quads;
  float4 mu, is, sc, sh;
  bn_load_channel_consts(gamma, beta, save_mean, save_istd, (int)q, mu, is, sc, sh);
  for (unsigned row = blockIdx.y; row < (unsigned)(N * OH); row += gridDim.y) {
  const unsigned n = row / (unsigned)OH, oh = row - n * (unsigned)OH;
  float4 best = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
  uchar4 bi = make_uchar4(255, 255, 255, 255);
  const float* xin = x + (size_t)n * H * W * C + 4 * q;
#pragma unroll
  for (int kh = 0; kh < 3; ++kh) {
    const int h = 2 * (int)oh - 1 + kh;
    if (h < 0 || h >= H) continue;
#pragma unroll
    for (int kw = 0; kw < 3; ++kw) {
      const int w = 2 * (int)ow - 1 + kw;
      if (w < 0 || w >= W) continue;
      const float4 z = f4_fma(bn_ld4(xin + (size_t)(unsigned)(h * W + w) * C), sc, sh);
      const unsigned char code = (unsigned char)(kh * 3 + kw);
      if (z.x > best.x || bi.x == 255) { best.x = z.x; bi.x = code; }
      if (z.y > best.y || bi.y == 255) { best.y = z.y; bi.y = code; }
      if (z.z > best.z || bi.z == 255) { best.z = z.z; bi.z = code; }
      if (z.w > best.w || bi.w == 255) { best.w = z.w; bi.w = code; }
    }
  }
  const size_t o = ((size_t)row * OW + ow) * C + 4 * q;
  bn_st4(yp + o, make_float4(fmaxf(best.x, 0.f), fmaxf(best.y, 0.f), fmaxf(best.z, 0.f), fmaxf(best.w, 0.f)));
  *reinterpret_cast<uchar4*>(pidx + o) = bi;
  }
}

static int bn_pool_apply_launch(const float* x, float* yp, unsigned char* pidx, const float* gamma, const float* beta,
                                const float* save_mean, const float* save_istd, int N, int H, int W, int C, int OH,
                                int OW, cudaStream_t s) {
  if ((int64_t)N * OH >= (1LL << 31) || (int64_t)H * W >= (1LL << 31) || (int64_t)OW * (C / 4) >= (1LL << 31))
    return (int)cudaErrorInvalidValue;
  const int64_t rows = (int64_t)N * OH;
  const int64_t per_row = (int64_t)OW * (C / 4);
  const int T = per_row % 256 == 0 || per_row % 256 > 128 ? 256 : 128;    // 56 x 16 quads: 7 blocks of 128, none half empty
  dim3 grid((unsigned)((per_row + T - 1) / T), (unsigned)(rows < 65535 ? rows : 65535));
  k_bn_pool_apply<<<grid, T, 0, s>>>(x, yp, pidx, gamma, beta, save_mean, save_istd, N, H, W, C, OH, OW);
  return (int)cudaGetLastError();
}

extern "C" int gk_bn_pool_fwd(const float* x, float* yp, unsigned char* pidx, const float* gamma, const float* beta,
                              float* running_mean, float* running_var, int64_t* num_batches, float* save_mean,
                              float* save_istd, double* accum, uint32_t* ticket, unsigned long long* bar, int N,
                              int H, int W, int C, float eps, float momentum, void* stream) {
  const int64_t M = (int64_t)N * H * W;
  const int OH = (H - 1) / 2 + 1, OW = (W - 1) / 2 + 1;
  BnGeom g;
  if (!bn_geometry(M, C, &g) || H < 2 || W < 2) return (int)cudaErrorInvalidValue;
  if (!bn_aligned(x) || !bn_aligned(yp) || !bn_aligned(gamma) || !bn_aligned(beta) || !bn_aligned(save_mean) ||
      !bn_aligned(save_istd) || (reinterpret_cast<uintptr_t>(pidx) & 3))
    return (int)cudaErrorMisalignedAddress;
  cudaStream_t s = (cudaStream_t)stream;
  if (bar != nullptr && gk_bn_fused_supported(M, C)) {
    // statistics through the cooperative kernel (y = NULL: phase 1 and the per-channel results only)
    int rc = gk_bn_fwd_fused(x, nullptr, nullptr, gamma, beta, running_mean, running_var, num_batches, save_mean,
                             save_istd, accum, bar, M, C, eps, momentum, 0, stream);
    if (rc) return rc;
  } else {
    k_bn_stats<<<g.grid, BN_T, 0, s>>>(x, M, C, g.rpb, g.qpb, accum, ticket, save_mean, save_istd, running_mean,
                                       running_var, num_batches, eps, momentum);
  }
  return bn_pool_apply_launch(x, yp, pidx, gamma, beta, save_mean, save_istd, N, H, W, C, OH, OW, s);
}

// Batch statistics from per-block partial sums that somebody else accumulated (the stem convolution's epilogue,
// csrc/gk_stem.cu): partials[p][0..C) sums, partials[p][C..2C) sums of squares.  Same finalisation as k_bn_stats.
__global__ void __launch_bounds__(256) k_bn_finalize_partials(const float* __restrict__ partials, int nparts, int64_t M, int C,
                                                               float* __restrict__ save_mean, float* __restrict__ save_istd,
                                                               float* running_mean, float* running_var, int64_t* num_batches,
                                                               float eps, float momentum) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < C) {
    double s = 0, ss = 0;
    for (int p = 0; p < nparts; ++p) {
      s += (double)__ldg(partials + (size_t)p * 2 * C + c);
      ss += (double)__ldg(partials + (size_t)p * 2 * C + C + c);
    }
    const double mean = s / (double)M;
    double var = ss / (double)M - mean * mean;
    if (var < 0) var = 0;
    save_mean[c] = (float)mean;
    save_istd[c] = (float)(1.0 / sqrt(var + (double)eps));
    if (running_mean != nullptr) {
      const double unbiased = M > 1 ? var * (double)M / (double)(M - 1) : var;
      running_mean[c] = (float)((1.0 - momentum) * (double)running_mean[c] + momentum * mean);
      running_var[c] = (float)((1.0 - momentum) * (double)running_var[c] + momentum * unbiased);
    }
  }
  if (c == 0 && num_batches != nullptr) *num_batches += 1;
}

// gk_bn_pool_fwd without the statistics pass: the sums come from `partials` ([nparts][2C] floats).
extern "C" int gk_bn_pool_fwd_sums(const float* x, float* yp, unsigned char* pidx, const float* gamma, const float* beta,
                                   float* running_mean, float* running_var, int64_t* num_batches, float* save_mean,
                                   float* save_istd, const float* partials, int nparts, int N, int H, int W, int C,
                                   float eps, float momentum, void* stream) {
  const int64_t M = (int64_t)N * H * W;
  const int OH = (H - 1) / 2 + 1, OW = (W - 1) / 2 + 1;
  if (C % 4 != 0 || C <= 0 || H < 2 || W < 2 || nparts <= 0 || partials == nullptr) return (int)cudaErrorInvalidValue;
  if (!bn_aligned(x) || !bn_aligned(yp) || !bn_aligned(gamma) || !bn_aligned(beta) || !bn_aligned(save_mean) ||
      !bn_aligned(save_istd) || (reinterpret_cast<uintptr_t>(pidx) & 3))
    return (int)cudaErrorMisalignedAddress;
  cudaStream_t s = (cudaStream_t)stream;
  k_bn_finalize_partials<<<(C + 255) / 256, 256, 0, s>>>(partials, nparts, M, C, save_mean, save_istd, running_mean,
                                                        running_var, num_batches, eps, momentum);
  return bn_pool_apply_launch(x, yp, pidx, gamma, beta, save_mean, save_istd, N, H, W, C, OH, OW, s);
}

// Full-resolution gradient of the max pool (dense gather, one thread per input pixel and channel quad): the
// BN backward then runs on it like on any other gradient.  Cheaper than gathering inside the cooperative
// kernel, which has one block per SM (16-32 warps) and would evaluate the windows twice.
// One thread per 2 x 2 block of input pixels and channel quad: the four pixels (2j, 2k), (2j, 2k+1), (2j+1, 2k),
// (2j+1, 2k+1) lie in the windows (oh, ow) with oh in {j, j+1}, ow in {k, k+1} and in no other, so four loads of
// (position code, gradient) serve four outputs (separately they take 1 + 2 + 2 + 4).  Grid: x = (column pair,
// channel quad), y = (image, row pair); 32-bit arithmetic only.  (The flat-index, one-pixel-per-thread version was
// instruction-bound: 62 us for 26 MB in, 103 MB out.)
__global__ void __launch_bounds__(256) k_pool_gather(BnPool P, float* __restrict__ g, int N, int C) {
  const unsigned quads = (unsigned)C >> 2;
  const unsigned W2 = ((unsigned)P.W + 1u) >> 1, H2 = ((unsigned)P.H + 1u) >> 1;
  const unsigned t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= W2 * quads) return;
  const unsigned k = t / quads, q = t - k * quads;
  for (unsigned row = blockIdx.y; row < (unsigned)N * H2; row += gridDim.y) {
    const unsigned n = row / H2, j = row - n * H2;
    float4 acc[2][2];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b) acc[a][b] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int dh = 0; dh < 2; ++dh) {
      const unsigned oh = j + dh;                          // window oh covers input rows 2*oh-1 .. 2*oh+1
      if (oh >= (unsigned)P.OH) continue;
#pragma unroll
      for (int dw = 0; dw < 2; ++dw) {
        const unsigned ow = k + dw;
        if (ow >= (unsigned)P.OW) continue;
        const size_t o = ((size_t)(n * (unsigned)P.OH + oh) * P.OW + ow) * C + 4 * q;
        const uchar4 id = *reinterpret_cast<const uchar4*>(P.idx + o);
        const float4 d = bn_ld4(P.dy + o);
        // pixel (2j + a, 2k + b) sits at window position (2j + a - (2 oh - 1), 2k + b - (2 ow - 1)); with oh = j + dh the
        // row offset is a + 1 - 2 dh: rows a = 0, 1 for dh = 0 (offsets 1, 2), only a = 1 for dh = 1 (offset 0)
#pragma unroll
        for (int a = dh; a < 2; ++a)
#pragma unroll
          for (int b = dw; b < 2; ++b) {
            const unsigned code = (unsigned)((a + 1 - 2 * dh) * 3 + (b + 1 - 2 * dw));
            acc[a][b].x += id.x == code ? d.x : 0.f; acc[a][b].y += id.y == code ? d.y : 0.f;
            acc[a][b].z += id.z == code ? d.z : 0.f; acc[a][b].w += id.w == code ? d.w : 0.f;
          }
      }
    }
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      const unsigned h = 2 * j + a;
      if (h >= (unsigned)P.H) continue;
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const unsigned w = 2 * k + b;
        if (w < (unsigned)P.W) bn_st4(g + ((size_t)(n * (unsigned)P.H + h) * P.W + w) * C + 4 * q, acc[a][b]);
      }
    }
  }
}

extern "C" int gk_bn_pool_bwd(const float* dyp, const unsigned char* pidx, const float* x, const float* gamma,
                              const float* beta, const float* save_mean, const float* save_istd, float* dx,
                              float* gfull, float* dgamma, float* dbeta, double* accum, unsigned long long* bar,
                              int N, int H, int W, int C, int flags, void* stream) {
  const int OH = (H - 1) / 2 + 1, OW = (W - 1) / 2 + 1;
  const int64_t M = (int64_t)N * H * W;
  BnPool P = {dyp, pidx, H, W, OH, OW};
  if (gfull == nullptr)      // gather inside the cooperative kernel (no extra buffer)
    return bn_bwd_fused_launch(nullptr, nullptr, x, nullptr, gamma, beta, save_mean, save_istd, dx, nullptr, dgamma,
                               dbeta, accum, bar, M, C, flags | 1, P, true, stream);
  if (!bn_aligned(gfull) || !bn_aligned(dyp) || (reinterpret_cast<uintptr_t>(pidx) & 3) || (C & 3))
    return (int)cudaErrorMisalignedAddress;
  if ((int64_t)N * H >= (1LL << 31) || (int64_t)W * (C / 4) >= (1LL << 31)) return (int)cudaErrorInvalidValue;
  {
    const int64_t rows = (int64_t)N * ((H + 1) / 2), per_row = (int64_t)((W + 1) / 2) * (C / 4);
    const int T = per_row % 256 == 0 || per_row % 256 > 128 ? 256 : 128;
    dim3 grid((unsigned)((per_row + T - 1) / T), (unsigned)(rows < 65535 ? rows : 65535));
    k_pool_gather<<<grid, T, 0, (cudaStream_t)stream>>>(P, gfull, N, C);
  }
  BnPool none = {nullptr, nullptr, 0, 0, 0, 0};
  return bn_bwd_fused_launch(gfull, nullptr, x, nullptr, gamma, beta, save_mean, save_istd, dx, nullptr, dgamma,
                             dbeta, accum, bar, M, C, flags | 1, none, false, stream);
}

// ===================================================== two batch norms into one activation (downsample blocks)
// y = relu( bnA(xa) + bnB(xb) ): the last BN of a bottleneck and the BN of its 1x1 downsample shortcut.  Done
// separately the shortcut's normalised activation is written and read back (and its gradient too); here both
// statistics are taken in one pass over xa and xb, and y / dxa / dxb come out of one kernel per direction.
// Accumulator rows per parity: forward [sum xa, sum xa^2, sum xb, sum xb^2], backward [sum g, sum g*xhat_a,
// sum g*xhat_b, unused]; scratch = 2 parities x 4 x C doubles + the usual 128 barrier words.
struct Bn2Side {
  const float* x;
  const float* gamma;
  const float* beta;
  float* save_mean;
  float* save_istd;
  float* running_mean;
  float* running_var;
  int64_t* num_batches;
  float eps, momentum;
};

__device__ __forceinline__ void bn2_consts(const double* s, const double* ss, double inv_m, float eps, float* mu,
                                           float* is, double* meanv, double* varv) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const double mean = s[j] * inv_m;
    double var = fma(ss[j], inv_m, -mean * mean);
    if (var < 0) var = 0;
    meanv[j] = mean;
    varv[j] = var;
    mu[j] = (float)mean;
    is[j] = bnf_rsqrt((float)var + eps);
  }
}

__device__ __forceinline__ void bn2_keep(const Bn2Side& S, int q, const float* mu, const float* is, const double* meanv,
                                         const double* varv, float4 rm, float4 rv, double unbias) {
  const float rmv[4] = {rm.x, rm.y, rm.z, rm.w}, rvv[4] = {rv.x, rv.y, rv.z, rv.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int c = 4 * q + j;
    S.save_mean[c] = mu[j];
    S.save_istd[c] = is[j];
    if (S.running_mean != nullptr) {
      S.running_mean[c] = (float)((1.0 - S.momentum) * (double)rmv[j] + S.momentum * meanv[j]);
      S.running_var[c] = (float)((1.0 - S.momentum) * (double)rvv[j] + S.momentum * (varv[j] * unbias));
    }
  }
}

template <int T_>
__global__ void __launch_bounds__(T_, 1)
k_bn2_fwd_fused(Bn2Side A, Bn2Side B, float* __restrict__ y, double* accum, unsigned long long* bar, int64_t M, int C,
                int rows_per_block, int qshift, int cached, double inv_m, double unbias, int flags,
                unsigned char* __restrict__ mask) {
  extern __shared__ float4 tile[];          // [cached][2][T_]: xa, xb
  __shared__ float4 red[2][T_];
  __shared__ unsigned long long s_gen;
  const int tid = threadIdx.x;
  const int qpb = 1 << qshift, ql = tid & (qpb - 1), rl = tid >> qshift, RL = T_ >> qshift;
  const int q = blockIdx.y * qpb + ql;
  if (tid == 0 && gridDim.x > 1) s_gen = bnf_ld_relaxed(bar + 2 * blockIdx.y + 1);
  const int relu = flags & 1;
  BNF_ROWS_DECL
  const float4 gmA = bn_ld4(A.gamma + 4 * q), btA = bn_ld4(A.beta + 4 * q);
  const float4 gmB = bn_ld4(B.gamma + 4 * q), btB = bn_ld4(B.beta + 4 * q);
  const bool keeper = blockIdx.x == 0 && rl == 0;
  float4 rmA = make_float4(0.f, 0.f, 0.f, 0.f), rvA = rmA, rmB = rmA, rvB = rmA;
  if (keeper && A.running_mean != nullptr) { rmA = bn_ld4(A.running_mean + 4 * q); rvA = bn_ld4(A.running_var + 4 * q); }
  if (keeper && B.running_mean != nullptr) { rmB = bn_ld4(B.running_mean + 4 * q); rvB = bn_ld4(B.running_var + 4 * q); }

  const int nc = n_it < cached ? n_it : cached;
  for (int i = 0; i < nc; ++i) {
    const int64_t o = (rbase + (int64_t)i * rstep) * C + 4 * q;
    bnf_cp_async16(&tile[(2 * i) * T_ + tid], A.x + o);
    bnf_cp_async16(&tile[(2 * i + 1) * T_ + tid], B.x + o);
  }
  float4 a1 = make_float4(0.f, 0.f, 0.f, 0.f), a2 = a1, b1 = a1, b2 = a1;
#pragma unroll 4
  for (int i = nc; i < n_it; ++i) {
    const int64_t o = (rbase + (int64_t)i * rstep) * C + 4 * q;
    const float4 xa = bn_ld4(A.x + o), xb = bn_ld4(B.x + o);
    a1 = f4_add(a1, xa); a2 = f4_fma(xa, xa, a2);
    b1 = f4_add(b1, xb); b2 = f4_fma(xb, xb, b2);
  }
  bnf_cp_async_wait_all();
#pragma unroll 4
  for (int i = 0; i < nc; ++i) {
    const float4 xa = tile[(2 * i) * T_ + tid], xb = tile[(2 * i + 1) * T_ + tid];
    a1 = f4_add(a1, xa); a2 = f4_fma(xa, xa, a2);
    b1 = f4_add(b1, xb); b2 = f4_fma(xb, xb, b2);
  }
  const bool alone = gridDim.x == 1;
  double sA[4], ssA[4], sB[4], ssB[4];
  bnf_reduce_rows(red, a1, a2, tid, rl, RL, qpb);
  if (alone) {
    const float4 ta = red[0][ql], tb = red[1][ql];
    sA[0] = ta.x; sA[1] = ta.y; sA[2] = ta.z; sA[3] = ta.w;
    ssA[0] = tb.x; ssA[1] = tb.y; ssA[2] = tb.z; ssA[3] = tb.w;
  }
  __syncthreads();
  bnf_reduce_rows(red, b1, b2, tid, rl, RL, qpb);
  const unsigned long long gen = alone ? 0ULL : s_gen;
  double* acc = accum + (gen & 1ULL) * 4 * (size_t)C;
  if (alone) {
    const float4 ta = red[0][ql], tb = red[1][ql];
    sB[0] = ta.x; sB[1] = ta.y; sB[2] = ta.z; sB[3] = ta.w;
    ssB[0] = tb.x; ssB[1] = tb.y; ssB[2] = tb.z; ssB[3] = tb.w;
  } else {
    if (rl == 0) {
      bnf_atomic_add8(acc, C, q, a1, a2);
      bnf_atomic_add8(acc + 2 * (size_t)C, C, q, b1, b2);
    }
    bnf_column_barrier(bar, gen);
    bnf_load_totals(acc, C, q, sA, ssA);
    bnf_load_totals(acc + 2 * (size_t)C, C, q, sB, ssB);
  }
  float muA[4], isA[4], muB[4], isB[4];
  double meanA[4], varA[4], meanB[4], varB[4];
  bn2_consts(sA, ssA, inv_m, A.eps, muA, isA, meanA, varA);
  bn2_consts(sB, ssB, inv_m, B.eps, muB, isB, meanB, varB);
  const float4 scA = make_float4(gmA.x * isA[0], gmA.y * isA[1], gmA.z * isA[2], gmA.w * isA[3]);
  const float4 shA = make_float4(fmaf(-muA[0], scA.x, btA.x), fmaf(-muA[1], scA.y, btA.y), fmaf(-muA[2], scA.z, btA.z),
                                 fmaf(-muA[3], scA.w, btA.w));
  const float4 scB = make_float4(gmB.x * isB[0], gmB.y * isB[1], gmB.z * isB[2], gmB.w * isB[3]);
  const float4 shB = make_float4(fmaf(-muB[0], scB.x, btB.x), fmaf(-muB[1], scB.y, btB.y), fmaf(-muB[2], scB.z, btB.z),
                                 fmaf(-muB[3], scB.w, btB.w));
#pragma unroll 4
  for (int i = n_it - 1; i >= 0; --i) {
    const int64_t o = (rbase + (int64_t)i * rstep) * C + 4 * q;
    const float4 xa = (i < nc) ? tile[(2 * i) * T_ + tid] : bn_ld4(A.x + o);
    const float4 xb = (i < nc) ? tile[(2 * i + 1) * T_ + tid] : bn_ld4(B.x + o);
    float4 v = f4_add(f4_fma(xa, scA, shA), f4_fma(xb, scB, shB));
    if (mask != nullptr) mask[(rbase + (int64_t)i * rstep) * (C >> 2) + q] = bn_mask_bits(v);
    if (relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
    bn_st4(y + o, v);
  }
  if (keeper) {
    if (!alone) {
      double* other = accum + ((gen + 1ULL) & 1ULL) * 4 * (size_t)C;
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int j = 0; j < 4; ++j) other[r * (size_t)C + 4 * q + j] = 0;
    }
    bn2_keep(A, q, muA, isA, meanA, varA, rmA, rvA, unbias);
    bn2_keep(B, q, muB, isB, meanB, varB, rmB, rvB, unbias);
    if (blockIdx.y == 0 && tid == 0) {
      if (A.num_batches != nullptr) *A.num_batches += 1;
      if (B.num_batches != nullptr) *B.num_batches += 1;
    }
  }
}

struct Bn2Back {
  const float* x;
  const float* gamma;
  const float* beta;
  const float* save_mean;
  const float* save_istd;
  float* dx;
  float* dgamma;
  float* dbeta;
};

// g = (dy [+ dy2]) masked by y > 0; dxa, dxb, dgamma/dbeta of both sides.  The masked gradient of the rows that do
// not fit in shared memory is parked in A.dx between the phases (same thread, same address).
template <int T_>
__global__ void __launch_bounds__(T_, 1)
k_bn2_bwd_fused(const float* __restrict__ dy, const float* __restrict__ dy2, const float* __restrict__ y, Bn2Back A,
                Bn2Back B, double* accum, unsigned long long* bar, int64_t M, int C, int rows_per_block, int qshift,
                int cached, double inv_m, int flags) {
  extern __shared__ float4 tile[];          // [cached][3][T_]: g, xa, xb
  __shared__ float4 red[2][T_];
  __shared__ unsigned long long s_gen;
  const int tid = threadIdx.x;
  const int qpb = 1 << qshift, ql = tid & (qpb - 1), rl = tid >> qshift, RL = T_ >> qshift;
  const int q = blockIdx.y * qpb + ql;
  if (tid == 0 && gridDim.x > 1) s_gen = bnf_ld_relaxed(bar + 2 * blockIdx.y + 1);
  const int relu = flags & 1;
  const bool mbits = (flags & 4) != 0;
  BNF_ROWS_DECL
  float4 muA, isA, scA, shA, muB, isB, scB, shB;
  bn_load_channel_consts(A.gamma, A.beta, A.save_mean, A.save_istd, q, muA, isA, scA, shA);
  bn_load_channel_consts(B.gamma, B.beta, B.save_mean, B.save_istd, q, muB, isB, scB, shB);

  const int nc = n_it < cached ? n_it : cached;
  for (int i = 0; i < nc; ++i) {
    const int64_t o = (rbase + (int64_t)i * rstep) * C + 4 * q;
    bnf_cp_async16(&tile[(3 * i) * T_ + tid], dy + o);
    bnf_cp_async16(&tile[(3 * i + 1) * T_ + tid], A.x + o);
    bnf_cp_async16(&tile[(3 * i + 2) * T_ + tid], B.x + o);
  }
  float4 sg = make_float4(0.f, 0.f, 0.f, 0.f), sa = sg, sb = sg;
  auto xhat = [](float4 x, float4 mu, float4 is) {
    return make_float4((x.x - mu.x) * is.x, (x.y - mu.y) * is.y, (x.z - mu.z) * is.z, (x.w - mu.w) * is.w);
  };
#pragma unroll 2
  for (int i = nc; i < n_it; ++i) {
    const int64_t o = (rbase + (int64_t)i * rstep) * C + 4 * q;
    const float4 xa = bn_ld4(A.x + o), xb = bn_ld4(B.x + o);
    float4 g = bn_ld4(dy + o);
    if (dy2 != nullptr) g = f4_add(g, bn_ld4(dy2 + o));
    if (relu) g = bn_mask_pos(g, bn_mask_src(y, o, rbase + (int64_t)i * rstep, C, q, mbits));
    bn_st4(A.dx + o, g);
    sg = f4_add(sg, g);
    sa = f4_fma(g, xhat(xa, muA, isA), sa);
    sb = f4_fma(g, xhat(xb, muB, isB), sb);
  }
  bnf_cp_async_wait_all();
#pragma unroll 2
  for (int i = 0; i < nc; ++i) {
    const int64_t o = (rbase + (int64_t)i * rstep) * C + 4 * q;
    const float4 xa = tile[(3 * i + 1) * T_ + tid], xb = tile[(3 * i + 2) * T_ + tid];
    float4 g = tile[(3 * i) * T_ + tid];
    if (dy2 != nullptr) g = f4_add(g, bn_ld4(dy2 + o));
    if (relu) g = bn_mask_pos(g, bn_mask_src(y, o, rbase + (int64_t)i * rstep, C, q, mbits));
    tile[(3 * i) * T_ + tid] = g;
    sg = f4_add(sg, g);
    sa = f4_fma(g, xhat(xa, muA, isA), sa);
    sb = f4_fma(g, xhat(xb, muB, isB), sb);
  }
  const bool alone = gridDim.x == 1;
  float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
  double tg[4], ta_[4], tb_[4], dummy[4];
  bnf_reduce_rows(red, sg, sa, tid, rl, RL, qpb);
  if (alone) {
    const float4 u = red[0][ql], v = red[1][ql];
    tg[0] = u.x; tg[1] = u.y; tg[2] = u.z; tg[3] = u.w;
    ta_[0] = v.x; ta_[1] = v.y; ta_[2] = v.z; ta_[3] = v.w;
  }
  __syncthreads();
  bnf_reduce_rows(red, sb, zero, tid, rl, RL, qpb);
  const unsigned long long gen = alone ? 0ULL : s_gen;
  double* acc = accum + (gen & 1ULL) * 4 * (size_t)C;
  if (alone) {
    const float4 u = red[0][ql];
    tb_[0] = u.x; tb_[1] = u.y; tb_[2] = u.z; tb_[3] = u.w;
  } else {
    if (rl == 0) {
      bnf_atomic_add8(acc, C, q, sg, sa);
      atomicAdd(acc + 2 * (size_t)C + 4 * q + 0, (double)sb.x); atomicAdd(acc + 2 * (size_t)C + 4 * q + 1, (double)sb.y);
      atomicAdd(acc + 2 * (size_t)C + 4 * q + 2, (double)sb.z); atomicAdd(acc + 2 * (size_t)C + 4 * q + 3, (double)sb.w);
    }
    bnf_column_barrier(bar, gen);
    bnf_load_totals(acc, C, q, tg, ta_);
    bnf_load_totals(acc + 2 * (size_t)C, C, q, tb_, dummy);
  }
  float mg[4], mga[4], mgb[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    mg[j] = (float)(tg[j] * inv_m);
    mga[j] = (float)(ta_[j] * inv_m);
    mgb[j] = (float)(tb_[j] * inv_m);
  }
#pragma unroll 2
  for (int i = n_it - 1; i >= 0; --i) {
    const int64_t o = (rbase + (int64_t)i * rstep) * C + 4 * q;
    float4 g, xa, xb;
    if (i < nc) {
      g = tile[(3 * i) * T_ + tid];
      xa = tile[(3 * i + 1) * T_ + tid];
      xb = tile[(3 * i + 2) * T_ + tid];
    } else {
      g = bn_ld4(A.dx + o);
      xa = bn_ld4(A.x + o);
      xb = bn_ld4(B.x + o);
    }
    float4 da, db;
    da.x = scA.x * (g.x - mg[0] - (xa.x - muA.x) * isA.x * mga[0]);
    da.y = scA.y * (g.y - mg[1] - (xa.y - muA.y) * isA.y * mga[1]);
    da.z = scA.z * (g.z - mg[2] - (xa.z - muA.z) * isA.z * mga[2]);
    da.w = scA.w * (g.w - mg[3] - (xa.w - muA.w) * isA.w * mga[3]);
    db.x = scB.x * (g.x - mg[0] - (xb.x - muB.x) * isB.x * mgb[0]);
    db.y = scB.y * (g.y - mg[1] - (xb.y - muB.y) * isB.y * mgb[1]);
    db.z = scB.z * (g.z - mg[2] - (xb.z - muB.z) * isB.z * mgb[2]);
    db.w = scB.w * (g.w - mg[3] - (xb.w - muB.w) * isB.w * mgb[3]);
    bn_st4(A.dx + o, da);
    bn_st4(B.dx + o, db);
  }
  if (blockIdx.x == 0 && rl == 0) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = 4 * q + j;
      if (!alone) {
        double* other = accum + ((gen + 1ULL) & 1ULL) * 4 * (size_t)C;
        other[c] = 0; other[C + c] = 0; other[2 * (size_t)C + c] = 0; other[3 * (size_t)C + c] = 0;
      }
      A.dbeta[c] = (float)tg[j];
      B.dbeta[c] = (float)tg[j];
      A.dgamma[c] = (float)ta_[j];
      B.dgamma[c] = (float)tb_[j];
    }
  }
}

namespace {
int bn2_prepare() {
  static bool ready[64] = {false};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= 64) return (int)cudaErrorInvalidDevice;
  if (!ready[dev]) {
    void* fns[4] = {(void*)k_bn2_fwd_fused<512>, (void*)k_bn2_bwd_fused<512>, (void*)k_bn2_fwd_fused<1024>,
                    (void*)k_bn2_bwd_fused<1024>};
    for (int i = 0; i < 4; ++i) {
      e = cudaFuncSetAttribute(fns[i], cudaFuncAttributeMaxDynamicSharedMemorySize,
                               (int)bnf_smem_budget(i < 2 ? 512 : 1024));
      if (e != cudaSuccess) return (int)e;
    }
    ready[dev] = true;
  }
  return 0;
}
}  // namespace

extern "C" int gk_bn2_fwd_fused(void* const* a, void* const* b, float eps_a, float mom_a, float eps_b, float mom_b,
                                float* y, unsigned char* mask, double* accum, unsigned long long* bar, int64_t M, int C,
                                int flags, void* stream) {
  BnfGeom g;
  if (!bnf_geometry(M, C, &g, true)) return (int)cudaErrorInvalidValue;
  for (int i = 0; i < 5; ++i)
    if (!bn_aligned(a[i]) || !bn_aligned(b[i])) return (int)cudaErrorMisalignedAddress;
  if (!bn_aligned(y) || !bn_aligned(accum)) return (int)cudaErrorMisalignedAddress;
  int rc = bn2_prepare();
  if (rc) return rc;
  Bn2Side A = {(const float*)a[0], (const float*)a[1], (const float*)a[2], (float*)a[3], (float*)a[4], (float*)a[5],
               (float*)a[6], (int64_t*)a[7], eps_a, mom_a};
  Bn2Side B = {(const float*)b[0], (const float*)b[1], (const float*)b[2], (float*)b[3], (float*)b[4], (float*)b[5],
               (float*)b[6], (int64_t*)b[7], eps_b, mom_b};
  const int T = g.threads;
  int cached = (int)(bnf_smem_budget(T) / ((size_t)T * 32));
  if (g.iters < cached) cached = g.iters;
  size_t smem = (size_t)cached * T * 32;
  double inv_m = 1.0 / (double)M, unbias = M > 1 ? (double)M / (double)(M - 1) : 1.0;
  void* args[] = {&A, &B, &y, &accum, &bar, &M, &C, &g.rpb, &g.qshift, &cached, &inv_m, &unbias, &flags, &mask};
  void* fn = T == 1024 ? (void*)k_bn2_fwd_fused<1024> : (void*)k_bn2_fwd_fused<512>;
  return bn_coop_result(cudaLaunchCooperativeKernel(fn, g.grid, dim3(T), args, smem,
                                          (cudaStream_t)stream));
}

extern "C" int gk_bn2_bwd_fused(const float* dy, const float* dy2, const float* y, void* const* a, void* const* b,
                                double* accum, unsigned long long* bar, int64_t M, int C, int flags, void* stream) {
  BnfGeom g;
  if (!bnf_geometry(M, C, &g, true)) return (int)cudaErrorInvalidValue;
  for (int i = 0; i < 8; ++i)
    if (!bn_aligned(a[i]) || !bn_aligned(b[i])) return (int)cudaErrorMisalignedAddress;
  if (!bn_aligned(dy) || !bn_aligned(dy2) || !bn_aligned(y) || !bn_aligned(accum)) return (int)cudaErrorMisalignedAddress;
  if ((flags & 1) && y == nullptr) return (int)cudaErrorInvalidValue;
  int rc = bn2_prepare();
  if (rc) return rc;
  Bn2Back A = {(const float*)a[0], (const float*)a[1], (const float*)a[2], (const float*)a[3], (const float*)a[4],
               (float*)a[5], (float*)a[6], (float*)a[7]};
  Bn2Back B = {(const float*)b[0], (const float*)b[1], (const float*)b[2], (const float*)b[3], (const float*)b[4],
               (float*)b[5], (float*)b[6], (float*)b[7]};
  const int T = g.threads;
  int cached = (int)(bnf_smem_budget(T) / ((size_t)T * 48));
  if (g.iters < cached) cached = g.iters;
  size_t smem = (size_t)cached * T * 48;
  double inv_m = 1.0 / (double)M;
  void* args[] = {&dy, &dy2, &y, &A, &B, &accum, &bar, &M, &C, &g.rpb, &g.qshift, &cached, &inv_m, &flags};
  void* fn = T == 1024 ? (void*)k_bn2_bwd_fused<1024> : (void*)k_bn2_bwd_fused<512>;
  return bn_coop_result(cudaLaunchCooperativeKernel(fn, g.grid, dim3(T), args, smem,
                                          (cudaStream_t)stream));
}
