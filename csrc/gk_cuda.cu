// gk_cuda.cu — sm_100a kernels of the sparsified-gradient chain (see include/gk_b200.h).
//
// Design notes (B200-first, nothing here is translated from the reference, which ships no
// native code at all — SURVEY.md §2.2):
//   * Every pass is HBM-bound, so the chain is organised around the number of times the
//     n-element arenas are touched: K1 (12-16 B/elem) + K2 (4) + K3 (4) for gaussian-k,
//     with all "decisions" (moments -> threshold -> refinement loop -> prefix sums) taken by
//     last-block-done epilogues on the device.  No host synchronisation anywhere, so the whole
//     chain can sit inside a CUDA graph together with forward/backward.
//   * Work is described by a block table (GkBlk): each CUDA block owns one contiguous,
//     16-byte aligned extent of the arena that belongs to exactly one selection segment.
//     One launch therefore serves the merged-buffer mode (1 segment), threshold groups and
//     the per-tensor mode (161 segments for ResNet-50) alike.
//   * Compaction is ordered and deterministic without spin-waiting: the counting pass leaves
//     per-block counts, its epilogue prefix-sums them, the compaction pass re-evaluates the
//     predicate and writes pair j of block b at prefix[b]+j.
//   * The compaction kernel IS the send side of the sparse allgather: it stores the 8-byte
//     (index,value) pairs straight into the receive slot of every destination rank (peer
//     mapped NVLink addresses) and publishes {count, seq} with a system-scope release store.
//     The receive kernel acquires those headers (bounded spin) and scatter-adds from local HBM.
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "../include/gk_b200.h"

#define T GK_THREADS
#define NWARP (GK_THREADS / 32)

// ----------------------------------------------------------------------------- helpers
__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ float4 ld4_stream(const float* p) {
  return __ldcs(reinterpret_cast<const float4*>(p));
}
__device__ __forceinline__ void st4_stream(float* p, float4 v) { __stcs(reinterpret_cast<float4*>(p), v); }
__device__ __forceinline__ uint32_t absbits(float x) { return __float_as_uint(x) & 0x7fffffffu; }

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ uint32_t warp_sum_u(uint32_t v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ uint32_t warp_incl_scan_u(uint32_t v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  return v;
}

// Ticket taken by every block when its global writes are done; true in the last block.
// The ticket word resets itself so the same control block serves every launch.
__device__ __forceinline__ bool last_block_done(uint32_t* ticket) {
  __shared__ int s_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = atomicAdd(ticket, 1u);
    s_last = (t == gridDim.x - 1);
    if (s_last) *ticket = 0;
  }
  __syncthreads();
  if (s_last) __threadfence();
  return s_last != 0;
}

__device__ __forceinline__ void st_release_sys_u64(uint64_t* p, uint64_t v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys_u64(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Magnitude key used by the radix select: larger key = selected first.
template <bool MINK>
__device__ __forceinline__ uint32_t sel_key(float x) {
  uint32_t b = absbits(x);
  return MINK ? (0x7fffffffu - b) : b;
}

// index of the gaussian refinement candidate reached after `a` halvings and `b` x1.5 steps
__device__ __forceinline__ int tri_idx(int a, int b) { return (a + b) * (a + b + 1) / 2 + a; }

// ----------------------------------------------------------------- K1 accumulate + moments
// Epilogue of K1, run by the last block: per-segment moments (fixed-order warp reduction of
// the per-block partials => deterministic), then the candidate thresholds of the counting
// pass, or the first radix digit for top-k.
__device__ void k1_epilogue(const GkChain& c) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int s = warp; s < c.nseg; s += NWARP) {
    const GkSeg sg = c.seg[s];
    double sum = 0, sq = 0, sa = 0;
    float mx = 0.f;
    for (int b = sg.blk_begin + lane; b < sg.blk_end; b += 32) {
      const double2 p0 = __ldcg(reinterpret_cast<const double2*>(c.part + 4 * (size_t)b));
      const double2 p1 = __ldcg(reinterpret_cast<const double2*>(c.part + 4 * (size_t)b + 2));
      sum += p0.x; sq += p0.y; sa += p1.x; mx = fmaxf(mx, (float)p1.y);
    }
    sum = warp_sum(sum); sq = warp_sum(sq); sa = warp_sum(sa); mx = warp_max(mx);
    if (lane == 0) {
      GkSegStat* st = c.stat + s;
      const double n = (double)sg.n;
      const double mean = sum / n;
      double var = (sg.n > 1) ? (sq - sum * mean) / (n - 1.0) : 0.0;  // unbiased, like torch.std
      if (var < 0) var = 0;
      // the reference converts the fp32 tensors mean/std to python floats (compression.py:367-369)
      const float meanf = (float)mean, stdf = (float)sqrt(var);
      const float meanabs = (float)(sa / n);
      st->mean = meanf; st->std = stdf; st->meanabs = meanabs; st->maxabs = mx;
      if (c.mode == GK_MODE_GAUSSIAN) {
        // right_thres = mu + |z| sigma in double (utils.py:136-138); the loop multiplies by
        // 0.5 (exact) or 1.5 in double, the comparison rounds the threshold to fp32.
        const double t0 = (double)meanf + sg.zabs * (double)stdf;
        const int L = c.nloops;
        double t15 = t0;
        for (int b = 0; b < L; ++b) {
          double t = t15;
          for (int a = 0; a + b < L; ++a) {
            st->cand[tri_idx(a, b)] = (float)t;
            t *= 0.5;
          }
          t15 *= 1.5;
        }
      } else if (c.mode == GK_MODE_REDSYNC) {
        // thres = mean|x| + ratio*(max|x| - mean|x|) in fp32 (compression.py:653-662); the
        // bisection of [0,1] visits a 3-level binary tree of ratios (heap order).
        const float d = __fsub_rn(mx, meanabs);
        const float ratios[7] = {0.5f, 0.25f, 0.75f, 0.125f, 0.375f, 0.625f, 0.875f};
        for (int j = 0; j < 7; ++j) st->cand[j] = __fadd_rn(meanabs, __fmul_rn(ratios[j], d));
      } else if (c.mode == GK_MODE_REDSYNCTRIM) {
        const float d = __fsub_rn(mx, meanabs);
        double r = 1.0 - 0.2;  // compression.py:711-712
        for (int j = 0; j < GK_NC; ++j) {
          st->cand[j] = __fadd_rn(meanabs, __fmul_rn((float)r, d));
          r = r - 0.2;
        }
      } else if (c.mode == GK_MODE_THRESHOLD) {
        st->cand[0] = sg.param;
      }
    }
    if (c.mode == GK_MODE_TOPK || c.mode == GK_MODE_MINK) {
      if (lane == 0) { c.stat[s].prefix = 0; c.stat[s].krem = (uint32_t)sg.k; }
    }
  }
}

// One radix-select step on segment s by one warp: find the digit whose bucket contains the
// krem-th largest key, update (prefix, krem), and clear the histogram for the next pass.
__device__ void radix_pick_warp(const GkChain& c, int s, int shift, int nbins, int lane) {
  uint32_t* h = c.hist + (size_t)s * GK_HIST_BINS;
  GkSegStat* st = c.stat + s;
  const uint32_t krem = st->krem;
  const int per = nbins / 32;  // bins per lane; lane 31 owns the highest digits
  uint32_t mysum = 0;
  for (int j = 0; j < per; ++j) mysum += __ldcg(h + lane * per + j);
  // suffix sums over lanes: above(l) = sum of lanes > l
  uint32_t incl = mysum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_down_sync(0xffffffffu, incl, o);
    if (lane + o < 32) incl += t;
  }
  const uint32_t above = incl - mysum;
  const bool mine = (above < krem) && (incl >= krem);
  const uint32_t ball = __ballot_sync(0xffffffffu, mine);
  if (ball == 0) {
    // fewer than krem keys matched the prefix (cannot happen when k <= n); take the lowest digit
    if (lane == 0) st->prefix = st->prefix << ((nbins == 2048) ? 11 : 9);
  } else if (mine) {
    uint32_t run = above;
    int digit = lane * per;
    for (int j = per - 1; j >= 0; --j) {
      const uint32_t v = __ldcg(h + lane * per + j);
      if (run + v >= krem) { digit = lane * per + j; break; }
      run += v;
    }
    const int width = (nbins == 2048) ? 11 : 9;
    st->prefix = (st->prefix << width) | (uint32_t)digit;
    st->krem = krem - run;  // rank inside the chosen bucket (1-based)
  }
  __syncwarp();
  for (int j = lane; j < nbins; j += 32) h[j] = 0;
}

template <bool HIST, bool MINK>
__global__ void __launch_bounds__(GK_THREADS) k_accumulate(GkChain c) {
  __shared__ uint32_t shist[HIST ? GK_HIST_BINS : 1];
  __shared__ double sred[3][NWARP];
  __shared__ float smax[NWARP];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const GkBlk B = c.blk[blockIdx.x];
  if (HIST) {
    for (int i = tid; i < GK_HIST_BINS; i += T) shist[i] = 0;
    __syncthreads();
  }
  const float* __restrict__ g = c.grad + B.start;
  float* __restrict__ r = c.resid + B.start;
  const bool user = c.use_residual != 0, zg = c.zero_grad != 0;
  double s = 0, q = 0, sa = 0;
  float mx = 0.f;
  const int n4 = B.len >> 2;
  const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
  for (int i = tid; i < n4; i += T) {
    float4 x = ld4_stream(g + 4 * i);
    if (user) {
      const float4 rv = ld4_stream(r + 4 * i);
      x.x += rv.x; x.y += rv.y; x.z += rv.z; x.w += rv.w;
    }
    st4(r + 4 * i, x);  // acc stays in L2 for the following passes
    if (zg) st4_stream(const_cast<float*>(g) + 4 * i, z4);
    const float e[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const double xd = (double)e[j];
      s += xd; q = fma(xd, xd, q); sa += fabs(xd);
      mx = fmaxf(mx, fabsf(e[j]));
      if (HIST) atomicAdd(&shist[sel_key<MINK>(e[j]) >> 20], 1u);
    }
  }
  for (int i = 4 * n4 + tid; i < B.len; i += T) {  // tail of a tensor whose size % 4 != 0
    float x = g[i];
    if (user) x += r[i];
    r[i] = x;
    if (zg) const_cast<float*>(g)[i] = 0.f;
    const double xd = (double)x;
    s += xd; q = fma(xd, xd, q); sa += fabs(xd);
    mx = fmaxf(mx, fabsf(x));
    if (HIST) atomicAdd(&shist[sel_key<MINK>(x) >> 20], 1u);
  }
  s = warp_sum(s); q = warp_sum(q); sa = warp_sum(sa); mx = warp_max(mx);
  if (lane == 0) { sred[0][warp] = s; sred[1][warp] = q; sred[2][warp] = sa; smax[warp] = mx; }
  __syncthreads();
  if (tid == 0) {
    double a = 0, b = 0, d = 0; float m = 0.f;
    for (int w = 0; w < NWARP; ++w) { a += sred[0][w]; b += sred[1][w]; d += sred[2][w]; m = fmaxf(m, smax[w]); }
    double* p = c.part + 4 * (size_t)blockIdx.x;
    p[0] = a; p[1] = b; p[2] = d; p[3] = (double)m;
  }
  if (HIST) {
    uint32_t* h = c.hist + (size_t)B.seg * GK_HIST_BINS;
    for (int i = tid; i < GK_HIST_BINS; i += T) {
      const uint32_t v = shist[i];
      if (v) atomicAdd(h + i, v);
    }
  }
  if (last_block_done(&c.ctrl->done[0])) {
    k1_epilogue(c);
    if (HIST) {
      __syncthreads();
      for (int sgi = warp; sgi < c.nseg; sgi += NWARP) radix_pick_warp(c, sgi, 20, GK_HIST_BINS, lane);
    }
  }
}

// ------------------------------------------------------------- K4 radix-select digit passes
// PASS 2: digit = key bits [19:9] of elements whose bits [30:20] equal the prefix.
// PASS 3: digit = key bits [8:0]  of elements whose bits [30:9]  equal the prefix.
template <int PASS, bool MINK>
__global__ void __launch_bounds__(GK_THREADS) k_radix_pass(GkChain c) {
  constexpr int NB = (PASS == 2) ? 2048 : 512;
  constexpr int SHIFT_MATCH = (PASS == 2) ? 20 : 9;
  constexpr int SHIFT_DIGIT = (PASS == 2) ? 9 : 0;
  __shared__ uint32_t shist[NB];
  __shared__ int s_any;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const GkBlk B = c.blk[blockIdx.x];
  for (int i = tid; i < NB; i += T) shist[i] = 0;
  if (tid == 0) s_any = 0;
  __syncthreads();
  const uint32_t prefix = c.stat[B.seg].prefix;
  const float* __restrict__ a = c.resid + B.start;
  const int n4 = B.len >> 2;
  bool any = false;
#pragma unroll 4
  for (int i = tid; i < n4; i += T) {
    const float4 x = ld4(a + 4 * i);
    const float e[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t key = sel_key<MINK>(e[j]);
      if ((key >> SHIFT_MATCH) == prefix) { atomicAdd(&shist[(key >> SHIFT_DIGIT) & (NB - 1)], 1u); any = true; }
    }
  }
  for (int i = 4 * n4 + tid; i < B.len; i += T) {
    const uint32_t key = sel_key<MINK>(a[i]);
    if ((key >> SHIFT_MATCH) == prefix) { atomicAdd(&shist[(key >> SHIFT_DIGIT) & (NB - 1)], 1u); any = true; }
  }
  if (any) s_any = 1;
  __syncthreads();
  if (s_any) {
    uint32_t* h = c.hist + (size_t)B.seg * GK_HIST_BINS;
    for (int i = tid; i < NB; i += T) {
      const uint32_t v = shist[i];
      if (v) atomicAdd(h + i, v);
    }
  }
  if (last_block_done(&c.ctrl->done[1])) {
    for (int s = warp; s < c.nseg; s += NWARP) {
      radix_pick_warp(c, s, SHIFT_DIGIT, NB, lane);
      if (PASS == 3) {
        __syncwarp();
        if (lane == 0) {
          GkSegStat* st = c.stat + s;
          st->tbits = st->prefix;  // full 31-bit key of the k-th element
          st->ties = st->krem;     // how many keys == T are still needed
          st->chosen = 0;
        }
      }
    }
  }
}

// ----------------------------------------------------------------- K2 counting + decisions
// Epilogue shared by the threshold-count pass and the (gt,eq) pass: per-segment totals, the
// reference's refinement loop replayed on the device, then the prefix sums K3 needs.
__device__ void k2_epilogue(const GkChain& c, bool topk) {
  __shared__ uint32_t sscan[2][T];
  __shared__ uint32_t stot[NWARP][GK_NC];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // phase 1: warp per segment
  for (int s = warp; s < c.nseg; s += NWARP) {
    const GkSeg sg = c.seg[s];
    GkSegStat* st = c.stat + s;
    uint32_t tot[GK_NC];
#pragma unroll
    for (int j = 0; j < GK_NC; ++j) tot[j] = 0;
    for (int b = sg.blk_begin + lane; b < sg.blk_end; b += 32) {
      const uint4* p = reinterpret_cast<const uint4*>(c.bcnt + (size_t)b * GK_NC);
#pragma unroll
      for (int v = 0; v < GK_NC / 4; ++v) {
        const uint4 u = __ldcg(p + v);
        tot[4 * v] += u.x; tot[4 * v + 1] += u.y; tot[4 * v + 2] += u.z; tot[4 * v + 3] += u.w;
      }
    }
#pragma unroll
    for (int j = 0; j < GK_NC; ++j) {
      const uint32_t w = warp_sum_u(tot[j]);
      if (lane == 0) stot[warp][j] = w;
    }
    __syncwarp();
    if (lane == 0) {
      const uint32_t* tot = stot[warp];
      const double k = (double)sg.k;
      int chosen = 0;
      uint32_t count = 0;
      if (topk) {
        const uint32_t ties = st->ties;
        count = tot[0] + (tot[1] < ties ? tot[1] : ties);
      } else {
        if (c.mode == GK_MODE_GAUSSIAN) {
          // compression.py:372-381: at most nloops evaluations; the indexes returned are
          // those of the LAST EVALUATED threshold.
          int a = 0, b = 0;
          for (int loop = 0; loop < c.nloops; ++loop) {
            chosen = tri_idx(a, b);
            const double cnt = (double)tot[chosen];
            if (cnt < 2.0 * k / 3.0) ++a;
            else if (cnt > 4.0 * k / 3.0) ++b;
            else break;
          }
        } else if (c.mode == GK_MODE_REDSYNC) {
          // compression.py:653-672
          double l = 0.0, r = 1.0;
          int node = 0;
          while (r - l > 0.2) {
            const double mid = l + (r - l) / 2;
            chosen = node;
            const double nnz = (double)tot[node];
            if (nnz > k && 2.0 * k > nnz) break;
            else if (nnz < k / 2.0) { r = mid; node = 2 * node + 1; }
            else { l = mid; node = 2 * node + 2; }
            if (node >= 7) break;
          }
        } else if (c.mode == GK_MODE_REDSYNCTRIM) {
          // compression.py:719-732: lower the ratio until at least k survive
          chosen = GK_NC - 1;
          for (int j = 0; j < GK_NC; ++j)
            if ((double)tot[j] >= k) { chosen = j; break; }
        } else {
          chosen = 0;
        }
        count = tot[chosen];
        st->thr = st->cand[chosen];
        st->ties = 0;
      }
      st->chosen = chosen;
      st->count = count;
    }
  }
  __syncthreads();
  __threadfence_block();
  // phase 2: exclusive scan of the per-segment counts -> out_base; total
  {
    const int per = (c.nseg + T - 1) / T;
    const int s0 = tid * per, s1 = min(c.nseg, s0 + per);
    uint32_t loc = 0;
    for (int s = s0; s < s1; ++s) loc += c.stat[s].count;
    sscan[0][tid] = loc;
    __syncthreads();
    if (warp == 0) {
      uint32_t carry = 0;
      for (int base = 0; base < T; base += 32) {
        const uint32_t v = sscan[0][base + lane];
        const uint32_t inc = warp_incl_scan_u(v, lane);
        sscan[0][base + lane] = carry + inc - v;
        carry += __shfl_sync(0xffffffffu, inc, 31);
      }
      if (lane == 0) c.ctrl->total = carry;
    }
    __syncthreads();
    uint32_t run = sscan[0][tid];
    for (int s = s0; s < s1; ++s) {
      const uint32_t cnt = c.stat[s].count;
      c.stat[s].out_base = run;
      run += cnt;
    }
  }
  __syncthreads();
  // phase 3: global exclusive scans over the blocks of (gt, eq); K3 subtracts the value at the
  // first block of its segment to make them segment-relative.
  {
    const int per = (c.nblk + T - 1) / T;
    const int b0 = tid * per, b1 = min(c.nblk, b0 + per);
    uint32_t lg = 0, le = 0;
    for (int b = b0; b < b1; ++b) {
      const int ch = topk ? 0 : c.stat[c.blk[b].seg].chosen;
      lg += __ldcg(c.bcnt + (size_t)b * GK_NC + ch);
      if (topk) le += __ldcg(c.bcnt + (size_t)b * GK_NC + 1);
    }
    sscan[0][tid] = lg; sscan[1][tid] = le;
    __syncthreads();
    if (warp < 2) {
      uint32_t carry = 0;
      for (int base = 0; base < T; base += 32) {
        const uint32_t v = sscan[warp][base + lane];
        const uint32_t inc = warp_incl_scan_u(v, lane);
        sscan[warp][base + lane] = carry + inc - v;
        carry += __shfl_sync(0xffffffffu, inc, 31);
      }
    }
    __syncthreads();
    uint32_t rg = sscan[0][tid], re = sscan[1][tid];
    for (int b = b0; b < b1; ++b) {
      const int ch = topk ? 0 : c.stat[c.blk[b].seg].chosen;
      c.boff[2 * (size_t)b] = rg;
      c.boff[2 * (size_t)b + 1] = re;
      rg += __ldcg(c.bcnt + (size_t)b * GK_NC + ch);
      if (topk) re += __ldcg(c.bcnt + (size_t)b * GK_NC + 1);
    }
  }
}

// Count |acc| > cand[j] for all candidates of the block's segment in one read of acc.
template <int NCAND>
__global__ void __launch_bounds__(GK_THREADS) k_count(GkChain c) {
  __shared__ float scand[GK_NC];
  __shared__ uint32_t swc[NWARP][GK_NC];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const GkBlk B = c.blk[blockIdx.x];
  if (tid < GK_NC) scand[tid] = c.stat[B.seg].cand[tid];
  __syncthreads();
  float cand[NCAND];
  uint32_t cnt[NCAND];
#pragma unroll
  for (int j = 0; j < NCAND; ++j) { cand[j] = scand[j]; cnt[j] = 0; }
  const float* __restrict__ a = c.resid + B.start;
  const int n4 = B.len >> 2;
#pragma unroll 4
  for (int i = tid; i < n4; i += T) {
    const float4 x = ld4(a + 4 * i);
    const float e[4] = {fabsf(x.x), fabsf(x.y), fabsf(x.z), fabsf(x.w)};
#pragma unroll
    for (int j = 0; j < NCAND; ++j)
      cnt[j] += (uint32_t)(e[0] > cand[j]) + (uint32_t)(e[1] > cand[j]) + (uint32_t)(e[2] > cand[j]) +
                (uint32_t)(e[3] > cand[j]);
  }
  for (int i = 4 * n4 + tid; i < B.len; i += T) {
    const float v = fabsf(a[i]);
#pragma unroll
    for (int j = 0; j < NCAND; ++j) cnt[j] += (uint32_t)(v > cand[j]);
  }
#pragma unroll
  for (int j = 0; j < NCAND; ++j) {
    const uint32_t w = warp_sum_u(cnt[j]);
    if (lane == 0) swc[warp][j] = w;
  }
  __syncthreads();
  if (tid < GK_NC) {
    uint32_t tot = 0;
    if (tid < NCAND)
      for (int w = 0; w < NWARP; ++w) tot += swc[w][tid];
    c.bcnt[(size_t)blockIdx.x * GK_NC + tid] = tot;
  }
  if (last_block_done(&c.ctrl->done[2])) k2_epilogue(c, false);
}

// top-k: per-block counts of key > T (column 0) and key == T (column 1)
template <bool MINK>
__global__ void __launch_bounds__(GK_THREADS) k_count_topk(GkChain c) {
  __shared__ uint32_t swc[NWARP][2];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const GkBlk B = c.blk[blockIdx.x];
  const uint32_t tb = c.stat[B.seg].tbits;
  uint32_t gt = 0, eq = 0;
  const float* __restrict__ a = c.resid + B.start;
  const int n4 = B.len >> 2;
#pragma unroll 4
  for (int i = tid; i < n4; i += T) {
    const float4 x = ld4(a + 4 * i);
    const float e[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t key = sel_key<MINK>(e[j]);
      gt += (uint32_t)(key > tb);
      eq += (uint32_t)(key == tb);
    }
  }
  for (int i = 4 * n4 + tid; i < B.len; i += T) {
    const uint32_t key = sel_key<MINK>(a[i]);
    gt += (uint32_t)(key > tb);
    eq += (uint32_t)(key == tb);
  }
  gt = warp_sum_u(gt); eq = warp_sum_u(eq);
  if (lane == 0) { swc[warp][0] = gt; swc[warp][1] = eq; }
  __syncthreads();
  if (tid < GK_NC) {
    uint32_t tot = 0;
    if (tid < 2)
      for (int w = 0; w < NWARP; ++w) tot += swc[w][tid];
    c.bcnt[(size_t)blockIdx.x * GK_NC + tid] = tot;
  }
  if (last_block_done(&c.ctrl->done[2])) k2_epilogue(c, true);
}

// ------------------------------------------------------- K3 select + compact + residual + send
template <bool TOPK, bool MINK>
__global__ void __launch_bounds__(GK_THREADS) k_compact_send(GkChain c, GkDst d) {
  __shared__ uint32_t swarp[NWARP];
  __shared__ uint32_t s_tile_total;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.x;
  const GkBlk B = c.blk[b];
  const GkSegStat* st = c.stat + B.seg;
  const float thr = st->thr;
  const uint32_t tb = st->tbits, ties = st->ties;
  const uint32_t seg_first = (uint32_t)c.seg[B.seg].blk_begin;
  const uint32_t base_gt = st->out_base + (c.boff[2 * (size_t)b] - c.boff[2 * (size_t)seg_first]);
  const uint32_t base_eq = TOPK ? (c.boff[2 * (size_t)b + 1] - c.boff[2 * (size_t)seg_first + 1]) : 0u;
  const uint32_t seqn = c.ctrl->seq + 1u;  // every block reads it before the last block bumps it
  const uint64_t poff = (uint64_t)(seqn & 1u) * d.parity_stride;
  const uint32_t cap = c.cap;
  float* __restrict__ a = c.resid + B.start;
  uint32_t run_gt = 0, run_eq = 0;  // selected before the current tile, inside this block
  bool wrote = false;
  const int ntile = (B.len + 4 * T - 1) / (4 * T);
  for (int t = 0; t < ntile; ++t) {
    const int i0 = t * 4 * T + tid * 4;
    float e[4];
    if (i0 + 3 < B.len) {
      const float4 x = ld4(a + i0);
      e[0] = x.x; e[1] = x.y; e[2] = x.z; e[3] = x.w;
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) e[j] = (i0 + j < B.len) ? a[i0 + j] : 0.f;
    }
    uint32_t fg = 0, fe = 0;  // bit j set: element j is gt / eq
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const bool valid = (i0 + j) < B.len;
      if (TOPK) {
        const uint32_t key = sel_key<MINK>(e[j]);
        if (valid && key > tb) fg |= 1u << j;
        if (valid && key == tb) fe |= 1u << j;
      } else {
        if (valid && fabsf(e[j]) > thr) fg |= 1u << j;
      }
    }
    const uint32_t mine = (uint32_t)__popc(fg) | ((uint32_t)__popc(fe) << 16);
    const uint32_t inc = warp_incl_scan_u(mine, lane);
    if (lane == 31) swarp[warp] = inc;
    __syncthreads();
    uint32_t wbase = 0, total = 0;
#pragma unroll
    for (int w = 0; w < NWARP; ++w) {
      const uint32_t v = swarp[w];
      if (w < warp) wbase += v;
      total += v;
    }
    if (mine) {
      const uint32_t excl = wbase + inc - mine;
      uint32_t gbefore = run_gt + (excl & 0xffffu);
      uint32_t ebefore = base_eq + run_eq + (excl >> 16);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const bool isg = (fg >> j) & 1u, ise = (fe >> j) & 1u;
        if (isg || (ise && ebefore < ties)) {
          const uint32_t pos = base_gt + gbefore + (ebefore < ties ? ebefore : ties);
          if (pos < cap) {
            const uint32_t idx = (uint32_t)(B.start + i0 + j);
            const uint64_t pair = (uint64_t)idx | ((uint64_t)__float_as_uint(e[j]) << 32);
            for (int q = 0; q < d.ndst; ++q) d.slot[q][poff + GK_SLOT_HDR + pos] = pair;
            a[i0 + j] = 0.f;  // sent => leaves the residual
            wrote = true;
          }
        }
        gbefore += isg ? 1u : 0u;
        ebefore += ise ? 1u : 0u;
      }
    }
    run_gt += total & 0xffffu;
    run_eq += total >> 16;
    __syncthreads();
  }
  (void)s_tile_total;
  if (wrote) __threadfence_system();  // pairs visible to the peers before the ticket
  if (last_block_done(&c.ctrl->done[3])) {
    if (tid == 0) {
      const uint32_t total = c.ctrl->total;
      const uint32_t sent = total < cap ? total : cap;
      __threadfence_system();
      const uint64_t hdr = (uint64_t)sent | ((uint64_t)seqn << 32);
      for (int q = 0; q < d.ndst; ++q) st_release_sys_u64(d.slot[q] + poff, hdr);
      c.ctrl->sent = sent;
      c.ctrl->overflow += total - sent;
      c.ctrl->sent_total += sent;
      c.ctrl->ring[seqn & 63u] = total;
      c.ctrl->seq = seqn;
    }
  }
}

// ------------------------------------------------------------------------ K5 receive side
template <bool UNSCATTER>
__global__ void __launch_bounds__(GK_THREADS) k_scatter(GkSrc s, GkCtrl* ctrl, float* __restrict__ grad,
                                                         float scale, int nelem) {
  __shared__ uint32_t scnt[GK_MAXP];
  __shared__ int s_bad;
  const int tid = threadIdx.x;
  const uint32_t seq = ctrl->seq;
  const uint64_t* base = s.arena + (uint64_t)(seq & 1u) * s.parity_stride;
  const uint32_t cap = (uint32_t)(s.slot_stride - GK_SLOT_HDR);
  if (tid == 0) s_bad = 0;
  __syncthreads();
  if (tid < s.nsrc) {
    const uint64_t* hp = base + (uint64_t)tid * s.slot_stride;
    uint64_t v = ld_acquire_sys_u64(hp);
    if (s.wait && !UNSCATTER) {
      const uint64_t t0 = globaltimer_ns();
      while ((uint32_t)(v >> 32) != seq) {
        if (globaltimer_ns() - t0 > s.timeout_ns) { s_bad = 1; break; }
        __nanosleep(100);
        v = ld_acquire_sys_u64(hp);
      }
    }
    const uint32_t n = (uint32_t)v;
    scnt[tid] = n < cap ? n : cap;
  }
  __syncthreads();
  if (s_bad) {
    if (tid == 0) atomicExch(&ctrl->err, 1u);
    return;
  }
  for (int r = 0; r < s.nsrc; ++r) {
    const uint64_t* pairs = base + (uint64_t)r * s.slot_stride + GK_SLOT_HDR;
    const uint32_t n = scnt[r];
    for (uint32_t j = blockIdx.x * T + tid; j < n; j += gridDim.x * T) {
      const uint64_t p = __ldcg(pairs + j);
      const uint32_t idx = (uint32_t)p;
      if (idx < (uint32_t)nelem) {
        if (UNSCATTER) grad[idx] = 0.f;
        else atomicAdd(grad + idx, __uint_as_float((uint32_t)(p >> 32)) * scale);
      }
    }
  }
}

// ------------------------------------------------------------------------- K6 fused SGD
__global__ void __launch_bounds__(GK_THREADS) k_sgd(float* __restrict__ p, const float* __restrict__ g,
                                                     float* __restrict__ m, const GkBlk* __restrict__ blk,
                                                     const GkHyper* __restrict__ hy,
                                                     const float* __restrict__ gscale) {
  const GkBlk B = blk[blockIdx.x];
  const GkHyper h = hy[B.seg];
  const float gs = gscale ? *gscale : 1.f;
  const bool has_m = (h.momentum != 0.f) && (m != nullptr);
  const float nlr = -h.lr, omd = 1.f - h.dampening;
  auto upd = [&](float pv, float gv, float& mv) -> float {
    if (gs != 1.f) gv = __fmul_rn(gv, gs);
    if (h.weight_decay != 0.f) gv = __fmaf_rn(h.weight_decay, pv, gv);
    float u = gv;
    if (has_m) {
      mv = h.first_step ? gv : __fmaf_rn(omd, gv, __fmul_rn(h.momentum, mv));
      u = h.nesterov ? __fmaf_rn(h.momentum, mv, gv) : mv;
    }
    return __fmaf_rn(nlr, u, pv);
  };
  float* pp = p + B.start; const float* gg = g + B.start; float* mm = has_m ? m + B.start : nullptr;
  const int n4 = B.len >> 2;
#pragma unroll 2
  for (int i = threadIdx.x; i < n4; i += T) {
    float4 pv = ld4(pp + 4 * i);
    const float4 gv = ld4(gg + 4 * i);
    float4 mv = has_m ? ld4(mm + 4 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
    pv.x = upd(pv.x, gv.x, mv.x); pv.y = upd(pv.y, gv.y, mv.y);
    pv.z = upd(pv.z, gv.z, mv.z); pv.w = upd(pv.w, gv.w, mv.w);
    st4(pp + 4 * i, pv);
    if (has_m) st4(mm + 4 * i, mv);
  }
  for (int i = 4 * n4 + threadIdx.x; i < B.len; i += T) {
    float mv = has_m ? mm[i] : 0.f;
    pp[i] = upd(pp[i], gg[i], mv);
    if (has_m) mm[i] = mv;
  }
}

// ------------------------------------------------------------------------------ K7 LARS
__global__ void __launch_bounds__(GK_THREADS) k_lars_norms(const float* __restrict__ p, const float* __restrict__ g,
                                                            const GkBlk* __restrict__ blk, const GkSeg* __restrict__ tens,
                                                            int ntens, const GkHyper* __restrict__ hy, double* part,
                                                            float* ratio, GkCtrl* ctrl) {
  __shared__ double sred[2][NWARP];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const GkBlk B = blk[blockIdx.x];
  double w2 = 0, g2 = 0;
  for (int i = tid; i < B.len; i += T) {
    const double a = (double)p[B.start + i], b = (double)g[B.start + i];
    w2 = fma(a, a, w2); g2 = fma(b, b, g2);
  }
  w2 = warp_sum(w2); g2 = warp_sum(g2);
  if (lane == 0) { sred[0][warp] = w2; sred[1][warp] = g2; }
  __syncthreads();
  if (tid == 0) {
    double a = 0, b = 0;
    for (int w = 0; w < NWARP; ++w) { a += sred[0][w]; b += sred[1][w]; }
    part[2 * (size_t)blockIdx.x] = a; part[2 * (size_t)blockIdx.x + 1] = b;
  }
  if (last_block_done(&ctrl->done[4])) {
    for (int t = warp; t < ntens; t += NWARP) {
      const GkSeg tg = tens[t];
      double a = 0, b = 0;
      for (int bb = tg.blk_begin + lane; bb < tg.blk_end; bb += 32) {
        a += __ldcg(part + 2 * (size_t)bb); b += __ldcg(part + 2 * (size_t)bb + 1);
      }
      a = warp_sum(a); b = warp_sum(b);
      if (lane == 0) {
        const GkHyper h = hy[tg.aux];
        const float wn = (float)sqrt(a), gn = (float)sqrt(b);
        float tr = 1.f;  // lars.py:92-101
        if (wn > 0.f && gn > 0.f) tr = h.lars_eeta * wn / (gn + h.weight_decay * wn + h.lars_eps);
        tr = fminf(fmaxf(tr, 0.f), 50.f);
        ratio[t] = h.lr * tr;  // scaled_lr
      }
    }
  }
}

__global__ void __launch_bounds__(GK_THREADS) k_lars_update(float* __restrict__ p, const float* __restrict__ g,
                                                             float* __restrict__ acc, const GkBlk* __restrict__ blk,
                                                             const GkSeg* __restrict__ tens,
                                                             const GkHyper* __restrict__ hy, const float* __restrict__ ratio) {
  const GkBlk B = blk[blockIdx.x];
  const GkHyper h = hy[tens[B.seg].aux];
  const float slr = ratio[B.seg];
  for (int i = threadIdx.x; i < B.len; i += T) {
    const int j = B.start + i;
    float dg = __fmaf_rn(h.weight_decay, p[j], g[j]);  // lars.py:112-113
    dg = fminf(fmaxf(dg, -10.f), 10.f);
    const float a = __fmaf_rn(slr, dg, __fmul_rn(h.momentum, acc[j]));  // lars.py:124-125
    acc[j] = a;
    p[j] = p[j] - a;
  }
}

// --------------------------------------------------------------------- K8 grad norm / scale
__global__ void __launch_bounds__(GK_THREADS) k_grad_norm(const float* __restrict__ g, const GkBlk* __restrict__ blk,
                                                           int nblk, double* part, float max_norm, float* out,
                                                           GkCtrl* ctrl) {
  __shared__ double sred[NWARP];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const GkBlk B = blk[blockIdx.x];
  double s = 0;
  const int n4 = B.len >> 2;
  for (int i = tid; i < n4; i += T) {
    const float4 x = ld4(g + B.start + 4 * i);
    s = fma((double)x.x, (double)x.x, s); s = fma((double)x.y, (double)x.y, s);
    s = fma((double)x.z, (double)x.z, s); s = fma((double)x.w, (double)x.w, s);
  }
  for (int i = 4 * n4 + tid; i < B.len; i += T) { const double x = g[B.start + i]; s = fma(x, x, s); }
  s = warp_sum(s);
  if (lane == 0) sred[warp] = s;
  __syncthreads();
  if (tid == 0) {
    double a = 0;
    for (int w = 0; w < NWARP; ++w) a += sred[w];
    part[blockIdx.x] = a;
  }
  if (last_block_done(&ctrl->done[5])) {
    double a = 0;
    for (int b = tid; b < nblk; b += T) a += __ldcg(part + b);
    a = warp_sum(a);
    __syncthreads();
    if (lane == 0) sred[warp] = a;
    __syncthreads();
    if (tid == 0) {
      double t = 0;
      for (int w = 0; w < NWARP; ++w) t += sred[w];
      const float norm = (float)sqrt(t);
      const float coef = max_norm / (norm + 1e-6f);  // torch.nn.utils.clip_grad_norm_
      out[0] = norm;
      out[1] = (max_norm > 0.f && coef < 1.f) ? coef : 1.f;
    }
  }
}

__global__ void __launch_bounds__(GK_THREADS) k_scale(float* __restrict__ g, const GkBlk* __restrict__ blk,
                                                       const float* __restrict__ scale) {
  const float s = *scale;
  if (s == 1.f) return;
  const GkBlk B = blk[blockIdx.x];
  for (int i = threadIdx.x; i < B.len; i += T) g[B.start + i] *= s;
}

// -------------------------------------------------------------------------- K9 random-k
// Keyed bijection of [0, n): 4-round Feistel network on the next even-bit-width power of two
// with cycle walking.  Distinct inputs give distinct outputs, so j = 0..k-1 yields k distinct
// uniformly scattered indices without touching the n-element tensor.
__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}
__device__ __forceinline__ uint32_t feistel_perm(uint32_t j, uint32_t n, uint64_t seed) {
  int bits = 2;
  while ((1ull << bits) < (uint64_t)n) bits += 2;
  const int half = bits / 2;
  const uint32_t mask = (1u << half) - 1u;
  uint32_t x = j;
  do {
    uint32_t l = x >> half, r = x & mask;
#pragma unroll
    for (int round = 0; round < 4; ++round) {
      const uint32_t f = mix32(r ^ (uint32_t)(seed >> (round * 8)) ^ (0x9e3779b9u * (round + 1))) & mask;
      const uint32_t nl = r;
      r = l ^ f;
      l = nl;
    }
    x = (l << half) | r;
  } while (x >= n);
  return x;
}

__global__ void __launch_bounds__(GK_THREADS) k_randomk(GkChain c, GkDst d, uint64_t seed) {
  // grid-stride over all (segment, j) outputs; GkSeg.aux holds the prefix sum of k (out base)
  const uint32_t seqn = c.ctrl->seq + 1u;
  const uint64_t poff = (uint64_t)(seqn & 1u) * d.parity_stride;
  bool wrote = false;
  for (int s = 0; s < c.nseg; ++s) {
    const GkSeg sg = c.seg[s];
    for (uint32_t j = blockIdx.x * T + threadIdx.x; j < (uint32_t)sg.k; j += gridDim.x * T) {
      const uint32_t li = feistel_perm(j, (uint32_t)sg.n, seed + 0x632be59bd9b4e019ull * (uint64_t)(s + 1) +
                                                              0x9e3779b97f4a7c15ull * (uint64_t)seqn);
      // segment-local index -> arena index: last block whose GkBlk.soff <= li
      int lo = sg.blk_begin, hi = sg.blk_end - 1;
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if ((uint32_t)c.blk[mid].soff <= li) lo = mid; else hi = mid - 1;
      }
      const GkBlk B = c.blk[lo];
      const uint32_t idx = (uint32_t)B.start + (li - (uint32_t)B.soff);
      const uint32_t pos = (uint32_t)sg.aux + j;
      if (pos < c.cap) {
        const float v = c.resid[idx];
        const uint64_t pair = (uint64_t)idx | ((uint64_t)__float_as_uint(v) << 32);
        for (int q = 0; q < d.ndst; ++q) d.slot[q][poff + GK_SLOT_HDR + pos] = pair;
        c.resid[idx] = 0.f;
        wrote = true;
      }
    }
  }
  if (wrote) __threadfence_system();
  if (last_block_done(&c.ctrl->done[3])) {
    if (threadIdx.x == 0) {
      uint32_t total = 0;
      for (int s = 0; s < c.nseg; ++s) total += (uint32_t)c.seg[s].k;
      const uint32_t sent = total < c.cap ? total : c.cap;
      __threadfence_system();
      const uint64_t hdr = (uint64_t)sent | ((uint64_t)seqn << 32);
      for (int q = 0; q < d.ndst; ++q) st_release_sys_u64(d.slot[q] + poff, hdr);
      c.ctrl->total = total; c.ctrl->sent = sent;
      c.ctrl->overflow += total - sent; c.ctrl->sent_total += sent;
      c.ctrl->ring[seqn & 63u] = total;
      c.ctrl->seq = seqn;
    }
  }
}

// ------------------------------------------------------------------- K10 bucket compressor
__global__ void __launch_bounds__(GK_THREADS) k_bucket_means(GkChain c, float* means) {
  __shared__ double sred[4][NWARP];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const GkBlk B = c.blk[blockIdx.x];
  double sp = 0, sn = 0, cp = 0, cn = 0;
  for (int i = tid; i < B.len; i += T) {
    const float x = c.grad[B.start + i];
    if (x >= 0.f) { sp += (double)x; cp += 1.0; } else { sn += (double)x; cn += 1.0; }
  }
  sp = warp_sum(sp); sn = warp_sum(sn); cp = warp_sum(cp); cn = warp_sum(cn);
  if (lane == 0) { sred[0][warp] = sp; sred[1][warp] = sn; sred[2][warp] = cp; sred[3][warp] = cn; }
  __syncthreads();
  if (tid == 0) {
    double a = 0, b = 0, e = 0, f = 0;
    for (int w = 0; w < NWARP; ++w) { a += sred[0][w]; b += sred[1][w]; e += sred[2][w]; f += sred[3][w]; }
    double* p = c.part + 4 * (size_t)blockIdx.x;
    p[0] = a; p[1] = b; p[2] = e; p[3] = f;
  }
  if (last_block_done(&c.ctrl->done[6])) {
    for (int s = warp; s < c.nseg; s += NWARP) {
      const GkSeg sg = c.seg[s];
      double a = 0, b = 0, e = 0, f = 0;
      for (int bb = sg.blk_begin + lane; bb < sg.blk_end; bb += 32) {
        const double* p = c.part + 4 * (size_t)bb;
        a += __ldcg(p); b += __ldcg(p + 1); e += __ldcg(p + 2); f += __ldcg(p + 3);
      }
      a = warp_sum(a); b = warp_sum(b); e = warp_sum(e); f = warp_sum(f);
      if (lane == 0) {
        means[2 * s] = e > 0 ? (float)(a / e) : 0.f;
        means[2 * s + 1] = f > 0 ? (float)(b / f) : 0.f;
      }
    }
  }
}

__global__ void __launch_bounds__(GK_THREADS) k_bucket_reconstruct(GkChain c, const float* __restrict__ lm,
                                                                    const float* __restrict__ gm) {
  const GkBlk B = c.blk[blockIdx.x];
  const float l0 = lm[2 * B.seg], l1 = lm[2 * B.seg + 1], g0 = gm[2 * B.seg], g1 = gm[2 * B.seg + 1];
  for (int i = threadIdx.x; i < B.len; i += T) {
    const float x = c.grad[B.start + i];
    c.grad[B.start + i] = (x >= 0.f) ? __fadd_rn(__fsub_rn(x, l0), g0) : __fadd_rn(__fsub_rn(x, l1), g1);
  }
}

// ----------------------------------------------------------------------- K13 / K14 helpers
__global__ void __launch_bounds__(GK_THREADS) k_copy(float* __restrict__ dst, const float* __restrict__ src, int64_t n) {
  const int64_t n4 = n >> 2;
  const int64_t stride = (int64_t)gridDim.x * T;
  if ((((uintptr_t)dst | (uintptr_t)src) & 15) == 0) {
    for (int64_t i = (int64_t)blockIdx.x * T + threadIdx.x; i < n4; i += stride) st4(dst + 4 * i, ld4(src + 4 * i));
    for (int64_t i = 4 * n4 + (int64_t)blockIdx.x * T + threadIdx.x; i < n; i += stride) dst[i] = src[i];
  } else {
    for (int64_t i = (int64_t)blockIdx.x * T + threadIdx.x; i < n; i += stride) dst[i] = src[i];
  }
}

__global__ void __launch_bounds__(GK_THREADS) k_histogram(const float* __restrict__ x, int64_t n, float lo, float inv_w,
                                                           int nbins, uint32_t* out) {
  extern __shared__ uint32_t sh[];
  for (int i = threadIdx.x; i < nbins; i += T) sh[i] = 0;
  __syncthreads();
  for (int64_t i = (int64_t)blockIdx.x * T + threadIdx.x; i < n; i += (int64_t)gridDim.x * T) {
    int bin = (int)floorf((x[i] - lo) * inv_w);
    bin = bin < 0 ? 0 : (bin >= nbins ? nbins - 1 : bin);
    atomicAdd(&sh[bin], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nbins; i += T)
    if (sh[i]) atomicAdd(out + i, sh[i]);
}

// =========================================================================== host entry points
static inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }
#define GK_RET() return (int)cudaGetLastError()

extern "C" {

const char* gk_version(void) { return "gk_b200 cuda sm_100a 0.1"; }

int gk_accumulate(const GkChain* c, void* stream) {
  if (c->nblk <= 0) return 0;
  if (c->mode == GK_MODE_TOPK) k_accumulate<true, false><<<c->nblk, T, 0, S(stream)>>>(*c);
  else if (c->mode == GK_MODE_MINK) k_accumulate<true, true><<<c->nblk, T, 0, S(stream)>>>(*c);
  else k_accumulate<false, false><<<c->nblk, T, 0, S(stream)>>>(*c);
  GK_RET();
}

int gk_count_pick(const GkChain* c, void* stream) {
  if (c->nblk <= 0) return 0;
  int nc = 1;
  if (c->mode == GK_MODE_GAUSSIAN) nc = c->nloops * (c->nloops + 1) / 2;
  else if (c->mode == GK_MODE_REDSYNC) nc = 7;
  else if (c->mode == GK_MODE_REDSYNCTRIM) nc = GK_NC;
  if (nc <= 1) k_count<1><<<c->nblk, T, 0, S(stream)>>>(*c);
  else if (nc <= 6) k_count<6><<<c->nblk, T, 0, S(stream)>>>(*c);
  else if (nc <= 7) k_count<7><<<c->nblk, T, 0, S(stream)>>>(*c);
  else k_count<GK_NC><<<c->nblk, T, 0, S(stream)>>>(*c);
  GK_RET();
}

int gk_topk_select(const GkChain* c, void* stream) {
  if (c->nblk <= 0) return 0;
  if (c->mode == GK_MODE_MINK) {
    k_radix_pass<2, true><<<c->nblk, T, 0, S(stream)>>>(*c);
    k_radix_pass<3, true><<<c->nblk, T, 0, S(stream)>>>(*c);
    k_count_topk<true><<<c->nblk, T, 0, S(stream)>>>(*c);
  } else {
    k_radix_pass<2, false><<<c->nblk, T, 0, S(stream)>>>(*c);
    k_radix_pass<3, false><<<c->nblk, T, 0, S(stream)>>>(*c);
    k_count_topk<false><<<c->nblk, T, 0, S(stream)>>>(*c);
  }
  GK_RET();
}

int gk_compact_send(const GkChain* c, const GkDst* dst, void* stream) {
  if (c->nblk <= 0) return 0;
  if (dst->ndst < 1 || dst->ndst > GK_MAXP) return (int)cudaErrorInvalidValue;
  if (c->mode == GK_MODE_TOPK) k_compact_send<true, false><<<c->nblk, T, 0, S(stream)>>>(*c, *dst);
  else if (c->mode == GK_MODE_MINK) k_compact_send<true, true><<<c->nblk, T, 0, S(stream)>>>(*c, *dst);
  else k_compact_send<false, false><<<c->nblk, T, 0, S(stream)>>>(*c, *dst);
  GK_RET();
}

int gk_compress(const GkChain* c, const GkDst* dst, void* stream) {
  int e = gk_accumulate(c, stream);
  if (e) return e;
  if (c->mode == GK_MODE_TOPK || c->mode == GK_MODE_MINK) e = gk_topk_select(c, stream);
  else e = gk_count_pick(c, stream);
  if (e) return e;
  return gk_compact_send(c, dst, stream);
}

int gk_scatter(const GkSrc* src, GkCtrl* ctrl, float* grad, float scale, int32_t nelem, void* stream) {
  if (src->nsrc < 1 || src->nsrc > GK_MAXP) return (int)cudaErrorInvalidValue;
  const uint64_t cap = src->slot_stride - GK_SLOT_HDR;
  int grid = (int)((cap + T - 1) / T);
  if (grid > 296) grid = 296;
  if (grid < 1) grid = 1;
  k_scatter<false><<<grid, T, 0, S(stream)>>>(*src, ctrl, grad, scale, nelem);
  GK_RET();
}

int gk_unscatter(const GkSrc* src, GkCtrl* ctrl, float* grad, int32_t nelem, void* stream) {
  if (src->nsrc < 1 || src->nsrc > GK_MAXP) return (int)cudaErrorInvalidValue;
  const uint64_t cap = src->slot_stride - GK_SLOT_HDR;
  int grid = (int)((cap + T - 1) / T);
  if (grid > 296) grid = 296;
  if (grid < 1) grid = 1;
  k_scatter<true><<<grid, T, 0, S(stream)>>>(*src, ctrl, grad, 0.f, nelem);
  GK_RET();
}

int gk_sgd(float* param, const float* grad, float* mom, const GkBlk* blk, int32_t nblk, const GkHyper* hyper,
           const float* grad_scale, void* stream) {
  if (nblk <= 0) return 0;
  k_sgd<<<nblk, T, 0, S(stream)>>>(param, grad, mom, blk, hyper, grad_scale);
  GK_RET();
}

int gk_lars(float* param, const float* grad, float* mom, const GkBlk* blk, int32_t nblk, const GkSeg* tens,
            int32_t ntens, const GkHyper* hyper, double* part, float* ratio, GkCtrl* ctrl, void* stream) {
  if (nblk <= 0) return 0;
  k_lars_norms<<<nblk, T, 0, S(stream)>>>(param, grad, blk, tens, ntens, hyper, part, ratio, ctrl);
  k_lars_update<<<nblk, T, 0, S(stream)>>>(param, grad, mom, blk, tens, hyper, ratio);
  GK_RET();
}

int gk_grad_norm(const float* grad, const GkBlk* blk, int32_t nblk, double* part, float max_norm, float* out,
                 GkCtrl* ctrl, void* stream) {
  if (nblk <= 0) return 0;
  k_grad_norm<<<nblk, T, 0, S(stream)>>>(grad, blk, nblk, part, max_norm, out, ctrl);
  GK_RET();
}

int gk_scale(float* grad, const GkBlk* blk, int32_t nblk, const float* scale, void* stream) {
  if (nblk <= 0) return 0;
  k_scale<<<nblk, T, 0, S(stream)>>>(grad, blk, scale);
  GK_RET();
}

int gk_randomk_send(const GkChain* c, const GkDst* dst, uint64_t seed, void* stream) {
  if (dst->ndst < 1 || dst->ndst > GK_MAXP) return (int)cudaErrorInvalidValue;
  k_randomk<<<148, T, 0, S(stream)>>>(*c, *dst, seed);
  GK_RET();
}

int gk_bucket_means(const GkChain* c, float* means, void* stream) {
  if (c->nblk <= 0) return 0;
  k_bucket_means<<<c->nblk, T, 0, S(stream)>>>(*c, means);
  GK_RET();
}

int gk_bucket_reconstruct(const GkChain* c, const float* lm, const float* gm, void* stream) {
  if (c->nblk <= 0) return 0;
  k_bucket_reconstruct<<<c->nblk, T, 0, S(stream)>>>(*c, lm, gm);
  GK_RET();
}

int gk_copy(float* dst, const float* src, int64_t n, void* stream) {
  if (n <= 0) return 0;
  int64_t blocks = (n / 4 + T - 1) / T;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  k_copy<<<(int)blocks, T, 0, S(stream)>>>(dst, src, n);
  GK_RET();
}

int gk_histogram(const float* x, int64_t n, float lo, float hi, int32_t nbins, uint32_t* out, void* stream) {
  if (n <= 0 || nbins <= 0 || nbins > 8192 || !(hi > lo)) return (int)cudaErrorInvalidValue;
  int64_t blocks = (n + T - 1) / T;
  if (blocks > 148 * 4) blocks = 148 * 4;
  k_histogram<<<(int)blocks, T, nbins * sizeof(uint32_t), S(stream)>>>(x, n, lo, (float)nbins / (hi - lo), nbins, out);
  GK_RET();
}

// ---- peer-memory plumbing for the P2P transport (set-up time only; these are the only
// entry points of the library that allocate).  The arena they hand out is written by peers'
// k_compact_send kernels over NVLink and read by the local k_scatter kernel.
int gk_ipc_alloc(uint64_t bytes, void** ptr, unsigned char* handle64) {
  cudaError_t e = cudaMalloc(ptr, bytes);
  if (e != cudaSuccess) return (int)e;
  e = cudaMemset(*ptr, 0, bytes);
  if (e != cudaSuccess) return (int)e;
  cudaIpcMemHandle_t h;
  e = cudaIpcGetMemHandle(&h, *ptr);
  if (e != cudaSuccess) return (int)e;
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64, &h, 64);
  return (int)cudaDeviceSynchronize();
}

int gk_ipc_open(const unsigned char* handle64, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  return (int)cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess);
}

int gk_ipc_close(void* ptr) { return (int)cudaIpcCloseMemHandle(ptr); }
int gk_ipc_free(void* ptr) { return (int)cudaFree(ptr); }

}  // extern "C"
