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
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "../include/gk_b200.h"

#define T GK_THREADS
#define NWARP (GK_THREADS / 32)
#define GK_SUB (NWARP * 1024)  /* elements per sub-tile: every warp owns a contiguous 1024-element slice */

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
__device__ __forceinline__ uint32_t warp_sum_u(uint32_t v) { return __reduce_add_sync(0xffffffffu, v); }
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
__device__ __forceinline__ bool last_block_done(uint32_t* ticket, bool system_scope = false) {
  __shared__ int s_last;
  // bar.sync makes the block's writes visible to thread 0; its (cumulative) fence then orders them
  // before the ticket for every observer — one fence per block instead of one per thread.
  __syncthreads();
  if (threadIdx.x == 0) {
    if (system_scope) __threadfence_system(); else __threadfence();
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

// 0xffffffff if a > b else 0 (one FSET; masks are then subtracted / and-ed, which is cheaper than the
// compare + select + add sequence the compiler emits for `cnt += (a > b)`)
__device__ __forceinline__ uint32_t gt_mask(float a, float b) {
  uint32_t m;
  asm("set.gt.u32.f32 %0, %1, %2;" : "=r"(m) : "f"(a), "f"(b));
  return m;
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
// Leader work of the K1 epilogue for one segment: moments -> stat, candidate thresholds.
__device__ void k1_finish_segment(const GkChain& c, int s, const GkSeg& sg, double sum, double sq, double sa, float mx) {
  GkSegStat* st = c.stat + s;
  for (int j = 0; j < GK_NC; ++j) st->tot[j] = 0;
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
  if (c.mode == GK_MODE_TOPK || c.mode == GK_MODE_MINK) {
    c.stat[s].prefix = 0; c.stat[s].krem = (uint32_t)sg.k;
  } else {
    // Sorted candidate order for the counting pass.  Candidates far in the tail ("high") are only
    // evaluated in rows where the row maximum exceeds the smallest of them; the split is a speed
    // heuristic (about 1 % expected exceedance), the counts are exact either way.
    const int nc = (c.mode == GK_MODE_GAUSSIAN) ? c.nloops * (c.nloops + 1) / 2
                 : (c.mode == GK_MODE_REDSYNC) ? 7 : (c.mode == GK_MODE_REDSYNCTRIM) ? GK_NC : 1;
    uint8_t ord[GK_NC];
    for (int j = 0; j < GK_NC; ++j) ord[j] = (uint8_t)j;
    for (int j = 1; j < nc; ++j) {          // insertion sort, <= 16 items
      const uint8_t key = ord[j];
      const float kv = st->cand[key];
      int i = j - 1;
      while (i >= 0 && st->cand[ord[i]] > kv) { ord[i + 1] = ord[i]; --i; }
      ord[i + 1] = key;
    }
    const float gate = (c.mode == GK_MODE_GAUSSIAN) ? fabsf(meanf) + 2.5f * stdf : 3.2f * meanabs;
    int nlow = 0;
    while (nlow < nc && !(st->cand[ord[nlow]] > gate)) ++nlow;
    if (c.mode == GK_MODE_THRESHOLD) nlow = 1;
    for (int j = 0; j < GK_NC; ++j) st->order[j] = ord[j];
    st->nlow = nlow;
  }
}

// Segments with at least this many table entries are reduced by the whole block, smaller ones by one warp.
#define GK_BIG_SEG 256

__device__ void k1_epilogue(const GkChain& c) {
  __shared__ double s_red[3][NWARP];
  __shared__ float s_mx[NWARP];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int s = 0; s < c.nseg; ++s) {
    const GkSeg sg = c.seg[s];
    const bool big = (sg.blk_end - sg.blk_begin) >= GK_BIG_SEG;   // uniform over the block
    if (!big && (s % NWARP) != warp) continue;
    const int first = big ? tid : lane, step = big ? T : 32;
    double sum = 0, sq = 0, sa = 0;
    float mx = 0.f;
#pragma unroll 4
    for (int b = sg.blk_begin + first; b < sg.blk_end; b += step) {
      const double2 p0 = __ldcg(reinterpret_cast<const double2*>(c.part + 4 * (size_t)b));
      const double2 p1 = __ldcg(reinterpret_cast<const double2*>(c.part + 4 * (size_t)b + 2));
      sum += p0.x; sq += p0.y; sa += p1.x; mx = fmaxf(mx, (float)p1.y);
    }
    sum = warp_sum(sum); sq = warp_sum(sq); sa = warp_sum(sa); mx = warp_max(mx);
    if (big) {
      if (lane == 0) { s_red[0][warp] = sum; s_red[1][warp] = sq; s_red[2][warp] = sa; s_mx[warp] = mx; }
      __syncthreads();
      if (tid == 0) {
        sum = sq = sa = 0; mx = 0.f;
        for (int w = 0; w < NWARP; ++w) { sum += s_red[0][w]; sq += s_red[1][w]; sa += s_red[2][w]; mx = fmaxf(mx, s_mx[w]); }
        k1_finish_segment(c, s, sg, sum, sq, sa, mx);
      }
      __syncthreads();
    } else if (lane == 0) {
      k1_finish_segment(c, s, sg, sum, sq, sa, mx);
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

// Persistent: a CUDA block walks the table entries blockIdx.x, blockIdx.x + gridDim.x, ... and keeps
// its partial moments in registers for as long as consecutive entries belong to the same segment (a
// "run"); only the last entry of a run pays for the block reduction, the others record zeros.  In the
// merged-buffer mode (one segment) that is one reduction and one ticket per block.
template <bool HIST, bool MINK>
__global__ void __launch_bounds__(GK_THREADS) k_accumulate(GkChain c) {
  __shared__ uint32_t shist[HIST ? GK_HIST_BINS : 1];
  __shared__ double sred[3][NWARP];
  __shared__ float smax[NWARP];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (HIST) {
    for (int i = tid; i < GK_HIST_BINS; i += T) shist[i] = 0;
    __syncthreads();
  }
  const bool user = c.use_residual != 0, zg = c.zero_grad != 0 && c.gsrc == nullptr;
  const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
  double ds = 0, dq = 0, da = 0;   // this thread's share of the current run
  float mx = 0.f;
  for (int eidx = blockIdx.x; eidx < c.nblk; eidx += gridDim.x) {
    const GkBlk B = c.blk[eidx];
    const int enext = eidx + gridDim.x;
    const bool last_of_run = (enext >= c.nblk) || (c.blk[enext].seg != B.seg);   // block-uniform
    const float* __restrict__ g = c.gsrc ? c.gsrc[eidx] : c.grad + B.start;
    float* __restrict__ r = c.resid + B.start;
    // fp32 partials over one entry (<= chunk/256 elements per thread), fp64 across entries
    float s = 0.f, q = 0.f, sa = 0.f;
    const int n4 = B.len >> 2;
    if (tid < GK_NC) c.bcnt[(size_t)eidx * GK_NC + tid] = 0;   // the counting pass accumulates with atomics
#pragma unroll 2
    for (int i = tid; i < n4; i += T) {
      float4 x = ld4_stream(g + 4 * i);
      if (user) {
        const float4 rv = ld4_stream(r + 4 * i);
        x.x += rv.x; x.y += rv.y; x.z += rv.z; x.w += rv.w;
      }
      st4(r + 4 * i, x);  // acc: default policy so that it can stay in L2 for the following passes
      if (zg) st4_stream(const_cast<float*>(g) + 4 * i, z4);
      s += (x.x + x.y) + (x.z + x.w);
      q = fmaf(x.x, x.x, q); q = fmaf(x.y, x.y, q); q = fmaf(x.z, x.z, q); q = fmaf(x.w, x.w, q);
      const float ax = fabsf(x.x), ay = fabsf(x.y), az = fabsf(x.z), aw = fabsf(x.w);
      sa += (ax + ay) + (az + aw);
      mx = fmaxf(fmaxf(mx, fmaxf(ax, ay)), fmaxf(az, aw));
      if (HIST) {
        atomicAdd(&shist[sel_key<MINK>(x.x) >> 20], 1u); atomicAdd(&shist[sel_key<MINK>(x.y) >> 20], 1u);
        atomicAdd(&shist[sel_key<MINK>(x.z) >> 20], 1u); atomicAdd(&shist[sel_key<MINK>(x.w) >> 20], 1u);
      }
    }
    for (int i = 4 * n4 + tid; i < B.len; i += T) {  // tail of a tensor whose size % 4 != 0
      float x = g[i];
      if (user) x += r[i];
      r[i] = x;
      if (zg) const_cast<float*>(g)[i] = 0.f;
      s += x; q = fmaf(x, x, q); sa += fabsf(x);
      mx = fmaxf(mx, fabsf(x));
      if (HIST) atomicAdd(&shist[sel_key<MINK>(x) >> 20], 1u);
    }
    ds += (double)s; dq += (double)q; da += (double)sa;
    double* p = c.part + 4 * (size_t)eidx;
    if (!last_of_run) {
      if (tid == 0) { p[0] = 0; p[1] = 0; p[2] = 0; p[3] = 0; }
      continue;
    }
    ds = warp_sum(ds); dq = warp_sum(dq); da = warp_sum(da); mx = warp_max(mx);
    if (lane == 0) { sred[0][warp] = ds; sred[1][warp] = dq; sred[2][warp] = da; smax[warp] = mx; }
    __syncthreads();
    if (tid == 0) {
      double a = 0, b = 0, d = 0; float m = 0.f;
      for (int w = 0; w < NWARP; ++w) { a += sred[0][w]; b += sred[1][w]; d += sred[2][w]; m = fmaxf(m, smax[w]); }
      p[0] = a; p[1] = b; p[2] = d; p[3] = (double)m;
    }
    if (HIST) {
      uint32_t* h = c.hist + (size_t)B.seg * GK_HIST_BINS;
      for (int i = tid; i < GK_HIST_BINS; i += T) {
        const uint32_t v = shist[i];
        if (v) { atomicAdd(h + i, v); shist[i] = 0; }
      }
    }
    ds = dq = da = 0; mx = 0.f;
    __syncthreads();   // sred / shist are reused by the next run
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
  const GkBlk B = c.blk[(PASS == 2) ? (c.nblk - 1 - (int)blockIdx.x) : (int)blockIdx.x];  // zig-zag
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
// Leader work of the K2 epilogue for one segment: replay the reference's refinement loop on the
// per-candidate totals and record the decision.
__device__ void k2_decide_segment(const GkChain& c, int s, const GkSeg& sg, const uint32_t* tot, bool topk) {
  GkSegStat* st = c.stat + s;
  const double k = (double)sg.k;
  int chosen = 0;
  uint32_t count = 0;
  if (topk) {
    const uint32_t ties = st->ties;
    count = tot[0] + (tot[1] < ties ? tot[1] : ties);
  } else {
    if (c.mode == GK_MODE_GAUSSIAN) {
      // compression.py:372-381: at most nloops evaluations; the indexes returned are those of the
      // LAST EVALUATED threshold.
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
    }
    count = tot[chosen];
    st->thr = st->cand[chosen];
    st->ties = 0;
  }
  st->chosen = chosen;
  st->count = count;
}

// Epilogue shared by the threshold-count pass and the (gt,eq) pass: per-segment totals, the
// reference's refinement loop replayed on the device, then the prefix sums K3 needs.
__device__ void k2_epilogue(const GkChain& c, bool topk) {
  __shared__ uint32_t sscan[2][T];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool dbg = (c.rsv & 4) != 0;
  if (dbg && tid == 0) c.ctrl->ring[40] = (uint32_t)globaltimer_ns();
  // phase 1: the counting pass already accumulated the per-candidate totals of every segment with
  // atomics (stat.tot); one thread per segment replays the refinement loop.  Thread t owns segments
  // [t*per, (t+1)*per) in phases 1 and 2, so nothing is read back from global memory in between.
  __shared__ int s_chosen0;
  const int sper = (c.nseg + T - 1) / T;
  const int s0 = tid * sper, s1 = min(c.nseg, s0 + sper);
  uint32_t loc = 0;
  for (int s = s0; s < s1; ++s) {
    const GkSeg sg = c.seg[s];
    uint32_t tot[GK_NC];
#pragma unroll
    for (int j = 0; j < GK_NC; ++j) tot[j] = __ldcg(&c.stat[s].tot[j]);   // 4-byte aligned only
    k2_decide_segment(c, s, sg, tot, topk);
    loc += c.stat[s].count;
    if (s == 0) s_chosen0 = c.stat[s].chosen;
  }
  if (dbg && tid == 0) c.ctrl->ring[41] = (uint32_t)globaltimer_ns();
  // phase 2: exclusive scan of the per-segment counts -> out_base; total
  {
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
  if (dbg && tid == 0) c.ctrl->ring[42] = (uint32_t)globaltimer_ns();
  // phase 3: global exclusive scans over the table entries of (gt, eq); K3 subtracts the value at the
  // first entry of its segment to make them segment-relative.  Each thread owns a contiguous run of
  // entries whose counts stay in registers between the two sweeps when the run is short.
  {
    const int per = (c.nblk + T - 1) / T;
    const int b0 = tid * per, b1 = min(c.nblk, b0 + per);
    const bool one_seg = c.nseg == 1;
    const int ch0 = topk ? 0 : s_chosen0;
    constexpr int KEEP = 16;
    uint32_t vg[KEEP], ve[KEEP];
    uint32_t lg = 0, le = 0;
    if (per <= KEEP) {
      // all loads are issued unconditionally on clamped indices (one round trip instead of `per`
      // serialised ones: a bounds check in front of each load would stop the compiler from batching them)
      int chs[KEEP];
#pragma unroll
      for (int i = 0; i < KEEP; ++i) {
        const int b = min(b0 + i, c.nblk - 1);
        chs[i] = (topk || one_seg) ? ch0 : c.blk[b].seg;
      }
      if (!(topk || one_seg)) {
#pragma unroll
        for (int i = 0; i < KEEP; ++i) chs[i] = c.stat[chs[i]].chosen;
      }
#pragma unroll
      for (int i = 0; i < KEEP; ++i) {
        const int b = min(b0 + i, c.nblk - 1);
        vg[i] = __ldcg(c.bcnt + (size_t)b * GK_NC + chs[i]);
        ve[i] = topk ? __ldcg(c.bcnt + (size_t)b * GK_NC + 1) : 0u;
      }
#pragma unroll
      for (int i = 0; i < KEEP; ++i) {
        const bool in = (i < per) && (b0 + i < b1);
        vg[i] = in ? vg[i] : 0u;
        ve[i] = in ? ve[i] : 0u;
        lg += vg[i]; le += ve[i];
      }
    } else {
      for (int b = b0; b < b1; ++b) {
        const int ch = topk ? 0 : (one_seg ? ch0 : c.stat[c.blk[b].seg].chosen);
        lg += __ldcg(c.bcnt + (size_t)b * GK_NC + ch);
        if (topk) le += __ldcg(c.bcnt + (size_t)b * GK_NC + 1);
      }
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
    if (per <= KEEP) {
#pragma unroll
      for (int i = 0; i < KEEP; ++i) {
        const int b = b0 + i;
        if (i < per && b < b1) {
          *reinterpret_cast<uint2*>(c.boff + 2 * (size_t)b) = make_uint2(rg, re);
          rg += vg[i]; re += ve[i];
        }
      }
    } else {
      for (int b = b0; b < b1; ++b) {
        const int ch = topk ? 0 : (one_seg ? ch0 : c.stat[c.blk[b].seg].chosen);
        c.boff[2 * (size_t)b] = rg;
        c.boff[2 * (size_t)b + 1] = re;
        rg += __ldcg(c.bcnt + (size_t)b * GK_NC + ch);
        if (topk) re += __ldcg(c.bcnt + (size_t)b * GK_NC + 1);
      }
    }
  }
  __syncthreads();
  if (dbg && tid == 0) c.ctrl->ring[43] = (uint32_t)globaltimer_ns();
}

// Persistent and barrier-free: a block walks the table entries (backwards, see below); inside an entry
// every warp owns a contiguous 1024-element slice per 8192-element sub-tile, counts its slice in
// registers and adds its (non-zero) per-candidate counts to the entry's row of bcnt with integer atomics
// (K1 zeroed the rows).  Integer adds commute, so the result is deterministic.
template <int NCAND>
__global__ void __launch_bounds__(GK_THREADS) k_count(GkChain c) {
  constexpr bool SKIP_ROWS = NCAND > 6;   // gaussian (<= 6 candidates): the smallest is exceeded in every row
  constexpr int NLOW = (NCAND + 1) / 2;    // candidates evaluated unconditionally
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int cur_seg = -1;
  float cs[NCAND];        // candidates of the current segment in ascending order
  uint32_t ord[NCAND];    // their original indices
#pragma unroll
  for (int j = 0; j < NCAND; ++j) { cs[j] = 0.f; ord[j] = j; }
  __shared__ uint32_t s_tot[GK_NC];
  if (tid < GK_NC) s_tot[tid] = 0;
  if ((c.rsv & 4) && blockIdx.x == 0 && tid == 0) c.ctrl->ring[44] = (uint32_t)globaltimer_ns();
  __syncthreads();
  uint32_t run[NCAND];     // this lane's counts since the last flush (current segment)
#pragma unroll
  for (int j = 0; j < NCAND; ++j) run[j] = 0;
  for (int it = blockIdx.x; it < c.nblk; it += gridDim.x) {
    // zig-zag: K1 walked the arena forwards, so its tail is what is still in L2 -> walk backwards here
    const int eidx = c.nblk - 1 - it;
    const GkBlk B = c.blk[eidx];
    if (B.seg != cur_seg) {   // block-uniform
      if (cur_seg >= 0) {     // segment change inside a block (multi-segment tables): flush straight to global
#pragma unroll
        for (int j = 0; j < NCAND; ++j) {
          const uint32_t w = warp_sum_u(run[j]);
          if (lane == 0 && w) atomicAdd(&c.stat[cur_seg].tot[ord[j]], w);
          run[j] = 0;
        }
      }
      cur_seg = B.seg;
      const GkSegStat* st = c.stat + B.seg;
#pragma unroll
      for (int j = 0; j < NCAND; ++j) { ord[j] = st->order[j]; cs[j] = st->cand[ord[j]]; }
    }
    uint32_t cnt[NCAND];
#pragma unroll
    for (int j = 0; j < NCAND; ++j) cnt[j] = 0;
    const float* __restrict__ a = c.resid + B.start;
    for (int t0 = 0; t0 < B.len; t0 += GK_SUB) {
      const int wbeg = t0 + warp * 1024;
      if (wbeg >= B.len) break;          // warp-uniform
      if (wbeg + 1024 <= B.len) {        // warp-uniform fast path: a full 1024-element slice
        const float* __restrict__ aw = a + wbeg + lane * 4;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          float4 v[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) v[i] = ld4(aw + (half * 4 + i) * 128);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float e0 = fabsf(v[i].x), e1 = fabsf(v[i].y), e2 = fabsf(v[i].z), e3 = fabsf(v[i].w);
            // Candidates are sorted ascending.  The lower half is evaluated for every element; the upper
            // half (far-tail thresholds) only in rows where some lane exceeds the smallest of them — one
            // row maximum + vote instead of 4 compares per candidate.  Exact either way.
            const float m = fmaxf(fmaxf(e0, e1), fmaxf(e2, e3));
            if (SKIP_ROWS && !__any_sync(0xffffffffu, m > cs[0])) continue;
#pragma unroll
            for (int j = 0; j < NLOW; ++j)
              cnt[j] -= (gt_mask(e0, cs[j]) + gt_mask(e1, cs[j])) + (gt_mask(e2, cs[j]) + gt_mask(e3, cs[j]));
            if (NLOW < NCAND && __any_sync(0xffffffffu, m > cs[NLOW < NCAND ? NLOW : 0])) {
#pragma unroll
              for (int j = NLOW; j < NCAND; ++j)
                cnt[j] -= (gt_mask(e0, cs[j]) + gt_mask(e1, cs[j])) + (gt_mask(e2, cs[j]) + gt_mask(e3, cs[j]));
            }
          }
        }
      } else {                            // ragged end of a tensor: plain scalar loop
        const int wend = min(wbeg + 1024, B.len);
        for (int i = wbeg + lane; i < wend; i += 32) {
          const float x = fabsf(a[i]);
#pragma unroll
          for (int j = 0; j < NCAND; ++j) cnt[j] += (uint32_t)(x > cs[j]);
        }
      }
    }
    uint32_t* row = c.bcnt + (size_t)eidx * GK_NC;
#pragma unroll
    for (int j = 0; j < NCAND; ++j) {
      run[j] += cnt[j];
      const uint32_t w = warp_sum_u(cnt[j]);
      if (lane == 0 && w) atomicAdd(row + ord[j], w);
    }
  }
  // segment totals: lanes -> warp (REDUX) -> block (smem atomics) -> one global atomic per candidate and block
  if (cur_seg >= 0) {
#pragma unroll
    for (int j = 0; j < NCAND; ++j) {
      const uint32_t w = warp_sum_u(run[j]);
      if (lane == 0 && w) atomicAdd(&s_tot[ord[j]], w);
    }
  }
  __syncthreads();
  if (tid < GK_NC && cur_seg >= 0 && s_tot[tid]) atomicAdd(&c.stat[cur_seg].tot[tid], s_tot[tid]);
  if (c.rsv & 1) return;   // experiment switch (tools/bench_chain.py): skip the epilogue
  if ((c.rsv & 4) && tid == 0) atomicMax(&c.ctrl->ring[45], (uint32_t)globaltimer_ns());
  if (last_block_done(&c.ctrl->done[2])) k2_epilogue(c, false);
}

// top-k: per-block counts of key > T (column 0) and key == T (column 1)
template <bool MINK>
__global__ void __launch_bounds__(GK_THREADS) k_count_topk(GkChain c) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int it = blockIdx.x; it < c.nblk; it += gridDim.x) {
    const int eidx = c.nblk - 1 - it;  // zig-zag (see k_count)
    const GkBlk B = c.blk[eidx];
    const uint32_t tb = c.stat[B.seg].tbits;
    uint32_t gt = 0, eq = 0;
    const float* __restrict__ a = c.resid + B.start;
    for (int t0 = 0; t0 < B.len; t0 += GK_SUB) {
      const int wbeg = t0 + warp * 1024;
      if (wbeg >= B.len) break;
      if (wbeg + 1024 <= B.len) {
        float4 v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = ld4(a + wbeg + (i * 32 + lane) * 4);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float e[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t key = sel_key<MINK>(e[j]);
            gt += (uint32_t)(key > tb);
            eq += (uint32_t)(key == tb);
          }
        }
      } else {
        for (int i = wbeg + lane; i < min(wbeg + 1024, B.len); i += 32) {
          const uint32_t key = sel_key<MINK>(a[i]);
          gt += (uint32_t)(key > tb);
          eq += (uint32_t)(key == tb);
        }
      }
    }
    gt = warp_sum_u(gt); eq = warp_sum_u(eq);
    uint32_t* row = c.bcnt + (size_t)eidx * GK_NC;
    if (lane == 0) {
      if (gt) { atomicAdd(row, gt); atomicAdd(&c.stat[B.seg].tot[0], gt); }
      if (eq) { atomicAdd(row + 1, eq); atomicAdd(&c.stat[B.seg].tot[1], eq); }
    }
  }
  if (last_block_done(&c.ctrl->done[2])) k2_epilogue(c, true);
}

// ------------------------------------------------------- K3 select + compact + residual + send
// Every warp owns a contiguous 1024-element slice of the block's 8192-element sub-tile and keeps it
// in registers (8 float4 per lane).  One block-level exchange of the per-warp counts gives every warp
// its output base; inside a warp the order is recovered with ballots, and only for the (rare) float4
// rows that contain a selected element.  One __syncthreads per sub-tile, no per-row scans.

template <bool TOPK, bool MINK>
__global__ void __launch_bounds__(GK_THREADS) k_compact_send(GkChain c, GkDst d) {
  __shared__ uint32_t s_wcnt[2][NWARP];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t seqn = c.ctrl->seq + 1u;  // every block reads it before the last block bumps it
  const uint64_t poff = (uint64_t)(seqn & 1u) * d.parity_stride;
  const uint32_t cap = c.cap;
  const uint32_t lt = (1u << lane) - 1u;
  int par = 0;
  for (int b = blockIdx.x; b < c.nblk; b += gridDim.x) {   // persistent over the table entries
  const GkBlk B = c.blk[b];
  const GkSegStat* st = c.stat + B.seg;
  const float thr = st->thr;
  const uint32_t tb = st->tbits, ties = st->ties;
  const uint32_t seg_first = (uint32_t)c.seg[B.seg].blk_begin;
  const uint2 off_b = *reinterpret_cast<const uint2*>(c.boff + 2 * (size_t)b);
  const uint2 off_s = *reinterpret_cast<const uint2*>(c.boff + 2 * (size_t)seg_first);
  const uint32_t base_gt = st->out_base + (off_b.x - off_s.x);
  const uint32_t base_eq = TOPK ? (off_b.y - off_s.y) : 0u;
  float* __restrict__ a = c.resid + B.start;
  uint32_t run_gt = 0, run_eq = 0;  // selected before the current sub-tile, inside this entry
  for (int t0 = 0; t0 < B.len; t0 += GK_SUB, par ^= 1) {
    const int wbeg = t0 + warp * 1024;
    float4 v[8];
    uint32_t fg = 0, fe = 0;  // bit 4*i+j: element j of row i is gt / eq
    if (wbeg + 1024 <= B.len) {  // warp-uniform fast path: the whole slice is in range
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = ld4(a + wbeg + (i * 32 + lane) * 4);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float e[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (TOPK) {
            const uint32_t key = sel_key<MINK>(e[j]);
            if (key > tb) fg |= 1u << (4 * i + j);
            if (key == tb) fe |= 1u << (4 * i + j);
          } else {
            fg |= gt_mask(fabsf(e[j]), thr) & (1u << (4 * i + j));
          }
        }
      }
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int i0 = wbeg + (i * 32 + lane) * 4;
        if (i0 + 3 < B.len) {
          v[i] = ld4(a + i0);
        } else {
          v[i].x = (i0 < B.len) ? a[i0] : 0.f;
          v[i].y = (i0 + 1 < B.len) ? a[i0 + 1] : 0.f;
          v[i].z = (i0 + 2 < B.len) ? a[i0 + 2] : 0.f;
          v[i].w = 0.f;
        }
        const float e[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const bool valid = (i0 + j) < B.len;
          if (TOPK) {
            const uint32_t key = sel_key<MINK>(e[j]);
            if (valid && key > tb) fg |= 1u << (4 * i + j);
            if (valid && key == tb) fe |= 1u << (4 * i + j);
          } else {
            if (valid && fabsf(e[j]) > thr) fg |= 1u << (4 * i + j);
          }
        }
      }
    }
    const uint32_t wsum = warp_sum_u((uint32_t)__popc(fg) | ((uint32_t)__popc(fe) << 16));
    if (lane == 0) s_wcnt[par][warp] = wsum;
    __syncthreads();
    uint32_t wbase = 0, total = 0;
#pragma unroll
    for (int w = 0; w < NWARP; ++w) {
      const uint32_t x = s_wcnt[par][w];
      if (w < warp) wbase += x;
      total += x;
    }
    if (wsum != 0) {  // warp-uniform
      uint32_t gb = run_gt + (wbase & 0xffffu);            // gt elements before the current row
      uint32_t eb = base_eq + run_eq + (wbase >> 16);      // eq elements before the current row (segment-wide)
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const uint32_t fgi = (fg >> (4 * i)) & 15u, fei = (fe >> (4 * i)) & 15u;
        if (__ballot_sync(0xffffffffu, (fgi | fei) != 0u) == 0u) continue;  // warp-uniform
        uint32_t gbefore = gb, ebefore = eb, tg = 0, te = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          // lanes hold 4 consecutive elements each: everything in lower lanes precedes this lane
          const uint32_t bg = __ballot_sync(0xffffffffu, (fgi >> j) & 1u);
          gbefore += __popc(bg & lt);
          tg += __popc(bg);
          if (TOPK) {
            const uint32_t be = __ballot_sync(0xffffffffu, (fei >> j) & 1u);
            ebefore += __popc(be & lt);
            te += __popc(be);
          }
        }
        if (fgi | fei) {
          const int i0 = wbeg + (i * 32 + lane) * 4;
          const float e[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const bool isg = (fgi >> j) & 1u, ise = (fei >> j) & 1u;
            if (isg || (ise && ebefore < ties)) {
              const uint32_t pos = base_gt + gbefore + (ebefore < ties ? ebefore : ties);
              if (pos < cap) {
                const uint32_t idx = (uint32_t)(B.start + i0 + j);
                const uint64_t pair = (uint64_t)idx | ((uint64_t)__float_as_uint(e[j]) << 32);
                for (int q = 0; q < d.ndst; ++q) d.slot[q][poff + GK_SLOT_HDR + pos] = pair;
                a[i0 + j] = 0.f;  // sent => leaves the residual
              }
            }
            gbefore += isg ? 1u : 0u;
            ebefore += ise ? 1u : 0u;
          }
        }
        gb += tg;
        eb += te;
      }
    }
    run_gt += total & 0xffffu;
    run_eq += total >> 16;
  }
  }  // entries
  // Every block: bar.sync + one gpu-scope fence by thread 0 + ticket.  The last block then issues ONE
  // system-scope fence before the release store of the header: fences are cumulative, so the pairs the
  // other blocks stored to peer memory (ordered before their tickets at gpu scope) are ordered before
  // the header at system scope (same pattern as NCCL's barrier + single-thread fence before a post).
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
// Slots [r_begin, r_end) are applied by one launch.  With ORDERED (one launch per source rank, issued in
// rank order) every coordinate receives its contributions in the same order on every GPU, so all replicas
// compute bit-identical averaged gradients no matter how the atomics of a single launch would have been
// scheduled; inside one rank's list the indices are unique, so plain read-modify-writes suffice.
template <bool UNSCATTER, bool ORDERED>
__global__ void __launch_bounds__(GK_THREADS) k_scatter(GkSrc s, GkCtrl* ctrl, float* __restrict__ grad,
                                                         float scale, int nelem, int r_begin, int r_end) {
  __shared__ uint32_t scnt[GK_MAXP];
  __shared__ int s_bad;
  const int tid = threadIdx.x;
  const uint32_t seq = ctrl->seq;
  const uint64_t* base = s.arena + (uint64_t)(seq & 1u) * s.parity_stride;
  const uint32_t cap = (uint32_t)(s.slot_stride - GK_SLOT_HDR);
  if (tid == 0) s_bad = 0;
  __syncthreads();
  if (tid >= r_begin && tid < r_end) {
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
  if (!UNSCATTER && ctrl->err) return;   // an earlier slot of this exchange timed out: leave grad alone
  for (int r = r_begin; r < r_end; ++r) {
    const uint64_t* pairs = base + (uint64_t)r * s.slot_stride + GK_SLOT_HDR;
    const uint32_t n = scnt[r];
    for (uint32_t j = blockIdx.x * T + tid; j < n; j += gridDim.x * T) {
      const uint64_t p = __ldcg(pairs + j);
      const uint32_t idx = (uint32_t)p;
      if (idx < (uint32_t)nelem) {
        const float v = __uint_as_float((uint32_t)(p >> 32)) * scale;
        if (UNSCATTER) grad[idx] = 0.f;
        else if (ORDERED) grad[idx] += v;
        else atomicAdd(grad + idx, v);
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
      }
    }
  }
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

// ---- gradient norm without caller-provided scratch -------------------------------------------
// Some callers cannot hand in scratch and must not allocate it either: a host application that
// carved up all device memory at start-up and links this library through the C ABI (it has no
// allocator left to ask), a monitoring or abort path that has to work while the device is out of
// memory, or code that runs while other streams are busy and therefore must avoid cudaMalloc /
// cudaFree, which synchronise the whole device.  For them the per-block partial sums and the
// last-block ticket of the norm kernel live in this fixed array instead.
//   * capacity: GK_NOALLOC_MAX_BLOCKS table entries (16384 -> 134 M elements at the default 8192-element
//     chunk); larger tables are rejected with GK_ERR_NOALLOC_TOO_LARGE and a message on stderr;
//   * cost: sizeof(GkNoAllocScratch) = 128 KiB + one GkCtrl of device memory on EVERY GPU that loads
//     the library, held until the process exits;
//   * the array is shared by all calls and keeps the previous call's contents.  The partial sums are
//     fully rewritten by each call; the ticket word must start at zero and is cleared on the stream
//     before every launch.  Two calls that overlap on the GPU (different streams, or different host
//     threads) race on it: callers must order them.
#define GK_NOALLOC_MAX_DEVICES 64   /* GK_NOALLOC_MAX_BLOCKS, GK_ERR_NOALLOC_TOO_LARGE: include/gk_b200.h */

struct GkNoAllocScratch {
  double part[GK_NOALLOC_MAX_BLOCKS];
  GkCtrl ctrl;
};

static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(GkNoAllocScratch)];

// In host code `g_noalloc_scratch` is only a handle for the runtime, not a device address; the address
// each device uses is looked up once per device (with that device current) and cached here.
static void* g_noalloc_addr[GK_NOALLOC_MAX_DEVICES];
static std::mutex g_noalloc_mutex;

static int gk_noalloc_scratch_ptr(GkNoAllocScratch** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= GK_NOALLOC_MAX_DEVICES) return (int)cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (g_noalloc_addr[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);   // address of the array on the CURRENT device
    if (e != cudaSuccess) return (int)e;
    g_noalloc_addr[dev] = p;
  }
  *out = reinterpret_cast<GkNoAllocScratch*>(g_noalloc_addr[dev]);
  return 0;
}

// =========================================================================== host entry points
static inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }

// Exactly one resident wave for a persistent kernel: SM count x the occupancy the runtime reports for
// that kernel (148 SMs on a B200).  Queried once per kernel and process; no allocation involved.
template <typename K>
static int gk_persistent_grid(K kernel, int* cache, int nblk) {
  if (*cache == 0) {
    int dev = 0, sms = 148, per = 4;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, kernel, T, 0) != cudaSuccess || per < 1) per = 4;
    const char* env = getenv("GK_BLOCKS_PER_SM");
    if (env && atoi(env) > 0 && atoi(env) < per) per = atoi(env);
    *cache = sms * per;
  }
  if (nblk <= *cache) return nblk;
  // every block gets the same number of table entries (+-1 at the very end): with ~4 entries per block a
  // plain "one resident wave" grid leaves some blocks with 25 % more work than others
  const int per_block = (nblk + *cache - 1) / *cache;
  return (nblk + per_block - 1) / per_block;
}
#define GK_LAUNCH_PERSISTENT(kernel, nblk, stream, ...)                         \
  do {                                                                          \
    static int cache_ = 0;                                                      \
    const int grid_ = gk_persistent_grid(kernel, &cache_, nblk);                \
    kernel<<<grid_, T, 0, S(stream)>>>(__VA_ARGS__);                            \
  } while (0)
#define GK_RET() return (int)cudaGetLastError()

extern "C" {

const char* gk_version(void) { return "gk_b200 cuda sm_100a 0.1"; }

int gk_accumulate(const GkChain* c, void* stream) {
  if (c->nblk <= 0) return 0;
  if (c->mode == GK_MODE_TOPK) GK_LAUNCH_PERSISTENT((k_accumulate<true, false>), c->nblk, stream, *c);
  else if (c->mode == GK_MODE_MINK) GK_LAUNCH_PERSISTENT((k_accumulate<true, true>), c->nblk, stream, *c);
  else GK_LAUNCH_PERSISTENT((k_accumulate<false, false>), c->nblk, stream, *c);
  GK_RET();
}

int gk_count_pick(const GkChain* c, void* stream) {
  if (c->nblk <= 0) return 0;
  int nc = 1;
  if (c->mode == GK_MODE_GAUSSIAN) nc = c->nloops * (c->nloops + 1) / 2;
  else if (c->mode == GK_MODE_REDSYNC) nc = 7;
  else if (c->mode == GK_MODE_REDSYNCTRIM) nc = GK_NC;
  if (nc <= 1) GK_LAUNCH_PERSISTENT(k_count<1>, c->nblk, stream, *c);
  else if (nc <= 6) GK_LAUNCH_PERSISTENT(k_count<6>, c->nblk, stream, *c);
  else if (nc <= 7) GK_LAUNCH_PERSISTENT(k_count<7>, c->nblk, stream, *c);
  else GK_LAUNCH_PERSISTENT(k_count<GK_NC>, c->nblk, stream, *c);
  GK_RET();
}

int gk_topk_select(const GkChain* c, void* stream) {
  if (c->nblk <= 0) return 0;
  if (c->mode == GK_MODE_MINK) {
    k_radix_pass<2, true><<<c->nblk, T, 0, S(stream)>>>(*c);
    k_radix_pass<3, true><<<c->nblk, T, 0, S(stream)>>>(*c);
    GK_LAUNCH_PERSISTENT(k_count_topk<true>, c->nblk, stream, *c);
  } else {
    k_radix_pass<2, false><<<c->nblk, T, 0, S(stream)>>>(*c);
    k_radix_pass<3, false><<<c->nblk, T, 0, S(stream)>>>(*c);
    GK_LAUNCH_PERSISTENT(k_count_topk<false>, c->nblk, stream, *c);
  }
  GK_RET();
}

int gk_compact_send(const GkChain* c, const GkDst* dst, void* stream) {
  if (c->nblk <= 0) return 0;
  if (dst->ndst < 1 || dst->ndst > GK_MAXP) return (int)cudaErrorInvalidValue;
  if (c->mode == GK_MODE_TOPK) GK_LAUNCH_PERSISTENT((k_compact_send<true, false>), c->nblk, stream, *c, *dst);
  else if (c->mode == GK_MODE_MINK) GK_LAUNCH_PERSISTENT((k_compact_send<true, true>), c->nblk, stream, *c, *dst);
  else GK_LAUNCH_PERSISTENT((k_compact_send<false, false>), c->nblk, stream, *c, *dst);
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
  if (src->nsrc <= 2) {
    // 0 + a + b is the same in either order: one launch with atomics is already bit-reproducible
    k_scatter<false, false><<<grid, T, 0, S(stream)>>>(*src, ctrl, grad, scale, nelem, 0, src->nsrc);
  } else {
    // three or more contributions per coordinate: fixed rank order, one (tiny) launch per source rank;
    // launch r only waits for rank r's header, so early arrivals are applied while later ones are in flight
    for (int r = 0; r < src->nsrc; ++r)
      k_scatter<false, true><<<grid, T, 0, S(stream)>>>(*src, ctrl, grad, scale, nelem, r, r + 1);
  }
  GK_RET();
}

int gk_unscatter(const GkSrc* src, GkCtrl* ctrl, float* grad, int32_t nelem, void* stream) {
  if (src->nsrc < 1 || src->nsrc > GK_MAXP) return (int)cudaErrorInvalidValue;
  const uint64_t cap = src->slot_stride - GK_SLOT_HDR;
  int grid = (int)((cap + T - 1) / T);
  if (grid > 296) grid = 296;
  if (grid < 1) grid = 1;
  k_scatter<true, false><<<grid, T, 0, S(stream)>>>(*src, ctrl, grad, 0.f, nelem, 0, src->nsrc);
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

int gk_grad_norm_static(const float* grad, const GkBlk* blk, int32_t nblk, float max_norm, float* out,
                        void* stream) {
  if (nblk <= 0) return 0;
  if (nblk > GK_NOALLOC_MAX_BLOCKS) {
    fprintf(stderr, "gk_grad_norm_static: table has %d entries, the built-in scratch holds %d "
                    "(use gk_grad_norm with caller-provided scratch)\n", nblk, GK_NOALLOC_MAX_BLOCKS);
    return GK_ERR_NOALLOC_TOO_LARGE;
  }
  GkNoAllocScratch* sc = nullptr;
  const int err = gk_noalloc_scratch_ptr(&sc);   // first call on a device: symbol look-up
  if (err) return err;
  // the array still holds whatever the previous call left; only the ticket has to start at zero
  cudaError_t e = cudaMemsetAsync(&sc->ctrl, 0, sizeof(GkCtrl), S(stream));
  if (e != cudaSuccess) return (int)e;
  k_grad_norm<<<nblk, T, 0, S(stream)>>>(grad, blk, nblk, sc->part, max_norm, out, &sc->ctrl);
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
