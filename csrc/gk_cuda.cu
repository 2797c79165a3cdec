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

// -------------------------------------------------------------- row maxima and gated reading
// K1 has every element of acc in registers once; it leaves behind, per 128-element row of every table
// entry, the largest |acc| bit pattern (GkChain.rowmax, n/128 words = 0.8 MB for ResNet-50).  All later
// passes of the chain look for elements ABOVE some threshold, and at density 0.001 only about one row in
// eight contains one: they read the row maxima and then only the rows that can contribute.  Counts and
// selections stay exact — a skipped row provably holds nothing above the gate.
#define GK_ROW 128           /* elements per row: one float4 per lane */
#define GK_RC 5              /* candidates evaluated per counting round (>= the widest round of any mode) */

__device__ __forceinline__ uint32_t ld_relaxed_gpu_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Gate for "|x| > thr" on row maxima: a row can hold such an element iff (int)rowmax_bits > gate.
// Negative thresholds select everything (gate -1); the comparison of the elements themselves is done in
// floating point, the gate only has to be conservative.
__device__ __forceinline__ int gate_of(float thr) { return (thr < 0.f) ? -1 : __float_as_int(thr); }

// One row of an entry as a float4 per lane; positions past the end of the entry read as NaN, which no
// comparison selects.
__device__ __forceinline__ float4 ld_row(const float* __restrict__ a, int row, int lane, int len) {
  const int i0 = row * GK_ROW + lane * 4;
  if (i0 + 3 < len) return ld4(a + i0);
  const float qn = __int_as_float(0x7fc00000);
  float4 v;
  v.x = (i0 < len) ? a[i0] : qn;
  v.y = (i0 + 1 < len) ? a[i0 + 1] : qn;
  v.z = (i0 + 2 < len) ? a[i0 + 2] : qn;
  v.w = qn;
  return v;
}

// ----------------------------------------------------------------- K2 search state machine
// The reference evaluates its candidate thresholds one after the other, each with a full pass and a host
// synchronisation (compression.py:372-381, :653-672, :719-732).  Here the search is replayed on the device
// on the totals known so far; when it needs a candidate that has not been counted yet, the next ROUND
// counts that candidate together with the others the search could ask for at the same depth:
//   gaussian (L evaluations): round a counts c*t0 for c = 0.5^a * 1.5^b, all b        (<= L rounds)
//   redsync:                  round d counts the nodes of depth d of the bisection tree (<= 3 rounds)
//   redsynctrim:              round r counts ratios 4r .. 4r+3                          (<= 4 rounds)
// In the regime the method is built for the first round settles it.  Returns true when the search is
// complete; otherwise st->rcand[0..nrc) / st->gate describe the next round.
__device__ bool k2_advance(const GkChain& c, const GkSeg& sg, GkSegStat* st) {
  const double k = (double)sg.k;
  const uint32_t known = st->known;
  int nrc = 0;
  uint8_t rc[GK_RC];
  if (c.mode == GK_MODE_GAUSSIAN) {
    int a = 0, b = 0, need = -1;
    for (int loop = 0; loop < c.nloops; ++loop) {
      const int idx = tri_idx(a, b);
      if (!((known >> idx) & 1u)) { need = a; break; }
      const double cnt = (double)__ldcg(&st->tot[idx]);
      if (cnt < 2.0 * k / 3.0) ++a;
      else if (cnt > 4.0 * k / 3.0) ++b;
      else break;
    }
    if (need >= 0)
      for (int bb = 0; need + bb < c.nloops && nrc < GK_RC; ++bb) rc[nrc++] = (uint8_t)tri_idx(need, bb);
  } else if (c.mode == GK_MODE_REDSYNC) {
    double l = 0.0, r = 1.0;
    int node = 0, need = -1;
    while (r - l > 0.2) {
      if (!((known >> node) & 1u)) { need = node; break; }
      const double mid = l + (r - l) / 2;
      const double nnz = (double)__ldcg(&st->tot[node]);
      if (nnz > k && 2.0 * k > nnz) break;
      else if (nnz < k / 2.0) { r = mid; node = 2 * node + 1; }
      else { l = mid; node = 2 * node + 2; }
      if (node >= 7) break;
    }
    if (need >= 0) {
      const int first = (need == 0) ? 0 : (need < 3 ? 1 : 3), cntn = first + 1;   // depth 0: {0}, 1: {1,2}, 2: {3..6}
      for (int j = 0; j < cntn && nrc < GK_RC; ++j) rc[nrc++] = (uint8_t)(first + j);
    }
  } else if (c.mode == GK_MODE_REDSYNCTRIM) {
    for (int j = 0; j < GK_NC; ++j) {
      if (!((known >> j) & 1u)) {
        for (int q = j; q < GK_NC && q < j + 4; ++q) rc[nrc++] = (uint8_t)q;
        break;
      }
      if ((double)__ldcg(&st->tot[j]) >= k) break;
    }
  } else {  // GK_MODE_THRESHOLD: one candidate, one round
    if (!(known & 1u)) rc[nrc++] = 0;
  }
  if (nrc == 0) return true;
  float cmin = st->cand[rc[0]];
  for (int j = 1; j < nrc; ++j) cmin = fminf(cmin, st->cand[rc[j]]);   // fminf ignores NaN candidates
  for (int j = 0; j < nrc; ++j) st->rcand[j] = rc[j];
  st->nrc = nrc;
  st->gate = gate_of(cmin);
  return false;
}

// ----------------------------------------------------------------- K1 accumulate + moments
// Epilogue of K1, run by the last block: per-segment moments (fixed-order warp reduction of
// the per-block partials => deterministic), then the candidate thresholds of the counting
// rounds, or the first radix digit for top-k.
// Leader work of the K1 epilogue for one segment: moments -> stat, candidate thresholds, first round.
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
  if (c.mode == GK_MODE_TOPK || c.mode == GK_MODE_MINK || c.mode == GK_MODE_DGC) {
    c.stat[s].prefix = 0; c.stat[s].krem = (uint32_t)sg.k;
    st->done = 0;
  } else {
    st->known = 0;
    st->nrc = 0;
    st->done = k2_advance(c, sg, st) ? 1 : 0;   // plans the first counting round
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
__device__ void radix_pick_warp_at(uint32_t* h, uint32_t* prefix_p, uint32_t* krem_p, int nbins, int lane) {
  const uint32_t krem = *krem_p;
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
    if (lane == 0) *prefix_p = *prefix_p << ((nbins == 2048) ? 11 : 9);
  } else if (mine) {
    uint32_t run = above;
    int digit = lane * per;
    for (int j = per - 1; j >= 0; --j) {
      const uint32_t v = __ldcg(h + lane * per + j);
      if (run + v >= krem) { digit = lane * per + j; break; }
      run += v;
    }
    const int width = (nbins == 2048) ? 11 : 9;
    *prefix_p = (*prefix_p << width) | (uint32_t)digit;
    *krem_p = krem - run;  // rank inside the chosen bucket (1-based)
  }
  __syncwarp();
  for (int j = lane; j < nbins; j += 32) h[j] = 0;
}

__device__ __forceinline__ void radix_pick_warp(const GkChain& c, int s, int shift, int nbins, int lane) {
  (void)shift;
  radix_pick_warp_at(c.hist + (size_t)s * GK_HIST_BINS, &c.stat[s].prefix, &c.stat[s].krem, nbins, lane);
}

// Persistent: a CUDA block walks the table entries blockIdx.x, blockIdx.x + gridDim.x, ... and keeps
// its partial moments in registers for as long as consecutive entries belong to the same segment (a
// "run"); only the last entry of a run pays for the block reduction, the others record zeros.  In the
// merged-buffer mode (one segment) that is one reduction and one ticket per block.
// A warp handles one 128-element row per step (rows warp, warp+8, ... of the entry), so the row maximum
// is one REDUX on the bit patterns.
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
    uint32_t* __restrict__ rmx = c.rowmax ? c.rowmax + (size_t)eidx * c.rpb : nullptr;
    // fp32 partials over one entry (<= chunk/256 elements per thread), fp64 across entries
    float s = 0.f, q = 0.f, sa = 0.f;
    const int nrow = (B.len + GK_ROW - 1) / GK_ROW;
    if (tid < GK_NC) c.bcnt[(size_t)eidx * GK_NC + tid] = 0;   // entries no counting round visits stay at zero
#pragma unroll 2
    for (int row = warp; row < nrow; row += NWARP) {
      const int i0 = row * GK_ROW + lane * 4;
      float4 x;
      int nv = 4;                       // valid elements of this lane's quad
      if (i0 + 3 < B.len) {
        x = ld4_stream(g + i0);
        if (user) {
          const float4 rv = ld4_stream(r + i0);
          x.x += rv.x; x.y += rv.y; x.z += rv.z; x.w += rv.w;
        }
        st4(r + i0, x);  // acc: default policy so that it can stay in L2 for the following passes
        if (zg) st4_stream(const_cast<float*>(g) + i0, z4);
      } else {                          // ragged end of a tensor whose size is not a multiple of 4
        nv = B.len - i0;
        nv = nv < 0 ? 0 : nv;
        float e[4] = {0.f, 0.f, 0.f, 0.f};
        for (int j = 0; j < nv; ++j) {
          float v = g[i0 + j];
          if (user) v += r[i0 + j];
          r[i0 + j] = v;
          if (zg) const_cast<float*>(g)[i0 + j] = 0.f;
          e[j] = v;
        }
        x = make_float4(e[0], e[1], e[2], e[3]);
      }
      s += (x.x + x.y) + (x.z + x.w);
      q = fmaf(x.x, x.x, q); q = fmaf(x.y, x.y, q); q = fmaf(x.z, x.z, q); q = fmaf(x.w, x.w, q);
      const float ax = fabsf(x.x), ay = fabsf(x.y), az = fabsf(x.z), aw = fabsf(x.w);
      sa += (ax + ay) + (az + aw);
      const float m4 = fmaxf(fmaxf(ax, ay), fmaxf(az, aw));
      mx = fmaxf(mx, m4);
      if (rmx) {
        // bit patterns of non-negative floats order like the floats (NaN sorts above everything: such a row
        // is always read); padding zeros cannot raise the maximum
        const uint32_t mb = max(max(absbits(x.x), absbits(x.y)), max(absbits(x.z), absbits(x.w)));
        const uint32_t rm = __reduce_max_sync(0xffffffffu, mb);
        if (lane == 0) rmx[row] = rm;
      }
      if (HIST) {
        if (nv == 4) {
          atomicAdd(&shist[sel_key<MINK>(x.x) >> 20], 1u); atomicAdd(&shist[sel_key<MINK>(x.y) >> 20], 1u);
          atomicAdd(&shist[sel_key<MINK>(x.z) >> 20], 1u); atomicAdd(&shist[sel_key<MINK>(x.w) >> 20], 1u);
        } else {
          const float e[4] = {x.x, x.y, x.z, x.w};
          for (int j = 0; j < nv; ++j) atomicAdd(&shist[sel_key<MINK>(e[j]) >> 20], 1u);
        }
      }
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
// One block per table entry; a row whose maximum lies below the prefix cannot match and is not read
// (largest-k only: for min-k the keys are reversed and every row is read).
template <int PASS, bool MINK>
__global__ void __launch_bounds__(GK_THREADS) k_radix_pass(GkChain c) {
  constexpr int NB = (PASS == 2) ? 2048 : 512;
  constexpr int SHIFT_MATCH = (PASS == 2) ? 20 : 9;
  constexpr int SHIFT_DIGIT = (PASS == 2) ? 9 : 0;
  __shared__ uint32_t shist[NB];
  __shared__ int s_any;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int eidx = (int)blockIdx.x;
  const GkBlk B = c.blk[eidx];
  const GkSegStat* st = c.stat + B.seg;
  const bool skip = st->done != 0;           // DGC: the sampled threshold was accepted, no exact select needed
  for (int i = tid; i < NB; i += T) shist[i] = 0;
  if (tid == 0) s_any = 0;
  __syncthreads();
  const uint32_t prefix = st->prefix;
  const float* __restrict__ a = c.resid + B.start;
  const uint32_t* __restrict__ rmx = (c.rowmax && !MINK) ? c.rowmax + (size_t)eidx * c.rpb : nullptr;
  const int nrow = skip ? 0 : (B.len + GK_ROW - 1) / GK_ROW;
  bool any = false;
  for (int row = warp; row < nrow; row += NWARP) {
    if (rmx && (rmx[row] >> SHIFT_MATCH) < prefix) continue;     // warp-uniform
    const float4 x = ld_row(a, row, lane, B.len);
    const float e[4] = {x.x, x.y, x.z, x.w};
    const int nv = B.len - (row * GK_ROW + lane * 4);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t key = sel_key<MINK>(e[j]);
      if (j < nv && (key >> SHIFT_MATCH) == prefix) { atomicAdd(&shist[(key >> SHIFT_DIGIT) & (NB - 1)], 1u); any = true; }
    }
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
      if (c.stat[s].done) continue;
      radix_pick_warp(c, s, SHIFT_DIGIT, NB, lane);
      if (PASS == 3) {
        __syncwarp();
        if (lane == 0) {
          GkSegStat* sst = c.stat + s;
          sst->tbits = sst->prefix;  // full 31-bit key of the k-th element
          sst->ties = sst->krem;     // how many keys == T are still needed
          sst->chosen = 0;
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
__device__ uint32_t k2_decide_segment(const GkChain& c, int s, const GkSeg& sg, const uint32_t* tot, bool topk,
                                      int* chosen_out = nullptr) {
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
  if (chosen_out) *chosen_out = chosen;
  return count;
}

// Epilogue shared by the threshold-count pass and the (gt,eq) pass: per-segment totals, the
// reference's refinement loop replayed on the device, then the prefix sums K3 needs.
__device__ void k2_epilogue(const GkChain& c, bool topk) {
  __shared__ uint32_t sscan[2][T];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
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
}


// Counting rounds.  One warp per table entry.  With ~13 MB read at density 0.001 the kernel is pure latency, so
// (a) the dependent memory round trips are kept to two before the rows arrive — the table entry, the entry's
// row maxima and (one segment) the search state in the first, the rows that can hold an element above the
// smallest candidate of the round in the second, up to GK_RB rows in flight through cp.async — and (b) the
// executed code path is kept short (rows are consumed from shared memory in a rolled loop: these kernels run
// once per step with a cold instruction cache, an unrolled body costs more in instruction fetches than in
// issue slots).  The entry's count for every candidate of the round goes to its row of bcnt (plain stores:
// one writer per cell), totals to stat.tot with integer atomics (commutative => deterministic).  Between
// rounds the blocks meet at a grid-wide barrier (cooperative launch) whose last arrival replays the searches
// (k2_advance) and, after the last round, computes the prefix sums K3 needs before it releases the others.
#define GK_RB 8   /* rows in flight per warp */

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  const uint32_t sa = (uint32_t)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}

// flags of up to 64 rows starting at row0: bit i set <=> row row0+i exists and may hold a selected element
__device__ __forceinline__ uint64_t row_flags(uint32_t r0, uint32_t r1, int row0, int nrow, int lane, bool gated, int gate) {
  const bool f0 = (row0 + lane < nrow) && (!gated || (int)r0 > gate);
  const bool f1 = (row0 + 32 + lane < nrow) && (!gated || (int)r1 > gate);
  return (uint64_t)__ballot_sync(0xffffffffu, f0) | ((uint64_t)__ballot_sync(0xffffffffu, f1) << 32);
}

// Bring the next (up to GK_RB) flagged rows into the warp's staging buffer; returns how many.  `m` loses the
// rows taken.  Rows that end inside a float4 are loaded with guarded scalar loads (NaN padding).
__device__ __forceinline__ int stage_rows(float4 (*buf)[32], const float* __restrict__ a, uint64_t& m, int row0,
                                          int len, int lane) {
  int nr = 0;
#pragma unroll 1
  for (; nr < GK_RB && m; ++nr) {
    const int rr = row0 + __ffsll((long long)m) - 1;
    m &= m - 1ull;
    const int i0 = rr * GK_ROW + lane * 4;
    if (i0 + 3 < len) cp_async16(&buf[nr][lane], a + i0);
    else buf[nr][lane] = ld_row(a, rr, lane, len);
  }
  cp_async_wait_all();
  __syncwarp();
  return nr;
}

// Prefix sums of one segment's per-entry counts for candidate column `col` (merged-buffer mode, thresholds):
// the short form of k2_epilogue — one decision, one column, no segment scan.
__device__ void k2_epilogue_one(const GkChain& c) {
  __shared__ uint32_t sscan1[T];
  __shared__ int s_col;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int per = (c.nblk + T - 1) / T;
  const int b0 = tid * per, b1 = min(c.nblk, b0 + per);
  if (tid == 0) {
    const GkSeg sg = c.seg[0];
    uint32_t tot[GK_NC];
#pragma unroll
    for (int j = 0; j < GK_NC; ++j) tot[j] = __ldcg(&c.stat[0].tot[j]);
    int chosen;
    const uint32_t count = k2_decide_segment(c, 0, sg, tot, false, &chosen);
    c.stat[0].out_base = 0;
    c.ctrl->total = count;
    s_col = chosen;
  }
  __syncthreads();
  const int col = s_col;
  // all loads of a thread's run are issued unconditionally on clamped indices: one round trip instead of `per`
  // serialised ones (a bounds check in front of each load stops the compiler from batching them)
  constexpr int KEEP = 16;
  uint32_t vg[KEEP];
  uint32_t loc = 0;
  if (per <= KEEP) {
#pragma unroll
    for (int i = 0; i < KEEP; ++i) vg[i] = __ldcg(c.bcnt + (size_t)min(b0 + i, c.nblk - 1) * GK_NC + col);
#pragma unroll
    for (int i = 0; i < KEEP; ++i) {
      vg[i] = (i < per && b0 + i < b1) ? vg[i] : 0u;
      loc += vg[i];
    }
  } else {
    for (int b = b0; b < b1; ++b) loc += __ldcg(c.bcnt + (size_t)b * GK_NC + col);
  }
  sscan1[tid] = loc;
  __syncthreads();
  if (warp == 0) {
    uint32_t carry = 0;
    for (int base = 0; base < T; base += 32) {
      const uint32_t v = sscan1[base + lane];
      const uint32_t inc = warp_incl_scan_u(v, lane);
      sscan1[base + lane] = carry + inc - v;
      carry += __shfl_sync(0xffffffffu, inc, 31);
    }
  }
  __syncthreads();
  uint32_t run = sscan1[tid];
  if (per <= KEEP) {
#pragma unroll
    for (int i = 0; i < KEEP; ++i) {
      if (i < per && b0 + i < b1) {
        *reinterpret_cast<uint2*>(c.boff + 2 * (size_t)(b0 + i)) = make_uint2(run, 0u);
        run += vg[i];
      }
    }
  } else {
    for (int b = b0; b < b1; ++b) {
      *reinterpret_cast<uint2*>(c.boff + 2 * (size_t)b) = make_uint2(run, 0u);
      run += __ldcg(c.bcnt + (size_t)b * GK_NC + col);
    }
  }
  __syncthreads();
}

template <int NR>
__device__ __forceinline__ void count_staged(const float4 (*buf)[32], int nr, const float (&cs)[GK_RC],
                                             uint32_t (&cnt)[GK_RC], int lane) {
#pragma unroll 1
  for (int u = 0; u < nr; ++u) {
    const float4 v = buf[u][lane];
    const float e0 = fabsf(v.x), e1 = fabsf(v.y), e2 = fabsf(v.z), e3 = fabsf(v.w);
#pragma unroll
    for (int j = 0; j < NR; ++j)
      cnt[j] += (uint32_t)(e0 > cs[j]) + (uint32_t)(e1 > cs[j]) + (uint32_t)(e2 > cs[j]) + (uint32_t)(e3 > cs[j]);
  }
}

__global__ void __launch_bounds__(GK_THREADS, 3) k_count_rounds(GkChain c) {
  __shared__ float4 s_rows[NWARP][GK_RB][32];
  __shared__ uint32_t s_tot[GK_RC], s_ci[GK_RC];
  __shared__ float s_cand[GK_NC];
  __shared__ int s_state[5];
  __shared__ uint32_t s_gen, s_word;
  __shared__ int s_last, s_nrc;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int noct = (c.nblk + NWARP - 1) / NWARP;
  // ctrl->rsv[0] = arrivals, rsv[1] = generation word: advances by 2 per barrier, bit 0 = "another round follows"
  uint32_t* bar = c.ctrl->rsv;
  if (tid == 0) s_gen = ld_relaxed_gpu_u32(bar + 1) & ~1u;
  const bool one_seg = c.nseg == 1;
  const bool dbg = (c.rsv & 4) != 0;               // tools/bench_chain.py --timeline: globaltimer stamps in ctrl->ring
  if (dbg && tid == 0 && blockIdx.x == 0) c.ctrl->ring[44] = (uint32_t)globaltimer_ns();
  for (int round = 0; round < GK_NC + 1; ++round) {
    if (tid < GK_RC) s_tot[tid] = 0;
    if (tid == 0) s_nrc = 0;
    if (one_seg) {
      // one segment: its search state is the same for every warp of the grid — one load per block, not one per
      // warp (thousands of warps asking one L2 sector for the same words serialise there)
      if (tid < GK_NC) s_cand[tid] = __ldcg(&c.stat->cand[tid]);
      if (tid == 32) {
        s_state[0] = __ldcg(&c.stat->done); s_state[1] = __ldcg(&c.stat->nrc); s_state[2] = __ldcg(&c.stat->gate);
        const uint2 w = __ldcg(reinterpret_cast<const uint2*>(c.stat->rcand));
        s_state[3] = (int)w.x; s_state[4] = (int)w.y;
      }
    }
    __syncthreads();
    for (int oct = blockIdx.x; oct < noct; oct += gridDim.x) {
      const int e = oct * NWARP + warp;
      if (e >= c.nblk) continue;                    // warp-uniform
      // round trip 1: the entry, its first 64 row maxima and — when there is only one segment, so that its
      // address is known — the search state; lane j holds candidate j
      const uint32_t* __restrict__ rmx = c.rowmax ? c.rowmax + (size_t)e * c.rpb : nullptr;
      uint32_t r0 = 0, r1 = 0;
      if (rmx) {
        if (lane < c.rpb) r0 = __ldcg(rmx + lane);
        if (32 + lane < c.rpb) r1 = __ldcg(rmx + 32 + lane);
      }
      const GkSegStat* st = c.stat;
      int done = 0, nrc = 0, gate = 0;
      uint2 rcw = make_uint2(0u, 0u);
      float cand_l = 0.f;
      if (one_seg) {
        done = s_state[0]; nrc = s_state[1]; gate = s_state[2];
        rcw = make_uint2((uint32_t)s_state[3], (uint32_t)s_state[4]);         // rcand[0..7]
        cand_l = s_cand[lane & (GK_NC - 1)];
      }
      const GkBlk B = c.blk[e];
      if (!one_seg) {                               // round trip 2 (several segments): this entry's segment
        st = c.stat + B.seg;
        done = __ldcg(&st->done); nrc = __ldcg(&st->nrc); gate = __ldcg(&st->gate);
        rcw = __ldcg(reinterpret_cast<const uint2*>(st->rcand));
        cand_l = __ldcg(&st->cand[lane & (GK_NC - 1)]);
      }
      if (done) continue;                           // this segment's search is complete (warp-uniform)
      float cs[GK_RC];
      uint32_t ci[GK_RC], cnt[GK_RC];
#pragma unroll
      for (int j = 0; j < GK_RC; ++j) {
        ci[j] = ((j < 4 ? rcw.x >> (8 * j) : rcw.y >> (8 * (j - 4))) & 0xffu) & (GK_NC - 1);
        const float cv = __shfl_sync(0xffffffffu, cand_l, (int)ci[j]);
        cs[j] = (j < nrc) ? cv : __int_as_float(0x7f800000);   // +inf: never exceeded
        cnt[j] = 0;
      }
      const float* __restrict__ a = c.resid + B.start;
      const int nrow = (B.len + GK_ROW - 1) / GK_ROW;
      // last round trip: the rows
      for (int row0 = 0; row0 < nrow; row0 += 64) {
        if (row0 && rmx) {                           // chunks above 8192 elements: further groups of 64 rows
          r0 = (row0 + lane < nrow) ? __ldcg(rmx + row0 + lane) : 0u;
          r1 = (row0 + 32 + lane < nrow) ? __ldcg(rmx + row0 + 32 + lane) : 0u;
        }
        uint64_t m = row_flags(r0, r1, row0, nrow, lane, rmx != nullptr, gate);
        while (m) {                                  // warp-uniform: m comes from ballots
          const int nr = stage_rows(s_rows[warp], a, m, row0, B.len, lane);
          if (nrc <= 1) count_staged<1>(s_rows[warp], nr, cs, cnt, lane);
          else if (nrc <= 3) count_staged<3>(s_rows[warp], nr, cs, cnt, lane);
          else count_staged<GK_RC>(s_rows[warp], nr, cs, cnt, lane);
          __syncwarp();                              // the buffer is refilled by the next batch
        }
      }
      uint32_t* brow = c.bcnt + (size_t)e * GK_NC;
#pragma unroll
      for (int j = 0; j < GK_RC; ++j) {
        if (j < nrc) {                              // warp-uniform
          const uint32_t w = warp_sum_u(cnt[j]);
          if (lane == 0) {
            brow[ci[j]] = w;
            if (one_seg) { if (w) atomicAdd(&s_tot[j], w); s_ci[j] = ci[j]; }
            else if (w) atomicAdd(&c.stat[B.seg].tot[ci[j]], w);
          }
        }
      }
      if (one_seg && lane == 0) s_nrc = nrc;
    }
    __syncthreads();
    if (one_seg && tid < s_nrc && s_tot[tid]) atomicAdd(&c.stat[0].tot[s_ci[tid]], s_tot[tid]);
    // ---- grid barrier; the last block to arrive advances every segment's search before releasing it
    __syncthreads();
    if (tid == 0) {
      __threadfence();
      if (dbg) atomicMax(&c.ctrl->ring[45], (uint32_t)globaltimer_ns());
      s_last = (atomicAdd(bar, 1u) == gridDim.x - 1u);
    }
    __syncthreads();
    const uint32_t gen_now = s_gen + 2u * (uint32_t)round;
    if (s_last) {
      __threadfence();
      if (dbg && tid == 0) c.ctrl->ring[46] = (uint32_t)globaltimer_ns();
      int more = 0;
      for (int s = tid; s < c.nseg; s += T) {
        GkSegStat* st = c.stat + s;
        if (st->done) continue;
        uint32_t known = st->known;
        for (int j = 0; j < st->nrc; ++j) known |= 1u << st->rcand[j];
        st->known = known;
        if (k2_advance(c, c.seg[s], st)) st->done = 1; else more = 1;
      }
      more = __syncthreads_or(more);
      if (dbg && tid == 0) c.ctrl->ring[47] = (uint32_t)globaltimer_ns();
      if (!more) {                                  // search complete everywhere: prefix sums for K3, then release
        __threadfence();
        if (one_seg) k2_epilogue_one(c); else k2_epilogue(c, false);
        __syncthreads();
      }
      if (dbg && tid == 0) c.ctrl->ring[48] = (uint32_t)globaltimer_ns();
      if (tid == 0) {
        bar[0] = 0;
        s_word = (gen_now + 2u) | (more ? 1u : 0u);
        __threadfence();
        st_release_gpu_u32(bar + 1, s_word);
      }
    } else if (tid == 0) {
      const uint64_t t0 = globaltimer_ns();
      uint32_t w;
      while (((w = ld_relaxed_gpu_u32(bar + 1)) & ~1u) == gen_now) {
        if (globaltimer_ns() - t0 > 2000000000ull) { atomicExch(&c.ctrl->err, 3u); w = gen_now + 2u; break; }
        __nanosleep(200);                           // hundreds of pollers on one L2 line: back off
      }
      __threadfence();
      s_word = w;
    }
    __syncthreads();
    if (!(s_word & 1u)) break;
  }
  if (dbg && tid == 0) atomicMax(&c.ctrl->ring[49], (uint32_t)globaltimer_ns());
}

// top-k / DGC: per-entry counts of key > T (column 0) and key == T (column 1), one warp per entry, rows below
// T skipped.  FIRST (DGC only) is the pass with the sampled threshold: its epilogue accepts the threshold when
// at most 4k/3 elements pass it (compression.py:598), otherwise the exact radix select runs.
template <bool MINK, bool FIRST>
__global__ void __launch_bounds__(GK_THREADS, 3) k_count_topk(GkChain c) {
  __shared__ float4 s_rows[NWARP][GK_RB][32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int e = (int)blockIdx.x * NWARP + warp;
  // round trip 1: entry + row maxima (+ the threshold key when there is one segment: one load per block);
  // last: the rows
  const bool one_seg = c.nseg == 1, active = e < c.nblk;
  __shared__ uint32_t s_tb;
  const uint32_t* __restrict__ rmx = (active && c.rowmax && !MINK) ? c.rowmax + (size_t)e * c.rpb : nullptr;
  uint32_t r0 = 0, r1 = 0;
  GkBlk B = {0, 0, 0, 0};
  if (active) {
    if (rmx) {
      if (lane < c.rpb) r0 = __ldcg(rmx + lane);
      if (32 + lane < c.rpb) r1 = __ldcg(rmx + 32 + lane);
    }
    B = c.blk[e];
  }
  if (one_seg) {
    if (tid == 0) s_tb = c.stat[0].tbits;
    __syncthreads();
  }
  if (active) {
    GkSegStat* st = c.stat + B.seg;
    const uint32_t tb = one_seg ? s_tb : st->tbits;
    uint32_t gt = 0, eq = 0;
    const float* __restrict__ a = c.resid + B.start;
    const int nrow = (B.len + GK_ROW - 1) / GK_ROW;
    for (int row0 = 0; row0 < nrow; row0 += 64) {
      if (row0 && rmx) {
        r0 = (row0 + lane < nrow) ? __ldcg(rmx + row0 + lane) : 0u;
        r1 = (row0 + 32 + lane < nrow) ? __ldcg(rmx + row0 + 32 + lane) : 0u;
      }
      // rows whose maximum key is >= tb: as a signed gate that is "> tb - 1" (tb <= 0x7fffffff)
      uint64_t m = row_flags(r0, r1, row0, nrow, lane, rmx != nullptr, (int)tb - 1);
      while (m) {
        const uint64_t mc = m;
        const int nr = stage_rows(s_rows[warp], a, m, row0, B.len, lane);
        uint64_t walk = mc;
#pragma unroll 1
        for (int u = 0; u < nr; ++u) {
          const int rr = row0 + __ffsll((long long)walk) - 1;
          walk &= walk - 1ull;
          const float4 v = s_rows[warp][u][lane];
          const float ev[4] = {v.x, v.y, v.z, v.w};
          const int nv = B.len - (rr * GK_ROW + lane * 4);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t key = sel_key<MINK>(ev[j]);
            gt += (uint32_t)(j < nv && key > tb);
            eq += (uint32_t)(j < nv && key == tb);
          }
        }
        __syncwarp();
      }
    }
    gt = warp_sum_u(gt); eq = warp_sum_u(eq);
    uint32_t* brow = c.bcnt + (size_t)e * GK_NC;
    if (lane == 0) {
      brow[0] = gt; brow[1] = eq;
      if (gt) atomicAdd(&st->tot[0], gt);
      if (eq) atomicAdd(&st->tot[1], eq);
    }
  }
  if (last_block_done(&c.ctrl->done[2])) {
    if (FIRST) {
      for (int s = tid; s < c.nseg; s += T) {
        GkSegStat* st = c.stat + s;
        const double k = (double)c.seg[s].k;
        const uint32_t passed = __ldcg(&st->tot[0]);
        st->tot[0] = 0; st->tot[1] = 0;             // the final (gt, eq) pass counts again
        if ((double)passed > 4.0 * k / 3.0) {       // too many: exact top-k of the whole segment instead
          st->done = 0;
        } else {
          st->done = 1;                             // keep the sampled threshold: every element above it
          st->ties = 0;
        }
      }
    } else {
      k2_epilogue(c, true);
    }
  }
}

// ------------------------------------------------------- K3 select + compact + residual + send
// Rows are skipped via the row maxima, the flagged ones staged through cp.async and consumed in rolled loops;
// the entry's first output position comes from K2's prefix sums and the elements' ranks from ballots, so the
// pairs leave in ascending index order.  Work is split two ways, because both shapes occur:
//   * the usual entry has a handful of flagged rows: ONE warp walks them in order (no block-level exchange;
//     all ~3 000 entries of ResNet-50 are in flight at once, the kernel is a few memory round trips long);
//   * in training the selected coordinates cluster in a few tensors whose entries have every row flagged: a
//     single warp walking 64 rows was the long tail of the first gated version (69 us in the step against 23
//     on uniformly spread data).  Entries with more than GK_K3_SOLO flagged rows are therefore left to the
//     whole block afterwards: warp w takes rows 8w .. 8w+7 of every 64-row group and one shared-memory
//     exchange of the per-warp counts gives the warps their bases.
// Entries are dealt to the warps of a block with a stride of the grid size, so that the entries of one tensor
// land in different blocks.
#define GK_K3_SOLO 16

struct K3Dec { float thr; uint32_t tb, ties; };

// which elements of the staged rows are selected: 4 bits per row in fg (above the threshold / key) and fe (ties)
template <bool TOPK, bool MINK>
__device__ __forceinline__ void k3_classify(const float4 (*buf)[32], int nr, uint64_t walk, int row0, int len, int lane,
                                            const K3Dec& dec, uint32_t& fg, uint32_t& fe) {
  fg = 0; fe = 0;
#pragma unroll 1
  for (int u = 0; u < nr; ++u) {
    const int rr = row0 + __ffsll((long long)walk) - 1;
    walk &= walk - 1ull;
    const float4 v = buf[u][lane];
    const int nv = len - (rr * GK_ROW + lane * 4);
    const float ev[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (TOPK) {
        const uint32_t key = sel_key<MINK>(ev[j]);
        if (j < nv && key > dec.tb) fg |= 1u << (4 * u + j);
        if (j < nv && key == dec.tb) fe |= 1u << (4 * u + j);
      } else {
        if (fabsf(ev[j]) > dec.thr) fg |= 1u << (4 * u + j);  // padding reads as NaN: never selected
      }
    }
  }
}

// store the selected elements of the staged rows; gb / eb = selected before the first of these rows
template <bool TOPK>
__device__ __forceinline__ void k3_emit(const float4 (*buf)[32], int nr, uint64_t walk, int row0, int lane, uint32_t fg,
                                        uint32_t fe, uint32_t& gb, uint32_t& eb, uint32_t base_gt, uint32_t ties,
                                        uint32_t cap, const GkDst& d, uint64_t poff, float* __restrict__ a, int start) {
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll 1
  for (int u = 0; u < nr; ++u) {
    const int rr = row0 + __ffsll((long long)walk) - 1;
    walk &= walk - 1ull;
    const uint32_t fgi = (fg >> (4 * u)) & 15u, fei = (fe >> (4 * u)) & 15u;
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
      const float4 v = buf[u][lane];
      const float ev[4] = {v.x, v.y, v.z, v.w};
      const int i0 = rr * GK_ROW + lane * 4;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const bool isg = (fgi >> j) & 1u, ise = (fei >> j) & 1u;
        if (isg || (ise && ebefore < ties)) {
          const uint32_t pos = base_gt + gbefore + (ebefore < ties ? ebefore : ties);
          if (pos < cap) {
            const uint32_t idx = (uint32_t)(start + i0 + j);
            const uint64_t pair = (uint64_t)idx | ((uint64_t)__float_as_uint(ev[j]) << 32);
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

template <bool TOPK, bool MINK>
__global__ void __launch_bounds__(GK_THREADS, 3) k_compact_send(GkChain c, GkDst d) {
  __shared__ float4 s_rows[NWARP][GK_RB][32];
  __shared__ uint32_t s_wcnt[2][NWARP];
  __shared__ uint32_t s_dec[6];
  __shared__ int s_dense[NWARP];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t seqn = c.ctrl->seq + 1u;  // every block reads it before the last block bumps it
  const uint64_t poff = (uint64_t)(seqn & 1u) * d.parity_stride;
  const uint32_t cap = c.cap;
  const bool one_seg = c.nseg == 1;
  const bool use_rmx = c.rowmax != nullptr && !(TOPK && MINK);
  const bool dbg = (c.rsv & 4) != 0;
  if (dbg && tid == 0 && blockIdx.x == 0) c.ctrl->ring[50] = (uint32_t)globaltimer_ns();
  // entry of warp w: w * gridDim.x + blockIdx.x — issue its loads before the block-level state is exchanged
  const int e = warp * (int)gridDim.x + (int)blockIdx.x;
  const bool active = e < c.nblk;
  uint32_t r0 = 0, r1 = 0;
  uint2 off_b = make_uint2(0u, 0u);
  GkBlk B = {0, 0, 0, 0};
  if (active) {
    if (use_rmx) {
      const uint32_t* q = c.rowmax + (size_t)e * c.rpb;
      if (lane < c.rpb) r0 = __ldcg(q + lane);
      if (32 + lane < c.rpb) r1 = __ldcg(q + 32 + lane);
    }
    off_b = __ldcg(reinterpret_cast<const uint2*>(c.boff + 2 * (size_t)e));
    B = c.blk[e];
  }
  if (one_seg) {                           // the single segment's decision: one load per block
    if (tid == 0) {
      const GkSegStat* st0 = c.stat;
      s_dec[0] = __float_as_uint(st0->thr); s_dec[1] = st0->tbits; s_dec[2] = st0->ties; s_dec[3] = st0->out_base;
      const uint2 o = __ldcg(reinterpret_cast<const uint2*>(c.boff));
      s_dec[4] = o.x; s_dec[5] = o.y;
    }
    __syncthreads();
  }
  // decision of the segment an entry belongs to
  auto decision = [&](const GkBlk& Bx, K3Dec& dec, uint32_t& out_base, uint2& off_s) {
    if (one_seg) {
      dec.thr = __uint_as_float(s_dec[0]); dec.tb = s_dec[1]; dec.ties = s_dec[2]; out_base = s_dec[3];
      off_s = make_uint2(s_dec[4], s_dec[5]);
    } else {
      const GkSegStat* st = c.stat + Bx.seg;
      dec.thr = st->thr; dec.tb = st->tbits; dec.ties = st->ties; out_base = st->out_base;
      off_s = __ldcg(reinterpret_cast<const uint2*>(c.boff + 2 * (size_t)c.seg[Bx.seg].blk_begin));
    }
  };
  // ---- solo pass: every warp its own entry, unless the entry has too many flagged rows
  bool dense = false;
  if (active) {
    K3Dec dec; uint32_t out_base; uint2 off_s;
    decision(B, dec, out_base, off_s);
    // rows that can hold a selected element: max > thr, or max key >= tb (> tb when no tie is taken)
    const int gate = TOPK ? (dec.ties ? (int)dec.tb - 1 : (int)dec.tb) : gate_of(dec.thr);
    const int nrow = (B.len + GK_ROW - 1) / GK_ROW;
    uint64_t m = row_flags(r0, r1, 0, nrow, lane, use_rmx, gate);
    dense = nrow > 64 || __popcll(m) > GK_K3_SOLO;
    if (!dense) {
      const uint32_t base_gt = out_base + (off_b.x - off_s.x);
      uint32_t gb = 0, eb = TOPK ? (off_b.y - off_s.y) : 0u;
      float* __restrict__ a = c.resid + B.start;
      while (m) {                                            // at most GK_K3_SOLO / GK_RB batches
        const uint64_t walk = m;
        const int nr = stage_rows(s_rows[warp], a, m, 0, B.len, lane);
        uint32_t fg, fe;
        k3_classify<TOPK, MINK>(s_rows[warp], nr, walk, 0, B.len, lane, dec, fg, fe);
        k3_emit<TOPK>(s_rows[warp], nr, walk, 0, lane, fg, fe, gb, eb, base_gt, dec.ties, cap, d, poff, a, B.start);
        __syncwarp();                                        // the staging buffer is refilled next
      }
    }
  }
  if (lane == 0) s_dense[warp] = dense ? 1 : 0;
  __syncthreads();
  // ---- cooperative pass: the whole block on every entry that was left over (block-uniform loop)
  int par = 0;
  for (int w2 = 0; w2 < NWARP; ++w2) {
    if (!s_dense[w2]) continue;
    const int e2 = w2 * (int)gridDim.x + (int)blockIdx.x;
    const GkBlk B2 = c.blk[e2];
    const uint2 ob = __ldcg(reinterpret_cast<const uint2*>(c.boff + 2 * (size_t)e2));
    K3Dec dec; uint32_t out_base; uint2 off_s;
    decision(B2, dec, out_base, off_s);
    const int gate = TOPK ? (dec.ties ? (int)dec.tb - 1 : (int)dec.tb) : gate_of(dec.thr);
    const uint32_t base_gt = out_base + (ob.x - off_s.x);
    const uint32_t base_eq = TOPK ? (ob.y - off_s.y) : 0u;
    float* __restrict__ a = c.resid + B2.start;
    const uint32_t* __restrict__ rmx = use_rmx ? c.rowmax + (size_t)e2 * c.rpb : nullptr;
    const int nrow = (B2.len + GK_ROW - 1) / GK_ROW;
    uint32_t run_gt = 0, run_eq = 0;       // selected in earlier 64-row groups of this entry
    for (int row0 = 0; row0 < nrow; row0 += 64) {
      uint32_t q0 = 0, q1 = 0;
      if (rmx) {
        q0 = (row0 + lane < nrow) ? __ldcg(rmx + row0 + lane) : 0u;
        q1 = (row0 + 32 + lane < nrow) ? __ldcg(rmx + row0 + 32 + lane) : 0u;
      }
      const uint64_t mall = row_flags(q0, q1, row0, nrow, lane, rmx != nullptr, gate);
      if (mall == 0ull) continue;                            // block-uniform: every warp sees all 64 flags
      uint64_t m = mall & (0xffull << (8 * warp));           // this warp's eight rows
      const uint64_t walk = m;
      const int nr = stage_rows(s_rows[warp], a, m, row0, B2.len, lane);
      uint32_t fg, fe;
      k3_classify<TOPK, MINK>(s_rows[warp], nr, walk, row0, B2.len, lane, dec, fg, fe);
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
      par ^= 1;
      if (wsum != 0) {                                       // warp-uniform
        uint32_t gb = run_gt + (wbase & 0xffffu);            // gt elements of the entry before this warp's rows
        uint32_t eb = base_eq + run_eq + (wbase >> 16);      // eq elements of the segment before them
        k3_emit<TOPK>(s_rows[warp], nr, walk, row0, lane, fg, fe, gb, eb, base_gt, dec.ties, cap, d, poff, a, B2.start);
      }
      __syncwarp();                                          // the staging buffer is refilled next
      run_gt += total & 0xffffu;
      run_eq += total >> 16;
    }
  }
  if (dbg && lane == 0) atomicMax(&c.ctrl->ring[51], (uint32_t)globaltimer_ns());
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
      if (dbg) c.ctrl->ring[52] = (uint32_t)globaltimer_ns();
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
        const float v = __fmul_rn(__uint_as_float((uint32_t)(p >> 32)), scale);   // never fused with the add
        if (UNSCATTER) grad[idx] = 0.f;
        else if (ORDERED) grad[idx] = __fadd_rn(grad[idx], v);
        else atomicAdd(grad + idx, v);
      }
    }
  }
}

// ------------------------------------------------------------------------- K6 fused SGD
// torch.optim.SGD arithmetic for one element (FMA placement included); shared by the materialising kernel
// (gradient read from the arena) and the folded one (gradient == 0 in the dense pass).
__device__ __forceinline__ float sgd_upd(const GkHyper& h, bool has_m, float gs, float pv, float gv, float& mv) {
  if (gs != 1.f) gv = __fmul_rn(gv, gs);
  if (h.weight_decay != 0.f) gv = __fmaf_rn(h.weight_decay, pv, gv);
  float u = gv;
  if (has_m) {
    mv = h.first_step ? gv : __fmaf_rn(1.f - h.dampening, gv, __fmul_rn(h.momentum, mv));
    u = h.nesterov ? __fmaf_rn(h.momentum, mv, gv) : mv;
  }
  return __fmaf_rn(-h.lr, u, pv);
}

__global__ void __launch_bounds__(GK_THREADS) k_sgd(float* __restrict__ p, const float* __restrict__ g,
                                                     float* __restrict__ m, const GkBlk* __restrict__ blk,
                                                     const GkHyper* __restrict__ hy,
                                                     const float* __restrict__ gscale) {
  const GkBlk B = blk[blockIdx.x];
  const GkHyper h = hy[B.seg];
  const float gs = gscale ? *gscale : 1.f;
  const bool has_m = (h.momentum != 0.f) && (m != nullptr);
  auto upd = [&](float pv, float gv, float& mv) -> float { return sgd_upd(h, has_m, gs, pv, gv, mv); };
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

// ------------------------------------------------- K6 folded: receive + momentum-SGD, no dense gradient
// The reference zeroes the group buffer, scatter-adds P lists, divides, and then runs SGD over dense
// tensors (distributed_optimizer.py:468-482, :543-546).  Only <= P*k coordinates of that gradient are
// non-zero, so the update splits into
//   phase A  the dense part with g == 0:  m = mu*m + (1-d)*wd*p ;  p -= lr*m     (16 B/element, and —
//            the point — independent of the peers: it runs while their pairs are still in flight);
//   wait     bounded acquire-spin on the P slot headers;
//   pass 1   every pair ORs its source rank into mask[idx] (the gradient arena, all zero by invariant);
//   barrier  grid-wide (cooperative launch);
//   pass 2   the pair of the LOWEST rank that holds a coordinate owns it: it sums the contributions in
//            rank order (binary search in the other ranks' index-sorted lists), applies the correction
//            m += (1-d)*g ; p -= lr*(1-d)*g (Nesterov / first step / no momentum accordingly) and clears
//            mask[idx] — so the arena is all zero again and zero_grad() has nothing to do.
// Every replica forms the same sums in the same order => bit-identical parameters on all ranks.  The
// result differs from the materialising path (k_scatter + k_sgd) only by the rounding of the split FMA
// on the <= P*k touched coordinates.
struct GkRecvUpdDev {
  float* p; float* m; uint32_t* mask; const GkBlk* blk; const GkHyper* hy; const uint8_t* gmap;
  GkCtrl* ctrl; uint32_t* bar; float scale; int nblk; int nelem; int skip_dense;
};

// Grid-wide barrier for a cooperative launch: bar[0] = arrival count, bar[1] = generation.  `gen` is the
// generation read by thread 0 before the block's first arrival.  Bounded like every wait of this library.
__device__ __forceinline__ void gk_grid_barrier(uint32_t* bar, uint32_t gen, uint32_t* err) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(bar, 1u) == gridDim.x - 1u) {
      bar[0] = 0;
      st_release_gpu_u32(bar + 1, gen + 1u);
    } else {
      const uint64_t t0 = globaltimer_ns();
      while (ld_relaxed_gpu_u32(bar + 1) == gen) {
        if (globaltimer_ns() - t0 > 2000000000ull) { atomicExch(err, 3u); break; }
        __nanosleep(400);
      }
    }
    __threadfence();
  }
  __syncthreads();
}

__global__ void __launch_bounds__(GK_THREADS) k_sgd_recv(GkRecvUpdDev u, GkSrc s) {
  __shared__ uint32_t scnt[GK_MAXP];
  __shared__ uint32_t s_gen;
  __shared__ int s_bad;
  const int tid = threadIdx.x;
  if (tid == 0) { s_gen = ld_relaxed_gpu_u32(u.bar + 1); s_bad = 0; }
  const bool dbg = (u.skip_dense & 4) != 0;        // tools/bench_chain.py --timeline
  if (dbg && tid == 0 && blockIdx.x == 0) u.ctrl->ring[53] = (uint32_t)globaltimer_ns();
  // ---- phase A: dense update with a zero gradient (same arithmetic as k_sgd given g == 0)
  if (!(u.skip_dense & 1)) {
    for (int e = blockIdx.x; e < u.nblk; e += gridDim.x) {
      const GkBlk B = u.blk[e];
      const GkHyper h = u.hy[B.seg];
      const bool has_m = (h.momentum != 0.f) && (u.m != nullptr);
      if (!has_m && h.weight_decay == 0.f) continue;     // p -= lr*0: nothing to do for this entry
      float* pp = u.p + B.start; float* mm = has_m ? u.m + B.start : nullptr;
      const int n4 = B.len >> 2;
#pragma unroll 4
      for (int i = tid; i < n4; i += T) {
        float4 pv = ld4(pp + 4 * i);
        float4 mv = has_m ? ld4(mm + 4 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
        pv.x = sgd_upd(h, has_m, 1.f, pv.x, 0.f, mv.x); pv.y = sgd_upd(h, has_m, 1.f, pv.y, 0.f, mv.y);
        pv.z = sgd_upd(h, has_m, 1.f, pv.z, 0.f, mv.z); pv.w = sgd_upd(h, has_m, 1.f, pv.w, 0.f, mv.w);
        st4(pp + 4 * i, pv);
        if (has_m) st4(mm + 4 * i, mv);
      }
      for (int i = 4 * n4 + tid; i < B.len; i += T) {
        float mv = has_m ? mm[i] : 0.f;
        pp[i] = sgd_upd(h, has_m, 1.f, pp[i], 0.f, mv);
        if (has_m) mm[i] = mv;
      }
    }
  }
  if (dbg && tid == 0) atomicMax(&u.ctrl->ring[54], (uint32_t)globaltimer_ns());
  // ---- wait for the peers' headers (their pairs were stored before the release of the header)
  const uint32_t seq = u.ctrl->seq;
  const uint64_t* base = s.arena + (uint64_t)(seq & 1u) * s.parity_stride;
  const uint32_t cap = (uint32_t)(s.slot_stride - GK_SLOT_HDR);
  __syncthreads();
  if (tid < s.nsrc) {
    const uint64_t* hp = base + (uint64_t)tid * s.slot_stride;
    uint64_t v = ld_acquire_sys_u64(hp);
    if (s.wait) {
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
  const bool bad = s_bad != 0;
  if (bad && tid == 0) atomicExch(&u.ctrl->err, 1u);
  const uint32_t gstride = gridDim.x * T, g0 = blockIdx.x * T + tid;
  // ---- pass 1: which ranks hold which coordinate
  if (!bad) {
    for (int r = 0; r < s.nsrc; ++r) {
      const uint64_t* pairs = base + (uint64_t)r * s.slot_stride + GK_SLOT_HDR;
      const uint32_t n = scnt[r];
      for (uint32_t j = g0; j < n; j += gstride) {
        const uint32_t idx = (uint32_t)__ldcg(pairs + j);
        if (idx < (uint32_t)u.nelem) atomicOr(u.mask + idx, 1u << r);
      }
    }
  }
  if (dbg && tid == 0) atomicMax(&u.ctrl->ring[55], (uint32_t)globaltimer_ns());
  gk_grid_barrier(u.bar, s_gen, &u.ctrl->err);
  if (dbg && tid == 0) atomicMax(&u.ctrl->ring[56], (uint32_t)globaltimer_ns());
  if (bad) return;
  // ---- pass 2: the lowest rank's pair owns the coordinate
  for (int r = 0; r < s.nsrc; ++r) {
    const uint64_t* pairs = base + (uint64_t)r * s.slot_stride + GK_SLOT_HDR;
    const uint32_t n = scnt[r];
    for (uint32_t j = g0; j < n; j += gstride) {
      const uint64_t pr = __ldcg(pairs + j);
      const uint32_t idx = (uint32_t)pr;
      if (idx >= (uint32_t)u.nelem) continue;
      // everything this coordinate may need is requested at once (p and m were written in phase A, possibly
      // by another block: read them at L2); most coordinates have one owner and use all of it
      uint32_t w = __ldcg(u.mask + idx);
      const float p_old = __ldcg(u.p + idx);
      const float m_old = u.m ? __ldcg(u.m + idx) : 0.f;
      const int gi = u.gmap ? (int)__ldg(u.gmap + (idx >> 6)) : 0;
      if (w == 0u || (int)(__ffs(w) - 1) != r) continue;
      float g = __fadd_rn(0.f, __fmul_rn(__uint_as_float((uint32_t)(pr >> 32)), u.scale));
      w &= w - 1u;
      while (w) {                                  // contributions of higher ranks, ascending
        const int q = __ffs(w) - 1;
        w &= w - 1u;
        const uint64_t* pq = base + (uint64_t)q * s.slot_stride + GK_SLOT_HDR;
        uint32_t lo = 0, hi = scnt[q];
        while (lo < hi) {
          const uint32_t mid = (lo + hi) >> 1;
          if ((uint32_t)__ldcg(pq + mid) < idx) lo = mid + 1; else hi = mid;
        }
        const uint64_t hit = (lo < scnt[q]) ? __ldcg(pq + lo) : ~0ull;
        if ((uint32_t)hit != idx) { atomicExch(&u.ctrl->err, 2u); continue; }   // list not index-sorted
        g = __fadd_rn(g, __fmul_rn(__uint_as_float((uint32_t)(hit >> 32)), u.scale));
      }
      const GkHyper h = u.hy[gi];
      const bool has_m = (h.momentum != 0.f) && (u.m != nullptr);
      float du = g;
      if (has_m) {
        const float dm = h.first_step ? g : __fmul_rn(1.f - h.dampening, g);
        u.m[idx] = __fadd_rn(m_old, dm);
        du = h.nesterov ? __fmaf_rn(h.momentum, dm, g) : dm;
      }
      u.p[idx] = __fmaf_rn(-h.lr, du, p_old);
      u.mask[idx] = 0u;
    }
  }
  if (dbg && tid == 0) atomicMax(&u.ctrl->ring[57], (uint32_t)globaltimer_ns());
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

// ------------------------------------------------------------------- K9 DGC sampling threshold
// compression.py:576-606: draw int(0.01*n) coordinates, take the k-th largest |acc| among them as the
// threshold, select everything above it; if more than 4k/3 pass, fall back to the exact top-k.  The sample
// is drawn with the keyed bijection of the random-k kernel (distinct coordinates, no n-element permutation);
// its k-th largest is found with the same 11+11+9-bit radix select as the exact top-k, on the sample only.
// With fewer than k sampled coordinates (density > 0.01, where the reference's torch.topk raises) the
// threshold is 0 and the exact top-k takes over.
__device__ __forceinline__ int dgc_sample_size(int n) { return (int)((double)n * 0.01); }

template <int PASS>
__global__ void __launch_bounds__(GK_THREADS) k_dgc_pass(GkChain c, uint64_t seed) {
  constexpr int NB = (PASS == 3) ? 512 : 2048;
  constexpr int SHIFT_MATCH = (PASS == 2) ? 20 : 9;        // PASS 1 matches everything
  constexpr int SHIFT_DIGIT = (PASS == 1) ? 20 : (PASS == 2 ? 9 : 0);
  __shared__ uint32_t shist[NB];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t seqn = c.ctrl->seq + 1u;
  for (int s = 0; s < c.nseg; ++s) {                       // block-uniform
    const GkSeg sg = c.seg[s];
    const int fk = dgc_sample_size(sg.n);
    if (fk < sg.k || fk < 1) continue;
    for (int i = tid; i < NB; i += T) shist[i] = 0;
    __syncthreads();
    const uint32_t prefix = (PASS == 1) ? 0u : c.stat[s].sprefix;
    const uint64_t sd = seed + 0x632be59bd9b4e019ull * (uint64_t)(s + 1) + 0x9e3779b97f4a7c15ull * (uint64_t)seqn;
    for (uint32_t j = blockIdx.x * T + tid; j < (uint32_t)fk; j += gridDim.x * T) {
      const uint32_t li = feistel_perm(j, (uint32_t)sg.n, sd);
      int lo = sg.blk_begin, hi = sg.blk_end - 1;          // segment-local index -> arena index
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if ((uint32_t)c.blk[mid].soff <= li) lo = mid; else hi = mid - 1;
      }
      const GkBlk B = c.blk[lo];
      const uint32_t key = absbits(c.resid[(uint32_t)B.start + (li - (uint32_t)B.soff)]);
      if (PASS == 1 || (key >> SHIFT_MATCH) == prefix) atomicAdd(&shist[(key >> SHIFT_DIGIT) & (NB - 1)], 1u);
    }
    __syncthreads();
    uint32_t* h = c.hist + (size_t)s * GK_HIST_BINS;
    for (int i = tid; i < NB; i += T) {
      const uint32_t v = shist[i];
      if (v) atomicAdd(h + i, v);
    }
    __syncthreads();
  }
  if (last_block_done(&c.ctrl->done[1])) {
    for (int s = warp; s < c.nseg; s += NWARP) {
      GkSegStat* st = c.stat + s;
      const GkSeg sg = c.seg[s];
      const int fk = dgc_sample_size(sg.n);
      if (fk < sg.k || fk < 1) {
        if (lane == 0) { st->tbits = 0; st->ties = 0; }
        continue;
      }
      if (PASS == 1 && lane == 0) { st->sprefix = 0; st->skrem = (uint32_t)sg.k; }
      __syncwarp();
      radix_pick_warp_at(c.hist + (size_t)s * GK_HIST_BINS, &st->sprefix, &st->skrem, NB, lane);
      __syncwarp();
      if (PASS == 3 && lane == 0) { st->tbits = st->sprefix; st->ties = 0; }
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

static int gk_sm_count() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0, n = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    sms = n > 0 ? n : 148;
  }
  return sms;
}

extern "C" {

const char* gk_version(void) { return "gk_b200 cuda sm_100a 0.1"; }

int gk_accumulate(const GkChain* c, void* stream) {
  if (c->nblk <= 0) return 0;
  if (c->rowmax != nullptr && (c->rpb <= 0 || c->rpb > 4096)) return (int)cudaErrorInvalidValue;
  if (c->mode == GK_MODE_TOPK || c->mode == GK_MODE_DGC) GK_LAUNCH_PERSISTENT((k_accumulate<true, false>), c->nblk, stream, *c);
  else if (c->mode == GK_MODE_MINK) GK_LAUNCH_PERSISTENT((k_accumulate<true, true>), c->nblk, stream, *c);
  else GK_LAUNCH_PERSISTENT((k_accumulate<false, false>), c->nblk, stream, *c);
  GK_RET();
}

int gk_count_pick(const GkChain* c, void* stream) {
  if (c->nblk <= 0) return 0;
  // the blocks meet at a grid barrier between counting rounds: one resident wave, cooperative launch
  static int cache = 0;
  const int noct = (c->nblk + NWARP - 1) / NWARP;
  const int grid = gk_persistent_grid(k_count_rounds, &cache, noct);
  GkChain cc = *c;
  void* args[] = {&cc};
  const cudaError_t e = cudaLaunchCooperativeKernel((const void*)k_count_rounds, dim3(grid), dim3(T), args, 0, S(stream));
  if (e != cudaSuccess) (void)cudaGetLastError();
  return (int)e;
}

int gk_topk_select(const GkChain* c, void* stream) {
  if (c->nblk <= 0) return 0;
  const int noct = (c->nblk + NWARP - 1) / NWARP;
  if (c->mode == GK_MODE_MINK) {
    k_radix_pass<2, true><<<c->nblk, T, 0, S(stream)>>>(*c);
    k_radix_pass<3, true><<<c->nblk, T, 0, S(stream)>>>(*c);
    k_count_topk<true, false><<<noct, T, 0, S(stream)>>>(*c);
  } else {
    k_radix_pass<2, false><<<c->nblk, T, 0, S(stream)>>>(*c);
    k_radix_pass<3, false><<<c->nblk, T, 0, S(stream)>>>(*c);
    k_count_topk<false, false><<<noct, T, 0, S(stream)>>>(*c);
  }
  GK_RET();
}

int gk_dgc_select(const GkChain* c, uint64_t seed, void* stream) {
  if (c->nblk <= 0) return 0;
  if (c->mode != GK_MODE_DGC) return (int)cudaErrorInvalidValue;
  const int noct = (c->nblk + NWARP - 1) / NWARP;
  const int grid = 4 * gk_sm_count();
  k_dgc_pass<1><<<grid, T, 0, S(stream)>>>(*c, seed);
  k_dgc_pass<2><<<grid, T, 0, S(stream)>>>(*c, seed);
  k_dgc_pass<3><<<grid, T, 0, S(stream)>>>(*c, seed);
  k_count_topk<false, true><<<noct, T, 0, S(stream)>>>(*c);      // accept the sampled threshold or not
  k_radix_pass<2, false><<<c->nblk, T, 0, S(stream)>>>(*c);        // no-ops for the segments that accepted it
  k_radix_pass<3, false><<<c->nblk, T, 0, S(stream)>>>(*c);
  k_count_topk<false, false><<<noct, T, 0, S(stream)>>>(*c);
  GK_RET();
}

int gk_compact_send(const GkChain* c, const GkDst* dst, void* stream) {
  if (c->nblk <= 0) return 0;
  if (dst->ndst < 1 || dst->ndst > GK_MAXP) return (int)cudaErrorInvalidValue;
  const int grid = (c->nblk + NWARP - 1) / NWARP;
  if (c->mode == GK_MODE_TOPK || c->mode == GK_MODE_DGC) k_compact_send<true, false><<<grid, T, 0, S(stream)>>>(*c, *dst);
  else if (c->mode == GK_MODE_MINK) k_compact_send<true, true><<<grid, T, 0, S(stream)>>>(*c, *dst);
  else k_compact_send<false, false><<<grid, T, 0, S(stream)>>>(*c, *dst);
  GK_RET();
}

int gk_compress(const GkChain* c, const GkDst* dst, void* stream) {
  if (c->mode == GK_MODE_DGC) return (int)cudaErrorInvalidValue;   // needs a seed: gk_dgc_compress
  int e = gk_accumulate(c, stream);
  if (e) return e;
  if (c->mode == GK_MODE_TOPK || c->mode == GK_MODE_MINK) e = gk_topk_select(c, stream);
  else e = gk_count_pick(c, stream);
  if (e) return e;
  return gk_compact_send(c, dst, stream);
}

int gk_dgc_compress(const GkChain* c, const GkDst* dst, uint64_t seed, void* stream) {
  int e = gk_accumulate(c, stream);
  if (e) return e;
  e = gk_dgc_select(c, seed, stream);
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

int gk_sgd_recv(const GkRecvUpd* u, const GkSrc* src, void* stream) {
  if (u->nblk <= 0) return 0;
  if (src->nsrc < 1 || src->nsrc > GK_MAXP) return (int)cudaErrorInvalidValue;
  // one resident wave: the blocks wait on one another at the grid barrier, so the launch is cooperative
  static int cache = 0;
  const int grid = gk_persistent_grid(k_sgd_recv, &cache, u->nblk);
  GkRecvUpdDev d{u->param, u->mom, u->mask, u->blk, u->hyper, u->gmap, u->ctrl, u->bar, u->scale,
                 u->nblk, u->nelem, u->skip_dense};
  GkSrc s = *src;
  void* args[] = {&d, &s};
  const cudaError_t e = cudaLaunchCooperativeKernel((const void*)k_sgd_recv, dim3(grid), dim3(T), args, 0, S(stream));
  if (e != cudaSuccess) (void)cudaGetLastError();
  return (int)e;
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
  k_randomk<<<gk_sm_count(), T, 0, S(stream)>>>(*c, *dst, seed);
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

