// gk_cpu.cpp — host twins of the sm_100a chain (include/gk_b200.h, gkcpu_* entry points).
//
// Same tables, same wire format, same selection rules as csrc/gk_cuda.cu, written as plain
// sequential C++ so that (a) the gloo/CPU mode of the framework (BASELINE.json config 1,
// 2 workers without a GPU) has a native hot path and (b) the GPU kernels have a second,
// independently written oracle besides the PyTorch restatement in ops/oracle.py.
// Semantics cited per function; the reference itself has no native code.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <utility>
#include <vector>

#include "../include/gk_b200.h"

namespace {

inline uint32_t fbits(float x) {
  uint32_t u;
  std::memcpy(&u, &x, 4);
  return u;
}
inline uint32_t absbits(float x) { return fbits(x) & 0x7fffffffu; }
inline uint32_t sel_key(float x, bool mink) {
  const uint32_t b = absbits(x);
  return mink ? (0x7fffffffu - b) : b;
}
inline int tri_idx(int a, int b) { return (a + b) * (a + b + 1) / 2 + a; }

struct Emit {
  const GkDst* d;
  uint64_t poff;
  uint32_t cap;
  uint32_t pos = 0;   // next output position
  uint32_t total = 0; // selected (pre-cap)
  // returns true when the pair was actually written (=> leaves the residual)
  bool put(uint32_t idx, float v) {
    ++total;
    if (pos >= cap) return false;
    const uint64_t pair = (uint64_t)idx | ((uint64_t)fbits(v) << 32);
    for (int q = 0; q < d->ndst; ++q) d->slot[q][poff + GK_SLOT_HDR + pos] = pair;
    ++pos;
    return true;
  }
};

void publish(const GkChain* c, const GkDst* d, uint64_t poff, uint32_t seqn, uint32_t total, uint32_t sent) {
  const uint64_t hdr = (uint64_t)sent | ((uint64_t)seqn << 32);
  for (int q = 0; q < d->ndst; ++q) d->slot[q][poff] = hdr;
  GkCtrl* ct = c->ctrl;
  ct->total = total;
  ct->sent = sent;
  ct->overflow += total - sent;
  ct->sent_total += sent;
  ct->ring[seqn & 63u] = total;
  ct->seq = seqn;
}

// acc = grad (+ resid) -> resid, optional zeroing of grad; returns moments of the segment
void accumulate_segment(const GkChain* c, const GkSeg& sg, double& sum, double& sq, double& sa, float& mx) {
  // four interleaved fp64 accumulators per moment (element i of a block goes to accumulator i % 4), added up in a
  // fixed order at the end: deterministic, and the adds no longer form one dependency chain
  double s4[4] = {0, 0, 0, 0}, q4[4] = {0, 0, 0, 0}, a4[4] = {0, 0, 0, 0};
  float m4[4] = {0.f, 0.f, 0.f, 0.f};
  for (int b = sg.blk_begin; b < sg.blk_end; ++b) {
    const GkBlk& B = c->blk[b];
    const float* __restrict src = c->gsrc ? c->gsrc[b] : c->grad + B.start;
    float* __restrict res = c->resid + B.start;
    const bool use_r = c->use_residual != 0;
    auto one = [&](int i, int lane) {
      float x = src[i];
      if (use_r) x += res[i];
      res[i] = x;
      const double xd = (double)x;
      s4[lane] += xd;
      q4[lane] += xd * xd;
      a4[lane] += std::fabs(xd);
      m4[lane] = std::max(m4[lane], std::fabs(x));
    };
    int i = 0;
    for (; i + 4 <= B.len; i += 4) {
      one(i, 0); one(i + 1, 1); one(i + 2, 2); one(i + 3, 3);
    }
    for (int l = 0; i < B.len; ++i, ++l) one(i, l);
    if (c->zero_grad && !c->gsrc) std::memset(c->grad + B.start, 0, sizeof(float) * (size_t)B.len);
  }
  sum = (s4[0] + s4[1]) + (s4[2] + s4[3]);
  sq = (q4[0] + q4[1]) + (q4[2] + q4[3]);
  sa = (a4[0] + a4[1]) + (a4[2] + a4[3]);
  mx = std::max(std::max(m4[0], m4[1]), std::max(m4[2], m4[3]));
}

uint32_t count_gt(const GkChain* c, const GkSeg& sg, float thr) {
  uint32_t n = 0;
  for (int b = sg.blk_begin; b < sg.blk_end; ++b) {
    const GkBlk& B = c->blk[b];
    for (int i = 0; i < B.len; ++i) n += std::fabs(c->resid[B.start + i]) > thr;
  }
  return n;
}


// One block of the momentum-SGD update with every per-group switch a template parameter, so that the loop body is
// straight-line code the compiler vectorises (AVX2 + FMA; std::fmaf is then one vfmadd, same rounding as before).
// WD weight decay, MOM momentum buffer in use, FIRST first step (momentum := g), NEST Nesterov, GRAD a gradient is
// read (else g = 0: the dense pass of the folded update), SCALE the gradient is multiplied by gs first.
template <bool WD, bool MOM, bool FIRST, bool NEST, bool GRAD, bool SCALE>
void sgd_span(float* __restrict p, float* __restrict m, const float* __restrict gin, float gs, int len, const GkHyper& h) {
  const float wd = h.weight_decay, mu = h.momentum, od = 1.f - h.dampening, nlr = -h.lr;
  for (int i = 0; i < len; ++i) {
    float g = GRAD ? gin[i] : 0.f;
    if (SCALE) g = g * gs;
    if (WD) g = std::fmaf(wd, p[i], g);
    float u = g;
    if (MOM) {
      float mm = m[i];
      mm = FIRST ? g : std::fmaf(od, g, mu * mm);
      m[i] = mm;
      u = NEST ? std::fmaf(mu, mm, g) : mm;
    }
    p[i] = std::fmaf(nlr, u, p[i]);
  }
}
using SpanFn = void (*)(float*, float*, const float*, float, int, const GkHyper&);
template <int B>
constexpr SpanFn span_fn() {
  return &sgd_span<(B & 1) != 0, (B & 2) != 0, (B & 4) != 0, (B & 8) != 0, (B & 16) != 0, (B & 32) != 0>;
}
template <int... I>
const SpanFn* span_table(std::integer_sequence<int, I...>) {
  static const SpanFn t[] = {span_fn<I>()...};
  return t;
}
inline void sgd_block(float* p, float* m, const float* g, float gs, int len, const GkHyper& h, bool has_m, bool grad) {
  static const SpanFn* table = span_table(std::make_integer_sequence<int, 64>{});
  const int bits = (h.weight_decay != 0.f ? 1 : 0) | (has_m ? 2 : 0) | (h.first_step ? 4 : 0) | (h.nesterov ? 8 : 0) |
                   (grad ? 16 : 0) | (grad && gs != 1.f ? 32 : 0);
  table[bits](p, m, g, gs, len, h);
}

// exact k largest (or smallest) keys of a segment, ties towards the lowest index: emits in table order
void emit_exact_topk(const GkChain* c, const GkSeg& sg, GkSegStat& st, Emit& em, bool mink) {
  std::vector<uint32_t> keys;
  keys.reserve(sg.n);
  for (int b = sg.blk_begin; b < sg.blk_end; ++b) {
    const GkBlk& B = c->blk[b];
    for (int i = 0; i < B.len; ++i) keys.push_back(sel_key(c->resid[B.start + i], mink));
  }
  const size_t kk = std::min<size_t>((size_t)sg.k, keys.size());
  std::vector<uint32_t> tmp(keys);
  std::nth_element(tmp.begin(), tmp.begin() + (kk - 1), tmp.end(), std::greater<uint32_t>());
  const uint32_t Tk = tmp[kk - 1];
  size_t ngt = 0;
  for (uint32_t v : keys) ngt += v > Tk;
  uint32_t ties = (uint32_t)(kk - ngt);
  st.tbits = Tk;
  st.ties = ties;
  st.chosen = 0;
  size_t pos = 0;
  for (int b = sg.blk_begin; b < sg.blk_end; ++b) {
    const GkBlk& B = c->blk[b];
    for (int i = 0; i < B.len; ++i, ++pos) {
      const uint32_t key = keys[pos];
      bool take = key > Tk;
      if (!take && key == Tk && ties > 0) { take = true; --ties; }
      if (take) {
        const int j = B.start + i;
        if (em.put((uint32_t)j, c->resid[j])) c->resid[j] = 0.f;
      }
    }
  }
}

static inline uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}
static inline uint32_t feistel_perm(uint32_t j, uint32_t n, uint64_t seed) {
  int bits = 2;
  while ((1ull << bits) < (uint64_t)n) bits += 2;
  const int half = bits / 2;
  const uint32_t mask = (1u << half) - 1u;
  uint32_t x = j;
  do {
    uint32_t l = x >> half, r = x & mask;
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


}  // namespace

extern "C" {

const char* gkcpu_version(void) { return "gk_b200 cpu 0.2"; }

int gkcpu_compress(const GkChain* c, const GkDst* dst) {
  if (dst->ndst < 1 || dst->ndst > GK_MAXP) return 1;
  const uint32_t seqn = c->ctrl->seq + 1u;
  Emit em{dst, (uint64_t)(seqn & 1u) * dst->parity_stride, c->cap};
  const bool mink = c->mode == GK_MODE_MINK;
  for (int s = 0; s < c->nseg; ++s) {
    const GkSeg& sg = c->seg[s];
    GkSegStat& st = c->stat[s];
    double sum, sq, sa;
    float mx;
    accumulate_segment(c, sg, sum, sq, sa, mx);
    const double n = (double)sg.n;
    const double mean = sum / n;
    double var = sg.n > 1 ? (sq - sum * mean) / (n - 1.0) : 0.0;
    if (var < 0) var = 0;
    st.mean = (float)mean;
    st.std = (float)std::sqrt(var);
    st.meanabs = (float)(sa / n);
    st.maxabs = mx;
    const double k = (double)sg.k;
    st.out_base = em.total;
    const uint32_t before = em.total;
    if (c->mode == GK_MODE_TOPK || mink) {
      emit_exact_topk(c, sg, st, em, mink);
    } else {
      float thr = 0.f;
      int chosen = 0;
      if (c->mode == GK_MODE_GAUSSIAN) {
        // compression.py:367-381
        const double t0 = (double)st.mean + sg.zabs * (double)st.std;
        double t15 = t0;
        for (int b = 0; b < c->nloops; ++b) {
          double t = t15;
          for (int a = 0; a + b < c->nloops; ++a) {
            st.cand[tri_idx(a, b)] = (float)t;
            t *= 0.5;
          }
          t15 *= 1.5;
        }
        int a = 0, b = 0;
        for (int loop = 0; loop < c->nloops; ++loop) {
          chosen = tri_idx(a, b);
          const double cnt = (double)count_gt(c, sg, st.cand[chosen]);
          if (cnt < 2.0 * k / 3.0) ++a;
          else if (cnt > 4.0 * k / 3.0) ++b;
          else break;
        }
        thr = st.cand[chosen];
      } else if (c->mode == GK_MODE_REDSYNC) {
        // compression.py:653-672
        const float d = mx - st.meanabs;
        const float ratios[7] = {0.5f, 0.25f, 0.75f, 0.125f, 0.375f, 0.625f, 0.875f};
        for (int j = 0; j < 7; ++j) st.cand[j] = st.meanabs + ratios[j] * d;
        double l = 0.0, r = 1.0;
        int node = 0;
        while (r - l > 0.2) {
          const double mid = l + (r - l) / 2;
          chosen = node;
          const double nnz = (double)count_gt(c, sg, st.cand[node]);
          if (nnz > k && 2.0 * k > nnz) break;
          else if (nnz < k / 2.0) { r = mid; node = 2 * node + 1; }
          else { l = mid; node = 2 * node + 2; }
          if (node >= 7) break;
        }
        thr = st.cand[chosen];
      } else if (c->mode == GK_MODE_REDSYNCTRIM) {
        // compression.py:711-732
        const float d = mx - st.meanabs;
        double r = 1.0 - 0.2;
        for (int j = 0; j < GK_NC; ++j) {
          st.cand[j] = st.meanabs + (float)r * d;
          r = r - 0.2;
        }
        chosen = GK_NC - 1;
        for (int j = 0; j < GK_NC; ++j)
          if ((double)count_gt(c, sg, st.cand[j]) >= k) { chosen = j; break; }
        thr = st.cand[chosen];
      } else {  // GK_MODE_THRESHOLD
        st.cand[0] = sg.param;
        thr = sg.param;
      }
      st.thr = thr;
      st.chosen = chosen;
      st.ties = 0;
      for (int b = sg.blk_begin; b < sg.blk_end; ++b) {
        const GkBlk& B = c->blk[b];
        for (int i = 0; i < B.len; ++i) {
          const int j = B.start + i;
          const float x = c->resid[j];
          if (std::fabs(x) > thr)
            if (em.put((uint32_t)j, x)) c->resid[j] = 0.f;
        }
      }
    }
    st.count = em.total - before;
  }
  publish(c, dst, em.poff, seqn, em.total, em.pos);
  return 0;
}

int gkcpu_scatter(const GkSrc* src, GkCtrl* ctrl, float* grad, float scale, int32_t nelem) {
  if (src->nsrc < 1 || src->nsrc > GK_MAXP) return 1;
  const uint32_t seq = ctrl->seq;
  const uint64_t* base = src->arena + (uint64_t)(seq & 1u) * src->parity_stride;
  const uint32_t cap = (uint32_t)(src->slot_stride - GK_SLOT_HDR);
  for (int r = 0; r < src->nsrc; ++r) {
    const uint64_t* slot = base + (uint64_t)r * src->slot_stride;
    if (src->wait && (uint32_t)(slot[0] >> 32) != seq) { ctrl->err = 1; return 2; }
    const uint32_t n = std::min((uint32_t)slot[0], cap);
    for (uint32_t j = 0; j < n; ++j) {
      const uint64_t p = slot[GK_SLOT_HDR + j];
      const uint32_t idx = (uint32_t)p;
      const uint32_t vb = (uint32_t)(p >> 32);
      float v;
      std::memcpy(&v, &vb, 4);
      if (idx < (uint32_t)nelem) grad[idx] += v * scale;
    }
  }
  return 0;
}

int gkcpu_unscatter(const GkSrc* src, GkCtrl* ctrl, float* grad, int32_t nelem) {
  if (src->nsrc < 1 || src->nsrc > GK_MAXP) return 1;
  const uint32_t seq = ctrl->seq;
  const uint64_t* base = src->arena + (uint64_t)(seq & 1u) * src->parity_stride;
  const uint32_t cap = (uint32_t)(src->slot_stride - GK_SLOT_HDR);
  for (int r = 0; r < src->nsrc; ++r) {
    const uint64_t* slot = base + (uint64_t)r * src->slot_stride;
    const uint32_t n = std::min((uint32_t)slot[0], cap);
    for (uint32_t j = 0; j < n; ++j) {
      const uint32_t idx = (uint32_t)slot[GK_SLOT_HDR + j];
      if (idx < (uint32_t)nelem) grad[idx] = 0.f;
    }
  }
  return 0;
}

int gkcpu_sgd(float* param, const float* grad, float* mom, const GkBlk* blk, int32_t nblk, const GkHyper* hyper,
              const float* grad_scale) {
  // torch.optim.SGD arithmetic (restated in the reference at dl_trainer.py:848-872)
  const float gs = grad_scale ? *grad_scale : 1.f;
  for (int b = 0; b < nblk; ++b) {
    const GkBlk& B = blk[b];
    const GkHyper& h = hyper[B.seg];
    const bool has_m = h.momentum != 0.f && mom != nullptr;
    sgd_block(param + B.start, has_m ? mom + B.start : nullptr, grad + B.start, gs, B.len, h, has_m, true);
  }
  return 0;
}

int gkcpu_sgd_recv(const GkRecvUpd* u, const GkSrc* src) {
  // Folded receive + SGD (twin of k_sgd_recv): dense pass with a zero gradient, then the received pairs are
  // summed per coordinate in rank order inside u->mask (read as floats here) and folded in.  Same result as
  // gkcpu_scatter + gkcpu_sgd up to the rounding of the split FMA on the touched coordinates.
  if (src->nsrc < 1 || src->nsrc > GK_MAXP) return 1;
  if (!u->skip_dense) {
    for (int b = 0; b < u->nblk; ++b) {
      const GkBlk& B = u->blk[b];
      const GkHyper& h = u->hyper[B.seg];
      const bool has_m = h.momentum != 0.f && u->mom != nullptr;
      sgd_block(u->param + B.start, has_m ? u->mom + B.start : nullptr, nullptr, 1.f, B.len, h, has_m, false);
    }
  }
  const uint32_t seq = u->ctrl->seq;
  const uint64_t* base = src->arena + (uint64_t)(seq & 1u) * src->parity_stride;
  const uint32_t cap = (uint32_t)(src->slot_stride - GK_SLOT_HDR);
  float* tot = reinterpret_cast<float*>(u->mask);
  uint32_t cnt[GK_MAXP];
  for (int r = 0; r < src->nsrc; ++r) {
    const uint64_t* slot = base + (uint64_t)r * src->slot_stride;
    if (src->wait && (uint32_t)(slot[0] >> 32) != seq) { u->ctrl->err = 1; return 2; }
    cnt[r] = std::min((uint32_t)slot[0], cap);
  }
  std::vector<uint32_t> touched;
  for (int r = 0; r < src->nsrc; ++r) {
    const uint64_t* slot = base + (uint64_t)r * src->slot_stride;
    for (uint32_t j = 0; j < cnt[r]; ++j) {
      const uint64_t p = slot[GK_SLOT_HDR + j];
      const uint32_t idx = (uint32_t)p;
      if (idx >= (uint32_t)u->nelem) continue;
      const uint32_t vb = (uint32_t)(p >> 32);
      float v;
      std::memcpy(&v, &vb, 4);
      tot[idx] = tot[idx] + v * u->scale;
      touched.push_back(idx);
    }
  }
  for (uint32_t idx : touched) {
    const float g = tot[idx];
    if (fbits(g) == 0u) continue;        // already folded in (coordinate sent by several ranks) or exact zero
    tot[idx] = 0.f;
    const GkHyper& h = u->hyper[u->gmap ? u->gmap[idx >> 6] : 0];
    const bool has_m = h.momentum != 0.f && u->mom != nullptr;
    float du = g;
    if (has_m) {
      const float dm = h.first_step ? g : (1.f - h.dampening) * g;
      u->mom[idx] = u->mom[idx] + dm;
      du = h.nesterov ? std::fmaf(h.momentum, dm, g) : dm;
    }
    u->param[idx] = std::fmaf(-h.lr, du, u->param[idx]);
  }
  for (uint32_t idx : touched) tot[idx] = 0.f;   // -0.0 leftovers
  return 0;
}

int gkcpu_lars(float* param, const float* grad, float* mom, const GkBlk* blk, int32_t nblk, const GkSeg* tens,
               int32_t ntens, const GkHyper* hyper, double* part, float* ratio) {
  (void)part;
  (void)nblk;
  for (int t = 0; t < ntens; ++t) {  // lars.py:83-132
    const GkSeg& tg = tens[t];
    const GkHyper& h = hyper[tg.aux];
    double w2 = 0, g2 = 0;
    for (int b = tg.blk_begin; b < tg.blk_end; ++b)
      for (int i = 0; i < blk[b].len; ++i) {
        const double a = param[blk[b].start + i], g = grad[blk[b].start + i];
        w2 += a * a;
        g2 += g * g;
      }
    const float wn = (float)std::sqrt(w2), gn = (float)std::sqrt(g2);
    float tr = 1.f;
    if (wn > 0.f && gn > 0.f) tr = h.lars_eeta * wn / (gn + h.weight_decay * wn + h.lars_eps);
    tr = std::min(std::max(tr, 0.f), 50.f);
    const float slr = h.lr * tr;
    ratio[t] = slr;
    for (int b = tg.blk_begin; b < tg.blk_end; ++b)
      for (int i = 0; i < blk[b].len; ++i) {
        const int j = blk[b].start + i;
        float dg = std::fmaf(h.weight_decay, param[j], grad[j]);
        dg = std::min(std::max(dg, -10.f), 10.f);
        const float a = std::fmaf(slr, dg, h.momentum * mom[j]);
        mom[j] = a;
        param[j] = param[j] - a;
      }
  }
  return 0;
}

int gkcpu_grad_norm(const float* grad, const GkBlk* blk, int32_t nblk, float max_norm, float* out) {
  double s = 0;
  for (int b = 0; b < nblk; ++b)
    for (int i = 0; i < blk[b].len; ++i) {
      const double x = grad[blk[b].start + i];
      s += x * x;
    }
  const float norm = (float)std::sqrt(s);
  const float coef = max_norm / (norm + 1e-6f);
  out[0] = norm;
  out[1] = (max_norm > 0.f && coef < 1.f) ? coef : 1.f;
  return 0;
}

int gkcpu_randomk_send(const GkChain* c, const GkDst* dst, uint64_t seed) {
  // compression.py:461-475 (k distinct random coordinates); same keyed bijection as the GPU
  if (dst->ndst < 1 || dst->ndst > GK_MAXP) return 1;
  const uint32_t seqn = c->ctrl->seq + 1u;
  const uint64_t poff = (uint64_t)(seqn & 1u) * dst->parity_stride;
  uint32_t total = 0;
  for (int s = 0; s < c->nseg; ++s) {
    const GkSeg& sg = c->seg[s];
    for (uint32_t j = 0; j < (uint32_t)sg.k; ++j) {
      const uint32_t li = feistel_perm(j, (uint32_t)sg.n, seed + 0x632be59bd9b4e019ull * (uint64_t)(s + 1) +
                                                              0x9e3779b97f4a7c15ull * (uint64_t)seqn);
      int lo = sg.blk_begin, hi = sg.blk_end - 1;
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if ((uint32_t)c->blk[mid].soff <= li) lo = mid; else hi = mid - 1;
      }
      const GkBlk& B = c->blk[lo];
      const uint32_t idx = (uint32_t)B.start + (li - (uint32_t)B.soff);
      const uint32_t pos = (uint32_t)sg.aux + j;
      if (pos < c->cap) {
        const uint64_t pair = (uint64_t)idx | ((uint64_t)fbits(c->resid[idx]) << 32);
        for (int q = 0; q < dst->ndst; ++q) dst->slot[q][poff + GK_SLOT_HDR + pos] = pair;
        c->resid[idx] = 0.f;
      }
    }
    total += (uint32_t)sg.k;
  }
  publish(c, dst, poff, seqn, total, std::min(total, c->cap));
  return 0;
}

int gkcpu_bucket_means(const GkChain* c, float* means) {
  // compression.py:243-259: mean of the non-negative entries and of the negative entries
  for (int s = 0; s < c->nseg; ++s) {
    const GkSeg& sg = c->seg[s];
    double sp = 0, sn = 0, cp = 0, cn = 0;
    for (int b = sg.blk_begin; b < sg.blk_end; ++b)
      for (int i = 0; i < c->blk[b].len; ++i) {
        const float x = c->grad[c->blk[b].start + i];
        if (x >= 0.f) { sp += x; cp += 1; } else { sn += x; cn += 1; }
      }
    means[2 * s] = cp > 0 ? (float)(sp / cp) : 0.f;
    means[2 * s + 1] = cn > 0 ? (float)(sn / cn) : 0.f;
  }
  return 0;
}

int gkcpu_bucket_reconstruct(const GkChain* c, const float* lm, const float* gm) {
  // compression.py:261-265 subtracts the local bucket mean, :305-312 adds the averaged one
  for (int b = 0; b < c->nblk; ++b) {
    const GkBlk& B = c->blk[b];
    for (int i = 0; i < B.len; ++i) {
      float& x = c->grad[B.start + i];
      x = (x >= 0.f) ? (x - lm[2 * B.seg]) + gm[2 * B.seg] : (x - lm[2 * B.seg + 1]) + gm[2 * B.seg + 1];
    }
  }
  return 0;
}

int gkcpu_dgc_compress(const GkChain* c, const GkDst* dst, uint64_t seed) {
  // compression.py:576-606 with the sample drawn by the keyed bijection of the random-k kernels (same
  // coordinates as csrc/gk_cuda.cu for the same seed): k-th largest |acc| of int(0.01 n) sampled coordinates
  // is the threshold; more than 4k/3 above it => exact top-k instead.
  if (dst->ndst < 1 || dst->ndst > GK_MAXP) return 1;
  const uint32_t seqn = c->ctrl->seq + 1u;
  Emit em{dst, (uint64_t)(seqn & 1u) * dst->parity_stride, c->cap};
  for (int s = 0; s < c->nseg; ++s) {
    const GkSeg& sg = c->seg[s];
    GkSegStat& st = c->stat[s];
    double sum, sq, sa;
    float mx;
    accumulate_segment(c, sg, sum, sq, sa, mx);
    const double n = (double)sg.n;
    const double mean = sum / n;
    double var = sg.n > 1 ? (sq - sum * mean) / (n - 1.0) : 0.0;
    if (var < 0) var = 0;
    st.mean = (float)mean; st.std = (float)std::sqrt(var); st.meanabs = (float)(sa / n); st.maxabs = mx;
    st.out_base = em.total;
    const uint32_t before = em.total;
    const int fk = (int)((double)sg.n * 0.01);
    uint32_t tb = 0;
    if (fk >= sg.k && fk >= 1) {
      const uint64_t sd = seed + 0x632be59bd9b4e019ull * (uint64_t)(s + 1) + 0x9e3779b97f4a7c15ull * (uint64_t)seqn;
      std::vector<uint32_t> keys((size_t)fk);
      for (int j = 0; j < fk; ++j) {
        const uint32_t li = feistel_perm((uint32_t)j, (uint32_t)sg.n, sd);
        int lo = sg.blk_begin, hi = sg.blk_end - 1;
        while (lo < hi) {
          const int mid = (lo + hi + 1) >> 1;
          if ((uint32_t)c->blk[mid].soff <= li) lo = mid; else hi = mid - 1;
        }
        const GkBlk& B = c->blk[lo];
        keys[(size_t)j] = absbits(c->resid[(uint32_t)B.start + (li - (uint32_t)B.soff)]);
      }
      std::nth_element(keys.begin(), keys.begin() + (sg.k - 1), keys.end(), std::greater<uint32_t>());
      tb = keys[(size_t)sg.k - 1];
    }
    size_t passed = 0;
    for (int b = sg.blk_begin; b < sg.blk_end; ++b)
      for (int i = 0; i < c->blk[b].len; ++i) passed += absbits(c->resid[c->blk[b].start + i]) > tb;
    if ((double)passed > 4.0 * (double)sg.k / 3.0) {
      emit_exact_topk(c, sg, st, em, false);
    } else {
      st.tbits = tb; st.ties = 0; st.chosen = 0;
      for (int b = sg.blk_begin; b < sg.blk_end; ++b) {
        const GkBlk& B = c->blk[b];
        for (int i = 0; i < B.len; ++i) {
          const int j = B.start + i;
          if (absbits(c->resid[j]) > tb)
            if (em.put((uint32_t)j, c->resid[j])) c->resid[j] = 0.f;
        }
      }
    }
    st.count = em.total - before;
  }
  publish(c, dst, em.poff, seqn, em.total, em.pos);
  return 0;
}

}  // extern "C"
