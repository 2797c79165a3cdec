/*
 * gk_b200.h — C ABI of the B200-native sparsified-gradient kernel chain.
 *
 * Two shared libraries implement this interface:
 *   libgk_cuda.so  (csrc/gk_cuda.cu, sm_100a)  — entry points prefixed gk_
 *   libgk_cpu.so   (csrc/gk_cpu.cpp, host)     — entry points prefixed gkcpu_
 * Both are plain C ABI (no torch / pybind dependency) and allocate nothing unless stated:
 * every buffer, scratch included, is passed in by the caller (the Python layer hands in
 * torch tensors).  All device entry points enqueue on the given stream and return
 * immediately; the return value is a cudaError_t cast to int (0 = success).
 *
 * The chain replaces the per-op torch pipeline of the reference
 * (compression.py:358-389 gaussian-k, :45-49 exact top-k, distributed_optimizer.py:416-430
 * send side, :449-485 receive side, dl_trainer.py:848-872 SGD math, lars.py:67-134).
 */
#ifndef GK_B200_H
#define GK_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GK_NC 16          /* max candidate thresholds evaluated in one counting pass      */
#define GK_MAXP 16        /* max ranks whose receive slots one send kernel can store into */
#define GK_HIST_BINS 2048 /* radix-select digit width (11 bits)                           */
#define GK_THREADS 256    /* threads per block of every chain kernel                      */
#define GK_SLOT_HDR 2     /* u64 words in front of every slot: [count | seq<<32, reserved]*/

/* selection modes (GkChain.mode) */
enum {
  GK_MODE_GAUSSIAN = 0,    /* compression.py:358-389 (nloops=3) and :405-435 (nloops=5) */
  GK_MODE_TOPK = 1,        /* compression.py:45-49 exact k largest |x|, ties -> lowest index */
  GK_MODE_REDSYNC = 2,     /* compression.py:644-677 */
  GK_MODE_REDSYNCTRIM = 3, /* compression.py:707-738 */
  GK_MODE_THRESHOLD = 4,   /* fixed threshold supplied in GkSeg.param (select |x| > param) */
  GK_MODE_MINK = 5,        /* compression.py:52-56 k smallest |x| */
  GK_MODE_DGC = 6          /* compression.py:576-606 threshold from a 1 % sample, exact top-k if > 4k/3 pass */
};

/* One block of work: a contiguous extent of the arena that belongs to one segment. */
typedef struct GkBlk {
  int32_t start; /* arena index of the first element (multiple of 4)      */
  int32_t len;   /* number of elements (<= chunk)                          */
  int32_t seg;   /* selection segment (compress tables) / param group or tensor (optimizer tables) */
  int32_t soff;  /* segment-local index of the block's first element (dense numbering of the segment) */
} GkBlk;

/* One selection segment = one tensor (threshold=0) or one merged group. */
typedef struct GkSeg {
  int32_t n;         /* real element count (sum of its blocks' len)           */
  int32_t k;         /* max(int(n*density),1), host-computed (compression.py:363) */
  int32_t blk_begin; /* its blocks are [blk_begin, blk_end) of the table       */
  int32_t blk_end;
  double zabs;       /* |Phi^-1(density/2)| (utils.py:136-138); gaussian only  */
  float param;       /* GK_MODE_THRESHOLD: the threshold                       */
  int32_t aux;       /* compress tables: prefix sum of k over earlier segments (random-k output base);
                        LARS tensor tables: index of the tensor's param group in the GkHyper array */
} GkSeg;

/* Per-segment device-side state written by the kernels. */
typedef struct GkSegStat {
  float mean, std, meanabs, maxabs; /* moments of acc = grad + residual            */
  float cand[GK_NC];                /* candidate thresholds of the counting pass   */
  float thr;                        /* chosen threshold (strict |x| > thr)         */
  uint32_t tbits;                   /* top-k: bit pattern of the k-th magnitude    */
  uint32_t ties;                    /* top-k: how many |x|==T elements to take     */
  int32_t chosen;                   /* index of the chosen candidate               */
  uint32_t count;                   /* elements selected in this segment (pre-cap) */
  uint32_t out_base;                /* first output position of this segment       */
  uint32_t prefix;                  /* radix-select: matched high bits so far      */
  uint32_t krem;                    /* radix-select: rank still to find            */
  uint8_t rcand[GK_NC];             /* candidates counted by the current round (first nrc entries)  */
  int32_t nrc;                      /* how many; 0 = no round planned                               */
  uint32_t tot[GK_NC];              /* per-candidate totals of the counting rounds (zeroed by K1's epilogue) */
  uint32_t known;                   /* bit j: tot[j] is final                                       */
  int32_t gate;                     /* rows whose maximum bit pattern is <= gate cannot contribute to the round */
  int32_t done;                     /* threshold search complete / DGC: sampled threshold accepted  */
  uint32_t sprefix, skrem;          /* DGC: radix-select state of the sample                        */
} GkSegStat;

/* Chain-wide control block (device memory, zero-initialised once by the caller). */
typedef struct GkCtrl {
  uint32_t done[8];  /* last-block-done tickets, self-resetting                 */
  uint32_t total;    /* pairs selected by the last compress (pre-cap)           */
  uint32_t sent;     /* pairs actually emitted (min(total, cap))                */
  uint32_t seq;      /* sequence number of the last completed send              */
  uint32_t err;      /* !=0: receive timed out / protocol error (sticky)        */
  uint32_t overflow; /* cumulative pairs withheld because of the cap            */
  uint32_t rsv[3];   /* grid barrier of the counting rounds: arrivals, generation, "another round" */
  uint64_t sent_total; /* cumulative pairs emitted (for logging)                */
  uint32_t ring[64];   /* ring[seq & 63] = pairs selected by send number seq (the reference keeps
                          this list on the host, distributed_optimizer.py:420)   */
} GkCtrl;

/* Everything one compress chain needs.  Passed by pointer from the host; the library
 * copies it by value into the kernel parameters, so it may live on the caller's stack. */
typedef struct GkChain {
  float* grad;        /* gradient arena (read; zeroed behind the read if zero_grad)  */
  float* resid;       /* residual arena (read, overwritten with acc, selected zeroed) */
  const GkBlk* blk;   /* [nblk] device                                               */
  GkSeg* seg;         /* [nseg] device (host rewrites k / zabs when density changes) */
  double* part;       /* [nblk*4] scratch: per-block sum, sumsq, sumabs, maxabs      */
  GkSegStat* stat;    /* [nseg] scratch                                              */
  uint32_t* bcnt;     /* [nblk*GK_NC] scratch: per-block counts per candidate        */
  uint32_t* boff;     /* [nblk*2] scratch: per-block (gt, eq) exclusive prefixes     */
  uint32_t* hist;     /* [nseg*GK_HIST_BINS] scratch (top-k only; zero-initialised)  */
  GkCtrl* ctrl;       /* [1] device                                                  */
  int32_t nblk, nseg;
  int32_t mode;         /* GK_MODE_*                                                 */
  int32_t nloops;       /* gaussian: 3 (gaussian) or 5 (gaussian2)                   */
  int32_t use_residual; /* 1: acc = grad + resid (error feedback); 0: acc = grad     */
  int32_t zero_grad;    /* 1: write 0 into grad after it is consumed                 */
  uint32_t cap;         /* max pairs emitted per send (receive-slot capacity)        */
  int32_t rsv;          /* experiment switches (tools/bench_chain.py); 0 in production */
  const float* const* gsrc; /* optional [nblk] device array: where K1 reads each table entry's gradient
                               from.  NULL = grad + start.  Lets K1 consume the tensors autograd
                               produced (p.grad stolen, no accumulate-into-arena kernels); those
                               sources are read only — zero_grad does not apply to them            */
  uint32_t* rowmax;   /* optional [nblk*rpb] scratch: K1 leaves the largest |acc| bit pattern of every
                         128-element row of every table entry here; the later passes read only the rows
                         that can hold an element above their threshold.  NULL = every row is read      */
  int32_t rpb;        /* rows reserved per table entry: ceil(chunk/128)                                 */
  int32_t rsv2;
} GkChain;

/* Where the compaction kernel stores this rank's pairs: one slot per destination.
 * A slot is GK_SLOT_HDR u64 header words followed by cap u64 pairs
 * (low 32 bits = int32 arena index, high 32 bits = fp32 value bits).
 * slot[d] is the parity-0 copy; parity p lives parity_stride u64 words further. */
typedef struct GkDst {
  uint64_t* slot[GK_MAXP];
  int32_t ndst;
  int32_t rsv;
  uint64_t parity_stride; /* 0 = single-buffered (all_gather transports) */
} GkDst;

/* Receive arena of one rank: nsrc slots of slot_stride u64 words each, per parity. */
typedef struct GkSrc {
  const uint64_t* arena; /* parity-0 slot 0                                   */
  uint64_t slot_stride;  /* = GK_SLOT_HDR + cap                                */
  uint64_t parity_stride;
  int32_t nsrc;
  int32_t wait;          /* 1: spin (bounded) until every slot's seq matches   */
  uint64_t timeout_ns;   /* bound of that spin; on expiry ctrl->err is set     */
} GkSrc;

/* Optimizer hyper-parameters of one param group (device array, rewritten by the host
 * whenever lr changes; dl_trainer.py:197-228 defines the values). */
typedef struct GkHyper {
  float lr, momentum, dampening, weight_decay;
  int32_t nesterov;
  int32_t first_step; /* 1 on the first step: momentum buffer := grad (torch.optim.SGD) */
  float lars_eeta, lars_eps;
} GkHyper;

/* ------------------------------------------------------------------ CUDA entry points */

/* K1  ef_accumulate_moments (+ first radix histogram when mode is top-k / min-k).
 * acc = grad (+ resid) -> resid; per-segment mean/std(unbiased)/mean|x|/max|x|; candidate
 * thresholds.  Replaces compression.py:365-370 and distributed_optimizer.py:375. */
int gk_accumulate(const GkChain* c, void* stream);

/* K2  threshold search on the device: counting rounds over the rows that can exceed the round's smallest
 * candidate, the reference's refinement loop replayed between rounds (grid barrier, cooperative launch),
 * then the prefix sums of the per-entry counts for K3.  Replaces compression.py:372-381 (and :653-672,
 * :719-732).  No host sync.  Uses ctrl->rsv as barrier state. */
int gk_count_pick(const GkChain* c, void* stream);

/* K4  exact top-k radix select (3 digit passes; the first is fused into gk_accumulate)
 * followed by the (gt, eq) counting pass.  Replaces compression.py:45-49. */
int gk_topk_select(const GkChain* c, void* stream);

/* K3  select + ordered compaction + residual update + send: stores the (index,value)
 * pairs into dst->slot[d] of every destination (peer-mapped memory for NVLink P2P),
 * then publishes {count, seq} with a system-scope release store.
 * Replaces compression.py:373-388 and distributed_optimizer.py:426-427. */
int gk_compact_send(const GkChain* c, const GkDst* dst, void* stream);

/* Convenience: the whole send-side chain for c->mode. */
int gk_compress(const GkChain* c, const GkDst* dst, void* stream);

/* K9  DGC sampling (compression.py:576-606), GK_MODE_DGC: k-th largest of a 1 % sample drawn with the keyed
 * bijection of gk_randomk_send -> threshold -> count -> (exact radix select if more than 4k/3 pass) ->
 * (gt, eq) counts + prefix sums.  gk_dgc_compress = gk_accumulate + gk_dgc_select + gk_compact_send. */
int gk_dgc_select(const GkChain* c, uint64_t seed, void* stream);
int gk_dgc_compress(const GkChain* c, const GkDst* dst, uint64_t seed, void* stream);

/* K5  receive: (optionally wait for all slots,) scatter-add value*scale into the dense
 * gradient arena.  Replaces distributed_optimizer.py:461-482. */
int gk_scatter(const GkSrc* src, GkCtrl* ctrl, float* grad, float scale, int32_t nelem_arena,
               void* stream);

/* zero_grad for a sparse gradient: clear exactly the coordinates the last receive wrote. */
int gk_unscatter(const GkSrc* src, GkCtrl* ctrl, float* grad, int32_t nelem_arena, void* stream);

/* K6  fused momentum-SGD over the parameter / gradient / momentum arenas
 * (torch.optim.SGD arithmetic, dl_trainer.py:848-872).  grad_scale may be NULL. */
int gk_sgd(float* param, const float* grad, float* mom, const GkBlk* blk, int32_t nblk,
           const GkHyper* hyper, const float* grad_scale, void* stream);

/* K6 (folded)  receive + momentum-SGD in ONE cooperative launch; no dense gradient is ever formed
 * (replaces distributed_optimizer.py:461-482 + :543-546 / dl_trainer.py:848-872 together).
 *   phase A: p, m updated as if the gradient were zero (weight decay + momentum; 16 B/element) — this does not
 *            depend on the peers and runs while their pairs are still arriving;
 *   then the bounded wait on the slot headers, and the <= P*k received pairs are folded in per coordinate,
 *   contributions summed in rank order (=> bit-identical replicas).
 * Requirements: every source list sorted by ascending index (the compaction kernels emit them that way when the
 * table entries are in arena order; random-k lists are not — use gk_scatter + gk_sgd there; a violation sets
 * ctrl->err = 2); `mask` is an all-zero uint32 view of nelem words (the gradient arena) and is all zero again
 * afterwards; `gmap[i >> 6]` = index into `hyper` of arena element i (NULL: group 0 everywhere); `bar` = 2
 * zero-initialised uint32 (grid barrier state, never reset); `blk[].seg` indexes `hyper` as in gk_sgd.
 * skip_dense != 0 leaves phase A out (momentum == 0 and weight_decay == 0 in every group). */
typedef struct GkRecvUpd {
  float* param;
  float* mom;           /* may be NULL (no momentum) */
  uint32_t* mask;
  const GkBlk* blk;
  const GkHyper* hyper;
  const uint8_t* gmap;
  GkCtrl* ctrl;         /* the chain's control block: seq of the exchange, err */
  uint32_t* bar;
  float scale;          /* 1 / world size */
  int32_t nblk, nelem, skip_dense;
} GkRecvUpd;
int gk_sgd_recv(const GkRecvUpd* u, const GkSrc* src, void* stream);

/* K7  LARS (lars.py:67-134): segmented ||w||, ||g|| then the trust-ratio update. */
int gk_lars(float* param, const float* grad, float* mom, const GkBlk* blk, int32_t nblk,
            const GkSeg* tens, int32_t ntens, const GkHyper* hyper, double* part, float* ratio,
            GkCtrl* ctrl, void* stream);

/* K8  global L2 norm of the gradient arena and the clip coefficient
 * min(1, max_norm/(norm+1e-6)) written to out[0..1] = {norm, coef}
 * (dist_trainer.py:80-85).  The coefficient is consumed by gk_sgd via grad_scale. */
int gk_grad_norm(const float* grad, const GkBlk* blk, int32_t nblk, double* part, float max_norm,
                 float* out, GkCtrl* ctrl, void* stream);

/* gk_grad_norm without caller-provided scratch.  (No counterpart in the reference: it exists for users
 * of this C ABI, not for parity.)
 * Same result as gk_grad_norm, but the per-block partial sums and the ticket live in a fixed
 * `static __device__` array inside the library, so neither the caller nor the library allocates
 * anything.  Meant for callers that cannot obtain scratch at call time: hosts that pre-carve all device
 * memory and talk to this library through the C ABI only, monitoring / abort paths that must work when
 * the device is out of memory, and code that must not call cudaMalloc/cudaFree because those
 * synchronise the whole device while other streams are running.
 *   - limit: at most GK_NOALLOC_MAX_BLOCKS = 16384 table entries (134 M elements at the default
 *     8192-element chunk); a larger table returns GK_ERR_NOALLOC_TOO_LARGE (-2), prints a message to
 *     stderr and launches nothing;
 *   - cost: 128 KiB + sizeof(GkCtrl) of device memory on every GPU that loads the library, until exit;
 *   - the first call on a device looks up the array's address (cudaGetSymbolAddress with that device
 *     current) and caches it; make that first call before starting any stream capture;
 *   - the array is shared by all calls and keeps the last call's contents; the ticket is cleared on
 *     `stream` before each launch, the partials are fully rewritten.  Calls that overlap on the GPU
 *     (other streams / threads) race on it — order them. */
#define GK_NOALLOC_MAX_BLOCKS 16384
#define GK_ERR_NOALLOC_TOO_LARGE (-2)
int gk_grad_norm_static(const float* grad, const GkBlk* blk, int32_t nblk, float max_norm, float* out,
                        void* stream);

/* y[i] *= *scale over the blocks (materialised clipping). */
int gk_scale(float* grad, const GkBlk* blk, int32_t nblk, const float* scale, void* stream);

/* K9  random-k: k distinct pseudo-random indices per segment from a keyed bijection of
 * [0,n) (compression.py:461-475); gathers values, zeroes the residual there, sends. */
int gk_randomk_send(const GkChain* c, const GkDst* dst, uint64_t seed, void* stream);

/* K10 bucket compressor (compression.py:243-265, 305-312): per-segment mean of the
 * non-negative and of the negative entries -> means[2*nseg]; then reconstruct. */
int gk_bucket_means(const GkChain* c, float* means, void* stream);
int gk_bucket_reconstruct(const GkChain* c, const float* local_means, const float* global_means,
                          void* stream);

/* K13 flat broadcast helper / generic: dst[i] = src[i] over n floats (peer pointers ok). */
int gk_copy(float* dst, const float* src, int64_t n, void* stream);

/* K14 on-device histogram of acc for the gradient-distribution study
 * (distributed_optimizer.py:422-424 dumped the whole tensor to disk instead). */
int gk_histogram(const float* x, int64_t n, float lo, float hi, int32_t nbins, uint32_t* out,
                 void* stream);

/* Fused training-mode batch norm on NHWC fp32 activations viewed as x[M = N*H*W][C], C % 4 == 0, every
 * pointer 16-byte aligned (csrc/gk_bn.cu).  The reference takes BN, ReLU and the residual add from
 * torch.nn as three to four separate passes (models/resnet.py via torchvision); here
 *   gk_bn_fwd:  y = [relu]( (x - mean) * istd * gamma + beta [+ res] )     (res may be NULL)
 *               save_mean / save_istd written, running stats (may be NULL) and *num_batches (may be NULL)
 *               updated with torch.nn.BatchNorm2d semantics (biased batch variance, unbiased running)
 *   gk_bn_bwd:  g = relu ? dy * (forward output > 0) : dy; dgamma = sum g*xhat, dbeta = sum g,
 *               dx = gamma*istd * (g - mean g - xhat * mean g*xhat); if dres != NULL it receives g (the
 *               residual branch's gradient) and y (the saved forward output) must be given; with y == NULL
 *               the ReLU mask is recomputed from x.
 * Scratch owned by the caller, per layer or per stream: accum = 2*C doubles and *ticket = one uint32, both
 * zero before the first call (the kernels leave them zero), coef = 2*C floats.  Two launches each, no host
 * synchronisation, no allocation. */
int gk_bn_fwd(const float* x, const float* res, float* y, const float* gamma, const float* beta,
              float* running_mean, float* running_var, int64_t* num_batches, float* save_mean, float* save_istd,
              double* accum, uint32_t* ticket, int64_t M, int C, float eps, float momentum, int relu, void* stream);
int gk_bn_bwd(const float* dy, const float* x, const float* y, const float* gamma, const float* beta,
              const float* save_mean, const float* save_istd, float* dx, float* dres, float* dgamma, float* dbeta,
              float* coef, double* accum, uint32_t* ticket, int64_t M, int C, int relu, void* stream);

/* Single-kernel variants of the two calls above (one cooperative launch each; the activation slab of
 * every block stays in shared memory between the statistics and the apply phase, so x / dy are read from
 * HBM once).  Available when gk_bn_fused_supported(M, C) returns 1 (C/4 has a power-of-two factor >= 8 and
 * the channel slabs fit the SM count).  Scratch, per layer: accum = 4*C doubles (two parities) and
 * bar = 128 uint64 (bar[2y], bar[2y+1]: arrival count and generation of channel slab y; bar[96..111]:
 * phase time stamps in ns when flags bit 1 is set; low word of bar[127]: error flag, set if a barrier wait
 * expired); both zero before the first call, never reset afterwards; one scratch serves forward and
 * backward of a layer.  flags: bit 0 = ReLU, bit 1 = time stamps. */
int gk_bn_fused_supported(int64_t M, int C);
/* gk_bn_bwd_fused with a second gradient of the same output (dy + dy2 is formed on the fly; dy2 may be NULL):
 * saves the separate add kernel when a block output feeds both a convolution and an identity shortcut. */
int gk_bn_bwd_fused2(const float* dy, const float* dy2, const float* x, const float* y, const float* gamma,
                     const float* beta, const float* save_mean, const float* save_istd, float* dx, float* dres,
                     float* dgamma, float* dbeta, double* accum, unsigned long long* bar, int64_t M, int C,
                     int flags, void* stream);
/* BN + ReLU + max pool 3x3 / stride 2 / padding 1 in one (the ResNet stem; torch.nn runs three modules and
 * writes the full-resolution activation twice).  x is [N][H][W][C]; the forward writes the pooled activation
 * yp[N][OH][OW][C] (OH = (H-1)/2+1) and pidx, one byte per pooled element: kh*3+kw of the first maximum of its
 * window; scratch as gk_bn_fwd (accum 2*C doubles + ticket).  The backward takes the pooled gradient and pidx
 * and produces dx, dgamma, dbeta; scratch as gk_bn_bwd_fused; needs gk_bn_fused_supported(N*H*W, C). */
/* gk_bn_pool_fwd without its statistics pass: the sums come from partials[nparts][2C] (sums, then sums of squares) */
int gk_bn_pool_fwd_sums(const float* x, float* yp, unsigned char* pidx, const float* gamma, const float* beta,
                        float* running_mean, float* running_var, int64_t* num_batches, float* save_mean,
                        float* save_istd, const float* partials, int nparts, int N, int H, int W, int C, float eps,
                        float momentum, void* stream);
int gk_bn_pool_fwd(const float* x, float* yp, unsigned char* pidx, const float* gamma, const float* beta,
                   float* running_mean, float* running_var, int64_t* num_batches, float* save_mean, float* save_istd,
                   double* accum, uint32_t* ticket, unsigned long long* bar, int N, int H, int W, int C, float eps,
                   float momentum, void* stream);
/* forward scratch: with bar != NULL (and a supported shape) the statistics come from the cooperative kernel and
 * accum / bar are the gk_bn_fwd_fused scratch; with bar == NULL accum / ticket are the gk_bn_fwd scratch.
 * backward: gfull = N*H*W*C floats receives the full-resolution pool gradient first (it may be the same
 * buffer as dx); gfull == NULL gathers inside the cooperative kernel instead (slower, no buffer). */
int gk_bn_pool_bwd(const float* dyp, const unsigned char* pidx, const float* x, const float* gamma, const float* beta,
                   const float* save_mean, const float* save_istd, float* dx, float* gfull, float* dgamma,
                   float* dbeta, double* accum, unsigned long long* bar, int N, int H, int W, int C, int flags,
                   void* stream);
/* Two batch norms into one activation: y = relu(bn_a(xa) + bn_b(xb)) — the last BN of a residual block and the
 * BN of its 1x1 downsample shortcut — in one cooperative kernel per direction (the shortcut's normalised
 * activation and its gradient are never materialised).  Pointer tables, 8 entries each:
 *   forward  a[] / b[] = { x, gamma, beta, save_mean, save_istd, running_mean|NULL, running_var|NULL, num_batches|NULL }
 *   backward a[] / b[] = { x, gamma, beta, save_mean, save_istd, dx, dgamma, dbeta }
 * dy2 may be NULL (second gradient of y, added on the fly); y is the saved forward output (ReLU mask).
 * Scratch per pair of layers: accum = 8*C doubles (two parities of four rows), bar as for gk_bn_fwd_fused;
 * needs gk_bn_fused_supported(M, C).  flags bit 0 = ReLU. */
int gk_bn2_fwd_fused(void* const* a, void* const* b, float eps_a, float mom_a, float eps_b, float mom_b, float* y,
                     unsigned char* mask, double* accum, unsigned long long* bar, int64_t M, int C, int flags,
                     void* stream);
/* ReLU bit masks: a forward call given mask != NULL also writes one byte per channel quad (M*C/4 bytes, bit j =
 * output j of the quad was positive); the backward calls take that buffer in place of y when flags has bit 2
 * (value 4) set — a sixteenth of the bytes of re-reading the activation.  gk_bn_fwd_fused_m is gk_bn_fwd_fused
 * with the extra mask argument. */
int gk_bn_fwd_fused_m(const float* x, const float* res, float* y, const float* gamma, const float* beta,
                      float* running_mean, float* running_var, int64_t* num_batches, float* save_mean,
                      float* save_istd, double* accum, unsigned long long* bar, int64_t M, int C, float eps,
                      float momentum, int flags, unsigned char* mask, void* stream);
int gk_bn2_bwd_fused(const float* dy, const float* dy2, const float* y, void* const* a, void* const* b, double* accum,
                     unsigned long long* bar, int64_t M, int C, int flags, void* stream);
/* Geometry knobs for benchmarks (defaults 9 = no cap, 128): cap on log2 of the channel slab in float4 quads, and the
 * largest per-block slab in KB for which one block takes a channel slab over all rows (no barrier). */
int gk_bn_set_tuning(int qcap_log2, int private_kb);
/* Activations of at least mb MB run the cooperative kernels with 1024 threads per block, smaller ones with 512
 * (default 40). */
int gk_bn_set_big_mb(int mb);
int gk_bn_set_pair_big_mb(int mb); /* the same threshold for the two-BN kernels (default 40) */
int gk_bn_fwd_fused(const float* x, const float* res, float* y, const float* gamma, const float* beta,
                    float* running_mean, float* running_var, int64_t* num_batches, float* save_mean,
                    float* save_istd, double* accum, unsigned long long* bar, int64_t M, int C, float eps,
                    float momentum, int flags, void* stream);
int gk_bn_bwd_fused(const float* dy, const float* x, const float* y, const float* gamma, const float* beta,
                    const float* save_mean, const float* save_istd, float* dx, float* dres, float* dgamma,
                    float* dbeta, double* accum, unsigned long long* bar, int64_t M, int C, int flags, void* stream);

/* Peer-memory plumbing of the P2P transport (set-up time only; the only allocating entry
 * points).  gk_ipc_alloc: cudaMalloc + zero + IPC handle (64 bytes) for the local receive arena;
 * gk_ipc_open maps a peer's arena into this process.  Replaces Horovod's allgather buffers
 * (distributed_optimizer.py:426-427). */
int gk_ipc_alloc(uint64_t bytes, void** ptr, unsigned char* handle64);
int gk_ipc_open(const unsigned char* handle64, void** ptr);
int gk_ipc_close(void* ptr);
int gk_ipc_free(void* ptr);

const char* gk_version(void);

/* ---- ResNet stem convolution on the tensor cores (csrc/gk_stem.cu: TMA + tcgen05.mma kind::tf32 + TMEM).
 * The reference runs this layer (torchvision resnet50 conv1, dl_trainer.py:187-196) through cuDNN.
 * Layouts: x [N, H, W, 3] fp32 NHWC; X the padded 2x2 space-to-depth image [N, H/2+3, W/2+3, 12]; w [64, 3, 7, 7];
 * wpk 12288 floats (operand order); y / gy [N, H/2, W/2, 64] NHWC; `err` an optional device uint32 that receives
 * the code of a pipeline wait that timed out (the kernels trap instead of hanging).  All pointers 16-byte aligned.
 * Return: 0, a cudaError_t, -2 (shape not covered) or -3 (the driver cannot encode the tensor map). */
int gk_stem_driver_ok(const void* probe);            /* 1 if the driver encodes the tensor maps used; probe = any device ptr */
int gk_stem_supported(int H, int W);                 /* forward: even sizes, W <= 332 */
int gk_stem_s2d(const void* x, void* X, int nimg, int H, int W, void* stream);
int gk_stem_s2d_layout(const void* x, void* X, int nimg, int H, int W, int nchw, void* stream); /* x [N,3,H,W] if nchw */
int gk_stem_pack_w(const void* w, void* wpk, void* stream);
int gk_stem_fwd(const void* X, const void* wpk, void* y, int nimg, int OH, int OW, void* err, void* stream);
/* ... and leave the per-channel sums / sums of squares of y in stats[gk_stem_stats_rows()][128] (one partial row per
 * CTA), accumulated in the epilogue: the batch statistics of the BatchNorm that follows (gk_bn_pool_fwd_sums). */
int gk_stem_stats_rows(void);
int gk_stem_fwd_stats(const void* X, const void* wpk, void* y, void* stats, int nimg, int OH, int OW, void* err,
                      void* stream);
int gk_stem_wgrad_supported(int OH, int OW);         /* OW % 8 == 0, OW <= 128 */
long long gk_stem_wgrad_workspace(void);             /* floats of scratch for gk_stem_wgrad */
int gk_stem_wgrad(const void* gy, const void* X, void* gw, void* work, int nimg, int OH, int OW, void* err,
                  void* stream);

/* ------------------------------------------------------------------- CPU twins (host) */
int gkcpu_compress(const GkChain* c, const GkDst* dst);
int gkcpu_dgc_compress(const GkChain* c, const GkDst* dst, uint64_t seed);
int gkcpu_scatter(const GkSrc* src, GkCtrl* ctrl, float* grad, float scale, int32_t nelem_arena);
int gkcpu_unscatter(const GkSrc* src, GkCtrl* ctrl, float* grad, int32_t nelem_arena);
int gkcpu_sgd(float* param, const float* grad, float* mom, const GkBlk* blk, int32_t nblk,
              const GkHyper* hyper, const float* grad_scale);
int gkcpu_sgd_recv(const GkRecvUpd* u, const GkSrc* src);
int gkcpu_lars(float* param, const float* grad, float* mom, const GkBlk* blk, int32_t nblk,
               const GkSeg* tens, int32_t ntens, const GkHyper* hyper, double* part, float* ratio);
int gkcpu_grad_norm(const float* grad, const GkBlk* blk, int32_t nblk, float max_norm, float* out);
int gkcpu_randomk_send(const GkChain* c, const GkDst* dst, uint64_t seed);
int gkcpu_bucket_means(const GkChain* c, float* means);
int gkcpu_bucket_reconstruct(const GkChain* c, const float* local_means, const float* global_means);
const char* gkcpu_version(void);

#ifdef __cplusplus
}
#endif
#endif /* GK_B200_H */
