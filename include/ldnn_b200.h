/* libldnn_b200 — C ABI of the sm_100a kernels and the native runtime.
 *
 * All entry points return 0 on success; on failure ldnn_last_error() describes why.
 * Pointers are device pointers unless noted; `stream` is a cudaStream_t passed as void*.
 * Citations (file:line) refer to the reference tree, see SURVEY.md.
 */
#ifndef LDNN_B200_H_
#define LDNN_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* ldnn_last_error(void);
int ldnn_version(void);
/* number of kernels this library has launched in the calling process (bench.py "gpu_launches") */
unsigned long long ldnn_launch_count(void);

/* ------------------------------------------------------------------ convolution (K-CONV-F/-D/-W)
 * Implicit-GEMM conv on tcgen05/TMEM fed by TMA; replaces cuDNN conv fwd/bwd
 * (Balanced All-Reduce/model.py:55,57,63,78; backward via trainer.py:208). */
typedef struct {
  int mode;            /* 0 fprop, 1 dgrad, 2 wgrad */
  int N, H, W;         /* batch and spatial size of the conv INPUT x */
  int Cin, Cout;       /* multiples of 64 (Cout may be 64) */
  int R, stride, pad;  /* square filter R x R */
  int block_n;         /* 64 / 128 / 256; 0 = auto */
  int splits;          /* split-K factor; 0 = auto */
  int accumulate;      /* dgrad: dx = result + addend */
  const void* x;       /* bf16|fp32 (see dtype) NHWC [N,H,W,Cin]   fprop/wgrad: input;  dgrad: OUTPUT dx */
  const void* w;       /* bf16|fp32 [Cout][R*R*Cin]     (KRSC)                                  */
  void* y;             /* bf16|fp32 NHWC [N,P,Q,Cout]   fprop: OUTPUT; dgrad/wgrad: input dy    */
  void* dw;            /* fp32 [Cout][R*R*Cin]     wgrad: OUTPUT                            */
  const void* addend;  /* dgrad + accumulate: bf16 tensor shaped like dx (may alias x)     */
  float* scratch;      /* split-K partial tiles (see ldnn_conv_query)                      */
  unsigned int* tickets; /* split-K arrival counters, zero-initialised once                */
  unsigned int a_lbo, a_sbo, b_lbo, b_sbo; /* debug: override smem-descriptor strides (0 = default) */
  int cluster;         /* fprop/dgrad: 2 = thread-block clusters of two M tiles, weight tile TMA-multicast to both
                          (needs block_n >= 128 and an even number of M tiles); x, w and y must then be this GPU's own
                          device memory (checked): TMA multicast loads here read only this GPU's memory, since on B200
                          a TMA multicast load from another GPU's memory hung the GPU.  -2 = CTA pair running one
                          tcgen05.mma.cta_group::2 (no multicast load).  0/1 = no cluster                        */
  int lite;            /* 1 = short pipeline (~96 KB smem) + 4 epilogue warps so two CTAs share an SM: for large
                          grids of short-K tiles (one CTA's prologue/epilogue hides under the other's mainloop) */
  int atomic_out;      /* wgrad: split-K partial tiles are red.add'ed straight into dw (caller zeroes dw
                          once per step) instead of going through scratch + ticket; not bit-reproducible */
  float* bn_sums;      /* fprop, optional: fp32 [2][Cout]; the epilogue ADDS per-channel sum and sum of
                          squares of the output (fused BN statistics); caller zeroes it per step        */
  /* dgrad, optional: fold the BN-backward reduction of the layer that owns dx into the epilogue
   * (valid only when this dgrad is the LAST contribution to dx): fuse_sums[2][Cin] += (sum dy, sum dy*xhat),
   * dy = dx * (fuse_out > 0 if fuse_relu).  fuse_out / fuse_y are bf16 tensors shaped like dx. */
  const void* fuse_out; const void* fuse_y; const float* fuse_mean; const float* fuse_invstd; float* fuse_sums;
  int fuse_relu;
  int dtype;           /* 0: bf16 activations/weights, tcgen05 kind::f16.  1: fp32 activations/weights (x, w, y, addend,
                          fuse_out, fuse_y are float), tcgen05 kind::tf32 — the reference's precision (fp32 storage,
                          TF32 convolutions: BalAR/trainer.py:203-210 through model.py:102-111, torch default)   */
} ldnn_conv_desc;

/* sizes a plan needs; covered_classes (dgrad): how many of the stride^2 output parity classes receive any
 * tap — fewer than stride^2 means part of dx is not written and the caller must pre-zero + accumulate */
int ldnn_conv_query(const ldnn_conv_desc* d, size_t* scratch_bytes, size_t* ticket_count, int* launches,
                    int* covered_classes);
/* launch grid the plan would use (for autotuners) */
int ldnn_conv_query_grid(const ldnn_conv_desc* d, int* gx, int* gy, int* gz);
int ldnn_conv_plan_create(const ldnn_conv_desc* d, void** plan);
int ldnn_conv_plan_run(void* plan, void* stream);
void ldnn_conv_plan_destroy(void* plan);

/* stem: x fp32 NCHW [N,C,H,W] -> bf16 im2col rows [N*P*Q][Kpad], k = (r*R+s)*C + c, zero padded.
 * (fused cast + layout change + patch gather for the 3-channel first conv, model.py:78) */
int ldnn_im2col_stem(const float* x, void* col, int N, int C, int H, int W, int R, int stride, int pad, int Kpad,
                     void* stream);

/* ------------------------------------------------------------------ batch norm (K-BN-F / K-BN-B)
 * Training-mode BN fused with residual add and ReLU (model.py:56,58,64,68-71,79-80). */
/* y: bf16 [M][C] conv output.  `chunks` CTAs each reduce M/chunks rows into partial[chunks][2][C]; the
 * last CTA to arrive (ticket) folds them in order, writes save_mean/save_invstd, scale/shift, and
 * updates running stats (momentum, unbiased variance) + num_batches_tracked. */
int ldnn_bn_fwd_stats(const void* y, int M, int C, int chunks, float* partial, unsigned int* ticket,
                      const float* gamma, const float* beta, float* running_mean, float* running_var,
                      long long* num_batches_tracked, float* save_mean, float* save_invstd, float* scale,
                      float* shift, float eps, float momentum, void* stream);
/* Same computation as ldnn_bn_fwd_stats, for callers that cannot allocate device memory and have no
 * workspace to pass: plug-ins / custom ops living inside a host whose allocator owns the whole GPU,
 * and steady-state loops that must not call cudaMalloc/cudaFree (both synchronise the device).
 * The partial sums and the ticket live in a fixed 1 MiB (+256 B) array inside the library, resident on
 * every GPU that loads it until process exit.  Limits: chunks*2*C <= 262144 floats, otherwise the call
 * is refused with an error.  It calls no cudaMalloc/cudaFree and no allocating library routine.
 * The first call on each device looks up the array's device address (one-time set-up, with that
 * device current): make that first call before any stream capture begins.  Every call on a device
 * shares the one array, and the array keeps the previous call's contents (the ticket is cleared in
 * stream order at the start of each call): two calls that overlap on the GPU race with each other —
 * keep them on one stream or order them yourself.  The per-device address cache inside the library is written
 * without a lock: two host threads making the FIRST call on a device at the same moment both look the address up and
 * store the same value, which is harmless.
 * This entry point has no counterpart in the reference. */
int ldnn_bn_fwd_stats_noalloc(const void* y, int M, int C, int chunks, const float* gamma, const float* beta,
                              float* running_mean, float* running_var, long long* num_batches_tracked,
                              float* save_mean, float* save_invstd, float* scale, float* shift, float eps,
                              float momentum, void* stream);
/* eval mode: scale/shift from the running statistics */
int ldnn_bn_eval_coeffs(const float* gamma, const float* beta, const float* running_mean, const float* running_var,
                        float* scale, float* shift, int C, float eps, void* stream);
/* out = relu(y*scale + shift + residual)   (residual may be NULL, relu 0/1) */
int ldnn_bn_fwd_apply(const void* y, const float* scale, const float* shift, const void* residual, void* out, int M,
                      int C, int relu, void* stream);
/* backward pass 1: dy = dout * (out > 0); dbeta = sum dy, dgamma = sum dy*xhat (ticketed like fwd_stats) */
int ldnn_bn_bwd_reduce(const void* dout, const void* out, const void* y, const float* save_mean,
                       const float* save_invstd, float* partial, unsigned int* ticket, int M, int C, int chunks,
                       int relu, float* dgamma, float* dbeta, void* stream);
/* backward pass 2: dy_raw = gamma*invstd*(dy - dbeta/M - xhat*dgamma/M); dres (optional) = dy or += dy */
int ldnn_bn_bwd_apply(const void* dout, const void* out, const void* y, const float* save_mean,
                      const float* save_invstd, const float* gamma, const float* dgamma, const float* dbeta,
                      void* dy_raw, void* dres, int M, int C, int relu, int res_accumulate, void* stream);
/* Fast path ("sums"): the conv epilogue accumulates per-channel sum / sum-of-squares into
 * ldnn_conv_desc.bn_sums, so BN forward is one launch (scale/shift derived per CTA; CTA 0 updates the
 * running statistics) and BN backward is reduce (red.add into sums[2][C]) + apply.  Summation order of
 * the fp32 atomics is not fixed; use the ticketed entry points above for bit-reproducible statistics.
 * training=0 normalises with the running statistics (sums ignored). */
int ldnn_bn_fwd_apply_sums(const void* y, const float* sums /* [2][C] */, const float* gamma, const float* beta,
                           float* running_mean, float* running_var, long long* num_batches_tracked,
                           float* save_mean, float* save_invstd, const void* residual, void* out, int M, int C,
                           int relu, int training, float eps, float momentum, void* stream);
int ldnn_bn_bwd_reduce_sums(const void* dout, const void* out, const void* y, const float* save_mean,
                            const float* save_invstd, float* sums /* [2][C], zeroed by the caller */, int M, int C,
                            int chunks, int relu, void* stream);
int ldnn_bn_bwd_apply_sums(const void* dout, const void* out, const void* y, const float* save_mean,
                           const float* save_invstd, const float* gamma, const float* sums, void* dy_raw, void* dres,
                           float* dgamma, float* dbeta, int M, int C, int relu, int res_accumulate, void* stream);
/* 3x3/s2/p1-style max pool on NHWC bf16 (ResNet-50 stem).  Forward also records, per output element, the
 * window position of its (first) maximum in idx (uint8 [N,P,Q,C], may be NULL for inference); backward
 * gathers dy through that map. */
int ldnn_maxpool_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, int k, int stride, int pad,
                     void* stream);
int ldnn_maxpool_bwd(const void* idx, const void* dy, void* dx, int N, int H, int W, int C, int k, int stride,
                     int pad, void* stream);

/* fp32-storage twins (suffix _f32): the same kernels instantiated for float activations — the BN / pool / stem side
 * of the TF32 path (ldnn_conv_desc.dtype = 1).  Same arguments; every `void*` activation tensor is float. */
int ldnn_im2col_stem_f32(const float* x, void* col, int N, int C, int H, int W, int R, int stride, int pad, int Kpad,
                         void* stream);
int ldnn_bn_fwd_stats_f32(const void* y, int M, int C, int chunks, float* partial, unsigned int* ticket,
                          const float* gamma, const float* beta, float* running_mean, float* running_var,
                          long long* num_batches_tracked, float* save_mean, float* save_invstd, float* scale,
                          float* shift, float eps, float momentum, void* stream);
int ldnn_bn_fwd_apply_f32(const void* y, const float* scale, const float* shift, const void* residual, void* out,
                          int M, int C, int relu, void* stream);
int ldnn_bn_bwd_reduce_f32(const void* dout, const void* out, const void* y, const float* save_mean,
                           const float* save_invstd, float* partial, unsigned int* ticket, int M, int C, int chunks,
                           int relu, float* dgamma, float* dbeta, void* stream);
int ldnn_bn_bwd_apply_f32(const void* dout, const void* out, const void* y, const float* save_mean,
                          const float* save_invstd, const float* gamma, const float* dgamma, const float* dbeta,
                          void* dy_raw, void* dres, int M, int C, int relu, int res_accumulate, void* stream);
int ldnn_bn_fwd_apply_sums_f32(const void* y, const float* sums, const float* gamma, const float* beta,
                               float* running_mean, float* running_var, long long* num_batches_tracked,
                               float* save_mean, float* save_invstd, const void* residual, void* out, int M, int C,
                               int relu, int training, float eps, float momentum, void* stream);
int ldnn_bn_bwd_reduce_sums_f32(const void* dout, const void* out, const void* y, const float* save_mean,
                                const float* save_invstd, float* sums, int M, int C, int chunks, int relu,
                                void* stream);
int ldnn_bn_bwd_apply_sums_f32(const void* dout, const void* out, const void* y, const float* save_mean,
                               const float* save_invstd, const float* gamma, const float* sums, void* dy_raw,
                               void* dres, float* dgamma, float* dbeta, int M, int C, int relu, int res_accumulate,
                               void* stream);
int ldnn_maxpool_fwd_f32(const void* x, void* y, void* idx, int N, int H, int W, int C, int k, int stride, int pad,
                         void* stream);
int ldnn_maxpool_bwd_f32(const void* idx, const void* dy, void* dx, int N, int H, int W, int C, int k, int stride,
                         int pad, void* stream);

/* ------------------------------------------------------------------ head (K-HEAD)
 * global-avg-pool + linear + softmax cross-entropy (+ correct count, dlogits, d_act, fc grads)
 * (model.py:99-100,108-110; main.py:52; trainer.py:207,212-216). */
int ldnn_head_fwd_bwd(const void* act /* bf16 [N][HW][C] */, int N, int HW, int C, int classes, const float* fc_w,
                      const float* fc_b, const long long* labels, float* pooled /* [N][C] */,
                      float* logits /* [N][classes] */, float* dlogits, float* loss_n /* [N] */,
                      float* correct_n /* [N] */, long long* pred /* [N] */, void* dact /* bf16 like act */,
                      float* dfc_w, float* dfc_b, float* loss_out /* mean loss */, float* correct_out /* count */,
                      int training, void* stream);

/* float activations (TF32 path) */
int ldnn_head_fwd_bwd_f32(const void* act, int N, int HW, int C, int classes, const float* fc_w, const float* fc_b,
                          const long long* labels, float* pooled, float* logits, float* dlogits, float* loss_n,
                          float* correct_n, long long* pred, void* dact, float* dfc_w, float* dfc_b, float* loss_out,
                          float* correct_out, int training, void* stream);

/* ------------------------------------------------------------------ optimizer (K-OPT)
 * Adam / SGD-momentum on flat fp32 slabs, also emits the bf16 compute copy
 * (trainer.py:205,210,218; main.py:53-54).  The same device functions sit in the tails of the
 * exchange kernels below. */
typedef struct {
  int kind;            /* 0 adam, 1 sgd-momentum */
  float lr, beta1, beta2, eps, momentum, weight_decay;
  float bc1, bc2;      /* adam bias corrections 1-beta^t */
  float grad_scale;    /* multiplies the (summed) gradient first */
} ldnn_optim;
int ldnn_optim_step(float* params, const float* grads, float* m, float* v, void* shadow_bf16, long long n,
                    const ldnn_optim* o, void* stream);
int ldnn_cast_bf16(const float* src, void* dst, long long n, void* stream);

/* ------------------------------------------------------------------ symmetric memory + exchange
 * One process per GPU; buffers are cudaMalloc'd here and mapped into every peer through CUDA IPC
 * (replaces torch.distributed/mpi4py transport for the gradient exchange:
 * BalAR/communication.py:4-31, BalR/communication.py:5-62, BalDR/communication.py:5-77). */
#define LDNN_MAX_RANKS 8
#define LDNN_IPC_HANDLE_BYTES 64
int ldnn_symm_alloc(size_t bytes, void** ptr, unsigned char* handle_out /* 64 B */);
int ldnn_symm_open(const unsigned char* handle, void** peer_ptr);
int ldnn_symm_close(void* peer_ptr);
int ldnn_symm_free(void* ptr);
int ldnn_enable_peer_access(int my_device, int peer_device);

typedef struct {
  int rank, world;
  int emulate;                         /* 1: one launch plays every rank (single-GPU tests), blockIdx.y = rank */
  int blocks;                          /* CTAs per rank */
  unsigned int epoch;                  /* strictly increasing per call; flags are never reset */
  unsigned long long timeout_cycles;   /* spin-wait watchdog; 0 = default */
  void* signal[LDNN_MAX_RANKS];        /* per-rank signal pad (uint32[LDNN_SIGNAL_WORDS]) */
  int* error_flag;                     /* set to nonzero on watchdog expiry (local) */
  int debug_delay_block;               /* tests: CTA (this - 1) of every rank idles debug_delay_cycles before it
                                          pushes (0 = off); exposes cross-CTA ordering assumptions            */
  unsigned int debug_delay_cycles;
} ldnn_comm;
#define LDNN_SIGNAL_WORDS 65536
/* pad layout (uint32 words): [0, 4096) barrier slots [block][rank] | [4096, 65472) ring chain flags |
 * [65472, 65536) user flag words, never touched by the kernels (the trainer's straggler announcements) */
#define LDNN_SIGNAL_USER_BASE (LDNN_SIGNAL_WORDS - 64)

/* mode bits for the exchange kernels */
#define LDNN_X_UPDATE 1   /* run the optimizer on the reduced gradient (fused tail) */
#define LDNN_X_SELFMIX 2  /* result_r = a*x_r + sum_p b_p x_p (self-weighted form), needs `recv` slab */

typedef struct {
  void* grads[LDNN_MAX_RANKS];   /* fp32 flat, symmetric: the data being exchanged */
  void* params[LDNN_MAX_RANKS];  /* fp32 flat, symmetric (UPDATE: updated slice is pushed to all peers) */
  void* recv[LDNN_MAX_RANKS];    /* fp32 flat, symmetric scratch (SELFMIX / gossip / ring staging) */
  float* m; float* v;            /* local optimizer moments */
  void* shadow;                  /* local bf16 copy of params (may be NULL) */
  long long n;                   /* elements */
  float weight[LDNN_MAX_RANKS];  /* b_p */
  float self_coef;               /* a */
  int mode;
  ldnn_optim opt;
  /* flag-free small/medium-payload rings (ldnn_ring_ll): per-rank symmetric staging pad of 3 phases x ll_phase4
   * float4 slots, every word initialised to the sentinel 0x80000000 (-0.0f) by the allocator */
  void* ll[LDNN_MAX_RANKS];
  long long ll_phase4;           /* slots per phase */
  long long ll_clear4;           /* slots the PREVIOUS ldnn_ring_ll call used (re-armed by this call) */
  unsigned int ll_epoch;         /* number of earlier ldnn_ring_ll calls on this pad (same on every rank) */
} ldnn_exchange;

/* K-AR: two-shot all-reduce (pull-reduce owned slice, fused tail, push to peers) */
int ldnn_allreduce_fused(const ldnn_comm* c, const ldnn_exchange* x, void* stream);
/* K-AR (NVLS): the same exchange with the reduce done inside NVSwitch: `mc_data` / `mc_params` are the MULTICAST
 * mappings (one address = the same offset of every rank's slab) of x->grads / x->params, offset like them.
 * multimem.ld_reduce on the owned slice, fused tail, multimem.st to all ranks.  Weights must be uniform, SELFMIX
 * is not available.  With c->emulate the multimem instructions are replaced by peer loads/stores (mc_* ignored).
 * Run on 2 / 4 / 8 B200 in round 2 (8 GPUs: 827 GB/s bus bandwidth at 512 MB vs 643 for the two-shot kernel). */
int ldnn_allreduce_nvls(const ldnn_comm* c, const ldnn_exchange* x, void* mc_data, void* mc_params, void* stream);
/* K-RING / K-DRING: reduce-scatter + all-gather around 1 or 2 (counter-rotating) rings */
int ldnn_ring_fused(const ldnn_comm* c, const ldnn_exchange* x, int num_rings, void* stream);
/* how ldnn_ring_fused cuts one rank's chunk into hand-off chains: J chains of sub4 float4 (host arithmetic only) */
int ldnn_ring_plan(long long n, int world, int num_rings, int blocks, int chains, int* J, long long* sub4);
/* K-RING / K-DRING for small and medium payloads: the same reduce-scatter + all-gather schedule with no flags, no
 * fences and no kernel-wide barrier — the payload travels through x->ll and announces itself (sentinel polling), one
 * dependency chain per element.  ldnn_ring_ll_slots = float4 slots one call uses per phase (pass it as the next call's
 * ll_clear4).  Hop cost ~ one NVLink store latency instead of store + release + flag. */
long long ldnn_ring_ll_slots(long long n, int world, int num_rings);
int ldnn_ring_ll(const ldnn_comm* c, const ldnn_exchange* x, int num_rings, void* stream);
/* compat gossip: out = c_self*x + sum_d c_d * x[(r-d) mod ws], d in offsets (1 or 2 entries) */
int ldnn_gossip_pull(const ldnn_comm* c, void* const* bufs /*symmetric fp32*/, void* const* recv /*symmetric*/,
                     long long n, float c_self, int n_off, const int* offsets, const float* coefs, void* stream);
/* K-BCAST: everyone copies `bytes` from rank `src`'s buffer */
int ldnn_broadcast(const ldnn_comm* c, void* const* bufs, size_t bytes, int src, void* stream);
/* K-MET: every rank's `count` floats land in slot [rank] of every peer's table */
int ldnn_allgather_small(const ldnn_comm* c, const float* mine, void* const* tables, int count, void* stream);
/* K-BAR: device-side barrier across ranks */
int ldnn_barrier(const ldnn_comm* c, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* LDNN_B200_H_ */
