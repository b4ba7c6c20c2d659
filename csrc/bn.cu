// Batch-norm (training + eval) fused with residual add and ReLU, on NHWC bf16  (K-BN-F / K-BN-B),
// the 3-channel stem's fused cast+im2col, and NHWC max-pool.
//
// Replaces ATen/cuDNN BatchNorm2d + F.relu + `out += shortcut(x)`
// (Balanced All-Reduce/model.py:56,58,64,68-71,79-80).  Statistics are reduced in two levels
// without atomics on data: every CTA writes per-channel partials for its pixel chunk, takes a
// ticket, and the last CTA to arrive folds the partials in a fixed order (deterministic), updates
// the running stats and emits scale/shift — so a BN forward is two launches (stats, apply) and a
// BN backward is two launches (reduce, apply) regardless of channel count.
#include "../include/ldnn_b200.h"
#include "common.cuh"
#include <atomic>

namespace ldnn {
extern std::atomic<unsigned long long> g_launches;

constexpr int kBnThreads = 256;

// Thread layout shared by the reduction kernels: a CTA owns a slab of 64 channels (blockIdx.x) and a
// chunk of rows (blockIdx.y); thread = 8 consecutive channels (one 16-byte load) x 32 row lanes, so
// layers with many channels and few pixels (layer3/4) still spread over many CTAs.
constexpr int kSlab = 64;
struct Lanes {
  int cg, pl, c0;
  __device__ Lanes() {
    cg = threadIdx.x & 7;
    pl = threadIdx.x >> 3;
    c0 = blockIdx.x * kSlab + cg * 8;
  }
};
constexpr int kRowLanes = kBnThreads / 8;   // 32

// Block-level fold of per-thread (a[8], b[8]) over the 32 row lanes -> partial[chunk][2][C] (this slab).
__device__ __forceinline__ void fold_to_partial(const Lanes& L, float (&a)[8], float (&b)[8], float* smem,
                                                float* partial, int C) {
  // smem layout: [lanes][2][64]
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    smem[(L.pl * 2 + 0) * kSlab + L.cg * 8 + i] = a[i];
    smem[(L.pl * 2 + 1) * kSlab + L.cg * 8 + i] = b[i];
  }
  __syncthreads();
  if (threadIdx.x < 2 * kSlab) {
    const int which = threadIdx.x / kSlab, c = threadIdx.x % kSlab;
    float s = 0.f;
#pragma unroll 8
    for (int l = 0; l < kRowLanes; ++l) s += smem[(l * 2 + which) * kSlab + c];
    partial[((size_t)blockIdx.y * 2 + which) * C + blockIdx.x * kSlab + c] = s;
  }
}

// "Last CTA" election.  One fence per CTA (thread 0, after the CTA barrier) publishes the whole CTA's
// partials before the ticket; the elected CTA fences once more before reading the others' partials.
__device__ __forceinline__ bool take_ticket_last(unsigned int* ticket) {
  __shared__ unsigned int s_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned int old = atomicAdd(ticket, 1u);
    s_last = (old == gridDim.x * gridDim.y - 1) ? 1u : 0u;
    if (s_last) { *ticket = 0; __threadfence(); }
  }
  __syncthreads();
  return s_last != 0;
}

// Last-CTA fold: out[0..2C) = sum over chunks of partial[chunk][0..2C), into smem.
// 2C columns x S slices of the chunk range run in parallel; every slice sums its chunks in order
// and the slices are combined in order, so the result does not depend on scheduling.
__device__ __forceinline__ void fold_partials(const float* __restrict__ partial, float* smem, int C, int chunks) {
  const int cols = 2 * C;
  const int S = max(1, kBnThreads / cols);              // slices (cols < 256 only)
  float* stage = smem + cols;                           // [S][cols] staging behind the result
  __syncthreads();
  for (int idx = threadIdx.x; idx < cols * S; idx += kBnThreads) {
    const int col = idx % cols, sl = idx / cols;
    const int k0 = (int)((long long)chunks * sl / S), k1 = (int)((long long)chunks * (sl + 1) / S);
    float acc = 0.f;
    int k = k0;
    for (; k + 8 <= k1; k += 8) {
      float t[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) t[j] = __ldcg(&partial[(size_t)(k + j) * cols + col]);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc += t[j];
    }
    for (; k < k1; ++k) acc += __ldcg(&partial[(size_t)k * cols + col]);
    stage[sl * cols + col] = acc;
  }
  __syncthreads();
  for (int col = threadIdx.x; col < cols; col += kBnThreads) {
    float acc = 0.f;
    for (int sl = 0; sl < S; ++sl) acc += stage[sl * cols + col];
    smem[col] = acc;
  }
  __syncthreads();
}

// ------------------------------------------------------------------ forward stats
template <typename T>
__global__ void __launch_bounds__(kBnThreads)
bn_stats_kernel(const T* __restrict__ y, float* __restrict__ partial, unsigned int* ticket, int M, int C,
                int rows_per_chunk, const float* __restrict__ gamma, const float* __restrict__ beta,
                float* running_mean, float* running_var, long long* nbt, float* save_mean, float* save_invstd,
                float* scale, float* shift, float eps, float momentum) {
  extern __shared__ float smem[];
  Lanes L;
  float s[8], q[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { s[i] = 0.f; q[i] = 0.f; }
  const int r0 = blockIdx.y * rows_per_chunk, r1 = min(M, r0 + rows_per_chunk);
#pragma unroll 4
  for (int r = r0 + L.pl; r < r1; r += kRowLanes) {
    float f[8];
    load8(y + (size_t)r * C + L.c0, f);
#pragma unroll
    for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] += f[i] * f[i]; }
  }
  fold_to_partial(L, s, q, smem, partial, C);
  if (!take_ticket_last(ticket)) return;
  // ---- last CTA: finalize (all 256 threads fold the partials; fixed order => deterministic) ----
  fold_partials(partial, smem, C, gridDim.y);
  const float inv_m = 1.f / (float)M;
  for (int c = threadIdx.x; c < C; c += kBnThreads) {
    const float sum = smem[c], sq = smem[C + c];
    const float mean = sum * inv_m;
    const float var = fmaxf(sq * inv_m - mean * mean, 0.f);
    const float invstd = rsqrtf(var + eps);
    save_mean[c] = mean;
    save_invstd[c] = invstd;
    const float g = gamma[c];
    scale[c] = g * invstd;
    shift[c] = beta[c] - mean * g * invstd;
    const float unbiased = (M > 1) ? var * (float)M / (float)(M - 1) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
  }
  if (threadIdx.x == 0 && nbt) *nbt += 1;
}

__global__ void bn_eval_coeffs_kernel(const float* gamma, const float* beta, const float* rm, const float* rv,
                                      float* scale, float* shift, int C, float eps) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < C) {
    const float is = rsqrtf(rv[c] + eps);
    scale[c] = gamma[c] * is;
    shift[c] = beta[c] - rm[c] * gamma[c] * is;
  }
}

// ------------------------------------------------------------------ forward apply
template <typename T>
__global__ void __launch_bounds__(256)
bn_apply_kernel(const T* __restrict__ y, const float* __restrict__ scale, const float* __restrict__ shift,
                const T* __restrict__ res, T* __restrict__ out, long long total8, int C,
                int relu) {
  const int c8 = C >> 3;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total8;
       i += (long long)gridDim.x * blockDim.x) {
    const int cg = (int)(i % c8);
    float f[8];
    load8(y + 8 * (size_t)i, f);
    const float4 s0 = reinterpret_cast<const float4*>(scale)[cg * 2], s1 = reinterpret_cast<const float4*>(scale)[cg * 2 + 1];
    const float4 h0 = reinterpret_cast<const float4*>(shift)[cg * 2], h1 = reinterpret_cast<const float4*>(shift)[cg * 2 + 1];
    const float sc[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
    const float sh[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] = f[k] * sc[k] + sh[k];
    if (res) {
      float r[8];
      load8(res + 8 * (size_t)i, r);
#pragma unroll
      for (int k = 0; k < 8; ++k) f[k] += r[k];
    }
    if (relu) {
#pragma unroll
      for (int k = 0; k < 8; ++k) f[k] = fmaxf(f[k], 0.f);
    }
    store8(out + 8 * (size_t)i, f);
  }
}

// ------------------------------------------------------------------ backward reduce
template <typename T>
__global__ void __launch_bounds__(kBnThreads)
bn_bwd_reduce_kernel(const T* __restrict__ dout, const T* __restrict__ outp,
                     const T* __restrict__ y, const float* __restrict__ mean,
                     const float* __restrict__ invstd, float* __restrict__ partial, unsigned int* ticket, int M, int C,
                     int rows_per_chunk, int relu, float* dgamma, float* dbeta) {
  extern __shared__ float smem[];
  Lanes L;
  float a[8], b[8], mu[8], is[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { a[i] = 0.f; b[i] = 0.f; mu[i] = mean[L.c0 + i]; is[i] = invstd[L.c0 + i]; }
  const int r0 = blockIdx.y * rows_per_chunk, r1 = min(M, r0 + rows_per_chunk);
#pragma unroll 2
  for (int r = r0 + L.pl; r < r1; r += kRowLanes) {
    const size_t off = (size_t)r * C + L.c0;
    float g[8], yy[8];
    load8(dout + off, g);
    load8(y + off, yy);
    if (relu) {
      float o[8];
      load8(outp + off, o);
#pragma unroll
      for (int i = 0; i < 8; ++i) g[i] = (o[i] > 0.f) ? g[i] : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] += g[i]; b[i] += g[i] * (yy[i] - mu[i]) * is[i]; }
  }
  fold_to_partial(L, a, b, smem, partial, C);
  if (!take_ticket_last(ticket)) return;
  fold_partials(partial, smem, C, gridDim.y);
  for (int c = threadIdx.x; c < C; c += kBnThreads) {
    dbeta[c] = smem[c];
    dgamma[c] = smem[C + c];
  }
}

// ------------------------------------------------------------------ backward apply
template <typename T>
__global__ void __launch_bounds__(256)
bn_bwd_apply_kernel(const T* __restrict__ dout, const T* __restrict__ outp,
                    const T* __restrict__ y, const float* __restrict__ mean,
                    const float* __restrict__ invstd, const float* __restrict__ gamma,
                    const float* __restrict__ dgamma, const float* __restrict__ dbeta,
                    T* __restrict__ dy_raw, T* __restrict__ dres, long long total8, int C,
                    int relu, int res_accumulate, float inv_m) {
  const int c8 = C >> 3;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total8;
       i += (long long)gridDim.x * blockDim.x) {
    const int c0 = (int)(i % c8) * 8;
    float g[8], yy[8];
    load8(dout + 8 * (size_t)i, g);
    load8(y + 8 * (size_t)i, yy);
    if (relu) {
      float o[8];
      load8(outp + 8 * (size_t)i, o);
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] = (o[k] > 0.f) ? g[k] : 0.f;
    }
    if (dres) {
      if (res_accumulate) {
        float r[8];
        load8(dres + 8 * (size_t)i, r);
        float t[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) t[k] = r[k] + g[k];
        store8(dres + 8 * (size_t)i, t);
      } else {
        store8(dres + 8 * (size_t)i, g);
      }
    }
    float o8[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int c = c0 + k;
      const float is = invstd[c];
      const float xh = (yy[k] - mean[c]) * is;
      o8[k] = gamma[c] * is * (g[k] - dbeta[c] * inv_m - xh * dgamma[c] * inv_m);
    }
    store8(dy_raw + 8 * (size_t)i, o8);
  }
}

// ------------------------------------------------------------------ "sums" variants (fast path)
// The conv epilogue already accumulated per-channel sum / sum-of-squares (conv_gemm.cuh), so the BN
// forward is ONE kernel: every CTA derives scale/shift for all channels into smem (CTA 0 also writes
// save_mean/save_invstd and the running statistics), then streams normalize + residual + ReLU.
// Backward likewise: `reduce` accumulates (sum dy, sum dy*xhat) with red.add (no ticket chain) and
// `apply` derives dgamma/dbeta + the per-channel coefficients in its prologue.
template <typename T>
__global__ void __launch_bounds__(256)
bn_apply_sums_kernel(const T* __restrict__ y, const float* __restrict__ sums,
                     const float* __restrict__ gamma, const float* __restrict__ beta, float* running_mean,
                     float* running_var, long long* nbt, float* save_mean, float* save_invstd,
                     const T* __restrict__ res, T* __restrict__ out, int rows_per_chunk,
                     int C, int M, int relu, int training, float eps, float momentum) {
  pdl_wait();
  pdl_launch_dependents();
  // CTA = (64-channel slab, row chunk): only this slab's 64 coefficients are derived in the prologue
  __shared__ float s_scale[kSlab], s_shift[kSlab];
  const float inv_m = 1.f / (float)M;
  if (threadIdx.x < kSlab) {
    const int c = blockIdx.x * kSlab + threadIdx.x;
    float mean, var;
    if (training) {
      mean = sums[c] * inv_m;
      var = fmaxf(sums[C + c] * inv_m - mean * mean, 0.f);
    } else {
      mean = running_mean[c];
      var = running_var[c];
    }
    const float invstd = rsqrtf(var + eps);
    const float g = gamma[c];
    s_scale[threadIdx.x] = g * invstd;
    s_shift[threadIdx.x] = beta[c] - mean * g * invstd;
    if (training && blockIdx.y == 0) {
      save_mean[c] = mean;
      save_invstd[c] = invstd;
      const float unbiased = (M > 1) ? var * (float)M / (float)(M - 1) : var;
      running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
      running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
    }
  }
  if (training && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0 && nbt) *nbt += 1;
  __syncthreads();
  Lanes L;
  float sc[8], sh[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) { sc[k] = s_scale[L.cg * 8 + k]; sh[k] = s_shift[L.cg * 8 + k]; }
  const int r0 = blockIdx.y * rows_per_chunk, r1 = min(M, r0 + rows_per_chunk);
#pragma unroll 4
  for (int r = r0 + L.pl; r < r1; r += kRowLanes) {
    const size_t off = (size_t)r * C + L.c0;
    float f[8];
    load8(y + off, f);
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] = f[k] * sc[k] + sh[k];
    if (res) {
      float rr[8];
      load8(res + off, rr);
#pragma unroll
      for (int k = 0; k < 8; ++k) f[k] += rr[k];
    }
    if (relu) {
#pragma unroll
      for (int k = 0; k < 8; ++k) f[k] = fmaxf(f[k], 0.f);
    }
    store8(out + off, f);
  }
}

template <typename T>
__global__ void __launch_bounds__(kBnThreads)
bn_bwd_reduce_atomic_kernel(const T* __restrict__ dout, const T* __restrict__ outp,
                            const T* __restrict__ y, const float* __restrict__ mean,
                            const float* __restrict__ invstd, float* __restrict__ sums, int M, int C,
                            int rows_per_chunk, int relu) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float smem[kRowLanes * 2 * kSlab];
  Lanes L;
  float a[8], b[8], mu[8], is[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { a[i] = 0.f; b[i] = 0.f; mu[i] = mean[L.c0 + i]; is[i] = invstd[L.c0 + i]; }
  const int r0 = blockIdx.y * rows_per_chunk, r1 = min(M, r0 + rows_per_chunk);
#pragma unroll 2
  for (int r = r0 + L.pl; r < r1; r += kRowLanes) {
    const size_t off = (size_t)r * C + L.c0;
    float g[8], yy[8];
    load8(dout + off, g);
    load8(y + off, yy);
    if (relu) {
      float o[8];
      load8(outp + off, o);
#pragma unroll
      for (int i = 0; i < 8; ++i) g[i] = (o[i] > 0.f) ? g[i] : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] += g[i]; b[i] += g[i] * (yy[i] - mu[i]) * is[i]; }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    smem[(L.pl * 2 + 0) * kSlab + L.cg * 8 + i] = a[i];
    smem[(L.pl * 2 + 1) * kSlab + L.cg * 8 + i] = b[i];
  }
  __syncthreads();
  if (threadIdx.x < 2 * kSlab) {
    const int which = threadIdx.x / kSlab, c = threadIdx.x % kSlab;
    float t = 0.f;
#pragma unroll 8
    for (int l = 0; l < kRowLanes; ++l) t += smem[(l * 2 + which) * kSlab + c];
    atomicAdd(&sums[(size_t)which * C + blockIdx.x * kSlab + c], t);
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
bn_bwd_apply_sums_kernel(const T* __restrict__ dout, const T* __restrict__ outp,
                         const T* __restrict__ y, const float* __restrict__ mean,
                         const float* __restrict__ invstd, const float* __restrict__ gamma,
                         const float* __restrict__ sums, T* __restrict__ dy_raw,
                         T* __restrict__ dres, float* dgamma, float* dbeta, int rows_per_chunk, int M,
                         int C, int relu, int res_accumulate, float inv_m) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float s_co[5][kSlab];     // mean | invstd | gamma*invstd | dbeta/M | dgamma/M  (this slab)
  if (threadIdx.x < kSlab) {
    const int c = blockIdx.x * kSlab + threadIdx.x;
    const float is = invstd[c], db = sums[c], dg = sums[C + c];
    s_co[0][threadIdx.x] = mean[c]; s_co[1][threadIdx.x] = is; s_co[2][threadIdx.x] = gamma[c] * is;
    s_co[3][threadIdx.x] = db * inv_m; s_co[4][threadIdx.x] = dg * inv_m;
    if (blockIdx.y == 0) { dbeta[c] = db; dgamma[c] = dg; }
  }
  __syncthreads();
  Lanes L;
  float mu[8], is[8], kk[8], kb[8], kg[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int j = L.cg * 8 + k;
    mu[k] = s_co[0][j]; is[k] = s_co[1][j]; kk[k] = s_co[2][j]; kb[k] = s_co[3][j]; kg[k] = s_co[4][j];
  }
  const int r0 = blockIdx.y * rows_per_chunk, r1 = min(M, r0 + rows_per_chunk);
#pragma unroll 2
  for (int r = r0 + L.pl; r < r1; r += kRowLanes) {
    const size_t off = (size_t)r * C + L.c0;
    float g[8], yy[8];
    load8(dout + off, g);
    load8(y + off, yy);
    if (relu) {
      float o[8];
      load8(outp + off, o);
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] = (o[k] > 0.f) ? g[k] : 0.f;
    }
    if (dres) {
      if (res_accumulate) {
        float rr[8], t[8];
        load8(dres + off, rr);
#pragma unroll
        for (int k = 0; k < 8; ++k) t[k] = rr[k] + g[k];
        store8(dres + off, t);
      } else {
        store8(dres + off, g);
      }
    }
    float o8[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) o8[k] = kk[k] * (g[k] - kb[k] - (yy[k] - mu[k]) * is[k] * kg[k]);
    store8(dy_raw + off, o8);
  }
}

// ------------------------------------------------------------------ stem im2col (fp32 NCHW -> bf16 rows)
template <typename T>
__global__ void __launch_bounds__(256)
im2col_stem_kernel(const float* __restrict__ x, T* __restrict__ col, int N, int C, int H, int W, int R,
                   int stride, int pad, int P, int Q, int Kpad) {
  pdl_wait();
  pdl_launch_dependents();
  // one thread per (pixel, 8 consecutive k)
  const int k8 = Kpad >> 3;
  const long long total = (long long)N * P * Q * k8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int kg = (int)(i % k8);
    const long long pix = i / k8;
    const int q = (int)(pix % Q), pp = (int)((pix / Q) % P), n = (int)(pix / ((long long)P * Q));
    float f[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int k = kg * 8 + j;
      float v = 0.f;
      if (k < R * R * C) {
        const int c = k % C, tap = k / C, r = tap / R, s = tap % R;
        const int h = pp * stride - pad + r, w = q * stride - pad + s;
        if (h >= 0 && h < H && w >= 0 && w < W) v = x[(((size_t)n * C + c) * H + h) * W + w];
      }
      f[j] = v;
    }
    store8(col + 8 * (size_t)i, f);
  }
}

// ------------------------------------------------------------------ max pool NHWC bf16
// Forward records the window position of the (first) maximum per output element, so backward is a
// gather: every input pixel looks at the <= ceil(k/stride)^2 windows that cover it and takes the
// gradient of those whose recorded position is this pixel (same tie-break as torch: first maximum).
template <typename T>
__global__ void maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y,
                                   uint8_t* __restrict__ idx, int N, int H, int W, int C, int k, int stride, int pad,
                                   int P, int Q) {
  pdl_wait();
  pdl_launch_dependents();
  const int c8 = C >> 3;
  const long long total = (long long)N * P * Q * c8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int cg = (int)(i % c8);
    const long long pix = i / c8;
    const int q = (int)(pix % Q), p = (int)((pix / Q) % P), n = (int)(pix / ((long long)P * Q));
    float m[8];
    uint8_t am[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { m[j] = -INFINITY; am[j] = 0; }
    for (int r = 0; r < k; ++r)
      for (int s = 0; s < k; ++s) {
        const int h = p * stride - pad + r, w = q * stride - pad + s;
        if (h < 0 || h >= H || w < 0 || w >= W) continue;
        float f[8];
        load8(x + (((size_t)n * H + h) * W + w) * C + cg * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (f[j] > m[j]) { m[j] = f[j]; am[j] = (uint8_t)(r * k + s); }
      }
    store8(y + 8 * (size_t)i, m);
    if (idx) {
      uint2 packed;
      packed.x = am[0] | (am[1] << 8) | (am[2] << 16) | ((uint32_t)am[3] << 24);
      packed.y = am[4] | (am[5] << 8) | (am[6] << 16) | ((uint32_t)am[7] << 24);
      reinterpret_cast<uint2*>(idx)[i] = packed;
    }
  }
}

template <typename T>
__global__ void maxpool_bwd_kernel(const uint8_t* __restrict__ idx, const T* __restrict__ dy,
                                   T* __restrict__ dx, int N, int H, int W, int C, int k, int stride,
                                   int pad, int P, int Q) {
  pdl_wait();
  pdl_launch_dependents();
  const int c8 = C >> 3;
  const long long total = (long long)N * H * W * c8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int cg = (int)(i % c8);
    const long long pix = i / c8;
    const int w = (int)(pix % W), h = (int)((pix / W) % H), n = (int)(pix / ((long long)H * W));
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    const int p_lo = max(0, (h + pad - k + stride) / stride), p_hi = min(P - 1, (h + pad) / stride);
    const int q_lo = max(0, (w + pad - k + stride) / stride), q_hi = min(Q - 1, (w + pad) / stride);
    for (int p = p_lo; p <= p_hi; ++p)
      for (int q = q_lo; q <= q_hi; ++q) {
        const int r = h - (p * stride - pad), s = w - (q * stride - pad);
        if (r < 0 || r >= k || s < 0 || s >= k) continue;
        const uint32_t me = (uint32_t)(r * k + s);
        const size_t o = (((size_t)n * P + p) * Q + q) * c8 + cg;
        const uint2 packed = reinterpret_cast<const uint2*>(idx)[o];
        float g[8];
        load8(dy + 8 * (size_t)o, g);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t a = ((j < 4 ? packed.x : packed.y) >> (8 * (j & 3))) & 0xffu;
          if (a == me) acc[j] += g[j];
        }
      }
    store8(dx + 8 * (size_t)i, acc);
  }
}

static size_t bn_smem_bytes(int C, int lanes) {
  const size_t fold = (size_t)lanes * 2 * kSlab;                               // per-lane partials of one slab
  const size_t fin = (size_t)2 * C + std::max<size_t>(kBnThreads, (size_t)2 * C);  // result + slice staging
  return std::max(fold, fin) * sizeof(float);
}

static int grid_for(long long work, int threads, int max_blocks = 148 * 8) {
  long long b = (work + threads - 1) / threads;
  return (int)std::max(1LL, std::min<long long>(b, max_blocks));
}

}  // namespace ldnn

using namespace ldnn;
#define COUNT_LAUNCH() g_launches.fetch_add(1, std::memory_order_relaxed)

template <typename T>
static int bn_fwd_stats_impl(const void* y, int M, int C, int chunks, float* partial, unsigned int* ticket,
                                 const float* gamma, const float* beta, float* running_mean, float* running_var,
                                 long long* nbt, float* save_mean, float* save_invstd, float* scale, float* shift,
                                 float eps, float momentum, void* stream) {
  LDNN_ARG(C % kSlab == 0 && C <= 2048, "bn: C=%d unsupported (need C%%64==0, C<=2048)", C);
  LDNN_ARG(chunks >= 1, "bn: chunks");
  const int rows = (M + chunks - 1) / chunks;
  const int lanes = kRowLanes;
  const size_t smem = bn_smem_bytes(C, lanes);
  bn_stats_kernel<T><<<dim3(C / kSlab, chunks), kBnThreads, smem, (cudaStream_t)stream>>>(
      (const T*)y, partial, ticket, M, C, rows, gamma, beta, running_mean, running_var, nbt, save_mean,
      save_invstd, scale, shift, eps, momentum);
  COUNT_LAUNCH();
  LDNN_CUDA(cudaGetLastError());
  return LDNN_OK;
}
extern "C" int ldnn_bn_fwd_stats(const void* y, int M, int C, int chunks, float* partial, unsigned int* ticket,
                                 const float* gamma, const float* beta, float* running_mean, float* running_var,
                                 long long* nbt, float* save_mean, float* save_invstd, float* scale, float* shift,
                                 float eps, float momentum, void* stream) {
  return bn_fwd_stats_impl<__nv_bfloat16>(y, M, C, chunks, partial, ticket, gamma, beta, running_mean, running_var, nbt, save_mean, save_invstd, scale, shift, eps, momentum, stream);
}
extern "C" int ldnn_bn_fwd_stats_f32(const void* y, int M, int C, int chunks, float* partial, unsigned int* ticket,
                                 const float* gamma, const float* beta, float* running_mean, float* running_var,
                                 long long* nbt, float* save_mean, float* save_invstd, float* scale, float* shift,
                                 float eps, float momentum, void* stream) {
  return bn_fwd_stats_impl<float>(y, M, C, chunks, partial, ticket, gamma, beta, running_mean, running_var, nbt, save_mean, save_invstd, scale, shift, eps, momentum, stream);
}

extern "C" int ldnn_bn_eval_coeffs(const float* gamma, const float* beta, const float* rm, const float* rv,
                                   float* scale, float* shift, int C, float eps, void* stream) {
  bn_eval_coeffs_kernel<<<(C + 255) / 256, 256, 0, (cudaStream_t)stream>>>(gamma, beta, rm, rv, scale, shift, C, eps);
  COUNT_LAUNCH();
  LDNN_CUDA(cudaGetLastError());
  return LDNN_OK;
}

template <typename T>
static int bn_fwd_apply_impl(const void* y, const float* scale, const float* shift, const void* residual,
                                 void* out, int M, int C, int relu, void* stream) {
  LDNN_ARG(C % 8 == 0, "bn: C %% 8");
  const long long total8 = (long long)M * C / 8;
  bn_apply_kernel<T><<<grid_for(total8, 256), 256, 0, (cudaStream_t)stream>>>(
      (const T*)y, scale, shift, (const T*)residual, (T*)out, total8, C, relu);
  COUNT_LAUNCH();
  LDNN_CUDA(cudaGetLastError());
  return LDNN_OK;
}
extern "C" int ldnn_bn_fwd_apply(const void* y, const float* scale, const float* shift, const void* residual,
                                 void* out, int M, int C, int relu, void* stream) {
  return bn_fwd_apply_impl<__nv_bfloat16>(y, scale, shift, residual, out, M, C, relu, stream);
}
extern "C" int ldnn_bn_fwd_apply_f32(const void* y, const float* scale, const float* shift, const void* residual,
                                 void* out, int M, int C, int relu, void* stream) {
  return bn_fwd_apply_impl<float>(y, scale, shift, residual, out, M, C, relu, stream);
}

template <typename T>
static int bn_bwd_reduce_impl(const void* dout, const void* out, const void* y, const float* save_mean,
                                  const float* save_invstd, float* partial, unsigned int* ticket, int M, int C,
                                  int chunks, int relu, float* dgamma, float* dbeta, void* stream) {
  LDNN_ARG(C % kSlab == 0 && C <= 2048, "bn: C=%d unsupported (need C%%64==0, C<=2048)", C);
  const int rows = (M + chunks - 1) / chunks;
  const int lanes = kRowLanes;
  const size_t smem = bn_smem_bytes(C, lanes);
  bn_bwd_reduce_kernel<T><<<dim3(C / kSlab, chunks), kBnThreads, smem, (cudaStream_t)stream>>>(
      (const T*)dout, (const T*)out, (const T*)y, save_mean, save_invstd, partial,
      ticket, M, C, rows, relu, dgamma, dbeta);
  COUNT_LAUNCH();
  LDNN_CUDA(cudaGetLastError());
  return LDNN_OK;
}
extern "C" int ldnn_bn_bwd_reduce(const void* dout, const void* out, const void* y, const float* save_mean,
                                  const float* save_invstd, float* partial, unsigned int* ticket, int M, int C,
                                  int chunks, int relu, float* dgamma, float* dbeta, void* stream) {
  return bn_bwd_reduce_impl<__nv_bfloat16>(dout, out, y, save_mean, save_invstd, partial, ticket, M, C, chunks, relu, dgamma, dbeta, stream);
}
extern "C" int ldnn_bn_bwd_reduce_f32(const void* dout, const void* out, const void* y, const float* save_mean,
                                  const float* save_invstd, float* partial, unsigned int* ticket, int M, int C,
                                  int chunks, int relu, float* dgamma, float* dbeta, void* stream) {
  return bn_bwd_reduce_impl<float>(dout, out, y, save_mean, save_invstd, partial, ticket, M, C, chunks, relu, dgamma, dbeta, stream);
}

template <typename T>
static int bn_bwd_apply_impl(const void* dout, const void* out, const void* y, const float* save_mean,
                                 const float* save_invstd, const float* gamma, const float* dgamma,
                                 const float* dbeta, void* dy_raw, void* dres, int M, int C, int relu,
                                 int res_accumulate, void* stream) {
  const long long total8 = (long long)M * C / 8;
  bn_bwd_apply_kernel<T><<<grid_for(total8, 256), 256, 0, (cudaStream_t)stream>>>(
      (const T*)dout, (const T*)out, (const T*)y, save_mean, save_invstd, gamma,
      dgamma, dbeta, (T*)dy_raw, (T*)dres, total8, C, relu, res_accumulate, 1.f / (float)M);
  COUNT_LAUNCH();
  LDNN_CUDA(cudaGetLastError());
  return LDNN_OK;
}
extern "C" int ldnn_bn_bwd_apply(const void* dout, const void* out, const void* y, const float* save_mean,
                                 const float* save_invstd, const float* gamma, const float* dgamma,
                                 const float* dbeta, void* dy_raw, void* dres, int M, int C, int relu,
                                 int res_accumulate, void* stream) {
  return bn_bwd_apply_impl<__nv_bfloat16>(dout, out, y, save_mean, save_invstd, gamma, dgamma, dbeta, dy_raw, dres, M, C, relu, res_accumulate, stream);
}
extern "C" int ldnn_bn_bwd_apply_f32(const void* dout, const void* out, const void* y, const float* save_mean,
                                 const float* save_invstd, const float* gamma, const float* dgamma,
                                 const float* dbeta, void* dy_raw, void* dres, int M, int C, int relu,
                                 int res_accumulate, void* stream) {
  return bn_bwd_apply_impl<float>(dout, out, y, save_mean, save_invstd, gamma, dgamma, dbeta, dy_raw, dres, M, C, relu, res_accumulate, stream);
}

static int apply_chunks(int M, int C) {
  // ~2-4 waves of CTAs over (C/64 slabs) x chunks, at least 32 rows (one pass of the row lanes) each
  const int slabs = C / kSlab;
  return std::max(1, std::min((M + 127) / 128, std::max(1, 592 / slabs)));
}

template <typename T>
static int bn_fwd_apply_sums_impl(const void* y, const float* sums, const float* gamma, const float* beta,
                                      float* running_mean, float* running_var, long long* nbt, float* save_mean,
                                      float* save_invstd, const void* residual, void* out, int M, int C, int relu,
                                      int training, float eps, float momentum, void* stream) {
  LDNN_ARG(C % kSlab == 0, "bn: C=%d unsupported (need C%%64==0)", C);
  const int chunks = apply_chunks(M, C);
  const int rows = (M + chunks - 1) / chunks;
  LDNN_CUDA(launch_pdl(bn_apply_sums_kernel<T>, dim3(C / kSlab, chunks), dim3(256), 0, (cudaStream_t)stream,
                       (const T*)y, sums, gamma, beta, running_mean, running_var, nbt, save_mean,
                       save_invstd, (const T*)residual, (T*)out, rows, C, M, relu, training, eps,
                       momentum));
  COUNT_LAUNCH();
  return LDNN_OK;
}
extern "C" int ldnn_bn_fwd_apply_sums(const void* y, const float* sums, const float* gamma, const float* beta,
                                      float* running_mean, float* running_var, long long* nbt, float* save_mean,
                                      float* save_invstd, const void* residual, void* out, int M, int C, int relu,
                                      int training, float eps, float momentum, void* stream) {
  return bn_fwd_apply_sums_impl<__nv_bfloat16>(y, sums, gamma, beta, running_mean, running_var, nbt, save_mean, save_invstd, residual, out, M, C, relu, training, eps, momentum, stream);
}
extern "C" int ldnn_bn_fwd_apply_sums_f32(const void* y, const float* sums, const float* gamma, const float* beta,
                                      float* running_mean, float* running_var, long long* nbt, float* save_mean,
                                      float* save_invstd, const void* residual, void* out, int M, int C, int relu,
                                      int training, float eps, float momentum, void* stream) {
  return bn_fwd_apply_sums_impl<float>(y, sums, gamma, beta, running_mean, running_var, nbt, save_mean, save_invstd, residual, out, M, C, relu, training, eps, momentum, stream);
}

template <typename T>
static int bn_bwd_reduce_sums_impl(const void* dout, const void* out, const void* y, const float* save_mean,
                                       const float* save_invstd, float* sums, int M, int C, int chunks, int relu,
                                       void* stream) {
  LDNN_ARG(C % kSlab == 0, "bn: C=%d unsupported (need C%%64==0)", C);
  const int rows = (M + chunks - 1) / chunks;
  LDNN_CUDA(launch_pdl(bn_bwd_reduce_atomic_kernel<T>, dim3(C / kSlab, chunks), dim3(kBnThreads), 0, (cudaStream_t)stream,
                       (const T*)dout, (const T*)out, (const T*)y, save_mean,
                       save_invstd, sums, M, C, rows, relu));
  COUNT_LAUNCH();
  return LDNN_OK;
}
extern "C" int ldnn_bn_bwd_reduce_sums(const void* dout, const void* out, const void* y, const float* save_mean,
                                       const float* save_invstd, float* sums, int M, int C, int chunks, int relu,
                                       void* stream) {
  return bn_bwd_reduce_sums_impl<__nv_bfloat16>(dout, out, y, save_mean, save_invstd, sums, M, C, chunks, relu, stream);
}
extern "C" int ldnn_bn_bwd_reduce_sums_f32(const void* dout, const void* out, const void* y, const float* save_mean,
                                       const float* save_invstd, float* sums, int M, int C, int chunks, int relu,
                                       void* stream) {
  return bn_bwd_reduce_sums_impl<float>(dout, out, y, save_mean, save_invstd, sums, M, C, chunks, relu, stream);
}

template <typename T>
static int bn_bwd_apply_sums_impl(const void* dout, const void* out, const void* y, const float* save_mean,
                                      const float* save_invstd, const float* gamma, const float* sums, void* dy_raw,
                                      void* dres, float* dgamma, float* dbeta, int M, int C, int relu,
                                      int res_accumulate, void* stream) {
  LDNN_ARG(C % kSlab == 0, "bn: C=%d unsupported (need C%%64==0)", C);
  const int chunks = apply_chunks(M, C);
  const int rows = (M + chunks - 1) / chunks;
  LDNN_CUDA(launch_pdl(bn_bwd_apply_sums_kernel<T>, dim3(C / kSlab, chunks), dim3(256), 0, (cudaStream_t)stream,
                       (const T*)dout, (const T*)out, (const T*)y, save_mean,
                       save_invstd, gamma, sums, (T*)dy_raw, (T*)dres, dgamma, dbeta, rows, M, C,
                       relu, res_accumulate, 1.f / (float)M));
  COUNT_LAUNCH();
  return LDNN_OK;
}
extern "C" int ldnn_bn_bwd_apply_sums(const void* dout, const void* out, const void* y, const float* save_mean,
                                      const float* save_invstd, const float* gamma, const float* sums, void* dy_raw,
                                      void* dres, float* dgamma, float* dbeta, int M, int C, int relu,
                                      int res_accumulate, void* stream) {
  return bn_bwd_apply_sums_impl<__nv_bfloat16>(dout, out, y, save_mean, save_invstd, gamma, sums, dy_raw, dres, dgamma, dbeta, M, C, relu, res_accumulate, stream);
}
extern "C" int ldnn_bn_bwd_apply_sums_f32(const void* dout, const void* out, const void* y, const float* save_mean,
                                      const float* save_invstd, const float* gamma, const float* sums, void* dy_raw,
                                      void* dres, float* dgamma, float* dbeta, int M, int C, int relu,
                                      int res_accumulate, void* stream) {
  return bn_bwd_apply_sums_impl<float>(dout, out, y, save_mean, save_invstd, gamma, sums, dy_raw, dres, dgamma, dbeta, M, C, relu, res_accumulate, stream);
}

template <typename T>
static int im2col_stem_impl(const float* x, void* col, int N, int C, int H, int W, int R, int stride, int pad,
                                int Kpad, void* stream) {
  LDNN_ARG(Kpad % 64 == 0 && Kpad >= R * R * C, "im2col: Kpad=%d", Kpad);
  const int P = (H + 2 * pad - R) / stride + 1, Q = (W + 2 * pad - R) / stride + 1;
  const long long total = (long long)N * P * Q * (Kpad / 8);
  LDNN_CUDA(launch_pdl(im2col_stem_kernel<T>, dim3(grid_for(total, 256)), dim3(256), 0, (cudaStream_t)stream, x,
                       (T*)col, N, C, H, W, R, stride, pad, P, Q, Kpad));
  COUNT_LAUNCH();
  return LDNN_OK;
}
extern "C" int ldnn_im2col_stem(const float* x, void* col, int N, int C, int H, int W, int R, int stride, int pad,
                                int Kpad, void* stream) {
  return im2col_stem_impl<__nv_bfloat16>(x, col, N, C, H, W, R, stride, pad, Kpad, stream);
}
extern "C" int ldnn_im2col_stem_f32(const float* x, void* col, int N, int C, int H, int W, int R, int stride, int pad,
                                int Kpad, void* stream) {
  return im2col_stem_impl<float>(x, col, N, C, H, W, R, stride, pad, Kpad, stream);
}

template <typename T>
static int maxpool_fwd_impl(const void* x, void* y, void* idx, int N, int H, int W, int C, int k, int stride,
                                int pad, void* stream) {
  LDNN_ARG(C % 8 == 0 && k * k <= 255, "maxpool: C=%d k=%d", C, k);
  const int P = (H + 2 * pad - k) / stride + 1, Q = (W + 2 * pad - k) / stride + 1;
  const long long total = (long long)N * P * Q * (C / 8);
  LDNN_CUDA(launch_pdl(maxpool_fwd_kernel<T>, dim3(grid_for(total, 256)), dim3(256), 0, (cudaStream_t)stream,
                       (const T*)x, (T*)y, (uint8_t*)idx, N, H, W, C, k, stride, pad, P, Q));
  COUNT_LAUNCH();
  return LDNN_OK;
}
extern "C" int ldnn_maxpool_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, int k, int stride,
                                int pad, void* stream) {
  return maxpool_fwd_impl<__nv_bfloat16>(x, y, idx, N, H, W, C, k, stride, pad, stream);
}
extern "C" int ldnn_maxpool_fwd_f32(const void* x, void* y, void* idx, int N, int H, int W, int C, int k, int stride,
                                int pad, void* stream) {
  return maxpool_fwd_impl<float>(x, y, idx, N, H, W, C, k, stride, pad, stream);
}

template <typename T>
static int maxpool_bwd_impl(const void* idx, const void* dy, void* dx, int N, int H, int W, int C, int k,
                                int stride, int pad, void* stream) {
  LDNN_ARG(idx != nullptr && C % 8 == 0, "maxpool backward needs the argmax map written by the forward pass");
  const int P = (H + 2 * pad - k) / stride + 1, Q = (W + 2 * pad - k) / stride + 1;
  const long long total = (long long)N * H * W * (C / 8);
  LDNN_CUDA(launch_pdl(maxpool_bwd_kernel<T>, dim3(grid_for(total, 256)), dim3(256), 0, (cudaStream_t)stream,
                       (const uint8_t*)idx, (const T*)dy, (T*)dx, N, H, W, C, k, stride, pad, P,
                       Q));
  COUNT_LAUNCH();
  return LDNN_OK;
}
extern "C" int ldnn_maxpool_bwd(const void* idx, const void* dy, void* dx, int N, int H, int W, int C, int k,
                                int stride, int pad, void* stream) {
  return maxpool_bwd_impl<__nv_bfloat16>(idx, dy, dx, N, H, W, C, k, stride, pad, stream);
}
extern "C" int ldnn_maxpool_bwd_f32(const void* idx, const void* dy, void* dx, int N, int H, int W, int C, int k,
                                int stride, int pad, void* stream) {
  return maxpool_bwd_impl<float>(idx, dy, dx, N, H, W, C, k, stride, pad, stream);
}

// ------------------------------------------------------------------ allocation-free BN statistics
// Some users of this library cannot give the BN-statistics kernel a workspace, and cannot obtain
// one: code that runs as a plug-in or custom op inside a host which has handed the whole GPU to its
// own caching allocator (so a cudaMalloc from here fails, or fragments the host's pool), and loops
// that must never hit cudaMalloc/cudaFree because those calls synchronise the whole device.
// For them the per-CTA partial sums and the arrival ticket live in the fixed array below, which
// every GPU that loads this library keeps resident until the process exits.
//
// Capacity: chunks * 2 * C floats must fit in kNoAllocPartialFloats (1 MiB: e.g. 128 chunks at
// C = 1024, or 64 chunks at C = 2048).  Larger requests are refused; use ldnn_bn_fwd_stats with a
// caller-owned workspace instead.
constexpr size_t kNoAllocTicketWords = 64;                 // 256 B header: word 0 is the ticket
constexpr size_t kNoAllocPartialFloats = 262144;           // 1 MiB of partial sums
static __device__ __align__(256) float g_bn_noalloc_scratch[kNoAllocTicketWords + kNoAllocPartialFloats];

namespace ldnn {
// The array's name is only a host-side handle; the device address is looked up once per device
// (with that device current) and cached here.
static float* g_bn_noalloc_ptr[64] = {nullptr};

static int bn_noalloc_scratch(float** out) {
  int dev = -1;
  LDNN_CUDA(cudaGetDevice(&dev));
  LDNN_ARG(dev >= 0 && dev < 64, "device ordinal %d out of range", dev);
  if (g_bn_noalloc_ptr[dev] == nullptr) {
    float* p = nullptr;
    LDNN_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_bn_noalloc_scratch));
    LDNN_ARG(p != nullptr, "cudaGetSymbolAddress returned a null scratch address");
    g_bn_noalloc_ptr[dev] = p;
  }
  *out = g_bn_noalloc_ptr[dev];
  return LDNN_OK;
}
}  // namespace ldnn

extern "C" int ldnn_bn_fwd_stats_noalloc(const void* y, int M, int C, int chunks, const float* gamma,
                                         const float* beta, float* running_mean, float* running_var,
                                         long long* nbt, float* save_mean, float* save_invstd, float* scale,
                                         float* shift, float eps, float momentum, void* stream) {
  LDNN_ARG(C % kSlab == 0 && C <= 2048, "bn: C=%d unsupported (need C%%64==0, C<=2048)", C);
  LDNN_ARG(chunks >= 1, "bn: chunks");
  const size_t need = (size_t)chunks * 2 * (size_t)C;
  LDNN_ARG(need <= kNoAllocPartialFloats,
           "bn_fwd_stats_noalloc: chunks*2*C = %zu floats exceeds the built-in scratch of %zu floats (1 MiB); "
           "lower `chunks` or call ldnn_bn_fwd_stats with your own workspace", need, kNoAllocPartialFloats);
  float* base = nullptr;
  int rc = bn_noalloc_scratch(&base);
  if (rc) return rc;
  unsigned int* ticket = reinterpret_cast<unsigned int*>(base);
  float* partial = base + kNoAllocTicketWords;
  cudaStream_t st = (cudaStream_t)stream;
  // The array keeps whatever the previous call left.  Partials are fully overwritten before they
  // are read; the ticket must start at zero, so it is cleared in stream order.
  LDNN_CUDA(cudaMemsetAsync(ticket, 0, kNoAllocTicketWords * sizeof(unsigned int), st));
  const int rows = (M + chunks - 1) / chunks;
  const int lanes = kRowLanes;
  const size_t smem = bn_smem_bytes(C, lanes);
  bn_stats_kernel<__nv_bfloat16><<<dim3(C / kSlab, chunks), kBnThreads, smem, st>>>((const __nv_bfloat16*)y, partial, ticket, M, C, rows, gamma, beta,
                                                   running_mean, running_var, nbt, save_mean, save_invstd, scale,
                                                   shift, eps, momentum);
  COUNT_LAUNCH();
  LDNN_CUDA(cudaGetLastError());
  return LDNN_OK;
}
