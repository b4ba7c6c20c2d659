// K-HEAD: global-average-pool + linear + softmax cross-entropy forward AND backward, with the
// correct-prediction count and the fc gradients, in two launches and no host sync
// (replaces Balanced All-Reduce/model.py:99-100,108-110; main.py:52; trainer.py:207,212-216 —
// upstream reads loss/accuracy back with three .item() calls per batch).
// K-OPT: Adam / SGD-momentum on flat fp32 slabs + bf16 compute copy (trainer.py:205,210,218).
#include "../include/ldnn_b200.h"
#include "common.cuh"
#include "optim.cuh"
#include <atomic>

namespace ldnn {
extern std::atomic<unsigned long long> g_launches;

constexpr int kMaxClasses = 1024;

// one CTA per sample
template <typename T>
__global__ void __launch_bounds__(256)
head_sample_kernel(const T* __restrict__ act, int HW, int C, int classes, const float* __restrict__ fc_w,
                   const float* __restrict__ fc_b, const long long* __restrict__ labels, float* __restrict__ pooled,
                   float* __restrict__ logits, float* __restrict__ dlogits, float* __restrict__ loss_n,
                   float* __restrict__ correct_n, long long* __restrict__ pred, T* __restrict__ dact,
                   int training, float inv_batch) {
  pdl_wait();
  pdl_launch_dependents();
  extern __shared__ float sm[];          // pooled[C] | logits[classes] | dl[classes]
  float* sp = sm;
  float* sl = sm + C;
  float* sd = sl + classes;
  const int n = blockIdx.x;
  const T* a = act + (size_t)n * HW * C;
  const float inv_hw = 1.f / (float)HW;
  for (int c = threadIdx.x * 2; c < C; c += blockDim.x * 2) {
    float s0 = 0.f, s1 = 0.f;
    for (int p = 0; p < HW; ++p) {
      float t0, t1;
      load2(a + (size_t)p * C + c, t0, t1);
      s0 += t0; s1 += t1;
    }
    sp[c] = s0 * inv_hw; sp[c + 1] = s1 * inv_hw;
    pooled[(size_t)n * C + c] = sp[c];
    pooled[(size_t)n * C + c + 1] = sp[c + 1];
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  for (int j = warp; j < classes; j += nwarps) {
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s += sp[c] * fc_w[(size_t)j * C + c];
    s = warp_sum(s);
    if (lane == 0) { sl[j] = s + fc_b[j]; logits[(size_t)n * classes + j] = sl[j]; }
  }
  __syncthreads();
  if (warp == 0) {
    float mx = -INFINITY; int arg = 0;
    for (int j = lane; j < classes; j += 32) if (sl[j] > mx) { mx = sl[j]; arg = j; }
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
      if (om > mx || (om == mx && oa < arg)) { mx = om; arg = oa; }
    }
    float se = 0.f;
    for (int j = lane; j < classes; j += 32) se += __expf(sl[j] - mx);
    se = warp_sum(se);
    const int lab = (int)labels[n];
    const float lse = mx + __logf(se);
    if (lane == 0) {
      loss_n[n] = lse - sl[lab];
      correct_n[n] = (arg == lab) ? 1.f : 0.f;
      pred[n] = arg;
    }
    for (int j = lane; j < classes; j += 32) {
      const float pj = __expf(sl[j] - lse);
      const float d = (pj - (j == lab ? 1.f : 0.f)) * inv_batch;
      sd[j] = d;
      dlogits[(size_t)n * classes + j] = d;
    }
  }
  if (!training) return;
  __syncthreads();
  // d_act[n, p, c] = (1/HW) * sum_j dl[j] * W[j, c]   (same for every pixel p)
  for (int c = threadIdx.x * 2; c < C; c += blockDim.x * 2) {
    float g0 = 0.f, g1 = 0.f;
    for (int j = 0; j < classes; ++j) {
      const float2 w = *reinterpret_cast<const float2*>(fc_w + (size_t)j * C + c);
      g0 += sd[j] * w.x; g1 += sd[j] * w.y;
    }
    for (int p = 0; p < HW; ++p) store2(dact + ((size_t)n * HW + p) * C + c, g0 * inv_hw, g1 * inv_hw);
  }
}

// ------------------------------------------------------------------ wide head (ImageNet-size fc)
// With ~1000 classes the per-sample kernel above re-reads the whole fc weight for every sample
// (1.7 ms for ResNet-50 at batch 32).  The wide path keeps the same math but tiles it so that the
// weight is read once per CTA: pool -> logits (warp per class) -> softmax/CE -> d_act (thread per
// channel, 8 samples per CTA) -> fc grads (channel tile x class group).
template <typename T>
__global__ void __launch_bounds__(256)
head_pool_kernel(const T* __restrict__ act, int HW, int C, float* __restrict__ pooled) {
  pdl_wait();
  pdl_launch_dependents();
  const int n = blockIdx.x;
  const T* a = act + (size_t)n * HW * C;
  const float inv_hw = 1.f / (float)HW;
  for (int c = threadIdx.x * 2; c < C; c += blockDim.x * 2) {
    float s0 = 0.f, s1 = 0.f;
    for (int p = 0; p < HW; ++p) {
      float t0, t1;
      load2(a + (size_t)p * C + c, t0, t1);
      s0 += t0; s1 += t1;
    }
    pooled[(size_t)n * C + c] = s0 * inv_hw;
    pooled[(size_t)n * C + c + 1] = s1 * inv_hw;
  }
}

// one warp per class; lanes stride over channels; 16 samples per pass
__global__ void __launch_bounds__(256)
head_logits_kernel(const float* __restrict__ pooled, const float* __restrict__ fc_w, const float* __restrict__ fc_b,
                   float* __restrict__ logits, int N, int C, int classes) {
  pdl_wait();
  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int j = blockIdx.x * 8 + warp;
  if (j >= classes) return;
  const float* w = fc_w + (size_t)j * C;
  for (int n0 = 0; n0 < N; n0 += 16) {
    float acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = 0.f;
    for (int c = lane; c < C; c += 32) {
      const float wv = w[c];
#pragma unroll
      for (int i = 0; i < 16; ++i)
        if (n0 + i < N) acc[i] += wv * pooled[(size_t)(n0 + i) * C + c];
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const float t = warp_sum(acc[i]);
      if (lane == 0 && n0 + i < N) logits[(size_t)(n0 + i) * classes + j] = t + fc_b[j];
    }
  }
}

__global__ void __launch_bounds__(256)
head_softmax_kernel(const float* __restrict__ logits, const long long* __restrict__ labels, float* __restrict__ dlogits,
                    float* __restrict__ loss_n, float* __restrict__ correct_n, long long* __restrict__ pred,
                    int classes, float inv_batch) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float s_val[8];
  __shared__ int s_arg[8];
  const int n = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float* lg = logits + (size_t)n * classes;
  float mx = -INFINITY; int arg = 0x7fffffff;
  for (int j = threadIdx.x; j < classes; j += blockDim.x) if (lg[j] > mx) { mx = lg[j]; arg = j; }
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, mx, o);
    const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (om > mx || (om == mx && oa < arg)) { mx = om; arg = oa; }
  }
  if (lane == 0) { s_val[warp] = mx; s_arg[warp] = arg; }
  __syncthreads();
  mx = s_val[0]; arg = s_arg[0];
  for (int k = 1; k < 8; ++k) if (s_val[k] > mx || (s_val[k] == mx && s_arg[k] < arg)) { mx = s_val[k]; arg = s_arg[k]; }
  __syncthreads();
  float se = 0.f;
  for (int j = threadIdx.x; j < classes; j += blockDim.x) se += __expf(lg[j] - mx);
  se = warp_sum(se);
  if (lane == 0) s_val[warp] = se;
  __syncthreads();
  se = 0.f;
  for (int k = 0; k < 8; ++k) se += s_val[k];
  const int lab = (int)labels[n];
  const float lse = mx + __logf(se);
  if (threadIdx.x == 0) { loss_n[n] = lse - lg[lab]; correct_n[n] = (arg == lab) ? 1.f : 0.f; pred[n] = arg; }
  for (int j = threadIdx.x; j < classes; j += blockDim.x)
    dlogits[(size_t)n * classes + j] = (__expf(lg[j] - lse) - (j == lab ? 1.f : 0.f)) * inv_batch;
}

// CTA = (128-channel tile, group of 8 samples); thread = one channel, 8 accumulators; the weight is
// streamed once per CTA (coalesced over channels), dlogits of the 8 samples are staged in smem.
template <typename T>
__global__ void __launch_bounds__(128)
head_dact_kernel(const float* __restrict__ dlogits, const float* __restrict__ fc_w, T* __restrict__ dact,
                 int N, int HW, int C, int classes) {
  pdl_wait();
  pdl_launch_dependents();
  extern __shared__ float s_dl[];                   // [8][classes]
  const int n0 = blockIdx.y * 8;
  for (int i = threadIdx.x; i < 8 * classes; i += blockDim.x) {
    const int nn = n0 + i / classes;
    s_dl[i] = nn < N ? dlogits[(size_t)nn * classes + i % classes] : 0.f;
  }
  __syncthreads();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll 4
  for (int j = 0; j < classes; ++j) {
    const float wv = fc_w[(size_t)j * C + c];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] += s_dl[i * classes + j] * wv;
  }
  const float inv_hw = 1.f / (float)HW;
  for (int i = 0; i < 8 && n0 + i < N; ++i) {
    for (int p = 0; p < HW; ++p) store1(dact + ((size_t)(n0 + i) * HW + p) * C + c, acc[i] * inv_hw);
  }
}

// CTA = (128-channel tile, group of 8 classes); also the batch totals (CTA 0,0)
__global__ void __launch_bounds__(128)
head_fcgrad_kernel(const float* __restrict__ pooled, const float* __restrict__ dlogits,
                   const float* __restrict__ loss_n, const float* __restrict__ correct_n, int N, int C, int classes,
                   float* __restrict__ dfc_w, float* __restrict__ dfc_b, float* __restrict__ loss_out,
                   float* __restrict__ correct_out, int training) {
  pdl_wait();
  pdl_launch_dependents();
  extern __shared__ float s_dl[];                   // [N][8]
  const int j0 = blockIdx.y * 8;
  if (training) {
    for (int i = threadIdx.x; i < N * 8; i += blockDim.x) {
      const int j = j0 + (i & 7);
      s_dl[i] = j < classes ? dlogits[(size_t)(i >> 3) * classes + j] : 0.f;
    }
    __syncthreads();
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < C) {
      float acc[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll 4
      for (int n = 0; n < N; ++n) {
        const float pv = pooled[(size_t)n * C + c];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += s_dl[n * 8 + j] * pv;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (j0 + j < classes) dfc_w[(size_t)(j0 + j) * C + c] = acc[j];
    }
    if (blockIdx.x == 0 && threadIdx.x < 8 && j0 + threadIdx.x < classes) {
      float sb = 0.f;
      for (int n = 0; n < N; ++n) sb += s_dl[n * 8 + threadIdx.x];
      dfc_b[j0 + threadIdx.x] = sb;
    }
  }
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x < 32) {
    float l = 0.f, k = 0.f;
    for (int n = threadIdx.x; n < N; n += 32) { l += loss_n[n]; k += correct_n[n]; }
    l = warp_sum(l); k = warp_sum(k);
    if (threadIdx.x == 0) { *loss_out = l / (float)N; *correct_out = k; }
  }
}

// ------------------------------------------------------------------ optimizer
__global__ void __launch_bounds__(256)
optim_kernel(float* __restrict__ params, const float* __restrict__ grads, float* __restrict__ m, float* __restrict__ v,
             __nv_bfloat16* __restrict__ shadow, long long n4, const ldnn_optim o) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
       i += (long long)gridDim.x * blockDim.x) {
    float4 p = reinterpret_cast<float4*>(params)[i];
    float4 g = reinterpret_cast<const float4*>(grads)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = o.kind == 0 ? reinterpret_cast<float4*>(v)[i] : make_float4(0, 0, 0, 0);
    optim_update4(p, g, mm, vv, o);
    reinterpret_cast<float4*>(params)[i] = p;
    reinterpret_cast<float4*>(m)[i] = mm;
    if (o.kind == 0) reinterpret_cast<float4*>(v)[i] = vv;
    if (shadow) store_bf16x4(shadow + 4 * i, p);
  }
}

__global__ void __launch_bounds__(256)
cast_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n4) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
       i += (long long)gridDim.x * blockDim.x)
    store_bf16x4(dst + 4 * i, reinterpret_cast<const float4*>(src)[i]);
}

}  // namespace ldnn

using namespace ldnn;

template <typename T>
static int head_fwd_bwd_impl(const void* act, int N, int HW, int C, int classes, const float* fc_w,
                             const float* fc_b, const long long* labels, float* pooled, float* logits,
                             float* dlogits, float* loss_n, float* correct_n, long long* pred, void* dact,
                             float* dfc_w, float* dfc_b, float* loss_out, float* correct_out, int training,
                             void* stream) {
  LDNN_ARG(C % 2 == 0 && classes <= kMaxClasses, "head: C=%d classes=%d", C, classes);
  if (classes > 64) {            // wide head: weight streamed once per CTA instead of once per sample
    cudaStream_t st = (cudaStream_t)stream;
    LDNN_ARG((size_t)8 * classes * sizeof(float) <= 48 * 1024, "head: classes=%d too many for the smem stage", classes);
    LDNN_CUDA(launch_pdl(head_pool_kernel<T>, dim3(N), dim3(256), 0, st, (const T*)act, HW, C, pooled));
    LDNN_CUDA(launch_pdl(head_logits_kernel, dim3((classes + 7) / 8), dim3(256), 0, st, (const float*)pooled, fc_w, fc_b,
                         logits, N, C, classes));
    LDNN_CUDA(launch_pdl(head_softmax_kernel, dim3(N), dim3(256), 0, st, (const float*)logits, labels, dlogits, loss_n,
                         correct_n, pred, classes, 1.f / (float)N));
    int launched = 3;
    if (training) {
      LDNN_CUDA(launch_pdl(head_dact_kernel<T>, dim3((C + 127) / 128, (N + 7) / 8), dim3(128),
                           (size_t)8 * classes * sizeof(float), st, (const float*)dlogits, fc_w, (T*)dact, N,
                           HW, C, classes));
      ++launched;
    }
    LDNN_CUDA(launch_pdl(head_fcgrad_kernel, dim3(training ? (C + 127) / 128 : 1, training ? (classes + 7) / 8 : 1),
                         dim3(128), (size_t)N * 8 * sizeof(float), st, (const float*)pooled, (const float*)dlogits,
                         (const float*)loss_n, (const float*)correct_n, N, C, classes, dfc_w, dfc_b, loss_out,
                         correct_out, training));
    g_launches.fetch_add(launched + 1);
    return LDNN_OK;
  }
  const size_t smem = (size_t)(C + 2 * classes) * sizeof(float);
  LDNN_CUDA(launch_pdl(head_sample_kernel<T>, dim3(N), dim3(256), smem, (cudaStream_t)stream, (const T*)act, HW,
                       C, classes, fc_w, fc_b, labels, pooled, logits, dlogits, loss_n, correct_n, pred,
                       (T*)dact, training, 1.f / (float)N));
  g_launches.fetch_add(1);
  // fc gradients + batch totals on a (channel tile x class group) grid
  LDNN_CUDA(launch_pdl(head_fcgrad_kernel, dim3(training ? (C + 127) / 128 : 1, training ? (classes + 7) / 8 : 1),
                       dim3(128), (size_t)N * 8 * sizeof(float), (cudaStream_t)stream, (const float*)pooled,
                       (const float*)dlogits, (const float*)loss_n, (const float*)correct_n, N, C, classes, dfc_w, dfc_b,
                       loss_out, correct_out, training));
  g_launches.fetch_add(1);
  return LDNN_OK;
}

extern "C" int ldnn_head_fwd_bwd(const void* act, int N, int HW, int C, int classes, const float* fc_w,
                                 const float* fc_b, const long long* labels, float* pooled, float* logits,
                                 float* dlogits, float* loss_n, float* correct_n, long long* pred, void* dact,
                                 float* dfc_w, float* dfc_b, float* loss_out, float* correct_out, int training,
                                 void* stream) {
  return head_fwd_bwd_impl<__nv_bfloat16>(act, N, HW, C, classes, fc_w, fc_b, labels, pooled, logits, dlogits, loss_n,
                                          correct_n, pred, dact, dfc_w, dfc_b, loss_out, correct_out, training, stream);
}
extern "C" int ldnn_head_fwd_bwd_f32(const void* act, int N, int HW, int C, int classes, const float* fc_w,
                                     const float* fc_b, const long long* labels, float* pooled, float* logits,
                                     float* dlogits, float* loss_n, float* correct_n, long long* pred, void* dact,
                                     float* dfc_w, float* dfc_b, float* loss_out, float* correct_out, int training,
                                     void* stream) {
  return head_fwd_bwd_impl<float>(act, N, HW, C, classes, fc_w, fc_b, labels, pooled, logits, dlogits, loss_n,
                                  correct_n, pred, dact, dfc_w, dfc_b, loss_out, correct_out, training, stream);
}

extern "C" int ldnn_optim_step(float* params, const float* grads, float* m, float* v, void* shadow_bf16, long long n,
                               const ldnn_optim* o, void* stream) {
  LDNN_ARG(n % 4 == 0, "optim: n=%lld must be a multiple of 4", n);
  LDNN_ARG(o->kind == 1 || v != nullptr, "adam needs v");
  const long long n4 = n / 4;
  const int blocks = (int)std::min<long long>((n4 + 255) / 256, 148 * 8);
  optim_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(params, grads, m, v, (__nv_bfloat16*)shadow_bf16, n4, *o);
  g_launches.fetch_add(1);
  LDNN_CUDA(cudaGetLastError());
  return LDNN_OK;
}

extern "C" int ldnn_cast_bf16(const float* src, void* dst, long long n, void* stream) {
  LDNN_ARG(n % 4 == 0, "cast: n %% 4");
  const long long n4 = n / 4;
  const int blocks = (int)std::min<long long>((n4 + 255) / 256, 148 * 8);
  cast_bf16_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst, n4);
  g_launches.fetch_add(1);
  LDNN_CUDA(cudaGetLastError());
  return LDNN_OK;
}
