// K-OPT device functions: one Adam / SGD-momentum update on 4 packed elements.
// Inlined into the standalone optimizer kernel AND into the tails of the exchange kernels
// (comm.cu), so "reduce -> scale -> update -> cast" is one pass over the data
// (reference: optimizer.step() / zero_grad() / StepLR, Balanced All-Reduce/trainer.py:205,210,218).
#pragma once
#include "../include/ldnn_b200.h"
#include <cuda_bf16.h>

namespace ldnn {

__device__ __forceinline__ void optim_update1(float& p, float g, float& m, float& v, const ldnn_optim& o) {
  g = g * o.grad_scale + o.weight_decay * p;
  if (o.kind == 0) {
    m = o.beta1 * m + (1.f - o.beta1) * g;
    v = o.beta2 * v + (1.f - o.beta2) * g * g;
    const float denom = sqrtf(v) * rsqrtf(o.bc2) + o.eps;
    p -= (o.lr / o.bc1) * (m / denom);
  } else {
    m = o.momentum * m + g;
    p -= o.lr * m;
  }
}

__device__ __forceinline__ void optim_update4(float4& p, const float4& g, float4& m, float4& v, const ldnn_optim& o) {
  optim_update1(p.x, g.x, m.x, v.x, o);
  optim_update1(p.y, g.y, m.y, v.y, o);
  optim_update1(p.z, g.z, m.z, v.z, o);
  optim_update1(p.w, g.w, m.w, v.w, o);
}

__device__ __forceinline__ void store_bf16x4(__nv_bfloat16* dst, const float4& p) {
  uint2 u;
  *reinterpret_cast<__nv_bfloat162*>(&u.x) = __floats2bfloat162_rn(p.x, p.y);
  *reinterpret_cast<__nv_bfloat162*>(&u.y) = __floats2bfloat162_rn(p.z, p.w);
  *reinterpret_cast<uint2*>(dst) = u;
}

}  // namespace ldnn
