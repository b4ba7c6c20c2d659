// Symmetric memory + the fused peer-to-peer exchange kernels (K-AR, K-RING, K-DRING, K-BCAST,
// K-BAR, K-MET).
//
// The reference exchanges gradients/weights once per global epoch, one tensor at a time, through
// torch.distributed all_reduce + a separate scale kernel (Balanced All-Reduce/communication.py:4-31)
// or through host-staged mpi4py Isend/Irecv + numpy round trips (Balanced Ring/communication.py:5-62,
// Balanced Double-Ring/communication.py:5-77).  Here every topology is ONE kernel over a flat fp32
// slab that loads/stores directly into the peers' buffers over NVLink/NVSwitch:
//   * system-scope release/acquire flags in a peer-mapped signal pad (never reset: the value is a
//     strictly increasing epoch, so there is no ABA across calls),
//   * 16-byte relaxed.sys vector loads/stores for the payload, reduction in fp32 registers,
//   * the per-rank weight (1/ws, n_r/sum n, or the local_weight forms), the Adam/SGD update and
//     the bf16 cast of the updated parameters run in the same pass (optim.cuh),
//   * every spin-wait carries a watchdog.
// `emulate=1` plays all ranks in one cooperative launch on one GPU (blockIdx.y = rank) so the
// exact kernel code is testable without multiple devices.
#include "../include/ldnn_b200.h"
#include "common.cuh"
#include "optim.cuh"
#include <atomic>
#include <cooperative_groups.h>

namespace ldnn {
extern std::atomic<unsigned long long> g_launches;

constexpr int kCommThreads = 512;
constexpr int kBarWords = 512 * LDNN_MAX_RANKS;        // barrier slots: [block][src rank]
constexpr int kUserWords = 64;                          // tail of the pad: host-visible flag words (straggler limit)
constexpr int kFlagWords = LDNN_SIGNAL_WORDS - kBarWords - kUserWords;

struct Peers { void* p[LDNN_MAX_RANKS]; };

// ------------------------------------------------------------------ system-scope primitives
__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_sys_f4(const float4* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_sys_f4(float4* p, const float4& v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

struct Ctx {
  int rank, world, block, blocks;
  unsigned long long timeout;
  int* err;
  unsigned int* sig[LDNN_MAX_RANKS];
};

__device__ __forceinline__ Ctx make_ctx(const ldnn_comm& c) {
  Ctx x;
  x.rank = c.emulate ? (int)blockIdx.y : c.rank;
  x.world = c.world;
  x.block = blockIdx.x;
  x.blocks = gridDim.x;
  x.timeout = c.timeout_cycles ? c.timeout_cycles : (1ull << 35);
  x.err = c.error_flag;
#pragma unroll
  for (int i = 0; i < LDNN_MAX_RANKS; ++i) x.sig[i] = reinterpret_cast<unsigned int*>(c.signal[i]);
  return x;
}

// Wait until *p has reached `seq` (wrap-safe).  Returns false on watchdog expiry.
__device__ __forceinline__ bool spin_until(const Ctx& x, const unsigned int* p, unsigned int seq) {
  const long long t0 = clock64();
  while ((int)(ld_acquire_sys(p) - seq) < 0) {
    if ((unsigned long long)(clock64() - t0) > x.timeout) {
      if (x.err) atomicExch(x.err, 1);
      return false;
    }
    __nanosleep(40);
  }
  return true;
}

// K-BAR: CTA `block` of every rank meets.  Release: all of this CTA's earlier writes (incl. peer
// stores) are visible to whoever observes the flag; acquire: symmetric.
__device__ __forceinline__ void comm_barrier(const Ctx& x, unsigned int seq) {
  __syncthreads();
  if ((int)threadIdx.x < x.world) {
    __threadfence_system();
    st_release_sys(x.sig[threadIdx.x] + x.block * LDNN_MAX_RANKS + x.rank, seq);
    spin_until(x, x.sig[x.rank] + x.block * LDNN_MAX_RANKS + threadIdx.x, seq);
  }
  __syncthreads();
}

// tests only: CTA `debug_delay_block - 1` of every rank idles for `debug_delay_cycles` before its push loop, which
// makes "a CTA reads what a DIFFERENT CTA of a peer has not pushed yet" deterministic instead of a rare race
__device__ __forceinline__ void debug_delay(const ldnn_comm& c, const Ctx& x) {
  if (c.debug_delay_block != 0 && c.debug_delay_block == x.block + 1) {
    const long long t0 = clock64();
    while ((unsigned long long)(clock64() - t0) < (unsigned long long)c.debug_delay_cycles) __nanosleep(200);
  }
}

__device__ __forceinline__ float4 f4_scale(const float4& a, float s) { return make_float4(a.x * s, a.y * s, a.z * s, a.w * s); }
__device__ __forceinline__ float4 f4_fma(const float4& a, float s, const float4& b) {
  return make_float4(fmaf(a.x, s, b.x), fmaf(a.y, s, b.y), fmaf(a.z, s, b.z), fmaf(a.w, s, b.w));
}

// ================================================================== K-AR: two-shot all-reduce
__global__ void __launch_bounds__(kCommThreads, 1)
allreduce_fused_kernel(const ldnn_comm c, const ldnn_exchange e) {
  const Ctx x = make_ctx(c);
  const int W = x.world, r = x.rank;
  const long long n4 = e.n >> 2;
  const long long per = (n4 + W - 1) / W;
  const long long lo = min((long long)r * per, n4), hi = min(lo + per, n4);
  const unsigned int seq = c.epoch * 4u;
  const bool update = (e.mode & LDNN_X_UPDATE) != 0, selfmix = (e.mode & LDNN_X_SELFMIX) != 0;

  comm_barrier(x, seq + 1);   // every rank's input slab is final
  debug_delay(c, x);

  float4* const my_params = reinterpret_cast<float4*>(e.params[r]);
  float4* const my_m = reinterpret_cast<float4*>(e.m);
  float4* const my_v = reinterpret_cast<float4*>(e.v);
  const long long stride = (long long)x.blocks * kCommThreads;
  for (long long i = lo + (long long)x.block * kCommThreads + threadIdx.x; i < hi; i += 2 * stride) {
    const long long i1 = i + stride;
    const bool two = i1 < hi;
    float4 a0[LDNN_MAX_RANKS], a1[LDNN_MAX_RANKS];
#pragma unroll
    for (int p = 0; p < LDNN_MAX_RANKS; ++p) {
      if (p < W) {
        const float4* src = reinterpret_cast<const float4*>(e.grads[p]);
        a0[p] = ld_sys_f4(src + i);
        if (two) a1[p] = ld_sys_f4(src + i1);
      }
    }
    float4 s0 = make_float4(0, 0, 0, 0), s1 = s0;
#pragma unroll
    for (int p = 0; p < LDNN_MAX_RANKS; ++p) {
      if (p < W) { s0 = f4_fma(a0[p], e.weight[p], s0); if (two) s1 = f4_fma(a1[p], e.weight[p], s1); }
    }
    if (update && !selfmix) {
      // fused tail: optimizer on the owned slice, then push the *updated parameters*
      float4 p0 = my_params[i], m0 = my_m[i], v0 = e.opt.kind == 0 ? my_v[i] : make_float4(0, 0, 0, 0);
      optim_update4(p0, s0, m0, v0, e.opt);
      my_m[i] = m0; if (e.opt.kind == 0) my_v[i] = v0;
      s0 = p0;
      if (two) {
        float4 p1 = my_params[i1], m1 = my_m[i1], v1 = e.opt.kind == 0 ? my_v[i1] : make_float4(0, 0, 0, 0);
        optim_update4(p1, s1, m1, v1, e.opt);
        my_m[i1] = m1; if (e.opt.kind == 0) my_v[i1] = v1;
        s1 = p1;
      }
    }
#pragma unroll
    for (int q = 0; q < LDNN_MAX_RANKS; ++q) {
      if (q < W) {
        float4* dst = reinterpret_cast<float4*>(selfmix ? e.recv[q] : (update ? e.params[q] : e.grads[q]));
        st_sys_f4(dst + i, s0);
        if (two) st_sys_f4(dst + i1, s1);
      }
    }
  }

  comm_barrier(x, seq + 2);   // CTA b of every rank has pushed its share: see the mapping note below

  // ---- local pass over the whole slab: self-mix, (per-rank) update, bf16 cast ----
  // comm_barrier joins CTA b of this rank with CTA b of every peer ONLY, so this pass may read just what those
  // CTAs pushed: it walks owner slice by owner slice with the push loop's own element-to-CTA mapping
  // (i = lo_p + b*T + tid + k*stride), not the flat i = b*T + tid + k*stride (lo_p is not a multiple of stride:
  // the flat walk would read elements pushed by some other CTA b' that may still be running).
  const bool cast = e.shadow != nullptr && (update || (e.mode & 4));
  if (selfmix || cast) {
    float4* g = reinterpret_cast<float4*>(e.grads[r]);
    const float4* rv = reinterpret_cast<const float4*>(e.recv[r]);
    __nv_bfloat16* sh = reinterpret_cast<__nv_bfloat16*>(e.shadow);
    for (int p = 0; p < W; ++p) {
     const long long lo_p = min((long long)p * per, n4), hi_p = min(lo_p + per, n4);
     for (long long i = lo_p + (long long)x.block * kCommThreads + threadIdx.x; i < hi_p; i += stride) {
      float4 val;
      if (selfmix) {
        val = f4_fma(g[i], e.self_coef, rv[i]);
        if (update) {
          float4 p0 = my_params[i], m0 = my_m[i], v0 = e.opt.kind == 0 ? my_v[i] : make_float4(0, 0, 0, 0);
          optim_update4(p0, val, m0, v0, e.opt);
          my_params[i] = p0; my_m[i] = m0; if (e.opt.kind == 0) my_v[i] = v0;
          val = p0;
        } else {
          g[i] = val;
        }
      } else {
        val = update ? my_params[i] : g[i];
      }
      if (cast) store_bf16x4(sh + 4 * i, val);
     }
    }
  }
}

// ================================================================== K-AR (NVLS): reduce in the switch
// Same ownership and tails as the two-shot kernel, but the reduce-scatter leg is ONE multimem.ld_reduce per
// 16 bytes on the multicast mapping of the slab (NVSwitch adds the W copies and returns the sum) and the
// all-gather leg is ONE multimem.st (the switch replicates the store to every rank).  Per GPU and direction
// that is ~S bytes on the links instead of 2*S*(W-1)/W.  The per-rank weight must be uniform (the switch can
// only add); the scale is applied to the sum.  MC=false is the same control flow with the multimem
// instructions spelled out as W peer loads / W peer stores: it runs in the single-GPU emulation (tests).
// Measured on 8 x B200: 760 / 806 / 827 GB/s bus bandwidth at 128 / 256 / 512 MB (two-shot 593 / 622 / 643); slower than the
// two-shot kernel at W = 2 — the Python side uses it from 4 ranks up.
__device__ __forceinline__ float4 mm_ld_reduce_f4(const float4* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void mm_st_f4(float4* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w) : "memory");
}

template <bool MC>
__global__ void __launch_bounds__(kCommThreads, 1)
allreduce_nvls_kernel(const ldnn_comm c, const ldnn_exchange e, float4* const mc_data, float4* const mc_params,
                      const float scale) {
  const Ctx x = make_ctx(c);
  const int W = x.world, r = x.rank;
  const long long n4 = e.n >> 2;
  const long long per = (n4 + W - 1) / W;
  const long long lo = min((long long)r * per, n4), hi = min(lo + per, n4);
  const unsigned int seq = c.epoch * 4u;
  const bool update = (e.mode & LDNN_X_UPDATE) != 0;

  comm_barrier(x, seq + 1);   // every rank's input slab is final
  debug_delay(c, x);

  float4* const my_params = reinterpret_cast<float4*>(e.params[r]);
  float4* const my_m = reinterpret_cast<float4*>(e.m);
  float4* const my_v = reinterpret_cast<float4*>(e.v);
  const long long stride = (long long)x.blocks * kCommThreads;
  for (long long i = lo + (long long)x.block * kCommThreads + threadIdx.x; i < hi; i += stride) {
    float4 s;
    if (MC) {
      s = mm_ld_reduce_f4(mc_data + i);
    } else {
      s = make_float4(0, 0, 0, 0);
#pragma unroll
      for (int p = 0; p < LDNN_MAX_RANKS; ++p)
        if (p < W) s = f4_fma(ld_sys_f4(reinterpret_cast<const float4*>(e.grads[p]) + i), 1.f, s);
    }
    s = f4_scale(s, scale);
    if (update) {
      float4 p0 = my_params[i], m0 = my_m[i], v0 = e.opt.kind == 0 ? my_v[i] : make_float4(0, 0, 0, 0);
      optim_update4(p0, s, m0, v0, e.opt);
      my_m[i] = m0; if (e.opt.kind == 0) my_v[i] = v0;
      s = p0;
    }
    if (MC) {
      mm_st_f4((update ? mc_params : mc_data) + i, s);
    } else {
#pragma unroll
      for (int q = 0; q < LDNN_MAX_RANKS; ++q)
        if (q < W) st_sys_f4(reinterpret_cast<float4*>(update ? e.params[q] : e.grads[q]) + i, s);
    }
  }

  comm_barrier(x, seq + 2);   // every owner's multicast stores have landed everywhere

  const bool cast = e.shadow != nullptr && (update || (e.mode & 4));
  if (cast) {
    const float4* src = update ? my_params : reinterpret_cast<const float4*>(e.grads[r]);
    __nv_bfloat16* sh = reinterpret_cast<__nv_bfloat16*>(e.shadow);
    for (int p = 0; p < W; ++p) {       // owner slice by owner slice, same element-to-CTA mapping as the pushes
      const long long lo_p = min((long long)p * per, n4), hi_p = min(lo_p + per, n4);
      for (long long i = lo_p + (long long)x.block * kCommThreads + threadIdx.x; i < hi_p; i += stride)
        store_bf16x4(sh + 4 * i, src[i]);
    }
  }
}

// ================================================================== K-RING / K-DRING: RS + AG rings
// Region k of the slab belongs to ring k; ring 0 rotates +1 (sends to r+1), ring 1 rotates -1.
// PUSH design: every hop is local loads -> remote *stores* into the next rank's buffer, followed by a
// release flag written into the next rank's signal pad; the consumer polls its OWN pad and reads the
// payload from its OWN memory.  No load ever crosses NVLink, so the ring is not exposed to the ~2.5 us
// peer-load latency (a pull ring needs megabytes of loads in flight to hide it).
//
// Message k of rank r (chunk indices mod W, written for direction +1; c0 = r):
//   k = 0          : send w_r*g[c0]                                 -> next.recv[c0]
//   1 <= k <= W-2  : wait(k-1); send recv[c] + w_r*g[c], c = c0-k   -> next.recv[c]
//   k = W-1        : wait(W-2); c = c0+1 is fully reduced here: (optimizer) -> own slab, -> next slab
//   W <= k <= 2W-3 : wait(k-1); forward the finished chunk c = c0-k+W from own slab -> next slab
//   end            : wait(2W-3) (last chunk has landed), bf16 cast of what arrived
// Flags: ONE word per chain, sig[kBarWords + ring*J + j] in the CONSUMER's pad; the hops of a chain are strictly
// sequential, so the word just counts them: value = epoch*64 + k + 1 after message k (never reset, wrap-safe compare).
//
// A CTA is four independent GROUPS of 128 threads (named barriers 1..4).  Every hop of a sub-chunk ends in
// "all stores issued -> system-scope release -> flag", during which its issuer sends nothing; with four
// chains per SM (and many short sub-chunks instead of a few long ones) another group is always streaming
// while one drains.  With two rings, groups 0,2 of EVERY CTA drive ring 0 and groups 1,3 ring 1: each ring
// gets all SMs of the launch (the two directions use different link directions and only share issue slots).
constexpr int kGroups = 4;
constexpr int kGroupThreads = kCommThreads / kGroups;

__device__ __forceinline__ void group_sync(int grp) {
  asm volatile("bar.sync %0, %1;" ::"r"(grp + 1), "n"(kGroupThreads) : "memory");
}

__global__ void __launch_bounds__(kCommThreads, 1)
ring_fused_kernel(const ldnn_comm c, const ldnn_exchange e, const int num_rings, const int J, const long long sub4) {
  const Ctx x = make_ctx(c);
  const int W = x.world, r = x.rank;
  const long long n4 = e.n >> 2;
  const unsigned int seq = c.epoch * 4u;
  const bool update = (e.mode & LDNN_X_UPDATE) != 0;
  const bool cast = e.shadow != nullptr && (update || (e.mode & 4));

  comm_barrier(x, seq + 1);     // every rank's slab is final and last call's staging is drained

  const int grp = threadIdx.x / kGroupThreads, gt = threadIdx.x % kGroupThreads;
  const int ring = (num_rings == 2) ? (grp & 1) : 0;
  const int groups_per_ring = x.blocks * kGroups / num_rings;                       // chains of this ring, per rank
  const int gir = (num_rings == 2) ? (x.block * (kGroups / 2) + (grp >> 1)) : (x.block * kGroups + grp);
  const int dir = (ring == 0) ? 1 : -1;
  const int next = ((r + dir) % W + W) % W;
  const long long reg_lo = n4 * ring / num_rings, reg_hi = n4 * (ring + 1) / num_rings;
  const long long chunk4 = (reg_hi - reg_lo + W - 1) / W;

  float4* const gr = reinterpret_cast<float4*>(e.grads[r]);
  float4* const pr = reinterpret_cast<float4*>(e.params[r]);
  const float4* const rv = reinterpret_cast<const float4*>(e.recv[r]);
  float4* const nrecv = reinterpret_cast<float4*>(e.recv[next]);
  float4* const nfinal = reinterpret_cast<float4*>(update ? e.params[next] : e.grads[next]);
  float4* const my_final = update ? pr : gr;
  float4* const my_m = reinterpret_cast<float4*>(e.m);
  float4* const my_v = reinterpret_cast<float4*>(e.v);
  __nv_bfloat16* const sh = reinterpret_cast<__nv_bfloat16*>(e.shadow);
  const unsigned int* const my_flags = x.sig[r] + kBarWords;
  unsigned int* const next_flags = x.sig[next] + kBarWords;
  const float wr = e.weight[r];
  const int nmsg = 2 * (W - 1);
  const unsigned int fseq = c.epoch * 64u;      // chain flag after message k of this call = fseq + k + 1  (k < 16)
  bool alive = true;

  for (int k = 0; k <= nmsg && W > 1; ++k) {
    // chunk handled by message k
    int cidx = (k <= W - 1) ? (r - dir * k) : (r - dir * (k - W));
    cidx = ((cidx % W) + W) % W;
    const long long c_lo = min(reg_lo + (long long)cidx * chunk4, reg_hi), c_hi = min(c_lo + chunk4, reg_hi);
    for (int j = gir; j < J; j += groups_per_ring) {
      const long long lo = min(c_lo + (long long)j * sub4, c_hi), hi = min(lo + sub4, c_hi);
      const int fidx = ring * J + j;
      if (k > 0) {
        if (gt == 0 && alive) alive = spin_until(x, my_flags + fidx, fseq + (unsigned int)k);
        group_sync(grp);
      }
      constexpr int U = 4;
      for (long long base = lo + gt; base < hi; base += (long long)U * kGroupThreads) {
        float4 a[U], g[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long long i = base + (long long)u * kGroupThreads;
          if (i >= hi) continue;
          if (k >= W) a[u] = my_final[i];                  // finished chunk that prev stored here
          else {
            g[u] = gr[i];
            if (k > 0) a[u] = rv[i];                       // partial sum that prev stored here
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long long i = base + (long long)u * kGroupThreads;
          if (i >= hi) continue;
          if (k < W - 1) {                                  // reduce-scatter hop
            const float4 acc = (k == 0) ? f4_scale(g[u], wr) : f4_fma(g[u], wr, a[u]);
            st_sys_f4(nrecv + i, acc);
          } else if (k == W - 1) {                          // owner: finish, fused tail, first all-gather hop
            float4 acc = f4_fma(g[u], wr, a[u]);
            if (update) {
              float4 p0 = pr[i], m0 = my_m[i], v0 = e.opt.kind == 0 ? my_v[i] : make_float4(0, 0, 0, 0);
              optim_update4(p0, acc, m0, v0, e.opt);
              my_m[i] = m0; if (e.opt.kind == 0) my_v[i] = v0;
              acc = p0;
            }
            my_final[i] = acc;
            if (cast) store_bf16x4(sh + 4 * i, acc);
            st_sys_f4(nfinal + i, acc);
          } else {                                          // all-gather hop (k == nmsg: last arrival, no send)
            if (cast) store_bf16x4(sh + 4 * i, a[u]);
            if (k < nmsg) st_sys_f4(nfinal + i, a[u]);
          }
        }
      }
      if (k < nmsg) {
        // the group's stores happen-before thread 0's release through the barrier; st.release.sys then orders
        // them (cumulatively) before the flag for the acquiring consumer
        group_sync(grp);
        if (gt == 0) st_release_sys(next_flags + fidx, fseq + (unsigned int)k + 1u);
      }
    }
  }
  if (W == 1 && (update || cast)) {
    // single rank: plain fused optimizer / cast over the whole slab
    for (long long i = (long long)x.block * kCommThreads + threadIdx.x; i < n4; i += (long long)x.blocks * kCommThreads) {
      float4 val = f4_scale(gr[i], e.weight[0]);
      if (update) {
        float4 p0 = pr[i], m0 = my_m[i], v0 = e.opt.kind == 0 ? my_v[i] : make_float4(0, 0, 0, 0);
        optim_update4(p0, val, m0, v0, e.opt);
        pr[i] = p0; my_m[i] = m0; if (e.opt.kind == 0) my_v[i] = v0;
        val = p0;
      }
      if (cast) store_bf16x4(sh + 4 * i, val);
    }
  }
  comm_barrier(x, seq + 3);   // nobody is still storing into a slab that is about to be reused
}

// ================================================================== K-RING / K-DRING, small and medium payloads
// The bulk kernel above pays, per hop, a system-scope release plus a flag round trip (~6 us): 2(W-1) dependent hops
// put its floor near 90 us at 8 GPUs.  Below a few tens of MB the rings therefore run FLAG-FREE: the payload itself
// says when it has arrived.  Every rank owns a staging pad of three phases; a phase is pre-filled with a sentinel
// (-0.0f, a bit pattern the sender never emits: it canonicalises -0 to +0, which sums identically).  A hop is
// one 16-byte store into the next rank's pad; the consumer polls the slot until none of its four words is the
// sentinel (each 32-bit word is single-copy atomic, so no ordering between words, no fence and no flag is needed),
// combines and stores onward.  Threads own element indices, so every element runs its own dependency chain around
// the ring with no CTA-wide synchronisation at all, and there is no kernel-wide barrier: finished chunks travel
// through the pad too (a peer never writes into another rank's gradient/parameter slab, which may still be in use).
// Phases: call t (e.ll_epoch, counted over calls of THIS kernel) uses phase t%3 and re-fills phase (t+2)%3 — the
// one call t-1 used — with the sentinel for call t+2.  A rank cannot finish call t+1 before every peer has started
// it (its last arrivals depend on all ranks' sends of that call), so nobody writes call t+2 data into a phase that
// is still being re-filled.  Same message schedule, chunk ownership, weights and fused tail as the bulk kernel.
__device__ __forceinline__ bool is_sentinel(float f) { return __float_as_uint(f) == 0x80000000u; }
__device__ __forceinline__ float4 no_sentinel(float4 v) {
  if (is_sentinel(v.x)) v.x = 0.f;
  if (is_sentinel(v.y)) v.y = 0.f;
  if (is_sentinel(v.z)) v.z = 0.f;
  if (is_sentinel(v.w)) v.w = 0.f;
  return v;
}
__device__ __forceinline__ bool poll_slot(const Ctx& x, const float4* p, float4& v) {
  v = ld_sys_f4(p);
  if (!(is_sentinel(v.x) | is_sentinel(v.y) | is_sentinel(v.z) | is_sentinel(v.w))) return true;
  const long long t0 = clock64();
  unsigned int ns = 32;
  for (;;) {
    // back off while waiting: a rank that is ahead of its upstream peer (uneven shards, skewed arrival) would otherwise
    // have tens of thousands of threads hammering L2 with probes next to its own backward kernels
    __nanosleep(ns);
    if (ns < 512) ns <<= 1;
    v = ld_sys_f4(p);
    if (!(is_sentinel(v.x) | is_sentinel(v.y) | is_sentinel(v.z) | is_sentinel(v.w))) return true;
    if ((unsigned long long)(clock64() - t0) > x.timeout) {
      if (x.err) atomicExch(x.err, 1);
      return false;
    }
  }
}

__global__ void __launch_bounds__(kCommThreads, 1)
ring_ll_kernel(const ldnn_comm c, const ldnn_exchange e, const int num_rings) {
  const Ctx x = make_ctx(c);
  const int W = x.world, r = x.rank;
  const long long n4 = e.n >> 2;
  const bool update = (e.mode & LDNN_X_UPDATE) != 0;
  const bool cast = e.shadow != nullptr && (update || (e.mode & 4));
  const unsigned int phase = e.ll_epoch % 3u, cphase = (e.ll_epoch + 2u) % 3u;
  const float4 kSent = make_float4(__uint_as_float(0x80000000u), __uint_as_float(0x80000000u),
                                   __uint_as_float(0x80000000u), __uint_as_float(0x80000000u));
  // re-arm the phase the previous call used (local stores; nobody touches it before call t+2)
  {
    float4* cl = reinterpret_cast<float4*>(e.ll[r]) + (long long)cphase * e.ll_phase4;
    for (long long i = (long long)x.block * kCommThreads + threadIdx.x; i < e.ll_clear4;
         i += (long long)x.blocks * kCommThreads)
      cl[i] = kSent;
  }
  // threads of a CTA split between the rings, so every CTA drives both directions
  const int tpr = kCommThreads / num_rings;                        // threads per ring in this CTA
  const int ring = threadIdx.x / tpr, t = threadIdx.x % tpr;
  const int dir = (ring == 0) ? 1 : -1;
  const int next = ((r + dir) % W + W) % W;
  const long long reg_lo = n4 * ring / num_rings, reg_hi = n4 * (ring + 1) / num_rings;
  const long long chunk4 = (reg_hi - reg_lo + W - 1) / W;
  const long long slot4 = ((n4 + num_rings - 1) / num_rings + W - 1) / W;       // slot stride: the larger ring's chunk
  const int nmsg = 2 * (W - 1);

  float4* const gr = reinterpret_cast<float4*>(e.grads[r]);
  float4* const pr = reinterpret_cast<float4*>(e.params[r]);
  float4* const my_final = update ? pr : gr;
  float4* const my_m = reinterpret_cast<float4*>(e.m);
  float4* const my_v = reinterpret_cast<float4*>(e.v);
  __nv_bfloat16* const sh = reinterpret_cast<__nv_bfloat16*>(e.shadow);
  const float4* const my_pad = reinterpret_cast<const float4*>(e.ll[r]) + (long long)phase * e.ll_phase4 +
                               (long long)ring * nmsg * slot4;
  float4* const next_pad = reinterpret_cast<float4*>(e.ll[next]) + (long long)phase * e.ll_phase4 +
                           (long long)ring * nmsg * slot4;
  const float wr = e.weight[r];
  bool alive = true;

  for (int k = 0; k <= nmsg && W > 1; ++k) {
    int cidx = (k <= W - 1) ? (r - dir * k) : (r - dir * (k - W));
    cidx = ((cidx % W) + W) % W;
    const long long c_lo = min(reg_lo + (long long)cidx * chunk4, reg_hi), c_hi = min(c_lo + chunk4, reg_hi);
    const long long len = c_hi - c_lo;
    constexpr int U = 4;                       // elements per thread in flight: polls and stores of a batch overlap
    for (long long base = (long long)x.block * tpr + t; base < len; base += (long long)U * x.blocks * tpr) {
      float4 a[U], g[U];
      // first probe of every slot of the batch and the local gradient loads go out together; in steady state the
      // upstream rank is a whole step ahead, so the probes hit and the spin loop below is not entered
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long l = base + (long long)u * x.blocks * tpr;
        if (l >= len) continue;
        if (k > 0) a[u] = ld_sys_f4(my_pad + (long long)(k - 1) * slot4 + l);
        if (k < W) g[u] = gr[c_lo + l];
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long l = base + (long long)u * x.blocks * tpr;
        if (l >= len) continue;
        if (k > 0 && (is_sentinel(a[u].x) | is_sentinel(a[u].y) | is_sentinel(a[u].z) | is_sentinel(a[u].w))) {
          if (alive) alive = poll_slot(x, my_pad + (long long)(k - 1) * slot4 + l, a[u]);
          if (!alive) a[u] = make_float4(0, 0, 0, 0);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long l = base + (long long)u * x.blocks * tpr;
        if (l >= len) continue;
        const long long i = c_lo + l;
        float4 out;
        if (k < W - 1) {                                    // reduce-scatter hop
          out = (k == 0) ? f4_scale(g[u], wr) : f4_fma(g[u], wr, a[u]);
        } else if (k == W - 1) {                            // owner: finish + fused tail
          out = f4_fma(g[u], wr, a[u]);
          if (update) {
            float4 p0 = pr[i], m0 = my_m[i], v0 = e.opt.kind == 0 ? my_v[i] : make_float4(0, 0, 0, 0);
            optim_update4(p0, out, m0, v0, e.opt);
            my_m[i] = m0; if (e.opt.kind == 0) my_v[i] = v0;
            out = p0;
          }
          out = no_sentinel(out);
          my_final[i] = out;
          if (cast) store_bf16x4(sh + 4 * i, out);
        } else {                                            // all-gather hop: keep and forward what arrived
          out = a[u];
          my_final[i] = out;
          if (cast) store_bf16x4(sh + 4 * i, out);
        }
        if (k < nmsg) st_sys_f4(next_pad + (long long)k * slot4 + l, no_sentinel(out));
      }
    }
  }
  if (W == 1 && (update || cast)) {
    for (long long i = (long long)x.block * kCommThreads + threadIdx.x; i < n4; i += (long long)x.blocks * kCommThreads) {
      float4 val = f4_scale(gr[i], e.weight[0]);
      if (update) {
        float4 p0 = pr[i], m0 = my_m[i], v0 = e.opt.kind == 0 ? my_v[i] : make_float4(0, 0, 0, 0);
        optim_update4(p0, val, m0, v0, e.opt);
        pr[i] = p0; my_m[i] = m0; if (e.opt.kind == 0) my_v[i] = v0;
        val = p0;
      }
      if (cast) store_bf16x4(sh + 4 * i, val);
    }
  }
}

// ================================================================== compat gossip (1- or 2-hop pull)
__global__ void __launch_bounds__(kCommThreads, 1)
gossip_pull_kernel(const ldnn_comm c, const Peers bufs, const Peers recv, const long long n4, const float c_self,
                   const int n_off, const int off0, const int off1, const float c0, const float c1) {
  const Ctx x = make_ctx(c);
  const int W = x.world, r = x.rank;
  const unsigned int seq = c.epoch * 4u;
  comm_barrier(x, seq + 1);
  const float4* mine = reinterpret_cast<const float4*>(bufs.p[r]);
  const float4* p0 = reinterpret_cast<const float4*>(bufs.p[((r - off0) % W + W) % W]);
  const float4* p1 = reinterpret_cast<const float4*>(bufs.p[((r - off1) % W + W) % W]);
  float4* out = reinterpret_cast<float4*>(recv.p[r]);
  const long long stride = (long long)x.blocks * kCommThreads;
#pragma unroll 4
  for (long long i = (long long)x.block * kCommThreads + threadIdx.x; i < n4; i += stride) {
    float4 acc = f4_scale(mine[i], c_self);
    if (n_off > 0) acc = f4_fma(ld_sys_f4(p0 + i), c0, acc);
    if (n_off > 1) acc = f4_fma(ld_sys_f4(p1 + i), c1, acc);
    out[i] = acc;
  }
  comm_barrier(x, seq + 2);   // every reader is done with the old values: safe to overwrite in place
  float4* dst = reinterpret_cast<float4*>(bufs.p[r]);
  for (long long i = (long long)x.block * kCommThreads + threadIdx.x; i < n4; i += stride) dst[i] = out[i];
}

// ================================================================== K-BCAST
__global__ void __launch_bounds__(kCommThreads, 1)
broadcast_kernel(const ldnn_comm c, const Peers bufs, const long long n16, const int src) {
  const Ctx x = make_ctx(c);
  const unsigned int seq = c.epoch * 4u;
  comm_barrier(x, seq + 1);
  if (x.rank != src) {
    const float4* s = reinterpret_cast<const float4*>(bufs.p[src]);
    float4* d = reinterpret_cast<float4*>(bufs.p[x.rank]);
    const long long stride = (long long)x.blocks * kCommThreads;
    for (long long i = (long long)x.block * kCommThreads + threadIdx.x; i < n16; i += 4 * stride) {
      float4 t[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) if (i + u * stride < n16) t[u] = ld_sys_f4(s + i + u * stride);
#pragma unroll
      for (int u = 0; u < 4; ++u) if (i + u * stride < n16) d[i + u * stride] = t[u];
    }
  }
  comm_barrier(x, seq + 2);
}

// ================================================================== K-MET
__global__ void __launch_bounds__(kCommThreads, 1)
allgather_small_kernel(const ldnn_comm c, const float* mine, const Peers tables, const int count) {
  const Ctx x = make_ctx(c);
  const unsigned int seq = c.epoch * 4u;
  const float* src = c.emulate ? mine + (size_t)x.rank * count : mine;
  comm_barrier(x, seq + 1);
  if (x.block == 0)
    for (int q = 0; q < x.world; ++q)
      for (int i = threadIdx.x; i < count; i += kCommThreads)
        reinterpret_cast<float*>(tables.p[q])[(size_t)x.rank * count + i] = src[i];
  comm_barrier(x, seq + 2);
}

__global__ void __launch_bounds__(kCommThreads, 1) barrier_kernel(const ldnn_comm c) {
  const Ctx x = make_ctx(c);
  comm_barrier(x, c.epoch * 4u + 1);
}

// ------------------------------------------------------------------ launch helper
template <typename... Args>
static int launch_comm(void (*kernel)(Args...), const ldnn_comm& c, cudaStream_t st, Args... args) {
  LDNN_ARG(c.world >= 1 && c.world <= LDNN_MAX_RANKS, "world=%d", c.world);
  LDNN_ARG(c.blocks >= 1 && c.blocks <= 512, "blocks=%d", c.blocks);
  dim3 grid(c.blocks, c.emulate ? c.world : 1);
  if (c.emulate) {
    void* params[] = {(void*)&args...};
    LDNN_CUDA(cudaLaunchCooperativeKernel((void*)kernel, grid, dim3(kCommThreads), params, 0, st));
  } else {
    kernel<<<grid, kCommThreads, 0, st>>>(args...);
    LDNN_CUDA(cudaGetLastError());
  }
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return LDNN_OK;
}

}  // namespace ldnn

using namespace ldnn;

// ------------------------------------------------------------------ symmetric memory (CUDA IPC)
extern "C" int ldnn_symm_alloc(size_t bytes, void** ptr, unsigned char* handle_out) {
  LDNN_CUDA(cudaMalloc(ptr, bytes));
  LDNN_CUDA(cudaMemset(*ptr, 0, bytes));
  if (handle_out) {
    cudaIpcMemHandle_t h;
    LDNN_CUDA(cudaIpcGetMemHandle(&h, *ptr));
    static_assert(sizeof(h) == LDNN_IPC_HANDLE_BYTES, "ipc handle size");
    memcpy(handle_out, &h, sizeof(h));
  }
  return LDNN_OK;
}
extern "C" int ldnn_symm_open(const unsigned char* handle, void** peer_ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  LDNN_CUDA(cudaIpcOpenMemHandle(peer_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return LDNN_OK;
}
extern "C" int ldnn_symm_close(void* peer_ptr) { LDNN_CUDA(cudaIpcCloseMemHandle(peer_ptr)); return LDNN_OK; }
extern "C" int ldnn_symm_free(void* ptr) { LDNN_CUDA(cudaFree(ptr)); return LDNN_OK; }
extern "C" int ldnn_enable_peer_access(int my_device, int peer_device) {
  int can = 0;
  LDNN_CUDA(cudaDeviceCanAccessPeer(&can, my_device, peer_device));
  LDNN_ARG(can, "device %d cannot access peer %d", my_device, peer_device);
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return LDNN_OK; }
  LDNN_CUDA(e);
  return LDNN_OK;
}

// ------------------------------------------------------------------ exchange entry points
extern "C" int ldnn_allreduce_fused(const ldnn_comm* c, const ldnn_exchange* x, void* stream) {
  LDNN_ARG(x->n % 4 == 0, "allreduce: n=%lld must be a multiple of 4", x->n);
  LDNN_ARG(!(x->mode & LDNN_X_UPDATE) || (x->m && (x->opt.kind == 1 || x->v)), "allreduce: UPDATE needs m (and v)");
  return launch_comm(allreduce_fused_kernel, *c, (cudaStream_t)stream, *c, *x);
}

extern "C" int ldnn_allreduce_nvls(const ldnn_comm* c, const ldnn_exchange* x, void* mc_data, void* mc_params,
                                   void* stream) {
  LDNN_ARG(x->n % 4 == 0, "nvls allreduce: n=%lld must be a multiple of 4", x->n);
  LDNN_ARG(!(x->mode & LDNN_X_SELFMIX), "nvls allreduce: self-weighted mixing needs the two-shot kernel");
  LDNN_ARG(!(x->mode & LDNN_X_UPDATE) || (x->m && (x->opt.kind == 1 || x->v)), "nvls allreduce: UPDATE needs m (and v)");
  for (int p = 1; p < c->world; ++p)
    LDNN_ARG(x->weight[p] == x->weight[0], "nvls allreduce: per-rank weights must be uniform (rank %d: %g vs %g)", p,
             (double)x->weight[p], (double)x->weight[0]);
  const bool mc = !c->emulate;
  LDNN_ARG(!mc || (mc_data && (!(x->mode & LDNN_X_UPDATE) || mc_params)), "nvls allreduce: multicast mapping missing");
  LDNN_ARG(!mc || (((uintptr_t)mc_data | (uintptr_t)mc_params) & 15) == 0, "nvls allreduce: multicast pointers must be 16-byte aligned");
  if (mc)
    return launch_comm(allreduce_nvls_kernel<true>, *c, (cudaStream_t)stream, *c, *x, (float4*)mc_data, (float4*)mc_params,
                       x->weight[0]);
  return launch_comm(allreduce_nvls_kernel<false>, *c, (cudaStream_t)stream, *c, *x, (float4*)nullptr, (float4*)nullptr,
                     x->weight[0]);
}

// sub-chunks of the bulk ring: one per hand-off chain.  Every group of 128 threads runs its own chains (4 groups per
// CTA).  Under load a hand-off (system-scope release, which waits for the group's remote stores to be acknowledged, +
// flag + the consumer's poll) costs ~10-13 us and is serial within a group, so the payload is cut into as FEW chains as
// keep all groups busy: `chains` per group and step (default 1; LDNN_RING_CHAINS), never shorter than 16 KB.  Measured
// at 128 MB: 8 x B200, 8 chains of 8 KB per group 158 GB/s, 2 of 32 KB 484 GB/s; 4 x B200, 4 / 2 / 1 chains per group
// 458 / 476 / 500 GB/s (profiles/ring_chains_4gpu.txt) — hand-off bound in every case, so fewer is better.
// Host-only arithmetic (no GPU needed): J chains of sub4 float4 each cover one rank's chunk of one ring; the flag
// table (one word per chain and ring) caps J.
extern "C" int ldnn_ring_plan(long long n, int world, int num_rings, int blocks, int chains, int* J_out,
                              long long* sub4_out) {
  LDNN_ARG(n >= 0 && n % 4 == 0 && world >= 1 && (num_rings == 1 || num_rings == 2) && blocks >= 1 && chains >= 1,
           "ring plan: n=%lld world=%d rings=%d blocks=%d chains=%d", n, world, num_rings, blocks, chains);
  const long long n4 = n / 4;
  const long long chunk4 = (n4 / num_rings + world) / world + 1;
  const int gpr = std::max(1, blocks * kGroups / num_rings);
  long long sub4 = std::max<long long>(1024, (chunk4 + (long long)chains * gpr - 1) / ((long long)chains * gpr));
  int J = (int)((chunk4 + sub4 - 1) / sub4);
  const int maxJ = kFlagWords / 2;
  if (J > maxJ) { J = maxJ; sub4 = (chunk4 + J - 1) / J; }
  if (J < 1) J = 1;
  *J_out = J; *sub4_out = sub4;
  return LDNN_OK;
}

extern "C" int ldnn_ring_fused(const ldnn_comm* c, const ldnn_exchange* x, int num_rings, void* stream) {
  LDNN_ARG(x->n % 4 == 0, "ring: n %% 4");
  LDNN_ARG(num_rings == 1 || num_rings == 2, "ring: num_rings=%d", num_rings);
  LDNN_ARG(!(x->mode & LDNN_X_SELFMIX), "ring: self-weighted mixing is only available on the all-reduce schedule");
  LDNN_ARG(c->blocks >= num_rings, "ring: need >= %d blocks", num_rings);
  LDNN_ARG(c->world == 1 || x->recv[c->emulate ? 0 : c->rank] != nullptr, "ring: needs the symmetric `recv` staging slab");
  static int chains = 0;
  if (chains == 0) { const char* e = getenv("LDNN_RING_CHAINS"); chains = e ? std::max(1, atoi(e)) : 1; }
  int J = 1;
  long long sub4 = 0;
  ldnn_ring_plan(x->n, c->world, num_rings, c->blocks, chains, &J, &sub4);
  return launch_comm(ring_fused_kernel, *c, (cudaStream_t)stream, *c, *x, num_rings, J, sub4);
}

extern "C" long long ldnn_ring_ll_slots(long long n, int world, int num_rings) {
  const long long n4 = n / 4;
  const long long slot4 = ((n4 + num_rings - 1) / num_rings + world - 1) / world;
  return (long long)num_rings * 2 * (world - 1) * slot4;
}

extern "C" int ldnn_ring_ll(const ldnn_comm* c, const ldnn_exchange* x, int num_rings, void* stream) {
  LDNN_ARG(x->n % 4 == 0, "ring_ll: n %% 4");
  LDNN_ARG(num_rings == 1 || num_rings == 2, "ring_ll: num_rings=%d", num_rings);
  LDNN_ARG(!(x->mode & LDNN_X_SELFMIX), "ring_ll: self-weighted mixing is only available on the all-reduce schedule");
  LDNN_ARG(!(x->mode & LDNN_X_UPDATE) || (x->m && (x->opt.kind == 1 || x->v)), "ring_ll: UPDATE needs m (and v)");
  LDNN_ARG(c->world == 1 || x->ll[c->emulate ? 0 : c->rank] != nullptr, "ring_ll: needs the symmetric staging pad `ll`");
  const long long used = ldnn_ring_ll_slots(x->n, c->world, num_rings);
  LDNN_ARG(used <= x->ll_phase4 && x->ll_clear4 <= x->ll_phase4,
           "ring_ll: payload needs %lld staging slots per phase, the pad holds %lld", used, x->ll_phase4);
  // few elements -> few CTAs (every element's chain is independent; idle CTAs would only add launch cost)
  ldnn_comm cc = *c;
  const long long slot4 = used / std::max(1, num_rings * 2 * (c->world - 1));
  const long long clear_ctas = (x->ll_clear4 + kCommThreads * 8 - 1) / (kCommThreads * 8);
  const long long want = std::max<long long>((slot4 * num_rings + kCommThreads - 1) / kCommThreads, clear_ctas);
  cc.blocks = (int)std::max<long long>(1, std::min<long long>(c->blocks, want));
  return launch_comm(ring_ll_kernel, cc, (cudaStream_t)stream, cc, *x, num_rings);
}

extern "C" int ldnn_gossip_pull(const ldnn_comm* c, void* const* bufs, void* const* recv, long long n, float c_self,
                                int n_off, const int* offsets, const float* coefs, void* stream) {
  LDNN_ARG(n % 4 == 0 && n_off >= 0 && n_off <= 2, "gossip: n=%lld n_off=%d", n, n_off);
  Peers b, r;
  for (int i = 0; i < LDNN_MAX_RANKS; ++i) { b.p[i] = i < c->world ? bufs[i] : nullptr; r.p[i] = i < c->world ? recv[i] : nullptr; }
  return launch_comm(gossip_pull_kernel, *c, (cudaStream_t)stream, *c, b, r, n / 4, c_self, n_off,
                     n_off > 0 ? offsets[0] : 0, n_off > 1 ? offsets[1] : 0, n_off > 0 ? coefs[0] : 0.f,
                     n_off > 1 ? coefs[1] : 0.f);
}

extern "C" int ldnn_broadcast(const ldnn_comm* c, void* const* bufs, size_t bytes, int src, void* stream) {
  LDNN_ARG(bytes % 16 == 0, "broadcast: bytes %% 16");
  Peers b;
  for (int i = 0; i < LDNN_MAX_RANKS; ++i) b.p[i] = i < c->world ? bufs[i] : nullptr;
  return launch_comm(broadcast_kernel, *c, (cudaStream_t)stream, *c, b, (long long)(bytes / 16), src);
}

extern "C" int ldnn_allgather_small(const ldnn_comm* c, const float* mine, void* const* tables, int count,
                                    void* stream) {
  Peers t;
  for (int i = 0; i < LDNN_MAX_RANKS; ++i) t.p[i] = i < c->world ? tables[i] : nullptr;
  return launch_comm(allgather_small_kernel, *c, (cudaStream_t)stream, *c, mine, t, count);
}

extern "C" int ldnn_barrier(const ldnn_comm* c, void* stream) {
  return launch_comm(barrier_kernel, *c, (cudaStream_t)stream, *c);
}
