// Implicit-GEMM convolution on tcgen05 / TMEM / TMA  (K-CONV-F / -D / -W).
//
// Replaces the reference's cuDNN convolutions (Balanced All-Reduce/model.py:55,57,63,78 and
// their autograd backward, trainer.py:208).  One warp-specialised kernel, three modes:
//
//   FPROP  y[pix, co]   = sum_{tap, ci} x[pix (+) tap, ci] * w[co, tap, ci]
//   DGRAD  dx[pix, ci]  = sum_{tap, co} dy[pix (-) tap, co] * w[co, tap, ci]      (B read MN-major: no transposed copy)
//   WGRAD  dw[co,tap,ci]= sum_{pix}     dy[pix, co] * x[pix (+) tap, ci]           (A and B read MN-major)
//
// Two element types, one kernel template (ET):
//   bf16  — activations NHWC bf16, weights KRSC bf16, tcgen05.mma kind::f16  (UMMA_K = 16)
//   float — activations NHWC fp32, weights KRSC fp32 read straight from the master slab, kind::tf32 (UMMA_K = 8):
//           the reference's own precision (fp32 storage, TF32 convolutions — torch's default cudnn.allow_tf32).
// A 128-byte swizzle row holds KE = 64 bf16 / 32 fp32 elements, so a pipeline stage carries one KE-wide K block:
// same bytes, same 4 MMAs per stage, same shared-memory descriptors in both precisions.
// Activations are NHWC, weights KRSC ("[Cout][R][S][Cin]").  There is no im2col
// buffer: for every filter tap the A tile is a *shifted 4-D TMA box* of the activation tensor
// (channels x w x h x n = KE x bw x bh x bn, bw*bh*bn pixels), and the TMA unit's out-of-bounds
// zero fill provides the padding.  Stride-2 convolutions use four parity views of the input
// (one tensor map per (h%2, w%2)), which turns them into the same shifted-box loads.
// Accumulators live in TMEM (128 lanes x BLOCK_N fp32 columns); a single elected thread issues
// tcgen05.mma; the epilogue warps read TMEM with tcgen05.ld and write bf16 NHWC / fp32 KRSC.
// Small-M / huge-K layers (layer3/4 at batch 64) use split-K: partial tiles go to an fp32
// scratch slab and the last CTA to arrive on a tile (atomic ticket) reduces them in a fixed
// order, so results are deterministic and no memset / atomics on the output are needed.
#pragma once
#include "ptx.cuh"
#include "common.cuh"

namespace ldnn {

enum ConvMode : int { MODE_FPROP = 0, MODE_DGRAD = 1, MODE_WGRAD = 2 };

struct TapEntry {
  int8_t map;   // which shifted-operand tensor map (parity view) this tap reads
  int8_t dh;    // box shift in h (in units of the view's rows)
  int8_t dw;    // box shift in w
  int8_t wtap;  // tap index inside the weight tensor (r*S+s)
};

// One output "class".  fprop / wgrad have a single class; a stride-2 dgrad has one class per output
// parity (h%2, w%2) — all of them run in ONE launch (CTAs find their class from blockIdx.x).
struct ConvClass {
  int m_begin;             // first blockIdx.x of this class
  int num_taps;
  int tiles_w, tiles_h;    // boxes along w and h of this class's pixel grid
  int Wd, Hd;              // extents of the pixel grid (store masks)
  int iters_total;         // fprop/dgrad: num_taps*kb_per_tap; wgrad: number of pixel blocks
  int iters_per_split;
  int splits;
  int pad_;
  long long out_base;      // element offset of pixel (0,0,0) of this class
  TapEntry taps[12];
};

struct ConvParams {
  int bw, bh, bn;          // pixel box (w, h, n); bw*bh*bn = 128 (fprop/dgrad M tile) or KE (wgrad K block)
  int Nd;                  // batch (store mask)
  int kb_per_tap;          // fprop/dgrad: reduction channels / KE
  int red_ch;              // fprop: Cin, dgrad: Cout (weight column stride per tap); wgrad: Cout (row mask)
  int n_total;             // total N of the GEMM (fprop: Cout, dgrad: Cin, wgrad: Cin)
  int num_classes;
  int max_splits;          // grid.z (fprop/dgrad) — classes may use fewer
  int accumulate;          // epilogue adds `addend` (same addressing as out)
  int atomic_out;          // wgrad split-K: red.add the fp32 tile straight into the (pre-zeroed) output
  long long sN, sH, sW;    // output strides in elements (fprop/dgrad); wgrad: sN = row stride (T*Cin)
  uint32_t a_lbo, a_sbo, b_lbo, b_sbo;   // smem descriptor strides (bytes)
  ConvClass cls[4];
};

// dgrad only: the BN-backward reduction of the layer that OWNS dx (dy = dx * (out > 0);
// sums += (sum dy, sum dy * xhat)) is folded into this epilogue when this dgrad is the last
// contribution to dx — one launch fewer per layer and dx is not re-read.
struct BnBwdFuse {
  const void* out;             // post-activation output of the owning layer (ReLU mask), same shape/type as dx
  const void* y;               // its pre-BN conv output
  const float* mean;
  const float* invstd;
  float* sums;                 // [2][C]
  int relu;
};

struct ConvMaps {
  CUtensorMap act[4];   // shifted operand (x for fprop/wgrad, dy for dgrad), one per parity view
  CUtensorMap other;    // weights (fprop/dgrad, 2-D) or dy (wgrad, 4-D)
};

// element-type traits of the MMA operands
template <typename ET> struct ElemTraits;
template <> struct ElemTraits<__nv_bfloat16> {
  static constexpr int kKE = 64;            // elements per 128-byte swizzle row = K block of a stage
  static constexpr int kUmmaK = 16;
  static constexpr bool kTf32 = false;
};
template <> struct ElemTraits<float> {
  static constexpr int kKE = 32;
  static constexpr int kUmmaK = 8;
  static constexpr bool kTf32 = true;
};

template <int BLOCK_N>
struct ConvCfg {
  static constexpr int kStages = (BLOCK_N == 256) ? 4 : (BLOCK_N == 128 ? 6 : 8);
  static constexpr int kABytes = 128 * 64 * 2;
  static constexpr int kBBytes = BLOCK_N * 64 * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;
  // 'lite' launch for large grids of short-K tiles: ~96 KB of stages and 4 epilogue warps, so two CTAs
  // share an SM and one CTA's prologue/epilogue hides under the other's mainloop
  static constexpr int kLiteStages = (BLOCK_N == 256) ? 2 : (BLOCK_N == 128 ? 3 : 4);
  static constexpr int kLiteSmemBytes = kLiteStages * kStageBytes + 1024 + 256;
  static constexpr int kLiteThreads = 64 + 32 * 4;
  static constexpr int kEpiWarps = 8;          // two warps per TMEM lane quarter, each takes half of the columns
  static constexpr int kThreads = 64 + 32 * kEpiWarps;
};

// CL = thread-block cluster size along the M-tile axis (1 or 2).  With CL = 2 the two CTAs work on
// neighbouring M tiles of the same N tile and K range: each loads HALF of every weight (B) tile and TMA
// multicasts it into both CTAs' shared memory, halving the L2 reads of the operand both need; a smem
// stage is recycled only when both CTAs' MMAs have drained it (multicast tcgen05.commit).
// LITE = short pipeline + 4 epilogue warps, two CTAs per SM (large grids of short-K tiles).
// CL = -2 ("pair"; run on hardware in round 2, both precisions): the two CTAs of a TPC run ONE tcgen05.mma.cta_group::2 of UMMA_M = 256:
// each CTA loads its own A box and only ITS HALF of the B tile (no multicast: half the B bytes enter each SM), the
// leader CTA issues the MMAs for both, each CTA's TMEM receives its own 128 rows and its epilogue is unchanged.
template <typename ET, int MODE, int BLOCK_N, int CL, bool LITE>
__global__ void __launch_bounds__(LITE ? 192 : 320, LITE ? 2 : 1)
conv_gemm_kernel(const __grid_constant__ ConvMaps maps, const ConvParams p, void* __restrict__ out,
                 const void* __restrict__ addend, float* __restrict__ scratch, unsigned int* __restrict__ tickets,
                 float* __restrict__ stats, const BnBwdFuse fuse) {
  using Cfg = ConvCfg<BLOCK_N>;
  using ETr = ElemTraits<ET>;
  constexpr int KE = ETr::kKE;                  // K block of a stage (and width of an MN-major box), in elements
  constexpr int kMnBox = 128 * KE;              // bytes of one MN-major box: KE (MN) x KE (K) elements = 128 B x KE rows
  constexpr bool PAIR = CL < 0;                 // cta_group::2 pair
  constexpr int CLN = CL < 0 ? -CL : CL;        // CTAs per cluster
  static_assert(!PAIR || (CLN == 2 && MODE != MODE_WGRAD && BLOCK_N >= 128 && !LITE), "pair: fprop/dgrad, N >= 128");
  constexpr int kStages = LITE ? Cfg::kLiteStages : Cfg::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kStages * Cfg::kStageBytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_full + 1);
  uint32_t* last_flag = tmem_ptr + 1;

  const int warp = threadIdx.x >> 5;
  int cid = 0;
  if (MODE == MODE_DGRAD) {
    while (cid + 1 < p.num_classes && (int)blockIdx.x >= p.cls[cid + 1].m_begin) ++cid;
  }
  const ConvClass& kc = p.cls[cid];
  const int m_tile = blockIdx.x - kc.m_begin, n_tile = blockIdx.y;
  int split = blockIdx.z, tap_fixed = 0;
  if (MODE == MODE_WGRAD) { tap_fixed = blockIdx.z / kc.splits; split = blockIdx.z % kc.splits; }
  if (split >= kc.splits) return;              // class with fewer splits than grid.z (whole CTA leaves)
  const int it_begin = split * kc.iters_per_split;
  const int it_end = min(it_begin + kc.iters_per_split, kc.iters_total);
  const int n0 = n_tile * BLOCK_N;

  if (warp == 0 && elect_one()) {
    for (int i = 0; i < 4; ++i) prefetch_tmap(&maps.act[i]);
    prefetch_tmap(&maps.other);
  }
  if (warp == 1) {
    if (elect_one()) {
      // pair: one multicast commit per phase frees a stage in both CTAs; multicast cluster: one commit per CTA
      for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], PAIR ? 1 : CLN); }
      mbar_init(tmem_full, 1);
      fence_mbar_init();
    }
    __syncwarp();
    if (!PAIR) tmem_alloc<BLOCK_N>(tmem_ptr);
  }
  if (PAIR) {
    cluster_sync_all();                          // both CTAs are running and their barriers exist
    if (warp == 1) tmem_alloc_2sm<BLOCK_N>(tmem_ptr);   // same warp of both CTAs; same column offset in both
    tc_fence_before();
    __syncthreads();
  } else {
    tc_fence_before();
    if (CLN > 1) cluster_sync_all(); else __syncthreads();     // peers' barriers exist before anyone signals them
  }
  tc_fence_after();
  const uint32_t crank = (CLN > 1) ? cluster_ctarank() : 0u;
  constexpr uint16_t kMcMask = (1u << CLN) - 1u;
  const uint32_t tmem_base = *tmem_ptr;
  // PDL: everything above overlapped the previous kernel's tail; from here on we touch its outputs
  pdl_wait();
  pdl_launch_dependents();

  // pixel-tile coordinates (fprop/dgrad: the M tile; wgrad: decoded per K block)
  auto decode = [&](int t, int& w0, int& h0, int& nn0) {
    int wt = t % kc.tiles_w;
    int ht = (t / kc.tiles_w) % kc.tiles_h;
    int nt = t / (kc.tiles_w * kc.tiles_h);
    w0 = wt * p.bw; h0 = ht * p.bh; nn0 = nt * p.bn;
  };

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      int stage = 0; uint32_t phase = 0;
      int w0 = 0, h0 = 0, nn0 = 0;
      if (MODE != MODE_WGRAD) decode(m_tile, w0, h0, nn0);
      for (int it = it_begin; it < it_end; ++it) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sA = smem + stage * Cfg::kStageBytes;
        uint8_t* sB = sA + Cfg::kABytes;
        if (PAIR) {
          // both CTAs' boxes (own A + own half of B, each) are credited to the LEADER's barrier; only it arms it
          if (crank == 0) mbar_expect_tx(&full_bar[stage], 2 * (Cfg::kABytes + Cfg::kBBytes / 2));
        } else {
          mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
        }
        if (PAIR) {
          const int tap = it / p.kb_per_tap, kb = it - tap * p.kb_per_tap;
          const TapEntry te = kc.taps[tap];
          tma_load_4d_2sm(sA, &maps.act[te.map], &full_bar[stage], kb * KE, w0 + te.dw, h0 + te.dh, nn0);
          if (MODE == MODE_FPROP) {
            // my BLOCK_N/2 weight rows, at the SAME smem offset in both CTAs (the MMA reads half of B from each)
            tma_load_2d_2sm(sB, &maps.other, &full_bar[stage], te.wtap * p.red_ch + kb * KE,
                            n0 + (int)crank * (BLOCK_N / 2));
          } else {
#pragma unroll
            for (int jj = 0; jj < BLOCK_N / (2 * KE); ++jj) {       // my half of the KE-wide ci boxes
              const int j = (int)crank * (BLOCK_N / (2 * KE)) + jj;
              tma_load_2d_2sm(sB + jj * kMnBox, &maps.other, &full_bar[stage], te.wtap * p.n_total + n0 + j * KE, kb * KE);
            }
          }
        } else if (MODE == MODE_FPROP || MODE == MODE_DGRAD) {
          const int tap = it / p.kb_per_tap, kb = it - tap * p.kb_per_tap;
          const TapEntry te = kc.taps[tap];
          tma_load_4d(sA, &maps.act[te.map], &full_bar[stage], kb * KE, w0 + te.dw, h0 + te.dh, nn0);
          if (MODE == MODE_FPROP) {
            // weights [Cout rows][taps*Cin cols], K-major: one box of KE cols x BLOCK_N rows
            if (CLN == 1) {
              tma_load_2d(sB, &maps.other, &full_bar[stage], te.wtap * p.red_ch + kb * KE, n0);
            } else {          // my half of the rows, delivered to both CTAs
              constexpr int kHalfRows = BLOCK_N / CLN;
              tma_load_2d_mc(sB + crank * kHalfRows * 128, &maps.other, &full_bar[stage],
                             te.wtap * p.red_ch + kb * KE, n0 + (int)crank * kHalfRows, kMcMask);
            }
          } else {
            // weights [Cout rows][taps*Cin cols]; B[n=ci, k=co] is MN-major: boxes of KE ci x KE co
#pragma unroll
            for (int j = 0; j < BLOCK_N / KE; ++j) {
              if (CLN == 1)
                tma_load_2d(sB + j * kMnBox, &maps.other, &full_bar[stage], te.wtap * p.n_total + n0 + j * KE, kb * KE);
              else if ((j % CLN) == (int)crank)      // boxes dealt round-robin to the cluster's CTAs
                tma_load_2d_mc(sB + j * kMnBox, &maps.other, &full_bar[stage], te.wtap * p.n_total + n0 + j * KE,
                               kb * KE, kMcMask);
            }
          }
        } else {
          decode(it, w0, h0, nn0);
          const TapEntry te = kc.taps[tap_fixed];
          // A = dy[pix][co]  (MN-major, 128/KE boxes of KE channels) ; B = x[pix (+) tap][ci] (MN-major)
#pragma unroll
          for (int j = 0; j < 128 / KE; ++j)
            tma_load_4d(sA + j * kMnBox, &maps.other, &full_bar[stage], m_tile * 128 + j * KE, w0, h0, nn0);
#pragma unroll
          for (int j = 0; j < BLOCK_N / KE; ++j)
            tma_load_4d(sB + j * kMnBox, &maps.act[te.map], &full_bar[stage], n0 + j * KE, w0 + te.dw, h0 + te.dh, nn0);
        }
        if (++stage == kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = make_idesc(ETr::kTf32 ? 2u : 1u, PAIR ? 256 : 128, BLOCK_N, MODE == MODE_WGRAD ? 1u : 0u,
                                          MODE == MODE_FPROP ? 0u : 1u);
    // one MMA consumes UMMA_K of K: 32 bytes along a K-major row, UMMA_K 128-byte rows of an MN-major box
    constexpr uint32_t kMnStep = ETr::kUmmaK * 128u;
    constexpr uint32_t a_kstep = (MODE == MODE_WGRAD) ? kMnStep : 32u;
    constexpr uint32_t b_kstep = (MODE == MODE_FPROP) ? 32u : kMnStep;
    // MN-major fp32 operands must sit in the 32-byte-atom flavour of the 128-byte swizzle (4-row atoms, SBO 512);
    // the host builds their tensor maps with CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B to match
    constexpr uint32_t kMnLayout = ETr::kTf32 ? kLayoutSw128Base32 : kLayoutSw128;
    constexpr uint32_t a_layout = (MODE == MODE_WGRAD) ? kMnLayout : kLayoutSw128;
    constexpr uint32_t b_layout = (MODE == MODE_FPROP) ? kLayoutSw128 : kMnLayout;
    int stage = 0; uint32_t phase = 0;
    for (int it = it_begin; it < it_end && !(PAIR && crank != 0); ++it) {      // pair: only the leader issues
      mbar_wait(&full_bar[stage], phase);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t sA = smem_u32(smem + stage * Cfg::kStageBytes);
        const uint32_t sB = sA + Cfg::kABytes;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint64_t da = make_smem_desc(sA + k * a_kstep, p.a_lbo, p.a_sbo, a_layout);
          const uint64_t db = make_smem_desc(sB + k * b_kstep, p.b_lbo, p.b_sbo, b_layout);
          const uint32_t acc = (it > it_begin || k > 0) ? 1u : 0u;
          if (ETr::kTf32) {
            if (PAIR) umma_tf32_2sm(tmem_base, da, db, idesc, acc);
            else umma_tf32(tmem_base, da, db, idesc, acc);
          } else {
            if (PAIR) umma_bf16_2sm(tmem_base, da, db, idesc, acc);
            else umma_bf16(tmem_base, da, db, idesc, acc);
          }
        }
        if (PAIR) {
          umma_commit_2sm_mc(&empty_bar[stage], kMcMask);                   // frees the stage in BOTH CTAs
          if (it == it_end - 1) umma_commit_2sm_mc(tmem_full, kMcMask);     // both epilogues may start
        } else {
          if (CLN == 1) umma_commit(&empty_bar[stage]);    // frees the smem slot when these MMAs retire
          else umma_commit_mc(&empty_bar[stage], kMcMask);   // ... in every CTA that multicasts into it
          if (it == it_end - 1) umma_commit(tmem_full);   // accumulator complete
        }
      }
      __syncwarp();
      if (++stage == kStages) { stage = 0; phase ^= 1; }
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int q = warp & 3;                    // TMEM lane quarter this warp may access
    const int row = q * 32 + lane_id();        // accumulator row == M index inside the tile
    const int et = (warp - 2) * 32 + lane_id();  // 0..255 epilogue thread id
    const int chalf = (warp - 2) >> 2;           // which half of the tile's columns this warp drains
    constexpr int epi_threads = LITE ? 128 : 256;
    constexpr int kColsPerWarp = BLOCK_N / (epi_threads >> 7);
    const int c_begin = chalf * kColsPerWarp, c_end = c_begin + kColsPerWarp;
    mbar_wait(tmem_full, 0);
    tc_fence_after();

    // ---- where does this row go? ----
    bool row_ok = true;
    long long row_off;
    if (MODE == MODE_WGRAD) {
      row_ok = (m_tile * 128 + row) < p.red_ch;   // Cout may be 64 (< one M tile)
      row_off = kc.out_base + (long long)(m_tile * 128 + row) * p.sN + (long long)kc.taps[tap_fixed].wtap * p.n_total;
    } else {
      int w0, h0, nn0;
      decode(m_tile, w0, h0, nn0);
      const int wl = row % p.bw, hl = (row / p.bw) % p.bh, nl = row / (p.bw * p.bh);
      const int w = w0 + wl, h = h0 + hl, n = nn0 + nl;
      row_ok = (w < kc.Wd) && (h < kc.Hd) && (n < p.Nd);
      row_off = kc.out_base + n * p.sN + h * p.sH + w * p.sW;
    }

    auto emit = [&](int col0, float (&v)[32]) {
      if (!row_ok) return;
      if (MODE == MODE_WGRAD) {
        float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(out) + row_off + n0 + col0);
        if (p.atomic_out && kc.splits > 1) {
#pragma unroll
          for (int i = 0; i < 8; ++i)
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + i), "f"(v[4 * i]),
                         "f"(v[4 * i + 1]), "f"(v[4 * i + 2]), "f"(v[4 * i + 3])
                         : "memory");
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
      } else {
        ET* o = reinterpret_cast<ET*>(out) + row_off + n0 + col0;
        const ET* a = reinterpret_cast<const ET*>(addend) + row_off + n0 + col0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float f[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = v[8 * i + j];
          if (p.accumulate) {
            float t[8];
            load8(a + 8 * i, t);
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] += t[j];
          }
          store8(o + 8 * i, f);
          if (MODE == MODE_DGRAD) {            // hand the stored (rounded) values back for the fused BN reduce
#pragma unroll
            for (int j = 0; j < 8; ++j) v[8 * i + j] = stored_value(o, f[j]);
          }
        }
      }
    };

    // ---- fused BN statistics (fprop): per-channel sum / sum of squares of the fp32 accumulators ----
    // Warp transpose-reduce (recursive halving, 31 shuffles per quantity): afterwards lane L holds the
    // total of column col0+L over this warp's 32 rows.  The 4 epilogue warps meet in (now idle)
    // pipeline smem and 2*BLOCK_N red.add's per CTA accumulate into stats[2][Cout].
    float* red = reinterpret_cast<float*>(smem);          // [BLOCK_N/32][4][2][32]
    const bool do_stats = (MODE == MODE_FPROP) && (stats != nullptr);
    const bool do_bwd = (MODE == MODE_DGRAD) && (fuse.sums != nullptr);
    float* const sum_dst = do_bwd ? fuse.sums : stats;
    auto col_reduce = [&](int col0, float (&a)[32], float (&b)[32]) {
      const uint32_t lane = lane_id();
#pragma unroll
      for (int off = 16; off >= 1; off >>= 1) {
        const bool upper = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < off; ++i) {
          const float keep_a = upper ? a[i + off] : a[i], send_a = upper ? a[i] : a[i + off];
          const float keep_b = upper ? b[i + off] : b[i], send_b = upper ? b[i] : b[i + off];
          a[i] = keep_a + __shfl_xor_sync(0xffffffffu, send_a, off);
          b[i] = keep_b + __shfl_xor_sync(0xffffffffu, send_b, off);
        }
      }
      float* dst = red + ((col0 >> 5) * 4 + q) * 64;
      dst[lane] = a[0];
      dst[32 + lane] = b[0];
    };
    auto col_stats = [&](int col0, const float (&v)[32]) {
      float a[32], b[32];
      if (MODE == MODE_DGRAD) {
        // dy = dx * (out > 0);  b = dy * xhat
        const ET* op = reinterpret_cast<const ET*>(fuse.out) + row_off + n0 + col0;
        const ET* yp = reinterpret_cast<const ET*>(fuse.y) + row_off + n0 + col0;
#pragma unroll
        for (int i8 = 0; i8 < 4; ++i8) {
          float o8[8], y8[8];
          if (row_ok) {
            load8(yp + 8 * i8, y8);
            if (fuse.relu) load8(op + 8 * i8, o8);
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int i = 8 * i8 + j;
            float g = row_ok ? v[i] : 0.f;
            if (row_ok && fuse.relu && !(o8[j] > 0.f)) g = 0.f;
            const float xh = row_ok ? (y8[j] - __ldg(fuse.mean + n0 + col0 + i)) * __ldg(fuse.invstd + n0 + col0 + i) : 0.f;
            a[i] = g; b[i] = g * xh;
          }
        }
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) { a[i] = row_ok ? v[i] : 0.f; b[i] = a[i] * a[i]; }
      }
      col_reduce(col0, a, b);
    };
    auto flush_stats = [&]() {
      asm volatile("bar.sync 2, %0;" ::"n"(epi_threads) : "memory");
      for (int idx = et; idx < 2 * BLOCK_N; idx += epi_threads) {
        const int which = idx / BLOCK_N, col = idx % BLOCK_N;
        const float* src = red + (col >> 5) * 4 * 64 + which * 32 + (col & 31);
        const float t = src[0] + src[64] + src[128] + src[192];
        atomicAdd(&sum_dst[(size_t)which * p.n_total + n0 + col], t);
      }
    };

    const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    if (kc.splits == 1 || (MODE == MODE_WGRAD && p.atomic_out)) {
#pragma unroll 1
      for (int c = c_begin; c < c_end; c += 32) {
        uint32_t r[32];
        tmem_ld_32x32(t_row + c, r);
        tmem_ld_wait();
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
        emit(c, v);
        if (do_stats || do_bwd) col_stats(c, v);
      }
      if (do_stats || do_bwd) flush_stats();
    } else {
      // ---- split-K: park the partial tile, take a ticket, last arriver reduces ----
      const int tiles_n = gridDim.y;
      const long long tile_lin = (MODE == MODE_WGRAD)
                                     ? ((long long)tap_fixed * gridDim.x + m_tile) * tiles_n + n_tile
                                     : (long long)blockIdx.x * tiles_n + n_tile;
      const int stride_splits = (MODE == MODE_WGRAD) ? kc.splits : p.max_splits;
      float4* part = reinterpret_cast<float4*>(scratch) + (tile_lin * stride_splits + split) * (BLOCK_N / 4) * 128;
#pragma unroll 1
      for (int c = c_begin; c < c_end; c += 32) {
        uint32_t r[32];
        tmem_ld_32x32(t_row + c, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 8; ++i)
          part[(c / 4 + i) * 128 + row] = make_float4(__uint_as_float(r[4 * i]), __uint_as_float(r[4 * i + 1]),
                                                      __uint_as_float(r[4 * i + 2]), __uint_as_float(r[4 * i + 3]));
      }
      asm volatile("bar.sync 1, %0;" ::"n"(epi_threads) : "memory");
      if (et == 0) {
        __threadfence();                       // one fence per CTA publishes every epilogue thread's partial
        const unsigned int old = atomicAdd(&tickets[tile_lin], 1u);
        const unsigned int last = (old == static_cast<unsigned int>(kc.splits - 1)) ? 1u : 0u;
        if (last) __threadfence();
        *last_flag = last;
      }
      asm volatile("bar.sync 1, %0;" ::"n"(epi_threads) : "memory");
      if (*last_flag) {
        const float4* base = reinterpret_cast<const float4*>(scratch) + tile_lin * stride_splits * (BLOCK_N / 4) * 128;
#pragma unroll 1
        for (int c = c_begin; c < c_end; c += 32) {
          float v[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = 0.f;
          // fixed summation order; 4 splits x 8 vectors in flight per thread
          int s = 0;
          for (; s + 4 <= kc.splits; s += 4) {
            float4 t[4][8];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const float4* ps = base + (long long)(s + u) * (BLOCK_N / 4) * 128;
#pragma unroll
              for (int i = 0; i < 8; ++i) t[u][i] = __ldcg(&ps[(c / 4 + i) * 128 + row]);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                v[4 * i] += t[u][i].x; v[4 * i + 1] += t[u][i].y; v[4 * i + 2] += t[u][i].z; v[4 * i + 3] += t[u][i].w;
              }
          }
          for (; s < kc.splits; ++s) {
            const float4* ps = base + (long long)s * (BLOCK_N / 4) * 128;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float4 t = __ldcg(&ps[(c / 4 + i) * 128 + row]);
              v[4 * i] += t.x; v[4 * i + 1] += t.y; v[4 * i + 2] += t.z; v[4 * i + 3] += t.w;
            }
          }
          emit(c, v);
          if (do_stats || do_bwd) col_stats(c, v);
        }
        if (do_stats || do_bwd) flush_stats();
        if (et == 0) tickets[tile_lin] = 0;   // ready for the next launch
      }
    }
    tc_fence_before();
  }

  if (CLN > 1) cluster_sync_all(); else __syncthreads();     // a peer may still signal my barriers until it is done
  if (warp == 1) {
    tc_fence_after();
    if (PAIR) tmem_dealloc_2sm<BLOCK_N>(tmem_base);
    else tmem_dealloc<BLOCK_N>(tmem_base);
  }
}

}  // namespace ldnn
