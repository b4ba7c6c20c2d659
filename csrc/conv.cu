// Host side of the implicit-GEMM convolution: tensor-map construction, tiling / split-K
// heuristics, launch plans.  Kernel: conv_gemm.cuh.
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <mutex>
#include <atomic>

#include "../include/ldnn_b200.h"
#include "common.cuh"
#include "conv_gemm.cuh"

namespace ldnn {

static thread_local char g_err[512] = "";
std::atomic<unsigned long long> g_launches{0};

bool pdl_enabled() {
  static int on = -1;
  if (on < 0) { const char* e = getenv("LDNN_PDL"); on = (e && e[0] == '0') ? 0 : 1; }
  return on == 1;
}

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// ---- cuTensorMapEncodeTiled through the runtime (no link-time libcuda dependency) ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// TMA element type of the fp32 (TF32) path.  FLOAT32 hands the raw fp32 bits to the tensor core, which reads
// the upper 19 (truncation); TFLOAT32 makes the TMA unit convert to TF32 while it copies.  LDNN_TF32_TMA=
// float32|tfloat32 picks one (default tfloat32).
static CUtensorMapDataType tf32_map_type() {
  static int kind = -1;
  if (kind < 0) { const char* e = getenv("LDNN_TF32_TMA"); kind = (e && (e[0] == 'f' || e[0] == 'F')) ? 0 : 1; }
  return kind ? CU_TENSOR_MAP_DATA_TYPE_TFLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
}

// bf16 (esize 2) or fp32 (esize 4) tensor, up to 4 dims, innermost contiguous, 128B swizzle, zero OOB fill.
// mn32 = the operand is read MN-major by a kind::tf32 MMA: 128-byte swizzle with 32-byte atoms (see ptx.cuh).
static int make_map(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_elems,
                    const uint32_t* box, int esize, bool mn32 = false) {
  EncodeTiledFn enc = get_encode();
  LDNN_ARG(enc != nullptr, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t gdim[4], gstr[3];
  cuuint32_t bx[4], es[4];
  for (int i = 0; i < rank; ++i) { gdim[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 1; i < rank; ++i) gstr[i - 1] = strides_elems[i] * (uint64_t)esize;
  LDNN_ARG((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base %p not 16-byte aligned", base);
  for (int i = 0; i < rank - 1; ++i) LDNN_ARG(gstr[i] % 16 == 0, "TMA stride %d = %llu B not a multiple of 16", i,
                                              (unsigned long long)gstr[i]);
  CUresult r = enc(m, esize == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : tf32_map_type(), rank, const_cast<void*>(base), gdim, gstr, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE,
                   (mn32 && esize == 4) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  LDNN_ARG(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with %d (rank %d dims %llu %llu box %u %u)", (int)r, rank,
           (unsigned long long)gdim[0], (unsigned long long)gdim[1], bx[0], bx[1]);
  return LDNN_OK;
}

// TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory
// hung the GPU.  The cluster = 2 variant multicasts the weight tile, so every tensor it maps must be plain device memory
// of the current device: not a peer mapping, not an IPC import, not host memory.
static int require_local_device_memory(const void* ptr, const char* what) {
  if (ptr == nullptr) return LDNN_OK;
  int cur = -1;
  LDNN_CUDA(cudaGetDevice(&cur));
  cudaPointerAttributes at;
  cudaError_t e = cudaPointerGetAttributes(&at, ptr);
  if (e != cudaSuccess) { cudaGetLastError(); at.type = cudaMemoryTypeUnregistered; at.device = -1; }
  LDNN_ARG(at.type == cudaMemoryTypeDevice && at.device == cur,
           "conv cluster=2 (TMA multicast of the weight tile) needs %s in this GPU's own device memory "
           "(pointer %p: memory type %d on device %d, current device %d); use cluster=0 or cluster=-2", what, ptr,
           (int)at.type, at.device, cur);
  return LDNN_OK;
}

static int next_pow2(int v) { int p = 1; while (p < v) p <<= 1; return p; }
static int cdiv(int a, int b) { return (a + b - 1) / b; }
static int floordiv(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }
static int posmod(int a, int b) { int m = a % b; return m < 0 ? m + b : m; }

struct Launch {
  ConvMaps maps;
  ConvParams p;
  dim3 grid;
  size_t scratch_floats = 0, tickets = 0;   // split-K workspace this launch needs
  size_t ticket_off = 0, scratch_off = 0;
  bool uses_split_scratch = false;
  int covered_classes = 1;                  // dgrad: parity classes that have taps
};
struct Plan {
  int mode, block_n, cluster = 1, lite = 0;
  int tf32 = 0;                             // fp32 storage + kind::tf32 (desc.dtype == 1)
  int pair = 0;                             // cluster == 2 run as a cta_group::2 pair (desc.cluster == -2)
  std::vector<Launch> launches;
  void* out;
  const void* addend;
  float* scratch;
  unsigned int* tickets;
  float* stats;
  BnBwdFuse fuse;
};

static void pixel_box(int Wd, int Hd, int Nd, int target, ConvParams& p, ConvClass& c) {
  p.bw = std::min(next_pow2(Wd), target);
  p.bh = std::min(next_pow2(Hd), target / p.bw);
  p.bn = target / (p.bw * p.bh);
  c.tiles_w = cdiv(Wd, p.bw);
  c.tiles_h = cdiv(Hd, p.bh);
  c.Wd = Wd; c.Hd = Hd; p.Nd = Nd;
}

static int pick_splits(int tiles, int iters, int requested) {
  int s = requested;
  if (s <= 0) {
    s = 1;
    if (tiles < 96) s = std::max(1, 148 / tiles);
    s = std::min(s, std::max(1, iters / 4));   // keep >= 4 k-iterations per CTA
    s = std::min(s, 32);
  }
  s = std::max(1, std::min(s, iters));
  int per = cdiv(iters, s);
  return cdiv(iters, per);                     // no empty split
}

// Builds every launch of a plan.  `dry` = only sizes (no tensor maps, pointers may be null).
static int build(const ldnn_conv_desc& d, Plan& plan, bool dry) {
  LDNN_ARG(d.mode >= 0 && d.mode <= 2, "bad conv mode %d", d.mode);
  LDNN_ARG(d.dtype == 0 || d.dtype == 1, "bad conv dtype %d (0 = bf16, 1 = fp32 storage / TF32 MMA)", d.dtype);
  const int esize = d.dtype == 1 ? 4 : 2;   // bytes per activation / weight element
  const int KE = 128 / esize;               // elements per 128-byte swizzle row = K block of a pipeline stage
  LDNN_ARG(d.Cin % 64 == 0 && d.Cout % 64 == 0, "Cin=%d / Cout=%d must be multiples of 64", d.Cin, d.Cout);
  LDNN_ARG(d.stride == 1 || d.stride == 2, "stride %d unsupported", d.stride);
  LDNN_ARG(d.R * d.R <= 12, "filter %dx%d too large for the tap table", d.R, d.R);
  const int T = d.R * d.R;
  const int P = (d.H + 2 * d.pad - d.R) / d.stride + 1, Q = (d.W + 2 * d.pad - d.R) / d.stride + 1;
  const int n_total = (d.mode == MODE_FPROP) ? d.Cout : d.Cin;
  // pixel tiles of the output-side grid (for the tile-count heuristics)
  const int grid_w = (d.mode == MODE_DGRAD) ? d.W : Q, grid_h = (d.mode == MODE_DGRAD) ? d.H : P;
  int bn_ = d.block_n;
  if (bn_ == 0) {
    if (d.mode == MODE_WGRAD) {
      bn_ = (n_total % 256 == 0) ? 256 : ((n_total % 128 == 0) ? 128 : 64);
    } else {
      bn_ = (n_total % 128 == 0) ? 128 : 64;
      // few output tiles: prefer narrower tiles (more CTAs, fewer K splits to fold) over split-K
      const long long m_tiles_est = cdiv((long long)d.N * grid_h * grid_w, 128);
      if (bn_ == 128 && m_tiles_est * (n_total / 128) < 64) bn_ = 64;
    }
  }
  LDNN_ARG(bn_ == 64 || bn_ == 128 || bn_ == 256, "block_n %d", bn_);
  LDNN_ARG(n_total % bn_ == 0, "N=%d not a multiple of block_n=%d", n_total, bn_);
  plan.mode = d.mode; plan.block_n = bn_; plan.tf32 = d.dtype == 1;
  plan.lite = d.lite ? 1 : 0;
  plan.cluster = ((d.cluster == 2 || d.cluster == -2) && d.mode != MODE_WGRAD && bn_ >= 128) ? 2 : 1;
  plan.pair = (d.cluster == -2 && plan.cluster == 2) ? 1 : 0;
  if (plan.cluster == 2 && !plan.pair && !dry) {
    int rc = require_local_device_memory(d.x, "x");
    if (!rc) rc = require_local_device_memory(d.w, "w");
    if (!rc) rc = require_local_device_memory(d.y, "y");
    if (rc) return rc;
  }
  LDNN_ARG(!(plan.pair && d.lite), "the CTA-pair variant has no lite form");
  plan.out = (d.mode == MODE_FPROP) ? d.y : (d.mode == MODE_DGRAD ? const_cast<void*>(d.x) : d.dw);
  plan.addend = d.addend ? d.addend : plan.out;
  plan.scratch = d.scratch; plan.tickets = d.tickets;
  plan.stats = (d.mode == MODE_FPROP) ? d.bn_sums : nullptr;
  memset(&plan.fuse, 0, sizeof(plan.fuse));
  if (d.mode == MODE_DGRAD && d.fuse_sums) {
    LDNN_ARG(d.fuse_out && d.fuse_y && d.fuse_mean && d.fuse_invstd, "dgrad BN-backward fusion needs out/y/mean/invstd");
    plan.fuse.out = d.fuse_out;
    plan.fuse.y = d.fuse_y;
    plan.fuse.mean = d.fuse_mean; plan.fuse.invstd = d.fuse_invstd; plan.fuse.sums = d.fuse_sums;
    plan.fuse.relu = d.fuse_relu;
  }

  const char* xb = reinterpret_cast<const char*>(d.x);
  const char* yb = reinterpret_cast<const char*>(d.y);

  // parity views of x for strided convs: view (hp,wp) has rows h = 2a+hp
  const bool mn_act = d.mode == MODE_WGRAD;            // operands read MN-major: wgrad x and dy, dgrad weights
  const bool mn_w = d.mode == MODE_DGRAD;
  auto x_view_map = [&](CUtensorMap* m, int hp, int wp, const uint32_t* box) -> int {
    const int s = d.stride;
    uint64_t dims[4] = {(uint64_t)d.Cin, (uint64_t)cdiv(d.W - wp, s), (uint64_t)cdiv(d.H - hp, s), (uint64_t)d.N};
    uint64_t str[4] = {1, (uint64_t)s * d.Cin, (uint64_t)s * d.W * d.Cin, (uint64_t)d.H * d.W * d.Cin};
    return make_map(m, xb + ((size_t)hp * d.W + wp) * d.Cin * esize, 4, dims, str, box, esize, mn_act);
  };
  auto y_map = [&](CUtensorMap* m, const uint32_t* box) -> int {
    uint64_t dims[4] = {(uint64_t)d.Cout, (uint64_t)Q, (uint64_t)P, (uint64_t)d.N};
    uint64_t str[4] = {1, (uint64_t)d.Cout, (uint64_t)Q * d.Cout, (uint64_t)P * Q * d.Cout};
    return make_map(m, yb, 4, dims, str, box, esize, mn_act);
  };
  auto w_map = [&](CUtensorMap* m, uint32_t box_rows) -> int {
    uint64_t dims[2] = {(uint64_t)T * d.Cin, (uint64_t)d.Cout};
    uint64_t str[2] = {1, (uint64_t)T * d.Cin};
    uint32_t box[2] = {(uint32_t)KE, box_rows};
    return make_map(m, d.w, 2, dims, str, box, esize, mn_w);
  };
  // forward-style tap table: input row = stride*p - pad + r
  auto fwd_taps = [&](ConvClass& c) {
    c.num_taps = T;
    for (int r = 0; r < d.R; ++r)
      for (int s = 0; s < d.R; ++s) {
        const int th = r - d.pad, tw = s - d.pad;
        TapEntry& te = c.taps[r * d.R + s];
        te.map = (int8_t)(posmod(th, d.stride) * 2 + posmod(tw, d.stride));
        if (d.stride == 1) te.map = 0;
        te.dh = (int8_t)floordiv(th, d.stride);
        te.dw = (int8_t)floordiv(tw, d.stride);
        te.wtap = (int8_t)(r * d.R + s);
      }
  };
  Launch L;
  ConvParams& p = L.p;
  memset(&p, 0, sizeof(p));
  p.accumulate = d.accumulate;
  p.n_total = n_total;
  p.num_classes = 1;
  // K-major SW128 tile: 8-row groups 1024 B apart; MN-major: KE-wide MN blocks (128 B x KE K-rows = 8192 B
  // bf16 / 4096 B fp32) apart, 8-row K groups 1024 B apart.
  const uint32_t mn_box = 128u * (uint32_t)KE;
  p.a_lbo = d.a_lbo ? d.a_lbo : (d.mode == MODE_WGRAD ? mn_box : 16u);
  const uint32_t mn_sbo = esize == 4 ? 512u : 1024u;   // K-row groups of an MN-major box: 8 rows (bf16) / 4 rows (fp32)
  p.a_sbo = d.a_sbo ? d.a_sbo : (d.mode == MODE_WGRAD ? mn_sbo : 1024u);
  p.b_lbo = d.b_lbo ? d.b_lbo : (d.mode == MODE_FPROP ? 16u : mn_box);
  p.b_sbo = d.b_sbo ? d.b_sbo : (d.mode == MODE_FPROP ? 1024u : mn_sbo);
  size_t scratch_floats = 0, tickets = 0;

  if (d.mode == MODE_FPROP) {
    ConvClass& c = p.cls[0];
    pixel_box(Q, P, d.N, 128, p, c);
    fwd_taps(c);
    p.kb_per_tap = d.Cin / KE; p.red_ch = d.Cin;
    c.iters_total = T * p.kb_per_tap;
    c.out_base = 0; p.sW = d.Cout; p.sH = (long long)Q * d.Cout; p.sN = (long long)P * Q * d.Cout;
    const int m_tiles = c.tiles_w * c.tiles_h * cdiv(d.N, p.bn), n_tiles = d.Cout / bn_;
    c.splits = pick_splits(m_tiles * n_tiles, c.iters_total, d.splits);
    c.iters_per_split = cdiv(c.iters_total, c.splits);
    p.max_splits = c.splits;
    if (c.splits > 1) { scratch_floats = (size_t)m_tiles * n_tiles * c.splits * 128 * bn_; tickets = (size_t)m_tiles * n_tiles; }
    L.grid = dim3(m_tiles, n_tiles, c.splits);
    if (!dry) {
      uint32_t box[4] = {(uint32_t)KE, (uint32_t)p.bw, (uint32_t)p.bh, (uint32_t)p.bn};
      const int nviews = d.stride == 1 ? 1 : 4;
      for (int v = 0; v < 4; ++v) {
        int rc = x_view_map(&L.maps.act[v], v < nviews ? v / 2 : 0, v < nviews ? v % 2 : 0, box);
        if (rc) return rc;
      }
      int rc = w_map(&L.maps.other, bn_ / plan.cluster);
      if (rc) return rc;
    }
    if (plan.cluster == 2) LDNN_ARG(m_tiles % 2 == 0, "cluster=2 needs an even number of M tiles (%d)", m_tiles);
  } else if (d.mode == MODE_DGRAD) {
    // one class per output parity (1 class for stride 1), all in one launch
    p.kb_per_tap = d.Cout / KE; p.red_ch = d.Cout;
    p.sW = (long long)d.stride * d.Cin; p.sH = (long long)d.stride * d.W * d.Cin; p.sN = (long long)d.H * d.W * d.Cin;
    const int n_tiles = d.Cin / bn_;
    int ncls = 0, m_total = 0, covered = 0;
    // all classes share one pixel box: size it from the largest class (parity 0,0)
    ConvClass tmp;
    pixel_box(cdiv(d.W, d.stride), cdiv(d.H, d.stride), d.N, 128, p, tmp);
    for (int hp = 0; hp < d.stride; ++hp)
      for (int wp = 0; wp < d.stride; ++wp) {
        ConvClass c;
        memset(&c, 0, sizeof(c));
        const int Hs = cdiv(d.H - hp, d.stride), Ws = cdiv(d.W - wp, d.stride);
        c.Wd = Ws; c.Hd = Hs; c.tiles_w = cdiv(Ws, p.bw); c.tiles_h = cdiv(Hs, p.bh);
        int nt = 0;
        for (int r = 0; r < d.R; ++r)
          for (int s = 0; s < d.R; ++s) {
            const int th = hp + d.pad - r, tw = wp + d.pad - s;
            if (posmod(th, d.stride) || posmod(tw, d.stride)) continue;
            TapEntry& te = c.taps[nt++];
            te.map = 0; te.dh = (int8_t)floordiv(th, d.stride); te.dw = (int8_t)floordiv(tw, d.stride);
            te.wtap = (int8_t)(r * d.R + s);
          }
        if (nt == 0) {
          LDNN_ARG(d.accumulate, "dgrad parity class (%d,%d) has no taps and accumulate=0", hp, wp);
          continue;
        }
        ++covered;
        c.num_taps = nt;
        c.iters_total = nt * p.kb_per_tap;
        c.out_base = ((long long)hp * d.W + wp) * d.Cin;
        c.m_begin = m_total;
        m_total += c.tiles_w * c.tiles_h * cdiv(d.N, p.bn);
        p.cls[ncls++] = c;
      }
    LDNN_ARG(ncls >= 1, "dgrad: no taps at all");
    if (plan.cluster == 2)
      for (int i = 0; i < ncls; ++i) {
        const int mt = (i + 1 < ncls ? p.cls[i + 1].m_begin : m_total) - p.cls[i].m_begin;
        LDNN_ARG(mt % 2 == 0, "cluster=2 needs an even number of M tiles per parity class (%d)", mt);
      }
    p.num_classes = ncls;
    L.covered_classes = covered;
    p.max_splits = 1;
    for (int i = 0; i < ncls; ++i) {
      ConvClass& c = p.cls[i];
      c.splits = pick_splits(m_total * n_tiles, c.iters_total, d.splits);
      c.iters_per_split = cdiv(c.iters_total, c.splits);
      p.max_splits = std::max(p.max_splits, c.splits);
    }
    if (p.max_splits > 1) { scratch_floats = (size_t)m_total * n_tiles * p.max_splits * 128 * bn_; tickets = (size_t)m_total * n_tiles; }
    L.grid = dim3(m_total, n_tiles, p.max_splits);
    if (!dry) {
      uint32_t box[4] = {(uint32_t)KE, (uint32_t)p.bw, (uint32_t)p.bh, (uint32_t)p.bn};
      for (int v = 0; v < 4; ++v) { int rc = y_map(&L.maps.act[v], box); if (rc) return rc; }
      int rc = w_map(&L.maps.other, KE);
      if (rc) return rc;
    }
  } else {
    LDNN_ARG(d.Cout % 128 == 0 || d.Cout == 64, "wgrad needs Cout %% 128 == 0 (or 64)");
    ConvClass& c = p.cls[0];
    pixel_box(Q, P, d.N, KE, p, c);
    fwd_taps(c);
    c.iters_total = c.tiles_w * c.tiles_h * cdiv(d.N, p.bn);
    p.sN = (long long)T * d.Cin;     // row stride of dw
    p.red_ch = d.Cout;               // rows that exist (store mask)
    p.atomic_out = d.atomic_out;
    const int m_tiles = cdiv(d.Cout, 128), n_tiles = d.Cin / bn_;
    c.splits = pick_splits(m_tiles * n_tiles * T, c.iters_total, d.splits);
    c.iters_per_split = cdiv(c.iters_total, c.splits);
    p.max_splits = c.splits;
    if (c.splits > 1 && !p.atomic_out) {
      scratch_floats = (size_t)m_tiles * n_tiles * T * c.splits * 128 * bn_;
      tickets = (size_t)m_tiles * n_tiles * T;
    }
    L.grid = dim3(m_tiles, n_tiles, T * c.splits);
    if (!dry) {
      uint32_t box[4] = {(uint32_t)KE, (uint32_t)p.bw, (uint32_t)p.bh, (uint32_t)p.bn};
      const int nviews = d.stride == 1 ? 1 : 4;
      for (int v = 0; v < 4; ++v) {
        int rc = x_view_map(&L.maps.act[v], v < nviews ? v / 2 : 0, v < nviews ? v % 2 : 0, box);
        if (rc) return rc;
      }
      int rc = y_map(&L.maps.other, box);
      if (rc) return rc;
    }
  }
  L.scratch_floats = scratch_floats; L.tickets = tickets;
  L.uses_split_scratch = scratch_floats > 0;
  plan.launches.push_back(L);
  return LDNN_OK;
}

template <typename ET, int MODE, int BN, int CL, bool LITE>
static int launch_one(const Launch& L, const Plan& pl, cudaStream_t st) {
  using Cfg = ConvCfg<BN>;
  constexpr int kSmem = LITE ? Cfg::kLiteSmemBytes : Cfg::kSmemBytes;
  static bool configured = false;
  if (!configured) {
    LDNN_CUDA(cudaFuncSetAttribute(conv_gemm_kernel<ET, MODE, BN, CL, LITE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   kSmem));
    configured = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = L.grid; cfg.stream = st;
  cfg.blockDim = dim3(LITE ? Cfg::kLiteThreads : Cfg::kThreads);
  cfg.dynamicSmemBytes = kSmem;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (CL > 1 || CL < 0) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = CL < 0 ? -CL : CL; attr[na].val.clusterDim.y = 1; attr[na].val.clusterDim.z = 1;
    ++na;
  }
  cfg.attrs = attr; cfg.numAttrs = na;
  LDNN_CUDA(cudaLaunchKernelEx(&cfg, conv_gemm_kernel<ET, MODE, BN, CL, LITE>, L.maps, L.p, pl.out, pl.addend,
                               pl.scratch ? pl.scratch + L.scratch_off : nullptr,
                               pl.tickets ? pl.tickets + L.ticket_off : nullptr, pl.stats, pl.fuse));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return LDNN_OK;
}

template <typename ET, int MODE, int BN>
static int launch_variant(const Launch& L, const Plan& pl, cudaStream_t st) {
  if (pl.lite) return launch_one<ET, MODE, BN, 1, true>(L, pl, st);
  return launch_one<ET, MODE, BN, 1, false>(L, pl, st);
}

template <typename ET, int MODE>
static int launch_bn(const Launch& L, const Plan& pl, cudaStream_t st) {
  if (pl.cluster == 2 && MODE != MODE_WGRAD) {      // cluster path: 128/256-wide tiles, full pipeline
    constexpr int M2 = MODE == MODE_WGRAD ? MODE_FPROP : MODE;
    if (pl.pair) {                                   // cta_group::2 pair (UMMA_M = 256)
      if (pl.block_n == 128) return launch_one<ET, M2, 128, -2, false>(L, pl, st);
      return launch_one<ET, M2, 256, -2, false>(L, pl, st);
    }
    if (pl.block_n == 128) return launch_one<ET, M2, 128, 2, false>(L, pl, st);
    return launch_one<ET, M2, 256, 2, false>(L, pl, st);
  }
  switch (pl.block_n) {
    case 64: return launch_variant<ET, MODE, 64>(L, pl, st);
    case 128: return launch_variant<ET, MODE, 128>(L, pl, st);
    default: return launch_variant<ET, MODE, 256>(L, pl, st);
  }
}

template <typename ET>
static int launch_mode(const Launch& L, const Plan& pl, cudaStream_t st) {
  if (pl.mode == MODE_FPROP) return launch_bn<ET, MODE_FPROP>(L, pl, st);
  if (pl.mode == MODE_DGRAD) return launch_bn<ET, MODE_DGRAD>(L, pl, st);
  return launch_bn<ET, MODE_WGRAD>(L, pl, st);
}

}  // namespace ldnn

using namespace ldnn;

extern "C" const char* ldnn_last_error(void) { return g_err; }
extern "C" int ldnn_version(void) { return 100; }
extern "C" unsigned long long ldnn_launch_count(void) { return g_launches.load(); }

extern "C" int ldnn_conv_query(const ldnn_conv_desc* d, size_t* scratch_bytes, size_t* ticket_count, int* launches,
                               int* covered_classes) {
  Plan pl;
  int rc = build(*d, pl, true);
  if (rc) return rc;
  if (scratch_bytes) *scratch_bytes = pl.launches.empty() ? 0 : pl.launches[0].scratch_floats * sizeof(float);
  if (ticket_count) *ticket_count = pl.launches.empty() ? 0 : pl.launches[0].tickets;
  if (launches) *launches = (int)pl.launches.size();
  if (covered_classes) *covered_classes = pl.launches.empty() ? 0 : pl.launches[0].covered_classes;
  return LDNN_OK;
}

extern "C" int ldnn_conv_query_grid(const ldnn_conv_desc* d, int* gx, int* gy, int* gz) {
  Plan pl;
  int rc = build(*d, pl, true);
  if (rc) return rc;
  *gx = (int)pl.launches[0].grid.x; *gy = (int)pl.launches[0].grid.y; *gz = (int)pl.launches[0].grid.z;
  return LDNN_OK;
}

extern "C" int ldnn_conv_plan_create(const ldnn_conv_desc* d, void** plan) {
  Plan* pl = new Plan();
  int rc = build(*d, *pl, false);
  if (rc) { delete pl; return rc; }
  bool need_scratch = false;
  for (auto& L : pl->launches) need_scratch |= L.uses_split_scratch;
  if (need_scratch && (!d->scratch || !d->tickets)) {
    delete pl;
    set_last_error("conv plan uses split-K but scratch/tickets were not provided");
    return LDNN_ERR_ARG;
  }
  *plan = pl;
  return LDNN_OK;
}

extern "C" int ldnn_conv_plan_run(void* plan, void* stream) {
  Plan* pl = reinterpret_cast<Plan*>(plan);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  for (const Launch& L : pl->launches) {
    const int rc = pl->tf32 ? launch_mode<float>(L, *pl, st) : launch_mode<__nv_bfloat16>(L, *pl, st);
    if (rc) return rc;
  }
  return LDNN_OK;
}

extern "C" void ldnn_conv_plan_destroy(void* plan) { delete reinterpret_cast<Plan*>(plan); }
