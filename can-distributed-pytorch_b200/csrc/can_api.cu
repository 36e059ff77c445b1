// extern "C" surface of libcan_b200.so (declared in include/can_b200.h).
// Host side: argument validation, TMA tensor-map encoding, launch configuration.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cudaTypedefs.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/can_b200.h"
#include "can_conv_tc.cuh"
#include "can_simple.cuh"
#include "can_wgrad_tc.cuh"
#include "can_aux.cuh"
#include "can_first_tc.cuh"
#include "can_data.cuh"

using namespace canb;

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
int cuda_fail(cudaError_t e, const char* what) {
  snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
  return static_cast<int>(e);
}
#define CAN_CHECK_LAUNCH(what)                                   \
  do {                                                           \
    cudaError_t e__ = cudaGetLastError();                        \
    if (e__ != cudaSuccess) return cuda_fail(e__, what);         \
  } while (0)

PFN_cuTensorMapEncodeTiled_v12000 get_encode() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  }
  return fn;
}

// bf16 tensor map with a 128-byte swizzled box whose innermost extent is 64 elements.
int make_tmap(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
              const uint32_t* box, bool swizzle128 = true, const uint32_t* elem_strides = nullptr, bool fp32 = false) {
  auto enc = get_encode();
  if (!enc) return fail(-100, "cuTensorMapEncodeTiled entry point not available");
  uint32_t estr[5] = {1, 1, 1, 1, 1};
  if (elem_strides)
    for (int i = 0; i < rank; ++i) estr[i] = elem_strides[i];
  CUresult r = enc(m, fp32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), dims, strides_bytes, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(-101, "cuTensorMapEncodeTiled failed (CUresult %d)", static_cast<int>(r));
  return 0;
}

int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v ? atoi(v) : dflt;
}

constexpr int kMaxSmem = 232448;  // 227 KB opt-in limit per CTA on sm_100

// SMs the one-CTA-per-SM kernels may fill: all of them, or CAN_B200_CTAS when set (the data-parallel trainer can leave a
// few to NCCL's all-reduce kernels, CAN_B200_DP_RESERVE_SMS; measured neutral to slightly negative at 2 GPUs, default off).
int avail_sms() {
  static int sm_count = 0;
  if (!sm_count) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sm_count <= 0) sm_count = 148;
  }
  const int e = env_int("CAN_B200_CTAS", 0);
  return e > 0 && e < sm_count ? e : sm_count;
}

// extra tensor maps of MODE_CTX2 (interpolation matrix, per-image tables); other modes pass copies of tmA
thread_local const CUtensorMap* g_tmW = nullptr;
thread_local const CUtensorMap* g_tmT = nullptr;
// output map of the staged (TMA store) epilogue; null -> direct stores
thread_local const CUtensorMap* g_tmO = nullptr;
thread_local const CUtensorMap* g_tmM = nullptr;   // mask map of the staged epilogue

// Extended launch: clusters of `cluster` CTAs (cta_group::2 pairs), and programmatic dependent launch for the kernels
// that call pdl_wait() before their first global read (CAN_B200_PDL=0 switches it off): inside a step ~100 dependent
// kernels follow each other, and every one of them spends its first microseconds on barrier set-up and TMEM allocation.
template <typename K, typename... Args>
int launch_ex(K kern, dim3 grid, int threads, int smem, cudaStream_t st, const char* what, int cluster, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(threads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (cluster > 1) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = cluster;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  if (env_int("CAN_B200_PDL", 0) != 0) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, args...);
  if (e == cudaSuccess) e = cudaGetLastError();   // (a fault of this launch under CUDA_LAUNCH_BLOCKING=1 shows up here)
  if (e != cudaSuccess) return cuda_fail(e, what);
  return 0;
}

template <int BN, int NSPLIT, int MODE, int PAIR = 0>
int launch_conv(const CUtensorMap& tmA, const CUtensorMap& tmB, ConvTcParams& p, dim3 grid, int stages_override,
                cudaStream_t st) {
  using L = ConvTcSmem<BN, gemm_planes(NSPLIT, MODE)>;
  constexpr int ctx = MODE;   // (ConvTcSmem::bytes sizes the table region from the mode)
  const int staged_bytes = p.out_stage_bytes + p.mask_bufs * BN * 256;
  // pair: each CTA holds half of B; the 128-wide pair tiles put two 64-channel k-chunks in a stage (see the kernel)
  // (PAIR == 2: the ring holds weight tiles only, three taps per stage at BN = 128; the A halo stages sit before it)
  constexpr int kStage = PAIR == 2 ? (BN == 128 ? 3 : 1) * (L::kBBytes / 2)
                         : PAIR    ? (BN == 128 ? 2 : 1) * (L::kABytes + L::kBBytes / 2)
                                   : L::kStageBytes;
  const int halo_a_bytes = PAIR == 2 ? p.n_a * p.halo_stride : 0;
  int stages = (kMaxSmem - L::bytes(0, ctx) - staged_bytes - halo_a_bytes) / kStage;
  if (stages > 8) stages = 8;
  if (stages_override > 0 && stages_override < stages) stages = stages_override;
  if (stages < 2) return fail(-20, "conv_tc: tile BN=%d nsplit=%d does not fit shared memory", BN, NSPLIT);
  p.stages = stages;
  const int smem = L::bytes(0, ctx) + stages * kStage + staged_bytes + halo_a_bytes;
  // eight epilogue warps for the narrow, epilogue-bound tiles of the bf16 path
  constexpr int EW = ((NSPLIT == 1 && BN <= 128 && (MODE == MODE_STORE || MODE == MODE_POOL || MODE == MODE_UNPOOL)) ||
                      MODE == MODE_CTX2 || MODE == MODE_CTX2_BWD) ? 8 : 4;   // (MODE_DFV with 8: measured slower)
  auto kern = conv_tc_kernel<BN, NSPLIT, MODE, 0, EW, PAIR>;
  static bool attr_done = false;  // per template instantiation
  static int attr_smem = 0;
  if (!attr_done || smem > attr_smem) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(conv_tc)");
    attr_done = true;
    attr_smem = kMaxSmem;
  }
  return launch_ex(kern, grid, 128 + 32 * EW, smem, st, "conv_tc launch", PAIR ? 2 : 1, tmA, tmB, g_tmW ? *g_tmW : tmA,
                   g_tmT ? *g_tmT : tmA, g_tmO ? *g_tmO : tmA, g_tmM ? *g_tmM : tmA, p);
}

// HALO variant (bf16 path, small Cin): resident weights + halo activation stages.
template <int BN, int MODE>
int launch_conv_halo(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmHalo, ConvTcParams& p,
                     dim3 grid, cudaStream_t st) {
  using L = ConvTcSmem<BN, 1>;
  const int tail = L::kTailBytes;
  const int staged_bytes = p.out_stage_bytes + p.mask_bufs * BN * 256;
  int stages = (kMaxSmem - 1024 - staged_bytes - p.w_bytes - tail) / p.halo_stride;
  if (stages > 8) stages = 8;
  if (stages < 2) return fail(-20, "conv_tc(halo): BN=%d does not fit shared memory", BN);
  p.stages = stages;
  const int smem = 1024 + staged_bytes + p.w_bytes + stages * p.halo_stride + tail;
  auto kern = conv_tc_kernel<BN, 1, MODE, 1, 8>;
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(conv_tc halo)");
    attr_done = true;
  }
  return launch_ex(kern, grid, 128 + 32 * 8, smem, st, "conv_tc(halo) launch", 1, tmA, tmB, tmHalo, tmA,
                   g_tmO ? *g_tmO : tmA, g_tmM ? *g_tmM : tmA, p);
}

template <int BN>
int dispatch_halo(int mode, const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& h, ConvTcParams& p, dim3 grid,
                  cudaStream_t st) {
  switch (mode) {
    case MODE_STORE: return launch_conv_halo<BN, MODE_STORE>(a, b, h, p, grid, st);
    case MODE_POOL: return launch_conv_halo<BN, MODE_POOL>(a, b, h, p, grid, st);
    case MODE_UNPOOL: return launch_conv_halo<BN, MODE_UNPOOL>(a, b, h, p, grid, st);
    default: return fail(-21, "conv_tc(halo): mode %d not available", mode);
  }
}

template <int BN, int NSPLIT>
int dispatch_mode(int mode, const CUtensorMap& a, const CUtensorMap& b, ConvTcParams& p, dim3 grid, int so,
                  cudaStream_t st) {
  switch (mode) {
    case MODE_STORE: return launch_conv<BN, NSPLIT, MODE_STORE>(a, b, p, grid, so, st);
    case MODE_POOL: return launch_conv<BN, NSPLIT, MODE_POOL>(a, b, p, grid, so, st);
    case MODE_UNPOOL: return launch_conv<BN, NSPLIT, MODE_UNPOOL>(a, b, p, grid, so, st);
    case MODE_DFV: return launch_conv<BN, NSPLIT, MODE_DFV>(a, b, p, grid, so, st);
    default: return fail(-21, "conv_tc: mode %d not available for BN=%d", mode, BN);
  }
}


struct WgradPlan {
  int TW, TH, tiles_x, tiles_y, total_patches, m_tiles, n_tiles, bn, ksplit, per, T, tap_groups, swap64, pix, hx;
};
WgradPlan wgrad_plan(int N, int H, int W, int Cin, int Cout, int taps, int nsplit, int dil = 1) {
  WgradPlan pl;
  pl.bn = Cin < 256 ? Cin : (nsplit > 1 ? 128 : 256);
  // 512-wide layers: nine taps in pairs (BN = 256) leave a half-loaded fifth tap group and 120 CTAs for 148 SMs;
  // three balanced groups of three taps at BN = 128 fill 144 (measured: 512->512 186 -> 172 us, 1024->512 369 -> 334 us;
  // the 256-wide layers are better off with BN = 256)
  if (nsplit == 1 && taps == 9 && Cin >= 512 && Cout >= 512 && env_int("CAN_B200_WGBN", 128) == 128) pl.bn = 128;
  // With the halo X tile (hx below) the 128-wide three-tap tiles stay under the L2->SM fill rate (18-20 KB per 384 cycles of
  // MMA) where the 256-wide two-tap ones need 40 KB per 512: every 3x3 layer with Cin >= 256 takes them (step 7.64 -> 7.49
  // ms, `profiles/r02_wgbn256_ab.txt`; CAN_B200_WGBN256=256 keeps BN = 256 for Cin = 256 / Cout < 512).
  if (nsplit == 1 && taps == 9 && Cin >= 256 && env_int("CAN_B200_WGBN256", 128) == 128) pl.bn = 128;
  // multi-tap tiles (bf16 path, BN >= 128): 64-80 KB stages would leave a 2-3 deep pipeline -> 32-pixel patches
  pl.pix = (nsplit == 1 && taps > 1 && pl.bn >= 128) ? 32 : 64;
  pl.TW = (W >= 16) ? 16 : 8;
  // 128-wide three-tap tiles: the three X patches of a stage as shifted views of one halo tile (8-pixel patch rows)
  pl.hx = (nsplit == 1 && taps == 9 && pl.bn == 128 && pl.pix == 32 && dil <= 2 && env_int("CAN_B200_WGHX", 1) != 0) ? 1 : 0;
  if (pl.hx) pl.TW = 8;
  pl.TH = pl.pix / pl.TW;
  pl.tiles_x = (W + pl.TW - 1) / pl.TW;
  pl.tiles_y = (H + pl.TH - 1) / pl.TH;
  pl.total_patches = pl.tiles_x * pl.tiles_y * N;
  pl.m_tiles = (Cout + 127) / 128;
  pl.n_tiles = Cin / pl.bn;
  // taps per CTA: as many TMEM accumulators as fit (bf16 path); the fp32-parity path keeps one
  // (balanced groups: 9 taps = 3+3+3 for BN=64 (one N=192 MMA per k-step) and BN=128, 2+2+2+2+1 for BN=256)
  pl.T = (nsplit == 1 && taps > 1) ? (pl.bn == 64 ? 3 : pl.bn == 128 ? 3 : 2) : 1;
  if (pl.T > taps) pl.T = taps;
  pl.tap_groups = (taps + pl.T - 1) / pl.T;
  // 64 -> 64 3x3 (bf16 path): operand roles swapped, all nine taps in one CTA (wgrad_tc_swap64_kernel)
  pl.swap64 = (nsplit == 1 && taps == 9 && Cin == 64 && Cout == 64 && env_int("CAN_B200_WGSWAP", 1) != 0) ? 1 : 0;
  if (pl.swap64) {
    pl.tap_groups = 1;
    const int sms = avail_sms();
    int ks = pl.total_patches < sms ? pl.total_patches : sms;
    pl.per = (pl.total_patches + ks - 1) / ks;
    pl.ksplit = (pl.total_patches + pl.per - 1) / pl.per;
    return pl;
  }
  const int tiles = pl.m_tiles * pl.n_tiles * pl.tap_groups;
  // split-K factor from a two-term cost model: MMA time at the wave efficiency of tiles*c CTAs on
  // the available SMs (one CTA per SM) + the fp32 partial tiles every split writes and the reduction re-reads
  // (measured: 11 splits of the 1024->512 layer wrote 155 MB = a third of that kernel's time).
  int ks = 1;
  double best = 1e30;
  const double flops = 2.0 * taps * static_cast<double>(Cin) * Cout * N * H * W;
  const double wbytes = 8.0 * taps * static_cast<double>(Cin) * Cout;  // 4 B written + 4 B read per split
  const int sms = avail_sms();
  for (int c = 1; c <= 160 && c <= pl.total_patches; ++c) {
    const int ctas = tiles * c;
    const int waves = (ctas + sms - 1) / sms;
    const double eff = static_cast<double>(ctas) / (static_cast<double>(sms) * waves);
    const double cost = flops / (1.1e15 * eff) + c * wbytes / 4.5e12 + 1.5e-6 * waves;
    if (cost < best - 1e-12) {
      best = cost;
      ks = c;
    }
  }
  pl.per = (pl.total_patches + ks - 1) / ks;
  pl.ksplit = (pl.total_patches + pl.per - 1) / pl.per;  // every split non-empty
  return pl;
}

thread_local const CUtensorMap* g_tmWS = nullptr;   // fp32 map of the split-K workspace (null -> per-lane stores)

template <int BN, int NSPLIT, int T, int PIX = 64, int HX = 0>
int launch_wgrad(const CUtensorMap& tmDZ, const CUtensorMap& tmX, WgradTcParams& p, dim3 grid, cudaStream_t st) {
  using L = WgradTcSmem<BN, NSPLIT, T, PIX>;
  const int stage_b = HX ? p.hx_stage : L::kStageBytes;
  int stages = (kMaxSmem - L::bytes(0)) / stage_b;
  if (stages > 8) stages = 8;
  if (stages < 2) return fail(-20, "wgrad_tc: tile BN=%d nsplit=%d does not fit shared memory", BN, NSPLIT);
  p.stages = stages;
  auto kern = wgrad_tc_kernel<BN, NSPLIT, T, PIX, HX>;
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(wgrad_tc)");
    attr_done = true;
  }
  p.ws_tma = g_tmWS ? 1 : 0;
  return launch_ex(kern, grid, 192, L::bytes(0) + stages * stage_b, st, "wgrad_tc launch", 1, tmDZ, tmX, g_tmWS ? *g_tmWS : tmDZ, p);
}

int make_act_tmap(CUtensorMap* m, const void* base, int coff, int C, int cs, int64_t plane, int nsplit, int N, int H,
                  int W, int TW, int TH) {
  const uint64_t dims[5] = {static_cast<uint64_t>(C), static_cast<uint64_t>(W), static_cast<uint64_t>(H),
                            static_cast<uint64_t>(N), static_cast<uint64_t>(nsplit)};
  const uint64_t csb = static_cast<uint64_t>(cs) * 2;
  const uint64_t pb = nsplit > 1 ? static_cast<uint64_t>(plane) * 2 : csb * W * H * N;
  const uint64_t str[4] = {csb, csb * W, csb * W * H, pb};
  const uint32_t box[5] = {64, static_cast<uint32_t>(TW), static_cast<uint32_t>(TH), 1, 1};
  return make_tmap(m, static_cast<const __nv_bfloat16*>(base) + coff, 5, dims, str, box);
}


// single-plane activation map whose 4th dimension is the 64-channel group: one TMA op brings `gbox`
// groups of a (TW x TH) patch, landing group-major ([group][pixel][64 ch]) as the MN-major operands want
int make_act_tmap_groups(CUtensorMap* m, const void* base, int coff, int C, int cs, int N, int H, int W, int TW, int TH,
                         int gbox) {
  const uint64_t dims[5] = {64, static_cast<uint64_t>(W), static_cast<uint64_t>(H), static_cast<uint64_t>(C / 64),
                            static_cast<uint64_t>(N)};
  const uint64_t csb = static_cast<uint64_t>(cs) * 2;
  const uint64_t str[4] = {csb, csb * W, 128, csb * W * H};
  const uint32_t box[5] = {64, static_cast<uint32_t>(TW), static_cast<uint32_t>(TH), static_cast<uint32_t>(gbox), 1};
  return make_tmap(m, static_cast<const __nv_bfloat16*>(base) + coff, 5, dims, str, box);
}

int launch_ctx_gemv(const float* Wm, const float* in, long long in_ns, long long in_ps, long long in_js,
                    const float* add, float* out, int N, cudaStream_t st) {
  // all images of a scale as one GEMM (a weight slab is read once per tile of positions, not once per image)
  const int pt = N >= 8 ? 64 : 32;   // positions per tile
  int tiles = 0;
  const int SS[4] = {1, 2, 3, 6};
  for (int j = 0; j < 4; ++j) tiles += ((N * SS[j] * SS[j] + pt - 1) / pt) * 8;
  if (pt == 64) ctx_pos_gemm_kernel<64><<<tiles, 256, 0, st>>>(Wm, in, in_ns, in_ps, in_js, add, out, N);
  else ctx_pos_gemm_kernel<32><<<tiles, 256, 0, st>>>(Wm, in, in_ns, in_ps, in_js, add, out, N);
  CAN_CHECK_LAUNCH("ctx_pos_gemm");
  return 0;
}

// cout, cin multiples of 32; taps <= 9
void launch_pack(const float* w, __nv_bfloat16* fwd, __nv_bfloat16* dgrad, int cout, int cin, int taps, int nsplit,
                 cudaStream_t st) {
  const dim3 grid(cin / 32, cout / 32);
  const size_t smem = 32 * (32 * taps + 1) * sizeof(float);
  if (taps == 9) pack_conv_weight_kernel_v3<9><<<grid, 256, smem, st>>>(w, fwd, dgrad, cout, cin, nsplit);
  else pack_conv_weight_kernel_v3<1><<<grid, 256, smem, st>>>(w, fwd, dgrad, cout, cin, nsplit);
}

}  // namespace

extern "C" {

const char* can_last_error_string(void) { return g_err; }
int can_version(void) { return 100; }

int can_conv_tc(const can_conv_args* a, void* stream) {
  if (!a) return fail(-1, "conv_tc: null args");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool ctx2 = a->mode == CAN_MODE_CTX2 || a->mode == CAN_MODE_CTX2_BWD;
  const bool ctx = a->mode == CAN_MODE_CTX || a->mode == CAN_MODE_CTX_BWD || ctx2;
  if (a->N <= 0 || a->H <= 0 || a->W <= 0) return fail(-2, "conv_tc: bad shape N=%d H=%d W=%d", a->N, a->H, a->W);
  if (a->Cin <= 0 || a->Cin % 64) return fail(-3, "conv_tc: Cin=%d must be a positive multiple of 64", a->Cin);
  if (a->taps != 1 && a->taps != 9) return fail(-4, "conv_tc: taps=%d (want 1 or 9)", a->taps);
  if (a->dil < 1 || a->dil > 4) return fail(-5, "conv_tc: dil=%d", a->dil);
  if (a->nsplit < 1 || a->nsplit > 3) return fail(-6, "conv_tc: nsplit=%d (want 1, 2 or 3)", a->nsplit);
  if (!a->x || !a->w) return fail(-7, "conv_tc: null x/w");
  if (a->x_cs % 8 || a->x_coff % 8 || a->out_cs % 8 || a->out_coff % 8)
    return fail(-8, "conv_tc: channel strides/offsets must be multiples of 8 (16-byte alignment)");
  if ((reinterpret_cast<uintptr_t>(a->x) & 15) || (reinterpret_cast<uintptr_t>(a->w) & 15) ||
      (reinterpret_cast<uintptr_t>(a->out) & 15))
    return fail(-9, "conv_tc: pointers must be 16-byte aligned");
  int bn = a->bn;
  if (ctx) {
    if (a->Cout != 2048 || a->Cin != 512 || a->taps != 1) return fail(-10, "conv_tc: CTX wants Cin=512 Cout=2048 taps=1");
    if (ctx2) {
      if (a->nsplit != 1 || !a->ctx_wint || !a->ctx_ttab || !a->out)
        return fail(-10, "conv_tc: CTX2 is the bf16 path (nsplit 1) and needs ctx_wint, ctx_ttab and out");
    } else if (!a->tabA || !a->tabB || !a->out) {
      return fail(-10, "conv_tc: CTX needs tabA/tabB/out");
    }
    if ((a->mode == CAN_MODE_CTX_BWD || a->mode == CAN_MODE_CTX2_BWD) && (!a->mask || a->out_cs != 4096))
      return fail(-10, "conv_tc: CTX_BWD needs the fi-gradient in `mask` and a 4096-channel output");
    bn = 256;
  } else if (a->mode == CAN_MODE_HEAD) {
    if (a->Cout != 64 || !a->head_w || !a->head_b || !a->head_out) return fail(-11, "conv_tc: HEAD wants Cout=64, head_w, head_out");
    bn = 64;
  } else {
    if (a->Cout != 64 && a->Cout != 128 && a->Cout != 256 && a->Cout != 512 && a->Cout != 1024)
      return fail(-12, "conv_tc: Cout=%d unsupported", a->Cout);
    if (!a->out) return fail(-12, "conv_tc: null out");
    if (bn == 0) bn = env_int("CAN_B200_BN", 0);
    if (bn == 0) bn = a->Cout < 256 ? a->Cout : (a->nsplit > 1 ? 128 : 256);
    if (a->nsplit == 3 && bn > 128) bn = 128;  // 3 planes x (A + B) must leave >= 2 pipeline stages
    if (bn > a->Cout) bn = a->Cout;
    if (bn != 64 && bn != 128 && bn != 256) return fail(-13, "conv_tc: bn=%d", bn);
  }
  if (a->mode == CAN_MODE_UNPOOL && (!a->mask || !a->arg || a->OH < 2 * a->H || a->OW < 2 * a->W))
    return fail(-14, "conv_tc: UNPOOL needs mask, arg and OH>=2H, OW>=2W");
  if (a->mode == CAN_MODE_DFV && (!a->mask || !a->add || !a->tabA)) return fail(-14, "conv_tc: DFV needs mask, add and tabA");
  if (a->mode < 0 || a->mode > CAN_MODE_CTX2_BWD) return fail(-15, "conv_tc: mode=%d", a->mode);

  ConvTcParams p;
  memset(&p, 0, sizeof(p));
  p.N = a->N; p.H = a->H; p.W = a->W;
  p.kchunks = a->Cin / 64;
  p.taps = a->taps;
  p.dil = a->dil;
  p.TW = (a->W >= 16) ? 16 : 8;
  // small-Cin layers on the bf16 path: resident weights + halo stages
  const bool halo = a->nsplit == 1 && a->taps == 9 && a->Cout == bn && (bn == 64 || bn == 128) &&
                    (a->mode == CAN_MODE_STORE || a->mode == CAN_MODE_POOL || a->mode == CAN_MODE_UNPOOL) &&
                    9 * p.kchunks * bn * 128 <= 147456 && env_int("CAN_B200_HALO", 1) != 0;
  // ... with ONE halo tile per (pixel tile, channel chunk): the patch is 8 pixels wide so that the 8-row groups of the
  // MMA's A operand are whole patch rows, one halo pitch apart (see the kernel).  Opt-in: CAN_B200_HALO1=1 for 64
  // output channels, 2 for every halo layer; the default stays one tile per x shift (measured
  // 1-2 % faster per 64-wide kernel, 12 % slower at 128 wide, nothing on the whole step: these layers are bound by the
  // MMA reading its operands from shared memory, not by L2->SM traffic)
  const int halo1_env = env_int("CAN_B200_HALO1", 0);
  p.halo1 = halo && a->mode != CAN_MODE_UNPOOL && (halo1_env == 2 || (halo1_env == 1 && bn == 64));
  if (p.halo1) p.TW = 8;
  // 3x3 layers that run as CTA pairs (below): the A operand comes from ONE halo tile per (pixel tile, channel chunk), the
  // nine taps being shifted views of it, while the weight tiles stream.  An SM takes ~64 B/cycle from L2; with a fresh
  // 16 KB A tile per tap and chunk a 128-wide pair tile needs 24 KB per 256 cycles of MMA (load-bound: 128->128 took
  // 222 us where its MMAs need ~130) and a 256-wide one 32 KB per 512 (just balanced).  The halo tile cuts the A traffic
  // by 6 (dil 1) / 4.7 (dil 2).  Tile = 8 x 16 pixels, as the 8-row groups of A must be whole patch rows.
  const bool halo_a = !halo && a->nsplit == 1 && a->taps == 9 && (bn == 128 || bn == 256) && a->W >= 8 && a->dil <= 2 &&
                      (a->mode == CAN_MODE_STORE || a->mode == CAN_MODE_POOL || a->mode == CAN_MODE_UNPOOL) &&
                      env_int("CAN_B200_PAIR", 1) != 0 && env_int("CAN_B200_HALOA", 1) != 0;
  if (halo_a) p.TW = 8;
  p.TH = 128 / p.TW;
  p.tiles_x = (a->W + p.TW - 1) / p.TW;
  p.tiles_y = (a->H + p.TH - 1) / p.TH;
  p.relu = a->relu;
  p.dbg = env_int("CAN_B200_DBG", 0);
  p.bias = a->bias;
  p.out = static_cast<__nv_bfloat16*>(a->out);
  p.out_plane = a->out_plane;
  p.out_cs = a->out_cs; p.out_coff = a->out_coff;
  p.out_f32 = a->out_f32;
  p.db = (a->mode == CAN_MODE_STORE || a->mode == CAN_MODE_UNPOOL || a->mode == CAN_MODE_DFV) ? a->db : nullptr;
  p.mask = static_cast<const __nv_bfloat16*>(a->mask);
  p.mask_cs = a->mask_cs; p.mask_coff = a->mask_coff; p.mask_plane = a->mask_plane;
  p.add = static_cast<const __nv_bfloat16*>(a->add);
  p.add_plane = a->add_plane; p.add_cs = a->add_cs; p.add_coff = a->add_coff;
  p.arg = a->arg; p.arg_cs = a->arg_cs;
  p.OH = a->OH; p.OW = a->OW;
  p.head_w = a->head_w; p.head_b = a->head_b; p.head_out = a->head_out;
  p.tabA = a->tabA; p.tabB = a->tabB;

  // activation map: (C, W, H, N, plane), box (64, TW, TH, 1, 1)
  CUtensorMap tmA, tmB;
  {
    const uint64_t dims[5] = {static_cast<uint64_t>(a->Cin), static_cast<uint64_t>(a->W), static_cast<uint64_t>(a->H),
                              static_cast<uint64_t>(a->N), static_cast<uint64_t>(a->nsplit)};
    const uint64_t cs = static_cast<uint64_t>(a->x_cs) * 2;
    const uint64_t plane = a->nsplit > 1 ? static_cast<uint64_t>(a->x_plane) * 2 : cs * a->W * a->H * a->N;
    const uint64_t str[4] = {cs, cs * a->W, cs * a->W * a->H, plane};
    const uint32_t box[5] = {64, static_cast<uint32_t>(p.TW), static_cast<uint32_t>(p.TH), 1, 1};
    const __nv_bfloat16* base = static_cast<const __nv_bfloat16*>(a->x) + a->x_coff;
    int rc = make_tmap(&tmA, base, 5, dims, str, box);
    if (rc) return rc;
  }
  {
    const uint64_t K = static_cast<uint64_t>(a->taps) * a->Cin;
    const uint64_t dims[3] = {K, static_cast<uint64_t>(a->Cout), static_cast<uint64_t>(a->nsplit)};
    const uint64_t str[2] = {K * 2, K * 2 * a->Cout};
    const uint32_t box[3] = {64, static_cast<uint32_t>(bn), 1};
    int rc = make_tmap(&tmB, a->w, 3, dims, str, box);
    if (rc) return rc;
  }
  // persistent CTAs: one per SM (or fewer when there are fewer tiles), striding over the tile list
  p.n_tiles = a->Cout / bn;
  p.num_tiles = p.tiles_x * p.tiles_y * a->N * p.n_tiles;
  int ctas = avail_sms();
  if (ctas > p.num_tiles) ctas = p.num_tiles;
  dim3 grid(ctas);
  const int so = a->stages > 0 ? a->stages : env_int("CAN_B200_STAGES", 0);
  const int ns = a->nsplit;
  g_tmW = nullptr;
  g_tmT = nullptr;
  // bf16 path, narrow tiles: the epilogue is store-bound when every lane writes 16 bytes to its own
  // line (measured 4-5 B/cycle/SM), so the tile is staged in shared memory and written by TMA
  CUtensorMap tmO, tmM;
  g_tmO = nullptr;
  g_tmM = nullptr;
  p.out_stage_bytes = 0;
  p.mask_bufs = 0;
  if (ns == 1 && a->mode == CAN_MODE_UNPOOL && (bn == 64 || bn == 128) && a->out && env_int("CAN_B200_TMASTORE", 1) != 0) {
    // max-unpool: the four positions of the 2x2 window are four (TH x TW) sub-tiles of the 2x upsampled map,
    // each written by one TMA store that walks the output with element stride 2 in x and y
    const uint64_t cs = static_cast<uint64_t>(a->out_cs) * 2;
    const uint64_t dimsU[4] = {static_cast<uint64_t>(a->Cout), static_cast<uint64_t>(a->OW), static_cast<uint64_t>(a->OH),
                               static_cast<uint64_t>(a->N)};
    const uint64_t strU[3] = {cs, cs * a->OW, cs * a->OW * a->OH};
    const uint32_t boxU[4] = {64, static_cast<uint32_t>(2 * p.TW), static_cast<uint32_t>(2 * p.TH), 1};
    const uint32_t estrU[4] = {1, 2, 2, 1};
    int rc = make_tmap(&tmO, static_cast<__nv_bfloat16*>(a->out) + a->out_coff, 4, dimsU, strU, boxU, true, estrU);
    if (rc) return rc;
    g_tmO = &tmO;
    p.out_stage_bytes = 2 * bn * 256;
  }
  // (halo-A pairs have the shared memory to spare: two 16 KB staging tiles; 256->256 data gradient 164 -> 156 us, the
  // other 256-wide launches unchanged, `profiles/r02_epi256_decomp.log`.  Without them it cost a ring stage: off.)
  const int stage256 = env_int("CAN_B200_STAGE256", halo_a ? 2 : 0);
  if (ns == 1 && a->mode == CAN_MODE_STORE && bn == 256 && a->out && !a->out_f32 && stage256 != 0 &&
      env_int("CAN_B200_TMASTORE", 1) != 0) {
    // widest tiles: 64-channel blocks through two 16 KB staging tiles, written by TMA (before the halo-A pairs this took a
    // fourth stage away from the operand ring and made the step 3 % slower)
    const uint64_t cs = static_cast<uint64_t>(a->out_cs) * 2;
    const uint64_t dimsO[4] = {static_cast<uint64_t>(a->Cout), static_cast<uint64_t>(a->W), static_cast<uint64_t>(a->H),
                               static_cast<uint64_t>(a->N)};
    const uint64_t strO[3] = {cs, cs * a->W, cs * a->W * a->H};
    const uint32_t boxO[4] = {64, static_cast<uint32_t>(p.TW), static_cast<uint32_t>(p.TH), 1};
    int rc = make_tmap(&tmO, static_cast<__nv_bfloat16*>(a->out) + a->out_coff, 4, dimsO, strO, boxO);
    if (rc) return rc;
    g_tmO = &tmO;
    p.out_bufs = stage256 == 2 ? 2 : 1;
    p.out_stage_bytes = p.out_bufs * 16384;
  }
  if (ns == 1 && a->mode == CAN_MODE_STORE && (bn == 64 || bn == 128) && a->out && env_int("CAN_B200_TMASTORE", 1) != 0) {
    const uint64_t cs = static_cast<uint64_t>(a->out_cs) * 2;
    const uint64_t dimsO[4] = {static_cast<uint64_t>(a->Cout), static_cast<uint64_t>(a->W), static_cast<uint64_t>(a->H),
                               static_cast<uint64_t>(a->N)};
    const uint64_t strO[3] = {cs, cs * a->W, cs * a->W * a->H};
    const uint32_t boxO[4] = {64, static_cast<uint32_t>(p.TW), static_cast<uint32_t>(p.TH), 1};
    int rc = make_tmap(&tmO, static_cast<__nv_bfloat16*>(a->out) + a->out_coff, 4, dimsO, strO, boxO);
    if (rc) return rc;
    g_tmO = &tmO;
    p.out_bufs = (bn == 64 && env_int("CAN_B200_OUTBUFS", 2) == 2) ? 2 : 1;
    p.out_stage_bytes = p.out_bufs * bn * 256;
    if (a->mask && a->mask_cs % 8 == 0 && a->mask_coff % 8 == 0 && (reinterpret_cast<uintptr_t>(a->mask) & 15) == 0 &&
        a->mode == CAN_MODE_STORE && env_int("CAN_B200_TMAMASK", 1) != 0) {
      const uint64_t mcs = static_cast<uint64_t>(a->mask_cs) * 2;
      const uint64_t strM[3] = {mcs, mcs * a->W, mcs * a->W * a->H};
      rc = make_tmap(&tmM, static_cast<const __nv_bfloat16*>(a->mask) + a->mask_coff, 4, dimsO, strM, boxO);
      if (rc) return rc;
      g_tmM = &tmM;
      // (halo-A pairs: 128->128 dgrad 217 -> 193 us with two; with the 30 KB halo stages of dil 2 a second buffer would
      // leave the weight ring two stages)
      p.mask_bufs = env_int("CAN_B200_MASKBUFS", (bn == 64 || (halo_a && a->dil == 1)) ? 2 : 1);
    }
  }
  // small-Cin layers on the bf16 path: resident weights + halo stages (L2->SM traffic / 3.6, / 9 with one halo tile)
  if (halo) {
    p.w_bytes = 9 * p.kchunks * bn * 128;
    const int halo_w = p.halo1 ? p.TW + 2 * a->dil : p.TW;
    p.halo_bytes = (p.TH + 2 * a->dil) * halo_w * 128;
    p.halo_stride = (p.halo_bytes + 1023) & ~1023;   // every stage starts on a swizzle-pattern boundary
    // staged output only while at least two halo stages still fit beside the resident weights
    while (p.mask_bufs > 0 &&
           1024 + p.out_stage_bytes + p.mask_bufs * bn * 256 + p.w_bytes + 2 * p.halo_stride + 8192 > kMaxSmem)
      --p.mask_bufs;
    if (!p.mask_bufs) g_tmM = nullptr;
    if (a->mode == CAN_MODE_STORE && p.out_bufs == 2 &&
        1024 + p.out_stage_bytes + p.mask_bufs * bn * 256 + p.w_bytes + 3 * p.halo_stride + 8192 > kMaxSmem) {
      p.out_bufs = 1;   // keep at least three halo stages
      p.out_stage_bytes = bn * 256;
    }
    if (p.out_stage_bytes && 1024 + p.out_stage_bytes + p.w_bytes + 2 * p.halo_stride + 8192 > kMaxSmem) {
      p.out_stage_bytes = 0;
      p.out_bufs = 0;
      g_tmO = nullptr;
    }
    CUtensorMap tmH;
    int rc = make_act_tmap(&tmH, a->x, a->x_coff, a->Cin, a->x_cs, a->x_plane, 1, a->N, a->H, a->W, halo_w,
                           p.TH + 2 * a->dil);
    if (rc) return rc;
    return bn == 64 ? dispatch_halo<64>(a->mode, tmA, tmB, tmH, p, grid, st)
                    : dispatch_halo<128>(a->mode, tmA, tmB, tmH, p, grid, st);
  }
  if (ctx2) {
    CUtensorMap tmW, tmT;
    int rc = make_act_tmap(&tmW, a->ctx_wint, 0, 64, 64, 0, 1, 1, a->H, a->W, p.TW, p.TH);
    if (rc) return rc;
    const uint64_t dimsT[5] = {64, 2048, static_cast<uint64_t>(a->N), 2, 1};
    const uint64_t rowb = 128, imgb = rowb * 2048, tabb = imgb * a->N;
    const uint64_t strT[4] = {rowb, imgb, tabb, tabb * 2};
    const uint32_t boxT[5] = {64, 256, 1, 1, 1};
    rc = make_tmap(&tmT, a->ctx_ttab, 5, dimsT, strT, boxT);
    if (rc) return rc;
    if (a->mode == CAN_MODE_CTX2_BWD && a->out && env_int("CAN_B200_TMASTORE", 1) != 0) {
      // the eight output segments of a tile leave through TMA stores of [pixel][16 ch] sub-tiles (2 x 32 KB staging)
      const uint64_t cs = static_cast<uint64_t>(a->out_cs) * 2;
      const uint64_t dimsD[4] = {4096, static_cast<uint64_t>(a->W), static_cast<uint64_t>(a->H), static_cast<uint64_t>(a->N)};
      const uint64_t strD[3] = {cs, cs * a->W, cs * a->W * a->H};
      const uint32_t boxD[4] = {16, static_cast<uint32_t>(p.TW), static_cast<uint32_t>(p.TH), 1};
      rc = make_tmap(&tmO, a->out, 4, dimsD, strD, boxD, false);
      if (rc) return rc;
      g_tmO = &tmO;
      p.out_stage_bytes = 65536;
    }
    if (a->mode == CAN_MODE_CTX2 && a->out_cs % 8 == 0 && a->out_coff % 8 == 0 && env_int("CAN_B200_TMASTORE", 1) != 0 &&
        env_int("CAN_B200_CTX2STAGE", 1) != 0) {
      // the tile's 64 fused channels leave through one TMA store of a swizzled [pixel][64 ch] staging tile (16 KB, fits
      // beside the four operand stages) instead of per-lane 32-byte stores
      const uint64_t cs = static_cast<uint64_t>(a->out_cs) * 2;
      const uint64_t dimsF[4] = {512, static_cast<uint64_t>(a->W), static_cast<uint64_t>(a->H), static_cast<uint64_t>(a->N)};
      const uint64_t strF[3] = {cs, cs * a->W, cs * a->W * a->H};
      const uint32_t boxF[4] = {64, static_cast<uint32_t>(p.TW), static_cast<uint32_t>(p.TH), 1};
      rc = make_tmap(&tmO, static_cast<__nv_bfloat16*>(a->out) + a->out_coff, 4, dimsF, strF, boxF);
      if (rc) return rc;
      g_tmO = &tmO;
      p.out_bufs = 1;
      p.out_stage_bytes = 16384;
    }
    g_tmW = &tmW;
    g_tmT = &tmT;
    rc = a->mode == CAN_MODE_CTX2 ? launch_conv<256, 1, MODE_CTX2>(tmA, tmB, p, grid, so, st)
                                  : launch_conv<256, 1, MODE_CTX2_BWD>(tmA, tmB, p, grid, so, st);
    g_tmW = nullptr;
    g_tmT = nullptr;
    g_tmO = nullptr;
    return rc;
  }
  if (ctx) {
    if (a->mode == CAN_MODE_CTX)
      return ns == 1   ? launch_conv<256, 1, MODE_CTX>(tmA, tmB, p, grid, so, st)
             : ns == 2 ? launch_conv<256, 2, MODE_CTX>(tmA, tmB, p, grid, so, st)
                       : launch_conv<256, 3, MODE_CTX>(tmA, tmB, p, grid, so, st);
    return ns == 1   ? launch_conv<256, 1, MODE_CTX_BWD>(tmA, tmB, p, grid, so, st)
           : ns == 2 ? launch_conv<256, 2, MODE_CTX_BWD>(tmA, tmB, p, grid, so, st)
                     : launch_conv<256, 3, MODE_CTX_BWD>(tmA, tmB, p, grid, so, st);
  }
  if (a->mode == CAN_MODE_HEAD) {
    return ns == 1   ? launch_conv<64, 1, MODE_HEAD>(tmA, tmB, p, grid, so, st)
           : ns == 2 ? launch_conv<64, 2, MODE_HEAD>(tmA, tmB, p, grid, so, st)
                     : launch_conv<64, 3, MODE_HEAD>(tmA, tmB, p, grid, so, st);
  }
  // bf16 path, 128- and 256-wide tiles: CTA pairs (cta_group::2).  Each CTA loads its own pixel tile and half of the
  // n-tile's weight rows: a third less shared-memory operand traffic per MMA at BN = 256 and every weight tile fetched
  // from L2 once per pair.  Measured (profiles/r02_pair_ab.txt): 512->512 169 -> 158 us, its data gradient 180 -> 164 us,
  // 256->256 179 -> 159 us, the training step 7.82 -> 7.46 ms.  CAN_B200_PAIR=0 switches them off.  The narrow HALO
  // layers and the weight-gradient kernels were built as pairs too and measured slower (same file): not kept.
  const int pix_tiles = p.tiles_x * p.tiles_y * a->N;
  const int m = a->mode;
  if (ns == 1 && (bn == 128 || bn == 256) && pix_tiles >= 2 && ctas >= 2 && env_int("CAN_B200_PAIR", 1) != 0 &&
      (bn == 256 || p.kchunks % 2 == 0) &&
      (m == CAN_MODE_STORE || m == CAN_MODE_POOL || m == CAN_MODE_UNPOOL || m == CAN_MODE_DFV)) {
    CUtensorMap tmBh;   // weight map with a box of half the n-tile (rank r loads rows [n0 + r bn/2, n0 + (r + 1) bn/2))
    const uint64_t K = static_cast<uint64_t>(a->taps) * a->Cin;
    const uint64_t dimsB[3] = {K, static_cast<uint64_t>(a->Cout), 1};
    const uint64_t strB[2] = {K * 2, K * 2 * a->Cout};
    const uint32_t boxB[3] = {64, static_cast<uint32_t>(bn / 2), 1};
    int rc = make_tmap(&tmBh, a->w, 3, dimsB, strB, boxB);
    if (rc) return rc;
    const int items = ((pix_tiles + 1) / 2) * p.n_tiles;
    int pairs = ctas / 2;
    if (pairs > items) pairs = items;
    const dim3 pg(2 * pairs);
    if (halo_a) {
      CUtensorMap tmH;
      const int halo_w = p.TW + 2 * a->dil, halo_h = p.TH + 2 * a->dil;
      p.halo_bytes = halo_h * halo_w * 128;
      p.halo_stride = (p.halo_bytes + 1023) & ~1023;
      p.n_a = 2;   // halo stages: the next chunk's tile lands while this one's nine taps are multiplied
      rc = make_act_tmap(&tmH, a->x, a->x_coff, a->Cin, a->x_cs, a->x_plane, 1, a->N, a->H, a->W, halo_w, halo_h);
      if (rc) return rc;
      g_tmW = &tmH;
      if (bn == 256) {
        rc = m == CAN_MODE_STORE  ? launch_conv<256, 1, MODE_STORE, 2>(tmA, tmBh, p, pg, so, st)
             : m == CAN_MODE_POOL ? launch_conv<256, 1, MODE_POOL, 2>(tmA, tmBh, p, pg, so, st)
                                  : launch_conv<256, 1, MODE_UNPOOL, 2>(tmA, tmBh, p, pg, so, st);
      } else {
        rc = m == CAN_MODE_STORE  ? launch_conv<128, 1, MODE_STORE, 2>(tmA, tmBh, p, pg, so, st)
             : m == CAN_MODE_POOL ? launch_conv<128, 1, MODE_POOL, 2>(tmA, tmBh, p, pg, so, st)
                                  : launch_conv<128, 1, MODE_UNPOOL, 2>(tmA, tmBh, p, pg, so, st);
      }
      g_tmW = nullptr;
      return rc;
    }
    CUtensorMap tmP, tmD;   // MODE_DFV: pooling-transpose matrix (A) and the pooled-map gradient rows (B) of the extra k-block
    if (m == CAN_MODE_DFV && bn == 256 && a->ctx_wint && a->ctx_ttab && env_int("CAN_B200_DFVKB", 1) != 0) {
      rc = make_act_tmap(&tmP, a->ctx_wint, 0, 64, 64, 0, 1, 1, a->H, a->W, p.TW, p.TH);
      if (rc) return rc;
      const uint64_t dimsD[3] = {64, static_cast<uint64_t>(a->Cout), static_cast<uint64_t>(a->N)};
      const uint64_t strD[2] = {128, 128ull * a->Cout};
      const uint32_t boxD[3] = {64, static_cast<uint32_t>(bn / 2), 1};
      rc = make_tmap(&tmD, a->ctx_ttab, 3, dimsD, strD, boxD);
      if (rc) return rc;
      p.dfv_kb = 1;
      g_tmW = &tmP;
      g_tmT = &tmD;
      rc = launch_conv<256, 1, MODE_DFV, 1>(tmA, tmBh, p, pg, so, st);
      g_tmW = nullptr;
      g_tmT = nullptr;
      return rc;
    }
    if (bn == 256) {
      switch (m) {
        case CAN_MODE_STORE: return launch_conv<256, 1, MODE_STORE, 1>(tmA, tmBh, p, pg, so, st);
        case CAN_MODE_POOL: return launch_conv<256, 1, MODE_POOL, 1>(tmA, tmBh, p, pg, so, st);
        case CAN_MODE_UNPOOL: return launch_conv<256, 1, MODE_UNPOOL, 1>(tmA, tmBh, p, pg, so, st);
        default: return launch_conv<256, 1, MODE_DFV, 1>(tmA, tmBh, p, pg, so, st);
      }
    }
    switch (m) {
      case CAN_MODE_STORE: return launch_conv<128, 1, MODE_STORE, 1>(tmA, tmBh, p, pg, so, st);
      case CAN_MODE_POOL: return launch_conv<128, 1, MODE_POOL, 1>(tmA, tmBh, p, pg, so, st);
      case CAN_MODE_UNPOOL: return launch_conv<128, 1, MODE_UNPOOL, 1>(tmA, tmBh, p, pg, so, st);
      default: return launch_conv<128, 1, MODE_DFV, 1>(tmA, tmBh, p, pg, so, st);
    }
  }
  if (bn == 64)
    return ns == 1   ? dispatch_mode<64, 1>(a->mode, tmA, tmB, p, grid, so, st)
           : ns == 2 ? dispatch_mode<64, 2>(a->mode, tmA, tmB, p, grid, so, st)
                     : dispatch_mode<64, 3>(a->mode, tmA, tmB, p, grid, so, st);
  if (bn == 128)
    return ns == 1   ? dispatch_mode<128, 1>(a->mode, tmA, tmB, p, grid, so, st)
           : ns == 2 ? dispatch_mode<128, 2>(a->mode, tmA, tmB, p, grid, so, st)
                     : dispatch_mode<128, 3>(a->mode, tmA, tmB, p, grid, so, st);
  if (ns == 3) return fail(-13, "conv_tc: bn=256 does not fit with 3 planes");
  return ns == 1 ? dispatch_mode<256, 1>(a->mode, tmA, tmB, p, grid, so, st)
                 : dispatch_mode<256, 2>(a->mode, tmA, tmB, p, grid, so, st);
}

int can_pack_conv_weight(const float* w, void* dst, int cout, int cin, int taps, int nsplit, int dgrad,
                         void* stream) {
  if (!w || !dst || cout <= 0 || cin <= 0 || taps <= 0 || nsplit < 1 || nsplit > 3)
    return fail(-1, "pack_conv_weight: bad arguments");
  if (cout % 32 == 0 && cin % 32 == 0 && taps <= 9) {
    __nv_bfloat16* d = static_cast<__nv_bfloat16*>(dst);
    launch_pack(w, dgrad ? nullptr : d, dgrad ? d : nullptr, cout, cin, taps, nsplit, static_cast<cudaStream_t>(stream));
  } else {
    const long long total = static_cast<long long>(cout) * cin * taps;
    const int blocks = static_cast<int>((total + 255) / 256 > 4096 ? 4096 : (total + 255) / 256);
    pack_conv_weight_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        w, static_cast<__nv_bfloat16*>(dst), cout, cin, taps, nsplit, dgrad);
  }
  CAN_CHECK_LAUNCH("pack_conv_weight");
  return 0;
}

int can_pack_ctx_w2(const float* w0, const float* w1, const float* w2, const float* w3, void* dst, int nsplit,
                    void* stream) {
  if (!w0 || !w1 || !w2 || !w3 || !dst || nsplit < 1 || nsplit > 3) return fail(-1, "pack_ctx_w2: bad arguments");
  pack_ctx_w2_kernel<<<2048, 256, 0, static_cast<cudaStream_t>(stream)>>>(w0, w1, w2, w3,
                                                                         static_cast<__nv_bfloat16*>(dst), nsplit);
  CAN_CHECK_LAUNCH("pack_ctx_w2");
  return 0;
}

int can_transpose_f32(const float* src, float* dst, int rows, int cols, void* stream) {
  if (!src || !dst || rows <= 0 || cols <= 0) return fail(-1, "transpose_f32: bad arguments");
  dim3 grid((cols + 31) / 32, (rows + 31) / 32), block(32, 8);
  transpose_f32_kernel<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(src, dst, rows, cols);
  CAN_CHECK_LAUNCH("transpose_f32");
  return 0;
}

int can_transpose_f32_multi(const float* const* srcs, float* dst, int n, int rows, int cols, void* stream) {
  if (!srcs || !dst || n <= 0 || n > 16 || rows <= 0 || cols <= 0) return fail(-1, "transpose_f32_multi: bad arguments (n=%d)", n);
  TransposeItems it;
  memset(&it, 0, sizeof(it));
  for (int i = 0; i < n; ++i) {
    if (!srcs[i]) return fail(-1, "transpose_f32_multi: null matrix %d", i);
    it.src[i] = srcs[i];
  }
  dim3 grid((cols + 31) / 32, (rows + 31) / 32, n), block(32, 8);
  transpose_f32_multi_kernel<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(it, dst, rows, cols);
  CAN_CHECK_LAUNCH("transpose_f32_multi");
  return 0;
}

int can_conv_first_fwd(const float* x, const float* w, const float* b, void* out, int64_t out_plane, int nsplit,
                       int N, int H, int W, void* stream) {
  if (!x || !w || !b || !out || N <= 0 || H <= 0 || W <= 0 || nsplit < 1 || nsplit > 3)
    return fail(-1, "conv_first_fwd: bad arguments");
  if (H > 65535 || N > 65535) return fail(-2, "conv_first_fwd: H or N exceeds grid limits");
  dim3 grid((W + 127) / 128, H, N);
  conv_first_kernel<<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(x, w, b, static_cast<__nv_bfloat16*>(out),
                                                                        out_plane, nsplit, N, H, W);
  CAN_CHECK_LAUNCH("conv_first_fwd");
  return 0;
}

int64_t can_ctx_pool_workspace_bytes(int N, int H) {
  return static_cast<int64_t>(N) * H * 12 * 512 * sizeof(float);
}

int can_ctx_pool_fwd(const void* fv, int64_t fv_plane, int nsplit, int cs, int coff, int N, int H, int W,
                     const float* w1t, const float* w2t, float* pooled, float* tabA, float* tabB, void* ws,
                     void* stream) {
  if (!fv || !w1t || !w2t || !pooled || !tabA || !tabB || !ws) return fail(-1, "ctx_pool_fwd: null pointer");
  if (N <= 0 || H <= 0 || W <= 0 || N > 65535) return fail(-2, "ctx_pool_fwd: bad shape");
  if (cs % 2 || coff % 2) return fail(-3, "ctx_pool_fwd: cs/coff must be even");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* R = static_cast<float*>(ws);
  // TMA-streamed rows kernel whenever the channel window is 16-byte aligned (always, from this package); the plain
  // per-row kernel otherwise
  const bool aligned16 = cs % 8 == 0 && coff % 8 == 0 && fv_plane % 8 == 0 && (reinterpret_cast<uintptr_t>(fv) & 15) == 0;
  if (aligned16 && W <= 16384 && H <= 65535) {
    // segment table of a row: window boundaries (sorted, unique) and the windows containing each segment
    PoolSegments seg;
    memset(&seg, 0, sizeof(seg));
    const int SS[4] = {1, 2, 3, 6};
    int ws12[12], we12[12], nb = 0, k = 0;
    for (int j = 0; j < 4; ++j)
      for (int i = 0; i < SS[j]; ++i, ++k) {
        ws12[k] = adaptive_start(i, W, SS[j]);
        we12[k] = adaptive_end(i, W, SS[j]);
        seg.bound[nb++] = ws12[k];
        seg.bound[nb++] = we12[k];
      }
    for (int a = 1; a < nb; ++a)
      for (int b = a; b > 0 && seg.bound[b - 1] > seg.bound[b]; --b) {
        const int t = seg.bound[b];
        seg.bound[b] = seg.bound[b - 1];
        seg.bound[b - 1] = t;
      }
    int u = 1;
    for (int a = 1; a < nb; ++a)
      if (seg.bound[a] != seg.bound[u - 1]) seg.bound[u++] = seg.bound[a];
    seg.nseg = u - 1;
    for (int sgi = 0; sgi < seg.nseg; ++sgi) {
      unsigned m = 0;
      for (int w12 = 0; w12 < 12; ++w12)
        if (seg.bound[sgi] >= ws12[w12] && seg.bound[sgi + 1] <= we12[w12]) m |= 1u << w12;
      seg.mask[sgi] = static_cast<unsigned short>(m);
    }
    {
      // persistent CTAs, producer warp + two sets of 8 consumer warps
      const int smem3 = pool_rows_smem_bytes(W);
      static thread_local bool attr_done = false;
      if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(ctx_pool_rows_kernel_v3, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             pool_rows_smem_bytes(16384));
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(ctx_pool_rows)");
        attr_done = true;
      }
      const int sms = avail_sms();
      const int NH = N * H;
      const int ctas = NH < sms ? NH : sms;
      CUtensorMap tm;
      const uint64_t dims[5] = {512, static_cast<uint64_t>(W), static_cast<uint64_t>(NH), static_cast<uint64_t>(nsplit), 1};
      const uint64_t str[4] = {static_cast<uint64_t>(cs) * 2, static_cast<uint64_t>(W) * cs * 2,
                               static_cast<uint64_t>(nsplit > 1 ? fv_plane : static_cast<int64_t>(NH) * W * cs) * 2,
                               static_cast<uint64_t>(nsplit > 1 ? fv_plane * nsplit : static_cast<int64_t>(NH) * W * cs) * 2};
      const uint32_t box[5] = {64, kPoolChunk, 1, 1, 1};
      int rc = make_tmap(&tm, static_cast<const __nv_bfloat16*>(fv) + coff, 5, dims, str, box, true);
      if (rc) return rc;
      ctx_pool_rows_kernel_v3<<<ctas, 544, smem3, st>>>(tm, nsplit, R, NH, W, seg);
    }
  } else {
    ctx_pool_rows_kernel<<<dim3(H, N), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(fv), fv_plane, nsplit, cs, coff,
                                                     R, H, W);
  }
  CAN_CHECK_LAUNCH("ctx_pool_rows");
  ctx_pool_finalize_kernel<<<dim3(50, N), 512, 0, st>>>(R, pooled, H, W);
  CAN_CHECK_LAUNCH("ctx_pool_finalize");
  int rc = launch_ctx_gemv(w1t, pooled, 50 * 512, 512, 0, nullptr, tabA, N, st);
  if (rc) return rc;
  return launch_ctx_gemv(w2t, tabA, 50 * 512, 512, 0, nullptr, tabB, N, st);
}

int64_t can_mse_count_workspace_bytes(int N) { return static_cast<int64_t>(N) * 4 + 16; }

int can_mse_count_fwd_bwd(const float* et, const float* gt, int N, int HW, float* loss, float* cnt_et, float* cnt_gt,
                          float* grad, float grad_scale, void* ws, void* stream) {
  if (!et || !gt || !loss || !cnt_et || !cnt_gt || !ws) return fail(-1, "mse_count: null pointer");
  if (N <= 0 || HW <= 0) return fail(-2, "mse_count: bad shape N=%d HW=%d", N, HW);
  float* part = static_cast<float*>(ws);
  unsigned int* counter = reinterpret_cast<unsigned int*>(part + N);
  mse_count_kernel<<<N, 1024, 0, static_cast<cudaStream_t>(stream)>>>(et, gt, N, HW, loss, cnt_et, cnt_gt, grad,
                                                                     grad_scale, part, counter);
  CAN_CHECK_LAUNCH("mse_count");
  return 0;
}

}  // extern "C"  (closed so the static scratch and its set-up helper sit at file scope)

// ---------------------------------------------------------------------------------------------
// Loss/count reduction for callers that must not allocate.
//
// Some hosts of this library cannot call cudaMalloc once they are running: an inference or
// training server that reserves the whole device up front and hands plugins nothing but their
// input and output buffers, or a step loop that cannot afford the device-wide synchronisation a
// cudaMalloc/cudaFree pair implies.  For them the scratch of the fused MSE+count kernel (one fp32
// partial per image plus the "blocks finished" counter) lives in the array below instead of a
// caller-provided workspace.  The array belongs to the loaded library: every GPU that loads
// libcan_b200.so holds these 4.25 KB until the process exits.
//
// `g_mse_static_scratch` is only a host-side stand-in for the device array; handing that name to a
// kernel from host code would pass a host address.  The device address is fetched once per device
// with cudaGetSymbolAddress (below) and that pointer is what the kernel receives.
// ---------------------------------------------------------------------------------------------
#define CAN_MSE_STATIC_MAX_IMAGES 1024
static __device__ __align__(256) unsigned char g_mse_static_scratch[CAN_MSE_STATIC_MAX_IMAGES * 4 + 256];

namespace {
constexpr int kMaxDevices = 64;
void* g_mse_static_ptr[kMaxDevices] = {};  // device address of the scratch, per device ordinal

// Set-up, once per device, with that device current.  Returns 0 and the device pointer, or an error code.
int mse_static_scratch(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "mse_count_static: cudaGetDevice");
  if (dev < 0 || dev >= kMaxDevices) return fail(-30, "mse_count_static: device ordinal %d not supported", dev);
  if (!g_mse_static_ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_mse_static_scratch);
    if (e != cudaSuccess || !p) return cuda_fail(e, "mse_count_static: cudaGetSymbolAddress");
    g_mse_static_ptr[dev] = p;
  }
  *out = g_mse_static_ptr[dev];
  return 0;
}
}  // namespace

extern "C" {

int can_mse_count_static_max_images(void) { return CAN_MSE_STATIC_MAX_IMAGES; }

int can_mse_count_fwd_bwd_static(const float* et, const float* gt, int N, int HW, float* loss, float* cnt_et,
                                 float* cnt_gt, float* grad, float grad_scale, void* stream) {
  if (!et || !gt || !loss || !cnt_et || !cnt_gt) return fail(-1, "mse_count_static: null pointer");
  if (N <= 0 || HW <= 0) return fail(-2, "mse_count_static: bad shape N=%d HW=%d", N, HW);
  if (N > CAN_MSE_STATIC_MAX_IMAGES)
    return fail(-3, "mse_count_static: N=%d exceeds the built-in scratch (at most %d images per call); use "
                    "can_mse_count_fwd_bwd with a caller-provided workspace", N, CAN_MSE_STATIC_MAX_IMAGES);
  void* scratch = nullptr;
  int rc = mse_static_scratch(&scratch);
  if (rc) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* part = static_cast<float*>(scratch);
  unsigned int* counter = reinterpret_cast<unsigned int*>(part + CAN_MSE_STATIC_MAX_IMAGES);
  // The array keeps whatever the previous call (of any thread, on any stream) left in it.  Only the
  // counter has to start at zero; the partials are fully rewritten before they are read.
  cudaError_t e = cudaMemsetAsync(counter, 0, sizeof(unsigned int), st);
  if (e != cudaSuccess) return cuda_fail(e, "mse_count_static: cudaMemsetAsync");
  mse_count_kernel<<<N, 1024, 0, st>>>(et, gt, N, HW, loss, cnt_et, cnt_gt, grad, grad_scale, part, counter);
  CAN_CHECK_LAUNCH("mse_count_static");
  return 0;
}

int can_zero_ranges(float* base, const int64_t* offsets, const int32_t* counts, int n, void* stream) {
  if (!base || !offsets || !counts || n <= 0 || n > ZeroRanges::kMax)
    return fail(-1, "zero_ranges: bad arguments (at most %d slices)", ZeroRanges::kMax);
  ZeroRanges r;
  memset(&r, 0, sizeof(r));
  r.n = n;
  for (int i = 0; i < n; ++i) {
    if (offsets[i] < 0 || counts[i] < 0) return fail(-2, "zero_ranges: negative offset / count in slice %d", i);
    r.off[i] = offsets[i];
    r.count[i] = counts[i];
  }
  zero_ranges_kernel<<<n, 256, 0, static_cast<cudaStream_t>(stream)>>>(base, r);
  CAN_CHECK_LAUNCH("zero_ranges");
  return 0;
}

int can_sgd_momentum_step(float* p, const float* g, float* buf, int64_t n, float lr, float momentum, float grad_scale,
                          int first_step, void* stream) {
  if (!p || !g || !buf || n <= 0) return fail(-1, "sgd_momentum_step: bad arguments");
  if ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(buf)) & 15)
    return fail(-2, "sgd_momentum_step: arenas must be 16-byte aligned");
  long long blocks = (n / 4 + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  sgd_momentum_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      p, g, buf, n, lr, momentum, grad_scale, first_step);
  CAN_CHECK_LAUNCH("sgd_momentum_step");
  return 0;
}

int can_bilinear_taps(int in_size, int out_size, int32_t* i0, int32_t* i1, float* l0, float* l1, void* stream) {
  if (in_size <= 0 || out_size <= 0 || !i0 || !i1 || !l0 || !l1) return fail(-1, "bilinear_taps: bad arguments");
  bilinear_taps_kernel<<<(out_size + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(in_size, out_size, i0,
                                                                                            i1, l0, l1);
  CAN_CHECK_LAUNCH("bilinear_taps");
  return 0;
}

int can_adaptive_windows(int in_size, int out_size, int32_t* start, int32_t* end, void* stream) {
  if (in_size <= 0 || out_size <= 0 || out_size > 1024 || !start || !end)
    return fail(-1, "adaptive_windows: bad arguments");
  adaptive_windows_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(in_size, out_size, start, end);
  CAN_CHECK_LAUNCH("adaptive_windows");
  return 0;
}

int can_nhwc_to_nchw_f32(const void* src, int64_t plane, int nsplit, int cs, int coff, float* dst, int N, int C, int H,
                         int W, void* stream) {
  if (!src || !dst || N <= 0 || C <= 0 || H <= 0 || W <= 0) return fail(-1, "nhwc_to_nchw_f32: bad arguments");
  nhwc_planes_to_nchw_f32_kernel<<<1184, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(src), plane, nsplit, cs, coff, dst, N, C, H, W);
  CAN_CHECK_LAUNCH("nhwc_to_nchw_f32");
  return 0;
}

int can_nchw_f32_to_nhwc(const float* src, void* dst, int64_t plane, int nsplit, int cs, int coff, int N, int C, int H,
                         int W, void* stream) {
  if (!src || !dst || N <= 0 || C <= 0 || H <= 0 || W <= 0) return fail(-1, "nchw_f32_to_nhwc: bad arguments");
  nchw_f32_to_nhwc_planes_kernel<<<1184, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      src, static_cast<__nv_bfloat16*>(dst), plane, nsplit, cs, coff, N, C, H, W);
  CAN_CHECK_LAUNCH("nchw_f32_to_nhwc");
  return 0;
}

int64_t can_conv_wgrad_workspace_bytes(int N, int H, int W, int Cin, int Cout, int taps, int nsplit) {
  if (N <= 0 || H <= 0 || W <= 0 || Cin <= 0 || Cin % 64 || Cout <= 0 || Cout % 64) return -1;
  // (the patch shape, hence the split count, can depend on the dilation: the larger of the two plans)
  const WgradPlan pl = wgrad_plan(N, H, W, Cin, Cout, taps, nsplit, 1), pl3 = wgrad_plan(N, H, W, Cin, Cout, taps, nsplit, 3);
  return static_cast<int64_t>(pl.ksplit > pl3.ksplit ? pl.ksplit : pl3.ksplit) * taps * Cout * Cin * sizeof(float);
}

}  // extern "C"
namespace {
int conv_wgrad_impl(const void* dz, int64_t dz_plane, int dz_cs, int dz_coff, const void* x, int64_t x_plane, int x_cs,
                    int x_coff, int nsplit, int N, int H, int W, int Cin, int Cout, int taps, int dil, float* grad,
                    int accumulate, float alpha, void* ws, int64_t ws_bytes, void* stream, int x_images);
}
extern "C" {

int can_conv_wgrad(const void* dz, int64_t dz_plane, int dz_cs, int dz_coff, const void* x, int64_t x_plane, int x_cs,
                   int x_coff, int nsplit, int N, int H, int W, int Cin, int Cout, int taps, int dil, float* grad,
                   int accumulate, float alpha, void* ws, int64_t ws_bytes, void* stream) {
  return conv_wgrad_impl(dz, dz_plane, dz_cs, dz_coff, x, x_plane, x_cs, x_coff, nsplit, N, H, W, Cin, Cout, taps, dil, grad,
                         accumulate, alpha, ws, ws_bytes, stream, 0);
}

int64_t can_ctx_ut_workspace_bytes(int B, int H, int W, int Cin) {
  if (B <= 0 || H <= 0 || W <= 0 || Cin <= 0 || Cin % 256) return -1;
  const WgradPlan pl = wgrad_plan(1, H, W, B * Cin, 64, 1, 1);
  return static_cast<int64_t>(pl.ksplit) * 64 * B * Cin * sizeof(float);
}

int can_ctx_ut(const void* wint, int wint_cs, const void* x, int x_cs, int x_coff, int B, int H, int W, int Cin, float* out,
               void* ws, int64_t ws_bytes, void* stream) {
  if (B <= 0 || Cin % 256) return fail(-2, "ctx_ut: B=%d Cin=%d (Cin must be a multiple of 256)", B, Cin);
  return conv_wgrad_impl(wint, 0, wint_cs, 0, x, 0, x_cs, x_coff, 1, 1, H, W, Cin, 64, 1, 1, out, 0, 1.f, ws, ws_bytes, stream, B);
}

}  // extern "C"
namespace {
// x_images == 0: the plain weight gradient.  x_images = B > 0 (single plane, taps == 1): dz is ONE image shared by the
// B images of x; grad is [Cout][B][Cin].
int conv_wgrad_impl(const void* dz, int64_t dz_plane, int dz_cs, int dz_coff, const void* x, int64_t x_plane, int x_cs,
                    int x_coff, int nsplit, int N, int H, int W, int Cin, int Cout, int taps, int dil, float* grad,
                    int accumulate, float alpha, void* ws, int64_t ws_bytes, void* stream, int x_images) {
  const int cin_img = Cin;
  if (x_images > 0) {
    if (nsplit != 1 || taps != 1 || N != 1) return fail(-8, "ctx_ut: single plane, 1x1, one dz image only");
    Cin = x_images * cin_img;   // n-tiles enumerate (image, cin tile); the result has B * Cin columns
  }
  if (!dz || !x || !grad || !ws) return fail(-1, "conv_wgrad: null pointer");
  if (N <= 0 || H <= 0 || W <= 0) return fail(-2, "conv_wgrad: bad shape");
  if (Cin <= 0 || Cin % 64 || Cout <= 0 || Cout % 64) return fail(-3, "conv_wgrad: Cin/Cout must be multiples of 64");
  if (Cout > 128 && Cout % 128) return fail(-3, "conv_wgrad: Cout must be 64 or a multiple of 128");
  if (Cin > 256 && Cin % 256) return fail(-3, "conv_wgrad: Cin must be 64, 128 or a multiple of 256");
  if (taps != 1 && taps != 9) return fail(-4, "conv_wgrad: taps=%d", taps);
  if (nsplit < 1 || nsplit > 3) return fail(-5, "conv_wgrad: nsplit=%d", nsplit);
  if (dz_cs % 8 || dz_coff % 8 || x_cs % 8 || x_coff % 8) return fail(-6, "conv_wgrad: strides/offsets must be multiples of 8");
  const WgradPlan pl = wgrad_plan(N, H, W, Cin, Cout, taps, nsplit, dil);
  const int64_t need = static_cast<int64_t>(pl.ksplit) * taps * Cout * Cin * sizeof(float);
  if (ws_bytes < need) return fail(-7, "conv_wgrad: workspace %lld < %lld bytes", (long long)ws_bytes, (long long)need);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CUtensorMap tmDZ, tmX;
  int rc = make_act_tmap(&tmDZ, dz, dz_coff, Cout, dz_cs, dz_plane, nsplit, N, H, W, pl.TW, pl.TH);
  if (rc) return rc;
  rc = make_act_tmap(&tmX, x, x_coff, Cin, x_cs, x_plane, nsplit, N, H, W, pl.TW, pl.TH);
  if (rc) return rc;
  WgradTcParams p;
  memset(&p, 0, sizeof(p));
  p.N = N; p.H = H; p.W = W; p.Cout = Cout; p.Cin = Cin; p.taps = taps; p.dil = dil;
  p.TW = pl.TW; p.TH = pl.TH; p.tiles_x = pl.tiles_x; p.tiles_y = pl.tiles_y;
  p.m_tiles = pl.m_tiles; p.n_tiles = pl.n_tiles;
  p.patches_per_split = pl.per; p.total_patches = pl.total_patches;
  p.m_groups = Cout >= 128 ? 2 : 1;
  p.ws = static_cast<float*>(ws);
  p.tap_groups = pl.tap_groups;
  p.dbg = env_int("CAN_B200_DBG", 0);
  dim3 grid(pl.m_tiles * pl.n_tiles * pl.tap_groups, pl.ksplit);
  CUtensorMap tmWS;
  g_tmWS = nullptr;
  if (!pl.swap64 && Cin % 32 == 0 && (reinterpret_cast<uintptr_t>(ws) & 15) == 0 && env_int("CAN_B200_WSTMA", 1) != 0) {
    const uint64_t dimsW[3] = {static_cast<uint64_t>(Cin), static_cast<uint64_t>(Cout), static_cast<uint64_t>(pl.ksplit) * taps};
    const uint64_t strW[2] = {static_cast<uint64_t>(Cin) * 4, static_cast<uint64_t>(Cin) * Cout * 4};
    const uint32_t boxW[3] = {32, 128, 1};
    rc = make_tmap(&tmWS, ws, 3, dimsW, strW, boxW, true, nullptr, true);
    if (rc) return rc;
    g_tmWS = &tmWS;
  }
  if (nsplit == 1 && !pl.swap64) {
    rc = make_act_tmap_groups(&tmDZ, dz, dz_coff, Cout, dz_cs, N, H, W, pl.TW, pl.TH, p.m_groups);
    if (rc) return rc;
    rc = make_act_tmap_groups(&tmX, x, x_coff, x_images > 0 ? cin_img : Cin, x_cs, x_images > 0 ? x_images : N, H, W,
                              pl.hx ? pl.TW + 2 * dil : pl.TW, pl.TH, pl.bn / 64);   // (hx: the box is the halo tile)
    if (rc) return rc;
    if (pl.hx) {
      p.hx_bytes = (pl.bn / 64) * (pl.TW + 2 * dil) * pl.TH * 128;
      p.hx_stage = 2 * pl.pix * 128 + ((p.hx_bytes + 1023) & ~1023);
    }
    p.x_img_tiles = x_images > 0 ? cin_img / pl.bn : 0;
  }
  if (pl.swap64) {
    using L = WgradSwapSmem;
    rc = make_act_tmap(&tmX, x, x_coff, Cin, x_cs, x_plane, nsplit, N, H, W, pl.TW, pl.TH + 2 * dil);   // halo tile
    if (rc) return rc;
    const int stage_b = 3 * (pl.TH + 2 * dil) * pl.TW * 128 + L::kTile;
    p.stages = (kMaxSmem - 2048) / stage_b;
    if (p.stages > 8) p.stages = 8;
    if (p.stages < 2) return fail(-20, "wgrad_tc_swap64: dil=%d does not fit shared memory", dil);
    static bool attr_done = false;
    if (!attr_done) {
      cudaError_t e = cudaFuncSetAttribute(wgrad_tc_swap64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem);
      if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(wgrad_tc_swap64)");
      attr_done = true;
    }
    rc = launch_ex(wgrad_tc_swap64_kernel, dim3(pl.ksplit), 192, 2048 + p.stages * stage_b, st, "wgrad_tc_swap64 launch", 1,
                   tmDZ, tmX, p);
    if (rc) return rc;
    return launch_ex(wgrad_reduce_kernel_t, dim3(Cout / 32, Cin), 32 * taps, 0, st, "wgrad_reduce_t", 1,
                     static_cast<const float*>(p.ws), grad, pl.ksplit, taps, Cout, Cin, accumulate, alpha);
  }
  if (pl.T > 1) {  // bf16 path, several taps per CTA
    if (pl.bn == 64) rc = launch_wgrad<64, 1, 3>(tmDZ, tmX, p, grid, st);
    else if (pl.hx) rc = launch_wgrad<128, 1, 3, 32, 1>(tmDZ, tmX, p, grid, st);
    else if (pl.bn == 128) rc = launch_wgrad<128, 1, 3, 32>(tmDZ, tmX, p, grid, st);
    else rc = launch_wgrad<256, 1, 2, 32>(tmDZ, tmX, p, grid, st);
  } else if (pl.bn == 64) {
    rc = nsplit == 1   ? launch_wgrad<64, 1, 1>(tmDZ, tmX, p, grid, st)
         : nsplit == 2 ? launch_wgrad<64, 2, 1>(tmDZ, tmX, p, grid, st)
                       : launch_wgrad<64, 3, 1>(tmDZ, tmX, p, grid, st);
  } else if (pl.bn == 128) {
    rc = nsplit == 1   ? launch_wgrad<128, 1, 1>(tmDZ, tmX, p, grid, st)
         : nsplit == 2 ? launch_wgrad<128, 2, 1>(tmDZ, tmX, p, grid, st)
                       : launch_wgrad<128, 3, 1>(tmDZ, tmX, p, grid, st);
  } else {
    rc = launch_wgrad<256, 1, 1>(tmDZ, tmX, p, grid, st);
  }
  if (rc) return rc;
  return launch_ex(wgrad_reduce_kernel_v3, dim3((Cin + 127) / 128, Cout), 32 * taps, 0, st, "wgrad_reduce", 1,
                   static_cast<const float*>(p.ws), grad, pl.ksplit, taps, Cout, Cin, accumulate, alpha);
}
}  // namespace

extern "C" {

int can_im2col_first(const float* x_nchw, void* out, int64_t out_plane, int nsplit, int N, int H, int W, void* stream) {
  if (!x_nchw || !out || N <= 0 || H <= 0 || W <= 0 || nsplit < 1 || nsplit > 3) return fail(-1, "im2col_first: bad arguments");
  if (H <= 65535 && N <= 65535) {
    im2col_first_kernel_v2<<<dim3((W + 127) / 128, H, N), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        x_nchw, static_cast<__nv_bfloat16*>(out), out_plane, nsplit, N, H, W);
  } else {
    im2col_first_kernel<<<148 * 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        x_nchw, static_cast<__nv_bfloat16*>(out), out_plane, nsplit, N, H, W);
  }
  CAN_CHECK_LAUNCH("im2col_first");
  return 0;
}

namespace {
constexpr int kFirstWgradCtas = 148 * 5;
int first_act_map(CUtensorMap* m, const void* base, int cs, int coff, int N, int H, int W, int rank, int box_px) {
  const uint64_t csb = static_cast<uint64_t>(cs) * 2;
  const uint64_t dims[4] = {64, static_cast<uint64_t>(W), static_cast<uint64_t>(N) * H, 1};
  const uint64_t str[3] = {csb, csb * W, csb * W * N * H};
  const uint32_t box[4] = {64, static_cast<uint32_t>(box_px), 1, 1};
  return make_tmap(m, static_cast<const __nv_bfloat16*>(base) + coff, rank, dims, str, box);
}
}  // namespace

int can_conv_first_tc_fwd(const float* x_nchw, const void* w_packed, const float* bias, void* out, int out_cs,
                          int out_coff, int N, int H, int W, void* stream) {
  if (!x_nchw || !w_packed || !bias || !out || N <= 0 || H <= 0 || W <= 0) return fail(-1, "conv_first_tc_fwd: bad arguments");
  if (out_cs % 8 || out_coff % 8 || (reinterpret_cast<uintptr_t>(out) & 15) || (reinterpret_cast<uintptr_t>(w_packed) & 15))
    return fail(-2, "conv_first_tc_fwd: output stride/offset must be multiples of 8 channels, pointers 16-byte aligned");
  if (static_cast<long long>(N) * H > 0x7fffffffLL) return fail(-3, "conv_first_tc_fwd: N*H too large");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CUtensorMap tmW, tmO;
  {
    const uint64_t dims[2] = {64, 64};
    const uint64_t str[1] = {128};
    const uint32_t box[2] = {64, 64};
    int rc = make_tmap(&tmW, w_packed, 2, dims, str, box);
    if (rc) return rc;
  }
  int rc = first_act_map(&tmO, out, out_cs, out_coff, N, H, W, 4, 128);
  if (rc) return rc;
  FirstFwdParams p;
  p.x = x_nchw; p.bias = bias; p.N = N; p.H = H; p.W = W;
  p.tiles_x = (W + 127) / 128;
  const long long tiles = static_cast<long long>(N) * H * p.tiles_x;
  if (tiles > 0x7fffffffLL) return fail(-3, "conv_first_tc_fwd: too many tiles");
  p.num_tiles = static_cast<int>(tiles);
  static thread_local bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(conv_first_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFirstFwdSmem);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(conv_first_tc)");
    attr_done = true;
  }
  const int ctas = p.num_tiles < 148 * 7 ? p.num_tiles : 148 * 7;
  conv_first_tc_kernel<<<ctas, 128, kFirstFwdSmem, st>>>(tmW, tmO, p);
  CAN_CHECK_LAUNCH("conv_first_tc");
  return 0;
}

int64_t can_conv_first_wgrad_workspace_bytes(void) { return static_cast<int64_t>(kFirstWgradCtas) * 64 * 32 * sizeof(float); }

int can_conv_first_wgrad(const void* dz, int dz_cs, int dz_coff, const float* x_nchw, int N, int H, int W, float* grad,
                         int accumulate, float alpha, void* ws, int64_t ws_bytes, void* stream) {
  if (!dz || !x_nchw || !grad || !ws || N <= 0 || H <= 0 || W <= 0) return fail(-1, "conv_first_wgrad: bad arguments");
  if (dz_cs % 8 || dz_coff % 8 || (reinterpret_cast<uintptr_t>(dz) & 15))
    return fail(-2, "conv_first_wgrad: dz stride/offset must be multiples of 8 channels, pointer 16-byte aligned");
  if (ws_bytes < can_conv_first_wgrad_workspace_bytes()) return fail(-3, "conv_first_wgrad: workspace too small");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CUtensorMap tmDZ;
  int rc = first_act_map(&tmDZ, dz, dz_cs, dz_coff, N, H, W, 3, 128);
  if (rc) return rc;
  FirstWgradParams p;
  p.x = x_nchw; p.N = N; p.H = H; p.W = W;
  p.tiles_x = (W + 127) / 128;
  const long long patches = static_cast<long long>(N) * H * p.tiles_x;
  if (patches > 0x7fffffffLL) return fail(-3, "conv_first_wgrad: too many patches");
  p.num_patches = static_cast<int>(patches);
  p.ws = static_cast<float*>(ws);
  static thread_local bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(wgrad_first_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFirstWgSmem);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(wgrad_first_tc)");
    attr_done = true;
  }
  const int ctas = p.num_patches < kFirstWgradCtas ? p.num_patches : kFirstWgradCtas;
  wgrad_first_tc_kernel<<<ctas, 128, kFirstWgSmem, st>>>(tmDZ, p);
  CAN_CHECK_LAUNCH("wgrad_first_tc");
  wgrad_first_reduce_kernel<<<64, 256, 0, st>>>(p.ws, grad, ctas, accumulate, alpha);
  CAN_CHECK_LAUNCH("wgrad_first_reduce");
  return 0;
}

int can_crowd_sample(const uint8_t* img, int H, int W, int C, const float* dmap, int flip, int ds, float* out_img,
                     float* out_dmap, void* stream) {
  if (!img || !out_img || H <= 0 || W <= 0 || (C != 1 && C != 3) || ds < 1) return fail(-1, "crowd_sample: bad arguments");
  const int rows = H / ds, cols = W / ds;
  if (rows <= 0 || cols <= 0 || rows * ds > 65535 || rows > 65535) return fail(-2, "crowd_sample: image %dx%d vs downsample %d", H, W, ds);
  if ((dmap == nullptr) != (out_dmap == nullptr)) return fail(-3, "crowd_sample: dmap and out_dmap go together");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int OH = rows * ds, OW = cols * ds;
  crowd_image_kernel<<<dim3((OW + 255) / 256, OH), 256, 0, st>>>(img, H, W, C, flip ? 1 : 0, out_img, OH, OW);
  CAN_CHECK_LAUNCH("crowd_image");
  if (dmap) {
    crowd_dmap_kernel<<<dim3((cols + 127) / 128, rows), 128, 0, st>>>(dmap, H, W, flip ? 1 : 0, out_dmap, rows, cols,
                                                                      static_cast<float>(ds) * static_cast<float>(ds));
    CAN_CHECK_LAUNCH("crowd_dmap");
  }
  return 0;
}

int64_t can_density_map_workspace_bytes(int npts) { return static_cast<int64_t>(npts > 0 ? npts : 1) * sizeof(HeadPar); }

int can_density_map(const double* points_xy, int npts, int H, int W, float* density, void* ws, int64_t ws_bytes,
                    void* stream) {
  if (!density || H <= 0 || W <= 0 || npts < 0 || (npts > 0 && (!points_xy || !ws))) return fail(-1, "density_map: bad arguments");
  if (npts == 2 || npts == 3)
    return fail(-3, "density_map: the reference is undefined for 2 or 3 heads (missing neighbours are inf)");
  if (ws_bytes < can_density_map_workspace_bytes(npts)) return fail(-2, "density_map: workspace too small");
  if ((H + 7) / 8 > 65535) return fail(-4, "density_map: image too tall");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  HeadPar* par = static_cast<HeadPar*>(ws);
  if (npts > 0) {
    density_sigma_kernel<<<(npts + 127) / 128, 128, 0, st>>>(points_xy, npts, H, W, par);
    CAN_CHECK_LAUNCH("density_sigma");
  }
  density_gather_kernel<<<dim3((W + 31) / 32, (H + 7) / 8), 256, 0, st>>>(par, npts, H, W, density);
  CAN_CHECK_LAUNCH("density_gather");
  return 0;
}

int can_bias_grad(const void* dz, int64_t plane, int nsplit, int cs, int coff, int C, int64_t npix, float* db,
                  void* stream) {
  if (!dz || !db || C <= 0 || C > 1024 || C % 2 || npix <= 0 || cs % 2 || coff % 2) return fail(-1, "bias_grad: bad arguments");
  long long blocks = (npix + 63) / 64;
  if (blocks > 148 * 4) blocks = 148 * 4;
  if (C % 8 == 0 && 256 % (C / 8) == 0 && cs % 8 == 0 && coff % 8 == 0) {
    long long b2 = (npix + 255) / 256;
    if (b2 > 148 * 8) b2 = 148 * 8;
    bias_grad_kernel_v2<<<static_cast<int>(b2), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(dz), plane, nsplit, cs, coff, C, npix, db);
  } else {
    bias_grad_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(dz), plane, nsplit, cs, coff, C, npix, db);
  }
  CAN_CHECK_LAUNCH("bias_grad");
  return 0;
}

int can_head_bwd(const float* dout, const void* y5, int64_t y_plane, int nsplit, const float* head_w, void* dz,
                 int64_t dz_plane, int64_t npix, float* d_head_w, float* d_head_b, void* stream) {
  if (!dout || !y5 || !head_w || !dz || !d_head_w || !d_head_b || npix <= 0) return fail(-1, "head_bwd: bad arguments");
  long long blocks = (npix + 63) / 64;
  if (blocks > 148 * 4) blocks = 148 * 4;
  head_bwd_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      dout, static_cast<const __nv_bfloat16*>(y5), y_plane, nsplit, head_w, static_cast<__nv_bfloat16*>(dz), dz_plane,
      npix, d_head_w, d_head_b);
  CAN_CHECK_LAUNCH("head_bwd");
  return 0;
}

int can_ctx_wint(void* out, int64_t out_plane, int nsplit, int N, int H, int W, void* stream) {
  if (!out || nsplit < 1 || nsplit > 3 || N <= 0 || H <= 0 || W <= 0) return fail(-1, "ctx_wint: bad arguments");
  ctx_wint_kernel<<<148 * 4, 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<__nv_bfloat16*>(out), out_plane,
                                                                         nsplit, N, H, W);
  CAN_CHECK_LAUNCH("ctx_wint");
  return 0;
}

int can_ctx_pint(void* out, int H, int W, void* stream) {
  if (!out || H <= 0 || W <= 0) return fail(-1, "ctx_pint: bad arguments");
  ctx_pint_kernel<<<148 * 4, 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<__nv_bfloat16*>(out), H, W);
  CAN_CHECK_LAUNCH("ctx_pint");
  return 0;
}

int can_ctx_dtab(const float* dpool, void* D, int N, void* stream) {
  if (!dpool || !D || N <= 0) return fail(-1, "ctx_dtab: bad arguments");
  ctx_dtab_kernel<<<148, 256, 0, static_cast<cudaStream_t>(stream)>>>(dpool, static_cast<__nv_bfloat16*>(D), N);
  CAN_CHECK_LAUNCH("ctx_dtab");
  return 0;
}

int can_ctx_pos_gemv(const float* Wm, const float* in, int64_t in_ns, int64_t in_ps, int64_t in_js, const float* add,
                     float* out, int N, void* stream) {
  if (!Wm || !in || !out || N <= 0 || N > 65535) return fail(-1, "ctx_pos_gemv: bad arguments");
  if (in_ps % 4 || in_ns % 4 || in_js % 4) return fail(-2, "ctx_pos_gemv: strides must be multiples of 4 floats");
  return launch_ctx_gemv(Wm, in, in_ns, in_ps, in_js, add, out, N, static_cast<cudaStream_t>(stream));
}

int can_ctx_outer(const float* A, int64_t a_ns, int64_t a_ps, int64_t a_js, const float* B, int N, float* dW,
                  float alpha, int accumulate, void* stream) {
  if (!A || !B || !dW || N <= 0) return fail(-1, "ctx_outer: bad arguments");
  ctx_outer_kernel<<<dim3(32, 4), 256, 0, static_cast<cudaStream_t>(stream)>>>(A, a_ns, a_ps, a_js, B, N, dW, alpha,
                                                                              accumulate);
  CAN_CHECK_LAUNCH("ctx_outer");
  return 0;
}

int can_pack_conv_weight_both(const float* w, void* dst_fwd, void* dst_dgrad, int cout, int cin, int taps, int nsplit,
                              void* stream) {
  if (!w || (!dst_fwd && !dst_dgrad) || cout <= 0 || cin <= 0 || taps <= 0 || taps > 9 || nsplit < 1 || nsplit > 3)
    return fail(-1, "pack_conv_weight_both: bad arguments");
  if (cout % 32 || cin % 32) return fail(-2, "pack_conv_weight_both: cout and cin must be multiples of 32");
  launch_pack(w, static_cast<__nv_bfloat16*>(dst_fwd), static_cast<__nv_bfloat16*>(dst_dgrad), cout, cin, taps, nsplit,
              static_cast<cudaStream_t>(stream));
  CAN_CHECK_LAUNCH("pack_conv_weight_both");
  return 0;
}

int can_pack_conv_weights_multi(const can_pack_item* items, int n, int nsplit, void* stream) {
  if (!items || n <= 0 || n > PackItems::kMax || nsplit < 1 || nsplit > 3)
    return fail(-1, "pack_conv_weights_multi: bad arguments (at most %d items)", PackItems::kMax);
  PackItems it;
  memset(&it, 0, sizeof(it));
  it.n = n;
  int blocks = 0;
  for (int i = 0; i < n; ++i) {
    const can_pack_item& a = items[i];
    if (!a.w || (!a.dst_fwd && !a.dst_dgrad) || a.cout <= 0 || a.cin <= 0 || a.cout % 32 || a.cin % 32 ||
        (a.taps != 9 && a.taps != 1))
      return fail(-2, "pack_conv_weights_multi: item %d: cout/cin must be positive multiples of 32, taps 9 or 1", i);
    it.first[i] = blocks;
    it.w[i] = a.w;
    it.fwd[i] = static_cast<__nv_bfloat16*>(a.dst_fwd);
    it.dgrad[i] = static_cast<__nv_bfloat16*>(a.dst_dgrad);
    it.cout[i] = a.cout;
    it.cin[i] = a.cin;
    it.taps[i] = a.taps;
    blocks += (a.cin / 32) * (a.cout / 32);
  }
  it.first[n] = blocks;
  const size_t smem = 32 * (32 * 9 + 1) * sizeof(float);
  pack_conv_weights_multi_kernel<<<blocks, 256, smem, static_cast<cudaStream_t>(stream)>>>(it, nsplit);
  CAN_CHECK_LAUNCH("pack_conv_weights_multi");
  return 0;
}

int can_ctx_ttab(const float* tabA, const float* tabB, void* T, int N, void* stream) {
  if (!tabA || !tabB || !T || N <= 0) return fail(-1, "ctx_ttab: bad arguments");
  ctx_ttab_kernel<<<148 * 4, 256, 0, static_cast<cudaStream_t>(stream)>>>(tabA, tabB, static_cast<__nv_bfloat16*>(T), N);
  CAN_CHECK_LAUNCH("ctx_ttab");
  return 0;
}

}  // extern "C"
