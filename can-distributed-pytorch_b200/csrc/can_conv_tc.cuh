// Implicit-GEMM convolution on tcgen05/TMEM fed by TMA (sm_100a).
//
//   D[pixel][cout] = sum_{tap, cin} X[pixel + off(tap)][cin] * Wp[cout][tap][cin]
//
// Persistent CTAs (one per SM) stride over 128-pixel (TH x TW spatial patch of one image) x BN-channel tiles:
//   warp 0        TMA producer: per k-block one 5-D box of the NHWC activation shifted by the
//                 tap offset (out-of-image pixels are zero-filled by TMA = conv padding) and
//                 one box of the K-major packed weights, both 128B-swizzled.
//   warp 1        allocates TMEM (2-4 accumulator stages) and issues tcgen05.mma (M=128, N=BN, K=16).
//   warps 2..2+EW epilogue (EW = 4 or 8): tcgen05.ld the fp32 accumulator, fuse bias / ReLU / 2x2 max-pool /
//                 1x1 density head / ReLU-mask / max-unpool scatter / contextual fusion.
//   warp 2+EW     output warp (staged epilogues): hands bf16 tiles staged in shared memory to TMA stores.
//   warp 3+EW     mask warp (staged dgrad epilogue): brings ReLU-mask tiles in by TMA.
// Producer, MMA issuer and the two helper warps each run on one lane chosen with elect.sync.
// NSPLIT > 1 is the fp32-parity mode: activations and weights are carried as NSPLIT bf16
// planes (x = hi + lo [+ lo2]) and the products hi*hi, hi*lo, lo*hi (...) are accumulated
// in the same fp32 TMEM tile.
#pragma once
#include "can_ptx.cuh"
#include "can_simple.cuh"

namespace canb {

enum ConvMode : int {
  MODE_STORE = 0,   // y = [relu](acc + bias) [* (mask > 0)]  -> NHWC bf16 planes (+ optional fp32 copy)
  MODE_POOL = 1,    // y = maxpool2x2(relu(acc + bias))        -> pooled NHWC (+ argmax codes)
  MODE_HEAD = 2,    // y = relu(acc + bias) stored, and density = y . head_w + head_b (BN == Cout == 64)
  MODE_UNPOOL = 3,  // g = acc * (mask > 0); scatter to the arg-max position of a 2x upsampled map
  MODE_CTX = 4,     // contextual fusion epilogue: fi from the stacked contrast GEMM
  MODE_CTX_BWD = 5, // same GEMM recomputed; writes dz_j and the direct ds_j (backward of the fusion)
  MODE_DFV = 6,     // dfv = (add - acc + pooled-gradient spread) * (mask > 0)
  MODE_CTX2 = 7,    // contextual fusion with the bilinear interpolation folded into the GEMM (bf16 path)
  MODE_CTX2_BWD = 8, // backward of the fusion on the same extended GEMM (bf16 path)
};

struct ConvTcParams {
  int N, H, W;         // spatial size of the GEMM's pixel grid (output == input grid, stride 1)
  int kchunks;         // Cin / 64
  int taps;            // 9 (3x3) or 1 (1x1)
  int dil;             // tap spacing (1 | 2); padding == dil
  int TW, TH;          // spatial tile, TW*TH == 128, TW in {8,16}
  int tiles_x, tiles_y;
  int n_tiles;         // Cout / BN
  int num_tiles;       // tiles_x * tiles_y * N * n_tiles (persistent CTAs stride over them)
  int stages;
  int dbg;             // CAN_B200_DBG bit flags (bring-up experiments only; 0 in production)
  int w_bytes;         // HALO: bytes of the resident weight region (9 * kchunks * BN * 128)
  int halo_bytes;      // HALO: bytes of one activation stage ((TH + 2 dil) * TW * 128; halo1: (TH + 2 dil) * (TW + 2 dil) * 128)
  int halo_stride;     // HALO: distance between stages (halo_bytes rounded up to 1024)
  int halo1;           // HALO: one halo tile per (pixel tile, channel chunk) serves all nine taps
  int n_a;             // PAIR == 2: number of A halo stages (each halo_stride bytes) in front of the weight ring
  int dfv_kb;          // MODE_DFV as a pair, BN = 256: the pooling-transpose term is one more k-block (A = tmW: can_ctx_pint
                       // tile, B = tmT: can_ctx_dtab rows) instead of thirteen table rows per pixel in the epilogue
  int out_stage_bytes; // > 0: the epilogue stages the bf16 tile in shared memory (BN * 256 B) and writes it with TMA (tmO)
  int out_bufs;        // MODE_STORE staged epilogue: 1 or 2 staging buffers of BN * 256 B (out_stage_bytes = their sum)
  int mask_bufs;       // staged epilogue: 0 = mask read from global memory, 1|2 = mask tiles (BN * 256 B each) brought in by TMA (tmM)
  int relu;
  const float* bias;   // [Cout] or null
  // primary output, NHWC bf16 planes
  __nv_bfloat16* out;
  long long out_plane;  // elements between planes
  int out_cs, out_coff; // pixel stride (channels) and channel offset
  float* out_f32;       // optional fp32 NHWC copy (same cs/coff addressing), may be null
  float* db;            // optional [Cout]: += column sums of the stored tile (bias gradient of the layer whose dZ this is)
  // mask source (dgrad): activation whose sign gates the gradient, NHWC bf16 (plane 0)
  const __nv_bfloat16* mask;
  int mask_cs, mask_coff;
  long long mask_plane;
  // additive source (MODE_DFV: upstream gradient that reaches fv through the concat)
  const __nv_bfloat16* add;
  long long add_plane;
  int add_cs, add_coff;
  // pool / unpool
  uint8_t* arg;         // [N][H/2][W/2][C] argmax codes (MODE_POOL writes if non-null, MODE_UNPOOL reads)
  int arg_cs;
  int OH, OW;           // MODE_UNPOOL: size of the upsampled map
  // head
  const float* head_w;  // [64]
  const float* head_b;  // [1]
  float* head_out;      // [N][H][W]
  // ctx
  const float* tabA;    // [N][50][512] a_j  = W_j1 pool_j
  const float* tabB;    // [N][50][512] b_j  = W_j2 a_j
};

// Planes that enter the GEMM.  The contextual GEMMs feed a sigmoid, where two planes (16
// mantissa bits) are already far below the fp32 tolerance; everything else uses all planes.
__host__ __device__ constexpr int gemm_planes(int nsplit, int mode) {
  return ((mode == MODE_CTX || mode == MODE_CTX_BWD) && nsplit > 2) ? 2 : nsplit;
}

template <int BN, int NSPLIT>
struct ConvTcSmem {
  static constexpr int kABytes = 128 * 128;       // 128 pixels x 64 ch x bf16
  static constexpr int kBBytes = BN * 128;        // BN rows x 64 k x bf16
  static constexpr int kStageBytes = NSPLIT * (kABytes + kBBytes);
  static constexpr int kCtxTabBytes = 2 * 50 * 64 * 4;
  static constexpr int kTailBytes = 1024 + 2 * BN * 4 + 4096;  // barriers + tmem ptr, bias (double-buffered), s_db[1024]
  // table region after the tail: MODE_CTX / MODE_CTX_BWD [2][50][64] fp32, MODE_DFV [50][BN] fp32
  static __host__ __device__ constexpr int tab_bytes(int mode) {
    return (mode == MODE_CTX || mode == MODE_CTX_BWD) ? kCtxTabBytes : (mode == MODE_DFV ? 50 * BN * 4 : 0);
  }
  static __host__ __device__ constexpr int bytes(int stages, int mode) {
    return 1024 /*align slack*/ + stages * kStageBytes + kTailBytes + tab_bytes(mode);
  }
};

template <int NSPLIT, int NV>
__device__ __forceinline__ void store_bf16_planes(__nv_bfloat16* base, long long plane, const float* v) {
  // NV (multiple of 8) values; base is 16B aligned.  Plane k holds bf16(v - sum of planes < k).
#pragma unroll
  for (int pl = 0; pl < NSPLIT; ++pl) {
    __nv_bfloat16* dst = base + pl * plane;
#pragma unroll
    for (int i = 0; i < NV; i += 8) {
      {
        float r[8];
        uint4 q;
#pragma unroll
        for (int e = 0; e < 8; ++e) r[e] = v[i + e];
        if (pl > 0) {
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            float acc = r[e];
#pragma unroll
            for (int pp = 0; pp < pl; ++pp) {
              float h = bf16_round(acc);
              acc = acc - h;
            }
            r[e] = acc;
          }
        }
        q.x = pack_bf16x2(r[0], r[1]);
        q.y = pack_bf16x2(r[2], r[3]);
        q.z = pack_bf16x2(r[4], r[5]);
        q.w = pack_bf16x2(r[6], r[7]);
        *reinterpret_cast<uint4*>(dst + i) = q;
      }
    }
  }
}


// fp32 accumulation inside the tensor core truncates: every tcgen05.mma that adds into a TMEM
// accumulator loses ~2^-25 of the accumulator's magnitude, biased, so a chain of n MMAs drifts
// by ~n * 3e-8 (measured: 1.5e-5 after 864 MMAs).  The fp32-parity path therefore spreads the
// dominant hi*hi products of successive k-blocks over NACC accumulators (short chains), keeps
// the small cross products (hi*lo, lo*hi, ...) in one more, and adds them up here in fp32 with
// round-to-nearest.  GS == 1 (bf16 path): one accumulator, one load.
template <int BN, int GS>
struct AccLayout {
  static constexpr int kNacc = (GS >= 2) ? (512 / BN - 1) : 1;   // hi*hi accumulators
  static constexpr int kCols = (GS >= 2) ? 512 : (BN < 32 ? 32 : BN);
  static constexpr int cols(int mode) { return (mode == MODE_CTX2 || mode == MODE_CTX2_BWD) ? 512 : kCols; }
  // two accumulator stages when they fit: the epilogue of tile i overlaps the MMAs of tile i+1
  // up to four accumulator stages: with narrow tiles an MMA tile lasts ~1 us, so any hiccup of the epilogue
  // (a barrier, a TMA hand-over) stalls the tensor pipe unless the MMA warp can run several tiles ahead
  static constexpr int acc_stages(int mode) { return cols(mode) * 4 <= 512 ? 4 : (cols(mode) * 2 <= 512 ? 2 : 1); }
};

template <int NV>
__device__ __forceinline__ void tmem_ld_cols(uint32_t taddr, uint32_t* r) {
  if (NV == 32) tmem_ld_32x32(taddr, r);
  else if (NV == 16) tmem_ld_32x16(taddr, r);
  else tmem_ld_32x8(taddr, r);
}

template <int BN, int GS, int NV>
__device__ __forceinline__ void load_acc(uint32_t taddr, int col, int nact, float* v) {
  uint32_t r[NV];
  tmem_ld_cols<NV>(taddr + col, r);
  tmem_ld_wait();
#pragma unroll
  for (int i = 0; i < NV; ++i) v[i] = __uint_as_float(r[i]);
  if (GS >= 2) {
    for (int a = 1; a <= nact; ++a) {
      // a < nact: further hi*hi accumulators; a == nact: the cross-product accumulator
      const int idx = (a < nact) ? a : AccLayout<BN, GS>::kNacc;
      tmem_ld_cols<NV>(taddr + idx * BN + col, r);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < NV; ++i) v[i] += __uint_as_float(r[i]);
    }
  }
}

// Sum each of 32 columns over the 32 lanes of the warp (lane = pixel): recursive halving, 31
// shuffles; on return lane l holds the sum of column l in v[0].  v[] is destroyed.
__device__ __forceinline__ float warp_column_sums(float* v, int lane) {
#pragma unroll
  for (int half = 16; half >= 1; half >>= 1) {
    const bool upper = (lane & half) != 0;
#pragma unroll
    for (int i = 0; i < half; ++i) {
      const float send = upper ? v[i] : v[i + half];
      const float keep = upper ? v[i + half] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, half);
    }
  }
  return v[0];
}

// bf16-path transcendental helpers (one MUFU each): sigmoid(-a) = 0.5 - 0.5 tanh(a/2), 1/x.
// The epilogue runs on four warps (one per scheduler), so every dependent multi-instruction IEEE
// division costs its full latency; measured: 380 us of a 533 us kernel before this change.
__device__ __forceinline__ float sigmoid_neg_fast(float a) {
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.5f * a));
  return fmaf(-0.5f, t, 0.5f);
}
__device__ __forceinline__ float rcp_fast(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// 32 consecutive bias values of the tile from shared memory as eight 16-byte reads.  Every lane reads the same address
// (a broadcast: one wavefront per instruction whatever its width), and the shared-memory port is what the narrow
// layers run out of (the MMA's operand reads take 61 % of its cycles at N = 64, `profiles/r01_conv64_ncu_full.txt`),
// so 32 scalar reads per warp and column group cost four times the wavefronts of these.  `s` is 16-byte aligned.
__device__ __forceinline__ void load_bias32(const float* s, float* b) {
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float4 t = reinterpret_cast<const float4*>(s)[i];
    b[4 * i] = t.x; b[4 * i + 1] = t.y; b[4 * i + 2] = t.z; b[4 * i + 3] = t.w;
  }
}

// HALO = 1 (small Cin, bf16 path): all weights of the CTA's output channels stay resident in shared
// memory and the activation patch is loaded with a (TH + 2 dil)-row halo once per (dx, channel chunk);
// the three dy taps are row-shifted views of that stage (row shifts are multiples of 1024 B, so the
// 128B swizzle phase is preserved).  L2->SM traffic per tile drops from 9 x (A + B) to 3 x halo.
// EW = number of epilogue warps (4 or 8).  The epilogue runs one warp per scheduler, i.e. latency
// bound; for narrow tiles (small K) it is the critical path, so EW = 8 lets two warps share each TMEM
// lane quarter, each taking half of the BN columns.
// PAIR = 1 (bf16 path): two CTAs of a cluster on the two SMs of a TPC work as a cta_group::2 pair.  One MMA covers two
// pixel tiles of the same n-tile (M = 256: each CTA's own 128-pixel A tile is its half of A) and each CTA holds only
// HALF of the n-tile's weight rows (rows [rank * BN/2, (rank + 1) * BN/2) of B): the tensor core of each SM reads
// A + B/2 instead of A + B from its shared memory, and the pair fetches every weight tile from L2 once instead of twice
// (wide tiles: 32 KB instead of 48 KB per CTA and k-block).  The even CTA issues every MMA; both CTAs load (their TMA
// completions are counted on the leader's barriers), run the epilogue of their own tile out of their own TMEM, and the
// MMA warp's commits arrive on the barriers of both.  No TMA multicast is involved.
template <int BN, int NSPLIT, int MODE, int HALO = 0, int EW = 4, int PAIR = 0>
__global__ void __launch_bounds__(128 + 32 * EW, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmT,
               const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmM,
               const ConvTcParams p) {
  constexpr int GS = gemm_planes(NSPLIT, MODE);  // planes that enter the GEMM
#ifdef CAN_DECOMP
  constexpr bool kDecomp = true;    // tools/gpu_decomp_mainloop.py: main-loop bring-up bits of CAN_B200_DBG compiled in
#else
  constexpr bool kDecomp = false;
#endif
  using L = ConvTcSmem<BN, GS>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1024-aligned, still a shared-space pointer
  const int stages = p.stages;
  uint8_t* s_out = smem;                                           // staged output tile (p.out_stage_bytes, 1024-aligned)
  uint8_t* s_mask = smem + p.out_stage_bytes;                       // staged mask tiles (p.mask_bufs x BN * 256 B)
  uint8_t* wbase = s_mask + p.mask_bufs * (BN * 256);               // HALO: resident weights
  uint8_t* ring = HALO ? wbase + p.w_bytes : wbase;                // stage ring
  // KPS = 64-channel k-chunks per pipeline stage.  The 128-wide pair tiles take two: one thread issues the loads and one
  // the MMAs, each pays a fixed price per stage hand-shake (barrier wait ~90 cycles + issue), and four N = 128 MMAs are
  // done in 256 cycles -- with one chunk per stage those two threads, not the tensor pipe, set the pace.
  constexpr int KPS = (PAIR == 1 && BN == 128) ? 2 : 1;
  // PAIR == 2: pairs whose A operand is one halo tile per (pixel tile, channel chunk) -- the nine taps are shifted views of
  // it, as in the single-tile HALO variant -- while the weight tiles stream through the ring, TPS taps per stage.
  constexpr bool HA = PAIR == 2;
  constexpr int TPS = HA ? (BN == 128 ? 3 : 1) : 1;
  const int stage_bytes = HALO ? p.halo_stride
                          : HA ? TPS * (L::kBBytes / 2)
                               : GS * KPS * (L::kABytes + (PAIR ? L::kBBytes / 2 : L::kBBytes));
  // single-halo-tile variant: not compiled into the max-unpool kernels (their epilogue sits on the register limit)
  const bool halo1 = HALO && MODE != MODE_UNPOOL && p.halo1;
  uint8_t* ringB = ring + (HA ? p.n_a * p.halo_stride : 0);          // PAIR == 2: weight ring behind the A halo stages
  uint8_t* tail = ringB + stages * stage_bytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tail);
  uint64_t* empty_bar = full_bar + 8;
  uint64_t* tmem_full_bar = empty_bar + 8;    // [4]
  uint64_t* tmem_empty_bar = tmem_full_bar + 4;  // [4]
  uint64_t* w_bar = tmem_empty_bar + 4;          // HALO: resident weights landed
  uint64_t* out_free_bar = w_bar + 1;            // [2] staged output: the TMA store has read this s_out buffer
  uint64_t* out_full_bar = out_free_bar + 2;     // [2] staged output: every epilogue thread has written its part of it
  uint64_t* mask_full_bar = out_full_bar + 2;    // [2] staged mask tile landed
  uint64_t* mask_free_bar = mask_full_bar + 2;   // [2] every epilogue thread is done with it
  uint64_t* fullA_bar = mask_free_bar + 2;       // [4] PAIR == 2: A halo stage landed (both CTAs' tiles, on the leader's)
  uint64_t* emptyA_bar = fullA_bar + 4;          // [4] PAIR == 2: the nine taps of that stage have been multiplied
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(emptyA_bar + 4);
  float* s_bias2 = reinterpret_cast<float*>(tail + 1024);  // [2][BN]
  float* s_db = s_bias2 + 2 * BN;   // [<=1024] per-CTA partial bias gradient (p.db)
  float* s_tab = s_db + 1024;       // MODE_CTX / MODE_CTX_BWD only: [2][50][64]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  constexpr int ACC_COLS = AccLayout<BN, GS>::cols(MODE);
  constexpr int ACC_STAGES = AccLayout<BN, GS>::acc_stages(MODE);
  const int base_kb = p.taps * p.kchunks;
  static_assert(!PAIR || (!HALO && NSPLIT == 1 && (MODE == MODE_STORE || MODE == MODE_POOL || MODE == MODE_UNPOOL || MODE == MODE_DFV)),
                "CTA pairs exist for the streamed (non-HALO) tiles of the bf16 path");
  static_assert(PAIR != 2 || (MODE != MODE_DFV && (BN == 128 || BN == 256)), "halo-A pairs: the 3x3 layers with 128/256-wide tiles");
  // Work items.  Plain: CTA b takes tiles b, b + grid, ...  Pair: the two CTAs walk the same list of tile PAIRS in
  // lockstep, rank r taking tile 2 q + r (an odd tile count leaves the last item's odd CTA a dead tile: it recomputes the
  // last tile and drops the result).
  const int rank = PAIR ? static_cast<int>(cluster_ctarank()) : 0;
  const bool leader = rank == 0;
  const int W0 = PAIR ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int WS = PAIR ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);
  const int pix_tiles = p.num_tiles / p.n_tiles;
  const int WN = PAIR ? ((pix_tiles + 1) >> 1) * p.n_tiles : p.num_tiles;
  // (n-tile fastest, as in the plain tile order: CTAs working at the same time share the activation patch in L2)
  auto tile_of = [&](int q) {
    return PAIR ? min(2 * (q / p.n_tiles) + rank, pix_tiles - 1) * p.n_tiles + q % p.n_tiles : q;
  };
  auto live_of = [&](int q) { return !PAIR || (2 * (q / p.n_tiles) + rank < pix_tiles); };
  constexpr int kBt = PAIR ? L::kBBytes / 2 : L::kBBytes;   // bytes of one weight tile in this CTA
  constexpr int kStageSz = HA ? TPS * kBt : GS * KPS * (L::kABytes + kBt);   // non-HALO stage: A tile(s) + this CTA's part of B
  // MODE_CTX2: two extra k-blocks whose A operand is the interpolation matrix Wint[pixel][64 positions]
  // and whose B operands are the per-image tables (-U-side of z, then s) -- see the epilogue.
  constexpr bool kCtx2 = (MODE == MODE_CTX2 || MODE == MODE_CTX2_BWD);
  constexpr bool kDfvX = MODE == MODE_DFV && PAIR == 1 && BN == 256;   // (see ConvTcParams::dfv_kb)
  const int num_kb = base_kb + (kCtx2 ? 2 : 0) + (kDfvX ? p.dfv_kb : 0);
  // Persistent CTA: tiles blockIdx.x, blockIdx.x + gridDim.x, ...; the n-tile index varies fastest so
  // that CTAs working at the same time share the activation patch in L2.
#define CAN_TILE_DECODE(tile)                                   \
  const int n_tile = (tile) % p.n_tiles;                        \
  int t_ = (tile) / p.n_tiles;                                  \
  const int tx_i = t_ % p.tiles_x;                              \
  t_ /= p.tiles_x;                                              \
  const int ty_i = t_ % p.tiles_y;                              \
  const int n_img = t_ / p.tiles_y;                             \
  const int x0 = tx_i * p.TW, y0 = ty_i * p.TH;                 \
  const int n0 = n_tile * BN;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (kCtx2) {
      tma_prefetch_desc(&tmW);
      tma_prefetch_desc(&tmT);
    }
    if (HA) tma_prefetch_desc(&tmW);
    if (kDfvX && p.dfv_kb) {
      tma_prefetch_desc(&tmW);
      tma_prefetch_desc(&tmT);
    }
    if (p.out_stage_bytes) tma_prefetch_desc(&tmO);
    if (p.mask_bufs) tma_prefetch_desc(&tmM);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < stages; ++s) {
        mbar_init(&full_bar[s], 1);
        mbar_init(&empty_bar[s], 1);
      }
      for (int a = 0; a < 4; ++a) {
        mbar_init(&tmem_full_bar[a], 1);
        // every epilogue thread arrives (pair: plus one arrival per epilogue warp of the odd CTA, on the leader's barrier)
        mbar_init(&tmem_empty_bar[a], 32 * EW + (PAIR ? EW : 0));
      }
      mbar_init(w_bar, 1);
      for (int a = 0; a < 4; ++a) {
        mbar_init(&fullA_bar[a], 1);
        mbar_init(&emptyA_bar[a], 1);
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(&out_free_bar[a], 1);
        mbar_init(&out_full_bar[a], 32 * EW);
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(&mask_full_bar[a], 1);
        mbar_init(&mask_free_bar[a], 32 * EW);
      }
      fence_mbar_init();
    }
    __syncwarp();
    if (PAIR) {
      tmem_alloc_pair(tmem_ptr_smem, ACC_COLS * ACC_STAGES);
      tmem_relinquish_pair();
    } else {
      tmem_alloc(tmem_ptr_smem, ACC_COLS * ACC_STAGES);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();   // the partner's barriers exist before anything signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  // everything above touched no global memory: with programmatic dependent launch it overlaps the previous kernel's tail
  pdl_launch_dependents();
  pdl_wait();

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      if (HALO) {
        // resident weights: every (tap, chunk) tile of this CTA's BN output channels, once
        mbar_expect_tx(w_bar, static_cast<uint32_t>(p.w_bytes));
        for (int kb = 0; kb < base_kb; ++kb) tma_load_3d(wbase + kb * L::kBBytes, &tmB, w_bar, kb * 64, 0, 0);
        for (int q_ = W0; q_ < WN; q_ += WS) {
        const int tile = tile_of(q_);
          CAN_TILE_DECODE(tile)
          (void)n0;
          for (int chunk = 0; chunk < p.kchunks; ++chunk) {
            for (int dxi = halo1 ? 2 : 0; dxi < 3; ++dxi) {   // halo1: a single tile, x halo included
              mbar_wait(&empty_bar[stage], phase ^ 1);
              mbar_expect_tx(&full_bar[stage], static_cast<uint32_t>(p.halo_bytes));
              tma_load_5d(ring + stage * stage_bytes, &tmW, &full_bar[stage], chunk * 64,
                          x0 + (halo1 ? -p.dil : (dxi - 1) * p.dil), y0 - p.dil, n_img, 0);
              if (++stage == stages) {
                stage = 0;
                phase ^= 1;
              }
            }
          }
        }
      } else if (HA) {
        // loads in the order the MMAs consume them: the halo tile of a chunk, then its nine weight tiles (TPS per stage)
        int kchunks64, dil, nst, nA;
        asm volatile("mov.u32 %0, %1;" : "=r"(kchunks64) : "r"(p.kchunks * 64));
        asm volatile("mov.u32 %0, %1;" : "=r"(dil) : "r"(p.dil));
        asm volatile("mov.u32 %0, %1;" : "=r"(nst) : "r"(stages));
        asm volatile("mov.u32 %0, %1;" : "=r"(nA) : "r"(p.n_a));
        const uint32_t fbA0 = mapa_shared(fullA_bar, 0), fbB0 = mapa_shared(full_bar, 0);
        const int cin = kchunks64;              // K stride between taps in the weight matrix
        const bool skipA = kDecomp && (p.dbg & 1024), skipB = kDecomp && (p.dbg & 2048);   // (CAN_DECOMP builds only)
        int sA = 0, sB = 0;
        uint32_t phA = 0, phB = 0;
        uint8_t* pa = ring;
        uint8_t* pb = ringB;
        for (int q_ = W0; q_ < WN; q_ += WS) {
          const int tile = tile_of(q_);
          CAN_TILE_DECODE(tile)
          const int nB = n0 + rank * (BN / 2);
          const int hx = x0 - dil, hy = y0 - dil;
          for (int c64 = 0; c64 < kchunks64; c64 += 64) {
            mbar_wait(&emptyA_bar[sA], phA ^ 1);
            if (leader) mbar_expect_tx(&fullA_bar[sA], skipA ? 0u : 2u * static_cast<uint32_t>(p.halo_bytes));
            if (!skipA) tma_load_5d_pair(pa, &tmW, fbA0 + 8 * sA, c64, hx, hy, n_img, 0);
            pa += p.halo_stride;
            if (++sA == nA) {
              sA = 0;
              phA ^= 1;
              pa = ring;
            }
            int kw = c64;                       // K coordinate of tap 0's tile of this chunk
#pragma unroll 1
            for (int t0 = 0; t0 < 9; t0 += TPS) {
              mbar_wait(&empty_bar[sB], phB ^ 1);
              if (leader) mbar_expect_tx(&full_bar[sB], skipB ? 0 : 2 * kStageSz);
#pragma unroll
              for (int j = 0; j < TPS; ++j, kw += cin)
                if (!skipB) tma_load_3d_pair(pb + j * kBt, &tmB, fbB0 + 8 * sB, kw, nB, 0);
              pb += kStageSz;
              if (++sB == nst) {
                sB = 0;
                phB ^= 1;
                pb = ringB;
              }
            }
          }
        }
      } else {
      // ONE thread issues every load, so whatever it spends per k-block is a floor under the whole main loop.  Measured
      // with loads, MMAs and epilogue all switched off (tools/gpu_decomp_mainloop.py): ~500 cycles per k-block when tap and
      // chunk came from a division per k-block and every address from a multiply -- as long as the four N = 256 MMAs of a
      // k-block take, and twice what the N = 128 ones do.  Hence: taps and chunks are nested counters, a tap's coordinates
      // are computed once per tap, stage / barrier addresses are running values, loop bounds live in registers.
      int kchunks64, taps, dil, nst;
      asm volatile("mov.u32 %0, %1;" : "=r"(kchunks64) : "r"(p.kchunks * 64));
      asm volatile("mov.u32 %0, %1;" : "=r"(taps) : "r"(p.taps));
      asm volatile("mov.u32 %0, %1;" : "=r"(dil) : "r"(p.dil));
      asm volatile("mov.u32 %0, %1;" : "=r"(nst) : "r"(stages));
      uint8_t* sa = ring;
      uint64_t* fbar = full_bar;
      uint64_t* ebar = empty_bar;
      const uint32_t fb0 = PAIR ? mapa_shared(full_bar, 0) : 0u;   // (shared::cluster addresses are linear inside a CTA's window)
      uint32_t fb = fb0;
      // (CAN_DECOMP builds only: bring-up experiments, the A / B tile of every k-block is not loaded)
      const bool skipA = kDecomp && (p.dbg & 1024), skipB = kDecomp && (p.dbg & 2048);
      const uint32_t tx_bytes = PAIR ? 2 * KPS * ((skipA ? 0 : L::kABytes) + (skipB ? 0 : L::kBBytes / 2))
                                     : (kDecomp && (p.dbg & (1024 | 2048)) && GS == 1 ? (skipA ? 0 : L::kABytes) + (skipB ? 0 : L::kBBytes)
                                                                           : L::kStageBytes);
      auto advance = [&]() {
        sa += kStageSz;
        ++fbar;
        ++ebar;
        fb += 8;
        if (++stage == nst) {
          stage = 0;
          phase ^= 1;
          sa = ring;
          fbar = full_bar;
          ebar = empty_bar;
          fb = fb0;
        }
      };
      for (int q_ = W0; q_ < WN; q_ += WS) {
        const int tile = tile_of(q_);
        CAN_TILE_DECODE(tile)
        const int nB = PAIR ? n0 + rank * (BN / 2) : n0;
        int kb64 = 0;   // 64 x the k-block index = the weight tile's K coordinate
        for (int tap = 0; tap < taps; ++tap) {
          int cx = x0, cy = y0;
          if (taps == 9) {
            const int r = tap / 3;
            cy += (r - 1) * dil;
            cx += (tap - 3 * r - 1) * dil;
          }
          for (int c64 = 0; c64 < kchunks64; c64 += 64 * KPS, kb64 += 64 * KPS) {
            uint8_t* sb = sa + GS * KPS * L::kABytes;
            mbar_wait(ebar, phase ^ 1);
            if (PAIR) {
              // both CTAs' tiles are counted on the leader's barrier (a completion that overtakes the leader's expect_tx
              // leaves the count transiently negative; the phase needs the leader's arrival too)
              if (leader) mbar_expect_tx(fbar, tx_bytes);
#pragma unroll
              for (int j = 0; j < KPS; ++j) {
                if (!skipA) tma_load_5d_pair(sa + j * L::kABytes, &tmA, fb, c64 + 64 * j, cx, cy, n_img, 0);
                if (!skipB) tma_load_3d_pair(sb + j * kBt, &tmB, fb, kb64 + 64 * j, nB, 0);
              }
            } else {
              mbar_expect_tx(fbar, tx_bytes);
#pragma unroll
              for (int pl = 0; pl < GS; ++pl) {
                if (!skipA || GS > 1) tma_load_5d(sa + pl * L::kABytes, &tmA, fbar, c64, cx, cy, n_img, pl);
                if (!skipB || GS > 1) tma_load_3d(sb + pl * L::kBBytes, &tmB, fbar, kb64, n0, pl);
              }
            }
            advance();
          }
        }
        if (kDfvX && p.dfv_kb) {
          // one extra k-block: A = the pooling-transpose matrix of this CTA's pixel tile, B = its half of the image's
          // pooled-map gradient rows
          mbar_wait(ebar, phase ^ 1);
          if (leader) mbar_expect_tx(fbar, 2 * kStageSz);
          tma_load_5d_pair(sa, &tmW, fb, 0, x0, y0, 0, 0);
          tma_load_3d_pair(sa + L::kABytes, &tmT, fb, 0, nB, n_img);
          advance();
        }
        if (kCtx2) {
          // two extra k-blocks: A = the interpolation matrix of this pixel tile, B = the per-image tables
#pragma unroll 1
          for (int e = 0; e < 2; ++e) {
            mbar_wait(ebar, phase ^ 1);
            mbar_expect_tx(fbar, L::kStageBytes);
            tma_load_5d(sa, &tmW, fbar, 0, x0, y0, 0, 0);
            tma_load_5d(sa + GS * L::kABytes, &tmT, fbar, 0, n0, n_img, e, 0);
            advance();
          }
        }
      }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (pair: the leader CTA's only) =====================
    if ((!PAIR || leader) && elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(PAIR ? 256 : 128, BN, 0, 0);
      auto mma = [&](uint32_t d, uint64_t da, uint64_t db, uint32_t acc) {
        if (PAIR) umma_bf16_pair(d, da, db, idesc, acc);
        else umma_bf16(d, da, db, idesc, acc);
      };
      auto commit = [&](uint64_t* bar) {
        if (PAIR) umma_commit_pair(bar);
        else umma_commit(bar);
      };
      int stage = 0;
      uint32_t phase = 0;
      constexpr int NACC = AccLayout<BN, GS>::kNacc;
      // non-HALO main loop: running descriptor / barrier addresses of the current stage
      const uint64_t da_base = umma_smem_desc(smem_u32(ring), 16, 1024);
      uint64_t da_run = da_base;
      uint64_t* fbar = full_bar;
      uint64_t* ebar = empty_bar;
      int nst;
      asm volatile("mov.u32 %0, %1;" : "=r"(nst) : "r"(stages));
      int sA_ = 0;          // PAIR == 2: A halo stage and its phase
      uint32_t phA_ = 0;
      int it = 0;  // tiles processed by this CTA
      for (int q_ = W0; q_ < WN; q_ += WS, ++it) {
        const int as = it % ACC_STAGES;
        const uint32_t aphase = static_cast<uint32_t>(it / ACC_STAGES) & 1u;
        mbar_wait(&tmem_empty_bar[as], aphase ^ 1);  // epilogue drained this accumulator stage (pair: in both CTAs)
        tc_fence_after();
        const uint32_t tacc = tmem_base + as * ACC_COLS;
        uint32_t used = 0;  // bit a: accumulator a already holds a partial sum
        if (HALO) {
          if (it == 0) mbar_wait(w_bar, 0);
          const uint64_t wdesc0 = umma_smem_desc(smem_u32(wbase), 16, 1024);
          const uint32_t dy_bytes = static_cast<uint32_t>(p.dil * p.TW * 128);
          const uint32_t tap_row_units = static_cast<uint32_t>((3 * p.kchunks * kBt) >> 4);  // dy -> dy+1
          uint32_t accum = 0;
          if (halo1) {
            // One (TH + 2 dil) x (TW + 2 dil) halo tile per channel chunk, TW = 8: the 8-row groups of A are the patch
            // rows, one halo pitch apart (SBO = pitch), and tap (dy, dx) starts (dy * pitch + dx) * dil pixels into
            // the tile.  Neither is a multiple of the 1024-byte swizzle pattern: the MMA applies the 128B swizzle to
            // the absolute shared-memory address, as the TMA write did (measured, tools/probe_umma_halo.cu), so any
            // 128-byte-aligned start reads the right rows.
            const uint32_t pitch = static_cast<uint32_t>(p.TW + 2 * p.dil);
            for (int chunk = 0; chunk < p.kchunks; ++chunk) {
              mbar_wait(&full_bar[stage], phase);
              tc_fence_after();
              const uint64_t da0 = umma_smem_desc(smem_u32(ring + stage * stage_bytes), 16, pitch * 128);
              const uint64_t db0 = wdesc0 + static_cast<uint32_t>((chunk * kBt) >> 4);
              const uint32_t tap_units = static_cast<uint32_t>((p.kchunks * kBt) >> 4);   // tap -> tap + 1
              const uint32_t dx_units = static_cast<uint32_t>(p.dil) * 8, dy_units = dx_units * pitch;
#pragma unroll
              for (int dyi = 0; dyi < 3; ++dyi) {
#pragma unroll
                for (int dxi = 0; dxi < 3; ++dxi) {
                  const uint64_t da = da0 + (dyi * dy_units + dxi * dx_units);
                  const uint64_t db = db0 + (dyi * 3 + dxi) * tap_units;
#pragma unroll
                  for (int k = 0; k < 4; ++k) {
                    mma(tacc, da + 2 * k, db + 2 * k, accum);
                    accum = 1;
                  }
                }
              }
              commit(&empty_bar[stage]);
              if (++stage == stages) {
                stage = 0;
                phase ^= 1;
              }
            }
          } else
          for (int chunk = 0; chunk < p.kchunks; ++chunk) {
            for (int dxi = 0; dxi < 3; ++dxi) {
              mbar_wait(&full_bar[stage], phase);
              tc_fence_after();
              // one descriptor per stage; taps and k-steps only add to its 14-bit address field (>> 4).
              // The issuing thread is alone: every ALU op it spends per MMA is exposed latency.
              const uint64_t da0 = umma_smem_desc(smem_u32(ring + stage * stage_bytes), 16, 1024);
              const uint64_t db0 = wdesc0 + static_cast<uint32_t>(((dxi * p.kchunks + chunk) * kBt) >> 4);
#pragma unroll
              for (int dyi = 0; dyi < 3; ++dyi) {
                const uint64_t da = da0 + static_cast<uint32_t>((dyi * dy_bytes) >> 4);
                const uint64_t db = db0 + static_cast<uint32_t>(dyi * tap_row_units);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  mma(tacc, da + 2 * k, db + 2 * k, accum);
                  accum = 1;
                }
              }
              commit(&empty_bar[stage]);
              if (++stage == stages) {
                stage = 0;
                phase ^= 1;
              }
            }
          }
        } else if (HA) {
          // A = the chunk's halo tile: 8-pixel patch rows one halo pitch apart (SBO = pitch), tap (dy, dx) starts
          // (dy * pitch + dx) * dil pixels into it (see the single-tile HALO variant above for why that is legal)
          const uint32_t pitch = static_cast<uint32_t>(p.TW + 2 * p.dil);
          const uint32_t dx_units = static_cast<uint32_t>(p.dil) * 8, dy_units = dx_units * pitch;
          const uint64_t dbB0 = umma_smem_desc(smem_u32(ringB), 16, 1024);
          uint32_t accum = 0;
          for (int chunk = 0; chunk < p.kchunks; ++chunk) {
            mbar_wait(&fullA_bar[sA_], phA_);
            tc_fence_after();
            const uint64_t da0 = umma_smem_desc(smem_u32(ring + sA_ * p.halo_stride), 16, pitch * 128);
#pragma unroll
            for (int dyi = 0; dyi < 3; ++dyi) {
#pragma unroll
              for (int dx0 = 0; dx0 < 3; dx0 += TPS) {
                mbar_wait(&full_bar[stage], phase);
                tc_fence_after();
                const uint64_t db0 = dbB0 + static_cast<uint32_t>(stage * (kStageSz >> 4));
#pragma unroll
                for (int j = 0; j < TPS; ++j) {
                  const uint64_t da = da0 + (dyi * dy_units + (dx0 + j) * dx_units);
                  const uint64_t db = db0 + static_cast<uint32_t>(j * (kBt >> 4));
#pragma unroll
                  for (int k = 0; k < 4; ++k) {
                    if (!(kDecomp && (p.dbg & 4096))) mma(tacc, da + 2 * k, db + 2 * k, accum);
                    accum = 1;
                  }
                }
                commit(&empty_bar[stage]);
                if (++stage == nst) {
                  stage = 0;
                  phase ^= 1;
                }
              }
            }
            commit(&emptyA_bar[sA_]);   // every tap of this halo tile has been issued: free it when they retire
            if (++sA_ == p.n_a) {
              sA_ = 0;
              phA_ ^= 1;
            }
          }
        } else {
        // (the issuing thread is alone too: the stage's descriptors and barrier addresses are running values, see the
        // producer's note; a descriptor's address field (>> 4, 14 bits) covers all of shared memory, so adding to it never
        // carries into the other fields)
        const int n_st = num_kb / KPS;   // stages per tile
        for (int kb = 0; kb < n_st; ++kb) {
          mbar_wait(fbar, phase);
          tc_fence_after();
          const uint64_t da0 = da_run;                                     // one descriptor pair per
          const uint64_t db0 = da_run + ((GS * KPS * L::kABytes) >> 4);    // stage (see HALO note)
#pragma unroll
          for (int j = 0; j < KPS; ++j)
#pragma unroll
          for (int pa = 0; pa < GS; ++pa) {
#pragma unroll
            for (int pb = 0; pb < GS; ++pb) {
              if (pa + pb < GS) {
                // CTX2: the last k-block (Wint x a-table) goes to the second 256-column region (S)
                const int acc = kCtx2 ? (kb == base_kb + 1 ? 1 : 0)
                                : (GS == 1)         ? 0
                                                    : ((pa + pb == 0) ? (kb % NACC) : NACC);
                const uint32_t accf = (used >> acc) & 1u;
                used |= 1u << acc;
                if (kDecomp && (p.dbg & 4096)) continue;   // (CAN_DECOMP builds: the loads and hand-shakes without the MMAs)
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  mma(tacc + acc * BN, da0 + (((j + pa) * L::kABytes + k * 32) >> 4),
                      db0 + ((j * kBt + pb * L::kBBytes + k * 32) >> 4), (k | j) ? 1u : accf);
              }
            }
          }
          commit(ebar);   // frees the smem slot when these MMAs retire (pair: in both CTAs)
          da_run += kStageSz >> 4;
          ++fbar;
          ++ebar;
          if (++stage == nst) {
            stage = 0;
            phase ^= 1;
            da_run = da_base;
            fbar = full_bar;
            ebar = empty_bar;
          }
        }
        }
        commit(&tmem_full_bar[as]);  // accumulator of this tile complete (pair: of both tiles, told to both CTAs)
      }
    }
  } else if (warp == 2 + EW) {
    // ===================== output warp (staged epilogue only) =====================
    // hands each staged tile to TMA and waits until it has been read, so no epilogue warp ever stalls on that
    // (one elect.sync for the whole warp: a second one in an else-branch would wait for the elected lane forever)
    const bool out_lane = p.out_stage_bytes && elect_one();
    if (MODE == MODE_UNPOOL && out_lane) {
      // four rounds per tile (one per position of the 2x2 window), two staging buffers; tmO strides by 2 in x and y
      int R = 0;
      for (int q_ = W0; q_ < WN; q_ += WS) {
        const int tile = tile_of(q_);
        CAN_TILE_DECODE(tile)
        for (int r = 0; r < 4; ++r, ++R) {
          const int b = R & 1;
          mbar_wait(&out_full_bar[b], static_cast<uint32_t>(R >> 1) & 1u);
          if (live_of(q_)) {   // (a pair's dead tile goes through the hand-shakes but stores nothing)
#pragma unroll
          for (int blk = 0; blk < BN / 64; ++blk)
            tma_store_4d(&tmO, s_out + b * (BN * 256) + blk * 16384, n0 + blk * 64, 2 * x0 + (r & 1), 2 * y0 + (r >> 1),
                         n_img);
          }
          bulk_commit_group();
          bulk_wait_group_read1();                      // the previous round's store has read its buffer
          if (R > 0) mbar_arrive(&out_free_bar[b ^ 1]);
        }
      }
      bulk_wait_group_read0();
      if (R > 0) mbar_arrive(&out_free_bar[(R - 1) & 1]);
    } else if (MODE == MODE_CTX2_BWD && out_lane) {
      // four rounds per tile, eight segment stores per round (tmO: box (16 ch, TW, TH), no swizzle)
      int R = 0;
      for (int q_ = W0; q_ < WN; q_ += WS) {
        const int tile = tile_of(q_);
        CAN_TILE_DECODE(tile)
        (void)n0;
        for (int r = 0; r < 4; ++r, ++R) {
          const int b = R & 1;
          mbar_wait(&out_full_bar[b], static_cast<uint32_t>(R >> 1) & 1u);
#pragma unroll
          for (int sg = 0; sg < 8; ++sg)   // sg < 4: dz_j -> channel j*512 + ..; sg >= 4: ds_j -> 2048 + j*512 + ..
            tma_store_4d(&tmO, s_out + b * 32768 + sg * 4096, (sg >> 2) * 2048 + (sg & 3) * 512 + n_tile * 64 + 16 * r, x0, y0,
                         n_img);
          bulk_commit_group();
          bulk_wait_group_read1();
          if (R > 0) mbar_arrive(&out_free_bar[b ^ 1]);
        }
      }
      bulk_wait_group_read0();
      if (R > 0) mbar_arrive(&out_free_bar[(R - 1) & 1]);
    } else if (MODE == MODE_CTX2 && out_lane) {
      // one 64-channel tile of the fused features per GEMM tile
      const int nb = p.out_bufs;
      int it = 0;
      for (int q_ = W0; q_ < WN; q_ += WS, ++it) {
        const int tile = tile_of(q_);
        CAN_TILE_DECODE(tile)
        (void)n0;
        const int b = it % nb;
        mbar_wait(&out_full_bar[b], static_cast<uint32_t>(it / nb) & 1u);
        tma_store_4d(&tmO, s_out + b * 16384, n_tile * 64, x0, y0, n_img);
        bulk_commit_group();
        bulk_wait_group_read0();
        mbar_arrive(&out_free_bar[b]);
      }
    } else if (MODE == MODE_STORE && BN == 256 && out_lane) {
      // BN = 256: four 64-channel blocks per tile, two 16 KB staging tiles (round = block)
      int R = 0;
      for (int q_ = W0; q_ < WN; q_ += WS) {
        const int tile = tile_of(q_);
        CAN_TILE_DECODE(tile)
        for (int blk = 0; blk < BN / 64; ++blk, ++R) {
          const int nb = p.out_bufs, b = R % nb;
          mbar_wait(&out_full_bar[b], static_cast<uint32_t>(R / nb) & 1u);
          // (a pair's dead tile goes through the hand-shakes but stores nothing: its threads wrote nothing to the buffer)
          if (live_of(q_)) tma_store_4d(&tmO, s_out + b * 16384, n0 + blk * 64, x0, y0, n_img);
          bulk_commit_group();
          if (nb == 1) {
            bulk_wait_group_read0();
            mbar_arrive(&out_free_bar[0]);
          } else {
            bulk_wait_group_read1();
            if (R > 0) mbar_arrive(&out_free_bar[b ^ 1]);
          }
        }
      }
      if (p.out_bufs == 2) {
        bulk_wait_group_read0();
        if (R > 0) mbar_arrive(&out_free_bar[(R - 1) & 1]);
      }
    } else if (out_lane) {
      const int nb = p.out_bufs;   // 1 or 2
      int it = 0;
      for (int q_ = W0; q_ < WN; q_ += WS, ++it) {
        const int tile = tile_of(q_);
        CAN_TILE_DECODE(tile)
        const int b = it % nb;
        mbar_wait(&out_full_bar[b], static_cast<uint32_t>(it / nb) & 1u);
        if (live_of(q_)) {
#pragma unroll
        for (int blk = 0; blk < BN / 64; ++blk)
          tma_store_4d(&tmO, s_out + b * (BN * 256) + blk * 16384, n0 + blk * 64, x0, y0, n_img);
        }
        bulk_commit_group();
        if (nb == 1) {
          bulk_wait_group_read0();
          mbar_arrive(&out_free_bar[0]);
        } else {
          bulk_wait_group_read1();                         // the previous tile's store has read its buffer
          if (it > 0) mbar_arrive(&out_free_bar[b ^ 1]);
        }
      }
      if (nb == 2 && it > 0) {
        bulk_wait_group_read0();
        mbar_arrive(&out_free_bar[(it - 1) & 1]);
      }
    }
  } else if (warp == 3 + EW) {
    // ===================== mask warp (staged epilogue only) =====================
    // brings the ReLU-mask tile of each output tile into shared memory, as far ahead as the buffers allow
    if (p.mask_bufs && elect_one()) {
      int it = 0;
      for (int q_ = W0; q_ < WN; q_ += WS, ++it) {
        const int tile = tile_of(q_);
        CAN_TILE_DECODE(tile)
        const int b = it % p.mask_bufs;
        const uint32_t ph = static_cast<uint32_t>(it / p.mask_bufs) & 1u;
        mbar_wait(&mask_free_bar[b], ph ^ 1u);
        mbar_expect_tx(&mask_full_bar[b], BN * 256);
#pragma unroll
        for (int blk = 0; blk < BN / 64; ++blk)
          tma_load_4d(s_mask + b * (BN * 256) + blk * 16384, &tmM, &mask_full_bar[b], n0 + blk * 64, x0, y0, n_img);
      }
    }
  } else {
    // ===================== epilogue (warps 2 .. 2+EW-1) =====================
    const int q = warp & 3;  // TMEM lane quarter this warp may read
    const int etid = threadIdx.x - 64;
    const int m = q * 32 + lane;
    const int ty = m / p.TW, tx = m - ty * p.TW;
    const int nact = min(AccLayout<BN, GS>::kNacc, num_kb);  // hi*hi accumulators actually written
    // columns of the tile handled by this warp (EW == 8: the two warps of a lane quarter split them)
    constexpr int CW = BN / (EW / 4);
    const int cbeg = ((warp - 2) >> 2) * CW, cend = cbeg + CW;
    // "this CTA's epilogue is done with accumulator stage a": the MMA issuer waits for it.  Pair: the issuer lives in the
    // leader CTA, so the odd CTA's warps arrive on the leader's barrier (one arrival per warp).
    auto release_acc = [&](int a) {
      if (PAIR && !leader) {
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(mapa_shared(&tmem_empty_bar[a], 0));
      } else {
        mbar_arrive(&tmem_empty_bar[a]);
      }
    };
    if (p.db) {
      for (int i = etid; i < 1024; i += 32 * EW) s_db[i] = 0.f;
      asm volatile("bar.sync 1, %0;" ::"n"(32 * EW) : "memory");
    }
    // bias gradient of narrow single-n-tile layers: every thread keeps the column sums of its own pixels in
    // registers over all tiles; the cross-lane reduction happens once per kernel instead of once per tile
    constexpr bool kDbRegMode = (CW == 32) && (MODE == MODE_STORE || MODE == MODE_UNPOOL);
    const bool db_reg = kDbRegMode && p.db && p.n_tiles == 1;
    float dbacc[kDbRegMode ? 32 : 1];
#pragma unroll
    for (int i = 0; i < (kDbRegMode ? 32 : 1); ++i) dbacc[i] = 0.f;
    int it = 0;
    for (int q_ = W0; q_ < WN; q_ += WS, ++it) {
        const int tile = tile_of(q_);
    CAN_TILE_DECODE(tile)
    const int as = it % ACC_STAGES;
    const uint32_t aphase = static_cast<uint32_t>(it / ACC_STAGES) & 1u;
    // single n-tile: the bias slice never changes -> staged once (tile 0), no global load + barrier per tile
    const bool bias_once = (p.n_tiles == 1) && MODE != MODE_CTX && MODE != MODE_CTX_BWD && MODE != MODE_DFV;
    float* s_bias = s_bias2 + (bias_once ? 0 : (it & 1) * BN);
    if (!bias_once || it == 0)
      for (int i = etid; i < BN; i += 32 * EW) s_bias[i] = p.bias ? p.bias[n0 + i] : 0.f;
    if (MODE == MODE_CTX || MODE == MODE_CTX_BWD) {
      asm volatile("bar.sync 1, %0;" ::"n"(32 * EW) : "memory");  // everyone finished reading the previous tile's tables
      // tables of this image for the 64 channels of this n-tile: [2][50][64]
      const float* ga = p.tabA + (static_cast<long long>(n_img) * 50) * 512 + n_tile * 64;
      const float* gb = p.tabB + (static_cast<long long>(n_img) * 50) * 512 + n_tile * 64;
      for (int i = etid; i < 50 * 16; i += 32 * EW) {
        const int pos = i >> 4, c4 = (i & 15) * 4;
        *reinterpret_cast<float4*>(s_tab + pos * 64 + c4) = *reinterpret_cast<const float4*>(ga + pos * 512 + c4);
        *reinterpret_cast<float4*>(s_tab + 3200 + pos * 64 + c4) =
            *reinterpret_cast<const float4*>(gb + pos * 512 + c4);
      }
    }
    if (MODE == MODE_DFV && !(kDfvX && p.dfv_kb)) {
      // dpool table of this image for the BN channels of this n-tile -> shared memory ([50][BN] fp32).  The
      // epilogue reads 13 of its rows per pixel and channel group; from global memory those reads miss the (small)
      // L1 next to a 200 KB shared-memory carve-out and pay the L2 latency 104 times per 32 channels.
      asm volatile("bar.sync 1, %0;" ::"n"(32 * EW) : "memory");  // everyone finished reading the previous tile's table
      const float* gt = p.tabA + (static_cast<long long>(n_img) * 50) * 512 + n0;
      constexpr int kVec = 50 * BN / 4, kPer = (kVec + 32 * EW - 1) / (32 * EW), kHalf = (kPer + 1) / 2;
#pragma unroll
      for (int h2 = 0; h2 < 2; ++h2) {   // loads first, then stores (two batches bound the registers)
        float4 tv[kHalf];
#pragma unroll
        for (int u = 0; u < kHalf; ++u) {
          const int i = etid + (h2 * kHalf + u) * 32 * EW;
          if (h2 * kHalf + u < kPer && i < kVec) tv[u] = __ldg(reinterpret_cast<const float4*>(gt + (i / (BN / 4)) * 512 + (i % (BN / 4)) * 4));
        }
#pragma unroll
        for (int u = 0; u < kHalf; ++u) {
          const int i = etid + (h2 * kHalf + u) * 32 * EW;
          if (h2 * kHalf + u < kPer && i < kVec) *reinterpret_cast<float4*>(s_tab + (i / (BN / 4)) * BN + (i % (BN / 4)) * 4) = tv[u];
        }
      }
    }
    if (!bias_once || it == 0) asm volatile("bar.sync 1, %0;" ::"n"(32 * EW) : "memory");

    const int y = y0 + ty, x = x0 + tx;
    const bool valid = live_of(q_) && (y < p.H) && (x < p.W);
    const uint32_t taddr = tmem_base + as * ACC_COLS + (static_cast<uint32_t>(q * 32) << 16);

    // Operands of the epilogue that do not depend on the accumulator (ReLU mask / arg-max codes /
    // upstream gradient of the first column group) are requested before waiting for the MMAs, and
    // inside the column loops one group ahead, so their latency is off the critical path.
    uint4 pre_m[4] = {};
    uint4 pre_a[2] = {};
    const __nv_bfloat16* pre_mbase = nullptr;
    const uint8_t* pre_abase = nullptr;
    if (MODE == MODE_STORE || MODE == MODE_UNPOOL || MODE == MODE_CTX2_BWD) {
      const long long pix_ = valid ? (static_cast<long long>(n_img) * p.H + y) * p.W + x : 0;
      if (p.mask && valid && !p.mask_bufs)
        pre_mbase = p.mask + pix_ * p.mask_cs + p.mask_coff +
                    (MODE == MODE_CTX2_BWD ? n_tile * 64 + ((warp - 2) >> 2) * (p.out_stage_bytes ? 8 : 64 / (EW / 4))
                                           : n0 + cbeg);
      if (MODE == MODE_UNPOOL && valid) pre_abase = p.arg + pix_ * p.arg_cs + n0 + cbeg;
      if (pre_mbase) {
        pre_m[0] = *reinterpret_cast<const uint4*>(pre_mbase);
        if (MODE != MODE_CTX2_BWD) {
#pragma unroll
          for (int i = 1; i < 4; ++i) pre_m[i] = *reinterpret_cast<const uint4*>(pre_mbase + 8 * i);
        }
      }
      if (pre_abase) {
        pre_a[0] = *reinterpret_cast<const uint4*>(pre_abase);
        pre_a[1] = *reinterpret_cast<const uint4*>(pre_abase + 16);
      }
    }

    mbar_wait(&tmem_full_bar[as], aphase);
    tc_fence_after();

    const bool kStaged = ((MODE == MODE_STORE && BN <= 128) || MODE == MODE_CTX2) && NSPLIT == 1 && p.out_stage_bytes > 0;
    const bool kBlockStaged = (MODE == MODE_STORE) && NSPLIT == 1 && BN == 256 && EW == 4 && p.out_stage_bytes > 0;
    const int ob_i = kStaged ? it % p.out_bufs : 0;     // staging buffer of this tile
    bool out_buf_ready = false;                         // (waited for as late as possible: just before the first write)
    const bool kMaskStaged = kStaged && p.mask_bufs > 0;
    const int mb_i = kMaskStaged ? it % p.mask_bufs : 0;
    if (kMaskStaged) mbar_wait(&mask_full_bar[mb_i], static_cast<uint32_t>(it / p.mask_bufs) & 1u);

    if (p.dbg & 16) {
      // bring-up experiment: no epilogue work at all
    } else if (MODE == MODE_STORE) {
      const long long pix = (static_cast<long long>(n_img) * p.H + y) * p.W + x;
      __nv_bfloat16* obase = p.out + pix * p.out_cs + p.out_coff + n0;
      const __nv_bfloat16* mbase = p.mask ? p.mask + pix * p.mask_cs + p.mask_coff + n0 : nullptr;
      float* fbase = p.out_f32 ? p.out_f32 + pix * p.out_cs + p.out_coff + n0 : nullptr;
#pragma unroll 1
      for (int c0 = cbeg; c0 < cend; c0 += 32) {
        float r[32];
        load_acc<BN, GS, 32>(taddr, c0, nact, r);
        if ((p.dbg & 64) || ((p.dbg & 128) && (warp & 3) == 1)) continue;   // bring-up experiments: accumulator reads only (all warps / the warp sharing the MMA issuer's scheduler)
        float v[32];
        load_bias32(s_bias + c0, v);
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          v[i] += r[i];
          if (p.relu) v[i] = fmaxf(v[i], 0.f);
        }
        // Staging-buffer waits are taken by EVERY thread (also those of out-of-image rows, which write nothing): a
        // thread that skipped them could run several tiles ahead and arrive twice in one phase of out_full.
        if (kStaged && !out_buf_ready) {   // the store that last used this buffer has read it
          mbar_wait(&out_free_bar[ob_i], (static_cast<uint32_t>(it / p.out_bufs) & 1u) ^ 1u);
          out_buf_ready = true;
        }
        if (kBlockStaged && (c0 & 63) == 0) {
          const int R = it * (BN / 64) + (c0 >> 6);
          mbar_wait(&out_free_bar[R % p.out_bufs], (static_cast<uint32_t>(R / p.out_bufs) & 1u) ^ 1u);
        }
        if (valid) {
          if (mbase) {
            uint4 mcur[4];
            if (kMaskStaged) {
              const uint8_t* mrow = s_mask + mb_i * (BN * 256) + (c0 >> 6) * 16384 + m * 128;
#pragma unroll
              for (int i = 0; i < 4; ++i)
                mcur[i] = *reinterpret_cast<const uint4*>(mrow + (((((c0 & 63) >> 3) + i) ^ (m & 7)) << 4));
            } else {
#pragma unroll
              for (int i = 0; i < 4; ++i) mcur[i] = pre_m[i];
              if (c0 + 32 < cend) {
#pragma unroll
                for (int i = 0; i < 4; ++i) pre_m[i] = *reinterpret_cast<const uint4*>(mbase + c0 + 32 + 8 * i);
              }
            }
#pragma unroll
            for (int i = 0; i < 32; i += 8) {
              const __nv_bfloat16* mb = reinterpret_cast<const __nv_bfloat16*>(&mcur[i >> 3]);
#pragma unroll
              for (int e = 0; e < 8; ++e)
                if (!(__bfloat162float(mb[e]) > 0.f)) v[i + e] = 0.f;
            }
          }
          if (kBlockStaged) {
            // BN = 256: 64-channel blocks go through two 16 KB staging tiles as they are produced (round = block)
            const int b = (it * (BN / 64) + (c0 >> 6)) % p.out_bufs;
            uint8_t* row = s_out + b * 16384 + m * 128;
#pragma unroll
            for (int i = 0; i < 32; i += 8) {
              const int chunk = ((c0 & 63) + i) >> 3;
              *reinterpret_cast<uint4*>(row + ((chunk ^ (m & 7)) << 4)) =
                  make_uint4(pack_bf16x2(v[i], v[i + 1]), pack_bf16x2(v[i + 2], v[i + 3]), pack_bf16x2(v[i + 4], v[i + 5]),
                             pack_bf16x2(v[i + 6], v[i + 7]));
            }
          } else if (kStaged) {
            // 16-byte chunks into the 128B-swizzled [64-ch block][pixel][64 ch] staging tile
            uint8_t* row = s_out + ob_i * (BN * 256) + (c0 >> 6) * 16384 + m * 128;
#pragma unroll
            for (int i = 0; i < 32; i += 8) {
              uint4 q;
              q.x = pack_bf16x2(v[i], v[i + 1]);
              q.y = pack_bf16x2(v[i + 2], v[i + 3]);
              q.z = pack_bf16x2(v[i + 4], v[i + 5]);
              q.w = pack_bf16x2(v[i + 6], v[i + 7]);
              const int chunk = ((c0 & 63) + i) >> 3;
              *reinterpret_cast<uint4*>(row + ((chunk ^ (m & 7)) << 4)) = q;
            }
          } else if (p.out && !(p.dbg & 256)) {   // (bit 256: bring-up experiment, all the arithmetic but no store)
            store_bf16_planes<NSPLIT, 32>(obase + c0, p.out_plane, v);
          }
          if (fbase) {
#pragma unroll
            for (int i = 0; i < 32; i += 4)
              *reinterpret_cast<float4*>(fbase + c0 + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
          }
        }
        if (kBlockStaged && (c0 & 63) == 32) {   // this 64-channel block is complete: hand it to the output warp
          fence_proxy_async();
          mbar_arrive(&out_full_bar[(it * (BN / 64) + (c0 >> 6)) % p.out_bufs]);
        }
        if (kDbRegMode && db_reg) {
#pragma unroll
          for (int i = 0; i < 32; ++i) dbacc[kDbRegMode ? i : 0] += valid ? v[i] : 0.f;
        } else if (p.db) {
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = valid ? v[i] : 0.f;
          atomicAdd(&s_db[n0 + c0 + lane], warp_column_sums(v, lane));
        }
      }
    } else if (MODE == MODE_POOL) {
      // 2x2 window = lanes {l, l^1, l^TW, l^TW^1}; the lane at the even/even corner writes.
      const int PH = p.H >> 1, PW = p.W >> 1;
      const int py = y >> 1, px = x >> 1;
      const bool writer = ((tx & 1) == 0) && ((ty & 1) == 0) && (py < PH) && (px < PW) && live_of(q_);
      const int pos = ((ty & 1) << 1) | (tx & 1);  // scan order inside the window
      const long long ppix = (static_cast<long long>(n_img) * PH + py) * PW + px;
      __nv_bfloat16* obase = p.out + ppix * p.out_cs + p.out_coff + n0;
      uint8_t* abase = p.arg ? p.arg + ppix * p.arg_cs + n0 : nullptr;
#pragma unroll 1
      for (int c0 = cbeg; c0 < cend; c0 += 32) {
        float r[32];
        load_acc<BN, GS, 32>(taddr, c0, nact, r);
        float v[32];
        load_bias32(s_bias + c0, v);
        uint32_t codes[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) codes[i] = 0;
        if constexpr (NSPLIT == 1) {
          // bf16 path: warp shuffles go through the shared-memory port these layers are bound by, so the window
          // maximum is taken on channel PAIRS (the values are compared at their stored precision anyway) and the
          // argmax comes from one 32-bit "I hold the maximum" word per lane: 35 shuffles per 32 channels, not 128
          // (measured against the per-channel exchange below: 64->64 271 -> 245 us, 128->128 277 -> 265 us, 256->256 171 -> 167 us).
          uint32_t pk[16], eq = 0;
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            // (sign bits cleared: a -0 from the ReLU would compare unequal to +0 bitwise below)
            const uint32_t mine = pack_bf16x2(fmaxf(r[2 * i] + v[2 * i], 0.f), fmaxf(r[2 * i + 1] + v[2 * i + 1], 0.f)) & 0x7fff7fffu;
            uint32_t m = bf16x2_max(mine, __shfl_xor_sync(0xffffffffu, mine, 1));
            m = bf16x2_max(m, __shfl_xor_sync(0xffffffffu, m, p.TW));
            const uint32_t x = mine ^ m;   // the maximum is one of the four bit patterns: equal value <=> equal bits
            eq |= ((x & 0xffffu) == 0u ? 1u : 0u) << (2 * i);
            eq |= ((x >> 16) == 0u ? 1u : 0u) << (2 * i + 1);
            pk[i] = m;
          }
          // eq words of the window in scan order (this lane is position `pos`; the writer is position 0)
          const uint32_t e1 = __shfl_xor_sync(0xffffffffu, eq, 1), e2 = __shfl_xor_sync(0xffffffffu, eq, p.TW),
                         e3 = __shfl_xor_sync(0xffffffffu, eq, p.TW | 1);
          if (writer) {
            // first position holding the maximum, for 32 channels at once: bit planes of the 2-bit code ...
            const uint32_t c1 = e1 & ~eq, c2 = e2 & ~(eq | e1), c3 = e3 & ~(eq | e1 | e2);
            const uint32_t lo = c1 | c3, hi = c2 | c3;
#pragma unroll
            for (int j = 0; j < 8; ++j) {   // ... spread to one byte per channel (bit k of a nibble -> bit 8 k)
              const uint32_t nl = (lo >> (4 * j)) & 0xFu, nh = (hi >> (4 * j)) & 0xFu;
              codes[j] = ((nl * 0x00204081u) & 0x01010101u) | (((nh * 0x00204081u) & 0x01010101u) << 1);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i)
              *reinterpret_cast<uint4*>(obase + c0 + 8 * i) = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
            if (abase) {
              *reinterpret_cast<uint4*>(abase + c0) = make_uint4(codes[0], codes[1], codes[2], codes[3]);
              *reinterpret_cast<uint4*>(abase + c0 + 16) = make_uint4(codes[4], codes[5], codes[6], codes[7]);
            }
          }
        } else {   // fp32 path: per-channel exchange on the full-precision values
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          // floor-mode pooling only writes windows that lie fully inside the image, so the
          // garbage computed for out-of-image tile pixels never reaches a written window.
          float a = fmaxf(r[i] + v[i], 0.f);
          float mx = fmaxf(a, __shfl_xor_sync(0xffffffffu, a, 1));
          mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, p.TW));
          int code = (a == mx) ? pos : 4;
          code = min(code, __shfl_xor_sync(0xffffffffu, code, 1));
          code = min(code, __shfl_xor_sync(0xffffffffu, code, p.TW));
          v[i] = mx;
          codes[i >> 2] |= static_cast<uint32_t>(code) << ((i & 3) * 8);
        }
        if (writer) {
          store_bf16_planes<NSPLIT, 32>(obase + c0, p.out_plane, v);
          if (abase) {
            *reinterpret_cast<uint4*>(abase + c0) = make_uint4(codes[0], codes[1], codes[2], codes[3]);
            *reinterpret_cast<uint4*>(abase + c0 + 16) = make_uint4(codes[4], codes[5], codes[6], codes[7]);
          }
        }
        }
      }
    } else if (MODE == MODE_HEAD) {
      const long long pix = (static_cast<long long>(n_img) * p.H + y) * p.W + x;
      __nv_bfloat16* obase = p.out ? p.out + pix * p.out_cs + p.out_coff + n0 : nullptr;
      float dot = 0.f;
#pragma unroll 1
      for (int c0 = cbeg; c0 < cend; c0 += 32) {
        float r[32];
        load_acc<BN, GS, 32>(taddr, c0, nact, r);
        float v[32];
        load_bias32(s_bias + c0, v);
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          v[i] = fmaxf(r[i] + v[i], 0.f);
          dot = fmaf(v[i], __ldg(p.head_w + c0 + i), dot);
        }
        if (valid && obase) store_bf16_planes<NSPLIT, 32>(obase + c0, p.out_plane, v);
      }
      if (valid) p.head_out[pix] = dot + __ldg(p.head_b);
    } else if (MODE == MODE_UNPOOL && NSPLIT == 1 && CW <= 64 && p.out_stage_bytes > 0) {
      // staged variant: gradients of this thread's CW channels stay in registers; one round per window position
      // writes a (TH x TW) sub-tile of the 2x upsampled map to shared memory, the output warp stores it with TMA
      const long long pix = valid ? (static_cast<long long>(n_img) * p.H + y) * p.W + x : 0;
      const __nv_bfloat16* mbase = p.mask + pix * p.mask_cs + p.mask_coff + n0;
      const uint8_t* abase = p.arg + pix * p.arg_cs + n0;
      constexpr int NI = CW / 32;
      float g[NI][32];
      uint32_t code[NI][8];
#pragma unroll
      for (int ci = 0; ci < NI; ++ci) {
        const int c0 = cbeg + ci * 32;
        float r[32];
        load_acc<BN, GS, 32>(taddr, c0, nact, r);
        uint4 mcur[4], a0 = make_uint4(0, 0, 0, 0), a1 = a0;
#pragma unroll
        for (int i = 0; i < 4; ++i) mcur[i] = make_uint4(0, 0, 0, 0);
        if (valid) {
          if (ci == 0) {
            a0 = pre_a[0];
            a1 = pre_a[1];
#pragma unroll
            for (int i = 0; i < 4; ++i) mcur[i] = pre_m[i];
          } else {
            a0 = *reinterpret_cast<const uint4*>(abase + c0);
            a1 = *reinterpret_cast<const uint4*>(abase + c0 + 16);
#pragma unroll
            for (int i = 0; i < 4; ++i) mcur[i] = *reinterpret_cast<const uint4*>(mbase + c0 + 8 * i);
          }
        }
        code[ci][0] = a0.x; code[ci][1] = a0.y; code[ci][2] = a0.z; code[ci][3] = a0.w;
        code[ci][4] = a1.x; code[ci][5] = a1.y; code[ci][6] = a1.z; code[ci][7] = a1.w;
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          const __nv_bfloat16* mb = reinterpret_cast<const __nv_bfloat16*>(&mcur[i >> 3]);
#pragma unroll
          for (int e = 0; e < 8; ++e) g[ci][i + e] = (__bfloat162float(mb[e]) > 0.f) ? r[i + e] : 0.f;   // mask 0 when !valid
        }
        if (kDbRegMode && db_reg) {
#pragma unroll
          for (int i = 0; i < 32; ++i) dbacc[kDbRegMode ? i : 0] += g[ci][i];
        } else if (p.db) {
          float t[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) t[i] = g[ci][i];
          atomicAdd(&s_db[n0 + c0 + lane], warp_column_sums(t, lane));
        }
      }
      tc_fence_before();
      release_acc(as);   // the accumulator is in registers now
#pragma unroll 1
      for (int wpos = 0; wpos < 4; ++wpos) {
        const int R = it * 4 + wpos, b = R & 1;
        mbar_wait(&out_free_bar[b], (static_cast<uint32_t>(R >> 1) & 1u) ^ 1u);
#pragma unroll
        for (int ci = 0; ci < NI; ++ci) {
          const int c0 = cbeg + ci * 32;
          uint8_t* row = s_out + b * (BN * 256) + (c0 >> 6) * 16384 + m * 128;
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const int cd = (code[ci][(i + e) >> 2] >> (((i + e) & 3) * 8)) & 0xff;
              v[e] = (cd == wpos) ? g[ci][i + e] : 0.f;
            }
            const int chunk = ((c0 & 63) + i) >> 3;
            *reinterpret_cast<uint4*>(row + ((chunk ^ (m & 7)) << 4)) =
                make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
          }
        }
        fence_proxy_async();
        mbar_arrive(&out_full_bar[b]);
      }
      continue;   // (accumulator already released above)
    } else if (MODE == MODE_UNPOOL) {
      // this GEMM runs on the pooled grid (H x W); gradients land on the OH x OW map
      const long long pix = (static_cast<long long>(n_img) * p.H + y) * p.W + x;
      const __nv_bfloat16* mbase = p.mask + pix * p.mask_cs + p.mask_coff + n0;
      const uint8_t* abase = p.arg + pix * p.arg_cs + n0;
#pragma unroll 1
      for (int c0 = cbeg; c0 < cend; c0 += 32) {
        float r[32];
        load_acc<BN, GS, 32>(taddr, c0, nact, r);
        float g[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) g[i] = 0.f;
        if (valid) {
          uint32_t code[8];
          uint4 mcur[4];
          {
            const uint4 a0 = pre_a[0], a1 = pre_a[1];
            code[0] = a0.x; code[1] = a0.y; code[2] = a0.z; code[3] = a0.w;
            code[4] = a1.x; code[5] = a1.y; code[6] = a1.z; code[7] = a1.w;
#pragma unroll
            for (int i = 0; i < 4; ++i) mcur[i] = pre_m[i];
            if (c0 + 32 < cend) {
              pre_a[0] = *reinterpret_cast<const uint4*>(abase + c0 + 32);
              pre_a[1] = *reinterpret_cast<const uint4*>(abase + c0 + 48);
#pragma unroll
              for (int i = 0; i < 4; ++i) pre_m[i] = *reinterpret_cast<const uint4*>(mbase + c0 + 32 + 8 * i);
            }
          }
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            const __nv_bfloat16* mb = reinterpret_cast<const __nv_bfloat16*>(&mcur[i >> 3]);
#pragma unroll
            for (int e = 0; e < 8; ++e)
              g[i + e] = (__bfloat162float(mb[e]) > 0.f) ? r[i + e] : 0.f;
          }
#pragma unroll
          for (int wpos = 0; wpos < 4; ++wpos) {
            const int oy = 2 * y + (wpos >> 1), ox = 2 * x + (wpos & 1);
            float v[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const int cd = (code[i >> 2] >> ((i & 3) * 8)) & 0xff;
              v[i] = (cd == wpos) ? g[i] : 0.f;
            }
            const long long opix = (static_cast<long long>(n_img) * p.OH + oy) * p.OW + ox;
            store_bf16_planes<NSPLIT, 32>(p.out + opix * p.out_cs + p.out_coff + n0 + c0, p.out_plane, v);
          }
        }
        if (kDbRegMode && db_reg) {
#pragma unroll
          for (int i = 0; i < 32; ++i) dbacc[kDbRegMode ? i : 0] += g[i];
        } else if (p.db) {
          atomicAdd(&s_db[n0 + c0 + lane], warp_column_sums(g, lane));
        }
      }
    } else if (MODE == MODE_CTX) {
      // BN == 256 columns = 4 scales x 64 channels of n-tile blockIdx.y; column j*64+c holds
      // (W_j2 fv)[c].  z_j = U(b_j) - W_j2 fv ; w_j = sigmoid(z_j) ; s_j = U(a_j)
      // fi = sum_j w_j s_j / (sum_j w_j + 1e-12)            (CANNet.py:47-84)
      const int SS[4] = {1, 2, 3, 6};
      const int OFF[4] = {0, 1, 5, 14};
      int pidx[4][4];
      float ly0[4], ly1[4], lx0[4], lx1[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        int yi0, yi1, xi0, xi1;
        bilinear_tap(min(y, p.H - 1), SS[j], p.H, yi0, yi1, ly0[j], ly1[j]);
        bilinear_tap(min(x, p.W - 1), SS[j], p.W, xi0, xi1, lx0[j], lx1[j]);
        pidx[j][0] = (OFF[j] + yi0 * SS[j] + xi0) * 64;
        pidx[j][1] = (OFF[j] + yi0 * SS[j] + xi1) * 64;
        pidx[j][2] = (OFF[j] + yi1 * SS[j] + xi0) * 64;
        pidx[j][3] = (OFF[j] + yi1 * SS[j] + xi1) * 64;
      }
      const long long pix = (static_cast<long long>(n_img) * p.H + y) * p.W + x;
      __nv_bfloat16* obase = p.out + pix * p.out_cs + p.out_coff + n_tile * 64;
      const float* sA = s_tab;
      const float* sB = s_tab + 3200;
#pragma unroll 1
      for (int c0 = 0; c0 < 64; c0 += 16) {
        float r[4][16];
#pragma unroll
        for (int j = 0; j < 4; ++j) load_acc<BN, GS, 16>(taddr, j * 64 + c0, nact, r[j]);
        float fi[16];
#pragma unroll
        for (int c = 0; c < 16; c += 4) {
          float num[4] = {0.f, 0.f, 0.f, 0.f}, den[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 a00 = *reinterpret_cast<const float4*>(sA + pidx[j][0] + c0 + c);
            const float4 a01 = *reinterpret_cast<const float4*>(sA + pidx[j][1] + c0 + c);
            const float4 a10 = *reinterpret_cast<const float4*>(sA + pidx[j][2] + c0 + c);
            const float4 a11 = *reinterpret_cast<const float4*>(sA + pidx[j][3] + c0 + c);
            const float4 b00 = *reinterpret_cast<const float4*>(sB + pidx[j][0] + c0 + c);
            const float4 b01 = *reinterpret_cast<const float4*>(sB + pidx[j][1] + c0 + c);
            const float4 b10 = *reinterpret_cast<const float4*>(sB + pidx[j][2] + c0 + c);
            const float4 b11 = *reinterpret_cast<const float4*>(sB + pidx[j][3] + c0 + c);
            const float av00[4] = {a00.x, a00.y, a00.z, a00.w}, av01[4] = {a01.x, a01.y, a01.z, a01.w};
            const float av10[4] = {a10.x, a10.y, a10.z, a10.w}, av11[4] = {a11.x, a11.y, a11.z, a11.w};
            const float bv00[4] = {b00.x, b00.y, b00.z, b00.w}, bv01[4] = {b01.x, b01.y, b01.z, b01.w};
            const float bv10[4] = {b10.x, b10.y, b10.z, b10.w}, bv11[4] = {b11.x, b11.y, b11.z, b11.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              // ATen order: h0*(w0*v00 + w1*v01) + h1*(w0*v10 + w1*v11)
              const float s = ly0[j] * (lx0[j] * av00[e] + lx1[j] * av01[e]) +
                              ly1[j] * (lx0[j] * av10[e] + lx1[j] * av11[e]);
              const float u = ly0[j] * (lx0[j] * bv00[e] + lx1[j] * bv01[e]) +
                              ly1[j] * (lx0[j] * bv10[e] + lx1[j] * bv11[e]);
              const float z = u - r[j][c + e];
              const float wgt = 1.f / (1.f + __expf(-z));
              num[e] = fmaf(wgt, s, num[e]);
              den[e] += wgt;
            }
          }
#pragma unroll
          for (int e = 0; e < 4; ++e) fi[c + e] = num[e] / (den[e] + 1e-12f);
        }
        if (valid) store_bf16_planes<NSPLIT, 16>(obase + c0, p.out_plane, fi);
      }
    } else if (MODE == MODE_CTX2) {
      // TMEM columns [0,256): -z_j = W_j2 fv - U(W_j2 a_j)   (8 k-blocks of fv + 1 k-block Wint x (-b tables))
      //              [256,512): s_j = U a_j                   (1 k-block Wint x a tables)
      // column j*64 + c for the 64 channels of n-tile blockIdx.y.  w_j = sigmoid(z_j) = 1/(1+exp(-z_j)).
      const long long pix = valid ? (static_cast<long long>(n_img) * p.H + y) * p.W + x : 0;
      __nv_bfloat16* obase = p.out + pix * p.out_cs + p.out_coff + n_tile * 64;
      constexpr int CH = 64 / (EW / 4);              // channels of the n-tile handled by this warp
      const int hb = ((warp - 2) >> 2) * CH;
#pragma unroll 1
      for (int c0 = hb; c0 < hb + CH; c0 += 16) {
        uint32_t zr[4][16], sr[4][16];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          tmem_ld_32x16(taddr + j * 64 + c0, zr[j]);
          tmem_ld_32x16(taddr + 256 + j * 64 + c0, sr[j]);
        }
        tmem_ld_wait();
        float fi[16];
#pragma unroll
        for (int c = 0; c < 16; ++c) {
          float num = 0.f, den = 0.f;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float wgt = sigmoid_neg_fast(__uint_as_float(zr[j][c]));
            num = fmaf(wgt, __uint_as_float(sr[j][c]), num);
            den += wgt;
          }
          fi[c] = num * rcp_fast(den + 1e-12f);
        }
        if (kStaged) {
          // two 16-byte chunks into the 128B-swizzled [pixel][64 ch] staging tile; the waits are taken by every thread
          if (!out_buf_ready) {
            mbar_wait(&out_free_bar[ob_i], (static_cast<uint32_t>(it / p.out_bufs) & 1u) ^ 1u);
            out_buf_ready = true;
          }
          if (valid) {
            uint8_t* row = s_out + ob_i * 16384 + m * 128;
#pragma unroll
            for (int i = 0; i < 16; i += 8)
              *reinterpret_cast<uint4*>(row + ((((c0 + i) >> 3) ^ (m & 7)) << 4)) =
                  make_uint4(pack_bf16x2(fi[i], fi[i + 1]), pack_bf16x2(fi[i + 2], fi[i + 3]), pack_bf16x2(fi[i + 4], fi[i + 5]),
                             pack_bf16x2(fi[i + 6], fi[i + 7]));
          }
        } else if (valid) {
          store_bf16_planes<NSPLIT, 16>(obase + c0, p.out_plane, fi);
        }
      }
    } else if (MODE == MODE_CTX2_BWD && NSPLIT == 1 && EW == 8 && p.out_stage_bytes > 0) {
      // Staged variant of the branch below: four rounds per tile, round s covers channels 16 s .. 16 s + 15 of
      // the n-tile (this warp: 8 of them), all eight output segments (dz_j, ds_j) of those channels go to a
      // [segment][pixel][16 ch] staging buffer and leave through eight TMA stores issued by the output warp.
      const long long pix = valid ? (static_cast<long long>(n_img) * p.H + y) * p.W + x : 0;
      const __nv_bfloat16* gbase = p.mask + pix * p.mask_cs + p.mask_coff + n_tile * 64;
      const int half = (warp - 2) >> 2;
#pragma unroll 1
      for (int s4 = 0; s4 < 4; ++s4) {
        const int c0 = 16 * s4 + 8 * half;
        uint32_t zr[4][8], sr[4][8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          tmem_ld_32x8(taddr + j * 64 + c0, zr[j]);
          tmem_ld_32x8(taddr + 256 + j * 64 + c0, sr[j]);
        }
        tmem_ld_wait();
        float g[8];
        {
          const uint4 gq = pre_m[0];
          if (valid && s4 < 3) pre_m[0] = *reinterpret_cast<const uint4*>(gbase + c0 + 16);
          const __nv_bfloat16* gb = reinterpret_cast<const __nv_bfloat16*>(&gq);
#pragma unroll
          for (int e = 0; e < 8; ++e) g[e] = valid ? __bfloat162float(gb[e]) : 0.f;
        }
        float dz[4][8], ds[4][8];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          float wj[4], num = 0.f, den = 0.f;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            wj[j] = sigmoid_neg_fast(__uint_as_float(zr[j][c]));
            num = fmaf(wj[j], __uint_as_float(sr[j][c]), num);
            den += wj[j];
          }
          const float dinv = rcp_fast(den + 1e-12f);
          const float fiv = num * dinv;
          const float gd = g[c] * dinv;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            ds[j][c] = gd * wj[j];
            dz[j][c] = gd * (__uint_as_float(sr[j][c]) - fiv) * wj[j] * (1.f - wj[j]);
          }
        }
        const int R = it * 4 + s4, b = R & 1;
        mbar_wait(&out_free_bar[b], (static_cast<uint32_t>(R >> 1) & 1u) ^ 1u);
        uint8_t* row = s_out + b * 32768 + m * 32 + half * 16;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          *reinterpret_cast<uint4*>(row + j * 4096) = make_uint4(pack_bf16x2(dz[j][0], dz[j][1]), pack_bf16x2(dz[j][2], dz[j][3]),
                                                                 pack_bf16x2(dz[j][4], dz[j][5]), pack_bf16x2(dz[j][6], dz[j][7]));
          *reinterpret_cast<uint4*>(row + (4 + j) * 4096) =
              make_uint4(pack_bf16x2(ds[j][0], ds[j][1]), pack_bf16x2(ds[j][2], ds[j][3]), pack_bf16x2(ds[j][4], ds[j][5]),
                         pack_bf16x2(ds[j][6], ds[j][7]));
        }
        fence_proxy_async();
        mbar_arrive(&out_full_bar[b]);
      }
    } else if (MODE == MODE_CTX2_BWD) {
      // Same TMEM contents as MODE_CTX2 (-z_j | s_j).  With D = sum_j w_j + 1e-12 and g = dL/dfi:
      //   ds_j (direct) = g w_j / D ;  dz_j = g (s_j - fi) / D * w_j (1 - w_j)
      // out (cs = 4096): channel j*512 + c <- dz_j ; 2048 + j*512 + c <- ds_j.
      const long long pix = valid ? (static_cast<long long>(n_img) * p.H + y) * p.W + x : 0;
      __nv_bfloat16* obase = p.out + pix * p.out_cs + n_tile * 64;
      const __nv_bfloat16* gbase = p.mask + pix * p.mask_cs + p.mask_coff + n_tile * 64;
      constexpr int CH = 64 / (EW / 4);              // channels of the n-tile handled by this warp
      const int hb = ((warp - 2) >> 2) * CH;
#pragma unroll 1
      for (int c0 = hb; c0 < hb + CH; c0 += 8) {
        uint32_t zr[4][8], sr[4][8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          tmem_ld_32x8(taddr + j * 64 + c0, zr[j]);
          tmem_ld_32x8(taddr + 256 + j * 64 + c0, sr[j]);
        }
        tmem_ld_wait();
        float g[8];
        {
          const uint4 gq = pre_m[0];
          if (valid && c0 + 8 < hb + CH) pre_m[0] = *reinterpret_cast<const uint4*>(gbase + c0 + 8);
          const __nv_bfloat16* gb = reinterpret_cast<const __nv_bfloat16*>(&gq);
#pragma unroll
          for (int e = 0; e < 8; ++e) g[e] = valid ? __bfloat162float(gb[e]) : 0.f;
        }
        float dz[4][8], ds[4][8];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          float wj[4], num = 0.f, den = 0.f;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            wj[j] = sigmoid_neg_fast(__uint_as_float(zr[j][c]));
            num = fmaf(wj[j], __uint_as_float(sr[j][c]), num);
            den += wj[j];
          }
          const float dinv = rcp_fast(den + 1e-12f);
          const float fiv = num * dinv;
          const float gd = g[c] * dinv;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            ds[j][c] = gd * wj[j];
            dz[j][c] = gd * (__uint_as_float(sr[j][c]) - fiv) * wj[j] * (1.f - wj[j]);
          }
        }
        if (valid) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            store_bf16_planes<NSPLIT, 8>(obase + j * 512 + c0, p.out_plane, dz[j]);
            store_bf16_planes<NSPLIT, 8>(obase + 2048 + j * 512 + c0, p.out_plane, ds[j]);
          }
        }
      }
    } else if (MODE == MODE_CTX_BWD) {
      // Backward of the fusion (CANNet.py:84) and of sigmoid(conv{S}_2(s_j - fv)) w.r.t. z_j.
      // With D = sum_j w_j + 1e-12 and g = dL/dfi:
      //   dL/ds_j (direct) = g w_j / D ;  dL/dw_j = g (s_j - fi) / D ;  dz_j = dL/dw_j w_j (1 - w_j)
      // out (cs = 4096): channel j*512 + c <- dz_j ; 2048 + j*512 + c <- direct ds_j.
      const int SS[4] = {1, 2, 3, 6};
      const int OFF[4] = {0, 1, 5, 14};
      int pidx[4][4];
      float ly0[4], ly1[4], lx0[4], lx1[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        int yi0, yi1, xi0, xi1;
        bilinear_tap(min(y, p.H - 1), SS[j], p.H, yi0, yi1, ly0[j], ly1[j]);
        bilinear_tap(min(x, p.W - 1), SS[j], p.W, xi0, xi1, lx0[j], lx1[j]);
        pidx[j][0] = (OFF[j] + yi0 * SS[j] + xi0) * 64;
        pidx[j][1] = (OFF[j] + yi0 * SS[j] + xi1) * 64;
        pidx[j][2] = (OFF[j] + yi1 * SS[j] + xi0) * 64;
        pidx[j][3] = (OFF[j] + yi1 * SS[j] + xi1) * 64;
      }
      const long long pix = valid ? (static_cast<long long>(n_img) * p.H + y) * p.W + x : 0;
      __nv_bfloat16* obase = p.out + pix * p.out_cs + n_tile * 64;
      const __nv_bfloat16* gbase = p.mask + pix * p.mask_cs + p.mask_coff + n_tile * 64;
      const float* sA = s_tab;
      const float* sB = s_tab + 3200;
#pragma unroll 1
      for (int c0 = 0; c0 < 64; c0 += 8) {
        float r[4][8];
#pragma unroll
        for (int j = 0; j < 4; ++j) load_acc<BN, GS, 8>(taddr, j * 64 + c0, nact, r[j]);
        float g[8];
        {
#pragma unroll
          for (int e = 0; e < 8; ++e) g[e] = 0.f;
#pragma unroll
          for (int pl = 0; pl < NSPLIT; ++pl) {
            const uint4 gq = *reinterpret_cast<const uint4*>(gbase + pl * p.mask_plane + c0);
            const __nv_bfloat16* gb = reinterpret_cast<const __nv_bfloat16*>(&gq);
#pragma unroll
            for (int e = 0; e < 8; ++e) g[e] += __bfloat162float(gb[e]);
          }
        }
        float dz[4][8], ds[4][8];
#pragma unroll
        for (int c = 0; c < 8; c += 4) {
          float sj[4][4], wj[4][4];
          float num[4] = {0.f, 0.f, 0.f, 0.f}, den[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 a00 = *reinterpret_cast<const float4*>(sA + pidx[j][0] + c0 + c);
            const float4 a01 = *reinterpret_cast<const float4*>(sA + pidx[j][1] + c0 + c);
            const float4 a10 = *reinterpret_cast<const float4*>(sA + pidx[j][2] + c0 + c);
            const float4 a11 = *reinterpret_cast<const float4*>(sA + pidx[j][3] + c0 + c);
            const float4 b00 = *reinterpret_cast<const float4*>(sB + pidx[j][0] + c0 + c);
            const float4 b01 = *reinterpret_cast<const float4*>(sB + pidx[j][1] + c0 + c);
            const float4 b10 = *reinterpret_cast<const float4*>(sB + pidx[j][2] + c0 + c);
            const float4 b11 = *reinterpret_cast<const float4*>(sB + pidx[j][3] + c0 + c);
            const float av00[4] = {a00.x, a00.y, a00.z, a00.w}, av01[4] = {a01.x, a01.y, a01.z, a01.w};
            const float av10[4] = {a10.x, a10.y, a10.z, a10.w}, av11[4] = {a11.x, a11.y, a11.z, a11.w};
            const float bv00[4] = {b00.x, b00.y, b00.z, b00.w}, bv01[4] = {b01.x, b01.y, b01.z, b01.w};
            const float bv10[4] = {b10.x, b10.y, b10.z, b10.w}, bv11[4] = {b11.x, b11.y, b11.z, b11.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float sv = ly0[j] * (lx0[j] * av00[e] + lx1[j] * av01[e]) +
                               ly1[j] * (lx0[j] * av10[e] + lx1[j] * av11[e]);
              const float u = ly0[j] * (lx0[j] * bv00[e] + lx1[j] * bv01[e]) +
                              ly1[j] * (lx0[j] * bv10[e] + lx1[j] * bv11[e]);
              const float z = u - r[j][c + e];
              const float wgt = 1.f / (1.f + __expf(-z));
              sj[j][e] = sv;
              wj[j][e] = wgt;
              num[e] = fmaf(wgt, sv, num[e]);
              den[e] += wgt;
            }
          }
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float dinv = 1.f / (den[e] + 1e-12f);
            const float fiv = num[e] * dinv;
            const float gd = g[c + e] * dinv;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              ds[j][c + e] = gd * wj[j][e];
              dz[j][c + e] = gd * (sj[j][e] - fiv) * wj[j][e] * (1.f - wj[j][e]);
            }
          }
        }
        if (valid) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            store_bf16_planes<NSPLIT, 8>(obase + j * 512 + c0, p.out_plane, dz[j]);
            store_bf16_planes<NSPLIT, 8>(obase + 2048 + j * 512 + c0, p.out_plane, ds[j]);
          }
        }
      }
    } else if (MODE == MODE_DFV) {
      // dfv[p][c] = ( add[p][c] - acc[p][c] + sum_j sum_{windows of scale j containing p}
      //               dpool_j[window][c] / area(window) ) * (fv[p][c] > 0)
      // acc = sum_j W_j2^T dz_j (this GEMM); add = gradient reaching fv through the concat;
      // the last term is the transpose of the adaptive average pooling (overlapping windows).
      const int SS[4] = {1, 2, 3, 6};
      const int OFF[4] = {0, 1, 5, 14};
      int spos[13];
      float swgt[13];
#pragma unroll
      for (int k = 0; k < 13; ++k) {
        spos[k] = 0;
        swgt[k] = 0.f;
      }
      if (!(kDfvX && p.dfv_kb)) {   // (with the extra k-block the accumulator already holds this term)
        const int yy = min(y, p.H - 1), xx = min(x, p.W - 1);
        int k = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int S = SS[j];
          int iy0 = 0, ix0 = 0;
          while (adaptive_end(iy0, p.H, S) <= yy) ++iy0;
          while (adaptive_end(ix0, p.W, S) <= xx) ++ix0;
          const bool y2 = (iy0 + 1 < S) && (adaptive_start(iy0 + 1, p.H, S) <= yy);
          const bool x2 = (ix0 + 1 < S) && (adaptive_start(ix0 + 1, p.W, S) <= xx);
          const int ny = (j == 0) ? 1 : 2, nx = (j == 0) ? 1 : 2;
#pragma unroll
          for (int a = 0; a < ny; ++a)
#pragma unroll
            for (int b = 0; b < nx; ++b) {
              const int iy = min(iy0 + a, S - 1), ix = min(ix0 + b, S - 1);
              const bool on = (a == 0 || y2) && (b == 0 || x2);
              const int hh = adaptive_end(iy, p.H, S) - adaptive_start(iy, p.H, S);
              const int ww = adaptive_end(ix, p.W, S) - adaptive_start(ix, p.W, S);
              spos[k] = (OFF[j] + iy * S + ix) * BN;   // row of the staged table
              swgt[k] = on ? 1.f / static_cast<float>(hh * ww) : 0.f;
              ++k;
            }
        }
      }
      const long long pix = valid ? (static_cast<long long>(n_img) * p.H + y) * p.W + x : 0;
      __nv_bfloat16* obase = p.out + pix * p.out_cs + p.out_coff + n0;
      const __nv_bfloat16* mbase = p.mask + pix * p.mask_cs + p.mask_coff + n0;
      const __nv_bfloat16* abase = p.add + pix * p.add_cs + p.add_coff + n0;
      const float* tab = s_tab;
      // upstream gradient (plane 0) and ReLU mask of the next column group are requested one group ahead
      uint4 nx_a[4] = {}, nx_m[4] = {};
      if (valid) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          nx_a[i] = *reinterpret_cast<const uint4*>(abase + cbeg + 8 * i);
          nx_m[i] = *reinterpret_cast<const uint4*>(mbase + cbeg + 8 * i);
        }
      }
#pragma unroll 1
      for (int c0 = cbeg; c0 < cend; c0 += 32) {
        float r[32];
        load_acc<BN, GS, 32>(taddr, c0, nact, r);
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = 0.f;
        uint4 cur_a[4], cur_m[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          cur_a[i] = nx_a[i];
          cur_m[i] = nx_m[i];
        }
        if (valid && c0 + 32 < cend) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            nx_a[i] = *reinterpret_cast<const uint4*>(abase + c0 + 32 + 8 * i);
            nx_m[i] = *reinterpret_cast<const uint4*>(mbase + c0 + 32 + 8 * i);
          }
        }
        if (valid) {
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = -r[i];
          if (!(kDfvX && p.dfv_kb))
#pragma unroll
          for (int k = 0; k < 13; ++k) {
            const float wk = swgt[k];
            const float* tp = tab + spos[k] + c0;
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
              const float4 t4 = *reinterpret_cast<const float4*>(tp + i);
              v[i] = fmaf(wk, t4.x, v[i]);
              v[i + 1] = fmaf(wk, t4.y, v[i + 1]);
              v[i + 2] = fmaf(wk, t4.z, v[i + 2]);
              v[i + 3] = fmaf(wk, t4.w, v[i + 3]);
            }
          }
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
#pragma unroll
            for (int pl = 0; pl < NSPLIT; ++pl) {
              const uint4 aq = pl == 0 ? cur_a[i >> 3] : *reinterpret_cast<const uint4*>(abase + pl * p.add_plane + c0 + i);
              const __nv_bfloat16* ab = reinterpret_cast<const __nv_bfloat16*>(&aq);
#pragma unroll
              for (int e = 0; e < 8; ++e) v[i + e] += __bfloat162float(ab[e]);
            }
            const uint4 mq = cur_m[i >> 3];
            const __nv_bfloat16* mb = reinterpret_cast<const __nv_bfloat16*>(&mq);
#pragma unroll
            for (int e = 0; e < 8; ++e)
              if (!(__bfloat162float(mb[e]) > 0.f)) v[i + e] = 0.f;
          }
          store_bf16_planes<NSPLIT, 32>(obase + c0, p.out_plane, v);
        }
        if (p.db) atomicAdd(&s_db[n0 + c0 + lane], warp_column_sums(v, lane));
      }
    }
    // this accumulator stage may be overwritten by the MMAs of the tile after next
    tc_fence_before();
    release_acc(as);
    if (kStaged) {
      // generic-proxy writes of s_out -> visible to the async proxy, then tell the output warp
      // (out-of-image pixels of edge tiles are clipped by the tensor map)
      fence_proxy_async();
      mbar_arrive(&out_full_bar[ob_i]);
      if (kMaskStaged) mbar_arrive(&mask_free_bar[mb_i]);
    }
    }  // tile loop
    if (kDbRegMode && db_reg) atomicAdd(&s_db[cbeg + lane], warp_column_sums(dbacc, lane));
    if (p.db) {
      asm volatile("bar.sync 1, %0;" ::"n"(32 * EW) : "memory");
      const int cout = p.n_tiles * BN;
      for (int i = etid; i < cout; i += 32 * EW) atomicAdd(p.db + i, s_db[i]);
    }
  }
#undef CAN_TILE_DECODE

  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();   // neither CTA leaves (or frees TMEM) while the other may still signal its barriers
  if (warp == 1) {
    tc_fence_after();
    if (PAIR) tmem_dealloc_pair(tmem_base, ACC_COLS * ACC_STAGES);
    else tmem_dealloc(tmem_base, ACC_COLS * ACC_STAGES);
  }
}

}  // namespace canb
