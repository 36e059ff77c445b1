// Weight-gradient GEMM on tcgen05/TMEM fed by TMA (sm_100a).
//
//   dW[co][tap][ci] = sum_{n,y,x} dZ[n,y,x,co] * X[n, y+dy(tap), x+dx(tap), ci]
//
// GEMM view: M = Cout tile (128), N = Cin tile (BN), K = pixels.  Both operands live in memory
// as NHWC (channels contiguous), i.e. they are MN-major for this GEMM: a TMA box of
// (64 channels x 64 pixels) lands as 64 rows of 128 B -- exactly the 128B-swizzled MN-major
// canonical layout (8-pixel groups 1024 B apart = SBO, 64-channel groups one box apart = LBO).
// One CTA owns (cout tile, cin tile, group of T taps) and a contiguous slice of the pixel
// patches (split-K).  The dZ patch is loaded once per stage and multiplied with the T shifted X
// patches into T TMEM accumulators, so dZ streams from HBM once per tap GROUP instead of once per
// tap (the large-spatial layers are HBM-bound otherwise).  Partial tiles go to an fp32 workspace,
// reduced deterministically by wgrad_reduce_kernel_v2.
#pragma once
#include "can_ptx.cuh"

namespace canb {

struct WgradTcParams {
  int N, H, W;
  int Cout, Cin;
  int taps, dil;
  int TW, TH;  // patch, TW*TH == 64 (32 for the PIX = 32 instantiations)
  int tiles_x, tiles_y;
  int m_tiles, n_tiles;
  int patches_per_split;
  int total_patches;
  int stages;
  int m_groups;  // 64-channel groups of dZ actually loaded (1 when Cout == 64)
  int x_img_tiles;  // > 0 (single plane): dZ is ONE image shared by B images of X; n_tile = image * x_img_tiles + cin tile,
                    // the result is [Cout][B][Cin]  (U^T of the contextual backward for the whole batch in one launch)
  int tap_groups;  // ceil(taps / T)
  int dbg;         // bring-up experiments only (bit 0: skip the partial-tile stores)
  int ws_tma;      // partial tiles leave through TMA stores (tmWS) instead of per-lane stores
  int hx_bytes;    // HX: bytes of the X halo tile of a stage ((BN / 64) groups x (TW + 2 dil) x TH pixels x 128 B)
  int hx_stage;    // HX: bytes of a stage (dZ patch + halo tile)
  float* ws;     // [split][tap][Cout][Cin]
};

template <int BN, int NSPLIT, int T, int PIX = 64>
struct WgradTcSmem {
  static constexpr int kGroupBytes = PIX * 128;           // PIX pixels x 64 ch x bf16 (PIX = 64, or 32 for deeper pipelines)
  static constexpr int kABytes = 2 * kGroupBytes;         // 128 cout
  static constexpr int kBBytes = (BN / 64) * kGroupBytes; // one tap
  static constexpr int kStageBytes = NSPLIT * (kABytes + T * kBBytes);
  static constexpr int kTmemCols = (T * BN <= 32) ? 32 : (T * BN <= 64) ? 64 : (T * BN <= 128) ? 128 : (T * BN <= 256) ? 256 : 512;
  static __host__ __device__ constexpr int bytes(int stages) { return 1024 + stages * kStageBytes + 1024; }
};

// HX = 1 (bf16 path, BN = 128, the three taps of a CTA share dy): the three X patches are column-shifted views of ONE halo
// tile ((TW + 2 dil) x TH pixels, TW = 8 so that the 8-pixel groups of the MN-major operand are whole patch rows one halo
// pitch apart) instead of three tiles -- a stage is 18-20 KB instead of 32 KB.  The kernel was bound by what an SM can take
// from L2 (ncu: 75 B/cycle/SM = 64 B per L2 clock, tensor pipe 86 %, `profiles/r02_ncu_wgrad_128_3_32_details.txt`).
template <int BN, int NSPLIT, int T, int PIX = 64, int HX = 0>
__global__ void __launch_bounds__(192, 1)
wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmDZ, const __grid_constant__ CUtensorMap tmX,
                const __grid_constant__ CUtensorMap tmWS, const WgradTcParams p) {
  using L = WgradTcSmem<BN, NSPLIT, T, PIX>;
  static_assert(T * BN <= 512, "accumulators of one tap group must fit TMEM");
  static_assert(!HX || (NSPLIT == 1 && T == 3 && PIX == 32), "halo X tiles: three taps of one row, 32-pixel patches");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1024-aligned, still a shared-space pointer
  const int stages = p.stages;
  const int stage_bytes = HX ? p.hx_stage : L::kStageBytes;
  uint8_t* tail = smem + stages * stage_bytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tail);
  uint64_t* empty_bar = full_bar + 8;
  uint64_t* tmem_full_bar = empty_bar + 8;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  int t = blockIdx.x;
  const int n_tile = t % p.n_tiles;
  t /= p.n_tiles;
  const int m_tile = t % p.m_tiles;
  const int tap0 = (t / p.m_tiles) * T;               // first tap of this CTA's group
  const int ntaps = min(T, p.taps - tap0);
  const int split = blockIdx.y;
  const int patch_begin = split * p.patches_per_split;
  const int patch_end = min(patch_begin + p.patches_per_split, p.total_patches);
  const int m_groups = p.m_groups;  // 1 or 2
  const uint32_t stage_tx = HX ? m_groups * L::kGroupBytes + p.hx_bytes : NSPLIT * (m_groups * L::kGroupBytes + ntaps * L::kBBytes);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmDZ);
    tma_prefetch_desc(&tmX);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < stages; ++s) {
        mbar_init(&full_bar[s], 1);
        mbar_init(&empty_bar[s], 1);
      }
      mbar_init(tmem_full_bar, 1);
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(tmem_ptr_smem, L::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  pdl_launch_dependents();   // (see can_ptx.cuh: the prologue above overlaps the previous kernel's tail)
  pdl_wait();

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      // this thread is alone on the critical path of the pipeline: tap offsets once, patch coordinates incrementally
      int tdx[T], tdy[T];
#pragma unroll
      for (int tt = 0; tt < T; ++tt) {
        const int tap = tap0 + tt;
        tdx[tt] = p.taps == 9 ? (tap % 3 - 1) * p.dil : 0;
        tdy[tt] = p.taps == 9 ? (tap / 3 - 1) * p.dil : 0;
      }
      int tx_i = patch_begin % p.tiles_x, ty_i = (patch_begin / p.tiles_x) % p.tiles_y,
          n_img = patch_begin / (p.tiles_x * p.tiles_y);
      for (int pid = patch_begin; pid < patch_end; ++pid) {
        const int x0 = tx_i * p.TW, y0 = ty_i * p.TH;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sa = smem + stage * stage_bytes;
        uint8_t* sb = sa + NSPLIT * L::kABytes;  // [plane][tap][BN/64 groups]
        mbar_expect_tx(&full_bar[stage], stage_tx);
        if (HX) {
          tma_load_5d(sa, &tmDZ, &full_bar[stage], 0, x0, y0, m_tile * 2, n_img);
          tma_load_5d(sb, &tmX, &full_bar[stage], 0, x0 - p.dil, y0 + tdy[0], n_tile * (BN / 64), n_img);   // (box: the halo tile)
        } else if (NSPLIT == 1) {
          // single-plane maps carry the 64-channel group as their 4th dimension: one TMA op per operand tile
          tma_load_5d(sa, &tmDZ, &full_bar[stage], 0, x0, y0, m_tile * 2, n_img);
          const int xg = p.x_img_tiles ? (n_tile % p.x_img_tiles) * (BN / 64) : n_tile * (BN / 64);
          const int xn = p.x_img_tiles ? n_tile / p.x_img_tiles : n_img;
#pragma unroll
          for (int tt = 0; tt < T; ++tt)
            if (tt < ntaps)
              tma_load_5d(sb + tt * L::kBBytes, &tmX, &full_bar[stage], 0, x0 + tdx[tt], y0 + tdy[tt], xg, xn);
        } else
#pragma unroll
        for (int pl = 0; pl < NSPLIT; ++pl) {
          for (int g = 0; g < m_groups; ++g)
            tma_load_5d(sa + pl * L::kABytes + g * L::kGroupBytes, &tmDZ, &full_bar[stage],
                        m_tile * 128 + g * 64, x0, y0, n_img, pl);
          for (int tt = 0; tt < ntaps; ++tt) {
            const int tap = tap0 + tt;
            int dy = 0, dx = 0;
            if (p.taps == 9) {
              dy = (tap / 3 - 1) * p.dil;
              dx = (tap % 3 - 1) * p.dil;
            }
#pragma unroll
            for (int g = 0; g < BN / 64; ++g)
              tma_load_5d(sb + (pl * T + tt) * L::kBBytes + g * L::kGroupBytes, &tmX, &full_bar[stage],
                          n_tile * BN + g * 64, x0 + dx, y0 + dy, n_img, pl);
          }
        }
        if (++stage == stages) {
          stage = 0;
          phase ^= 1;
        }
        if (++tx_i == p.tiles_x) {
          tx_i = 0;
          if (++ty_i == p.tiles_y) {
            ty_i = 0;
            ++n_img;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      uint32_t accum = 0;
      for (int pid = patch_begin; pid < patch_end; ++pid) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * stage_bytes);
        // one descriptor pair per stage; planes, taps and k-steps only add to the address field (>> 4)
        const uint64_t da0 = umma_smem_desc(sa, L::kGroupBytes, 1024);
        if (HX) {
          // B = the halo tile: 8-pixel groups are patch rows one halo pitch apart (SBO), 64-channel groups one whole tile
          // apart (LBO); tap dx starts dx * dil pixels into the rows; a k-step of 16 pixels is two rows
          const uint32_t pitch = static_cast<uint32_t>(p.TW + 2 * p.dil) * 128u;
          const uint64_t dbh = umma_smem_desc(sa + L::kABytes, pitch * p.TH, pitch);
          constexpr uint32_t idesc_h = umma_idesc_bf16(128, BN, 1, 1);
#pragma unroll
          for (int tt = 0; tt < T; ++tt) {
#pragma unroll
            for (int k = 0; k < PIX / 16; ++k)
              umma_bf16(tmem_base + tt * BN, da0 + ((k * 2048) >> 4), dbh + ((tt * p.dil * 128 + k * 2 * pitch) >> 4), idesc_h,
                        (accum | k) ? 1u : 0u);
          }
          accum = 1;
          umma_commit(&empty_bar[stage]);
          if (++stage == stages) {
            stage = 0;
            phase ^= 1;
          }
          continue;
        }
        const uint64_t db0 = umma_smem_desc(sa + NSPLIT * L::kABytes, L::kGroupBytes, 1024);
        // Several taps per instruction: the X tiles of consecutive taps are consecutive 64-channel
        // groups in shared memory, exactly what one wider MMA (N = taps * BN <= 256) expects, and a
        // tcgen05.mma costs ~60 cycles + 0.63 cycles per N column (measured), so fewer, wider MMAs win.
        constexpr int kTapsPerMma = (256 / BN) < T ? (256 / BN) : T;
        for (int tt = 0; tt < ntaps; tt += kTapsPerMma) {
          const int tc = min(kTapsPerMma, ntaps - tt);
          const uint32_t idesc_n = umma_idesc_bf16(128, tc * BN, 1, 1);
#pragma unroll
          for (int pa = 0; pa < NSPLIT; ++pa) {
#pragma unroll
            for (int pb = 0; pb < NSPLIT; ++pb) {
              if (pa + pb < NSPLIT) {
#pragma unroll
                for (int k = 0; k < PIX / 16; ++k) {  // 16 pixels (= 2 swizzle atoms of 8 rows) per MMA
                  umma_bf16(tmem_base + tt * BN, da0 + ((pa * L::kABytes + k * 2048) >> 4),
                            db0 + (((pb * T + tt) * L::kBBytes + k * 2048) >> 4), idesc_n,
                            (accum | (pa + pb) | k) ? 1u : 0u);
                }
              }
            }
          }
        }
        accum = 1;
        umma_commit(&empty_bar[stage]);
        if (++stage == stages) {
          stage = 0;
          phase ^= 1;
        }
      }
      umma_commit(tmem_full_bar);
    }
  } else {
    const int q = warp & 3;
    const int co = m_tile * 128 + q * 32 + lane;
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    const bool valid = (co < p.Cout) && (patch_end > patch_begin);
    if (p.ws_tma) {
      // The operand ring is idle now: 32-column slabs of the accumulator go through two 16 KB staging tiles
      // ([128 rows][32 fp32], 128B swizzle) and leave as TMA stores -- a row-per-lane store pattern (every lane its own
      // 128-byte line) runs at 4-5 B/cycle/SM and cost 6 % of all weight-gradient time.  tmWS: (Cin, Cout, split*tap),
      // box (32, 128, 1): rows past Cout are clipped by the map.
      const int row = q * 32 + lane, etid = threadIdx.x - 64;
      int chunk = 0;
      for (int tt = 0; tt < ntaps; ++tt) {
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 32, ++chunk) {
          uint32_t r[32];
          tmem_ld_32x32(taddr + tt * BN + c0, r);
          tmem_ld_wait();
          uint8_t* tile = smem + (chunk & 1) * 16384 + row * 128;
#pragma unroll
          for (int i = 0; i < 32; i += 4)
            *reinterpret_cast<uint4*>(tile + (((i >> 2) ^ (row & 7)) << 4)) = make_uint4(r[i], r[i + 1], r[i + 2], r[i + 3]);
          fence_proxy_async();
          if (etid == 0) bulk_wait_group_read0();   // every earlier store (incl. the last user of the other tile) has read
          asm volatile("bar.sync 1, 128;" ::: "memory");
          if (etid == 0 && patch_end > patch_begin) {
            tma_store_3d(&tmWS, smem + (chunk & 1) * 16384, n_tile * BN + c0, m_tile * 128, split * p.taps + tap0 + tt);
            bulk_commit_group();
          }
        }
      }
      if (etid == 0) bulk_wait_group_read0();
    } else
    for (int tt = 0; tt < ntaps; ++tt) {
      float* dst =
          p.ws + ((static_cast<long long>(split) * p.taps + tap0 + tt) * p.Cout + co) * p.Cin + n_tile * BN;
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + tt * BN + c0, r);
        tmem_ld_wait();
        if (valid && !(p.dbg & 1)) {
#pragma unroll
          for (int i = 0; i < 32; i += 4)
            *reinterpret_cast<float4*>(dst + c0 + i) = make_float4(__uint_as_float(r[i]), __uint_as_float(r[i + 1]),
                                                                   __uint_as_float(r[i + 2]), __uint_as_float(r[i + 3]));
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, L::kTmemCols);
  }
}

// Weight gradient of a 64 -> 64 3x3 layer (bf16 path), operand roles swapped.
//
// With M = Cout = 64 the kernel above spends half of every M = 128 MMA on padding and needs three
// CTAs (tap groups) that each re-read the dZ patch.  Here the nine shifted X patches are the A
// operand: two taps (2 x 64 input channels, consecutive 8 KB tiles in shared memory = the two
// 64-channel groups of an MN-major M = 128 operand) per MMA, dZ (N = 64 output channels) is B,
//   D_pair[(tap, ci)][co] += X_tap[pixel][ci]^T dZ[pixel][co]
// so all nine taps (five pair accumulators, 320 TMEM columns) live in one CTA, every MMA row is
// useful (the ninth tap's partner rows are ignored), and a patch costs one dZ + nine X tiles.
// Partial results go to ws[split][tap][ci][co] (note: ci before co), reduced by wgrad_reduce_kernel_t.
// The X patches of the three taps that share a column shift dx are row-shifted views of ONE halo
// tile ((TH + 2 dil) rows, row = TW pixels = a multiple of 1024 B, so the swizzle phase is kept):
// a stage holds three halo tiles + the dZ patch (44 KB for dil = 1) instead of ten 8 KB tiles.
struct WgradSwapSmem {
  static constexpr int kTile = 64 * 128;                  // 64 pixels x 64 ch x bf16
  static __host__ __device__ constexpr int halo_bytes(int dil) { return (64 + 2 * dil * 16) * 128; }  // TW*TH == 64, TW <= 16
  static __host__ __device__ constexpr int stage_bytes(int dil) { return 3 * halo_bytes(dil) + kTile; }
  static __host__ __device__ constexpr int bytes(int stages, int dil) { return 1024 + stages * stage_bytes(dil) + 1024; }
};

__global__ void __launch_bounds__(192, 1)
wgrad_tc_swap64_kernel(const __grid_constant__ CUtensorMap tmDZ, const __grid_constant__ CUtensorMap tmX,
                       const WgradTcParams p) {
  using L = WgradSwapSmem;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1024-aligned, still a shared-space pointer
  const int stages = p.stages;
  const int row_bytes = p.TW * 128;                       // one patch row (1024 or 2048 B)
  const int halo_bytes = (p.TH + 2 * p.dil) * row_bytes;
  const int stage_bytes = 3 * halo_bytes + L::kTile;
  uint8_t* tail = smem + stages * stage_bytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tail);
  uint64_t* empty_bar = full_bar + 8;
  uint64_t* tmem_full_bar = empty_bar + 8;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int split = blockIdx.x;
  const int patch_begin = split * p.patches_per_split;
  const int patch_end = min(patch_begin + p.patches_per_split, p.total_patches);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmDZ);
    tma_prefetch_desc(&tmX);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < stages; ++s) {
        mbar_init(&full_bar[s], 1);
        mbar_init(&empty_bar[s], 1);
      }
      mbar_init(tmem_full_bar, 1);
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(tmem_ptr_smem, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  pdl_launch_dependents();   // (see can_ptx.cuh: the prologue above overlaps the previous kernel's tail)
  pdl_wait();

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int pid = patch_begin; pid < patch_end; ++pid) {
        int q = pid;
        const int tx_i = q % p.tiles_x;
        q /= p.tiles_x;
        const int ty_i = q % p.tiles_y;
        const int n_img = q / p.tiles_y;
        const int x0 = tx_i * p.TW, y0 = ty_i * p.TH;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sx = smem + stage * stage_bytes;   // 3 halo tiles (dx = -dil, 0, +dil), then dZ
        mbar_expect_tx(&full_bar[stage], static_cast<uint32_t>(stage_bytes));
#pragma unroll
        for (int dxi = 0; dxi < 3; ++dxi)
          tma_load_5d(sx + dxi * halo_bytes, &tmX, &full_bar[stage], 0, x0 + (dxi - 1) * p.dil, y0 - p.dil, n_img, 0);
        tma_load_5d(sx + 3 * halo_bytes, &tmDZ, &full_bar[stage], 0, x0, y0, n_img, 0);
        if (++stage == stages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 1, 1);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t accum = 0;
      for (int pid = patch_begin; pid < patch_end; ++pid) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t sx = smem_u32(smem + stage * stage_bytes);
        const uint32_t dyb = static_cast<uint32_t>(p.dil * row_bytes);           // one tap row (dy) down
        const uint64_t db0 = umma_smem_desc(sx + 3 * halo_bytes, L::kTile, 1024);
        // pairs 0..2: taps (dy=-1, dx) and (dy=0, dx) of halo tile dx: the two "64-channel groups" are dyb apart
        // pair 3: taps (dy=+1, dx=-1) and (dy=+1, dx=0): one halo tile apart;  pair 4: tap (dy=+1, dx=+1) alone
#pragma unroll
        for (int pair = 0; pair < 5; ++pair) {
          const uint32_t a_start = pair < 3 ? sx + pair * halo_bytes : sx + (pair - 3) * 2 * halo_bytes + 2 * dyb;
          // (pair 4: the partner rows are ignored; distance 0 keeps their reads inside this stage)
          const uint64_t da0 =
              umma_smem_desc(a_start, pair < 3 ? dyb : (pair == 3 ? static_cast<uint32_t>(halo_bytes) : 0u), 1024);
#pragma unroll
          for (int k = 0; k < 4; ++k)   // 16 pixels per MMA
            umma_bf16(tmem_base + pair * 64, da0 + ((k * 2048) >> 4), db0 + ((k * 2048) >> 4), idesc,
                      (accum | k) ? 1u : 0u);
        }
        accum = 1;
        umma_commit(&empty_bar[stage]);
        if (++stage == stages) {
          stage = 0;
          phase ^= 1;
        }
      }
      umma_commit(tmem_full_bar);
    }
  } else {
    const int q = warp & 3;
    const int row = q * 32 + lane;            // (tap within pair, ci)
    const int unit = row >> 6, ci = row & 63;
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    const bool any = patch_end > patch_begin;
#pragma unroll 1
    for (int pair = 0; pair < 5; ++pair) {
      // pair -> taps (see the MMA issuer): tap = dyi * 3 + dxi
      const int tap = pair < 3 ? unit * 3 + pair : (pair == 3 ? 6 + unit : (unit == 0 ? 8 : 9));
      float* dst = p.ws + ((static_cast<long long>(split) * 9 + tap) * 64 + ci) * 64;
#pragma unroll 1
      for (int c0 = 0; c0 < 64; c0 += 32) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + pair * 64 + c0, r);
        tmem_ld_wait();
        if (tap < 9 && any) {
#pragma unroll
          for (int i = 0; i < 32; i += 4)
            *reinterpret_cast<float4*>(dst + c0 + i) = make_float4(__uint_as_float(r[i]), __uint_as_float(r[i + 1]),
                                                                   __uint_as_float(r[i + 2]), __uint_as_float(r[i + 3]));
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// grad[co][ci][tap] = (accumulate ? grad : 0) + alpha * sum_s ws[s][tap][ci][co]   (ws of wgrad_tc_swap64_kernel)
// grid (cout/32, cin), block 32*taps
__global__ void wgrad_reduce_kernel_t(const float* __restrict__ ws, float* __restrict__ grad, int splits, int taps,
                                      int cout, int cin, int accumulate, float alpha) {
  __shared__ float s_t[32 * 9];
  pdl_launch_dependents();
  pdl_wait();
  const int ci = blockIdx.y, co0 = blockIdx.x * 32;
  const int c = threadIdx.x & 31, tap = threadIdx.x >> 5;
  const long long slab = static_cast<long long>(taps) * cout * cin;
  const float* src = ws + (static_cast<long long>(tap) * cin + ci) * cout + co0 + c;
  float acc = 0.f;
  int sp = 0;
  for (; sp + 4 <= splits; sp += 4) {
    const float a0 = src[(sp + 0) * slab], a1 = src[(sp + 1) * slab], a2 = src[(sp + 2) * slab],
                a3 = src[(sp + 3) * slab];
    acc += (a0 + a1) + (a2 + a3);
  }
  for (; sp < splits; ++sp) acc += src[sp * slab];
  s_t[c * taps + tap] = acc;
  __syncthreads();
  const int co = co0 + threadIdx.x / taps, tp = threadIdx.x % taps;
  float* g = grad + (static_cast<long long>(co) * cin + ci) * taps + tp;
  *g = (accumulate ? *g : 0.f) + alpha * s_t[threadIdx.x];
}

}  // namespace canb
