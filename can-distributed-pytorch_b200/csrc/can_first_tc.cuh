// conv1_1 (3 -> 64, 3x3, pad 1) on the tensor cores without an im2col buffer in HBM (bf16 path).
//
// The reference's first layer (model/CANNet.py:21,120-134: make_layers, cfg[0]) has K = 3*3*3 = 27.
// An im2col buffer padded to 64 channels costs 403 MB written + 403 MB read per 4 x 768x1024 step in
// the forward pass and another 403 MB read in the weight gradient; both kernels below rebuild the
// 27-wide patch rows in shared memory from the fp32 NCHW image (38 MB) instead:
//
//   * thread = pixel: 27 fp32 taps (k = ci*9 + ky*3 + kx, the order of the OIHW weight) from a staged
//     3 x 3 x (P+2) window -> bf16 -> four 16-byte chunks of a 128-byte row, written with the 128B
//     swizzle (chunk ^ (row & 7)), k = 27..31 zero.  That shared-memory image is at the same time
//       - a K-major  A[M = pixel][K = k]  operand (forward:  out = A W^T), and
//       - an MN-major B[K = pixel][N = k] operand (weight gradient: dW = dZ^T A),
//     of which the MMAs only touch K (resp. N) = 32.
//   * several small CTAs per SM (no intra-CTA pipeline): the phases of different CTAs overlap.
//
// Both kernels are HBM-bound by design: forward = image read + one bf16 activation write; weight
// gradient = one dZ read.
#pragma once
#include "can_ptx.cuh"
#include "can_simple.cuh"

namespace canb {

// Builds rows [0, P) of the patch matrix for pixels (n, y, x0 .. x0+P-1).  s_in: [3][3][P+4] floats.
// Called by all threads of the CTA; ends with the writes fenced for the async proxy (no barrier).
template <int P>
__device__ __forceinline__ void first_build_rows(const float* __restrict__ x, int n, int y, int x0, int H, int W,
                                                 float* s_in, uint8_t* s_a) {
  constexpr int PW = P + 4;
  static_assert(P == 128, "one thread per pixel, blockDim.x == P");
  {
    // all loads first, then the stores: a load-store-load-store sequence would expose the memory latency
    // nine times per tile (in-order issue stalls at every dependent store)
    const int t = threadIdx.x;
    float v[9], e[9];
#pragma unroll
    for (int r = 0; r < 9; ++r) {   // r = ci * 3 + window row; coalesced along x
      const int yy = y + (r % 3) - 1;
      const float* src = x + ((static_cast<long long>(n) * 3 + r / 3) * H + yy) * W + x0 - 1;
      const bool row_ok = yy >= 0 && yy < H;
      const int xx = x0 + t - 1, xe = x0 + P + t - 1;
      v[r] = (row_ok && xx >= 0 && xx < W) ? __ldg(src + t) : 0.f;
      e[r] = (t < 2 && row_ok && xe < W) ? __ldg(src + P + t) : 0.f;   // the two extra halo columns
    }
#pragma unroll
    for (int r = 0; r < 9; ++r) {
      s_in[r * PW + t] = v[r];
      if (t < 2) s_in[r * PW + P + t] = e[r];
    }
  }
  __syncthreads();
  for (int t = threadIdx.x; t < P; t += blockDim.x) {
    const bool inside = (x0 + t) < W;
    uint32_t pk[16];
#pragma unroll
    for (int k2 = 0; k2 < 16; ++k2) {
      float v[2];
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int k = 2 * k2 + e;
        v[e] = (k < 27 && inside) ? s_in[((k / 9) * 3 + (k % 9) / 3) * PW + t + (k % 3)] : 0.f;
      }
      pk[k2] = pack_bf16x2(v[0], v[1]);
    }
    uint8_t* row = s_a + t * 128;
#pragma unroll
    for (int c = 0; c < 4; ++c)
      *reinterpret_cast<uint4*>(row + ((c ^ (t & 7)) << 4)) = make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
  }
  fence_proxy_async();
}

struct FirstFwdParams {
  const float* x;      // [N][3][H][W] fp32
  const float* bias;   // [64]
  int N, H, W;
  int tiles_x;         // ceil(W / 128)
  int num_tiles;       // N * H * tiles_x
};

// tmW: packed weight [64 co][64 k] bf16 (k >= 27 zero), box (64, 64), 128B swizzle
// tmO: output activation (64 ch, W, N*H), box (64, 128, 1), 128B swizzle
constexpr int kFirstFwdSmem = 1024 + 16384 /*A, then the staged output tile*/ + 8192 /*W*/ + 9 * 132 * 4 + 64 * 4 + 64;
__global__ void __launch_bounds__(128)
conv_first_tc_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmO,
                     const FirstFwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1024-aligned, still a shared-space pointer
  uint8_t* s_a = smem;
  uint8_t* s_w = smem + 16384;
  uint8_t* s_o = s_a;   // the output tile is staged over the A tile once the MMAs have consumed it
  float* s_in = reinterpret_cast<float*>(smem + 24576);
  float* s_bias = s_in + 9 * 132;
  uint64_t* w_bar = reinterpret_cast<uint64_t*>(s_bias + 64);
  uint64_t* mma_bar = w_bar + 1;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(mma_bar + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmO);
    mbar_init(w_bar, 1);
    mbar_init(mma_bar, 1);
    fence_mbar_init();
  }
  if (threadIdx.x < 64) s_bias[threadIdx.x] = p.bias[threadIdx.x];
  if (warp == 0) {
    tmem_alloc(tmem_ptr_smem, 64);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  if (threadIdx.x == 0) {
    mbar_expect_tx(w_bar, 8192);
    tma_load_2d(s_w, &tmW, w_bar, 0, 0);
  }
  constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 0, 0);
  const int m = threadIdx.x;   // pixel of the tile = TMEM lane = row of A and of the staged output
  const uint32_t taddr = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);
  int it = 0;
  for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
    const int tx = tile % p.tiles_x;
    const int row_id = tile / p.tiles_x;   // n * H + y
    const int y = row_id % p.H, n = row_id / p.H;
    const int x0 = tx * 128;
    if (threadIdx.x == 0) bulk_wait_group_read0();   // the previous tile's store has read s_o (= s_a); ordered
    first_build_rows<128>(p.x, n, y, x0, p.H, p.W, s_in, s_a);   // before the row writes by the barrier inside
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) {
      if (it == 0) mbar_wait(w_bar, 0);
      tc_fence_after();
      const uint64_t da = umma_smem_desc(smem_u32(s_a), 16, 1024);
      const uint64_t db = umma_smem_desc(smem_u32(s_w), 16, 1024);
      umma_bf16(tmem_base, da, db, idesc, 0u);
      umma_bf16(tmem_base, da + 2, db + 2, idesc, 1u);   // K = 16 .. 31
      umma_commit(mma_bar);
    }
    mbar_wait(mma_bar, static_cast<uint32_t>(it) & 1u);
    tc_fence_after();
#pragma unroll
    for (int c0 = 0; c0 < 64; c0 += 32) {
      uint32_t r[32];
      tmem_ld_32x32(taddr + c0, r);
      tmem_ld_wait();
      uint8_t* orow = s_o + m * 128;
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = fmaxf(__uint_as_float(r[i + e]) + s_bias[c0 + i + e], 0.f);
        const int chunk = (c0 + i) >> 3;
        *reinterpret_cast<uint4*>(orow + ((chunk ^ (m & 7)) << 4)) =
            make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
      }
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) {
      tma_store_4d(&tmO, s_o, 0, x0, row_id, 0);   // pixels past W are clipped by the map
      bulk_commit_group();
    }
  }
  if (threadIdx.x == 0) bulk_wait_group_read0();
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 64);
  }
  (void)lane;
}

struct FirstWgradParams {
  const float* x;      // [N][3][H][W] fp32
  int N, H, W;
  int tiles_x;         // ceil(W / 128)
  int num_patches;     // N * H * tiles_x
  float* ws;           // [gridDim.x][64 co][32 k] fp32 partial sums
};

// tmDZ: dZ of conv1_1's output (64 ch, W, N*H), box (64, 128, 1), 128B swizzle (an MN-major A tile: [128 px][64 co])
constexpr int kFirstWgSmem = 1024 + 16384 /*dZ*/ + 16384 /*B*/ + 9 * 132 * 4 + 64;
__global__ void __launch_bounds__(128)
wgrad_first_tc_kernel(const __grid_constant__ CUtensorMap tmDZ, const FirstWgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1024-aligned, still a shared-space pointer
  uint8_t* s_dz = smem;            // [128 px][64 co]; M rows 64..127 of the MMA re-read rows 0..63 (group distance 0), ignored
  uint8_t* s_b = smem + 16384;     // [128 px][32 k (+ unused half)]
  float* s_in = reinterpret_cast<float*>(smem + 32768);
  uint64_t* dz_bar = reinterpret_cast<uint64_t*>(s_in + 9 * 132);
  uint64_t* mma_bar = dz_bar + 1;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(mma_bar + 1);
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmDZ);
    mbar_init(dz_bar, 1);
    mbar_init(mma_bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(tmem_ptr_smem, 32);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  constexpr uint32_t idesc = umma_idesc_bf16(128, 32, 1, 1);
  int it = 0;
  for (int patch = blockIdx.x; patch < p.num_patches; patch += gridDim.x, ++it) {
    const int tx = patch % p.tiles_x;
    const int row_id = patch / p.tiles_x;
    const int y = row_id % p.H, n = row_id / p.H;
    const int x0 = tx * 128;
    // previous patch's MMAs have finished reading s_dz / s_b
    if (it > 0) mbar_wait(mma_bar, static_cast<uint32_t>(it - 1) & 1u);
    if (threadIdx.x == 0) {
      mbar_expect_tx(dz_bar, 16384);
      tma_load_3d(s_dz, &tmDZ, dz_bar, 0, x0, row_id);   // pixels past W arrive as zeros
    }
    first_build_rows<128>(p.x, n, y, x0, p.H, p.W, s_in, s_b);
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) {
      mbar_wait(dz_bar, static_cast<uint32_t>(it) & 1u);
      tc_fence_after();
      const uint64_t da = umma_smem_desc(smem_u32(s_dz), 0, 1024);
      const uint64_t db = umma_smem_desc(smem_u32(s_b), 0, 1024);
#pragma unroll
      for (int k = 0; k < 8; ++k)   // 16 pixels per MMA
        umma_bf16(tmem_base, da + ((k * 2048) >> 4), db + ((k * 2048) >> 4), idesc, (it | k) ? 1u : 0u);
      umma_commit(mma_bar);
    }
  }
  if (it > 0) mbar_wait(mma_bar, static_cast<uint32_t>(it - 1) & 1u);
  tc_fence_after();
  if (threadIdx.x < 64) {   // TMEM lane = output channel
    uint32_t r[32];
    tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(warp * 32) << 16), r);
    tmem_ld_wait();
    float* dst = p.ws + (static_cast<long long>(blockIdx.x) * 64 + threadIdx.x) * 32;
#pragma unroll
    for (int i = 0; i < 32; i += 4)
      *reinterpret_cast<float4*>(dst + i) = it > 0 ? make_float4(__uint_as_float(r[i]), __uint_as_float(r[i + 1]),
                                                                 __uint_as_float(r[i + 2]), __uint_as_float(r[i + 3]))
                                                   : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 32);
  }
}

// grad[co][k] (k < 27, the OIHW layout of frontend.0.weight) = (accumulate ? grad : 0) + alpha * sum_s ws[s][co][k]
__global__ void wgrad_first_reduce_kernel(const float* __restrict__ ws, float* __restrict__ grad, int splits,
                                          int accumulate, float alpha) {
  const int co = blockIdx.x, k = threadIdx.x & 31, part = threadIdx.x >> 5;   // 8 parts x 32 k
  __shared__ float s_p[8][32];
  float acc = 0.f;
  for (int s = part; s < splits; s += 8) acc += ws[(static_cast<long long>(s) * 64 + co) * 32 + k];
  s_p[part][k] = acc;
  __syncthreads();
  if (part == 0 && k < 27) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += s_p[i][k];
    float* g = grad + co * 27 + k;
    *g = (accumulate ? *g : 0.f) + alpha * t;
  }
}

}  // namespace canb
