// Second-generation HBM-bound helpers (coalesced, vectorised): bias gradient, split-K wgrad
// reduction into the parameter layout, and weight packing for both GEMM orientations.
#pragma once
#include "can_ptx.cuh"
#include "can_simple.cuh"

namespace canb {

// db[c] += sum over pixels of dz[pix][c]   (dz: NHWC bf16 planes, C % 8 == 0, C <= 1024)
// 256 threads = G channel groups (8 channels = one 16-byte load each) x PL pixel lanes.
// Each block owns a contiguous pixel range, keeps 4 loads in flight per thread, combines the
// pixel lanes in shared memory and issues one atomicAdd per channel.
__global__ void __launch_bounds__(256)
bias_grad_kernel_v2(const __nv_bfloat16* __restrict__ dz, long long plane, int nsplit, int cs, int coff, int C,
                    long long npix, float* __restrict__ db) {
  __shared__ float s_part[256 * 8];
  const int G = C >> 3;
  const int PL = 256 / G;          // G in {8,16,32,64,128} -> PL in {32,...,2}
  const int cg = threadIdx.x % G, pl = threadIdx.x / G;
  const long long per = (npix + gridDim.x - 1) / gridDim.x;
  const long long p0 = blockIdx.x * per, p1 = min(p0 + per, npix);
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
  if (pl < PL) {
    const __nv_bfloat16* base = dz + coff + cg * 8;
    for (int s = 0; s < nsplit; ++s) {
      const __nv_bfloat16* b = base + s * plane;
      long long px = p0 + pl;
      for (; px + 3LL * PL < p1; px += 4LL * PL) {
        uint4 q[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) q[u] = *reinterpret_cast<const uint4*>(b + (px + static_cast<long long>(u) * PL) * cs);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q[u]);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            acc[2 * e] += __low2float(h[e]);
            acc[2 * e + 1] += __high2float(h[e]);
          }
        }
      }
      for (; px < p1; px += PL) {
        const uint4 q = *reinterpret_cast<const uint4*>(b + px * cs);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          acc[2 * e] += __low2float(h[e]);
          acc[2 * e + 1] += __high2float(h[e]);
        }
      }
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) s_part[threadIdx.x * 8 + e] = acc[e];
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += 256) {
    const int g = c >> 3, e = c & 7;
    float t = 0.f;
    for (int l = 0; l < PL; ++l) t += s_part[(l * G + g) * 8 + e];
    atomicAdd(db + c, t);
  }
}

// Reduction of the split-K partials into the parameter layout:
//   grad[co][ci][tap] = (accumulate ? grad : 0) + alpha * sum_s ws[s][tap][co][ci]
// grid (ceil(cin/128), cout), block 32*taps: every thread owns four consecutive ci of one tap (16-byte loads,
// up to four splits in flight), the block transposes 128 ci x taps through shared memory and writes
// 128*taps contiguous floats.
__global__ void wgrad_reduce_kernel_v3(const float* __restrict__ ws, float* __restrict__ grad, int splits, int taps,
                                       int cout, int cin, int accumulate, float alpha) {
  __shared__ float s_t[128 * 9];
  pdl_launch_dependents();
  pdl_wait();
  const int co = blockIdx.y, ci0 = blockIdx.x * 128;
  const int lane = threadIdx.x & 31, tap = threadIdx.x >> 5;
  const int c4 = lane * 4;
  const int nci = min(128, cin - ci0);
  const long long slab = static_cast<long long>(taps) * cout * cin;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (c4 < nci) {
    const float* src = ws + (static_cast<long long>(tap) * cout + co) * cin + ci0 + c4;
    int sp = 0;
    for (; sp + 4 <= splits; sp += 4) {
      const float4 a0 = *reinterpret_cast<const float4*>(src + (sp + 0) * slab);
      const float4 a1 = *reinterpret_cast<const float4*>(src + (sp + 1) * slab);
      const float4 a2 = *reinterpret_cast<const float4*>(src + (sp + 2) * slab);
      const float4 a3 = *reinterpret_cast<const float4*>(src + (sp + 3) * slab);
      acc.x += (a0.x + a1.x) + (a2.x + a3.x);
      acc.y += (a0.y + a1.y) + (a2.y + a3.y);
      acc.z += (a0.z + a1.z) + (a2.z + a3.z);
      acc.w += (a0.w + a1.w) + (a2.w + a3.w);
    }
    for (; sp < splits; ++sp) {
      const float4 a = *reinterpret_cast<const float4*>(src + sp * slab);
      acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
    }
  }
  s_t[(c4 + 0) * taps + tap] = acc.x;
  s_t[(c4 + 1) * taps + tap] = acc.y;
  s_t[(c4 + 2) * taps + tap] = acc.z;
  s_t[(c4 + 3) * taps + tap] = acc.w;
  __syncthreads();
  float* g = grad + (static_cast<long long>(co) * cin + ci0) * taps;
  const int n = nci * taps;
  for (int j = threadIdx.x; j < n; j += blockDim.x) g[j] = (accumulate ? g[j] : 0.f) + alpha * s_t[j];
}

// Pack one conv weight [cout][cin][taps] fp32 into the K-major bf16 planes of BOTH GEMM orientations in one pass
// (either destination may be null):   fwd [pl][co][tap][ci]     dgrad [pl][ci][taps-1-tap][co]
// Tile = 32 cin x 32 cout (grid (cin/32, cout/32) for the single-weight kernel), block 256; all loads of a thread are
// issued before the first dependent shared-memory store (a load-store loop paid the memory latency nine times).
template <int TAPS>
__device__ __forceinline__ void pack_conv_weight_tile(const float* __restrict__ w, __nv_bfloat16* __restrict__ fwd,
                                                      __nv_bfloat16* __restrict__ dgrad, int cout, int cin, int nsplit,
                                                      int ci0, int co0, float* s_w) {
  constexpr int taps = TAPS;
  constexpr int row = 32 * TAPS + 1;
  constexpr int per_row = 32 * TAPS, tot = 32 * per_row;
  constexpr int NL = (tot + 255) / 256;   // 36 (3x3) or 4 (1x1)
  {
    float v[NL];
#pragma unroll
    for (int u = 0; u < NL; ++u) {
      const int i = threadIdx.x + u * 256;
      const int r = i / per_row, o = i - r * per_row;
      v[u] = (i < tot) ? __ldg(w + (static_cast<long long>(co0 + r) * cin + ci0) * taps + o) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < NL; ++u) {
      const int i = threadIdx.x + u * 256;
      const int r = i / per_row, o = i - r * per_row;
      if (i < tot) s_w[r * row + o] = v[u];
    }
  }
  __syncthreads();
  const long long total = static_cast<long long>(cout) * cin * taps;
  if (fwd) {
    for (int idx = threadIdx.x; idx < 32 * taps * 4; idx += 256) {
      const int qd = idx & 3, rt = idx >> 2;
      const int r = rt / taps, tap = rt - r * taps;
      float v[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] = s_w[r * row + (qd * 8 + e) * taps + tap];
      __nv_bfloat16* d = fwd + (static_cast<long long>(co0 + r) * taps + tap) * cin + ci0 + qd * 8;
      for (int pl = 0; pl < nsplit; ++pl) {
        *reinterpret_cast<uint4*>(d + pl * total) = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]),
                                                               pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] -= bf16_round(v[e]);
      }
    }
  }
  if (dgrad) {
    for (int idx = threadIdx.x; idx < 32 * taps * 4; idx += 256) {
      const int qd = idx & 3, rt = idx >> 2;
      const int ci = rt / taps, tap = rt - ci * taps;
      float v[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] = s_w[(qd * 8 + e) * row + ci * taps + tap];
      __nv_bfloat16* d = dgrad + (static_cast<long long>(ci0 + ci) * taps + (taps - 1 - tap)) * cout + co0 + qd * 8;
      for (int pl = 0; pl < nsplit; ++pl) {
        *reinterpret_cast<uint4*>(d + pl * total) = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]),
                                                               pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] -= bf16_round(v[e]);
      }
    }
  }
}

template <int TAPS>
__global__ void __launch_bounds__(256)
pack_conv_weight_kernel_v3(const float* __restrict__ w, __nv_bfloat16* __restrict__ fwd,
                           __nv_bfloat16* __restrict__ dgrad, int cout, int cin, int nsplit) {
  extern __shared__ float s_w[];  // [32 co][32*TAPS + 1]
  pack_conv_weight_tile<TAPS>(w, fwd, dgrad, cout, cin, nsplit, blockIdx.x * 32, blockIdx.y * 32, s_w);
}

// Every conv weight of the network in ONE launch (a training step re-packs all of them after the optimizer update:
// sixteen launches of 2-10 us each otherwise, most of them too small to fill the GPU).  Item i owns the blocks
// [first[i], first[i + 1]) = its (cin/32) x (cout/32) tiles; taps is 9 or 1.
struct PackItems {
  static constexpr int kMax = 24;
  int n;
  int first[kMax + 1];
  const float* w[kMax];
  __nv_bfloat16* fwd[kMax];
  __nv_bfloat16* dgrad[kMax];
  int cout[kMax], cin[kMax], taps[kMax];
};
__global__ void __launch_bounds__(256)
pack_conv_weights_multi_kernel(const __grid_constant__ PackItems it, int nsplit) {
  extern __shared__ float s_w[];  // [32 co][32*9 + 1]
  int i = 0;
  while (i + 1 < it.n && static_cast<int>(blockIdx.x) >= it.first[i + 1]) ++i;
  const int t = blockIdx.x - it.first[i];
  const int tiles_x = it.cin[i] / 32;
  const int ci0 = (t % tiles_x) * 32, co0 = (t / tiles_x) * 32;
  if (it.taps[i] == 9) pack_conv_weight_tile<9>(it.w[i], it.fwd[i], it.dgrad[i], it.cout[i], it.cin[i], nsplit, ci0, co0, s_w);
  else pack_conv_weight_tile<1>(it.w[i], it.fwd[i], it.dgrad[i], it.cout[i], it.cin[i], nsplit, ci0, co0, s_w);
}

// im2col of the 3-channel image (second generation): out[pl][n][y][x][k], k = ci*9 + r*3 + s for
// k < 27, zero for 27 <= k < 64.  One block = one image row segment of 128 pixels: the 3 x 3 x 130
// input window is staged in shared memory with coalesced loads; eight threads cooperate on one
// pixel so that every warp store covers 512 contiguous bytes.
__global__ void __launch_bounds__(256)
im2col_first_kernel_v2(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, long long out_plane, int nsplit,
                       int N, int H, int W) {
  __shared__ float s_in[3][3][132];
  const int x0 = blockIdx.x * 128, y = blockIdx.y, n = blockIdx.z;
  for (int i = threadIdx.x; i < 9 * 130; i += 256) {
    const int c = i % 130, rr = (i / 130) % 3, ci = i / 390;
    const int yy = y + rr - 1, xx = x0 + c - 1;
    s_in[ci][rr][c] = (yy >= 0 && yy < H && xx >= 0 && xx < W)
                          ? __ldg(x + ((static_cast<long long>(n) * 3 + ci) * H + yy) * W + xx)
                          : 0.f;
  }
  __syncthreads();
  const int kg = threadIdx.x & 7;
#pragma unroll 1
  for (int pp = threadIdx.x >> 3; pp < 128; pp += 32) {
    const int xq = x0 + pp;
    if (xq >= W) break;
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int k = kg * 8 + e;
      float val = 0.f;
      if (k < 27) {
        const int ci = k / 9, r = (k % 9) / 3, sx = k % 3;
        val = s_in[ci][r][pp + sx];
      }
      v[e] = val;
    }
    __nv_bfloat16* o = out + ((static_cast<long long>(n) * H + y) * W + xq) * 64 + kg * 8;
    for (int pl = 0; pl < nsplit; ++pl) {
      uint4 q;
      q.x = pack_bf16x2(v[0], v[1]);
      q.y = pack_bf16x2(v[2], v[3]);
      q.z = pack_bf16x2(v[4], v[5]);
      q.w = pack_bf16x2(v[6], v[7]);
      *reinterpret_cast<uint4*>(o + pl * out_plane) = q;
      if (pl + 1 < nsplit) {
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] -= bf16_round(v[e]);
      }
    }
  }
}

// Segment table of one image row for K1 (built on the host, passed by value): the 12 x-windows of
// the scales 1,2,3,6 overlap when W % S != 0, so the row is cut at every window boundary into
// disjoint segments and `mask[s]` says which windows contain segment s.
struct PoolSegments {
  int nseg;
  int bound[26];
  unsigned short mask[26];
};

// K1 stage A: persistent CTAs stream image rows through shared memory with TMA
// (two 256-channel x kPoolChunk-pixel boxes per stage, kPoolStages stages), so ~160 KB per SM are in
// flight at no register cost -- register-staged versions were latency-bound at 1.1 TB/s
// (31 % warps active, long-scoreboard stalls; profiles/r01_k1_pool_rows_v3.txt).  Rows a point of the
// (unaligned) channel window cannot be loaded by TMA fall back to ctx_pool_rows_kernel (can_simple.cuh).
//   warp 16      producer: runs ahead, throttled by the per-stage "empty" barriers
//   warps 0-7    consumer set 0: even stages;   warps 8-15  consumer set 1: odd stages
// In a consumer set thread t owns channels 2t, 2t+1: straight-line code over the 32 staged pixels
// (independent 4-byte shared-memory reads, two running sums) that flushes the running sums into the
// windows containing the current segment wherever a window boundary falls.  At the end of a row set 1
// hands its 12 window sums to set 0 through shared memory (double-buffered, one named barrier per row).
//   tm: bf16 map {512 ch, W, N*H, nsplit, 1}, box {64, kPoolChunk, 1, 1, 1}, 128B swizzle: a stage is eight
//       [32 pixels][64 channels] tiles of 4 KB, 16-byte chunk c of pixel p stored at chunk c ^ (p & 7) -- the box shape
//       every other kernel of this library uses
//   R[n][y][12][512] fp32
constexpr int kPoolChunk = 32;    // pixels per stage (32 KB)
constexpr int kPoolStages = 5;
constexpr int kPoolXchgBytes = 2 * 12 * 512 * 4;   // set 1 -> set 0 hand-over, two rows
__host__ __device__ inline int pool_rows_smem_bytes(int W) {
  return kPoolStages * kPoolChunk * 1024 + kPoolXchgBytes + 8 * ((W + kPoolChunk - 1) / kPoolChunk) + 1024;
}
__global__ void __launch_bounds__(544, 1)
ctx_pool_rows_kernel_v3(const __grid_constant__ CUtensorMap tm, int nsplit, float* __restrict__ R, int NH, int W,
                        const __grid_constant__ PoolSegments seg) {
  extern __shared__ __align__(128) unsigned char s_raw[];
  constexpr int kChunk = kPoolChunk, kStages = kPoolStages;
  unsigned char* s_al = s_raw + ((1024u - (smem_u32(s_raw) & 1023u)) & 1023u);   // swizzled TMA tiles: 1024-byte aligned
  __nv_bfloat16* buf = reinterpret_cast<__nv_bfloat16*>(s_al);  // [kStages][2 halves][kChunk][256]
  float* s_xchg = reinterpret_cast<float*>(s_al + kStages * kChunk * 1024);                     // [2][12][512]
  uint32_t* s_endmask = reinterpret_cast<uint32_t*>(s_al + kStages * kChunk * 1024 + kPoolXchgBytes);  // [chunks]
  const int chunks_per_row = (W + kChunk - 1) / kChunk;
  uint32_t* s_segstart = s_endmask + chunks_per_row;                                            // [chunks]
  __shared__ uint64_t full_bar[kStages], empty_bar[kStages];
  __shared__ unsigned short s_segmask[26];
  if (threadIdx.x < 26) s_segmask[threadIdx.x] = threadIdx.x < seg.nseg ? seg.mask[threadIdx.x] : 0;
  // bit p of s_endmask[c]: a segment ends after pixel c * kChunk + p;  s_segstart[c]: segment of its first pixel
  for (int c = threadIdx.x; c < chunks_per_row; c += blockDim.x) {
    uint32_t m = 0;
    int first = 0;
    for (int b = 1; b <= seg.nseg; ++b) {
      const int e = seg.bound[b] - 1 - c * kChunk;
      if (e >= 0 && e < kChunk) m |= 1u << e;
      if (b < seg.nseg && seg.bound[b] <= c * kChunk) first = b;
    }
    s_endmask[c] = m;
    s_segstart[c] = first;
  }
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tm);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 8);
    }
    fence_mbar_init();
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5;
  // A CTA takes a CONTIGUOUS block of rows, so that neighbouring loads share 2 MB pages.  With rows b, b + 148, ... per
  // CTA almost every TMA load landed in another page, and on a large feature map (>= ~0.5 GiB) whose pages no kernel
  // had READ yet the first launch ended in an "unspecified launch failure" about once per 30 000 stages -- with this
  // assignment never in 36 fresh-process runs (profiles/r02_k1_cold_page_fault.txt; cause below the CUDA level not
  // found).  Every row still gets its own record, so results do not depend on the assignment.
  const int rows_per_cta = (NH + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int row_beg = min(static_cast<int>(blockIdx.x) * rows_per_cta, NH), row_end = min(row_beg + rows_per_cta, NH);
  if (warp == 16) {
    // ---- producer
    if (threadIdx.x == 512) {
      int sg = 0;
      uint32_t ph = 0;
      for (int rowi = row_beg; rowi < row_end; ++rowi)
        for (int pl = 0; pl < nsplit; ++pl)
          for (int ch = 0; ch < chunks_per_row; ++ch) {
            mbar_wait(&empty_bar[sg], ph ^ 1u);
            __nv_bfloat16* dst = buf + sg * kChunk * 512;
            mbar_expect_tx(&full_bar[sg], kChunk * 1024u);
#pragma unroll
            for (int g = 0; g < 8; ++g)   // eight 64-channel groups of 32 pixels
              tma_load_5d(dst + g * kChunk * 64, &tm, &full_bar[sg], g * 64, ch * kChunk, rowi, pl, 0);
            if (++sg == kStages) {
              sg = 0;
              ph ^= 1u;
            }
          }
    }
    return;
  }
  // ---- consumers
  const int set = warp >> 3;                  // 0 or 1
  const int tid = threadIdx.x & 255;
  const int t2 = tid * 2;
  // channels 2 tid, 2 tid + 1: group tid >> 5, 16-byte chunk (tid & 31) >> 2 of the pixel's 128-byte row, 4 bytes inside it
  const int soff = (tid >> 5) * kChunk * 64 + (tid & 3) * 2;   // (in bf16 elements; the chunk is XOR-ed with the pixel below)
  const int c16 = (tid & 31) >> 2;
  const int lane = threadIdx.x & 31;
  float acc[12][2];
  float cur0 = 0.f, cur1 = 0.f;   // running sums of segment sgi (may span stages)
  int sgi = 0;
  auto flush = [&]() {
    const unsigned m = s_segmask[sgi];
#pragma unroll
    for (int k = 0; k < 12; ++k) {
      if (m & (1u << k)) {
        acc[k][0] += cur0;
        acc[k][1] += cur1;
      }
    }
    cur0 = cur1 = 0.f;
  };
  int sg = 0, it = 0, rowpar = 0;
  uint32_t ph = 0;
  for (int rowi = row_beg; rowi < row_end; ++rowi, rowpar ^= 1) {
#pragma unroll
    for (int k = 0; k < 12; ++k) acc[k][0] = acc[k][1] = 0.f;
    for (int pl = 0; pl < nsplit; ++pl)
      for (int ch = 0; ch < chunks_per_row; ++ch, ++it) {
        if ((it & 1) == set) {
          const int first = static_cast<int>(s_segstart[ch]);
          if (first != sgi) {   // the other set took the stages in between (or a new row / plane began)
            flush();
            sgi = first;
          }
          const uint32_t em = s_endmask[ch];
          mbar_wait(&full_bar[sg], ph);
          const __nv_bfloat16* b = buf + sg * kChunk * 512 + soff;
          // pixels past W are zero-filled by TMA
#pragma unroll
          for (int g = 0; g < kChunk / 8; ++g) {
            __nv_bfloat162 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u)   // pixel g * 8 + u: (pixel & 7) == u
              v[u] = *reinterpret_cast<const __nv_bfloat162*>(b + (g * 8 + u) * 64 + ((c16 ^ u) << 3));
            if ((em >> (g * 8)) & 0xffu) {
#pragma unroll
              for (int u = 0; u < 8; ++u) {
                cur0 += __low2float(v[u]);
                cur1 += __high2float(v[u]);
                if (em & (1u << (g * 8 + u))) {
                  flush();
                  ++sgi;
                }
              }
            } else {   // no boundary inside these 8 pixels
              float a0 = 0.f, a1 = 0.f;
#pragma unroll
              for (int u = 0; u < 8; u += 2) {
                cur0 += __low2float(v[u]);
                cur1 += __high2float(v[u]);
                a0 += __low2float(v[u + 1]);
                a1 += __high2float(v[u + 1]);
              }
              cur0 += a0;
              cur1 += a1;
            }
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(&empty_bar[sg]);   // this warp is done with the stage
        }
        if (++sg == kStages) {
          sg = 0;
          ph ^= 1u;
        }
      }
    flush();   // (adds zeros when the row's last stage was this set's: its boundary bit already flushed)
    // row finished: set 1 hands over, set 0 adds and writes the 12 window sums
    float* xc = s_xchg + rowpar * 12 * 512 + t2;
    if (set == 1) {
#pragma unroll
      for (int k = 0; k < 12; ++k) *reinterpret_cast<float2*>(xc + k * 512) = make_float2(acc[k][0], acc[k][1]);
    }
    asm volatile("bar.sync 1, 512;" ::: "memory");
    if (set == 0) {
      float* o = R + static_cast<long long>(rowi) * 12 * 512 + t2;
#pragma unroll
      for (int k = 0; k < 12; ++k) {
        const float2 other = *reinterpret_cast<const float2*>(xc + k * 512);
        *reinterpret_cast<float2*>(o + k * 512) = make_float2(acc[k][0] + other.x, acc[k][1] + other.y);
      }
    }
  }
}

}  // namespace canb
