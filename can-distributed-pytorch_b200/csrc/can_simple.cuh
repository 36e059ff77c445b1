// HBM-/latency-bound kernels of the CANNet step: weight packing, the 3->64 first
// convolution, the multi-scale pooling pass (K1) and its tiny 1x1 tables, the fused
// MSE-sum + per-image count reduction, fused SGD-momentum, and small helpers.
#pragma once
#include "can_ptx.cuh"

namespace canb {

// --------------------------------------------------------------------------------
// split-precision helpers: value = plane0 + plane1 (+ plane2), each bf16
// --------------------------------------------------------------------------------
__device__ __forceinline__ void split_store(__nv_bfloat16* dst, long long plane, int nsplit, float v) {
  float acc = v;
  for (int pl = 0; pl < nsplit; ++pl) {
    __nv_bfloat16 h = __float2bfloat16_rn(acc);
    dst[pl * plane] = h;
    acc -= __bfloat162float(h);
  }
}
__device__ __forceinline__ float split_load(const __nv_bfloat16* src, long long plane, int nsplit) {
  float v = 0.f;
  for (int pl = nsplit - 1; pl >= 0; --pl) v += __bfloat162float(src[pl * plane]);
  return v;
}

// ATen align_corners=True source index (UpSample.h:271-295,442-476), fp32 bit-exact.
__device__ __forceinline__ void bilinear_tap(int dst, int in_size, int out_size, int& i0, int& i1, float& l0,
                                                 float& l1) {
  float scale = (out_size > 1) ? __fdiv_rn(static_cast<float>(in_size - 1), static_cast<float>(out_size - 1)) : 0.f;
  float src = __fmul_rn(scale, static_cast<float>(dst));
  i0 = min(static_cast<int>(floorf(src)), in_size - 1);
  l1 = fminf(fmaxf(__fsub_rn(src, static_cast<float>(i0)), 0.f), 1.f);
  i1 = i0 + (i0 < in_size - 1 ? 1 : 0);
  l0 = __fsub_rn(1.f, l1);
}
// --------------------------------------------------------------------------------
// weight packing: OIHW fp32 -> K-major bf16 planes
//   fwd  : dst[pl][co][tap][ci]          = W[co][ci][tap]
//   dgrad: dst[pl][ci][taps-1-tap][co]   = W[co][ci][tap]   (transposed + spatially flipped)
// --------------------------------------------------------------------------------
__global__ void pack_conv_weight_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ dst, int cout,
                                        int cin, int taps, int nsplit, int dgrad) {
  const long long total = static_cast<long long>(cout) * cin * taps;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    // i indexes the destination so that writes coalesce
    long long o;
    int co, ci, tap;
    if (!dgrad) {
      ci = i % cin;
      tap = (i / cin) % taps;
      co = i / (static_cast<long long>(cin) * taps);
      o = i;
    } else {
      co = i % cout;
      const int tflip = (i / cout) % taps;
      ci = i / (static_cast<long long>(cout) * taps);
      tap = taps - 1 - tflip;
      o = i;
    }
    const float v = w[(static_cast<long long>(co) * cin + ci) * taps + tap];
    split_store(dst + o, total, nsplit, v);
  }
}

// stacked contrast weights for the fused contextual GEMM:
//   dst[pl][(cc*4 + j)*64 + c][k] = W_j2[cc*64 + c][k]     (row order: channel chunk, scale, channel)
__global__ void pack_ctx_w2_kernel(const float* __restrict__ w0, const float* __restrict__ w1,
                                   const float* __restrict__ w2, const float* __restrict__ w3,
                                   __nv_bfloat16* __restrict__ dst, int nsplit) {
  const long long total = 2048LL * 512;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int k = i & 511;
    const int row = i >> 9;
    const int c = row & 63, j = (row >> 6) & 3, cc = row >> 8;
    const float* w = j == 0 ? w0 : (j == 1 ? w1 : (j == 2 ? w2 : w3));
    split_store(dst + i, total, nsplit, w[(cc * 64 + c) * 512 + k]);
  }
}

// fp32 transpose of a [rows][cols] matrix (used for W^T tables of the tiny 1x1 convs)
__global__ void transpose_f32_kernel(const float* __restrict__ src, float* __restrict__ dst, int rows, int cols) {
  __shared__ float tile[32][33];
  const int bx = blockIdx.x * 32, by = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int r = by + i, c = bx + threadIdx.x;
    tile[i][threadIdx.x] = (r < rows && c < cols) ? src[static_cast<long long>(r) * cols + c] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = bx + i, r = by + threadIdx.x;
    if (r < rows && c < cols) dst[static_cast<long long>(c) * rows + r] = tile[threadIdx.x][i];
  }
}

struct TransposeItems {
  const float* src[16];
};
// n matrices of one shape in one launch (blockIdx.z = matrix); dst = [n][cols][rows]
__global__ void transpose_f32_multi_kernel(TransposeItems it, float* __restrict__ dst, int rows, int cols) {
  __shared__ float tile[32][33];
  const float* __restrict__ src = it.src[blockIdx.z];
  float* __restrict__ out = dst + static_cast<long long>(blockIdx.z) * rows * cols;
  const int bx = blockIdx.x * 32, by = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int r = by + i, c = bx + threadIdx.x;
    tile[i][threadIdx.x] = (r < rows && c < cols) ? src[static_cast<long long>(r) * cols + c] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = bx + i, r = by + threadIdx.x;
    if (r < rows && c < cols) out[static_cast<long long>(c) * rows + r] = tile[threadIdx.x][i];
  }
}

// --------------------------------------------------------------------------------
// first convolution 3 -> 64, 3x3 pad 1, bias + ReLU.  K = 27 is too thin for the tensor
// pipe; this is a direct fp32 conv bound by the 64-channel bf16 store (128 B / pixel).
// in: NCHW fp32 (the caller's image tensor); out: NHWC bf16 planes
// --------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
conv_first_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                  __nv_bfloat16* __restrict__ out, long long out_plane, int nsplit, int N, int H, int W) {
  __shared__ float sw[27][64];
  __shared__ float sb[64];
  for (int i = threadIdx.x; i < 27 * 64; i += blockDim.x) {
    const int co = i & 63, k = i >> 6;  // k = ci*9 + tap
    sw[k][co] = w[co * 27 + k];
  }
  if (threadIdx.x < 64) sb[threadIdx.x] = b[threadIdx.x];
  __syncthreads();
  const int xq = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y, n = blockIdx.z;
  if (xq >= W) return;
  float in[27];
#pragma unroll
  for (int ci = 0; ci < 3; ++ci)
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int yy = y + r - 1, xx = xq + s - 1;
        in[ci * 9 + r * 3 + s] = (yy >= 0 && yy < H && xx >= 0 && xx < W)
                                     ? __ldg(x + ((static_cast<long long>(n) * 3 + ci) * H + yy) * W + xx)
                                     : 0.f;
      }
  __nv_bfloat16* o = out + ((static_cast<long long>(n) * H + y) * W + xq) * 64;
#pragma unroll 1
  for (int c0 = 0; c0 < 64; c0 += 8) {
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = sb[c0 + e];
#pragma unroll
    for (int k = 0; k < 27; ++k) {
      const float4 wa = *reinterpret_cast<const float4*>(&sw[k][c0]);
      const float4 wb = *reinterpret_cast<const float4*>(&sw[k][c0 + 4]);
      acc[0] = fmaf(in[k], wa.x, acc[0]);
      acc[1] = fmaf(in[k], wa.y, acc[1]);
      acc[2] = fmaf(in[k], wa.z, acc[2]);
      acc[3] = fmaf(in[k], wa.w, acc[3]);
      acc[4] = fmaf(in[k], wb.x, acc[4]);
      acc[5] = fmaf(in[k], wb.y, acc[5]);
      acc[6] = fmaf(in[k], wb.z, acc[6]);
      acc[7] = fmaf(in[k], wb.w, acc[7]);
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = fmaxf(acc[e], 0.f);
    for (int pl = 0; pl < nsplit; ++pl) {
      uint4 q;
      q.x = pack_bf16x2(acc[0], acc[1]);
      q.y = pack_bf16x2(acc[2], acc[3]);
      q.z = pack_bf16x2(acc[4], acc[5]);
      q.w = pack_bf16x2(acc[6], acc[7]);
      *reinterpret_cast<uint4*>(o + pl * out_plane + c0) = q;
      if (pl + 1 < nsplit) {
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] -= bf16_round(acc[e]);
      }
    }
  }
}

// --------------------------------------------------------------------------------
// adaptive-pool window tables (ATen AdaptivePooling.h:32-38), integer exact
// --------------------------------------------------------------------------------
__host__ __device__ inline int adaptive_start(int i, int in_size, int out_size) {
  return (i / out_size) * in_size + ((i % out_size) * in_size) / out_size;
}
__host__ __device__ inline int adaptive_end(int i, int in_size, int out_size) {
  return 1 + ((i + 1) * in_size - 1) / out_size;
}

// K1 stage A: one CTA per (row y, image n); thread t owns channels 2t, 2t+1 of the 512.
// Writes row sums over the 12 x-windows (scales 1,2,3,6): R[n][y][12][512] fp32.
// Reads fv exactly once (HBM/L2-bound).  Windows may overlap (in % S != 0), so the row is
// cut at every window boundary into disjoint segments and each window accumulates the
// segments it covers.
__global__ void __launch_bounds__(256)
ctx_pool_rows_kernel(const __nv_bfloat16* __restrict__ fv, long long fv_plane, int nsplit, int cs, int coff,
                     float* __restrict__ R, int H, int W) {
  __shared__ int s_bound[26];
  __shared__ int s_nseg;
  __shared__ int s_ws[12], s_we[12];
  if (threadIdx.x == 0) {
    const int SS[4] = {1, 2, 3, 6};
    int nb = 0, k = 0;
    for (int j = 0; j < 4; ++j)
      for (int i = 0; i < SS[j]; ++i, ++k) {
        s_ws[k] = adaptive_start(i, W, SS[j]);
        s_we[k] = adaptive_end(i, W, SS[j]);
        s_bound[nb++] = s_ws[k];
        s_bound[nb++] = s_we[k];
      }
    // sort + unique (24 entries)
    for (int a = 1; a < nb; ++a) {
      int v = s_bound[a], b = a - 1;
      while (b >= 0 && s_bound[b] > v) {
        s_bound[b + 1] = s_bound[b];
        --b;
      }
      s_bound[b + 1] = v;
    }
    int u = 1;
    for (int a = 1; a < nb; ++a)
      if (s_bound[a] != s_bound[u - 1]) s_bound[u++] = s_bound[a];
    s_nseg = u - 1;
  }
  __syncthreads();
  const int y = blockIdx.x, n = blockIdx.y;
  const int c = threadIdx.x * 2;
  const __nv_bfloat16* row = fv + (static_cast<long long>(n) * H + y) * W * cs + coff + c;
  float acc0[12], acc1[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) acc0[k] = acc1[k] = 0.f;
  const int nseg = s_nseg;
  for (int sgi = 0; sgi < nseg; ++sgi) {
    const int xs = s_bound[sgi], xe = s_bound[sgi + 1];
    float a0 = 0.f, a1 = 0.f;
    for (int pl = 0; pl < nsplit; ++pl) {
      const __nv_bfloat16* rp = row + pl * fv_plane;
      int x = xs;
      for (; x + 8 <= xe; x += 8) {  // 8 independent loads in flight
        __nv_bfloat162 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u)
          v[u] = *reinterpret_cast<const __nv_bfloat162*>(rp + static_cast<long long>(x + u) * cs);
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          a0 += __low2float(v[u]);
          a1 += __high2float(v[u]);
        }
      }
      for (; x < xe; ++x) {
        const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(rp + static_cast<long long>(x) * cs);
        a0 += __low2float(v);
        a1 += __high2float(v);
      }
    }
#pragma unroll
    for (int k = 0; k < 12; ++k) {
      if (xs >= s_ws[k] && xe <= s_we[k]) {
        acc0[k] += a0;
        acc1[k] += a1;
      }
    }
  }
  float* o = R + ((static_cast<long long>(n) * H + y) * 12) * 512 + c;
#pragma unroll
  for (int k = 0; k < 12; ++k) *reinterpret_cast<float2*>(o + k * 512) = make_float2(acc0[k], acc1[k]);
}

// position p in [0,50): scale index j, window row i, window col l, index of the x-window in the 12-list
__device__ __forceinline__ void ctx_pos_decode(int p, int& j, int& S, int& wy, int& wx, int& xwin) {
  if (p < 1) { j = 0; S = 1; p -= 0; xwin = 0; }
  else if (p < 5) { j = 1; S = 2; p -= 1; xwin = 1; }
  else if (p < 14) { j = 2; S = 3; p -= 5; xwin = 3; }
  else { j = 3; S = 6; p -= 14; xwin = 6; }
  wy = p / S;
  wx = p - wy * S;
  xwin += wx;
}

// K1 stage B: pooled[n][p][c] = sum_{y in ywin(p)} R[n][y][xwin(p)][c] / area   (grid: 50 x N, 512 threads)
__global__ void __launch_bounds__(512)
ctx_pool_finalize_kernel(const float* __restrict__ R, float* __restrict__ pooled, int H, int W) {
  const int p = blockIdx.x, n = blockIdx.y, c = threadIdx.x;
  int j, S, wy, wx, xwin;
  ctx_pos_decode(p, j, S, wy, wx, xwin);
  const int ys = adaptive_start(wy, H, S), ye = adaptive_end(wy, H, S);
  const int xs = adaptive_start(wx, W, S), xe = adaptive_end(wx, W, S);
  float acc = 0.f;
  int y = ys;
  for (; y + 8 <= ye; y += 8) {  // 8 independent loads in flight
    float v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = R[((static_cast<long long>(n) * H + y + u) * 12 + xwin) * 512 + c];
#pragma unroll
    for (int u = 0; u < 8; ++u) acc += v[u];
  }
  for (; y < ye; ++y) acc += R[((static_cast<long long>(n) * H + y) * 12 + xwin) * 512 + c];
  pooled[(static_cast<long long>(n) * 50 + p) * 512 + c] = acc / static_cast<float>((ye - ys) * (xe - xs));
}

// --------------------------------------------------------------------------------
// fused MSE(sum) + per-image counts + dL/det   (train_eval_utils.py:20,37,83)
// one CTA per image; ws = [N*3 partials | 1 counter]; the last CTA to finish adds the
// per-image losses in index order, so the scalar is bit-reproducible run to run.
// out: loss[1], cnt_et[N], cnt_gt[N], grad[N*HW] (optional) = 2 (et - gt) * grad_scale
// --------------------------------------------------------------------------------
__device__ __forceinline__ float block_reduce_sum(float v, float* sbuf) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) sbuf[w] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  v = (threadIdx.x < nw) ? sbuf[threadIdx.x] : 0.f;
  if (w == 0) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  }
  return v;  // valid in thread 0
}

__global__ void __launch_bounds__(1024)
mse_count_kernel(const float* __restrict__ et, const float* __restrict__ gt, int N, int HW, float* __restrict__ loss,
                 float* __restrict__ cnt_et, float* __restrict__ cnt_gt, float* __restrict__ grad,
                 float grad_scale, float* __restrict__ ws_partial, unsigned int* __restrict__ ws_counter) {
  __shared__ float sbuf[32];
  __shared__ bool s_last;
  const int n = blockIdx.x;
  const float* e = et + static_cast<long long>(n) * HW;
  const float* g = gt + static_cast<long long>(n) * HW;
  float sl = 0.f, se = 0.f, sg = 0.f;
  for (int i = threadIdx.x; i < HW; i += blockDim.x) {
    const float a = e[i], b = g[i];
    const float d = a - b;
    sl = fmaf(d, d, sl);
    se += a;
    sg += b;
    if (grad) grad[static_cast<long long>(n) * HW + i] = 2.f * d * grad_scale;
  }
  sl = block_reduce_sum(sl, sbuf);
  se = block_reduce_sum(se, sbuf);
  sg = block_reduce_sum(sg, sbuf);
  if (threadIdx.x == 0) {
    ws_partial[n] = sl;
    cnt_et[n] = se;
    cnt_gt[n] = sg;
    __threadfence();
    const unsigned int done = atomicAdd(ws_counter, 1u);
    s_last = (done == static_cast<unsigned int>(N - 1));
  }
  __syncthreads();
  if (s_last && threadIdx.x == 0) {
    __threadfence();
    float tot = 0.f;
    for (int i = 0; i < N; ++i) tot += reinterpret_cast<volatile float*>(ws_partial)[i];
    loss[0] = tot;
    *ws_counter = 0u;  // ready for the next launch on this stream
  }
}

// --------------------------------------------------------------------------------
// fused SGD with momentum over a flat fp32 arena (torch.optim.SGD: dampening 0, no
// nesterov, weight_decay 0; train.py:63,125-126).  first != 0: buf = g.
// grad_scale folds the data-parallel mean (1/world) into the update.
// --------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
sgd_momentum_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ buf, long long n,
                    float lr, float momentum, float grad_scale, int first) {
  const long long n4 = n >> 2;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4; i += stride) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    const float4 gv = reinterpret_cast<const float4*>(g)[i];
    float4 bv;
    if (first) {
      bv = make_float4(gv.x * grad_scale, gv.y * grad_scale, gv.z * grad_scale, gv.w * grad_scale);
    } else {
      bv = reinterpret_cast<float4*>(buf)[i];
      bv.x = fmaf(momentum, bv.x, gv.x * grad_scale);
      bv.y = fmaf(momentum, bv.y, gv.y * grad_scale);
      bv.z = fmaf(momentum, bv.z, gv.z * grad_scale);
      bv.w = fmaf(momentum, bv.w, gv.w * grad_scale);
    }
    pv.x = fmaf(-lr, bv.x, pv.x);
    pv.y = fmaf(-lr, bv.y, pv.y);
    pv.z = fmaf(-lr, bv.z, pv.z);
    pv.w = fmaf(-lr, bv.w, pv.w);
    reinterpret_cast<float4*>(buf)[i] = bv;
    reinterpret_cast<float4*>(p)[i] = pv;
  }
  // tail
  for (long long i = (n4 << 2) + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += stride) {
    const float gs = g[i] * grad_scale;
    const float b = first ? gs : fmaf(momentum, buf[i], gs);
    buf[i] = b;
    p[i] = fmaf(-lr, b, p[i]);
  }
}

// Zero up to kMax slices of one fp32 buffer in one launch (the bias / head gradient slices of the flat gradient arena:
// they are accumulated with atomics, everything else in the arena is overwritten every step, so an 83 MB memset per
// step would write 82.9 MB too many).  blockIdx.x = slice.
struct ZeroRanges {
  static constexpr int kMax = 32;
  int n;
  long long off[kMax];
  int count[kMax];
};
__global__ void __launch_bounds__(256)
zero_ranges_kernel(float* __restrict__ base, const __grid_constant__ ZeroRanges r) {
  float* p = base + r.off[blockIdx.x];
  for (int i = threadIdx.x; i < r.count[blockIdx.x]; i += blockDim.x) p[i] = 0.f;
}

// --------------------------------------------------------------------------------
// backward helpers
// --------------------------------------------------------------------------------
// im2col of the 3-channel image for the first conv's weight gradient (and, optionally, its
// forward on the tensor cores): out[pl][n][y][x][k], k = ci*9 + r*3 + s for k < 27, zero for
// 27 <= k < 64.  With it, conv1_1 is a 1x1 GEMM with K = 64.
__global__ void __launch_bounds__(256)
im2col_first_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, long long out_plane, int nsplit,
                    int N, int H, int W) {
  const long long total = static_cast<long long>(N) * H * W * 8;  // 8 threads per pixel, 8 k each
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int kg = i & 7;
    const long long pix = i >> 3;
    const int xq = pix % W;
    const int y = (pix / W) % H;
    const int n = pix / (static_cast<long long>(W) * H);
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int k = kg * 8 + e;
      float val = 0.f;
      if (k < 27) {
        const int ci = k / 9, r = (k % 9) / 3, sx = k % 3;
        const int yy = y + r - 1, xx = xq + sx - 1;
        if (yy >= 0 && yy < H && xx >= 0 && xx < W) val = __ldg(x + ((static_cast<long long>(n) * 3 + ci) * H + yy) * W + xx);
      }
      v[e] = val;
    }
    __nv_bfloat16* o = out + pix * 64 + kg * 8;
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

// db[c] += sum over pixels of dZ[pix][c]  (dZ: NHWC planes).  blockDim = 256; C in {64..1024}.
// Each block owns a contiguous pixel range; per-channel partials are combined in shared memory
// and added to db with one atomic per channel per block.
__global__ void __launch_bounds__(256)
bias_grad_kernel(const __nv_bfloat16* __restrict__ dz, long long plane, int nsplit, int cs, int coff, int C,
                 long long npix, float* __restrict__ db) {
  __shared__ float s_acc[1024];
  for (int i = threadIdx.x; i < C; i += blockDim.x) s_acc[i] = 0.f;
  __syncthreads();
  const int pairs = C >> 1;                     // bf16x2 per thread
  const int lanes_c = pairs < 256 ? pairs : 256;  // threads along channels
  const int rows = 256 / lanes_c;               // pixel rows processed concurrently
  const int tc = threadIdx.x % lanes_c, tr = threadIdx.x / lanes_c;
  const long long per = (npix + gridDim.x - 1) / gridDim.x;
  const long long p0 = blockIdx.x * per, p1 = min(p0 + per, npix);
  for (int c2 = tc; c2 < pairs; c2 += lanes_c) {
    float a0 = 0.f, a1 = 0.f;
    for (long long px = p0 + tr; px < p1; px += rows) {
      for (int pl = 0; pl < nsplit; ++pl) {
        const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(dz + pl * plane + px * cs + coff + c2 * 2);
        a0 += __low2float(v);
        a1 += __high2float(v);
      }
    }
    atomicAdd(&s_acc[c2 * 2], a0);
    atomicAdd(&s_acc[c2 * 2 + 1], a1);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += blockDim.x) atomicAdd(db + i, s_acc[i]);
}

// backward of output_layer (1x1, 64 -> 1; CANNet.py:17,90) fused with the ReLU mask of the last
// backend conv:  dZ5[pix][c] = dOut[pix] * hw[c] * (y5[pix][c] > 0)
//                d_hw[c]    += sum_pix dOut[pix] * y5[pix][c];   d_hb += sum_pix dOut[pix]
// 256 threads = 8 pixels x 32 channel pairs per step.
__global__ void __launch_bounds__(256)
head_bwd_kernel(const float* __restrict__ dout, const __nv_bfloat16* __restrict__ y5, long long y_plane, int nsplit,
                const float* __restrict__ hw, __nv_bfloat16* __restrict__ dz, long long dz_plane, long long npix,
                float* __restrict__ d_hw, float* __restrict__ d_hb) {
  __shared__ float s_hw[64];
  __shared__ float s_acc[65];
  if (threadIdx.x < 64) s_hw[threadIdx.x] = hw[threadIdx.x];
  if (threadIdx.x < 65) s_acc[threadIdx.x] = 0.f;
  __syncthreads();
  const int c2 = threadIdx.x & 31, tr = threadIdx.x >> 5;
  const long long per = (npix + gridDim.x - 1) / gridDim.x;
  const long long p0 = blockIdx.x * per, p1 = min(p0 + per, npix);
  float a0 = 0.f, a1 = 0.f, ab = 0.f;
  const float w0 = s_hw[c2 * 2], w1 = s_hw[c2 * 2 + 1];
  for (long long px = p0 + tr; px < p1; px += 8) {
    const float g = dout[px];
    float y0 = 0.f, y1 = 0.f;
    for (int pl = nsplit - 1; pl >= 0; --pl) {
      const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(y5 + pl * y_plane + px * 64 + c2 * 2);
      y0 += __low2float(v);
      y1 += __high2float(v);
    }
    a0 = fmaf(g, y0, a0);
    a1 = fmaf(g, y1, a1);
    if (c2 == 0) ab += g;
    float z0 = y0 > 0.f ? g * w0 : 0.f, z1 = y1 > 0.f ? g * w1 : 0.f;
    for (int pl = 0; pl < nsplit; ++pl) {
      const __nv_bfloat162 q = __floats2bfloat162_rn(z0, z1);
      *reinterpret_cast<__nv_bfloat162*>(dz + pl * dz_plane + px * 64 + c2 * 2) = q;
      z0 -= __low2float(q);
      z1 -= __high2float(q);
    }
  }
  atomicAdd(&s_acc[c2 * 2], a0);
  atomicAdd(&s_acc[c2 * 2 + 1], a1);
  if (c2 == 0) atomicAdd(&s_acc[64], ab);
  __syncthreads();
  if (threadIdx.x < 64) atomicAdd(d_hw + threadIdx.x, s_acc[threadIdx.x]);
  if (threadIdx.x == 64) atomicAdd(d_hb, s_acc[64]);
}

// --------------------------------------------------------------------------------
// contextual-module tables: tiny per-position 1x1 convs and their gradients
// --------------------------------------------------------------------------------
// out[n][OFF_j + p][r] = (add ? add(n, OFF_j + p, r) : 0) + sum_k Wm[j][r][k] * in(n, OFF_j + p, k)
//   Wm : [4][512 rows][512] row-major (conv weights as stored -> forward; their transposes -> backward)
//   in(n, pos, k) = in[n*in_ns + pos*in_ps + j*in_js + k]   (same addressing for `add`)
// The positions of ALL images of a scale are the rows of one [N * S^2 x 512] x [512 x 512]^T product, so a weight slab
// is read once per PT positions instead of once per image (batch 16: a per-image GEMV re-read 268 MB from L2 and took
// 37 us per launch).  Every output is summed in k order whatever the batch size, so an image's tables do not depend on
// the batch it is part of.  Tile = PT positions x 64 output rows, K in steps of 32 through shared
// memory (k-major, padded), 256 threads x (PT/16 positions x 4 rows), fp32 FMA in k order.  PT = 64 for large batches:
// with 2 positions per thread the kernel is bound by its own shared-memory reads (one 16-byte read per 4 FMAs).
// grid: sum_j ceil(N S_j^2 / PT) * 8 tiles, scale-major.
template <int PT>
__global__ void __launch_bounds__(256)
ctx_pos_gemm_kernel(const float* __restrict__ Wm, const float* __restrict__ in, long long in_ns, long long in_ps,
                    long long in_js, const float* __restrict__ add, float* __restrict__ out, int N) {
  constexpr int RT = 64, KT = 32, PP = PT / 16;   // PP positions per thread
  __shared__ __align__(16) float s_a[KT][PT + 4];
  __shared__ __align__(16) float s_w[KT][RT + 4];
  const int SS[4] = {1, 2, 3, 6};
  const int OFF[4] = {0, 1, 5, 14};
  int t = blockIdx.x, j = 0, S2 = 1, M = N;
  for (; j < 4; ++j) {
    S2 = SS[j] * SS[j];
    M = N * S2;
    const int tiles = ((M + PT - 1) / PT) * (512 / RT);
    if (t < tiles) break;
    t -= tiles;
  }
  if (j == 4) return;
  const int rt = t % (512 / RT), mt = t / (512 / RT);
  const int m0 = mt * PT, r0 = rt * RT;
  const int tid = threadIdx.x;
  // loaders: 8 threads per row of 32 k (one float4 each); 32 rows per pass
  const int lr = tid >> 3, lk = (tid & 7) * 4;
  const float* a_src[PT / 32];
  bool a_ok[PT / 32];
#pragma unroll
  for (int u = 0; u < PT / 32; ++u) {
    const int am = m0 + lr + 32 * u;
    a_ok[u] = am < M;
    a_src[u] = in + (a_ok[u] ? (am / S2) * in_ns + (OFF[j] + am % S2) * in_ps : 0) + j * in_js + lk;
  }
  const float* w_src0 = Wm + (static_cast<long long>(j) * 512 + r0 + lr) * 512 + lk;   // rows lr and lr + 32
  const float* w_src1 = w_src0 + 32 * 512;
  const int tx = tid & 15, ty = tid >> 4;               // rows 4 tx .. 4 tx + 3, positions PP ty .. PP ty + PP - 1
  float acc[PP][4];
#pragma unroll
  for (int i = 0; i < PP; ++i)
#pragma unroll
    for (int e = 0; e < 4; ++e) acc[i][e] = 0.f;
  float4 av[PT / 32];
#pragma unroll
  for (int u = 0; u < PT / 32; ++u)
    av[u] = a_ok[u] ? __ldg(reinterpret_cast<const float4*>(a_src[u])) : make_float4(0.f, 0.f, 0.f, 0.f);
  float4 w0 = __ldg(reinterpret_cast<const float4*>(w_src0)), w1 = __ldg(reinterpret_cast<const float4*>(w_src1));
  for (int k0 = 0; k0 < 512; k0 += KT) {
    __syncthreads();   // the previous step's reads are done
#pragma unroll
    for (int u = 0; u < PT / 32; ++u) {
      s_a[lk][lr + 32 * u] = av[u].x; s_a[lk + 1][lr + 32 * u] = av[u].y;
      s_a[lk + 2][lr + 32 * u] = av[u].z; s_a[lk + 3][lr + 32 * u] = av[u].w;
    }
    s_w[lk][lr] = w0.x; s_w[lk + 1][lr] = w0.y; s_w[lk + 2][lr] = w0.z; s_w[lk + 3][lr] = w0.w;
    s_w[lk][lr + 32] = w1.x; s_w[lk + 1][lr + 32] = w1.y; s_w[lk + 2][lr + 32] = w1.z; s_w[lk + 3][lr + 32] = w1.w;
    __syncthreads();
    if (k0 + KT < 512) {   // next step's operands in flight while this one is multiplied
#pragma unroll
      for (int u = 0; u < PT / 32; ++u)
        if (a_ok[u]) av[u] = __ldg(reinterpret_cast<const float4*>(a_src[u] + k0 + KT));
      w0 = __ldg(reinterpret_cast<const float4*>(w_src0 + k0 + KT));
      w1 = __ldg(reinterpret_cast<const float4*>(w_src1 + k0 + KT));
    }
#pragma unroll
    for (int kk = 0; kk < KT; ++kk) {
      float a[PP];
      if (PP == 4) {
        const float4 a4 = *reinterpret_cast<const float4*>(&s_a[kk][4 * ty]);
        a[0] = a4.x; a[1] = a4.y; a[PP - 2] = a4.z; a[PP - 1] = a4.w;
      } else {
        const float2 a2 = *reinterpret_cast<const float2*>(&s_a[kk][2 * ty]);
        a[0] = a2.x; a[1] = a2.y;
      }
      const float4 w4 = *reinterpret_cast<const float4*>(&s_w[kk][4 * tx]);
#pragma unroll
      for (int i = 0; i < PP; ++i) {
        acc[i][0] = fmaf(a[i], w4.x, acc[i][0]); acc[i][1] = fmaf(a[i], w4.y, acc[i][1]);
        acc[i][2] = fmaf(a[i], w4.z, acc[i][2]); acc[i][3] = fmaf(a[i], w4.w, acc[i][3]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < PP; ++i) {
    const int m = m0 + PP * ty + i;
    if (m < M) {
      const int n = m / S2, pos = OFF[j] + m % S2;
      float4 o = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
      if (add) {
        const float4 ad = *reinterpret_cast<const float4*>(add + n * in_ns + pos * in_ps + j * in_js + r0 + 4 * tx);
        o.x += ad.x; o.y += ad.y; o.z += ad.z; o.w += ad.w;
      }
      *reinterpret_cast<float4*>(out + (static_cast<long long>(n) * 50 + pos) * 512 + r0 + 4 * tx) = o;
    }
  }
}

// dW[j][r][c] (+)= alpha * sum_{n, p in scale j} A(n, OFF_j + p, r) * B[n][OFF_j + p][c]
//   A(n, pos, r) = A[n*a_ns + pos*a_ps + j*a_js + r];  B: [N][50][512]
// grid (32 row blocks of 16, 4 scales); 256 threads, 2 columns each.
__global__ void __launch_bounds__(256)
ctx_outer_kernel(const float* __restrict__ A, long long a_ns, long long a_ps, long long a_js,
                 const float* __restrict__ B, int N, float* __restrict__ dW, float alpha, int accumulate) {
  __shared__ float s_a[36][16];
  const int SS[4] = {1, 2, 3, 6};
  const int OFF[4] = {0, 1, 5, 14};
  const int j = blockIdx.y;
  const int S2 = SS[j] * SS[j];
  const int r0 = blockIdx.x * 16;
  const int c = threadIdx.x;
  float acc0[16], acc1[16];
#pragma unroll
  for (int r = 0; r < 16; ++r) acc0[r] = acc1[r] = 0.f;
  for (int n = 0; n < N; ++n) {
    __syncthreads();
    for (int i = threadIdx.x; i < S2 * 16; i += 256)
      s_a[i >> 4][i & 15] = A[n * a_ns + (OFF[j] + (i >> 4)) * a_ps + j * a_js + r0 + (i & 15)];
    __syncthreads();
    const float* bp = B + (static_cast<long long>(n) * 50 + OFF[j]) * 512;
    constexpr int kU = 12;   // positions of loads in flight (the loop is bound by the L2 latency of the B rows)
    for (int p = 0; p < S2; p += kU) {
      float b0[kU], b1[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const bool on = (p + u) < S2;
        b0[u] = on ? bp[(p + u) * 512 + c] : 0.f;
        b1[u] = on ? bp[(p + u) * 512 + c + 256] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        if (p + u >= S2) break;
        const int pp = min(p + u, S2 - 1);
#pragma unroll
        for (int r = 0; r < 16; ++r) {
          const float a = s_a[pp][r];
          acc0[r] = fmaf(a, b0[u], acc0[r]);
          acc1[r] = fmaf(a, b1[u], acc1[r]);
        }
      }
    }
  }
#pragma unroll
  for (int r = 0; r < 16; ++r) {
    float* d = dW + (static_cast<long long>(j) * 512 + r0 + r) * 512;
    d[c] = (accumulate ? d[c] : 0.f) + alpha * acc0[r];
    d[c + 256] = (accumulate ? d[c + 256] : 0.f) + alpha * acc1[r];
  }
}

// Bilinear interpolation matrix of the four scales as an NHWC "activation":
//   out[pl][n][y][x][pos] = weight of pooled position pos (0..49, scale-major) in pixel (y, x);
//   positions 50..63 are zero.  U^T applied to a full-resolution tensor is then a GEMM with
//   K = pixels on the tensor cores (can_conv_wgrad with this tensor as the "dz" operand).
__global__ void __launch_bounds__(256)
ctx_wint_kernel(__nv_bfloat16* __restrict__ out, long long out_plane, int nsplit, int N, int H, int W) {
  const long long total = static_cast<long long>(N) * H * W * 64;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int pos = i & 63;
    const long long pix = i >> 6;
    const int x = pix % W;
    const int y = (pix / W) % H;
    float wv = 0.f;
    if (pos < 50) {
      int j, S, wy, wx, xwin;
      ctx_pos_decode(pos, j, S, wy, wx, xwin);
      int yi0, yi1, xi0, xi1;
      float ly0, ly1, lx0, lx1;
      bilinear_tap(y, S, H, yi0, yi1, ly0, ly1);
      bilinear_tap(x, S, W, xi0, xi1, lx0, lx1);
      const float wyv = (wy == yi0 ? ly0 : 0.f) + ((wy == yi1 && yi1 != yi0) ? ly1 : 0.f);
      const float wxv = (wx == xi0 ? lx0 : 0.f) + ((wx == xi1 && xi1 != xi0) ? lx1 : 0.f);
      wv = wyv * wxv;
    }
    split_store(out + i, out_plane, nsplit, wv);
  }
}

// Transpose of the adaptive average pooling of the four scales as an NHWC "activation", negated:
//   out[y][x][pos] = -1 / area(window pos)  if pixel (y, x) lies in window pos (0..49, scale-major; windows overlap when the
//   size does not divide), else 0; positions 50..63 are zero.  A operand of MODE_DFV's extra k-block:
//   acc[pixel][c] -= sum_pos (1 / area) [pixel in pos] dpool[pos][c]   (CANNet.py:42,51,60,70 backwards).
__global__ void __launch_bounds__(256)
ctx_pint_kernel(__nv_bfloat16* __restrict__ out, int H, int W) {
  const long long total = static_cast<long long>(H) * W * 64;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int pos = i & 63;
    const long long pix = i >> 6;
    const int x = pix % W;
    const int y = pix / W;
    float wv = 0.f;
    if (pos < 50) {
      int j, S, wy, wx, xwin;
      ctx_pos_decode(pos, j, S, wy, wx, xwin);
      const int ys = adaptive_start(wy, H, S), ye = adaptive_end(wy, H, S);
      const int xs = adaptive_start(wx, W, S), xe = adaptive_end(wx, W, S);
      if (y >= ys && y < ye && x >= xs && x < xe) wv = -1.f / static_cast<float>((ye - ys) * (xe - xs));
    }
    out[i] = __float2bfloat16_rn(wv);
  }
}

// B operand of that k-block (K-major, 64 positions per row): D[n][c][pos] = dpool[n][pos][c], positions 50..63 zero.
// One thread per 8 positions of a row (one 16-byte store).
__global__ void __launch_bounds__(256)
ctx_dtab_kernel(const float* __restrict__ dpool, __nv_bfloat16* __restrict__ D, int N) {
  const long long total = static_cast<long long>(N) * 512 * 8;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int p0 = static_cast<int>(i & 7) * 8;
    const int c = static_cast<int>(i >> 3) & 511;
    const long long n = i >> 12;
    const float* __restrict__ tab = dpool + n * 50 * 512 + c;
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = (p0 + e < 50) ? tab[(p0 + e) * 512] : 0.f;
    uint4 q;
    q.x = pack_bf16x2(v[0], v[1]);
    q.y = pack_bf16x2(v[2], v[3]);
    q.z = pack_bf16x2(v[4], v[5]);
    q.w = pack_bf16x2(v[6], v[7]);
    *reinterpret_cast<uint4*>(D + i * 8) = q;
  }
}

// Per-image B operands of the two extra k-blocks of MODE_CTX2 (K-major, 64 positions per row):
//   T[0][n][(cc*4 + j)*64 + c][pos] = -tabB[n][pos][cc*64 + c]  if pos belongs to scale j, else 0
//   T[1][n][(cc*4 + j)*64 + c][pos] =  tabA[n][pos][cc*64 + c]  if pos belongs to scale j, else 0
// (positions 50..63 are zero).  512 KB per image.
__global__ void __launch_bounds__(256)
ctx_ttab_kernel(const float* __restrict__ tabA, const float* __restrict__ tabB, __nv_bfloat16* __restrict__ T, int N) {
  // one thread per 8 consecutive positions of a row (one 16-byte store); a row (which, n, cc, j, c) holds the table values
  // of scale j only, so three quarters of the threads just write zeros
  const int OFF[5] = {0, 1, 5, 14, 50};
  const long long total = 2LL * N * 2048 * 8;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int p0 = static_cast<int>(i & 7) * 8;
    const int row = static_cast<int>(i >> 3) & 2047;
    const long long img = i >> 14;            // which * N + n
    const int n = static_cast<int>(img % N);
    const int which = static_cast<int>(img / N);
    const int c = row & 63, j = (row >> 6) & 3, cc = row >> 8;
    const float* __restrict__ tab = (which == 0 ? tabB : tabA) + static_cast<long long>(n) * 50 * 512 + cc * 64 + c;
    const float sgn = which == 0 ? -1.f : 1.f;
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int pos = p0 + e;
      v[e] = (pos >= OFF[j] && pos < OFF[j + 1]) ? sgn * tab[pos * 512] : 0.f;
    }
    uint4 q;
    q.x = pack_bf16x2(v[0], v[1]);
    q.y = pack_bf16x2(v[2], v[3]);
    q.z = pack_bf16x2(v[4], v[5]);
    q.w = pack_bf16x2(v[6], v[7]);
    *reinterpret_cast<uint4*>(T + i * 8) = q;
  }
}

// --------------------------------------------------------------------------------
// index-table probes (bit-exact tests of the device arithmetic against the oracle)
// --------------------------------------------------------------------------------
__global__ void bilinear_taps_kernel(int in_size, int out_size, int* i0, int* i1, float* l0, float* l1) {
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= out_size) return;
  bilinear_tap(d, in_size, out_size, i0[d], i1[d], l0[d], l1[d]);
}
__global__ void adaptive_windows_kernel(int in_size, int out_size, int* start, int* end) {
  const int i = threadIdx.x;
  if (i >= out_size) return;
  start[i] = adaptive_start(i, in_size, out_size);
  end[i] = adaptive_end(i, in_size, out_size);
}

// NHWC bf16 planes -> NCHW fp32 (debug / parity probes of intermediate activations)
__global__ void nhwc_planes_to_nchw_f32_kernel(const __nv_bfloat16* __restrict__ src, long long plane, int nsplit,
                                               int cs, int coff, float* __restrict__ dst, int N, int C, int H,
                                               int W) {
  const long long total = static_cast<long long>(N) * C * H * W;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int x = i % W;
    const int y = (i / W) % H;
    const int c = (i / (static_cast<long long>(W) * H)) % C;
    const int n = i / (static_cast<long long>(W) * H * C);
    dst[i] = split_load(src + ((static_cast<long long>(n) * H + y) * W + x) * cs + coff + c, plane, nsplit);
  }
}
// NCHW fp32 -> NHWC bf16 planes
__global__ void nchw_f32_to_nhwc_planes_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                               long long plane, int nsplit, int cs, int coff, int N, int C, int H,
                                               int W) {
  const long long total = static_cast<long long>(N) * C * H * W;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = i % C;
    const int x = (i / C) % W;
    const int y = (i / (static_cast<long long>(C) * W)) % H;
    const int n = i / (static_cast<long long>(C) * W * H);
    const float v = src[((static_cast<long long>(n) * C + c) * H + y) * W + x];
    split_store(dst + ((static_cast<long long>(n) * H + y) * W + x) * cs + coff + c, plane, nsplit, v);
  }
}

}  // namespace canb
