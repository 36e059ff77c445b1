// Per-sample data path of the reference (model/CrowdDataset.py:38-66) on the GPU: /255, gray -> RGB, horizontal
// flip, cv2.resize(INTER_LINEAR) of the image to a multiple of 8 and of the density map to 1/8 (x 64), ImageNet
// normalisation -- NCHW fp32 out, which is what conv1_1 reads.  HBM-bound and tiny (2.4 MB in, 9.4 MB out for a
// 768x1024 image); the arithmetic mirrors the oracle operation by operation (no FMA contraction: explicit _rn
// intrinsics), so the image path is bit-exact against cv2's float64 pipeline + torchvision's float32 normalize.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace canb {

// cv2 INTER_LINEAR: source index and fractional weight of destination index d (half-pixel centres, clamped)
__device__ __forceinline__ void cv_linear_coeff(int d, int n_dst, int n_src, int& s, double& f) {
  const double scale = static_cast<double>(n_src) / static_cast<double>(n_dst);
  f = __dadd_rn(__dmul_rn(static_cast<double>(d) + 0.5, scale), -0.5);
  const double fl = floor(f);
  s = static_cast<int>(fl);
  f = __dadd_rn(f, -fl);
  if (s < 0) {
    s = 0;
    f = 0.0;
  }
  if (s >= n_src - 1) {
    s = n_src - 1;
    f = 0.0;
  }
}

// img: uint8 [H][W][C] (C = 1 or 3); out: float [3][OH][OW].  256 threads per block.
// u / 255 in double (the reference's `img / 255`) comes from a 256-entry table built with exactly that division
// (a double division per tap and channel made the kernel FP64-bound: 81 us for a 1539x2051 sample).
__global__ void __launch_bounds__(256)
crowd_image_kernel(const uint8_t* __restrict__ img, int H, int W, int C, int flip, float* __restrict__ out, int OH, int OW) {
  __shared__ double s_u[256];
  s_u[threadIdx.x] = static_cast<double>(threadIdx.x) / 255.0;
  __syncthreads();
  const int ox = blockIdx.x * blockDim.x + threadIdx.x, oy = blockIdx.y;
  if (ox >= OW) return;
  const float mean[3] = {0.485f, 0.456f, 0.406f}, stdv[3] = {0.229f, 0.224f, 0.225f};
  int sx = ox, sy = oy;
  double fx = 0.0, fy = 0.0;
  const bool same = (OH == H) && (OW == W);   // cv2.resize copies when the size does not change
  if (!same) {
    cv_linear_coeff(ox, OW, W, sx, fx);
    cv_linear_coeff(oy, OH, H, sy, fy);
  }
  const int sx1 = min(sx + 1, W - 1), sy1 = min(sy + 1, H - 1);
  // the flip (CrowdDataset.py:48-50) happens before the resize: column x of the flipped image is column W-1-x
  const int cx0 = flip ? W - 1 - sx : sx, cx1 = flip ? W - 1 - sx1 : sx1;
  const double a1 = fx, a0 = __dadd_rn(1.0, -fx), b1 = fy, b0 = __dadd_rn(1.0, -fy);
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const int cc = (C == 1) ? 0 : c;
    const double p00 = s_u[img[(static_cast<long long>(sy) * W + cx0) * C + cc]];
    double v = p00;
    if (!same) {
      const double p01 = s_u[img[(static_cast<long long>(sy) * W + cx1) * C + cc]];
      const double p10 = s_u[img[(static_cast<long long>(sy1) * W + cx0) * C + cc]];
      const double p11 = s_u[img[(static_cast<long long>(sy1) * W + cx1) * C + cc]];
      const double r0 = __dadd_rn(__dmul_rn(p00, a0), __dmul_rn(p01, a1));   // horizontal pass, row sy
      const double r1 = __dadd_rn(__dmul_rn(p10, a0), __dmul_rn(p11, a1));   // horizontal pass, row sy + 1
      v = __dadd_rn(__dmul_rn(r0, b0), __dmul_rn(r1, b1));                   // vertical pass
    }
    const float xf = static_cast<float>(v);                                   // torch.tensor(img, dtype=torch.float)
    out[(static_cast<long long>(c) * OH + oy) * OW + ox] = __fdiv_rn(__fsub_rn(xf, mean[c]), stdv[c]);
  }
}

// dmap: float [H][W]; out: float [OH][OW] = cv2.resize(dmap, (OW, OH)) * scale   (float32 pipeline)
__global__ void crowd_dmap_kernel(const float* __restrict__ dmap, int H, int W, int flip, float* __restrict__ out, int OH,
                                  int OW, float scale) {
  const int ox = blockIdx.x * blockDim.x + threadIdx.x, oy = blockIdx.y;
  if (ox >= OW) return;
  int sx = ox, sy = oy;
  double fxd = 0.0, fyd = 0.0;
  const bool same = (OH == H) && (OW == W);
  if (!same) {
    cv_linear_coeff(ox, OW, W, sx, fxd);
    cv_linear_coeff(oy, OH, H, sy, fyd);
  }
  const int sx1 = min(sx + 1, W - 1), sy1 = min(sy + 1, H - 1);
  const int cx0 = flip ? W - 1 - sx : sx, cx1 = flip ? W - 1 - sx1 : sx1;
  const float a1 = static_cast<float>(fxd), a0 = __fsub_rn(1.f, a1), b1 = static_cast<float>(fyd), b0 = __fsub_rn(1.f, b1);
  float v = dmap[static_cast<long long>(sy) * W + cx0];
  if (!same) {
    const float p00 = v, p01 = dmap[static_cast<long long>(sy) * W + cx1];
    const float p10 = dmap[static_cast<long long>(sy1) * W + cx0], p11 = dmap[static_cast<long long>(sy1) * W + cx1];
    const float r0 = __fadd_rn(__fmul_rn(p00, a0), __fmul_rn(p01, a1));
    const float r1 = __fadd_rn(__fmul_rn(p10, a0), __fmul_rn(p11, a1));
    v = __fadd_rn(__fmul_rn(r0, b0), __fmul_rn(r1, b1));
  }
  out[static_cast<long long>(oy) * OW + ox] = __fmul_rn(v, scale);
}

}  // namespace canb

namespace canb {

// Ground-truth density maps (data_preparation/k_nearest_gaussian_kernel.py:14-54): one truncated, normalised, separable
// Gaussian per annotated head, sigma = 0.1 x (sum of the distances to the three nearest other heads).
// head_par[i] = { row, col, radius (or -1: skipped), -0.5 / sigma^2, sum_k exp(-0.5 k^2 / sigma^2) }
struct HeadPar {
  int r, c, R, pad;
  double coef, sum;
};

// one thread per head: brute-force 4 smallest distances (the head itself included, as KDTree.query(k=4) does)
__global__ void density_sigma_kernel(const double* __restrict__ pts, int n, int H, int W, HeadPar* __restrict__ par) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double px = pts[2 * i], py = pts[2 * i + 1];
  double sigma;
  if (n == 1) {
    sigma = (static_cast<double>(H) + static_cast<double>(W)) / 2.0 / 2.0 / 2.0;   // np.average(shape) / 2 / 2
  } else {
    double best[4] = {1e300, 1e300, 1e300, 1e300};
    for (int j = 0; j < n; ++j) {
      const double dx = px - pts[2 * j], dy = py - pts[2 * j + 1];
      double d = sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
#pragma unroll
      for (int k = 0; k < 4; ++k) {   // insertion into the sorted quadruple
        if (d < best[k]) {
          const double t = best[k];
          best[k] = d;
          d = t;
        }
      }
    }
    sigma = __dmul_rn(__dadd_rn(__dadd_rn(best[1], best[2]), best[3]), 0.1);
  }
  HeadPar hp;
  hp.r = static_cast<int>(py);
  hp.c = static_cast<int>(px);
  hp.pad = 0;
  // the reference tests int(pt) < size only (truncation toward zero): negative indices address from the far side, as
  // numpy's `pt2d[int(pt[1]), int(pt[0])] = 1.` does; further out it raises, here the head is skipped
  bool inside = hp.r < H && hp.c < W && hp.r >= -H && hp.c >= -W;
  if (hp.r < 0) hp.r += H;
  if (hp.c < 0) hp.c += W;
  if (sigma <= 1e-15) {
    // four or more coincident annotations: scipy's gaussian_filter leaves the impulse untouched for sigma <= 1e-15
    hp.R = inside ? 0 : -1;
    hp.coef = 0.0;
    hp.sum = 1.0;
    par[i] = hp;
    return;
  }
  const int R = static_cast<int>(__dadd_rn(__dmul_rn(4.0, sigma), 0.5));
  hp.R = inside ? R : -1;
  hp.coef = -0.5 / __dmul_rn(sigma, sigma);
  double s = 0.0;
  for (int k = -R; k <= R; ++k) s += exp(__dmul_rn(hp.coef, static_cast<double>(k) * static_cast<double>(k)));
  hp.sum = s;
  par[i] = hp;
}

// one thread per pixel, heads in annotation order (the float32 running sum of the reference's `density +=`)
__global__ void __launch_bounds__(256)
density_gather_kernel(const HeadPar* __restrict__ par, int n, int H, int W, float* __restrict__ density) {
  __shared__ HeadPar s_par[128];
  const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
  const int bx0 = blockIdx.x * 32, bx1 = bx0 + 31, by0 = blockIdx.y * 8, by1 = by0 + 7;
  float acc = 0.f;
  for (int base = 0; base < n; base += 128) {
    const int cnt = min(128, n - base);
    __syncthreads();
    if (threadIdx.x < cnt) s_par[threadIdx.x] = par[base + threadIdx.x];
    __syncthreads();
    for (int i = 0; i < cnt; ++i) {
      const int R = s_par[i].R, r = s_par[i].r, c = s_par[i].c;
      // whole block outside this head's support (uniform branch)
      if (R < 0 || r + R < by0 || r - R > by1 || c + R < bx0 || c - R > bx1) continue;
      const int dy = y - r, dx = x - c;
      if (dy < -R || dy > R || dx < -R || dx > R) continue;
      const double coef = s_par[i].coef, sum = s_par[i].sum;
      const double wy = exp(__dmul_rn(coef, static_cast<double>(dy * dy))) / sum;
      const double wx = exp(__dmul_rn(coef, static_cast<double>(dx * dx))) / sum;
      const double col = static_cast<double>(static_cast<float>(wy));   // first pass (rows) stored as float32
      acc = __fadd_rn(acc, static_cast<float>(__dmul_rn(col, wx)));     // second pass, float32 image, float32 sum
    }
  }
  if (x < W && y < H) density[static_cast<long long>(y) * W + x] = acc;
}

}  // namespace canb
