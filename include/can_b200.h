/* can_b200.h -- C ABI of the B200-native CANNet step kernels (libcan_b200.so).
 *
 * The reference (zgzhengSEU/CAN-distributed-pytorch) has no FFI of its own: every FLOP
 * goes through torch.nn / torch.nn.functional.  Each entry point below therefore cites
 * the reference call site(s) it replaces.  All pointers are DEVICE pointers owned by
 * the caller (torch tensors on the Python side); the library never allocates or frees
 * memory visible to the caller -- scratch comes from caller-provided workspaces sized
 * by the *_workspace_bytes() queries.  Every launch is asynchronous on `stream`
 * (a cudaStream_t passed as void*; NULL = legacy default stream).
 *
 * Return value: 0 = ok; negative = invalid argument (nothing was launched);
 * positive = cudaError_t.  can_last_error_string() describes the last failure of the
 * calling thread.
 *
 * Activation layout ("NHWC planes"): bf16 [plane][N][H][W][cs] where the tensor occupies
 * channels [coff, coff+C) of a pixel stride `cs`.  nsplit = 1 is the bf16 path; nsplit = 2
 * (3) is the fp32-parity path: value = plane0 + plane1 (+ plane2), each plane bf16, the
 * products hi*hi + hi*lo + lo*hi (...) accumulated in fp32 on the tensor cores.
 */
#ifndef CAN_B200_H_
#define CAN_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* can_last_error_string(void);
int can_version(void);

/* ---- modes of the tcgen05 implicit-GEMM kernel -------------------------------------- */
enum {
  CAN_MODE_STORE = 0,  /* y = [relu](acc + bias) [* (mask > 0)]                              */
  CAN_MODE_POOL = 1,   /* y = maxpool2x2(relu(acc + bias)), optional argmax codes            */
  CAN_MODE_HEAD = 2,   /* y = relu(acc + bias) and density = y . head_w + head_b (Cout == 64) */
  CAN_MODE_UNPOOL = 3, /* g = acc * (mask > 0) scattered to the argmax slot of the 2x map    */
  CAN_MODE_CTX = 4,    /* contextual fusion epilogue                                         */
  CAN_MODE_CTX_BWD = 5,/* backward of the fusion: writes dz_j and the direct ds_j (4096 channels)  */
  CAN_MODE_DFV = 6,    /* dfv = (add - acc + adaptive-pool^T spread) * (mask > 0)              */
  CAN_MODE_CTX2 = 7,   /* CTX with the bilinear interpolation folded into the GEMM (bf16 path)   */
  CAN_MODE_CTX2_BWD = 8/* CTX_BWD on the same extended GEMM (bf16 path)                          */
};

typedef struct can_conv_args {
  int32_t N, H, W;       /* pixel grid of the GEMM (output grid == input grid; stride 1)     */
  int32_t Cin, Cout;     /* Cin % 64 == 0; Cout in {64,128,256,512} (2048 for CAN_MODE_CTX)   */
  int32_t taps;          /* 9 = 3x3, 1 = 1x1                                                  */
  int32_t dil;           /* dilation == padding: 1 (frontend) or 2 (backend)                  */
  int32_t nsplit;        /* 1 = bf16; 2 or 3 = split-bf16 planes (3 = the fp32-parity path)    */
  int32_t mode;          /* CAN_MODE_*                                                        */
  int32_t relu;          /* CAN_MODE_STORE only                                               */
  int32_t bn;            /* N-tile override (0 = auto): 64 | 128 | 256                        */
  int32_t stages;        /* pipeline-depth override (0 = auto)                                */
  const void* x;         /* input activation, NHWC planes                                     */
  int64_t x_plane;
  int32_t x_cs, x_coff;
  const void* w;         /* packed weights [plane][Cout][taps][Cin] bf16 (can_pack_conv_weight) */
  const float* bias;     /* [Cout] fp32 or NULL                                               */
  void* out;             /* output NHWC planes (pooled grid for POOL, 2x grid for UNPOOL)     */
  int64_t out_plane;
  int32_t out_cs, out_coff;
  float* out_f32;        /* optional fp32 NHWC copy of the output (STORE)                     */
  const void* mask;      /* STORE/UNPOOL: activation whose sign gates the result (plane 0)    */
  int32_t mask_cs, mask_coff;
  int64_t mask_plane;    /* CTX_BWD: `mask` carries dL/dfi as NHWC planes                     */
  const void* add;       /* DFV: gradient that reaches fv through the concat, NHWC planes     */
  int64_t add_plane;
  int32_t add_cs, add_coff;
  uint8_t* arg;          /* POOL: argmax codes out (may be NULL); UNPOOL: codes in            */
  int32_t arg_cs;
  int32_t OH, OW;        /* UNPOOL: size of the upsampled map                                 */
  const float* head_w;   /* HEAD: [64]                                                        */
  const float* head_b;  /* HEAD: [1] (device pointer: no host sync to read the bias)             */
  float* head_out;       /* HEAD: [N][H][W] fp32 density                                      */
  const float* tabA;     /* CTX: [N][50][512] a_j = conv{S}_1(pool_S(fv))                     */
  const float* tabB;     /* CTX: [N][50][512] b_j = conv{S}_2(a_j)                            */
  const void* ctx_wint;  /* CTX2: can_ctx_wint(nsplit 1, N 1) interpolation matrix [H][W][64] bf16; DFV: can_ctx_pint (optional) */
  const void* ctx_ttab;  /* CTX2: can_ctx_ttab tables [2][N][2048][64] bf16; DFV: can_ctx_dtab (optional)                */
  float* db;             /* STORE/UNPOOL/DFV (dgrad use): [Cout] += per-channel sums of the stored
                            tile = bias gradient of the layer whose dZ this launch produces; NULL = off */
} can_conv_args;

/* The dense contractions of the step on tcgen05/TMEM fed by TMA.
 *   forward 3x3 / dilated 3x3 + bias + ReLU      model/CANNet.py:14-16 via make_layers :104-121
 *   + fused MaxPool2d(2,2)                       model/CANNet.py:112
 *   + fused output_layer (1x1, 64->1)            model/CANNet.py:17,90
 *   dgrad (same kernel on dgrad-packed weights, ReLU mask / max-unpool fused):
 *                                                autograd of the above, utils/train_eval_utils.py:38
 *   CAN_MODE_CTX: contrast 1x1 convs + sigmoid + weighted fusion, written into the channel
 *   half [512,1024) of the backend input (no torch.cat):   model/CANNet.py:45-49,54-58,63-67,
 *                                                75-81,84,87                                 */
int can_conv_tc(const can_conv_args* args, void* stream);

/* conv weight [Cout][Cin][kh][kw] fp32 -> packed bf16 planes.
 * dgrad = 0: dst[pl][co][tap][ci]; dgrad = 1: dst[pl][ci][taps-1-tap][co] (transposed, flipped).
 * Replaces cuDNN's internal filter transforms for nn.Conv2d (model/CANNet.py:114).          */
int can_pack_conv_weight(const float* w, void* dst, int cout, int cin, int taps, int nsplit, int dgrad,
                         void* stream);
/* Both packings in one pass over the fp32 weight (either destination may be NULL); the training
 * step re-packs every weight after each optimizer update.                                        */
int can_pack_conv_weight_both(const float* w, void* dst_fwd, void* dst_dgrad, int cout, int cin, int taps,
                              int nsplit, void* stream);
/* The same for up to 24 weights in one launch (all 3x3 / 1x1 conv weights of the network; cout and cin multiples of
 * 32, taps 9 or 1).  A training step calls it once after the optimizer update (utils/train_eval_utils.py:52).      */
typedef struct can_pack_item {
  const float* w;    /* [cout][cin][taps] fp32 */
  void* dst_fwd;     /* [plane][cout][taps][cin] bf16, or NULL */
  void* dst_dgrad;   /* [plane][cin][taps][cout] bf16 (taps flipped), or NULL */
  int cout, cin, taps;
} can_pack_item;
int can_pack_conv_weights_multi(const can_pack_item* items, int n, int nsplit, void* stream);
/* conv{1,2,3,6}_2.weight -> stacked [plane][2048][512] in (channel-chunk, scale, channel) row
 * order for CAN_MODE_CTX (model/CANNet.py:19,21,23,25).                                      */
int can_pack_ctx_w2(const float* w1_2, const float* w2_2, const float* w3_2, const float* w6_2, void* dst,
                    int nsplit, void* stream);
int can_transpose_f32(const float* src, float* dst, int rows, int cols, void* stream);
/* The same for n <= 16 matrices of one shape in ONE launch: srcs is a HOST array of n device pointers, dst receives the
 * n transposes back to back ([n][cols][rows]).  (The eight W^T tables of conv{1,2,3,6}_{1,2}, model/CANNet.py:18-25, that
 * the backward of the contextual module needs after every optimizer step.)                                          */
int can_transpose_f32_multi(const float* const* srcs, float* dst, int n, int rows, int cols, void* stream);

/* First convolution 3->64 + bias + ReLU on the caller's NCHW fp32 image
 * (frontend.0 + frontend.1, model/CANNet.py:14,40). out: NHWC planes, cs = 64.               */
int can_conv_first_fwd(const float* x_nchw, const float* w, const float* b, void* out, int64_t out_plane,
                       int nsplit, int N, int H, int W, void* stream);

/* K1: multi-scale adaptive average pooling (S = 1,2,3,6) in ONE pass over fv, then the
 * per-position 1x1 convs.  Replaces 4x F.adaptive_avg_pool2d + conv{S}_1 (+ conv{S}_2 on the
 * <=36 pooled positions, valid because a bias-free 1x1 conv commutes with bilinear upsampling):
 * model/CANNet.py:42-43,51-52,60-61,70-72.
 *   w1 / w2 : [4][512(co)][512(ci)] fp32 stacks of the conv{S}_1 / conv{S}_2 weights as stored
 *   pooled, tabA, tabB : [N][50][512] fp32; positions ordered S=1 | S=2 (4) | S=3 (9) | S=6 (36)
 *   ws : can_ctx_pool_workspace_bytes(N, H) bytes                                            */
int64_t can_ctx_pool_workspace_bytes(int N, int H);
int can_ctx_pool_fwd(const void* fv, int64_t fv_plane, int nsplit, int cs, int coff, int N, int H, int W,
                     const float* w1, const float* w2, float* pooled, float* tabA, float* tabB, void* ws,
                     void* stream);

/* Fused MSELoss(reduction='sum') + per-image counts + dL/d(et)
 * (utils/train_eval_utils.py:20,37,83; test.py:32).  grad may be NULL (inference).
 * ws: can_mse_count_workspace_bytes(N) bytes, zero-initialised once by the caller.           */
int64_t can_mse_count_workspace_bytes(int N);
int can_mse_count_fwd_bwd(const float* et, const float* gt, int N, int HW, float* loss, float* cnt_et,
                          float* cnt_gt, float* grad, float grad_scale, void* ws, void* stream);

/* The same fused loss / count / gradient reduction for callers that cannot allocate device memory.
 * (No counterpart in the reference: this entry point exists only for hosts of this library.)
 * Who needs it: a server or plugin host that reserves the whole GPU at start-up and gives a plugin
 * only its input and output buffers, or a step loop that must avoid the device-wide stall of
 * cudaMalloc/cudaFree.  Same arguments and results as can_mse_count_fwd_bwd minus the workspace: the
 * scratch is a 4.25 KB array built into the library (held by every GPU that loads it, until exit).
 *   - limit: N <= can_mse_count_static_max_images() (1024) images per call; larger N is rejected
 *     with a negative return code and an explanatory can_last_error_string().
 *   - the first call on each device looks the array's device address up (cudaGetSymbolAddress, with
 *     that device current); make that first call before starting any stream capture.
 *   - no cudaMalloc / cudaFree / allocating library routine is called, on the first call or later.
 *   - all calls on a device share the one array: calls whose kernels can overlap on the GPU
 *     (different streams, or different host threads) race on it -- serialise them.               */
int can_mse_count_static_max_images(void);
int can_mse_count_fwd_bwd_static(const float* et, const float* gt, int N, int HW, float* loss, float* cnt_et,
                                 float* cnt_gt, float* grad, float grad_scale, void* stream);

/* Zero n <= 32 slices [offsets[i], offsets[i] + counts[i]) (in floats) of one fp32 buffer in one launch: the gradient
 * slices that backward accumulates into (biases, head) -- the part of `optimizer.zero_grad()`
 * (utils/train_eval_utils.py:29) that the fused step still needs; `offsets` / `counts` are HOST arrays.            */
int can_zero_ranges(float* base, const int64_t* offsets, const int32_t* counts, int n, void* stream);

/* torch.optim.SGD(momentum, dampening 0, no nesterov, wd 0) over a flat fp32 arena
 * (train.py:63,125-126; utils/train_eval_utils.py:52).  g is multiplied by grad_scale first. */
int can_sgd_momentum_step(float* p, const float* g, float* buf, int64_t n, float lr, float momentum,
                          float grad_scale, int first_step, void* stream);

/* Weight gradient of a 3x3 / dilated 3x3 / 1x1 convolution on tcgen05 (both operands MN-major,
 * split-K over pixel patches, deterministic reduction):
 *   grad[co][ci][tap] (+)= alpha * sum_{n,y,x} dz[n,y,x,co] * x[n, y+dy, x+dx, ci]
 * Replaces cuDNN backward-filter under loss.backward() (utils/train_eval_utils.py:38) for the
 * layers built at model/CANNet.py:14-16,19,21,23,25.  grad is fp32 in the parameter's own
 * [Cout][Cin][kh][kw] layout; ws: can_conv_wgrad_workspace_bytes(...) bytes of scratch.        */
int64_t can_conv_wgrad_workspace_bytes(int N, int H, int W, int Cin, int Cout, int taps, int nsplit);
int can_conv_wgrad(const void* dz, int64_t dz_plane, int dz_cs, int dz_coff, const void* x, int64_t x_plane,
                   int x_cs, int x_coff, int nsplit, int N, int H, int W, int Cin, int Cout, int taps, int dil,
                   float* grad, int accumulate, float alpha, void* ws, int64_t ws_bytes, void* stream);

/* im2col of the 3-channel image: out[pl][n][y][x][64], k = ci*9 + r*3 + s (27 used, rest zero),
 * which turns frontend.0 (model/CANNet.py:14) into a K=64 1x1 GEMM for its weight gradient.    */
int can_im2col_first(const float* x_nchw, void* out, int64_t out_plane, int nsplit, int N, int H, int W,
                     void* stream);

/* frontend.0 (model/CANNet.py:14: Conv2d(3, 64, 3, padding=1) + ReLU) on the tensor cores without an
 * im2col buffer: the 27-wide patch rows are rebuilt in shared memory from the fp32 NCHW image.
 * bf16 path only (single-plane output).  w_packed: the [64][64] bf16 matrix of
 * can_pack_conv_weight(w padded to [64][64][1][1], nsplit 1); out: NHWC bf16 with pixel stride out_cs
 * and channel offset out_coff.  Replaces the cuDNN forward of that layer (model/CANNet.py:40).      */
int can_conv_first_tc_fwd(const float* x_nchw, const void* w_packed, const float* bias, void* out, int out_cs,
                          int out_coff, int N, int H, int W, void* stream);

/* Weight gradient of frontend.0 from the image itself (no im2col buffer):
 *   grad[co][ci][ky][kx] (+)= alpha * sum_{n,y,x} dz[n,y,x,co] * x[n,ci,y+ky-1,x+kx-1]  (bf16 products, fp32 sums)
 * dz: NHWC bf16 (single plane), ws: can_conv_first_wgrad_workspace_bytes() bytes of scratch.
 * Replaces cuDNN backward-filter for that layer under loss.backward() (utils/train_eval_utils.py:38). */
int64_t can_conv_first_wgrad_workspace_bytes(void);
int can_conv_first_wgrad(const void* dz, int dz_cs, int dz_coff, const float* x_nchw, int N, int H, int W, float* grad,
                         int accumulate, float alpha, void* ws, int64_t ws_bytes, void* stream);

/* One sample of the reference's data path (model/CrowdDataset.py:38-66) on the device:
 *   out_img  [3][H/ds*ds][W/ds*ds] fp32 = normalize( cv2.resize( flip( gray->RGB( img / 255 ) ) ) )   (ImageNet mean/std)
 *   out_dmap [H/ds][W/ds]          fp32 = cv2.resize( flip( dmap ), (W/ds, H/ds) ) * ds^2             (optional: dmap may be NULL)
 * img: uint8 [H][W][C], C = 1 or 3 (device memory); cv2.resize = INTER_LINEAR, half-pixel centres, a plain copy when the size
 * does not change.  The image path is bit-exact against cv2 (float64) + torchvision normalize (float32).        */
int can_crowd_sample(const uint8_t* img, int H, int W, int C, const float* dmap, int flip, int ds, float* out_img,
                     float* out_dmap, void* stream);

/* Ground-truth density map of one image (data_preparation/k_nearest_gaussian_kernel.py:14-54): for every annotated head
 * (points_xy = [n][2] doubles, (col, row), device memory) a unit impulse at (int(row), int(col)) blurred by
 * scipy.ndimage.gaussian_filter(sigma, mode='constant') with sigma = 0.1 * (distances to the three nearest other heads)
 * (one head: mean(H, W) / 4); heads outside the image are skipped; float32 accumulation in annotation order.
 * density: [H][W] fp32.  n = 2 or 3 is rejected (the reference itself fails there).  ws: can_density_map_workspace_bytes(n). */
int64_t can_density_map_workspace_bytes(int npts);
int can_density_map(const double* points_xy, int npts, int H, int W, float* density, void* ws, int64_t ws_bytes,
                    void* stream);

/* db[c] += sum_pixels dz[pixel][c]  (bias gradient of every conv; db must be initialised).    */
int can_bias_grad(const void* dz, int64_t plane, int nsplit, int cs, int coff, int C, int64_t npix, float* db,
                  void* stream);

/* Backward of output_layer (model/CANNet.py:17,90) fused with the ReLU mask of backend.10:
 * dz[pix][c] = dout[pix] * head_w[c] * (y5 > 0); d_head_w += sum dout*y5; d_head_b += sum dout. */
int can_head_bwd(const float* dout, const void* y5, int64_t y_plane, int nsplit, const float* head_w, void* dz,
                 int64_t dz_plane, int64_t npix, float* d_head_w, float* d_head_b, void* stream);

/* Backward of the contextual module's small pieces (autograd of model/CANNet.py:42-84).
 * can_ctx_wint: the bilinear-interpolation matrix of the four scales as an NHWC tensor
 *   out[pl][n][y][x][pos], pos < 50 (scale-major), zero for 50..63; U^T of a full-resolution tensor
 *   is then a tensor-core GEMM over pixels (can_conv_wgrad with this tensor as `dz`).
 * can_ctx_pos_gemv: out[n][pos][r] = add(n,pos,r) + sum_k Wm[scale(pos)][r][k] * in(n,pos,k), with
 *   in(n,pos,k) = in[n*in_ns + pos*in_ps + scale*in_js + k]; Wm = [4][512][512] (weights as stored
 *   for the forward 1x1 convs on pooled positions, their transposes for the backward).
 * can_ctx_outer: dW[scale][r][c] (+)= alpha * sum_{n,pos in scale} A(n,pos,r) * B[n][pos][c]
 *   (weight gradients of conv{S}_1 and the pooled-position part of conv{S}_2).                  */
int can_ctx_wint(void* out, int64_t out_plane, int nsplit, int N, int H, int W, void* stream);
/* U^T of a full-resolution tensor for the whole batch in one launch (bf16 path):
 *   out[pos][b][c] = sum_{y,x} wint[y][x][pos] * x[b][y][x][c],  pos < 64, c < Cin (multiple of 256)
 * wint: can_ctx_wint output for N = 1 (pixel stride wint_cs = 64); x: NHWC bf16, B images.  The tensor-core GEMM of
 * can_conv_wgrad with the interpolation matrix as the shared `dz` operand; ws: can_ctx_ut_workspace_bytes(). */
int64_t can_ctx_ut_workspace_bytes(int B, int H, int W, int Cin);
int can_ctx_ut(const void* wint, int wint_cs, const void* x, int x_cs, int x_coff, int B, int H, int W, int Cin, float* out,
               void* ws, int64_t ws_bytes, void* stream);
/* CTX2 operands: T[0][n][(cc*4+j)*64+c][pos] = -tabB[n][pos][cc*64+c], T[1] = +tabA (zero outside scale j). */
int can_ctx_ttab(const float* tabA, const float* tabB, void* T, int N, void* stream);
/* MODE_DFV with the transpose of the adaptive average pooling (the gradient of model/CANNet.py:42,51,60,70) as one more
 * k-block of its GEMM instead of thirteen table rows per pixel in the epilogue (bf16 path, CTA pairs, 256-wide tiles):
 *   can_ctx_pint: A operand, [H][W][64] bf16, -1/area(window) where the pixel lies in pooled position pos (image-
 *                 independent: build once per feature-map size); pass it as can_conv_args.ctx_wint;
 *   can_ctx_dtab: B operand, [N][512][64] bf16 = the pooled-map gradient dpool [N][50][512] fp32 transposed; pass it as
 *                 can_conv_args.ctx_ttab (tabA must still be given: small shapes keep the epilogue form).          */
int can_ctx_pint(void* out, int H, int W, void* stream);
int can_ctx_dtab(const float* dpool, void* D, int N, void* stream);
int can_ctx_pos_gemv(const float* Wm, const float* in, int64_t in_ns, int64_t in_ps, int64_t in_js,
                     const float* add, float* out, int N, void* stream);
int can_ctx_outer(const float* A, int64_t a_ns, int64_t a_ps, int64_t a_js, const float* B, int N, float* dW,
                  float alpha, int accumulate, void* stream);

/* Index-table probes: the exact device arithmetic of the pooling windows and bilinear taps
 * (ATen AdaptivePooling.h:32-38, UpSample.h:271-295,442-476 as used by model/CANNet.py:42-76). */
int can_bilinear_taps(int in_size, int out_size, int32_t* i0, int32_t* i1, float* l0, float* l1, void* stream);
int can_adaptive_windows(int in_size, int out_size, int32_t* start, int32_t* end, void* stream);

/* Layout probes / boundary transposes between the caller's NCHW fp32 and NHWC planes.        */
int can_nhwc_to_nchw_f32(const void* src, int64_t plane, int nsplit, int cs, int coff, float* dst, int N, int C,
                         int H, int W, void* stream);
int can_nchw_f32_to_nhwc(const float* src, void* dst, int64_t plane, int nsplit, int cs, int coff, int N, int C,
                         int H, int W, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CAN_B200_H_ */
