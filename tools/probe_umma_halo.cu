// Which shared-memory rows does tcgen05.mma read for a K-major SWIZZLE_128B A operand whose start address is NOT
// 1024-byte aligned and/or whose 8-row groups are not 1024 bytes apart?  (Question behind a single-halo-tile conv:
// the x shift of a tap is +128 B on the start address, the group stride is the halo pitch.)
//
// A in smem: rows of 128 B written the way a TMA SWIZZLE_128B box lands on a 1024-aligned buffer (16-byte chunk c of
// row r at r*128 + ((c ^ (r & 7)) << 4)); every chunk j of row r carries (r / 16, r % 16, j, 0...).  B = 16 x 64
// selection matrix (same layout), so D[m][n] = A_row(m)[n] for n < 16: D tells which row/chunk the MMA read.
//
// build: nvcc -gencode arch=compute_100a,code=sm_100a -I can-distributed-pytorch_b200/csrc -o tools/_bin/probe_umma_halo tools/probe_umma_halo.cu
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "can_ptx.cuh"

using namespace canb;

constexpr int kRows = 320;

struct Variant { int shift_rows, sbo, base_off; };

__global__ void __launch_bounds__(128) probe_kernel(Variant v, float* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024 - (smem_u32(smem_raw) & 1023)) & 1023);
  uint8_t* sA = smem;                    // kRows * 128
  uint8_t* sB = smem + kRows * 128;      // 16 * 128 (1024-aligned: 320 * 128 = 40960)
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int t = threadIdx.x, warp = t >> 5;
  for (int i = t; i < kRows * 64; i += 128) {
    const int r = i >> 6, c = i & 63, chunk = c >> 3, e = c & 7;
    const float val = e == 0 ? static_cast<float>(r / 16) : e == 1 ? static_cast<float>(r % 16) : e == 2 ? static_cast<float>(chunk) : 0.f;
    *reinterpret_cast<__nv_bfloat16*>(sA + r * 128 + ((chunk ^ (r & 7)) << 4) + e * 2) = __float2bfloat16_rn(val);
  }
  for (int i = t; i < 16 * 64; i += 128) {
    const int r = i >> 6, c = i & 63, chunk = c >> 3, e = c & 7;
    *reinterpret_cast<__nv_bfloat16*>(sB + r * 128 + ((chunk ^ (r & 7)) << 4) + e * 2) = __float2bfloat16_rn(c == r ? 1.f : 0.f);
  }
  if (t == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(&tmem_slot, 32);
    tmem_relinquish();
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  if (t == 0) {
    uint64_t da = umma_smem_desc(smem_u32(sA) + v.shift_rows * 128, 16, v.sbo);
    da |= static_cast<uint64_t>(v.base_off & 7) << 49;
    const uint64_t db = umma_smem_desc(smem_u32(sB), 16, 1024);
    umma_bf16(tmem_base, da, db, umma_idesc_bf16(128, 16, 0, 0), 0);
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  uint32_t r[16];
  tmem_ld_32x16(tmem_base + (static_cast<uint32_t>(warp * 32) << 16), r);
  tmem_ld_wait();
  for (int i = 0; i < 16; ++i) out[t * 16 + i] = __uint_as_float(r[i]);
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 32);
}

int main() {
  const Variant vs[] = {{0, 1024, 0}, {1, 1024, 0}, {1, 1024, 1}, {0, 1280, 0}, {1, 1280, 0}, {1, 1280, 1}, {2, 1280, 0},
                        {2, 1280, 2}, {10, 1280, 0}, {10, 1280, 2}, {11, 1280, 3}, {11, 1280, 0}, {1, 2048, 0}, {1, 2048, 1},
                        {3, 1536, 0}, {3, 1536, 3}, {8, 1024, 0}, {16, 2048, 0}};
  float* d_out = nullptr;
  if (cudaMalloc(&d_out, 128 * 16 * sizeof(float)) != cudaSuccess) return 1;
  const int smem = 1024 + kRows * 128 + 16 * 128;
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  std::vector<float> h(128 * 16);
  for (const Variant& v : vs) {
    cudaMemset(d_out, 0xff, 128 * 16 * sizeof(float));
    probe_kernel<<<1, 128, smem>>>(v, d_out);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
      printf("shift %d sbo %d base_off %d: CUDA error %s\n", v.shift_rows, v.sbo, v.base_off, cudaGetErrorString(e));
      return 2;
    }
    cudaMemcpy(h.data(), d_out, h.size() * sizeof(float), cudaMemcpyDeviceToHost);
    int ok = 0, row_ok = 0;
    for (int m = 0; m < 128; ++m) {
      const int want = v.shift_rows + (m / 8) * (v.sbo / 128) + (m % 8);
      const float* d = &h[m * 16];
      const bool c0 = d[0] == want / 16 && d[1] == want % 16 && d[2] == 0.f;
      const bool c1 = d[8] == want / 16 && d[9] == want % 16 && d[10] == 1.f;
      ok += c0 && c1;
      row_ok += d[0] == want / 16 && d[1] == want % 16;
    }
    printf("shift %2d sbo %4d base_off %d: %3d/128 rows exact (row id of chunk 0 right in %3d)", v.shift_rows, v.sbo, v.base_off,
           ok, row_ok);
    if (ok != 128) {
      printf("   first rows [row,chunk | row,chunk]:");
      for (int m = 0; m < 18; ++m)
        printf(" %d:%d,%d|%d,%d", m, static_cast<int>(h[m * 16] * 16 + h[m * 16 + 1]), static_cast<int>(h[m * 16 + 2]),
               static_cast<int>(h[m * 16 + 8] * 16 + h[m * 16 + 9]), static_cast<int>(h[m * 16 + 10]));
    }
    printf("\n");
  }
  return 0;
}
