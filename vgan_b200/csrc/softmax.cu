// Producer of the mask and its backward (Generator.py:18-22: `upper_softmax`), one warp per row:
//   p = softmax(h) over the D features of a row,  U = p where p < 1/D, else 1
//   dh = p * (g - sum_f g_f p_f),  g = dU where p < 1/D, else 0   (the `1` branch is a constant)
// HBM-bound: the forward reads H once from HBM (the second and third sweep of a row hit L1) and writes U; the
// backward reads H and dU and writes dH.  The row maximum and 1/sum are kept ([N] floats each) so that the backward
// recomputes p instead of reading a saved copy.
#include "common.cuh"
#include "ptx.cuh"

namespace vg {

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__global__ void __launch_bounds__(256)
upper_softmax_kernel(const float* __restrict__ H, int64_t N, int64_t D, float thr, float* __restrict__ U,
                     float* __restrict__ row_max, float* __restrict__ row_inv_sum) {
  const int lane = threadIdx.x & 31;
  const int64_t r = static_cast<int64_t>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (r >= N) return;
  const float* h = H + r * D;
  float m = -INFINITY;
  for (int64_t f = lane; f < D; f += 32) m = fmaxf(m, h[f]);
  m = warp_max(m);
  float s = 0.f;
  for (int64_t f = lane; f < D; f += 32) s += expf(h[f] - m);
  s = warp_sum(s);
  const float inv = 1.f / s;
  float* u = U + r * D;
  for (int64_t f = lane; f < D; f += 32) {
    const float p = expf(h[f] - m) * inv;
    u[f] = p < thr ? p : 1.f;
  }
  if (lane == 0) {
    row_max[r] = m;
    row_inv_sum[r] = inv;
  }
}

__global__ void __launch_bounds__(256)
upper_softmax_bwd_kernel(const float* __restrict__ H, const float* __restrict__ row_max,
                         const float* __restrict__ row_inv_sum, const float* __restrict__ dU, int64_t N, int64_t D,
                         float thr, float* __restrict__ dH) {
  const int lane = threadIdx.x & 31;
  const int64_t r = static_cast<int64_t>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (r >= N) return;
  const float* h = H + r * D;
  const float* g = dU + r * D;
  const float m = row_max[r], inv = row_inv_sum[r];
  float dot = 0.f;
  for (int64_t f = lane; f < D; f += 32) {
    const float p = expf(h[f] - m) * inv;
    if (p < thr) dot = fmaf(g[f], p, dot);
  }
  dot = warp_sum(dot);
  float* o = dH + r * D;
  for (int64_t f = lane; f < D; f += 32) {
    const float p = expf(h[f] - m) * inv;
    o[f] = p * ((p < thr ? g[f] : 0.f) - dot);
  }
}

int run_upper_softmax(const float* H, int64_t N, int64_t D, float* U, float* row_max, float* row_inv_sum,
                      cudaStream_t st) {
  const float thr = static_cast<float>(1.0 / static_cast<double>(D));
  upper_softmax_kernel<<<static_cast<unsigned>((N + 7) / 8), 256, 0, st>>>(H, N, D, thr, U, row_max, row_inv_sum); count_launch();
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

int run_upper_softmax_bwd(const float* H, const float* row_max, const float* row_inv_sum, const float* dU, int64_t N,
                          int64_t D, float* dH, cudaStream_t st) {
  const float thr = static_cast<float>(1.0 / static_cast<double>(D));
  upper_softmax_bwd_kernel<<<static_cast<unsigned>((N + 7) / 8), 256, 0, st>>>(H, row_max, row_inv_sum, dU, N, D, thr,
                                                                             dH); count_launch();
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

}  // namespace vg
