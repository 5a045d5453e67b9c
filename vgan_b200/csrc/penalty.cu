// K3 — column-max penalty  weight * mean_d(1 - max_i U[i,d])  and its gradient, with the
// lowest-row tie rule of CPU torch.topk pinned.
// Replaces /root/reference/src/models/Mmd_loss_constrained.py:50.  Also the loss finaliser.
#include "common.cuh"
#include "ptx.cuh"

namespace vg {

__device__ __forceinline__ uint32_t float_key(float v) {      // order-preserving float -> uint
  const uint32_t b = __float_as_uint(v);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float key_float(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

__global__ void __launch_bounds__(256)
colmax_kernel(const float* __restrict__ U, int64_t N, int64_t D, int rows_per_block,
              unsigned long long* __restrict__ cmax) {
  const int64_t r0 = static_cast<int64_t>(blockIdx.x) * rows_per_block;
  const int64_t r1 = min(r0 + rows_per_block, N);
  for (int64_t d = threadIdx.x; d < D; d += blockDim.x) {
    unsigned long long best = 0ull;
    for (int64_t r = r0; r < r1; ++r) {
      // larger value wins; on equal values the LOWER row index wins
      const unsigned long long key = (static_cast<unsigned long long>(float_key(U[r * D + d])) << 32) |
                                     static_cast<unsigned long long>(0xffffffffu - static_cast<uint32_t>(r));
      best = key > best ? key : best;
    }
    atomicMax(&cmax[d], best);
  }
}

__global__ void __launch_bounds__(256)
penalty_kernel(const unsigned long long* __restrict__ cmax, int64_t D, float weight,
               float* __restrict__ penalty, float* __restrict__ dU, int64_t row_begin, int64_t rows) {
  __shared__ float red[8];
  float part = 0.f;
  for (int64_t d = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; d < D;
       d += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const unsigned long long key = cmax[d];
    const float vmax = key_float(static_cast<uint32_t>(key >> 32));
    const int64_t row = 0xffffffffu - static_cast<uint32_t>(key & 0xffffffffu);
    part += 1.f - vmax;
    if (dU && row >= row_begin && row < row_begin + rows)
      atomicAdd(&dU[(row - row_begin) * D + d], -weight / static_cast<float>(D));
  }
  part = warp_sum(part);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = part;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += red[w];
    atomicAdd(penalty, weight * t / static_cast<float>(D));
  }
}

__global__ void finalize_kernel(const double* __restrict__ sums, double nx, double ny,
                                const float* __restrict__ penalty, float* __restrict__ loss) {
  const double mmd2 = sums[0] / (nx * nx) - 2.0 * sums[1] / (nx * ny) + sums[2] / (ny * ny);
  loss[0] = static_cast<float>(mmd2 + (penalty ? static_cast<double>(penalty[0]) : 0.0));
}

int run_penalty(const float* U, int64_t N, int64_t D, float weight, unsigned long long* cmax,
                float* penalty, float* dU, int64_t row_begin, int64_t rows, cudaStream_t st) {
  VG_CUDA_CHECK(cudaMemsetAsync(cmax, 0, sizeof(unsigned long long) * D, st));
  const int rpb = 64;
  colmax_kernel<<<static_cast<unsigned>((N + rpb - 1) / rpb), 256, 0, st>>>(U, N, D, rpb, cmax);
  const unsigned blocks = static_cast<unsigned>((D + 255) / 256 < 64 ? (D + 255) / 256 : 64);
  penalty_kernel<<<blocks, 256, 0, st>>>(cmax, D, weight, penalty, dU, row_begin, rows);
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

int run_finalize(const double* sums, int64_t Nx, int64_t Ny, const float* penalty, float* loss,
                 cudaStream_t st) {
  finalize_kernel<<<1, 1, 0, st>>>(sums, static_cast<double>(Nx), static_cast<double>(Ny), penalty, loss);
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

}  // namespace vg
