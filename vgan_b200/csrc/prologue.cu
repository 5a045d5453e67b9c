// K2 — HBM-bound prologue kernels: stacked column statistics, operand packing
// (Y = U*X masking, optional centering, bf16 cast, row norms), closed-form bandwidth and
// the device-side parameter block.  Replaces vstack / the pow-sum prologue of
// torch.cdist / `fake_subspaces*batch` / RBF.get_bandwidth
// (/root/reference/src/models/Mmd_loss_constrained.py:16-22,43; src/vgan.py:616).
#include "common.cuh"
#include "ptx.cuh"

namespace vg {

struct PackArgs {
  const float* X;
  const float* U;   // mask (Y = U*X) when Y == nullptr
  const float* Y;
  int64_t Nx, Ny, D, Dp, Nxp, Nyp;
  int center, fp32;
  __nv_bfloat16* zq;
  float* zf;
  float* nrm;
  float* mean;
  double* stats;    // [0] sum |z|^2 (raw), [2+f] raw column sums
};

__device__ __forceinline__ float load_stacked(const PackArgs& a, int64_t zrow, int64_t f) {
  // zrow indexes the un-padded stack [X ; Y]
  if (zrow < a.Nx) return a.X[zrow * a.D + f];
  const int64_t y = zrow - a.Nx;
  return a.Y ? a.Y[y * a.D + f] : a.U[y * a.D + f] * a.X[y * a.D + f];
}

// column sums + sum of squares of the raw stacked sample, double accumulation
__global__ void __launch_bounds__(256) colstats_kernel(PackArgs a, int rows_per_block) {
  const int64_t n = a.Nx + a.Ny;
  const int64_t r0 = static_cast<int64_t>(blockIdx.x) * rows_per_block;
  const int64_t r1 = min(r0 + rows_per_block, n);
  float sq = 0.f;
  for (int64_t f = threadIdx.x; f < a.D; f += blockDim.x) {
    float cs = 0.f;
    for (int64_t r = r0; r < r1; ++r) {
      const float v = load_stacked(a, r, f);
      cs += v;
      sq = fmaf(v, v, sq);
    }
    atomicAdd(&a.stats[2 + f], static_cast<double>(cs));
  }
  sq = warp_sum(sq);
  __shared__ float red[8];
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = sq;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t += red[w];
    atomicAdd(&a.stats[0], t);
  }
}

__global__ void __launch_bounds__(256) mean_kernel(PackArgs a) {
  const int64_t f = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (f >= a.Dp) return;
  const double n = static_cast<double>(a.Nx + a.Ny);
  a.mean[f] = (a.center && f < a.D) ? static_cast<float>(a.stats[2 + f] / n) : 0.f;
}

// one warp per stored (padded) row
__global__ void __launch_bounds__(256) pack_kernel(PackArgs a) {
  const int lane = threadIdx.x & 31;
  const int64_t r = static_cast<int64_t>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (r >= a.Nxp + a.Nyp) return;
  const bool in_x = r < a.Nxp;
  const int64_t local = in_x ? r : r - a.Nxp;
  const bool valid = local < (in_x ? a.Nx : a.Ny);
  const int64_t zrow = in_x ? local : a.Nx + local;
  float nrm = 0.f;
  for (int64_t f = lane; f < a.Dp; f += 32) {
    float z = 0.f;
    if (valid && f < a.D) z = load_stacked(a, zrow, f) - a.mean[f];
    if (a.fp32) {
      a.zf[r * a.Dp + f] = z;
      nrm = fmaf(z, z, nrm);
    } else {
      const __nv_bfloat16 q = __float2bfloat16_rn(z);
      a.zq[r * a.Dp + f] = q;
      const float zr = __bfloat162float(q);
      nrm = fmaf(zr, zr, nrm);     // norm of the value the tensor core will see
    }
  }
  nrm = warp_sum(nrm);
  if (lane == 0) a.nrm[r] = valid ? nrm : kPadNorm;
}

struct ScaleArgs {
  float mults[VGAN_B200_MAX_SCALES];
  int n_scales, frozen;
  int64_t n, D;
  const double* stats;
  float* bandwidth;
  float* inv_scales;
};

// bandwidth = sum_ij |zi-zj|^2 / (n^2-n) = (2n sum|z|^2 - 2 |sum z|^2) / (n^2-n)
__global__ void __launch_bounds__(256) scales_kernel(ScaleArgs a) {
  __shared__ double red[256];
  double acc = 0.0;
  if (!a.frozen) {
    for (int64_t f = threadIdx.x; f < a.D; f += 256) acc += a.stats[2 + f] * a.stats[2 + f];
  }
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    float bw;
    if (!a.frozen) {
      const double n = static_cast<double>(a.n);
      bw = static_cast<float>((2.0 * n * a.stats[0] - 2.0 * red[0]) / (n * n - n));
      *a.bandwidth = bw;
    } else {
      bw = *a.bandwidth;
    }
    // the reference multiplies the 0-dim bandwidth with the fp32 multiplier vector -> fp32
    for (int m = 0; m < a.n_scales; ++m) a.inv_scales[m] = 1.0f / (bw * a.mults[m]);
  }
}

struct SetupArgs {
  const float* inv_scales;
  int n_scales;
  int64_t Nx, Ny, rows;
  const float* nrm;
  float* nrms;
  float* par;
};

// parameter block + norms scaled for the exp2 argument (see common.cuh)
__global__ void __launch_bounds__(256) setup_kernel(SetupArgs a) {
  const int ref = a.n_scales - 1;
  const float c0 = kLog2e * a.inv_scales[ref];
  if (blockIdx.x == 0 && threadIdx.x < kParFloats) {
    const int t = threadIdx.x;
    float v = 0.f;
    if (t == 0) v = c0;
    else if (t == 1) v = -2.f * c0;
    else if (t >= 2 && t < 10) v = (t - 2 < a.n_scales) ? a.inv_scales[t - 2] / a.inv_scales[ref] : 0.f;
    else if (t >= 10 && t < 18) v = (t - 10 < a.n_scales) ? a.inv_scales[t - 10] / static_cast<float>(a.Nx) : 0.f;
    else if (t >= 18 && t < 26) v = (t - 18 < a.n_scales) ? -a.inv_scales[t - 18] / static_cast<float>(a.Ny) : 0.f;
    a.par[t] = v;
  }
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < a.rows) a.nrms[i] = c0 * a.nrm[i];
}

// ------------------------------------------------------------------ host launchers
int run_prologue(const float* X, const float* U, const float* Y, const vgan_b200_layout_t& L,
                 int flags, void* ws, cudaStream_t st) {
  uint8_t* base = static_cast<uint8_t*>(ws);
  PackArgs a;
  a.X = X; a.U = U; a.Y = Y;
  a.Nx = L.Nx; a.Ny = L.Ny; a.D = L.D; a.Dp = L.Dp; a.Nxp = L.Nxp; a.Nyp = L.Nyp;
  a.center = (flags & VGAN_B200_CENTER) ? 1 : 0;
  a.fp32 = (flags & VGAN_B200_PREC_MASK) == VGAN_B200_PREC_FP32;
  a.zq = reinterpret_cast<__nv_bfloat16*>(base + L.off_zq);
  a.zf = reinterpret_cast<float*>(base + L.off_zf);
  a.nrm = reinterpret_cast<float*>(base + L.off_nrm);
  a.mean = reinterpret_cast<float*>(base + L.off_mean);
  a.stats = reinterpret_cast<double*>(base + L.off_stats);
  VG_CUDA_CHECK(cudaMemsetAsync(a.stats, 0, sizeof(double) * (2 + L.Dp), st));
  const int64_t n = L.Nx + L.Ny;
  const int rpb = 64;
  colstats_kernel<<<static_cast<unsigned>((n + rpb - 1) / rpb), 256, 0, st>>>(a, rpb);
  mean_kernel<<<static_cast<unsigned>((L.Dp + 255) / 256), 256, 0, st>>>(a);
  pack_kernel<<<static_cast<unsigned>((L.Nxp + L.Nyp + 7) / 8), 256, 0, st>>>(a);
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

int run_scales(const vgan_b200_layout_t& L, void* ws, float* bandwidth, int frozen,
               const float* mults_host, int n_scales, float* inv_scales, cudaStream_t st) {
  ScaleArgs a;
  for (int m = 0; m < VGAN_B200_MAX_SCALES; ++m) a.mults[m] = m < n_scales ? mults_host[m] : 1.f;
  a.n_scales = n_scales; a.frozen = frozen; a.n = L.Nx + L.Ny; a.D = L.D;
  a.stats = reinterpret_cast<const double*>(static_cast<uint8_t*>(ws) + L.off_stats);
  a.bandwidth = bandwidth; a.inv_scales = inv_scales;
  scales_kernel<<<1, 256, 0, st>>>(a);
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

int run_setup(const vgan_b200_layout_t& L, void* ws, const float* inv_scales, int n_scales,
              float* par, cudaStream_t st) {
  uint8_t* base = static_cast<uint8_t*>(ws);
  SetupArgs a;
  a.inv_scales = inv_scales; a.n_scales = n_scales; a.Nx = L.Nx; a.Ny = L.Ny;
  a.rows = L.Nxp + L.Nyp;
  a.nrm = reinterpret_cast<const float*>(base + L.off_nrm);
  a.nrms = reinterpret_cast<float*>(base + L.off_nrms);
  a.par = par;
  setup_kernel<<<static_cast<unsigned>((a.rows + 255) / 256), 256, 0, st>>>(a);
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

}  // namespace vg
