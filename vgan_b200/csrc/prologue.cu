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
  int center, fp32, split;
  __nv_bfloat16* zq;
  __nv_bfloat16* zb;    // split operands (BF16X3): zq = row operand [hi | hi | lo], zb = column operand [hi | lo | hi],
                        // both [rows, 3 Dp]: sum_k zq[i,k] zb[j,k] = hi_i.hi_j + hi_i.lo_j + lo_i.hi_j
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
    } else if (a.split) {
      const __nv_bfloat16 hi = __float2bfloat16_rn(z);
      const float zh = __bfloat162float(hi);
      const __nv_bfloat16 lo = __float2bfloat16_rn(z - zh);
      const float zr = zh + __bfloat162float(lo);          // the value the three products represent (16 bits)
      const int64_t o = r * 3 * a.Dp + f;
      a.zq[o] = hi; a.zq[o + a.Dp] = hi; a.zq[o + 2 * a.Dp] = lo;
      a.zb[o] = hi; a.zb[o + a.Dp] = lo; a.zb[o + 2 * a.Dp] = hi;
      nrm = fmaf(zr, zr, nrm);
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

// ---- masked fast path (Y = U*X, Nx == Ny, D % 4 == 0, 16-byte aligned inputs): X and U are read ONCE per pass
// for both halves of the stack (the generic kernels read X twice), with 16-byte loads.
__global__ void __launch_bounds__(256) colstats_masked_kernel(PackArgs a, int rows_per_block) {
  const int64_t r0 = static_cast<int64_t>(blockIdx.x) * rows_per_block;
  const int64_t r1 = min(r0 + rows_per_block, a.Nx);
  const int64_t D4 = a.D >> 2;
  float sq = 0.f;
  for (int64_t f4 = threadIdx.x; f4 < D4; f4 += blockDim.x) {
    float4 cs = make_float4(0.f, 0.f, 0.f, 0.f);
    const float4* xp = reinterpret_cast<const float4*>(a.X) + r0 * D4 + f4;
    const float4* up = reinterpret_cast<const float4*>(a.U) + r0 * D4 + f4;
#pragma unroll 4
    for (int64_t r = r0; r < r1; ++r, xp += D4, up += D4) {
      const float4 x = __ldg(xp), u = __ldg(up);
      const float4 y = make_float4(u.x * x.x, u.y * x.y, u.z * x.z, u.w * x.w);
      cs.x += x.x + y.x; cs.y += x.y + y.y; cs.z += x.z + y.z; cs.w += x.w + y.w;
      sq = fmaf(x.x, x.x, fmaf(x.y, x.y, fmaf(x.z, x.z, fmaf(x.w, x.w, sq))));
      sq = fmaf(y.x, y.x, fmaf(y.y, y.y, fmaf(y.z, y.z, fmaf(y.w, y.w, sq))));
    }
    atomicAdd(&a.stats[2 + 4 * f4 + 0], static_cast<double>(cs.x));
    atomicAdd(&a.stats[2 + 4 * f4 + 1], static_cast<double>(cs.y));
    atomicAdd(&a.stats[2 + 4 * f4 + 2], static_cast<double>(cs.z));
    atomicAdd(&a.stats[2 + 4 * f4 + 3], static_cast<double>(cs.w));
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

__device__ __forceinline__ uint2 pack4_bf16(float a, float b, float c, float d, float& nrm) {
  const __nv_bfloat16 q0 = __float2bfloat16_rn(a), q1 = __float2bfloat16_rn(b), q2 = __float2bfloat16_rn(c),
                      q3 = __float2bfloat16_rn(d);
  const float r0 = __bfloat162float(q0), r1 = __bfloat162float(q1), r2 = __bfloat162float(q2), r3 = __bfloat162float(q3);
  nrm = fmaf(r0, r0, fmaf(r1, r1, fmaf(r2, r2, fmaf(r3, r3, nrm))));   // norm of the values the tensor core will see
  uint2 o;
  o.x = static_cast<uint32_t>(__bfloat16_as_ushort(q0)) | (static_cast<uint32_t>(__bfloat16_as_ushort(q1)) << 16);
  o.y = static_cast<uint32_t>(__bfloat16_as_ushort(q2)) | (static_cast<uint32_t>(__bfloat16_as_ushort(q3)) << 16);
  return o;
}

// one warp per sample row r: writes stored rows r (X half) and Nxp + r (Y half), bf16 path
__global__ void __launch_bounds__(256) pack_masked_kernel(PackArgs a) {
  const int lane = threadIdx.x & 31;
  const int64_t r = static_cast<int64_t>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (r >= a.Nxp) return;
  const bool valid = r < a.Nx;
  const int64_t D4 = a.D >> 2, Dp4 = a.Dp >> 2;
  float nx = 0.f, ny = 0.f;
  uint2* zx = reinterpret_cast<uint2*>(a.zq + r * a.Dp);
  uint2* zy = reinterpret_cast<uint2*>(a.zq + (a.Nxp + r) * a.Dp);
  for (int64_t f4 = lane; f4 < Dp4; f4 += 32) {
    uint2 ox = make_uint2(0u, 0u), oy = make_uint2(0u, 0u);
    if (valid && f4 < D4) {
      const float4 x = __ldg(reinterpret_cast<const float4*>(a.X) + r * D4 + f4);
      const float4 u = __ldg(reinterpret_cast<const float4*>(a.U) + r * D4 + f4);
      const float4 m = *reinterpret_cast<const float4*>(a.mean + 4 * f4);
      ox = pack4_bf16(x.x - m.x, x.y - m.y, x.z - m.z, x.w - m.w, nx);
      oy = pack4_bf16(u.x * x.x - m.x, u.y * x.y - m.y, u.z * x.z - m.z, u.w * x.w - m.w, ny);
    }
    zx[f4] = ox;
    zy[f4] = oy;
  }
  nx = warp_sum(nx);
  ny = warp_sum(ny);
  if (lane == 0) {
    a.nrm[r] = valid ? nx : kPadNorm;
    a.nrm[a.Nxp + r] = valid ? ny : kPadNorm;
  }
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
  a.split = (!a.fp32 && (flags & VGAN_B200_BF16X3)) ? 1 : 0;
  a.zq = reinterpret_cast<__nv_bfloat16*>(base + L.off_zq);
  a.zb = a.zq + static_cast<size_t>(L.Nxp + L.Nyp) * 3 * L.Dp;
  a.zf = reinterpret_cast<float*>(base + L.off_zf);
  a.nrm = reinterpret_cast<float*>(base + L.off_nrm);
  a.mean = reinterpret_cast<float*>(base + L.off_mean);
  a.stats = reinterpret_cast<double*>(base + L.off_stats);
  VG_CUDA_CHECK(cudaMemsetAsync(a.stats, 0, sizeof(double) * (2 + L.Dp), st));
  const int64_t n = L.Nx + L.Ny;
  const int rpb = 64;
  const bool masked_fast = !Y && U && !a.fp32 && !a.split && L.Nx == L.Ny && (L.D & 3) == 0 &&
                           ((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(U)) & 15) == 0;
  if (masked_fast) {
    const int rpb2 = 32;
    colstats_masked_kernel<<<static_cast<unsigned>((L.Nx + rpb2 - 1) / rpb2), 256, 0, st>>>(a, rpb2);
    mean_kernel<<<static_cast<unsigned>((L.Dp + 255) / 256), 256, 0, st>>>(a);
    pack_masked_kernel<<<static_cast<unsigned>((L.Nxp + 7) / 8), 256, 0, st>>>(a);
  } else {
    colstats_kernel<<<static_cast<unsigned>((n + rpb - 1) / rpb), 256, 0, st>>>(a, rpb);
    mean_kernel<<<static_cast<unsigned>((L.Dp + 255) / 256), 256, 0, st>>>(a);
    pack_kernel<<<static_cast<unsigned>((L.Nxp + L.Nyp + 7) / 8), 256, 0, st>>>(a);
  }
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
