// fp32 CUDA-core variant of K1 (the "fp32 path", 1e-4 parity) and the dense kernel-matrix
// kernel.  Same unit decomposition as the tcgen05 kernel but with 64-row tiles, 64-wide
// feature slices and register-tiled fp32 FMAs; also the natural path for the tiny-D
// encodings of VGAN.fit (D = latent = 1..64, /root/reference/src/vgan.py:268-287) where a
// tensor-core tile would be mostly padding.
// Replaces /root/reference/src/models/Mmd_loss_constrained.py:24-26, 42-50 + autograd.
#include <cstring>

#include "common.cuh"
#include "ptx.cuh"

namespace vg {

constexpr int kST = 64;     // tile edge
constexpr int kSK = 16;     // K step

struct SimtParams {
  int tiles_x, tiles_y;     // 64-row tiles in X / Y regions
  int nslices;              // Dp / 64
  int gx_begin, gx_end, gy_begin, gy_end;
  int xx_part, xx_nparts, n_xx_units;
  int pv, n_scales;
  int64_t Nx, Ny, D, Dp, row_begin, row_end;
  const float* zf;
  const float* nrm;         // unscaled |z|^2 (1e30 on pads)
  const float* inv_scales;
  double* sums;
  double* colsum;           // [D] column sums of the dY rows written (after dY_mul), or nullptr
  float* dY;
  const float* dY_mul;
  float* dX;
};

__global__ void __launch_bounds__(256) mmd2_simt_kernel(const SimtParams p) {
  __shared__ float As[kSK][kST + 4];
  __shared__ float Bs[kSK][kST + 4];
  __shared__ float Ps[kST][kST + 1];
  __shared__ float Vs[kST][kST + 4];
  __shared__ float ni_s[kST], nj_s[kST];
  __shared__ double red[8][2];

  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  // ---- unit decode (mirrors mmd_tc.cu)
  const int ngx = p.gx_end - p.gx_begin, ngy = p.gy_end - p.gy_begin;
  const int n_grad_units = (ngx + ngy) * p.nslices;
  int ti, slice, j_begin, j_end, pv, row_is_x, sym;
  if (static_cast<int>(blockIdx.x) < n_grad_units) {
    const int t = blockIdx.x / p.nslices;
    slice = blockIdx.x - t * p.nslices;
    row_is_x = t < ngx;
    ti = row_is_x ? p.gx_begin + t : p.tiles_x + p.gy_begin + (t - ngx);
    j_begin = 0; j_end = p.tiles_x + p.tiles_y; pv = p.pv; sym = 0;
  } else {
    const int x = blockIdx.x - n_grad_units;
    ti = p.xx_part + x * p.xx_nparts; slice = 0; row_is_x = 1;
    j_begin = ti; j_end = p.tiles_x; pv = 0; sym = 1;
  }
  float inv_s[VGAN_B200_MAX_SCALES];
#pragma unroll
  for (int m = 0; m < VGAN_B200_MAX_SCALES; ++m) inv_s[m] = m < p.n_scales ? p.inv_scales[m] : 0.f;
  const float cx = 1.f / static_cast<float>(p.Nx), cy = -1.f / static_cast<float>(p.Ny);
  const int64_t row0 = static_cast<int64_t>(ti) * kST;
  if (tid < kST) ni_s[tid] = p.nrm[row0 + tid];

  float O[4][4] = {};
  float rs[4] = {};
  double acc_x = 0.0, acc_y = 0.0;

  for (int jt = j_begin; jt < j_end; ++jt) {
    const int64_t col0 = static_cast<int64_t>(jt) * kST;
    const bool col_is_x = jt < p.tiles_x;
    const float cj = col_is_x ? cx : cy;
    __syncthreads();                      // previous iteration done with Ps/Vs/nj_s
    if (tid < kST) nj_s[tid] = p.nrm[col0 + tid];
    // ---- Gram tile, fp32 register tiling
    float acc[4][4] = {};
    for (int64_t k0 = 0; k0 < p.Dp; k0 += kSK) {
#pragma unroll
      for (int l = 0; l < 4; ++l) {
        const int e = tid + l * 256;       // 0..1023 -> (row = e/16, k = e%16)
        const int r = e >> 4, k = e & 15;
        As[k][r] = p.zf[(row0 + r) * p.Dp + k0 + k];
        Bs[k][r] = p.zf[(col0 + r) * p.Dp + k0 + k];
      }
      __syncthreads();
#pragma unroll
      for (int k = 0; k < kSK; ++k) {
        const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
        const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
        const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
      }
      __syncthreads();
    }
    // ---- kernel values
    float ksum = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float d2 = fmaxf(ni_s[ty * 4 + i] + nj_s[tx * 4 + j] - 2.f * acc[i][j], 0.f);
        float kk = 0.f, ww = 0.f;
        for (int m = 0; m < p.n_scales; ++m) {
          const float e = exp2f(-d2 * inv_s[m] * kLog2e);
          kk += e;
          ww = fmaf(e, inv_s[m], ww);
        }
        ksum += kk;
        ww *= cj;
        rs[i] += ww;
        Ps[ty * 4 + i][tx * 4 + j] = ww;
      }
    }
    const float fac = (sym && jt != ti) ? 2.f : 1.f;
    if (col_is_x) acc_x += static_cast<double>(ksum * fac); else acc_y += static_cast<double>(ksum);
    if (pv) {
      // ---- O[64 x 64 slice] += P[64 x 64] V[64 x 64 slice]
#pragma unroll
      for (int l = 0; l < 16; ++l) {
        const int e = tid + l * 256;       // 0..4095 -> (j = e/64, f = e%64)
        Vs[e >> 6][e & 63] = p.zf[(col0 + (e >> 6)) * p.Dp + slice * kST + (e & 63)];
      }
      __syncthreads();
#pragma unroll 8
      for (int j = 0; j < kST; ++j) {
        const float4 v = *reinterpret_cast<const float4*>(&Vs[j][tx * 4]);
        const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float pij = Ps[ty * 4 + i][j];
#pragma unroll
          for (int f = 0; f < 4; ++f) O[i][f] = fmaf(pij, vv[f], O[i][f]);
        }
      }
    }
  }

  // ---- block sums
  for (int o = 16; o > 0; o >>= 1) {
    acc_x += __shfl_xor_sync(0xffffffffu, acc_x, o);
    acc_y += __shfl_xor_sync(0xffffffffu, acc_y, o);
  }
  if ((tid & 31) == 0) { red[tid >> 5][0] = acc_x; red[tid >> 5][1] = acc_y; }
  __syncthreads();
  if (tid == 0 && slice == 0) {
    double sx = 0, sy = 0;
    for (int w = 0; w < 8; ++w) { sx += red[w][0]; sy += red[w][1]; }
    if (row_is_x) atomicAdd(&p.sums[0], sx);
    else { atomicAdd(&p.sums[1], sx); atomicAdd(&p.sums[2], sy); }
  }
  if (!pv) return;
  // row sums of P: reduce over the 16 tx lanes that share ty
#pragma unroll
  for (int i = 0; i < 4; ++i)
    for (int o = 8; o > 0; o >>= 1) rs[i] += __shfl_xor_sync(0xffffffffu, rs[i], o);
  const float ci4 = 4.f * (row_is_x ? cx : cy);
  const int64_t n_region = row_is_x ? p.Nx : p.Ny;
  float csum[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t rr = static_cast<int64_t>(row_is_x ? ti : ti - p.tiles_x) * kST + ty * 4 + i;
    if (rr >= n_region || rr < p.row_begin || rr >= p.row_end) continue;
    float* out = (row_is_x ? p.dX : p.dY) + (rr - p.row_begin) * p.D;
    const float* mul = (!row_is_x && p.dY_mul) ? p.dY_mul + (rr - p.row_begin) * p.D : nullptr;
    const float* zrow = p.zf + (row0 + ty * 4 + i) * p.Dp;
#pragma unroll
    for (int f = 0; f < 4; ++f) {
      const int64_t ff = static_cast<int64_t>(slice) * kST + tx * 4 + f;
      if (ff < p.D) {
        float g = ci4 * (O[i][f] - rs[i] * zrow[ff]);
        if (mul) g *= mul[ff];
        out[ff] = g;
        csum[f] += g;
      }
    }
  }
  // column sums of the Y rows written: the two ty halves of a warp first, then one atomic per (warp, column)
  if (p.colsum && !row_is_x) {
#pragma unroll
    for (int f = 0; f < 4; ++f) {
      const float t = csum[f] + __shfl_xor_sync(0xffffffffu, csum[f], 16);
      const int64_t ff = static_cast<int64_t>(slice) * kST + tx * 4 + f;
      if ((tid & 16) == 0 && ff < p.D) atomicAdd(p.colsum + ff, static_cast<double>(t));
    }
  }
}

int launch_simt(const SimtParams& p, cudaStream_t st) {
  const int ngrad = (p.gx_end - p.gx_begin) + (p.gy_end - p.gy_begin);
  const int grid = ngrad * p.nslices + p.n_xx_units;
  if (grid <= 0) return VGAN_B200_OK;
  mmd2_simt_kernel<<<grid, 256, 0, st>>>(p); count_launch();
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

// ------------------------------------------------------------------ dense K[n,n]
// K_ij = sum_m exp(-(|zi-zj|^2 + eps) * inv_s[m]); direct differences in fp32.
__global__ void __launch_bounds__(256)
matrix_kernel(const float* __restrict__ Z, int64_t n, int64_t D, const float* __restrict__ inv_scales,
              int n_scales, float eps, float* __restrict__ K) {
  __shared__ float zi[16][65], zj[16][65];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int64_t i = static_cast<int64_t>(blockIdx.y) * 16 + ty, j = static_cast<int64_t>(blockIdx.x) * 16 + tx;
  float d2 = 0.f;
  for (int64_t k0 = 0; k0 < D; k0 += 64) {
    for (int e = threadIdx.x; e < 16 * 64; e += 256) {
      const int r = e >> 6, k = e & 63;
      const int64_t gi = static_cast<int64_t>(blockIdx.y) * 16 + r, gj = static_cast<int64_t>(blockIdx.x) * 16 + r;
      zi[r][k] = (gi < n && k0 + k < D) ? Z[gi * D + k0 + k] : 0.f;
      zj[r][k] = (gj < n && k0 + k < D) ? Z[gj * D + k0 + k] : 0.f;
    }
    __syncthreads();
#pragma unroll 16
    for (int k = 0; k < 64; ++k) {
      const float df = zi[ty][k] - zj[tx][k];
      d2 = fmaf(df, df, d2);
    }
    __syncthreads();
  }
  if (i < n && j < n) {
    float kk = 0.f;
    for (int m = 0; m < n_scales; ++m) kk += exp2f(-(d2 + eps) * inv_scales[m] * kLog2e);
    K[i * n + j] = kk;
  }
}

int launch_matrix(const float* Z, int64_t n, int64_t D, const float* inv_scales, int n_scales,
                  float eps, float* K, cudaStream_t st) {
  dim3 grid(static_cast<unsigned>((n + 15) / 16), static_cast<unsigned>((n + 15) / 16));
  matrix_kernel<<<grid, 256, 0, st>>>(Z, n, D, inv_scales, n_scales, eps, K); count_launch();
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

}  // namespace vg

namespace vg {
int launch_simt_from_abi(const vgan_b200_layout_t& L, void* ws, int64_t row_begin, int64_t row_end, int xx_part,
                         int xx_nparts, float* dY, const float* dY_mul, float* dX, cudaStream_t st) {
  uint8_t* base = static_cast<uint8_t*>(ws);
  SimtParams p;
  memset(&p, 0, sizeof(p));
  const bool need_dx = (L.flags & VGAN_B200_NEED_DX) != 0 && dX != nullptr;
  p.tiles_x = static_cast<int>(L.Nxp / kST);
  p.tiles_y = static_cast<int>(L.Nyp / kST);
  p.pv = dY ? 1 : 0;
  p.nslices = p.pv ? static_cast<int>(L.Dp / kST) : 1;
  const int64_t ye = row_end < L.Ny ? row_end : L.Ny, xe = row_end < L.Nx ? row_end : L.Nx;
  p.gy_begin = static_cast<int>(row_begin / kST);
  p.gy_end = ye > row_begin ? static_cast<int>((ye + kST - 1) / kST) : p.gy_begin;
  if (need_dx) {
    p.gx_begin = static_cast<int>(row_begin / kST);
    p.gx_end = xe > row_begin ? static_cast<int>((xe + kST - 1) / kST) : p.gx_begin;
    p.n_xx_units = 0;
  } else {
    p.n_xx_units = xx_part < p.tiles_x ? (p.tiles_x - xx_part + xx_nparts - 1) / xx_nparts : 0;
  }
  p.xx_part = xx_part; p.xx_nparts = xx_nparts; p.n_scales = L.n_scales;
  p.Nx = L.Nx; p.Ny = L.Ny; p.D = L.D; p.Dp = L.Dp; p.row_begin = row_begin; p.row_end = row_end;
  p.zf = reinterpret_cast<const float*>(base + L.off_zf);
  p.nrm = reinterpret_cast<const float*>(base + L.off_nrm);
  p.inv_scales = reinterpret_cast<const float*>(base + L.off_par) + kParInvS;
  p.sums = reinterpret_cast<double*>(base + L.off_sums);
  p.colsum = p.pv ? reinterpret_cast<double*>(base + L.off_colsum) : nullptr;
  p.dY = dY; p.dY_mul = dY_mul; p.dX = dX;
  return launch_simt(p, st);
}
}  // namespace vg
