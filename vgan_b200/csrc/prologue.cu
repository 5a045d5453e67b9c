// K2 — HBM-bound prologue / epilogue kernels of one evaluation:
//   stats    per-block column sums and sums of squares of the rows in range (+ column maxima of U)
//   scales   closed-form bandwidth, 1/scales, parameter block, centring mean, common-mode constants, penalty value
//   pack     masking Y = U*X, centring, bf16 cast (or hi/lo split), row norms, exp2-scaled norms
//   finalize loss value, penalty gradient
// Replaces vstack / the pow-sum prologue of torch.cdist / `fake_subspaces*batch` / RBF.get_bandwidth / topk penalty
// (/root/reference/src/models/Mmd_loss_constrained.py:16-22,43,50; src/vgan.py:616).
#include "common.cuh"
#include "ptx.cuh"

namespace vg {

struct PackArgs {
  const float* X;
  const float* U;   // mask (Y = U*X) when Y == nullptr
  const float* Y;
  int64_t Nx, Ny, D, Dp, Nxp, Nyp, D_u;
  int64_t rb, re;   // row range of X and of Y handled by this call
  int center, fp32, split, ones;   // ones: constant-1 features at columns D .. D + ones - 1 of every valid row
  __nv_bfloat16* zq;
  __nv_bfloat16* zb;    // split operands (BF16X3): zq = row operand [hi | hi | lo], zb = column operand [hi | lo | hi],
                        // both [rows, 3 Dp]: sum_k zq[i,k] zb[j,k] = hi_i.hi_j + hi_i.lo_j + lo_i.hi_j
  float* zf;
  float* nrm;
  float* nrms;
  const float* mean;
  const float* par;
  double* stats;    // [0] sum |x|^2, [1] sum |y|^2, [2 + f] column sums of X, [2 + Dp + f] column sums of Y (raw)
  unsigned long long* cmax;   // packed column maxima of U, or nullptr
};

__device__ __forceinline__ uint32_t float_key(float v) {      // order-preserving float -> uint
  const uint32_t b = __float_as_uint(v);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float key_float(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}
// larger value wins; on equal values the LOWER row index wins (CPU torch.topk's choice)
__device__ __forceinline__ unsigned long long max_key(float v, int64_t row) {
  return (static_cast<unsigned long long>(float_key(v)) << 32) |
         static_cast<unsigned long long>(0xffffffffu - static_cast<uint32_t>(row));
}

// 2-D thread map of the statistics kernels (round 2: one thread per column walking 32-64 rows serially left 20 of 256
// threads busy at D = 20 and cost 30 us for 40 KB).  1024 threads: column c = tid % cp of the pass (cp = power of two
// <= 64 that covers the column items), row lane tid / cp (>= 16 of them); a thread strides the block's rows by the row
// lanes and the lanes meet through shared memory before the block's atomics.  Few, fat blocks: every block ends in
// atomics on the SAME addresses, which serialise at ~60 ns each (1024 blocks: 65 us at N = 65536, D = 256), so the
// grid is at most 64-128 blocks and the bytes in flight come from the 1024 threads.
constexpr int kStatThreads = 1024;
__host__ __device__ inline int pow2_cols(int64_t cols) {
  int p = 1;
  while (p < cols && p < 64) p <<= 1;
  return p;
}
// rows per block of a statistics kernel: at most max_blocks blocks, at least 4 strides of the row lanes, a multiple of
// the row lanes
static int64_t stat_rows_per_block(int64_t rows, int rl_n, int max_blocks) {
  int64_t rpb = (rows + max_blocks - 1) / max_blocks;
  if (rpb < 4 * rl_n) rpb = 4 * rl_n;
  return (rpb + rl_n - 1) / rl_n * rl_n;
}
__device__ __forceinline__ float block_sum_1024(float v, float* red) {     // red[32]; result valid in thread 0
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.f;
  if (threadIdx.x < 32) t = warp_sum(red[threadIdx.x]);
  return t;
}

// ---- generic statistics: rows q of the range's stack [X rows rb..xe ; Y rows rb..ye), and the column maxima of the
// U rows that belong to the block's X rows (cpu = pow2_cols(D_u))
__global__ void __launch_bounds__(kStatThreads) colstats_kernel(PackArgs a, int rows_per_block, int cp, int cpu) {
  __shared__ float part[2][kStatThreads];
  __shared__ unsigned long long kpart[kStatThreads];
  __shared__ float red[32];
  const int64_t xe = min(a.re, a.Nx), ye = min(a.re, a.Ny);
  const int64_t nX = max(xe - a.rb, static_cast<int64_t>(0)), nY = max(ye - a.rb, static_cast<int64_t>(0));
  const int64_t q0 = static_cast<int64_t>(blockIdx.x) * rows_per_block;
  const int64_t q1 = min(q0 + rows_per_block, nX + nY);
  const int64_t qx1 = min(q1, nX), qy0 = max(q0, nX);       // X rows [q0, qx1), Y rows [qy0, q1) of the stack
  const int tid = threadIdx.x;
  {
    const int c = tid & (cp - 1), rl = tid / cp, rl_n = kStatThreads / cp;
    float sqx = 0.f, sqy = 0.f;
    for (int64_t f0 = 0; f0 < a.D; f0 += cp) {
      const int64_t f = f0 + c;
      float cx = 0.f, cy = 0.f;
      if (f < a.D) {
#pragma unroll 4
        for (int64_t q = q0 + rl; q < qx1; q += rl_n) {
          const float v = a.X[(a.rb + q) * a.D + f];
          cx += v;
          sqx = fmaf(v, v, sqx);
        }
#pragma unroll 4
        for (int64_t q = qy0 + rl; q < q1; q += rl_n) {
          const int64_t y = a.rb + (q - nX);
          const float v = a.Y ? a.Y[y * a.D + f] : a.U[y * a.D + f] * a.X[y * a.D + f];
          cy += v;
          sqy = fmaf(v, v, sqy);
        }
      }
      __syncthreads();           // the previous column pass has been read
      part[0][tid] = cx;
      part[1][tid] = cy;
      __syncthreads();
      if (rl == 0 && f < a.D) {  // the row lanes meet
        for (int k = c + cp; k < kStatThreads; k += cp) { cx += part[0][k]; cy += part[1][k]; }
        if (q0 < nX) atomicAdd(&a.stats[2 + f], static_cast<double>(cx));
        if (q1 > nX) atomicAdd(&a.stats[2 + a.Dp + f], static_cast<double>(cy));
      }
    }
    const float tx = block_sum_1024(sqx, red);
    const float ty = block_sum_1024(sqy, red);
    if (tid == 0) {
      if (tx != 0.f) atomicAdd(&a.stats[0], static_cast<double>(tx));
      if (ty != 0.f) atomicAdd(&a.stats[1], static_cast<double>(ty));
    }
  }
  if (a.cmax && q0 < qx1) {
    const int c = tid & (cpu - 1), rl = tid / cpu, rl_n = kStatThreads / cpu;
    for (int64_t d0 = 0; d0 < a.D_u; d0 += cpu) {
      const int64_t d = d0 + c;
      unsigned long long best = 0ull;
      if (d < a.D_u) {
#pragma unroll 4
        for (int64_t q = q0 + rl; q < qx1; q += rl_n) {
          const unsigned long long key = max_key(a.U[(a.rb + q) * a.D_u + d], a.rb + q);
          best = key > best ? key : best;
        }
      }
      __syncthreads();
      kpart[tid] = best;
      __syncthreads();
      if (rl == 0 && d < a.D_u) {
        for (int k = c + cpu; k < kStatThreads; k += cpu) best = kpart[k] > best ? kpart[k] : best;
        atomicMax(&a.cmax[d], best);
      }
    }
  }
}

// ---- masked fast path (Y = U*X, Nx == Ny, D % 4 == 0, 16-byte aligned inputs): X and U are read ONCE per pass
// for both halves of the stack, with 16-byte loads; the column maxima of U ride along.  cp = pow2_cols(D / 4).
__global__ void __launch_bounds__(kStatThreads) colstats_masked_kernel(PackArgs a, int rows_per_block, int cp) {
  __shared__ __align__(16) unsigned char buf[2 * kStatThreads * 16];     // sums, then keys: 32 KiB
  __shared__ float red[32];
  float4* pcx = reinterpret_cast<float4*>(buf);
  float4* pcy = pcx + kStatThreads;
  unsigned long long* pkey = reinterpret_cast<unsigned long long*>(buf);   // [4][1024]
  const int tid = threadIdx.x;
  const int c = tid & (cp - 1), rl = tid / cp, rl_n = kStatThreads / cp;
  const int64_t r0 = a.rb + static_cast<int64_t>(blockIdx.x) * rows_per_block;
  const int64_t r1 = min(r0 + rows_per_block, min(a.re, a.Nx));
  const int64_t D4 = a.D >> 2;
  float sqx = 0.f, sqy = 0.f;
  for (int64_t f0 = 0; f0 < D4; f0 += cp) {
    const int64_t f4 = f0 + c;
    float4 cx = make_float4(0.f, 0.f, 0.f, 0.f), cy = cx;
    unsigned long long m0 = 0ull, m1 = 0ull, m2 = 0ull, m3 = 0ull;
    if (f4 < D4) {
      const float4* xp = reinterpret_cast<const float4*>(a.X) + (r0 + rl) * D4 + f4;
      const float4* up = reinterpret_cast<const float4*>(a.U) + (r0 + rl) * D4 + f4;
      const int64_t step = static_cast<int64_t>(rl_n) * D4;
#pragma unroll 4
      for (int64_t r = r0 + rl; r < r1; r += rl_n, xp += step, up += step) {
        const float4 x = __ldg(xp), u = __ldg(up);
        const float4 y = make_float4(u.x * x.x, u.y * x.y, u.z * x.z, u.w * x.w);
        cx.x += x.x; cx.y += x.y; cx.z += x.z; cx.w += x.w;
        cy.x += y.x; cy.y += y.y; cy.z += y.z; cy.w += y.w;
        sqx = fmaf(x.x, x.x, fmaf(x.y, x.y, fmaf(x.z, x.z, fmaf(x.w, x.w, sqx))));
        sqy = fmaf(y.x, y.x, fmaf(y.y, y.y, fmaf(y.z, y.z, fmaf(y.w, y.w, sqy))));
        if (a.cmax) {
          unsigned long long k;
          k = max_key(u.x, r); m0 = k > m0 ? k : m0;
          k = max_key(u.y, r); m1 = k > m1 ? k : m1;
          k = max_key(u.z, r); m2 = k > m2 ? k : m2;
          k = max_key(u.w, r); m3 = k > m3 ? k : m3;
        }
      }
    }
    const bool owner = rl == 0 && f4 < D4;
    __syncthreads();             // the previous column pass has been read
    pcx[tid] = cx; pcy[tid] = cy;
    __syncthreads();
    if (owner) {                 // the row lanes meet
      for (int k = c + cp; k < kStatThreads; k += cp) {
        const float4 x = pcx[k], y = pcy[k];
        cx.x += x.x; cx.y += x.y; cx.z += x.z; cx.w += x.w;
        cy.x += y.x; cy.y += y.y; cy.z += y.z; cy.w += y.w;
      }
      double* sx = a.stats + 2 + 4 * f4;
      double* sy = a.stats + 2 + a.Dp + 4 * f4;
      atomicAdd(sx + 0, static_cast<double>(cx.x)); atomicAdd(sx + 1, static_cast<double>(cx.y));
      atomicAdd(sx + 2, static_cast<double>(cx.z)); atomicAdd(sx + 3, static_cast<double>(cx.w));
      atomicAdd(sy + 0, static_cast<double>(cy.x)); atomicAdd(sy + 1, static_cast<double>(cy.y));
      atomicAdd(sy + 2, static_cast<double>(cy.z)); atomicAdd(sy + 3, static_cast<double>(cy.w));
    }
    if (a.cmax) {                // (block-uniform)
      __syncthreads();           // the sums have been read: the buffer now carries the keys
      pkey[tid] = m0; pkey[kStatThreads + tid] = m1; pkey[2 * kStatThreads + tid] = m2; pkey[3 * kStatThreads + tid] = m3;
      __syncthreads();
      if (owner) {
        for (int k = c + cp; k < kStatThreads; k += cp) {
          m0 = pkey[k] > m0 ? pkey[k] : m0;
          m1 = pkey[kStatThreads + k] > m1 ? pkey[kStatThreads + k] : m1;
          m2 = pkey[2 * kStatThreads + k] > m2 ? pkey[2 * kStatThreads + k] : m2;
          m3 = pkey[3 * kStatThreads + k] > m3 ? pkey[3 * kStatThreads + k] : m3;
        }
        atomicMax(a.cmax + 4 * f4 + 0, m0); atomicMax(a.cmax + 4 * f4 + 1, m1);
        atomicMax(a.cmax + 4 * f4 + 2, m2); atomicMax(a.cmax + 4 * f4 + 3, m3);
      }
    }
  }
  const float tx = block_sum_1024(sqx, red);
  const float ty = block_sum_1024(sqy, red);
  if (tid == 0) {
    atomicAdd(&a.stats[0], static_cast<double>(tx));
    atomicAdd(&a.stats[1], static_cast<double>(ty));
  }
}

// ------------------------------------------------------------------ scales
struct ScaleArgs {
  float mults[VGAN_B200_MAX_SCALES];
  int n_scales, frozen, center, common_mode, ones;
  int64_t Nx, Ny, D, Dp, Nxp, Nyp, D_u;
  float weight;
  const double* stats;
  const unsigned long long* cmax;
  float* bandwidth;
  float* par;
  float* mean;
  float* cm;
  float* misc;
};

__device__ __forceinline__ double block_sum_256d(double v, double* red) {   // red[8]; result valid in every thread
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();               // the previous sum has been read
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0.0;
#pragma unroll
  for (int w = 0; w < 8; ++w) t += red[w];
  return t;
}

// bandwidth = sum_ij |zi-zj|^2 / (n^2-n) = (2n sum|z|^2 - 2 |sum z|^2) / (n^2-n)   (Mmd_loss_constrained.py:16-22)
__global__ void __launch_bounds__(256) scales_kernel(ScaleArgs a) {
  __shared__ double red[8];
  __shared__ float sh[16];
  const double* csx = a.stats + 2;
  const double* csy = a.stats + 2 + a.Dp;
  double axx = 0.0, ayy = 0.0, axy = 0.0;
  for (int64_t f = threadIdx.x; f < a.D; f += 256) {
    const double x = csx[f], y = csy[f];
    axx += x * x; ayy += y * y; axy += x * y;
  }
  axx = block_sum_256d(axx, red);
  ayy = block_sum_256d(ayy, red);
  axy = block_sum_256d(axy, red);
  const double nx = static_cast<double>(a.Nx), ny = static_cast<double>(a.Ny), n = nx + ny;
  const double sqx = a.stats[0], sqy = a.stats[1];
  __shared__ float sh_inv[VGAN_B200_MAX_SCALES];
  __shared__ double sh_w[4][VGAN_B200_MAX_SCALES];
  if (threadIdx.x == 0) {
    float bw;
    if (!a.frozen) {
      bw = static_cast<float>((2.0 * n * (sqx + sqy) - 2.0 * (axx + 2.0 * axy + ayy)) / (n * n - n));
      *a.bandwidth = bw;
    } else {
      bw = *a.bandwidth;
    }
    sh[4] = bw;
    // the reference multiplies the 0-dim bandwidth with the fp32 multiplier vector -> fp32
    for (int m = 0; m < VGAN_B200_MAX_SCALES; ++m) sh_inv[m] = m < a.n_scales ? 1.0f / (bw * a.mults[m]) : 0.f;
  }
  if (threadIdx.x < kParFloats) a.par[threadIdx.x] = 0.f;
  __syncthreads();
  // common mode: w0_ab = w(mean squared distance between block a and block b); one (block pair, scale) per thread
  if (threadIdx.x < 4 * VGAN_B200_MAX_SCALES) {
    const int rc = threadIdx.x / VGAN_B200_MAX_SCALES, m = threadIdx.x % VGAN_B200_MAX_SCALES;
    double w = 0.0;
    if (a.common_mode && m < a.n_scales) {
      const double mx = sqx / nx, my = sqy / ny;
      const double d2 = rc == 0 ? 2.0 * (mx - axx / (nx * nx))
                                : (rc == 3 ? 2.0 * (my - ayy / (ny * ny)) : mx + my - 2.0 * axy / (nx * ny));
      w = exp(-fmax(d2, 0.0) * sh_inv[m]) * sh_inv[m];
    }
    sh_w[rc][m] = w;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const float bw = sh[4];
    float inv_s[VGAN_B200_MAX_SCALES];
    for (int m = 0; m < VGAN_B200_MAX_SCALES; ++m) inv_s[m] = sh_inv[m];
    const int ref = a.n_scales - 1;
    const float c0 = kLog2e * inv_s[ref];
    float* par = a.par;
    par[0] = c0;
    par[1] = -2.f * c0;
    const float cx = 1.f / static_cast<float>(a.Nx), cy = -1.f / static_cast<float>(a.Ny);
    for (int m = 0; m < a.n_scales; ++m) {
      par[2 + m] = inv_s[m] / inv_s[ref];
      par[kParCoefX + m] = inv_s[m] * cx;
      par[kParCoefY + m] = inv_s[m] * cy;
      par[kParInvS + m] = inv_s[m];
    }
    par[kParBw] = bw;
    float w0[2][2];
    for (int r = 0; r < 2; ++r)
      for (int c = 0; c < 2; ++c) {
        double w = 0.0;
        for (int m = 0; m < a.n_scales; ++m) w += sh_w[2 * r + c][m];
        w0[r][c] = static_cast<float>(w);
      }
    const float npx = static_cast<float>(a.Nxp - a.Nx), npy = static_cast<float>(a.Nyp - a.Ny);
    for (int r = 0; r < 2; ++r) {
      par[kParCw0 + 2 * r + 0] = cx * w0[r][0];
      par[kParCw0 + 2 * r + 1] = cy * w0[r][1];
      // the row sums the kernel forms in FP32 include the padding columns (P' = -c_j w0 there); read from the ones
      // column of O (VGAN_B200_ONES) they do not
      par[kParB + r] = (w0[r][0] - w0[r][1]) + (a.ones ? 0.f : cx * w0[r][0] * npx + cy * w0[r][1] * npy);
    }
    sh[0] = w0[0][0]; sh[1] = w0[0][1]; sh[2] = w0[1][0]; sh[3] = w0[1][1];
  }
  __syncthreads();
  // centring mean and the common-mode gradient vectors A_a[f] = w0_ax mu_x[f] - w0_ay mu_y[f] (means of the STORED rows)
  for (int64_t f = threadIdx.x; f < a.Dp; f += 256) {
    float mean = 0.f, ax = 0.f, ay = 0.f;
    if (f < a.D) {
      const double x = csx[f], y = csy[f];
      const double mu = a.center ? (x + y) / n : 0.0;
      const double mux = x / nx - mu, muy = y / ny - mu;
      mean = static_cast<float>(mu);
      ax = static_cast<float>(sh[0] * mux - sh[1] * muy);
      ay = static_cast<float>(sh[2] * mux - sh[3] * muy);
    }
    a.mean[f] = mean;
    a.cm[f] = ax;
    a.cm[a.Dp + f] = ay;
  }
  // penalty value  weight * mean_d(1 - max_i U[i,d])   (Mmd_loss_constrained.py:50)
  double pen = 0.0;
  const bool has_pen = a.cmax && a.weight != 0.f;
  if (has_pen)
    for (int64_t d = threadIdx.x; d < a.D_u; d += 256)
      pen += 1.0 - static_cast<double>(key_float(static_cast<uint32_t>(a.cmax[d] >> 32)));
  pen = block_sum_256d(pen, red);
  if (threadIdx.x == 0)
    a.misc[kMiscPenalty / 4] = has_pen ? static_cast<float>(a.weight * pen / static_cast<double>(a.D_u)) : 0.f;
}

// ------------------------------------------------------------------ pack
// one warp per stored (padded) row of the range's stack: stored X rows [x0, x1), stored Y rows [y0, y1)
// (indices inside their padded regions)
__global__ void __launch_bounds__(256) pack_kernel(PackArgs a, int64_t x0, int64_t x1, int64_t y0, int64_t y1) {
  const int lane = threadIdx.x & 31;
  const int64_t q = static_cast<int64_t>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  const int64_t nX = x1 - x0;
  if (q >= nX + (y1 - y0)) return;
  const bool in_x = q < nX;
  const int64_t local = in_x ? x0 + q : y0 + (q - nX);
  const int64_t r = in_x ? local : a.Nxp + local;            // stored row
  const bool valid = local < (in_x ? a.Nx : a.Ny);
  const float c0 = a.par[0];
  float nrm = 0.f;
  for (int64_t f = lane; f < a.Dp; f += 32) {
    float z = 0.f;
    if (valid && f < a.D) {
      const float raw = in_x ? a.X[local * a.D + f]
                             : (a.Y ? a.Y[local * a.D + f] : a.U[local * a.D + f] * a.X[local * a.D + f]);
      z = raw - a.mean[f];
    } else if (valid && f < a.D + a.ones) {
      z = 1.f;                       // VGAN_B200_ONES (split: hi = 1, lo = 0)
    }
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
      const __nv_bfloat16 qv = __float2bfloat16_rn(z);
      a.zq[r * a.Dp + f] = qv;
      const float zr = __bfloat162float(qv);
      nrm = fmaf(zr, zr, nrm);     // norm of the value the tensor core will see
    }
  }
  nrm = warp_sum(nrm);
  if (lane == 0) {
    nrm = valid ? nrm : kPadNorm;
    a.nrm[r] = nrm;
    a.nrms[r] = c0 * nrm;
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

// masked fast path: one warp per sample row r in [r0, r1): writes stored rows r (X half) and Nxp + r (Y half), bf16
__global__ void __launch_bounds__(256) pack_masked_kernel(PackArgs a, int64_t r0, int64_t r1) {
  const int lane = threadIdx.x & 31;
  const int64_t r = r0 + static_cast<int64_t>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (r >= r1) return;
  const bool valid = r < a.Nx;
  const int64_t D4 = a.D >> 2, Dp4 = a.Dp >> 2;
  const float c0 = a.par[0];
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
    } else if (valid && a.ones && f4 == D4) {
      ox = pack4_bf16(1.f, 1.f, 1.f, 0.f, nx);      // VGAN_B200_ONES: D % 4 == 0 here, the three ones fill this group
      oy = pack4_bf16(1.f, 1.f, 1.f, 0.f, ny);
    }
    zx[f4] = ox;
    zy[f4] = oy;
  }
  nx = warp_sum(nx);
  ny = warp_sum(ny);
  if (lane == 0) {
    nx = valid ? nx : kPadNorm;
    ny = valid ? ny : kPadNorm;
    a.nrm[r] = nx;
    a.nrm[a.Nxp + r] = ny;
    a.nrms[r] = c0 * nx;
    a.nrms[a.Nxp + r] = c0 * ny;
  }
}

// ------------------------------------------------------------------ finalize
// loss = S_xx/Nx^2 - 2 S_xy/(Nx Ny) + S_yy/Ny^2 + penalty; penalty gradient -weight/D_u at (argmax row, d)
__global__ void __launch_bounds__(256)
finalize_kernel(const double* __restrict__ sums, double nx, double ny, const unsigned long long* __restrict__ cmax,
                int64_t D_u, float weight, float* __restrict__ dU, int64_t row_begin, int64_t rows,
                double* __restrict__ colsum, float* __restrict__ misc, float* __restrict__ loss) {
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    const double mmd2 = sums[0] / (nx * nx) - 2.0 * sums[1] / (nx * ny) + sums[2] / (ny * ny);
    const float l = static_cast<float>(mmd2 + static_cast<double>(misc[kMiscPenalty / 4]));
    misc[kMiscLoss / 4] = l;
    if (loss) loss[0] = l;
  }
  if (!cmax || weight == 0.f) return;
  const float g = -weight / static_cast<float>(D_u);
  for (int64_t d = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; d < D_u;
       d += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t row = 0xffffffffu - static_cast<uint32_t>(cmax[d] & 0xffffffffu);
    if (dU && row >= row_begin && row < row_begin + rows) atomicAdd(&dU[(row - row_begin) * D_u + d], g);
    if (colsum) colsum[d] += static_cast<double>(g);
  }
}

// ------------------------------------------------------------------ host launchers
static PackArgs make_pack_args(const vgan_b200_layout_t& L, void* ws, const float* X, const float* U, const float* Y,
                               int64_t rb, int64_t re) {
  uint8_t* base = static_cast<uint8_t*>(ws);
  PackArgs a;
  a.X = X; a.U = U; a.Y = Y;
  a.Nx = L.Nx; a.Ny = L.Ny; a.D = L.D; a.Dp = L.Dp; a.Nxp = L.Nxp; a.Nyp = L.Nyp; a.D_u = L.D_u;
  a.rb = rb; a.re = re;
  a.center = (L.flags & VGAN_B200_CENTER) ? 1 : 0;
  a.fp32 = (L.flags & VGAN_B200_PREC_MASK) == VGAN_B200_PREC_FP32;
  a.split = (!a.fp32 && (L.flags & VGAN_B200_BF16X3)) ? 1 : 0;
  a.ones = (L.flags & VGAN_B200_ONES) ? 3 : 0;
  a.zq = reinterpret_cast<__nv_bfloat16*>(base + L.off_zq);
  a.zb = a.zq + static_cast<size_t>(L.Nxp + L.Nyp) * 3 * L.Dp;
  a.zf = reinterpret_cast<float*>(base + L.off_zf);
  a.nrm = reinterpret_cast<float*>(base + L.off_nrm);
  a.nrms = reinterpret_cast<float*>(base + L.off_nrms);
  a.mean = reinterpret_cast<const float*>(base + L.off_mean);
  a.par = reinterpret_cast<const float*>(base + L.off_par);
  a.stats = reinterpret_cast<double*>(base + L.off_stats);
  a.cmax = (L.D_u > 0 && U) ? reinterpret_cast<unsigned long long*>(base + L.off_cmax) : nullptr;
  return a;
}

static bool masked_fast(const PackArgs& a) {
  return !a.Y && a.U && !a.fp32 && !a.split && a.Nx == a.Ny && (a.D & 3) == 0 && (a.D_u == 0 || a.D_u == a.D) &&
         ((reinterpret_cast<uintptr_t>(a.X) | reinterpret_cast<uintptr_t>(a.U)) & 15) == 0;
}

int run_stats(const vgan_b200_layout_t& L, void* ws, const float* X, const float* U, const float* Y, int64_t rb,
              int64_t re, cudaStream_t st) {
  PackArgs a = make_pack_args(L, ws, X, U, Y, rb, re);
  VG_CUDA_CHECK(cudaMemsetAsync(static_cast<uint8_t*>(ws) + L.off_acc, 0, L.acc_bytes, st));
  int64_t nX = (re < L.Nx ? re : L.Nx) - rb, nY = (re < L.Ny ? re : L.Ny) - rb;
  if (nX < 0) nX = 0;
  if (nY < 0) nY = 0;
  const int64_t rows = nX + nY;
  if (rows <= 0) return VGAN_B200_OK;       // an empty shard only takes part in the exchanges
  // at most 64 (small problems) or 128 blocks: every block ends in 2 + 2-3 D atomics on the same addresses
  if (masked_fast(a)) {
    const int cp = pow2_cols(L.D >> 2);
    const int64_t rpb = stat_rows_per_block(nX, kStatThreads / cp, nX < 16384 ? 64 : 128);
    colstats_masked_kernel<<<static_cast<unsigned>((nX + rpb - 1) / rpb), kStatThreads, 0, st>>>(
        a, static_cast<int>(rpb), cp);
    count_launch();
  } else {
    const int cp = pow2_cols(L.D), cpu = pow2_cols(L.D_u > 0 ? L.D_u : 1);
    const int rl_max = kStatThreads / (cp < cpu ? cp : cpu);
    const int64_t rpb = stat_rows_per_block(rows, rl_max, rows < 16384 ? 64 : 128);
    colstats_kernel<<<static_cast<unsigned>((rows + rpb - 1) / rpb), kStatThreads, 0, st>>>(a, static_cast<int>(rpb),
                                                                                           cp, cpu);
    count_launch();
  }
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

int run_scales(const vgan_b200_layout_t& L, void* ws, float* bandwidth, int frozen, const float* mults_host,
               float weight, cudaStream_t st) {
  uint8_t* base = static_cast<uint8_t*>(ws);
  ScaleArgs a;
  for (int m = 0; m < VGAN_B200_MAX_SCALES; ++m) a.mults[m] = m < L.n_scales ? mults_host[m] : 1.f;
  a.n_scales = L.n_scales; a.frozen = frozen;
  a.center = (L.flags & VGAN_B200_CENTER) ? 1 : 0;
  a.common_mode = (L.flags & VGAN_B200_PREC_MASK) == VGAN_B200_PREC_BF16;   // the tcgen05 kernels (bf16, bf16x3)
  a.ones = (L.flags & VGAN_B200_ONES) ? 3 : 0;
  a.Nx = L.Nx; a.Ny = L.Ny; a.D = L.D; a.Dp = L.Dp; a.Nxp = L.Nxp; a.Nyp = L.Nyp; a.D_u = L.D_u;
  a.weight = weight;
  a.stats = reinterpret_cast<const double*>(base + L.off_stats);
  a.cmax = L.D_u > 0 ? reinterpret_cast<const unsigned long long*>(base + L.off_cmax) : nullptr;
  a.bandwidth = bandwidth;
  a.par = reinterpret_cast<float*>(base + L.off_par);
  a.mean = reinterpret_cast<float*>(base + L.off_mean);
  a.cm = reinterpret_cast<float*>(base + L.off_cm);
  a.misc = reinterpret_cast<float*>(base + L.off_misc);
  scales_kernel<<<1, 256, 0, st>>>(a); count_launch();
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

int run_pack(const vgan_b200_layout_t& L, void* ws, const float* X, const float* U, const float* Y, int64_t rb,
             int64_t re, cudaStream_t st) {
  PackArgs a = make_pack_args(L, ws, X, U, Y, rb, re);
  // stored rows of each block in range; the call that reaches the end of a block also writes its zero padding
  const int64_t x0 = rb < L.Nxp ? rb : L.Nxp, x1 = re >= L.Nx ? L.Nxp : (re > x0 ? re : x0);
  const int64_t y0 = rb < L.Nyp ? rb : L.Nyp, y1 = re >= L.Ny ? L.Nyp : (re > y0 ? re : y0);
  if (masked_fast(a)) {
    if (x1 > x0) {
      pack_masked_kernel<<<static_cast<unsigned>((x1 - x0 + 7) / 8), 256, 0, st>>>(a, x0, x1);
      count_launch();
    }
  } else {
    const int64_t rows = (x1 - x0) + (y1 - y0);
    if (rows > 0) {
      pack_kernel<<<static_cast<unsigned>((rows + 7) / 8), 256, 0, st>>>(a, x0, x1, y0, y1);
      count_launch();
    }
  }
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

int run_finalize(const vgan_b200_layout_t& L, void* ws, float weight, float* dU, int64_t row_begin, int64_t rows,
                 int colsum_penalty, float* loss, cudaStream_t st) {
  uint8_t* base = static_cast<uint8_t*>(ws);
  const bool pen = L.D_u > 0 && weight != 0.f;
  const unsigned blocks = pen ? static_cast<unsigned>((L.D_u + 255) / 256 < 64 ? (L.D_u + 255) / 256 : 64) : 1u;
  finalize_kernel<<<blocks, 256, 0, st>>>(
      reinterpret_cast<const double*>(base + L.off_sums), static_cast<double>(L.Nx), static_cast<double>(L.Ny),
      pen ? reinterpret_cast<const unsigned long long*>(base + L.off_cmax) : nullptr, L.D_u, weight, dU, row_begin, rows,
      (pen && colsum_penalty) ? reinterpret_cast<double*>(base + L.off_colsum) : nullptr,
      reinterpret_cast<float*>(base + L.off_misc), loss); count_launch();
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

}  // namespace vg
