// Permutation test of the unbiased MMD^2 statistic on the device: replaces
// torch_two_sample's Cython permutation_test_mat (an O(n_perm n^2) scalar loop on the host), reached from
// MMDStatistic.pval at /root/reference/src/vgan.py:425-428, 702-705.
// For a relabeling with membership vector m (m_i = 1: point i belongs to sample 1)
//   aKa = sum_ij m_i m_j K_ij,  aK1 = sum_i m_i rowsum_i,  d1 = sum_i m_i K_ii
//   k11 = aKa - d1,  k12 = aK1 - aKa,  k22 = (total - 2 aK1 + aKa) - (trace - d1)
//   stat = a00 k11 + a11 k22 + 2 a01 k12,   a00 = 1/(n1(n1-1)), a11 = 1/(n2(n2-1)), a01 = -1/(n1 n2)
// Only aKa is O(n^2) per relabeling.  One CTA evaluates 8 relabelings at once: their membership bits are packed
// into one byte per point in shared memory, so every K_ij (read once per CTA, L2 resident: n = 1000 -> 4 MB)
// feeds 8 conditional accumulations.  The fp32 partial sums of one row (n / 256 terms per thread) are flushed into
// fp64 accumulators after every row.
#include "common.cuh"

namespace vg {

constexpr int kPermPerCta = 8;
constexpr int kPermThreads = 256;

// rowsum[i] = sum_j K_ij, diag[i] = K_ii, tot[0] += total, tot[1] += trace   (one warp per row)
__global__ void perm_rowstats_kernel(const float* __restrict__ K, int64_t n, double* __restrict__ rowsum,
                                     double* __restrict__ diag, double* __restrict__ tot) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (i >= n) return;
  const int lane = threadIdx.x & 31;
  const float* row = K + i * n;
  double s = 0.0;
  for (int64_t j = lane; j < n; j += 32) s += static_cast<double>(row[j]);
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) {
    const double d = static_cast<double>(row[i]);
    rowsum[i] = s;
    diag[i] = d;
    atomicAdd(&tot[0], s);
    atomicAdd(&tot[1], d);
  }
}

__device__ __forceinline__ double block_sum(double v, double* red) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0.0;
  for (int w = 0; w < kPermThreads / 32; ++w) t += red[w];
  return t;
}

__global__ void __launch_bounds__(kPermThreads)
perm_stat_kernel(const float* __restrict__ K, int64_t n, const uint8_t* __restrict__ member, int64_t n_perm,
                 const double* __restrict__ rowsum, const double* __restrict__ diag, const double* __restrict__ tot,
                 double a00, double a11, double a01, double* __restrict__ stats) {
  extern __shared__ uint8_t bits[];          // bits[i]: bit b = membership of point i under relabeling p0 + b
  __shared__ double red[kPermThreads / 32];
  const int64_t p0 = static_cast<int64_t>(blockIdx.x) * kPermPerCta;
  const int np = static_cast<int>(n_perm - p0 < kPermPerCta ? n_perm - p0 : kPermPerCta);
  for (int64_t i = threadIdx.x; i < n; i += kPermThreads) {
    unsigned m = 0;
    for (int b = 0; b < np; ++b) m |= (member[(p0 + b) * n + i] != 0 ? 1u : 0u) << b;
    bits[i] = static_cast<uint8_t>(m);
  }
  __syncthreads();

  double aKa[kPermPerCta], aK1[kPermPerCta], d1[kPermPerCta];
  float part[kPermPerCta];
#pragma unroll
  for (int b = 0; b < kPermPerCta; ++b) { aKa[b] = 0.0; aK1[b] = 0.0; d1[b] = 0.0; part[b] = 0.f; }

  for (int64_t i = 0; i < n; ++i) {
    const unsigned mi = bits[i];             // block-uniform
    if (mi != 0) {
      const float* row = K + i * n;
      for (int64_t j = threadIdx.x; j < n; j += kPermThreads) {
        const unsigned both = mi & bits[j];
        const float kv = row[j];
#pragma unroll
        for (int b = 0; b < kPermPerCta; ++b) part[b] += ((both >> b) & 1u) ? kv : 0.f;
      }
#pragma unroll
      for (int b = 0; b < kPermPerCta; ++b) { aKa[b] += static_cast<double>(part[b]); part[b] = 0.f; }
    }
  }
  for (int64_t i = threadIdx.x; i < n; i += kPermThreads) {
    const unsigned mi = bits[i];
    const double rs = rowsum[i], dg = diag[i];
#pragma unroll
    for (int b = 0; b < kPermPerCta; ++b)
      if ((mi >> b) & 1u) { aK1[b] += rs; d1[b] += dg; }
  }
  const double total = tot[0], trace = tot[1];
#pragma unroll
  for (int b = 0; b < kPermPerCta; ++b) {
    const double sa = block_sum(aKa[b], red);
    const double s1 = block_sum(aK1[b], red);
    const double sd = block_sum(d1[b], red);
    if (threadIdx.x == 0 && b < np) {
      const double k11 = sa - sd;
      const double k12 = s1 - sa;
      const double k22 = (total - 2.0 * s1 + sa) - (trace - sd);
      stats[p0 + b] = a00 * k11 + a11 * k22 + 2.0 * a01 * k12;
    }
  }
}

size_t perm_scratch_bytes(int64_t n) { return static_cast<size_t>(2 * n + 2) * sizeof(double); }

int run_permutation_stats(const float* K, int64_t n, int64_t n1, const uint8_t* member, int64_t n_perm,
                          void* scratch, double* stats, cudaStream_t st) {
  double* rowsum = static_cast<double*>(scratch);
  double* diag = rowsum + n;
  double* tot = diag + n;
  VG_CUDA_CHECK(cudaMemsetAsync(tot, 0, 2 * sizeof(double), st));
  const int rows_per_cta = 256 / 32;
  perm_rowstats_kernel<<<static_cast<unsigned>((n + rows_per_cta - 1) / rows_per_cta), 256, 0, st>>>(K, n, rowsum,
                                                                                                      diag, tot); count_launch();
  VG_CUDA_CHECK(cudaGetLastError());
  const int64_t n2 = n - n1;
  const double a00 = 1.0 / (static_cast<double>(n1) * static_cast<double>(n1 - 1));
  const double a11 = 1.0 / (static_cast<double>(n2) * static_cast<double>(n2 - 1));
  const double a01 = -1.0 / (static_cast<double>(n1) * static_cast<double>(n2));
  const size_t smem = static_cast<size_t>(n);
  if (smem > 48 * 1024)
    VG_CUDA_CHECK(cudaFuncSetAttribute(perm_stat_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       static_cast<int>(smem)));
  perm_stat_kernel<<<static_cast<unsigned>((n_perm + kPermPerCta - 1) / kPermPerCta), kPermThreads, smem, st>>>(
      K, n, member, n_perm, rowsum, diag, tot, a00, a11, a01, stats); count_launch();
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

}  // namespace vg
