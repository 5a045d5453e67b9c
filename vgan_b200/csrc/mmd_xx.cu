// K1s — kernel sums only (no gradient GEMM) on tcgen05, N = 256: the S_xx block of the no-KL path
// (Mmd_loss_constrained.py:46, needed for the loss value only) walks the upper triangle of X x X.
// One step = one row tile I against TWO column tiles (2q, 2q+1) that share every Z_I chunk:
//   S[128 x 256] = Z_I Z_{J,J+1}^T   4 x tcgen05.mma M=128 N=256 K=16 per 48 KiB stage (A 16 + B 32 KiB)
// An N=128 Gram stream is bound by the ISSUE rate of its one thread (~48 cycles per MMA + ~160 per
// stage against 256 execution cycles); at N=256 the same issue cost buys 512 execution cycles and the
// A operand is read once for two tiles.  TMEM = 2 x 256 columns (double-buffered S), no O.
// Epilogue = tc_epi math without P: u = c0 d2, t = 2^-u, ladder, k summed with tile weight
// 0 (below the diagonal / outside X), 1 (diagonal tile), 2 (above).
// Warps: 0-7 epilogue (warp%4 = lane quadrant, warp/4 = which of the two column tiles),
// 8 and 10 TMA producers (alternate stages), 9 MMA issuer, 11 TMEM allocator + column norms.
// Work partition: the (row tile, column pair) steps of this shard's row tiles form one flat sequence
// that the grid's CTAs cut into equal contiguous ranges (a range may span several row tiles), so the
// grid can be any size — all SMs when the kernel runs alone, the SMs the persistent gradient clusters
// leave free when it runs beside them.
#include "common.cuh"
#include "ptx.cuh"

namespace vg {

constexpr int kXStages = 4;
constexpr int kXStageBytes = 3 * kChunkBytes;      // Z_I chunk [128 x 64] + Z_J pair chunk [256 x 64]
constexpr int kXThreads = 384;

struct __align__(8) XxBarriers {
  uint64_t ring_full[kXStages], ring_empty[kXStages];
  uint64_t s_full[2], s_empty[2];
  uint64_t nj_full[2];
  uint32_t tmem_base;
  float red[8];
};

constexpr int kXSmemNj = kXStages * kXStageBytes;
constexpr int kXSmemBar = kXSmemNj + 2 * 2 * kTile * 4;
constexpr int kXSmemTotal = kXSmemBar + static_cast<int>(sizeof(XxBarriers));
static_assert(kXSmemTotal <= 232448, "shared memory budget");

// walks the flat step sequence: unit u = row tile xx_part + u * xx_nparts, column pairs [ti / 2, q_end)
struct XxIter {
  int ti, q, q_end, nparts;
  __device__ __forceinline__ void advance() {
    if (++q == q_end) {
      ti += nparts;
      q = ti >> 1;
    }
  }
};

__host__ __device__ inline long long xx_total_steps(int xx_part, int xx_nparts, int n_units, int tiles_x) {
  const int q_end = (tiles_x + 1) >> 1;
  long long t = 0;
  for (int u = 0; u < n_units; ++u) t += q_end - ((xx_part + u * xx_nparts) >> 1);
  return t;
}

// sum over 32 TMEM columns of k = sum_m exp(-d2 / s_m) for one row
template <bool LADDER>
__device__ __forceinline__ float epi_ksum_32(uint32_t taddr, const float* __restrict__ njp, float ni, float m2c0,
                                             const float (&ratio)[VGAN_B200_MAX_SCALES], int n_scales) {
  uint32_t v[32];
  tmem_ld_32x32(taddr, v);
  tmem_ld_wait();
  if constexpr (LADDER) {
    const uint64_t m2 = f2_pack(m2c0, m2c0), ni2 = f2_pack(ni, ni);
    uint64_t ks2 = f2_pack(0.f, 0.f);
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const float4 nj4 = *reinterpret_cast<const float4*>(njp + c * 4);
      const uint64_t nj2[2] = {f2_pack(nj4.x, nj4.y), f2_pack(nj4.z, nj4.w)};
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const uint64_t s2 = f2_pack(__uint_as_float(v[c * 4 + 2 * q]), __uint_as_float(v[c * 4 + 2 * q + 1]));
        float u0, u1;
        f2_unpack(f2_fma(s2, m2, f2_add(ni2, nj2[q])), u0, u1);
        u0 = fmaxf(u0, 0.f);
        u1 = fmaxf(u1, 0.f);
        const uint64_t t1 = f2_pack(ex2_approx(-u0), ex2_approx(-u1));
        const uint64_t t2 = f2_mul(t1, t1), t4 = f2_mul(t2, t2), t8 = f2_mul(t4, t4), t16 = f2_mul(t8, t8);
        ks2 = f2_add(ks2, f2_add(f2_add(t1, t2), f2_add(f2_add(t4, t8), t16)));
      }
    }
    float a, b;
    f2_unpack(ks2, a, b);
    return a + b;
  } else {
    float ks = 0.f;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const float4 nj4 = *reinterpret_cast<const float4*>(njp + c * 4);
      const float njv[4] = {nj4.x, nj4.y, nj4.z, nj4.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float uarg = fmaf(__uint_as_float(v[c * 4 + e]), m2c0, ni + njv[e]);
        uarg = fmaxf(uarg, 0.f);
        for (int m = 0; m < n_scales; ++m) ks += ex2_approx(-uarg * ratio[m]);
      }
    }
    return ks;
  }
}

template <bool LADDER>
__global__ void __launch_bounds__(kXThreads, 1)
mmd2_xx_kernel(const __grid_constant__ CUtensorMap zmap, const __grid_constant__ CUtensorMap zmap256,
               const TcParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  XxBarriers* bar = reinterpret_cast<XxBarriers*>(smem + kXSmemBar);
  float* nj_smem = reinterpret_cast<float*>(smem + kXSmemNj);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // this CTA's range of the flat step sequence, and the (row tile, column pair) it starts at
  const int q_end = (p.tiles_x + 1) >> 1;              // column-tile PAIRS [ti / 2, q_end) cover tiles ti .. tiles_x-1
  const long long total = xx_total_steps(p.xx_part, p.xx_nparts, p.n_xx_units, p.tiles_x);
  const long long s_begin = total * blockIdx.x / gridDim.x, s_end = total * (blockIdx.x + 1) / gridDim.x;
  const int nq = static_cast<int>(s_end - s_begin);
  XxIter it0{p.xx_part, p.xx_part >> 1, q_end, p.xx_nparts};
  {
    long long skip = s_begin;
    while (skip >= q_end - (it0.ti >> 1)) {
      skip -= q_end - (it0.ti >> 1);
      it0.ti += p.xx_nparts;
    }
    it0.q = (it0.ti >> 1) + static_cast<int>(skip);
  }
  const unsigned long long gt_begin = global_timer_ns();

  if (threadIdx.x == 0) {
    for (int s = 0; s < kXStages; ++s) {
      mbar_init(&bar->ring_full[s], 1);
      mbar_init(&bar->ring_empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&bar->s_full[b], 1);
      mbar_init(&bar->s_empty[b], 256);
      mbar_init(&bar->nj_full[b], 1);
    }
    fence_barrier_init();
  }
  if (warp == 11) {
    tmem_alloc(&bar->tmem_base, 512);
    tmem_relinquish();
  }
  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&zmap);
    tma_prefetch_desc(&zmap256);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;

  if (warp == 8 || warp == 10) {
    // ===================== TMA producers =====================
    const uint32_t pj = warp == 8 ? 0u : 1u;
    uint32_t s = 0, ph = 0;
    XxIter w = it0;
    for (int it = 0; it < nq; ++it, w.advance()) {
      const int jrow = w.q * 2 * kTile;
      const int ti = w.ti;
      for (int kc = 0; kc < p.gchunks; ++kc) {
        if ((s & 1u) == pj) {
          mbar_wait(&bar->ring_empty[s], ph ^ 1);
          uint8_t* dst = smem + s * kXStageBytes;
          if (lane == 0) {
            mbar_arrive_expect_tx(&bar->ring_full[s], kXStageBytes);
            tma_load_2d(dst, &zmap, &bar->ring_full[s], kc * kChunk, ti * kTile);
          } else if (lane == 1) {
            tma_load_2d(dst + kChunkBytes, &zmap256, &bar->ring_full[s], kc * kChunk, jrow);
          }
          __syncwarp();
        }
        if (++s == kXStages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 11) {
    // ===================== scaled column norms of the tile pair =====================
    if (lane == 0) {
      XxIter w = it0;
      for (int it = 0; it < nq; ++it, w.advance()) {
        const int b = it & 1;
        mbar_wait(&bar->s_empty[b], ((it >> 1) & 1) ^ 1);
        mbar_arrive_expect_tx(&bar->nj_full[b], 2 * kTile * 4);
        bulk_load_1d(nj_smem + b * 2 * kTile, p.nrms + static_cast<int64_t>(w.q) * 2 * kTile, 2 * kTile * 4,
                     &bar->nj_full[b]);
      }
    }
  } else if (warp == 9) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = make_idesc_bf16(kTile, 2 * kTile, 0, 0);
    constexpr uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
    const uint32_t ring_lo = (smem_u32(smem) & 0x3FFFFu) >> 4;
    uint32_t s = 0, ph = 0;
    for (int it = 0; it < nq; ++it) {
      const int b = it & 1;
      mbar_wait(&bar->s_empty[b], ((it >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem + b * 2 * kTile;
      for (int kc = 0; kc < p.gchunks; ++kc) {
        mbar_wait(&bar->ring_full[s], ph);
        tc_fence_after();
        const uint32_t a_lo = ring_lo + s * (kXStageBytes >> 4);
        const uint32_t b_lo = a_lo + (kChunkBytes >> 4);
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < kChunk / 16; ++kk) {
            const uint64_t adesc = (static_cast<uint64_t>(desc_hi) << 32) | (a_lo + kk * 2);
            const uint64_t bdesc = (static_cast<uint64_t>(desc_hi) << 32) | (b_lo + kk * 2);
            umma_bf16_ss(d_tmem, adesc, bdesc, idesc, (kc | kk) != 0);
          }
          umma_commit(&bar->ring_empty[s]);
          if (kc == p.gchunks - 1) umma_commit(&bar->s_full[b]);
        }
        __syncwarp();
        if (++s == kXStages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp < 8) {
    // ===================== epilogue =====================
    const int q = warp & 3, g = warp >> 2;
    const int row = q * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(q * 32) << 16;
    const float m2c0 = p.par[1];
    float ratio[VGAN_B200_MAX_SCALES];
#pragma unroll
    for (int m = 0; m < VGAN_B200_MAX_SCALES; ++m) ratio[m] = p.par[2 + m];
    float acc = 0.f, ni = 0.f;
    XxIter w = it0;
    int ti = -1;
    for (int it = 0; it < nq; ++it, w.advance()) {
      const int b = it & 1;
      if (w.ti != ti) {
        ti = w.ti;
        ni = p.nrms[static_cast<int64_t>(ti) * kTile + row];
      }
      const int jt = w.q * 2 + g;                      // this warp's column tile
      const float wgt = (jt < ti || jt >= p.tiles_x) ? 0.f : (jt == ti ? 1.f : 2.f);
      mbar_wait(&bar->nj_full[b], (it >> 1) & 1);
      mbar_wait(&bar->s_full[b], (it >> 1) & 1);
      tc_fence_after();
      const float* njp = nj_smem + b * 2 * kTile + g * kTile;
      float ks = 0.f;
#pragma unroll 1
      for (int cc = 0; cc < 4; ++cc)
        ks += epi_ksum_32<LADDER>(tmem + lane_addr + b * 2 * kTile + g * kTile + cc * 32, njp + cc * 32, ni, m2c0,
                                  ratio, p.n_scales);
      tc_fence_before();
      mbar_arrive(&bar->s_empty[b]);
      acc = fmaf(wgt, ks, acc);
    }
    acc = warp_sum(acc);
    if (lane == 0) bar->red[warp] = acc;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (threadIdx.x == 0) {
      double sx = 0.0;
      for (int wv = 0; wv < 8; ++wv) sx += bar->red[wv];
      atomicAdd(&p.sums[0], sx);
    }
  }
  tc_fence_before();
  __syncthreads();
#ifdef VGAN_B200_DEBUG
  if (p.timeline && threadIdx.x == 0) {
    atomicMin(&p.timeline[2], gt_begin);
    atomicMax(&p.timeline[3], global_timer_ns());
  }
#else
  (void)gt_begin;
#endif
  if (warp == 11) tmem_dealloc(tmem, 512);
}

// max_ctas > 0 bounds the grid (the SMs a concurrently running kernel leaves free); 0 = one CTA per SM
int launch_xx(vgan_b200_ctx* ctx, const TcParams& p, const CUtensorMap& map, const CUtensorMap& map256, int ladder,
              cudaStream_t st, int max_ctas) {
  if (p.n_xx_units <= 0) return VGAN_B200_OK;
  auto kern = ladder ? mmd2_xx_kernel<true> : mmd2_xx_kernel<false>;
  VG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kXSmemTotal));
  const int n_sms = ctx->n_sms;
  const long long total = xx_total_steps(p.xx_part, p.xx_nparts, p.n_xx_units, p.tiles_x);
  long long grid = max_ctas > 0 ? max_ctas : n_sms;
  if (grid > total) grid = total;
  if (grid < 1) return VGAN_B200_OK;
  kern<<<static_cast<unsigned>(grid), kXThreads, kXSmemTotal, st>>>(map, map256, p); count_launch();
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

}  // namespace vg
