// K1 — fused MMD^2 forward+backward on tcgen05 tensor cores (bf16 operands, fp32 TMEM
// accumulation).  Replaces RBF.forward + MMDLossConstrained.forward + autograd backward
// (/root/reference/src/models/Mmd_loss_constrained.py:24-26, 42-50).
//
// One CTA = one work unit = (row tile I of 128 stacked rows, feature slice p <= 256 wide):
//   for every column tile J (128 stacked rows):
//     S   = Z_I Z_J^T                    tcgen05.mma, K streamed in 64-wide TMA chunks
//     d2  = |z_i|^2 + |z_j|^2 - 2 S      epilogue warps, straight out of TMEM
//     t   = exp(-d2/s_4); k = t+t^2+t^4+t^8+t^16 ; w = sum_m t^(2^q)/s_m   (x2 ladder)
//     sums += k (S_xx / S_xy / S_yy by tile type) ; P = c_j w  -> bf16 -> smem
//     O_I[:, slice p] += P Z_J[:, slice p]          second tcgen05.mma, O stays in TMEM
//   dZ_I[:, slice p] = 4 c_i (O - rowsum(P) z_i)
// The pairwise matrix only ever exists as a 128x128 TMEM tile.
//
// Warp roles (640 threads): warps 0-15 epilogue (warp%4 = TMEM lane quadrant, warp/4 = column quarter of 32),
// warp 16 = TMA producer (Gram operands and V slices), warp 17 = MMA issuer, warp 18 = column norms,
// warp 19 = TMEM allocator.  This kernel serves D_p <= 256, where the epilogue (FP32 + MUFU issue) is the bound:
// with 8 epilogue warps (2 per scheduler) the issue slots were 49 % busy (ncu, profiles/r1d_base_d64_summary.txt:
// fixed-latency dependency stalls with nothing else to issue); 4 warps per scheduler hide them.  setmaxnreg moves
// registers from the 4 service warps (96 -> 56) to the epilogue warps (96 -> 104).
#include "common.cuh"
#include "ptx.cuh"
#include "tc_epi.cuh"

namespace vg {

constexpr int kStagesMax = 5;
constexpr int kStageBytes = 2 * kChunkBytes;   // Gram: Z_I chunk + Z_J chunk;  PV: two V chunks (128 features)
constexpr int kPBytes = 2 * kChunkBytes;                     // 128 x 128 bf16
constexpr int kNumEpiWarps = 16;
constexpr int kNumEpiThreads = kNumEpiWarps * 32;
constexpr int kThreads = kNumEpiThreads + 128;
constexpr int kWarpTma = kNumEpiWarps, kWarpMma = kNumEpiWarps + 1, kWarpNorms = kNumEpiWarps + 2,
              kWarpAlloc = kNumEpiWarps + 3;

struct __align__(8) TcBarriers {
  uint64_t ring_full[kStagesMax], ring_empty[kStagesMax];
  uint64_t s_full[2], s_empty[2];
  uint64_t nj_full[2];
  uint64_t p_full[2], p_empty[2];
  uint64_t o_full;
  uint32_t tmem_base;
  float red[kNumEpiWarps][2];
};

// SPLIT (fp32-class results from split bf16 operands z = hi + lo, BF16X3) in ONE pass: the K = 3 D_p Gram
// [hi|hi|lo] x [hi|lo|hi] is computed once per column tile, the epilogue parks BOTH bf16(P) and the remainder
// P - bf16(P), and the gradient GEMM issues its three parts O += P_hi V_hi + P_hi V_lo + P_lo V_hi back to back
// (three passes of the plain kernel recomputed the Gram and the whole epilogue three times: 3x the time where the
// epilogue is the bound, i.e. at every width this kernel serves).  Four P buffers instead of two, three ring stages
// instead of five.
template <bool SPLIT> struct TcCfg {
  static constexpr int Stages = SPLIT ? 3 : 5;
  static constexpr int PBufs = SPLIT ? 4 : 2;        // [hi 0, hi 1] (+ [lo 0, lo 1])
  static constexpr int PvParts = SPLIT ? 3 : 1;
  static constexpr int SmemRing = 0;
  static constexpr int SmemP = SmemRing + Stages * kStageBytes;
  static constexpr int SmemNj = SmemP + PBufs * kPBytes;
  static constexpr int SmemBar = SmemNj + 2 * kTile * 4;
  static constexpr int SmemTotal = SmemBar + static_cast<int>(sizeof(TcBarriers));
  static_assert(SmemTotal <= 232448, "shared memory budget");
};

struct Unit {
  int ti;        // stacked row tile
  int slice;     // feature slice
  int j_begin, j_end;
  int pv;        // gradient GEMM on
  int row_is_x;  // row tile lies in the X region
  int sym;       // S_xx-only unit walking the upper triangle: off-diagonal tiles count twice
};

__device__ __forceinline__ Unit decode_unit(const TcParams& p) {
  Unit u;
  const int ngx = p.gx_end - p.gx_begin;
  const int ngy = p.gy_end - p.gy_begin;
  const int n_grad_units = (ngx + ngy) * p.nslices;
  int b = blockIdx.x;
  if (b < n_grad_units) {
    int t = b / p.nslices;
    u.slice = b - t * p.nslices;
    u.row_is_x = t < ngx;
    u.ti = u.row_is_x ? p.gx_begin + t : p.tiles_x + p.gy_begin + (t - ngx);
    u.j_begin = 0;
    u.j_end = p.tiles_x + p.tiles_y;
    u.pv = p.pv;
    u.sym = 0;
  } else {
    int x = b - n_grad_units;
    u.ti = p.xx_part + x * p.xx_nparts;
    u.slice = 0;
    u.row_is_x = 1;
    u.j_begin = u.ti;      // upper triangle of the symmetric X x X block
    u.j_end = p.tiles_x;
    u.pv = 0;
    u.sym = 1;
  }
  return u;
}

template <bool LADDER, bool SPLIT>
__global__ void __launch_bounds__(kThreads, 1)
mmd2_tc_kernel(const __grid_constant__ CUtensorMap zmap, const __grid_constant__ CUtensorMap zmapB,
               const __grid_constant__ CUtensorMap zmapV, const TcParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];   // SWIZZLE_128B tiles need 1024-byte alignment
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  using Cfg = TcCfg<SPLIT>;
  constexpr int kStages = Cfg::Stages, kSmemRing = Cfg::SmemRing, kSmemP = Cfg::SmemP, kSmemNj = Cfg::SmemNj,
                kSmemBar = Cfg::SmemBar, kPvParts = Cfg::PvParts;
  TcBarriers* bar = reinterpret_cast<TcBarriers*>(smem + kSmemBar);
  float* nj_smem = reinterpret_cast<float*>(smem + kSmemNj);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const Unit u = decode_unit(p);
  const int n_tiles = u.j_end - u.j_begin;
  const int slice_chunks =
      min(p.slice_chunks, p.kchunks - u.slice * p.slice_chunks);  // chunks in this slice

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&bar->ring_full[s], 1);
      mbar_init(&bar->ring_empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&bar->s_full[b], 1);
      mbar_init(&bar->s_empty[b], kNumEpiThreads);
      mbar_init(&bar->nj_full[b], 1);
      mbar_init(&bar->p_full[b], kNumEpiThreads);
      mbar_init(&bar->p_empty[b], 1);
    }
    mbar_init(&bar->o_full, 1);
    fence_barrier_init();
  }
  if (warp == kWarpAlloc) {
    tmem_alloc(&bar->tmem_base, 512);
    tmem_relinquish();
  }
  if (warp == kWarpTma && lane == 0) {
    tma_prefetch_desc(&zmap);
    tma_prefetch_desc(&zmapB);
    tma_prefetch_desc(&zmapV);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;
  const uint32_t tmem_o = tmem + 256;

  if (warp >= kNumEpiWarps) setmaxnreg_dec<56>();
  if (warp == kWarpTma) {
    // ===================== TMA producer: one ring for Gram operands and V slices ==========
    // Stages are issued in exactly the order the MMA thread consumes them:
    //   Gram(0), Gram(1), PV(0), Gram(2), PV(1), ...
    if (lane == 0) {
      uint32_t s = 0, ph = 0;
      const int pv_stages = (slice_chunks + 1) >> 1;
      for (int n = 0; n <= n_tiles; ++n) {
        if (n < n_tiles) {
          const int jt = u.j_begin + n;
          for (int kc = 0; kc < p.gchunks; ++kc) {
            mbar_wait(&bar->ring_empty[s], ph ^ 1);
            uint8_t* dst = smem + kSmemRing + s * kStageBytes;
            mbar_arrive_expect_tx(&bar->ring_full[s], kStageBytes);
            tma_load_2d(dst, &zmap, &bar->ring_full[s], kc * kChunk, u.ti * kTile);
            tma_load_2d(dst + kChunkBytes, &zmapB, &bar->ring_full[s], kc * kChunk, jt * kTile);
            if (++s == kStages) { s = 0; ph ^= 1; }
          }
        }
        if (u.pv && n >= 1) {
          const int jt = u.j_begin + n - 1;
          for (int part = 0; part < kPvParts; ++part) {
            // V operand of this part: the hi block (chunk 0) or the lo block (chunk 2 kchunks) of the row-operand array
            const int vbase = SPLIT ? (part == 1 ? 2 * p.kchunks : 0) : p.v_chunk0;
            for (int h = 0; h < pv_stages; ++h) {
              const int nch = min(2, slice_chunks - 2 * h);
              mbar_wait(&bar->ring_empty[s], ph ^ 1);
              uint8_t* dst = smem + kSmemRing + s * kStageBytes;
              mbar_arrive_expect_tx(&bar->ring_full[s], nch * kChunkBytes);
              for (int c = 0; c < nch; ++c)
                tma_load_2d(dst + c * kChunkBytes, &zmapV, &bar->ring_full[s],
                            (vbase + u.slice * p.slice_chunks + 2 * h + c) * kChunk, jt * kTile);
              if (++s == kStages) { s = 0; ph ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == kWarpNorms) {
    // ===================== bulk-copy producer: scaled column norms =====================
    if (lane == 0) {
      for (int n = 0; n < n_tiles; ++n) {
        const int jt = u.j_begin + n;
        const int b = n & 1;
        mbar_wait(&bar->s_empty[b], ((n >> 1) & 1) ^ 1);   // epilogue done with nj[b]
        mbar_arrive_expect_tx(&bar->nj_full[b], kTile * 4);
        bulk_load_1d(nj_smem + b * kTile, p.nrms + static_cast<int64_t>(jt) * kTile, kTile * 4,
                     &bar->nj_full[b]);
      }
    }
  } else if (warp == kWarpMma) {
    // ===================== MMA issuer =====================
    // The whole warp runs the (warp-uniform) control flow so that addresses and descriptors live
    // in uniform registers; one elected lane issues the tcgen05 instructions.
    constexpr uint32_t idesc_gram = make_idesc_bf16(kTile, kTile, 0, 0);
    // descriptor high word: SBO = 1024 B, version 1, SWIZZLE_128B; low word: address and LBO
    constexpr uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
    constexpr uint32_t lbo_v = (static_cast<uint32_t>(kChunkBytes) >> 4) << 16;   // MN-major V: next 64 features
    const uint32_t ring_lo = (smem_u32(smem + kSmemRing) & 0x3FFFFu) >> 4;
    const uint32_t p_lo0 = (smem_u32(smem + kSmemP) & 0x3FFFFu) >> 4;
    const int pv_stages = (slice_chunks + 1) >> 1;
    uint32_t s = 0, ph = 0;
    for (int n = 0; n <= n_tiles; ++n) {
      if (n < n_tiles) {
        // ---- Gram(n): S[b] = Z_I Z_J^T over all K chunks
        const int b = n & 1;
        mbar_wait(&bar->s_empty[b], ((n >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem + b * kTile;
        for (int kc = 0; kc < p.gchunks; ++kc) {
          mbar_wait(&bar->ring_full[s], ph);
          tc_fence_after();
          const uint32_t a_lo = ring_lo + s * (kStageBytes >> 4);
          const uint32_t b_lo = a_lo + (kChunkBytes >> 4);
          if (elect_one()) {
#pragma unroll
            for (int kk = 0; kk < kChunk / 16; ++kk) {
              const uint64_t adesc = (static_cast<uint64_t>(desc_hi) << 32) | (a_lo + kk * 2);
              const uint64_t bdesc = (static_cast<uint64_t>(desc_hi) << 32) | (b_lo + kk * 2);
              umma_bf16_ss(d_tmem, adesc, bdesc, idesc_gram, (kc | kk) != 0);
            }
            umma_commit(&bar->ring_empty[s]);
          }
          __syncwarp();
          if (++s == kStages) { s = 0; ph ^= 1; }
        }
        if (elect_one()) umma_commit(&bar->s_full[b]);
        __syncwarp();
      }
      if (u.pv && n >= 1) {
        // ---- PV(n-1): O += P Z_J[:, slice]   (two 64-feature chunks = 128 columns of O per stage)
        const int m = n - 1, pb = m & 1;
        mbar_wait(&bar->p_full[pb], (m >> 1) & 1);
        tc_fence_after();
        for (int part = 0; part < kPvParts; ++part) {
          // A operand: bf16(P) (parts 0, 1) or the remainder tile (part 2: buffers 2, 3)
          const uint32_t p_lo = p_lo0 + (pb + (part == 2 ? 2 : 0)) * (kPBytes >> 4);
          for (int h = 0; h < pv_stages; ++h) {
            const int nch = min(2, slice_chunks - 2 * h);
            const uint32_t idesc_pv = make_idesc_bf16(kTile, nch * kChunk, 0, 1);
            mbar_wait(&bar->ring_full[s], ph);
            tc_fence_after();
            const uint32_t v_lo = ring_lo + s * (kStageBytes >> 4);
            if (elect_one()) {
#pragma unroll
              for (int kk = 0; kk < kTile / 16; ++kk) {
                const uint64_t adesc = (static_cast<uint64_t>(desc_hi) << 32) |
                                       (p_lo + (kk >> 2) * (kChunkBytes >> 4) + (kk & 3) * 2);
                const uint64_t bdesc = (static_cast<uint64_t>(desc_hi) << 32) | (v_lo + kk * (2048 >> 4)) | lbo_v;
                umma_bf16_ss(tmem_o + h * 2 * kChunk, adesc, bdesc, idesc_pv, (m | part | kk) != 0);
              }
              umma_commit(&bar->ring_empty[s]);
            }
            __syncwarp();
            if (++s == kStages) { s = 0; ph ^= 1; }
          }
        }
        if (elect_one()) umma_commit(&bar->p_empty[pb]);
        __syncwarp();
      }
    }
    if (u.pv && elect_one()) umma_commit(&bar->o_full);
    __syncwarp();
  } else if (warp < kNumEpiWarps) {
    // ===================== epilogue warps =====================
    setmaxnreg_inc<104>();   // 512 x 8 registers <= the 128 x 40 the service warps released
    const int q = warp & 3, g = warp >> 2;                 // lane quadrant, column quarter
    const int row = q * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(q * 32) << 16;
    const float c0 = p.par[0], m2c0 = p.par[1];
    const float ni = p.nrms[static_cast<int64_t>(u.ti) * kTile + row];
    float coef_x[VGAN_B200_MAX_SCALES], coef_y[VGAN_B200_MAX_SCALES], ratio[VGAN_B200_MAX_SCALES];
#pragma unroll
    for (int m = 0; m < VGAN_B200_MAX_SCALES; ++m) {
      ratio[m] = p.par[2 + m];
      coef_x[m] = p.par[kParCoefX + m];
      coef_y[m] = p.par[kParCoefY + m];
    }
    // common-mode level c_j w0 of (this row block, X columns) / (this row block, Y columns); unused without PV
    const float cw0_x = u.pv ? p.par[kParCw0 + (u.row_is_x ? 0 : 2)] : 0.f;
    const float cw0_y = u.pv ? p.par[kParCw0 + (u.row_is_x ? 0 : 2) + 1] : 0.f;
    (void)c0;
    float acc_x = 0.f, acc_y = 0.f;   // sum of k over X / Y columns (weighted by symmetry)
    float rsum = 0.f;                 // sum_j P_ij over this thread's column half

    for (int n = 0; n < n_tiles; ++n) {
      const int jt = u.j_begin + n;
      const int b = n & 1;
      const bool col_is_x = jt < p.tiles_x;
      const float fac = (u.sym && jt != u.ti) ? 2.f : 1.f;
      float cf[VGAN_B200_MAX_SCALES];
#pragma unroll
      for (int m = 0; m < VGAN_B200_MAX_SCALES; ++m) cf[m] = col_is_x ? coef_x[m] : coef_y[m];
      mbar_wait(&bar->nj_full[b], (n >> 1) & 1);
      mbar_wait(&bar->s_full[b], (n >> 1) & 1);
      tc_fence_after();
      uint8_t* p_row = smem + kSmemP + b * kPBytes + (g >> 1) * kChunkBytes + row * 128;
      const float* njp = nj_smem + b * kTile + g * 32;
      float ksum = 0.f;
      uint4 pk[4], pk_lo[4];
      epi_half_tile<LADDER>(tmem + lane_addr + b * kTile + g * 32, njp, ni, m2c0, cf, ratio, p.n_scales, ksum, rsum, pk,
                            p.p_lo != 0, col_is_x ? cw0_x : cw0_y, SPLIT ? pk_lo : nullptr);
      if (u.pv) {
        mbar_wait(&bar->p_empty[b], ((n >> 1) & 1) ^ 1);   // PV(n-2) done with P[b]
#pragma unroll
        for (int i = 0; i < 4; ++i)                          // 128B-swizzled K-major A operand
          *reinterpret_cast<uint4*>(p_row + ((((g & 1) * 4 + i) ^ (row & 7)) << 4)) = pk[i];
        if constexpr (SPLIT) {
#pragma unroll
          for (int i = 0; i < 4; ++i)                        // the remainder tile, two buffers further on
            *reinterpret_cast<uint4*>(p_row + 2 * kPBytes + ((((g & 1) * 4 + i) ^ (row & 7)) << 4)) = pk_lo[i];
        }
      }
      tc_fence_before();
      mbar_arrive(&bar->s_empty[b]);
      if (u.pv) {
        fence_proxy_async_smem();
        mbar_arrive(&bar->p_full[b]);
      }
      if (col_is_x) acc_x = fmaf(ksum, fac, acc_x); else acc_y += ksum;
#ifdef VGAN_B200_DEBUG
      if (p.dbg) atomicAdd(&p.dbg[static_cast<int64_t>(blockIdx.x) * p.dbg_stride + n], ksum);
#endif
    }

    // ---- block sums (only slice 0 owns them)
    acc_x = warp_sum(acc_x);
    acc_y = warp_sum(acc_y);
    if (lane == 0) {
      bar->red[warp][0] = acc_x;
      bar->red[warp][1] = acc_y;
    }
    asm volatile("bar.sync 1, %0;" ::"n"(kNumEpiThreads) : "memory");
    if (threadIdx.x == 0 && u.slice == 0 && !(p.pass_flags & kPassNoSums)) {
      double sx = 0.0, sy = 0.0;
      for (int wv = 0; wv < kNumEpiWarps; ++wv) {
        sx += bar->red[wv][0];
        sy += bar->red[wv][1];
      }
      if (u.row_is_x) {
        atomicAdd(&p.sums[0], sx);            // X rows x X cols -> S_xx (x Y cols is S_xy again)
      } else {
        atomicAdd(&p.sums[1], sx);            // Y rows x X cols -> S_xy
        atomicAdd(&p.sums[2], sy);            // Y rows x Y cols -> S_yy
      }
    }

    // ---- gradient rows: dZ_i[slice] = 4 c_i (O_i - rowsum(P)_i z_i)  (x dY_mul)
    if (u.pv) {
      mbar_wait(&bar->o_full, 0);                       // every PV has completed: the P buffers are dead
      tc_fence_after();
      float* rs_smem = reinterpret_cast<float*>(smem + kSmemP);   // [4 column quarters][128 rows] row sums of P
      rs_smem[g * kTile + row] = rsum;
      asm volatile("bar.sync 1, %0;" ::"n"(kNumEpiThreads) : "memory");
      const float rtot = rs_smem[row] + rs_smem[kTile + row] + rs_smem[2 * kTile + row] + rs_smem[3 * kTile + row];
      if (g < 2)                                        // two column halves of the slice, as in the other kernels
        write_grad_rows(p, tmem_o, lane_addr, g, row, u.ti, u.row_is_x, u.slice, slice_chunks, rtot, false, true);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kWarpAlloc) tmem_dealloc(tmem, 512);
}

int launch_tc(const TcParams& p, const CUtensorMap& map, const CUtensorMap& mapB, const CUtensorMap& mapV, int ladder,
              cudaStream_t st, int split) {
  const int ngrad = (p.gx_end - p.gx_begin) + (p.gy_end - p.gy_begin);
  const int grid = ngrad * p.nslices + p.n_xx_units;
  if (grid <= 0) return VGAN_B200_OK;
  if (split) {
    auto kern = ladder ? mmd2_tc_kernel<true, true> : mmd2_tc_kernel<false, true>;
    VG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TcCfg<true>::SmemTotal));
    kern<<<grid, kThreads, TcCfg<true>::SmemTotal, st>>>(map, mapB, mapV, p);
  } else {
    auto kern = ladder ? mmd2_tc_kernel<true, false> : mmd2_tc_kernel<false, false>;
    VG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TcCfg<false>::SmemTotal));
    kern<<<grid, kThreads, TcCfg<false>::SmemTotal, st>>>(map, mapB, mapV, p);
  }
  count_launch();
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

}  // namespace vg
