// K1, cluster variant for D_p > 256 (2..4 feature slices): the slice CTAs of one row tile form a
// thread-block cluster and SHARE the pairwise tile instead of recomputing it.
//   CTA c of the cluster (= feature slice c) computes S = Z_I Z_J^T and P = c_j w for the column
//   tiles t = c, c+CS, c+2CS, ... ; every P tile (bf16, 32 KiB) is written to a slot in the
//   producer's shared memory and bulk-copied over DSMEM into the same slot of every peer; each
//   CTA then accumulates O[:, slice c] += P_t Z_Jt[:, slice c] for ALL tiles t in order.
// Executed flops drop from (4 Gram + 1 PV) to (1 Gram + 1 PV) per slice at D = 1024, and the
// L2->SM operand traffic by 3x.  Same math, layouts and descriptors as mmd_tc.cu.
//
// Protocol (t = global column-tile index, slot = t % 3, writer = t % CS):
//   p_full[slot]    (count 1, per CTA)  : tile t has landed in this CTA's slot
//                                         local: plain arrive; peers: remote arrive.expect_tx + bulk copy
//   slot_free[slot] (count CS, in the WRITER of tile t+3): every CTA finished PV(t)
//                                         (tcgen05.commit multicast to the next writer of that slot)
// The MMA stream of every CTA is  G(0) | for m: [1/CS of the next Gram] [PV(m - LAG)]  so that all
// dependencies point backwards in the common tile order (no deadlock) and Gram and PV stages
// interleave in the one TMA ring.
#include "common.cuh"
#include "ptx.cuh"
#include "tc_epi.cuh"

namespace vg {

constexpr int kCStages = 4;                       // 4 x 32 KiB operand ring
constexpr int kCStageBytes = 2 * kChunkBytes;
constexpr int kPSlots = 3;                        // 3 x 32 KiB P tiles
constexpr int kCPBytes = 2 * kChunkBytes;
constexpr int kLag = 1;                           // PV(t) is issued kLag periods after tile t's turn
constexpr int kCThreads = 384;

struct __align__(8) ClBarriers {
  uint64_t ring_full[kCStages], ring_empty[kCStages];
  uint64_t s_full[2], s_empty[2];
  uint64_t nj_full[2];
  uint64_t p_full[kPSlots], slot_free[kPSlots];
  uint64_t o_full;
  uint32_t tmem_base;
  float red[8][4];
};

constexpr int kCSmemRing = 0;
constexpr int kCSmemP = kCSmemRing + kCStages * kCStageBytes;
constexpr int kCSmemNj = kCSmemP + kPSlots * kCPBytes;
constexpr int kCSmemBar = kCSmemNj + 2 * kTile * 4;
constexpr int kCSmemTotal = kCSmemBar + static_cast<int>(sizeof(ClBarriers));
static_assert(kCSmemTotal <= 232448, "shared memory budget");

template <bool LADDER>
__global__ void __launch_bounds__(kCThreads, 1)
mmd2_tc_cluster_kernel(const __grid_constant__ CUtensorMap zmap, const TcParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  ClBarriers* bar = reinterpret_cast<ClBarriers*>(smem + kCSmemBar);
  float* nj_smem = reinterpret_cast<float*>(smem + kCSmemNj);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int CS = p.cs;                                 // cluster size (<= 4 feature slices per cluster)
  const int c = static_cast<int>(cluster_ctarank());
  const int grp = (blockIdx.x / CS) % p.ngroups;       // D_p > 1024: several clusters per row tile
  const int slice = grp * CS + c;                      // this CTA's feature slice (may be past the end)
  // ---- unit decode: one cluster per gradient row tile
  const int ngx = p.gx_end - p.gx_begin;
  const int unit = blockIdx.x / (CS * p.ngroups);
  const int row_is_x = unit < ngx;
  const int ti = row_is_x ? p.gx_begin + unit : p.tiles_x + p.gy_begin + (unit - ngx);
  const int T = p.tiles_x + p.tiles_y;                 // column tiles
  const int K = (T + CS - 1) / CS;                     // Gram tiles per CTA (the tail is padded)
  const int TT = K * CS;                               // padded tile count; tiles >= T carry P = 0
  const int slice_chunks = max(0, min(p.slice_chunks, p.kchunks - slice * p.slice_chunks));
  const int pv_stages = (slice_chunks + 1) >> 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kCStages; ++s) {
      mbar_init(&bar->ring_full[s], 1);
      mbar_init(&bar->ring_empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&bar->s_full[b], 1);
      mbar_init(&bar->s_empty[b], 256);
      mbar_init(&bar->nj_full[b], 1);
    }
    for (int s = 0; s < kPSlots; ++s) {
      mbar_init(&bar->p_full[s], 1);
      mbar_init(&bar->slot_free[s], CS);
    }
    mbar_init(&bar->o_full, 1);
    fence_barrier_init();
  }
  if (warp == 11) {
    tmem_alloc(&bar->tmem_base, 512);
    tmem_relinquish();
  }
  if (warp == 8 && lane == 0) tma_prefetch_desc(&zmap);
  tc_fence_before();
  cluster_sync_all();          // peers' barriers are initialised before any remote arrive / copy
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;
  const uint32_t tmem_o = tmem + 256;

  if (warp == 8) {
    // ===================== TMA producer (mirrors the MMA stream order exactly) =====================
    if (lane == 0) {
      uint32_t s = 0, ph = 0;
      auto gram_stages = [&](int k, int kc0, int kc1) {
        const int jt = min(CS * k + c, T - 1);
        for (int kc = kc0; kc < kc1; ++kc) {
          mbar_wait(&bar->ring_empty[s], ph ^ 1);
          uint8_t* dst = smem + kCSmemRing + s * kCStageBytes;
          mbar_arrive_expect_tx(&bar->ring_full[s], kCStageBytes);
          tma_load_2d(dst, &zmap, &bar->ring_full[s], kc * kChunk, ti * kTile);
          tma_load_2d(dst + kChunkBytes, &zmap, &bar->ring_full[s], kc * kChunk, jt * kTile);
          if (++s == kCStages) { s = 0; ph ^= 1; }
        }
      };
      auto pv_stages_load = [&](int t) {
        const int jt = min(t, T - 1);
        for (int h = 0; h < pv_stages; ++h) {
          const int nch = min(2, slice_chunks - 2 * h);
          mbar_wait(&bar->ring_empty[s], ph ^ 1);
          uint8_t* dst = smem + kCSmemRing + s * kCStageBytes;
          mbar_arrive_expect_tx(&bar->ring_full[s], nch * kChunkBytes);
          for (int cc = 0; cc < nch; ++cc)
            tma_load_2d(dst + cc * kChunkBytes, &zmap, &bar->ring_full[s],
                        (slice * p.slice_chunks + 2 * h + cc) * kChunk, jt * kTile);
          if (++s == kCStages) { s = 0; ph ^= 1; }
        }
      };
      gram_stages(0, 0, p.kchunks);
      for (int m = 0; m < TT + kLag; ++m) {
        const int kn = 1 + m / CS, q = m % CS;
        if (kn < K) gram_stages(kn, (q * p.kchunks) / CS, ((q + 1) * p.kchunks) / CS);
        const int tc = m - kLag;
        if (tc >= 0 && tc < TT) pv_stages_load(tc);
      }
    }
  } else if (warp == 10) {
    // ===================== bulk-copy producer: scaled column norms of this CTA's tiles ============
    if (lane == 0) {
      for (int k = 0; k < K; ++k) {
        const int jt = min(CS * k + c, T - 1);
        const int b = k & 1;
        mbar_wait(&bar->s_empty[b], ((k >> 1) & 1) ^ 1);
        mbar_arrive_expect_tx(&bar->nj_full[b], kTile * 4);
        bulk_load_1d(nj_smem + b * kTile, p.nrms + static_cast<int64_t>(jt) * kTile, kTile * 4,
                     &bar->nj_full[b]);
      }
    }
  } else if (warp == 9) {
    // ===================== MMA issuer (warp-uniform; one elected lane issues) =====================
    constexpr uint32_t idesc_gram = make_idesc_bf16(kTile, kTile, 0, 0);
    constexpr uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
    constexpr uint32_t lbo_v = (static_cast<uint32_t>(kChunkBytes) >> 4) << 16;
    const uint32_t ring_lo = (smem_u32(smem + kCSmemRing) & 0x3FFFFu) >> 4;
    const uint32_t p_lo0 = (smem_u32(smem + kCSmemP) & 0x3FFFFu) >> 4;
    uint32_t s = 0, ph = 0, pf_mask = 0;
    long long t_pf = 0, t_ring = 0, t_se = 0;
    const long long t_begin = clock64();
    auto gram = [&](int k, int kc0, int kc1) {
      const int b = k & 1;
      const uint32_t d_tmem = tmem + b * kTile;
      for (int kc = kc0; kc < kc1; ++kc) {
        if (kc == 0) {
          const long long t0 = clock64();
          mbar_wait(&bar->s_empty[b], ((k >> 1) & 1) ^ 1);
          t_se += clock64() - t0;
          tc_fence_after();
        }
        {
          const long long t0 = clock64();
          mbar_wait(&bar->ring_full[s], ph);
          t_ring += clock64() - t0;
        }
        tc_fence_after();
        const uint32_t a_lo = ring_lo + s * (kCStageBytes >> 4);
        const uint32_t b_lo = a_lo + (kChunkBytes >> 4);
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < kChunk / 16; ++kk) {
            const uint64_t adesc = (static_cast<uint64_t>(desc_hi) << 32) | (a_lo + kk * 2);
            const uint64_t bdesc = (static_cast<uint64_t>(desc_hi) << 32) | (b_lo + kk * 2);
            umma_bf16_ss(d_tmem, adesc, bdesc, idesc_gram, (kc | kk) != 0);
          }
          umma_commit(&bar->ring_empty[s]);
          if (kc == p.kchunks - 1) umma_commit(&bar->s_full[b]);
        }
        __syncwarp();
        if (++s == kCStages) { s = 0; ph ^= 1; }
      }
    };
    auto pv = [&](int t) {
      const int slot = t % kPSlots;
      {
        const long long t0 = clock64();
        mbar_wait(&bar->p_full[slot], (pf_mask >> slot) & 1);
        t_pf += clock64() - t0;
      }
      pf_mask ^= 1u << slot;
      tc_fence_after();
      const uint32_t p_lo = p_lo0 + slot * (kCPBytes >> 4);
      for (int h = 0; h < pv_stages; ++h) {
        const int nch = min(2, slice_chunks - 2 * h);
        const uint32_t idesc_pv = make_idesc_bf16(kTile, nch * kChunk, 0, 1);
        {
          const long long t0 = clock64();
          mbar_wait(&bar->ring_full[s], ph);
          t_ring += clock64() - t0;
        }
        tc_fence_after();
        const uint32_t v_lo = ring_lo + s * (kCStageBytes >> 4);
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < kTile / 16; ++kk) {
            const uint64_t adesc = (static_cast<uint64_t>(desc_hi) << 32) |
                                   (p_lo + (kk >> 2) * (kChunkBytes >> 4) + (kk & 3) * 2);
            const uint64_t bdesc = (static_cast<uint64_t>(desc_hi) << 32) | (v_lo + kk * (2048 >> 4)) | lbo_v;
            umma_bf16_ss(tmem_o + h * 2 * kChunk, adesc, bdesc, idesc_pv, (t | kk) != 0);
          }
          umma_commit(&bar->ring_empty[s]);
        }
        __syncwarp();
        if (++s == kCStages) { s = 0; ph ^= 1; }
      }
      // the slot may be refilled (by the writer of tile t + kPSlots) once every CTA got here
      if (t + kPSlots < TT && elect_one())
        umma_commit_mc(&bar->slot_free[slot], static_cast<uint16_t>(1u << ((t + kPSlots) % CS)));
      __syncwarp();
    };
    gram(0, 0, p.kchunks);
    for (int m = 0; m < TT + kLag; ++m) {
      const int kn = 1 + m / CS, q = m % CS;
      if (kn < K) gram(kn, (q * p.kchunks) / CS, ((q + 1) * p.kchunks) / CS);
      const int tc = m - kLag;
      if (tc >= 0 && tc < TT) pv(tc);
    }
    if (elect_one()) umma_commit(&bar->o_full);
    __syncwarp();
    if (p.dbg && lane == 0) {
      float* d = p.dbg + static_cast<int64_t>(blockIdx.x) * p.dbg_stride;
      d[0] = static_cast<float>(clock64() - t_begin);
      d[1] = static_cast<float>(t_pf);
      d[2] = static_cast<float>(t_ring);
      d[3] = static_cast<float>(t_se);
    }
  } else if (warp < 8) {
    // ===================== epilogue warps: this CTA's own column tiles =====================
    const int q = warp & 3, g = warp >> 2;
    const int row = q * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(q * 32) << 16;
    const float m2c0 = p.par[1];
    const float ni = p.nrms[static_cast<int64_t>(ti) * kTile + row];
    float coef_x[VGAN_B200_MAX_SCALES], coef_y[VGAN_B200_MAX_SCALES], ratio[VGAN_B200_MAX_SCALES];
#pragma unroll
    for (int m = 0; m < VGAN_B200_MAX_SCALES; ++m) {
      ratio[m] = p.par[2 + m];
      coef_x[m] = p.par[10 + m];
      coef_y[m] = p.par[18 + m];
    }
    float acc_x = 0.f, acc_y = 0.f, rsum = 0.f;
    uint32_t sf_mask = 0;
    long long e_sfull = 0, e_comp = 0, e_slot = 0, e_pub = 0;
    for (int k = 0; k < K; ++k) {
      const int t = CS * k + c;
      const bool valid = t < T;
      const int jt = min(t, T - 1);
      const int b = k & 1;
      const bool col_is_x = jt < p.tiles_x;
      float cf[VGAN_B200_MAX_SCALES];
#pragma unroll
      for (int m = 0; m < VGAN_B200_MAX_SCALES; ++m) cf[m] = valid ? (col_is_x ? coef_x[m] : coef_y[m]) : 0.f;
      long long tq = clock64();
      mbar_wait(&bar->nj_full[b], (k >> 1) & 1);
      mbar_wait(&bar->s_full[b], (k >> 1) & 1);
      e_sfull += clock64() - tq; tq = clock64();
      tc_fence_after();
      const float* njp = nj_smem + b * kTile + g * 64;
      float ksum = 0.f;
      uint4 pk[8];
      epi_tile_64<LADDER>(tmem + lane_addr + b * kTile + g * 64, njp, ni, m2c0, cf, ratio, p.n_scales, ksum, rsum, pk);
      tc_fence_before();
      mbar_arrive(&bar->s_empty[b]);          // S and the norms are free; P is parked in registers
      if (valid) { if (col_is_x) acc_x += ksum; else acc_y += ksum; }
      // ---- publish the P tile: own slot first, then a bulk copy into the same slot of every peer
      const int slot = t % kPSlots;
      e_comp += clock64() - tq; tq = clock64();
      if (t >= kPSlots) {
        mbar_wait(&bar->slot_free[slot], (sf_mask >> slot) & 1);   // every CTA finished PV(t - kPSlots)
        sf_mask ^= 1u << slot;
      }
      e_slot += clock64() - tq; tq = clock64();
      uint8_t* p_slot = smem + kCSmemP + slot * kCPBytes;
      uint8_t* p_row = p_slot + g * kChunkBytes + row * 128;
#pragma unroll
      for (int i = 0; i < 8; ++i)
        *reinterpret_cast<uint4*>(p_row + ((i ^ (row & 7)) << 4)) = pk[i];
      fence_proxy_async_smem();
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (threadIdx.x == 0) {
        mbar_arrive(&bar->p_full[slot]);
        const uint32_t src = smem_u32(p_slot), fb = smem_u32(&bar->p_full[slot]);
        for (int d = 0; d < CS; ++d) {
          if (d == c) continue;
          const uint32_t rbar = mapa_u32(fb, d);
          mbar_remote_arrive_expect_tx(rbar, kCPBytes);
          bulk_copy_to_peer(mapa_u32(src, d), src, kCPBytes, rbar);
        }
      }
      e_pub += clock64() - tq;
    }
    if (p.dbg && threadIdx.x == 0) {
      float* d = p.dbg + static_cast<int64_t>(blockIdx.x) * p.dbg_stride;
      d[4] = static_cast<float>(e_sfull);
      d[5] = static_cast<float>(e_comp);
      d[6] = static_cast<float>(e_slot);
      d[7] = static_cast<float>(e_pub);
    }
    // ---- partial sums of this CTA's tiles
    acc_x = warp_sum(acc_x);
    acc_y = warp_sum(acc_y);
    if (lane == 0) { bar->red[warp][0] = acc_x; bar->red[warp][1] = acc_y; }
    asm volatile("bar.sync 1, 256;" ::: "memory");   // all epilogue warps are done with the norms
    nj_smem[g * kTile + row] = rsum;
    if (threadIdx.x == 0 && grp == 0) {
      double sx = 0.0, sy = 0.0;
      for (int wv = 0; wv < 8; ++wv) { sx += bar->red[wv][0]; sy += bar->red[wv][1]; }
      if (row_is_x) {
        atomicAdd(&p.sums[0], sx);
      } else {
        atomicAdd(&p.sums[1], sx);
        atomicAdd(&p.sums[2], sy);
      }
    }
    mbar_wait(&bar->o_full, 0);      // every PV of this CTA has completed: its P slots are dead
    tc_fence_after();
  }

  // ===================== cluster-wide row sums of P, then the gradient rows =====================
  float* rsx = reinterpret_cast<float*>(smem + kCSmemP);       // [CS][128], reuses P slot 0
  tc_fence_before();
  cluster_sync_all();      // #1: no CTA of the cluster reads a P slot any more
  if (warp < 4) {
    const int row = warp * 32 + lane;
    const float tot = nj_smem[row] + nj_smem[kTile + row];     // both column halves of this CTA's tiles
    const uint32_t a = smem_u32(rsx + c * kTile + row);
    for (int d = 0; d < CS; ++d) st_cluster_f32(mapa_u32(a, d), tot);
  }
  cluster_sync_all();      // #2: all partial row sums have arrived
  if (warp < 8) {
    const int q = warp & 3, g = warp >> 2;
    const int row = q * 32 + lane;
    float rtot = 0.f;
    for (int d = 0; d < CS; ++d) rtot += rsx[d * kTile + row];
    tc_fence_after();
    write_grad_rows(p, tmem_o, static_cast<uint32_t>(q * 32) << 16, g, row, ti, row_is_x, slice, slice_chunks, rtot);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 11) tmem_dealloc(tmem, 512);
}

int launch_tc_cluster(const TcParams& p, const CUtensorMap& map, int ladder, cudaStream_t st) {
  const int ngrad = (p.gx_end - p.gx_begin) + (p.gy_end - p.gy_begin);
  if (ngrad <= 0) return VGAN_B200_OK;
  auto kern = ladder ? mmd2_tc_cluster_kernel<true> : mmd2_tc_cluster_kernel<false>;
  VG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kCSmemTotal));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(static_cast<unsigned>(ngrad * p.cs * p.ngroups));
  cfg.blockDim = dim3(kCThreads);
  cfg.dynamicSmemBytes = kCSmemTotal;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = static_cast<unsigned>(p.cs);
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  VG_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, map, p));
  return VGAN_B200_OK;
}

}  // namespace vg
