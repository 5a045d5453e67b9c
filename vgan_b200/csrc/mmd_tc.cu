// K1 — fused MMD^2 forward+backward on tcgen05 tensor cores (bf16 operands, fp32 TMEM
// accumulation).  Replaces RBF.forward + MMDLossConstrained.forward + autograd backward
// (/root/reference/src/models/Mmd_loss_constrained.py:24-26, 42-50).
//
// One CTA = one work unit = (row tile I of 128 stacked rows, a contiguous range of column tiles J):
//   for every column tile J of the range (128 stacked rows):
//     S   = Z_I Z_J^T                    tcgen05.mma; Z_I stays in shared memory for the whole unit, Z_J streams by TMA
//     d2  = |z_i|^2 + |z_j|^2 - 2 S      epilogue warps, straight out of TMEM
//     t   = exp(-d2/s_4); k = t+t^2+t^4+t^8+t^16 ; w = sum_m t^(2^q)/s_m   (x2 ladder)
//     sums += k (S_xx / S_xy / S_yy by tile type) ; P = c_j w  -> bf16 -> smem
//     O_I += P Z_J                       second tcgen05.mma, O stays in TMEM
//   dZ_I = 4 c_i (O - rowsum(P) z_i)
// The pairwise matrix only ever exists as a 128x128 TMEM tile.
//
// Units: gradient units (a Y row tile, or an X row tile with NEED_DX, against ALL columns — cut into `col_split`
// column ranges when there are fewer row tiles than SMs, the ranges adding up through red.global.add) and S_xx units
// (an X row tile against a chunk of its upper-triangle columns, kernel sums only: no weights, no P, no second GEMM).
// The S_xx chunks are short (<= xx_chunk tiles) so that they pack the tail of the grid behind the long gradient units.
//
// Warp roles (640 threads): warps 0-15 epilogue in two groups of 8 that alternate over the column tiles (warp%4 = TMEM
// lane quadrant, (warp/4)%2 = column half of 64, warp/8 = group = parity of the tile), warp 16 = TMA producer (column operand and V slices), warp 17 = MMA issuer, warp 18 = column norms,
// warp 19 = TMEM allocator.  This kernel serves D_p <= 256, where the epilogue (FP32 + MUFU issue) is the bound:
// with 8 epilogue warps (2 per scheduler) the issue slots were 49 % busy (ncu, profiles/r1d_base_d64_summary.txt:
// fixed-latency dependency stalls with nothing else to issue); 4 warps per scheduler hide them.  setmaxnreg moves
// registers from the 4 service warps (96 -> 56) to the epilogue warps (96 -> 104).
//
// Shared memory: the row operand Z_I is loaded ONCE per unit (round 2: it used to travel through the ring with every
// column tile — a third of the L2 -> SM bytes, and at D_p = 256 those bytes were the bound: 192 KiB per tile at the
// measured 54 B/clk/SM = 3 550 cycles against 2 260 of epilogue math).  The ring then carries the column operand
// (two 64-feature chunks per 32 KiB stage) and the V chunks of the gradient GEMM.
#include "common.cuh"
#include "ptx.cuh"
#include "tc_epi.cuh"

namespace vg {

constexpr int kStagesMax = 5;
constexpr int kStageBytes = 2 * kChunkBytes;   // two 64-feature chunks of a column tile (Gram B operand or V)
constexpr int kPBytes = 2 * kChunkBytes;                     // 128 x 128 bf16
constexpr int kNumEpiWarps = 16;
constexpr int kNumEpiThreads = kNumEpiWarps * 32;
constexpr int kThreads = kNumEpiThreads + 128;
constexpr int kWarpTma = kNumEpiWarps, kWarpMma = kNumEpiWarps + 1, kWarpNorms = kNumEpiWarps + 2,
              kWarpAlloc = kNumEpiWarps + 3;
constexpr int kSmemLimit = 232448;
constexpr int kEpiGroupThreads = kNumEpiThreads / 2;   // two epilogue groups of 8 warps alternate over the column tiles
constexpr int kNjBufs = 4;   // column-norm buffers: the norms of tile n + 2 (same epilogue group) load while tile n is computed
constexpr int kLag = 3;   // the Gram stream runs this many column tiles ahead of the gradient GEMM (see the MMA warp)

struct __align__(8) TcBarriers {
  uint64_t ring_full[kStagesMax], ring_empty[kStagesMax];
  uint64_t s_full[2], s_empty[2];
  uint64_t nj_full[kNjBufs], nj_empty[kNjBufs];
  uint64_t p_full[2], p_empty[2];
  uint64_t o_full, zi_full, zi_ready;
  uint32_t tmem_base;
  float red[kNumEpiWarps][2];
};
static_assert(sizeof(TcBarriers) <= 1024, "barrier block");

// SPLIT (fp32-class results from split bf16 operands z = hi + lo, BF16X3) in ONE pass: the K = 3 D_p Gram
// [hi|hi|lo] x [hi|lo|hi] is computed once per column tile, the epilogue parks BOTH bf16(P) and the remainder
// P - bf16(P), and the gradient GEMM issues its three parts O += P_hi V_hi + P_hi V_lo + P_lo V_hi back to back
// (three passes of the plain kernel recomputed the Gram and the whole epilogue three times).  Four P buffers
// instead of two.
//
// Shared-memory geometry (bytes): [barriers 1 KiB][column norms 2 KiB][P buffers][Z_I resident][ring stages].
// The row operand is resident whenever it fits beside the P buffers and at least two ring stages: always on the plain
// bf16 path (<= 4 chunks), and with split operands at D_p = 64 (hi and lo chunk: the VGAN encodings).  Otherwise
// (split, D_p >= 128) the row operand travels through the ring with the column operand as before.
struct TcGeom {
  int res;         // Z_I resident
  int zi_chunks;   // 16 KiB chunks of the resident row operand
  int stages;      // ring stages of 32 KiB
  int off_p, off_zi, off_ring, total;
};
__host__ __device__ inline TcGeom tc_geom(bool split, int kchunks) {
  TcGeom g;
  const int pbufs = split ? 4 : 2;
  g.res = split ? (kchunks == 1) : (kchunks <= kSliceMax / kChunk);   // (wider: forward-only calls, no gradient GEMM)
  g.zi_chunks = g.res ? (split ? 2 * kchunks : kchunks) : 0;
  g.off_p = 1024 + kNjBufs * kTile * 4;
  g.off_zi = g.off_p + pbufs * kPBytes;
  g.off_ring = g.off_zi + g.zi_chunks * kChunkBytes;
  int st = (kSmemLimit - g.off_ring) / kStageBytes;
  g.stages = st > kStagesMax ? kStagesMax : st;
  g.total = g.off_ring + g.stages * kStageBytes;
  return g;
}

struct Unit {
  int ti;        // stacked row tile
  int slice;     // feature slice (always 0 here: D_p <= 256)
  int j_begin, j_end;
  int pv;        // gradient GEMM on
  int row_is_x;  // row tile lies in the X region
  int sym;       // S_xx-only unit walking the upper triangle: off-diagonal tiles count twice
  int first;     // the range starts at the unit's first column (adds the common-mode remainder)
};

// S_xx chunks of X row tile t (columns t .. tiles_x - 1 in pieces of at most xx_chunk tiles)
__host__ __device__ inline int xx_chunks_of(int tiles_x, int t, int xx_chunk) {
  return (tiles_x - t + xx_chunk - 1) / xx_chunk;
}

__device__ __forceinline__ Unit decode_unit(const TcParams& p) {
  Unit u;
  const int ngx = p.gx_end - p.gx_begin;
  const int ngy = p.gy_end - p.gy_begin;
  const int T = p.tiles_x + p.tiles_y;
  const int n_grad_units = (ngx + ngy) * p.col_split;
  int b = blockIdx.x;
  u.slice = 0;
  if (b < n_grad_units) {
    const int t = b / p.col_split, r = b - t * p.col_split;
    u.row_is_x = t < ngx;
    u.ti = u.row_is_x ? p.gx_begin + t : p.tiles_x + p.gy_begin + (t - ngx);
    u.j_begin = static_cast<int>(static_cast<long long>(T) * r / p.col_split);
    u.j_end = static_cast<int>(static_cast<long long>(T) * (r + 1) / p.col_split);
    u.pv = p.pv;
    u.sym = 0;
    u.first = r == 0;
  } else {
    int x = b - n_grad_units;
    int t = p.xx_part;
    int nch = 1;
    for (; t < p.tiles_x; t += p.xx_nparts) {
      nch = xx_chunks_of(p.tiles_x, t, p.xx_chunk);
      if (x < nch) break;
      x -= nch;
    }
    if (t >= p.tiles_x) { t = p.tiles_x - 1; x = 0; nch = 1; }    // (not reached: the grid is sized by the same walk)
    const int len = p.tiles_x - t;
    u.ti = t;
    u.row_is_x = 1;
    u.j_begin = t + static_cast<int>(static_cast<long long>(len) * x / nch);   // upper triangle of the symmetric X x X block
    u.j_end = t + static_cast<int>(static_cast<long long>(len) * (x + 1) / nch);
    u.pv = 0;
    u.sym = 1;
    u.first = 0;
  }
  return u;
}

// Epilogue of one unit, two groups of 8 warps: group n & 1 owns column tile n (S[n & 1], P[n & 1], the norms buffer
// n & 1); a warp covers 32 rows x 64 columns in two steps of 32.  The groups only meet in the FP32 pipe: while one sits
// in its barrier waits, tcgen05.ld latency and the fence behind its P stores, the other one computes (measured against
// all 16 warps sharing every tile, same box: N = 16384, D = 128 506 -> 454 us, N = 65536, D = 64 7.65 -> 6.72 ms).
// PV = false: kernel sums only (S_xx chunks, forward-only calls).
template <bool LADDER, bool SPLIT, bool PV, bool ONES>
__device__ __forceinline__ void epilogue_tiles(const TcParams& p, const Unit& u, TcBarriers* bar, uint8_t* p_smem,
                                               const float* nj_smem, uint32_t tmem, int warp, int lane, float& acc_x,
                                               float& acc_y, float& rsum) {
  const int q = warp & 3, hcol = (warp >> 2) & 1, grp = warp >> 3;
  const int row = q * 32 + lane;
  const uint32_t lane_addr = static_cast<uint32_t>(q * 32) << 16;
  const float m2c0 = p.par[1];
  const float ni = ONES ? 0.f : p.nrms[static_cast<int64_t>(u.ti) * kTile + row];   // (ONES: inside the Gram)
  float coef_x[VGAN_B200_MAX_SCALES], coef_y[VGAN_B200_MAX_SCALES], ratio[VGAN_B200_MAX_SCALES];
#pragma unroll
  for (int m = 0; m < VGAN_B200_MAX_SCALES; ++m) {
    ratio[m] = p.par[2 + m];
    coef_x[m] = PV ? p.par[kParCoefX + m] : 0.f;
    coef_y[m] = PV ? p.par[kParCoefY + m] : 0.f;
  }
  const float cw0_x = PV ? p.par[kParCw0 + (u.row_is_x ? 0 : 2)] : 0.f;
  const float cw0_y = PV ? p.par[kParCw0 + (u.row_is_x ? 0 : 2) + 1] : 0.f;
  const int n_tiles = u.j_end - u.j_begin;
  const int b = grp;
  const uint32_t taddr = tmem + lane_addr + b * kTile + hcol * 64;
  uint8_t* p_row = p_smem + b * kPBytes + hcol * kChunkBytes + row * 128;
  for (int n = grp; n < n_tiles; n += 2) {
    const int jt = u.j_begin + n;
    const uint32_t par = (n >> 1) & 1;
    const bool col_is_x = jt < p.tiles_x;
    const int nb = n & (kNjBufs - 1);
    const float* njp = nj_smem + nb * kTile + hcol * 64;
    mbar_wait(&bar->nj_full[nb], (n / kNjBufs) & 1);
    mbar_wait(&bar->s_full[b], par);
    tc_fence_after();
    float ksum = 0.f;
    if constexpr (PV) {
      float cf[VGAN_B200_MAX_SCALES];
#pragma unroll
      for (int m = 0; m < VGAN_B200_MAX_SCALES; ++m) cf[m] = col_is_x ? coef_x[m] : coef_y[m];
      const float cw0 = col_is_x ? cw0_x : cw0_y;
      uint4 pk[4], pk_lo[4];
      {
        uint32_t v[32];
        tmem_ld_32x32(taddr, v);
        tmem_ld_wait();
        epi_core_32<LADDER, SPLIT ? kPBoth : kPHi, ONES>(v, njp, ni, m2c0, cf, ratio, p.n_scales, ksum, rsum, pk, cw0, pk_lo);
      }
      mbar_wait(&bar->p_empty[b], par ^ 1);              // PV(n-2) done with P[b]
#pragma unroll
      for (int i = 0; i < 4; ++i) {                        // 128B-swizzled K-major A operand
        *reinterpret_cast<uint4*>(p_row + ((i ^ (row & 7)) << 4)) = pk[i];
        if constexpr (SPLIT) *reinterpret_cast<uint4*>(p_row + 2 * kPBytes + ((i ^ (row & 7)) << 4)) = pk_lo[i];
      }
      {
        uint32_t v[32];
        tmem_ld_32x32(taddr + 32, v);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(&bar->s_empty[b]);     // all of S[b] is in registers: the Gram two tiles on may overwrite it
        epi_core_32<LADDER, SPLIT ? kPBoth : kPHi, ONES>(v, njp + 32, ni, m2c0, cf, ratio, p.n_scales, ksum, rsum, pk, cw0,
                                                   pk_lo);
      }
      mbar_arrive(&bar->nj_empty[nb]);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        *reinterpret_cast<uint4*>(p_row + (((4 + i) ^ (row & 7)) << 4)) = pk[i];
        if constexpr (SPLIT) *reinterpret_cast<uint4*>(p_row + 2 * kPBytes + (((4 + i) ^ (row & 7)) << 4)) = pk_lo[i];
      }
      fence_proxy_async_smem();
      mbar_arrive(&bar->p_full[b]);
      if (col_is_x) acc_x += ksum; else acc_y += ksum;
    } else {
      uint32_t v[32];
      tmem_ld_32x32(taddr, v);
      tmem_ld_wait();
      epi_k_32<LADDER, ONES>(v, njp, ni, m2c0, ratio, p.n_scales, ksum);
      tmem_ld_32x32(taddr + 32, v);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&bar->s_empty[b]);
      epi_k_32<LADDER, ONES>(v, njp + 32, ni, m2c0, ratio, p.n_scales, ksum);
      mbar_arrive(&bar->nj_empty[nb]);
      const float fac = (u.sym && jt != u.ti) ? 2.f : 1.f;
      if (col_is_x) acc_x = fmaf(ksum, fac, acc_x); else acc_y += ksum;
    }
  }
}
template <bool LADDER, bool SPLIT>
__global__ void __launch_bounds__(kThreads, 1)
mmd2_tc_kernel(const __grid_constant__ CUtensorMap zmap, const __grid_constant__ CUtensorMap zmapB,
               const __grid_constant__ CUtensorMap zmapV, const TcParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];   // SWIZZLE_128B tiles need 1024-byte alignment
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  constexpr int kPvParts = SPLIT ? 3 : 1;
  const TcGeom geo = tc_geom(SPLIT, p.kchunks);
  const int kStages = geo.stages;
  TcBarriers* bar = reinterpret_cast<TcBarriers*>(smem);
  float* nj_smem = reinterpret_cast<float*>(smem + 1024);
  uint8_t* p_smem = smem + geo.off_p;
  uint8_t* zi_smem = smem + geo.off_zi;
  uint8_t* ring = smem + geo.off_ring;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const Unit u = decode_unit(p);
  const int n_tiles = u.j_end - u.j_begin;
  const int slice_chunks = p.kchunks;                      // one slice: D_p <= 256
  const int g_stages = (p.kchunks + 1) >> 1;               // resident form: column-operand stages per tile

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStagesMax; ++s) {
      mbar_init(&bar->ring_full[s], 1);
      mbar_init(&bar->ring_empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&bar->s_full[b], 1);
      mbar_init(&bar->s_empty[b], kEpiGroupThreads);
      mbar_init(&bar->p_full[b], kEpiGroupThreads);
      mbar_init(&bar->p_empty[b], 1);
    }
    for (int b = 0; b < kNjBufs; ++b) {
      mbar_init(&bar->nj_full[b], 1);
      mbar_init(&bar->nj_empty[b], kEpiGroupThreads);
    }
    mbar_init(&bar->o_full, 1);
    mbar_init(&bar->zi_full, 1);
    mbar_init(&bar->zi_ready, kTile);
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
    // ===================== TMA producer: the resident row operand, then one ring for column operand and V ==========
    // Stages are issued in exactly the order the MMA thread consumes them:
    //   Gram(0), Gram(1), Gram(2), PV(0), Gram(3), PV(1), ...
    if (lane == 0) {
      uint32_t s = 0, ph = 0;
      const int pv_stages = (slice_chunks + 1) >> 1;
      if (geo.res) {
        mbar_arrive_expect_tx(&bar->zi_full, geo.zi_chunks * kChunkBytes);
        if constexpr (SPLIT) {      // hi chunk 0 and lo chunk 2 kchunks of the row-operand array [hi | hi | lo]
          tma_load_2d(zi_smem, &zmap, &bar->zi_full, 0, u.ti * kTile);
          tma_load_2d(zi_smem + kChunkBytes, &zmap, &bar->zi_full, 2 * p.kchunks * kChunk, u.ti * kTile);
        } else {
          for (int kc = 0; kc < p.kchunks; ++kc)
            tma_load_2d(zi_smem + kc * kChunkBytes, &zmap, &bar->zi_full, kc * kChunk, u.ti * kTile);
        }
      }
      for (int n = 0; n < n_tiles + kLag; ++n) {
        if (u.pv && n >= kLag) {
          // ---- V operand of PV(n - kLag)
          const int jt = u.j_begin + n - kLag;
          if (SPLIT && geo.res) {
            // one stage [V_hi | V_lo] (64 features each) serves the three gradient parts
            mbar_wait(&bar->ring_empty[s], ph ^ 1);
            uint8_t* dst = ring + s * kStageBytes;
            mbar_arrive_expect_tx(&bar->ring_full[s], kStageBytes);
            tma_load_2d(dst, &zmapV, &bar->ring_full[s], 0, jt * kTile);
            tma_load_2d(dst + kChunkBytes, &zmapV, &bar->ring_full[s], 2 * p.kchunks * kChunk, jt * kTile);
            if (++s == kStages) { s = 0; ph ^= 1; }
          } else {
            for (int part = 0; part < kPvParts; ++part) {
              // V operand of this part: the hi block (chunk 0) or the lo block (chunk 2 kchunks) of the row-operand array
              const int vbase = SPLIT ? (part == 1 ? 2 * p.kchunks : 0) : p.v_chunk0;
              for (int h = 0; h < pv_stages; ++h) {
                const int nch = min(2, slice_chunks - 2 * h);
                mbar_wait(&bar->ring_empty[s], ph ^ 1);
                uint8_t* dst = ring + s * kStageBytes;
                mbar_arrive_expect_tx(&bar->ring_full[s], nch * kChunkBytes);
                for (int c = 0; c < nch; ++c)
                  tma_load_2d(dst + c * kChunkBytes, &zmapV, &bar->ring_full[s], (vbase + 2 * h + c) * kChunk,
                              jt * kTile);
                if (++s == kStages) { s = 0; ph ^= 1; }
              }
            }
          }
        }
        if (n < n_tiles) {
          // ---- operands of Gram(n)
          const int jt = u.j_begin + n;
          if (geo.res) {
            // column operand, two chunks per stage (split operands: [hi | lo] of the column array = its chunks 0, 1)
            const int bchunks = SPLIT ? 2 : p.kchunks;
            for (int h = 0; h < (SPLIT ? 1 : g_stages); ++h) {
              const int nch = min(2, bchunks - 2 * h);
              mbar_wait(&bar->ring_empty[s], ph ^ 1);
              uint8_t* dst = ring + s * kStageBytes;
              mbar_arrive_expect_tx(&bar->ring_full[s], nch * kChunkBytes);
              for (int c = 0; c < nch; ++c)
                tma_load_2d(dst + c * kChunkBytes, &zmapB, &bar->ring_full[s], (2 * h + c) * kChunk, jt * kTile);
              if (++s == kStages) { s = 0; ph ^= 1; }
            }
          } else {
            for (int kc = 0; kc < p.gchunks; ++kc) {
              mbar_wait(&bar->ring_empty[s], ph ^ 1);
              uint8_t* dst = ring + s * kStageBytes;
              mbar_arrive_expect_tx(&bar->ring_full[s], kStageBytes);
              tma_load_2d(dst, &zmap, &bar->ring_full[s], kc * kChunk, u.ti * kTile);
              tma_load_2d(dst + kChunkBytes, &zmapB, &bar->ring_full[s], kc * kChunk, jt * kTile);
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
        const int b = n & (kNjBufs - 1);
        mbar_wait(&bar->nj_empty[b], ((n / kNjBufs) & 1) ^ 1);   // epilogue done with nj[b]
        mbar_arrive_expect_tx(&bar->nj_full[b], kTile * 4);
        bulk_load_1d(nj_smem + b * kTile, p.nrms + static_cast<int64_t>(jt) * kTile, kTile * 4,
                     &bar->nj_full[b]);
      }
    }
  } else if (warp == kWarpMma) {
    // ===================== MMA issuer =====================
    // The whole warp runs the (warp-uniform) control flow so that addresses and descriptors live
    // in uniform registers; one elected lane issues the tcgen05 instructions.
    // Order: PV(n - 3), Gram(n).  The epilogue releases S[b] as soon as its columns are in registers, so Gram(n) only
    // needs epilogue(n - 2) to have STARTED; issued behind PV(n - 3) it completes two epilogue tiles before it is
    // needed (with PV(n - 1) in front of it — the in-order issuer then blocks until epilogue(n - 1) has ENDED — S(n + 1)
    // had one tile of slack, and at D_p = 256, where a tile carries 2 048 tensor cycles and four ring stages behind
    // the L2 latency, the epilogue waited for it).
    constexpr uint32_t idesc_gram = make_idesc_bf16(kTile, kTile, 0, 0);
    // descriptor high word: SBO = 1024 B, version 1, SWIZZLE_128B; low word: address and LBO
    constexpr uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
    constexpr uint32_t lbo_v = (static_cast<uint32_t>(kChunkBytes) >> 4) << 16;   // MN-major V: next 64 features
    const uint32_t ring_lo = (smem_u32(ring) & 0x3FFFFu) >> 4;
    const uint32_t zi_lo = (smem_u32(zi_smem) & 0x3FFFFu) >> 4;
    const uint32_t p_lo0 = (smem_u32(p_smem) & 0x3FFFFu) >> 4;
    const int pv_stages = (slice_chunks + 1) >> 1;
    uint32_t s = 0, ph = 0;
    if (geo.res) {
      mbar_wait(p.ones ? &bar->zi_ready : &bar->zi_full, 0);     // (ones: the row norms have been written into Z_I)
      tc_fence_after();
    }
    // 4 k-steps of 16 over one 64-feature chunk pair
    auto gram_chunk = [&](uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, bool first) {
#pragma unroll
      for (int kk = 0; kk < kChunk / 16; ++kk) {
        const uint64_t adesc = (static_cast<uint64_t>(desc_hi) << 32) | (a_lo + kk * 2);
        const uint64_t bdesc = (static_cast<uint64_t>(desc_hi) << 32) | (b_lo + kk * 2);
        umma_bf16_ss(d_tmem, adesc, bdesc, idesc_gram, !(first && kk == 0));
      }
    };
    for (int n = 0; n < n_tiles + kLag; ++n) {
      if (u.pv && n >= kLag) {
        // ---- PV(n - kLag): O += P Z_J   (two 64-feature chunks = 128 columns of O per stage)
        const int m = n - kLag, pb = m & 1;
        mbar_wait(&bar->p_full[pb], (m >> 1) & 1);
        tc_fence_after();
        if (SPLIT && geo.res) {
          // stage = [V_hi | V_lo], 64 features each: O(64 columns) += P_hi V_hi + P_lo V_hi + P_hi V_lo
          constexpr uint32_t idesc_pv = make_idesc_bf16(kTile, kChunk, 0, 1);
          mbar_wait(&bar->ring_full[s], ph);
          tc_fence_after();
          const uint32_t v_lo = ring_lo + s * (kStageBytes >> 4);
          if (elect_one()) {
#pragma unroll
            for (int part = 0; part < 3; ++part) {
              const uint32_t p_lo = p_lo0 + (pb + (part == 1 ? 2 : 0)) * (kPBytes >> 4);
              const uint32_t vv = v_lo + (part == 2 ? (kChunkBytes >> 4) : 0);
#pragma unroll
              for (int kk = 0; kk < kTile / 16; ++kk) {
                const uint64_t adesc = (static_cast<uint64_t>(desc_hi) << 32) |
                                       (p_lo + (kk >> 2) * (kChunkBytes >> 4) + (kk & 3) * 2);
                const uint64_t bdesc = (static_cast<uint64_t>(desc_hi) << 32) | (vv + kk * (2048 >> 4)) | lbo_v;
                umma_bf16_ss(tmem_o, adesc, bdesc, idesc_pv, (m | part | kk) != 0);
              }
            }
            umma_commit(&bar->ring_empty[s]);
          }
          __syncwarp();
          if (++s == kStages) { s = 0; ph ^= 1; }
        } else {
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
        }
        if (elect_one()) umma_commit(&bar->p_empty[pb]);
        __syncwarp();
      }
      if (n < n_tiles) {
        // ---- Gram(n): S[b] = Z_I Z_J^T over all K chunks
        const int b = n & 1;
        mbar_wait(&bar->s_empty[b], ((n >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem + b * kTile;
        if (geo.res) {
          if constexpr (SPLIT) {
            // stage = [B_hi | B_lo]; resident = [A_hi | A_lo]:  hi hi + lo hi + hi lo
            mbar_wait(&bar->ring_full[s], ph);
            tc_fence_after();
            const uint32_t b_lo = ring_lo + s * (kStageBytes >> 4);
            if (elect_one()) {
              gram_chunk(d_tmem, zi_lo, b_lo, true);
              gram_chunk(d_tmem, zi_lo + (kChunkBytes >> 4), b_lo, false);
              gram_chunk(d_tmem, zi_lo, b_lo + (kChunkBytes >> 4), false);
              umma_commit(&bar->ring_empty[s]);
            }
            __syncwarp();
            if (++s == kStages) { s = 0; ph ^= 1; }
          } else {
            for (int h = 0; h < g_stages; ++h) {
              const int nch = min(2, p.kchunks - 2 * h);
              mbar_wait(&bar->ring_full[s], ph);
              tc_fence_after();
              const uint32_t b_lo = ring_lo + s * (kStageBytes >> 4);
              const uint32_t a_lo = zi_lo + 2 * h * (kChunkBytes >> 4);
              if (elect_one()) {
                gram_chunk(d_tmem, a_lo, b_lo, h == 0);
                if (nch == 2) gram_chunk(d_tmem, a_lo + (kChunkBytes >> 4), b_lo + (kChunkBytes >> 4), false);
                umma_commit(&bar->ring_empty[s]);
              }
              __syncwarp();
              if (++s == kStages) { s = 0; ph ^= 1; }
            }
          }
        } else {
          for (int kc = 0; kc < p.gchunks; ++kc) {
            mbar_wait(&bar->ring_full[s], ph);
            tc_fence_after();
            const uint32_t a_lo = ring_lo + s * (kStageBytes >> 4);
            if (elect_one()) {
              gram_chunk(d_tmem, a_lo, a_lo + (kChunkBytes >> 4), kc == 0);
              umma_commit(&bar->ring_empty[s]);
            }
            __syncwarp();
            if (++s == kStages) { s = 0; ph ^= 1; }
          }
        }
        if (elect_one()) umma_commit(&bar->s_full[b]);
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
    float acc_x = 0.f, acc_y = 0.f;   // sum of k over X / Y columns (weighted by symmetry)
    float rsum = 0.f;                 // sum_j P_ij over this thread's column quarter
    const bool ones = p.ones != 0 && geo.res;
    if (ones && warp < 4) {
      // VGAN_B200_ONES: columns D .. D + 2 of the resident row operand (1.0 in every stored row) become the three bf16
      // pieces of -(|z_i|^2 - 6) / 2 (the norm includes the three ones; the column operand keeps its ones), so that
      //   S' = z_i.z_j - (|z_i|^2 - 6) / 2   and   u = -2 c0 S' + c0 |z_j|^2 = c0 d2
      // with no row norm in the epilogue.  Split operands: the hi chunk is patched, the lo chunk is 0 there.
      mbar_wait(&bar->zi_full, 0);
      const float a = -0.5f * (p.nrm[static_cast<int64_t>(u.ti) * kTile + row] - 6.f);
      const __nv_bfloat16 h0 = __float2bfloat16_rn(a);
      const float r1 = a - __bfloat162float(h0);
      const __nv_bfloat16 h1 = __float2bfloat16_rn(r1);
      const __nv_bfloat16 h2 = __float2bfloat16_rn(r1 - __bfloat162float(h1));
      const __nv_bfloat16 piece[3] = {h0, h1, h2};
#pragma unroll
      for (int e = 0; e < 3; ++e) {
        const int col = static_cast<int>(p.D) + e;
        const int kc = col >> 6, cc = col & 63;                      // chunk, column inside the 128-byte swizzle row
        *reinterpret_cast<__nv_bfloat16*>(zi_smem + kc * kChunkBytes + row * 128 + (((cc >> 3) ^ (row & 7)) << 4) +
                                          (cc & 7) * 2) = piece[e];
      }
      fence_proxy_async_smem();
      mbar_arrive(&bar->zi_ready);
    }
    if (ones) {
      if (u.pv) epilogue_tiles<LADDER, SPLIT, true, true>(p, u, bar, p_smem, nj_smem, tmem, warp, lane, acc_x, acc_y, rsum);
      else epilogue_tiles<LADDER, SPLIT, false, true>(p, u, bar, p_smem, nj_smem, tmem, warp, lane, acc_x, acc_y, rsum);
    } else {
      if (u.pv) epilogue_tiles<LADDER, SPLIT, true, false>(p, u, bar, p_smem, nj_smem, tmem, warp, lane, acc_x, acc_y, rsum);
      else epilogue_tiles<LADDER, SPLIT, false, false>(p, u, bar, p_smem, nj_smem, tmem, warp, lane, acc_x, acc_y, rsum);
    }

    if (u.pv && warp < 8) prefetch_grad_rows(p, u.ti, u.row_is_x, u.slice, slice_chunks);   // inputs of the gradient rows -> L2
    // ---- block sums
    acc_x = warp_sum(acc_x);
    acc_y = warp_sum(acc_y);
    if (lane == 0) {
      bar->red[warp][0] = acc_x;
      bar->red[warp][1] = acc_y;
    }
    asm volatile("bar.sync 1, %0;" ::"n"(kNumEpiThreads) : "memory");
    if (threadIdx.x == 0 && !(p.pass_flags & kPassNoSums)) {
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

    // ---- gradient rows: dZ_i = 4 c_i (O_i - rowsum(P)_i z_i)  (x dY_mul)
    if (u.pv) {
      mbar_wait(&bar->o_full, 0);                       // every PV has completed: the P buffers are dead
      tc_fence_after();
      float* rs_smem = reinterpret_cast<float*>(p_smem);   // [4 column quarters][128 rows] row sums of P
      rs_smem[g * kTile + row] = rsum;
      asm volatile("bar.sync 1, %0;" ::"n"(kNumEpiThreads) : "memory");
      float rtot = rs_smem[row] + rs_smem[kTile + row] + rs_smem[2 * kTile + row] + rs_smem[3 * kTile + row];
      if (ones && warp < 8) {       // the row sums of the ROUNDED P' (all three parts with split operands): column D of O
        rtot = __uint_as_float(tmem_ld_32x32_x1(tmem_o + lane_addr + static_cast<uint32_t>(p.D)));
        tmem_ld_wait();
      }
      if (warp < 8)       // O through the ring (dead: every stage was consumed by an MMA that has completed)
        write_grad_rows(p, tmem_o, reinterpret_cast<float*>(ring), reinterpret_cast<float*>(p_smem + 4096), u.ti,
                        u.row_is_x, u.slice, slice_chunks, rtot, p.col_split > 1, u.first != 0);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kWarpAlloc) tmem_dealloc(tmem, 512);
}

// Cuts of the grid (host): the gradient units are cut into column ranges while there are fewer of them than SMs
// (every range at least 4 column tiles), the S_xx rows into chunks short enough to pack the tail.
static void plan_units(TcParams& p, int n_sms) {
  const int ngrad = (p.gx_end - p.gx_begin) + (p.gy_end - p.gy_begin);
  const int T = p.tiles_x + p.tiles_y;
  p.col_split = 1;
  if (p.pv && ngrad > 0 && !(p.pass_flags & kPassAccumulate)) {
    int cs = n_sms / ngrad;
    const int cap = T / (T <= 16 ? 2 : 4);        // at least 4 column tiles per range (2 on the smallest problems)
    if (cs > cap) cs = cap;
    if (cs > 1) p.col_split = cs;
  }
  // S_xx chunk length: about a quarter of the average work per SM, 8 .. 64 tiles
  long long xx_tiles = 0;
  for (int t = p.xx_part; p.n_xx_units > 0 && t < p.tiles_x; t += p.xx_nparts) xx_tiles += p.tiles_x - t;
  const long long per_sm = (static_cast<long long>(ngrad) * T + xx_tiles) / (n_sms > 0 ? n_sms : 1);
  int c = static_cast<int>(per_sm / 4);
  if (c < 8) c = 8;
  if (c > 64) c = 64;
  p.xx_chunk = c;
}

int launch_tc(const TcParams& p_in, const CUtensorMap& map, const CUtensorMap& mapB, const CUtensorMap& mapV,
              int ladder, cudaStream_t st, int split, int n_sms) {
  TcParams p = p_in;
  plan_units(p, n_sms);
  const int ngrad = (p.gx_end - p.gx_begin) + (p.gy_end - p.gy_begin);
  int n_xx = 0;
  if (p.n_xx_units > 0)
    for (int t = p.xx_part; t < p.tiles_x; t += p.xx_nparts) n_xx += xx_chunks_of(p.tiles_x, t, p.xx_chunk);
  const int grid = ngrad * p.col_split + n_xx;
  if (grid <= 0) return VGAN_B200_OK;
  if (p.col_split > 1) {
    // column ranges of a unit add up in the output rows
    const int64_t ye = p.row_end < p.Ny ? p.row_end : p.Ny, xe = p.row_end < p.Nx ? p.row_end : p.Nx;
    if (p.dY && ye > p.row_begin)
      VG_CUDA_CHECK(cudaMemsetAsync(p.dY, 0, static_cast<size_t>(ye - p.row_begin) * p.D * sizeof(float), st));
    if (p.dX && p.gx_end > p.gx_begin && xe > p.row_begin)
      VG_CUDA_CHECK(cudaMemsetAsync(p.dX, 0, static_cast<size_t>(xe - p.row_begin) * p.D * sizeof(float), st));
  }
  const TcGeom geo = tc_geom(split != 0, p.kchunks);
  if (split) {
    auto kern = ladder ? mmd2_tc_kernel<true, true> : mmd2_tc_kernel<false, true>;
    VG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
    kern<<<grid, kThreads, geo.total, st>>>(map, mapB, mapV, p);
  } else {
    auto kern = ladder ? mmd2_tc_kernel<true, false> : mmd2_tc_kernel<false, false>;
    VG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
    kern<<<grid, kThreads, geo.total, st>>>(map, mapB, mapV, p);
  }
  count_launch();
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

}  // namespace vg
