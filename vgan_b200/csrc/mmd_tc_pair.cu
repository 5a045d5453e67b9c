// K1, CTA-pair variant (cta_group::2) of the cluster kernel, for D_p = 2 or 4 chunks per slice.
// Cluster = 2 row tiles x CS feature slices.  CTAs 2c (leader) and 2c+1 form the tcgen05 pair of
// slice c; they own DIFFERENT row tiles (I0, I1) and share every B operand:
//   Gram  M=256 = [Z_I0 ; Z_I1] chunks (each CTA loads its own 16 KiB), B = Z_J chunk, each CTA
//         loads only HALF of its 128 rows (8 KiB)
//   PV    M=256 = [P(I0,t) ; P(I1,t)], B = Z_Jt[:, slice c], each CTA holds half of the features
// P tiles are exchanged among the CTAs of the SAME row tile exactly as in mmd_tc_cluster.cu.
//
// What bounds this kernel is not the tensor pipe, L2 or shared memory but the ISSUE stream of the
// single thread that feeds the tensor pipe (tools/gpu_cluster_timers.py, gpu_mma_rate.py,
// gpu_tma_rate.py): a tcgen05.mma costs ~48 issue cycles against 64 execution cycles at N=128,
// every mbarrier wait ~100 more.  Hence TWO issuing warps with independent operand rings:
//   warp 9  (leader)  Gram stream   G ring: kG2Stages x 32 KiB (A chunk 16 KiB + B chunk 16 KiB, N = 256)
//   warp 10 (leader)  PV stream     V ring: kVStages x 16 KiB (64 rows of Z_J x this CTA's half slice)
// The two streams only meet through the S buffers, the P slots and TMEM; the tensor pipe interleaves
// them in arrival order, so no static Gram/PV schedule (and no "lag") is needed.
// TMA issue is latency-bound per thread as well (~650 cycles per wait/expect/issue round trip):
// two producer warps alternate over the G stages, a third feeds the V ring.
// PERSISTENT, stream-K style: the grid is exactly the number of clusters that can be resident (15 on a
// B200: 8-CTA clusters fit 120 of the 148 SMs); the work = (row-tile pair, slice group) units x K rounds of
// CS column tiles is cut into equal contiguous ranges, so every cluster finishes at the same time whatever
// the unit count (no wave quantisation, also on the small shards of a multi-GPU run), nothing is left
// pending in the hardware queue (the S_xx kernel can then run on the 28 SMs clusters cannot use), and a
// unit cut between two clusters accumulates its gradient rows with red.global.add (dY zeroed by the host).
// Only the leader issues MMAs; the peer's epilogue arrives on the leader's s_empty, the peer's warp 9
// forwards its p_full to the leader (p_pair), and every tcgen05.commit is multicast to both CTAs.
// Slot loop of the P tiles (what bounds the kernel: tools/gpu_cluster_timers.py, and 8 820 instead of 11 180 cycles
// per round when the exchange moves 16 bytes): a tile is PARKED as soon as this pair's own PV of the slot's previous
// tile is done (own_free) and stored to L2 at once; only the multicast load into the peers' slots waits until every
// pair has finished with the previous tile (slot_free, waited by the publisher), so the store is off the loop.
// S[b] is released right after its columns reach registers (the column norms have their own barrier, nj_empty).
#include <cstdlib>

#include "common.cuh"
#include "ptx.cuh"
#include "tc_epi.cuh"

// In-kernel wait timers (tools/gpu_cluster_timers.py): compiled in only with -DVGAN_B200_TIMERS (make TIMERS=1);
// the product build keeps clock reads out of the single-thread issue loops.
#ifdef VGAN_B200_TIMERS
#define VG_CLK() clock64()
#else
#define VG_CLK() 0ll
#endif

namespace vg {

constexpr int kGStagesMax = 4;                     // barrier slots of the Gram operand ring
// ONE Gram issue covers the pair's next TWO column tiles (N = 256: the leader holds
// the 128 rows of tile k, the peer those of tile k + 1; D = both S buffers at once).  The row operand [Z_I0; Z_I1] is
// then streamed once per two column tiles instead of once per tile (-21 % of all L2 -> SM bytes at D_p = 1024), a Gram
// instruction runs 128 instead of 64 cycles, and a ring stage (A chunk 16 KiB + B chunk 16 KiB) feeds 512 instead of
// 256 tensor cycles, so the same 96 KiB in flight cover twice the time.  S is then double-buffered only against the PV
// stream, not against its own Gram.  Ranges are padded to an even number of rounds with zero-coefficient tiles.
#ifndef VG_G2_STAGES
#define VG_G2_STAGES 3
#endif
#ifndef VG_G2_SLOTS
#define VG_G2_SLOTS 3
#endif
constexpr int kG2Stages = VG_G2_STAGES;
constexpr int kG2StageBytes = 2 * kChunkBytes;                // 32 KiB
constexpr int kVStages = 2;                        // V ring: half tiles (64 rows of Z_J)
constexpr int kVStageBytes = kChunkBytes;          // 2 boxes of [64 x 64] bf16
constexpr int kQPBytes = 2 * kChunkBytes;
constexpr int kQThreads = 512;   // warps: 0-7 epilogue (192 registers), 8-15 service (64 registers): 8/11 G producers,
                                 // 9 Gram MMA (peer: P forwarder), 10 PV MMA, 12 V producer, 13 P publisher, 14 norms
// VG_G2_STAGES / VG_G2_SLOTS are compile-time tuning knobs (same shared-memory total either way).  Measured
// (tools/gpu_cluster_timers.py, cycles per round of 4 column tiles): 3 stages + 3 P slots 10 840; 2 stages + 4 slots
// 12 710 — the Gram warp then waits 6 600 cycles per round for operands, and the slot wait does not shrink because
// the in-order PV stream of every pair stalls on whichever pair is late.
constexpr int kQSlotsMax = 4;

struct __align__(8) PairBarriers {
  uint64_t g_full[kGStagesMax], g_empty[kGStagesMax];
  uint64_t v_full[kVStages], v_empty[kVStages];
  uint64_t s_full[2], s_empty[2];
  uint64_t nj_full[2], nj_empty[2];
  uint64_t p_full[kQSlotsMax], slot_free[kQSlotsMax], p_pair[kQSlotsMax], pub_full[kQSlotsMax], own_free[kQSlotsMax];
  uint64_t o_full;
  uint32_t tmem_base;
  float red[8][4];
};

struct PairCfg {
  static constexpr int GS = kG2Stages;              // Gram ring stages
  static constexpr int GSB = kG2StageBytes;         // bytes per stage
  static constexpr int QS = VG_G2_SLOTS;            // P slots, 32 KiB each
  static constexpr int SmemG = 0;
  static constexpr int SmemV = SmemG + GS * GSB;
  static constexpr int SmemP = SmemV + kVStages * kVStageBytes;
  static constexpr int SmemNj = SmemP + QS * kQPBytes;
  static constexpr int SmemBar = SmemNj + 2 * kTile * 4;
  static constexpr int SmemTotal = SmemBar + static_cast<int>(sizeof(PairBarriers));
  static_assert(SmemTotal <= 232448, "shared memory budget");
  static_assert(SmemV % 1024 == 0 && SmemP % 1024 == 0, "operand tiles need 1024-byte alignment");
  static_assert(QS * kQPBytes <= kXchgBytesPerSm, "exchange scratch per SM");
};

#define VG_PAIR_DECODE_RANGE \
    long long w_lo, w_hi; \
    if (pi < n_full) { w_lo = (u_first + pi) * static_cast<long long>(K); w_hi = w_lo + K; } \
    else if (pi == n_full && has_head) { w_lo = w_begin; w_hi = head_end; } \
    else { w_lo = tail_begin; w_hi = w_end; } \
    const int unit_g = static_cast<int>(w_lo / K); \
    const int r0 = static_cast<int>(w_lo - static_cast<long long>(unit_g) * K); \
    const int Kv = static_cast<int>(w_hi - w_lo);      /* rounds that count */ \
    const int Kp = (Kv + 1) & ~1;                      /* rounds that run (even: the extra one has zero weights) */ \
    const bool partial = Kv < K; \
    const int TTp = Kp * CS; \
    const int tile0 = r0 * CS; \
    const int grp = unit_g % p.ngroups; \
    const int slice = grp * CS + c; \
    int unit = 2 * (unit_g / p.ngroups) + hh; \
    const bool dup = unit >= ngrad; \
    if (dup) unit = ngrad - 1; \
    const int row_is_x = unit < ngx; \
    const int ti = row_is_x ? p.gx_begin + unit : p.tiles_x + p.gy_begin + (unit - ngx); \
    const int slice_chunks = max(0, min(p.slice_chunks, p.kchunks - slice * p.slice_chunks)); \
    const int my_chunks = slice_chunks >> 1; \
    (void)0

template <bool LADDER>
__global__ void __launch_bounds__(kQThreads, 1)
mmd2_tc_pair_kernel(const __grid_constant__ CUtensorMap zmap, const __grid_constant__ CUtensorMap zmap64,
                    const __grid_constant__ CUtensorMap vmap64, const TcParams p) {
  // zmap: row operand (128-row boxes), zmap64: Gram column operand (64-row boxes), vmap64: V operand of the gradient
  // GEMM (64-row boxes); one array on the plain bf16 path, the split arrays of prologue.cu with BF16X3
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  using Cfg = PairCfg;
  constexpr int kQSlots = Cfg::QS;
  constexpr int kQSmemG = Cfg::SmemG, kQSmemV = Cfg::SmemV, kQSmemP = Cfg::SmemP, kQSmemNj = Cfg::SmemNj,
                kQSmemBar = Cfg::SmemBar;
  PairBarriers* bar = reinterpret_cast<PairBarriers*>(smem + kQSmemBar);
  float* nj_smem = reinterpret_cast<float*>(smem + kQSmemNj);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int CS = p.cs;                                 // feature slices per cluster; cluster size = 2 CS
  const int rank = static_cast<int>(cluster_ctarank());
  const int c = rank >> 1;                             // pair index = slice within the group
  const int hh = rank & 1;                             // 0: leader (row tile I0), 1: peer (row tile I1)
  const int csz = 2 * CS;
  const int ngx = p.gx_end - p.gx_begin;
  const int ngrad = ngx + (p.gy_end - p.gy_begin);
  const uint16_t pair_mask = static_cast<uint16_t>(3u << (2 * c));
  const int T = p.tiles_x + p.tiles_y;                 // column tiles
  const int K = (T + CS - 1) / CS;                     // rounds of CS column tiles per unit (the tail is padded)
  // ---- stream-K partition: units x K rounds, equal contiguous ranges per cluster
  const int n_units = ((ngrad + 1) / 2) * p.ngroups;   // (row-tile pair, slice group)
  const long long W = static_cast<long long>(n_units) * K;
  const int n_clusters = static_cast<int>(gridDim.x) / csz;
  const int vc = static_cast<int>(blockIdx.x) / csz;
  const long long w_begin = W * vc / n_clusters, w_end = W * (vc + 1) / n_clusters;
  // Order of the ranges of this cluster: its FULL units first, every one swept from column round 0, so that all
  // clusters walk the column operand in step (one HBM fetch per Z_J tile serves them all through L2: Z is
  // twice the L2 at the bench shape); the cut units (head, tail) last.
  const long long u_first = (w_begin + K - 1) / K, u_last = w_end / K;
  const bool one_piece = u_first > u_last;              // the whole range lies inside one unit
  const int n_full = one_piece ? 0 : static_cast<int>(u_last - u_first);
  const long long head_end = one_piece ? w_end : u_first * K;
  const long long tail_begin = one_piece ? w_end : u_last * K;
  const bool has_head = head_end > w_begin;
  const int n_parts = n_full + (has_head ? 1 : 0) + (w_end > tail_begin ? 1 : 0);
#ifdef VGAN_B200_DEBUG
  const uint32_t pbytes = (p.dbg_flags & 1) ? 16u : static_cast<uint32_t>(kQPBytes);   // bit 0: measurement only
  const int dbg_flags = p.dbg_flags;
#else
  constexpr uint32_t pbytes = static_cast<uint32_t>(kQPBytes);
  constexpr int dbg_flags = 0;
#endif

  if (threadIdx.x == 0) {
    for (int s = 0; s < kGStagesMax; ++s) {
      mbar_init(&bar->g_full[s], 1);
      mbar_init(&bar->g_empty[s], 1);
    }
    for (int s = 0; s < kVStages; ++s) {
      mbar_init(&bar->v_full[s], 1);
      mbar_init(&bar->v_empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&bar->s_full[b], 1);
      mbar_init(&bar->s_empty[b], hh == 0 ? 512 : 256);   // leader: both epilogues
      mbar_init(&bar->nj_full[b], 1);
      mbar_init(&bar->nj_empty[b], 256);                  // this CTA's epilogue threads
    }
    for (int s = 0; s < kQSlots; ++s) {
      mbar_init(&bar->p_full[s], 1);
      mbar_init(&bar->slot_free[s], CS);
      mbar_init(&bar->p_pair[s], 1);
      mbar_init(&bar->pub_full[s], 1);
      mbar_init(&bar->own_free[s], 1);
    }
    mbar_init(&bar->o_full, 1);
    fence_barrier_init();
  }
  if (warp == 11) {
    tmem_alloc_2cta(&bar->tmem_base, 512);
    tmem_relinquish_2cta();
  }
  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&zmap);
    tma_prefetch_desc(&zmap64);
    tma_prefetch_desc(&vmap64);
  }
  tc_fence_before();
  cluster_sync_all();          // peers' barriers are initialised before any remote arrive / copy
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;
  const uint32_t tmem_o = tmem + 256;


  // ---- pipeline state that persists across the work ranges of this CTA (each warp uses its role's part)
  uint32_t rs = 0, rph = 0;            // ring stage / phase: G ring (warps 8, 11, 9), V ring (warps 12, 10)
  uint32_t kg = 0;                     // Gram tiles of this CTA so far: S / norm buffer kg & 1, parity (kg >> 1) & 1
  uint32_t pf_mask = 0, sf_mask = 0, pub_mask = 0, of_mask = 0;   // per-slot phase parities of p_full / slot_free / pub_full / own_free
  uint32_t part = 0;                   // parity of o_full
  long long tm0 = 0, tm1 = 0, tm2 = 0, tm3 = 0;      // debug timers (meaning depends on the role)
  const long long t_begin = VG_CLK();
  unsigned long long gt_begin;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_begin));
  // resident now: the host makes the S_xx launch wait until every persistent CTA has counted itself in
  if (p.started && threadIdx.x == 0) atomicAdd(p.started, 1u);
  long long tm_tail = 0, tm_t1 = 0, tm_t2 = 0, tm_t3 = 0;   // tail: barrier #1, row-sum exchange + #2, gradient rows
  float* rsx = reinterpret_cast<float*>(smem + kQSmemP);       // [CS][128] row-sum exchange, reuses P slot 0

  if (warp >= 8) {
    // ===================== service warps (TMA, MMA issue, P exchange): few registers =====================
    setmaxnreg_dec<64>();
    for (int pi = 0; pi < n_parts; ++pi, ++part) {
    VG_PAIR_DECODE_RANGE;
    (void)partial; (void)dup; (void)row_is_x; (void)Kv;
  if (warp == 8 || warp == 11) {
    // ===================== G producers: Gram operand stages, in the Gram warp's order =====================
    const uint32_t pj = warp == 8 ? 0u : 1u;
    {
      uint32_t it = 0;
      for (int k = 0; k < Kp; k += 2) {
        const int jt = min(tile0 + CS * (k + hh) + c, T - 1);   // leader: column tile of round k, peer: of round k + 1
        for (int kc = 0; kc < p.gchunks; ++kc, ++it) {
          if ((it & 1u) == pj) {
            mbar_wait(&bar->g_empty[rs], rph ^ 1);
            uint8_t* dst = smem + kQSmemG + rs * kG2StageBytes;
            if (dbg_flags & 2) {            // measurement only: no operand loads
              if (lane == 0 && hh == 0) mbar_arrive(&bar->g_full[rs]);
            } else if (lane == 0) {
              if (hh == 0) mbar_arrive_expect_tx(&bar->g_full[rs], 2 * kG2StageBytes);
              tma_load_2d_2cta(dst, &zmap, &bar->g_full[rs], kc * kChunk, ti * kTile);
            } else if (lane <= 2) {           // the 128 rows of this CTA's column tile as two 64-row boxes
              tma_load_2d_2cta(dst + kChunkBytes + (lane - 1) * (kChunkBytes / 2), &zmap64, &bar->g_full[rs],
                               kc * kChunk, jt * kTile + 64 * (lane - 1));
            }
            __syncwarp();
          }
          if (++rs == kG2Stages) { rs = 0; rph ^= 1; }
        }
      }
    }
  } else if (warp == 12) {
    // ===================== V producer: half tiles of Z_J[:, this CTA's half of the slice] =====================
    if (my_chunks > 0) {
      for (int t = 0; t < TTp; ++t) {
        const int jt = min(tile0 + t, T - 1);
        for (int vh = 0; vh < 2 * p.v_parts; ++vh) {     // split operands, first pass: V_hi then V_lo against the same P tile
          const int h = vh & 1;
          const int vbase = vh < 2 ? p.v_chunk0 : p.v_chunk1;
          mbar_wait(&bar->v_empty[rs], rph ^ 1);
          uint8_t* dst = smem + kQSmemV + rs * kVStageBytes;
          if (dbg_flags & 4) {            // measurement only: no V loads
            if (lane == 0 && hh == 0) mbar_arrive(&bar->v_full[rs]);
          } else {
            if (lane == 0 && hh == 0) mbar_arrive_expect_tx(&bar->v_full[rs], slice_chunks * (kChunkBytes / 2));
            if (lane < my_chunks)
              tma_load_2d_2cta(dst + lane * (kChunkBytes / 2), &vmap64, &bar->v_full[rs],
                               (vbase + slice * p.slice_chunks + hh * my_chunks + lane) * kChunk,
                               jt * kTile + 64 * h);
          }
          __syncwarp();
          if (++rs == kVStages) { rs = 0; rph ^= 1; }
        }
      }
    }
  } else if (warp == 14) {
    // ===================== bulk-copy producer: scaled column norms of this CTA's tiles ============
    if (lane == 0) {
      for (int k = 0; k < Kp; ++k, ++kg) {
        const int jt = min(tile0 + CS * k + c, T - 1);
        const uint32_t b = kg & 1;
        mbar_wait(&bar->nj_empty[b], ((kg >> 1) & 1) ^ 1);
        mbar_arrive_expect_tx(&bar->nj_full[b], kTile * 4);
        bulk_load_1d(nj_smem + b * kTile, p.nrms + static_cast<int64_t>(jt) * kTile, kTile * 4,
                     &bar->nj_full[b]);
      }
    }
  } else if (warp == 13) {
    // ===================== publisher: sends this CTA's P tiles to the other slices of its row tile =======
    // (kept off the epilogue warps).  pub_full[slot] counts only this CTA's own tiles (waiting on p_full would
    // alias: it advances through the peers' tiles in between, and a parity wait cannot tell phases two apart)
    uint16_t peers = 0;                                   // the other slices of MY row tile
    for (int d = 0; d < CS; ++d)
      if (d != c) peers |= static_cast<uint16_t>(1u << (2 * d + hh));
    const uint32_t smid = sm_id();
    if (smid >= static_cast<uint32_t>(kXchgSms)) __trap();
    uint8_t* xbase = p.xchg + static_cast<size_t>(smid) * kXchgBytesPerSm;
    for (int k = 0; k < Kp; ++k) {
      const int t = CS * k + c;
      const int slot = t % kQSlots;
      mbar_wait(&bar->pub_full[slot], (pub_mask >> slot) & 1);
      pub_mask ^= 1u << slot;
      if (lane == 0 && peers) {
        const uint32_t src = smem_u32(smem + kQSmemP + slot * kQPBytes);
        if (dbg_flags & 8) {
          // DSMEM route: one bulk copy per peer (~12 B/clk per SM: 3 x 32 KiB take ~8000 cycles)
          if (t >= kQSlots) mbar_wait(&bar->slot_free[slot], (sf_mask >> slot) & 1);
          for (int d = 0; d < CS; ++d)
            if (d != c) {
              const uint32_t peer = static_cast<uint32_t>(2 * d + hh);
              bulk_copy_to_peer(mapa_u32(src, peer), src, pbytes, mapa_u32(smem_u32(&bar->p_full[slot]), peer));
            }
        } else {
          // L2 route: bulk store to this SM's scratch slot, then ONE multicast bulk load into the same slot of
          // every peer.  The scratch slot is free for the same reason the shared-memory slot is (slot_free).
          // The store to L2 does not wait for the receivers: the tile was parked as soon as THIS pair's PV(t - slots)
          // had finished (own_free).  Only the multicast load into the peers' slots waits until every pair is done
          // with the slot's previous tile, so store + commit are off the slot loop's critical path.
          uint8_t* g = xbase + slot * kQPBytes;
          bulk_store_global(g, src, pbytes);
          bulk_commit_wait_all();
          if (t >= kQSlots) mbar_wait(&bar->slot_free[slot], (sf_mask >> slot) & 1);   // every CTA finished PV(t - kQSlots)
          bulk_load_multicast(src, g, pbytes, smem_u32(&bar->p_full[slot]), peers);
        }
      }
      if (t >= kQSlots) sf_mask ^= 1u << slot;
      __syncwarp();
    }
  } else if (warp == 9 && hh == 1) {
    // ===================== peer CTA: forward "my P tile t is in place" to the leader's PV warp =========
    // P tiles written by another slice land by bulk copy: the RECEIVER arms its own barrier (no remote
    // arrive): the first kQSlots tiles of the range here, tile t + kQSlots right after the wait for tile t
    if (lane == 0)
      for (int t = 0; t < kQSlots && t < TTp; ++t)
        if (t % CS != c) mbar_arrive_expect_tx(&bar->p_full[t % kQSlots], pbytes);
    const uint32_t pp0 = mapa_u32(smem_u32(&bar->p_pair[0]), static_cast<uint32_t>(rank - 1));
    for (int t = 0; t < TTp; ++t) {
      const int slot = t % kQSlots;
      mbar_wait(&bar->p_full[slot], (pf_mask >> slot) & 1);
      pf_mask ^= 1u << slot;
      if (lane == 0) {
        if (t + kQSlots < TTp && (t + kQSlots) % CS != c) mbar_arrive_expect_tx(&bar->p_full[slot], pbytes);
        mbar_remote_arrive(pp0 + slot * 8);
      }
      __syncwarp();
    }
  } else if (warp == 9) {
    // ===================== leader CTA: Gram MMA stream (warp-uniform; one lane issues) ==========
    constexpr uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
    const uint32_t g_lo = (smem_u32(smem + kQSmemG) & 0x3FFFFu) >> 4;
    {
      constexpr uint32_t idesc_gram2 = make_idesc_bf16(2 * kTile, 2 * kTile, 0, 0);
      for (int k = 0; k < Kp; k += 2, kg += 2) {       // kg is even here: tile k -> S[0], tile k + 1 -> S[1]
        {
          const long long t0 = VG_CLK();
          const uint32_t par = ((kg >> 1) & 1) ^ 1;
          mbar_wait(&bar->s_empty[0], par);              // both CTAs' epilogues released S[0] and S[1]
          mbar_wait(&bar->s_empty[1], par);
          tm1 += VG_CLK() - t0;
        }
        tc_fence_after();
        for (int kc = 0; kc < p.gchunks; ++kc) {
          {
            const long long t0 = VG_CLK();
            mbar_wait(&bar->g_full[rs], rph);
            tm0 += VG_CLK() - t0;
          }
          tc_fence_after();
          const uint32_t a_lo = g_lo + rs * (kG2StageBytes >> 4);
          const uint32_t b_lo = a_lo + (kChunkBytes >> 4);
          if (elect_one()) {
#pragma unroll
            for (int kk = 0; kk < kChunk / 16; ++kk) {
              const uint64_t adesc = (static_cast<uint64_t>(desc_hi) << 32) | (a_lo + kk * 2);
              const uint64_t bdesc = (static_cast<uint64_t>(desc_hi) << 32) | (b_lo + kk * 2);
              umma_bf16_ss_2cta(tmem, adesc, bdesc, idesc_gram2, (kc | kk) != 0);
            }
            umma_commit_2cta_mc(&bar->g_empty[rs], pair_mask);
            if (kc == p.gchunks - 1) {
              umma_commit_2cta_mc(&bar->s_full[0], pair_mask);
              umma_commit_2cta_mc(&bar->s_full[1], pair_mask);
            }
          }
          __syncwarp();
          if (++rs == kG2Stages) { rs = 0; rph ^= 1; }
        }
      }
    }
  } else if (warp == 10 && hh == 0) {
    // ===================== leader CTA: PV MMA stream =====================
    constexpr uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
    constexpr uint32_t lbo_v = (static_cast<uint32_t>(kChunkBytes / 2) >> 4) << 16;   // next 64-feature block
    const uint32_t idesc_pv = make_idesc_bf16(2 * kTile, slice_chunks * kChunk, 0, 1);
    const uint32_t v_lo0 = (smem_u32(smem + kQSmemV) & 0x3FFFFu) >> 4;
    const uint32_t p_lo0 = (smem_u32(smem + kQSmemP) & 0x3FFFFu) >> 4;
    if (lane == 0)                                       // arm the first remote tiles of the range (see warp 9, peer)
      for (int t = 0; t < kQSlots && t < TTp; ++t)
        if (t % CS != c) mbar_arrive_expect_tx(&bar->p_full[t % kQSlots], pbytes);
    for (int t = 0; t < TTp; ++t) {
      const int slot = t % kQSlots;
      {
        const long long t0 = VG_CLK();
        mbar_wait(&bar->p_full[slot], (pf_mask >> slot) & 1);   // my P(I0, t)
        if (lane == 0 && t + kQSlots < TTp && (t + kQSlots) % CS != c)
          mbar_arrive_expect_tx(&bar->p_full[slot], pbytes);    // arm the slot for its next (remote) tile
        mbar_wait(&bar->p_pair[slot], (pf_mask >> slot) & 1);   // the peer's P(I1, t)
        tm2 += VG_CLK() - t0;
      }
      pf_mask ^= 1u << slot;
      tc_fence_after();
      const uint32_t p_lo = p_lo0 + slot * (kQPBytes >> 4);
      if (my_chunks > 0) {
        for (int vh = 0; vh < 2 * p.v_parts; ++vh) {
          const int h = vh & 1;
          {
            const long long t0 = VG_CLK();
            mbar_wait(&bar->v_full[rs], rph);
            tm3 += VG_CLK() - t0;
          }
          tc_fence_after();
          const uint32_t v_lo = v_lo0 + rs * (kVStageBytes >> 4);
          const uint32_t pa_lo = p_lo + h * (kChunkBytes >> 4);
          if (elect_one()) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint64_t adesc = (static_cast<uint64_t>(desc_hi) << 32) | (pa_lo + j * 2);
              const uint64_t bdesc = (static_cast<uint64_t>(desc_hi) << 32) | (v_lo + j * (2048 >> 4)) | lbo_v;
              umma_bf16_ss_2cta(tmem_o, adesc, bdesc, idesc_pv, (t | vh | j) != 0);
            }
            umma_commit_2cta_mc(&bar->v_empty[rs], pair_mask);
          }
          __syncwarp();
          if (++rs == kVStages) { rs = 0; rph ^= 1; }
        }
      }
      // both CTAs of the pair are done with slot: tell the two writers (rows I0 and I1) of tile t + kQSlots
      if (t + kQSlots < TTp && elect_one()) {
        umma_commit_2cta_mc(&bar->slot_free[slot], static_cast<uint16_t>(3u << (2 * ((t + kQSlots) % CS))));
        // the next tile of this slot is written by this pair: its epilogues may park it once OUR PV(t) is done
        if ((t + kQSlots) % CS == c) umma_commit_2cta_mc(&bar->own_free[slot], pair_mask);
      }
      __syncwarp();
    }
    if (elect_one()) umma_commit_2cta_mc(&bar->o_full, pair_mask);
    __syncwarp();
  }
    tc_fence_before();
    cluster_sync_all();      // #1, #2, #3 of the epilogue warps below
    cluster_sync_all();
    cluster_sync_all();
    }
  } else {
    // ===================== epilogue warps: this CTA's own column tiles (register-hungry) =====================
    setmaxnreg_inc<192>();
    for (int pi = 0; pi < n_parts; ++pi, ++part) {
    VG_PAIR_DECODE_RANGE;
    (void)TTp; (void)my_chunks;
    {
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
      coef_x[m] = p.par[kParCoefX + m];
      coef_y[m] = p.par[kParCoefY + m];
    }
    // common-mode level c_j w0 of (this row block, X columns) / (this row block, Y columns)
    const float cw0_x = p.par[kParCw0 + (row_is_x ? 0 : 2)], cw0_y = p.par[kParCw0 + (row_is_x ? 0 : 2) + 1];
    float acc_x = 0.f, acc_y = 0.f, rsum = 0.f;
    // One column tile = three steps.  load: wait for S[b] (and the norms), pull this thread's 64 columns into registers
    // and release S[b] at once (before the math), so the next Gram into the buffer overlaps the tile's whole epilogue;
    // compute: kernel values, sums, P as bf16; park: P into its shared-memory slot for the PV warp and the publisher.
    auto load_S = [&](uint32_t b, uint32_t par, uint32_t (&va)[32], uint32_t (&vb)[32]) {
      const long long tq = VG_CLK();
      mbar_wait(&bar->nj_full[b], par);
      mbar_wait(&bar->s_full[b], par);
      tm0 += VG_CLK() - tq;
      tc_fence_after();
      const uint32_t taddr = tmem + lane_addr + b * kTile + g * 64;
      tmem_ld_32x32(taddr, va);
      tmem_ld_32x32(taddr + 32, vb);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&bar->s_empty[b]);
      if (hh == 1) mbar_remote_arrive(mapa_u32(smem_u32(&bar->s_empty[b]), static_cast<uint32_t>(rank - 1)));
    };
    auto compute = [&](int k, uint32_t b, const uint32_t (&va)[32], const uint32_t (&vb)[32], uint4 (&pk)[8]) {
      const long long tq = VG_CLK();
      const int tg = tile0 + CS * k + c;                 // column tile
      const bool valid = tg < T && k < Kv;               // padded rounds (tail of a unit, even-round padding): zero weights
      const bool col_is_x = min(tg, T - 1) < p.tiles_x;
      float cf[VGAN_B200_MAX_SCALES];
#pragma unroll
      for (int m = 0; m < VGAN_B200_MAX_SCALES; ++m) cf[m] = valid ? (col_is_x ? coef_x[m] : coef_y[m]) : 0.f;
      const float cw0 = valid ? (col_is_x ? cw0_x : cw0_y) : 0.f;
      const float* njp = nj_smem + b * kTile + g * 64;
      float ksum = 0.f;
      uint4 h0[4], h1[4];
      epi_compute_32<LADDER>(va, njp, ni, m2c0, cf, ratio, p.n_scales, ksum, rsum, h0, p.p_lo != 0, cw0);
      epi_compute_32<LADDER>(vb, njp + 32, ni, m2c0, cf, ratio, p.n_scales, ksum, rsum, h1, p.p_lo != 0, cw0);
#pragma unroll
      for (int i = 0; i < 4; ++i) { pk[i] = h0[i]; pk[4 + i] = h1[i]; }
      mbar_arrive(&bar->nj_empty[b]);                    // the column norms of this tile are no longer read
      if (valid) { if (col_is_x) acc_x += ksum; else acc_y += ksum; }
      tm1 += VG_CLK() - tq;
    };
    auto park = [&](int k, const uint4 (&pk)[8]) {
      long long tq = VG_CLK();
      const int t = CS * k + c;                          // tile index within the range (slot protocol)
      const int slot = t % kQSlots;
      if (t >= kQSlots) {
        mbar_wait(&bar->own_free[slot], (of_mask >> slot) & 1);    // this pair finished PV(t - kQSlots)
        of_mask ^= 1u << slot;
      }
      tm2 += VG_CLK() - tq; tq = VG_CLK();
      uint8_t* p_slot = smem + kQSmemP + slot * kQPBytes;
      uint8_t* p_row = p_slot + g * kChunkBytes + row * 128;
#pragma unroll
      for (int i = 0; i < 8; ++i)
        *reinterpret_cast<uint4*>(p_row + ((i ^ (row & 7)) << 4)) = pk[i];
      fence_proxy_async_smem();
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (threadIdx.x == 0) {
        mbar_arrive(&bar->p_full[slot]);     // own PV warp / forwarder
        mbar_arrive(&bar->pub_full[slot]);   // publisher warp
      }
      tm3 += VG_CLK() - tq;
    };
    {
      // both S buffers fill together: S[1] is pulled (and released) right after the math of S[0], before P(k) waits
      // for its slot, so the next two-tile Gram starts one tile's math after s_full instead of a whole tile later
      for (int k = 0; k < Kp; k += 2, kg += 2) {
        const uint32_t par = (kg >> 1) & 1;
        uint4 pk0[8], pk1[8];
        uint32_t va[32], vb[32];
        load_S(0, par, va, vb);
        compute(k, 0, va, vb, pk0);
        load_S(1, par, va, vb);
        park(k, pk0);
        compute(k + 1, 1, va, vb, pk1);
        park(k + 1, pk1);
      }
    }
    if (!dup) prefetch_grad_rows(p, ti, row_is_x, slice, slice_chunks);    // inputs of the gradient rows -> L2
    // ---- partial sums of this CTA's tiles
    acc_x = warp_sum(acc_x);
    acc_y = warp_sum(acc_y);
    if (lane == 0) { bar->red[warp][0] = acc_x; bar->red[warp][1] = acc_y; }
    asm volatile("bar.sync 1, 256;" ::: "memory");   // all epilogue warps are done with the norms
    nj_smem[g * kTile + row] = rsum;
    if (threadIdx.x == 0 && grp == 0 && !dup && !(p.pass_flags & kPassNoSums)) {
      double sx = 0.0, sy = 0.0;
      for (int wv = 0; wv < 8; ++wv) { sx += bar->red[wv][0]; sy += bar->red[wv][1]; }
      if (row_is_x) {
        atomicAdd(&p.sums[0], sx);
      } else {
        atomicAdd(&p.sums[1], sx);
        atomicAdd(&p.sums[2], sy);
      }
    }
    mbar_wait(&bar->o_full, part & 1);   // every PV of this range has completed: its P slots are dead
    tc_fence_after();
    }

    // ===================== cluster-wide row sums of P, then the gradient rows of this range =====================
    const long long tq_tail = VG_CLK();
    tc_fence_before();
    cluster_sync_all();      // #1: no CTA of the cluster reads a P slot any more
    const long long tq_s1 = VG_CLK();
    if (warp < 4) {
      const int row = warp * 32 + lane;
      const float tot = nj_smem[row] + nj_smem[kTile + row];     // both column halves of this CTA's tiles
      const uint32_t a = smem_u32(rsx + c * kTile + row);
      for (int d = 0; d < CS; ++d) st_cluster_f32(mapa_u32(a, static_cast<uint32_t>(2 * d + hh)), tot);
    }
    cluster_sync_all();      // #2: all partial row sums have arrived
    const long long tq_s2 = VG_CLK();
    if (warp < 8) {
      const int row = (warp & 3) * 32 + lane;
      float rtot = 0.f;
      for (int d = 0; d < CS; ++d) rtot += rsx[d * kTile + row];
      tc_fence_after();
      if (!dup)           // O through P slots 1 and 2 (dead since cluster barrier #1; slot 0 holds the row-sum exchange)
        write_grad_rows(p, tmem_o, reinterpret_cast<float*>(smem + kQSmemP + kQPBytes),
                        reinterpret_cast<float*>(smem + kQSmemP + 8192), ti, row_is_x, slice, slice_chunks, rtot, partial,
                        r0 == 0);
    }
    const long long tq_w = VG_CLK();
    tc_fence_before();
    cluster_sync_all();      // #3: O, the norm buffers and P slot 0 may be reused by the next range
    tm_tail += VG_CLK() - tq_tail;
    tm_t1 += tq_s1 - tq_tail; tm_t2 += tq_s2 - tq_s1; tm_t3 += tq_w - tq_s2;
    }
  }

#ifdef VGAN_B200_DEBUG
  if (p.dbg && lane == 0 && (warp == 0 || ((warp == 9 || warp == 10) && hh == 0))) {
    float* d = p.dbg + static_cast<int64_t>(blockIdx.x) * p.dbg_stride;
    const int o = warp == 0 ? 4 : (warp == 9 ? 0 : 8);
    d[o + 0] = static_cast<float>(tm0);
    d[o + 1] = static_cast<float>(tm1);
    d[o + 2] = static_cast<float>(tm2);
    d[o + 3] = static_cast<float>(tm3);
    if (warp == 9) {
      d[12] = static_cast<float>(VG_CLK() - t_begin);
      unsigned long long gt_end;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_end));
      d[13] = static_cast<float>(gt_end - gt_begin);
    }
    if (warp == 0) {
      d[14] = static_cast<float>(tm_tail);
      d[16] = static_cast<float>(tm_t1); d[17] = static_cast<float>(tm_t2); d[18] = static_cast<float>(tm_t3);
    }
  }
  if (p.timeline && threadIdx.x == 0) {
    atomicMin(&p.timeline[0], gt_begin);
    atomicMax(&p.timeline[1], global_timer_ns());
  }
#else
  (void)tm0; (void)tm1; (void)tm2; (void)tm3; (void)tm_tail; (void)tm_t1; (void)tm_t2; (void)tm_t3; (void)t_begin; (void)gt_begin;
#endif
  if (warp == 11) tmem_dealloc_2cta(tmem, 512);
}

int launch_tc_pair(vgan_b200_ctx* ctx, const TcParams& p, const CUtensorMap& map, const CUtensorMap& map64,
                   const CUtensorMap& mapV64, int ladder, cudaStream_t st, int* resident_ctas) {
  if (resident_ctas) *resident_ctas = 0;
  const int ngrad = (p.gx_end - p.gx_begin) + (p.gy_end - p.gy_begin);
  if (ngrad <= 0) return VGAN_B200_OK;
  auto kern = ladder ? mmd2_tc_pair_kernel<true> : mmd2_tc_pair_kernel<false>;
  const int smem_bytes = PairCfg::SmemTotal;
  VG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  const int n_units = ((ngrad + 1) / 2) * p.ngroups;
  cudaLaunchConfig_t cfg{};
  cfg.blockDim = dim3(kQThreads);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = static_cast<unsigned>(2 * p.cs);
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  // persistent grid: as many clusters as can be resident at once (cached per cluster size in the context; the same
  // for both instantiations: same block size and shared memory)
  int& ncl = ctx->resident_clusters[2 * p.cs];
  if (ncl == 0) {
    cfg.gridDim = dim3(static_cast<unsigned>(2 * p.cs * 64));
    VG_CUDA_CHECK(cudaOccupancyMaxActiveClusters(&ncl, kern, &cfg));
    if (ncl < 1) ncl = 1;
  }
  int cap = ncl;
  if (p.max_clusters > 0 && p.max_clusters < cap) cap = p.max_clusters;   // vgan_b200_set_cluster_limit
  const int n_clusters = cap < n_units ? cap : n_units;
  cfg.gridDim = dim3(static_cast<unsigned>(n_clusters * 2 * p.cs));
  if (resident_ctas) *resident_ctas = (n_clusters < ncl ? n_clusters : ncl) * 2 * p.cs;
  VG_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, map, map64, mapV64, p));
  count_launch();
  return VGAN_B200_OK;
}

}  // namespace vg
