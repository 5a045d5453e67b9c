// Shared epilogue math of the tcgen05 K1 kernels: 32 TMEM columns of S -> kernel values.
#pragma once
#include "common.cuh"
#include "ptx.cuh"

namespace vg {

// What the epilogue parks for the gradient GEMM: bf16(P'), the remainder P' - bf16(P') (third pass of the three-pass
// split-operand form), or both tiles at once (single-pass split operands, mmd_tc.cu)
enum : int { kPHi = 0, kPLo = 1, kPBoth = 2 };

// Kernel values of 32 columns already in registers (the math is spelled out below the function).
// Ladder form, nested so that t^16 is never formed (3 squarings, 16 packed ops per column pair instead of 17):
//   k = t (1 + t (1 + t^2 (1 + t^4 (1 + t^8))))                    = t + t^2 + t^4 + t^8 + t^16
//   w = t (c4 + t (c3 + t^2 (c2 + t^4 (c1 + t^8 c0)))) - c_j w0    = sum_m c_m t^(2^(4-m)) - c_j w0
// (scale index 4 is the widest kernel, t itself; index 0 the narrowest, t^16)
// ONES (VGAN_B200_ONES): the Gram already carries -|z_i|^2 / 2 (three bf16 pieces in the ones columns of the resident row
// operand), so u = fma(S, -2 c0, nj) without the row norm, and the row sum of P comes out of the ones column of O:
// 14 packed ops per column pair.
template <bool LADDER, int PMODE, bool ONES = false>
__device__ __forceinline__ void epi_core_32(const uint32_t (&v)[32], const float* __restrict__ njp, float ni,
                                            float m2c0, const float (&cf)[VGAN_B200_MAX_SCALES],
                                            const float (&ratio)[VGAN_B200_MAX_SCALES], int n_scales,
                                            float& ksum, float& rsum, uint4 (&pk)[4], float cw0, uint4* pk_lo) {
  // cw0 = c_j w0: the common-mode level of this (row block, column block) pair.  P' = c_j (w - w0) is what gets rounded
  // to bf16 and multiplied; the closed-form remainder is added back in write_grad_rows.  It rides in the last FMA of
  // the weight sum, so it costs no instruction.
  if constexpr (LADDER) {
    // two columns per issue slot (packed f32x2)
    const uint64_t m2 = f2_pack(m2c0, m2c0), ni2 = f2_pack(ni, ni), mcw = f2_pack(-cw0, -cw0), one = f2_pack(1.f, 1.f);
    const uint64_t c0 = f2_pack(cf[0], cf[0]), c1 = f2_pack(cf[1], cf[1]), c2 = f2_pack(cf[2], cf[2]),
                   c3 = f2_pack(cf[3], cf[3]), c4 = f2_pack(cf[4], cf[4]);
    uint64_t ks2 = f2_pack(ksum, 0.f), rs2 = f2_pack(rsum, 0.f);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const float4 nja = *reinterpret_cast<const float4*>(njp + c * 8);
      const float4 njb = *reinterpret_cast<const float4*>(njp + c * 8 + 4);
      const uint64_t nj2[4] = {f2_pack(nja.x, nja.y), f2_pack(nja.z, nja.w), f2_pack(njb.x, njb.y), f2_pack(njb.z, njb.w)};
      uint32_t pw[4], pl[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint64_t s2 = f2_pack(__uint_as_float(v[c * 8 + 2 * q]), __uint_as_float(v[c * 8 + 2 * q + 1]));
        float u0, u1;
        f2_unpack(ONES ? f2_fma(s2, m2, nj2[q]) : f2_fma(s2, m2, f2_add(ni2, nj2[q])), u0, u1);
        // no clamp at 0 here (the reference's cdist clamps): the norms are those of the values the tensor core
        // multiplies, so u of a coincident pair is 0 +- 1e-6 and t = 1 +- 1e-6 — far inside the 2e-3 / 1e-4 budgets,
        // and two FMNMX per column pair are saved; padding has u = +huge
        const uint64_t t1 = f2_pack(ex2_approx(-u0), ex2_approx(-u1));
        const uint64_t t2 = f2_mul(t1, t1), t4 = f2_mul(t2, t2), t8 = f2_mul(t4, t4);
        ks2 = f2_fma(t1, f2_fma(t1, f2_fma(t2, f2_fma(t4, f2_add(t8, one), one), one), one), ks2);
        const uint64_t w2 = f2_fma(t1, f2_fma(t1, f2_fma(t2, f2_fma(t4, f2_fma(t8, c0, c1), c2), c3), c4), mcw);
        if constexpr (!ONES) rs2 = f2_add(rs2, w2);
        float w0, w1;
        f2_unpack(w2, w0, w1);
        pw[q] = pack_bf16x2(w0, w1);
        if constexpr (PMODE != kPHi) {
          pl[q] = pack_bf16x2(w0 - __uint_as_float(pw[q] << 16), w1 - __uint_as_float(pw[q] & 0xffff0000u));
          if constexpr (PMODE == kPLo) pw[q] = pl[q];      // split-operand pass 3: the remainder of P after its bf16 rounding
        }
      }
      pk[c].x = pw[0]; pk[c].y = pw[1]; pk[c].z = pw[2]; pk[c].w = pw[3];
      if constexpr (PMODE == kPBoth) { pk_lo[c].x = pl[0]; pk_lo[c].y = pl[1]; pk_lo[c].z = pl[2]; pk_lo[c].w = pl[3]; }
    }
    float a, b;
    f2_unpack(ks2, a, b);
    ksum = a + b;
    f2_unpack(rs2, a, b);
    rsum = a + b;
  } else {
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const float4 nja = *reinterpret_cast<const float4*>(njp + c * 8);
    const float4 njb = *reinterpret_cast<const float4*>(njp + c * 8 + 4);
    const float njv[8] = {nja.x, nja.y, nja.z, nja.w, njb.x, njb.y, njb.z, njb.w};
    float w[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      float uarg = fmaf(__uint_as_float(v[c * 8 + e]), m2c0, ONES ? njv[e] : ni + njv[e]);
      uarg = fmaxf(uarg, 0.f);
      float kk = 0.f, ww = -cw0;
      for (int m = 0; m < n_scales; ++m) {
        const float em = ex2_approx(-uarg * ratio[m]);
        kk += em;
        ww = fmaf(em, cf[m], ww);
      }
      ksum += kk;
      w[e] = ww;
      rsum += w[e];
    }
    uint32_t pw[4], pl[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      pw[q] = pack_bf16x2(w[2 * q], w[2 * q + 1]);
      pl[q] = pack_bf16x2(w[2 * q] - __uint_as_float(pw[q] << 16), w[2 * q + 1] - __uint_as_float(pw[q] & 0xffff0000u));
      if constexpr (PMODE == kPLo) pw[q] = pl[q];
    }
    pk[c].x = pw[0]; pk[c].y = pw[1]; pk[c].z = pw[2]; pk[c].w = pw[3];
    if constexpr (PMODE == kPBoth) { pk_lo[c].x = pl[0]; pk_lo[c].y = pl[1]; pk_lo[c].z = pl[2]; pk_lo[c].w = pl[3]; }
  }
  }
}

// The same with the parked tile chosen at run time (pair kernel: plo = third pass of the three-pass split form)
template <bool LADDER>
__device__ __forceinline__ void epi_compute_32(const uint32_t (&v)[32], const float* __restrict__ njp, float ni,
                                               float m2c0, const float (&cf)[VGAN_B200_MAX_SCALES],
                                               const float (&ratio)[VGAN_B200_MAX_SCALES], int n_scales,
                                               float& ksum, float& rsum, uint4 (&pk)[4], bool plo = false,
                                               float cw0 = 0.f) {
  if (plo) epi_core_32<LADDER, kPLo>(v, njp, ni, m2c0, cf, ratio, n_scales, ksum, rsum, pk, cw0, nullptr);
  else epi_core_32<LADDER, kPHi>(v, njp, ni, m2c0, cf, ratio, n_scales, ksum, rsum, pk, cw0, nullptr);
}

// Sum over 32 columns already in registers of k alone (units without a gradient GEMM): 10 packed ops per column pair
template <bool LADDER, bool ONES = false>
__device__ __forceinline__ void epi_k_32(const uint32_t (&v)[32], const float* __restrict__ njp, float ni, float m2c0,
                                         const float (&ratio)[VGAN_B200_MAX_SCALES], int n_scales, float& ksum) {
  if constexpr (LADDER) {
    const uint64_t m2 = f2_pack(m2c0, m2c0), ni2 = f2_pack(ni, ni), one = f2_pack(1.f, 1.f);
    uint64_t ks2 = f2_pack(ksum, 0.f);
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const float4 nj4 = *reinterpret_cast<const float4*>(njp + c * 4);
      const uint64_t nj2[2] = {f2_pack(nj4.x, nj4.y), f2_pack(nj4.z, nj4.w)};
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const uint64_t s2 = f2_pack(__uint_as_float(v[c * 4 + 2 * q]), __uint_as_float(v[c * 4 + 2 * q + 1]));
        float u0, u1;
        f2_unpack(ONES ? f2_fma(s2, m2, nj2[q]) : f2_fma(s2, m2, f2_add(ni2, nj2[q])), u0, u1);
        const uint64_t t1 = f2_pack(ex2_approx(-u0), ex2_approx(-u1));
        const uint64_t t2 = f2_mul(t1, t1), t4 = f2_mul(t2, t2), t8 = f2_mul(t4, t4);
        ks2 = f2_fma(t1, f2_fma(t1, f2_fma(t2, f2_fma(t4, f2_add(t8, one), one), one), one), ks2);
      }
    }
    float a, b;
    f2_unpack(ks2, a, b);
    ksum = a + b;
  } else {
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const float4 nj4 = *reinterpret_cast<const float4*>(njp + c * 4);
      const float njv[4] = {nj4.x, nj4.y, nj4.z, nj4.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float uarg = fmaf(__uint_as_float(v[c * 4 + e]), m2c0, ONES ? njv[e] : ni + njv[e]);
        uarg = fmaxf(uarg, 0.f);
        for (int m = 0; m < n_scales; ++m) ksum += ex2_approx(-uarg * ratio[m]);
      }
    }
  }
}

// The math of epi_core_32 (one thread = one row of the tile, 32 columns of S already in registers):
//   u = c0*d2 = fma(S, -2c0, ni + nj),  t = 2^-u,  k = sum_m t^(2^q),  w = sum_m coef_m t^(2^q)
// (ladder) or the general n_scales form; k is added to ksum, w to rsum, and the 32 values of P = w come back as bf16
// packed into 4 x uint4 (8 columns each, ascending).

// L2 prefetch of what write_grad_rows will read from global memory (the fp32 input rows of the unit's row tile, its
// slice of the features): 256 threads, thread = (row, half of the slice's 128-byte lines).  Issued when the epilogue has
// finished its last tile, so the lines arrive while the last gradient GEMMs drain; read cold under the full L2 / HBM
// load of the other CTAs the gradient rows cost 36 000 - 53 000 cycles per work range in the pair kernel (a handful of
// dependent DRAM round trips of several thousand cycles each, tools/gpu_cluster_timers.py).
__device__ __forceinline__ void prefetch_grad_rows(const TcParams& p, int ti, int row_is_x, int slice, int slice_chunks) {
  const int t = threadIdx.x & 255;
  const int r = t & (kTile - 1), half = t >> 7;
  const int64_t region_row = static_cast<int64_t>(row_is_x ? ti : ti - p.tiles_x) * kTile + r;
  const int64_t n_region = row_is_x ? p.Nx : p.Ny;
  if (!(region_row < n_region && region_row >= p.row_begin && region_row < p.row_end)) return;
  const int f_slice = slice * p.slice_chunks * kChunk;
  const int lines = slice_chunks * kChunk / 32;          // 128-byte lines of the slice per row
  const float* a0 = row_is_x ? p.Xf : (p.Yf ? p.Yf : p.Uf);
  const float* a1 = (!row_is_x && !p.Yf) ? p.Xf : nullptr;                  // masked rows: U and X
  const float* a2 = (!row_is_x && p.dY_mul && p.dY_mul != a1 + p.row_begin * p.D) ? p.dY_mul - p.row_begin * p.D : nullptr;
  for (int l = half; l < lines; l += 2) {
    const int64_t f = f_slice + l * 32;
    if (f >= p.D) break;
    const int64_t off = region_row * p.D + f;
    asm volatile("prefetch.global.L2 [%0];" ::"l"(a0 + off));
    if (a1) asm volatile("prefetch.global.L2 [%0];" ::"l"(a1 + off));
    if (a2) asm volatile("prefetch.global.L2 [%0];" ::"l"(a2 + off));
  }
}

// dZ rows of one unit: out[row, f] = 4 c_i (O'[row, f] + A[f] - (rtot' + B) z[row, f]) (x mul), f in the slice, where
// O' and rtot' come from P' = c_j (w_ij - w0) and A, B (scales_kernel) are the closed-form common-mode remainder
// sum_j c_j w0 (z_j - z_i).  The expression is linear in (O', rtot'): a unit cut into column ranges adds the ranges'
// results (accumulate), and only the range that starts at column round 0 (`first_range`) adds A and B.
// Column sums of the Y rows written go to p.colsum (the D-vector of the step: colsum_i dL/dU[i,:] on the masked path).
//
// Called by 256 threads = warps 0-7 of the CTA (warp % 4 = TMEM lane quadrant), with their own named barrier.
// TMEM hands a thread ONE ROW of O; read that way from global memory (round 1: a thread walked its row of X, U, the
// multiplier and the output in 16-byte steps) a warp touches 32 rows per instruction — the gradient rows of one work
// range took 68 000 cycles in the pair kernel, 37 % of the kernel at N = 4096 (tools/gpu_cluster_timers.py).  Now O
// goes through shared memory, 128 columns at a time (`stage`: [128 rows][128 columns] fp32, 16-byte chunks XOR-swizzled
// by the row so that both the row-wise writes and the row-contiguous reads are conflict-free), and a warp then handles
// one row at a time with its lanes along the features: every global access is a contiguous 512-byte segment, the
// per-feature constants stay in registers, and the column sums accumulate per lane (no warp transposes).
//   stage: 64 KiB, aux: (1 + 8) x 128 floats; both dead operand memory of the caller.  rtot: row sum of P' of the
//   calling thread's row (row = (warp % 4) * 32 + lane).
__device__ __forceinline__ void write_grad_rows(const TcParams& p, uint32_t tmem_o, float* stage, float* aux, int ti,
                                                int row_is_x, int slice, int slice_chunks, float rtot,
                                                bool accumulate, bool first_range) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = warp & 3, g = warp >> 2;
  const int row = q * 32 + lane;
  const uint32_t lane_addr = static_cast<uint32_t>(q * 32) << 16;
  const int slice_w = slice_chunks * kChunk;
  const int64_t tile_row0 = static_cast<int64_t>(row_is_x ? ti : ti - p.tiles_x) * kTile;
  const int64_t n_region = row_is_x ? p.Nx : p.Ny;
  const float ci4 = 4.f * (row_is_x ? 1.f / static_cast<float>(p.Nx) : -1.f / static_cast<float>(p.Ny));
  float* out_base = row_is_x ? p.dX : p.dY;
  const float* mul_base = (!row_is_x && p.dY_mul) ? p.dY_mul : nullptr;
  const bool cmode = first_range && !(p.pass_flags & kPassNoRowsum);
  const float* cmv = p.cm + (row_is_x ? 0 : p.Dp);
  accumulate = accumulate || (p.pass_flags & kPassAccumulate);
  const int f_slice = slice * p.slice_chunks * kChunk;
  // 16-byte accesses: rows are 16-byte aligned when D % 4 == 0 and the bases are
  const bool vec = (p.D & 3) == 0 && (reinterpret_cast<uintptr_t>(p.Xf) & 15) == 0 &&
                   (!p.Yf || (reinterpret_cast<uintptr_t>(p.Yf) & 15) == 0) &&
                   (!p.Uf || (reinterpret_cast<uintptr_t>(p.Uf) & 15) == 0) &&
                   (reinterpret_cast<uintptr_t>(out_base) & 15) == 0 &&
                   (!mul_base || (reinterpret_cast<uintptr_t>(mul_base) & 15) == 0);
  const bool want_colsum = p.colsum != nullptr && !row_is_x;
  float* rt = aux;                    // [128] row totals (with B on the first range)
  float* csp = aux + kTile;           // [8 warps][128] column-sum partials
  if (g == 0) {
    float t = (p.pass_flags & kPassNoRowsum) ? 0.f : rtot;
    if (cmode) t += p.par[kParB + (row_is_x ? 0 : 1)];
    rt[row] = t;
  }
  float4* stage4 = reinterpret_cast<float4*>(stage);
#if defined(VGAN_B200_DEBUG) && defined(VGAN_B200_TIMERS)
#define VG_WG_CLK() clock64()
  long long wg_a = 0, wg_b = 0, wg_c = 0;
#else
#define VG_WG_CLK() 0ll
#endif
  for (int c0 = 0; c0 < slice_w; c0 += 128) {
    const long long wg_t0 = VG_WG_CLK();
    // ---- phase A: this thread's row of O, columns c0 + 64 g + [0, 64), TMEM -> swizzled shared memory
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int col = c0 + g * 64 + h * 32;
      if (col < slice_w) {              // (warp-uniform)
        uint32_t v[32];
        tmem_ld_32x32(tmem_o + lane_addr + col, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int chunk = g * 16 + h * 8 + j;
          stage4[row * 32 + (chunk ^ (row & 31))] =
              make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]),
                          __uint_as_float(v[4 * j + 3]));
        }
      }
    }
    asm volatile("bar.sync 2, 256;" ::: "memory");
    const long long wg_t1 = VG_WG_CLK();
    // ---- phase B: warp w takes rows w, w + 8, ...; lane l the features f0 .. f0 + 3 of every row
    const int f0 = f_slice + c0 + 4 * lane;
    const bool lane_on = c0 + 4 * lane < slice_w && f0 < p.D;
    const bool lane_vec = lane_on && vec && f0 + 4 <= p.D;
    float mean4[4] = {0.f, 0.f, 0.f, 0.f}, av4[4] = {0.f, 0.f, 0.f, 0.f}, cs[4] = {0.f, 0.f, 0.f, 0.f};
    if (lane_on) {
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if (f0 + e < p.D) {
          mean4[e] = p.mean[f0 + e];
          if (cmode) av4[e] = cmv[f0 + e];
        }
    }
    if (lane_vec) {
      // four rows per step, every global load of the step issued before its first use (a row at a time the loop ran
      // at one memory latency per row).  Registers decide whether the loads really go out together: the multiplier of
      // the masked path IS the X row (dL/dU = dL/dY * X), so it is not loaded twice, and O comes from shared memory
      // only when it is used.
      const float* pa = row_is_x ? p.Xf : (p.Yf ? p.Yf : p.Xf);            // primary input row
      const float* pb = (!row_is_x && !p.Yf) ? p.Uf : nullptr;               // masked rows: z = U * X - mean
      const bool mul_is_pa = mul_base && pb && mul_base == p.Xf + p.row_begin * p.D;
      const float* pm = (mul_base && !mul_is_pa) ? mul_base - p.row_begin * p.D : nullptr;
      for (int rr = 0; rr < kTile / 8; rr += 4) {
        float4 va[4], vb[4], vm[4];
        bool ok[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int64_t region_row = tile_row0 + (rr + k) * 8 + warp;
          // (rows outside the range are never dereferenced: the callers of a multi-GPU job pass pointers that are
          // only valid for their own rows)
          ok[k] = region_row < n_region && region_row >= p.row_begin && region_row < p.row_end;   // warp-uniform
          va[k] = make_float4(0.f, 0.f, 0.f, 0.f);
          vb[k] = vm[k] = make_float4(1.f, 1.f, 1.f, 1.f);
          if (ok[k]) {
            const int64_t off = region_row * p.D + f0;
            va[k] = __ldg(reinterpret_cast<const float4*>(pa + off));
            if (pb) vb[k] = __ldg(reinterpret_cast<const float4*>(pb + off));
            if (pm) vm[k] = __ldg(reinterpret_cast<const float4*>(pm + off));
          }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (!ok[k]) continue;
          const int r = (rr + k) * 8 + warp;
          const float4 o = stage4[r * 32 + (lane ^ (r & 31))];
          const float rtr = rt[r];
          const float4 m = mul_is_pa ? va[k] : vm[k];
          // unrounded z_i = input row - centring mean
          float4 gv;
          gv.x = ci4 * ((o.x + av4[0]) - rtr * (va[k].x * vb[k].x - mean4[0])) * m.x;
          gv.y = ci4 * ((o.y + av4[1]) - rtr * (va[k].y * vb[k].y - mean4[1])) * m.y;
          gv.z = ci4 * ((o.z + av4[2]) - rtr * (va[k].z * vb[k].z - mean4[2])) * m.z;
          gv.w = ci4 * ((o.w + av4[3]) - rtr * (va[k].w * vb[k].w - mean4[3])) * m.w;
          cs[0] += gv.x; cs[1] += gv.y; cs[2] += gv.z; cs[3] += gv.w;
          float4* out = reinterpret_cast<float4*>(out_base + (tile_row0 + r - p.row_begin) * p.D + f0);
          if (accumulate) atomicAdd(out, gv);     // unit shared between CTAs (column ranges / stream-K cut) or a BF16X3 pass
          else *out = gv;
        }
      }
    } else if (lane_on) {
      // D % 4 != 0 (or misaligned bases): the same four rows per step with 4-byte accesses (rows are not 16-byte aligned)
      const float* pa = row_is_x ? p.Xf : (p.Yf ? p.Yf : p.Xf);
      const float* pb = (!row_is_x && !p.Yf) ? p.Uf : nullptr;
      const bool mul_is_pa = mul_base && pb && mul_base == p.Xf + p.row_begin * p.D;
      const float* pm = (mul_base && !mul_is_pa) ? mul_base - p.row_begin * p.D : nullptr;
      const int ne = min(4, static_cast<int>(p.D) - f0);          // valid features of this lane (>= 1)
      for (int rr = 0; rr < kTile / 8; rr += 4) {
        float va[4][4], vb[4][4], vm[4][4];
        bool ok[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int64_t region_row = tile_row0 + (rr + k) * 8 + warp;
          ok[k] = region_row < n_region && region_row >= p.row_begin && region_row < p.row_end;   // warp-uniform
          const int64_t off = region_row * p.D + f0;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            va[k][e] = 0.f; vb[k][e] = 1.f; vm[k][e] = 1.f;
            if (ok[k] && e < ne) {
              va[k][e] = __ldg(pa + off + e);
              if (pb) vb[k][e] = __ldg(pb + off + e);
              if (pm) vm[k][e] = __ldg(pm + off + e);
            }
          }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (!ok[k]) continue;
          const int r = (rr + k) * 8 + warp;
          const float4 o = stage4[r * 32 + (lane ^ (r & 31))];
          const float ov[4] = {o.x, o.y, o.z, o.w};
          const float rtr = rt[r];
          float* out = out_base + (tile_row0 + r - p.row_begin) * p.D + f0;
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (e < ne) {
              const float gval = ci4 * ((ov[e] + av4[e]) - rtr * (va[k][e] * vb[k][e] - mean4[e])) *
                                 (mul_is_pa ? va[k][e] : vm[k][e]);
              cs[e] += gval;
              if (accumulate) atomicAdd(out + e, gval);
              else out[e] = gval;
            }
        }
      }
    }
    const long long wg_t2 = VG_WG_CLK();
    if (want_colsum) {      // (CTA-uniform)
#pragma unroll
      for (int e = 0; e < 4; ++e) csp[warp * kTile + 4 * lane + e] = cs[e];
      asm volatile("bar.sync 2, 256;" ::: "memory");
      if (threadIdx.x < kTile) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += csp[w * kTile + threadIdx.x];
        const int f = f_slice + c0 + static_cast<int>(threadIdx.x);
        if (c0 + static_cast<int>(threadIdx.x) < slice_w && f < p.D) atomicAdd(p.colsum + f, static_cast<double>(t));
      }
    }
    asm volatile("bar.sync 2, 256;" ::: "memory");      // the stage (and the partials) may be rewritten
#if defined(VGAN_B200_DEBUG) && defined(VGAN_B200_TIMERS)
    wg_a += wg_t1 - wg_t0; wg_b += wg_t2 - wg_t1; wg_c += clock64() - wg_t2;
#else
    (void)wg_t0; (void)wg_t1; (void)wg_t2;
#endif
  }
#if defined(VGAN_B200_DEBUG) && defined(VGAN_B200_TIMERS)
  if (p.dbg && threadIdx.x == 0) {
    float* d = p.dbg + static_cast<int64_t>(blockIdx.x) * p.dbg_stride;
    d[20] += static_cast<float>(wg_a); d[21] += static_cast<float>(wg_b); d[22] += static_cast<float>(wg_c);
  }
#endif
}

}  // namespace vg
