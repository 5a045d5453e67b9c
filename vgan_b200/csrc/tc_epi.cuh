// Shared epilogue math of the tcgen05 K1 kernels: 32 TMEM columns of S -> kernel values.
#pragma once
#include "common.cuh"
#include "ptx.cuh"

namespace vg {

// What the epilogue parks for the gradient GEMM: bf16(P'), the remainder P' - bf16(P') (third pass of the three-pass
// split-operand form), or both tiles at once (single-pass split operands, mmd_tc.cu)
enum : int { kPHi = 0, kPLo = 1, kPBoth = 2 };

// Kernel values of 32 columns already in registers (see epi_half_tile for the math).
// Ladder form, nested so that t^16 is never formed (3 squarings, 16 packed ops per column pair instead of 17):
//   k = t (1 + t (1 + t^2 (1 + t^4 (1 + t^8))))                    = t + t^2 + t^4 + t^8 + t^16
//   w = t (c4 + t (c3 + t^2 (c2 + t^4 (c1 + t^8 c0)))) - c_j w0    = sum_m c_m t^(2^(4-m)) - c_j w0
// (scale index 4 is the widest kernel, t itself; index 0 the narrowest, t^16)
template <bool LADDER, int PMODE>
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
        f2_unpack(f2_fma(s2, m2, f2_add(ni2, nj2[q])), u0, u1);
        // no clamp at 0 here (the reference's cdist clamps): the norms are those of the values the tensor core
        // multiplies, so u of a coincident pair is 0 +- 1e-6 and t = 1 +- 1e-6 — far inside the 2e-3 / 1e-4 budgets,
        // and two FMNMX per column pair are saved; padding has u = +huge
        const uint64_t t1 = f2_pack(ex2_approx(-u0), ex2_approx(-u1));
        const uint64_t t2 = f2_mul(t1, t1), t4 = f2_mul(t2, t2), t8 = f2_mul(t4, t4);
        ks2 = f2_fma(t1, f2_fma(t1, f2_fma(t2, f2_fma(t4, f2_add(t8, one), one), one), one), ks2);
        const uint64_t w2 = f2_fma(t1, f2_fma(t1, f2_fma(t2, f2_fma(t4, f2_fma(t8, c0, c1), c2), c3), c4), mcw);
        rs2 = f2_add(rs2, w2);
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
      float uarg = fmaf(__uint_as_float(v[c * 8 + e]), m2c0, ni + njv[e]);
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
template <bool LADDER>
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
        f2_unpack(f2_fma(s2, m2, f2_add(ni2, nj2[q])), u0, u1);
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
        float uarg = fmaf(__uint_as_float(v[c * 4 + e]), m2c0, ni + njv[e]);
        uarg = fmaxf(uarg, 0.f);
        for (int m = 0; m < n_scales; ++m) ksum += ex2_approx(-uarg * ratio[m]);
      }
    }
  }
}

// One thread = one row of the tile.  Reads S[row, col0 .. col0+31] from TMEM, forms
//   u = c0*d2 = fma(S, -2c0, ni + nj),  t = 2^-u,  k = sum_m t^(2^q),  w = sum_m coef_m t^(2^q)
// (ladder) or the general n_scales form, adds k to ksum and w to rsum, and returns the 32 values of
// P = w as bf16 packed into 4 x uint4 (8 columns each, ascending).
template <bool LADDER, int PMODE>
__device__ __forceinline__ void epi_half_tile(uint32_t taddr, const float* __restrict__ njp, float ni,
                                              float m2c0, const float (&cf)[VGAN_B200_MAX_SCALES],
                                              const float (&ratio)[VGAN_B200_MAX_SCALES], int n_scales,
                                              float& ksum, float& rsum, uint4 (&pk)[4], float cw0, uint4* pk_lo) {
  uint32_t v[32];
  tmem_ld_32x32(taddr, v);
  tmem_ld_wait();
  epi_core_32<LADDER, PMODE>(v, njp, ni, m2c0, cf, ratio, n_scales, ksum, rsum, pk, cw0, pk_lo);
}

// Column sums over the 32 lanes of a warp of 32 per-lane values: lane l returns sum over lanes of v[l].
// Butterfly: each step halves the values a lane holds (16 + 8 + 4 + 2 + 1 = 31 shuffles).
__device__ __forceinline__ float warp_transpose_sum(float (&v)[32], int lane) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    const bool up = (lane & s) != 0;
#pragma unroll
    for (int i = 0; i < s; ++i) {
      const float send = up ? v[i] : v[i + s];
      const float keep = up ? v[i + s] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
  return v[0];
}

// dZ rows of one unit: out[row, f] = 4 c_i (O'[row, f] + A[f] - (rtot' + B) z[row, f]) (x mul), f in the slice, where
// O' and rtot' come from P' = c_j (w_ij - w0) and A, B (scales_kernel) are the closed-form common-mode remainder
// sum_j c_j w0 (z_j - z_i).  The expression is linear in (O', rtot'): a unit cut into column ranges adds the ranges'
// results (accumulate), and only the range that starts at column round 0 (`first_range`) adds A and B.
// Thread = (row, column half g); O columns of the slice live at tmem_o.  Column sums of the Y rows written go to
// p.colsum (the D-vector of the step: colsum_i dL/dU[i,:] on the masked path).
__device__ __forceinline__ void write_grad_rows(const TcParams& p, uint32_t tmem_o, uint32_t lane_addr, int g,
                                                int row, int ti, int row_is_x, int slice, int slice_chunks,
                                                float rtot, bool accumulate, bool first_range) {
  const int lane = threadIdx.x & 31;
  const int slice_w = slice_chunks * kChunk;
  const int64_t region_row = static_cast<int64_t>(row_is_x ? ti : ti - p.tiles_x) * kTile + row;
  const int64_t n_region = row_is_x ? p.Nx : p.Ny;
  const bool row_ok = region_row < n_region && region_row >= p.row_begin && region_row < p.row_end;
  const float ci4 = 4.f * (row_is_x ? 1.f / static_cast<float>(p.Nx) : -1.f / static_cast<float>(p.Ny));
  float* out = (row_is_x ? p.dX : p.dY) + (region_row - p.row_begin) * p.D;
  const float* mul = (!row_is_x && p.dY_mul) ? p.dY_mul + (region_row - p.row_begin) * p.D : nullptr;
  // unrounded z_i = input row - centring mean (rows outside the range are never dereferenced: the callers of a
  // multi-GPU job pass pointers that are only valid for their own rows)
  const float* xrow = p.Xf + region_row * p.D;
  const float* yrow = p.Yf ? p.Yf + region_row * p.D : nullptr;
  const float* urow = p.Uf ? p.Uf + region_row * p.D : nullptr;
  const bool zvec = (p.D & 3) == 0 && (reinterpret_cast<uintptr_t>(p.Xf) & 15) == 0 &&
                    (!p.Yf || (reinterpret_cast<uintptr_t>(p.Yf) & 15) == 0) &&
                    (!p.Uf || (reinterpret_cast<uintptr_t>(p.Uf) & 15) == 0);
  if (p.pass_flags & kPassNoRowsum) rtot = 0.f;
  const bool cmode = first_range && !(p.pass_flags & kPassNoRowsum);
  if (cmode) rtot += p.par[kParB + (row_is_x ? 0 : 1)];
  const float* cmv = p.cm + (row_is_x ? 0 : p.Dp);
  accumulate = accumulate || (p.pass_flags & kPassAccumulate);
  const int f_slice = slice * p.slice_chunks * kChunk;
  const bool vec_ok = (p.D & 3) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0 &&
                      (!mul || (reinterpret_cast<uintptr_t>(mul) & 15) == 0);
  const bool want_colsum = p.colsum != nullptr && !row_is_x;       // warp-uniform
  for (int cc = 0; cc < slice_chunks; ++cc) {
    const int col = g * (slice_w / 2) + cc * 32;
    uint32_t v[32];
    tmem_ld_32x32(tmem_o + lane_addr + col, v);
    tmem_ld_wait();
    float cs[32];
#pragma unroll
    for (int e = 0; e < 32; ++e) cs[e] = 0.f;
    if (row_ok) {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int f0 = f_slice + col + c * 8;
        float zv[8];
        if (zvec && f0 + 8 <= p.D) {
          float4 a0, a1;
          if (row_is_x) {
            a0 = *reinterpret_cast<const float4*>(xrow + f0); a1 = *reinterpret_cast<const float4*>(xrow + f0 + 4);
          } else if (yrow) {
            a0 = *reinterpret_cast<const float4*>(yrow + f0); a1 = *reinterpret_cast<const float4*>(yrow + f0 + 4);
          } else {
            const float4 x0 = *reinterpret_cast<const float4*>(xrow + f0), x1 = *reinterpret_cast<const float4*>(xrow + f0 + 4);
            const float4 u0 = *reinterpret_cast<const float4*>(urow + f0), u1 = *reinterpret_cast<const float4*>(urow + f0 + 4);
            a0 = make_float4(u0.x * x0.x, u0.y * x0.y, u0.z * x0.z, u0.w * x0.w);
            a1 = make_float4(u1.x * x1.x, u1.y * x1.y, u1.z * x1.z, u1.w * x1.w);
          }
          const float4 m0 = *reinterpret_cast<const float4*>(p.mean + f0), m1 = *reinterpret_cast<const float4*>(p.mean + f0 + 4);
          zv[0] = a0.x - m0.x; zv[1] = a0.y - m0.y; zv[2] = a0.z - m0.z; zv[3] = a0.w - m0.w;
          zv[4] = a1.x - m1.x; zv[5] = a1.y - m1.y; zv[6] = a1.z - m1.z; zv[7] = a1.w - m1.w;
        } else {
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const int f = f0 + e;
            float raw = 0.f, mu = 0.f;
            if (f < p.D) {
              raw = row_is_x ? xrow[f] : (yrow ? yrow[f] : urow[f] * xrow[f]);
              mu = p.mean[f];
            }
            zv[e] = raw - mu;
          }
        }
        float av[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (cmode) {
          const float4 a0 = *reinterpret_cast<const float4*>(cmv + f0), a1 = *reinterpret_cast<const float4*>(cmv + f0 + 4);
          av[0] = a0.x; av[1] = a0.y; av[2] = a0.z; av[3] = a0.w; av[4] = a1.x; av[5] = a1.y; av[6] = a1.z; av[7] = a1.w;
        }
        float gv[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          gv[e] = ci4 * ((__uint_as_float(v[c * 8 + e]) + av[e]) - rtot * zv[e]);
        }
        if (vec_ok && f0 + 8 <= p.D) {
          // 16-byte accesses (rows are 16-byte aligned when D % 4 == 0)
          if (mul) {
            const float4 m0 = *reinterpret_cast<const float4*>(mul + f0), m1 = *reinterpret_cast<const float4*>(mul + f0 + 4);
            gv[0] *= m0.x; gv[1] *= m0.y; gv[2] *= m0.z; gv[3] *= m0.w;
            gv[4] *= m1.x; gv[5] *= m1.y; gv[6] *= m1.z; gv[7] *= m1.w;
          }
          if (accumulate) {     // unit shared between CTAs (column ranges / stream-K cut) or a BF16X3 pass: red.v4.f32
            atomicAdd(reinterpret_cast<float4*>(out + f0), make_float4(gv[0], gv[1], gv[2], gv[3]));
            atomicAdd(reinterpret_cast<float4*>(out + f0 + 4), make_float4(gv[4], gv[5], gv[6], gv[7]));
          } else {
            *reinterpret_cast<float4*>(out + f0) = make_float4(gv[0], gv[1], gv[2], gv[3]);
            *reinterpret_cast<float4*>(out + f0 + 4) = make_float4(gv[4], gv[5], gv[6], gv[7]);
          }
#pragma unroll
          for (int e = 0; e < 8; ++e) cs[c * 8 + e] = gv[e];
        } else {
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const int f = f0 + e;
            if (f < p.D) {
              float gval = gv[e];
              if (mul) gval *= mul[f];
              if (accumulate) atomicAdd(out + f, gval);   // unit shared between clusters (stream-K cut) / BF16X3 pass
              else out[f] = gval;
              cs[c * 8 + e] = gval;
            }
          }
        }
      }
    }
    if (want_colsum) {
      const float tot = warp_transpose_sum(cs, lane);
      const int f = f_slice + col + lane;
      if (f < p.D) atomicAdd(p.colsum + f, static_cast<double>(tot));
    }
  }
}

}  // namespace vg
