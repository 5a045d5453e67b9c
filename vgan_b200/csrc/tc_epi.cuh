// Shared epilogue math of the tcgen05 K1 kernels: 32 TMEM columns of S -> kernel values.
#pragma once
#include "common.cuh"
#include "ptx.cuh"

namespace vg {

// Kernel values of 32 columns already in registers (see epi_half_tile for the math).
template <bool LADDER>
__device__ __forceinline__ void epi_compute_32(const uint32_t (&v)[32], const float* __restrict__ njp, float ni,
                                                    float m2c0, const float (&cf)[VGAN_B200_MAX_SCALES],
                                                    const float (&ratio)[VGAN_B200_MAX_SCALES], int n_scales,
                                                    float& ksum, float& rsum, uint4 (&pk)[4], bool plo = false) {
  if constexpr (LADDER) {
    // two columns per issue slot (packed f32x2): ~11.5 instead of ~19.5 instructions per pair
    const uint64_t m2 = f2_pack(m2c0, m2c0), ni2 = f2_pack(ni, ni);
    const uint64_t c0 = f2_pack(cf[0], cf[0]), c1 = f2_pack(cf[1], cf[1]), c2 = f2_pack(cf[2], cf[2]),
                   c3 = f2_pack(cf[3], cf[3]), c4 = f2_pack(cf[4], cf[4]);
    uint64_t ks2 = f2_pack(0.f, 0.f), rs2 = f2_pack(0.f, 0.f);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const float4 nja = *reinterpret_cast<const float4*>(njp + c * 8);
      const float4 njb = *reinterpret_cast<const float4*>(njp + c * 8 + 4);
      const uint64_t nj2[4] = {f2_pack(nja.x, nja.y), f2_pack(nja.z, nja.w), f2_pack(njb.x, njb.y), f2_pack(njb.z, njb.w)};
      uint32_t pw[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint64_t s2 = f2_pack(__uint_as_float(v[c * 8 + 2 * q]), __uint_as_float(v[c * 8 + 2 * q + 1]));
        float u0, u1;
        f2_unpack(f2_fma(s2, m2, f2_add(ni2, nj2[q])), u0, u1);
        // no clamp at 0 here (the reference's cdist clamps): the norms are those of the values the tensor core
        // multiplies, so u of a coincident pair is 0 +- 1e-6 and t = 1 +- 1e-6 — far inside the 2e-3 / 1e-4 budgets,
        // and two FMNMX per column pair (8 % of the epilogue's instructions) are saved; padding has u = +huge
        const uint64_t t1 = f2_pack(ex2_approx(-u0), ex2_approx(-u1));
        const uint64_t t2 = f2_mul(t1, t1), t4 = f2_mul(t2, t2), t8 = f2_mul(t4, t4), t16 = f2_mul(t8, t8);
        ks2 = f2_add(ks2, f2_add(f2_add(t1, t2), f2_add(f2_add(t4, t8), t16)));
        // scale index 4 is the widest kernel (t1), index 0 the narrowest (t16)
        const uint64_t w2 = f2_fma(t16, c0, f2_fma(t8, c1, f2_fma(t4, c2, f2_fma(t2, c3, f2_mul(t1, c4)))));
        rs2 = f2_add(rs2, w2);
        float w0, w1;
        f2_unpack(w2, w0, w1);
        pw[q] = pack_bf16x2(w0, w1);
        if (plo)      // split-operand pass 3: the remainder of P after its bf16 rounding
          pw[q] = pack_bf16x2(w0 - __uint_as_float(pw[q] << 16), w1 - __uint_as_float(pw[q] & 0xffff0000u));
      }
      pk[c].x = pw[0]; pk[c].y = pw[1]; pk[c].z = pw[2]; pk[c].w = pw[3];
    }
    float a, b;
    f2_unpack(ks2, a, b);
    ksum += a + b;
    f2_unpack(rs2, a, b);
    rsum += a + b;
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
      float kk = 0.f, ww = 0.f;
      for (int m = 0; m < n_scales; ++m) {
        const float em = ex2_approx(-uarg * ratio[m]);
        kk += em;
        ww = fmaf(em, cf[m], ww);
      }
      ksum += kk;
      w[e] = ww;
      rsum += w[e];
    }
    uint32_t pw[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      pw[q] = pack_bf16x2(w[2 * q], w[2 * q + 1]);
      if (plo)
        pw[q] = pack_bf16x2(w[2 * q] - __uint_as_float(pw[q] << 16), w[2 * q + 1] - __uint_as_float(pw[q] & 0xffff0000u));
    }
    pk[c].x = pw[0]; pk[c].y = pw[1]; pk[c].z = pw[2]; pk[c].w = pw[3];
  }
  }
}

// One thread = one row of the tile.  Reads S[row, col0 .. col0+31] from TMEM, forms
//   u = c0*d2 = fma(S, -2c0, ni + nj),  t = 2^-u,  k = sum_m t^(2^q),  w = sum_m coef_m t^(2^q)
// (ladder) or the general n_scales form, adds k to ksum and w to rsum, and returns the 32 values of
// P = w as bf16 packed into 4 x uint4 (8 columns each, ascending).
template <bool LADDER>
__device__ __forceinline__ void epi_half_tile(uint32_t taddr, const float* __restrict__ njp, float ni,
                                              float m2c0, const float (&cf)[VGAN_B200_MAX_SCALES],
                                              const float (&ratio)[VGAN_B200_MAX_SCALES], int n_scales,
                                              float& ksum, float& rsum, uint4 (&pk)[4], bool plo = false) {
  uint32_t v[32];
  tmem_ld_32x32(taddr, v);
  tmem_ld_wait();
  epi_compute_32<LADDER>(v, njp, ni, m2c0, cf, ratio, n_scales, ksum, rsum, pk, plo);
}

// 64 columns: both TMEM loads are issued before the single wait, so their latencies overlap
template <bool LADDER>
__device__ __forceinline__ void epi_tile_64(uint32_t taddr, const float* __restrict__ njp, float ni, float m2c0,
                                            const float (&cf)[VGAN_B200_MAX_SCALES],
                                            const float (&ratio)[VGAN_B200_MAX_SCALES], int n_scales,
                                            float& ksum, float& rsum, uint4 (&pk)[8], bool plo = false) {
  uint32_t va[32], vb[32];
  tmem_ld_32x32(taddr, va);
  tmem_ld_32x32(taddr + 32, vb);
  tmem_ld_wait();
  uint4 h0[4], h1[4];
  epi_compute_32<LADDER>(va, njp, ni, m2c0, cf, ratio, n_scales, ksum, rsum, h0, plo);
  epi_compute_32<LADDER>(vb, njp + 32, ni, m2c0, cf, ratio, n_scales, ksum, rsum, h1, plo);
#pragma unroll
  for (int i = 0; i < 4; ++i) { pk[i] = h0[i]; pk[4 + i] = h1[i]; }
}

// dZ rows of one unit: out[row, f] = 4 c_i (O[row, f] - rtot z[row, f]) (x mul), f in the slice.
// The expression is linear in (O, rtot): a unit cut into column ranges adds the ranges' results (accumulate).
// Thread = (row, column half g); O columns of the slice live at tmem_o.
__device__ __forceinline__ void write_grad_rows(const TcParams& p, uint32_t tmem_o, uint32_t lane_addr, int g,
                                                int row, int ti, int row_is_x, int slice, int slice_chunks,
                                                float rtot, bool accumulate = false) {
  const int slice_w = slice_chunks * kChunk;
  const int64_t region_row = static_cast<int64_t>(row_is_x ? ti : ti - p.tiles_x) * kTile + row;
  const int64_t n_region = row_is_x ? p.Nx : p.Ny;
  const bool row_ok = region_row < n_region && region_row >= p.row_begin && region_row < p.row_end;
  const float ci4 = 4.f * (row_is_x ? 1.f / static_cast<float>(p.Nx) : -1.f / static_cast<float>(p.Ny));
  float* out = (row_is_x ? p.dX : p.dY) + (region_row - p.row_begin) * p.D;
  const float* mul = (!row_is_x && p.dY_mul) ? p.dY_mul + (region_row - p.row_begin) * p.D : nullptr;
  const int64_t zoff = (static_cast<int64_t>(ti) * kTile + row) * p.zq_stride;
  const __nv_bfloat16* zrow = p.zq + zoff;
  const __nv_bfloat16* zrow_lo = p.zq_lo ? p.zq_lo + zoff : nullptr;
  if (p.pass_flags & kPassNoRowsum) rtot = 0.f;
  accumulate = accumulate || (p.pass_flags & kPassAccumulate);
  const int f_slice = slice * p.slice_chunks * kChunk;
  const bool vec_ok = !accumulate && (p.D & 3) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0 &&
                      (!mul || (reinterpret_cast<uintptr_t>(mul) & 15) == 0);
  for (int cc = 0; cc < slice_chunks; ++cc) {
    const int col = g * (slice_w / 2) + cc * 32;
    uint32_t v[32];
    tmem_ld_32x32(tmem_o + lane_addr + col, v);
    tmem_ld_wait();
    if (row_ok) {
      const uint4* z4 = reinterpret_cast<const uint4*>(zrow + f_slice + col);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const uint4 zz = z4[c];
        const uint32_t zw[4] = {zz.x, zz.y, zz.z, zz.w};
        uint4 zl = make_uint4(0u, 0u, 0u, 0u);
        if (zrow_lo) zl = reinterpret_cast<const uint4*>(zrow_lo + f_slice + col)[c];
        const uint32_t zlw[4] = {zl.x, zl.y, zl.z, zl.w};
        float gv[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float z = __uint_as_float((e & 1) ? (zw[e >> 1] & 0xffff0000u) : (zw[e >> 1] << 16)) +
                          __uint_as_float((e & 1) ? (zlw[e >> 1] & 0xffff0000u) : (zlw[e >> 1] << 16));
          gv[e] = ci4 * (__uint_as_float(v[c * 8 + e]) - rtot * z);
        }
        const int f0 = f_slice + col + c * 8;
        if (vec_ok && f0 + 8 <= p.D) {
          // 16-byte accesses (rows are 16-byte aligned when D % 4 == 0)
          if (mul) {
            const float4 m0 = *reinterpret_cast<const float4*>(mul + f0), m1 = *reinterpret_cast<const float4*>(mul + f0 + 4);
            gv[0] *= m0.x; gv[1] *= m0.y; gv[2] *= m0.z; gv[3] *= m0.w;
            gv[4] *= m1.x; gv[5] *= m1.y; gv[6] *= m1.z; gv[7] *= m1.w;
          }
          *reinterpret_cast<float4*>(out + f0) = make_float4(gv[0], gv[1], gv[2], gv[3]);
          *reinterpret_cast<float4*>(out + f0 + 4) = make_float4(gv[4], gv[5], gv[6], gv[7]);
        } else {
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const int f = f0 + e;
            if (f < p.D) {
              float gval = gv[e];
              if (mul) gval *= mul[f];
              if (accumulate) atomicAdd(out + f, gval);   // unit shared between clusters (stream-K cut)
              else out[f] = gval;
            }
          }
        }
      }
    }
  }
}

}  // namespace vg
