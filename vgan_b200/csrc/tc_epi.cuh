// Shared epilogue math of the tcgen05 K1 kernels: 32 TMEM columns of S -> kernel values.
#pragma once
#include "common.cuh"
#include "ptx.cuh"

namespace vg {

// One thread = one row of the tile.  Reads S[row, col0 .. col0+31] from TMEM, forms
//   u = c0*d2 = fma(S, -2c0, ni + nj),  t = 2^-u,  k = sum_m t^(2^q),  w = sum_m coef_m t^(2^q)
// (ladder) or the general n_scales form, adds k to ksum and w to rsum, and returns the 32 values of
// P = w as bf16 packed into 4 x uint4 (8 columns each, ascending).
template <bool LADDER>
__device__ __forceinline__ void epi_half_tile(uint32_t taddr, const float* __restrict__ njp, float ni,
                                              float m2c0, const float (&cf)[VGAN_B200_MAX_SCALES],
                                              const float (&ratio)[VGAN_B200_MAX_SCALES], int n_scales,
                                              float& ksum, float& rsum, uint4 (&pk)[4]) {
  uint32_t v[32];
  tmem_ld_32x32(taddr, v);
  tmem_ld_wait();
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
      if (LADDER) {
        const float t1 = ex2_approx(-uarg);
        const float t2 = t1 * t1, t4 = t2 * t2, t8 = t4 * t4, t16 = t8 * t8;
        ksum += (t1 + t2) + (t4 + t8) + t16;
        // scale index 4 is the widest kernel (t1), index 0 the narrowest (t16)
        w[e] = fmaf(t16, cf[0], fmaf(t8, cf[1], fmaf(t4, cf[2], fmaf(t2, cf[3], t1 * cf[4]))));
      } else {
        float kk = 0.f, ww = 0.f;
        for (int m = 0; m < n_scales; ++m) {
          const float em = ex2_approx(-uarg * ratio[m]);
          kk += em;
          ww = fmaf(em, cf[m], ww);
        }
        ksum += kk;
        w[e] = ww;
      }
      rsum += w[e];
    }
    pk[c].x = pack_bf16x2(w[0], w[1]);
    pk[c].y = pack_bf16x2(w[2], w[3]);
    pk[c].z = pack_bf16x2(w[4], w[5]);
    pk[c].w = pack_bf16x2(w[6], w[7]);
  }
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
  const __nv_bfloat16* zrow = p.zq + (static_cast<int64_t>(ti) * kTile + row) * p.Dp;
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
        float gv[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float z = __uint_as_float((e & 1) ? (zw[e >> 1] & 0xffff0000u) : (zw[e >> 1] << 16));
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
