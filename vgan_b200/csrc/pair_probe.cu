// Debug aid (vgan_b200_pair_probe): the cta_group::2 building blocks on ONE CTA pair, with raw
// outputs, before they are used in K1:
//   S_h = A_h B^T        M = 256 (128 rows from each CTA), N = 128 (64 B rows from each CTA)
//   O_h = bf16(S_h) V    M = 256, N = 256 (128 V columns from each CTA, MN-major, LBO between
//                        the two 64-column blocks)
// A0, A1, B: bf16 [128,128];  V: bf16 [128,256];  S_h float [128,128];  O_h float [128,256].
#include "common.cuh"
#include "ptx.cuh"

namespace vg {

constexpr int kPairSmem = 2 * kChunkBytes /*A*/ + kChunkBytes /*B half: 2 x 8 KiB*/ +
                          2 * kChunkBytes /*V half*/ + 2 * kChunkBytes /*P*/ + 256;

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
pair_probe_kernel(const __grid_constant__ CUtensorMap mapA0, const __grid_constant__ CUtensorMap mapA1,
                  const __grid_constant__ CUtensorMap mapBh, const __grid_constant__ CUtensorMap mapV,
                  float* __restrict__ S_out, float* __restrict__ O_out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;                       // 2 K-chunks [128 x 64]
  uint8_t* sB = smem + 2 * kChunkBytes;     // 2 K-chunks [64 x 64] (this CTA's half of the N rows)
  uint8_t* sV = smem + 3 * kChunkBytes;     // 2 feature chunks [128 j x 64 f] (this CTA's 128 features)
  uint8_t* sP = smem + 5 * kChunkBytes;     // 2 K-major atoms [128 x 64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 7 * kChunkBytes);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);
  const int warp = threadIdx.x >> 5;
  const uint32_t h = cluster_ctarank();     // 0 = leader
  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_init(&bars[2], 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc_2cta(tmem_slot, 512);
    tmem_relinquish_2cta();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (threadIdx.x == 0) {
    // every CTA loads its own operands; all bytes are counted on the leader's barrier
    if (h == 0) mbar_arrive_expect_tx(&bars[0], 2 * (2 * kChunkBytes + kChunkBytes + 2 * kChunkBytes));
    for (int c = 0; c < 2; ++c) {
      tma_load_2d_2cta(sA + c * kChunkBytes, h == 0 ? &mapA0 : &mapA1, &bars[0], c * kChunk, 0);
      tma_load_2d_2cta(sB + c * (kChunkBytes / 2), &mapBh, &bars[0], c * kChunk, static_cast<int>(h) * 64);
      tma_load_2d_2cta(sV + c * kChunkBytes, &mapV, &bars[0], (static_cast<int>(h) * 2 + c) * kChunk, 0);
    }
    if (h == 0) {
      mbar_wait(&bars[0], 0);
      tc_fence_after();
      constexpr uint32_t idesc = make_idesc_bf16(256, 128, 0, 0);
      for (int c = 0; c < 2; ++c)
        for (int kk = 0; kk < 4; ++kk)
          umma_bf16_ss_2cta(tmem, make_sdesc(smem_u32(sA) + c * kChunkBytes + kk * 32, 0, 1024),
                            make_sdesc(smem_u32(sB) + c * (kChunkBytes / 2) + kk * 32, 0, 1024), idesc,
                            (c | kk) != 0);
      umma_commit_2cta_mc(&bars[1], 0x3);
    }
  }
  mbar_wait(&bars[1], 0);
  tc_fence_after();
  const int row = threadIdx.x;
  const uint32_t lane_addr = static_cast<uint32_t>(warp * 32) << 16;
  float* S_mine = S_out + static_cast<size_t>(h) * 128 * 128;
  for (int cc = 0; cc < 4; ++cc) {
    uint32_t v[32];
    tmem_ld_32x32(tmem + lane_addr + cc * 32, v);
    tmem_ld_wait();
    for (int e = 0; e < 32; ++e) S_mine[row * 128 + cc * 32 + e] = __uint_as_float(v[e]);
    for (int c = 0; c < 4; ++c) {
      uint4 pk;
      pk.x = pack_bf16x2(__uint_as_float(v[c * 8 + 0]), __uint_as_float(v[c * 8 + 1]));
      pk.y = pack_bf16x2(__uint_as_float(v[c * 8 + 2]), __uint_as_float(v[c * 8 + 3]));
      pk.z = pack_bf16x2(__uint_as_float(v[c * 8 + 4]), __uint_as_float(v[c * 8 + 5]));
      pk.w = pack_bf16x2(__uint_as_float(v[c * 8 + 6]), __uint_as_float(v[c * 8 + 7]));
      const int j = cc * 32 + c * 8;
      const int atom = j >> 6, chunk16 = (j & 63) >> 3;
      *reinterpret_cast<uint4*>(sP + atom * kChunkBytes + row * 128 + ((chunk16 ^ (row & 7)) << 4)) = pk;
    }
  }
  tc_fence_before();
  fence_proxy_async_smem();
  cluster_sync_all();            // both CTAs' P tiles are in place
  if (h == 0 && threadIdx.x == 0) {
    tc_fence_after();
    constexpr uint32_t idesc_pv = make_idesc_bf16(256, 256, 0, 1);
    for (int kk = 0; kk < 8; ++kk)
      umma_bf16_ss_2cta(tmem + 128, make_sdesc(smem_u32(sP) + (kk >> 2) * kChunkBytes + (kk & 3) * 32, 0, 1024),
                        make_sdesc(smem_u32(sV) + kk * 2048, kChunkBytes, 1024), idesc_pv, kk != 0);
    umma_commit_2cta_mc(&bars[2], 0x3);
  }
  mbar_wait(&bars[2], 0);
  tc_fence_after();
  float* O_mine = O_out + static_cast<size_t>(h) * 128 * 256;
  for (int cc = 0; cc < 8; ++cc) {
    uint32_t v[32];
    tmem_ld_32x32(tmem + 128 + lane_addr + cc * 32, v);
    tmem_ld_wait();
    for (int e = 0; e < 32; ++e) O_mine[row * 256 + cc * 32 + e] = __uint_as_float(v[e]);
  }
  tc_fence_before();
  cluster_sync_all();
  if (warp == 0) tmem_dealloc_2cta(tmem, 512);
}

int launch_pair_probe(const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& bh, const CUtensorMap& v,
                      float* S_out, float* O_out, cudaStream_t st) {
  VG_CUDA_CHECK(cudaFuncSetAttribute(pair_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kPairSmem));
  pair_probe_kernel<<<2, 128, kPairSmem, st>>>(a0, a1, bh, v, S_out, O_out);
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

}  // namespace vg
