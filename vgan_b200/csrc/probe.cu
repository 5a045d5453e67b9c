// Debug aid (vgan_b200_umma_probe): one Gram tile and one gradient tile through exactly
// the descriptors K1 uses — K-major SW128 A/B for S = A B^T, then bf16(S) written to shared
// memory in the swizzled K-major A layout and B re-read as an MN-major operand spanning two
// 64-column blocks (LBO) for O = bf16(S) B.  Raw S and O are returned so a test can tell
// which descriptor is wrong.  A, B: bf16 [128,128] row-major.
#include "common.cuh"
#include "ptx.cuh"

namespace vg {

constexpr int kProbeSmem = 4 * kChunkBytes /*A0 A1 B0 B1*/ + 2 * kChunkBytes /*P*/ + 1024 + 256;

__global__ void __launch_bounds__(128, 1)
umma_probe_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                  float* __restrict__ S_out, float* __restrict__ O_out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* sA = smem;                       // 2 chunks [128 x 64]
  uint8_t* sB = smem + 2 * kChunkBytes;     // 2 chunks [128 x 64]
  uint8_t* sP = smem + 4 * kChunkBytes;     // 2 K-major atoms [128 x 64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 6 * kChunkBytes);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);
  const int warp = threadIdx.x >> 5;

  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_init(&bars[2], 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(tmem_slot, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(&bars[0], 4 * kChunkBytes);
    for (int c = 0; c < 2; ++c) {
      tma_load_2d(sA + c * kChunkBytes, &mapA, &bars[0], c * kChunk, 0);
      tma_load_2d(sB + c * kChunkBytes, &mapB, &bars[0], c * kChunk, 0);
    }
    mbar_wait(&bars[0], 0);
    tc_fence_after();
    constexpr uint32_t idesc = make_idesc_bf16(128, 128, 0, 0);
    for (int c = 0; c < 2; ++c)
      for (int kk = 0; kk < 4; ++kk)
        umma_bf16_ss(tmem, make_sdesc(smem_u32(sA) + c * kChunkBytes + kk * 32, 0, 1024),
                     make_sdesc(smem_u32(sB) + c * kChunkBytes + kk * 32, 0, 1024), idesc,
                     (c | kk) != 0);
    umma_commit(&bars[1]);
  }
  mbar_wait(&bars[1], 0);
  tc_fence_after();
  const int row = threadIdx.x;
  const uint32_t lane_addr = static_cast<uint32_t>(warp * 32) << 16;
  for (int cc = 0; cc < 4; ++cc) {
    uint32_t v[32];
    tmem_ld_32x32(tmem + lane_addr + cc * 32, v);
    tmem_ld_wait();
    for (int e = 0; e < 32; ++e) S_out[row * 128 + cc * 32 + e] = __uint_as_float(v[e]);
    for (int c = 0; c < 4; ++c) {
      uint4 pk;
      pk.x = pack_bf16x2(__uint_as_float(v[c * 8 + 0]), __uint_as_float(v[c * 8 + 1]));
      pk.y = pack_bf16x2(__uint_as_float(v[c * 8 + 2]), __uint_as_float(v[c * 8 + 3]));
      pk.z = pack_bf16x2(__uint_as_float(v[c * 8 + 4]), __uint_as_float(v[c * 8 + 5]));
      pk.w = pack_bf16x2(__uint_as_float(v[c * 8 + 6]), __uint_as_float(v[c * 8 + 7]));
      const int j = cc * 32 + c * 8;             // first column of this 16-byte chunk
      const int atom = j >> 6, chunk16 = (j & 63) >> 3;
      *reinterpret_cast<uint4*>(sP + atom * kChunkBytes + row * 128 + ((chunk16 ^ (row & 7)) << 4)) = pk;
    }
  }
  tc_fence_before();
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) {
    tc_fence_after();
    constexpr uint32_t idesc_pv = make_idesc_bf16(128, 128, 0, 1);
    for (int kk = 0; kk < 8; ++kk)
      umma_bf16_ss(tmem + 128,
                   make_sdesc(smem_u32(sP) + (kk >> 2) * kChunkBytes + (kk & 3) * 32, 0, 1024),
                   make_sdesc(smem_u32(sB) + kk * 2048, kChunkBytes, 1024), idesc_pv, kk != 0);
    umma_commit(&bars[2]);
  }
  mbar_wait(&bars[2], 0);
  tc_fence_after();
  for (int cc = 0; cc < 4; ++cc) {
    uint32_t v[32];
    tmem_ld_32x32(tmem + 128 + lane_addr + cc * 32, v);
    tmem_ld_wait();
    for (int e = 0; e < 32; ++e) O_out[row * 128 + cc * 32 + e] = __uint_as_float(v[e]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

int launch_probe(const CUtensorMap& mapA, const CUtensorMap& mapB, float* S_out, float* O_out,
                 cudaStream_t st) {
  VG_CUDA_CHECK(cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     kProbeSmem));
  umma_probe_kernel<<<1, 128, kProbeSmem, st>>>(mapA, mapB, S_out, O_out);
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

}  // namespace vg

// ------------------------------------------------------------------ MMA issue-rate probe
// Issues `iters` x 4 tcgen05.mma (M=128, N=n, K=16) on resident shared-memory operands (no TMA
// traffic) and returns the elapsed SM cycles: the achievable tensor rate for SS operands.
namespace vg {
__global__ void __launch_bounds__(384, 1)
umma_rate_kernel(const __grid_constant__ CUtensorMap zmap, int n, int iters, int b_mn, int tma_warps, int tma_tiles,
                 long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar, tbar[8];
  __shared__ uint32_t tmem_slot;
  __shared__ volatile int done;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 6 * kChunkBytes / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    for (int i = 0; i < 8; ++i) mbar_init(&tbar[i], 1);
    done = 0;
    fence_barrier_init();
  }
  if (warp == 0) { tmem_alloc(&tmem_slot, 512); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (warp == 0) {
    const uint32_t idesc = make_idesc_bf16(128, n, 0, b_mn);
    const uint32_t a0 = smem_u32(smem), b0 = smem_u32(smem) + 2 * kChunkBytes;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const uint64_t ad = make_sdesc(a0 + (it & 1) * kChunkBytes + kk * 32, 0, 1024);
          const uint64_t bd = b_mn ? make_sdesc(b0 + kk * 2048, kChunkBytes, 1024)
                                   : make_sdesc(b0 + kk * 32, 0, 1024);
          umma_bf16_ss(tmem + (it & 1) * 256, ad, bd, idesc, 1);
        }
      }
      __syncwarp();
    }
    if (elect_one()) umma_commit(&bar);
    __syncwarp();
    mbar_wait(&bar, 0);
    if (threadIdx.x == 0) { out[2 * blockIdx.x] = clock64() - t0; done = 1; }
  } else if (warp >= 4 && warp < 4 + tma_warps && (threadIdx.x & 31) == 0) {
    // concurrent TMA traffic into spare shared memory (beyond the 6 operand chunks): each warp keeps one
    // 16 KiB box in flight and re-issues it as soon as it lands, until the MMA loop is finished
    const int w = warp - 4;
    uint8_t* dst = smem + (6 + w) * kChunkBytes;
    long long n_loads = 0;
    uint32_t ph = 0;
    int tile = (blockIdx.x * 7 + w) % tma_tiles, kc = w;
    while (!done) {
      mbar_arrive_expect_tx(&tbar[w], kChunkBytes);
      tma_load_2d(dst, &zmap, &tbar[w], (kc & 15) * kChunk, tile * kTile);
      mbar_wait(&tbar[w], ph);
      ph ^= 1;
      ++n_loads; ++kc;
      if ((kc & 15) == 0) tile = (tile + 1) % tma_tiles;
    }
    atomicAdd(reinterpret_cast<unsigned long long*>(&out[2 * blockIdx.x + 1]), static_cast<unsigned long long>(n_loads));
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

int launch_rate_probe(const CUtensorMap& zmap, int n, int iters, int b_mn, int tma_warps, int tma_tiles, int grid,
                      long long* out, cudaStream_t st) {
  const int bytes = (6 + 8) * kChunkBytes;
  VG_CUDA_CHECK(cudaFuncSetAttribute(umma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  VG_CUDA_CHECK(cudaMemsetAsync(out, 0, sizeof(long long) * 2 * grid, st));
  umma_rate_kernel<<<grid, 384, bytes, st>>>(zmap, n, iters, b_mn, tma_warps, tma_tiles, out);
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}
}  // namespace vg
