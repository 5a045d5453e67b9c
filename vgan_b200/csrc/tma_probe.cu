// Measurement aid (vgan_b200_tma_rate_probe): L2 -> shared-memory TMA throughput of one CTA per SM
// as a function of the bytes kept in flight, with the access patterns of K1's operand ring.  No MMA:
// a consumer thread releases every stage as soon as it lands.  Used to size the ring.
//   mode 0: every CTA streams the SAME tiles in the same order (the column operand Z_J)
//   mode 1: every CTA re-reads its OWN 128-row tile (the row operand Z_I, always an L2 hit)
//   mode 2: each stage = one private chunk + one shared chunk (a Gram stage)
#include "common.cuh"
#include "ptx.cuh"

namespace vg {

__global__ void __launch_bounds__(512, 1)
tma_rate_kernel(const __grid_constant__ CUtensorMap map, int tiles, int kchunks, int stages, int loads, int box_bytes,
                int iters, int mode, int nprod, long long* __restrict__ cycles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t full[16], empty[16];
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    fence_barrier_init();
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool lane_par = mode >= 4;
  mode &= 3;
  const int stage_bytes = loads * box_bytes;
  const int own = blockIdx.x % tiles;
  const long long t0 = clock64();
  // producers = warps [0, nprod), consumers = warps [nprod, 2 nprod); worker w handles iterations i = w (mod nprod)
  if (warp < nprod && (lane == 0 || lane_par)) {
    int s = warp % stages, ph = 0;                 // nprod divides stages: each producer owns fixed stages
    for (int i = warp; i < iters; i += nprod) {
      mbar_wait(&empty[s], ph ^ 1);
      if (lane == 0) mbar_arrive_expect_tx(&full[s], stage_bytes);
      __syncwarp();
      uint8_t* dst = smem + s * stage_bytes;
      if (lane_par) {
        // one load per lane, a single warp instruction
        if (lane < loads) {
          const int l = lane;
          const bool shared_load = mode == 0 || (mode == 2 && (l & 1));
          const int q = (mode == 2 && loads > 1) ? (i * loads + l) >> 1 : i * loads + l;
          const int kc = q & (kchunks - 1), tile = (q / kchunks) & (tiles - 1);
          tma_load_2d(dst + l * box_bytes, &map, &full[s], kc * kChunk, (shared_load ? tile : own) * kTile);
        }
        __syncwarp();
      } else {
        for (int l = 0; l < loads; ++l) {
          const bool shared_load = mode == 0 || (mode == 2 && (l & 1));
          const int q = (mode == 2 && loads > 1) ? (i * loads + l) >> 1 : i * loads + l;
          const int kc = q & (kchunks - 1), tile = (q / kchunks) & (tiles - 1);   // powers of two
          tma_load_2d(dst + l * box_bytes, &map, &full[s], kc * kChunk, (shared_load ? tile : own) * kTile);
        }
      }
      s += nprod;
      if (s >= stages) { s -= stages; ph ^= 1; }
    }
  } else if (warp >= nprod && warp < 2 * nprod && lane == 0) {
    int s = (warp - nprod) % stages, ph = 0;
    for (int i = warp - nprod; i < iters; i += nprod) {
      mbar_wait(&full[s], ph);
      mbar_arrive(&empty[s]);
      s += nprod;
      if (s >= stages) { s -= stages; ph ^= 1; }
    }
    if (warp == nprod) cycles[blockIdx.x] = clock64() - t0;
  }
}

int launch_tma_rate(const CUtensorMap& map, int tiles, int kchunks, int stages, int loads, int box_rows, int iters,
                    int mode, int grid, int nprod, long long* cycles, cudaStream_t st) {
  const int box_bytes = box_rows * kChunk * 2;
  const int bytes = stages * loads * box_bytes;
  if (stages < 1 || stages > 16 || bytes > 220 * 1024) {
    set_error("tma_rate_probe: stages in [1,16], stages*loads*box <= 220 KiB");
    return VGAN_B200_ERR_ARG;
  }
  VG_CUDA_CHECK(cudaFuncSetAttribute(tma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  tma_rate_kernel<<<grid, 64 * nprod, bytes, st>>>(map, tiles, kchunks, stages, loads, box_bytes, iters, mode, nprod, cycles);
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

}  // namespace vg
