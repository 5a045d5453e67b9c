// Measurement / debugging entry points.  Compiled ONLY into libvgan_b200_dbg.so (`make DEBUG=1`, -DVGAN_B200_DEBUG),
// which tools/ and one GPU test load; the product library has none of them and no switch that changes results.
#ifdef VGAN_B200_DEBUG
#include "common.cuh"

namespace vg {
int launch_rate_probe(const CUtensorMap& zmap, int n, int iters, int b_mn, int tma_warps, int tma_tiles, int grid,
                      long long* out, cudaStream_t st);
int launch_tma_rate(const CUtensorMap& map, int tiles, int kchunks, int stages, int loads, int box_rows, int iters,
                    int mode, int grid, int nprod, long long* cycles, cudaStream_t st);
int launch_pair_probe(const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, float*, float*,
                      cudaStream_t);
int launch_probe(const CUtensorMap& mapA, const CUtensorMap& mapB, float* S_out, float* O_out, cudaStream_t st);
}  // namespace vg

using namespace vg;

extern "C" {

/* when buf != NULL, the next bf16 mmd2_fwd_bwd calls on this context add the kernel sum of every (work unit, column
 * tile) / the pair kernel's wait timers to buf[unit * stride + ...] (device floats) */
VGAN_B200_API int vgan_b200_debug_tile_sums(vgan_b200_ctx* ctx, float* buf, int stride) {
  ctx->dbg = buf;
  ctx->dbg_stride = stride;
  return VGAN_B200_OK;
}

/* buf: 4 device uint64, caller presets {~0, 0, ~0, 0}; the D > 256 kernels record the %globaltimer (ns) of their first
 * CTA start / last CTA end: [0,1] gradient kernel, [2,3] S_xx kernel */
VGAN_B200_API int vgan_b200_debug_timeline(vgan_b200_ctx* ctx, unsigned long long* buf) {
  ctx->timeline = buf;
  return VGAN_B200_OK;
}

/* measurement switches that BREAK results (TcParams::dbg_flags: 1 = 16-byte P exchange, 2 = no Gram operand loads,
 * 4 = no V loads, 8 = DSMEM exchange) and the S_xx scheduling mode (0 overlap, 1 serial, 2 S_xx first) */
VGAN_B200_API int vgan_b200_debug_variant(vgan_b200_ctx* ctx, int dbg_flags, int xx_mode) {
  ctx->dbg_flags = dbg_flags;
  ctx->xx_mode = xx_mode;
  return VGAN_B200_OK;
}

/* one Gram tile and one gradient tile through the tcgen05 pipeline with raw outputs. A,B bf16 [128,128] row-major;
 * S_out float [128,128] = A B^T;  O_out float [128,128] = bf16(S) B */
VGAN_B200_API int vgan_b200_umma_probe(vgan_b200_ctx* ctx, const void* A, const void* B, float* S_out, float* O_out,
                                       void* stream) {
  CUtensorMap ma, mb;
  if (int e = get_tmap_bf16_2d(ctx, &ma, A, 128, 128)) return e;
  if (int e = get_tmap_bf16_2d(ctx, &mb, B, 128, 128)) return e;
  return launch_probe(ma, mb, S_out, O_out, static_cast<cudaStream_t>(stream));
}

/* cta_group::2 Gram + gradient tile on one CTA pair. A0,A1,B bf16 [128,128], V bf16 [128,256];
 * S_out float [2,128,128] = A_h B^T;  O_out float [2,128,256] = bf16(S_h) V */
VGAN_B200_API int vgan_b200_pair_probe(vgan_b200_ctx* ctx, const void* A0, const void* A1, const void* B, const void* V,
                                       float* S_out, float* O_out, void* stream) {
  CUtensorMap a0, a1, bh, v;
  if (int e = get_tmap_bf16_2d(ctx, &a0, A0, 128, 128)) return e;
  if (int e = get_tmap_bf16_2d(ctx, &a1, A1, 128, 128)) return e;
  if (int e = get_tmap_bf16_2d(ctx, &bh, B, 128, 128, 64)) return e;
  if (int e = get_tmap_bf16_2d(ctx, &v, V, 128, 256)) return e;
  return launch_pair_probe(a0, a1, bh, v, S_out, O_out, static_cast<cudaStream_t>(stream));
}

/* L2 -> shared TMA throughput with `stages` x `loads` boxes ([box_rows x 64] bf16) in flight per CTA */
VGAN_B200_API int vgan_b200_tma_rate_probe(vgan_b200_ctx* ctx, const void* Z, int64_t rows, int64_t cols, int stages,
                                           int loads, int box_rows, int iters, int mode, int grid, int nprod,
                                           long long* cycles_out, void* stream) {
  if (!Z || !cycles_out || rows % kTile || cols % kChunk || (box_rows != 64 && box_rows != 128)) {
    set_error("tma_rate_probe: bad arguments");
    return VGAN_B200_ERR_ARG;
  }
  CUtensorMap map;
  if (int e = get_tmap_bf16_2d(ctx, &map, Z, static_cast<uint64_t>(rows), static_cast<uint64_t>(cols), box_rows)) return e;
  return launch_tma_rate(map, static_cast<int>(rows / kTile), static_cast<int>(cols / kChunk), stages, loads, box_rows,
                         iters, mode, grid, nprod < 1 ? 1 : (nprod > 8 ? 8 : nprod), cycles_out,
                         static_cast<cudaStream_t>(stream));
}

/* SM cycles for iters x 4 tcgen05.mma (M=128, N=n, K=16) on resident operands while TMA warps stream boxes */
VGAN_B200_API int vgan_b200_umma_rate_probe(vgan_b200_ctx* ctx, int n, int iters, int b_mn, const void* Z, int64_t rows,
                                            int tma_warps, int grid, long long* out, void* stream) {
  if (!Z || rows % kTile || tma_warps < 0 || tma_warps > 8 || grid < 1 || !out) {
    set_error("umma_rate_probe: bad arguments");
    return VGAN_B200_ERR_ARG;
  }
  CUtensorMap map;
  if (int e = get_tmap_bf16_2d(ctx, &map, Z, static_cast<uint64_t>(rows), 1024)) return e;
  return launch_rate_probe(map, n, iters, b_mn, tma_warps, static_cast<int>(rows / kTile), grid, out,
                           static_cast<cudaStream_t>(stream));
}

}  // extern "C"
#endif  // VGAN_B200_DEBUG
