// Shared constants, the context, the parameter block and host helpers of the vgan_b200 kernels.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>

#include "../../include/vgan_b200.h"

namespace vg {

constexpr int kTile = 128;          // rows per tile (TMEM lanes) and columns per step
constexpr int kChunk = 64;          // bf16 elements per K chunk = one 128-byte swizzle row
constexpr int kChunkBytes = kTile * kChunk * 2;   // 16 KiB
constexpr int kSliceMax = 256;      // output features owned by one CTA (TMEM columns of O)
constexpr float kPadNorm = 1e30f;   // |z|^2 of padding rows: makes every exp underflow to 0
constexpr float kLog2e = 1.4426950408889634f;

// Device-resident parameter block written by scales_kernel (no host sync needed), floats:
// par[0] = c0 = log2(e) * inv_s[ref]       par[1] = -2 c0
// par[2 + m]  = inv_s[m] / inv_s[ref]      (general path exponent ratios)
// par[10 + m] = cx * inv_s[m]              (P coefficients for X columns, cx = +1/Nx)
// par[18 + m] = cy * inv_s[m]              (P coefficients for Y columns, cy = -1/Ny)
// par[26 + 2 a + b] = c_b * w0_ab          (common-mode level of row block a (0 = X, 1 = Y) x column block b)
// par[30 + a] = B_a                        (added to the row sum of P' of a row of block a)
// par[32 + m] = inv_s[m]                   par[40] = bandwidth
constexpr int kParCoefX = 10, kParCoefY = 18, kParCw0 = 26, kParB = 30, kParInvS = 32, kParBw = 40;
constexpr int kPassNoSums = 1, kPassNoRowsum = 2, kPassAccumulate = 4;
constexpr int kParFloats = 64;
// misc words of the accumulator block (layout.off_misc): float penalty, float loss, uint32 "CTAs started"
constexpr int kMiscPenalty = 0, kMiscLoss = 4, kMiscStarted = 8, kMiscBytes = 64;
constexpr int kXxMinFreeSms = 8;         // fewer SMs left free by the gradient clusters: S_xx runs over all SMs at low priority instead
// P-tile exchange scratch (pair kernel): 3 slots of one bf16 [128,128] tile per SM, indexed by %smid so that a
// region is only ever rewritten by the CTA resident on that SM (stays in L2, never streams to HBM)
constexpr int kXchgSms = 256;
constexpr int kXchgBytesPerSm = 4 * 2 * kChunkBytes;   // up to 4 P tiles of 32 KiB per SM
constexpr size_t kXchgTotalBytes = static_cast<size_t>(kXchgSms) * kXchgBytesPerSm;

struct TcParams {
  int tiles_x, tiles_y;       // 128-row tiles in the X / Y regions of the stacked operand
  int kchunks;                // Dp / 64 (feature chunks: slices, O columns, V operand)
  int gchunks;                // K chunks of the Gram GEMM: kchunks, or 3 * kchunks with split operands (BF16X3)
  int v_chunk0;               // chunk offset of the V operand inside its array (split: hi block 0, lo block 2 * kchunks)
  int v_parts, v_chunk1;      // pair kernel: 2 = every P tile multiplies a second V block (at v_chunk1) as well
  int p_lo;                   // 1: the P operand is the bf16 remainder P - bf16(P) (third pass of BF16X3)
  int pass_flags;             // kPassNoSums | kPassNoRowsum | kPassAccumulate (passes 2, 3 of BF16X3)
  int nslices, slice_chunks;  // output feature slices per row tile, chunks per (full) slice
  int cs, ngroups;            // pair kernel: feature slices per cluster, slice groups (clusters) per row-tile pair
  int gx_begin, gx_end;       // X row tiles that need a gradient (NEED_DX), else empty
  int gy_begin, gy_end;       // Y row tiles of this shard
  int xx_part, xx_nparts;     // S_xx-only units: X tiles t with t % nparts == part
  int n_xx_units;
  int ones;                   // 3: the stored rows carry three features 1.0 behind the D real ones (VGAN_B200_ONES)
  int col_split, xx_chunk;    // mmd_tc.cu (set by its launcher): column ranges per gradient unit, tiles per S_xx chunk
  int pv;                     // 1: run the gradient GEMM and write dX/dY
  int n_scales;
  int max_clusters;           // pair kernel: cap on the persistent clusters (0 = none)
  int64_t Nx, Ny, D, Dp;
  int64_t row_begin, row_end; // gradient row range (rows of Y, and of X with NEED_DX)
  const __nv_bfloat16* zq;    // stacked operand, row stride zq_stride
  int64_t zq_stride;
  // the UNROUNDED rows for the -rowsum(P) z_i term of the gradient (fp32 inputs, row stride D): z_i = x_i - mean for
  // X rows, (y_i or u_i * x_i) - mean for Y rows.  Reading z_i in bf16 here was the whole gradient error (1.6e-3 =
  // the relative L2 error of one bf16-rounded array, profiles/parity_r2_before.json).
  const float* Xf;
  const float* Uf;
  const float* Yf;
  const float* mean;
  const float* nrms;          // scaled norms
  const float* nrm;           // unscaled norms |z_i|^2 of the stored rows
  const float* par;           // parameter block
  const float* cm;            // [2][Dp] common-mode gradient vectors A_x, A_y
  double* sums;
  double* colsum;             // [D] column sums of the gradient rows written (after dY_mul), or nullptr
  float* dY;
  const float* dY_mul;
  float* dX;
  unsigned* started;          // pair kernel: CTAs that have begun executing (the S_xx launch waits on it), or nullptr
  uint8_t* xchg;              // P-tile exchange scratch, kXchgSms x kXchgBytesPerSm (pair kernel), L2 resident
#ifdef VGAN_B200_DEBUG
  float* dbg;                 // debug: per-(unit, tile) kernel sums / wait timers, or nullptr
  int dbg_stride;
  unsigned long long* timeline;   // [0,1] = first start / last end (globaltimer ns) of the pair kernel, [2,3] = S_xx kernel
  int dbg_flags;              // measurement only (wrong results): see mmd_tc_pair.cu
#endif
};

#define VG_CUDA_CHECK(expr)                                                          \
  do {                                                                               \
    cudaError_t _e = (expr);                                                         \
    if (_e != cudaSuccess) {                                                         \
      vg::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                    __LINE__);                                                       \
      return VGAN_B200_ERR_CUDA;                                                     \
    }                                                                                \
  } while (0)

void set_error(const char* fmt, ...);
// kernels launched through the calling thread's current context (vgan_b200_launch_count): every launcher calls
// count_launch() after a launch; the entry point that is running points the thread at its context's counter
extern thread_local unsigned long long* tl_launch_counter;
inline void count_launch(int n = 1) {
  if (tl_launch_counter) *tl_launch_counter += static_cast<unsigned long long>(n);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);
typedef CUresult (*StreamWaitValueFn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);

}  // namespace vg

// The opaque context of the C ABI: everything that outlives a call.  One per (host thread, device, use).
struct vgan_b200_ctx {
  int device = 0;
  int n_sms = 0;
  unsigned long long launches = 0;          // kernels launched through this context (vgan_b200_launch_count)
  int cluster_limit = 0;                    // vgan_b200_set_cluster_limit
  int resident_clusters[9] = {0};           // pair kernel: clusters of size i that fit the device (0 = not asked yet)
  vg::EncodeTiledFn encode = nullptr;
  vg::StreamWaitValueFn stream_wait = nullptr;
  // the D > 256 path forks the gradient kernel and the S_xx kernel onto these and joins them into the caller's stream
  cudaStream_t side = nullptr, hi = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_join_hi = nullptr;
  // tensor maps are pure functions of (base, rows, cols, box rows): a training loop re-uses one workspace
  struct MapEntry { const void* base; uint64_t rows, cols; uint32_t box_rows; CUtensorMap map; };
  static constexpr int kMaps = 16;
  MapEntry maps[kMaps];
  int n_maps = 0, next_map = 0;
#ifdef VGAN_B200_DEBUG
  float* dbg = nullptr;
  int dbg_stride = 0;
  unsigned long long* timeline = nullptr;
  int dbg_flags = 0;
  int xx_mode = 0;
#endif
};

namespace vg {

// bf16 [rows, cols] row-major, box = 64 columns x box_rows rows, 128-byte swizzle (cached in the context)
int get_tmap_bf16_2d(vgan_b200_ctx* ctx, CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols,
                     uint32_t box_rows = 128);

// kernels' host launchers (defined in their own .cu files)
// map = row operand (A), mapB = column operand of the Gram, mapV = value operand of the gradient GEMM; the three are
// the same array on the plain bf16 path and differ with split operands (BF16X3, see prologue.cu)
int launch_tc(const TcParams& p, const CUtensorMap& map, const CUtensorMap& mapB, const CUtensorMap& mapV, int ladder,
              cudaStream_t st, int split, int n_sms);   // split: BF16X3 in one pass (D_p <= 256)
int launch_xx(vgan_b200_ctx* ctx, const TcParams& p, const CUtensorMap& map, const CUtensorMap& map256, int ladder,
              cudaStream_t st, int max_ctas = 0);
int launch_tc_pair(vgan_b200_ctx* ctx, const TcParams& p, const CUtensorMap& map, const CUtensorMap& mapB64,
                   const CUtensorMap& mapV64, int ladder, cudaStream_t st, int* resident_ctas = nullptr);

}  // namespace vg
