// Shared constants, the parameter block and host helpers of the vgan_b200 kernels.
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

// Device-resident parameter block written by setup_kernel (no host sync needed).
// par[0] = c0 = log2(e) * inv_s[ref]       par[1] = -2 c0
// par[2 + m]  = inv_s[m] / inv_s[ref]      (general path exponent ratios)
// par[10 + m] = cx * inv_s[m]              (P coefficients for X columns, cx = +1/Nx)
// par[18 + m] = cy * inv_s[m]              (P coefficients for Y columns, cy = -1/Ny)
constexpr int kPassNoSums = 1, kPassNoRowsum = 2, kPassAccumulate = 4;
constexpr int kParFloats = 32;
constexpr int kParStartedOffset = 512;   // byte offset in the (1 KiB) parameter block of the pair kernel's "CTAs started" counter
constexpr int kXxMinFreeSms = 8;         // fewer SMs left free by the gradient clusters: S_xx runs over all SMs at low priority instead
// P-tile exchange scratch (pair kernel): 3 slots of one bf16 [128,128] tile per SM, indexed by %smid so that a
// region is only ever rewritten by the CTA resident on that SM (stays in L2, never streams to HBM)
constexpr int kXchgSms = 256;
constexpr int kXchgBytesPerSm = 4 * 2 * kChunkBytes;   // up to 4 P tiles of 32 KiB per SM

struct TcParams {
  int tiles_x, tiles_y;       // 128-row tiles in the X / Y regions of the stacked operand
  int kchunks;                // Dp / 64 (feature chunks: slices, O columns, V operand)
  int gchunks;                // K chunks of the Gram GEMM: kchunks, or 3 * kchunks with split operands (BF16X3)
  int v_chunk0;               // chunk offset of the V operand inside its array (split: hi block 0, lo block 2 * kchunks)
  int p_lo;                   // 1: the P operand is the bf16 remainder P - bf16(P) (third pass of BF16X3)
  int pass_flags;             // kPassNoSums | kPassNoRowsum | kPassAccumulate (passes 2, 3 of BF16X3)
  int nslices, slice_chunks;  // output feature slices per row tile, chunks per (full) slice
  int cs, ngroups;            // cluster kernel: CTAs per cluster, slice groups (clusters) per row tile
  int gx_begin, gx_end;       // X row tiles that need a gradient (NEED_DX), else empty
  int gy_begin, gy_end;       // Y row tiles of this shard
  int xx_part, xx_nparts;     // S_xx-only units: X tiles t with t % nparts == part
  int n_xx_units;
  int pv;                     // 1: run the gradient GEMM and write dX/dY
  int n_scales;
  int64_t Nx, Ny, D, Dp;
  int64_t row_begin, row_end; // gradient row range (rows of Y, and of X with NEED_DX)
  const __nv_bfloat16* zq;    // stacked operand (for the final z_i read), row stride zq_stride
  const __nv_bfloat16* zq_lo; // split operands: the low halves (z_i = hi + lo), same stride; else nullptr
  int64_t zq_stride;
  const float* nrms;          // scaled norms
  const float* par;           // parameter block
  double* sums;
  float* dY;
  const float* dY_mul;
  float* dX;
  unsigned* started;          // pair kernel: CTAs that have begun executing (the S_xx launch waits on it), or nullptr
  uint8_t* xchg;              // P-tile exchange scratch, kXchgSms x kXchgBytesPerSm (cluster kernels), L2 resident
  float* dbg;                 // debug: per-(unit, tile) kernel sums, or nullptr
  int dbg_stride;
  unsigned long long* timeline;   // debug: [0,1] = first start / last end (globaltimer ns) of the pair kernel, [2,3] = S_xx kernel
  int dbg_flags;              // measurement only: bit 0 = exchange 16 bytes per P tile instead of 32 KiB (wrong results)
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
int check_device();   // 0 when the current device is compute capability 10.x

// kernels' host launchers (defined in their own .cu files)
// map = row operand (A), mapB = column operand of the Gram, mapV = value operand of the gradient GEMM; the three are
// the same array on the plain bf16 path and differ with split operands (BF16X3, see prologue.cu)
int launch_tc(const TcParams& p, const CUtensorMap& map, const CUtensorMap& mapB, const CUtensorMap& mapV, int ladder,
              cudaStream_t st);
int launch_xx(const TcParams& p, const CUtensorMap& map, const CUtensorMap& map256, int ladder, cudaStream_t st,
              int max_ctas = 0);
int launch_tc_cluster(const TcParams& p, const CUtensorMap& map, int ladder, cudaStream_t st);
int launch_tc_pair(const TcParams& p, const CUtensorMap& map, const CUtensorMap& mapB64, const CUtensorMap& mapV64,
                   int ladder, cudaStream_t st, int* resident_ctas = nullptr);
int launch_probe(const CUtensorMap& mapA, const CUtensorMap& mapB, float* S_out, float* O_out,
                 cudaStream_t st);
int make_tmap_bf16_2d(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows = 128);

}  // namespace vg
