// C ABI of vgan_b200 (include/vgan_b200.h): context, argument checking, workspace layout, TMA
// descriptor creation and kernel launches.  No torch types, no host synchronisation, no global mutable state.
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <new>

#include "common.cuh"

namespace vg {

static thread_local char g_err[512] = "";
thread_local unsigned long long* tl_launch_counter = nullptr;

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// makes the context's device current for the duration of a call (tensors may live on a device that is not the
// caller's current one) and restores the previous one
struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
    if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
  }
  ~DeviceGuard() {
    int cur = -1;
    if (prev >= 0 && cudaGetDevice(&cur) == cudaSuccess && cur != prev) cudaSetDevice(prev);
  }
};

#define VG_ENTER(ctx)                                                                   \
  if (!(ctx)) {                                                                         \
    set_error("null context (vgan_b200_create first)");                                 \
    return VGAN_B200_ERR_ARG;                                                           \
  }                                                                                     \
  DeviceGuard guard_((ctx)->device);                                                    \
  if (!guard_.ok) {                                                                     \
    set_error("cannot make device %d current", (ctx)->device);                          \
    return VGAN_B200_ERR_DEVICE;                                                        \
  }                                                                                     \
  tl_launch_counter = &(ctx)->launches;

int get_tmap_bf16_2d(vgan_b200_ctx* ctx, CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols,
                     uint32_t box_rows) {
  for (int i = 0; i < ctx->n_maps; ++i) {
    const vgan_b200_ctx::MapEntry& e = ctx->maps[i];
    if (e.base == base && e.rows == rows && e.cols == cols && e.box_rows == box_rows) {
      *map = e.map;
      return VGAN_B200_OK;
    }
  }
  if (!ctx->encode) {
    set_error("cuTensorMapEncodeTiled not available from the driver");
    return VGAN_B200_ERR_CUDA;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(kChunk), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = ctx->encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box,
                           estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d", static_cast<int>(r));
    return VGAN_B200_ERR_CUDA;
  }
  vgan_b200_ctx::MapEntry& e = ctx->maps[ctx->next_map];
  e.base = base; e.rows = rows; e.cols = cols; e.box_rows = box_rows; e.map = *map;
  ctx->next_map = (ctx->next_map + 1) % vgan_b200_ctx::kMaps;
  if (ctx->n_maps < vgan_b200_ctx::kMaps) ++ctx->n_maps;
  return VGAN_B200_OK;
}

// launchers defined in the kernel files
int run_stats(const vgan_b200_layout_t&, void*, const float*, const float*, const float*, int64_t, int64_t, cudaStream_t);
int run_scales(const vgan_b200_layout_t&, void*, float*, int, const float*, float, cudaStream_t);
int run_pack(const vgan_b200_layout_t&, void*, const float*, const float*, const float*, int64_t, int64_t, cudaStream_t);
int run_finalize(const vgan_b200_layout_t&, void*, float, float*, int64_t, int64_t, int, float*, cudaStream_t);
int run_upper_softmax(const float*, int64_t, int64_t, float*, float*, float*, cudaStream_t);
int run_upper_softmax_bwd(const float*, const float*, const float*, const float*, int64_t, int64_t, float*, cudaStream_t);
int launch_matrix(const float*, int64_t, int64_t, const float*, int, float, float*, cudaStream_t);
int run_subspace_dist(const void*, int, int64_t, int64_t, float, void*, uint8_t*, long long*, int*, cudaStream_t);
int run_subspace_dist_static(const void*, int, int64_t, int64_t, float, uint8_t*, long long*, int*, cudaStream_t);
size_t perm_scratch_bytes(int64_t n);
int run_permutation_stats(const float*, int64_t, int64_t, const uint8_t*, int64_t, void*, double*, cudaStream_t);
int launch_simt_from_abi(const vgan_b200_layout_t& L, void* ws, int64_t row_begin, int64_t row_end, int xx_part,
                         int xx_nparts, float* dY, const float* dY_mul, float* dX, cudaStream_t st);

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static int fill_layout(int64_t Nx, int64_t Ny, int64_t D, int64_t D_u, int n_scales, int flags,
                       vgan_b200_layout_t* L) {
  if (!L || Nx < 1 || Ny < 1 || D < 1 || D_u < 0) {
    set_error("workspace_layout: need Nx, Ny, D >= 1 (got %lld, %lld, %lld)", (long long)Nx, (long long)Ny,
              (long long)D);
    return VGAN_B200_ERR_ARG;
  }
  if (Nx + Ny > (1ll << 30) || D > (1 << 20) || D_u > (1 << 20)) {
    set_error("workspace_layout: sizes out of range");
    return VGAN_B200_ERR_ARG;
  }
  if (n_scales < 1 || n_scales > VGAN_B200_MAX_SCALES) {
    set_error("workspace_layout: n_scales outside [1, %d]", VGAN_B200_MAX_SCALES);
    return VGAN_B200_ERR_ARG;
  }
  if ((flags & VGAN_B200_LADDER) && n_scales != 5) {
    set_error("workspace_layout: LADDER needs exactly 5 scales");
    return VGAN_B200_ERR_ARG;
  }
  memset(L, 0, sizeof(*L));
  L->Nx = Nx; L->Ny = Ny; L->D = D; L->D_u = D_u;
  L->flags = flags & ~VGAN_B200_ONES; L->n_scales = n_scales;
  const bool fp32 = (flags & VGAN_B200_PREC_MASK) == VGAN_B200_PREC_FP32;
  L->Dp = (D + kChunk - 1) / kChunk * kChunk;
  // bf16, more than one feature slice: an even chunk count, so that every slice is 2 or 4 chunks wide and the
  // CTA-pair kernel (each CTA of a pair holds half a slice) serves every width; at most one extra zero chunk
  if (!fp32 && L->Dp > kSliceMax) L->Dp = (D + 2 * kChunk - 1) / (2 * kChunk) * (2 * kChunk);
  // D = 62 .. 64 on the plain bf16 path: one more (zero) chunk, so that the ones below fit — the second chunk costs the
  // D <= 256 kernel 1 % (N = 16384: D = 100 417 us against D = 32 411 us), the ones save 7 %
  if (!fp32 && !(flags & VGAN_B200_BF16X3) && L->Dp == kChunk && D + 3 > L->Dp) L->Dp = 2 * kChunk;
  // three constant-1 features in the padding (see VGAN_B200_ONES): only where mmd_tc.cu keeps the row operand resident
  // (plain bf16 up to D_p = 256; split operands at D_p = 64) and the padding has room for them
  {
    const bool split_ops = !fp32 && (flags & VGAN_B200_BF16X3);
    if (!fp32 && D + 3 <= L->Dp && L->Dp <= kSliceMax && (!split_ops || L->Dp == kChunk)) L->flags |= VGAN_B200_ONES;
  }
  L->Nxp = (Nx + kTile - 1) / kTile * kTile;
  L->Nyp = (Ny + kTile - 1) / kTile * kTile;
  const size_t rows = static_cast<size_t>(L->Nxp + L->Nyp);
  size_t off = 0;
  L->off_par = off;    off = align_up(off + kParFloats * 4, 1024);
  L->off_acc = off;
  L->off_sums = off;   off += 4 * 8;
  L->off_colsum = off; off = align_up(off + static_cast<size_t>(L->Dp) * 8, 64);
  L->off_stats = off;  off = align_up(off + static_cast<size_t>(2 + 2 * L->Dp) * 8, 64);
  L->off_cmax = off;   off = align_up(off + static_cast<size_t>(D_u > 0 ? D_u : 1) * 8, 64);
  L->off_misc = off;   off += kMiscBytes;
  L->acc_bytes = off - L->off_acc;
  off = align_up(off, 1024);
  L->off_mean = off;   off = align_up(off + L->Dp * 4, 1024);
  L->off_cm = off;     off = align_up(off + 2 * L->Dp * 4, 1024);
  L->off_nrm = off;    off = align_up(off + rows * 4, 1024);
  L->off_nrms = off;   off = align_up(off + rows * 4, 1024);
  L->off_xchg = off;
  if (!fp32 && L->Dp > kSliceMax) off = align_up(off + kXchgTotalBytes, 1024);
  L->off_zq = off;
  const bool split = !fp32 && (flags & VGAN_B200_BF16X3);
  L->off_zf = off;     off = align_up(off + rows * L->Dp * (fp32 ? 4 : (split ? 12 : 2)), 1024);
  L->total = off;
  return VGAN_B200_OK;
}

static int check_call(const vgan_b200_layout_t* L, const void* ws, const char* what) {
  if (!L || !ws) {
    set_error("%s: null layout or workspace", what);
    return VGAN_B200_ERR_ARG;
  }
  if (reinterpret_cast<uintptr_t>(ws) % 1024) {
    set_error("%s: workspace must be 1024-byte aligned", what);
    return VGAN_B200_ERR_ARG;
  }
  return VGAN_B200_OK;
}

static int check_inputs(const vgan_b200_layout_t* L, const float* X, const float* U, const float* Y,
                        int64_t row_begin, int64_t row_end, const char* what) {
  if (!X || (!U && !Y)) {
    set_error("%s: X and one of U / Y must be non-null", what);
    return VGAN_B200_ERR_ARG;
  }
  if (!Y && L->Nx != L->Ny) {
    set_error("%s: masking Y = U*X needs Ny == Nx (got %lld vs %lld)", what, (long long)L->Ny, (long long)L->Nx);
    return VGAN_B200_ERR_ARG;
  }
  if (!Y && L->D_u != 0 && L->D_u != L->D) {
    set_error("%s: masking Y = U*X needs U of width D (D_u %lld, D %lld)", what, (long long)L->D_u, (long long)L->D);
    return VGAN_B200_ERR_ARG;
  }
  if (row_begin < 0 || row_end < row_begin) {
    set_error("%s: bad row range [%lld, %lld)", what, (long long)row_begin, (long long)row_end);
    return VGAN_B200_ERR_ARG;
  }
  return VGAN_B200_OK;
}

static int ensure_aux(vgan_b200_ctx* ctx) {
  if (ctx->side) return VGAN_B200_OK;
  int lo_pri = 0, hi_pri = 0;
  VG_CUDA_CHECK(cudaDeviceGetStreamPriorityRange(&lo_pri, &hi_pri));
  VG_CUDA_CHECK(cudaStreamCreateWithPriority(&ctx->side, cudaStreamNonBlocking, lo_pri));
  VG_CUDA_CHECK(cudaStreamCreateWithPriority(&ctx->hi, cudaStreamNonBlocking, hi_pri));
  VG_CUDA_CHECK(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
  VG_CUDA_CHECK(cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
  VG_CUDA_CHECK(cudaEventCreateWithFlags(&ctx->ev_join_hi, cudaEventDisableTiming));
  return VGAN_B200_OK;
}

// K1 of one shard on the tcgen05 kernels
static int run_tc(vgan_b200_ctx* ctx, const vgan_b200_layout_t& L, void* ws, const float* X, const float* U,
                  const float* Y, int64_t row_begin, int64_t row_end, int xx_part, int xx_nparts, float* dY,
                  const float* dY_mul, float* dX, cudaStream_t st) {
  uint8_t* base = static_cast<uint8_t*>(ws);
  const int flags = L.flags;
  const bool need_dx = (flags & VGAN_B200_NEED_DX) != 0 && dX != nullptr;
  const int64_t Nx = L.Nx, Ny = L.Ny, D = L.D;
  TcParams p;
  memset(&p, 0, sizeof(p));
  p.tiles_x = static_cast<int>(L.Nxp / kTile);
  p.tiles_y = static_cast<int>(L.Nyp / kTile);
  p.kchunks = static_cast<int>(L.Dp / kChunk);
  p.slice_chunks = kSliceMax / kChunk;
  if (p.slice_chunks > p.kchunks) p.slice_chunks = p.kchunks;
  p.pv = dY ? 1 : 0;
  p.nslices = p.pv ? (p.kchunks + p.slice_chunks - 1) / p.slice_chunks : 1;
  const int64_t ye = row_end < Ny ? row_end : Ny, xe = row_end < Nx ? row_end : Nx;
  p.gy_begin = static_cast<int>(row_begin / kTile);
  p.gy_end = ye > row_begin ? static_cast<int>((ye + kTile - 1) / kTile) : p.gy_begin;
  if (need_dx) {
    p.gx_begin = static_cast<int>(row_begin / kTile);
    p.gx_end = xe > row_begin ? static_cast<int>((xe + kTile - 1) / kTile) : p.gx_begin;
    p.n_xx_units = 0;
  } else {
    p.gx_begin = p.gx_end = 0;
    p.n_xx_units = xx_part < p.tiles_x ? (p.tiles_x - xx_part + xx_nparts - 1) / xx_nparts : 0;
  }
  p.xx_part = xx_part; p.xx_nparts = xx_nparts;
  p.n_scales = L.n_scales;
  p.max_clusters = ctx->cluster_limit;
  p.v_parts = 1;
  p.ones = (flags & VGAN_B200_ONES) ? 3 : 0;
  p.Nx = Nx; p.Ny = Ny; p.D = D; p.Dp = L.Dp;
  p.row_begin = row_begin; p.row_end = row_end;
  p.zq = reinterpret_cast<const __nv_bfloat16*>(base + L.off_zq);
  p.nrms = reinterpret_cast<const float*>(base + L.off_nrms);
  p.nrm = reinterpret_cast<const float*>(base + L.off_nrm);
  p.par = reinterpret_cast<const float*>(base + L.off_par);
  p.cm = reinterpret_cast<const float*>(base + L.off_cm);
  p.sums = reinterpret_cast<double*>(base + L.off_sums);
  p.colsum = p.pv ? reinterpret_cast<double*>(base + L.off_colsum) : nullptr;
  p.dY = dY; p.dY_mul = dY_mul; p.dX = dX;
  p.xchg = L.Dp > kSliceMax ? base + L.off_xchg : nullptr;
  p.started = nullptr;
#ifdef VGAN_B200_DEBUG
  p.dbg = ctx->dbg; p.dbg_stride = ctx->dbg_stride;
  p.timeline = ctx->timeline;
  p.dbg_flags = ctx->dbg_flags;
#endif
  // ---- operand arrays and their tensor maps.  Plain bf16: one array [rows, Dp] serves as row operand, column operand
  // and V.  BF16X3: row operand [hi | hi | lo] and column operand [hi | lo | hi], both [rows, 3 Dp]; V = the hi or lo
  // block of the row-operand array.
  const bool split = (flags & VGAN_B200_BF16X3) != 0;
  const uint64_t rows = static_cast<uint64_t>(L.Nxp + L.Nyp);
  const uint64_t width = static_cast<uint64_t>(split ? 3 * L.Dp : L.Dp);
  const __nv_bfloat16* zb = split ? p.zq + rows * width : p.zq;
  p.gchunks = split ? 3 * p.kchunks : p.kchunks;
  p.zq_stride = static_cast<int64_t>(width);
  p.Xf = X; p.Uf = Y ? nullptr : U; p.Yf = Y;
  p.mean = reinterpret_cast<const float*>(base + L.off_mean);
  const int ladder = (flags & VGAN_B200_LADDER) ? 1 : 0;
  p.cs = p.nslices < 4 ? p.nslices : 4;
  p.ngroups = (p.nslices + p.cs - 1) / p.cs;
  // (D_p = 256 stays on mmd_tc.cu: on the CTA-pair kernel, as clusters of one pair, it measured 11.3 against 9.7 ms
  // at N = 65536 and 217 against 82 us at N = 4096 — profiles/r2n7_pair_at_256.txt)
  const bool multi_slice = p.pv && p.nslices >= 2;
  CUtensorMap map, mapB, mapV, mapB64, mapV64, mapB256;
  if (int e = get_tmap_bf16_2d(ctx, &map, p.zq, rows, width)) return e;
  if (int e = get_tmap_bf16_2d(ctx, &mapB, zb, rows, width)) return e;
  mapV = map;
  if (multi_slice) {
    if (int e = get_tmap_bf16_2d(ctx, &mapB64, zb, rows, width, 64)) return e;
    if (int e = get_tmap_bf16_2d(ctx, &mapV64, p.zq, rows, width, 64)) return e;
    if (p.n_xx_units > 0)
      if (int e = get_tmap_bf16_2d(ctx, &mapB256, zb, rows, width, 256)) return e;
  }

  // one pass of the tcgen05 kernels over this shard (BF16X3 runs three: see below)
  auto run_pass = [&](TcParams& p, bool zero_outputs) -> int {
    if (!multi_slice) return launch_tc(p, map, mapB, mapV, ladder, st, 0, ctx->n_sms);
    // D_p > 256: the slice CTAs of two row tiles run as one thread-block cluster (mmd_tc_pair.cu) and share every
    // P tile; the S_xx-only units (no gradient GEMM) run as their own sums-only kernel (mmd_xx.cu) on a low-priority
    // side stream, the cluster kernel on a high-priority one.  8-CTA clusters occupy 120 of the 148 SMs; the S_xx
    // kernel gets the other 28.  The order is enforced: the persistent CTAs count themselves in at their first
    // instruction, the side stream holds the S_xx launch (cuStreamWaitValue32, no SM occupied) until all of them are
    // resident, and the S_xx grid is sized to the SMs they leave free.  (Launched freely, the S_xx CTAs were
    // dispatched first and the clusters started 1-9 ms late.)  The two streams and three events live in the context.
    int xx_mode = 0;
#ifdef VGAN_B200_DEBUG
    xx_mode = ctx->xx_mode;             // 1: serial, 2: S_xx kernel dispatched first (measurement)
#endif
    const bool overlap = p.n_xx_units > 0 && xx_mode != 1;
    if (overlap)
      if (int e = ensure_aux(ctx)) return e;
    cudaStream_t main_st = st;
    if (zero_outputs) {
      // stream-K cuts a unit between two clusters: those rows are accumulated with red.global.add.  Zeroed BEFORE
      // the fork: a memset in front of the gradient kernel on its own stream lets the S_xx kernel start first and
      // occupy SMs the persistent clusters need (measured: the clusters then start up to 9 ms late).
      if (ye > row_begin)
        VG_CUDA_CHECK(cudaMemsetAsync(dY, 0, static_cast<size_t>(ye - row_begin) * D * sizeof(float), st));
      if (need_dx && xe > row_begin)
        VG_CUDA_CHECK(cudaMemsetAsync(dX, 0, static_cast<size_t>(xe - row_begin) * D * sizeof(float), st));
    }
    // (not while the caller captures a CUDA graph: the plain low-priority launch is used there)
    cudaStreamCaptureStatus capturing = cudaStreamCaptureStatusNone;
    VG_CUDA_CHECK(cudaStreamIsCapturing(st, &capturing));
    StreamWaitValueFn stream_wait =
        (overlap && xx_mode == 0 && capturing == cudaStreamCaptureStatusNone) ? ctx->stream_wait : nullptr;
    if (stream_wait) {
      p.started = reinterpret_cast<unsigned*>(base + L.off_misc + kMiscStarted);
      VG_CUDA_CHECK(cudaMemsetAsync(p.started, 0, sizeof(unsigned), st));
    }
    int resident_ctas = 0;
    if (overlap) {
      VG_CUDA_CHECK(cudaEventRecord(ctx->ev_fork, st));
      VG_CUDA_CHECK(cudaStreamWaitEvent(ctx->hi, ctx->ev_fork, 0));
      main_st = ctx->hi;
      if (xx_mode == 2) {
        VG_CUDA_CHECK(cudaStreamWaitEvent(ctx->side, ctx->ev_fork, 0));
        if (int e = launch_xx(ctx, p, map, mapB256, ladder, ctx->side)) return e;
      }
    }
    if (int e = launch_tc_pair(ctx, p, map, mapB64, mapV64, ladder, main_st, &resident_ctas)) return e;
    if (p.n_xx_units == 0) return VGAN_B200_OK;
    if (!overlap) return launch_xx(ctx, p, map, mapB256, ladder, st);
    if (xx_mode != 2) {
      VG_CUDA_CHECK(cudaStreamWaitEvent(ctx->side, ctx->ev_fork, 0));
      int free_sms = 0;
      if (stream_wait && resident_ctas > 0) free_sms = ctx->n_sms - resident_ctas;
      if (free_sms >= kXxMinFreeSms) {
        CUresult r = stream_wait(reinterpret_cast<CUstream>(ctx->side), reinterpret_cast<CUdeviceptr>(p.started),
                                 static_cast<cuuint32_t>(resident_ctas), CU_STREAM_WAIT_VALUE_GEQ);
        if (r != CUDA_SUCCESS) free_sms = 0;   // stream memory operations unavailable: plain low-priority launch
      } else {
        free_sms = 0;       // no room beside the clusters: plain low-priority launch over all SMs
      }
      if (int e = launch_xx(ctx, p, map, mapB256, ladder, ctx->side, free_sms)) return e;
    }
    VG_CUDA_CHECK(cudaEventRecord(ctx->ev_join, ctx->side));
    VG_CUDA_CHECK(cudaEventRecord(ctx->ev_join_hi, ctx->hi));
    VG_CUDA_CHECK(cudaStreamWaitEvent(st, ctx->ev_join, 0));
    VG_CUDA_CHECK(cudaStreamWaitEvent(st, ctx->ev_join_hi, 0));
    return VGAN_B200_OK;
  };
  if (!split) return run_pass(p, true);
  // BF16X3, D_p <= 256: ONE pass — the K = 3 D_p Gram once per column tile, both bf16(P) and P - bf16(P) parked, the
  // three gradient parts issued back to back (mmd_tc.cu, TcCfg<true>)
  if (!multi_slice) return launch_tc(p, map, mapB, mapV, ladder, st, 1, ctx->n_sms);
  // BF16X3, D_p > 256: O = P_hi (V_hi + V_lo) + P_lo V_hi in TWO passes into zeroed outputs.  Pass 1 multiplies every
  // parked P tile with both V blocks (one Gram and one epilogue for two of the three parts) and also produces the
  // kernel sums, the rowsum(P) z_i term and the common-mode remainder, with the S_xx units beside it; pass 2 parks
  // the remainder P - bf16(P) against V_hi.  (Three passes until round 2: 214 ms at the bench shape.)
  if (p.pv) {
    if (ye > row_begin)
      VG_CUDA_CHECK(cudaMemsetAsync(dY, 0, static_cast<size_t>(ye - row_begin) * D * sizeof(float), st));
    if (need_dx && xe > row_begin)
      VG_CUDA_CHECK(cudaMemsetAsync(dX, 0, static_cast<size_t>(xe - row_begin) * D * sizeof(float), st));
  }
  p.pass_flags = kPassAccumulate;
  p.v_chunk0 = 0; p.v_chunk1 = 2 * p.kchunks; p.v_parts = p.pv ? 2 : 1; p.p_lo = 0;
  if (int e = run_pass(p, false)) return e;
  if (!p.pv) return VGAN_B200_OK;
  p.n_xx_units = 0;
  p.pass_flags = kPassAccumulate | kPassNoSums | kPassNoRowsum;
  p.v_chunk0 = 0; p.v_parts = 1; p.p_lo = 1;
  return run_pass(p, false);
}

}  // namespace vg

using namespace vg;

extern "C" {

int vgan_b200_abi_version(void) { return VGAN_B200_ABI_VERSION; }
const char* vgan_b200_last_error(void) { return g_err; }

int vgan_b200_create(int device, vgan_b200_ctx** out) {
  if (!out) {
    set_error("create: null output pointer");
    return VGAN_B200_ERR_ARG;
  }
  *out = nullptr;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count < 1) {
    cudaGetLastError();
    set_error("no CUDA device available (vgan_b200 has no CPU fallback)");
    return VGAN_B200_ERR_DEVICE;
  }
  if (device < 0 || device >= count) {
    set_error("create: device ordinal %d out of range (%d devices)", device, count);
    return VGAN_B200_ERR_DEVICE;
  }
  int major = 0;
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
  if (major != 10) {
    set_error("device compute capability %d.x is not sm_100 (B200); no other backend exists", major);
    return VGAN_B200_ERR_DEVICE;
  }
  vgan_b200_ctx* ctx = new (std::nothrow) vgan_b200_ctx();
  if (!ctx) {
    set_error("create: out of host memory");
    return VGAN_B200_ERR_ARG;
  }
  ctx->device = device;
  cudaDeviceGetAttribute(&ctx->n_sms, cudaDevAttrMultiProcessorCount, device);
  void* ptr = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
      q == cudaDriverEntryPointSuccess)
    ctx->encode = reinterpret_cast<EncodeTiledFn>(ptr);
  ptr = nullptr;
  if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
      q == cudaDriverEntryPointSuccess)
    ctx->stream_wait = reinterpret_cast<StreamWaitValueFn>(ptr);
  *out = ctx;
  return VGAN_B200_OK;
}

int vgan_b200_destroy(vgan_b200_ctx* ctx) {
  if (!ctx) return VGAN_B200_OK;
  {
    DeviceGuard guard(ctx->device);
    if (ctx->side) cudaStreamDestroy(ctx->side);
    if (ctx->hi) cudaStreamDestroy(ctx->hi);
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
    if (ctx->ev_join_hi) cudaEventDestroy(ctx->ev_join_hi);
  }
  delete ctx;
  return VGAN_B200_OK;
}

int vgan_b200_set_cluster_limit(vgan_b200_ctx* ctx, int max_clusters) {
  if (!ctx || max_clusters < 0) {
    set_error("set_cluster_limit: null context or negative limit");
    return VGAN_B200_ERR_ARG;
  }
  ctx->cluster_limit = max_clusters;
  return VGAN_B200_OK;
}

int vgan_b200_launch_count(vgan_b200_ctx* ctx, unsigned long long* out) {
  if (!ctx || !out) {
    set_error("launch_count: null pointer");
    return VGAN_B200_ERR_ARG;
  }
  *out = ctx->launches;
  return VGAN_B200_OK;
}

int vgan_b200_workspace_layout(int64_t Nx, int64_t Ny, int64_t D, int64_t D_u, int n_scales, int flags,
                               vgan_b200_layout_t* out) {
  return fill_layout(Nx, Ny, D, D_u, n_scales, flags, out);
}

int vgan_b200_stats(vgan_b200_ctx* ctx, const vgan_b200_layout_t* L, void* ws, const float* X, const float* U,
                    const float* Y, int64_t row_begin, int64_t row_end, void* stream) {
  VG_ENTER(ctx);
  if (int e = check_call(L, ws, "stats")) return e;
  if (int e = check_inputs(L, X, U, Y, row_begin, row_end, "stats")) return e;
  return run_stats(*L, ws, X, U, Y, row_begin, row_end, static_cast<cudaStream_t>(stream));
}

int vgan_b200_scales(vgan_b200_ctx* ctx, const vgan_b200_layout_t* L, void* ws, float* bandwidth, int frozen,
                     const float* mults_host, float weight, void* stream) {
  VG_ENTER(ctx);
  if (int e = check_call(L, ws, "scales")) return e;
  if (!bandwidth || !mults_host) {
    set_error("scales: null pointer");
    return VGAN_B200_ERR_ARG;
  }
  return run_scales(*L, ws, bandwidth, frozen, mults_host, weight, static_cast<cudaStream_t>(stream));
}

int vgan_b200_pack(vgan_b200_ctx* ctx, const vgan_b200_layout_t* L, void* ws, const float* X, const float* U,
                   const float* Y, int64_t row_begin, int64_t row_end, void* stream) {
  VG_ENTER(ctx);
  if (int e = check_call(L, ws, "pack")) return e;
  if (int e = check_inputs(L, X, U, Y, row_begin, row_end, "pack")) return e;
  return run_pack(*L, ws, X, U, Y, row_begin, row_end, static_cast<cudaStream_t>(stream));
}

int vgan_b200_mmd2_fwd_bwd(vgan_b200_ctx* ctx, const vgan_b200_layout_t* L, void* ws, const float* X, const float* U,
                           const float* Y, int64_t row_begin, int64_t row_end, int xx_part, int xx_nparts, float* dY,
                           const float* dY_mul, float* dX, void* stream) {
  VG_ENTER(ctx);
  if (int e = check_call(L, ws, "mmd2_fwd_bwd")) return e;
  if (dY)
    if (int e = check_inputs(L, X, U, Y, row_begin, row_end, "mmd2_fwd_bwd")) return e;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (row_begin < 0 || row_begin % kTile || row_end < row_begin || xx_nparts < 1 || xx_part < 0 ||
      xx_part >= xx_nparts) {
    set_error("mmd2_fwd_bwd: bad shard (row_begin %lld must be a multiple of 128, row_end %lld, xx %d/%d)",
              (long long)row_begin, (long long)row_end, xx_part, xx_nparts);
    return VGAN_B200_ERR_ARG;
  }
  const bool need_dx = (L->flags & VGAN_B200_NEED_DX) != 0;
  if (need_dx && !dX) {
    set_error("mmd2_fwd_bwd: the workspace was laid out with NEED_DX but dX is null");
    return VGAN_B200_ERR_ARG;
  }
  if ((need_dx || dY_mul) && !dY) {
    set_error("mmd2_fwd_bwd: dY is null but a gradient output was requested");
    return VGAN_B200_ERR_ARG;
  }
  if ((L->flags & VGAN_B200_PREC_MASK) == VGAN_B200_PREC_FP32)
    return launch_simt_from_abi(*L, ws, row_begin, row_end, xx_part, xx_nparts, dY, dY_mul, dX, st);
  return run_tc(ctx, *L, ws, X, U, Y, row_begin, row_end, xx_part, xx_nparts, dY, dY_mul, dX, st);
}

int vgan_b200_finalize(vgan_b200_ctx* ctx, const vgan_b200_layout_t* L, void* ws, float weight, float* dU,
                       int64_t dU_row_begin, int64_t dU_rows, int colsum_penalty, float* loss, void* stream) {
  VG_ENTER(ctx);
  if (int e = check_call(L, ws, "finalize")) return e;
  return run_finalize(*L, ws, weight, dU, dU_row_begin, dU_rows, colsum_penalty, loss,
                      static_cast<cudaStream_t>(stream));
}

int vgan_b200_mmd2_eval(vgan_b200_ctx* ctx, const vgan_b200_eval_t* a) {
  if (!a || !a->layout) {
    set_error("mmd2_eval: null arguments");
    return VGAN_B200_ERR_ARG;
  }
  const vgan_b200_layout_t* L = a->layout;
  const int64_t n_rows = L->Nx > L->Ny ? L->Nx : L->Ny;
  if (int e = vgan_b200_stats(ctx, L, a->ws, a->X, a->U, a->Y, 0, n_rows, a->stream)) return e;
  if (int e = vgan_b200_scales(ctx, L, a->ws, a->bandwidth, a->frozen, a->mults_host, a->weight, a->stream)) return e;
  if (int e = vgan_b200_pack(ctx, L, a->ws, a->X, a->U, a->Y, 0, n_rows, a->stream)) return e;
  if (int e = vgan_b200_mmd2_fwd_bwd(ctx, L, a->ws, a->X, a->U, a->Y, 0, n_rows, 0, 1, a->dY, a->dY_mul, a->dX, a->stream))
    return e;
  return vgan_b200_finalize(ctx, L, a->ws, a->weight, a->dU_penalty, 0, L->Nx,
                            (a->dU_penalty != nullptr && a->dU_penalty == a->dY) ? 1 : 0, a->loss, a->stream);
}

int vgan_b200_upper_softmax(vgan_b200_ctx* ctx, const float* H, int64_t N, int64_t D, float* U, float* row_max,
                            float* row_inv_sum, void* stream) {
  VG_ENTER(ctx);
  if (!H || !U || !row_max || !row_inv_sum || N < 1 || D < 1) {
    set_error("upper_softmax: null pointer or empty input");
    return VGAN_B200_ERR_ARG;
  }
  return run_upper_softmax(H, N, D, U, row_max, row_inv_sum, static_cast<cudaStream_t>(stream));
}

int vgan_b200_upper_softmax_bwd(vgan_b200_ctx* ctx, const float* H, const float* row_max, const float* row_inv_sum,
                                const float* dU, int64_t N, int64_t D, float* dH, void* stream) {
  VG_ENTER(ctx);
  if (!H || !row_max || !row_inv_sum || !dU || !dH || N < 1 || D < 1) {
    set_error("upper_softmax_bwd: null pointer or empty input");
    return VGAN_B200_ERR_ARG;
  }
  return run_upper_softmax_bwd(H, row_max, row_inv_sum, dU, N, D, dH, static_cast<cudaStream_t>(stream));
}

int vgan_b200_mmd2_matrix(vgan_b200_ctx* ctx, const float* Z, int64_t n, int64_t D, const float* inv_scales,
                          int n_scales, float eps, float* K_out, void* stream) {
  VG_ENTER(ctx);
  if (!Z || !inv_scales || !K_out || n < 1 || D < 1 || n_scales < 1 || n > 65536) {
    set_error("mmd2_matrix: null pointer or bad size (n <= 65536)");
    return VGAN_B200_ERR_ARG;
  }
  return launch_matrix(Z, n, D, inv_scales, n_scales, eps, K_out, static_cast<cudaStream_t>(stream));
}

int vgan_b200_permutation_scratch_bytes(int64_t n, size_t* out) {
  if (!out || n < 1) {
    set_error("permutation_scratch_bytes: bad arguments");
    return VGAN_B200_ERR_ARG;
  }
  *out = perm_scratch_bytes(n);
  return VGAN_B200_OK;
}

int vgan_b200_permutation_stats(vgan_b200_ctx* ctx, const float* K, int64_t n, int64_t n_1, const uint8_t* member,
                                int64_t n_perm, void* scratch, double* stats, void* stream) {
  VG_ENTER(ctx);
  if (!K || !member || !scratch || !stats || n_1 < 2 || n - n_1 < 2 || n > 16384 || n_perm < 1) {
    set_error("permutation_stats: null pointer or bad size (n_1 >= 2, n - n_1 >= 2, n <= 16384, n_perm >= 1)");
    return VGAN_B200_ERR_ARG;
  }
  return run_permutation_stats(K, n, n_1, member, n_perm, scratch, stats, static_cast<cudaStream_t>(stream));
}

int vgan_b200_subspace_scratch_bytes(int64_t n, int64_t D, size_t* out) {
  if (!out || n < 1 || D < 1) {
    set_error("subspace_scratch_bytes: bad arguments");
    return VGAN_B200_ERR_ARG;
  }
  *out = static_cast<size_t>(n) * ((D + 63) / 64) * 8 + static_cast<size_t>(n) * 8;
  return VGAN_B200_OK;
}

int vgan_b200_subspace_dist(vgan_b200_ctx* ctx, const void* U, int in_is_bool, int64_t n, int64_t D, float threshold,
                            void* scratch, uint8_t* unique_out, long long* counts_out, int* n_unique, void* stream) {
  VG_ENTER(ctx);
  if (!U || !scratch || !unique_out || !counts_out || !n_unique || n < 1 || D < 1 || n > 65536) {
    set_error("subspace_dist: null pointer or bad size (1 <= n <= 65536)");
    return VGAN_B200_ERR_ARG;
  }
  return run_subspace_dist(U, in_is_bool, n, D, threshold, scratch, unique_out, counts_out, n_unique,
                           static_cast<cudaStream_t>(stream));
}

int vgan_b200_subspace_dist_static(vgan_b200_ctx* ctx, const void* U, int in_is_bool, int64_t n, int64_t D,
                                   float threshold, uint8_t* unique_out, long long* counts_out, int* n_unique,
                                   void* stream) {
  VG_ENTER(ctx);   // makes the context's device current: the static array's address is asked for on that device
  if (!U || !unique_out || !counts_out || !n_unique || n < 1 || D < 1 || n > 65536) {
    set_error("subspace_dist_static: null pointer or bad size (1 <= n <= 65536)");
    return VGAN_B200_ERR_ARG;
  }
  return run_subspace_dist_static(U, in_is_bool, n, D, threshold, unique_out, counts_out, n_unique,
                                  static_cast<cudaStream_t>(stream));
}

}  // extern "C"
