// C ABI of vgan_b200 (include/vgan_b200.h): argument checking, workspace layout, TMA
// descriptor creation and kernel launches.  No torch types, no host synchronisation.
#include <cstdarg>
#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace vg {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) {
    set_error("no CUDA device available (vgan_b200 has no CPU fallback)");
    return VGAN_B200_ERR_DEVICE;
  }
  int major = 0;
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  if (major != 10) {
    set_error("device compute capability %d.x is not sm_100 (B200); no other backend exists", major);
    return VGAN_B200_ERR_DEVICE;
  }
  return VGAN_B200_OK;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

typedef CUresult (*StreamWaitValueFn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);

static StreamWaitValueFn get_stream_wait() {
  static StreamWaitValueFn fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<StreamWaitValueFn>(ptr);
  }
  return fn;
}

// bf16 [rows, cols] row-major, box = 64 columns x 128 rows, 128-byte swizzle
int make_tmap_bf16_2d(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows) {
  EncodeTiledFn enc = get_encode();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled not available from the driver");
    return VGAN_B200_ERR_CUDA;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(kChunk), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box,
                   estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d", static_cast<int>(r));
    return VGAN_B200_ERR_CUDA;
  }
  return VGAN_B200_OK;
}

// launchers defined in the kernel files
int run_prologue(const float*, const float*, const float*, const vgan_b200_layout_t&, int, void*, cudaStream_t);
int run_scales(const vgan_b200_layout_t&, void*, float*, int, const float*, int, float*, cudaStream_t);
int run_setup(const vgan_b200_layout_t&, void*, const float*, int, float*, cudaStream_t);
int run_penalty(const float*, int64_t, int64_t, float, unsigned long long*, float*, float*, int64_t, int64_t,
                cudaStream_t);
int run_finalize(const double*, int64_t, int64_t, const float*, float*, cudaStream_t);
int launch_matrix(const float*, int64_t, int64_t, const float*, int, float, float*, cudaStream_t);
int run_subspace_dist(const void*, int, int64_t, int64_t, float, void*, uint8_t*, long long*, int*, cudaStream_t);
size_t perm_scratch_bytes(int64_t n);
int run_permutation_stats(const float*, int64_t, int64_t, const uint8_t*, int64_t, void*, double*, cudaStream_t);
struct SimtParams;
int launch_rate_probe(const CUtensorMap& zmap, int n, int iters, int b_mn, int tma_warps, int tma_tiles, int grid,
                      long long* out, cudaStream_t st);
int launch_tma_rate(const CUtensorMap& map, int tiles, int kchunks, int stages, int loads, int box_rows, int iters,
                    int mode, int grid, int nprod, long long* cycles, cudaStream_t st);
int launch_pair_probe(const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, float*, float*,
                      cudaStream_t);
int launch_simt_from_abi(const vgan_b200_layout_t& L, void* ws, int64_t row_begin, int64_t row_end, int xx_part,
                         int xx_nparts, const float* inv_scales, int n_scales, int flags, double* sums,
                         float* dY, const float* dY_mul, float* dX, cudaStream_t st);

extern int g_pair_variant;
static unsigned long long* g_timeline = nullptr;
static float* g_dbg = nullptr;
static int g_dbg_stride = 0;

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static int fill_layout(int64_t Nx, int64_t Ny, int64_t D, int64_t D_u, int flags, vgan_b200_layout_t* L) {
  if (!L || Nx < 1 || Ny < 1 || D < 1 || D_u < 0) {
    set_error("workspace_layout: need Nx, Ny, D >= 1 (got %lld, %lld, %lld)", (long long)Nx, (long long)Ny,
              (long long)D);
    return VGAN_B200_ERR_ARG;
  }
  if (Nx + Ny > (1ll << 30) || D > (1 << 20)) {
    set_error("workspace_layout: sizes out of range");
    return VGAN_B200_ERR_ARG;
  }
  memset(L, 0, sizeof(*L));
  L->Nx = Nx; L->Ny = Ny; L->D = D;
  L->Dp = (D + kChunk - 1) / kChunk * kChunk;
  // bf16, more than one feature slice: an even chunk count, so that every slice is 2 or 4 chunks wide and the
  // CTA-pair kernel (each CTA of a pair holds half a slice) serves every width; at most one extra zero chunk
  if ((flags & VGAN_B200_PREC_MASK) != VGAN_B200_PREC_FP32 && L->Dp > kSliceMax)
    L->Dp = (D + 2 * kChunk - 1) / (2 * kChunk) * (2 * kChunk);
  L->Nxp = (Nx + kTile - 1) / kTile * kTile;
  L->Nyp = (Ny + kTile - 1) / kTile * kTile;
  const size_t rows = static_cast<size_t>(L->Nxp + L->Nyp);
  const bool fp32 = (flags & VGAN_B200_PREC_MASK) == VGAN_B200_PREC_FP32;
  // small buffers first so that their offsets depend on (Nx, Ny, D) only; the operand next;
  // the penalty scratch (the only part that depends on D_u and precision) last.
  size_t off = 0;
  L->off_par = off;   off = align_up(off + kParFloats * 4, 1024);
  L->off_stats = off; off = align_up(off + (2 + L->Dp) * 8, 1024);
  L->off_mean = off;  off = align_up(off + L->Dp * 4, 1024);
  L->off_nrm = off;   off = align_up(off + rows * 4, 1024);
  L->off_nrms = off;  off = align_up(off + rows * 4, 1024);
  L->off_xchg = off;  // before the operand: must not depend on D_u (callers pass D_u = 0 to recompute offsets)
  if (!fp32 && L->Dp > kSliceMax) off = align_up(off + static_cast<size_t>(kXchgSms) * kXchgBytesPerSm, 1024);
  L->off_zq = off;
  const bool split = !fp32 && (flags & VGAN_B200_BF16X3);
  L->off_zf = off;    off = align_up(off + rows * L->Dp * (fp32 ? 4 : (split ? 12 : 2)), 1024);
  L->off_cmax = off;  off = align_up(off + static_cast<size_t>(D_u > 0 ? D_u : 1) * 8, 1024);

  L->total = off;
  return VGAN_B200_OK;
}

}  // namespace vg

using namespace vg;

extern "C" {

int vgan_b200_abi_version(void) { return VGAN_B200_ABI_VERSION; }
const char* vgan_b200_last_error(void) { return g_err; }

int vgan_b200_workspace_layout(int64_t Nx, int64_t Ny, int64_t D, int64_t D_u, int flags,
                               vgan_b200_layout_t* out) {
  return fill_layout(Nx, Ny, D, D_u, flags, out);
}

int vgan_b200_prologue(const float* X, const float* U, const float* Y, int64_t Nx, int64_t Ny, int64_t D,
                       int flags, void* ws, void* stream) {
  if (int e = check_device()) return e;
  vgan_b200_layout_t L;
  if (int e = fill_layout(Nx, Ny, D, 0, flags, &L)) return e;
  if (!X || !ws || (!U && !Y)) {
    set_error("prologue: X, ws and one of U / Y must be non-null");
    return VGAN_B200_ERR_ARG;
  }
  if (!Y && Nx != Ny) {
    set_error("prologue: masking Y = U*X needs Ny == Nx (got %lld vs %lld)", (long long)Ny, (long long)Nx);
    return VGAN_B200_ERR_ARG;
  }
  if (reinterpret_cast<uintptr_t>(ws) % 1024) {
    set_error("prologue: workspace must be 1024-byte aligned");
    return VGAN_B200_ERR_ARG;
  }
  return run_prologue(X, U, Y, L, flags, ws, static_cast<cudaStream_t>(stream));
}

int vgan_b200_scales(void* ws, int64_t Nx, int64_t Ny, int64_t D, float* bandwidth, int frozen,
                     const float* mults_host, int n_scales, float* inv_scales, void* stream) {
  if (int e = check_device()) return e;
  vgan_b200_layout_t L;
  if (int e = fill_layout(Nx, Ny, D, 0, 0, &L)) return e;
  if (!ws || !bandwidth || !mults_host || !inv_scales || n_scales < 1 || n_scales > VGAN_B200_MAX_SCALES) {
    set_error("scales: null pointer or n_scales outside [1, %d]", VGAN_B200_MAX_SCALES);
    return VGAN_B200_ERR_ARG;
  }
  return run_scales(L, ws, bandwidth, frozen, mults_host, n_scales, inv_scales, static_cast<cudaStream_t>(stream));
}

}  // extern "C"

extern "C" {

int vgan_b200_mmd2_fwd_bwd(void* ws, int64_t Nx, int64_t Ny, int64_t D, int64_t row_begin, int64_t row_end,
                           int xx_part, int xx_nparts, const float* inv_scales, int n_scales, int flags,
                           double* sums, float* dY, const float* dY_mul, float* dX, void* stream) {
  if (int e = check_device()) return e;
  vgan_b200_layout_t L;
  if (int e = fill_layout(Nx, Ny, D, 0, flags, &L)) return e;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (!ws || !inv_scales || !sums || n_scales < 1 || n_scales > VGAN_B200_MAX_SCALES) {
    set_error("mmd2_fwd_bwd: null pointer or n_scales outside [1, %d]", VGAN_B200_MAX_SCALES);
    return VGAN_B200_ERR_ARG;
  }
  if (row_begin < 0 || row_begin % kTile || row_end <= row_begin || xx_nparts < 1 || xx_part < 0 ||
      xx_part >= xx_nparts) {
    set_error("mmd2_fwd_bwd: bad shard (row_begin %lld must be a multiple of 128, row_end %lld, xx %d/%d)",
              (long long)row_begin, (long long)row_end, xx_part, xx_nparts);
    return VGAN_B200_ERR_ARG;
  }
  const bool need_dx = (flags & VGAN_B200_NEED_DX) != 0;
  if (need_dx && !dX) {
    set_error("mmd2_fwd_bwd: NEED_DX set but dX is null");
    return VGAN_B200_ERR_ARG;
  }
  if ((need_dx || dY_mul) && !dY) {
    set_error("mmd2_fwd_bwd: dY is null but a gradient output was requested");
    return VGAN_B200_ERR_ARG;
  }
  if ((flags & VGAN_B200_LADDER) && n_scales != 5) {
    set_error("mmd2_fwd_bwd: LADDER needs exactly 5 scales");
    return VGAN_B200_ERR_ARG;
  }
  uint8_t* base = static_cast<uint8_t*>(ws);
  if ((flags & VGAN_B200_PREC_MASK) == VGAN_B200_PREC_FP32)
    return launch_simt_from_abi(L, ws, row_begin, row_end, xx_part, xx_nparts, inv_scales, n_scales, flags,
                                sums, dY, dY_mul, dX, st);

  float* par = reinterpret_cast<float*>(base + L.off_par);
  if (int e = run_setup(L, ws, inv_scales, n_scales, par, st)) return e;
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
  p.n_scales = n_scales;
  p.Nx = Nx; p.Ny = Ny; p.D = D; p.Dp = L.Dp;
  p.row_begin = row_begin; p.row_end = row_end;
  p.zq = reinterpret_cast<const __nv_bfloat16*>(base + L.off_zq);
  p.nrms = reinterpret_cast<const float*>(base + L.off_nrms);
  p.par = par;
  p.sums = sums; p.dY = dY; p.dY_mul = dY_mul; p.dX = dX;
  p.xchg = L.Dp > kSliceMax ? base + L.off_xchg : nullptr;
  p.dbg = g_dbg; p.dbg_stride = g_dbg_stride;
  p.timeline = g_timeline;
  p.started = nullptr;
  p.dbg_flags = g_pair_variant >> 8;
  // ---- operand arrays and their tensor maps.  Plain bf16: one array [rows, Dp] serves as row operand, column operand
  // and V.  BF16X3: row operand [hi | hi | lo] and column operand [hi | lo | hi], both [rows, 3 Dp]; V = the hi or lo
  // block of the row-operand array.
  const bool split = (flags & VGAN_B200_BF16X3) != 0;
  const uint64_t rows = static_cast<uint64_t>(L.Nxp + L.Nyp);
  const uint64_t width = static_cast<uint64_t>(split ? 3 * L.Dp : L.Dp);
  const __nv_bfloat16* zb = split ? p.zq + rows * width : p.zq;
  p.gchunks = split ? 3 * p.kchunks : p.kchunks;
  p.zq_stride = static_cast<int64_t>(width);
  p.zq_lo = split ? p.zq + 2 * L.Dp : nullptr;
  const int ladder = (flags & VGAN_B200_LADDER) ? 1 : 0;
  p.cs = p.nslices < 4 ? p.nslices : 4;
  p.ngroups = (p.nslices + p.cs - 1) / p.cs;
  const bool multi_slice = p.pv && p.nslices >= 2 && !(flags & VGAN_B200_NO_CLUSTER);
  if (split && multi_slice && (flags & VGAN_B200_NO_PAIR)) {
    set_error("mmd2_fwd_bwd: BF16X3 is not available on the NO_PAIR fallback kernel");
    return VGAN_B200_ERR_UNSUPPORTED;
  }
  CUtensorMap map, mapB, mapV, mapB64, mapV64, mapB256;
  if (int e = make_tmap_bf16_2d(&map, p.zq, rows, width)) return e;
  if (int e = make_tmap_bf16_2d(&mapB, zb, rows, width)) return e;
  mapV = map;
  if (multi_slice) {
    if (int e = make_tmap_bf16_2d(&mapB64, zb, rows, width, 64)) return e;
    if (int e = make_tmap_bf16_2d(&mapV64, p.zq, rows, width, 64)) return e;
    if (p.n_xx_units > 0)
      if (int e = make_tmap_bf16_2d(&mapB256, zb, rows, width, 256)) return e;
  }

  // one pass of the tcgen05 kernels over this shard (BF16X3 runs three: see below)
  auto run_pass = [&](TcParams& p, bool zero_outputs) -> int {
  if (p.pv && p.nslices >= 2 && !(flags & VGAN_B200_NO_CLUSTER)) {
    // D_p > 256: the slice CTAs of two row tiles run as one thread-block cluster (mmd_tc_pair.cu) and share every
    // P tile; the S_xx-only units (no gradient GEMM) run as their own sums-only kernel (mmd_xx.cu) on a low-priority
    // side stream, the cluster kernel on a high-priority one.  8-CTA clusters occupy 120 of the 148 SMs; the S_xx
    // kernel gets the other 28.  The order is enforced: the persistent CTAs count themselves in at their first
    // instruction, the side stream holds the S_xx launch (cuStreamWaitValue32, no SM occupied) until all of them are
    // resident, and the S_xx grid is sized to the SMs they leave free.  (Launched freely, the S_xx CTAs were
    // dispatched first and the clusters started 1-9 ms late: tools/gpu_timeline.py.)
    // internal streams and events, one set per device, created on first use and kept (see the header: calls that
    // share a device must come from one host thread at a time)
    struct Aux { cudaStream_t side, hi; cudaEvent_t ev_fork, ev_join, ev_join_hi; };
    static Aux aux_of_device[64] = {};
    int dev_id = 0;
    VG_CUDA_CHECK(cudaGetDevice(&dev_id));
    if (dev_id < 0 || dev_id >= 64) {
      set_error("mmd2_fwd_bwd: device ordinal %d out of range", dev_id);
      return VGAN_B200_ERR_DEVICE;
    }
    Aux& aux = aux_of_device[dev_id];
    cudaStream_t& side = aux.side;
    cudaStream_t& hi = aux.hi;
    cudaEvent_t& ev_fork = aux.ev_fork;
    cudaEvent_t& ev_join = aux.ev_join;
    cudaEvent_t& ev_join_hi = aux.ev_join_hi;
    static const int xx_mode = getenv("VGAN_B200_XX_MODE") ? atoi(getenv("VGAN_B200_XX_MODE")) : 0;   // 1: serial (measurement)
    const bool overlap = p.n_xx_units > 0 && xx_mode != 1;
    if (overlap && !side) {
      int lo_pri = 0, hi_pri = 0;
      VG_CUDA_CHECK(cudaDeviceGetStreamPriorityRange(&lo_pri, &hi_pri));
      VG_CUDA_CHECK(cudaStreamCreateWithPriority(&side, cudaStreamNonBlocking, lo_pri));
      VG_CUDA_CHECK(cudaStreamCreateWithPriority(&hi, cudaStreamNonBlocking, hi_pri));
      VG_CUDA_CHECK(cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming));
      VG_CUDA_CHECK(cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming));
      VG_CUDA_CHECK(cudaEventCreateWithFlags(&ev_join_hi, cudaEventDisableTiming));
    }
    cudaStream_t main_st = st;
    const bool pair_ok = (p.kchunks % 4 == 0 || p.kchunks % 4 == 2) && !(flags & VGAN_B200_NO_PAIR);
    if (pair_ok && zero_outputs) {
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
        (overlap && pair_ok && xx_mode == 0 && capturing == cudaStreamCaptureStatusNone) ? get_stream_wait() : nullptr;
    if (stream_wait) {
      p.started = reinterpret_cast<unsigned*>(base + L.off_par + kParStartedOffset);
      VG_CUDA_CHECK(cudaMemsetAsync(p.started, 0, sizeof(unsigned), st));
    }
    int resident_ctas = 0;
    if (overlap) {
      VG_CUDA_CHECK(cudaEventRecord(ev_fork, st));
      VG_CUDA_CHECK(cudaStreamWaitEvent(hi, ev_fork, 0));
      main_st = hi;
      if (xx_mode == 2) {      // measurement: S_xx kernel dispatched first
        VG_CUDA_CHECK(cudaStreamWaitEvent(side, ev_fork, 0));
        if (int e = launch_xx(p, map, mapB256, ladder, side)) return e;
      }
    }
    // full 128/256-wide slices: CTA pairs (cta_group::2) halve the shared-memory operand traffic
    if (pair_ok) {
      if (int e = launch_tc_pair(p, map, mapB64, mapV64, ladder, main_st, &resident_ctas)) return e;
    } else {
      if (int e = launch_tc_cluster(p, map, ladder, main_st)) return e;
    }
    if (p.n_xx_units == 0) return VGAN_B200_OK;
    if (!overlap) return launch_xx(p, map, mapB256, ladder, st);
    if (xx_mode != 2) {
      VG_CUDA_CHECK(cudaStreamWaitEvent(side, ev_fork, 0));
      int free_sms = 0;
      if (stream_wait && resident_ctas > 0) {
        int dev = 0, n_sms = 0;
        VG_CUDA_CHECK(cudaGetDevice(&dev));
        VG_CUDA_CHECK(cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev));
        free_sms = n_sms - resident_ctas;
      }
      if (free_sms >= kXxMinFreeSms) {
        CUresult r = stream_wait(reinterpret_cast<CUstream>(side), reinterpret_cast<CUdeviceptr>(p.started),
                                 static_cast<cuuint32_t>(resident_ctas), CU_STREAM_WAIT_VALUE_GEQ);
        if (r != CUDA_SUCCESS) free_sms = 0;   // stream memory operations unavailable: plain low-priority launch
      } else {
        free_sms = 0;       // no room beside the clusters: plain low-priority launch over all SMs
      }
      if (int e = launch_xx(p, map, mapB256, ladder, side, free_sms)) return e;
    }
    VG_CUDA_CHECK(cudaEventRecord(ev_join, side));
    VG_CUDA_CHECK(cudaEventRecord(ev_join_hi, hi));
    VG_CUDA_CHECK(cudaStreamWaitEvent(st, ev_join, 0));
    VG_CUDA_CHECK(cudaStreamWaitEvent(st, ev_join_hi, 0));
    return VGAN_B200_OK;
  }
  return launch_tc(p, map, mapB, mapV, ladder, st);
  };
  if (!split) return run_pass(p, true);
  // BF16X3: O = P_hi V_hi + P_hi V_lo + P_lo V_hi, one pass each into zeroed outputs; pass 1 also produces the kernel
  // sums and the rowsum(P) z_i term, the S_xx units run once beside it
  if (p.pv) {
    if (ye > row_begin)
      VG_CUDA_CHECK(cudaMemsetAsync(dY, 0, static_cast<size_t>(ye - row_begin) * D * sizeof(float), st));
    if (need_dx && xe > row_begin)
      VG_CUDA_CHECK(cudaMemsetAsync(dX, 0, static_cast<size_t>(xe - row_begin) * D * sizeof(float), st));
  }
  p.pass_flags = kPassAccumulate;
  p.v_chunk0 = 0; p.p_lo = 0;
  if (int e = run_pass(p, false)) return e;
  if (!p.pv) return VGAN_B200_OK;
  p.n_xx_units = 0;
  p.pass_flags = kPassAccumulate | kPassNoSums | kPassNoRowsum;
  p.v_chunk0 = 2 * p.kchunks; p.p_lo = 0;
  if (int e = run_pass(p, false)) return e;
  p.v_chunk0 = 0; p.p_lo = 1;
  return run_pass(p, false);
}

int vgan_b200_colmax_penalty(const float* U, int64_t N, int64_t D_u, float weight, void* scratch,
                             float* penalty, float* dU, int64_t dU_row_begin, int64_t dU_rows, void* stream) {
  if (int e = check_device()) return e;
  if (!U || !scratch || !penalty || N < 1 || D_u < 1) {
    set_error("colmax_penalty: null pointer or empty U");
    return VGAN_B200_ERR_ARG;
  }
  return run_penalty(U, N, D_u, weight, static_cast<unsigned long long*>(scratch), penalty, dU, dU_row_begin,
                     dU_rows, static_cast<cudaStream_t>(stream));
}

int vgan_b200_finalize_loss(const double* sums, int64_t Nx, int64_t Ny, const float* penalty, float* loss,
                            void* stream) {
  if (int e = check_device()) return e;
  if (!sums || !loss || Nx < 1 || Ny < 1) {
    set_error("finalize_loss: null pointer");
    return VGAN_B200_ERR_ARG;
  }
  return run_finalize(sums, Nx, Ny, penalty, loss, static_cast<cudaStream_t>(stream));
}

int vgan_b200_mmd2_matrix(const float* Z, int64_t n, int64_t D, const float* inv_scales, int n_scales,
                          float eps, float* K_out, void* stream) {
  if (int e = check_device()) return e;
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

int vgan_b200_permutation_stats(const float* K, int64_t n, int64_t n_1, const uint8_t* member, int64_t n_perm,
                                void* scratch, double* stats, void* stream) {
  if (int e = check_device()) return e;
  if (!K || !member || !scratch || !stats || n_1 < 2 || n - n_1 < 2 || n > 200000 || n_perm < 1) {
    set_error("permutation_stats: null pointer or bad size (n_1 >= 2, n - n_1 >= 2, n <= 200000, n_perm >= 1)");
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

int vgan_b200_subspace_dist(const void* U, int in_is_bool, int64_t n, int64_t D, float threshold, void* scratch,
                            uint8_t* unique_out, long long* counts_out, int* n_unique, void* stream) {
  if (int e = check_device()) return e;
  if (!U || !scratch || !unique_out || !counts_out || !n_unique || n < 1 || D < 1 || n > (1 << 20)) {
    set_error("subspace_dist: null pointer or bad size (1 <= n <= 2^20)");
    return VGAN_B200_ERR_ARG;
  }
  return run_subspace_dist(U, in_is_bool, n, D, threshold, scratch, unique_out, counts_out, n_unique,
                           static_cast<cudaStream_t>(stream));
}

int vgan_b200_debug_tile_sums(float* buf, int stride) {
  g_dbg = buf;
  g_dbg_stride = stride;
  return VGAN_B200_OK;
}

int vgan_b200_debug_timeline(unsigned long long* buf) {
  g_timeline = buf;
  return VGAN_B200_OK;
}

int vgan_b200_debug_variant(int v) {
  g_pair_variant = v;
  return VGAN_B200_OK;
}

int vgan_b200_pair_probe(const void* A0, const void* A1, const void* B, const void* V, float* S_out, float* O_out,
                         void* stream) {
  if (int e = check_device()) return e;
  CUtensorMap a0, a1, bh, v;
  if (int e = make_tmap_bf16_2d(&a0, A0, 128, 128)) return e;
  if (int e = make_tmap_bf16_2d(&a1, A1, 128, 128)) return e;
  if (int e = make_tmap_bf16_2d(&bh, B, 128, 128, 64)) return e;
  if (int e = make_tmap_bf16_2d(&v, V, 128, 256)) return e;
  return launch_pair_probe(a0, a1, bh, v, S_out, O_out, static_cast<cudaStream_t>(stream));
}

int vgan_b200_tma_rate_probe(const void* Z, int64_t rows, int64_t cols, int stages, int loads, int box_rows,
                             int iters, int mode, int grid, int nprod, long long* cycles_out, void* stream) {
  if (int e = check_device()) return e;
  if (!Z || !cycles_out || rows % kTile || cols % kChunk || (box_rows != 64 && box_rows != 128)) {
    set_error("tma_rate_probe: bad arguments");
    return VGAN_B200_ERR_ARG;
  }
  CUtensorMap map;
  if (int e = make_tmap_bf16_2d(&map, Z, static_cast<uint64_t>(rows), static_cast<uint64_t>(cols), box_rows)) return e;
  return launch_tma_rate(map, static_cast<int>(rows / kTile), static_cast<int>(cols / kChunk), stages, loads, box_rows,
                         iters, mode, grid, nprod < 1 ? 1 : (nprod > 8 ? 8 : nprod), cycles_out,
                         static_cast<cudaStream_t>(stream));
}

int vgan_b200_umma_rate_probe(int n, int iters, int b_mn, const void* Z, int64_t rows, int tma_warps, int grid,
                              long long* out, void* stream) {
  if (int e = check_device()) return e;
  if (!Z || rows % kTile || tma_warps < 0 || tma_warps > 8 || grid < 1 || !out) {
    set_error("umma_rate_probe: bad arguments");
    return VGAN_B200_ERR_ARG;
  }
  CUtensorMap map;
  if (int e = make_tmap_bf16_2d(&map, Z, static_cast<uint64_t>(rows), 1024)) return e;
  return launch_rate_probe(map, n, iters, b_mn, tma_warps, static_cast<int>(rows / kTile), grid, out,
                           static_cast<cudaStream_t>(stream));
}

int vgan_b200_umma_probe(const void* A, const void* B, float* S_out, float* O_out, void* stream) {
  if (int e = check_device()) return e;
  if (!A || !B || !S_out || !O_out) {
    set_error("umma_probe: null pointer");
    return VGAN_B200_ERR_ARG;
  }
  CUtensorMap ma, mb;
  if (int e = make_tmap_bf16_2d(&ma, A, 128, 128)) return e;
  if (int e = make_tmap_bf16_2d(&mb, B, 128, 128)) return e;
  return launch_probe(ma, mb, S_out, O_out, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
