/* vgan_b200 — C ABI of the B200-native (sm_100a) squared-MMD forward+backward path.
 *
 * Drop-in boundary for V-GAN's loss module.  The reference has no FFI (it is pure
 * Python/torch); the calls these entry points replace are
 *   RBF.get_bandwidth               /root/reference/src/models/Mmd_loss_constrained.py:16-22
 *   RBF.forward                     /root/reference/src/models/Mmd_loss_constrained.py:24-26
 *   MMDLossConstrained.forward      /root/reference/src/models/Mmd_loss_constrained.py:42-50
 *   its autograd backward (EuclideanDistBackward0 ... MulBackward)  — SURVEY.md 3.3
 *   the masking  Y = U * X          /root/reference/src/vgan.py:616 (also :275,:312)
 *   tts.MMDStatistic(...)(x, y, alphas, ret_matrix=True)            /root/reference/src/vgan.py:419-428
 *
 * Conventions: extern "C"; every function returns 0 on success or a negative
 * vgan_b200_status; vgan_b200_last_error() gives the text (thread local).  All data
 * pointers are DEVICE pointers borrowed for the call unless a name ends in _host; all
 * work is enqueued on `stream` (a cudaStream_t passed as void*); nothing synchronises
 * the host.  No CPU fallback exists: without an sm_100 device every call fails.
 * Concurrency: one host thread at a time per device (vgan_b200_mmd2_fwd_bwd forks onto two internal streams per
 * device for D > 256 and joins them back into `stream` before it returns); different devices are independent —
 * the intended deployment is one process per GPU.
 */
#ifndef VGAN_B200_H
#define VGAN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VGAN_B200_ABI_VERSION 1
#define VGAN_B200_MAX_SCALES 8

typedef enum {
  VGAN_B200_OK = 0,
  VGAN_B200_ERR_ARG = -1,       /* bad shape / null pointer / misalignment          */
  VGAN_B200_ERR_DEVICE = -2,    /* no CUDA device, or not compute capability 10.x   */
  VGAN_B200_ERR_CUDA = -3,      /* a CUDA runtime/driver call failed                */
  VGAN_B200_ERR_UNSUPPORTED = -4
} vgan_b200_status;

/* flags for vgan_b200_prologue / vgan_b200_mmd2_fwd_bwd */
enum {
  VGAN_B200_PREC_BF16 = 0,      /* bf16 operands, fp32 accumulate, tcgen05 path (2e-3) */
  VGAN_B200_PREC_FP32 = 1,      /* fp32 CUDA-core path (1e-4)                           */
  VGAN_B200_PREC_MASK = 1,
  VGAN_B200_CENTER    = 2,      /* subtract the stacked column mean (distance invariant) */
  VGAN_B200_NEED_DX   = 4,      /* also produce dL/dX (VGAN.fit, vgan.py:283)           */
  VGAN_B200_LADDER    = 8,      /* scales are s, s/2, s/4, s/8, s/16: 1 ex2 + 4 squarings */
  VGAN_B200_BF16X3    = 64,     /* with PREC_BF16: fp32-class results on the tensor cores.  Every operand value is split
                                   z = hi + lo (two bf16), the Gram is hi.hi + hi.lo + lo.hi in ONE K = 3 D GEMM, the
                                   gradient GEMM runs three passes (P_hi V_hi, P_hi V_lo, P_lo V_hi): ~6x the bf16 cost */
  VGAN_B200_NO_PAIR   = 32,     /* D > 256: 1-CTA cluster kernel instead of the cta_group::2 pair kernel */
  VGAN_B200_NO_CLUSTER = 16     /* D > 256: independent slice CTAs (Gram recomputed) instead of the cluster kernel */
};

/* Layout of the caller-provided workspace (all offsets in bytes from `ws`).
 * Z is the stacked sample [X ; Y] (Mmd_loss_constrained.py:43): X rows at [0,Nx),
 * zero rows up to Nxp (multiple of 128), Y rows at [Nxp, Nxp+Ny), zero rows up to
 * Nxp+Nyp; Dp = D rounded up to 64 with zero columns (to 128 on the bf16 path when D > 256). */
typedef struct {
  int64_t Nx, Ny, D;
  int64_t Dp, Nxp, Nyp;
  size_t off_zq;      /* bf16  [Nxp+Nyp, Dp]  (PREC_BF16); with BF16X3 two arrays [Nxp+Nyp, 3 Dp] back to back:
                         row operand [hi | hi | lo], then column operand [hi | lo | hi]                 */
  size_t off_zf;      /* float [Nxp+Nyp, Dp]  (PREC_FP32; same offset, one of the two exists) */
  size_t off_nrm;     /* float [Nxp+Nyp]  |z_i|^2 of the stored (rounded) rows; 1e30 on pads */
  size_t off_nrms;    /* float [Nxp+Nyp]  the same, scaled for the exp2 argument             */
  size_t off_mean;    /* float [Dp]   stacked column mean (zeros unless CENTER)              */
  size_t off_stats;   /* double[2+Dp] sum|z|^2, spare, column sums of the stored fp32 rows   */
  size_t off_cmax;    /* uint64[D_u]  packed (value,row) column maxima of U                   */
  size_t off_par;     /* float[32]    device parameter block of K1 (exp2 scale, coefficients) */
  size_t off_xchg;    /* bytes: P-tile exchange scratch of the D_p > 256 kernels (L2 resident), else empty  */
  size_t total;       /* bytes to allocate                                                   */
} vgan_b200_layout_t;

int vgan_b200_abi_version(void);
const char* vgan_b200_last_error(void);

/* Fills *out for the given sizes and precision (flags & PREC_MASK). D_u = columns of U. */
int vgan_b200_workspace_layout(int64_t Nx, int64_t Ny, int64_t D, int64_t D_u, int flags,
                               vgan_b200_layout_t* out);

/* K2 prologue.  Builds the stacked operand in `ws` from X [Nx,D] and either Y [Ny,D]
 * or, when Y == NULL, the mask U [Nx,D] (then Ny must equal Nx and Y = U*X is formed on
 * the fly — vgan.py:616).  Also leaves sum|z|^2 and the column sums for the closed-form
 * bandwidth (Mmd_loss_constrained.py:16-22).  fp32 inputs, row stride D. */
int vgan_b200_prologue(const float* X, const float* U, const float* Y,
                       int64_t Nx, int64_t Ny, int64_t D, int flags, void* ws, void* stream);

/* Bandwidth + scales.  If frozen == 0: *bandwidth (device float) = sum_ij d2 / (n^2-n)
 * from the prologue statistics, else *bandwidth is read.  Then
 * inv_scales[m] = 1 / fp32(bandwidth * mults_host[m]) (the reference's rounding,
 * Mmd_loss_constrained.py:26) for m < n_scales, all on the device, no host sync. */
int vgan_b200_scales(void* ws, int64_t Nx, int64_t Ny, int64_t D, float* bandwidth, int frozen,
                     const float* mults_host, int n_scales, float* inv_scales, void* stream);

/* K1: fused forward+backward over the pairwise matrix, which never reaches HBM.
 *   sums[0..2] += S_xx, S_xy, S_yy of this shard (double, atomically accumulated; the
 *                 caller zeroes them);  S_xx is taken from X-row tiles t with
 *                 t % xx_nparts == xx_part, S_xy/S_yy from Y rows [row_begin,row_end).
 *   dY[i - row_begin, :] = dL/dy_i  (x dY_mul[i - row_begin, :] when dY_mul != NULL,
 *                 which yields dL/dU = dL/dY * X directly) for i in [row_begin,row_end);
 *                 row stride D.  NULL skips the gradient GEMM (forward only).
 *   dX (NEED_DX): dL/dx_i for the same row range of X.
 * row_begin must be a multiple of 128 (row_end need not be). */
int vgan_b200_mmd2_fwd_bwd(void* ws, int64_t Nx, int64_t Ny, int64_t D,
                           int64_t row_begin, int64_t row_end, int xx_part, int xx_nparts,
                           const float* inv_scales, int n_scales, int flags,
                           double* sums, float* dY, const float* dY_mul, float* dX,
                           void* stream);

/* K3: weight * mean_d(1 - max_i U[i,d]) added to *penalty (device float) and its
 * gradient -weight/D_u added at (argmax row, d), lowest row on ties
 * (Mmd_loss_constrained.py:50).  dU may be NULL. */
int vgan_b200_colmax_penalty(const float* U, int64_t N, int64_t D_u, float weight,
                             void* scratch /* >= 8*D_u bytes, e.g. ws + off_cmax */,
                             float* penalty, float* dU, int64_t dU_row_begin, int64_t dU_rows,
                             void* stream);

/* loss[0] = sums[0]/Nx^2 - 2 sums[1]/(Nx Ny) + sums[2]/Ny^2 + penalty[0]  (device). */
int vgan_b200_finalize_loss(const double* sums, int64_t Nx, int64_t Ny, const float* penalty,
                            float* loss, void* stream);

/* Dense kernel matrix K[n,n] = sum_m exp(-(d2 + eps) * inv_scales[m]) for small n
 * (RBF.forward parity; MMDStatistic(..., ret_matrix=True) with inv_scales = alphas,
 * eps = 1e-5).  Z fp32 [n,D]. */
int vgan_b200_mmd2_matrix(const float* Z, int64_t n, int64_t D, const float* inv_scales,
                          int n_scales, float eps, float* K_out, void* stream);

/* Permutation test of the unbiased MMD^2 statistic (torch_two_sample MMDStatistic.pval ->
 * permutation_test.pyx::permutation_test_mat, reached from vgan.py:425-428 / 702-705).  K float [n,n] kernel
 * matrix (vgan_b200_mmd2_matrix), member uint8 [n_perm, n]: row p marks with 1 the n_1 points that relabeling p
 * assigns to sample 1 (row with the first n_1 ones = the observed labels).  stats double [n_perm] receives
 * a00 (sum K11 - tr K11) + a11 (sum K22 - tr K22) + 2 a01 sum K12 of every relabeling; the p-value is the fraction
 * of rows whose statistic is >= the observed one.  scratch >= vgan_b200_permutation_scratch_bytes(n). */
int vgan_b200_permutation_scratch_bytes(int64_t n, size_t* out);
int vgan_b200_permutation_stats(const float* K, int64_t n, int64_t n_1, const uint8_t* member, int64_t n_perm,
                                void* scratch, double* stats, void* stream);

/* Subspace distribution (vgan.py:372-382 / 649-659): unique rows of the boolean mask matrix
 * (U[i,d] >= threshold for float U, U[i,d] != 0 when in_is_bool) in np.unique(axis=0) order, and how often each
 * occurs.  unique_out uint8 [n, D] (the first *n_unique rows are written), counts_out int64 [n], n_unique device
 * int; scratch >= vgan_b200_subspace_scratch_bytes(n, D).  Replaces generate_subspaces' threshold
 * (vgan.py:369 / 646) + np.unique on the host. */
int vgan_b200_subspace_scratch_bytes(int64_t n, int64_t D, size_t* out);
int vgan_b200_subspace_dist(const void* U, int in_is_bool, int64_t n, int64_t D, float threshold, void* scratch,
                            uint8_t* unique_out, long long* counts_out, int* n_unique, void* stream);

/* Debug aid: one Gram tile and one gradient tile through the tcgen05 pipeline with raw
 * outputs, used by tests to localise descriptor/layout errors. A,B bf16 [128,128] row-major;
 * S_out float [128,128] = A B^T;  O_out float [128,128] = bf16(S) B. */
/* Debug aid: when buf != NULL, the next bf16 mmd2_fwd_bwd calls add the kernel sum of every
 * (work unit, column tile) to buf[unit * stride + tile] (device floats).  NULL switches it off. */
int vgan_b200_debug_tile_sums(float* buf, int stride);

/* Measurement aid: when buf != NULL (4 device uint64, caller presets {~0, 0, ~0, 0}) the D > 256 kernels record the
 * %globaltimer (ns) of their first CTA start / last CTA end: [0,1] gradient kernel, [2,3] S_xx kernel. */
int vgan_b200_debug_timeline(unsigned long long* buf);

/* Test / measurement aid for the CTA-pair kernel: bits 0-7 = cap on the number of persistent clusters (0 = as many
 * as fit; a small cap forces the stream-K partition to cut units between clusters on small problems); bits 8+ =
 * measurement switches that break results (see TcParams::dbg_flags).  0 = default. */
int vgan_b200_debug_variant(int v);

/* Debug aid: cta_group::2 Gram + gradient tile on one CTA pair. A0,A1,B bf16 [128,128], V bf16 [128,256];
 * S_out float [2,128,128] = A_h B^T;  O_out float [2,128,256] = bf16(S_h) V. */
int vgan_b200_pair_probe(const void* A0, const void* A1, const void* B, const void* V, float* S_out, float* O_out,
                         void* stream);

/* Measurement aid: L2 -> shared TMA throughput with `stages` x `loads` boxes ([box_rows x 64] bf16) in flight per
 * CTA, `grid` CTAs, `nprod` producer (and consumer) threads; mode 0 = all CTAs stream the same tiles, 1 = each re-reads its own tile, 2 = one of each per
 * stage.  Z bf16 [rows, cols]; cycles_out[grid] = SM cycles of each CTA for `iters` stages. */
int vgan_b200_tma_rate_probe(const void* Z, int64_t rows, int64_t cols, int stages, int loads, int box_rows,
                             int iters, int mode, int grid, int nprod, long long* cycles_out, void* stream);

/* Measurement aid: SM cycles for iters x 4 tcgen05.mma (M=128, N=n, K=16) on resident operands in each of `grid`
 * CTAs while `tma_warps` warps per CTA stream 16 KiB TMA boxes of Z (bf16 [rows,1024]) into spare shared memory:
 * out[2b] = cycles of CTA b, out[2b+1] = boxes it loaded meanwhile.  Shows how operand reads and TMA fills
 * share the shared-memory bandwidth. */
int vgan_b200_umma_rate_probe(int n, int iters, int b_mn, const void* Z, int64_t rows, int tma_warps, int grid,
                              long long* out, void* stream);

int vgan_b200_umma_probe(const void* A, const void* B, float* S_out, float* O_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VGAN_B200_H */
