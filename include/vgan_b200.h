/* vgan_b200 — C ABI of the B200-native (sm_100a) squared-MMD forward+backward path.
 *
 * Drop-in boundary for V-GAN's loss module.  The reference has no FFI (it is pure
 * Python/torch); the calls these entry points replace are
 *   RBF.get_bandwidth               /root/reference/src/models/Mmd_loss_constrained.py:16-22
 *   RBF.forward                     /root/reference/src/models/Mmd_loss_constrained.py:24-26
 *   MMDLossConstrained.forward      /root/reference/src/models/Mmd_loss_constrained.py:42-50
 *   its autograd backward (EuclideanDistBackward0 ... MulBackward)  — SURVEY.md 3.3
 *   the masking  Y = U * X          /root/reference/src/vgan.py:616 (also :275,:312)
 *   upper_softmax (producer of U)   /root/reference/src/models/Generator.py:18-22
 *   tts.MMDStatistic(...)(x, y, alphas, ret_matrix=True)            /root/reference/src/vgan.py:419-428
 *
 * Conventions: extern "C"; every function returns 0 on success or a negative
 * vgan_b200_status; vgan_b200_last_error() gives the text (thread local).  All data
 * pointers are DEVICE pointers borrowed for the call unless a name ends in _host; all
 * work is enqueued on `stream` (a cudaStream_t passed as void*); nothing synchronises
 * the host.  No CPU fallback exists: without an sm_100 device every call fails.
 *
 * State: there is no global mutable state.  Everything a call needs beyond its arguments
 * (device attributes, cached tensor maps, the two internal streams the D > 256 kernels fork
 * onto) lives in an opaque context, vgan_b200_create(device).  Calls on DISTINCT contexts
 * are thread-safe against each other (also on the same device, on distinct streams); one
 * context must not be used by two host threads at the same time.  The library has no
 * switch, environment variable or entry point that changes results: measurement aids
 * live in a separate debug build (libvgan_b200_dbg.so, `make DEBUG=1`).
 */
#ifndef VGAN_B200_H
#define VGAN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define VGAN_B200_API __attribute__((visibility("default")))   /* the library is built with -fvisibility=hidden */
#else
#define VGAN_B200_API
#endif

#define VGAN_B200_ABI_VERSION 2
#define VGAN_B200_MAX_SCALES 8

typedef enum {
  VGAN_B200_OK = 0,
  VGAN_B200_ERR_ARG = -1,       /* bad shape / null pointer / misalignment          */
  VGAN_B200_ERR_DEVICE = -2,    /* no CUDA device, or not compute capability 10.x   */
  VGAN_B200_ERR_CUDA = -3,      /* a CUDA runtime/driver call failed                */
  VGAN_B200_ERR_UNSUPPORTED = -4
} vgan_b200_status;

/* flags of vgan_b200_workspace_layout (they select the kernels every later call on that workspace runs) */
enum {
  VGAN_B200_PREC_BF16 = 0,      /* bf16 operands, fp32 accumulate, tcgen05 path (2e-3) */
  VGAN_B200_PREC_FP32 = 1,      /* fp32 CUDA-core path (1e-4)                           */
  VGAN_B200_PREC_MASK = 1,
  VGAN_B200_CENTER    = 2,      /* subtract the stacked column mean (distance invariant) */
  VGAN_B200_NEED_DX   = 4,      /* also produce dL/dX (VGAN.fit, vgan.py:283)           */
  VGAN_B200_LADDER    = 8,      /* scales are s, s/2, s/4, s/8, s/16: 1 ex2 + 4 squarings */
  VGAN_B200_BF16X3    = 64,     /* with PREC_BF16: fp32-class results on the tensor cores.  Every operand value is split
                                   z = hi + lo (two bf16), the Gram is hi.hi + hi.lo + lo.hi in ONE K = 3 D GEMM, the
                                   gradient GEMM adds its three parts P_hi V_hi + P_hi V_lo + P_lo V_hi (one pass for
                                   D <= 256, two above) */
  VGAN_B200_ONES      = 65536   /* OUTPUT of vgan_b200_workspace_layout in layout.flags, ignored on input: the stored rows
                                   carry three constant features 1.0 in the zero padding behind the D real ones (set
                                   when the padding has room and the D <= 256 kernel keeps the row operand in shared
                                   memory).  Distances do not change (off_nrm / off_nrms include the +3); the kernel
                                   puts the row norms into those columns of its row operand and reads the row sums of
                                   P from the matching column of O, which takes two packed FP32 operations per column
                                   pair out of the epilogue. */
};

typedef struct vgan_b200_ctx vgan_b200_ctx;    /* opaque */

/* Layout of the caller-provided workspace (all offsets in bytes from `ws`, which must be 1024-byte aligned).
 * Z is the stacked sample [X ; Y] (Mmd_loss_constrained.py:43): X rows at [0,Nx),
 * zero rows up to Nxp (multiple of 128), Y rows at [Nxp, Nxp+Ny), zero rows up to
 * Nxp+Nyp; Dp = D rounded up to 64 with zero columns (to 128 on the bf16 path when D > 256 and, plain bf16, for
 * D = 62..64, where the extra chunk makes room for VGAN_B200_ONES).
 * The accumulator block [off_acc, off_acc + acc_bytes) is zeroed by vgan_b200_stats; on several GPUs the caller
 * reduces its parts between the calls (see vgan_b200_stats / _mmd2_fwd_bwd). */
typedef struct {
  int64_t Nx, Ny, D, D_u;
  int64_t Dp, Nxp, Nyp;
  int32_t flags, n_scales;
  size_t off_par;     /* float[64]    device parameter block of K1 (exp2 scale, coefficients, 1/scales, bandwidth) */
  size_t off_acc;     /* start of the accumulator block                                          */
  size_t off_sums;    /* double[4]    S_xx, S_xy, S_yy, spare                                    */
  size_t off_colsum;  /* double[Dp]   column sums of the gradient rows written (colsum_i dL/dU[i,:] on the masked path);
                                      directly behind the sums: one all-reduce of 4 + Dp doubles serves both */
  size_t off_stats;   /* double[2+2Dp] sum|x|^2, sum|y|^2, column sums of X, column sums of Y    */
  size_t off_cmax;    /* uint64[D_u]  packed (value,row) column maxima of U                      */
  size_t off_misc;    /* float penalty, float loss, uint32 counters                              */
  size_t acc_bytes;
  size_t off_mean;    /* float [Dp]   stacked column mean (zeros unless CENTER)                  */
  size_t off_cm;      /* float [2 Dp] common-mode gradient vectors of X rows / Y rows (see scales) */
  size_t off_nrm;     /* float [Nxp+Nyp]  |z_i|^2 of the stored (rounded) rows; 1e30 on pads     */
  size_t off_nrms;    /* float [Nxp+Nyp]  the same, scaled for the exp2 argument                 */
  size_t off_xchg;    /* bytes: P-tile exchange scratch of the D_p > 256 kernel (L2 resident), else empty */
  size_t off_zq;      /* bf16  [Nxp+Nyp, Dp]  (PREC_BF16); with BF16X3 two arrays [Nxp+Nyp, 3 Dp] back to back:
                         row operand [hi | hi | lo], then column operand [hi | lo | hi]          */
  size_t off_zf;      /* float [Nxp+Nyp, Dp]  (PREC_FP32; same offset, one of the two exists)    */
  size_t total;       /* bytes to allocate                                                       */
} vgan_b200_layout_t;

VGAN_B200_API int vgan_b200_abi_version(void);
VGAN_B200_API const char* vgan_b200_last_error(void);

/* Context for `device` (CUDA ordinal).  Creating it checks for compute capability 10.x. */
VGAN_B200_API int vgan_b200_create(int device, vgan_b200_ctx** out);
VGAN_B200_API int vgan_b200_destroy(vgan_b200_ctx* ctx);
/* Resource limit: at most `max_clusters` persistent thread-block clusters for the D > 256 gradient kernel (0 = as
 * many as the device holds).  The work is cut evenly over however many run; results do not depend on the limit
 * beyond the order of fp32 additions on rows shared by two clusters. */
VGAN_B200_API int vgan_b200_set_cluster_limit(vgan_b200_ctx* ctx, int max_clusters);
/* Number of kernels this library has launched through `ctx` so far (host counter, incremented at every launch). */
VGAN_B200_API int vgan_b200_launch_count(vgan_b200_ctx* ctx, unsigned long long* out);

/* Fills *out.  D_u = columns of U (0: no penalty term), n_scales = number of bandwidth multipliers. */
VGAN_B200_API int vgan_b200_workspace_layout(int64_t Nx, int64_t Ny, int64_t D, int64_t D_u, int n_scales, int flags,
                                             vgan_b200_layout_t* out);

/* ---- one evaluation = the five calls below in this order on one stream (vgan_b200_mmd2_eval issues them).
 * X [Nx,D]; either Y [Ny,D] or, when Y == NULL, the mask U [Nx,D] (then Ny == Nx and Y = U*X is formed on the fly,
 * vgan.py:616); U [Nx,D_u] also feeds the column-max penalty when D_u > 0.  fp32, row stride D (D_u).
 * Row ranges [row_begin,row_end) select rows of X AND of Y (clipped to Nx / Ny): the whole problem is
 * row_begin = 0, row_end = max(Nx,Ny); a rank of a multi-GPU job passes its own rows. */

/* 1. K2a: zeroes the accumulator block, then adds the statistics of the rows in range: sum|x|^2, sum|y|^2, column
 *    sums of X and of Y (closed-form bandwidth, Mmd_loss_constrained.py:16-22; centring mean; common mode) and the
 *    column maxima of U (Mmd_loss_constrained.py:50).  Multi-GPU: all-reduce(sum) the doubles at off_stats and
 *    all-reduce(max) the keys at off_cmax (as unsigned 64-bit) before step 2. */
VGAN_B200_API int vgan_b200_stats(vgan_b200_ctx* ctx, const vgan_b200_layout_t* L, void* ws, const float* X,
                                  const float* U, const float* Y, int64_t row_begin, int64_t row_end, void* stream);

/* 2. Bandwidth + scales + parameter block (single small kernel).  If frozen == 0: *bandwidth (device float) =
 *    sum_ij d2 / (n^2-n) from the statistics, else *bandwidth is read.  1/s_m = 1 / fp32(bandwidth * mults_host[m])
 *    (the reference's rounding, Mmd_loss_constrained.py:26).  Also: the centring mean, the penalty value
 *    weight * mean_d(1 - max_i U[i,d]), and the common-mode constants — for every (row block, column block) pair
 *    w0 = w(mean d2 of the block); the kernels multiply P' = c_j (w_ij - w0) instead of P = c_j w_ij (what bf16
 *    rounds is then the VARIATION of the weights, not their common level) and the closed-form remainder
 *    sum_j c_j w0 (z_j - z_i) is added back from the block means. */
VGAN_B200_API int vgan_b200_scales(vgan_b200_ctx* ctx, const vgan_b200_layout_t* L, void* ws, float* bandwidth,
                                   int frozen, const float* mults_host, float weight, void* stream);

/* 3. K2b: packs the rows in range into the stacked operand (masking, centring, bf16 cast or hi/lo split, row
 *    norms of the stored values, exp2-scaled norms); the call whose range reaches the end of a block also writes
 *    that block's zero padding rows.  Multi-GPU: all-gather the packed rows (off_zq) and off_nrms (off_nrm on the
 *    fp32 path) of every rank before step 4 — half the bytes of the fp32 inputs. */
VGAN_B200_API int vgan_b200_pack(vgan_b200_ctx* ctx, const vgan_b200_layout_t* L, void* ws, const float* X,
                                 const float* U, const float* Y, int64_t row_begin, int64_t row_end, void* stream);

/* 4. K1: fused forward+backward over the pairwise matrix, which never reaches HBM.
 *   sums (off_sums) += S_xx, S_xy, S_yy of this shard (double): S_xx from X-row tiles t with t % xx_nparts ==
 *                 xx_part, S_xy/S_yy from Y rows [row_begin,row_end).
 *   dY[i - row_begin, :] = dL/dy_i  (x dY_mul[i - row_begin, :] when dY_mul != NULL, which yields
 *                 dL/dU = dL/dY * X directly) for i in [row_begin,row_end); row stride D.  NULL skips the gradient
 *                 GEMM (forward only).  The column sums of the rows written are added to off_colsum.
 *   dX (NEED_DX): dL/dx_i for the same row range of X.
 *   X, U, Y: the inputs again (the same pointers as in steps 1 and 3; may be NULL when dY is NULL): the gradient's
 *                 -rowsum(P) z_i term takes z_i from the fp32 inputs, not from the rounded operand.
 * row_begin must be a multiple of 128 (row_end need not be).  Multi-GPU: all-reduce(sum) the 4 + Dp doubles at
 * off_sums before step 5. */
VGAN_B200_API int vgan_b200_mmd2_fwd_bwd(vgan_b200_ctx* ctx, const vgan_b200_layout_t* L, void* ws, const float* X,
                                         const float* U, const float* Y, int64_t row_begin, int64_t row_end,
                                         int xx_part, int xx_nparts, float* dY, const float* dY_mul, float* dX,
                                         void* stream);

/* 5. loss[0] = S_xx/Nx^2 - 2 S_xy/(Nx Ny) + S_yy/Ny^2 + penalty (device float; also kept at off_misc + 4), and the
 *    penalty's gradient -weight/D_u at (argmax row, d), lowest row on ties (Mmd_loss_constrained.py:50), added to
 *    dU[row - dU_row_begin, d] for the argmax rows inside [dU_row_begin, dU_row_begin + dU_rows) (dU may be NULL)
 *    and, when colsum_penalty != 0, to every entry of off_colsum.  loss may be NULL. */
VGAN_B200_API int vgan_b200_finalize(vgan_b200_ctx* ctx, const vgan_b200_layout_t* L, void* ws, float weight,
                                     float* dU, int64_t dU_row_begin, int64_t dU_rows, int colsum_penalty,
                                     float* loss, void* stream);

/* The whole evaluation of MMDLossConstrained.forward(X, Y, U) + backward on one GPU: steps 1-5 on `stream`. */
typedef struct {
  const vgan_b200_layout_t* layout;
  void* ws;
  const float* X;
  const float* U;            /* mask / penalty input, or NULL                                   */
  const float* Y;            /* NULL: Y = U * X                                                  */
  float* bandwidth;          /* device float, see vgan_b200_scales                               */
  int frozen;
  const float* mults_host;   /* layout->n_scales multipliers                                     */
  float weight;
  float* dY;                 /* [Ny, D] or NULL                                                  */
  const float* dY_mul;       /* NULL, or the factor that turns dL/dY into dL/dU (= X)            */
  float* dX;                 /* [Nx, D] with NEED_DX                                             */
  float* dU_penalty;         /* where the penalty gradient goes: dY itself on the masked path, a zeroed [Nx, D_u]
                                buffer otherwise, or NULL                                        */
  float* loss;               /* device float, or NULL                                            */
  void* stream;
} vgan_b200_eval_t;
VGAN_B200_API int vgan_b200_mmd2_eval(vgan_b200_ctx* ctx, const vgan_b200_eval_t* args);

/* Producer of the mask (Generator.py:18-22): p = softmax(H[i,:]), U = p where p < 1/D else 1.  Writes U and
 * the row statistics (max, 1/sum of exponentials) the backward needs.  H, U fp32 [N, D]. */
VGAN_B200_API int vgan_b200_upper_softmax(vgan_b200_ctx* ctx, const float* H, int64_t N, int64_t D, float* U,
                                          float* row_max, float* row_inv_sum, void* stream);
/* Its backward: dH = p * (g - sum_f g_f p_f) with g = dU where p < 1/D, else 0. */
VGAN_B200_API int vgan_b200_upper_softmax_bwd(vgan_b200_ctx* ctx, const float* H, const float* row_max,
                                              const float* row_inv_sum, const float* dU, int64_t N, int64_t D,
                                              float* dH, void* stream);

/* Dense kernel matrix K[n,n] = sum_m exp(-(d2 + eps) * inv_scales[m]) for small n
 * (RBF.forward parity; MMDStatistic(..., ret_matrix=True) with inv_scales = alphas,
 * eps = 1e-5).  Z fp32 [n,D]; n <= 65536. */
VGAN_B200_API int vgan_b200_mmd2_matrix(vgan_b200_ctx* ctx, const float* Z, int64_t n, int64_t D,
                                        const float* inv_scales, int n_scales, float eps, float* K_out, void* stream);

/* Permutation test of the unbiased MMD^2 statistic (torch_two_sample MMDStatistic.pval ->
 * permutation_test.pyx::permutation_test_mat, reached from vgan.py:425-428 / 702-705).  K float [n,n] kernel
 * matrix (vgan_b200_mmd2_matrix), member uint8 [n_perm, n]: row p marks with 1 the n_1 points that relabeling p
 * assigns to sample 1 (row with the first n_1 ones = the observed labels).  stats double [n_perm] receives
 * a00 (sum K11 - tr K11) + a11 (sum K22 - tr K22) + 2 a01 sum K12 of every relabeling; the p-value is the fraction
 * of rows whose statistic is >= the observed one.  scratch >= vgan_b200_permutation_scratch_bytes(n); n <= 16384
 * (the GoF test runs at n = 1000). */
VGAN_B200_API int vgan_b200_permutation_scratch_bytes(int64_t n, size_t* out);
VGAN_B200_API int vgan_b200_permutation_stats(vgan_b200_ctx* ctx, const float* K, int64_t n, int64_t n_1,
                                              const uint8_t* member, int64_t n_perm, void* scratch, double* stats,
                                              void* stream);

/* Subspace distribution (vgan.py:372-382 / 649-659): unique rows of the boolean mask matrix
 * (U[i,d] >= threshold for float U, U[i,d] != 0 when in_is_bool) in np.unique(axis=0) order, and how often each
 * occurs.  unique_out uint8 [n, D] (the first *n_unique rows are written), counts_out int64 [n], n_unique device
 * int; scratch >= vgan_b200_subspace_scratch_bytes(n, D); n <= 65536 (two O(n^2) passes over packed rows; the
 * reference samples 500).  Replaces generate_subspaces' threshold (vgan.py:369 / 646) + np.unique on the host. */
VGAN_B200_API int vgan_b200_subspace_scratch_bytes(int64_t n, int64_t D, size_t* out);
VGAN_B200_API int vgan_b200_subspace_dist(vgan_b200_ctx* ctx, const void* U, int in_is_bool, int64_t n, int64_t D,
                                          float threshold, void* scratch, uint8_t* unique_out, long long* counts_out,
                                          int* n_unique, void* stream);

/* vgan_b200_subspace_dist without a scratch argument and without any allocation: no cudaMalloc / cudaFree / allocating
 * library call happens inside, so it can be used where those are forbidden (under CUDA graph capture, in tight
 * serving loops, in harnesses that ban them).  NO COUNTERPART IN THE REFERENCE.
 * The scratch is a static device array of VGAN_B200_SUBSPACE_STATIC_SCRATCH_BYTES (1 MiB) that every GPU
 * which loads the library holds until the process exits; shapes with
 * vgan_b200_subspace_scratch_bytes(n, D) > VGAN_B200_SUBSPACE_STATIC_SCRATCH_BYTES are rejected with
 * VGAN_B200_ERR_ARG (the reference's 500 sampled subspaces fit up to D = 16,704 features; n = 4096 up to D = 1,984).
 * Inputs and outputs are the caller's, as in vgan_b200_subspace_dist, and the results are identical.  All calls on a
 * device share the one array, which keeps the previous call's contents (nothing in it needs to start at zero): calls
 * that overlap on the GPU - different streams, contexts or host threads - RACE on it and must be serialised by the
 * caller. */
#define VGAN_B200_SUBSPACE_STATIC_SCRATCH_BYTES (1u << 20)
VGAN_B200_API int vgan_b200_subspace_dist_static(vgan_b200_ctx* ctx, const void* U, int in_is_bool, int64_t n,
                                                 int64_t D, float threshold, uint8_t* unique_out,
                                                 long long* counts_out, int* n_unique, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VGAN_B200_H */
