// Subspace statistics on the device: the unique rows of the boolean mask matrix
// `generator(noise) >= 1/D` with their multiplicities, in the lexicographic row order numpy's
// np.unique(axis=0, return_counts=True) returns them in (/root/reference/src/vgan.py:372-382, 649-659).
// Rows are packed MSB-first into 64-bit words, so row order == unsigned lexicographic word order.
// n is the number of sampled subspaces (500 in the reference): two O(n^2 W) rank passes, no sort.
#include <mutex>

#include "common.cuh"

namespace vg {

constexpr size_t kSubspaceStaticScratchBytes = VGAN_B200_SUBSPACE_STATIC_SCRATCH_BYTES;   // 1 MiB

__global__ void subspace_pack_kernel(const void* __restrict__ U, int in_is_bool, int64_t n, int64_t D, float thr,
                                     int W, unsigned long long* __restrict__ packed) {
  const int64_t i = blockIdx.x;
  for (int w = threadIdx.x; w < W; w += blockDim.x) {
    unsigned long long bits = 0;
    for (int b = 0; b < 64; ++b) {
      const int64_t d = static_cast<int64_t>(w) * 64 + b;
      bool on = false;
      if (d < D)
        on = in_is_bool ? (static_cast<const uint8_t*>(U)[i * D + d] != 0)
                        : (static_cast<const float*>(U)[i * D + d] >= thr);   // torch.greater_equal(u, 1/D)
      bits = (bits << 1) | (on ? 1ull : 0ull);
    }
    packed[i * W + w] = bits;
  }
}

__device__ __forceinline__ int row_cmp(const unsigned long long* a, const unsigned long long* b, int W) {
  for (int w = 0; w < W; ++w) {
    if (a[w] != b[w]) return a[w] < b[w] ? -1 : 1;
  }
  return 0;
}

// pass 1: first[i] = no earlier equal row; cnt[i] = number of rows equal to row i
__global__ void subspace_count_kernel(const unsigned long long* __restrict__ packed, int64_t n, int W,
                                      int* __restrict__ first, int* __restrict__ cnt) {
  const int64_t i = blockIdx.x;
  const unsigned long long* ri = packed + i * W;
  int c = 0, earlier = 0;
  for (int64_t j = threadIdx.x; j < n; j += blockDim.x) {
    if (row_cmp(ri, packed + j * W, W) == 0) {
      ++c;
      if (j < i) earlier = 1;
    }
  }
  __shared__ int sc, se;
  if (threadIdx.x == 0) { sc = 0; se = 0; }
  __syncthreads();
  atomicAdd(&sc, c);
  atomicOr(&se, earlier);
  __syncthreads();
  if (threadIdx.x == 0) { cnt[i] = sc; first[i] = se ? 0 : 1; }
}

// pass 2: rank of every first occurrence among the distinct rows; scatter row + count
__global__ void subspace_scatter_kernel(const unsigned long long* __restrict__ packed, int64_t n, int64_t D, int W,
                                        const int* __restrict__ first, const int* __restrict__ cnt,
                                        uint8_t* __restrict__ unique_out, long long* __restrict__ counts_out,
                                        int* __restrict__ n_unique) {
  const int64_t i = blockIdx.x;
  if (!first[i]) return;
  const unsigned long long* ri = packed + i * W;
  int r = 0;
  for (int64_t j = threadIdx.x; j < n; j += blockDim.x)
    if (first[j] && row_cmp(packed + j * W, ri, W) < 0) ++r;
  __shared__ int sr;
  if (threadIdx.x == 0) sr = 0;
  __syncthreads();
  atomicAdd(&sr, r);
  __syncthreads();
  const int rank = sr;
  for (int64_t d = threadIdx.x; d < D; d += blockDim.x)
    unique_out[static_cast<int64_t>(rank) * D + d] = static_cast<uint8_t>((ri[d >> 6] >> (63 - (d & 63))) & 1ull);
  if (threadIdx.x == 0) {
    counts_out[rank] = cnt[i];
    atomicAdd(n_unique, 1);
  }
}

int run_subspace_dist(const void* U, int in_is_bool, int64_t n, int64_t D, float thr, void* scratch,
                      uint8_t* unique_out, long long* counts_out, int* n_unique, cudaStream_t st) {
  const int W = static_cast<int>((D + 63) / 64);
  unsigned long long* packed = static_cast<unsigned long long*>(scratch);
  int* first = reinterpret_cast<int*>(packed + n * W);
  int* cnt = first + n;
  VG_CUDA_CHECK(cudaMemsetAsync(n_unique, 0, sizeof(int), st));
  subspace_pack_kernel<<<static_cast<unsigned>(n), 64, 0, st>>>(U, in_is_bool, n, D, thr, W, packed); count_launch();
  subspace_count_kernel<<<static_cast<unsigned>(n), 128, 0, st>>>(packed, n, W, first, cnt); count_launch();
  subspace_scatter_kernel<<<static_cast<unsigned>(n), 128, 0, st>>>(packed, n, D, W, first, cnt, unique_out,
                                                                    counts_out, n_unique); count_launch();
  VG_CUDA_CHECK(cudaGetLastError());
  return VGAN_B200_OK;
}

// ---- no-allocation variant (vgan_b200_subspace_dist_static) ---------------------------------------------------------
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
// it): the same three kernels with their scratch in a file-scope device array instead of a caller buffer.  Every GPU
// that loads the library holds these kStaticScratchBytes until the process exits.  The array is shared by every call
// on a device and keeps the last call's contents; nothing in it has to start at zero (the pack kernel writes every
// packed word and the count kernel every first/cnt word of rows 0..n-1 before a later kernel reads it; only the
// caller's n_unique is cleared).  Calls that overlap on the GPU (two streams, two contexts, two host threads) race
// on it: serialise them.
static __device__ __align__(256) unsigned char g_subspace_scratch[kSubspaceStaticScratchBytes];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_scratch_mutex;
void* g_scratch_addr[kMaxDevices] = {nullptr};   // device address of g_subspace_scratch on device i (0 = not asked yet)
}  // namespace

// In host code `g_subspace_scratch` is a host-side stand-in of the symbol, not a device address: the address the
// kernels get comes from cudaGetSymbolAddress, asked once per device with that device current (the entry point's
// device guard has made it current).
static int static_scratch_address(void** out) {
  int dev = -1;
  VG_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    set_error("subspace_dist_static: device index %d outside the table of %d devices", dev, kMaxDevices);
    return VGAN_B200_ERR_ARG;
  }
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (!g_scratch_addr[dev]) {
    void* ptr = nullptr;
    VG_CUDA_CHECK(cudaGetSymbolAddress(&ptr, g_subspace_scratch));
    if (!ptr) {
      set_error("subspace_dist_static: cudaGetSymbolAddress returned a null address on device %d", dev);
      return VGAN_B200_ERR_CUDA;
    }
    g_scratch_addr[dev] = ptr;
  }
  *out = g_scratch_addr[dev];
  return VGAN_B200_OK;
}

int run_subspace_dist_static(const void* U, int in_is_bool, int64_t n, int64_t D, float thr, uint8_t* unique_out,
                             long long* counts_out, int* n_unique, cudaStream_t st) {
  const size_t need = static_cast<size_t>(n) * ((D + 63) / 64) * 8 + static_cast<size_t>(n) * 8;
  if (need > kSubspaceStaticScratchBytes) {
    set_error("subspace_dist_static: n = %lld, D = %lld needs %zu bytes of scratch, the static array holds %zu "
              "(use vgan_b200_subspace_dist with a caller buffer)",
              static_cast<long long>(n), static_cast<long long>(D), need, kSubspaceStaticScratchBytes);
    return VGAN_B200_ERR_ARG;
  }
  void* scratch = nullptr;
  const int rc = static_scratch_address(&scratch);
  if (rc != VGAN_B200_OK) return rc;
  return run_subspace_dist(U, in_is_bool, n, D, thr, scratch, unique_out, counts_out, n_unique, st);
}

}  // namespace vg
