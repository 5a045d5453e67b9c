// Measured ceilings of the pipes the MMD^2 kernels are bounded by: MUFU ex2 rate, scalar and packed
// FP32 issue rate, tcgen05 kind::f16 / kind::tf32 rate of one CTA per SM, L2 -> SM read bandwidth.  Stand-alone
// binary (NOT part of libvgan_b200.so): `make -C tools/peaks && tools/peaks/peaks > gpurun_out/peaks.json`.
// Every figure is (operations counted in the kernel) / (CUDA-event time of a launch that runs ~0.2-0.5 s), so the
// clock is whatever the chip sustains under that load; the SM clock seen by the kernel (clock64 / globaltimer) is
// printed next to it.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../vgan_b200/csrc/ptx.cuh"

using namespace vg;

#define CK(x)                                                                               \
  do {                                                                                      \
    cudaError_t e_ = (x);                                                                   \
    if (e_ != cudaSuccess) {                                                                \
      fprintf(stderr, "%s failed: %s (%s:%d)\n", #x, cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(1);                                                                              \
    }                                                                                       \
  } while (0)

struct ClockRec { unsigned long long cycles, ns; };

__device__ __forceinline__ void clock_begin(unsigned long long& c, unsigned long long& t) {
  c = clock64();
  t = global_timer_ns();
}
__device__ __forceinline__ void clock_end(ClockRec* rec, unsigned long long c, unsigned long long t) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    rec->cycles = clock64() - c;
    rec->ns = global_timer_ns() - t;
  }
}

// ---- MUFU: 8 independent ex2 chains per thread
__global__ void __launch_bounds__(512) k_ex2(int iters, float* sink, ClockRec* rec) {
  unsigned long long c, t;
  clock_begin(c, t);
  float x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = 0.1f * (threadIdx.x & 7) + 0.01f * i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
#pragma unroll
    for (int i = 0; i < 8; ++i) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += x[i];
  if (s == 123.456f) sink[0] = s;
  clock_end(rec, c, t);
}

// ---- scalar FFMA: 8 chains
__global__ void __launch_bounds__(512) k_ffma(int iters, float* sink, ClockRec* rec) {
  unsigned long long c, t;
  clock_begin(c, t);
  float x[8];
  const float a = 0.999f, b = 0.001f;
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = 0.1f * (threadIdx.x & 7) + 0.01f * i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(a), "f"(b));
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += x[i];
  if (s == 123.456f) sink[0] = s;
  clock_end(rec, c, t);
}

// ---- packed FFMA2 (fma.rn.f32x2): 8 chains of register pairs
__global__ void __launch_bounds__(512) k_ffma2(int iters, float* sink, ClockRec* rec) {
  unsigned long long c, t;
  clock_begin(c, t);
  uint64_t x[8];
  const uint64_t a = f2_pack(0.999f, 0.998f), b = f2_pack(0.001f, 0.002f);
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = f2_pack(0.1f * (threadIdx.x & 7) + 0.01f * i, 0.3f);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(x[i]) : "l"(a), "l"(b));
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float lo, hi;
    f2_unpack(x[i], lo, hi);
    s += lo + hi;
  }
  if (s == 123.456f) sink[0] = s;
  clock_end(rec, c, t);
}

// ---- the ladder epilogue's instruction mix per column pair: 2 ex2 + 17 packed FP32 ops (tc_epi.cuh), 4 chains
__global__ void __launch_bounds__(512) k_epimix(int iters, float* sink, ClockRec* rec) {
  unsigned long long c, t;
  clock_begin(c, t);
  uint64_t acc[4], ws[4];
  const uint64_t m2 = f2_pack(-0.002f, -0.002f), nn = f2_pack(0.7f, 0.7f), cf = f2_pack(0.3f, 0.3f);
#pragma unroll
  for (int i = 0; i < 4; ++i) { acc[i] = f2_pack(0.f, 0.f); ws[i] = f2_pack(0.01f * threadIdx.x, 0.5f + i); }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float u0, u1;
      f2_unpack(f2_fma(ws[i], m2, f2_add(nn, nn)), u0, u1);                        // 2
      const uint64_t t1 = f2_pack(ex2_approx(-u0), ex2_approx(-u1));               // 2 MUFU
      const uint64_t t2 = f2_mul(t1, t1), t4 = f2_mul(t2, t2), t8 = f2_mul(t4, t4), t16 = f2_mul(t8, t8);   // 4
      acc[i] = f2_add(acc[i], f2_add(f2_add(t1, t2), f2_add(f2_add(t4, t8), t16)));                          // 5
      const uint64_t w2 = f2_fma(t16, cf, f2_fma(t8, cf, f2_fma(t4, cf, f2_fma(t2, cf, f2_mul(t1, cf)))));   // 5
      ws[i] = f2_add(ws[i], w2);                                                   // 1
    }
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float lo, hi;
    f2_unpack(acc[i], lo, hi);
    s += lo + hi;
    f2_unpack(ws[i], lo, hi);
    s += lo + hi;
  }
  if (s == 123.456f) sink[0] = s;
  clock_end(rec, c, t);
}

// ---- L2 -> SM: every CTA streams the same buffer (smaller than L2) with 16-byte loads that bypass L1
__global__ void __launch_bounds__(512) k_l2read(const uint4* buf, size_t per_thread, int passes, uint32_t* sink,
                                                ClockRec* rec) {
  unsigned long long c, t;
  clock_begin(c, t);
  uint32_t acc = 0;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const uint4* p0 = buf + static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  for (int p = 0; p < passes; ++p) {
    for (size_t k = 0; k < per_thread; k += 4) {          // per_thread is a multiple of 4
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = __ldcg(p0 + (k + u) * stride);
#pragma unroll
      for (int u = 0; u < 4; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
  }
  if (acc == 0x12345678u) sink[0] = acc;
  clock_end(rec, c, t);
}

// ---- tcgen05: one CTA per SM, M = 128, N = 256 per instruction, operands resident in shared memory
// (A 16 KiB + B 32 KiB per K chunk of 128 bytes, 2 chunk sets), accumulating into TMEM columns [0, 256).
// kind 0 = f16 (bf16 operands, K = 16 per instruction), kind 1 = tf32 (K = 8 per instruction).
template <int KIND>
__global__ void __launch_bounds__(128, 1) k_umma(int batches, ClockRec* rec) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar[2];
  __shared__ uint32_t tmem_base;
  unsigned long long c, t;
  clock_begin(c, t);
  // random-looking operand bits with small exponents (no inf/nan): the power draw depends on the data
  uint32_t* w = reinterpret_cast<uint32_t*>(smem);
  for (int i = threadIdx.x; i < 2 * 48 * 1024 / 4; i += blockDim.x) {
    uint32_t h = (i * 2654435761u) ^ (blockIdx.x * 40503u);
    h ^= h >> 13; h *= 0x5bd1e995u; h ^= h >> 15;
    if (KIND == 0) w[i] = (h & 0x807f807fu) | 0x3f003f00u;       // two bf16 in [0.5, 1) with random sign / mantissa
    else w[i] = (h & 0x807fe000u) | 0x3f000000u;                 // tf32 in [0.5, 1)
  }
  if (threadIdx.x == 0) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    fence_barrier_init();
  }
  fence_proxy_async_smem();
  if (threadIdx.x < 32) {
    tmem_alloc(&tmem_base, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base;
  if (threadIdx.x < 32) {
    // instruction descriptor: D fp32; A/B format 1 = bf16 (kind::f16) or 2 = tf32 (kind::tf32); K-major both
    constexpr uint32_t fmt = KIND == 0 ? 1u : 2u;
    constexpr uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | (static_cast<uint32_t>(256 >> 3) << 17) |
                               (static_cast<uint32_t>(128 >> 4) << 24);
    constexpr uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
    const uint32_t lo0 = (smem_u32(smem) & 0x3FFFFu) >> 4;
    for (int b = 0; b < batches; ++b) {
      if (b >= 2) mbar_wait(&bar[b & 1], ((b >> 1) - 1) & 1);     // batch b - 2 has completed
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int s = 0; s < 8; ++s) {                             // 8 stages x 4 instructions = 32 per batch
          const uint32_t a_lo = lo0 + (s & 1) * ((48 * 1024) >> 4);
          const uint32_t b_lo = a_lo + ((16 * 1024) >> 4);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            const uint64_t adesc = (static_cast<uint64_t>(desc_hi) << 32) | (a_lo + kk * 2);
            const uint64_t bdesc = (static_cast<uint64_t>(desc_hi) << 32) | (b_lo + kk * 2);
            if (KIND == 0) {
              umma_bf16_ss(tmem, adesc, bdesc, idesc, 1u);
            } else {
              asm volatile(
                  "{\n\t.reg .pred p;\n\t"
                  "setp.ne.b32 p, %4, 0;\n\t"
                  "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n"
                  ::"r"(tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(1u)
                  : "memory");
            }
          }
        }
        umma_commit(&bar[b & 1]);
      }
      __syncwarp();
    }
    // drain
    for (int b = batches - 2 < 0 ? 0 : batches - 2; b < batches; ++b) mbar_wait(&bar[b & 1], (b >> 1) & 1);
    tc_fence_after();
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tmem, 512);
  clock_end(rec, c, t);
}

__global__ void __launch_bounds__(512, 1) k_dummy(int* x) {
  extern __shared__ uint8_t dsm[];
  if (x && threadIdx.x == 0) x[0] = dsm[0];
}

struct Timed { double ms; double mhz; };

template <typename F>
static Timed timed(F launch, ClockRec* d_rec) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  launch();                                   // warm-up
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  launch();
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms = 0.f;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  ClockRec r;
  CK(cudaMemcpy(&r, d_rec, sizeof(r), cudaMemcpyDeviceToHost));
  return {ms, r.ns ? 1e3 * static_cast<double>(r.cycles) / static_cast<double>(r.ns) : 0.0};
}

int main() {
  int dev = 0, sms = 0;
  CK(cudaSetDevice(dev));
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  float* sink;
  ClockRec* rec;
  CK(cudaMalloc(&sink, 64));
  CK(cudaMalloc(&rec, sizeof(ClockRec)));
  const int blocks = sms * 4, threads = 512;          // 2048 threads per SM
  const double nthreads = static_cast<double>(blocks) * threads;
  printf("{\"sms\": %d", sms);

  {
    const int iters = 1 << 16;
    Timed r = timed([&] { k_ex2<<<blocks, threads>>>(iters, sink, rec); }, rec);
    const double ops = nthreads * iters * 16.0;
    printf(", \"mufu_ex2\": {\"per_s\": %.4e, \"per_clk_sm\": %.3f, \"sm_mhz\": %.0f, \"ms\": %.2f}", ops / (r.ms * 1e-3),
           ops / (r.ms * 1e-3) / (r.mhz * 1e6) / sms, r.mhz, r.ms);
  }
  {
    const int iters = 1 << 15;
    Timed r = timed([&] { k_ffma<<<blocks, threads>>>(iters, sink, rec); }, rec);
    const double ops = nthreads * iters * 32.0;
    printf(", \"ffma\": {\"lane_ops_per_s\": %.4e, \"lanes_per_clk_sm\": %.2f, \"sm_mhz\": %.0f, \"ms\": %.2f}",
           ops / (r.ms * 1e-3), ops / (r.ms * 1e-3) / (r.mhz * 1e6) / sms, r.mhz, r.ms);
  }
  {
    const int iters = 1 << 15;
    Timed r = timed([&] { k_ffma2<<<blocks, threads>>>(iters, sink, rec); }, rec);
    const double inst = nthreads * iters * 32.0;       // thread-level packed instructions (2 lanes each)
    printf(", \"ffma2\": {\"thread_inst_per_s\": %.4e, \"thread_inst_per_clk_sm\": %.2f, \"warp_inst_per_clk_sm\": %.3f, "
           "\"sm_mhz\": %.0f, \"ms\": %.2f}",
           inst / (r.ms * 1e-3), inst / (r.ms * 1e-3) / (r.mhz * 1e6) / sms, inst / 32 / (r.ms * 1e-3) / (r.mhz * 1e6) / sms,
           r.mhz, r.ms);
  }
  {
    const int iters = 1 << 13;
    Timed r = timed([&] { k_epimix<<<blocks, threads>>>(iters, sink, rec); }, rec);
    const double pairs = nthreads * iters * 4.0 * 2.0;  // kernel pairs (matrix elements) evaluated
    printf(", \"ladder_epilogue_mix\": {\"pairs_per_s\": %.4e, \"pairs_per_clk_sm\": %.3f, \"sm_mhz\": %.0f, \"ms\": %.2f, "
           "\"note\": \"2 ex2 + 17 packed fp32 per 2 pairs, no tcgen05.ld / bf16 pack / smem store\"}",
           pairs / (r.ms * 1e-3), pairs / (r.ms * 1e-3) / (r.mhz * 1e6) / sms, r.mhz, r.ms);
  }
  {
    const size_t bytes = 48ull << 20;                   // well inside the 126 MB L2
    uint4* buf;
    CK(cudaMalloc(&buf, bytes));
    CK(cudaMemset(buf, 1, bytes));
    uint32_t* isink = reinterpret_cast<uint32_t*>(sink);
    const int passes = 400;
    const int l2_blocks = sms * 4;
    const size_t stride = static_cast<size_t>(l2_blocks) * threads;
    const size_t per_thread = bytes / 16 / stride / 4 * 4;
    Timed r = timed([&] { k_l2read<<<l2_blocks, threads>>>(buf, per_thread, passes, isink, rec); }, rec);
    const double moved = static_cast<double>(passes) * static_cast<double>(per_thread * stride) * 16.0;
    printf(", \"l2_read\": {\"GBps\": %.1f, \"bytes_per_clk_sm\": %.1f, \"sm_mhz\": %.0f, \"ms\": %.2f, \"buffer_MiB\": 48}",
           moved / (r.ms * 1e-3) / 1e9, moved / (r.ms * 1e-3) / (r.mhz * 1e6) / sms, r.mhz, r.ms);
    CK(cudaFree(buf));
  }
  {
    const int smem = 2 * 48 * 1024;
    CK(cudaFuncSetAttribute(k_umma<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    CK(cudaFuncSetAttribute(k_umma<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const int batches = 12000;                           // x 32 instructions of 128 cycles: ~50 M cycles at full rate
    for (int kind = 0; kind < 2; ++kind) {
      Timed r = timed([&] {
        if (kind == 0) k_umma<0><<<sms, 128, smem>>>(batches, rec);
        else k_umma<1><<<sms, 128, smem>>>(batches, rec);
      }, rec);
      const double kper = kind == 0 ? 16.0 : 8.0;
      const double flops = static_cast<double>(sms) * batches * 32.0 * 2.0 * 128.0 * 256.0 * kper;
      printf(", \"%s\": {\"tflops\": %.1f, \"flops_per_clk_sm\": %.0f, \"sm_mhz\": %.0f, \"ms\": %.2f, "
             "\"shape\": \"cta_group::1 M128 N256 K%d, operands resident in smem\"}",
             kind == 0 ? "umma_bf16" : "umma_tf32", flops / (r.ms * 1e-3) / 1e12,
             flops / (r.ms * 1e-3) / (r.mhz * 1e6) / sms, r.mhz, r.ms, static_cast<int>(kper));
    }
  }
  // resident clusters of the gradient kernel's shape (512 threads, ~225 KiB shared memory) per cluster size
  {
    const int smem = 225 * 1024;
    CK(cudaFuncSetAttribute(k_dummy, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    printf(", \"resident_clusters_512thr_225KiB\": {");
    const int sizes[5] = {1, 2, 4, 6, 8};
    for (int i = 0; i < 5; ++i) {
      cudaLaunchConfig_t cfg{};
      cfg.blockDim = dim3(512);
      cfg.gridDim = dim3(static_cast<unsigned>(sizes[i] * 64 * 6));
      cfg.dynamicSmemBytes = smem;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = static_cast<unsigned>(sizes[i]);
      attr[0].val.clusterDim.y = 1;
      attr[0].val.clusterDim.z = 1;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      int ncl = 0;
      cudaError_t e = cudaOccupancyMaxActiveClusters(&ncl, k_dummy, &cfg);
      if (e != cudaSuccess) { ncl = -1; cudaGetLastError(); }
      printf("%s\"%d\": %d", i ? ", " : "", sizes[i], ncl);
    }
    printf("}");
  }
  printf("}\n");
  return 0;
}
