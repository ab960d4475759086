// Issue / pipe rates of the SIMT instructions the sparse KL half-step is made of, on sm_100a:
// FFMA (3 registers), FFMA2 (fma.rn.f32x2), FHFMA (fma.rn.f32.f16: fp16 x fp16 + fp32), HADD2.F32 (half -> float),
// LDS.128 / LDS.64 / LDS.32 under the gather patterns considered for the shared-memory slab.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o simt_rate simt_rate.cu && ./simt_rate
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

constexpr int ITERS = 4096;
constexpr int NACC = 8;

__global__ void k_ffma(float* out, float a, float b) {
  float acc[NACC];
  for (int i = 0; i < NACC; ++i) acc[i] = threadIdx.x + i;
  float x = a + threadIdx.x, y = b;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) acc[i] = fmaf(acc[i], x, y);
  }
  float s = 0;
  for (int i = 0; i < NACC; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__device__ __forceinline__ unsigned long long ffma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long d;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__global__ void k_ffma2(float* out, float a, float b) {
  unsigned long long acc[NACC];
  for (int i = 0; i < NACC; ++i) { float2 v = make_float2(threadIdx.x + i, i); acc[i] = *reinterpret_cast<unsigned long long*>(&v); }
  float2 xv = make_float2(a + threadIdx.x, a), yv = make_float2(b, b + 1);
  unsigned long long x = *reinterpret_cast<unsigned long long*>(&xv), y = *reinterpret_cast<unsigned long long*>(&yv);
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) acc[i] = ffma2(acc[i], x, y);
  }
  float s = 0;
  for (int i = 0; i < NACC; ++i) { float2 v = *reinterpret_cast<float2*>(&acc[i]); s += v.x + v.y; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__device__ __forceinline__ float fhfma(unsigned short a, unsigned short b, float c) {
  asm volatile("fma.rn.f32.f16 %0, %1, %2, %0;" : "+f"(c) : "h"(a), "h"(b));
  return c;
}
__global__ void k_fhfma(float* out, const uint32_t* in) {
  float acc[NACC];
  for (int i = 0; i < NACC; ++i) acc[i] = threadIdx.x + i;
  const uint32_t p = in[threadIdx.x], q = in[threadIdx.x + 32];
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; i += 2) {
      acc[i] = fhfma(p & 0xffff, q & 0xffff, acc[i]);
      acc[i + 1] = fhfma(p >> 16, q >> 16, acc[i + 1]);
    }
  }
  float s = 0;
  for (int i = 0; i < NACC; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// half -> float conversions feeding FFMA: the cost of unpacking a gathered fp16 operand
__global__ void k_cvt_ffma(float* out, const uint32_t* in) {
  float acc[NACC];
  for (int i = 0; i < NACC; ++i) acc[i] = threadIdx.x + i;
  uint32_t p = in[threadIdx.x];
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; i += 2) {
      const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&p));
      acc[i] = fmaf(acc[i], f.x, f.y);
      acc[i + 1] = fmaf(acc[i + 1], f.y, f.x);
      p += 0x00010001u;   // keeps the conversion inside the loop
    }
  }
  float s = 0;
  for (int i = 0; i < NACC; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// shared-memory gathers.  smem = 200 KB of words; pattern: lane group of `grp` lanes shares a random record
// (as lanes of one row share a gene), record stride `rec_words`, each lane reads VEC words at its slot.
template <int VEC>
__global__ void k_lds(float* out, const int* recs, int nrec_pool, int grp, int rec_words, int lane_words) {
  extern __shared__ __align__(16) uint32_t sm[];
  for (int i = threadIdx.x; i < 50 * 1024; i += blockDim.x) sm[i] = i;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane / grp, sub = lane % grp;
  uint32_t acc = 0;
  int idx = (warp * 8 + g) % nrec_pool;
  for (int it = 0; it < ITERS; ++it) {
    const int rec = recs[(idx + it * 8) % nrec_pool];     // L1-resident index stream, same for the lanes of a group
    const uint32_t* p = sm + rec * rec_words + sub * lane_words;
    if (VEC == 4) { const uint4 v = *reinterpret_cast<const uint4*>(p); acc += v.x ^ v.y ^ v.z ^ v.w; }
    if (VEC == 2) { const uint2 v = *reinterpret_cast<const uint2*>(p); acc += v.x ^ v.y; }
    if (VEC == 1) { acc += p[0]; }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

int main() {
  int dev = 0;
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, dev));
  const int sms = prop.multiProcessorCount;
  int clk_khz = 0;
  cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, dev);
  printf("%s, %d SMs, nominal %d MHz\n", prop.name, sms, clk_khz / 1000);
  const int threads = 256, blocks = sms * 8;
  float* out;
  CK(cudaMalloc(&out, (size_t)blocks * threads * 4));
  uint32_t* in;
  CK(cudaMalloc(&in, 4096));
  std::vector<uint32_t> hin(1024, 0x3c003c00u);
  CK(cudaMemcpy(in, hin.data(), 4096, cudaMemcpyHostToDevice));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  auto report = [&](const char* name, float ms, double ops_per_thread) {
    const double total = ops_per_thread * blocks * threads;
    printf("%-34s %8.3f ms  %8.1f lane-ops/ns/SM  (%.1f lane-ops/clk/SM at 1.9 GHz)\n", name, ms,
           total / (ms * 1e6) / sms, total / (ms * 1e6) / sms / 1.9);
  };
#define RUN(name, ops, ...)                                   \
  do {                                                        \
    __VA_ARGS__; CK(cudaDeviceSynchronize());                 \
    cudaEventRecord(e0); __VA_ARGS__; cudaEventRecord(e1);    \
    CK(cudaDeviceSynchronize());                              \
    float ms; cudaEventElapsedTime(&ms, e0, e1);              \
    report(name, ms, ops);                                    \
  } while (0)
  RUN("FFMA (3-reg)", (double)ITERS * NACC, k_ffma<<<blocks, threads>>>(out, 1.0001f, 0.5f));
  RUN("FFMA2 (counted as 2 FMA)", (double)ITERS * NACC * 2, k_ffma2<<<blocks, threads>>>(out, 1.0001f, 0.5f));
  RUN("FHFMA (f16 x f16 + f32)", (double)ITERS * NACC, k_fhfma<<<blocks, threads>>>(out, in));
  RUN("HADD2.F32 x2 + FFMA x2 (4 instr)", (double)ITERS * NACC * 2, k_cvt_ffma<<<blocks, threads>>>(out, in));

  // LDS patterns: one CTA of 1024 threads per SM, 200 KB of dynamic shared memory
  const int nrec_pool = 1 << 16;
  std::vector<int> hrec(nrec_pool);
  uint32_t s = 12345;
  int* recs;
  CK(cudaMalloc(&recs, nrec_pool * 4));
  const int smem = 200 * 1024;
  CK(cudaFuncSetAttribute(k_lds<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  CK(cudaFuncSetAttribute(k_lds<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  CK(cudaFuncSetAttribute(k_lds<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  struct Pat { const char* name; int vec, grp, rec_words, lane_words; };
  const Pat pats[] = {
      {"LDS.128 grp 8  rec 128B (conflict-free)", 4, 8, 32, 4},
      {"LDS.128 grp 4  rec 64B", 4, 4, 16, 4},
      {"LDS.128 grp 32 rec 512B (one record)", 4, 32, 128, 4},
      {"LDS.64  grp 8  rec 64B", 2, 8, 16, 2},
      {"LDS.64  grp 16 rec 128B", 2, 16, 32, 2},
      {"LDS.32  grp 8  rec 32B", 1, 8, 8, 1},
      {"LDS.32  grp 4  rec 80B stride 5 words", 1, 4, 20, 5},
      {"LDS.32  grp 32 rec 128B (one record)", 1, 32, 32, 1},
      {"LDS.32  grp 1  random words", 1, 1, 1, 1},
  };
  for (const Pat& p : pats) {
    const int nrec = 50 * 1024 / p.rec_words - 1;
    for (int i = 0; i < nrec_pool; ++i) { s = s * 1664525u + 1013904223u; hrec[i] = (s >> 8) % nrec; }
    CK(cudaMemcpy(recs, hrec.data(), nrec_pool * 4, cudaMemcpyHostToDevice));
    const int lb = sms, lt = 1024;
    float ms = 0;
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      if (p.vec == 4) k_lds<4><<<lb, lt, smem>>>(out, recs, nrec_pool, p.grp, p.rec_words, p.lane_words);
      if (p.vec == 2) k_lds<2><<<lb, lt, smem>>>(out, recs, nrec_pool, p.grp, p.rec_words, p.lane_words);
      if (p.vec == 1) k_lds<1><<<lb, lt, smem>>>(out, recs, nrec_pool, p.grp, p.rec_words, p.lane_words);
      cudaEventRecord(e1);
      CK(cudaDeviceSynchronize());
      cudaEventElapsedTime(&ms, e0, e1);
    }
    const double bytes = (double)ITERS * lb * lt * p.vec * 4;
    printf("%-42s %8.3f ms  %7.1f B/ns/SM (%.1f B/clk/SM at 1.9 GHz), %.2f clk per warp-load\n", p.name, ms,
           bytes / (ms * 1e6) / sms, bytes / (ms * 1e6) / sms / 1.9, ms * 1e6 * 1.9 / ((double)ITERS * lt / 32));
  }
  return 0;
}
