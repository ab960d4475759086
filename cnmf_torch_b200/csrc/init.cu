// 'random' initialisation of a stack of replicates, drawn on the device straight into the engine's layout.
//
// run_nmf(init='random') (external nmf-torch, called at /root/reference/src/cnmf/cnmf.py:819 with
// random_state = the ledger's nmf_seed, cnmf.py:885) seeds torch with the replicate's seed and draws
//     H = |sqrt(mean(X)/K) * randn(N, K)|   then   W = |sqrt(mean(X)/K) * randn(K, G)|
// on the device it computes on.  Under use_gpu=True that is torch's CUDA generator: Philox4_32_10 through cuRAND's
// device API, one subsequence per thread of a grid-stride kernel whose geometry depends on the element count and
// on the device (ATen/native/cuda/DistributionTemplates.h: calc_execution_policy +
// distribution_elementwise_grid_stride_kernel with curand_normal4, unroll 4).  This kernel restates that geometry
// and calls the same cuRAND device functions, so every value is bit-identical to what
//     g = torch.Generator('cuda'); g.manual_seed(seed); torch.randn((N, K), generator=g, device='cuda'); torch.randn((K, G), ...)
// yields on the same GPU (tests/test_gpu_init.py pins it against torch itself) -- for all replicates of a stack in
// ONE launch, written transposed into Ht / Wall with the scale and |.| applied, instead of two torch launches per
// replicate plus a permuting copy (1200 launches, 10.6 ms of a 52 ms end-to-end step at cfg-2).
#include <curand_kernel.h>

#include "common.cuh"

namespace cnmf {

constexpr int IR_BLOCK = 256;  // torch's block_size_bound

struct InitGeom {
  uint32_t grid;       // virtual blocks torch would launch for this element count
  uint64_t offset_inc; // Philox offset consumed (counter_offset)
};

static InitGeom init_geom(int64_t numel, int sms, int max_threads_per_sm) {
  InitGeom g;
  uint64_t blocks = (static_cast<uint64_t>(numel) + IR_BLOCK - 1) / IR_BLOCK;
  const uint64_t cap = static_cast<uint64_t>(sms) * (max_threads_per_sm / IR_BLOCK);
  g.grid = static_cast<uint32_t>(blocks < cap ? blocks : cap);
  g.offset_inc = ((static_cast<uint64_t>(numel) - 1) / (static_cast<uint64_t>(IR_BLOCK) * g.grid * 4) + 1) * 4;
  return g;
}

// blockIdx.y = replicate; blockIdx.x = virtual block of the H draw, then of the W draw
__global__ void __launch_bounds__(IR_BLOCK) init_random_kernel(const int64_t* __restrict__ seeds, int K, int Kp,
                                                               int64_t n_total, int64_t n_local, int64_t row_offset,
                                                               int64_t G, float avg, float* __restrict__ Ht, int64_t ldh,
                                                               float* __restrict__ Wall, int64_t ldw, uint32_t gridH,
                                                               uint32_t gridW, uint64_t offW) {
  const int r = blockIdx.y;
  uint32_t vb = blockIdx.x;
  const bool isW = vb >= gridH;
  if (isW) vb -= gridH;
  const int64_t numel = isW ? static_cast<int64_t>(K) * G : n_total * K;
  const int64_t T = static_cast<int64_t>(isW ? gridW : gridH) * IR_BLOCK;
  const int64_t idx = static_cast<int64_t>(vb) * IR_BLOCK + threadIdx.x;
  curandStatePhilox4_32_10_t st;
  curand_init(static_cast<unsigned long long>(seeds[r]), static_cast<unsigned long long>(idx), isW ? offW : 0ull, &st);
  const int64_t rounded = ((numel - 1) / (T * 4) + 1) * T * 4;
  for (int64_t li0 = idx; li0 < rounded; li0 += T * 4) {
    const float4 z = curand_normal4(&st);
#pragma unroll
    for (int ii = 0; ii < 4; ++ii) {
      const int64_t li = li0 + T * ii;
      if (li >= numel) continue;
      const float v = fabsf((&z.x)[ii] * avg);   // randn * avg, then abs: two fp32 ops in torch, the same two here
      if (isW) {
        const int64_t k = li / G, j = li - k * G;
        Wall[(static_cast<int64_t>(r) * Kp + k) * ldw + j] = v;
      } else {
        const int64_t i = li / K, k = li - i * K;
        const int64_t il = i - row_offset;      // row-sharded engines keep their block of the global draw
        if (il >= 0 && il < n_local) Ht[(static_cast<int64_t>(r) * Kp + k) * ldh + il] = v;
      }
    }
  }
}

}  // namespace cnmf

using namespace cnmf;

extern "C" int cnmf_b200_init_random(const int64_t* seeds, int32_t R, int32_t K, int32_t Kp, int64_t n_total,
                                     int64_t n_local, int64_t row_offset, int64_t n_genes, float avg, float* Ht,
                                     int64_t ldh, float* Wall, int64_t ldw, void* stream) {
  CNMF_NVTX("cnmf_b200_init_random");
  CNMF_CHECK_ARG(seeds && Ht && Wall, "init_random: null pointer");
  CNMF_CHECK_ARG(R > 0 && K > 0 && Kp >= K && n_total > 0 && n_local > 0 && n_genes > 0 && row_offset >= 0 &&
                     row_offset + n_local <= n_total && ldh >= n_local && ldw >= n_genes,
                 "init_random: bad shape");
  CNMF_CHECK_ARG(n_total * K < (1ll << 31) && static_cast<int64_t>(K) * n_genes < (1ll << 31),
                 "init_random: a tensor of 2^31 or more elements is drawn by torch in 32-bit-indexable pieces; "
                 "draw it with torch (engine.make_init)");
  CNMF_CHECK_ARG(R <= 65535, "init_random: at most 65535 replicates per call");
  int dev = 0, sms = 0, tpsm = 0;
  CNMF_CUDA(cudaGetDevice(&dev));
  CNMF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  CNMF_CUDA(cudaDeviceGetAttribute(&tpsm, cudaDevAttrMaxThreadsPerMultiProcessor, dev));
  const InitGeom gh = init_geom(n_total * K, sms, tpsm), gw = init_geom(static_cast<int64_t>(K) * n_genes, sms, tpsm);
  dim3 grid(gh.grid + gw.grid, static_cast<unsigned>(R));
  init_random_kernel<<<grid, IR_BLOCK, 0, (cudaStream_t)stream>>>(seeds, K, Kp, n_total, n_local, row_offset, n_genes,
                                                                  avg, Ht, ldh, Wall, ldw, gh.grid, gw.grid,
                                                                  gh.offset_inc);
  CNMF_LAUNCH_CHECK();
  return 0;
}
