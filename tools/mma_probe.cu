// Micro-benchmark (run on the GPU box): cost of the mode-N tcgen05.mma stream in isolation.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I include -o /tmp/mma_probe tools/mma_probe.cu
// One CTA per SM, one thread issues `reps` tiles of (groups x terms x ks) kind::f16 MMAs with
// M = 128, N = n, K = 16 from no-swizzle K-major shared-memory operands, commit + wait per tile.
// Variants: N, same / alternating accumulator, operand layout (LBO/SBO), optional smem traffic.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "../multimodal-brain-pattern-identification_xai_b200/csrc/melspec_chain.cuh"
using namespace eegspec;

__device__ __forceinline__ uint64_t desc_of(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t swz) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)(lbo >> 4) << 16;
  d |= (uint64_t)(sbo >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)swz << 61;
  return d;
}

struct P { int n, mmas_per_tile, reps, alt, lbo, sbo, swz, kstep_bytes, noise; };

__global__ void __launch_bounds__(256, 1) probe(P p, long long* out) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ unsigned long long bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;  // fp16 1.0
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(c_smem_u32(&tmem_base_s)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  if (tid == 0) { c_mbar_init(c_smem_u32(&bar), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::); }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::);
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::);
  const uint32_t tmem = tmem_base_s;
  const uint32_t idesc = (1u << 4) | ((uint32_t)(p.n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  if (warp == 0) {
    if (lane == 0) {
      const uint32_t a0 = c_smem_u32(smem), b0 = a0 + 64 * 1024;
      long long t0 = clock64();
      for (int r = 0; r < p.reps; ++r) {
        for (int i = 0; i < p.mmas_per_tile; ++i) {
          const int ks = i & 3, term = (i >> 2) % 3, grp = i / 12;
          const uint32_t aa = a0 + ((grp * 2 + (term == 1)) & 3) * 16384 + ks * p.kstep_bytes;
          const uint32_t bb = b0 + ((grp * 2 + (term == 2)) & 3) * 16384 + ks * p.kstep_bytes;
          const uint32_t d = tmem + (p.alt ? (r & 1) * 224 : 0) + (grp & 1) * (p.n <= 112 ? 112 : 0);
          c_umma(d, desc_of(aa, p.lbo, p.sbo, p.swz), desc_of(bb, p.lbo, p.sbo, p.swz), (i % 12) != 0, idesc);
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(c_smem_u32(&bar)) : "memory");
        c_mbar_wait(c_smem_u32(&bar), r & 1);
      }
      long long t1 = clock64();
      if (blockIdx.x == 0) out[0] = t1 - t0;
    }
  } else if (p.noise) {
    // other warps hammer shared memory (LDS/STS on the upper 32 KB) while the MMAs run
    volatile float* z = reinterpret_cast<volatile float*>(smem + 128 * 1024);
    float acc = 0.f;
    for (int it = 0; it < p.noise; ++it) {
      acc += z[(tid * 33 + it) & 8191];
      z[(tid + it * 32) & 8191] = acc;
    }
    if (acc == 123.f) out[1] = 1;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::);
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}

int main() {
  long long* d;
  cudaMalloc(&d, 64);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
  struct { const char* name; P p; } cases[] = {
      {"N112 K-major noswz LBO128 SBO1024 (kernel)", {112, 24, 50, 1, 128, 1024, 0, 256, 0}},
      {"same, single accumulator", {112, 24, 50, 0, 128, 1024, 0, 256, 0}},
      {"N128", {128, 24, 50, 0, 128, 1024, 0, 256, 0}},
      {"N224 (12 per tile)", {224, 12, 50, 0, 128, 1024, 0, 256, 0}},
      {"N256 (12 per tile)", {256, 12, 50, 0, 128, 1024, 0, 256, 0}},
      {"N64", {64, 24, 50, 0, 128, 1024, 0, 256, 0}},
      {"N16", {16, 24, 50, 0, 128, 1024, 0, 256, 0}},
      {"N8", {8, 24, 50, 0, 128, 1024, 0, 256, 0}},
  };
  for (auto& c : cases) {
    cudaMemset(d, 0, 64);
    probe<<<148, 256, 160 * 1024>>>(c.p, d);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[2];
    cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    const double per = (double)h[0] / (c.p.reps * c.p.mmas_per_tile);
    printf("%-48s %s  %8.1f cycles/MMA  floor %5.1f  per tile %8.0f\n", c.name, cudaGetErrorString(e), per,
           128.0 * c.p.n / 256.0, (double)h[0] / c.p.reps);
  }
  return 0;
}
