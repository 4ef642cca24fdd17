// Micro-benchmark 2 (GPU box): tcgen05.mma with the A operand in tensor memory (TS form).
//   (1) correctness of the assumed TMEM A layout (lane = row, 32-bit column = 2 consecutive K)
//   (2) cycles per MMA for M = 128, K = 16, N in {104, 112, 64, 48}, A from TMEM, B from smem
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include "../multimodal-brain-pattern-identification_xai_b200/csrc/melspec_chain.cuh"
using namespace eegspec;

__device__ __forceinline__ void umma_ts(uint32_t d, uint32_t a_tmem, uint64_t b_desc, uint32_t acc, uint32_t idesc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n"
      :
      : "r"(d), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(acc)
      : "memory");
}

struct P { int n, mmas, reps, check, noise, ldcols; };

__global__ void __launch_bounds__(640, 1) probe(P p, const __half* A, const __half* B, float* D, long long* out) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ unsigned long long bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < 64 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  __syncthreads();
  // B (n x 16, K-major no-swizzle core matrices)
  for (int i = tid; i < p.n * 16; i += blockDim.x) {
    const int n = i / 16, k = i % 16;
    *reinterpret_cast<__half*>(smem + c_umma_offset(n, k)) = B[i];
  }
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
  const uint32_t a_col = 448;
  // A row = tid: 16 halves -> 8 columns
  if (tid < 128) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(A + tid * 16);
    const uint32_t addr = tmem + ((uint32_t)(32 * warp) << 16) + a_col;
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};\n" ::"r"(addr), "r"(src[0]),
                 "r"(src[1]), "r"(src[2]), "r"(src[3]), "r"(src[4]), "r"(src[5]), "r"(src[6]), "r"(src[7])
                 : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::);
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::);
  const uint32_t idesc = (1u << 4) | ((uint32_t)(p.n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  if (warp >= 4 && p.noise) {
    // epilogue-like traffic: every warp streams ldcols accumulator columns of its lane quadrant, repeatedly
    const uint32_t base = tmem + ((uint32_t)(32 * (warp & 3)) << 16);
    float acc = 0.f;
    long long t0 = clock64();
    for (int it = 0; it < p.noise; ++it) {
      for (int c = 0; c < p.ldcols; c += 16) {
        float re[8], im[8];
        c_tmem_ld8x2(base + c, base + c + 8, re, im);
        acc += re[0] + im[7];
      }
    }
    long long t1 = clock64();
    if (blockIdx.x == 0 && tid == 128) out[1] = t1 - t0;
    if (acc == 123.456f) out[2] = 1;
  }
  if (tid == 0) {
    const uint64_t bd = c_umma_desc(c_smem_u32(smem));
    long long t0 = clock64();
    for (int r = 0; r < p.reps; ++r) {
      for (int i = 0; i < p.mmas; ++i) umma_ts(tmem + (i & 1) * 224, tmem + a_col, bd, p.check ? 0 : (i > 1), idesc);
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(c_smem_u32(&bar)) : "memory");
      c_mbar_wait(c_smem_u32(&bar), r & 1);
    }
    long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
  }
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::);
  if (p.check && blockIdx.x == 0 && tid < 128) {
    for (int c = 0; c < p.n; c += 8) {
      float v[8];
      c_tmem_ld8(tmem + ((uint32_t)(32 * warp) << 16) + c, v);
      for (int i = 0; i < 8; ++i) D[tid * p.n + c + i] = v[i];
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::);
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}

int main() {
  const int NMAX = 128;
  std::vector<__half> hA(128 * 16), hB(NMAX * 16);
  srand(1);
  for (auto& v : hA) v = __float2half((float)(rand() % 7 - 3));
  for (auto& v : hB) v = __float2half((float)(rand() % 7 - 3));
  __half *dA, *dB; float* dD; long long* dO;
  cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dB, hB.size() * 2); cudaMalloc(&dD, 128 * NMAX * 4); cudaMalloc(&dO, 64);
  cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  for (int n : {112, 104, 64, 48}) {
    P p{n, 2, 1, 1, 0, 0};
    cudaMemset(dD, 0, 128 * NMAX * 4);
    probe<<<1, 128, 64 * 1024>>>(p, dA, dB, dD, dO);
    cudaError_t e = cudaDeviceSynchronize();
    std::vector<float> hD(128 * n);
    cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int m = 0; m < 128; ++m)
      for (int c = 0; c < n; ++c) {
        float want = 0.f;
        for (int k = 0; k < 16; ++k) want += __half2float(hA[m * 16 + k]) * __half2float(hB[c * 16 + k]);
        if (want != hD[m * n + c]) ++bad;
      }
    printf("TS check N=%d: %s, %d mismatches of %d\n", n, cudaGetErrorString(e), bad, 128 * n);
    if (e != cudaSuccess) return 1;
  }
  for (int n : {112, 104, 64, 48, 224}) {
    P p{n, 96, 20, 0, 0, 0};
    probe<<<148, 128, 64 * 1024>>>(p, dA, dB, dD, dO);
    cudaError_t e = cudaDeviceSynchronize();
    long long h;
    cudaMemcpy(&h, dO, 8, cudaMemcpyDeviceToHost);
    printf("TS N=%d: %s %8.1f cycles/MMA (floor %.0f)\n", n, cudaGetErrorString(e), (double)h / (p.reps * p.mmas), 128.0 * n / 256);
    if (e != cudaSuccess) return 1;
  }
  // MMA stream next to 16 warps streaming accumulator columns (the epilogue's tcgen05.ld traffic)
  for (int nw : {4, 8, 16}) {
    for (int mm : {0, 96}) {
      P p{104, mm ? mm : 1, mm ? 40 : 1, 0, 60, 208};
      cudaMemset(dO, 0, 64);
      probe<<<148, 128 + 32 * nw, 64 * 1024>>>(p, dA, dB, dD, dO);
      cudaError_t e = cudaDeviceSynchronize();
      long long h[2];
      cudaMemcpy(h, dO, 16, cudaMemcpyDeviceToHost);
      const double bytes = (double)nw * 60 * 208 * 32 * 4;
      printf("noise warps %2d, MMAs %s: %s  %8.1f cycles/MMA;  ld stream %.1f B/clk (%lld cycles)\n", nw, mm ? "on " : "off",
             cudaGetErrorString(e), mm ? (double)h[0] / (p.reps * p.mmas) : 0.0, bytes / (double)h[1], h[1]);
    }
  }
  return 0;
}
