// FP64 pipe probe (B200): dependent-chain latency and one-warp issue interval of DADD / DMUL / DFMA.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_probe tools/fp64_latency_probe.cu && ./fp64_probe
// Row f4 (csrc/classifier_eeg.cu) is a chain y -> y a1 -> z0 -> next y of unfused float64 operations; this is the
// measurement its roofline paragraph in DESIGN.md quotes.
#include <cstdio>
#include <cuda_runtime.h>

template <int OP, int ILP>
__global__ void chain(double* out, long long* cycles, int iters, double a, double b) {
  double v[ILP];
  for (int i = 0; i < ILP; ++i) v[i] = a + i + threadIdx.x;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 16; ++u) {
#pragma unroll
      for (int i = 0; i < ILP; ++i) {
        if (OP == 0) v[i] = __dadd_rn(v[i], b);
        if (OP == 1) v[i] = __dmul_rn(v[i], b);
        if (OP == 2) v[i] = fma(v[i], b, a);
      }
    }
  }
  long long t1 = clock64();
  double s = 0;
  for (int i = 0; i < ILP; ++i) s += v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int OP, int ILP>
void run(const char* name, int warps) {
  double* out; long long* cyc;
  cudaMalloc(&out, sizeof(double) * 1024 * 148); cudaMalloc(&cyc, 8);
  const int iters = 2000;
  chain<OP, ILP><<<148, 32 * warps>>>(out, cyc, iters, 1.0, 1.0000001);
  chain<OP, ILP><<<148, 32 * warps>>>(out, cyc, iters, 1.0, 1.0000001);
  long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  printf("%-5s ILP %2d, %2d warps/SM: %7.2f cycles per dependent step, %6.2f cycles per warp instruction\n", name, ILP, warps,
         (double)c / (iters * 16.0), (double)c / (iters * 16.0 * ILP));
  cudaFree(out); cudaFree(cyc);
}

// The recurrence of row f4 itself: one direct-form-II-transposed step per sample, unfused, SciPy's order.
struct Coef { double b[13], a[13]; };
template <int N, int MODE>   // MODE 0: coefficients from the kernel parameter (constant bank), 1: copied to registers first
__global__ void df2t(double* out, long long* cycles, int iters, const __grid_constant__ Coef c, double x0) {
  double z[N - 1];
  for (int i = 0; i < N - 1; ++i) z[i] = 0.0;
  double b[N], a[N];
  for (int i = 0; i < N; ++i) { b[i] = c.b[i]; a[i] = c.a[i]; }
  double x = x0 + threadIdx.x, acc = 0.0;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const double* bb = MODE ? b : c.b;
      const double* aa = MODE ? a : c.a;
      const double y = __dadd_rn(z[0], __dmul_rn(bb[0], x));
#pragma unroll
      for (int i = 0; i < N - 2; ++i) {
        const double t = ((i + 1) & 1) ? z[i + 1] : __dadd_rn(z[i + 1], __dmul_rn(x, bb[i + 1]));
        z[i] = __dsub_rn(t, __dmul_rn(y, aa[i + 1]));
      }
      z[N - 2] = __dsub_rn(__dmul_rn(x, bb[N - 1]), __dmul_rn(y, aa[N - 1]));
      acc += y;
      x += 0.25;
    }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}
template <int N, int MODE>
void run_df2t(const char* name, int warps) {
  double* out; long long* cyc;
  cudaMalloc(&out, sizeof(double) * 1024 * 148); cudaMalloc(&cyc, 8);
  Coef c;
  for (int i = 0; i < 13; ++i) { c.b[i] = (i & 1) ? 0.0 : 1e-3 * (i + 1); c.a[i] = i ? 0.05 / (i + 1) : 1.0; }
  const int iters = 500;
  df2t<N, MODE><<<148, 32 * warps>>>(out, cyc, iters, c, 1.0);
  df2t<N, MODE><<<148, 32 * warps>>>(out, cyc, iters, c, 1.0);
  long long cy; cudaMemcpy(&cy, cyc, 8, cudaMemcpyDeviceToHost);
  printf("df2t N=%d %s, %2d warps/SM: %7.1f cycles per sample\n", N, name, warps, (double)cy / (iters * 8.0));
  cudaFree(out); cudaFree(cyc);
}

int main() {
  run_df2t<11, 0>("coefficients in the constant bank", 1); run_df2t<11, 1>("coefficients in registers", 1);
  run_df2t<11, 0>("coefficients in the constant bank", 4); run_df2t<11, 0>("coefficients in the constant bank", 8);
  run_df2t<13, 0>("coefficients in the constant bank", 4); run_df2t<13, 0>("coefficients in the constant bank", 8);
  run_df2t<13, 0>("coefficients in the constant bank", 16);
  run<0, 1>("DADD", 1); run<1, 1>("DMUL", 1); run<2, 1>("DFMA", 1);
  run<0, 2>("DADD", 1); run<0, 4>("DADD", 1); run<0, 8>("DADD", 1); run<0, 16>("DADD", 1);
  run<0, 8>("DADD", 4); run<0, 8>("DADD", 8); run<0, 8>("DADD", 16); run<2, 8>("DFMA", 16);
  return 0;
}
