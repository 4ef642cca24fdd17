// Row (f)2: the spectrogram-condition front end of the DiffEEG denoiser -- the consumer of the path.
//
// Replaces, for the condition branch of DiffEEG.forward,
//   /root/reference/root/src/models/diffEEG.py:47-53   ConvTranspose2d(C -> H2, kernel (3, 3), stride (1, 8), padding (1, 2))
//   /root/reference/root/src/models/diffEEG.py:108-110 spec = upsample1(spectrogram); relu; flatten(start_dim=2);
//                                                       F.interpolate(size=T, mode="linear", align_corners=False)
//   /root/reference/root/src/models/diffEEG.py:136-145 within-class mixing (training), partner index made explicit.
//
// What the reference materialises: the transposed convolution blows the (C, 33, 2000) image up to
// (H2, 33, 15 991), flattens it to 527 703 samples per output channel and then keeps TWO neighbours of
// it per output sample (linear interpolation down to T = 2000): 4 000 of 527 703 values, 0.76 %.
// Stride 8 against a 3-wide kernel also means that a convolution output column t has at most ONE
// kernel column (kt = (t + 2) mod 8, if < 3); the other 5 / 8 of the columns hold relu(bias) only.
//
// So the operator is a gather: per output sample j two flattened positions i0, i0 + 1 -> (f, t) ->
// kt, t' = (t + 2 - kt) / 8, and a contraction over C x 3 image values in[c, f + 1 - kf, t'] with the
// weight slice w[c, :, kf, kt].  3 / 8 x 4 000 positions x 57 values: the kernel reads 18 % of the image
// (as 32-byte sectors: about half of its bytes) instead of all of it, writes (H2, T), and never forms the
// 135 MB intermediate per record.  The contraction (K = 57, N = H2 = 16 in the shipped config) is
// 11 MFLOP per record next to 2.7 MB of sector traffic: HBM-bound gather, FFMA is the right pipe.
//
// Interpolation indices are ATen's, bit for bit: scale = float(in_len) / float(T),
// src = scale * (j + 0.5f) - 0.5f in fp32 (no contraction), i0 = int(src), lambda = src - i0.
#include "common.cuh"
#include "internal.h"

namespace eegspec {

constexpr int kCondJ = 128;         // output samples per CTA
constexpr int kCondThreads = 256;
constexpr int kCondKw = 3, kCondKh = 3, kCondStrideT = 8, kCondPadT = 2, kCondPadF = 1;

struct CondArgs {
  const float* spec;      // (B, C, F, Tc)
  const float* weight;    // (C, H2, 3, 3)  ConvTranspose2d layout
  const float* bias;      // (H2)
  const int* partner;     // (B) or null
  float* out;             // (B, H2, T)
  int n_ch, n_freqs, n_cols, h2, out_len;
  int t_out;              // (n_cols - 1) * 8 - 2 * 2 + 3: columns of the (never materialised) convolution output
  int in_len;             // n_freqs * t_out
  int k_pad;              // n_ch * 3 rounded up to an odd number (bank-conflict-free rows)
  float scale, alpha;
};

__global__ void __launch_bounds__(kCondThreads) spec_condition_kernel(const CondArgs a) {
  extern __shared__ __align__(16) float sm[];
  const int K = a.n_ch * kCondKh;
  float* A = sm;                                        // [2 kCondJ][k_pad]    gathered image values per position
  float* W = A + 2 * kCondJ * a.k_pad;                  // [3][K][h2]           w[c][h][kf][kt] -> W[kt][c * 3 + kf][h]
  __shared__ float lam[kCondJ];
  __shared__ int kts[2 * kCondJ];                       // kernel column of position p = which * kCondJ + jj, -1 = bias only
  __shared__ int fs[2 * kCondJ], tqs[2 * kCondJ];
  __shared__ int live[2 * kCondJ], n_live;              // positions that read the image (3 / 8 of them)

  const int tid = threadIdx.x;
  const int tiles = (a.out_len + kCondJ - 1) / kCondJ;
  const long long rec = blockIdx.x / tiles;
  const int j0 = (int)(blockIdx.x - rec * tiles) * kCondJ;
  const long long img = (long long)a.n_ch * a.n_freqs * a.n_cols;
  const float* sa = a.spec + rec * img;
  const float* sb = a.partner ? a.spec + (long long)a.partner[rec] * img : nullptr;
  const bool mix = sb != nullptr && sb != sa;

  if (tid == 0) n_live = 0;
  for (int i = tid; i < kCondKw * K * a.h2; i += kCondThreads) {
    const int h = i % a.h2, k = (i / a.h2) % K, kt = i / (a.h2 * K);
    const int c = k / kCondKh, kf = k - c * kCondKh;
    W[i] = __ldg(a.weight + (((long long)c * a.h2 + h) * kCondKh + kf) * kCondKw + kt);
  }
  __syncthreads();
  // ---- positions of this CTA's samples (ATen's linear / align_corners=False arithmetic in fp32)
  {
    const int jj = tid & (kCondJ - 1), which = tid >> 7;          // 256 threads = 2 x kCondJ positions
    const int j = j0 + jj;
    float src = __fsub_rn(__fmul_rn(a.scale, (float)j + 0.5f), 0.5f);
    src = src < 0.f ? 0.f : src;
    int i0 = (int)src;
    if (i0 > a.in_len - 1) i0 = a.in_len - 1;
    if (which == 0) lam[jj] = src - (float)i0;
    const int i = i0 + ((which && i0 < a.in_len - 1) ? 1 : 0);
    const int f = i / a.t_out, t = i - f * a.t_out;
    const int kt = (t + kCondPadT) % kCondStrideT;
    const int tq = (t + kCondPadT) / kCondStrideT;
    const bool is_live = j < a.out_len && kt < kCondKw && tq < a.n_cols;
    kts[tid] = is_live ? kt : -1;
    fs[tid] = f;
    tqs[tid] = tq;
    if (is_live) live[atomicAdd(&n_live, 1)] = tid;
  }
  __syncthreads();
  // ---- gather: A[p][c * 3 + kf] = in[c, f + 1 - kf, t'] (0 outside the image).  A thread owns ONE k (its channel /
  // kernel-row offset is computed once) and walks the live positions four apart: the loads of the unrolled
  // iterations are independent and in flight together.
  {
    const int k = tid & 63, lane_p = tid >> 6;                    // 64 k slots (K <= 64 served per pass) x 4 position lanes
    const int nl = n_live;
    for (int k0 = 0; k0 < K; k0 += 64) {
      const int kk = k0 + k;
      if (kk < K) {
        const int c = kk / kCondKh, kf = kk - c * kCondKh;
        const long long ch_off = (long long)c * a.n_freqs * a.n_cols;
        const int df = kCondPadF - kf;
#pragma unroll 4
        for (int u = lane_p; u < nl; u += 4) {
          const int p = live[u];
          const int f = fs[p] + df;
          float v = 0.f;
          if (f >= 0 && f < a.n_freqs) {
            const long long off = ch_off + (long long)f * a.n_cols + tqs[p];
            v = __ldg(sa + off);
            if (mix) v = a.alpha * v + (1.f - a.alpha) * __ldg(sb + off);     // diffEEG.py:143
          }
          A[p * a.k_pad + kk] = v;
        }
      }
    }
  }
  __syncthreads();
  // ---- contraction + relu + interpolation: thread = (sample jj, output channels hg, hg + 2, ...), 4 at a time
  const int jj = tid & (kCondJ - 1), hg = tid >> 7;
  const int j = j0 + jj;
  if (j >= a.out_len) return;
  const int kt0 = kts[jj], kt1 = kts[kCondJ + jj];
  const float l1 = lam[jj], l0 = 1.f - l1;
  const float* A0 = A + jj * a.k_pad;
  const float* A1 = A + (kCondJ + jj) * a.k_pad;
  for (int h0 = 4 * hg; h0 < a.h2; h0 += 8) {
    float v0[4], v1[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) v0[i] = v1[i] = (h0 + i < a.h2) ? __ldg(a.bias + h0 + i) : 0.f;
    if (kt0 >= 0) {
      const float* w = W + (long long)kt0 * K * a.h2 + h0;
      for (int k = 0; k < K; ++k) {
        const float x = A0[k];
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (h0 + i < a.h2) v0[i] = fmaf(x, w[k * a.h2 + i], v0[i]);
      }
    }
    if (kt1 >= 0) {
      const float* w = W + (long long)kt1 * K * a.h2 + h0;
      for (int k = 0; k < K; ++k) {
        const float x = A1[k];
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (h0 + i < a.h2) v1[i] = fmaf(x, w[k * a.h2 + i], v1[i]);
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (h0 + i < a.h2)
        a.out[(rec * a.h2 + h0 + i) * a.out_len + j] = l0 * fmaxf(v0[i], 0.f) + l1 * fmaxf(v1[i], 0.f);
  }
}

int spec_condition_launch(const float* spec, long long batch, const eegspec_cond_params* p, const float* weight,
                          const float* bias, const int* partner, float* out, cudaStream_t st) {
  if (!spec || !weight || !bias || !out || !p || batch <= 0) return EEGSPEC_EINVAL;
  if (p->n_channels <= 0 || p->n_freqs <= 0 || p->n_cols <= 0 || p->out_channels <= 0 || p->out_length <= 0)
    return EEGSPEC_EINVAL;
  if (p->n_channels > 64 || p->out_channels > 256) return EEGSPEC_EUNSUPPORTED;
  CondArgs a;
  a.spec = spec;
  a.weight = weight;
  a.bias = bias;
  a.partner = partner;
  a.out = out;
  a.n_ch = p->n_channels;
  a.n_freqs = p->n_freqs;
  a.n_cols = p->n_cols;
  a.h2 = p->out_channels;
  a.out_len = p->out_length;
  const long long t_out = (long long)(p->n_cols - 1) * kCondStrideT - 2 * kCondPadT + kCondKw;
  const long long in_len = t_out * p->n_freqs;   // frequency rows: (F - 1) * 1 - 2 * 1 + 3 = F
  if (t_out <= 0 || in_len > (1LL << 30)) return EEGSPEC_EUNSUPPORTED;
  a.t_out = (int)t_out;
  a.in_len = (int)in_len;
  a.k_pad = (p->n_channels * kCondKh) | 1;
  a.scale = (float)a.in_len / (float)a.out_len;   // ATen: area_pixel_compute_scale<float>
  a.alpha = p->mix_alpha;
  const int K = a.n_ch * kCondKh;
  const size_t smem = sizeof(float) * ((size_t)2 * kCondJ * a.k_pad + (size_t)kCondKw * K * a.h2);
  if (smem > 200 * 1024) return EEGSPEC_EUNSUPPORTED;
  if (smem > 48 * 1024) {
    const cudaError_t e = cudaFuncSetAttribute(spec_condition_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  const long long ctas = batch * ((a.out_len + kCondJ - 1) / kCondJ);
  if (ctas > 0x7fffffffLL) return EEGSPEC_EUNSUPPORTED;
  spec_condition_kernel<<<(unsigned)ctas, kCondThreads, smem, st>>>(a);
  count_launch();
  return (int)cudaGetLastError();
}

}  // namespace eegspec
