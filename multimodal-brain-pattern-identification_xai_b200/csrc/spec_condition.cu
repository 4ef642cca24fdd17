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
// 11 MFLOP per record: FFMA is the right pipe.
//
// Kernel shape: persistent CTAs (the transposed weights are staged in shared memory once per CTA) walk
// tiles of 32 output samples.  Per tile: (1) the 64 positions are classified and the LIVE ones compacted,
// (2) the image values of the live positions are gathered into rows of A, (3) the contraction runs over the
// live rows only (a warp that mixes live and bias-only positions would pay for both), (4) the two
// neighbours of every sample are blended from V (live) or relu(bias).
//
// Two sources for the image values: the (B, C, F, Tc) image itself, or -- fused with compute_stft --
// the knots of stft_image_kernel's interpolation (F x n_frames {value, delta} per row, 321 KB per record
// instead of 5 MB).  A pixel is then evaluated exactly as phase 3 of that kernel writes it, so both
// sources give the same bits; the image path is bound by random 32-byte DRAM sectors, the knots path
// reads an L2-resident array.
//
// Interpolation indices are ATen's, bit for bit: scale = float(in_len) / float(T),
// src = scale * (j + 0.5f) - 0.5f in fp32 (no contraction), i0 = int(src), lambda = src - i0.
#include "common.cuh"
#include "internal.h"

namespace eegspec {

#ifndef EEGSPEC_COND_J
#define EEGSPEC_COND_J 32
#endif
constexpr int kCondJ = EEGSPEC_COND_J;   // output samples per tile: 32 keeps 7 CTAs (56 warps) of gather loads in flight per SM
constexpr int kCondThreads = 256;
constexpr int kCondKw = 3, kCondKh = 3, kCondStrideT = 8, kCondPadT = 2, kCondPadF = 1;

struct CondArgs {
  const float* spec;      // (B, C, F, Tc) image, or null when `knots` is given
  const float2* knots;    // (B, C, F, n_frames) {normalised value, delta}: the image as compute_stft's kernel holds it
  int n_frames, D;        // knots: image column t' sits at t' * (n_frames - 1) = i * D + rem, D = Tc - 1
  float inv_d;
  const float* weight;    // (C, H2, 3, 3)  ConvTranspose2d layout
  const float* bias;      // (H2)
  const int* partner;     // (B) or null
  float* out;             // (B, H2, T)
  int n_ch, n_freqs, n_cols, h2, out_len;
  int t_out;              // (n_cols - 1) * 8 - 2 * 2 + 3: columns of the (never materialised) convolution output
  int in_len;             // n_freqs * t_out
  long long n_tiles;      // batch * ceil(out_len / J): the CTAs walk them round-robin
  int k_pad;              // n_ch * 3 rounded up to an odd number (bank-conflict-free rows)
  int v_pad;              // h2 | 1
  int w_pad;              // floats per kernel column of the staged weights, = 8 mod 32
  float scale, alpha;
};

// Per live position (one that reads the image): where it sits and what it needs.
struct CondLive {
  int p_kt;      // position slot (which * kCondJ + jj) | kernel column << 16
  int f;         // frequency row of the convolution output
  int base;      // f * row length + image column t' (knots: + the left knot of t's bracket): offset inside a channel
  float fr;      // knots: rem / D of the bracket, exactly phase 3's fraction
};

template <int J>
__global__ void __launch_bounds__(kCondThreads) spec_condition_kernel(const CondArgs a) {
  extern __shared__ __align__(16) float sm[];
  const int K = a.n_ch * kCondKh;
  const int h2p = (a.h2 + 3) & ~3, HQ = h2p >> 2;
  float* A = sm;                                        // [2 J][k_pad]  gathered image values, one row per LIVE position
  float* V = A + 2 * J * a.k_pad;                  // [2 J][v_pad]  relu(conv) of the live positions
  float* W = V + 2 * J * a.v_pad;                  // [3][w_pad]         w[c][h][kf][kt] -> W[kt][(c * 3 + kf) * h2p + h]
  __shared__ float lam[J];
  __shared__ float rb[256];                             // relu(bias): all a bias-only position contributes
  __shared__ int slot[2 * J];                      // position -> row of A / V, -1 = bias only
  __shared__ CondLive live[2 * J];
  __shared__ int n_live;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tiles = (a.out_len + J - 1) / J;
  const long long img = (long long)a.n_ch * a.n_freqs * (a.knots ? a.n_frames : a.n_cols);   // elements per record

  // weights, once per (persistent) CTA: a warp per (kt, k) row, lanes over h (the row's index arithmetic is warp-uniform)
  for (int row = warp; row < kCondKw * K; row += kCondThreads / 32) {
    const int kt = row / K, k = row - kt * K;
    const int c = k / kCondKh, kf = k - c * kCondKh;
    const float* src = a.weight + ((long long)c * a.h2 * kCondKh + kf) * kCondKw + kt;
    float* dst = W + kt * a.w_pad + k * h2p;
    for (int h = lane; h < h2p; h += 32) dst[h] = h < a.h2 ? __ldg(src + (long long)h * kCondKh * kCondKw) : 0.f;
  }
  if (tid < a.h2) {
    const float b = __ldg(a.bias + tid);
    rb[tid] = b < 0.f ? 0.f : b;
  }
  for (long long tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
  const long long rec = tile / tiles;
  const int j0 = (int)(tile - rec * tiles) * J;
  const long long rec_b = a.partner ? (long long)a.partner[rec] : rec;
  const float* sa = a.spec ? a.spec + rec * img : nullptr;
  const float* sb = a.spec ? a.spec + rec_b * img : nullptr;
  const float2* ka = a.knots ? a.knots + rec * img : nullptr;
  const float2* kb = a.knots ? a.knots + rec_b * img : nullptr;
  const bool mix = rec_b != rec;
  __syncthreads();                     // the staged weights; the previous tile's readers of slot / live / V
  if (tid == 0) n_live = 0;
  __syncthreads();
  // ---- positions of this CTA's samples (ATen's linear / align_corners=False arithmetic in fp32)
  for (int p = tid; p < 2 * J; p += kCondThreads) {
    const int jj = p & (J - 1), which = p / J;
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
    int s = -1;
    if (is_live) {
      s = atomicAdd(&n_live, 1);
      CondLive l;
      l.p_kt = p | (kt << 16);
      l.f = f;
      int col = tq;
      l.fr = 0.f;
      if (a.knots) {
        // image column t' sits at t' * (n_frames - 1) = i * D + rem between knots i and i + 1 (stft_image_kernel, phase 3)
        const unsigned num = (unsigned)tq * (unsigned)(a.n_frames - 1);
        int ki = (int)(num / (unsigned)a.D), rem = (int)(num - (unsigned)ki * (unsigned)a.D);
        if (ki >= a.n_frames - 1) { ki = a.n_frames - 2; rem = a.D; }
        col = ki;
        l.fr = (float)rem * a.inv_d;
      }
      l.base = f * (a.knots ? a.n_frames : a.n_cols) + col;
      live[s] = l;
    }
    slot[p] = s;
  }
  __syncthreads();
  const int nl = n_live;
  // ---- gather: A[u][c * 3 + kf] = in[c, f + 1 - kf, t'] (0 outside the image) for live position u.  A thread owns
  // ONE k (its channel / kernel-row offset is computed once) and walks the live positions four apart: the loads of the
  // unrolled iterations are independent and in flight together.
  {
    const int k = tid & 63, lane_p = tid >> 6;                    // 64 k slots (K <= 64 served per pass) x 4 position lanes
    const int row_len = a.knots ? a.n_frames : a.n_cols;
    for (int k0 = 0; k0 < K; k0 += 64) {
      const int kk = k0 + k;
      if (kk < K) {
        const int c = kk / kCondKh, kf = kk - c * kCondKh;
        const int df = kCondPadF - kf;
        const int ch_off = (c * a.n_freqs + df) * row_len;        // + CondLive::base = the element inside the record
        if (a.knots) {
          // the image sample is evaluated from its two knots exactly like stft_image_kernel's phase 3 writes it:
          // sat(value + delta * (rem / D)), NaN rows stay NaN
#pragma unroll 4
          for (int u = lane_p; u < nl; u += 4) {
            const CondLive l = live[u];
            const int f = l.f + df;
            float v = 0.f;
            if (f >= 0 && f < a.n_freqs) {
              const int off = ch_off + l.base;
              const float2 kn = __ldg(ka + off);
              v = fmaf(kn.y, l.fr, kn.x);
              v = v != v ? v : __saturatef(v);
              if (mix) {
                const float2 k2 = __ldg(kb + off);
                float v2 = fmaf(k2.y, l.fr, k2.x);
                v2 = v2 != v2 ? v2 : __saturatef(v2);
                v = a.alpha * v + (1.f - a.alpha) * v2;                         // diffEEG.py:143
              }
            }
            A[u * a.k_pad + kk] = v;
          }
        } else {
#pragma unroll 4
          for (int u = lane_p; u < nl; u += 4) {
            const CondLive l = live[u];
            const int f = l.f + df;
            float v = 0.f;
            if (f >= 0 && f < a.n_freqs) {
              const int off = ch_off + l.base;
              v = __ldg(sa + off);
              if (mix) v = a.alpha * v + (1.f - a.alpha) * __ldg(sb + off);     // diffEEG.py:143
            }
            A[u * a.k_pad + kk] = v;
          }
        }
      }
    }
  }
  __syncthreads();
  // ---- contraction + relu over the LIVE positions only: item = (live position u, four output channels).  Lanes of a
  // warp take consecutive u (rows of A an odd stride apart: no bank conflict) and read at most three weight rows
  // (one per kernel column, w_pad puts them in different banks): broadcasts.
  for (int it = tid; it < nl * HQ; it += kCondThreads) {
    const int hq = it / nl, u = it - hq * nl;
    const int kt = live[u].p_kt >> 16;
    const float* Ar = A + u * a.k_pad;
    const float4* w = reinterpret_cast<const float4*>(W + kt * a.w_pad) + hq;
    const int h0 = 4 * hq;
    float v0 = h0 < a.h2 ? __ldg(a.bias + h0) : 0.f, v1 = h0 + 1 < a.h2 ? __ldg(a.bias + h0 + 1) : 0.f;
    float v2 = h0 + 2 < a.h2 ? __ldg(a.bias + h0 + 2) : 0.f, v3 = h0 + 3 < a.h2 ? __ldg(a.bias + h0 + 3) : 0.f;
#pragma unroll 3
    for (int k = 0; k < K; ++k) {
      const float x = Ar[k];
      const float4 ww = w[k * HQ];
      v0 = fmaf(x, ww.x, v0);
      v1 = fmaf(x, ww.y, v1);
      v2 = fmaf(x, ww.z, v2);
      v3 = fmaf(x, ww.w, v3);
    }
    float* Vr = V + u * a.v_pad + h0;                                           // relu that keeps NaN, like torch's
    Vr[0] = v0 < 0.f ? 0.f : v0;
    if (h0 + 1 < a.h2) Vr[1] = v1 < 0.f ? 0.f : v1;
    if (h0 + 2 < a.h2) Vr[2] = v2 < 0.f ? 0.f : v2;
    if (h0 + 3 < a.h2) Vr[3] = v3 < 0.f ? 0.f : v3;
  }
  __syncthreads();
  // ---- interpolation between the two neighbours of every sample: thread = (sample jj, every (256 / J)-th output channel)
  const int jj = tid & (J - 1), hg = tid / J;
  const int j = j0 + jj;
  if (j >= a.out_len) continue;
  const int u0 = slot[jj], u1 = slot[J + jj];
  const float l1 = lam[jj], l0 = 1.f - l1;
  const float* V0 = V + (u0 < 0 ? 0 : u0) * a.v_pad;
  const float* V1 = V + (u1 < 0 ? 0 : u1) * a.v_pad;
  float* o = a.out + (rec * a.h2) * a.out_len + j;
  for (int h = hg; h < a.h2; h += kCondThreads / J) {
    const float x0 = u0 < 0 ? rb[h] : V0[h], x1 = u1 < 0 ? rb[h] : V1[h];
    o[(long long)h * a.out_len] = l0 * x0 + l1 * x1;
  }
  }
}

static int cond_launch(const float* spec, const float* knots, int n_frames, long long batch, const eegspec_cond_params* p,
                       const float* weight, const float* bias, const int* partner, float* out, cudaStream_t st);

int spec_condition_launch(const float* spec, long long batch, const eegspec_cond_params* p, const float* weight,
                          const float* bias, const int* partner, float* out, cudaStream_t st) {
  if (!spec) return EEGSPEC_EINVAL;
  return cond_launch(spec, nullptr, 0, batch, p, weight, bias, partner, out, st);
}

int spec_condition_knots_launch(const float* knots, int n_frames, long long batch, const eegspec_cond_params* p,
                                const float* weight, const float* bias, const int* partner, float* out, cudaStream_t st) {
  if (!knots || n_frames < 2) return EEGSPEC_EINVAL;
  if (reinterpret_cast<uintptr_t>(knots) & 7) return EEGSPEC_EALIGN;
  if (p && ((long long)p->n_cols * n_frames >= (1LL << 31) || p->n_cols < 2)) return EEGSPEC_EUNSUPPORTED;
  return cond_launch(nullptr, knots, n_frames, batch, p, weight, bias, partner, out, st);
}

static int cond_launch(const float* spec, const float* knots, int n_frames, long long batch, const eegspec_cond_params* p,
                       const float* weight, const float* bias, const int* partner, float* out, cudaStream_t st) {
  if (!weight || !bias || !out || !p || batch <= 0) return EEGSPEC_EINVAL;
  if (p->n_channels <= 0 || p->n_freqs <= 0 || p->n_cols <= 0 || p->out_channels <= 0 || p->out_length <= 0)
    return EEGSPEC_EINVAL;
  if (p->n_channels > 64 || p->out_channels > 256) return EEGSPEC_EUNSUPPORTED;
  CondArgs a;
  a.spec = spec;
  a.knots = reinterpret_cast<const float2*>(knots);
  a.n_frames = n_frames;
  a.D = p->n_cols - 1;
  a.inv_d = 1.0f / (float)(p->n_cols - 1);   // the same float stft_image_launch hands its kernel
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
  // the gather addresses one record's image (or knots) with 32-bit element offsets
  if ((long long)p->n_channels * p->n_freqs * (knots ? n_frames : p->n_cols) >= (1LL << 31)) return EEGSPEC_EUNSUPPORTED;
  a.t_out = (int)t_out;
  a.in_len = (int)in_len;
  a.k_pad = (p->n_channels * kCondKh) | 1;
  a.v_pad = p->out_channels | 1;
  const int h2p = (p->out_channels + 3) & ~3;
  a.w_pad = ((p->n_channels * kCondKh * h2p + 31) & ~31) + 8;
  a.scale = (float)a.in_len / (float)a.out_len;   // ATen: area_pixel_compute_scale<float>
  a.alpha = p->mix_alpha;
  constexpr int J = kCondJ;
  const size_t smem = sizeof(float) * ((size_t)2 * J * (a.k_pad + a.v_pad) + (size_t)kCondKw * a.w_pad);
  if (smem > 200 * 1024) return EEGSPEC_EUNSUPPORTED;
  a.n_tiles = batch * ((a.out_len + J - 1) / J);
  int dev = 0, n_sm = 0, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
  // set even below 48 KB: the static arrays count against the default limit
  cudaError_t e = cudaFuncSetAttribute(spec_condition_kernel<J>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, spec_condition_kernel<J>, kCondThreads, smem);
  if (e != cudaSuccess) return (int)e;
  if (per_sm < 1) return EEGSPEC_EUNSUPPORTED;
  // persistent CTAs: the weights are staged once per CTA, the tiles are walked round-robin
  const long long ctas = a.n_tiles < (long long)n_sm * per_sm ? a.n_tiles : (long long)n_sm * per_sm;
  spec_condition_kernel<J><<<(unsigned)ctas, kCondThreads, smem, st>>>(a);
  count_launch();
  return (int)cudaGetLastError();
}

}  // namespace eegspec
