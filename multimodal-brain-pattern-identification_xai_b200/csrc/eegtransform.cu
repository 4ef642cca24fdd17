// Row (f)1: the raw-EEG transformer in front of the spectrogram path.
//
// Replaces _EEGTransformer.transform (/root/reference/root/src/data/dataset.py:73-104): channel
// selection (:88-90) or the "magic 8" differences (:106-123), clip to +-1024 / NaN -> 0 / divide
// by 32 (:92-94), a 4th-order Butterworth low-pass applied as a CAUSAL IIR along time with
// SciPy's lfilter (:125-131, direct form II transposed, zero initial state) and decimation
// x[::downsample] (:101-102).
//
// The recurrence is sequential in time, so it is parallelised as a block scan over chunks of
// 16 * downsample samples.  The filter is linear in (state, input):
//   * every thread = (chunk, output channel) runs the recurrence over its chunk from a ZERO state
//     (fp32, samples from a shared-memory tile loaded with coalesced float4s) and keeps the 16
//     decimated outputs and the final state;
//   * the true state at the start of a chunk follows from s_{k+1} = M s_k + z_k (M = the 4 x 4
//     zero-input transition over one chunk, computed on the host in double precision); ten
//     chunks per block, the carry runs through the blocks of a record in double precision;
//   * the zero-input response of that state at the decimated positions (a 4 x 16 table G) is
//     added to the zero-state outputs.
// One CTA per record, 10 warps (chunks) x 32 lanes (channels).  Algorithmic bytes per record:
// 4 * L * n_ch in + 4 * ceil(L / ds) * n_out out (HBM bound; 0.95 MB for (10000, 20) -> (2000, 19)).
#include <math.h>
#include <string.h>

#include "common.cuh"
#include "internal.h"

namespace eegspec {

constexpr int kXfP = 16;        // decimated outputs per chunk
constexpr int kXfChunks = 10;   // chunks per block = warps per CTA
constexpr int kXfThreads = 32 * kXfChunks;

struct XfArgs {
  const float* x;
  float* out;
  long long rec_stride, length, l_out;
  int n_ch, n_out, ds, chunk, normalize, filter, out_channel_major;
  float clip_abs, scale;
  float b[EEGSPEC_XF_ORDER + 1], a[EEGSPEC_XF_ORDER + 1];
  float g[EEGSPEC_XF_ORDER][kXfP];                    // zero-input response of unit state i at sample m * ds
  double m[EEGSPEC_XF_ORDER][EEGSPEC_XF_ORDER];       // state after one chunk of zero input: s' = M s
  int sel_a[EEGSPEC_MAX_XF_CH], sel_b[EEGSPEC_MAX_XF_CH];
};

// DS > 0: compile-time decimation with normalize / filter / difference fixed by NORM / FILT / DIFF
// (straight-line recurrence: ~16 instructions per sample instead of ~42 with run-time switches,
// the kernel was issue bound at 65 % with them); DS == 0: everything read from the arguments.
template <int DS, bool NORM, bool FILT, bool DIFF>
__global__ void __launch_bounds__(kXfThreads) eeg_transform_kernel(const XfArgs a) {
  extern __shared__ __align__(16) float xs[];                       // [block samples][n_ch]
  __shared__ float zs[kXfChunks][EEGSPEC_XF_ORDER][32];             // zero-state final state of chunk k, channel c
  __shared__ double carry[EEGSPEC_XF_ORDER][32];                    // state at the start of the block

  const int tid = threadIdx.x, c = tid & 31, k = tid >> 5;
  const long long rec = blockIdx.x;
  const float* xr = a.x + rec * a.rec_stride;
  const int blk_samples = kXfChunks * a.chunk;
  const bool live = c < a.n_out;
  const int ca = live ? a.sel_a[c] : 0, cb = live ? a.sel_b[c] : -1;
  const int ds = DS ? DS : a.ds;
  const bool norm = DS ? NORM : (a.normalize != 0), filt = DS ? FILT : (a.filter != 0);
  if (tid < 32)
    for (int i = 0; i < EEGSPEC_XF_ORDER; ++i) carry[i][tid] = 0.0;

  const float b0 = a.b[0], b1 = a.b[1], b2 = a.b[2], b3 = a.b[3], b4 = a.b[4];
  const float a1 = a.a[1], a2 = a.a[2], a3 = a.a[3], a4 = a.a[4];

  for (long long t0 = 0; t0 < a.length; t0 += blk_samples) {
    __syncthreads();   // the previous block's tile and states have been consumed
    {   // tile load: blk_samples x n_ch floats, contiguous in the time-major record; zeros beyond the
        // end.  cp.async keeps all of a thread's 16-byte requests in flight at once (a plain
        // load -> store loop exposed one DRAM latency per request: 0.55 ms per 1 024 records).
      const long long n_valid = min((long long)blk_samples, a.length - t0) * a.n_ch;
      const float* src = xr + t0 * a.n_ch;
      const uint32_t dst = static_cast<uint32_t>(__cvta_generic_to_shared(xs));
      const int n4 = blk_samples * a.n_ch / 4;
      for (int i = tid; i < n4; i += kXfThreads) {
        const int bytes = (4LL * i < n_valid) ? 16 : 0;   // src-size 0: the 16 bytes are zero-filled
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst + 16u * i),
                     "l"(bytes ? src + 4LL * i : src), "r"(bytes)
                     : "memory");
      }
      asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();

    // ---- zero-state pass over this thread's chunk
    float yz[kXfP];
    float z0 = 0.f, z1 = 0.f, z2 = 0.f, z3 = 0.f;
    {
      const float* xp = xs + (long long)k * a.chunk * a.n_ch;
      const float* xq = xs + (long long)k * a.chunk * a.n_ch + (cb >= 0 ? cb : ca) - ca;   // subtrahend, same indexing
      const int n_ch = a.n_ch;
      const bool diff = DS ? DIFF : (cb >= 0);
      const float clip = a.clip_abs, scale = a.scale;
#pragma unroll
      for (int m = 0; m < kXfP; ++m) {
#pragma unroll
        for (int r = 0; r < (DS ? DS : 8); ++r) {
          if (!DS && r >= ds) break;
          float v = xp[ca];
          if (diff) v -= xq[ca];
          xp += n_ch;
          xq += n_ch;
          if (norm)   // np.clip propagates NaN, nan_to_num zeroes it (dataset.py:93-94)
            v = (v != v) ? 0.f : fminf(fmaxf(v, -clip), clip) * scale;
          float y = v;
          if (filt) {
            y = fmaf(b0, v, z0);
            z0 = fmaf(b1, v, fmaf(-a1, y, z1));
            z1 = fmaf(b2, v, fmaf(-a2, y, z2));
            z2 = fmaf(b3, v, fmaf(-a3, y, z3));
            z3 = fmaf(b4, v, -a4 * y);
          }
          if (r == 0) yz[m] = y;
        }
      }
    }
    if (filt) {
      zs[k][0][c] = z0; zs[k][1][c] = z1; zs[k][2][c] = z2; zs[k][3][c] = z3;
      __syncthreads();
      // ---- true state at the start of chunk k: s_{j+1} = M s_j + z_j from the block's carry
      double s[EEGSPEC_XF_ORDER];
#pragma unroll
      for (int i = 0; i < EEGSPEC_XF_ORDER; ++i) s[i] = carry[i][c];
      const int upto = (k == kXfChunks - 1) ? kXfChunks : k;   // the last warp also produces the next carry
      double s_k[EEGSPEC_XF_ORDER] = {s[0], s[1], s[2], s[3]};
      for (int j = 0; j < upto; ++j) {
        double nx[EEGSPEC_XF_ORDER];
#pragma unroll
        for (int i = 0; i < EEGSPEC_XF_ORDER; ++i)
          nx[i] = a.m[i][0] * s[0] + a.m[i][1] * s[1] + a.m[i][2] * s[2] + a.m[i][3] * s[3] + (double)zs[j][i][c];
#pragma unroll
        for (int i = 0; i < EEGSPEC_XF_ORDER; ++i) s[i] = nx[i];
        if (j == k - 1) {
#pragma unroll
          for (int i = 0; i < EEGSPEC_XF_ORDER; ++i) s_k[i] = s[i];
        }
      }
      __syncthreads();   // every warp has read the old carry
      if (k == kXfChunks - 1) {
#pragma unroll
        for (int i = 0; i < EEGSPEC_XF_ORDER; ++i) carry[i][c] = s[i];
      }
      const float f0 = (float)s_k[0], f1 = (float)s_k[1], f2 = (float)s_k[2], f3 = (float)s_k[3];
#pragma unroll
      for (int m = 0; m < kXfP; ++m)
        yz[m] += a.g[0][m] * f0 + a.g[1][m] * f1 + a.g[2][m] * f2 + a.g[3][m] * f3;
    }
    // ---- store the decimated outputs of this chunk
    if (live) {
      const long long o0 = t0 / ds + (long long)k * kXfP;   // t0 and the chunk are multiples of ds
      if (a.out_channel_major) {
        float* o = a.out + (rec * a.n_out + c) * a.l_out + o0;
#pragma unroll
        for (int m = 0; m < kXfP; ++m)
          if (o0 + m < a.l_out) o[m] = yz[m];
      } else {
        float* o = a.out + (rec * a.l_out + o0) * a.n_out + c;
#pragma unroll
        for (int m = 0; m < kXfP; ++m)
          if (o0 + m < a.l_out) o[(long long)m * a.n_out] = yz[m];
      }
    }
  }
}

// ------------------------------------------------------------------------------------- host
// Direct form II transposed, one step (double precision): the recurrence SciPy's lfilter runs.
static double df2t_step(const double* b, const double* a, double* z, double v) {
  const double y = b[0] * v + z[0];
  for (int i = 0; i < EEGSPEC_XF_ORDER - 1; ++i) z[i] = b[i + 1] * v - a[i + 1] * y + z[i + 1];
  z[EEGSPEC_XF_ORDER - 1] = b[EEGSPEC_XF_ORDER] * v - a[EEGSPEC_XF_ORDER] * y;
  return y;
}

void xform_params_default(eegspec_xform_params* p) {
  memset(p, 0, sizeof(*p));
  p->n_out = 19;
  for (int c = 0; c < EEGSPEC_MAX_XF_CH; ++c) { p->sel_a[c] = c < 19 ? c : 0; p->sel_b[c] = -1; }
  p->normalize = 1;
  p->clip_abs = 1024.f;
  p->scale = 1.f / 32.f;
  p->filter = 1;
  // scipy.signal.butter(4, 20 / 100, btype="low"): bilinear transform of the 4th-order Butterworth
  // prototype, evaluated here in double precision (two biquads multiplied out).
  {
    const double wc = tan(M_PI * 0.2 / 2.0);   // pre-warped cutoff, fs = 2
    double bb[5] = {1, 0, 0, 0, 0}, aa[5] = {1, 0, 0, 0, 0};
    int nb = 1;
    for (int q = 0; q < 2; ++q) {
      const double th = M_PI * (2.0 * q + 1.0) / 8.0;          // pole angles of the prototype
      const double k1 = 2.0 * sin(th);                          // s^2 + k1 s + 1
      const double w2 = wc * wc, d = 1.0 + k1 * wc + w2;
      const double sb[3] = {w2 / d, 2.0 * w2 / d, w2 / d};
      const double sa[3] = {1.0, 2.0 * (w2 - 1.0) / d, (1.0 - k1 * wc + w2) / d};
      double tb[5] = {0, 0, 0, 0, 0}, ta[5] = {0, 0, 0, 0, 0};
      for (int i = 0; i < nb; ++i)
        for (int j = 0; j < 3; ++j) { tb[i + j] += bb[i] * sb[j]; ta[i + j] += aa[i] * sa[j]; }
      memcpy(bb, tb, sizeof(bb));
      memcpy(aa, ta, sizeof(aa));
      nb += 2;
    }
    for (int i = 0; i < 5; ++i) { p->b[i] = bb[i]; p->a[i] = aa[i]; }
  }
  p->downsample = 5;
  p->out_channel_major = 0;
}

long long xform_out_length(long long length, const eegspec_xform_params* p) {
  const int ds = p->downsample < 1 ? 1 : p->downsample;
  return (length + ds - 1) / ds;
}

int eeg_transform_launch(const float* x, long long batch, long long length, int n_ch, long long rec_stride,
                         const eegspec_xform_params* p, float* out, cudaStream_t st) {
  if (!x || !out || !p || batch <= 0 || length <= 0 || n_ch <= 0) return EEGSPEC_EINVAL;
  if (p->n_out < 1 || p->n_out > EEGSPEC_MAX_XF_CH) return EEGSPEC_EINVAL;
  const int ds = p->downsample < 1 ? 1 : p->downsample;
  if (ds > 8) return EEGSPEC_EUNSUPPORTED;
  if ((n_ch & 3) || (rec_stride & 3) || (reinterpret_cast<uintptr_t>(x) & 15)) return EEGSPEC_EALIGN;   // 16-byte cp.async
  for (int c = 0; c < p->n_out; ++c)
    if (p->sel_a[c] < 0 || p->sel_a[c] >= n_ch || p->sel_b[c] >= n_ch) return EEGSPEC_EINVAL;
  if (p->filter && fabs(p->a[0] - 1.0) > 1e-12) return EEGSPEC_EINVAL;

  XfArgs a;
  a.x = x;
  a.out = out;
  a.rec_stride = rec_stride;
  a.length = length;
  a.l_out = xform_out_length(length, p);
  a.n_ch = n_ch;
  a.n_out = p->n_out;
  a.ds = ds;
  a.chunk = kXfP * ds;
  a.normalize = p->normalize;
  a.filter = p->filter;
  a.out_channel_major = p->out_channel_major;
  a.clip_abs = p->clip_abs;
  a.scale = p->scale;
  for (int i = 0; i <= EEGSPEC_XF_ORDER; ++i) { a.b[i] = (float)p->b[i]; a.a[i] = (float)p->a[i]; }
  for (int c = 0; c < EEGSPEC_MAX_XF_CH; ++c) { a.sel_a[c] = p->sel_a[c]; a.sel_b[c] = p->sel_b[c]; }
  for (int i = 0; i < EEGSPEC_XF_ORDER; ++i) {   // unit state i, zero input, one chunk
    double z[EEGSPEC_XF_ORDER] = {0, 0, 0, 0};
    z[i] = 1.0;
    for (int n = 0; n < a.chunk; ++n) {
      const double y = p->filter ? df2t_step(p->b, p->a, z, 0.0) : 0.0;
      if (n % ds == 0) a.g[i][n / ds] = (float)y;
    }
    for (int r = 0; r < EEGSPEC_XF_ORDER; ++r) a.m[r][i] = z[r];
  }
  const size_t smem = sizeof(float) * (size_t)kXfChunks * a.chunk * n_ch;
  if (smem > 200 * 1024) return EEGSPEC_EUNSUPPORTED;
  bool any_diff = false, all_diff = true;
  for (int c = 0; c < p->n_out; ++c) { any_diff |= p->sel_b[c] >= 0; all_diff &= p->sel_b[c] >= 0; }
  const bool fast = ds == 5 && p->normalize && p->filter && (any_diff == all_diff);
  void (*kernel)(const XfArgs) = !fast ? eeg_transform_kernel<0, false, false, false>
                                 : all_diff ? eeg_transform_kernel<5, true, true, true>
                                            : eeg_transform_kernel<5, true, true, false>;
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return EEGSPEC_EUNSUPPORTED;
  if (smem > 48 * 1024) {   // cheap and idempotent; three instantiations, so no per-device cache
    const cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  if (batch > 0x7fffffffLL) return EEGSPEC_EUNSUPPORTED;
  kernel<<<(unsigned)batch, kXfThreads, smem, st>>>(a);
  count_launch();
  return (int)cudaGetLastError();
}

}  // namespace eegspec
