// Mode R: fused STFT -> log1p|Z| -> time resample -> per-row min-max image kernel.
//
// Replaces /root/reference/root/src/utils/DiffEEG_utils.py:43-74 for all (record, channel)
// rows of a batch in one launch.  One CTA owns one row (or a column slice of one row when the
// batch is too small to fill 148 SMs):
//
//   phase 1  warp-per-frame radix-2 FFT.  A real N-point frame is packed as N/2 complex
//            points, NPL points per lane; butterflies whose partner lives in another lane use
//            warp shuffles, the rest stay in registers.  The Hann window (and SciPy's
//            1/sum(w) 'spectrum' scale) is folded into the load, zero boundary/padding is a
//            predicated load.  log1p|Z| lands transposed in shared memory as S[freq][frame].
//   phase 2  per frequency row: min/max over the *interpolated* columns (evaluated only at
//            the first/last column of every linear segment when up-sampling), then the knots
//            are normalised in place, so that
//   phase 3  the image write is a pure lerp of two shared-memory knots per pixel, emitted as
//            128-bit streaming stores, one contiguous 512 B line per warp.
//
// HBM traffic is the algorithmic minimum: the row is read once (frames overlap in L1/L2),
// the image is written once, nothing else touches DRAM.
#include "common.cuh"
#include "internal.h"

namespace eegspec {

struct StftArgs {
  const float* x;
  float* out;
  float* ws;
  long long rows, L, row_stride, n_frames, ws_stride, q_step;
  int hop, half, n_slots, n_pad, gather, split, logmag_only, use_ws;
  int out_cols, D, r_step, small_index;
  float inv_d, eps;
};

// Left knot and remainder of output column j:  j*(n-1) = i*D + rem, clamped to the last segment.
__device__ __forceinline__ void bracket_of(const StftArgs& a, long long j, long long& i, int& rem) {
  if (a.small_index) {
    unsigned num = (unsigned)j * (unsigned)(a.n_frames - 1);
    unsigned q = num / (unsigned)a.D;
    i = q;
    rem = (int)(num - q * (unsigned)a.D);
  } else {
    long long num = j * (a.n_frames - 1);
    i = num / a.D;
    rem = (int)(num - i * a.D);
  }
  if (i >= a.n_frames - 1) {
    i = a.n_frames - 2;
    rem = a.D;
  }
}

// Interpolated (un-normalised) value of column j on knot row Sf.
__device__ __forceinline__ float column_value(const StftArgs& a, const float* Sf, long long j) {
  long long i;
  int rem;
  bracket_of(a, j, i, rem);
  const long long lo = a.gather ? 2 * j : i;
  const float s0 = Sf[lo], s1 = Sf[lo + 1];
  return fmaf(s1 - s0, (float)rem * a.inv_d, s0);
}

template <int LOG2N>
__device__ __forceinline__ void fft_phase(const StftArgs& a, const float* __restrict__ xrow,
                                          float* S, const float* win, const float2* tw,
                                          const float2* utw, float2* zs_all) {
  constexpr int N = 1 << LOG2N, P = N / 2, LOG2P = LOG2N - 1;
  constexpr int LPF = P < 32 ? P : 32;       // lanes per frame
  constexpr int NPL = P < 32 ? 1 : P / 32;   // complex points per lane
  constexpr int FPW = 32 / LPF;              // frames per warp pass
  constexpr int NSH = LOG2P - (NPL > 1 ? (LOG2P - 5) : 0);  // shuffle stages
  constexpr bool HOIST = (NPL == 1);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int le = lane % LPF, grp = lane / LPF;
  float2* zs = zs_all + warp * (FPW * P) + grp * P;

  // Loop-invariant per-lane constants (register resident on the 1-point-per-lane path).
  float2 h_win = make_float2(0.f, 0.f), h_utw = make_float2(0.f, 0.f);
  float2 h_tw[HOIST ? NSH : 1];
  if constexpr (HOIST) {
    h_win = make_float2(win[2 * le], win[2 * le + 1]);
    h_utw = utw[le];
#pragma unroll
    for (int s = 0; s < NSH; ++s) {
      const int h = (P / 2) >> s;
      h_tw[s] = (le & h) ? tw[(le % h) * (P / (2 * h))] : make_float2(1.f, 0.f);
    }
  }

  for (int base = warp * FPW; base < a.n_slots; base += nwarps * FPW) {
    const int slot = base + grp;
    const bool live = slot < a.n_slots;
    long long frame = slot;
    if (a.gather) {
      int rem;
      bracket_of(a, slot >> 1, frame, rem);
      frame += (slot & 1);
    }
    const long long s0 = frame * a.hop - a.half;

    float2 z[NPL];
#pragma unroll
    for (int r = 0; r < NPL; ++r) {
      const int e = le * NPL + r;
      const long long i0 = s0 + 2 * e, i1 = i0 + 1;
      const float v0 = (live && i0 >= 0 && i0 < a.L) ? __ldg(xrow + i0) : 0.f;
      const float v1 = (live && i1 >= 0 && i1 < a.L) ? __ldg(xrow + i1) : 0.f;
      const float2 w = HOIST ? h_win : make_float2(win[2 * e], win[2 * e + 1]);
      z[r] = make_float2(v0 * w.x, v1 * w.y);
    }

    // Decimation-in-frequency radix-2; element e ends up holding Z[bitrev(e)].
#pragma unroll
    for (int st = 0; st < LOG2P; ++st) {
      const int h = (P / 2) >> st;
      if (h >= NPL) {
        const int lm = h / NPL;
        const bool lower = (le & lm) != 0;
#pragma unroll
        for (int r = 0; r < NPL; ++r) {
          float2 o;
          o.x = __shfl_xor_sync(0xffffffffu, z[r].x, lm);
          o.y = __shfl_xor_sync(0xffffffffu, z[r].y, lm);
          float2 t;
          t.x = lower ? o.x - z[r].x : z[r].x + o.x;
          t.y = lower ? o.y - z[r].y : z[r].y + o.y;
          float2 w;
          if constexpr (HOIST) {
            w = h_tw[st];
          } else {
            const int m = (le * NPL + r) % h;
            w = lower ? tw[m * (P / (2 * h))] : make_float2(1.f, 0.f);
          }
          z[r] = cmul(t, w);
        }
      } else {
#pragma unroll
        for (int r = 0; r < NPL; ++r) {
          if ((r & h) == 0) {
            const float2 p = z[r], q = z[r + h];
            z[r] = make_float2(p.x + q.x, p.y + q.y);
            const float2 d = make_float2(p.x - q.x, p.y - q.y);
            z[r + h] = cmul(d, tw[(r % h) * (P / (2 * h))]);
          }
        }
      }
    }

#pragma unroll
    for (int r = 0; r < NPL; ++r) {
      const unsigned e = le * NPL + r;
      zs[__brev(e) >> (32 - LOG2P)] = z[r];
    }
    __syncwarp();

    // Real-input untangle: X[k] = ((Z[k] + conj Z[P-k]) - i e^{-2 pi i k/N} (Z[k] - conj Z[P-k])) / 2
#pragma unroll
    for (int r = 0; r < NPL; ++r) {
      const int k = le + LPF * r;
      const float2 zk = zs[k], zp = zs[(P - k) & (P - 1)];
      const float2 w = HOIST ? h_utw : utw[k];
      const float ar = zk.x + zp.x, ai = zk.y - zp.y;
      const float br = zk.x - zp.x, bi = zk.y + zp.y;
      const float ur = fmaf(w.x, br, w.y * bi), ui = fmaf(w.x, bi, -w.y * br);
      const float xr = 0.5f * (ar + ui), xi = 0.5f * (ai - ur);
      const float mag = sqrtf(fmaf(xr, xr, xi * xi));
      if (live) {
        S[(long long)k * a.n_pad + slot] = log1pf(mag);
        // Nyquist bin X[P] = Re Z[0] - Im Z[0]; log1p is applied in a separate pass.
        if (k == 0) S[(long long)P * a.n_pad + slot] = fabsf(zk.x - zk.y);
      }
    }
    __syncwarp();
  }
}

constexpr int kStftThreads = 256;

// MODE 0: knots in shared memory (the normal case); 1: knots in a global workspace slice
// (very long rows); 2: log-magnitude only, knots written straight to the output.
template <int LOG2N, int MODE>
__global__ void __launch_bounds__(kStftThreads)
stft_image_kernel(const StftArgs a) {
  constexpr int N = 1 << LOG2N, P = N / 2, F = P + 1;
  constexpr int LPF = P < 32 ? P : 32;
  constexpr int FPW = 32 / LPF;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float2* tw = reinterpret_cast<float2*>(smem_raw);                 // P/2 (>=1)
  float2* utw = tw + (P / 2 > 0 ? P / 2 : 1);                       // P
  float2* zs = utw + P;                                             // warps * FPW * P
  float2* mm = zs + (kStftThreads / 32) * FPW * P;                  // F  (min, 1/(max-min+eps))
  float* win = reinterpret_cast<float*>(mm + F);                    // N
  float* S_sm = win + N;                                            // F * n_pad (when on chip)

  const long long row = blockIdx.x / a.split;
  const int part = blockIdx.x % a.split;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = kStftThreads / 32;
  const float* __restrict__ xrow = a.x + row * a.row_stride;

  float* S;
  if constexpr (MODE == 2) S = a.out + row * (long long)F * a.n_frames;
  else if constexpr (MODE == 1) S = a.ws + (long long)blockIdx.x * a.ws_stride;
  else S = S_sm;

  // Tables.  Periodic Hann scaled by 1/sum(w) = 2/N (scipy scaling='spectrum', mode 'stft').
  for (int j = tid; j < N; j += kStftThreads)
    win[j] = (0.5f - 0.5f * cospif(2.0f * j / N)) * (2.0f / N);
  for (int k = tid; k < P / 2; k += kStftThreads)
    tw[k] = make_float2(cospif(2.0f * k / P), -sinpif(2.0f * k / P));
  for (int k = tid; k < P; k += kStftThreads)
    utw[k] = make_float2(cospif(2.0f * k / N), sinpif(2.0f * k / N));
  __syncthreads();

  fft_phase<LOG2N>(a, xrow, S, win, tw, utw, zs);
  __syncthreads();

  for (int s = tid; s < a.n_slots; s += kStftThreads) {
    float* p = S + (long long)P * a.n_pad + s;
    *p = log1pf(*p);
  }
  if constexpr (MODE == 2) return;
  __syncthreads();

  // Phase 2: per-frequency min/max of the interpolated row, then normalise the knots.
  const bool ends_only = !a.gather && 2 * (a.n_frames - 1) < a.out_cols;
  const long long n_items = ends_only ? 2 * (a.n_frames - 1) : a.out_cols;
  const long long nm1 = a.n_frames - 1;
  for (int f = warp; f < F; f += nwarps) {
    float* Sf = S + (long long)f * a.n_pad;
    float lo = __int_as_float(0x7f800000), hi = __int_as_float(0xff800000);
    for (long long t = lane; t < n_items; t += 32) {
      long long j = t;
      if (ends_only) {
        const long long seg = t >> 1;
        if ((t & 1) == 0) j = (seg * a.D + nm1 - 1) / nm1;
        else j = (seg == nm1 - 1) ? a.D : ((seg + 1) * a.D + nm1 - 1) / nm1 - 1;
      }
      const float v = column_value(a, Sf, j);
      lo = min_nan(lo, v);
      hi = max_nan(hi, v);
    }
    lo = warp_min_nan(lo);
    hi = warp_max_nan(hi);
    const float inv = 1.0f / (hi - lo + a.eps);
    __syncwarp();
    for (int s = lane; s < a.n_slots; s += 32) Sf[s] = (Sf[s] - lo) * inv;
  }
  __syncthreads();

  // Phase 3: image write.  A thread keeps the four brackets of its column quad in registers
  // and walks the frequency rows; a warp emits one contiguous 512 B line per row.
  const int Q = a.out_cols >> 2;
  const int q0 = (int)((long long)part * Q / a.split), q1 = (int)((long long)(part + 1) * Q / a.split);
  float* __restrict__ orow = a.out + row * (long long)F * a.out_cols;
  for (int q = q0 + tid; q < q1; q += kStftThreads) {
    int lo_i[4];
    float fr[4];
    {
      long long i;
      int rem;
      bracket_of(a, 4LL * q, i, rem);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        lo_i[c] = a.gather ? 2 * (4 * q + c) : (int)i;
        fr[c] = (float)rem * a.inv_d;
        i += a.q_step;
        rem += a.r_step;
        if (rem >= a.D) { rem -= a.D; ++i; }
        if (i >= nm1) { i = nm1 - 1; rem = a.D; }
      }
    }
    float* o = orow + 4LL * q;
    const float* Sf = S;
#pragma unroll 4
    for (int f = 0; f < F; ++f, Sf += a.n_pad, o += a.out_cols) {
      float v[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float s0 = Sf[lo_i[c]], s1 = Sf[lo_i[c] + 1];
        v[c] = max_nan(min_nan(fmaf(s1 - s0, fr[c], s0), 1.0f), 0.0f);
      }
      st_stream_f4(o, make_float4(v[0], v[1], v[2], v[3]));
    }
  }
}

// ------------------------------------------------------------------------------------- host
static size_t stft_fixed_smem(int nperseg) {
  const int N = nperseg, P = N / 2, F = P + 1;
  const int LPF = P < 32 ? P : 32, FPW = 32 / LPF;
  size_t b = 0;
  b += sizeof(float2) * (P / 2 > 0 ? P / 2 : 1);
  b += sizeof(float2) * P;
  b += sizeof(float2) * (kStftThreads / 32) * FPW * P;
  b += sizeof(float2) * F;
  b += sizeof(float) * N;
  return b;
}

static constexpr size_t kSmemBudget = 200 * 1024;

int stft_plan_make(long long length, int nperseg, int noverlap, long long out_cols,
                   eegspec_stft_plan* plan) {
  if (!plan || length <= 0 || nperseg <= 0) return EEGSPEC_EINVAL;
  if (noverlap < 0 || noverlap >= nperseg) return EEGSPEC_ENOVERLAP;
  if (length < nperseg) return EEGSPEC_ESHORT;
  if (nperseg & 1) return EEGSPEC_EUNSUPPORTED;
  const int hop = nperseg - noverlap, half = nperseg / 2;
  const long long ext = length + 2LL * half;
  // SciPy: nadd = (-(len - nperseg) % nstep) % nperseg  with Python's non-negative modulo.
  long long m = (-(ext - nperseg)) % hop;
  if (m < 0) m += hop;
  const long long nadd = m % nperseg;
  plan->hop = hop;
  plan->half = half;
  plan->nadd = (int32_t)nadd;
  plan->n_freqs = nperseg / 2 + 1;
  plan->n_frames = (ext + nadd - nperseg) / hop + 1;
  plan->gather = (out_cols > 0 && plan->n_frames > 2 * out_cols + 2) ? 1 : 0;
  plan->n_slots = plan->gather ? 2 * out_cols : plan->n_frames;
  plan->reserved = 0;
  return EEGSPEC_OK;
}

long long interp_bracket(long long j, long long n_frames, long long out_cols) {
  if (n_frames < 2 || out_cols < 2) return 0;
  long long i = (j * (n_frames - 1)) / (out_cols - 1);
  return i > n_frames - 2 ? n_frames - 2 : i;
}

static int check_pow2(int nperseg) {
  if (nperseg < 16 || nperseg > 1024 || (nperseg & (nperseg - 1))) return EEGSPEC_EUNSUPPORTED;
  return EEGSPEC_OK;
}

struct StftLaunchShape {
  int split, use_ws, n_pad;
  size_t smem, ws_per_cta;
  long long grid;
};

static StftLaunchShape stft_shape(long long rows, int nperseg, const eegspec_stft_plan& pl) {
  StftLaunchShape s;
  const int F = pl.n_freqs;
  s.n_pad = (int)(pl.n_slots | 1);
  const size_t s_bytes = sizeof(float) * (size_t)F * s.n_pad;
  const size_t fixed = stft_fixed_smem(nperseg);
  s.use_ws = (fixed + s_bytes > kSmemBudget) ? 1 : 0;
  s.smem = fixed + (s.use_ws ? 0 : s_bytes);
  s.ws_per_cta = s.use_ws ? s_bytes : 0;
  // Small batches: slice the columns of a row over several CTAs so that all 148 SMs write.
  int resident = (int)((220 * 1024) / (s.smem + 1024));
  if (resident > 2048 / kStftThreads) resident = 2048 / kStftThreads;
  if (resident < 1) resident = 1;
  const long long want = 148LL * resident;
  long long split = rows >= want ? 1 : (want + rows - 1) / rows;
  if (split > 10) split = 10;
  s.split = (int)split;
  s.grid = rows * split;
  return s;
}

size_t stft_image_workspace(long long rows, long long length, int nperseg, int noverlap,
                            long long out_cols) {
  eegspec_stft_plan pl;
  if (stft_plan_make(length, nperseg, noverlap, out_cols, &pl) != EEGSPEC_OK) return 0;
  if (check_pow2(nperseg) != EEGSPEC_OK || rows <= 0) return 0;
  const StftLaunchShape s = stft_shape(rows, nperseg, pl);
  return s.ws_per_cta * (size_t)s.grid;
}

template <int LOG2N, int MODE>
static int launch_mode(const StftArgs& a, long long grid, size_t smem, cudaStream_t st) {
  static bool configured[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute(stft_image_kernel<LOG2N, MODE>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)(kSmemBudget + 8 * 1024));
    if (e != cudaSuccess) return (int)e;
    configured[dev] = true;
  }
  stft_image_kernel<LOG2N, MODE><<<(unsigned)grid, kStftThreads, smem, st>>>(a);
  count_launch();
  return (int)cudaGetLastError();
}

template <int LOG2N>
static int launch_one(const StftArgs& a, long long grid, size_t smem, cudaStream_t st) {
  if (a.logmag_only) return launch_mode<LOG2N, 2>(a, grid, smem, st);
  if (a.use_ws) return launch_mode<LOG2N, 1>(a, grid, smem, st);
  return launch_mode<LOG2N, 0>(a, grid, smem, st);
}

static int dispatch(int nperseg, const StftArgs& a, long long grid, size_t smem, cudaStream_t st) {
  switch (nperseg) {
    case 16: return launch_one<4>(a, grid, smem, st);
    case 32: return launch_one<5>(a, grid, smem, st);
    case 64: return launch_one<6>(a, grid, smem, st);
    case 128: return launch_one<7>(a, grid, smem, st);
    case 256: return launch_one<8>(a, grid, smem, st);
    case 512: return launch_one<9>(a, grid, smem, st);
    case 1024: return launch_one<10>(a, grid, smem, st);
    default: return EEGSPEC_EUNSUPPORTED;
  }
}

int stft_image_launch(const float* x, long long rows, long long length, long long row_stride,
                      int nperseg, int noverlap, long long out_cols, float eps, float* out,
                      void* workspace, size_t workspace_bytes, cudaStream_t st) {
  if (!x || !out || rows <= 0 || row_stride < length) return EEGSPEC_EINVAL;
  if (out_cols < 4 || (out_cols & 3) || out_cols > (1 << 24)) return EEGSPEC_EUNSUPPORTED;
  if (reinterpret_cast<uintptr_t>(out) & 15) return EEGSPEC_EALIGN;
  eegspec_stft_plan pl;
  int rc = stft_plan_make(length, nperseg, noverlap, out_cols, &pl);
  if (rc != EEGSPEC_OK) return rc;
  if ((rc = check_pow2(nperseg)) != EEGSPEC_OK) return rc;
  if (rows > (1LL << 27)) return EEGSPEC_EUNSUPPORTED;
  const StftLaunchShape s = stft_shape(rows, nperseg, pl);
  if (s.use_ws) {
    if (!workspace || workspace_bytes < s.ws_per_cta * (size_t)s.grid) return EEGSPEC_EWORKSPACE;
    if (reinterpret_cast<uintptr_t>(workspace) & 15) return EEGSPEC_EALIGN;
  }
  StftArgs a;
  a.x = x;
  a.out = out;
  a.ws = static_cast<float*>(workspace);
  a.rows = rows;
  a.L = length;
  a.row_stride = row_stride;
  a.n_frames = pl.n_frames;
  a.ws_stride = (long long)(s.ws_per_cta / sizeof(float));
  a.hop = pl.hop;
  a.half = pl.half;
  a.n_slots = (int)pl.n_slots;
  a.n_pad = s.n_pad;
  a.gather = pl.gather;
  a.split = s.split;
  a.logmag_only = 0;
  a.use_ws = s.use_ws;
  a.out_cols = (int)out_cols;
  a.D = (int)out_cols - 1;
  a.q_step = (pl.n_frames - 1) / a.D;
  a.r_step = (int)((pl.n_frames - 1) % a.D);
  a.small_index = ((double)out_cols * (double)pl.n_frames < 4.0e9) ? 1 : 0;
  a.inv_d = 1.0f / (float)a.D;
  a.eps = eps;
  return dispatch(nperseg, a, s.grid, s.smem, st);
}

int stft_logmag_launch(const float* x, long long rows, long long length, long long row_stride,
                       int nperseg, int noverlap, float* out, cudaStream_t st) {
  if (!x || !out || rows <= 0 || row_stride < length) return EEGSPEC_EINVAL;
  eegspec_stft_plan pl;
  int rc = stft_plan_make(length, nperseg, noverlap, 0, &pl);
  if (rc != EEGSPEC_OK) return rc;
  if ((rc = check_pow2(nperseg)) != EEGSPEC_OK) return rc;
  if (pl.n_frames > (1LL << 30) || rows > (1LL << 27)) return EEGSPEC_EUNSUPPORTED;
  StftArgs a = {};
  a.x = x;
  a.out = out;
  a.rows = rows;
  a.L = length;
  a.row_stride = row_stride;
  a.n_frames = pl.n_frames;
  a.hop = pl.hop;
  a.half = pl.half;
  a.n_slots = (int)pl.n_frames;
  a.n_pad = (int)pl.n_frames;
  a.split = 1;
  a.logmag_only = 1;
  a.out_cols = 4;
  a.D = 3;
  a.inv_d = 1.f / 3.f;
  return dispatch(nperseg, a, rows, stft_fixed_smem(nperseg), st);
}

}  // namespace eegspec
