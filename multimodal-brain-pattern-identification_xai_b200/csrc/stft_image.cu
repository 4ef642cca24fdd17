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
#include <atomic>

#include "common.cuh"
#include "internal.h"
#include "stft_thread_fft.cuh"

namespace eegspec {

struct StftArgs {
  const float* x;
  float* out;
  float* ws;
  long long rows, L, row_stride, n_frames, ws_stride, q_step, q32_step;
  int hop, half, n_slots, n_pad, gather, split, logmag_only, use_ws;
  int out_cols, D, r_step, r32_step, small_index, ends_only, n_items;
  float inv_d, eps;
  int stage;   // thread-per-frame path: row staged in shared memory (aliases the knot array)
  int knots_only;   // stop after phase 2 and write the normalised knots {value, delta} instead of the image
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

// ceil(a / b) for the first / last column of a linear segment
__device__ __forceinline__ long long ceil_div(const StftArgs& a, long long num, long long den) {
  if (a.small_index) return ((unsigned)num + (unsigned)den - 1u) / (unsigned)den;
  return (num + den - 1) / den;
}

// STRIDE = floats between consecutive knots of one frequency row (2: {value, delta} pairs).
template <int LOG2N, int STRIDE>
__device__ __forceinline__ void fft_phase(const StftArgs& a, const float* __restrict__ xrow,
                                          float* S, const float* win, const float2* tw,
                                          const float2* utw, float2* zs_all) {
  constexpr int N = 1 << LOG2N, P = N / 2, LOG2P = LOG2N - 1;
  constexpr int LPF = P < 32 ? P : 32;       // lanes per frame
  constexpr int NPL = P < 32 ? 1 : P / 32;   // complex points per lane
  constexpr int FPW = 32 / LPF;              // frames per warp pass
  constexpr int NSH = NPL > 1 ? 5 : LOG2P;   // shuffle stages
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
      // opaque to the optimiser: keeps the twiddles in registers instead of re-deriving them per frame
      asm volatile("" : "+f"(h_tw[s].x), "+f"(h_tw[s].y));
    }
    asm volatile("" : "+f"(h_win.x), "+f"(h_win.y), "+f"(h_utw.x), "+f"(h_utw.y));
  }

  // Per-lane output rows (k = le + LPF*r) are loop invariant: hoist their base pointers.
  float* Sk[NPL];
#pragma unroll
  for (int r = 0; r < NPL; ++r) Sk[r] = S + (long long)(le + LPF * r) * a.n_pad * STRIDE;
  float* Snyq = S + (long long)P * a.n_pad * STRIDE;
  const bool small = a.L < (1LL << 30) && !a.gather;   // 32-bit sample indices

  // Sample fetch for one slot: two values per complex point, zero outside [0, L) (SciPy's zero
  // boundary extension and end padding).  Unsigned compares fold the two range checks into one.
  const int lane_off = 2 * le * NPL - a.half;
  auto fetch = [&](int slot, float (&v0)[NPL], float (&v1)[NPL]) {
    if (small) {
      const int s0 = slot * a.hop + lane_off;
      const unsigned len = slot < a.n_slots ? (unsigned)a.L : 0u;
#pragma unroll
      for (int r = 0; r < NPL; ++r) {
        const int i0 = s0 + 2 * r;
        v0[r] = (unsigned)i0 < len ? __ldg(xrow + i0) : 0.f;
        v1[r] = (unsigned)(i0 + 1) < len ? __ldg(xrow + i0 + 1) : 0.f;
      }
    } else {
      const bool live = slot < a.n_slots;
      long long frame = slot;
      if (a.gather) {
        int rem;
        bracket_of(a, slot >> 1, frame, rem);
        frame += (slot & 1);
      }
      const long long s0 = frame * a.hop - a.half;
#pragma unroll
      for (int r = 0; r < NPL; ++r) {
        const long long i0 = s0 + 2 * (le * NPL + r), i1 = i0 + 1;
        v0[r] = (live && i0 >= 0 && i0 < a.L) ? __ldg(xrow + i0) : 0.f;
        v1[r] = (live && i1 >= 0 && i1 < a.L) ? __ldg(xrow + i1) : 0.f;
      }
    }
  };

  const int stride_slots = nwarps * FPW;
  float nv0[NPL], nv1[NPL];
  fetch(warp * FPW + grp, nv0, nv1);
  for (int base = warp * FPW; base < a.n_slots; base += stride_slots) {
    const int slot = base + grp;
    const bool live = slot < a.n_slots;

    float2 z[NPL];
#pragma unroll
    for (int r = 0; r < NPL; ++r) {
      const int e = le * NPL + r;
      const float2 w = HOIST ? h_win : make_float2(win[2 * e], win[2 * e + 1]);
      z[r] = make_float2(nv0[r] * w.x, nv1[r] * w.y);
    }
    // Software pipeline: the next frame's samples are in flight while this one is transformed.
    fetch(slot + stride_slots, nv0, nv1);   // out-of-range slots read as zeros

    // Decimation-in-frequency radix-2; element e ends up holding Z[bitrev(e)].
#pragma unroll
    for (int st = 0; st < LOG2P; ++st) {
      const int h = (P / 2) >> st;
      if (h >= NPL) {
        const int lm = h / NPL;
        const bool lower = (le & lm) != 0;
        const float sgn = lower ? -1.f : 1.f;
#pragma unroll
        for (int r = 0; r < NPL; ++r) {
          float2 o;
          o.x = __shfl_xor_sync(0xffffffffu, z[r].x, lm);
          o.y = __shfl_xor_sync(0xffffffffu, z[r].y, lm);
          // upper lane: own + partner; lower lane: partner - own
          const float2 t = make_float2(fmaf(z[r].x, sgn, o.x), fmaf(z[r].y, sgn, o.y));
          if (h == 1 && NPL == 1) {   // last stage of the 1-point path: twiddle is 1
            z[r] = t;
          } else {
            float2 w;
            if constexpr (HOIST) {
              w = h_tw[st];
            } else {
              const int m = (le * NPL + r) % h;
              w = lower ? tw[m * (P / (2 * h))] : make_float2(1.f, 0.f);
            }
            z[r] = cmul(t, w);
          }
        }
      } else {
#pragma unroll
        for (int r = 0; r < NPL; ++r) {
          if ((r & h) == 0) {
            const float2 p = z[r], q = z[r + h];
            z[r] = make_float2(p.x + q.x, p.y + q.y);
            const float2 d = make_float2(p.x - q.x, p.y - q.y);
            z[r + h] = (r % h == 0) ? d : cmul(d, tw[(r % h) * (P / (2 * h))]);
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
      const float mag = sqrt_pos(fmaf(xr, xr, xi * xi));
      if (live) {
        Sk[r][slot * STRIDE] = log1p_pos(mag);
        // Nyquist bin X[P] = Re Z[0] - Im Z[0]; log1p is applied in a separate pass.
        if (k == 0) Snyq[slot * STRIDE] = fabsf(zk.x - zk.y);
      }
    }
    __syncwarp();
  }
}

// Phase 1 for nperseg <= 64: one thread per frame, everything in registers
// (stft_thread_fft.cuh).  Interior frames of 16-byte aligned rows are fetched as float4.
template <int LOG2N>
__device__ __forceinline__ void frame_fetch(const StftArgs& a, const float* __restrict__ xrow, int slot,
                                            float (&v)[1 << LOG2N]) {
  constexpr int N = 1 << LOG2N;
  const bool vec_ok = !a.gather && ((a.hop | a.half) & 3) == 0 &&
                      (reinterpret_cast<uintptr_t>(xrow) & 15) == 0;
  long long frame = slot;
  if (a.gather) {
    int rem;
    bracket_of(a, slot >> 1, frame, rem);
    frame += (slot & 1);
  }
  const long long s0 = frame * a.hop - a.half;
  if (vec_ok && s0 >= 0 && s0 + N <= a.L) {
    const float4* src = reinterpret_cast<const float4*>(xrow + s0);
#pragma unroll
    for (int i = 0; i < N / 4; ++i) {
      const float4 t = __ldg(src + i);
      v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
    }
  } else {   // SciPy's zero boundary extension / end padding
#pragma unroll
    for (int i = 0; i < N; ++i) {
      const long long n = s0 + i;
      v[i] = (n >= 0 && n < a.L) ? __ldg(xrow + n) : 0.f;
    }
  }
}

// Row staging for the thread-per-frame path.  The zero-extended row (SciPy boundary + padding)
// is copied into shared memory with coalesced 16-byte cp.async; every 32-sample block is padded
// to 36 floats so that the per-thread frame reads (lane pitch = hop samples) are conflict-free
// 128-bit loads when hop = 32.  e = sample index in the extended row (n = e - half).
__device__ __forceinline__ int stage_index(int e) { return e + ((e >> 5) << 2); }

__device__ __forceinline__ void stage_row(const StftArgs& a, const float* __restrict__ xrow, float* stage,
                                          int n_ext) {
  const int L = (int)a.L;
  for (int c = threadIdx.x; 4 * c < n_ext; c += blockDim.x) {
    const int e0 = 4 * c, n0 = e0 - a.half;
    float* dst = stage + stage_index(e0);
    if (n0 >= 0 && n0 + 4 <= L) {
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)),
                   "l"(xrow + n0)
                   : "memory");
    } else {
      float4 z;
      z.x = (n0 >= 0 && n0 < L) ? __ldg(xrow + n0) : 0.f;
      z.y = (n0 + 1 >= 0 && n0 + 1 < L) ? __ldg(xrow + n0 + 1) : 0.f;
      z.z = (n0 + 2 >= 0 && n0 + 2 < L) ? __ldg(xrow + n0 + 2) : 0.f;
      z.w = (n0 + 3 >= 0 && n0 + 3 < L) ? __ldg(xrow + n0 + 3) : 0.f;
      *reinterpret_cast<float4*>(dst) = z;
    }
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

template <int LOG2N>
__device__ __forceinline__ void frame_from_stage(const StftArgs& a, const float* stage, int slot,
                                                 float (&v)[1 << LOG2N]) {
  constexpr int N = 1 << LOG2N;
  const int e0 = slot * a.hop;
#pragma unroll
  for (int i = 0; i < N / 4; ++i) {
    const float4 t = *reinterpret_cast<const float4*>(stage + stage_index(e0 + 4 * i));
    v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
  }
}

// v holds the samples of slot threadIdx.x (fetched by the caller ahead of the table set-up).
template <int LOG2N, int STRIDE>
__device__ __forceinline__ void fft_thread_phase(const StftArgs& a, const float* __restrict__ xrow,
                                                 float* S, float (&v)[1 << LOG2N]) {
  const long long kstride = (long long)a.n_pad * STRIDE;
  if ((int)threadIdx.x < a.n_slots) {
    float* Sc = S + (long long)threadIdx.x * STRIDE;
    frame_fft_mag<LOG2N>(v, [&](int k, float mag) { Sc[k * kstride] = log1p_pos(mag); });
  }
#pragma unroll 1
  for (int slot = threadIdx.x + blockDim.x; slot < a.n_slots; slot += blockDim.x) {   // long rows only
    float w[1 << LOG2N];
    frame_fetch<LOG2N>(a, xrow, slot, w);
    float* Sc = S + (long long)slot * STRIDE;
    frame_fft_mag<LOG2N>(w, [&](int k, float mag) { Sc[k * kstride] = log1p_pos(mag); });
  }
}

constexpr int kStftMaxThreads = 512;
constexpr int kStftThreadFftMax = 320;   // thread-per-frame path: 2 CTAs of 320 threads x <= 102 registers

// MODE 0: knots in shared memory (the normal case); 1: knots in a global workspace slice
// (very long rows); 2: log-magnitude only, knots written straight to the output.
template <int LOG2N, int MODE>
__global__ void __launch_bounds__((LOG2N <= 6 ? kStftThreadFftMax : kStftMaxThreads), (LOG2N <= 7 ? 2 : 1))
stft_image_kernel(const StftArgs a) {
  constexpr int N = 1 << LOG2N, P = N / 2, F = P + 1;
  constexpr int LPF = P < 32 ? P : 32;
  constexpr int FPW = 32 / LPF;
  constexpr int STRIDE = MODE == 2 ? 1 : 2;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int nthreads = blockDim.x, nwarps = nthreads >> 5;
  constexpr bool kThreadFft = LOG2N <= 6;   // tables live in the instruction stream
  float2* tw = reinterpret_cast<float2*>(smem_raw);                 // P/2 (>=1)
  float2* utw = tw + (kThreadFft ? 0 : (P / 2 > 0 ? P / 2 : 1));    // P
  float2* zs = utw + (kThreadFft ? 0 : P);                          // warps * FPW * P
  float2* ends = zs + (kThreadFft ? 0 : nwarps * FPW * P);          // one per segment when ends_only
  float* win = reinterpret_cast<float*>(ends + (a.ends_only ? a.n_items : 0));  // N
  int* rowflag = reinterpret_cast<int*>(win + (kThreadFft ? 0 : N));  // F + 1 (NaN rows, any)
  float* part_lo = reinterpret_cast<float*>(rowflag + F + 1 + ((F + 1) & 1));   // nthreads: partial row minima
  float* part_hi = part_lo + nthreads;                              // nthreads: partial row maxima
  float* S_sm = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(part_hi + nthreads) + 15) & ~uintptr_t(15));  // F*n_pad*STRIDE, 16 B aligned

  const long long row = blockIdx.x / a.split;
  const int part = blockIdx.x % a.split;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float* __restrict__ xrow = a.x + row * a.row_stride;

  float* S;
  if constexpr (MODE == 2) S = a.out + row * (long long)F * a.n_frames;
  else if constexpr (MODE == 1) S = a.ws + (long long)blockIdx.x * a.ws_stride;
  else S = S_sm;

  // Thread-per-frame path: the frame's samples are requested first so that their DRAM latency
  // hides behind the table set-up; the row a later CTA will own is pulled into L2 meanwhile.
  float v_first[kThreadFft ? N : 1];
  if constexpr (kThreadFft) {
    if (MODE == 0 && a.stage) {
      stage_row(a, xrow, S_sm, (a.n_slots - 1) * a.hop + N);
      __syncthreads();
      if (tid < a.n_slots) frame_from_stage<LOG2N>(a, S_sm, tid, v_first);
      __syncthreads();   // the knots are about to overwrite the staged samples
    } else if (tid < a.n_slots) frame_fetch<LOG2N>(a, xrow, tid, v_first);
    const long long ahead = row + (long long)(2 * 148) / a.split;
    if (part == 0 && ahead < a.rows && (long long)tid * 32 < a.L)
      asm volatile("prefetch.global.L2 [%0];" ::"l"(a.x + ahead * a.row_stride + (long long)tid * 32));
  }

  auto setup_tables = [&]() {
  // Tables.  Periodic Hann scaled by 1/sum(w) = 2/N (scipy scaling='spectrum', mode 'stft').
    if constexpr (!kThreadFft) {
      for (int j = tid; j < N; j += nthreads)
        win[j] = (0.5f - 0.5f * cospif(2.0f * j / N)) * (2.0f / N);
      for (int k = tid; k < P / 2; k += nthreads)
        tw[k] = make_float2(cospif(2.0f * k / P), -sinpif(2.0f * k / P));
      for (int k = tid; k < P; k += nthreads)
        utw[k] = make_float2(cospif(2.0f * k / N), sinpif(2.0f * k / N));
    }
    if constexpr (MODE != 2) {
      for (int f = tid; f <= F; f += nthreads) rowflag[f] = 0;
      // First and last output column of every linear segment, as (knot slot, fraction): the
      // extrema of a piecewise-linear row over the sampled columns can only sit there.
      if (a.ends_only) {
        // ends_only implies n_frames - 1 < out_cols / 2, so everything fits 32-bit integers.
        // Segment seg covers columns [ceil(seg*D/nm1), ceil((seg+1)*D/nm1) - 1]; each thread walks a
        // contiguous run of segments, carrying (seg*D) div/mod nm1 forward without dividing again.
        const int nm1 = (int)a.n_frames - 1, D = a.D;
        const int per = (nm1 + nthreads - 1) / nthreads;
        const int seg0 = tid * per, seg1 = min(nm1, seg0 + per);
        if (seg0 < seg1) {
          const int dq = (int)((unsigned)D / (unsigned)nm1), dr = D - dq * nm1;
          int q = (int)((unsigned)(seg0 * D) / (unsigned)nm1), r = seg0 * D - q * nm1;   // seg0 * D < 2^31
          int jf = q + (r ? 1 : 0);                       // first column of seg0
          for (int seg = seg0; seg < seg1; ++seg) {
            q += dq; r += dr;
            if (r >= nm1) { r -= nm1; ++q; }
            const int jn = q + (r ? 1 : 0);               // first column of seg + 1
            const int jl = (seg == nm1 - 1) ? D : jn - 1; // last column of seg
            // column j sits at fraction (j*nm1 - seg*D)/D inside segment seg
            ends[seg] = make_float2((float)(jf * nm1 - seg * D) * a.inv_d, (float)(jl * nm1 - seg * D) * a.inv_d);
            jf = jn;
          }
        }
      }
    }
  };
  if constexpr (!kThreadFft) setup_tables();
  if constexpr (!kThreadFft) __syncthreads();   // the thread path reads no shared table in phase 1

  const long long nm1 = a.n_frames - 1;
  if constexpr (kThreadFft) {
    fft_thread_phase<LOG2N, STRIDE>(a, xrow, S, v_first);
    setup_tables();   // segment ends: first needed in phase 2, kept out of the FFT's register budget
    if constexpr (MODE == 2) return;
  } else {
    fft_phase<LOG2N, STRIDE>(a, xrow, S, win, tw, utw, zs);
    __syncthreads();
    for (int s = tid; s < a.n_slots; s += nthreads) {
      float* p = S + ((long long)P * a.n_pad + s) * STRIDE;
      *p = log1p_pos(*p);
    }
    if constexpr (MODE == 2) return;
  }
  __syncthreads();

  // Phase 2: per-frequency min/max of the interpolated row; then every knot becomes the pair
  // {normalised value, normalised delta to the next knot} so that phase 3 is one FFMA per pixel.
  if (a.ends_only && nthreads >= F) {
    // Up-sampling (the usual case).  Thread = (frequency row f, run of consecutive segments):
    // running min / max in two registers, lanes of a warp on consecutive rows (conflict-free:
    // the knot row pitch is an odd number of float2).
    float2* S2 = reinterpret_cast<float2*>(S);
    const int nseg = a.n_items, groups = nthreads / F;
    const int chunk = (nseg + groups - 1) / groups;
    const int f = tid % F, c = tid / F;
    const int seg0 = min(nseg, c * chunk), seg1 = c < groups ? min(nseg, seg0 + chunk) : seg0;
    float2* K = S2 + (long long)f * a.n_pad;
    float lo = __int_as_float(0x7f800000), hi = __int_as_float(0xff800000);
    if (seg0 < seg1) {
      float s0 = K[seg0].x;
#pragma unroll 4
      for (int seg = seg0; seg < seg1; ++seg) {
        const float2 e = ends[seg];                 // fractions of the first / last column
        const float s1 = K[seg + 1].x, d = s1 - s0;
        const float va = fmaf(d, e.x, s0), vb = fmaf(d, e.y, s0);
        lo = min_nan(lo, min_nan(va, vb));
        hi = max_nan(hi, max_nan(va, vb));
        K[seg].y = d;
        s0 = s1;
      }
      if (seg1 == nseg) K[nseg].y = 0.f;
    }
    part_lo[tid] = lo;
    part_hi[tid] = hi;
    __syncthreads();
    if (seg0 < seg1) {
      for (int g = 0; g < groups; ++g) {
        lo = min_nan(lo, part_lo[g * F + f]);
        hi = max_nan(hi, part_hi[g * F + f]);
      }
      const float inv = 1.0f / (hi - lo + a.eps);
      if (c == 0 && !(inv == inv)) { rowflag[f] = 1; rowflag[F] = 1; }
      const int last = seg1 == nseg ? nseg + 1 : seg1;
#pragma unroll 4
      for (int seg = seg0; seg < last; ++seg) {
        const float2 k = K[seg];
        K[seg] = make_float2((k.x - lo) * inv, k.y * inv);
      }
    }
  } else
  for (int f = warp; f < F; f += nwarps) {
    float* Sf = S + (long long)f * a.n_pad * STRIDE;
    float lo = __int_as_float(0x7f800000), hi = __int_as_float(0xff800000);
    if (a.ends_only) {
      for (int seg = lane; seg < a.n_items; seg += 32) {   // n_items = segments here
        const float2 e = ends[seg];                          // fractions of first / last column
        const float s0 = Sf[seg * STRIDE], d = Sf[(seg + 1) * STRIDE] - s0;
        const float va = fmaf(d, e.x, s0), vb = fmaf(d, e.y, s0);
        lo = min_nan(lo, min_nan(va, vb));
        hi = max_nan(hi, max_nan(va, vb));
      }
    } else {
      long long i;
      int rem;
      bracket_of(a, lane, i, rem);
      for (int j = lane; j < a.out_cols; j += 32) {
        const long long o = (a.gather ? 2LL * j : i) * STRIDE;
        const float s0 = Sf[o], s1 = Sf[o + STRIDE];
        const float v = fmaf(s1 - s0, (float)rem * a.inv_d, s0);
        lo = min_nan(lo, v);
        hi = max_nan(hi, v);
        i += a.q32_step;
        rem += a.r32_step;
        if (rem >= a.D) { rem -= a.D; ++i; }
        if (i >= nm1) { i = nm1 - 1; rem = a.D; }
      }
    }
    lo = warp_min_nan(lo);
    hi = warp_max_nan(hi);
    const float inv = 1.0f / (hi - lo + a.eps);
    if (lane == 0 && !(inv == inv)) { rowflag[f] = 1; rowflag[F] = 1; }
    // deltas first (they read only the values), then the values in place
    for (int s = lane; s < a.n_slots - 1; s += 32)
      Sf[s * STRIDE + 1] = (Sf[(s + 1) * STRIDE] - Sf[s * STRIDE]) * inv;
    if (lane == 0) Sf[(a.n_slots - 1) * STRIDE + 1] = 0.f;
    __syncwarp();
    for (int s = lane; s < a.n_slots; s += 32) Sf[s * STRIDE] = (Sf[s * STRIDE] - lo) * inv;
  }
  __syncthreads();

  if constexpr (MODE == 0) {
    if (a.knots_only) {
      // The image is a pure function of these F x n_frames knot pairs (pixel = sat(value + delta * fraction)): a
      // consumer that only samples the image (row (f)2, spec_condition.cu) reads 17 KB per row instead of 264 KB.
      // Rows poisoned by a NaN sample are marked {NaN, 0}.
      float2* ko = reinterpret_cast<float2*>(a.out) + row * (long long)F * a.n_frames;
      const float2* S2 = reinterpret_cast<const float2*>(S);
      const int nf = (int)a.n_frames;
      for (int f = warp; f < F; f += nwarps) {
        const bool bad = rowflag[f] != 0;
        for (int s = lane; s < nf; s += 32)
          ko[f * nf + s] = bad ? make_float2(__int_as_float(0x7fc00000), 0.f) : S2[(long long)f * a.n_pad + s];
      }
      return;
    }
  }

  // Phase 3: image write.  A thread keeps the four brackets of its column quad in registers
  // and walks the frequency rows; a warp emits one contiguous 512 B line per row.
  const int Q = a.out_cols >> 2;
  const int q0 = (int)((unsigned)part * (unsigned)Q / (unsigned)a.split);
  const int q1 = (int)((unsigned)(part + 1) * (unsigned)Q / (unsigned)a.split);
  float* __restrict__ orow = a.out + row * (long long)F * a.out_cols;
  for (int q = q0 + tid; q < q1; q += nthreads) {
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
    const float2* Kf = reinterpret_cast<const float2*>(S);
#pragma unroll 4
    for (int f = 0; f < F; ++f, Kf += a.n_pad, o += a.out_cols) {
      float v[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float2 k = Kf[lo_i[c]];
        v[c] = __saturatef(fmaf(k.y, fr[c], k.x));   // FFMA.SAT: [0,1] like the reference
      }
      st_stream_f4(o, make_float4(v[0], v[1], v[2], v[3]));
    }
  }
  // Rows poisoned by a NaN sample: torch.min/max propagate NaN, the whole row is NaN.  The
  // saturating FFMA above flushed them to 0, so they are rewritten here (rare path).
  if (rowflag[F]) {
    const float qnan = __int_as_float(0x7fc00000);
    for (int f = 0; f < F; ++f) {
      if (!rowflag[f]) continue;
      float* o = orow + (long long)f * a.out_cols;
      for (int q = q0 + tid; q < q1; q += nthreads) st_stream_f4(o + 4 * q, make_float4(qnan, qnan, qnan, qnan));
    }
  }
}

// ---------------------------------------------------------------------- decimating plans (very long rows)
// A 1-hour record (BASELINE configs[3]: 720 000 samples, 22 501 frames) is sampled at 2 000 columns: the plan keeps
// two frames per column (`gather`), 4 000 slots, which do not fit one CTA's shared memory -- the row kernel above
// then keeps its knots in a global workspace and every CTA of a row transforms all 4 000 frames (0.5 ms for 20
// rows, 2 % of the HBM roofline).  Here the COLUMNS are cut into chunks instead: a CTA transforms the frames of
// its 128 columns (thread = slot), writes the un-normalised pixels and folds their minima / maxima into the
// row's (row, frequency) slots with atomics (min / max commute: deterministic); a second small kernel
// normalises in place.  Same arithmetic per pixel as the row kernel up to the order of the normalisation:
// fma(s1 - s0, fraction, s0), then (v - lo) / (hi - lo + eps), saturated; a NaN anywhere in a frequency row
// makes the whole row NaN like torch.min / max.
constexpr int kLongCols = 128;   // columns per CTA (two slots each)

__device__ __forceinline__ unsigned int f2ord(float f) {
  const unsigned int b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned int u) {
  return __uint_as_float((u & 0x80000000u) ? (u ^ 0x80000000u) : ~u);
}

// mm: (rows, F, 4) {ordered min, ordered max, NaN flag, unused}, preset to {0xffffffff, 0, 0, 0}
template <int LOG2N>
__global__ void __launch_bounds__(2 * kLongCols) stft_long_pixels_kernel(const StftArgs a, unsigned int* __restrict__ mm) {
  constexpr int N = 1 << LOG2N, F = N / 2 + 1, PITCH = 2 * kLongCols + 1;
  __shared__ float S[F * PITCH];
  const int tid = threadIdx.x, lane = tid & 31;
  const long long row = blockIdx.y;
  const int j0 = blockIdx.x * kLongCols;
  const float* __restrict__ xrow = a.x + row * a.row_stride;
  {
    const int j = j0 + (tid >> 1);
    float v[N];
    if (j < a.out_cols) {
      frame_fetch<LOG2N>(a, xrow, 2 * j + (tid & 1), v);
      frame_fft_mag<LOG2N>(v, [&](int k, float mag) { S[k * PITCH + tid] = log1p_pos(mag); });
    }
  }
  __syncthreads();
  const int jl = tid & (kLongCols - 1), half = tid >> 7;   // thread = (column, every second frequency row)
  const int j = j0 + jl;
  const bool live = j < a.out_cols;
  float fr = 0.f;
  if (live) {
    long long i;
    int rem;
    bracket_of(a, j, i, rem);
    fr = (float)rem * a.inv_d;
  }
  float* o = a.out + row * (long long)F * a.out_cols + j;
  for (int k = half; k < F; k += 2) {
    float v = 0.f;
    bool nan = false;
    float lo = __int_as_float(0x7f800000), hi = __int_as_float(0xff800000);
    if (live) {
      const float s0 = S[k * PITCH + 2 * jl], s1 = S[k * PITCH + 2 * jl + 1];
      v = fmaf(s1 - s0, fr, s0);
      o[(long long)k * a.out_cols] = v;
      nan = v != v;
      if (!nan) lo = hi = v;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
      lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, d));
      hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, d));
    }
    const bool any_nan = __any_sync(0xffffffffu, nan);
    if (lane == 0) {
      unsigned int* m = mm + (row * F + k) * 4;
      if (lo <= hi) {
        atomicMin(m, f2ord(lo));
        atomicMax(m + 1, f2ord(hi));
      }
      if (any_nan) atomicOr(m + 2, 1u);
    }
  }
}

__global__ void __launch_bounds__(256) stft_long_normalise_kernel(float* __restrict__ out, const unsigned int* __restrict__ mm,
                                                                  int out_cols, float eps) {
  const long long rf = blockIdx.y;                       // row * F + frequency
  const unsigned int* m = mm + rf * 4;
  const float lo = ord2f(m[0]), hi = ord2f(m[1]);
  const float inv = 1.0f / (hi - lo + eps);
  const bool bad = m[2] != 0u || !(inv == inv);
  float4* o = reinterpret_cast<float4*>(out + rf * out_cols);
  for (int q = blockIdx.x * 256 + threadIdx.x; 4 * q < out_cols; q += gridDim.x * 256) {
    float4 v = o[q];
    if (bad) {
      v.x = v.y = v.z = v.w = __int_as_float(0x7fc00000);
    } else {
      v.x = __saturatef((v.x - lo) * inv); v.y = __saturatef((v.y - lo) * inv);
      v.z = __saturatef((v.z - lo) * inv); v.w = __saturatef((v.w - lo) * inv);
    }
    o[q] = v;
  }
}

__global__ void stft_long_preset_kernel(unsigned int* mm, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) reinterpret_cast<uint4*>(mm)[i] = make_uint4(0xffffffffu, 0u, 0u, 0u);
}

template <int LOG2N>
static int stft_long_launch(const StftArgs& a, void* workspace, cudaStream_t st) {
  constexpr int F = (1 << LOG2N) / 2 + 1;
  unsigned int* mm = static_cast<unsigned int*>(workspace);
  const long long n = a.rows * F;
  stft_long_preset_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(mm, n);
  count_launch();
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return (int)e;
  stft_long_pixels_kernel<LOG2N><<<dim3((unsigned)((a.out_cols + kLongCols - 1) / kLongCols), (unsigned)a.rows), 2 * kLongCols, 0, st>>>(a, mm);
  count_launch();
  if ((e = cudaGetLastError()) != cudaSuccess) return (int)e;
  const int gx = (a.out_cols / 4 + 255) / 256;
  stft_long_normalise_kernel<<<dim3((unsigned)(gx < 1 ? 1 : gx), (unsigned)n), 256, 0, st>>>(a.out, mm, a.out_cols, a.eps);
  count_launch();
  return (int)cudaGetLastError();
}

// ------------------------------------------------------------------------------------- host
static size_t stft_fixed_smem(int nperseg, int threads, long long n_ends) {
  const int N = nperseg, P = N / 2, F = P + 1;
  const int LPF = P < 32 ? P : 32, FPW = 32 / LPF;
  size_t b = 0;
  if (N > 64) {   // the thread-per-frame path keeps no tables in shared memory
    b += sizeof(float2) * (P / 2 > 0 ? P / 2 : 1);
    b += sizeof(float2) * P;
    b += sizeof(float2) * (threads / 32) * FPW * P;
    b += sizeof(float) * N;
  }
  b += sizeof(float2) * n_ends;
  b += sizeof(int) * (F + 1 + ((F + 1) & 1));
  b += sizeof(float) * 2 * threads;   // partial row minima / maxima
  b += 16;                            // the knot array starts 16-byte aligned
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
  int split, use_ws, n_pad, threads, ends_only, n_items;
  size_t smem, ws_per_cta;
  long long grid;
};

static StftLaunchShape stft_shape(long long rows, int nperseg, const eegspec_stft_plan& pl,
                                  long long out_cols) {
  StftLaunchShape s;
  const int F = pl.n_freqs;
  s.n_pad = (int)(pl.n_slots | 1);
  s.ends_only = (!pl.gather && 2 * (pl.n_frames - 1) < out_cols &&
                 out_cols * pl.n_frames < (1LL << 31)) ? 1 : 0;
  s.n_items = s.ends_only ? (int)(pl.n_frames - 1) : (int)out_cols;   // segments | columns
  const size_t s_bytes = sizeof(float2) * (size_t)F * s.n_pad;   // {value, delta} knot pairs
  // 256-thread CTAs when four of them fit on an SM, otherwise 512 threads so that two CTAs
  // still keep 32 warps resident.
  s.threads = 256;
  size_t fixed = stft_fixed_smem(nperseg, s.threads, s.ends_only ? s.n_items : 0);
  if (nperseg <= 64) {
    // thread-per-frame FFT: one thread per slot up to the register-bound CTA size
    long long t = (pl.n_slots + 31) / 32 * 32;
    s.threads = (int)(t < 128 ? 128 : (t > kStftThreadFftMax ? kStftThreadFftMax : t));
    fixed = stft_fixed_smem(nperseg, s.threads, s.ends_only ? s.n_items : 0);
  } else if (fixed + s_bytes > 54 * 1024) {
    s.threads = kStftMaxThreads;
    fixed = stft_fixed_smem(nperseg, s.threads, s.ends_only ? s.n_items : 0);
  }
  s.use_ws = (fixed + s_bytes > kSmemBudget) ? 1 : 0;
  s.smem = fixed + (s.use_ws ? 0 : s_bytes);
  s.ws_per_cta = s.use_ws ? s_bytes : 0;
  // Small batches: slice the columns of a row over several CTAs so that all 148 SMs write.
  int resident = (int)((220 * 1024) / (s.smem + 1024));
  const int reg_threads = nperseg <= 64 ? 640 : 1024;             // ~100 | 64 registers per thread
  if (resident > reg_threads / s.threads) resident = reg_threads / s.threads;
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
  const StftLaunchShape s = stft_shape(rows, nperseg, pl, out_cols);
  return s.ws_per_cta * (size_t)s.grid;
}

template <int LOG2N, int MODE>
static int launch_mode(const StftArgs& a, long long grid, size_t smem, int threads, cudaStream_t st) {
  static std::atomic<bool> configured[64];   // idempotent set-up; atomic so that concurrent callers do not race on the flag
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute(stft_image_kernel<LOG2N, MODE>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)(kSmemBudget + 8 * 1024));
    if (e != cudaSuccess) return (int)e;
    configured[dev] = true;
  }
  stft_image_kernel<LOG2N, MODE><<<(unsigned)grid, threads, smem, st>>>(a);
  count_launch();
  return (int)cudaGetLastError();
}

template <int LOG2N>
static int launch_one(const StftArgs& a, long long grid, size_t smem, int threads, cudaStream_t st) {
  if (a.logmag_only) return launch_mode<LOG2N, 2>(a, grid, smem, threads, st);
  if (a.use_ws) return launch_mode<LOG2N, 1>(a, grid, smem, threads, st);
  return launch_mode<LOG2N, 0>(a, grid, smem, threads, st);
}

static int dispatch(int nperseg, const StftArgs& a, long long grid, size_t smem, int threads,
                    cudaStream_t st) {
  switch (nperseg) {
    case 16: return launch_one<4>(a, grid, smem, threads, st);
    case 32: return launch_one<5>(a, grid, smem, threads, st);
    case 64: return launch_one<6>(a, grid, smem, threads, st);
    case 128: return launch_one<7>(a, grid, smem, threads, st);
    case 256: return launch_one<8>(a, grid, smem, threads, st);
    case 512: return launch_one<9>(a, grid, smem, threads, st);
    case 1024: return launch_one<10>(a, grid, smem, threads, st);
    default: return EEGSPEC_EUNSUPPORTED;
  }
}

static int stft_launch_common(const float* x, long long rows, long long length, long long row_stride, int nperseg,
                              int noverlap, long long out_cols, float eps, float* out, void* workspace,
                              size_t workspace_bytes, bool knots_only, cudaStream_t st);

int stft_image_launch(const float* x, long long rows, long long length, long long row_stride,
                      int nperseg, int noverlap, long long out_cols, float eps, float* out,
                      void* workspace, size_t workspace_bytes, cudaStream_t st) {
  return stft_launch_common(x, rows, length, row_stride, nperseg, noverlap, out_cols, eps, out, workspace,
                            workspace_bytes, false, st);
}

// knots: (rows, nperseg/2+1, n_frames) float2 {normalised value, normalised delta to the next knot}.  Served when the
// knot array of a row fits shared memory and every frame is a knot (no decimating plan); EUNSUPPORTED otherwise.
int stft_knots_launch(const float* x, long long rows, long long length, long long row_stride, int nperseg,
                      int noverlap, long long out_cols, float eps, float* knots, cudaStream_t st) {
  return stft_launch_common(x, rows, length, row_stride, nperseg, noverlap, out_cols, eps, knots, nullptr, 0, true, st);
}

static int stft_launch_common(const float* x, long long rows, long long length, long long row_stride, int nperseg,
                              int noverlap, long long out_cols, float eps, float* out, void* workspace,
                              size_t workspace_bytes, bool knots_only, cudaStream_t st) {
  if (!x || !out || rows <= 0 || row_stride < length) return EEGSPEC_EINVAL;
  if (out_cols < 4 || (out_cols & 3) || out_cols > (1 << 24)) return EEGSPEC_EUNSUPPORTED;
  if (reinterpret_cast<uintptr_t>(out) & 15) return EEGSPEC_EALIGN;
  eegspec_stft_plan pl;
  int rc = stft_plan_make(length, nperseg, noverlap, out_cols, &pl);
  if (rc != EEGSPEC_OK) return rc;
  if ((rc = check_pow2(nperseg)) != EEGSPEC_OK) return rc;
  if (rows > (1LL << 27)) return EEGSPEC_EUNSUPPORTED;
  StftLaunchShape s = stft_shape(rows, nperseg, pl, out_cols);
  if (knots_only) {
    if (s.use_ws || pl.gather || pl.n_frames < 2 || pl.n_frames > (1 << 20)) return EEGSPEC_EUNSUPPORTED;
    s.split = 1;            // one CTA per row writes the row's knots
    s.grid = rows;
  }
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
  a.knots_only = knots_only ? 1 : 0;
  {
    // Row staging (thread-per-frame path): one pass over the slots, 16-byte alignment everywhere,
    // and the padded extended row has to fit the knot array it aliases.
    const long long n_ext = (pl.n_slots - 1) * (long long)pl.hop + nperseg;
    const long long stage_floats = n_ext + ((n_ext >> 5) << 2) + 8;
    a.stage = (nperseg <= 64 && !s.use_ws && !pl.gather && pl.n_slots <= s.threads && (pl.hop & 3) == 0 &&
               (pl.half & 3) == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0 && (row_stride & 3) == 0 &&
               length < (1LL << 30) &&
               stage_floats * sizeof(float) <= sizeof(float2) * (size_t)pl.n_freqs * s.n_pad) ? 1 : 0;
  }
  a.out_cols = (int)out_cols;
  a.D = (int)out_cols - 1;
  a.q_step = (pl.n_frames - 1) / a.D;
  a.r_step = (int)((pl.n_frames - 1) % a.D);
  a.q32_step = (32 * (pl.n_frames - 1)) / a.D;
  a.r32_step = (int)((32 * (pl.n_frames - 1)) % a.D);
  a.ends_only = s.ends_only;
  a.n_items = s.n_items;
  a.small_index = ((double)out_cols * (double)pl.n_frames < 4.0e9) ? 1 : 0;
  a.inv_d = 1.0f / (float)a.D;
  a.eps = eps;
  // decimating plans on the thread-per-frame FFT: columns cut into chunks (stft_long_*), the workspace the row kernel
  // would have needed (checked above) holds the (row, frequency) minima / maxima
  if (pl.gather && s.use_ws && !knots_only && rows <= 65535 && rows * pl.n_freqs <= 65535 &&
      workspace_bytes >= (size_t)rows * pl.n_freqs * 16) {
    if (nperseg == 64) return stft_long_launch<6>(a, workspace, st);
    if (nperseg == 32) return stft_long_launch<5>(a, workspace, st);
    if (nperseg == 16) return stft_long_launch<4>(a, workspace, st);
  }
  return dispatch(nperseg, a, s.grid, s.smem, s.threads, st);
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
  return dispatch(nperseg, a, rows, stft_fixed_smem(nperseg, 256, 0), 256, st);   // 256 <= kStftThreadFftMax
}

}  // namespace eegspec
