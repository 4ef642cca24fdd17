// Mode R, phase 1 for nperseg <= 64: one thread transforms one whole frame in registers.
//
// A real N-point frame (N = 16, 32, 64) is packed as P = N/2 complex points, run through a fully
// unrolled decimation-in-frequency radix-2 FFT whose twiddles are compile-time literals, and
// untangled into the N/2+1 one-sided bins.  Compared with the warp-per-frame path (shuffles,
// shared-memory untangle) this needs ~6x fewer issue slots per frame: no shuffles, no shared
// memory, no table loads.  The Hann window and SciPy's 1/sum(w) 'spectrum' scale
// (/root/reference/root/src/utils/DiffEEG_utils.py:46-51 -> scipy.signal.stft) are literals too.
#pragma once
#include <type_traits>

#include "common.cuh"

namespace eegspec {

// cos / sin (2 pi i / 64) for any integer i, evaluated at compile time only.
__host__ __device__ constexpr double cos64(int i) {
  constexpr double t[33] = {1.0, 0.9951847266721969, 0.9807852804032304, 0.9569403357322088, 0.9238795325112867, 0.881921264348355, 0.8314696123025452, 0.773010453362737, 0.7071067811865476, 0.6343932841636455, 0.5555702330196023, 0.4713967368259978, 0.38268343236508984, 0.29028467725446233, 0.19509032201612833, 0.09801714032956077, 0.0, -0.09801714032956065, -0.1950903220161282, -0.29028467725446216, -0.3826834323650897, -0.4713967368259977, -0.555570233019602, -0.6343932841636454, -0.7071067811865475, -0.773010453362737, -0.8314696123025453, -0.8819212643483549, -0.9238795325112867, -0.9569403357322088, -0.9807852804032304, -0.9951847266721968, -1.0};
  i &= 63;
  return t[i <= 32 ? i : 64 - i];
}
__host__ __device__ constexpr double sin64(int i) {
  constexpr double t[33] = {0.0, 0.0980171403295606, 0.19509032201612825, 0.29028467725446233, 0.3826834323650898, 0.47139673682599764, 0.5555702330196022, 0.6343932841636455, 0.7071067811865475, 0.773010453362737, 0.8314696123025452, 0.8819212643483549, 0.9238795325112867, 0.9569403357322089, 0.9807852804032304, 0.9951847266721968, 1.0, 0.9951847266721969, 0.9807852804032304, 0.9569403357322089, 0.9238795325112867, 0.881921264348355, 0.8314696123025455, 0.7730104533627371, 0.7071067811865476, 0.6343932841636455, 0.5555702330196022, 0.47139673682599786, 0.3826834323650899, 0.2902846772544624, 0.1950903220161286, 0.09801714032956083, 0.0};
  i &= 63;
  return i <= 32 ? t[i] : -t[64 - i];
}

// Compile-time loop: f(std::integral_constant<int, I>) for I = B .. E-1.
template <int B, int E, typename F>
__device__ __forceinline__ void static_for(F&& f) {
  if constexpr (B < E) {
    f(std::integral_constant<int, B>{});
    static_for<B + 1, E>(f);
  }
}

__host__ __device__ constexpr int brev_bits(int x, int bits) {
  int r = 0;
  for (int b = 0; b < bits; ++b) r |= ((x >> b) & 1) << (bits - 1 - b);
  return r;
}

// In-register FFT of one windowed real frame; v[N] holds the raw samples on entry.
// emit(k, |X[k]|) is called for k = 0 .. N/2 with the SciPy-scaled magnitude.
template <int LOG2N, typename Emit>
__device__ __forceinline__ void frame_fft_mag(const float (&v)[1 << LOG2N], Emit&& emit) {
  constexpr int N = 1 << LOG2N, P = N / 2, LOG2P = LOG2N - 1, TS = 64 / N;
  float zr[P], zi[P];
  // Periodic Hann / sum(w) = (0.5 - 0.5 cos(2 pi j / N)) * 2 / N; the untangle's factor 1/2 is
  // folded in as well (exact: a power of two), bins 0 and N/2 undo it below.
  static_for<0, P>([&](auto R) {
    constexpr int r = decltype(R)::value;
    constexpr float w0 = (float)((0.5 - 0.5 * cos64(2 * r * TS)) / N);
    constexpr float w1 = (float)((0.5 - 0.5 * cos64((2 * r + 1) * TS)) / N);
    zr[r] = v[2 * r] * w0;
    zi[r] = v[2 * r + 1] * w1;
  });
  // Decimation in frequency: element e ends up holding Z[bitrev(e)].
  static_for<0, LOG2P>([&](auto ST) {
    constexpr int h = P >> (decltype(ST)::value + 1);
    static_for<0, P>([&](auto I) {
      constexpr int i = decltype(I)::value;
      if constexpr ((i & h) == 0) {
        constexpr int j = i + h;
        constexpr int q = (i & (h - 1)) * (P / (2 * h));   // twiddle W_P^q = exp(-2 pi i q / P)
        const float tr = zr[i] - zr[j], ti = zi[i] - zi[j];
        zr[i] += zr[j];
        zi[i] += zi[j];
        if constexpr (q == 0) {
          zr[j] = tr;
          zi[j] = ti;
        } else if constexpr (4 * q == P) {           // -i
          zr[j] = ti;
          zi[j] = -tr;
        } else if constexpr (8 * q == P) {           // (1 - i) / sqrt 2
          constexpr float c = 0.70710678118654752f;
          zr[j] = (tr + ti) * c;
          zi[j] = (ti - tr) * c;
        } else if constexpr (8 * q == 3 * P) {       // (-1 - i) / sqrt 2
          constexpr float c = 0.70710678118654752f;
          zr[j] = (ti - tr) * c;
          zi[j] = -(tr + ti) * c;
        } else {
          constexpr float c = (float)cos64(q * 2 * TS), s = (float)sin64(q * 2 * TS);
          zr[j] = fmaf(tr, c, ti * s);
          zi[j] = fmaf(ti, c, -tr * s);
        }
      }
    });
  });
  // Real-input untangle (frame scaled by 1/2 above):
  //   X[k] = S + T, X[P-k] = conj(S - T),  S = Z[k] + conj Z[P-k],  T = -i w^k (Z[k] - conj Z[P-k])
  {
    const float a0 = zr[0], b0 = zi[0];
    emit(0, 2.f * fabsf(a0 + b0));
    emit(P, 2.f * fabsf(a0 - b0));
  }
  static_for<1, P / 2 + 1>([&](auto K) {
    constexpr int k = decltype(K)::value;
    constexpr int ia = brev_bits(k, LOG2P), ib = brev_bits(P - k, LOG2P);
    const float ar = zr[ia], ai = zi[ia], br = zr[ib], bi = zi[ib];
    const float sr = ar + br, si = ai - bi, dr = ar - br, di = ai + bi;
    constexpr float c = (float)cos64(k * TS), s = (float)sin64(k * TS);   // w^k = c - i s
    const float tr = fmaf(c, di, -s * dr), ti = -fmaf(c, dr, s * di);
    const float xr = sr + tr, xi = si + ti;
    emit(k, sqrt_mufu(fmaf(xr, xr, xi * xi)));
    if constexpr (k != P - k) {
      const float yr = sr - tr, yi = si - ti;
      emit(P - k, sqrt_mufu(fmaf(yr, yr, yi * yi)));
    }
  });
}

}  // namespace eegspec
