// Row (f)4: the classifier-side EEG preprocessing -- the tensor the multimodal model's EEG branch (and the
// SHAP GradientExplainer, XAI_Multimodality.py:2283-2290) receives.
//
// Replaces CombinedDataset.process_eeg, /root/reference/root/jupyter_notebooks/XAI_Multimodality.py:1148-1164,
// with its helpers :1212-1286 (the same code as /root/reference/root/src/utils/data_utils.py:183-249):
//   waves = eeg.values.T                                      (20, n)   :1151
//   butter_bandpass_filter(0.5-20 Hz @ 200 Hz, order 5)       causal lfilter along time, zero state   :1156
//   handle_nan                                                NaN -> nanmean of the (filtered) row     :1157
//   calculate_differential_signals                            18 montage differences appended -> (38, n)
//   denoise_filter                                            band-pass order 6, 4-tap forward mean (np.roll on the
//                                                             FLATTENED array), every 4th sample -> (38, ~n/4)
//   normalize                                                 per-row z-score, population std + 1e-6
//   select_and_map_channels                                   rows 0..18 and 19..36 (EKG + the first 17 differences:
//                                                             the differences start at row 20, the index at 19)
//   pad_or_truncate(3000)                                     -> (1, 37, 3000) fp32
//
// What bounds it.  The two Butterworth band-passes are 10th / 12th order DIRECT-FORM recurrences with poles at
// radius 0.995 / 0.998, run by the reference in float64.  They are so ill-conditioned (states ~1e12 times larger
// than the output) that a 1e-12 relative change of the input moves the z-scored output by 0.08: the reference's
// result is defined by its rounding sequence.  So this kernel executes SciPy's lfilter operation for operation
// (df2t_step below: unfused multiplies and adds in SciPy's order -- bit-identical recurrences), takes the montage
// differences BEFORE the second filter like the reference (filter(a) - filter(b) is not filter(a - b) at this
// conditioning), and cannot cut a row into chunks.  20 first-filter and 38 second-filter recurrences per record,
// ~2 000 double-precision operations per sample: FP64-pipe bound (64 lanes per SM), not HBM bound.
//   kernel 1  thread = (record, parquet column), one pass over time: first filter + NaN fill (a NaN turns the rest
//             of a filtered row into NaN, so the fill value -- the nanmean of the row -- is the mean of the prefix,
//             known on the spot); the filled float64 rows go to the workspace.
//   kernel 2  thread = (record, row of the 38-row array): channel rows read their own value, difference rows
//             subtract their two channels; second filter, 4-tap mean + decimation, float4 stores into the workspace.
//   kernel 3  CTA = (record, output row): z-score with the mean / std from the float64 moments kernel 2 leaves per
//             row (one pass over the decimated rows), zero padding to 3000.
// Round 2 ran kernels 1 + 2 as ONE warp per record (lanes = columns, then lanes = rows, differences by shuffle):
// 237 instructions per sample on 1.7 warps per scheduler, with 20 / 32 and 38 / 64 of the lanes busy -- 5.0 ms per
// 1 024 records, bound by the latency of its own dependent float64 chain.  One thread per recurrence keeps every
// lane busy and puts 58 independent recurrences per record on the machine instead of one warp.
#include <math.h>
#include <string.h>

#include "common.cuh"
#include "internal.h"

namespace eegspec {

constexpr int kClsMaxTaps = EEGSPEC_CLS_MAX_TAPS;   // 13

struct ClsFilterArgs {
  const float* x;        // (batch, n, n_ch) time-major
  double* y1;            // workspace: (batch, n, n_ch) first-filter output after the NaN fill, float64 like the reference's
  float* dec;            // workspace: (batch, n_rows_all, dec_pitch) decimated second-filter output (in-row terms of the 4-tap mean)
  double* head;          // workspace: (batch, n_rows_all, 8): [0..2] first second-filter outputs of the row (for the wrap),
                         //            [4], [5] sum and sum of squares of its decimated float32 samples
  int* status;           // (batch) or null: 1 = a filtered row is all NaN (the reference drops it)
  long long batch, rec_stride, n, n_dec, dec_pitch;
  int n_ch, n_rows_all;  // n_rows_all = n_ch + n_pairs rows of the reference's stacked array
  int src_a[64], src_b[64];   // row r: its minuend / subtrahend channel (src_b < 0: a channel row)
  double b1[kClsMaxTaps], a1[kClsMaxTaps], b2[kClsMaxTaps], a2[kClsMaxTaps];
};

// One direct-form-II-transposed step, in EXACTLY the operation order of SciPy's lfilter (scipy/signal/
// _lfilter.c.in, double_filt): y = z0 + b0 x;  z_i = (z_{i+1} + x b_{i+1}) - y a_{i+1};  z_last = x b - y a, every
// product and sum rounded on its own (no FMA).  This is not pedantry: the reference's direct-form 10th / 12th order
// band-passes are so ill-conditioned (states ~1e12 times larger than the output) that a 1e-12 relative change of
// the input moves the z-scored output by 0.08 -- the reference's result is defined by its rounding sequence, and
// fused multiply-adds gave 4e-2 of disagreement.  With the same sequence the recurrences are bit-identical to
// SciPy's.  SKIP: odd numerator taps are zero (Butterworth band-pass); x * 0 + z == z exactly, so they are dropped.
template <int N, bool SKIP>
__device__ __forceinline__ double df2t_step(const double (&b)[kClsMaxTaps], const double (&a)[kClsMaxTaps],
                                            double (&z)[kClsMaxTaps - 1], double x) {
  const double y = __dadd_rn(z[0], __dmul_rn(b[0], x));
#pragma unroll
  for (int i = 0; i < N - 2; ++i) {
    const double t = (SKIP && ((i + 1) & 1)) ? z[i + 1] : __dadd_rn(z[i + 1], __dmul_rn(x, b[i + 1]));
    z[i] = __dsub_rn(t, __dmul_rn(y, a[i + 1]));
  }
  z[N - 2] = __dsub_rn((SKIP && ((N - 1) & 1)) ? 0.0 : __dmul_rn(x, b[N - 1]), __dmul_rn(y, a[N - 1]));
  return y;
}

// np.nanmean(row) = np.sum(row with NaN -> 0) / count, and np.sum over a contiguous row is NumPy's PAIRWISE
// summation: blocks of <= 128 elements summed with 8 strided accumulators, combined as ((r0+r1)+(r2+r3))+((r4+r5)+
// (r6+r7)), a block remainder added one by one, blocks combined by recursive halving (left half rounded down to a
// multiple of 8).  The fill value feeds the ill-conditioned second filter, so it has to be that sum to the last bit.
// The tree only depends on the row length: its traversal is uniform across the warp.  Every block starts at a
// multiple of 8, so accumulator index = sample index mod 8.
constexpr int kPwBlock = 128, kPwDepth = 40;
struct PairwiseWalk {            // uniform: where in NumPy's recursion the current sample sits
  long long len[kPwDepth];       // frame: length of the segment
  unsigned char right[kPwDepth]; // frame: 1 = its left half is done (value on the lane's stack), now in the right half
  int depth;                     // frames in use; the top frame is the current leaf
  long long leaf_left;           // samples of the current leaf still to come
  __device__ void descend() {
    while (len[depth - 1] > kPwBlock) {
      long long n2 = len[depth - 1] / 2;
      n2 -= n2 % 8;
      len[depth] = n2;
      right[depth] = 0;
      ++depth;
    }
    leaf_left = len[depth - 1];
  }
};
__device__ __forceinline__ double pw_combine(const double (&r)[8]) {
  return __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])), __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
}

constexpr int kClsHead = 8;       // doubles per row in the `head` workspace
constexpr int kClsAhead = 3;      // blocks of eight samples the first filter's loads run ahead
constexpr int kClsThreads = 64;   // two warps per CTA: 58 recurrences per record spread over as many SM partitions as possible

// First filter: thread = (record, parquet column).  One pass over time: band-pass, NaN fill (a NaN turns the rest of
// the filtered row into NaN, so the fill value -- the nanmean of the row -- is the mean of the prefix, known on the
// spot), the filled float64 row goes to y1[rec][t][ch] for the second kernel.
template <int N1, bool SKIP>
__global__ void __launch_bounds__(kClsThreads) classifier_eeg_first_kernel(const __grid_constant__ ClsFilterArgs a) {
  const long long g = (long long)blockIdx.x * kClsThreads + threadIdx.x;
  const int nc = a.n_ch;
  if (g >= a.batch * nc) return;
  const long long rec = g / nc;
  const int ch = (int)(g - rec * nc);
  const long long n = a.n;
  const float* xp = a.x + rec * a.rec_stride + ch;
  double* yp = a.y1 + rec * n * nc + ch;
  double z1[kClsMaxTaps - 1];
#pragma unroll
  for (int i = 0; i < kClsMaxTaps - 1; ++i) z1[i] = 0.0;
  double fill = 0.0;
  bool filled = false;
  // pairwise-sum state of this thread's first-filter output (see above)
  PairwiseWalk walk;
  walk.depth = 1;
  walk.len[0] = n;
  walk.right[0] = 0;
  walk.descend();
  double pr[8], pres = 0.0, pstack[kPwDepth];
  int psp = 0;
  bool in_rest = false;          // uniform: the leaf's last (len % 8) samples are added to pres one by one
#pragma unroll
  for (int j = 0; j < 8; ++j) pr[j] = 0.0;

  // Loads run kClsAhead blocks of eight samples ahead of the recurrence: one block is ~600 cycles of arithmetic for a
  // lone warp (74 cycles per step, tools/fp64_latency_probe.cu), less than a DRAM round trip -- with one block of
  // look-ahead the kernel waited for memory on every block (255 cycles per sample, now 160; deeper changes nothing).
  float xn[kClsAhead][8];
#pragma unroll
  for (int d = 0; d < kClsAhead; ++d)
#pragma unroll
    for (int k = 0; k < 8; ++k) xn[d][k] = 8 * d + k < n ? __ldg(xp + (long long)(8 * d + k) * nc) : 0.f;
  const float* xq = xp + 8LL * kClsAhead * nc;   // the block kClsAhead after the current one
  for (long long t0 = 0; t0 < n; t0 += 8, xq += 8LL * nc) {
    float xc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      xc[k] = xn[0][k];
#pragma unroll
      for (int d = 0; d + 1 < kClsAhead; ++d) xn[d][k] = xn[d + 1][k];
    }
    if (t0 + 8 * (kClsAhead + 1) <= n) {
#pragma unroll
      for (int k = 0; k < 8; ++k) xn[kClsAhead - 1][k] = __ldg(xq + (long long)k * nc);
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) xn[kClsAhead - 1][k] = t0 + 8 * kClsAhead + k < n ? __ldg(xq + (long long)k * nc) : 0.f;
    }
    if (!in_rest && walk.leaf_left < 8) {   // uniform: fewer than 8 samples left in the leaf -> the sequential rest
      pres = pw_combine(pr);
      in_rest = true;
    }
    // Fast path (every block of a NaN-free row whose length is a multiple of 8): eight steps, ONE NaN test -- a NaN
    // output makes the state NaN, so every later output is NaN too and the last one tells.
    double yv[8];
    bool have_y = false;
    if (!filled && !in_rest && t0 + 8 <= n) {
#pragma unroll
      for (int k = 0; k < 8; ++k) yv[k] = df2t_step<N1, SKIP>(a.b1, a.a1, z1, (double)xc[k]);
      have_y = true;
      if (yv[7] == yv[7]) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          pr[k] = __dadd_rn(pr[k], yv[k]);
          yp[(long long)k * nc] = yv[k];
        }
        yp += 8LL * nc;
        goto walk_step;
      }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (t0 + k < n) {                       // uniform
        double y1;
        if (!filled) {
          y1 = have_y ? yv[k] : df2t_step<N1, SKIP>(a.b1, a.a1, z1, (double)xc[k]);
          if (y1 != y1) {   // the rest of this filtered row is NaN: nanmean = pairwise sum of the prefix / its length
            filled = true;
            const long long t = t0 + k;
            if (t == 0) {
              if (a.status) a.status[rec] = 1;
            } else {
              double cur = in_rest ? pres : pw_combine(pr);
              int sp = psp;
              for (int d = walk.depth - 2; d >= 0; --d)     // up the recursion: finished left halves are added
                if (walk.right[d]) cur = __dadd_rn(pstack[--sp], cur);
              fill = cur / (double)t;
            }
            y1 = fill;
          } else if (in_rest) {
            pres = __dadd_rn(pres, y1);
          } else {
            pr[k] = __dadd_rn(pr[k], y1);
          }
        } else {
          y1 = fill;
        }
        yp[(long long)k * nc] = y1;
      }
    }
    yp += 8LL * nc;
  walk_step:
    // ---- pairwise walk (uniform): these 8 samples (or the row's last few) are consumed
    {
      const long long used = min(8LL, min(walk.leaf_left, n - t0));
      walk.leaf_left -= used;
      if (walk.leaf_left == 0 && t0 + used < n) {   // leaf complete and more samples follow: climb, then descend
        double cur = in_rest ? pres : pw_combine(pr);
        in_rest = false;
#pragma unroll
        for (int j = 0; j < 8; ++j) pr[j] = 0.0;
        --walk.depth;                                // pop the leaf
        while (walk.depth > 0 && walk.right[walk.depth - 1]) {   // finished a right half: left + right
          cur = __dadd_rn(pstack[--psp], cur);
          --walk.depth;
        }
        // finished the left half of the top frame: keep its sum, go into the right half
        pstack[psp++] = cur;
        long long n2 = walk.len[walk.depth - 1] / 2;
        n2 -= n2 % 8;
        walk.right[walk.depth - 1] = 1;
        walk.len[walk.depth] = walk.len[walk.depth - 1] - n2;
        walk.right[walk.depth] = 0;
        ++walk.depth;
        walk.descend();
      }
    }
  }
}

// Second filter: thread = (record, row of the reference's 38-row stacked array).  A channel row reads its own
// filled first-filter output, a difference row subtracts its two channels (BEFORE the filter, like the reference);
// band-pass, 4-tap mean + decimation, float4 stores into the workspace.
template <int N2, bool SKIP>
__global__ void __launch_bounds__(kClsThreads) classifier_eeg_second_kernel(const __grid_constant__ ClsFilterArgs a) {
  const long long g = (long long)blockIdx.x * kClsThreads + threadIdx.x;
  if (g >= a.batch * a.n_rows_all) return;
  const long long rec = g / a.n_rows_all;
  const int r = (int)(g - rec * a.n_rows_all);
  const long long n = a.n;
  const int nc = a.n_ch;
  const int sa = a.src_a[r], sb = a.src_b[r];
  const bool diff = sb >= 0;
  const double* pa = a.y1 + rec * n * nc + sa;
  const double* pb = a.y1 + rec * n * nc + (diff ? sb : sa);
  float* dp = a.dec + (rec * a.n_rows_all + r) * a.dec_pitch;
  double* hp = a.head + (rec * a.n_rows_all + r) * kClsHead;
  double s1 = 0.0, s2 = 0.0;   // moments of the row as the finish kernel will read it (float32 values), for the z-score
  double z2[kClsMaxTaps - 1];
#pragma unroll
  for (int i = 0; i < kClsMaxTaps - 1; ++i) z2[i] = 0.0;

  // The next eight inputs are loaded while the current eight run through the recurrence; the subtraction of a
  // difference row waits until the values are used (doing it at load time would stall on the loads right away).
  double an[8], bn[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    an[k] = k < n ? pa[(long long)k * nc] : 0.0;
    bn[k] = (diff && k < n) ? pb[(long long)k * nc] : 0.0;
  }
  pa += 8LL * nc;
  pb += 8LL * nc;
  float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
  for (long long t0 = 0; t0 < n; t0 += 8, pa += 8LL * nc, pb += 8LL * nc) {
    double uc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) uc[k] = diff ? __dsub_rn(an[k], bn[k]) : an[k];
    if (t0 + 16 <= n) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        an[k] = pa[(long long)k * nc];
        if (diff) bn[k] = pb[(long long)k * nc];
      }
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const bool in = t0 + 8 + k < n;
        an[k] = in ? pa[(long long)k * nc] : 0.0;
        bn[k] = (diff && in) ? pb[(long long)k * nc] : 0.0;
      }
    }
    double acc[2] = {0.0, 0.0};   // the two decimated samples of this iteration
    if (t0 + 8 <= n) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const double y2 = df2t_step<N2, SKIP>(a.b2, a.a2, z2, uc[k]);
        if (t0 == 0 && k < 3) hp[k] = y2;
        acc[k >> 2] = __dadd_rn(acc[k >> 2], y2);
      }
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (t0 + k < n) {                       // uniform
          const double y2 = df2t_step<N2, SKIP>(a.b2, a.a2, z2, uc[k]);
          if (t0 + k < 3) hp[t0 + k] = y2;
          acc[k >> 2] = __dadd_rn(acc[k >> 2], y2);
        }
      }
    }
    // decimated samples m = t0 / 4 and m + 1 exist when their first term does (y[:, 0:-1:4]: t < n - 1); terms
    // beyond the row's end come from the NEXT row of the flattened array and are added by the finish kernel
    const long long m = t0 >> 2;            // even
    const bool e0 = t0 < n - 1, e1 = t0 + 4 < n - 1;
    if (e0) {
      const float v0 = (float)(acc[0] * 0.25), v1 = (float)(acc[1] * 0.25);
      s1 += (double)v0;
      s2 = fma((double)v0, (double)v0, s2);
      if (e1) {
        s1 += (double)v1;
        s2 = fma((double)v1, (double)v1, s2);
      }
      if ((m & 2) == 0) { q.x = v0; q.y = v1; }
      else { q.z = v0; q.w = v1; }
      const long long m_last = e1 ? m + 1 : m;
      if ((m & 2) || m_last == a.n_dec - 1) *reinterpret_cast<float4*>(dp + (m & ~3LL)) = q;
    }
  }
  hp[4] = s1;
  hp[5] = s2;
}

struct ClsFinishArgs {
  const float* dec;
  const double* head;
  float* out;            // (batch, 1, n_rows, fixed_length)
  long long n, n_dec, dec_pitch;
  int n_rows_all, n_rows, fixed_length;
  float eps;
};

__global__ void __launch_bounds__(256) classifier_eeg_finish_kernel(const ClsFinishArgs a) {
  // Output row o is row o of the reference's stacked array: rows 0 .. n_feat - 1 are kept, then
  // select_and_map_channels continues at index n_feat (= 19: EKG and the first 17 differences) -- one sequence.
  const int r = blockIdx.x;
  const long long rec = blockIdx.y;
  const float* d = a.dec + (rec * a.n_rows_all + r) * a.dec_pitch;
  const double* h = a.head + (rec * a.n_rows_all + r) * kClsHead;
  // the 4-tap mean of the last decimated sample may reach past the row's end: np.roll rolled the FLATTENED
  // array, so those terms are the first samples of the next row
  const long long t_last = 4 * (a.n_dec - 1);
  const int spill = (int)(t_last + 4 - a.n);       // 0 .. 2 terms taken from the next row
  double wrap = 0.0;
  for (int k = 0; k < spill; ++k) wrap += h[kClsHead + k];
  wrap *= 0.25;
  // mean and population standard deviation from the moments the second-filter kernel left (float64 sums of the
  // float32 samples; the wrap only touches the last sample).  NumPy makes two passes; the moments agree with it
  // to ~1e-13 relative on these rows (zero-mean band-passed signals), far inside the 1e-4 bar on the z-score.
  const double v_last = (double)d[a.n_dec - 1];
  const double s1 = h[4] + wrap;
  const double s2 = h[5] + ((v_last + wrap) * (v_last + wrap) - v_last * v_last);
  const double mean = s1 / (double)a.n_dec;
  const double var = s2 / (double)a.n_dec - mean * mean;
  const double sd = sqrt(var > 0.0 ? var : 0.0);
  const double inv = 1.0 / (sd + (double)a.eps);
  float* op = a.out + (rec * a.n_rows + r) * a.fixed_length;
  for (int m = threadIdx.x; m < a.fixed_length; m += blockDim.x) {
    float v = 0.f;
    if (m < a.n_dec) v = (float)(((double)d[m] + (m == a.n_dec - 1 ? wrap : 0.0) - mean) * inv);
    op[m] = v;
  }
}

// butter(5 / 6, [0.5, 20] / 100, btype='band') of SciPy, the reference's configuration (XAI_Multimodality.py:205,
// :1156, :1268); tests/test_oracle_classifier.py checks them against scipy.signal.butter.
void cls_eeg_params_default(eegspec_cls_eeg_params* p) {
  memset(p, 0, sizeof(*p));
  static const double b1[11] = {0.001151186385343785, 0.0, -0.005755931926718925, 0.0, 0.01151186385343785, 0.0,
                                -0.01151186385343785, 0.0, 0.005755931926718925, 0.0, -0.001151186385343785};
  static const double a1[11] = {1.0, -7.984802744723613, 28.780856801970636, -61.73788169912945, 87.3622028993991,
                                -85.26668263284571, 58.15374986122678, -27.370865501025044, 8.508305070326013,
                                -1.5772751257927395, 0.13239307063386374};
  static const double b2[13] = {0.00029907903377907595, 0.0, -0.0017944742026744557, 0.0, 0.004486185506686139, 0.0,
                                -0.005981580675581519, 0.0, 0.004486185506686139, 0.0, -0.0017944742026744557, 0.0,
                                0.00029907903377907595};
  static const double a2[13] = {1.0, -9.590565799057455, 42.26925892991835, -113.28827916089676, 205.76645819090106,
                                -266.95101519796117, 253.7409420393535, -178.0830856181304, 91.60031314619462,
                                -33.67762836365999, 8.400738521120699, -1.276480884387838, 0.08934419660572798};
  p->n_taps1 = 11;
  p->n_taps2 = 13;
  for (int i = 0; i < 11; ++i) { p->b1[i] = b1[i]; p->a1[i] = a1[i]; }
  for (int i = 0; i < 13; ++i) { p->b2[i] = b2[i]; p->a2[i] = a2[i]; }
  // CFG.map_features over CFG.feature_to_index (XAI_Multimodality.py:211-224)
  static const int pa[18] = {0, 4, 5, 6, 0, 1, 2, 3, 11, 15, 16, 17, 11, 12, 13, 14, 8, 9};
  static const int pb[18] = {4, 5, 6, 7, 1, 2, 3, 7, 15, 16, 17, 18, 12, 13, 14, 18, 9, 10};
  p->n_pairs = 18;
  for (int i = 0; i < 18; ++i) { p->pair_a[i] = pa[i]; p->pair_b[i] = pb[i]; }
  p->n_features = 19;
  p->fixed_length = 3000;
  p->eps = 1e-6f;
}

static long long cls_n_dec(long long n) { return n >= 2 ? (n - 2) / 4 + 1 : 0; }   // len(range(0, n - 1, 4))

size_t cls_eeg_workspace(long long batch, long long length, int n_ch) {
  if (batch <= 0 || length < 2 || n_ch <= 0) return 0;
  const long long pitch = (cls_n_dec(length) + 3) & ~3LL;
  const long long rows = n_ch + EEGSPEC_MAX_PAIRS;      // an upper bound: the pair count is a run-time parameter
  return (size_t)batch * rows * ((size_t)pitch * sizeof(float) + kClsHead * sizeof(double)) +
         (size_t)batch * (size_t)length * n_ch * sizeof(double);                     // + the float64 first-filter rows
}

int cls_eeg_launch(const float* x, long long batch, long long length, int n_ch, long long rec_stride,
                   const eegspec_cls_eeg_params* p, float* out, int* status, void* workspace, size_t workspace_bytes,
                   cudaStream_t st) {
  if (!x || !out || !p || batch <= 0 || n_ch <= 0) return EEGSPEC_EINVAL;
  if (length < 2) return EEGSPEC_ESHORT;
  if (rec_stride < length * n_ch) return EEGSPEC_EINVAL;
  if (p->n_taps1 != 11 || p->n_taps2 != 13) return EEGSPEC_EUNSUPPORTED;   // orders 5 and 6, the reference's
  if (p->n_pairs < 1 || p->n_pairs > EEGSPEC_MAX_PAIRS || p->n_features < 1 || p->n_features > n_ch) return EEGSPEC_EINVAL;
  if (n_ch > 32) return EEGSPEC_EUNSUPPORTED;                               // rows of the stacked array: n_ch + n_pairs <= 64
  if (p->fixed_length < 1) return EEGSPEC_EINVAL;
  if (fabs(p->a1[0] - 1.0) > 1e-12 || fabs(p->a2[0] - 1.0) > 1e-12) return EEGSPEC_EINVAL;
  for (int i = 0; i < p->n_pairs; ++i)
    if (p->pair_a[i] < 0 || p->pair_a[i] >= n_ch || p->pair_b[i] < 0 || p->pair_b[i] >= n_ch) return EEGSPEC_EINVAL;
  if (batch > 65535) return EEGSPEC_EUNSUPPORTED;
  const size_t need = cls_eeg_workspace(batch, length, n_ch);
  if (!workspace || workspace_bytes < need) return EEGSPEC_EWORKSPACE;
  if (reinterpret_cast<uintptr_t>(workspace) & 15) return EEGSPEC_EALIGN;
  const int n_rows_all = n_ch + p->n_pairs, n_rows = p->n_features + p->n_pairs;
  // output row o is row o of the stacked array and its 4-tap mean may reach into row o + 1
  if (n_rows > n_rows_all - 1) return EEGSPEC_EINVAL;

  ClsFilterArgs f;
  f.x = x;
  f.rec_stride = rec_stride;
  f.n = length;
  f.n_dec = cls_n_dec(length);
  f.dec_pitch = (f.n_dec + 3) & ~3LL;
  f.n_ch = n_ch;
  f.n_rows_all = n_rows_all;
  for (int r = 0; r < 64; ++r) {
    f.src_a[r] = r < n_ch ? r : (r < n_rows_all ? p->pair_a[r - n_ch] : 0);
    f.src_b[r] = r < n_ch ? -1 : (r < n_rows_all ? p->pair_b[r - n_ch] : -1);
  }
  f.batch = batch;
  f.y1 = static_cast<double*>(workspace);                                     // 8-byte data first
  f.head = f.y1 + (size_t)batch * (size_t)length * n_ch;
  f.dec = reinterpret_cast<float*>(f.head + (size_t)batch * n_rows_all * kClsHead);
  f.status = status;
  bool skip = true;
  for (int i = 0; i < kClsMaxTaps; ++i) {
    f.b1[i] = i < p->n_taps1 ? p->b1[i] : 0.0;
    f.a1[i] = i < p->n_taps1 ? p->a1[i] : 0.0;
    f.b2[i] = i < p->n_taps2 ? p->b2[i] : 0.0;
    f.a2[i] = i < p->n_taps2 ? p->a2[i] : 0.0;
    if ((i & 1) && (f.b1[i] != 0.0 || f.b2[i] != 0.0)) skip = false;
  }
  if (status) {
    const cudaError_t e = cudaMemsetAsync(status, 0, sizeof(int) * (size_t)batch, st);
    if (e != cudaSuccess) return (int)e;
  }
  const unsigned g1 = (unsigned)((batch * n_ch + kClsThreads - 1) / kClsThreads);
  const unsigned g2 = (unsigned)((batch * n_rows_all + kClsThreads - 1) / kClsThreads);
  if (skip) classifier_eeg_first_kernel<11, true><<<g1, kClsThreads, 0, st>>>(f);
  else classifier_eeg_first_kernel<11, false><<<g1, kClsThreads, 0, st>>>(f);
  count_launch();
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return (int)e;
  if (skip) classifier_eeg_second_kernel<13, true><<<g2, kClsThreads, 0, st>>>(f);
  else classifier_eeg_second_kernel<13, false><<<g2, kClsThreads, 0, st>>>(f);
  count_launch();
  if ((e = cudaGetLastError()) != cudaSuccess) return (int)e;

  ClsFinishArgs g;
  g.dec = f.dec;
  g.head = f.head;
  g.out = out;
  g.n = length;
  g.n_dec = f.n_dec;
  g.dec_pitch = f.dec_pitch;
  g.n_rows_all = n_rows_all;
  g.n_rows = n_rows;
  g.fixed_length = p->fixed_length;
  g.eps = p->eps;
  classifier_eeg_finish_kernel<<<dim3((unsigned)n_rows, (unsigned)batch), 256, 0, st>>>(g);
  count_launch();
  return (int)cudaGetLastError();
}

}  // namespace eegspec
