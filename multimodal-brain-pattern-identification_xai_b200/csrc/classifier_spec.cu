// Row (f)3: the classifier-side post-processing of the Kaggle spectrograms -- the spectrogram-branch input of the
// multimodal model (and of the LIME / saliency experiments).
//
// Replaces the numeric chain of CombinedDataset.process_spectrogram,
// /root/reference/root/jupyter_notebooks/XAI_Multimodality.py:1183-1202 with its helpers :1212-1222, :1286-1307
// (the same code as /root/reference/root/src/utils/data_utils.py:129-181), on the (400, 300) window that
// :1173-1183 cut out of the parquet frame (pure indexing: done by the caller, transform-side helper):
//   handle_nan            NaN -> nanmean of its ROW (over the 300 columns)                          :1184
//   baseline_correction   subtract the COLUMN mean (over the 400 rows)                              :1185
//   apply_notch_filter    iirnotch(60 Hz, Q 30, fs 200) + filtfilt along axis 0 (odd padding of 9,
//                         initial state lfilter_zi * first sample, forward then backward)           :1186
//   smooth_spectrogram    gaussian_filter(sigma 1): 9 taps, reflect, axis 0 then axis 1             :1187
//   normalize_signal      (x - min) / (max - min + 1e-6) over the whole image                       :1188
//   resample_spectrogram  skimage resize to (400, 300): the identity here (same shape)              :1189
//   np.tile x 3, float32, (3, 400, 300)                                                             :1190-1202
//
// Row-wise, column-wise and whole-image steps alternate, and one image is 480 KB (more than an SM's shared
// memory), so the chain is four small HBM-bound kernels with fp32 intermediates:
//   rowfill   warp per row: nanmean of the row (+ "row is all NaN" status: the reference drops such rows)
//   columns   CTA = (record, 32 columns): NaN fill, column mean, the notch recurrences forward and backward
//             (lane = column, double precision), vertical Gaussian -> workspace image
//   rows      CTA = (record, 8 rows): horizontal Gaussian -> plane 0 of the output (not yet normalised),
//             per-CTA min / max
//   planes    min / max of the record from the partials, normalise plane 0 in place, write planes 1 and 2
// Algorithmic bytes per record: 4 * 400 * 300 in + 3 * 4 * 400 * 300 out = 1.92 MB; the intermediates add
// 2.4 MB of (mostly L2-resident) traffic.
#include <math.h>
#include <string.h>

#include "common.cuh"
#include "internal.h"

namespace eegspec {

constexpr int kSpecColBlock = 32;     // columns per CTA of the column kernel
constexpr int kSpecPad = 9;           // filtfilt padlen = 3 * max(len(a), len(b))
constexpr int kSpecRowBlock = 8;      // rows per CTA of the row kernel
constexpr int kSpecMaxRadius = 8;

struct SpecArgs {
  const float* win;      // (batch, rows, cols), NaNs allowed
  float* rowfill;        // workspace (batch, rows)
  float* mid;            // workspace (batch, rows, cols)
  float2* partial;       // workspace (batch, row blocks): min, max
  float* out;            // (batch, 3, rows, cols)
  int* status;           // (batch) or null
  int rows, cols, radius, n_row_blocks;
  float eps;
  float g[kSpecMaxRadius + 1];   // Gaussian taps 0 .. radius (normalised)
  double nb[3], na[3], zi[2];
};

__global__ void __launch_bounds__(256) cls_spec_rowfill_kernel(const SpecArgs a) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int r = blockIdx.x * 8 + warp;
  const long long rec = blockIdx.y;
  if (r >= a.rows) return;
  const float* p = a.win + (rec * a.rows + r) * a.cols;
  double s = 0.0;
  int cnt = 0;
  for (int c = lane; c < a.cols; c += 32) {
    const float v = __ldg(p + c);
    if (v == v) { s += (double)v; ++cnt; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  }
  if (lane == 0) {
    a.rowfill[rec * a.rows + r] = cnt ? (float)(s / (double)cnt) : 0.f;
    if (cnt == 0 && a.status) a.status[rec] = 1;     // the reference drops the row and resizes: refused
  }
}

__device__ __forceinline__ int reflect_index(int i, int n) {   // scipy.ndimage mode='reflect': (d c b a | a b c d | d c b a)
  while (i < 0 || i >= n) i = i < 0 ? -i - 1 : 2 * n - 1 - i;
  return i;
}

__global__ void __launch_bounds__(256) cls_spec_column_kernel(const SpecArgs a) {
  extern __shared__ __align__(16) float S[];         // [rows + 2 * kSpecPad][kSpecColBlock + 1]
  constexpr int P = kSpecColBlock + 1;
  const int tid = threadIdx.x;
  const int c0 = blockIdx.x * kSpecColBlock;
  const long long rec = blockIdx.y;
  const int nc = min(kSpecColBlock, a.cols - c0);
  const float* w = a.win + rec * (long long)a.rows * a.cols;
  const float* fill = a.rowfill + rec * a.rows;
  float* X = S + kSpecPad * P;                        // data rows start after the left pad
  for (int i = tid; i < a.rows * kSpecColBlock; i += 256) {
    const int r = i / kSpecColBlock, c = i - r * kSpecColBlock;
    float v = 0.f;
    if (c < nc) {
      v = __ldg(w + (long long)r * a.cols + c0 + c);
      if (v != v) v = fill[r];                        // handle_nan
    }
    X[r * P + c] = v;
  }
  __syncthreads();
  if (tid < kSpecColBlock && tid < nc) {
    const int c = tid, n = a.rows;
    double s = 0.0;
    for (int r = 0; r < n; ++r) s += (double)X[r * P + c];
    const double mean = s / (double)n;                // baseline_correction
    auto x_at = [&](int r) { return (double)X[r * P + c] - mean; };
    // scipy.signal.filtfilt(method='pad', padtype='odd', padlen=9): odd extension, forward, backward.
    // The forward output replaces the slab in place (S row e = extended sample e; data sample r sits in S row
    // r + 9).  Left extension: reads data rows 9 .. 1 (S rows 18 .. 10) while writing S rows 0 .. 8.  Main part:
    // reads and then writes the same row.  Right extension: needs data rows n - 2 .. n - 10, whose S rows the main
    // part has already overwritten -- so those nine values are taken into registers first.
    const double x_first = x_at(0), x_last = x_at(n - 1);
    double tail[kSpecPad];
#pragma unroll
    for (int k = 0; k < kSpecPad; ++k) tail[k] = x_at(n - 2 - k);
    const double b0 = a.nb[0], b1 = a.nb[1], b2 = a.nb[2], a1 = a.na[1], a2 = a.na[2];
    const int ne = n + 2 * kSpecPad;
    float* Y = S;
    const double e0 = 2.0 * x_first - x_at(kSpecPad);
    double z0 = a.zi[0] * e0, z1 = a.zi[1] * e0;
    auto step = [&](double x) {
      const double y = fma(b0, x, z0);
      z0 = fma(b1, x, fma(-a1, y, z1));
      z1 = fma(b2, x, -a2 * y);
      return y;
    };
    for (int e = 0; e < kSpecPad; ++e) Y[e * P + c] = (float)step(2.0 * x_first - x_at(kSpecPad - e));
    for (int r = 0; r < n; ++r) Y[(r + kSpecPad) * P + c] = (float)step(x_at(r));
#pragma unroll
    for (int k = 0; k < kSpecPad; ++k) Y[(n + kSpecPad + k) * P + c] = (float)step(2.0 * x_last - tail[k]);
    const double y_last = (double)Y[(ne - 1) * P + c];
    z0 = a.zi[0] * y_last;
    z1 = a.zi[1] * y_last;
    for (int e = ne - 1; e >= 0; --e) Y[e * P + c] = (float)step((double)Y[e * P + c]);
  }
  __syncthreads();
  // vertical Gaussian (axis 0), reflect at the image's first / last row; the filtered image is S rows 9 .. 9 + rows
  float* o = a.mid + rec * (long long)a.rows * a.cols;
  for (int i = tid; i < a.rows * kSpecColBlock; i += 256) {
    const int r = i / kSpecColBlock, c = i - r * kSpecColBlock;
    if (c >= nc) continue;
    float acc = a.g[0] * X[r * P + c];
    for (int k = 1; k <= a.radius; ++k)
      acc += a.g[k] * (X[reflect_index(r - k, a.rows) * P + c] + X[reflect_index(r + k, a.rows) * P + c]);
    o[(long long)r * a.cols + c0 + c] = acc;
  }
}

__global__ void __launch_bounds__(256) cls_spec_row_kernel(const SpecArgs a) {
  extern __shared__ __align__(16) float R[];         // [kSpecRowBlock][cols]
  __shared__ float red_lo[8], red_hi[8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int r0 = blockIdx.x * kSpecRowBlock;
  const long long rec = blockIdx.y;
  const int nr = min(kSpecRowBlock, a.rows - r0);
  const float* m = a.mid + (rec * a.rows + r0) * (long long)a.cols;
  for (int i = tid; i < nr * a.cols; i += 256) R[i] = m[i];
  __syncthreads();
  float lo = INFINITY, hi = -INFINITY;
  float* o = a.out + (rec * 3 * a.rows + r0) * (long long)a.cols;   // plane 0
  for (int i = tid; i < nr * a.cols; i += 256) {
    const int r = i / a.cols, c = i - r * a.cols;
    const float* row = R + r * a.cols;
    float acc = a.g[0] * row[c];
    for (int k = 1; k <= a.radius; ++k)
      acc += a.g[k] * (row[reflect_index(c - k, a.cols)] + row[reflect_index(c + k, a.cols)]);
    o[i] = acc;
    lo = fminf(lo, acc);
    hi = fmaxf(hi, acc);
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) {
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, s));
    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, s));
  }
  if (lane == 0) { red_lo[warp] = lo; red_hi[warp] = hi; }
  __syncthreads();
  if (tid == 0) {
    for (int w = 1; w < 8; ++w) { lo = fminf(lo, red_lo[w]); hi = fmaxf(hi, red_hi[w]); }
    a.partial[rec * a.n_row_blocks + blockIdx.x] = make_float2(lo, hi);
  }
}

__global__ void __launch_bounds__(256) cls_spec_planes_kernel(const SpecArgs a) {
  __shared__ float s_lo, s_hi;
  const long long rec = blockIdx.y;
  if (threadIdx.x < 32) {
    float lo = INFINITY, hi = -INFINITY;
    for (int i = threadIdx.x; i < a.n_row_blocks; i += 32) {
      const float2 p = a.partial[rec * a.n_row_blocks + i];
      lo = fminf(lo, p.x);
      hi = fmaxf(hi, p.y);
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
      lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, s));
      hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, s));
    }
    if (threadIdx.x == 0) { s_lo = lo; s_hi = hi; }
  }
  __syncthreads();
  const float lo = s_lo, inv = 1.f / (s_hi - s_lo + a.eps);
  const long long plane = (long long)a.rows * a.cols;
  float* o = a.out + rec * 3 * plane;
  const long long n4 = plane / 4;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
    float4 v = reinterpret_cast<const float4*>(o)[i];
    v.x = (v.x - lo) * inv; v.y = (v.y - lo) * inv; v.z = (v.z - lo) * inv; v.w = (v.w - lo) * inv;
    reinterpret_cast<float4*>(o)[i] = v;
    st_stream_f4(o + plane + 4 * i, v);
    st_stream_f4(o + 2 * plane + 4 * i, v);
  }
}

void cls_spec_params_default(eegspec_cls_spec_params* p) {
  memset(p, 0, sizeof(*p));
  p->rows = 400;   // CFG.image_size (XAI_Multimodality.py:246)
  p->cols = 300;
  // scipy.signal.iirnotch(60, 30, fs=200) and lfilter_zi of it (:1294-1297)
  p->notch_b[0] = 0.9695312529087462; p->notch_b[1] = 0.5992032674528753; p->notch_b[2] = 0.9695312529087462;
  p->notch_a[0] = 1.0; p->notch_a[1] = 0.5992032674528753; p->notch_a[2] = 0.9390625058174924;
  p->notch_zi[0] = 0.0304687470912538; p->notch_zi[1] = 0.0304687470912538;
  p->gauss_sigma = 1.0f;   // :1299
  p->gauss_truncate = 4.0f;
  p->eps = 1e-6f;          // :1292
}

size_t cls_spec_workspace(long long batch, const eegspec_cls_spec_params* p) {
  if (!p || batch <= 0 || p->rows <= 0 || p->cols <= 0) return 0;
  const size_t blocks = (size_t)(p->rows + kSpecRowBlock - 1) / kSpecRowBlock;
  const size_t img = ((size_t)p->rows * p->cols * sizeof(float) * (size_t)batch + 255) & ~(size_t)255;
  const size_t fill = ((size_t)p->rows * sizeof(float) * (size_t)batch + 255) & ~(size_t)255;
  return img + fill + blocks * sizeof(float2) * (size_t)batch;
}

int cls_spec_launch(const float* win, long long batch, const eegspec_cls_spec_params* p, float* out, int* status,
                    void* workspace, size_t workspace_bytes, cudaStream_t st) {
  if (!win || !out || !p || batch <= 0) return EEGSPEC_EINVAL;
  if (p->rows < 2 * kSpecPad + 2 || p->cols < 1 || !(p->gauss_sigma > 0.f)) return EEGSPEC_EINVAL;
  if (batch > 65535 || (((long long)p->rows * p->cols) & 3)) return EEGSPEC_EUNSUPPORTED;
  if (fabs(p->notch_a[0] - 1.0) > 1e-12) return EEGSPEC_EINVAL;
  const int radius = (int)(p->gauss_truncate * p->gauss_sigma + 0.5f);      // scipy.ndimage: int(truncate * sd + 0.5)
  if (radius < 0 || radius > kSpecMaxRadius) return EEGSPEC_EUNSUPPORTED;
  const size_t need = cls_spec_workspace(batch, p);
  if (!workspace || workspace_bytes < need) return EEGSPEC_EWORKSPACE;
  if ((reinterpret_cast<uintptr_t>(workspace) | reinterpret_cast<uintptr_t>(out)) & 15) return EEGSPEC_EALIGN;
  SpecArgs a;
  a.win = win;
  a.out = out;
  a.status = status;
  a.rows = p->rows;
  a.cols = p->cols;
  a.radius = radius;
  a.n_row_blocks = (p->rows + kSpecRowBlock - 1) / kSpecRowBlock;
  a.eps = p->eps;
  {
    double w[kSpecMaxRadius + 1], sum = 0.0;
    for (int k = 0; k <= radius; ++k) { w[k] = exp(-0.5 * k * k / ((double)p->gauss_sigma * p->gauss_sigma)); sum += k ? 2.0 * w[k] : w[k]; }
    for (int k = 0; k <= kSpecMaxRadius; ++k) a.g[k] = k <= radius ? (float)(w[k] / sum) : 0.f;
  }
  for (int i = 0; i < 3; ++i) { a.nb[i] = p->notch_b[i]; a.na[i] = p->notch_a[i]; }
  a.zi[0] = p->notch_zi[0];
  a.zi[1] = p->notch_zi[1];
  char* ws = static_cast<char*>(workspace);
  const size_t img = ((size_t)p->rows * p->cols * sizeof(float) * (size_t)batch + 255) & ~(size_t)255;
  const size_t fill = ((size_t)p->rows * sizeof(float) * (size_t)batch + 255) & ~(size_t)255;
  a.mid = reinterpret_cast<float*>(ws);
  a.rowfill = reinterpret_cast<float*>(ws + img);
  a.partial = reinterpret_cast<float2*>(ws + img + fill);
  if (status) {
    const cudaError_t e = cudaMemsetAsync(status, 0, sizeof(int) * (size_t)batch, st);
    if (e != cudaSuccess) return (int)e;
  }
  cudaError_t e;
  cls_spec_rowfill_kernel<<<dim3((unsigned)((p->rows + 7) / 8), (unsigned)batch), 256, 0, st>>>(a);
  count_launch();
  if ((e = cudaGetLastError()) != cudaSuccess) return (int)e;
  const size_t smem_c = sizeof(float) * (size_t)(p->rows + 2 * kSpecPad) * (kSpecColBlock + 1);
  if (smem_c > 200 * 1024) return EEGSPEC_EUNSUPPORTED;
  if (smem_c > 48 * 1024) {
    e = cudaFuncSetAttribute(cls_spec_column_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_c);
    if (e != cudaSuccess) return (int)e;
  }
  cls_spec_column_kernel<<<dim3((unsigned)((p->cols + kSpecColBlock - 1) / kSpecColBlock), (unsigned)batch), 256, smem_c, st>>>(a);
  count_launch();
  if ((e = cudaGetLastError()) != cudaSuccess) return (int)e;
  const size_t smem_r = sizeof(float) * (size_t)kSpecRowBlock * p->cols;
  if (smem_r > 48 * 1024) return EEGSPEC_EUNSUPPORTED;
  cls_spec_row_kernel<<<dim3((unsigned)a.n_row_blocks, (unsigned)batch), 256, smem_r, st>>>(a);
  count_launch();
  if ((e = cudaGetLastError()) != cudaSuccess) return (int)e;
  cls_spec_planes_kernel<<<dim3(8, (unsigned)batch), 256, 0, st>>>(a);
  count_launch();
  return (int)cudaGetLastError();
}

}  // namespace eegspec
