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
//   columns   thread = (record, column), streaming global memory at full occupancy: NaN fill, column mean, the
//             notch recurrences forward and backward (double precision), vertical Gaussian in a register window
//             behind the backward sweep -> workspace image
//   rows      CTA = (record, 16 rows): horizontal Gaussian -> plane 0 of the output (not yet normalised),
//             per-CTA min / max
//   planes    min / max of the record from the partials, normalise plane 0 in place, write planes 1 and 2
// Algorithmic bytes per record: 4 * 400 * 300 in + 3 * 4 * 400 * 300 out = 1.92 MB; the intermediates add
// 2.4 MB of (mostly L2-resident) traffic.
#include <math.h>
#include <string.h>

#include <mutex>

#include "common.cuh"
#include "internal.h"

namespace eegspec {

constexpr int kSpecPad = 9;           // filtfilt padlen = 3 * max(len(a), len(b))
constexpr int kSpecRowBlock = 16;     // rows per CTA of the row kernel
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

// thread = (record, column): everything along axis 0 in one thread, straight out of global memory.  A warp covers
// 32 consecutive columns, so every row access is one coalesced 128-byte line; 300 K threads per 1 024 records give
// the latency hiding that a shared-memory slab with one recurrence warp per CTA (the first version: 1.5 ms per
// 1 024 records) could not.  Pass 1: forward notch over the odd extension of the NaN-filled column, output to the
// workspace image, column sum on the way.  Pass 2: backward notch over that (in place, descending rows) with the
// vertical Gaussian running four rows behind it in a register window; the column mean comes off at the end.
template <int RADIUS>
__global__ void __launch_bounds__(128) cls_spec_column_kernel(const SpecArgs a) {
  const int c = blockIdx.x * 128 + threadIdx.x;
  const long long rec = blockIdx.y;
  if (c >= a.cols) return;
  const int n = a.rows;
  const float* w = a.win + rec * (long long)n * a.cols + c;
  const float* fill = a.rowfill + rec * n;
  float* y = a.mid + rec * (long long)n * a.cols + c;        // forward output, then the final column
  const long long ld = a.cols;
  auto load = [&](int r) {
    const float v = __ldg(w + r * ld);
    return v == v ? v : __ldg(fill + r);                     // handle_nan
  };
  // baseline_correction subtracts the column mean BEFORE the filters.  Everything after it is linear with unit DC
  // gain (filtfilt with steady-state initial conditions over an odd extension maps a constant to itself, the
  // Gaussian taps sum to 1), so filters(x - mean) = filters(x) - mean: the mean is accumulated during the forward
  // pass and subtracted from the final column -- one pass over the image less (float64 recurrences: the two orders
  // agree to ~1e-12 of the column's scale).
  double s = 0.0;
  auto x_at = [&](int r) { return (double)load(r); };
  // scipy.signal.filtfilt(method='pad', padtype='odd', padlen=9)
  const double x_first = x_at(0), x_last = x_at(n - 1);
  const double b0 = a.nb[0], b1 = a.nb[1], b2 = a.nb[2], a1 = a.na[1], a2 = a.na[2];
  const double e0 = 2.0 * x_first - x_at(kSpecPad);
  double z0 = a.zi[0] * e0, z1 = a.zi[1] * e0;
  auto step = [&](double x) {
    const double o = fma(b0, x, z0);
    z0 = fma(b1, x, fma(-a1, o, z1));
    z1 = fma(b2, x, -a2 * o);
    return o;
  };
  for (int e = 0; e < kSpecPad; ++e) step(2.0 * x_first - x_at(kSpecPad - e));     // left extension: outputs not needed
#pragma unroll 4
  for (int r = 0; r < n; ++r) {
    const double x = x_at(r);
    s += x;
    y[r * ld] = (float)step(x);
  }
  const float mean = (float)(s / (double)n);
  double tail[kSpecPad];                                     // forward outputs over the right extension
#pragma unroll
  for (int k = 0; k < kSpecPad; ++k) tail[k] = step(2.0 * x_last - x_at(n - 2 - k));
  // backward pass: starts at the end of the extension with zi * (last forward output)
  z0 = a.zi[0] * tail[kSpecPad - 1];
  z1 = a.zi[1] * tail[kSpecPad - 1];
#pragma unroll
  for (int k = kSpecPad - 1; k >= 0; --k) step(tail[k]);
  // Rows n - 1 .. 0 come out of the backward recurrence one per step; W[k] = f[r + k] (k = 0 .. 2 RADIUS) is the
  // register window behind it and the Gaussian of row r + RADIUS is complete once f[r] is known.  Reflection
  // (scipy.ndimage 'reflect'): below the last row f[n - 1 + j] = f[n - j] is filled in when row n - 1 is emitted
  // and then simply shifts along; above the first row the recurrence has ended and f[-j] = f[j - 1] is taken
  // from the window itself (index 2 j - 1 after the shift).
  float W[2 * RADIUS + 1];
#pragma unroll
  for (int k = 0; k <= 2 * RADIUS; ++k) W[k] = 0.f;
  for (int r = n - 1; r >= -RADIUS; --r) {
    float f = 0.f;
    if (r >= 0) f = (float)step((double)y[r * ld]);
#pragma unroll
    for (int k = 2 * RADIUS; k > 0; --k) W[k] = W[k - 1];
    if (r < 0) {
#pragma unroll
      for (int k = 1; k <= 2 * RADIUS; k += 2)
        if (k == -2 * r - 1) f = W[k];
    }
    W[0] = f;
    const int rc = r + RADIUS;               // centre of the window: W[RADIUS] = f[rc]
    if (rc <= n - 1) {
      if (rc == n - 1) {
#pragma unroll
        for (int k = 1; k <= RADIUS; ++k) W[RADIUS + k] = W[RADIUS - k + 1];
      }
      float acc = a.g[0] * W[RADIUS];
#pragma unroll
      for (int k = 1; k <= RADIUS; ++k) acc += a.g[k] * (W[RADIUS - k] + W[RADIUS + k]);
      y[rc * ld] = acc - mean;               // row rc was consumed RADIUS steps ago: never read again
    }
  }
}

// CTA = (record, kSpecRowBlock rows); thread = column (and column + 256 ...), looping over the rows: no divisions,
// reflected neighbour columns resolved once per thread.
template <int RADIUS>
__global__ void __launch_bounds__(256) cls_spec_row_kernel(const SpecArgs a) {
  extern __shared__ __align__(16) float R[];         // [kSpecRowBlock][cols]
  __shared__ float red_lo[8], red_hi[8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int r0 = blockIdx.x * kSpecRowBlock;
  const long long rec = blockIdx.y;
  const int nr = min(kSpecRowBlock, a.rows - r0);
  const float* m = a.mid + (rec * a.rows + r0) * (long long)a.cols;
  const int n_el = nr * a.cols;
  if (!(a.cols & 3)) {
    for (int i = tid; i < (n_el >> 2); i += 256) reinterpret_cast<float4*>(R)[i] = reinterpret_cast<const float4*>(m)[i];
  } else {
    for (int i = tid; i < n_el; i += 256) R[i] = m[i];
  }
  __syncthreads();
  float lo = INFINITY, hi = -INFINITY;
  float* o = a.out + (rec * 3 * a.rows + r0) * (long long)a.cols;   // plane 0
  for (int c = tid; c < a.cols; c += 256) {
    int nb_l[RADIUS], nb_r[RADIUS];
#pragma unroll
    for (int k = 1; k <= RADIUS; ++k) { nb_l[k - 1] = reflect_index(c - k, a.cols); nb_r[k - 1] = reflect_index(c + k, a.cols); }
    for (int r = 0; r < nr; ++r) {
      const float* row = R + r * a.cols;
      float acc = a.g[0] * row[c];
#pragma unroll
      for (int k = 1; k <= RADIUS; ++k) acc += a.g[k] * (row[nb_l[k - 1]] + row[nb_r[k - 1]]);
      o[r * a.cols + c] = acc;
      lo = fminf(lo, acc);
      hi = fmaxf(hi, acc);
    }
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
  if (radius != 4) return EEGSPEC_EUNSUPPORTED;   // sigma 1, truncate 4: the reference's (the kernels unroll the 9 taps)
  cls_spec_column_kernel<4><<<dim3((unsigned)((p->cols + 127) / 128), (unsigned)batch), 128, 0, st>>>(a);
  count_launch();
  if ((e = cudaGetLastError()) != cudaSuccess) return (int)e;
  const size_t smem_r = sizeof(float) * (size_t)kSpecRowBlock * p->cols;
  if (smem_r > 48 * 1024) return EEGSPEC_EUNSUPPORTED;
  cls_spec_row_kernel<4><<<dim3((unsigned)a.n_row_blocks, (unsigned)batch), 256, smem_r, st>>>(a);
  count_launch();
  if ((e = cudaGetLastError()) != cudaSuccess) return (int)e;
  cls_spec_planes_kernel<<<dim3(8, (unsigned)batch), 256, 0, st>>>(a);
  count_launch();
  return (int)cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------------------
// No-allocation variant (no counterpart in the reference; added for callers
// that cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that forbid it).  The scratch the
// four kernels above pass between them lives in a file-scope device array instead of a caller workspace: every
// GPU that loads the library holds EEGSPEC_CLS_SPEC_STATIC_BYTES until exit.  All calls share it (it keeps the
// last call's contents; every byte a call reads it has written first, so nothing needs clearing beyond the status
// array that cls_spec_launch already resets): calls that overlap on the GPU race.
static __device__ __align__(256) unsigned char g_cls_spec_static_scratch[EEGSPEC_CLS_SPEC_STATIC_BYTES];

int cls_spec_static_launch(const float* win, long long batch, const eegspec_cls_spec_params* p, float* out, int* status,
                           cudaStream_t st) {
  if (!win || !out || !p || batch <= 0) return EEGSPEC_EINVAL;
  if (p->rows <= 0 || p->cols <= 0) return EEGSPEC_EINVAL;
  // overflow-safe: a batch whose workspace does not fit in 2^40 bytes is beyond the limit whatever the product says
  if ((double)p->rows * (double)p->cols * 4.0 * (double)batch > (double)EEGSPEC_CLS_SPEC_STATIC_BYTES)
    return EEGSPEC_ESTATICLIMIT;
  if (cls_spec_workspace(batch, p) > (size_t)EEGSPEC_CLS_SPEC_STATIC_BYTES) return EEGSPEC_ESTATICLIMIT;
  // In host code the array's name is a host-side stand-in, not a device address: ask the runtime for the address
  // on the CURRENT device, once per device.
  static std::mutex mu;
  static void* addr[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= 64) return EEGSPEC_EUNSUPPORTED;
  void* ptr = nullptr;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
      e = cudaGetSymbolAddress(&ptr, g_cls_spec_static_scratch);
      if (e != cudaSuccess) return (int)e;
      if (!ptr) return EEGSPEC_EWORKSPACE;
      addr[dev] = ptr;
    }
    ptr = addr[dev];
  }
  return cls_spec_launch(win, batch, p, out, status, ptr, (size_t)EEGSPEC_CLS_SPEC_STATIC_BYTES, st);
}

}  // namespace eegspec
