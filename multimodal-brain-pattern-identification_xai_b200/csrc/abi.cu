// extern "C" surface of libeegspec.so (declared in include/eegspec.h).
#include <atomic>
#include <mutex>

#include "internal.h"

namespace eegspec {

static std::atomic<long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

// Montage of the north_star, as channel indices into EEG_COLUMNS
// (/root/reference/root/config/config.yml:34-54) for the first 16 map_features pairs
// (/root/reference/root/config/config.yml:529-545): chains LL, LP, RP, RR.
//   Fp1=0 F3=1 C3=2 P3=3 F7=4 T3=5 T5=6 O1=7 Fz=8 Cz=9 Pz=10 Fp2=11 F4=12 C4=13 P4=14 F8=15
//   T4=16 T6=17 O2=18 EKG=19
static const int kPairA[16] = {0, 4, 5, 6, 0, 1, 2, 3, 11, 15, 16, 17, 11, 12, 13, 14};
static const int kPairB[16] = {4, 5, 6, 7, 1, 2, 3, 7, 15, 16, 17, 18, 12, 13, 14, 18};

// ---- host-buffer pipeline -------------------------------------------------------------------
constexpr int kPipeDepth = 3;
struct HostPipe {
  bool ready = false;
  cudaStream_t stream[kPipeDepth] = {};
  void* buf[kPipeDepth][3] = {};   // in, out, workspace
  size_t cap[kPipeDepth][3] = {};
};
static std::mutex g_pipe_mu[64];   // one per device: host-buffer calls on different devices do not serialise
static HostPipe g_pipes[64];

// Every early return of a host-buffer call goes through here: asynchronous copies into / out of the
// caller's buffers may still be in flight on the pipeline streams, and the caller is about to free them.
static int pipe_drain(HostPipe& p, int status) {
  for (int s = 0; s < kPipeDepth; ++s) {
    const cudaError_t e = cudaStreamSynchronize(p.stream[s]);
    if (status == EEGSPEC_OK && e != cudaSuccess) status = (int)e;
  }
  return status;
}

// Restores the calling thread's current device on scope exit (the pipelines select `device`).
struct DeviceGuard {
  int prev = -1;
  DeviceGuard() { if (cudaGetDevice(&prev) != cudaSuccess) prev = -1; }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

static cudaError_t pipe_reserve(HostPipe& p, int slot, int which, size_t bytes) {
  if (bytes <= p.cap[slot][which]) return cudaSuccess;
  if (p.buf[slot][which]) {
    cudaError_t e = cudaFree(p.buf[slot][which]);
    if (e != cudaSuccess) return e;
    p.buf[slot][which] = nullptr;
    p.cap[slot][which] = 0;
  }
  cudaError_t e = cudaMalloc(&p.buf[slot][which], bytes);
  if (e == cudaSuccess) p.cap[slot][which] = bytes;
  return e;
}

static int pipe_open(int device, HostPipe** out) {
  if (device < 0 || device >= 64) return EEGSPEC_EINVAL;
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return (int)e;
  HostPipe& p = g_pipes[device];
  if (!p.ready) {
    for (int i = 0; i < kPipeDepth; ++i) {
      e = cudaStreamCreateWithFlags(&p.stream[i], cudaStreamNonBlocking);
      if (e != cudaSuccess) return (int)e;
    }
    p.ready = true;
  }
  *out = &p;
  return EEGSPEC_OK;
}

// Channel-major records: the rows of the electrodes some pair of the montage reads, as runs of
// consecutive rows (HMS montage over EEG_COLUMNS: rows 0-7 and 11-18 of 20; Fz, Cz, Pz and EKG are
// never read).  Returns 0 runs when every row is used or the record is time-major (one flat copy).
static int used_channel_runs(const eegspec_mel_params* p, int n_ch, int* start, int* len) {
  if (p->time_major || n_ch > 256) return 0;
  bool used[256] = {};
  const int n_pairs = p->n_chains * p->pairs_per_chain;
  for (int i = 0; i < n_pairs && i < EEGSPEC_MAX_PAIRS; ++i) {
    if (p->pair_a[i] < 0 || p->pair_a[i] >= n_ch || p->pair_b[i] < 0 || p->pair_b[i] >= n_ch) return 0;
    used[p->pair_a[i]] = used[p->pair_b[i]] = true;
  }
  int n = 0, n_used = 0;
  for (int c = 0; c < n_ch; ++c) {
    if (!used[c]) continue;
    ++n_used;
    if (n && start[n - 1] + len[n - 1] == c) ++len[n - 1];
    else if (n == EEGSPEC_MAX_PAIRS) return 0;
    else { start[n] = c; len[n] = 1; ++n; }
  }
  return n_used == n_ch ? 0 : n;
}

}  // namespace eegspec

using namespace eegspec;

extern "C" {

int eegspec_abi_version(void) { return EEGSPEC_ABI_VERSION; }

const char* eegspec_strerror(int status) {
  switch (status) {
    case EEGSPEC_OK: return "ok";
    case EEGSPEC_EINVAL: return "invalid argument (null pointer, non-positive size or bad stride)";
    case EEGSPEC_ENOVERLAP: return "noverlap must be less than nperseg.";
    case EEGSPEC_EUNSUPPORTED: return "parameter combination not implemented by the CUDA path (no fallback exists)";
    case EEGSPEC_ESHORT: return "record shorter than one window / the crop length";
    case EEGSPEC_EWORKSPACE: return "workspace missing or too small";
    case EEGSPEC_EALIGN: return "pointer not 16-byte aligned";
    case EEGSPEC_ESTATICLIMIT: return "request exceeds the built-in scratch of this no-allocation entry point (use the workspace variant)";
    default: break;
  }
  if (status > 0) return cudaGetErrorString(static_cast<cudaError_t>(status));
  return "unknown eegspec status";
}

int eegspec_stft_plan_make(int64_t length, int nperseg, int noverlap, int64_t out_cols,
                           eegspec_stft_plan* plan) {
  return stft_plan_make(length, nperseg, noverlap, out_cols, plan);
}

int64_t eegspec_interp_bracket(int64_t j, int64_t n_frames, int64_t out_cols) {
  return interp_bracket(j, n_frames, out_cols);
}

size_t eegspec_stft_image_workspace(int64_t rows, int64_t length, int nperseg, int noverlap,
                                    int64_t out_cols) {
  return stft_image_workspace(rows, length, nperseg, noverlap, out_cols);
}

int eegspec_stft_image_f32(const float* x, int64_t rows, int64_t length, int64_t row_stride,
                           int nperseg, int noverlap, int64_t out_cols, float eps, float* out,
                           void* workspace, size_t workspace_bytes, void* stream) {
  return stft_image_launch(x, rows, length, row_stride, nperseg, noverlap, out_cols, eps, out,
                           workspace, workspace_bytes, static_cast<cudaStream_t>(stream));
}

int eegspec_stft_logmag_f32(const float* x, int64_t rows, int64_t length, int64_t row_stride,
                            int nperseg, int noverlap, float* out, void* stream) {
  return stft_logmag_launch(x, rows, length, row_stride, nperseg, noverlap, out,
                            static_cast<cudaStream_t>(stream));
}

void eegspec_mel_params_default(eegspec_mel_params* p) {
  if (!p) return;
  *p = eegspec_mel_params{};
  p->sr = 200;
  p->n_fft = 1024;
  p->win_length = 128;
  p->hop = 0;
  p->n_mels = 128;
  p->fmin = 0.f;
  p->fmax = 20.f;
  p->crop = 10000;
  p->n_time_div = 256;
  p->amin = 1e-10f;
  p->top_db = 80.f;
  p->db_offset = 40.f;
  p->db_scale = 40.f;
  p->pad_mode = 0;
  p->n_chains = 4;
  p->pairs_per_chain = 4;
  for (int i = 0; i < 16; ++i) {
    p->pair_a[i] = kPairA[i];
    p->pair_b[i] = kPairB[i];
  }
  p->time_major = 0;
  p->engine = 0;
}

int eegspec_melspec_geometry(const eegspec_mel_params* p, int32_t* n_frames, int32_t* n_time) {
  MelGeometry g;
  const int rc = melspec_geometry(p, &g);
  if (rc != EEGSPEC_OK) return rc;
  if (n_frames) *n_frames = g.n_frames;
  if (n_time) *n_time = g.n_time;
  return EEGSPEC_OK;
}

size_t eegspec_melspec_workspace(int64_t batch, const eegspec_mel_params* p) {
  return melspec_workspace(batch, p);
}

size_t eegspec_melspec_workspace_default(int64_t batch, const eegspec_mel_params* p) {
  return melspec_workspace_default(batch, p);
}

int eegspec_melspec_f32(const float* eeg, int64_t batch, int32_t n_ch, int64_t length,
                        int64_t rec_stride, const eegspec_mel_params* p, float* out,
                        void* workspace, size_t workspace_bytes, void* stream) {
  return melspec_launch(eeg, batch, n_ch, length, rec_stride, p, out, workspace, workspace_bytes,
                        static_cast<cudaStream_t>(stream));
}

int eegspec_melspec_sliding_f32(const float* eeg, int32_t n_ch, int64_t total_length, int64_t stride,
                                int64_t n_windows, const eegspec_mel_params* p, float* out, void* workspace,
                                size_t workspace_bytes, void* stream) {
  return melspec_sliding_launch(eeg, n_ch, total_length, stride, n_windows, p, out, workspace, workspace_bytes,
                                static_cast<cudaStream_t>(stream));
}

int eegspec_stft_image_host_f32(const float* x, int64_t rows, int64_t length, int nperseg,
                                int noverlap, int64_t out_cols, float eps, float* out,
                                int device) {
  if (!x || !out || rows <= 0 || length <= 0) return EEGSPEC_EINVAL;
  eegspec_stft_plan pl;
  int rc = stft_plan_make(length, nperseg, noverlap, out_cols, &pl);
  if (rc != EEGSPEC_OK) return rc;
  if (device < 0 || device >= 64) return EEGSPEC_EINVAL;
  std::lock_guard<std::mutex> lock(g_pipe_mu[device]);
  DeviceGuard restore;
  HostPipe* pipe = nullptr;
  if ((rc = pipe_open(device, &pipe)) != EEGSPEC_OK) return rc;

  const size_t in_row = (size_t)length * sizeof(float);
  const size_t out_row = (size_t)pl.n_freqs * out_cols * sizeof(float);
  // ~128 MiB of image per chunk keeps three chunks in flight without hogging HBM.
  int64_t chunk = (int64_t)((128ull << 20) / out_row);
  if (chunk < 1) chunk = 1;
  if (chunk > rows) chunk = rows;
  const size_t ws_bytes = stft_image_workspace(chunk, length, nperseg, noverlap, out_cols);

  int n = 0;
  for (int64_t r0 = 0; r0 < rows; r0 += chunk, ++n) {
    const int s = n % kPipeDepth;
    const int64_t nr = rows - r0 < chunk ? rows - r0 : chunk;
    cudaError_t e;
    if ((e = pipe_reserve(*pipe, s, 0, in_row * chunk)) != cudaSuccess) return pipe_drain(*pipe, (int)e);
    if ((e = pipe_reserve(*pipe, s, 1, out_row * chunk)) != cudaSuccess) return pipe_drain(*pipe, (int)e);
    if (ws_bytes && (e = pipe_reserve(*pipe, s, 2, ws_bytes)) != cudaSuccess) return pipe_drain(*pipe, (int)e);
    cudaStream_t st = pipe->stream[s];
    e = cudaMemcpyAsync(pipe->buf[s][0], x + r0 * length, in_row * nr, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return pipe_drain(*pipe, (int)e);
    rc = stft_image_launch(static_cast<const float*>(pipe->buf[s][0]), nr, length, length, nperseg,
                           noverlap, out_cols, eps, static_cast<float*>(pipe->buf[s][1]),
                           pipe->buf[s][2], pipe->cap[s][2], st);
    if (rc != EEGSPEC_OK) return pipe_drain(*pipe, rc);
    e = cudaMemcpyAsync(out + r0 * (int64_t)pl.n_freqs * out_cols, pipe->buf[s][1], out_row * nr,
                        cudaMemcpyDeviceToHost, st);
    if (e != cudaSuccess) return pipe_drain(*pipe, (int)e);
  }
  return pipe_drain(*pipe, EEGSPEC_OK);
}

int eegspec_melspec_host_f32(const float* eeg, int64_t batch, int32_t n_ch, int64_t length,
                             const eegspec_mel_params* p, float* out, int device) {
  if (!eeg || !out || batch <= 0 || n_ch <= 0 || length <= 0) return EEGSPEC_EINVAL;
  MelGeometry g;
  int rc = melspec_geometry(p, &g);
  if (rc != EEGSPEC_OK) return rc;
  if (length < p->crop) return EEGSPEC_ESHORT;            // everything melspec_launch would refuse is refused
  for (int i = 0; i < p->n_chains * p->pairs_per_chain; ++i)   // BEFORE the first byte is copied
    if (p->pair_a[i] < 0 || p->pair_a[i] >= n_ch || p->pair_b[i] < 0 || p->pair_b[i] >= n_ch) return EEGSPEC_EINVAL;
  if (device < 0 || device >= 64) return EEGSPEC_EINVAL;
  std::lock_guard<std::mutex> lock(g_pipe_mu[device]);
  DeviceGuard restore;
  HostPipe* pipe = nullptr;
  if ((rc = pipe_open(device, &pipe)) != EEGSPEC_OK) return rc;

  const size_t in_rec = (size_t)n_ch * length * sizeof(float);
  const size_t out_rec = (size_t)p->n_mels * g.n_time * p->n_chains * sizeof(float);
  int64_t chunk = 64;
  if (chunk > batch) chunk = batch;
  const size_t ws_bytes = melspec_workspace(chunk, p);
  int run_start[EEGSPEC_MAX_PAIRS], run_len[EEGSPEC_MAX_PAIRS];
  const int n_runs = used_channel_runs(p, n_ch, run_start, run_len);

  int n = 0;
  for (int64_t r0 = 0; r0 < batch; r0 += chunk, ++n) {
    const int s = n % kPipeDepth;
    const int64_t nb = batch - r0 < chunk ? batch - r0 : chunk;
    cudaError_t e;
    if ((e = pipe_reserve(*pipe, s, 0, in_rec * chunk)) != cudaSuccess) return pipe_drain(*pipe, (int)e);
    if ((e = pipe_reserve(*pipe, s, 1, out_rec * chunk)) != cudaSuccess) return pipe_drain(*pipe, (int)e);
    if ((e = pipe_reserve(*pipe, s, 2, ws_bytes)) != cudaSuccess) return pipe_drain(*pipe, (int)e);
    cudaStream_t st = pipe->stream[s];
    if (n_runs == 0) {
      e = cudaMemcpyAsync(pipe->buf[s][0], eeg + r0 * (int64_t)n_ch * length, in_rec * nb,
                          cudaMemcpyHostToDevice, st);
      if (e != cudaSuccess) return pipe_drain(*pipe, (int)e);
    }
    for (int r = 0; r < n_runs; ++r) {   // only the electrode rows the montage reads: same layout, gaps left stale
      e = cudaMemcpy2DAsync(static_cast<float*>(pipe->buf[s][0]) + (int64_t)run_start[r] * length, in_rec,
                            eeg + (r0 * (int64_t)n_ch + run_start[r]) * length, in_rec,
                            (size_t)run_len[r] * length * sizeof(float), (size_t)nb, cudaMemcpyHostToDevice, st);
      if (e != cudaSuccess) return pipe_drain(*pipe, (int)e);
    }
    rc = melspec_launch(static_cast<const float*>(pipe->buf[s][0]), nb, n_ch, length,
                        (long long)n_ch * length, p, static_cast<float*>(pipe->buf[s][1]),
                        pipe->buf[s][2], pipe->cap[s][2], st);
    if (rc != EEGSPEC_OK) return pipe_drain(*pipe, rc);
    e = cudaMemcpyAsync(out + r0 * (int64_t)(out_rec / sizeof(float)), pipe->buf[s][1],
                        out_rec * nb, cudaMemcpyDeviceToHost, st);
    if (e != cudaSuccess) return pipe_drain(*pipe, (int)e);
  }
  return pipe_drain(*pipe, EEGSPEC_OK);
}

int eegspec_release(int device) {
  if (device < 0 || device >= 64) return EEGSPEC_EINVAL;
  std::lock_guard<std::mutex> lock(g_pipe_mu[device]);
  HostPipe& p = g_pipes[device];
  if (!p.ready) return EEGSPEC_OK;
  DeviceGuard restore;
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return (int)e;
  int status = pipe_drain(p, EEGSPEC_OK);
  for (int s = 0; s < kPipeDepth; ++s) {
    for (int w = 0; w < 3; ++w) {
      if (p.buf[s][w] && (e = cudaFree(p.buf[s][w])) != cudaSuccess && status == EEGSPEC_OK) status = (int)e;
      p.buf[s][w] = nullptr;
      p.cap[s][w] = 0;
    }
    if ((e = cudaStreamDestroy(p.stream[s])) != cudaSuccess && status == EEGSPEC_OK) status = (int)e;
    p.stream[s] = nullptr;
  }
  p.ready = false;
  return status;
}

int eegspec_melspec_host_bytes(int64_t batch, int32_t n_ch, int64_t length, const eegspec_mel_params* p,
                               int64_t* h2d_bytes, int64_t* d2h_bytes) {
  if (!p || batch < 0 || n_ch <= 0 || length <= 0) return EEGSPEC_EINVAL;
  MelGeometry g;
  const int rc = melspec_geometry(p, &g);
  if (rc != EEGSPEC_OK) return rc;
  int run_start[EEGSPEC_MAX_PAIRS], run_len[EEGSPEC_MAX_PAIRS];
  const int n_runs = used_channel_runs(p, n_ch, run_start, run_len);
  int64_t rows = n_runs ? 0 : n_ch;
  for (int r = 0; r < n_runs; ++r) rows += run_len[r];
  if (h2d_bytes) *h2d_bytes = batch * rows * length * (int64_t)sizeof(float);
  if (d2h_bytes) *d2h_bytes = batch * (int64_t)p->n_mels * g.n_time * p->n_chains * (int64_t)sizeof(float);
  return EEGSPEC_OK;
}

void eegspec_xform_params_default(eegspec_xform_params* p) {
  if (p) eegspec::xform_params_default(p);
}

int64_t eegspec_xform_out_length(int64_t length, const eegspec_xform_params* p) {
  return (p && length > 0) ? eegspec::xform_out_length(length, p) : 0;
}

int eegspec_eeg_transform_f32(const float* x, int64_t batch, int64_t length, int32_t n_ch, int64_t rec_stride,
                              const eegspec_xform_params* p, float* out, void* stream) {
  return eegspec::eeg_transform_launch(x, batch, length, n_ch, rec_stride, p, out, static_cast<cudaStream_t>(stream));
}

// Records per pass of the transform -> STFT pipeline.  Passes small enough for the transformed rows to stay
// in L2 (24 MB, ~150 records) were measured SLOWER than large ones: 1.20 ms against 1.05 ms per 1 024
// records -- sixteen short launches lose more to ramp-up and tails than the 6 % of HBM traffic the
// intermediate costs (2 x 152 KB of 5.8 MB per record).  So a pass is as large as a 320 MB workspace allows.
static int64_t xf_stft_chunk(int64_t batch, int64_t l_out, const eegspec_xform_params* p) {
  const int64_t per_rec = (int64_t)p->n_out * l_out * (int64_t)sizeof(float);
  int64_t chunk = (320LL << 20) / (per_rec > 0 ? per_rec : 1);
  if (chunk < 1) chunk = 1;
  return chunk < batch ? chunk : batch;
}

size_t eegspec_eeg_transform_stft_workspace(int64_t batch, int64_t length, const eegspec_xform_params* p, int nperseg,
                                            int noverlap, int64_t out_cols) {
  if (!p || batch <= 0 || length <= 0) return 0;
  const int64_t l_out = eegspec::xform_out_length(length, p);
  const int64_t chunk = xf_stft_chunk(batch, l_out, p);
  const size_t mid = ((size_t)chunk * p->n_out * l_out * sizeof(float) + 255) & ~(size_t)255;
  return mid + stft_image_workspace(chunk * p->n_out, l_out, nperseg, noverlap, out_cols);
}

int eegspec_eeg_transform_stft_f32(const float* x, int64_t batch, int64_t length, int32_t n_ch, int64_t rec_stride,
                                   const eegspec_xform_params* p, int nperseg, int noverlap, int64_t out_cols,
                                   float eps, float* out, void* workspace, size_t workspace_bytes, void* stream) {
  if (!x || !out || !p || batch <= 0 || length <= 0) return EEGSPEC_EINVAL;
  const size_t need = eegspec_eeg_transform_stft_workspace(batch, length, p, nperseg, noverlap, out_cols);
  if (!workspace || workspace_bytes < need) return EEGSPEC_EWORKSPACE;
  if (reinterpret_cast<uintptr_t>(workspace) & 15) return EEGSPEC_EALIGN;
  eegspec_stft_plan pl;
  const int64_t l_out = eegspec::xform_out_length(length, p);
  int rc = stft_plan_make(l_out, nperseg, noverlap, out_cols, &pl);
  if (rc != EEGSPEC_OK) return rc;
  eegspec_xform_params q = *p;
  q.out_channel_major = 1;   // rows of L_out samples: the layout the STFT kernel reads
  const int64_t chunk = xf_stft_chunk(batch, l_out, p);
  const size_t mid_bytes = ((size_t)chunk * p->n_out * l_out * sizeof(float) + 255) & ~(size_t)255;
  float* mid = static_cast<float*>(workspace);
  void* ws2 = static_cast<char*>(workspace) + mid_bytes;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  for (int64_t r0 = 0; r0 < batch; r0 += chunk) {
    const int64_t nb = batch - r0 < chunk ? batch - r0 : chunk;
    rc = eeg_transform_launch(x + r0 * rec_stride, nb, length, n_ch, rec_stride, &q, mid, st);
    if (rc != EEGSPEC_OK) return rc;
    rc = stft_image_launch(mid, nb * p->n_out, l_out, l_out, nperseg, noverlap, out_cols, eps,
                           out + r0 * (int64_t)p->n_out * pl.n_freqs * out_cols, ws2, workspace_bytes - mid_bytes, st);
    if (rc != EEGSPEC_OK) return rc;
  }
  return EEGSPEC_OK;
}

int eegspec_spec_condition_f32(const float* spec, int64_t batch, const eegspec_cond_params* p, const float* weight,
                               const float* bias, const int32_t* partner, float* out, void* stream) {
  return spec_condition_launch(spec, batch, p, weight, bias, partner, out, static_cast<cudaStream_t>(stream));
}

void eegspec_cls_eeg_params_default(eegspec_cls_eeg_params* p) {
  if (p) cls_eeg_params_default(p);
}

size_t eegspec_classifier_eeg_workspace(int64_t batch, int64_t length, int32_t n_ch) {
  return cls_eeg_workspace(batch, length, n_ch);
}

int eegspec_classifier_eeg_f32(const float* x, int64_t batch, int64_t length, int32_t n_ch, int64_t rec_stride,
                               const eegspec_cls_eeg_params* p, float* out, int32_t* status, void* workspace,
                               size_t workspace_bytes, void* stream) {
  return cls_eeg_launch(x, batch, length, n_ch, rec_stride, p, out, status, workspace, workspace_bytes,
                        static_cast<cudaStream_t>(stream));
}

void eegspec_cls_spec_params_default(eegspec_cls_spec_params* p) {
  if (p) cls_spec_params_default(p);
}

size_t eegspec_classifier_spec_workspace(int64_t batch, const eegspec_cls_spec_params* p) {
  return cls_spec_workspace(batch, p);
}

int eegspec_classifier_spec_f32(const float* win, int64_t batch, const eegspec_cls_spec_params* p, float* out,
                                int32_t* status, void* workspace, size_t workspace_bytes, void* stream) {
  return cls_spec_launch(win, batch, p, out, status, workspace, workspace_bytes, static_cast<cudaStream_t>(stream));
}

int eegspec_classifier_spec_static_f32(const float* win, int64_t batch, const eegspec_cls_spec_params* p, float* out,
                                       int32_t* status, void* stream) {
  return cls_spec_static_launch(win, batch, p, out, status, static_cast<cudaStream_t>(stream));
}

int eegspec_stft_knots_f32(const float* x, int64_t rows, int64_t length, int64_t row_stride, int nperseg, int noverlap,
                           int64_t out_cols, float eps, float* knots, void* stream) {
  return stft_knots_launch(x, rows, length, row_stride, nperseg, noverlap, out_cols, eps, knots,
                           static_cast<cudaStream_t>(stream));
}

int eegspec_spec_condition_knots_f32(const float* knots, int32_t n_frames, int64_t batch, const eegspec_cond_params* p,
                                     const float* weight, const float* bias, const int32_t* partner, float* out,
                                     void* stream) {
  return spec_condition_knots_launch(knots, n_frames, batch, p, weight, bias, partner, out,
                                     static_cast<cudaStream_t>(stream));
}

int64_t eegspec_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

}  // extern "C"
