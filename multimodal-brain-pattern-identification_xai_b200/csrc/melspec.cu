// Mode N: bipolar montage -> NaN fill -> Hann STFT power -> mel -> dB -> (x+40)/40 -> chain mean.
//
// Front half follows the reference (pair list /root/reference/root/config/config.yml:529-545,
// differencing /root/reference/root/src/utils/data_utils.py:183-215, middle crop
// /root/reference/root/src/utils/cfg_utils.py:369-371, NaN fill cfg_utils.py:380-385); the mel/dB
// back half is the librosa recipe BASELINE.json names (see oracle/melspec_oracle.py).
//
// Engine 1 (this file): exact fp32 SIMT.
//   The 128-sample Hann window sits in a 1024-point frame and the mel bank (0..20 Hz) only
//   touches rFFT bins 1..102, so the transform is a pruned DFT, not an FFT: fold the symmetric
//   window about its centre (even part E[m], odd part O[m], m < 64) and the frame spectrum is
//   two small real matrix products  Re = E . cos, Im = -O . sin  with a constant
//   64 x 104 table pair.  |X|^2 is independent of where the window sits in the frame.
//   kernel 1  one CTA per (record, pair): difference + NaN fill into shared memory, then two
//             tiles of 128 frames: build E/O, register-tiled FFMA product (4 frames x 13 bins x
//             {cos,sin} per thread), power, 2-tap sparse mel, 10 log10 -> scratch; pair maximum.
//   kernel 2  dB referencing (needs the pair maximum), top_db clip, (x+40)/40, mean over the
//             pairs of a chain, one float4 (4 chains) per (mel, time) pixel.
#include <atomic>
#include <math.h>

#include <mutex>
#include <vector>

#include "melspec_common.cuh"

namespace eegspec {

__global__ void __launch_bounds__(kMelThreads, 1) melspec_pair_kernel(const MelArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* cos_s = reinterpret_cast<float*>(smem_raw);              // kFold*128
  float* sin_s = cos_s + kFold * kBinGroups * kGroupPitch;        // kFold*128
  float* tile = sin_s + kFold * kBinGroups * kGroupPitch;         // E | O  (2*kFold*kTileFrames)
  float* y = tile + 2 * kFold * kTileFrames;                      // crop
  float* wh = y + ((a.crop + 3) & ~3);                            // kFold
  float* mw0 = wh + kFold;                                        // n_mels
  float* mw1 = mw0 + kMaxMels;
  int* mk0 = reinterpret_cast<int*>(mw1 + kMaxMels);
  __shared__ float red_a[kMelThreads / 32], red_b[kMelThreads / 32];
  __shared__ float bcast[2];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int pair = blockIdx.x;
  const long long rec = blockIdx.y;
  const float* ea = a.eeg + rec * a.rec_stride + a.pair_a[pair] * a.ch_stride + a.crop_off * a.t_stride;
  const float* eb = a.eeg + rec * a.rec_stride + a.pair_b[pair] * a.ch_stride + a.crop_off * a.t_stride;

  for (int i = tid; i < kFold * kBinGroups * kGroupPitch; i += kMelThreads) {
    cos_s[i] = a.t.cos_t[i];
    sin_s[i] = a.t.sin_t[i];
  }
  for (int i = tid; i < kFold; i += kMelThreads) wh[i] = a.t.win_half[i];
  for (int i = tid; i < a.n_mels; i += kMelThreads) {
    mw0[i] = a.t.mel_w0[i];
    mw1[i] = a.t.mel_w1[i];
    mk0[i] = a.t.mel_k0[i];
  }

  // Montage difference + NaN statistics (cfg_utils.py:380-385 applied to the difference).
  float sum = 0.f, nans = 0.f;
  for (int i = tid; i < a.crop; i += kMelThreads) {
    const float d = __ldg(ea + i * a.t_stride) - __ldg(eb + i * a.t_stride);
    y[i] = d;
    if (d != d) nans += 1.f; else sum += d;
  }
  sum = warp_sum(sum);
  nans = warp_sum(nans);
  if (lane == 0) { red_a[warp] = sum; red_b[warp] = nans; }
  __syncthreads();
  if (warp == 0) {
    float s = lane < kMelThreads / 32 ? red_a[lane] : 0.f;
    float c = lane < kMelThreads / 32 ? red_b[lane] : 0.f;
    s = warp_sum(s);
    c = warp_sum(c);
    if (lane == 0) { bcast[0] = s; bcast[1] = c; }
  }
  __syncthreads();
  {
    const float n_nan = bcast[1];
    if (n_nan > 0.f) {
      const bool all_nan = n_nan >= (float)a.crop;
      const float fill = all_nan ? 0.f : bcast[0] / ((float)a.crop - n_nan);
      for (int i = tid; i < a.crop; i += kMelThreads) {
        const float d = y[i];
        if (all_nan) y[i] = 0.f; else if (d != d) y[i] = fill;
      }
    }
  }
  __syncthreads();

  float* E = tile;
  float* O = tile + kFold * kTileFrames;
  float* Pw = tile;  // [kBinSlots][kTileFrames], aliases E/O once the product is done
  const int fg = lane, bg = warp;          // frames 4*fg.., bins 13*bg..
  float vmax = 0.f;
  float* qbase = a.q + ((rec * a.n_pairs + pair) * (long long)a.n_mels) * a.n_time;

  const int n_tiles = (a.n_time + kTileFrames - 1) / kTileFrames;
  for (int tl = 0; tl < n_tiles; ++tl) {
    const int f0 = tl * kTileFrames;
    // Fold the windowed frame about its centre sample f*hop.
    for (int i = tid; i < kFold * kTileFrames; i += kMelThreads) {
      const int m = i / kTileFrames, t = i % kTileFrames;
      const int f = f0 + t;
      float e = 0.f, o = 0.f;
      if (f < a.n_frames) {
        const int c = f * a.hop;
        const float yp = padded_sample(y, c + m, a.crop, a.pad_mode);
        if (m == 0) {
          e = wh[0] * yp;
        } else {
          const float ym = padded_sample(y, c - m, a.crop, a.pad_mode);
          e = wh[m] * (yp + ym);
          o = wh[m] * (yp - ym);
        }
      }
      E[i] = e;
      O[i] = o;
    }
    __syncthreads();

    float ac[4][kBinsPerGroup], as[4][kBinsPerGroup];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int k = 0; k < kBinsPerGroup; ++k) { ac[r][k] = 0.f; as[r][k] = 0.f; }

#pragma unroll 2
    for (int m = 0; m < kFold; ++m) {
      const float4 e4 = *reinterpret_cast<const float4*>(E + m * kTileFrames + 4 * fg);
      const float4 o4 = *reinterpret_cast<const float4*>(O + m * kTileFrames + 4 * fg);
      const float ev[4] = {e4.x, e4.y, e4.z, e4.w}, ov[4] = {o4.x, o4.y, o4.z, o4.w};
      float cv[kGroupPitch], sv[kGroupPitch];
#pragma unroll
      for (int v = 0; v < kGroupPitch / 4; ++v) {
        const float4 c4 = *reinterpret_cast<const float4*>(cos_s + (m * kBinGroups + bg) * kGroupPitch + 4 * v);
        const float4 s4 = *reinterpret_cast<const float4*>(sin_s + (m * kBinGroups + bg) * kGroupPitch + 4 * v);
        cv[4 * v] = c4.x; cv[4 * v + 1] = c4.y; cv[4 * v + 2] = c4.z; cv[4 * v + 3] = c4.w;
        sv[4 * v] = s4.x; sv[4 * v + 1] = s4.y; sv[4 * v + 2] = s4.z; sv[4 * v + 3] = s4.w;
      }
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int k = 0; k < kBinsPerGroup; ++k) {
          ac[r][k] = fmaf(ev[r], cv[k], ac[r][k]);
          as[r][k] = fmaf(ov[r], sv[k], as[r][k]);
        }
    }
    __syncthreads();  // every warp is done reading E/O before Pw overwrites them

#pragma unroll
    for (int k = 0; k < kBinsPerGroup; ++k) {
      float4 p;
      p.x = fmaf(ac[0][k], ac[0][k], as[0][k] * as[0][k]);
      p.y = fmaf(ac[1][k], ac[1][k], as[1][k] * as[1][k]);
      p.z = fmaf(ac[2][k], ac[2][k], as[2][k] * as[2][k]);
      p.w = fmaf(ac[3][k], ac[3][k], as[3][k] * as[3][k]);
      *reinterpret_cast<float4*>(Pw + (bg * kBinsPerGroup + k) * kTileFrames + 4 * fg) = p;
    }
    __syncthreads();

    // Sparse mel projection (<= 2 adjacent taps per filter), dB numerator, running max.
    {
      const int t = tid & (kTileFrames - 1);
      const int f = f0 + t;
      const int halves = kMelThreads / kTileFrames;
      const int per = (a.n_mels + halves - 1) / halves;
      const int m0 = (tid / kTileFrames) * per;
      const int m1 = min(a.n_mels, m0 + per);
      if (f < a.n_frames) {
        for (int m = m0; m < m1; ++m) {
          const int k0 = mk0[m];
          const float v = fmaf(mw1[m], Pw[(k0 + 1) * kTileFrames + t], mw0[m] * Pw[k0 * kTileFrames + t]);
          vmax = fmaxf(vmax, v);
          if (f < a.n_time)
            qbase[(long long)m * a.n_time + f] = 3.0102999566398120f * __log2f(fmaxf(a.amin, v));
        }
      }
    }
    __syncthreads();
  }

  // Frames beyond the kept width (frame 256 of 257) only contribute to the pair maximum.
  const int n_extra = a.n_frames - n_tiles * kTileFrames;
  if (n_extra > 0) {
    float* Px = tile;  // [n_extra][kBinSlots]
    for (int i = tid; i < n_extra * kBinSlots; i += kMelThreads) {
      const int x = i / kBinSlots, ks = i % kBinSlots;
      const int c = (n_tiles * kTileFrames + x) * a.hop;
      const int g = ks / kBinsPerGroup, kk = ks % kBinsPerGroup;
      float re = 0.f, im = 0.f;
      for (int m = 0; m < kFold; ++m) {
        const float yp = padded_sample(y, c + m, a.crop, a.pad_mode);
        const float ym = m ? padded_sample(y, c - m, a.crop, a.pad_mode) : 0.f;
        const float e = wh[m] * (yp + ym), o = m ? wh[m] * (yp - ym) : 0.f;
        re = fmaf(e, cos_s[(m * kBinGroups + g) * kGroupPitch + kk], re);
        im = fmaf(o, sin_s[(m * kBinGroups + g) * kGroupPitch + kk], im);
      }
      Px[i] = fmaf(re, re, im * im);
    }
    __syncthreads();
    for (int i = tid; i < n_extra * a.n_mels; i += kMelThreads) {
      const int x = i / a.n_mels, m = i % a.n_mels;
      const int k0 = mk0[m];
      vmax = fmaxf(vmax, fmaf(mw1[m], Px[x * kBinSlots + k0 + 1], mw0[m] * Px[x * kBinSlots + k0]));
    }
  }

  vmax = warp_max_nan(vmax);
  __syncthreads();
  if (lane == 0) red_a[warp] = vmax;
  __syncthreads();
  if (tid == 0) {
    float v = red_a[0];
    for (int w = 1; w < kMelThreads / 32; ++w) v = max_nan(v, red_a[w]);
    a.refdb[rec * a.n_pairs + pair] = 3.0102999566398120f * __log2f(fmaxf(a.amin, v));
  }
}

// out[rec][mel][t][chain] = mean over the chain's pairs of (max(q - ref, -top_db) + off)/scale
__global__ void __launch_bounds__(256) melspec_combine_kernel(const MelArgs a, long long n_pix) {
  const long long pix = blockIdx.x * 256LL + threadIdx.x;  // (rec, mel, t)
  if (pix >= n_pix) return;
  const int per_rec = a.n_mels * a.n_time;
  const long long rec = pix / per_rec;
  const int mt = (int)(pix - rec * per_rec);
  const float* q = a.q + rec * a.n_pairs * (long long)per_rec + mt;
  const float* ref = a.refdb + rec * a.n_pairs;
  float* o = a.out + pix * a.n_chains;
  const float inv_ppc = 1.0f / (float)a.ppc;
  for (int c = 0; c < a.n_chains; c += 4) {
    float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) {
      if (c + cc < a.n_chains) {
        float s = 0.f;
        for (int p = 0; p < a.ppc; ++p) {
          const int pr = (c + cc) * a.ppc + p;
          const float db = fmaxf(__ldg(q + (long long)pr * per_rec) - ref[pr], -a.top_db);
          s += (db + a.db_offset) * a.inv_db_scale;
        }
        v[cc] = s * inv_ppc;
      }
    }
    if (a.n_chains - c >= 4 && (a.n_chains & 3) == 0) {
      st_stream_f4(o + c, make_float4(v[0], v[1], v[2], v[3]));
    } else {
      for (int cc = 0; cc < 4 && c + cc < a.n_chains; ++cc) o[c + cc] = v[cc];
    }
  }
}

// ------------------------------------------------------------------------------------- host
static double hz_to_mel(double f) {
  const double f_sp = 200.0 / 3, min_log_hz = 1000.0, min_log_mel = min_log_hz / f_sp;
  const double logstep = log(6.4) / 27.0;
  return f >= min_log_hz ? min_log_mel + log(f / min_log_hz) / logstep : f / f_sp;
}
static double mel_to_hz(double m) {
  const double f_sp = 200.0 / 3, min_log_hz = 1000.0, min_log_mel = min_log_hz / f_sp;
  const double logstep = log(6.4) / 27.0;
  return m >= min_log_mel ? min_log_hz * exp(logstep * (m - min_log_mel)) : f_sp * m;
}

// librosa.filters.mel(htk=False, norm='slaney', dtype=float32), row-major (n_mels, n_fft/2+1).
static void slaney_filterbank(const eegspec_mel_params& p, std::vector<float>& w) {
  const int nb = p.n_fft / 2 + 1, nm = p.n_mels;
  std::vector<double> fftf(nb), melf(nm + 2);
  const double val = 1.0 / (p.n_fft * (1.0 / p.sr));
  for (int k = 0; k < nb; ++k) fftf[k] = k * val;
  const double m_lo = hz_to_mel(p.fmin), m_hi = hz_to_mel(p.fmax);
  const double step = (m_hi - m_lo) / (nm + 1);
  for (int i = 0; i < nm + 2; ++i) melf[i] = mel_to_hz(i == nm + 1 ? m_hi : i * step + m_lo);
  w.assign((size_t)nm * nb, 0.f);
  for (int i = 0; i < nm; ++i) {
    const double fd0 = melf[i + 1] - melf[i], fd1 = melf[i + 2] - melf[i + 1];
    const double enorm = 2.0 / (melf[i + 2] - melf[i]);
    for (int k = 0; k < nb; ++k) {
      const double lower = -(melf[i] - fftf[k]) / fd0;
      const double upper = (melf[i + 2] - fftf[k]) / fd1;
      const float tri = (float)fmax(0.0, fmin(lower, upper));
      w[(size_t)i * nb + k] = (float)((double)tri * enorm);
    }
  }
}

struct MelHostTables {
  MelGeometry g;
  std::vector<float> cos_t, sin_t, win_half, w0, w1;
  std::vector<int> k0;
  std::vector<unsigned short> tc_tab, tc3_tab;
  bool chain_ok = false;
};

static int validate(const eegspec_mel_params* p) {
  if (!p) return EEGSPEC_EINVAL;
  if (p->sr <= 0 || p->n_fft <= 0 || p->n_mels <= 0 || p->crop <= 0 || p->n_time_div <= 0)
    return EEGSPEC_EINVAL;
  if (p->n_chains <= 0 || p->pairs_per_chain <= 0 ||
      p->n_chains * p->pairs_per_chain > EEGSPEC_MAX_PAIRS)
    return EEGSPEC_EINVAL;
  if (!(p->fmax > p->fmin) || p->fmin < 0 || p->fmax > p->sr * 0.5f) return EEGSPEC_EINVAL;
  if (!(p->db_scale != 0.f) || !(p->amin > 0.f)) return EEGSPEC_EINVAL;
  if (p->win_length != 2 * kFold || p->n_fft < p->win_length) return EEGSPEC_EUNSUPPORTED;
  if (p->n_mels > kMaxMels || p->crop > kMaxCrop || p->crop < p->win_length) return EEGSPEC_EUNSUPPORTED;
  if (p->pad_mode != 0 && p->pad_mode != 1) return EEGSPEC_EUNSUPPORTED;
  if (p->engine < 0 || p->engine > 3) return EEGSPEC_EUNSUPPORTED;
  return EEGSPEC_OK;
}

// Frame geometry only (cheap: runs on every call; the filter bank is built once per parameter set).
static int frame_geometry(const eegspec_mel_params& p, MelGeometry& g) {
  const int hop = p.hop > 0 ? p.hop : p.crop / p.n_time_div;
  if (hop <= 0) return EEGSPEC_EINVAL;
  g.hop = hop;
  g.n_frames = 1 + p.crop / hop;  // centred: 1 + (crop + n_fft - n_fft)//hop
  g.n_time = (g.n_frames / 32) * 32;
  g.n_pairs = p.n_chains * p.pairs_per_chain;
  g.k_lo = g.k_hi = 0;
  if (g.n_time <= 0) return EEGSPEC_ESHORT;
  if (g.n_frames - g.n_time > 2 * kFold * kTileFrames / kBinSlots) return EEGSPEC_EUNSUPPORTED;
  return EEGSPEC_OK;
}

static int build_host_tables(const eegspec_mel_params& p, MelHostTables& h) {
  const int grc = frame_geometry(p, h.g);
  if (grc != EEGSPEC_OK) return grc;

  std::vector<float> fb;
  slaney_filterbank(p, fb);
  const int nb = p.n_fft / 2 + 1;
  int k_lo = nb, k_hi = -1;
  for (int i = 0; i < p.n_mels; ++i)
    for (int k = 0; k < nb; ++k)
      if (fb[(size_t)i * nb + k] > 0.f) { k_lo = k < k_lo ? k : k_lo; k_hi = k > k_hi ? k : k_hi; }
  if (k_hi < 0) { k_lo = 0; k_hi = 0; }
  if (k_hi - k_lo + 1 > kBinSlots - 1) return EEGSPEC_EUNSUPPORTED;
  h.g.k_lo = k_lo;
  h.g.k_hi = k_hi;

  h.k0.assign(p.n_mels, 0);
  h.w0.assign(p.n_mels, 0.f);
  h.w1.assign(p.n_mels, 0.f);
  for (int i = 0; i < p.n_mels; ++i) {
    int first = -1, last = -1;
    for (int k = 0; k < nb; ++k)
      if (fb[(size_t)i * nb + k] > 0.f) { if (first < 0) first = k; last = k; }
    if (first < 0) continue;                       // empty filter: weight 0 on slot 0
    if (last - first > 1) return EEGSPEC_EUNSUPPORTED;  // more than two taps per filter
    h.k0[i] = first - k_lo;
    h.w0[i] = fb[(size_t)i * nb + first];
    h.w1[i] = last > first ? fb[(size_t)i * nb + last] : 0.f;
  }

  const size_t tsz = (size_t)kFold * kBinGroups * kGroupPitch;
  h.cos_t.assign(tsz, 0.f);
  h.sin_t.assign(tsz, 0.f);
  const double theta = 2.0 * M_PI / p.n_fft;
  for (int m = 0; m < kFold; ++m)
    for (int ks = 0; ks <= k_hi - k_lo; ++ks) {
      const int g = ks / kBinsPerGroup, kk = ks % kBinsPerGroup;
      // reduce k*m modulo n_fft before the trig call: exact in integers
      const long long km = ((long long)(k_lo + ks) * m) % p.n_fft;
      h.cos_t[((size_t)m * kBinGroups + g) * kGroupPitch + kk] = (float)cos(theta * km);
      h.sin_t[((size_t)m * kBinGroups + g) * kGroupPitch + kk] = (float)sin(theta * km);
    }
  h.win_half.assign(kFold, 0.f);
  for (int m = 0; m < kFold; ++m)
    h.win_half[m] = (float)(0.5 - 0.5 * cos(2.0 * M_PI * (kFold + m) / (2.0 * kFold)));
  {  // engine 2: the same tables as [bin slot][m] doubles, split into fp16 halves
    const int nb_used = k_hi - k_lo + 1;
    std::vector<double> cd((size_t)nb_used * kFold), sd((size_t)nb_used * kFold);
    for (int ks = 0; ks < nb_used; ++ks)
      for (int m = 0; m < kFold; ++m) {
        const long long km = ((long long)(k_lo + ks) * m) % p.n_fft;
        cd[(size_t)ks * kFold + m] = cos(theta * km);
        sd[(size_t)ks * kFold + m] = sin(theta * km);
      }
    h.tc_tab.assign(kTcTableHalves, 0);
    melspec_tc_build_tables(cd.data(), sd.data(), nb_used, h.tc_tab.data());
    // engine 3 folds the window into the tables; E[0] = y[c] + y[c] is halved there
    for (int ks = 0; ks < nb_used; ++ks)
      for (int m = 0; m < kFold; ++m) {
        const double w = (0.5 - 0.5 * cos(2.0 * M_PI * (kFold + m) / (2.0 * kFold))) * (m == 0 ? 0.5 : 1.0);
        cd[(size_t)ks * kFold + m] *= w;
        sd[(size_t)ks * kFold + m] *= w;
      }
    h.tc3_tab.assign(kTcTableHalves, 0);
    melspec_tc_build_tables(cd.data(), sd.data(), nb_used, h.tc3_tab.data());
    h.chain_ok = melspec_chain_bank_matches(p, k_lo, h.k0.data(), h.w0.data(), h.w1.data());
  }
  return EEGSPEC_OK;
}

int melspec_geometry(const eegspec_mel_params* p, MelGeometry* g) {
  int rc = validate(p);
  if (rc != EEGSPEC_OK) return rc;
  MelGeometry tmp;
  rc = frame_geometry(*p, tmp);
  if (rc == EEGSPEC_OK && g) *g = tmp;
  return rc;
}

struct TableKey {
  int dev, sr, n_fft, win_length, n_mels;
  float fmin, fmax;
  bool operator==(const TableKey& o) const {
    return dev == o.dev && sr == o.sr && n_fft == o.n_fft && win_length == o.win_length &&
           n_mels == o.n_mels && fmin == o.fmin && fmax == o.fmax;
  }
};
struct TableEntry { TableKey key; MelTables dev; MelGeometry g; bool chain_ok; };

static std::mutex g_table_mu;
static std::vector<TableEntry> g_tables;

template <typename T>
static cudaError_t upload(T** dst, const std::vector<T>& src) {
  cudaError_t e = cudaMalloc(reinterpret_cast<void**>(dst), src.size() * sizeof(T));
  if (e != cudaSuccess) return e;
  return cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice);
}

// Device tables for (device, filterbank parameters); built once, kept for the process lifetime.
static int device_tables(const eegspec_mel_params& p, MelTables* out, MelGeometry* g, bool* chain_ok) {
  int dev = 0;
  cudaError_t ce = cudaGetDevice(&dev);
  if (ce != cudaSuccess) return (int)ce;
  int rc = frame_geometry(p, *g);
  if (rc != EEGSPEC_OK) return rc;
  const TableKey key{dev, p.sr, p.n_fft, p.win_length, p.n_mels, p.fmin, p.fmax};
  std::lock_guard<std::mutex> lock(g_table_mu);
  for (const TableEntry& e : g_tables)
    if (e.key == key) {
      *out = e.dev;
      *chain_ok = e.chain_ok;
      g->k_lo = e.g.k_lo;
      g->k_hi = e.g.k_hi;
      return EEGSPEC_OK;
    }
  MelHostTables h;   // first use of this bank on this device: build and upload (milliseconds)
  if ((rc = build_host_tables(p, h)) != EEGSPEC_OK) return rc;
  *g = h.g;
  *chain_ok = h.chain_ok;
  TableEntry e;
  e.key = key;
  e.g = h.g;
  if ((ce = upload(&e.dev.cos_t, h.cos_t)) != cudaSuccess) return (int)ce;
  if ((ce = upload(&e.dev.sin_t, h.sin_t)) != cudaSuccess) return (int)ce;
  if ((ce = upload(&e.dev.win_half, h.win_half)) != cudaSuccess) return (int)ce;
  if ((ce = upload(&e.dev.mel_k0, h.k0)) != cudaSuccess) return (int)ce;
  if ((ce = upload(&e.dev.mel_w0, h.w0)) != cudaSuccess) return (int)ce;
  if ((ce = upload(&e.dev.mel_w1, h.w1)) != cudaSuccess) return (int)ce;
  if ((ce = upload(&e.dev.tc_tab, h.tc_tab)) != cudaSuccess) return (int)ce;
  if ((ce = upload(&e.dev.tc3_tab, h.tc3_tab)) != cudaSuccess) return (int)ce;
  e.chain_ok = h.chain_ok;
  g_tables.push_back(e);
  *out = e.dev;
  return EEGSPEC_OK;
}

constexpr long long kMelChunk = 256;  // records per scratch pass

// ---------------------------------------------------------------- time-major records (parquet layout)
// The reference reads records as (L, 20) rows of a parquet frame.  Engine 3 wants electrode rows
// (TMA row copies), so a time-major batch first goes through this gather: the crop window of the
// electrodes some pair reads, as compact channel-major rows in the workspace.  One tile of 256
// samples x n_ch per CTA through shared memory (row pitch n_ch + 1: a warp reading one electrode of
// 32 consecutive samples touches 32 banks); loads and stores are full 128-byte lines.
constexpr int kGatherTile = 256;
struct GatherRows { int n, r[2 * EEGSPEC_MAX_PAIRS]; };

__global__ void __launch_bounds__(kGatherTile) gather_rows_kernel(const float* __restrict__ x, long long rec_stride,
                                                                  int n_ch, int t0, int crop, const GatherRows rows,
                                                                  float* __restrict__ out) {
  extern __shared__ float tile[];   // [kGatherTile][n_ch + 1]
  const int tid = threadIdx.x, s0 = blockIdx.x * kGatherTile;
  const long long rec = blockIdx.y;
  const int n_valid = min(kGatherTile, crop - s0);
  const float* src = x + rec * rec_stride + (long long)(t0 + s0) * n_ch;
  const int n_el = n_valid * n_ch, pitch = n_ch + 1;
  if (!(n_ch & 3) && !(reinterpret_cast<uintptr_t>(src) & 15)) {
    for (int e = 4 * tid; e < n_el; e += 4 * kGatherTile) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(src + e));
      const int s = e / n_ch, c = e - s * n_ch;   // the four stay in one sample row (n_ch % 4 == 0)
      float* d = tile + s * pitch + c;
      d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
    }
  } else {
    for (int e = tid; e < n_el; e += kGatherTile) {
      const int s = e / n_ch;
      tile[s * pitch + (e - s * n_ch)] = __ldg(src + e);
    }
  }
  __syncthreads();
  if (tid < n_valid) {
    float* o = out + rec * (long long)rows.n * crop + s0 + tid;
    for (int u = 0; u < rows.n; ++u) o[(long long)u * crop] = tile[tid * pitch + rows.r[u]];
  }
}

// Distinct electrodes of the montage in ascending order; pair indices rewritten to positions in it.
static int compact_rows(const eegspec_mel_params& p, int* rows, int* pair_a, int* pair_b) {
  const int n_pairs = p.n_chains * p.pairs_per_chain;
  int n = 0;
  for (int pass = 0; pass < 2; ++pass)
    for (int i = 0; i < n_pairs; ++i) {
      const int c = pass ? p.pair_b[i] : p.pair_a[i];
      bool seen = false;
      for (int j = 0; j < n; ++j) seen |= rows[j] == c;
      if (!seen) rows[n++] = c;
    }
  for (int i = 1; i < n; ++i)   // insertion sort: <= 64 entries
    for (int j = i; j > 0 && rows[j - 1] > rows[j]; --j) { const int t = rows[j]; rows[j] = rows[j - 1]; rows[j - 1] = t; }
  for (int i = 0; i < n_pairs; ++i)
    for (int j = 0; j < n; ++j) {
      if (rows[j] == p.pair_a[i]) pair_a[i] = j;
      if (rows[j] == p.pair_b[i]) pair_b[i] = j;
    }
  return n;
}

constexpr long long kGatherChunk = 1024;   // records per gather + engine-3 pass

static size_t gather_scratch_per_record(const eegspec_mel_params& p) {
  int rows[2 * EEGSPEC_MAX_PAIRS], pa[EEGSPEC_MAX_PAIRS], pb[EEGSPEC_MAX_PAIRS];
  return (size_t)compact_rows(p, rows, pa, pb) * (size_t)p.crop * sizeof(float);
}

static size_t per_record_scratch(const eegspec_mel_params& p, const MelGeometry& g) {
  const size_t q = (size_t)g.n_pairs * p.n_mels * g.n_time * sizeof(float);
  const size_t r = (size_t)g.n_pairs * sizeof(float);
  return q + ((r + 15) & ~(size_t)15);
}

size_t melspec_workspace(long long batch, const eegspec_mel_params* p) {
  MelGeometry g;
  if (batch <= 0 || melspec_geometry(p, &g) != EEGSPEC_OK) return 0;
  const long long chunk = batch < kMelChunk ? batch : kMelChunk;
  size_t need = per_record_scratch(*p, g) * (size_t)chunk;
  if (p->time_major) {   // room for the gathered electrode rows of one engine-3 pass (+ the row list)
    const size_t tm = gather_scratch_per_record(*p) * (size_t)(batch < kGatherChunk ? batch : kGatherChunk);
    if (tm > need) need = tm;
  }
  return need;
}

size_t melspec_workspace_default(long long batch, const eegspec_mel_params* p) {
  MelGeometry g;
  if (batch <= 0 || melspec_geometry(p, &g) != EEGSPEC_OK || !p->time_major) return 0;
  return gather_scratch_per_record(*p) * (size_t)(batch < kGatherChunk ? batch : kGatherChunk);
}

static size_t pair_kernel_smem(const eegspec_mel_params& p) {
  size_t f = 0;
  f += 2 * (size_t)kFold * kBinGroups * kGroupPitch;   // cos, sin
  f += 2 * (size_t)kFold * kTileFrames;                // E | O
  f += (size_t)((p.crop + 3) & ~3);                    // y
  f += kFold + 2 * kMaxMels;                           // wh, mw0, mw1
  return f * sizeof(float) + kMaxMels * sizeof(int);
}

int melspec_launch(const float* eeg, long long batch, int n_ch, long long length,
                   long long rec_stride, const eegspec_mel_params* p, float* out, void* workspace,
                   size_t workspace_bytes, cudaStream_t st) {
  if (rec_stride < (long long)n_ch * length) return EEGSPEC_EINVAL;
  return melspec_launch_strided(eeg, batch, n_ch, length, rec_stride, p && p->time_major ? 1 : length, p, out,
                                workspace, workspace_bytes, st);
}

// Sliding windows over one long record (BASELINE configs[3]): window w is a "record" that starts
// w * stride samples into the series, so consecutive records overlap and nothing is materialised --
// channel-major the electrode rows stay `total_length` apart, time-major a sample row is n_ch wide.
int melspec_sliding_launch(const float* eeg, int n_ch, long long total_length, long long stride, long long n_windows,
                           const eegspec_mel_params* p, float* out, void* workspace, size_t workspace_bytes,
                           cudaStream_t st) {
  if (!p) return EEGSPEC_EINVAL;
  if (stride <= 0 || n_windows <= 0 || total_length <= 0) return EEGSPEC_EINVAL;
  if (total_length < p->crop || (n_windows - 1) * stride + p->crop > total_length) return EEGSPEC_ESHORT;
  return melspec_launch_strided(eeg, n_windows, n_ch, p->crop, p->time_major ? stride * n_ch : stride,
                                p->time_major ? 1 : total_length, p, out, workspace, workspace_bytes, st);
}

// `length` samples of every record are addressed as eeg[rec * rec_stride + ch * ch_stride + t * t_stride]
// with t_stride = n_ch for time-major records and 1 otherwise; records may overlap.
int melspec_launch_strided(const float* eeg, long long batch, int n_ch, long long length, long long rec_stride,
                           long long ch_stride, const eegspec_mel_params* p, float* out, void* workspace,
                           size_t workspace_bytes, cudaStream_t st) {
  int rc = validate(p);
  if (rc != EEGSPEC_OK) return rc;
  if (!eeg || !out || batch <= 0 || n_ch <= 0) return EEGSPEC_EINVAL;
  if (length < p->crop) return EEGSPEC_ESHORT;
  if (rec_stride <= 0 || ch_stride <= 0) return EEGSPEC_EINVAL;
  if (reinterpret_cast<uintptr_t>(out) & 15) return EEGSPEC_EALIGN;
  const int n_pairs = p->n_chains * p->pairs_per_chain;
  for (int i = 0; i < n_pairs; ++i)
    if (p->pair_a[i] < 0 || p->pair_a[i] >= n_ch || p->pair_b[i] < 0 || p->pair_b[i] >= n_ch)
      return EEGSPEC_EINVAL;

  MelArgs a;
  MelGeometry g;
  bool chain_ok = false;
  if ((rc = device_tables(*p, &a.t, &g, &chain_ok)) != EEGSPEC_OK) return rc;
  a.rec_stride = rec_stride;
  a.ch_stride = ch_stride;
  a.t_stride = p->time_major ? n_ch : 1;
  a.length = length;
  a.crop = p->crop;
  a.crop_off = (int)((length - p->crop) / 2);   // cfg_utils.py:369-371
  a.hop = g.hop;
  a.n_frames = g.n_frames;
  a.n_time = g.n_time;
  a.n_mels = p->n_mels;
  a.n_pairs = n_pairs;
  a.n_chains = p->n_chains;
  a.ppc = p->pairs_per_chain;
  a.pad_mode = p->pad_mode;
  a.amin = p->amin;
  a.top_db = p->top_db;
  a.db_offset = p->db_offset;
  a.inv_db_scale = 1.0f / p->db_scale;
  for (int i = 0; i < EEGSPEC_MAX_PAIRS; ++i) {
    a.pair_a[i] = i < n_pairs ? p->pair_a[i] : 0;
    a.pair_b[i] = i < n_pairs ? p->pair_b[i] : 0;
  }

  // Time-major records: gather the electrode rows into the workspace, then engine 3 on those.
  if ((p->engine == 0 || p->engine == 3) && p->time_major && chain_ok && !(p->crop & 3) && workspace &&
      !(reinterpret_cast<uintptr_t>(workspace) & 15)) {
    GatherRows rows;
    MelArgs g3 = a;
    rows.n = compact_rows(*p, rows.r, g3.pair_a, g3.pair_b);
    const int n_rows = rows.n;
    const size_t per_rec = (size_t)n_rows * p->crop * sizeof(float);
    long long pass = (long long)(workspace_bytes / per_rec);
    if (pass > kGatherChunk) pass = kGatherChunk;
    if (pass >= 1 && (size_t)kGatherTile * (n_ch + 1) * sizeof(float) <= 48 * 1024) {
      float* scratch = static_cast<float*>(workspace);
      cudaError_t e;
      eegspec_mel_params p3 = *p;
      p3.time_major = 0;
      g3.q = nullptr;
      g3.refdb = nullptr;
      g3.rec_stride = (long long)n_rows * p->crop;
      g3.ch_stride = p->crop;
      g3.t_stride = 1;
      g3.length = p->crop;
      g3.crop_off = 0;
      int rc3 = EEGSPEC_OK;
      for (long long r0 = 0; r0 < batch && rc3 == EEGSPEC_OK; r0 += pass) {
        const long long nb = batch - r0 < pass ? batch - r0 : pass;
        gather_rows_kernel<<<dim3((unsigned)((p->crop + kGatherTile - 1) / kGatherTile), (unsigned)nb), kGatherTile,
                             (size_t)kGatherTile * (n_ch + 1) * sizeof(float), st>>>(
            eeg + r0 * rec_stride, rec_stride, n_ch, a.crop_off, p->crop, rows, scratch);
        count_launch();
        if ((e = cudaGetLastError()) != cudaSuccess) return (int)e;
        g3.eeg = scratch;
        g3.out = out + r0 * (long long)p->n_mels * g.n_time * p->n_chains;
        rc3 = melspec_chain_launch(g3, p3, nb, st);
        if (rc3 == EEGSPEC_EUNSUPPORTED && r0 == 0) break;   // geometry not served: the scratch engines below
      }
      if (rc3 != EEGSPEC_EUNSUPPORTED) return rc3;
    }
  }

  // Engine 3 (default when the geometry is the north-star one): one persistent launch, no scratch.
  if (p->engine == 0 || p->engine == 3) {
    a.eeg = eeg;
    a.out = out;
    a.q = nullptr;
    a.refdb = nullptr;
    const int rc3 = chain_ok ? melspec_chain_launch(a, *p, batch, st) : EEGSPEC_EUNSUPPORTED;
    if (rc3 != EEGSPEC_EUNSUPPORTED || p->engine == 3) return rc3;
  }

  const long long chunk = batch < kMelChunk ? batch : kMelChunk;
  const size_t need = per_record_scratch(*p, g) * (size_t)chunk;
  if (!workspace || workspace_bytes < need) return EEGSPEC_EWORKSPACE;
  if (reinterpret_cast<uintptr_t>(workspace) & 15) return EEGSPEC_EALIGN;

  static std::atomic<bool> configured[64];   // idempotent set-up; atomic so that concurrent callers do not race on the flag
  int dev = 0;
  cudaGetDevice(&dev);
  const size_t smem = pair_kernel_smem(*p);
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute(melspec_pair_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    if (e != cudaSuccess) return (int)e;
    configured[dev] = true;
  }
  if (smem > 220 * 1024) return EEGSPEC_EUNSUPPORTED;

  const size_t q_bytes = (size_t)chunk * n_pairs * p->n_mels * g.n_time * sizeof(float);
  for (long long r0 = 0; r0 < batch; r0 += chunk) {
    const long long nb = batch - r0 < chunk ? batch - r0 : chunk;
    a.eeg = eeg + r0 * rec_stride;
    a.q = static_cast<float*>(workspace);
    a.refdb = reinterpret_cast<float*>(static_cast<char*>(workspace) + q_bytes);
    a.out = out + r0 * (long long)p->n_mels * g.n_time * p->n_chains;
    cudaError_t e;
    if (p->engine == 1) {
      melspec_pair_kernel<<<dim3(n_pairs, (unsigned)nb), kMelThreads, smem, st>>>(a);
      count_launch();
      e = cudaGetLastError();
    } else {   // tcgen05 engine with scratch
      const int rc2 = melspec_tc_launch_pair(a, n_pairs, nb, st);
      if (rc2 < 0) return rc2;
      e = static_cast<cudaError_t>(rc2);
    }
    if (e != cudaSuccess) return (int)e;
    const long long n_pix = nb * p->n_mels * g.n_time;
    melspec_combine_kernel<<<(unsigned)((n_pix + 255) / 256), 256, 0, st>>>(a, n_pix);
    count_launch();
    e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
  }
  return EEGSPEC_OK;
}

}  // namespace eegspec
