/*
 * eegspec.h -- C ABI of the B200-native EEG -> spectrogram transform (libeegspec.so).
 *
 * Plain C: pointers, sizes, POD structs; no torch / C++ types cross this boundary.
 * Every entry point returns an int status: 0 = ok, negative = argument error (EEGSPEC_E*),
 * positive = a cudaError_t raised by the launch.  Nothing throws across the ABI.  All device
 * entry points are asynchronous on the CUDA stream that is passed in (a cudaStream_t cast to
 * void*; NULL = the legacy default stream) and never synchronise.  The library is stateless
 * apart from per-device constant tables for the mel transform that are built once, under a lock.
 *
 * The reference (KC-decoder/Multimodal-Brain-Pattern-Identification_XAI) is pure Python and
 * defines no FFI of its own; each entry point below names the reference code it replaces.
 */
#ifndef EEGSPEC_H_
#define EEGSPEC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EEGSPEC_ABI_VERSION 1

enum {
  EEGSPEC_OK = 0,
  EEGSPEC_EINVAL = -1,        /* null pointer, non-positive size, bad stride                  */
  EEGSPEC_ENOVERLAP = -2,     /* noverlap >= nperseg (SciPy: "noverlap must be less than ...") */
  EEGSPEC_EUNSUPPORTED = -3,  /* legal for the reference, not implemented here (no fallback)   */
  EEGSPEC_ESHORT = -4,        /* record shorter than one window / the crop length              */
  EEGSPEC_EWORKSPACE = -5,    /* workspace missing or too small                                */
  EEGSPEC_EALIGN = -6,        /* output pointer / row pitch not 16-byte aligned                */
  EEGSPEC_ESTATICLIMIT = -7   /* request exceeds the built-in scratch of a *_static_* entry    */
};

int eegspec_abi_version(void);
/* Static string for a status returned by any function below (cudaGetErrorString for > 0). */
const char *eegspec_strerror(int status);

/* ------------------------------------------------------------------------------------------
 * Mode R: the reference's compute_stft(signal, config)
 *   /root/reference/root/src/utils/DiffEEG_utils.py:27-76  (called from
 *   /root/reference/root/src/training/DiffEEG_train.py:131,209)
 * A "row" is one (record, channel) series; rows = B*C.
 * ------------------------------------------------------------------------------------------ */

/* Integer framing facts of scipy.signal.stft(boundary='zeros', padded=True) as used at
 * DiffEEG_utils.py:46-51 -- what the reference gets implicitly from SciPy.  Host only. */
typedef struct eegspec_stft_plan {
  int32_t hop;            /* nperseg - noverlap                                         */
  int32_t half;           /* nperseg / 2 zeros prepended and appended                   */
  int32_t nadd;           /* extra zeros appended to reach a whole number of hops       */
  int32_t n_freqs;        /* nperseg / 2 + 1                                            */
  int64_t n_frames;       /* STFT columns                                               */
  int64_t n_slots;        /* frames the kernel actually evaluates (<= n_frames)         */
  int32_t gather;         /* 1: only the frames bracketing an output column are evaluated */
  int32_t reserved;
} eegspec_stft_plan;

int eegspec_stft_plan_make(int64_t length, int nperseg, int noverlap, int64_t out_cols,
                           eegspec_stft_plan *plan);

/* Left knot index np.interp uses for output column j (DiffEEG_utils.py:57-61). Host only. */
int64_t eegspec_interp_bracket(int64_t j, int64_t n_frames, int64_t out_cols);

/* Scratch bytes eegspec_stft_image_f32 needs for this shape (0 when everything fits on chip). */
size_t eegspec_stft_image_workspace(int64_t rows, int64_t length, int nperseg, int noverlap,
                                    int64_t out_cols);

/* x: device, rows series of `length` fp32, consecutive rows `row_stride` elements apart.
 * out: device, (rows, nperseg/2+1, out_cols) fp32 contiguous, 16-byte aligned.
 * Replaces DiffEEG_utils.py:43-74 for all rows at once: STFT -> log1p|Z| -> linear
 * resample to out_cols (2000 in the reference) -> per-(row,freq) min-max with eps (1e-8). */
int eegspec_stft_image_f32(const float *x, int64_t rows, int64_t length, int64_t row_stride,
                           int nperseg, int noverlap, int64_t out_cols, float eps, float *out,
                           void *workspace, size_t workspace_bytes, void *stream);

/* The plotting twin compute_stft_single (DiffEEG_utils.py:504-509): log1p|Z| only.
 * out: device, (rows, nperseg/2+1, n_frames) fp32 contiguous. */
int eegspec_stft_logmag_f32(const float *x, int64_t rows, int64_t length, int64_t row_stride,
                            int nperseg, int noverlap, float *out, void *stream);

/* ------------------------------------------------------------------------------------------
 * Mode N: montage differences -> NaN fill -> Hann/STFT -> mel -> dB -> (x+40)/40 -> chain mean
 *   front half replaces /root/reference/root/config/config.yml:529-545 (pair list),
 *   /root/reference/root/src/utils/data_utils.py:183-215 (differencing),
 *   /root/reference/root/src/utils/cfg_utils.py:369-371 (crop), :380-385 (NaN fill);
 *   back half (mel / dB) is specified by BASELINE.json only -- see DESIGN.md.
 * ------------------------------------------------------------------------------------------ */
#define EEGSPEC_MAX_PAIRS 32

typedef struct eegspec_mel_params {
  int32_t sr;            /* 200                                                          */
  int32_t n_fft;         /* 1024                                                         */
  int32_t win_length;    /* 128                                                          */
  int32_t hop;           /* 0 => crop / n_time_div (10000 // 256 = 39)                   */
  int32_t n_mels;        /* 128                                                          */
  float fmin;            /* 0                                                            */
  float fmax;            /* 20                                                           */
  int32_t crop;          /* 10000: middle window, offset (L - crop)//2                   */
  int32_t n_time_div;    /* 256: hop = crop // n_time_div; width = (n_frames//32)*32     */
  float amin;            /* 1e-10                                                        */
  float top_db;          /* 80                                                           */
  float db_offset;       /* 40                                                           */
  float db_scale;        /* 40                                                           */
  int32_t pad_mode;      /* 0 = constant zeros (librosa >= 0.10), 1 = reflect (< 0.10)   */
  int32_t n_chains;      /* 4                                                            */
  int32_t pairs_per_chain; /* 4                                                          */
  int32_t pair_a[EEGSPEC_MAX_PAIRS]; /* channel index of the minuend, chain-major        */
  int32_t pair_b[EEGSPEC_MAX_PAIRS]; /* channel index of the subtrahend                  */
  int32_t time_major;    /* 0: record is (n_ch, L) channel-major; 1: (L, n_ch)           */
  int32_t engine;        /* 0 = default (3 when the geometry allows, else 2), 1 = fp32 SIMT DFT,
                          * 2 = tcgen05 DFT + scratch, 3 = persistent clustered tcgen05, no scratch */
} eegspec_mel_params;

/* Fills the defaults above (HMS montage LL, LP, RP, RR in EEG_COLUMNS order). */
void eegspec_mel_params_default(eegspec_mel_params *p);

/* Output geometry for these parameters: frames per pair and kept time columns. */
int eegspec_melspec_geometry(const eegspec_mel_params *p, int32_t *n_frames, int32_t *n_time);

/* Scratch bytes that serve EVERY engine for `batch` records (the generic engines 1 / 2 keep
 * 16 x n_mels x n_time fp32 per record for up to 256 records at a time). */
size_t eegspec_melspec_workspace(int64_t batch, const eegspec_mel_params *p);

/* Scratch bytes the DEFAULT engine needs when it serves the geometry: 0 for channel-major records,
 * the gathered electrode rows of one pass for time-major ones.  A call that turns out to need more
 * (a geometry only the generic engines serve) returns EEGSPEC_EWORKSPACE and changes nothing; the
 * caller then retries with eegspec_melspec_workspace() bytes. */
size_t eegspec_melspec_workspace_default(int64_t batch, const eegspec_mel_params *p);

/* eeg: device fp32, `batch` records `rec_stride` elements apart, each n_ch x length
 * (or length x n_ch when time_major).  out: device (batch, n_mels, n_time, n_chains) fp32. */
int eegspec_melspec_f32(const float *eeg, int64_t batch, int32_t n_ch, int64_t length,
                        int64_t rec_stride, const eegspec_mel_params *p, float *out,
                        void *workspace, size_t workspace_bytes, void *stream);

/* Sliding windows over ONE long record (BASELINE.json configs[3]: a 1-hour EEG cut into overlapping
 * 50-s windows).  Window w is samples [w*stride, w*stride + p->crop) of the series -- the window
 * /root/reference/root/src/utils/cfg_utils.py:369-371 cuts around one label offset, taken at every
 * stride.  eeg: device fp32, (n_ch, total_length) channel-major, or (total_length, n_ch) when
 * p->time_major.  The windows are read in place (nothing is materialised: 57.6 MB in for a 1-hour
 * record instead of 1.4 GB of copies).  out: device (n_windows, n_mels, n_time, n_chains) fp32.
 * Needs (n_windows - 1) * stride + p->crop <= total_length; workspace as eegspec_melspec_f32 with
 * batch = n_windows. */
int eegspec_melspec_sliding_f32(const float *eeg, int32_t n_ch, int64_t total_length, int64_t stride,
                                int64_t n_windows, const eegspec_mel_params *p, float *out,
                                void *workspace, size_t workspace_bytes, void *stream);

/* ------------------------------------------------------------------------------------------
 * Host-buffer front doors (what the reference's caller has: CPU tensors from a DataLoader).
 * x / out are HOST pointers (pinned for full PCIe speed; pageable works, slower).  The call
 * streams record chunks through device buffers on `device` (H2D, kernels, D2H overlapped on
 * internal streams) and returns after the last byte of `out` has landed.
 * ------------------------------------------------------------------------------------------ */
int eegspec_stft_image_host_f32(const float *x, int64_t rows, int64_t length, int nperseg,
                                int noverlap, int64_t out_cols, float eps, float *out,
                                int device);

int eegspec_melspec_host_f32(const float *eeg, int64_t batch, int32_t n_ch, int64_t length,
                             const eegspec_mel_params *p, float *out, int device);

/* Frees what the host-buffer front doors keep per device between calls (three streams and their staging
 * buffers; they only grow otherwise).  Waits for the device's pipelines to drain first.  Safe to call at any
 * time from any thread; the next host-buffer call on that device sets the pipeline up again. */
int eegspec_release(int device);

/* Bytes eegspec_melspec_host_f32 moves over PCIe for these arguments.  Channel-major records are
 * copied row-wise: only the electrode rows some pair of the montage reads travel (16 of the 20
 * EEG_COLUMNS rows for the HMS montage), the others are never touched on either side. */
int eegspec_melspec_host_bytes(int64_t batch, int32_t n_ch, int64_t length, const eegspec_mel_params *p,
                               int64_t *h2d_bytes, int64_t *d2h_bytes);

/* ------------------------------------------------------------------------------------------
 * Row (f)1 -- the step in front of the path: raw-EEG transformer
 *   replaces _EEGTransformer.transform, /root/reference/root/src/data/dataset.py:73-104
 *   (selection :88-90 or "magic 8" differences :106-123, clip +-1024 / NaN -> 0 / divide by 32
 *   :92-94, Butterworth low-pass as a causal IIR along time :125-131, decimation :101-102).
 * ------------------------------------------------------------------------------------------ */
#define EEGSPEC_MAX_XF_CH 32
#define EEGSPEC_XF_ORDER 4

typedef struct eegspec_xform_params {
  int32_t n_out;                       /* output channels (19, or 8 for the magic-8 montage)   */
  int32_t sel_a[EEGSPEC_MAX_XF_CH];    /* output channel c = x[:, sel_a[c]] - x[:, sel_b[c]],   */
  int32_t sel_b[EEGSPEC_MAX_XF_CH];    /*   sel_b[c] < 0: plain selection x[:, sel_a[c]]        */
  int32_t normalize;                   /* clip to +-clip_abs, NaN -> 0, multiply by scale       */
  float clip_abs;                      /* 1024                                                  */
  float scale;                         /* 1 / 32                                                */
  int32_t filter;                      /* 1: apply the IIR b / a below (a[0] == 1), 0: skip     */
  double b[EEGSPEC_XF_ORDER + 1];      /* scipy.signal.butter(4, 0.2) numerator                 */
  double a[EEGSPEC_XF_ORDER + 1];      /* denominator                                           */
  int32_t downsample;                  /* keep every downsample-th sample (1 = all), <= 8       */
  int32_t out_channel_major;           /* 0: out is (batch, L_out, n_out) like the reference;
                                        * 1: (batch, n_out, L_out), the layout compute_stft eats */
  int32_t mu_law;                      /* 1: sign(y) log(1 + mu |y|) / log(1 + mu) after the filter
                                        *    (dataset.py:98-99, 133-147)                          */
  float mu;                            /* 1 (the reference calls _quantize_data(x, 1))            */
} eegspec_xform_params;

/* Defaults of CFG.dataset["eeg"] (cfg_utils.py:97-106): 19 selected channels, normalize,
 * 4th-order 20 Hz Butterworth at 200 Hz (coefficients in double precision), downsample 5. */
void eegspec_xform_params_default(eegspec_xform_params *p);

/* Output samples per record: ceil(length / downsample). */
int64_t eegspec_xform_out_length(int64_t length, const eegspec_xform_params *p);

/* x: device fp32, `batch` records of (length, n_ch) TIME-major (the reference's (L, C) arrays),
 * consecutive records `rec_stride` elements apart, 16-byte aligned, rec_stride % 4 == 0, and either
 * n_ch % 4 == 0 or length % 4 == 0 (the reference's (10000, 19) windows qualify; else EEGSPEC_EALIGN).
 * out: device fp32 (batch, L_out, n_out) or (batch, n_out, L_out), see out_channel_major. */
int eegspec_eeg_transform_f32(const float *x, int64_t batch, int64_t length, int32_t n_ch,
                              int64_t rec_stride, const eegspec_xform_params *p, float *out,
                              void *stream);

/* Row (f)1 fused with the path: raw (batch, length, n_ch) time-major EEG -> the compute_stft image of
 * the transformed signal, (batch, n_out, nperseg/2+1, out_cols) fp32 -- what
 * compute_stft(EEGDataset[i]["x"]) produces in /root/reference/root/src/training/DiffEEG_train.py:128-131
 * from the (C, 2000) rows dataset.py:213-228 hands out.  Two launches per pass of up to ~2 000 records; the
 * transformed rows (n_out x L_out fp32 per record, 152 KB = 3 % of the image) live in the caller's
 * workspace between them (eegspec_eeg_transform_stft_workspace() bytes). */
size_t eegspec_eeg_transform_stft_workspace(int64_t batch, int64_t length, const eegspec_xform_params *p,
                                            int nperseg, int noverlap, int64_t out_cols);
int eegspec_eeg_transform_stft_f32(const float *x, int64_t batch, int64_t length, int32_t n_ch,
                                   int64_t rec_stride, const eegspec_xform_params *p, int nperseg,
                                   int noverlap, int64_t out_cols, float eps, float *out,
                                   void *workspace, size_t workspace_bytes, void *stream);

/* ------------------------------------------------------------------------------------------
 * Row (f)2 -- the consumer of the path: spectrogram-condition front end of the DiffEEG denoiser
 *   replaces /root/reference/root/src/models/diffEEG.py:47-53 (ConvTranspose2d(C -> H2, kernel (3, 3),
 *   stride (1, 8), padding (1, 2))), :108-110 (relu -> flatten(start_dim=2) -> F.interpolate(size=T,
 *   mode="linear", align_corners=False)) and :136-145 (within-class mixing, training mode) with the
 *   random permutation made explicit: partner[b] is the record b is mixed with.
 * The 527 703-sample intermediate per output channel is never formed: each output sample gathers the
 * two convolution outputs the interpolation keeps.
 * ------------------------------------------------------------------------------------------ */
typedef struct eegspec_cond_params {
  int32_t n_channels;    /* C: input channels of the image (19 or 8)                      */
  int32_t n_freqs;       /* F: 33                                                         */
  int32_t n_cols;        /* image columns, 2000                                           */
  int32_t out_channels;  /* H2 = hidden_channels / 2 (16 in the shipped config)           */
  int32_t out_length;    /* T = x.shape[-1] of the denoiser input, 2000                   */
  float mix_alpha;       /* 0.5 (diffEEG.py:142); only read when partner != NULL          */
} eegspec_cond_params;

/* spec: device (batch, C, F, n_cols) fp32; weight: device (C, H2, 3, 3), the ConvTranspose2d layout;
 * bias: device (H2); partner: device int32 (batch) or NULL; out: device (batch, H2, T) fp32. */
int eegspec_spec_condition_f32(const float *spec, int64_t batch, const eegspec_cond_params *p,
                               const float *weight, const float *bias, const int32_t *partner,
                               float *out, void *stream);

/* The same operator WITHOUT the image: compute_stft's kernel stops after its normalisation and hands over the
 * F x n_frames knot pairs {value, delta} every pixel of the image is a linear function of
 * (pixel = sat(value + delta * fraction), DiffEEG_utils.py:57-61, 72-74); the condition kernel evaluates the
 * 85 500 pixels it needs from them.  Per record of the DiffEEG batch: 152 KB of signal in, 321 KB of knots through
 * L2, 128 KB out -- instead of writing a 5 MB image and gathering 2.7 MB of it back.  Bit-identical to
 * eegspec_spec_condition_f32(eegspec_stft_image_f32(x)).
 *   knots: device, (rows, nperseg/2+1, n_frames) pairs of fp32, 8-byte aligned, n_frames from eegspec_stft_plan_make.
 *   EEGSPEC_EUNSUPPORTED when a row's knots do not fit shared memory or the plan decimates (very long rows). */
int eegspec_stft_knots_f32(const float *x, int64_t rows, int64_t length, int64_t row_stride, int nperseg,
                           int noverlap, int64_t out_cols, float eps, float *knots, void *stream);
int eegspec_spec_condition_knots_f32(const float *knots, int32_t n_frames, int64_t batch,
                                     const eegspec_cond_params *p, const float *weight, const float *bias,
                                     const int32_t *partner, float *out, void *stream);

/* ------------------------------------------------------------------------------------------
 * Rows (f)4 / (f)3 -- the classifier-side preprocessing of the multimodal model's two inputs
 *   (CombinedDataset, /root/reference/root/jupyter_notebooks/XAI_Multimodality.py:1114-1307; the helpers
 *   are the same code as /root/reference/root/src/utils/data_utils.py:129-249).
 * ------------------------------------------------------------------------------------------ */
#define EEGSPEC_CLS_MAX_TAPS 13

/* process_eeg (:1148-1164): band-pass -> NaN fill -> 18 montage differences -> band-pass + 4-tap mean +
 * decimate 4 -> per-row z-score -> rows 0..18 and 19..36 of the 38-row array -> pad / truncate to 3000. */
typedef struct eegspec_cls_eeg_params {
  int32_t n_taps1;                         /* 11: butter(5, [0.5, 20] Hz @ 200 Hz, 'band')  (:1156, :1253)        */
  double b1[EEGSPEC_CLS_MAX_TAPS], a1[EEGSPEC_CLS_MAX_TAPS];
  int32_t n_taps2;                         /* 13: butter(6, ...) of denoise_filter           (:1268)               */
  double b2[EEGSPEC_CLS_MAX_TAPS], a2[EEGSPEC_CLS_MAX_TAPS];
  int32_t n_pairs;                         /* 18 map_features pairs                           (:211-217)            */
  int32_t pair_a[EEGSPEC_MAX_PAIRS], pair_b[EEGSPEC_MAX_PAIRS];
  int32_t n_features;                      /* 19 = len(feature_to_index): rows kept AND the start index of the
                                            * "differential" rows in select_and_map_channels  (:1278-1282)          */
  int32_t fixed_length;                    /* 3000                                            (:245)                */
  float eps;                               /* 1e-6                                            (:1276)               */
} eegspec_cls_eeg_params;

void eegspec_cls_eeg_params_default(eegspec_cls_eeg_params *p);
size_t eegspec_classifier_eeg_workspace(int64_t batch, int64_t length, int32_t n_ch);

/* x: device fp32 (batch, length, n_ch) time-major -- eeg.values of the parquet frames (n_ch = 20 EEG_COLUMNS),
 * records `rec_stride` elements apart.  out: device (batch, 1, n_features + n_pairs, fixed_length) fp32.
 * status: device int32 (batch) or NULL; set to 1 for a record in which a channel starts with NaN (its filtered
 * row is all NaN: the reference drops the row and shifts every later channel -- refused, result undefined). */
int eegspec_classifier_eeg_f32(const float *x, int64_t batch, int64_t length, int32_t n_ch, int64_t rec_stride,
                               const eegspec_cls_eeg_params *p, float *out, int32_t *status,
                               void *workspace, size_t workspace_bytes, void *stream);

/* process_spectrogram (:1183-1202) on the (rows, cols) = (400, 300) window cut out by :1173-1183. */
typedef struct eegspec_cls_spec_params {
  int32_t rows, cols;                      /* 400, 300 = CFG.image_size                                           */
  double notch_b[3], notch_a[3];           /* iirnotch(60, 30, fs=200)                        (:1294-1297)        */
  double notch_zi[2];                      /* lfilter_zi(notch_b, notch_a): filtfilt's initial state             */
  float gauss_sigma, gauss_truncate;       /* 1.0, 4.0 (scipy.ndimage default) -> 9 taps      (:1299-1301)        */
  float eps;                               /* 1e-6                                            (:1292)             */
} eegspec_cls_spec_params;

void eegspec_cls_spec_params_default(eegspec_cls_spec_params *p);
size_t eegspec_classifier_spec_workspace(int64_t batch, const eegspec_cls_spec_params *p);

/* win: device fp32 (batch, rows, cols), NaNs allowed.  out: device (batch, 3, rows, cols) fp32, the three
 * planes identical (np.tile).  status: device int32 (batch) or NULL; 1 = a row is entirely NaN (the reference
 * drops it and resizes the image with scikit-image -- refused). */
int eegspec_classifier_spec_f32(const float *win, int64_t batch, const eegspec_cls_spec_params *p, float *out,
                                int32_t *status, void *workspace, size_t workspace_bytes, void *stream);

/* No-allocation variant of eegspec_classifier_spec_f32.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  Same inputs, outputs and results as
 * eegspec_classifier_spec_f32, but the scratch between its four kernels is a device array built into the
 * library (EEGSPEC_CLS_SPEC_STATIC_BYTES, held by every GPU that loads the library until exit) instead of a
 * caller workspace, and the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).
 * Limit: eegspec_classifier_spec_workspace(batch, p) <= EEGSPEC_CLS_SPEC_STATIC_BYTES, i.e. at most 17 windows
 * of (400, 300) per call; beyond it the call returns EEGSPEC_ESTATICLIMIT and launches nothing.  Every call
 * shares the one array: calls that overlap on the GPU (different streams, different threads) race -- serialise
 * them.  The array keeps the last call's contents; nothing in it needs to start at zero. */
#define EEGSPEC_CLS_SPEC_STATIC_BYTES (8u << 20)
int eegspec_classifier_spec_static_f32(const float *win, int64_t batch, const eegspec_cls_spec_params *p, float *out,
                                       int32_t *status, void *stream);

/* Number of kernel launches issued by this library since load (bench bookkeeping). */
int64_t eegspec_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* EEGSPEC_H_ */
