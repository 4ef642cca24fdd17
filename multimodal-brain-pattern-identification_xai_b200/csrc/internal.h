// Internal C++ interface between the kernel translation units and the C ABI (abi.cu).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "../../include/eegspec.h"

namespace eegspec {

void count_launch();

// stft_image.cu
int stft_plan_make(long long length, int nperseg, int noverlap, long long out_cols,
                   eegspec_stft_plan* plan);
long long interp_bracket(long long j, long long n_frames, long long out_cols);
size_t stft_image_workspace(long long rows, long long length, int nperseg, int noverlap,
                            long long out_cols);
int stft_image_launch(const float* x, long long rows, long long length, long long row_stride,
                      int nperseg, int noverlap, long long out_cols, float eps, float* out,
                      void* workspace, size_t workspace_bytes, cudaStream_t st);
int stft_knots_launch(const float* x, long long rows, long long length, long long row_stride, int nperseg,
                      int noverlap, long long out_cols, float eps, float* knots, cudaStream_t st);
int stft_logmag_launch(const float* x, long long rows, long long length, long long row_stride,
                       int nperseg, int noverlap, float* out, cudaStream_t st);

// melspec.cu
struct MelGeometry {
  int hop, n_frames, n_time, n_pairs;
  int k_lo, k_hi;  // inclusive range of rFFT bins with a non-zero mel weight
};
int melspec_geometry(const eegspec_mel_params* p, MelGeometry* g);
size_t melspec_workspace(long long batch, const eegspec_mel_params* p);
size_t melspec_workspace_default(long long batch, const eegspec_mel_params* p);
int melspec_launch(const float* eeg, long long batch, int n_ch, long long length,
                   long long rec_stride, const eegspec_mel_params* p, float* out, void* workspace,
                   size_t workspace_bytes, cudaStream_t st);
int melspec_launch_strided(const float* eeg, long long batch, int n_ch, long long length, long long rec_stride,
                           long long ch_stride, const eegspec_mel_params* p, float* out, void* workspace,
                           size_t workspace_bytes, cudaStream_t st);
int melspec_sliding_launch(const float* eeg, int n_ch, long long total_length, long long stride, long long n_windows,
                           const eegspec_mel_params* p, float* out, void* workspace, size_t workspace_bytes,
                           cudaStream_t st);

// spec_condition.cu
int spec_condition_launch(const float* spec, long long batch, const eegspec_cond_params* p, const float* weight,
                          const float* bias, const int* partner, float* out, cudaStream_t st);

// classifier_eeg.cu / classifier_spec.cu
void cls_eeg_params_default(eegspec_cls_eeg_params* p);
size_t cls_eeg_workspace(long long batch, long long length, int n_ch);
int cls_eeg_launch(const float* x, long long batch, long long length, int n_ch, long long rec_stride,
                   const eegspec_cls_eeg_params* p, float* out, int* status, void* workspace, size_t workspace_bytes,
                   cudaStream_t st);
void cls_spec_params_default(eegspec_cls_spec_params* p);
size_t cls_spec_workspace(long long batch, const eegspec_cls_spec_params* p);
int cls_spec_launch(const float* win, long long batch, const eegspec_cls_spec_params* p, float* out, int* status,
                    void* workspace, size_t workspace_bytes, cudaStream_t st);
int cls_spec_static_launch(const float* win, long long batch, const eegspec_cls_spec_params* p, float* out, int* status,
                           cudaStream_t st);

int spec_condition_knots_launch(const float* knots, int n_frames, long long batch, const eegspec_cond_params* p,
                                const float* weight, const float* bias, const int* partner, float* out, cudaStream_t st);

// eegtransform.cu
void xform_params_default(eegspec_xform_params* p);
long long xform_out_length(long long length, const eegspec_xform_params* p);
int eeg_transform_launch(const float* x, long long batch, long long length, int n_ch, long long rec_stride,
                         const eegspec_xform_params* p, float* out, cudaStream_t st);

}  // namespace eegspec
