// Declarations shared by the two mode-N engines (melspec.cu: fp32 SIMT, melspec_tc.cu: tcgen05).
#pragma once
#include "common.cuh"
#include "internal.h"

namespace eegspec {

constexpr int kFold = 64;       // win_length / 2
constexpr int kBinGroups = 8;   // one warp per bin group
constexpr int kBinsPerGroup = 13;
constexpr int kBinSlots = kBinGroups * kBinsPerGroup;  // 104 >= bins with mel weight
constexpr int kGroupPitch = 16;                        // table row pitch per group (float4 x 4)
constexpr int kTileFrames = 128;
constexpr int kMelThreads = 256;
constexpr int kMaxMels = 256;
constexpr int kMaxCrop = 12288;

struct MelTables {            // device-resident constant tables for one parameter set
  float* cos_t;               // [kFold][kBinGroups][kGroupPitch]
  float* sin_t;
  float* win_half;            // [kFold]  w[W/2 + m]
  int* mel_k0;                // [n_mels] first tap, as a bin slot (bin - k_lo)
  float* mel_w0;              // [n_mels]
  float* mel_w1;              // [n_mels]
  unsigned short* tc_tab;     // engine 2: fp16 {cos hi, cos lo, sin hi, sin lo} in UMMA layout
  unsigned short* tc3_tab;    // engine 3: the same with the Hann window folded in (row 0 halved)
};

struct MelArgs {
  const float* eeg;
  float* q;       // [batch][pairs][n_mels][n_time]  10 log10(max(amin, mel))
  float* refdb;   // [batch][pairs]                  10 log10(max(amin, max mel))
  float* out;     // [batch][n_mels][n_time][n_chains]
  MelTables t;
  long long rec_stride, ch_stride, t_stride, length;
  int crop, crop_off, hop, n_frames, n_time, n_mels, n_pairs, n_chains, ppc, pad_mode;
  float amin, top_db, db_offset, inv_db_scale;
  int pair_a[EEGSPEC_MAX_PAIRS], pair_b[EEGSPEC_MAX_PAIRS];
};

__device__ __forceinline__ float padded_sample(const float* y, int i, int n, int pad_mode) {
  if (i < 0) {
    if (pad_mode == 0) return 0.f;
    i = -i;
  } else if (i >= n) {
    if (pad_mode == 0) return 0.f;
    i = 2 * (n - 1) - i;
  }
  return (i >= 0 && i < n) ? y[i] : 0.f;
}


// Engine 2 (melspec_tc.cu): launches the tcgen05 pair kernel for one scratch chunk.
size_t melspec_tc_smem_bytes(int crop);
int melspec_tc_launch_pair(const MelArgs& a, int n_pairs, long long n_records, cudaStream_t st);
constexpr int kTcTableHalves = 4 * 112 * kFold;
void melspec_tc_build_tables(const double* cos_d, const double* sin_d, int n_bins, unsigned short* out);

// Engine 3 (melspec_ts.cu): persistent clustered kernel, no scratch.
bool melspec_chain_bank_matches(const eegspec_mel_params& p, int k_lo, const int* k0, const float* w0,
                                const float* w1);
int melspec_chain_launch(const MelArgs& m, const eegspec_mel_params& p, long long batch, cudaStream_t st);

}  // namespace eegspec
