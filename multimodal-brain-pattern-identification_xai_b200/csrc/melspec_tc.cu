// Mode N, engine 2: the pruned DFT on the 5th-generation tensor cores (tcgen05 + TMEM).
//
// Same algorithm as melspec.cu (folded 128-sample Hann frame -> bins 1..102 of the 1024-point
// spectrum -> |X|^2 -> 2-tap mel -> dB numerator), but the two real products
//     Re[frame][bin] = sum_m E[frame][m] cos(2 pi bin m / 1024)
//     Im[frame][bin] = sum_m O[frame][m] sin(2 pi bin m / 1024)          (m < 64)
// run as tcgen05.mma kind::f16 with fp32 accumulation in tensor memory.  fp16 alone is far too
// coarse for a -80 dB spectrogram, so both operands are split into a high and a low half
//     E = Eh + El,   cos = Ch + Cl        (each half an fp16; |x - xh - xl| <= 2^-22 |x|)
// and three products are accumulated into the same TMEM tile: Eh.Ch + El.Ch + Eh.Cl.  The dropped
// El.Cl term is 2^-22 relative, so the result carries ~fp32 accuracy (tests: <= 1e-3 on the
// (dB+40)/40 image, measured ~1e-5).  Frames are pre-scaled by an exact power of two so that the
// pair's largest sample lands in [2^10, 2^11): no fp16 overflow for any finite input, and the low
// halves stay well inside the fp16 subnormal range relative to the pair maximum.
//
// One CTA per (record, pair), 256 threads:
//   all threads   difference + NaN fill + max|y| into shared memory
//   per 128-frame tile:
//     all threads fold/window/scale/split -> four 128x64 fp16 operand tiles in the UMMA K-major
//                 no-swizzle core-matrix layout (8 rows x 16 B contiguous), 128-bit stores
//     one thread  24 x tcgen05.mma (M=128, N=112, K=16), tcgen05.commit -> mbarrier
//     all threads tcgen05.ld the 128 x 224 fp32 accumulators (lane = frame), power -> smem
//     all threads sparse mel, 10 log10, store to scratch, running max      (as engine 1)
#include <atomic>
#include <cuda_fp16.h>

#include "melspec_common.cuh"

namespace eegspec {

constexpr int kTcBins = 112;                         // MMA N: bin slots padded to a multiple of 16
constexpr int kTcTileBytes = kTileFrames * kFold * 2;  // one 128x64 fp16 operand tile: 16 KB
constexpr int kTcTabBytes = kTcBins * kFold * 2;       // one 112x64 fp16 table: 14 KB
constexpr int kTcCols = 256;                           // TMEM columns allocated (2 x 112 used)

// Byte offset of element (row, k) in a K-major no-swizzle operand: core matrices of 8 rows x 16 B,
// the eight K-chunks of a row group contiguous (LBO = 128 B), row groups 1 KB apart (SBO).
__host__ __device__ constexpr int umma_offset(int row, int k) {
  return (row >> 3) * 1024 + (k >> 3) * 128 + (row & 7) * 16 + (k & 7) * 2;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);        // start address
  d |= (uint64_t)(128 >> 4) << 16;                   // leading byte offset: next K core matrix
  d |= (uint64_t)(1024 >> 4) << 32;                  // stride byte offset: next 8-row group
  d |= (uint64_t)1 << 46;                            // descriptor version (sm_100)
  return d;                                          // base offset 0, layout SWIZZLE_NONE
}

// kind::f16, A = B = fp16 K-major, D = fp32, M = 128, N = 112.
constexpr uint32_t kIdesc = (1u << 4) | ((uint32_t)(kTcBins >> 3) << 17) | ((uint32_t)(kTileFrames >> 4) << 24);

__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t"
      "}\n"
      :
      : "r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(kIdesc), "r"(accumulate), "r"(0), "r"(0), "r"(0), "r"(0)
      : "memory");
}

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "DONE:\n\t"
      "}\n"
      :
      : "r"(bar), "r"(parity)
      : "memory");
}

// Two 8-column TMEM loads (Re and Im of the same bins) and the wait that makes the registers
// valid.  The registers are threaded through the wait as "+r" operands so that the compiler
// cannot schedule a use of them ahead of tcgen05.wait::ld.
__device__ __forceinline__ void tmem_ld8x2(uint32_t addr_re, uint32_t addr_im, float (&re)[8], float (&im)[8]) {
  uint32_t a[8], b[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
               : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3]), "=r"(a[4]), "=r"(a[5]), "=r"(a[6]), "=r"(a[7])
               : "r"(addr_re));
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
               : "=r"(b[0]), "=r"(b[1]), "=r"(b[2]), "=r"(b[3]), "=r"(b[4]), "=r"(b[5]), "=r"(b[6]), "=r"(b[7])
               : "r"(addr_im));
  asm volatile("tcgen05.wait::ld.sync.aligned;\n"
               : "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(a[4]), "+r"(a[5]), "+r"(a[6]), "+r"(a[7]),
                 "+r"(b[0]), "+r"(b[1]), "+r"(b[2]), "+r"(b[3]), "+r"(b[4]), "+r"(b[5]), "+r"(b[6]), "+r"(b[7])
               :
               : "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) { re[i] = __uint_as_float(a[i]); im[i] = __uint_as_float(b[i]); }
}

// Split x into fp16 high and low halves, packed for two consecutive k.
__device__ __forceinline__ void split2(float x0, float x1, uint32_t& hi, uint32_t& lo) {
  const __half h0 = __float2half_rn(x0), h1 = __float2half_rn(x1);
  const __half l0 = __float2half_rn(x0 - __half2float(h0)), l1 = __float2half_rn(x1 - __half2float(h1));
  hi = (uint32_t)__half_as_ushort(h0) | ((uint32_t)__half_as_ushort(h1) << 16);
  lo = (uint32_t)__half_as_ushort(l0) | ((uint32_t)__half_as_ushort(l1) << 16);
}

__global__ void __launch_bounds__(kMelThreads, 1) melspec_pair_tc_kernel(const MelArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char* tab = smem_raw;                                   // Ch | Cl | Sh | Sl
  unsigned char* tiles = tab + 4 * kTcTabBytes;                    // Eh | El | Oh | Ol
  float* y = reinterpret_cast<float*>(tiles + 4 * kTcTileBytes);   // crop
  float* wh = y + ((a.crop + 3) & ~3);                             // kFold  window half
  float* whs = wh + kFold;                                         // kFold  window half * scale
  float* mw0 = whs + kFold;
  float* mw1 = mw0 + kMaxMels;
  int* mk0 = reinterpret_cast<int*>(mw1 + kMaxMels);
  __shared__ float red_a[kMelThreads / 32], red_b[kMelThreads / 32], red_c[kMelThreads / 32];
  __shared__ float bcast[4];
  __shared__ __align__(8) unsigned long long mbar;
  __shared__ uint32_t tmem_base_s;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int pair = blockIdx.x;
  const long long rec = blockIdx.y;
  const float* ea = a.eeg + rec * a.rec_stride + a.pair_a[pair] * a.ch_stride + a.crop_off * a.t_stride;
  const float* eb = a.eeg + rec * a.rec_stride + a.pair_b[pair] * a.ch_stride + a.crop_off * a.t_stride;

  // TMEM allocation (one warp), mbarrier init (one thread).
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(kTcCols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  if (tid == 32) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&mbar)), "r"(1));
    asm volatile("fence.mbarrier_init.release.cluster;" ::);
  }

  // Constant tables: fp16 hi/lo DFT matrices (already in UMMA layout), window, mel taps.
  {
    const uint4* src = reinterpret_cast<const uint4*>(a.t.tc_tab);
    uint4* dst = reinterpret_cast<uint4*>(tab);
    for (int i = tid; i < 4 * kTcTabBytes / 16; i += kMelThreads) dst[i] = __ldg(src + i);
  }
  for (int i = tid; i < kFold; i += kMelThreads) wh[i] = a.t.win_half[i];
  for (int i = tid; i < a.n_mels; i += kMelThreads) {
    mw0[i] = a.t.mel_w0[i];
    mw1[i] = a.t.mel_w1[i];
    mk0[i] = a.t.mel_k0[i];
  }

  // Montage difference + NaN statistics (cfg_utils.py:380-385 applied to the difference).
  float sum = 0.f, nans = 0.f, amax = 0.f;
  for (int i = tid; i < a.crop; i += kMelThreads) {
    const float d = __ldg(ea + i * a.t_stride) - __ldg(eb + i * a.t_stride);
    y[i] = d;
    if (d != d) nans += 1.f; else { sum += d; amax = fmaxf(amax, fabsf(d)); }
  }
  sum = warp_sum(sum);
  nans = warp_sum(nans);
  amax = warp_max_nan(amax);
  if (lane == 0) { red_a[warp] = sum; red_b[warp] = nans; red_c[warp] = amax; }
  __syncthreads();
  if (warp == 0) {
    float s = lane < kMelThreads / 32 ? red_a[lane] : 0.f;
    float c = lane < kMelThreads / 32 ? red_b[lane] : 0.f;
    float m = lane < kMelThreads / 32 ? red_c[lane] : 0.f;
    s = warp_sum(s);
    c = warp_sum(c);
    m = warp_max_nan(m);
    if (lane == 0) {
      const bool all_nan = c >= (float)a.crop;
      const float fill = (c > 0.f && !all_nan) ? s / ((float)a.crop - c) : 0.f;
      if (all_nan) m = 0.f;
      m = fmaxf(m, fabsf(fill));
      // power-of-two scale: m * 2^e in [2^10, 2^11); exact, undone on the power spectrum
      int ex = 0;
      if (m > 0.f && m < __int_as_float(0x7f800000)) frexpf(m, &ex); else ex = 11;
      int e = 11 - ex;
      e = max(-60, min(60, e));
      bcast[0] = fill;
      bcast[1] = c;
      bcast[2] = __int_as_float((127 + e) << 23);        // 2^e
      bcast[3] = __int_as_float((127 - e) << 23);        // 2^-e
    }
  }
  __syncthreads();
  {
    const float n_nan = bcast[1];
    if (n_nan > 0.f) {
      const bool all_nan = n_nan >= (float)a.crop;
      const float fill = bcast[0];
      for (int i = tid; i < a.crop; i += kMelThreads) {
        const float d = y[i];
        if (all_nan) y[i] = 0.f; else if (d != d) y[i] = fill;
      }
    }
  }
  const float scale = bcast[2], descale = bcast[3];
  for (int i = tid; i < kFold; i += kMelThreads) whs[i] = wh[i] * scale;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // tables -> visible to the MMA
  asm volatile("tcgen05.fence::before_thread_sync;" ::);
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::);
  const uint32_t tmem_base = tmem_base_s;
  const uint32_t bar = smem_u32(&mbar);

  float* Pw = reinterpret_cast<float*>(tiles);   // [kTcBins][kTileFrames] fp32, aliases the operands
  float vmax = 0.f;
  float* qbase = a.q + ((rec * a.n_pairs + pair) * (long long)a.n_mels) * a.n_time;
  const int n_tiles = (a.n_time + kTileFrames - 1) / kTileFrames;

  for (int tl = 0; tl < n_tiles; ++tl) {
    const int f0 = tl * kTileFrames;
    // ---- operand build: item = (frame t, octet o of the fold index), 8 fp16 = one 16-byte chunk
    for (int it = tid; it < kTileFrames * (kFold / 8); it += kMelThreads) {
      const int t = it & (kTileFrames - 1), o = it / kTileFrames;
      const int f = f0 + t;
      float ev[8], ov[8];
      if (f < a.n_frames) {
        const int c = f * a.hop;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int m = 8 * o + j;
          const float yp = padded_sample(y, c + m, a.crop, a.pad_mode);
          const float ym = m ? padded_sample(y, c - m, a.crop, a.pad_mode) : 0.f;
          ev[j] = whs[m] * (yp + ym);
          ov[j] = m ? whs[m] * (yp - ym) : 0.f;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) { ev[j] = 0.f; ov[j] = 0.f; }
      }
      uint4 eh, el, oh, ol;
      split2(ev[0], ev[1], eh.x, el.x); split2(ev[2], ev[3], eh.y, el.y);
      split2(ev[4], ev[5], eh.z, el.z); split2(ev[6], ev[7], eh.w, el.w);
      split2(ov[0], ov[1], oh.x, ol.x); split2(ov[2], ov[3], oh.y, ol.y);
      split2(ov[4], ov[5], oh.z, ol.z); split2(ov[6], ov[7], oh.w, ol.w);
      const int off = umma_offset(t, 8 * o);
      *reinterpret_cast<uint4*>(tiles + 0 * kTcTileBytes + off) = eh;
      *reinterpret_cast<uint4*>(tiles + 1 * kTcTileBytes + off) = el;
      *reinterpret_cast<uint4*>(tiles + 2 * kTcTileBytes + off) = oh;
      *reinterpret_cast<uint4*>(tiles + 3 * kTcTileBytes + off) = ol;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::);
    __syncthreads();

    // ---- one thread issues the 24 MMAs of this tile
    if (tid == 0) {
      asm volatile("tcgen05.fence::after_thread_sync;" ::);
      const uint32_t tab0 = smem_u32(tab), til0 = smem_u32(tiles);
#pragma unroll
      for (int part = 0; part < 2; ++part) {            // 0: Re = E.cos, 1: Im = O.sin
        const uint32_t d = tmem_base + part * kTcBins;
        const uint32_t a_hi = til0 + (2 * part) * kTcTileBytes, a_lo = a_hi + kTcTileBytes;
        const uint32_t b_hi = tab0 + (2 * part) * kTcTabBytes, b_lo = b_hi + kTcTabBytes;
        uint32_t acc = 0;
#pragma unroll
        for (int term = 0; term < 3; ++term) {          // hi.hi + lo.hi + hi.lo
          const uint32_t aa = term == 1 ? a_lo : a_hi, bb = term == 2 ? b_lo : b_hi;
#pragma unroll
          for (int ks = 0; ks < kFold / 16; ++ks) {     // K = 16 per instruction = 2 core matrices
            umma_f16(d, umma_desc(aa + ks * 256), umma_desc(bb + ks * 256), acc);
            acc = 1;
          }
        }
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
    }
    mbar_wait(bar, tl & 1);
    asm volatile("tcgen05.fence::after_thread_sync;" ::);

    // ---- accumulators -> power spectrum in shared memory (operands are dead now)
    {
      const int q = warp & 3, half = warp >> 2;            // TMEM lane quarter, column half
      const int t = 32 * q + lane;
      const uint32_t lane_addr = tmem_base + ((uint32_t)(32 * q) << 16);
      const float ds2 = descale * descale;
#pragma unroll
      for (int c8 = 0; c8 < kTcBins / 16; ++c8) {
        const int col = half * (kTcBins / 2) + 8 * c8;
        float re[8], im[8];
        tmem_ld8x2(lane_addr + col, lane_addr + kTcBins + col, re, im);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          Pw[(col + j) * kTileFrames + t] = fmaf(re[j], re[j], im[j] * im[j]) * ds2;
      }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::);
    __syncthreads();

    // ---- sparse mel projection, dB numerator, running max (identical to engine 1)
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

  // Frames beyond the kept width (frame 256 of 257) only feed the pair maximum: plain fp32 DFT
  // with the twiddles rebuilt from their fp16 halves (2^-22 accurate).
  const int n_extra = a.n_frames - n_tiles * kTileFrames;
  if (n_extra > 0) {
    float* Px = reinterpret_cast<float*>(tiles);  // [n_extra][kTcBins]
    const __half* Ch = reinterpret_cast<const __half*>(tab);
    const __half* Cl = reinterpret_cast<const __half*>(tab + kTcTabBytes);
    const __half* Sh = reinterpret_cast<const __half*>(tab + 2 * kTcTabBytes);
    const __half* Sl = reinterpret_cast<const __half*>(tab + 3 * kTcTabBytes);
    for (int i = tid; i < n_extra * kTcBins; i += kMelThreads) {
      const int x = i / kTcBins, ks = i % kTcBins;
      const int c = (n_tiles * kTileFrames + x) * a.hop;
      float re = 0.f, im = 0.f;
      for (int m = 0; m < kFold; ++m) {
        const float yp = padded_sample(y, c + m, a.crop, a.pad_mode);
        const float ym = m ? padded_sample(y, c - m, a.crop, a.pad_mode) : 0.f;
        const float e = wh[m] * (yp + ym), o = m ? wh[m] * (yp - ym) : 0.f;
        const int off = umma_offset(ks, m) >> 1;
        re = fmaf(e, __half2float(Ch[off]) + __half2float(Cl[off]), re);
        im = fmaf(o, __half2float(Sh[off]) + __half2float(Sl[off]), im);
      }
      Px[i] = fmaf(re, re, im * im);
    }
    __syncthreads();
    for (int i = tid; i < n_extra * a.n_mels; i += kMelThreads) {
      const int x = i / a.n_mels, m = i % a.n_mels;
      const int k0 = mk0[m];
      vmax = fmaxf(vmax, fmaf(mw1[m], Px[x * kTcBins + k0 + 1], mw0[m] * Px[x * kTcBins + k0]));
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
  if (warp == 0) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTcCols));
  }
}

size_t melspec_tc_smem_bytes(int crop) {
  size_t b = 4 * (size_t)kTcTabBytes + 4 * (size_t)kTcTileBytes;
  b += sizeof(float) * (size_t)((crop + 3) & ~3);
  b += sizeof(float) * (2 * kFold + 2 * kMaxMels) + sizeof(int) * kMaxMels;
  return b + 128;
}

int melspec_tc_launch_pair(const MelArgs& a, int n_pairs, long long n_records, cudaStream_t st) {
  static std::atomic<bool> configured[64];   // idempotent set-up; atomic so that concurrent callers do not race on the flag
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute(melspec_pair_tc_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    if (e != cudaSuccess) return (int)e;
    configured[dev] = true;
  }
  const size_t smem = melspec_tc_smem_bytes(a.crop);
  if (smem > 220 * 1024) return EEGSPEC_EUNSUPPORTED;
  melspec_pair_tc_kernel<<<dim3(n_pairs, (unsigned)n_records), kMelThreads, smem, st>>>(a);
  count_launch();
  return (int)cudaGetLastError();
}

// Host: fp16 hi/lo split of the folded DFT tables, laid out for the MMA (see umma_offset).
void melspec_tc_build_tables(const double* cos_d, const double* sin_d, int n_bins,
                             unsigned short* out /* 4 * kTcBins * kFold halves */) {
  for (int i = 0; i < 4 * kTcBins * kFold; ++i) out[i] = 0;
  for (int ks = 0; ks < n_bins && ks < kTcBins; ++ks)
    for (int m = 0; m < kFold; ++m) {
      const int off = umma_offset(ks, m) >> 1;
      for (int part = 0; part < 2; ++part) {
        const double v = part == 0 ? cos_d[(size_t)ks * kFold + m] : sin_d[(size_t)ks * kFold + m];
        const __half hi = __float2half_rn((float)v);
        const __half lo = __float2half_rn((float)(v - (double)__half2float(hi)));
        out[(2 * part) * kTcBins * kFold + off] = __half_as_ushort(hi);
        out[(2 * part + 1) * kTcBins * kFold + off] = __half_as_ushort(lo);
      }
    }
}

}  // namespace eegspec
