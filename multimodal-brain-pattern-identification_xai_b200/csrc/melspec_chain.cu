// Mode N, engine 3: persistent, clustered, scratch-free montage -> mel image kernel.
//
// Same arithmetic as engine 2 (melspec_tc.cu: folded 128-sample Hann frame -> pruned DFT on
// tcgen05 with split-fp16 operands and fp32 TMEM accumulators -> |X|^2 -> 2-tap mel -> dB), but
//   * a work item is (record, chain); a cluster of two CTAs owns it, CTA h taking frames
//     128h .. 128h+127 of each of the chain's pairs (one M = 128 MMA tile per pair);
//   * every thread owns one frame (one TMEM lane) and 64 mels and keeps the chain sum acc[64]
//     and the previous pair's log-mel values qprev[64] in registers; the mel taps are literals
//     (mel_default_epilogue.h, generated) so that all register indices are static;
//   * power_to_db(ref = max) needs the pair maximum over both halves: the two CTAs exchange it
//     through distributed shared memory and the fold of pair p is deferred to the epilogue of
//     pair p+1, so the exchange latency is never on the critical path;
//   * the epilogue of tile j-1 runs while the tensor pipe works on tile j (accumulators are
//     double-buffered in TMEM), and the next pair's electrode rows arrive by cp.async.bulk.
// No scratch, no second kernel: the image is written once, from registers.
//
// Front half follows the reference (pair list /root/reference/root/config/config.yml:529-545,
// differencing /root/reference/root/src/utils/data_utils.py:183-215, crop
// /root/reference/root/src/utils/cfg_utils.py:369-371, NaN fill cfg_utils.py:380-385).
#include <cuda_fp16.h>
#include <stdlib.h>

#include "mel_default_epilogue.h"
#include "melspec_chain.cuh"

namespace eegspec {

// ------------------------------------------------------------------------------ the kernel
// Thread roles (NQ = mel parts per frame):  warps 0 .. 4 NQ - 1 compute -- thread = (frame t,
// part q): builds 64/NQ fold indices of frame t's operands and owns mels [q, q+1) * 128/NQ of it;
// the last warp is the control warp: one lane issues the TMA prefetch, the tcgen05.mma stream
// and the cross-CTA exchange of the pair maximum.  Compute warps never meet in a CTA-wide
// barrier except once per tile (y complete); everything else is mbarrier arrive / wait.
constexpr int kCMaxWarps = 16;

struct ChainShared {
  unsigned long long mma_bar, load_bar, ops_bar, epi_bar, diff_bar, xch_bar[2];
  uint32_t tmem_base;
  float w_amax[2][kCMaxWarps];      // per-warp max|y| of pair j (NaN if any NaN), [j & 1]
  float w_vmax[2][kCMaxWarps];      // per-warp log2 mel maximum of tile j (unscaled units), [j & 1]
  float xch[2][2];                  // [exchange & 1][cluster rank]: log2 of the pair maximum
  float red[kCMaxWarps];
  float px[kCMaxExtra * 128];       // power spectrum of the extra frames [frame][bin slot]
  int mel_k0[kMaxMels];
  float mel_w0[kMaxMels], mel_w1[kMaxMels];
};

template <int NQ>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128 * NQ + 128, 1)
melspec_chain_kernel(const ChainArgs a) {
  constexpr int kCompute = 128 * NQ, kWarps = 4 * NQ, kMels = 128 / NQ, kOct = 8 / NQ;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char* tab = smem_raw;                                       // Ch | Cl | Sh | Sl
  unsigned char* tiles = tab + 4 * kCTabBytes;                         // Eh | El | Oh | Ol
  float* ybuf = reinterpret_cast<float*>(tiles + 4 * kCTileBytes);     // 2 x kCYFloats
  float* stage = ybuf + 2 * kCYFloats;                                 // 2 x kCStageFloats
  unsigned char* xt = reinterpret_cast<unsigned char*>(stage + 2 * kCStageFloats);   // Exh | Exl | Oxh | Oxl
  __shared__ ChainShared sh;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int h = (int)c_cluster_rank(), peer = h ^ 1;
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const int n_items = (int)a.n_items;
  const int n_my_items = n_items > cluster_id ? (n_items - cluster_id + n_clusters - 1) / n_clusters : 0;
  const int J = n_my_items * a.ppc;
  if (J == 0) return;   // both CTAs of the cluster agree

  const uint32_t mma_bar = c_smem_u32(&sh.mma_bar), load_bar = c_smem_u32(&sh.load_bar);
  const uint32_t ops_bar = c_smem_u32(&sh.ops_bar), epi_bar = c_smem_u32(&sh.epi_bar);
  const uint32_t diff_bar = c_smem_u32(&sh.diff_bar);
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(c_smem_u32(&sh.tmem_base)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  if (tid == 32) {
    c_mbar_init(mma_bar, 1);
    c_mbar_init(load_bar, 1);
    c_mbar_init(ops_bar, kWarps);
    c_mbar_init(epi_bar, kWarps);
    c_mbar_init(diff_bar, kWarps);
    c_mbar_init(c_smem_u32(&sh.xch_bar[0]), 2);
    c_mbar_init(c_smem_u32(&sh.xch_bar[1]), 2);
    asm volatile("fence.mbarrier_init.release.cluster;" ::);
  }
  {
    const uint4* src = reinterpret_cast<const uint4*>(a.tab);
    uint4* dst = reinterpret_cast<uint4*>(tab);
    for (int i = tid; i < 4 * kCTabBytes / 16; i += blockDim.x) dst[i] = __ldg(src + i);
    float4* yz = reinterpret_cast<float4*>(ybuf);
    for (int i = tid; i < 2 * kCYFloats / 4; i += blockDim.x) yz[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    uint4* xz = reinterpret_cast<uint4*>(xt);
    for (int i = tid; i < 4 * kCXTileBytes / 16; i += blockDim.x) xz[i] = make_uint4(0u, 0u, 0u, 0u);
    for (int i = tid; i < a.n_mels; i += blockDim.x) {
      sh.mel_k0[i] = a.mel_k0[i];
      sh.mel_w0[i] = a.mel_w0[i];
      sh.mel_w1[i] = a.mel_w1[i];
    }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::);
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::);
  c_cluster_arrive();   // the peer's mbarriers are initialised before anyone arrives on them remotely
  c_cluster_wait();
  const uint32_t tmem_base = sh.tmem_base;

  const int seg_lo = a.seg_lo[h], seg_n = a.seg_n[h];
  const int n_extra = h ? min(a.n_frames - a.n_time, kCMaxExtra) : 0;

  auto pair_of = [&](int j, int& rec, int& chain) {
    const int item = cluster_id + (j / a.ppc) * n_clusters;
    rec = item / a.n_chains;
    chain = item - rec * a.n_chains;
    return chain * a.ppc + j % a.ppc;
  };

  // ======================================================================= control warp group
  // The register file is partitioned per scheduler (16 384 each): the control warp group gives
  // its registers back so that the compute warps can hold their accumulators (setmaxnreg).
  if (warp >= kWarps) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
   if (warp == kWarps) {
    auto issue_load = [&](int j) {   // both electrode rows of pair j -> staging (TMA bulk copies)
      int rec, chain;
      const int pr = pair_of(j, rec, chain);
      const float* base = a.eeg + rec * a.rec_stride + a.crop_off + seg_lo;
      const uint32_t bytes = (uint32_t)seg_n * 4u;
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      c_mbar_expect_tx(load_bar, 2 * bytes);
      c_bulk_g2s(c_smem_u32(stage), base + a.pair_a[pr] * a.ch_stride, bytes, load_bar);
      c_bulk_g2s(c_smem_u32(stage + kCStageFloats), base + a.pair_b[pr] * a.ch_stride, bytes, load_bar);
    };
    if (lane == 0) issue_load(0);
    c_mbar_wait(diff_bar, 0);                 // staging consumed by diff(0)
    if (lane == 0 && J > 1) issue_load(1);
    for (int j = 0; j <= J; ++j) {
      if (j < J) {                            // ---- the MMA stream of tile j
        c_mbar_wait(ops_bar, j & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::);
        if (lane == 0) {
          uint32_t tab0 = c_smem_u32(tab), til0 = c_smem_u32(tiles), xt0 = c_smem_u32(xt);
#pragma unroll 1
          for (int grp = 0; grp < (n_extra > 0 ? 4 : 2); ++grp) {   // Re, Im, extra Re, extra Im
            const int part = grp & 1;
            const bool big = grp < 2;
            // big: D[frame][bin] = tiles (A) x tables (B); extra: D'[bin][frame] = tables (A) x xt (B)
            const uint32_t d = big ? tmem_base + (j & 1) * kCAccCols + part * kCBins
                                   : tmem_base + kCXTmemCol + (j & 1) * 2 * kCXCols + part * kCXCols;
            const uint32_t t_hi = tab0 + (2 * part) * kCTabBytes, t_lo = t_hi + kCTabBytes;
            const uint32_t f_hi = big ? til0 + (2 * part) * kCTileBytes : xt0 + (2 * part) * kCXTileBytes;
            const uint32_t f_lo = f_hi + (big ? kCTileBytes : kCXTileBytes);
            const uint32_t idesc = big ? kCIdesc : kCIdescX;
            uint32_t accum = 0;
#pragma unroll 1
            for (int term = 0; term < 3; ++term) {   // hi.hi + lo.hi + hi.lo  (frames x tables)
              const uint64_t fd = c_umma_desc(term == 1 ? f_lo : f_hi), td = c_umma_desc(term == 2 ? t_lo : t_hi);
#pragma unroll
              for (int ks = 0; ks < kFold / 16; ++ks) {
                const uint64_t fdk = fd + (uint64_t)(ks * 16), tdk = td + (uint64_t)(ks * 16);
                c_umma(d, big ? fdk : tdk, big ? tdk : fdk, accum, idesc);
                accum = 1;
              }
            }
          }
          asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(mma_bar) : "memory");
        }
        __syncwarp();
      }
      if (j >= 1) {                           // ---- publish the maximum of tile j-1 to both CTAs
        const int jt = j - 1;
        c_mbar_wait(epi_bar, jt & 1);
        if (lane == 0) {
          float r_true = sh.w_vmax[jt & 1][0];   // per-warp log2 maxima, already in unscaled units
          for (int w = 1; w < kWarps; ++w) r_true = fmaxf(r_true, sh.w_vmax[jt & 1][w]);
          float* slot = &sh.xch[jt & 1][h];
          *slot = r_true;
          c_st_peer_f32(c_smem_u32(slot), peer, r_true);
          const uint32_t xb = c_smem_u32(&sh.xch_bar[jt & 1]);
          c_mbar_arrive_cluster(xb, (uint32_t)h);
          c_mbar_arrive_cluster(xb, (uint32_t)peer);
        }
        __syncwarp();
      }
      if (j + 1 < J) {                        // ---- staging is free once diff(j+1) is done
        c_mbar_wait(diff_bar, (j + 1) & 1);
        if (lane == 0 && j + 2 < J) issue_load(j + 2);
      }
    }
   }
  } else {
    if constexpr (NQ == 4) asm volatile("setmaxnreg.inc.sync.aligned.u32 112;");   // pool: 128 x (96 - 24) freed
    else asm volatile("setmaxnreg.inc.sync.aligned.u32 240;");
    // ============================================================================= compute warps
    const int t_frame = tid & (kTileFrames - 1), part = tid >> 7;
    auto cta_max_nan = [&](float v) {
      v = warp_max_nan(v);
      if (lane == 0) sh.red[warp] = v;
      c_named_bar(1, kCompute);
      float r = sh.red[0];
#pragma unroll
      for (int w = 1; w < kWarps; ++w) r = max_nan(r, sh.red[w]);
      c_named_bar(1, kCompute);
      return r;
    };
    auto cta_sum = [&](float v) {
      v = warp_sum(v);
      if (lane == 0) sh.red[warp] = v;
      c_named_bar(1, kCompute);
      float r = 0.f;
#pragma unroll
      for (int w = 0; w < kWarps; ++w) r += sh.red[w];
      c_named_bar(1, kCompute);
      return r;
    };
    // difference of pair j's electrode rows into ybuf[j & 1]; per-warp max|y| (NaN if any NaN)
    auto diff_pass = [&](int j) {
      c_mbar_wait(load_bar, j & 1);
      const float4* sa = reinterpret_cast<const float4*>(stage);
      const float4* sb = reinterpret_cast<const float4*>(stage + kCStageFloats);
      float4* yd = reinterpret_cast<float4*>(ybuf + (j & 1) * kCYFloats + kCMargin);
      float amax = 0.f;
      for (int i = tid; i < (seg_n >> 2); i += kCompute) {
        const float4 x = sa[i], y = sb[i];
        const float4 d = make_float4(x.x - y.x, x.y - y.y, x.z - y.z, x.w - y.w);
        yd[i] = d;
        amax = max_nan(max_nan(amax, fabsf(d.x)), max_nan(fabsf(d.y), max_nan(fabsf(d.z), fabsf(d.w))));
      }
      amax = warp_max_nan(amax);
      if (lane == 0) sh.w_amax[j & 1][warp] = amax;
      __syncwarp();
      if (lane == 0) c_mbar_arrive(diff_bar);
    };

    float acc[kMels], qprev[kMels];
#pragma unroll
    for (int i = 0; i < kMels; ++i) { acc[i] = 0.f; qprev[i] = 0.f; }
    float csum = 0.f;
    int e_jm1 = 0, e_jm2 = 0;

    auto flush = [&](int j_last, float t_fold) {   // write the item whose last tile is j_last
      int rec, chain;
      pair_of(j_last, rec, chain);
      float* o = a.out + (((long long)rec * a.n_mels + kMels * part) * a.n_time + 128 * h + t_frame) * a.n_chains + chain;
      const int mstride = a.n_time * a.n_chains;
#pragma unroll
      for (int i = 0; i < kMels; ++i)
        o[i * mstride] = fmaf(a.out_gain, acc[i] + fmaxf(qprev[i], t_fold) - csum, a.out_bias);
    };
    auto fold_constants = [&](int n, float& t_out) {   // exchange n: threshold and reference of tile n
      c_mbar_wait_cluster(c_smem_u32(&sh.xch_bar[n & 1]), (uint32_t)((n >> 1) & 1));
      const float r_pair = fmaxf(fmaxf(sh.xch[n & 1][0], sh.xch[n & 1][1]), a.log2_amin);
      const float off = 2.f * (float)e_jm2;
      t_out = fmaxf(a.log2_amin, r_pair - a.top_log2) + off;
      csum += r_pair + off;
    };

    diff_pass(0);
    for (int j = 0; j <= J; ++j) {
      int e_j = 0;
      if (j < J) {
        c_named_bar(1, kCompute);   // diff(j) is complete in every warp: y and w_amax are final
        float* yb = ybuf + (j & 1) * kCYFloats;
        float amax = sh.w_amax[j & 1][0];
#pragma unroll
        for (int w = 1; w < kWarps; ++w) amax = max_nan(amax, sh.w_amax[j & 1][w]);
        if (amax != amax) {
          // NaN samples (rare).  cfg_utils.py:380-385 on the difference: nanmean over the whole
          // crop, all-NaN -> 0.  Each CTA re-reads both full rows, so no exchange is needed and
          // both halves get the identical fill value.
          int rec, chain;
          const int pr = pair_of(j, rec, chain);
          const float* ra = a.eeg + rec * a.rec_stride + a.pair_a[pr] * a.ch_stride + a.crop_off;
          const float* rb = a.eeg + rec * a.rec_stride + a.pair_b[pr] * a.ch_stride + a.crop_off;
          float s_ = 0.f, c_ = 0.f;
          for (int i = tid; i < a.crop; i += kCompute) {
            const float d = __ldg(ra + i) - __ldg(rb + i);
            if (d != d) c_ += 1.f; else s_ += d;
          }
          s_ = cta_sum(s_);
          c_ = cta_sum(c_);
          const bool all_nan = c_ >= (float)a.crop;
          const float fill = all_nan ? 0.f : s_ / ((float)a.crop - c_);
          float m = 0.f;
          for (int i = tid; i < seg_n; i += kCompute) {
            float d = yb[kCMargin + i];
            if (all_nan) d = 0.f; else if (d != d) d = fill;
            yb[kCMargin + i] = d;
            m = fmaxf(m, fabsf(d));
          }
          amax = cta_max_nan(m);
        }
        if (a.pad_mode == 1) {   // reflect padding (librosa < 0.10); zero padding is the resident state
          for (int k = tid; k < 2 * kCMargin + 128; k += kCompute) {
            const int idx = k < kCMargin ? k : seg_n + k;          // left margin | right margin
            const int i = seg_lo + idx - kCMargin;                 // crop coordinate
            const int src = i < 0 ? -i : (i >= a.crop ? 2 * (a.crop - 1) - i : -1);
            const int sidx = src - seg_lo;
            if (idx < kCYFloats) yb[idx] = (src >= 0 && sidx >= 0 && sidx < seg_n) ? yb[kCMargin + sidx] : 0.f;
          }
          c_named_bar(1, kCompute);
        }
        {   // exact power-of-two pre-scale: amax * 2^e in [2^10, 2^11)
          int ex = 0;
          if (amax > 0.f && amax < __int_as_float(0x7f800000)) { frexpf(amax, &ex); e_j = 11 - ex; }
          e_j = max(-60, min(60, e_j));
        }
        const float s = __int_as_float((127 + e_j) << 23);

        if (j >= 1) c_mbar_wait(mma_bar, (j - 1) & 1);   // tile j-1 is done: operand tiles are free
        asm volatile("tcgen05.fence::after_thread_sync;" ::);
        {   // operand build: E = s(y[c+m] + y[c-m]), O = s(y[c+m] - y[c-m]) as fp16 hi/lo
          const float* yc = yb + kCMargin + (128 * h + t_frame) * a.hop - seg_lo;
#pragma unroll
          for (int o = 0; o < kOct; ++o) {
            const int m0 = (64 / NQ) * part + 8 * o;
            float ev[8], ov[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              const float yp = yc[m0 + q], yms = yc[-(m0 + q)] * s;
              ev[q] = fmaf(yp, s, yms);
              ov[q] = fmaf(yp, s, -yms);
            }
            uint4 eh, el, oh, ol;
            c_split2(ev[0], ev[1], eh.x, el.x); c_split2(ev[2], ev[3], eh.y, el.y);
            c_split2(ev[4], ev[5], eh.z, el.z); c_split2(ev[6], ev[7], eh.w, el.w);
            c_split2(ov[0], ov[1], oh.x, ol.x); c_split2(ov[2], ov[3], oh.y, ol.y);
            c_split2(ov[4], ov[5], oh.z, ol.z); c_split2(ov[6], ov[7], oh.w, ol.w);
            const int off = c_umma_offset(t_frame, m0);
            *reinterpret_cast<uint4*>(tiles + 0 * kCTileBytes + off) = eh;
            *reinterpret_cast<uint4*>(tiles + 1 * kCTileBytes + off) = el;
            *reinterpret_cast<uint4*>(tiles + 2 * kCTileBytes + off) = oh;
            *reinterpret_cast<uint4*>(tiles + 3 * kCTileBytes + off) = ol;
          }
        }
        // frames beyond the kept width (they only feed the pair maximum) are rows of a small
        // second operand: one (frame, fold index) element per thread
        if (tid < n_extra * kFold) {
          const int x = tid >> 6, m = tid & (kFold - 1);
          const float* yc = yb + kCMargin + (a.n_time + x) * a.hop - seg_lo;
          const float yp = yc[m], yms = yc[-m] * s;
          const float e = fmaf(yp, s, yms), o = fmaf(yp, s, -yms);
          const float eh = __uint_as_float(__float_as_uint(e) & 0xFFFFE000u);
          const float oh = __uint_as_float(__float_as_uint(o) & 0xFFFFE000u);
          const int off = c_umma_offset(x, m);
          *reinterpret_cast<__half*>(xt + 0 * kCXTileBytes + off) = __float2half_rn(eh);
          *reinterpret_cast<__half*>(xt + 1 * kCXTileBytes + off) = __float2half_rn(e - eh);
          *reinterpret_cast<__half*>(xt + 2 * kCXTileBytes + off) = __float2half_rn(oh);
          *reinterpret_cast<__half*>(xt + 3 * kCXTileBytes + off) = __float2half_rn(o - oh);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::);
        __syncwarp();
        if (lane == 0) c_mbar_arrive(ops_bar);
      }

      if (j >= 1) {   // ---- epilogue of tile j-1 (the tensor pipe is busy with tile j meanwhile)
        const int jt = j - 1;
        if (j == J) c_mbar_wait(mma_bar, jt & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::);
        float t_prev = 0.f;   // clip threshold of tile j-2, in that tile's scaled log2 units
        if (j >= 2) fold_constants(j - 2, t_prev);
        float t_fold = t_prev;
        if (jt % a.ppc == 0) {   // first pair of an item: write the finished one, start afresh
          if (jt > 0) flush(jt - 1, t_prev);
#pragma unroll
          for (int i = 0; i < kMels; ++i) { acc[i] = 0.f; qprev[i] = 0.f; }
          csum = 0.f;
          t_fold = 0.f;
        }
        const uint32_t lane_addr = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16) + (jt & 1) * kCAccCols;
        float vmax = 0.f;
        float re[8], im[8];
#define CH_POW(s, i) const float p##s = fmaf(re[i], re[i], im[i] * im[i]);
#define CH_MEL_TAIL(i)                       \
  vmax = fmaxf(vmax, v);                     \
  acc[i] += fmaxf(qprev[i], t_fold);         \
  qprev[i] = c_lg2(v);
#define CH_MEL1(i, sa, w0) { const float v = (w0) * p##sa; CH_MEL_TAIL(i) }
#define CH_MEL2(i, sa, sb, w0, w1) { const float v = fmaf(w1, p##sb, (w0) * p##sa); CH_MEL_TAIL(i) }
#define CH_LOAD(c) c_tmem_ld8x2(lane_addr + col0 + 8 * (c), lane_addr + kCBins + col0 + 8 * (c), re, im);
#define CH_PART(TAG)                                                \
  {                                                                 \
    constexpr int col0 = MELGEN_COL0_##TAG;                         \
    MELGEN_EPILOGUE_##TAG(CH_LOAD, CH_POW, CH_MEL1, CH_MEL2)        \
  }
        if constexpr (NQ == 2) {
          if (part == 0) CH_PART(P2_0) else CH_PART(P2_1)
        } else {
          if (part == 0) CH_PART(P4_0) else if (part == 1) CH_PART(P4_1) else if (part == 2) CH_PART(P4_2) else CH_PART(P4_3)
        }
#undef CH_PART
#undef CH_LOAD
#undef CH_POW
#undef CH_MEL_TAIL
#undef CH_MEL1
#undef CH_MEL2

        // extra frames: lane = bin slot, column = frame; power -> shared -> mel maximum (warps 0-3)
        if (n_extra > 0 && part == 0) {
          const uint32_t xaddr = tmem_base + ((uint32_t)(32 * warp) << 16) + kCXTmemCol + (jt & 1) * 2 * kCXCols;
          float xr[4], xi[4];
          c_tmem_ld4(xaddr, xr);
          c_tmem_ld4(xaddr + kCXCols, xi);
#pragma unroll
          for (int x = 0; x < kCMaxExtra; ++x) sh.px[x * 128 + tid] = fmaf(xr[x], xr[x], xi[x] * xi[x]);
          c_named_bar(2, 128);
          for (int i = tid; i < n_extra * a.n_mels; i += 128) {
            const int x = i / a.n_mels, m = i % a.n_mels;
            const int k0 = sh.mel_k0[m];
            vmax = fmaxf(vmax, fmaf(sh.mel_w1[m], sh.px[x * 128 + k0 + 1], sh.mel_w0[m] * sh.px[x * 128 + k0]));
          }
        }
        vmax = warp_max_nan(vmax);
        if (lane == 0) sh.w_vmax[jt & 1][warp] = c_lg2(vmax) - 2.f * (float)e_jm1;
        asm volatile("tcgen05.fence::before_thread_sync;" ::);
        __syncwarp();
        if (lane == 0) c_mbar_arrive(epi_bar);
      }

      if (j + 1 < J) diff_pass(j + 1);
      e_jm2 = e_jm1;
      e_jm1 = e_j;
    }

    // ---- the last pair's maximum has been exchanged: write the last item
    float t_last;
    fold_constants(J - 1, t_last);
    flush(J - 1, t_last);
  }

  asm volatile("tcgen05.fence::before_thread_sync;" ::);
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  c_cluster_arrive();   // nobody exits while its peer may still write or arrive remotely
  c_cluster_wait();
}

// ------------------------------------------------------------------------------------- host
size_t melspec_chain_smem_bytes() {
  return 4 * (size_t)kCTabBytes + 4 * (size_t)kCTileBytes + sizeof(float) * (2 * (size_t)kCYFloats + 2 * (size_t)kCStageFloats) +
         4 * (size_t)kCXTileBytes;
}

// The generated epilogue is only valid for the bank it was generated from.
bool melspec_chain_bank_matches(const eegspec_mel_params& p, int k_lo, const int* k0, const float* w0,
                                const float* w1) {
  if (p.sr != MELGEN_SR || p.n_fft != MELGEN_NFFT || p.n_mels != MELGEN_NMELS || k_lo != MELGEN_KLO) return false;
  for (int m = 0; m < MELGEN_NMELS; ++m) {
    if (k0[m] != kGenK0[m]) return false;
    if (fabsf(w0[m] - kGenW0[m]) > 2e-6f * fabsf(kGenW0[m])) return false;
    if (fabsf(w1[m] - kGenW1[m]) > 2e-6f * fabsf(kGenW1[m]) + 1e-12f) return false;
  }
  return true;
}

// Returns EEGSPEC_EUNSUPPORTED when this geometry / layout is not served by engine 3.
int melspec_chain_launch(const MelArgs& m, const eegspec_mel_params& p, long long batch, cudaStream_t st) {
  if (p.time_major || m.n_time != 2 * kTileFrames || m.n_frames < m.n_time) return EEGSPEC_EUNSUPPORTED;
  if (m.n_frames - m.n_time > kCMaxExtra || m.hop > 40 || m.hop < 1 || (m.crop & 7)) return EEGSPEC_EUNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(m.eeg) & 15) || (m.rec_stride & 3) || (m.ch_stride & 3) || (m.crop_off & 3))
    return EEGSPEC_EUNSUPPORTED;
  ChainArgs a;
  a.tl = nullptr;
  a.tl_j0 = 0;
  a.stagger_ns = 0;
  a.eeg = m.eeg;
  a.out = m.out;
  a.tab = m.t.tc3_tab;
  a.mel_k0 = m.t.mel_k0;
  a.mel_w0 = m.t.mel_w0;
  a.mel_w1 = m.t.mel_w1;
  a.rec_stride = m.rec_stride;
  a.ch_stride = m.ch_stride;
  a.n_items = batch * m.n_chains;
  a.crop = m.crop;
  a.crop_off = m.crop_off;
  a.hop = m.hop;
  a.n_frames = m.n_frames;
  a.n_mels = m.n_mels;
  a.n_time = m.n_time;
  a.n_chains = m.n_chains;
  a.ppc = m.ppc;
  a.pad_mode = m.pad_mode;
  const int last = (m.n_frames < m.n_time + kCMaxExtra ? m.n_frames : m.n_time + kCMaxExtra) - 1;
  auto up4 = [](int v) { return (v + 3) & ~3; };
  a.seg_lo[0] = 0;
  a.seg_n[0] = (up4((kTileFrames - 1) * m.hop + 64) < m.crop ? up4((kTileFrames - 1) * m.hop + 64) : m.crop);
  a.seg_lo[1] = (kTileFrames * m.hop - 63) & ~3;
  const int hi1 = up4(last * m.hop + 64) < m.crop ? up4(last * m.hop + 64) : m.crop;
  a.seg_n[1] = hi1 - a.seg_lo[1];
  for (int h = 0; h < 2; ++h)
    if (a.seg_n[h] <= 0 || a.seg_n[h] > kCStageFloats || a.seg_lo[h] < 0) return EEGSPEC_EUNSUPPORTED;
  if (a.seg_lo[1] + a.seg_n[1] < m.crop && last * m.hop + 64 > a.seg_lo[1] + a.seg_n[1]) return EEGSPEC_EUNSUPPORTED;
  a.log2_amin = log2f(m.amin);
  a.top_log2 = m.top_db / kLog2To10Log10;
  a.out_bias = m.db_offset * m.inv_db_scale;
  a.out_gain = kLog2To10Log10 * m.inv_db_scale / (float)m.ppc;
  for (int i = 0; i < EEGSPEC_MAX_PAIRS; ++i) { a.pair_a[i] = m.pair_a[i]; a.pair_b[i] = m.pair_b[i]; }

  static bool configured[64] = {};
  static int sm_count[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return EEGSPEC_EUNSUPPORTED;
  const size_t smem = melspec_chain_smem_bytes();
  if (!configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute(melspec_chain_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(melspec_chain_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    e = cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return (int)e;
    configured[dev] = true;
  }
  long long clusters = sm_count[dev] / 2;
  if (clusters < 1) clusters = 1;
  if (clusters > a.n_items) clusters = a.n_items;
  if ((a.n_items + clusters - 1) / clusters * a.ppc > (1LL << 30)) return EEGSPEC_EUNSUPPORTED;
  // EEGSPEC_CHAIN_NQ = 2 | 4 selects the lockstep kernel of this file (2 or 4 mel parts per frame);
  // the default is the role-split kernel (melspec_roles.cu).
  static const int nq = [] { const char* v = getenv("EEGSPEC_CHAIN_NQ"); return v ? (v[0] == '2' ? 2 : 4) : 0; }();
  // EEGSPEC_ENGINE3 = roles forces the older role-split kernel (operands in shared memory).
  static const bool force_roles = [] { const char* v = getenv("EEGSPEC_ENGINE3"); return v && v[0] == 'r'; }();
  if (nq == 0 && !force_roles && a.n_chains == 4 && a.ppc == 4 && a.n_mels == 128) {
    // default: operands in tensor memory, work item = record (melspec_ts.cu)
    ChainArgs b = a;
    b.n_items = batch;
    static const unsigned stagger = [] { const char* v = getenv("EEGSPEC_STAGGER_NS"); return v ? (unsigned)atoi(v) : 2500u; }();
    b.stagger_ns = stagger;
    long long cl = sm_count[dev] / 2;
    if (cl < 1) cl = 1;
    if (cl > batch) cl = batch;
    if ((batch + cl - 1) / cl * 16 > (1LL << 28)) return EEGSPEC_EUNSUPPORTED;
    const int rc = melspec_ts_run(b, cl, st);
    if (rc == 0) count_launch();
    return rc;
  }
  if (nq == 0) {
    const int rc = melspec_roles_run(a, clusters, st);
    if (rc == 0) count_launch();
    return rc;
  }
  if (nq == 2) melspec_chain_kernel<2><<<(unsigned)(2 * clusters), 128 * 2 + 128, smem, st>>>(a);
  else melspec_chain_kernel<4><<<(unsigned)(2 * clusters), 128 * 4 + 128, smem, st>>>(a);
  count_launch();
  return (int)cudaGetLastError();
}

}  // namespace eegspec
