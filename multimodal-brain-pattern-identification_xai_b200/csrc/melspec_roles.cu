// Mode N, engine 3 "roles": the persistent clustered montage -> mel image kernel with the work of
// a tile split over specialised warp groups that run concurrently on different tiles.
//
// Same arithmetic and data flow as melspec_chain.cu (folded 128-sample Hann frame -> pruned DFT
// on tcgen05 with split-fp16 operands and fp32 TMEM accumulators -> |X|^2 -> 2-tap mel -> dB,
// work item = (record, chain), cluster of two CTAs = the two halves of the 256 kept frames,
// deferred fold of the pair maximum exchanged through distributed shared memory), but instead
// of 16 warps walking build -> epilogue -> difference in lockstep:
//
//   4 NQ warps   epilogue   thread = (frame, 128/NQ mels): tcgen05.ld, power, mel taps, lg2, sum;
//                           keeps the chain sum acc[] in registers (setmaxnreg.inc)
//   8 warps      build      electrode difference, NaN statistics, max|y|, fp16 hi/lo operand
//                           tiles of the next pair -- LSU / conversion work that overlaps the
//                           epilogue group's FMA / MUFU work on the previous pair
//   1 warp       mma        one lane issues the tcgen05.mma stream of a tile and its commit
//   1 warp       loader     one lane issues the cp.async.bulk (TMA) prefetch of the electrode rows
//   1 warp       publisher  one lane reduces the per-warp maxima of a finished tile and exchanges
//                           them with the peer CTA (st.shared::cluster + remote mbarrier arrive)
//
// All hand-offs are mbarriers; the only CTA-scope barriers are named barriers inside one group.
//
// Deferred fold without register cost.  power_to_db clips every pair image at max - top_db (and
// at amin); the clip level is only known once both halves of the pair have been reduced and
// exchanged, one tile later.  The lockstep kernel carries the pair's log values in registers
// until then (and spills); here a thread parks them in tensor memory instead: it writes its
// log-mel values over accumulator columns that only it has read (tcgen05.st), and the next
// tile's epilogue reads them back (tcgen05.ld), folds them with the now known clip level into
// the chain sum and only then hands the accumulator buffer back to the MMA warp.
//
// Front half follows the reference (pair list /root/reference/root/config/config.yml:529-545,
// differencing /root/reference/root/src/utils/data_utils.py:183-215, crop
// /root/reference/root/src/utils/cfg_utils.py:369-371, NaN fill cfg_utils.py:380-385).
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "mel_default_epilogue.h"
#include "melspec_chain.cuh"

namespace eegspec {

constexpr int kRBuildWarps = 8, kRBuild = 32 * kRBuildWarps;
constexpr int kRMaxEpiWarps = 16;
constexpr int kTlTiles = 64, kTlSlots = 32;
constexpr int kRYFloats = kCStageFloats + 2 * kCMargin + 128;

struct RolesShared {
  unsigned long long load_bar, stage_free, ops_bar, mma_bar[2], tmem_free[2], epi_done[2], xch_bar[2];
  uint32_t tmem_base;
  int e_tile[4];                    // power-of-two pre-scale exponent of tile j, [j & 3]
  float w_amax[kRBuildWarps];       // per-warp max|y| of the pair being differenced (NaN if any NaN)
  float w_vmax[2][kRMaxEpiWarps];      // per-warp log2 mel maximum of tile j (unscaled units), [j & 1]
  float xch[2][2];                  // [exchange & 1][cluster rank]: log2 of the pair maximum
  float red[kRBuildWarps];
  float px[kCMaxExtra * 128];       // power spectrum of the extra frames [frame][bin slot]
  int mel_k0[kMaxMels];
  float mel_w0[kMaxMels], mel_w1[kMaxMels];
};

template <int NQ>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128 * NQ + kRBuild + 128, 1)
melspec_roles_kernel(const ChainArgs a) {
  constexpr int kREpiWarps = 4 * NQ, kREpi = 128 * NQ, kMels = 128 / NQ;
  constexpr int kRMmaWarp = kREpiWarps + kRBuildWarps, kRLoadWarp = kRMmaWarp + 1, kRPubWarp = kRMmaWarp + 2;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char* tab = smem_raw;                                       // Ch | Cl | Sh | Sl
  unsigned char* tiles = tab + 4 * kCTabBytes;                         // Eh | El | Oh | Ol
  float* ybuf = reinterpret_cast<float*>(tiles + 4 * kCTileBytes);     // kRYFloats
  float* stage = ybuf + kRYFloats;                                     // 2 x kCStageFloats
  unsigned char* xt = reinterpret_cast<unsigned char*>(stage + 2 * kCStageFloats);   // Exh | Exl | Oxh | Oxl
  __shared__ RolesShared sh;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int h = (int)c_cluster_rank(), peer = h ^ 1;
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const int n_items = (int)a.n_items;
  // debug timeline: clock stamps of the first kTlTiles tiles of CTAs 0 and 1 (a.tl is normally null)
#define TL(slot) do { if (a.tl != nullptr && blockIdx.x < 2 && lane == 0 && j >= a.tl_j0 && j < a.tl_j0 + kTlTiles) a.tl[((blockIdx.x * kTlTiles + (j - a.tl_j0)) * kTlSlots) + (slot)] = clock64(); } while (0)
  const int n_my_items = n_items > cluster_id ? (n_items - cluster_id + n_clusters - 1) / n_clusters : 0;
  const int J = n_my_items * a.ppc;
  if (J == 0) return;   // both CTAs of the cluster agree

  const uint32_t load_bar = c_smem_u32(&sh.load_bar), stage_free = c_smem_u32(&sh.stage_free);
  const uint32_t ops_bar = c_smem_u32(&sh.ops_bar);
  const uint32_t mma_bar0 = c_smem_u32(&sh.mma_bar[0]), tmem_free0 = c_smem_u32(&sh.tmem_free[0]);
  const uint32_t xch_bar0 = c_smem_u32(&sh.xch_bar[0]), epi_done0 = c_smem_u32(&sh.epi_done[0]);
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(c_smem_u32(&sh.tmem_base)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  if (tid == 32) {
    c_mbar_init(load_bar, 1);
    c_mbar_init(stage_free, kRBuildWarps);
    c_mbar_init(ops_bar, kRBuildWarps);
    for (int b = 0; b < 2; ++b) {
      c_mbar_init(mma_bar0 + 8 * b, 1);
      c_mbar_init(tmem_free0 + 8 * b, kREpiWarps);
      c_mbar_init(epi_done0 + 8 * b, kREpiWarps);
      c_mbar_init(xch_bar0 + 8 * b, 2);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::);
  }
  {
    const uint4* src = reinterpret_cast<const uint4*>(a.tab);
    uint4* dst = reinterpret_cast<uint4*>(tab);
    for (int i = tid; i < 4 * kCTabBytes / 16; i += blockDim.x) dst[i] = __ldg(src + i);
    float4* yz = reinterpret_cast<float4*>(ybuf);
    for (int i = tid; i < kRYFloats / 4; i += blockDim.x) yz[i] = make_float4(0.f, 0.f, 0.f, 0.f);
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

  // ======================================================================= single-lane roles
  // The register file is partitioned per scheduler: this warp group gives its registers back so
  // that the epilogue group can hold its accumulators (setmaxnreg).
  if (warp >= kRMmaWarp) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
    if (warp == kRMmaWarp) {
      // ---- the tcgen05.mma stream: tile j as soon as its operands are built and the
      //      accumulator buffer (j & 1) has been handed back (tile j - 2 folded, during epilogue j - 1)
      for (int j = 0; j < J; ++j) {
        c_mbar_wait(ops_bar, j & 1);
        TL(5);
        if (j >= 2) c_mbar_wait(tmem_free0 + 8 * (j & 1), ((j - 2) >> 1) & 1);
        TL(6);
        asm volatile("tcgen05.fence::after_thread_sync;" ::);
        if (lane == 0) {
          const uint32_t tab0 = c_smem_u32(tab), til0 = c_smem_u32(tiles), xt0 = c_smem_u32(xt);
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
          asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(mma_bar0 + 8 * (j & 1))
                       : "memory");
        }
        TL(7);
        __syncwarp();
      }
    } else if (warp == kRLoadWarp) {
      // ---- TMA prefetch: both electrode rows of pair j + 1 once the difference of pair j has
      //      consumed the staging buffers
      if (lane == 0) {
        auto issue_load = [&](int j) {
          int rec, chain;
          const int pr = pair_of(j, rec, chain);
          const float* base = a.eeg + rec * a.rec_stride + a.crop_off + seg_lo;
          const uint32_t bytes = (uint32_t)seg_n * 4u;
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          c_mbar_expect_tx(load_bar, 2 * bytes);
          c_bulk_g2s(c_smem_u32(stage), base + a.pair_a[pr] * a.ch_stride, bytes, load_bar);
          c_bulk_g2s(c_smem_u32(stage + kCStageFloats), base + a.pair_b[pr] * a.ch_stride, bytes, load_bar);
        };
        issue_load(0);
        for (int j = 0; j + 1 < J; ++j) {
          c_mbar_wait(stage_free, j & 1);
          issue_load(j + 1);
        }
      }
    } else if (warp == kRPubWarp) {
      // ---- publish the maximum of every finished tile to both CTAs of the cluster
      if (lane == 0) {
        for (int j = 0; j < J; ++j) {
          c_mbar_wait(epi_done0 + 8 * (j & 1), (j >> 1) & 1);
          TL(14);
          float r_true = sh.w_vmax[j & 1][0];   // per-warp log2 maxima, already in unscaled units
          for (int w = 1; w < kREpiWarps; ++w) r_true = fmaxf(r_true, sh.w_vmax[j & 1][w]);
          float* slot = &sh.xch[j & 1][h];
          *slot = r_true;
          c_st_peer_f32(c_smem_u32(slot), peer, r_true);
          const uint32_t xb = xch_bar0 + 8 * (j & 1);
          c_mbar_arrive_cluster(xb, (uint32_t)h);
          c_mbar_arrive_cluster(xb, (uint32_t)peer);
          TL(15);
        }
      }
    }
  } else if (warp >= kREpiWarps) {
    // ================================================================================ build group
    // thread = (frame t, half of the fold indices): difference + statistics of pair j, then its
    // fp16 hi/lo operand tiles.  Runs one tile ahead of the epilogue group.
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
    const int btid = tid - kREpi, bwarp = warp - kREpiWarps;
    const int t_frame = btid & (kTileFrames - 1), part = btid >> 7;
    auto grp_max_nan = [&](float v) {
      v = warp_max_nan(v);
      if (lane == 0) sh.red[bwarp] = v;
      c_named_bar(1, kRBuild);
      float r = sh.red[0];
#pragma unroll
      for (int w = 1; w < kRBuildWarps; ++w) r = max_nan(r, sh.red[w]);
      c_named_bar(1, kRBuild);
      return r;
    };
    auto grp_sum = [&](float v) {
      v = warp_sum(v);
      if (lane == 0) sh.red[bwarp] = v;
      c_named_bar(1, kRBuild);
      float r = 0.f;
#pragma unroll
      for (int w = 0; w < kRBuildWarps; ++w) r += sh.red[w];
      c_named_bar(1, kRBuild);
      return r;
    };

    for (int j = 0; j < J; ++j) {
      c_named_bar(1, kRBuild);   // every warp has finished reading y / w_amax of pair j - 1
      if (bwarp == 0) TL(0);
      // ---- difference of pair j's electrode rows into y; per-warp max|y| (NaN if any NaN)
      c_mbar_wait(load_bar, j & 1);
      if (bwarp == 0) TL(1);
      {
        const float4* sa = reinterpret_cast<const float4*>(stage);
        const float4* sb = reinterpret_cast<const float4*>(stage + kCStageFloats);
        float4* yd = reinterpret_cast<float4*>(ybuf + kCMargin);
        float amax = 0.f;
        for (int i = btid; i < (seg_n >> 2); i += kRBuild) {
          const float4 x = sa[i], y = sb[i];
          const float4 d = make_float4(x.x - y.x, x.y - y.y, x.z - y.z, x.w - y.w);
          yd[i] = d;
          amax = max_nan(max_nan(amax, fabsf(d.x)), max_nan(fabsf(d.y), max_nan(fabsf(d.z), fabsf(d.w))));
        }
        amax = warp_max_nan(amax);
        if (lane == 0) sh.w_amax[bwarp] = amax;
        __syncwarp();
        if (lane == 0) c_mbar_arrive(stage_free);   // the loader may fetch pair j + 1
      }
      c_named_bar(1, kRBuild);   // y and w_amax are final
      if (bwarp == 0) TL(2);
      float amax = sh.w_amax[0];
#pragma unroll
      for (int w = 1; w < kRBuildWarps; ++w) amax = max_nan(amax, sh.w_amax[w]);
      if (amax != amax) {
        // NaN samples (rare).  cfg_utils.py:380-385 on the difference: nanmean over the whole
        // crop, all-NaN -> 0.  Each CTA re-reads both full rows, so no exchange is needed and
        // both halves get the identical fill value.
        int rec, chain;
        const int pr = pair_of(j, rec, chain);
        const float* ra = a.eeg + rec * a.rec_stride + a.pair_a[pr] * a.ch_stride + a.crop_off;
        const float* rb = a.eeg + rec * a.rec_stride + a.pair_b[pr] * a.ch_stride + a.crop_off;
        float s_ = 0.f, c_ = 0.f;
        for (int i = btid; i < a.crop; i += kRBuild) {
          const float d = __ldg(ra + i) - __ldg(rb + i);
          if (d != d) c_ += 1.f; else s_ += d;
        }
        s_ = grp_sum(s_);
        c_ = grp_sum(c_);
        const bool all_nan = c_ >= (float)a.crop;
        const float fill = all_nan ? 0.f : s_ / ((float)a.crop - c_);
        float m = 0.f;
        for (int i = btid; i < seg_n; i += kRBuild) {
          float d = ybuf[kCMargin + i];
          if (all_nan) d = 0.f; else if (d != d) d = fill;
          ybuf[kCMargin + i] = d;
          m = fmaxf(m, fabsf(d));
        }
        amax = grp_max_nan(m);
      }
      if (a.pad_mode == 1) {   // reflect padding (librosa < 0.10); zero padding is the resident state
        for (int k = btid; k < 2 * kCMargin + 128; k += kRBuild) {
          const int idx = k < kCMargin ? k : seg_n + k;          // left margin | right margin
          const int i = seg_lo + idx - kCMargin;                 // crop coordinate
          const int src = i < 0 ? -i : (i >= a.crop ? 2 * (a.crop - 1) - i : -1);
          const int sidx = src - seg_lo;
          if (idx < kRYFloats) ybuf[idx] = (src >= 0 && sidx >= 0 && sidx < seg_n) ? ybuf[kCMargin + sidx] : 0.f;
        }
        c_named_bar(1, kRBuild);
      }
      int e_j = 0;   // exact power-of-two pre-scale: amax * 2^e in [2^10, 2^11)
      {
        int ex = 0;
        if (amax > 0.f && amax < __int_as_float(0x7f800000)) { frexpf(amax, &ex); e_j = 11 - ex; }
        e_j = max(-60, min(60, e_j));
      }
      const float s = __int_as_float((127 + e_j) << 23);
      if (btid == 0) sh.e_tile[j & 3] = e_j;

      if (j >= 1) c_mbar_wait(mma_bar0 + 8 * ((j - 1) & 1), ((j - 1) >> 1) & 1);   // operand tiles are free
      if (bwarp == 0) TL(3);
      {   // operand build: E = s(y[c+m] + y[c-m]), O = s(y[c+m] - y[c-m]) as fp16 hi/lo
        const float* yc = ybuf + kCMargin + (128 * h + t_frame) * a.hop - seg_lo;
#pragma unroll
        for (int o = 0; o < 4; ++o) {
          const int m0 = 32 * part + 8 * o;
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
      if (btid < n_extra * kFold) {
        const int x = btid >> 6, m = btid & (kFold - 1);
        const float* yc = ybuf + kCMargin + (a.n_time + x) * a.hop - seg_lo;
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
      __syncwarp();
      if (lane == 0) c_mbar_arrive(ops_bar);
      if (bwarp == 0) TL(4);
    }
  } else {
    // ============================================================================= epilogue group
    // thread = (frame t = TMEM lane, part q): owns mels [q, q + 1) * 128 / NQ of frame 128 h + t
    // and keeps their chain sum acc[] in registers; the log-mel values of the tile in flight are
    // parked in accumulator columns only this thread reads (bins exclusive to its part).
    if constexpr (NQ == 4) asm volatile("setmaxnreg.inc.sync.aligned.u32 88;");    // launch: 72 x 896
    else asm volatile("setmaxnreg.inc.sync.aligned.u32 160;");                     // launch: 96 x 640
    const int t_frame = tid & (kTileFrames - 1), part = tid >> 7;
    // Parking columns of log-value group g (8 mels): the part's load chunk 1 + g / 2, Re half for
    // even g, Im half for odd g.  Those bins are read by this part only, and the generated
    // epilogue has consumed the chunk by the time the group is stored (one group late).
    const uint32_t qcol = NQ == 4 ? (uint32_t)(MELGEN_COL0_P4_1 * part) : (uint32_t)(MELGEN_COL0_P2_1 * part);
    static_assert(MELGEN_COL0_P4_2 == 2 * MELGEN_COL0_P4_1 && MELGEN_COL0_P4_3 == 3 * MELGEN_COL0_P4_1, "part pitch");
    auto park_col = [](int g) { return (uint32_t)(8 * (1 + g / 2) + (g & 1) * kCBins); };
    const uint32_t lane_base = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16);
    float acc[kMels];
#pragma unroll
    for (int i = 0; i < kMels; ++i) acc[i] = 0.f;
    float csum = 0.f;   // sum over the item's pairs of (reference + 2 e)
    int e_prev = 0;

    // exchange n is complete: fold the parked log values of tile n into the chain sum with the
    // pair's clip level, then hand accumulator buffer n & 1 back to the MMA warp.
    auto fold_tile = [&](int n) {
      c_mbar_wait_cluster(xch_bar0 + 8 * (n & 1), (uint32_t)((n >> 1) & 1));
      { const int j = n + 1; if (warp == 0) TL(9); }
      const float r_pair = fmaxf(fmaxf(sh.xch[n & 1][0], sh.xch[n & 1][1]), a.log2_amin);
      const float off = 2.f * (float)e_prev;          // e_prev = pre-scale exponent of tile n
      const float t_clip = fmaxf(a.log2_amin, r_pair - a.top_log2) + off;
      csum += r_pair + off;
      const uint32_t qaddr = lane_base + (n & 1) * kCAccCols + qcol;
#pragma unroll
      for (int g = 0; g < kMels / 8; ++g) {
        float q[8];
        c_tmem_ld8(qaddr + park_col(g), q);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[8 * g + i] += fmaxf(q[i], t_clip);
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::);
      __syncwarp();
      if (lane == 0) c_mbar_arrive(tmem_free0 + 8 * (n & 1));
    };
    auto flush = [&](int j_last) {   // write the item whose last tile is j_last
      int rec, chain;
      pair_of(j_last, rec, chain);
      float* o = a.out + (((long long)rec * a.n_mels + kMels * part) * a.n_time + 128 * h + t_frame) * a.n_chains + chain;
      const int mstride = a.n_time * a.n_chains;
#pragma unroll
      for (int i = 0; i < kMels; ++i) o[i * mstride] = fmaf(a.out_gain, acc[i] - csum, a.out_bias);
    };

    for (int j = 0; j < J; ++j) {
      c_mbar_wait(mma_bar0 + 8 * (j & 1), (j >> 1) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::);
      const int e_j = sh.e_tile[j & 3];
      if (warp == 0) TL(8);
      if (j >= 1) fold_tile(j - 1);
      if (warp == 0) TL(10);
      if (j % a.ppc == 0) {   // first pair of an item: write the finished one, start afresh
        if (j > 0) flush(j - 1);
#pragma unroll
        for (int i = 0; i < kMels; ++i) acc[i] = 0.f;
        csum = 0.f;
      }
      if (warp == 0) TL(11);
      const uint32_t lane_addr = lane_base + (j & 1) * kCAccCols;
      const uint32_t qaddr = lane_addr + qcol;
      float vmax = 0.f;
      float re[8], im[8], qb[16];
#define CH_POW(s, i) const float p##s = fmaf(re[i], re[i], im[i] * im[i]);
#define CH_MEL_TAIL(i)                                                                                   \
  vmax = fmaxf(vmax, v);                                                                                 \
  qb[(i) & 15] = c_lg2(v);                                                                               \
  if (((i) & 7) == 7 && (i) >= 15) CH_PARK(((i) - 15) / 8)
#define CH_PARK(g)                                                                                               \
  c_tmem_st8(qaddr + park_col(g), qb[(8 * (g)) & 15], qb[(8 * (g) + 1) & 15], qb[(8 * (g) + 2) & 15],             \
             qb[(8 * (g) + 3) & 15], qb[(8 * (g) + 4) & 15], qb[(8 * (g) + 5) & 15], qb[(8 * (g) + 6) & 15],      \
             qb[(8 * (g) + 7) & 15]);
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
      CH_PARK(kMels / 8 - 1)
      c_tmem_wait_st();
      if (warp == 0) TL(12);
#undef CH_PART
#undef CH_LOAD
#undef CH_POW
#undef CH_MEL_TAIL
#undef CH_PARK
#undef CH_MEL1
#undef CH_MEL2

      // extra frames: lane = bin slot, column = frame; power -> shared -> mel maximum (warps 0-3)
      if (n_extra > 0 && part == 0) {
        const uint32_t xaddr = tmem_base + ((uint32_t)(32 * warp) << 16) + kCXTmemCol + (j & 1) * 2 * kCXCols;
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
        c_named_bar(2, 128);   // px is rewritten by the next tile
      }
      vmax = warp_max_nan(vmax);
      if (lane == 0) sh.w_vmax[j & 1][warp] = c_lg2(vmax) - 2.f * (float)e_j;
      __syncwarp();
      if (lane == 0) c_mbar_arrive(epi_done0 + 8 * (j & 1));   // the publisher may exchange the maximum of tile j
      if (warp == 0) TL(13);
      e_prev = e_j;
    }

    // ---- the last pair's maximum has been exchanged: fold it and write the last item
    fold_tile(J - 1);
    flush(J - 1);
  }

  asm volatile("tcgen05.fence::before_thread_sync;" ::);
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  c_cluster_arrive();   // nobody exits while its peer may still write or arrive remotely
  c_cluster_wait();
}

// ------------------------------------------------------------------------------------- host
size_t melspec_roles_smem_bytes() {
  return 4 * (size_t)kCTabBytes + 4 * (size_t)kCTileBytes + sizeof(float) * ((size_t)kRYFloats + 2 * (size_t)kCStageFloats) +
         4 * (size_t)kCXTileBytes;
}

// Launches the role-split kernel on `clusters` clusters of two CTAs (arguments prepared by
// melspec_chain_launch, which owns the geometry checks).  EEGSPEC_ROLES_NQ = 2 | 4 picks the
// number of mel parts per frame (8 or 16 epilogue warps); 4 is the default.
int melspec_roles_run(const ChainArgs& a, long long clusters, cudaStream_t st) {
  static bool configured[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return EEGSPEC_EUNSUPPORTED;
  const size_t smem = melspec_roles_smem_bytes();
  if (!configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute(melspec_roles_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(melspec_roles_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    configured[dev] = true;
  }
  static const int nq = [] { const char* v = getenv("EEGSPEC_ROLES_NQ"); return (v && v[0] == '2') ? 2 : 4; }();
  // EEGSPEC_TIMELINE=<file>: debug only -- synchronous launch that dumps clock stamps of the
  // pipeline hand-offs of CTAs 0 and 1 (kTlTiles tiles from EEGSPEC_TIMELINE_J0) as raw int64.
  const char* tl_path = getenv("EEGSPEC_TIMELINE");
  ChainArgs b = a;
  const size_t tl_bytes = sizeof(long long) * 2 * kTlTiles * kTlSlots;
  if (tl_path) {
    const char* j0 = getenv("EEGSPEC_TIMELINE_J0");
    b.tl_j0 = j0 ? atoi(j0) : 0;
    if (cudaMalloc(&b.tl, tl_bytes) != cudaSuccess) return EEGSPEC_EWORKSPACE;
    cudaMemsetAsync(b.tl, 0, tl_bytes, st);
  }
  if (nq == 2) melspec_roles_kernel<2><<<(unsigned)(2 * clusters), 128 * 2 + kRBuild + 128, smem, st>>>(b);
  else melspec_roles_kernel<4><<<(unsigned)(2 * clusters), 128 * 4 + kRBuild + 128, smem, st>>>(b);
  if (tl_path) {
    std::vector<long long> hbuf(2 * kTlTiles * kTlSlots);
    cudaStreamSynchronize(st);
    cudaMemcpy(hbuf.data(), b.tl, tl_bytes, cudaMemcpyDeviceToHost);
    cudaFree(b.tl);
    if (FILE* f = fopen(tl_path, "wb")) { fwrite(hbuf.data(), 1, tl_bytes, f); fclose(f); }
  }
  return (int)cudaGetLastError();
}

}  // namespace eegspec
