// Mode N, engine 3 "ts": the persistent clustered montage -> mel image kernel with the frame
// operand of the pruned DFT in TENSOR MEMORY (tcgen05.mma, A from TMEM, tables from shared memory).
//
// Why: with both operands in shared memory an M = 128, K = 16 tcgen05.mma costs ~126 cycles for
// any N <= 128 (the 4 KB frame operand is read at ~32 B/clk; tools/mma_probe.cu), so the 24 MMAs
// of a tile took 3 000 cycles and the tensor pipe was the bottleneck of the earlier role-split
// kernel (git history) at 15 % "active".  With the frame operand in tensor memory the same MMA
// runs at its N / 2 floor (59 cycles at N = 112, tools/mma_probe2.cu).
//
// Arithmetic: folded 128-sample Hann frame, split-fp16 operands, fp32
// TMEM accumulators, |X|^2 -> 2-tap mel -> lg2, deferred clip of the pair image with the maximum
// exchanged between the two CTAs of the cluster, log values parked in tensor memory), and:
//
//   * work item = record: the cluster walks the 16 pairs of a record (CTA h = frames 128 h ..
//     128 h + 127), so the four chain images of a pixel meet in one thread.  Chains 0-2 are
//     stashed planar in the record's own output rows (full 128-byte lines, L2 resident) and the
//     last chain is written together with them as one float4 per pixel: every global store is a
//     full sector (the per-chain 4-byte stores at a 16-byte pitch of the older kernels cost
//     104 M partial sectors per 1 024 records and a third of the L1 -> L2 write bandwidth).
//   * the build group writes the operand K-step by K-step into a two-slot ring in tensor
//     memory (tcgen05.st) and the MMA warp consumes each slot with six MMAs
//     (Re: Eh.Ch + El.Ch + Eh.Cl, Im: Oh.Sh + Ol.Sh + Oh.Sl), so operand build and tensor pipe
//     overlap within a tile and no operand tile lives in shared memory;
//   * N = 104 (bins 1..102 + 2 zero rows): 2 x 208 accumulator columns + 64 ring columns +
//     32 columns for the frames beyond the kept width = 512;
//   * the frames beyond the kept width only feed the pair maximum: one fp16 term
//     (tables as A from shared memory, frames as B, N = 8) is accurate enough for that;
//   * the pair maximum is exchanged by the epilogue warps themselves (shared-memory atomic max
//     on both CTAs + remote mbarrier arrive), no publisher hop.
//
// Front half follows the reference (pair list /root/reference/root/config/config.yml:529-545,
// differencing /root/reference/root/src/utils/data_utils.py:183-215, crop
// /root/reference/root/src/utils/cfg_utils.py:369-371, NaN fill cfg_utils.py:380-385).
#include <atomic>
#ifdef EEGSPEC_TIMELINE   // debug builds only (tools/timeline_roles.py): clock stamps of the pipeline hand-offs
#include <stdio.h>
#include <stdlib.h>

#include <vector>
#endif

#include <math.h>

#include <mutex>

#include "mel_default_epilogue.h"
#include "melspec_chain.cuh"

namespace eegspec {

constexpr int kTBins = 104;                          // MMA N
constexpr int kTAcc = 2 * kTBins;                    // accumulator columns per buffer: Re | Im
constexpr int kTRingCol = 2 * kTAcc;                 // 416: ring of 2 slots x {Eh, El, Oh, Ol} x 8 columns
constexpr int kTXCol = kTRingCol + 64;               // 480: [buffer][Re | Im][8] of the extra frames
constexpr int kTXN = 8;                              // MMA N of the extra-frame product
constexpr int kTXTileBytes = kTXN * kFold * 2;       // 1 024: 8 rows x 64 fp16
constexpr int kTBuildWarps = 8, kTBuild = 32 * kTBuildWarps;
constexpr int kTEpiWarps = 16, kTEpi = 32 * kTEpiWarps, kTMels = 32;
constexpr int kTMmaWarp = kTEpiWarps + kTBuildWarps, kTLoadWarp = kTMmaWarp + 1, kTDiffWarp0 = kTMmaWarp + 2;
constexpr int kTDiff = 64;   // two difference warps
constexpr int kTThreads = kTEpi + kTBuild + 128;
constexpr int kTYFloats = kCStageFloats + 2 * kCMargin + 128;
constexpr int kTlTilesT = 64, kTlSlotsT = 40;
constexpr uint32_t kTIdesc = (1u << 4) | ((uint32_t)(kTBins >> 3) << 17) | ((uint32_t)(kTileFrames >> 4) << 24);
constexpr uint32_t kTIdescX = (1u << 4) | ((uint32_t)(kTXN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);

static_assert(kTXCol + 4 * kTXN <= 512, "tensor memory budget");
static_assert(MELGEN_COL0_P4_3 + 8 * MELGEN_CHUNKS_P4_3 <= kTBins, "the generated epilogue stays inside N");

struct TsShared {
  unsigned long long load_bar[2], stage_free, y_ready[2], y_free[2], ops_bar[2], afree_bar[2], mma_bar[2], tmem_free[2], xch_bar[2];
  uint32_t tmem_base;
  unsigned int xmax[4];             // this CTA's maximum of tile j (order-preserving uint of log2), [j & 3]
  __align__(16) float pv[2][kTEpiWarps];   // the peer's per-warp maxima of tile j (log2), [j & 1], written by st.async
  int e_tile[4];                    // power-of-two pre-scale exponent of tile j, [j & 3]
  float w_amax[2][2];               // per diff warp max|y| of the pair in y buffer j & 1 (NaN if any NaN)
  float amax_fin[2];                // max|y| after the NaN fill, [j & 1]
  float red[2];
  float4 xw[128];                   // extra frames: mel taps owned by table row r: {w0, w1} of up to two mels with k0 = bin(r)
};

// tcgen05.mma, A operand in tensor memory; the shared-memory descriptor of B is passed as two words
// (only the low one, the address field, differs between the MMAs of this kernel)
__device__ __forceinline__ void t_umma_ts(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_lo, uint32_t b_hi, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 bd;\n\tsetp.ne.b32 p, %5, 0;\n\tmov.b64 bd, {%2, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], bd, %4, p;\n\t}\n"
      :
      : "r"(d_tmem), "r"(a_tmem), "r"(b_lo), "r"(b_hi), "r"(kTIdesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ long long t_globaltimer() {
  long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ bool t_elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}\n" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void t_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void t_tmem_st4(uint32_t addr, uint32_t v0, uint32_t v1, uint32_t v2, uint32_t v3) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};\n" ::"r"(addr), "r"(v0), "r"(v1), "r"(v2),
               "r"(v3)
               : "memory");
}
// 2 x 8 TMEM columns of this thread's lane, one wait
__device__ __forceinline__ void t_tmem_ld8x2(uint32_t a0, uint32_t a1, float (&v)[16]) {
  uint32_t r[16];
#define T_LD8(o, addr)                                                                                              \
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"                    \
               : "=r"(r[o]), "=r"(r[o + 1]), "=r"(r[o + 2]), "=r"(r[o + 3]), "=r"(r[o + 4]), "=r"(r[o + 5]),         \
                 "=r"(r[o + 6]), "=r"(r[o + 7])                                                                     \
               : "r"(addr));
  T_LD8(0, a0) T_LD8(8, a1)
#undef T_LD8
  asm volatile("tcgen05.wait::ld.sync.aligned;\n"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                 "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
               :
               : "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
// order-preserving float <-> uint (larger float = larger uint; 0 is below every float)
__device__ __forceinline__ unsigned int t_f2ord(float f) {
  const unsigned int b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float t_ord2f(unsigned int u) {
  return __uint_as_float((u & 0x80000000u) ? (u ^ 0x80000000u) : ~u);
}
// Asynchronous 4-byte store into the peer CTA's shared memory that completes 4 transaction bytes
// on the peer's mbarrier at the same offset as `local_bar`: data and signal travel together, no
// cluster-scope release fence (which is a MEMBAR.ALL.GPU waiting on every store in flight).
__device__ __forceinline__ uint32_t t_mapa(uint32_t local_addr, uint32_t peer_rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(peer_rank));
  return r;
}
// raddr / rbar: shared::cluster addresses in the peer CTA (t_mapa, computed once)
__device__ __forceinline__ void t_st_async_peer(uint32_t raddr, uint32_t rbar, float v) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(raddr),
               "r"(__float_as_uint(v)), "r"(rbar)
               : "memory");
}

// L2 residency hints.  The stash (finished chains parked in the record's own output rows) is re-read one record
// later by the transposition; without hints the 819 MB input stream evicts part of it first (round 1: 1.87 GB of
// DRAM traffic per 1 024 records against 1.36 GB algorithmic).  Electrode rows are read with evict_first (each
// is used by at most four pairs within a few tiles), stash stores ask for evict_last, final pixels evict_first.
__device__ __forceinline__ uint64_t t_policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t t_policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void t_st_f4_hint(float* ptr, float4 v, uint64_t pol) {
  asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(ptr), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w), "l"(pol)
               : "memory");
}
__device__ __forceinline__ void t_bulk_g2s_hint(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar, uint64_t pol) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar), "l"(pol)
      : "memory");
}

// Wait with an explicit back-off.  c_mbar_wait compiles to TRYWAIT / NANOSLEEP.SYNCS / PHASECHK / BRA and
// a waiting warp goes round that loop every ~20 ns: sixteen epilogue warps in the exchange wait and eight
// build warps waiting for a ring slot executed 40 % of the kernel's warp instructions there.
__device__ __forceinline__ void t_mbar_wait_sleep(uint32_t bar, uint32_t parity, uint32_t ns) {
  asm volatile(
      "{\n\t.reg .pred P1;\n\t"
      "TW_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, %2;\n\t"
      "@P1 bra TD_%=;\n\t"
      "nanosleep.u32 %3;\n\t"
      "bra TW_%=;\n\t"
      "TD_%=:\n\t}\n"
      :
      : "r"(bar), "r"(parity), "r"(0x989680u), "r"(ns)
      : "memory");
}

template <bool kTimeline>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kTThreads, 1) melspec_ts_kernel(const ChainArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char* tab = smem_raw;                                       // Ch | Cl | Sh | Sl
  float* ybuf = reinterpret_cast<float*>(tab + 4 * kCTabBytes);        // kTYFloats
  float* stage = ybuf + 2 * kTYFloats;                                 // [j & 1]{row a | row b} x kCStageFloats (ybuf is double buffered too)
  unsigned char* xt = reinterpret_cast<unsigned char*>(stage + 4 * kCStageFloats);   // [j & 1]{Exh | Oxh}
  unsigned char* xtab = xt + 4 * kTXTileBytes;   // extra frames: {cos hi | sin hi} with row r = bin r - r / 32 (128 rows each)
  __shared__ TsShared sh;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int h = (int)c_cluster_rank(), peer = h ^ 1;
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const int n_items = (int)a.n_items;   // records
  const int n_my_items = n_items > cluster_id ? (n_items - cluster_id + n_clusters - 1) / n_clusters : 0;
  // geometry this kernel is launched for (melspec_ts_run checks it): literals, not runtime divisors
  constexpr int kPpc = 4, tpi = 16, kNMels = 128, kNTime = 2 * kTileFrames;   // tiles per item = 4 chains x 4 pairs
  const int J = n_my_items * tpi;
  if (J == 0) return;   // both CTAs of the cluster agree
#define TL(slot)                                                                                             \
  do {                                                                                                       \
    if (kTimeline && blockIdx.x < 2 && lane == 0 && j >= a.tl_j0 && j < a.tl_j0 + kTlTilesT)                 \
      a.tl[((blockIdx.x * kTlTilesT + (j - a.tl_j0)) * kTlSlotsT) + (slot)] = t_globaltimer();                                  \
  } while (0)

  const uint32_t load_bar0 = c_smem_u32(&sh.load_bar[0]), stage_free = c_smem_u32(&sh.stage_free);
  const uint32_t ops_bar0 = c_smem_u32(&sh.ops_bar[0]), afree0 = c_smem_u32(&sh.afree_bar[0]);
  const uint32_t mma_bar0 = c_smem_u32(&sh.mma_bar[0]), tmem_free0 = c_smem_u32(&sh.tmem_free[0]);
  const uint32_t xch_bar0 = c_smem_u32(&sh.xch_bar[0]);
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(c_smem_u32(&sh.tmem_base)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  if (tid == 32) {
    c_mbar_init(load_bar0, 1);
    c_mbar_init(load_bar0 + 8, 1);
    c_mbar_init(stage_free, 2);
    for (int b = 0; b < 2; ++b) {
      c_mbar_init(ops_bar0 + 8 * b, kTBuildWarps);
      c_mbar_init(c_smem_u32(&sh.y_ready[b]), 2);
      c_mbar_init(c_smem_u32(&sh.y_free[b]), kTBuildWarps);
      c_mbar_init(afree0 + 8 * b, 1);
      c_mbar_init(mma_bar0 + 8 * b, 1);
      c_mbar_init(tmem_free0 + 8 * b, kTEpiWarps);
      c_mbar_init(xch_bar0 + 8 * b, kTEpiWarps);   // + 4 kTEpiWarps transaction bytes from the peer
    }
    for (int i = 0; i < 4; ++i) sh.xmax[i] = 0u;
    asm volatile("fence.mbarrier_init.release.cluster;" ::);
  }
  {
    const uint4* src = reinterpret_cast<const uint4*>(a.tab);
    uint4* dst = reinterpret_cast<uint4*>(tab);
    for (int i = tid; i < 4 * kCTabBytes / 16; i += blockDim.x) dst[i] = __ldg(src + i);
    float4* yz = reinterpret_cast<float4*>(ybuf);
    for (int i = tid; i < 2 * kTYFloats / 4; i += blockDim.x) yz[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    uint4* xz = reinterpret_cast<uint4*>(xt);
    for (int i = tid; i < 4 * kTXTileBytes / 16; i += blockDim.x) xz[i] = make_uint4(0u, 0u, 0u, 0u);
    // Extra-frame operand: the hi tables again with every 32nd row repeating the next warp's first
    // bin (row r holds bin r - r / 32), so that a lane finds its right neighbour's power with one
    // shuffle and no warp needs another warp's lanes.  xw[r]: the taps of the (at most two) mels
    // whose first bin is bin(r); rows 31, 63, 95, 127 only lend their value to lane 30.
    for (int i = tid; i < 2 * 128 * 8; i += blockDim.x) {
      const int t = i >> 10, r = (i >> 3) & 127, k8 = i & 7, b = r - (r >> 5);
      const uint4 v = b < kCBins ? *reinterpret_cast<const uint4*>(a.tab + (size_t)(2 * t) * (kCTabBytes / 2) + c_umma_offset(b, 8 * k8) / 2)
                                 : make_uint4(0u, 0u, 0u, 0u);
      *reinterpret_cast<uint4*>(xtab + t * 16384 + c_umma_offset(r, 8 * k8)) = v;
    }
    if (tid < 128) {
      const int b = tid - (tid >> 5);
      float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
      if ((tid & 31) != 31) {
        int found = 0;
        for (int m = 0; m < a.n_mels; ++m)
          if (a.mel_k0[m] == b) {
            if (found == 0) { w.x = a.mel_w0[m]; w.y = a.mel_w1[m]; }
            else { w.z = a.mel_w0[m]; w.w = a.mel_w1[m]; }
            ++found;
          }
      }
      sh.xw[tid] = w;
    }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::);
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::);
  c_cluster_arrive();   // the peer's mbarriers and maxima are initialised before anyone touches them remotely
  c_cluster_wait();
  const uint32_t tmem_base = sh.tmem_base;

  const int seg_lo = a.seg_lo[h], seg_n = a.seg_n[h];
  const int n_extra = h ? min(a.n_frames - a.n_time, kCMaxExtra) : 0;

  // Clusters walk the chains of a record in rotated order (first chain = cluster_id mod 4): all
  // CTAs run in lockstep, and without the rotation every SM would store a finished chain at the
  // same moment (a 9.5 MB burst into L2 every fourth tile, then nothing).
  const int rot = cluster_id & 3;
  auto pair_of = [&](int j, int& rec) {   // pair index of tile j and its record
    const int q = j / tpi;
    rec = cluster_id + q * n_clusters;
    const int k = j - q * tpi;
    return ((k + rot * kPpc) & (tpi - 1));
  };

  // ======================================================================= single-lane roles
  if (warp >= kTMmaWarp) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");   // 512 x 88 + 256 x 56 + 128 x 40 = the 896 x 72 of the launch
    if (warp == kTMmaWarp) {
      // ---- the tcgen05.mma stream: K-step ks of tile j as soon as its ring slot is written; the
      //      accumulator buffer (j & 1) must have been handed back (tile j - 2 folded)
      // Every operand of the 24 MMAs of a tile is loop invariant except the accumulator buffer:
      // ring slot and barrier parities depend on the K-step only (4 steps per tile), the table
      // descriptors differ in their low word (14-bit address field) only.
      const uint32_t tab0 = c_smem_u32(tab), xt0 = c_smem_u32(xt), xtab0 = c_smem_u32(xtab);
      const uint32_t desc_hi = (uint32_t)(c_umma_desc(0) >> 32), lo_base = (uint32_t)c_umma_desc(0);
      auto dlo = [&](uint32_t addr) { return lo_base | ((addr >> 4) & 0x3FFFu); };
      uint32_t tlo[4][kFold / 16];   // [Ch, Cl, Sh, Sl][K-step]
#pragma unroll
      for (int t = 0; t < 4; ++t)
#pragma unroll
        for (int ks = 0; ks < kFold / 16; ++ks) tlo[t][ks] = dlo(tab0 + t * kCTabBytes + ks * 256);
      const bool leader = t_elect_one();
      bool seen_prev = true;
      auto observe = [&](int j, int where) {   // timeline only: when does the MMA warp first see tile j - 1 complete?
        if (!kTimeline || seen_prev || j < 1) return;
        uint32_t ok;
        asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.u32 %0, 1, 0, P1;\n\t}\n"
                     : "=r"(ok) : "r"(mma_bar0 + 8 * ((j - 1) & 1)), "r"((uint32_t)(((j - 1) >> 1) & 1)) : "memory");
        if (ok) {
          seen_prev = true;
          const int jj = j;
          { const int j = jj - 1; TL(32); if (blockIdx.x < 2 && lane == 0 && j >= a.tl_j0 && j < a.tl_j0 + kTlTilesT) a.tl[((blockIdx.x * kTlTilesT + (j - a.tl_j0)) * kTlSlotsT) + 33] = where; }
        }
      };
      for (int j = 0; j < J; ++j) {
        const uint32_t d_re = tmem_base + (j & 1) * kTAcc, d_im = d_re + kTBins;
        seen_prev = false;
        observe(j, 0);
#pragma unroll
        for (int ks = 0; ks < kFold / 16; ++ks) {
          constexpr int dummy = 0; (void)dummy;
          const int slot = ks & 1;                                  // (4 j + ks) & 1
          c_mbar_wait(ops_bar0 + 8 * slot, (uint32_t)(ks >> 1));    // ((4 j + ks) >> 1) & 1
          observe(j, 1 + 2 * ks);
          if (ks == 0) {
            TL(5);
            if (j >= 2) c_mbar_wait(tmem_free0 + 8 * (j & 1), ((j - 2) >> 1) & 1);
            TL(6);
            observe(j, 10);
          }
          TL(24 + ks);
          asm volatile("tcgen05.fence::after_thread_sync;" ::);
          if (leader) {
            const uint32_t ring = tmem_base + kTRingCol + 32 * slot;   // Eh +0, El +8, Oh +16, Ol +24
            t_umma_ts(d_re, ring + 0, tlo[0][ks], desc_hi, ks > 0);
            t_umma_ts(d_im, ring + 16, tlo[2][ks], desc_hi, ks > 0);
            t_umma_ts(d_re, ring + 8, tlo[0][ks], desc_hi, 1);
            t_umma_ts(d_im, ring + 24, tlo[2][ks], desc_hi, 1);
            t_umma_ts(d_re, ring + 0, tlo[1][ks], desc_hi, 1);
            t_umma_ts(d_im, ring + 16, tlo[3][ks], desc_hi, 1);
            t_commit(afree0 + 8 * slot);
            if (ks == kFold / 16 - 1) {
              if (n_extra > 0) {
                // D'[bin][frame] = table (A, shared memory, M = 128) x extra frames (B, N = 8), hi terms only
                const uint32_t dx = tmem_base + kTXCol + (j & 1) * 2 * kTXN;
                const uint32_t xb = xt0 + (j & 1) * 2 * kTXTileBytes;
#pragma unroll
                for (int k2 = 0; k2 < kFold / 16; ++k2) {
                  c_umma(dx, ((uint64_t)desc_hi << 32) | dlo(xtab0 + k2 * 256), ((uint64_t)desc_hi << 32) | dlo(xb + k2 * 256),
                         k2 > 0, kTIdescX);
                  c_umma(dx + kTXN, ((uint64_t)desc_hi << 32) | dlo(xtab0 + 16384 + k2 * 256),
                         ((uint64_t)desc_hi << 32) | dlo(xb + kTXTileBytes + k2 * 256), k2 > 0, kTIdescX);
                }
              }
              t_commit(mma_bar0 + 8 * (j & 1));
            }
          }
          __syncwarp();
          TL(28 + ks);
          observe(j, 2 + 2 * ks);
        }
        TL(7);
      }
    } else if (warp == kTLoadWarp) {
      // ---- TMA prefetch: both electrode rows of pair j + 1 once the difference of pair j has
      //      consumed the staging buffers
      if (lane == 0) {
        const uint64_t pol_stream = t_policy_evict_first();
        auto issue_load = [&](int j) {
          int rec;
          const int pr = pair_of(j, rec);
          const float* base = a.eeg + rec * a.rec_stride + a.crop_off + seg_lo;
          const uint32_t bytes = (uint32_t)seg_n * 4u;
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          const uint32_t load_bar = load_bar0 + 8 * (j & 1);
          float* st = stage + (j & 1) * 2 * kCStageFloats;
          c_mbar_expect_tx(load_bar, 2 * bytes);
          t_bulk_g2s_hint(c_smem_u32(st), base + a.pair_a[pr] * a.ch_stride, bytes, load_bar, pol_stream);
          t_bulk_g2s_hint(c_smem_u32(st + kCStageFloats), base + a.pair_b[pr] * a.ch_stride, bytes, load_bar, pol_stream);
        };
        // Stagger the clusters by a quarter of a chain each: all CTAs do identical work and would
        // otherwise store their finished chains in the same instant (L2 write bursts).
        for (int k = 0; k < (cluster_id & 3); ++k) asm volatile("nanosleep.u32 %0;" ::"r"(a.stagger_ns));
        // two staging buffers: the rows of pair j + 2 are requested as soon as the difference of pair j
        // has left its buffer, a whole tile before they are needed (with one buffer the ~0.9 us of the
        // copy sat on the difference warps' loop, which sets the kernel's period)
        issue_load(0);
        if (J > 1) issue_load(1);
        for (int j = 0; j + 2 < J; ++j) {
          c_mbar_wait(stage_free, j & 1);
          issue_load(j + 2);
        }
      }
    } else {
      // ---- difference warps (2 x 32 threads): electrode difference of pair j into y buffer
      //      j & 1, max|y| (NaN if any NaN), NaN fill, reflect margins -- one tile ahead of the
      //      build group, which only reads y
      const int dtid = tid - 32 * kTDiffWarp0, dw = warp - kTDiffWarp0;
      const uint32_t y_ready0 = c_smem_u32(&sh.y_ready[0]), y_free0 = c_smem_u32(&sh.y_free[0]);
      auto pair_max_nan = [&](float v) {   // over both warps
        v = warp_max_nan(v);
        if (lane == 0) sh.red[dw] = v;
        c_named_bar(7, kTDiff);
        const float r = max_nan(sh.red[0], sh.red[1]);
        c_named_bar(7, kTDiff);
        return r;
      };
      auto pair_sum = [&](float v) {
        v = warp_sum(v);
        if (lane == 0) sh.red[dw] = v;
        c_named_bar(7, kTDiff);
        const float r = sh.red[0] + sh.red[1];
        c_named_bar(7, kTDiff);
        return r;
      };
      for (int j = 0; j < J; ++j) {
        float* yb = ybuf + (j & 1) * kTYFloats;
        c_mbar_wait(load_bar0 + 8 * (j & 1), (j >> 1) & 1);
        if (dw == 0) TL(0);
        if (j >= 2) c_mbar_wait(y_free0 + 8 * (j & 1), ((j - 2) >> 1) & 1);   // the build group has read pair j - 2
        if (dw == 0) TL(1);
        float amax = 0.f;
        {
          const float4* sa = reinterpret_cast<const float4*>(stage + (j & 1) * 2 * kCStageFloats);
          const float4* sb = sa + kCStageFloats / 4;
          float4* yd = reinterpret_cast<float4*>(yb + kCMargin);
#pragma unroll 4
          for (int i = dtid; i < (seg_n >> 2); i += kTDiff) {
            const float4 x = sa[i], y = sb[i];
            const float4 d = make_float4(x.x - y.x, x.y - y.y, x.z - y.z, x.w - y.w);
            yd[i] = d;
            amax = max_nan(max_nan(amax, fabsf(d.x)), max_nan(fabsf(d.y), max_nan(fabsf(d.z), fabsf(d.w))));
          }
          __syncwarp();
          if (lane == 0) c_mbar_arrive(stage_free);   // the loader may fetch pair j + 1
        }
        amax = pair_max_nan(amax);
        if (amax != amax) {
          // NaN samples (rare).  cfg_utils.py:380-385 on the difference: nanmean over the whole
          // crop, all-NaN -> 0.  Each CTA re-reads both full rows, so no exchange is needed and
          // both halves get the identical fill value.
          int rec;
          const int pr = pair_of(j, rec);
          const float* ra = a.eeg + rec * a.rec_stride + a.pair_a[pr] * a.ch_stride + a.crop_off;
          const float* rb = a.eeg + rec * a.rec_stride + a.pair_b[pr] * a.ch_stride + a.crop_off;
          float s_ = 0.f, c_ = 0.f;
          for (int i = dtid; i < a.crop; i += kTDiff) {
            const float d = __ldg(ra + i) - __ldg(rb + i);
            if (d != d) c_ += 1.f; else s_ += d;
          }
          s_ = pair_sum(s_);
          c_ = pair_sum(c_);
          const bool all_nan = c_ >= (float)a.crop;
          const float fill = all_nan ? 0.f : s_ / ((float)a.crop - c_);
          float m = 0.f;
          for (int i = dtid; i < seg_n; i += kTDiff) {
            float d = yb[kCMargin + i];
            if (all_nan) d = 0.f; else if (d != d) d = fill;
            yb[kCMargin + i] = d;
            m = fmaxf(m, fabsf(d));
          }
          amax = pair_max_nan(m);
        }
        if (a.pad_mode == 1) {   // reflect padding (librosa < 0.10); zero padding is the resident state
          for (int k = dtid; k < 2 * kCMargin + 128; k += kTDiff) {
            const int idx = k < kCMargin ? k : seg_n + k;          // left margin | right margin
            const int i = seg_lo + idx - kCMargin;                 // crop coordinate
            const int src = i < 0 ? -i : (i >= a.crop ? 2 * (a.crop - 1) - i : -1);
            const int sidx = src - seg_lo;
            if (idx < kTYFloats) yb[idx] = (src >= 0 && sidx >= 0 && sidx < seg_n) ? yb[kCMargin + sidx] : 0.f;
          }
        }
        if (dtid == 0) sh.amax_fin[j & 1] = amax;
        __syncwarp();
        if (lane == 0) c_mbar_arrive(y_ready0 + 8 * (j & 1));
        if (dw == 0) TL(2);
      }
    }
  } else if (warp >= kTEpiWarps) {
    // ================================================================================ build group
    // thread = (frame t = TMEM lane, half p of a K-step): difference + statistics of pair j, then
    // its fp16 hi/lo operand K-step by K-step into the tensor-memory ring.
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
    const int btid = tid - kTEpi, bwarp = warp - kTEpiWarps;
    const int t_frame = btid & (kTileFrames - 1), part = btid >> 7;
    const uint32_t ring_addr = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16) + kTRingCol + 4 * part;
    const uint32_t y_ready0 = c_smem_u32(&sh.y_ready[0]), y_free0 = c_smem_u32(&sh.y_free[0]);

    for (int j = 0; j < J; ++j) {
      c_mbar_wait(y_ready0 + 8 * (j & 1), (j >> 1) & 1);   // y buffer j & 1 and its max|y| are final
      const float* yb = ybuf + (j & 1) * kTYFloats;
      const float amax = sh.amax_fin[j & 1];
      int e_j = 0;   // exact power-of-two pre-scale: amax * 2^e in [2^10, 2^11)
      {
        const int ef = (int)((__float_as_uint(amax) >> 23) & 0xffu);   // biased exponent; 0 = zero / subnormal, 255 = inf / NaN
        if (ef > 0 && ef < 255) e_j = 137 - ef;                        // 11 - (ef - 126)
        e_j = max(-60, min(60, e_j));
      }
      const float s = __int_as_float((127 + e_j) << 23);
      if (btid == 0) sh.e_tile[j & 3] = e_j;

      const float* yc = yb + kCMargin + (128 * h + t_frame) * a.hop - seg_lo;
#pragma unroll 1
      for (int ks = 0; ks < kFold / 16; ++ks) {
        const int g = 4 * j + ks, slot = ks & 1;   // slot = g & 1
        // E = s (y[c+m] + y[c-m]), O = s (y[c+m] - y[c-m]) for 8 fold indices, as fp16 hi/lo:
        // computed while the MMAs of step g - 2 still read the ring slot
        uint4 eh, el, oh, ol;
        {
          const int m0 = 16 * ks + 8 * part;
          float ev[8], ov[8];
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float yp = yc[m0 + q], yms = yc[-(m0 + q)] * s;
            ev[q] = fmaf(yp, s, yms);
            ov[q] = fmaf(yp, s, -yms);
          }
          c_split2(ev[0], ev[1], eh.x, el.x); c_split2(ev[2], ev[3], eh.y, el.y);
          c_split2(ev[4], ev[5], eh.z, el.z); c_split2(ev[6], ev[7], eh.w, el.w);
          c_split2(ov[0], ov[1], oh.x, ol.x); c_split2(ov[2], ov[3], oh.y, ol.y);
          c_split2(ov[4], ov[5], oh.z, ol.z); c_split2(ov[6], ov[7], oh.w, ol.w);
        }
        if (ks == kFold / 16 - 1) {   // last read of y buffer j & 1 (the extra frames were read at ks == 0)
          __syncwarp();
          if (lane == 0) c_mbar_arrive(y_free0 + 8 * (j & 1));
        }
        if (ks >= 2 && bwarp == 0) TL(16 + 4 * (ks - 2));
        if (g >= 2) t_mbar_wait_sleep(afree0 + 8 * slot, (uint32_t)(((ks >> 1) + 1) & 1), 32);   // ((g - 2) >> 1) & 1: the MMAs of step g - 2 have read the slot
        if (ks == 0 && bwarp == 0) TL(3);
        if (ks >= 2 && bwarp == 0) TL(17 + 4 * (ks - 2));
        asm volatile("tcgen05.fence::after_thread_sync;" ::);
        if (ks == 0 && btid < n_extra * kFold) {
          // frames beyond the kept width (they only feed the pair maximum): rows of a small B
          // operand in shared memory, hi halves only; the MMAs of tile j - 2 that read this
          // buffer completed before the commit waited for above.
          const int x = btid >> 6, m = btid & (kFold - 1);
          const float* yx = yb + kCMargin + (a.n_time + x) * a.hop - seg_lo;
          const float yp = yx[m], yms = yx[-m] * s;
          const float e = fmaf(yp, s, yms), o = fmaf(yp, s, -yms);
          const int off = (j & 1) * 2 * kTXTileBytes + c_umma_offset(x, m);
          *reinterpret_cast<__half*>(xt + off) = __float2half_rn(e);
          *reinterpret_cast<__half*>(xt + kTXTileBytes + off) = __float2half_rn(o);
        }
        {
          const uint32_t ra = ring_addr + 32 * slot;
          t_tmem_st4(ra + 0, eh.x, eh.y, eh.z, eh.w);
          t_tmem_st4(ra + 8, el.x, el.y, el.z, el.w);
          t_tmem_st4(ra + 16, oh.x, oh.y, oh.z, oh.w);
          t_tmem_st4(ra + 24, ol.x, ol.y, ol.z, ol.w);
        }
        c_tmem_wait_st();
        if (ks >= 2 && bwarp == 0) TL(18 + 4 * (ks - 2));
        asm volatile("tcgen05.fence::before_thread_sync;" ::);
        if (ks == kFold / 16 - 1) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) c_mbar_arrive(ops_bar0 + 8 * slot);
        if (ks >= 2 && bwarp == 0) TL(19 + 4 * (ks - 2));
      }
      if (bwarp == 0) TL(4);
    }
  } else {
    // ============================================================================= epilogue group
    // thread = (frame t = TMEM lane, part q): owns mels [32 q, 32 q + 32) of frame 128 h + t,
    // keeps their chain sum acc[] in registers and parks the log-mel values of the tile in
    // flight in accumulator columns only it reads.
    asm volatile("setmaxnreg.inc.sync.aligned.u32 88;");    // launch: 72 x 896
    const int t_frame = tid & (kTileFrames - 1), part = tid >> 7;
    const uint32_t qcol = (uint32_t)(MELGEN_COL0_P4_1 * part);
    static_assert(MELGEN_COL0_P4_2 == 2 * MELGEN_COL0_P4_1 && MELGEN_COL0_P4_3 == 3 * MELGEN_COL0_P4_1, "part pitch");
    auto park_col = [](int g) { return (uint32_t)(8 * (1 + g / 2) + (g & 1) * kTBins); };
    const uint32_t lane_base = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16);
    const uint64_t pol_keep = t_policy_evict_last(), pol_done = t_policy_evict_first();
    const uint32_t peer_pv0 = t_mapa(c_smem_u32(&sh.pv[0][warp]), (uint32_t)peer), peer_xch0 = t_mapa(xch_bar0, (uint32_t)peer);
    float acc[kTMels];
#pragma unroll
    for (int i = 0; i < kTMels; ++i) acc[i] = 0.f;
    float csum = 0.f;   // sum over the chain's pairs of (reference + 2 e)
    int e_prev = 0;

    // exchange n is complete: fold the parked log values of tile n into the chain sum with the
    // pair's clip level, then hand accumulator buffer n & 1 back to the MMA warp.
    auto fold_tile = [&](int n) {
      t_mbar_wait_sleep(xch_bar0 + 8 * (n & 1), (uint32_t)((n >> 1) & 1), 64);
      { const int j = n; if (warp == 0) TL(9); }
      float r_pair = fmaxf(t_ord2f(sh.xmax[n & 3]), a.log2_amin);
      {
        const float4* pv = reinterpret_cast<const float4*>(sh.pv[n & 1]);
#pragma unroll
        for (int i = 0; i < kTEpiWarps / 4; ++i) {
          const float4 v = pv[i];
          r_pair = fmaxf(fmaxf(r_pair, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
        }
      }
      const float off = 2.f * (float)e_prev;          // e_prev = pre-scale exponent of tile n
      const float t_clip = fmaxf(a.log2_amin, r_pair - a.top_log2) + off;
      csum += r_pair + off;
      const uint32_t qaddr = lane_base + (n & 1) * kTAcc + qcol;
#pragma unroll
      for (int g = 0; g < kTMels / 16; ++g) {
        float q[16];
        t_tmem_ld8x2(qaddr + park_col(2 * g), qaddr + park_col(2 * g + 1), q);
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[16 * g + i] += fmaxf(q[i], t_clip);
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::);
      __syncwarp();
      if (lane == 0) c_mbar_arrive(tmem_free0 + 8 * (n & 1));
      if (tid == 0) sh.xmax[(n + 2) & 3] = 0u;   // every warp has read that slot (see the ordering note at the exchange)
    };
    // Output rows of this thread: mel rows [32 part, 32 part + 32) of the record, frames of this
    // half: one 2 KB segment (128 frames x 4 chains) per mel row.
    auto row_seg = [&](int rec) {
      return a.out + (((long long)rec * kNMels + kTMels * part) * kNTime + 128 * h) * 4;
    };
    constexpr long long mstride = (long long)kNTime * 4;
    // The chain whose last tile is j_last is complete.  It is stashed in the record's own output
    // rows, blocked so that every store is a full-line 16-byte-per-lane store: the segment of
    // row 4 g + c holds, for chain c, [frame][mels 4 g .. 4 g + 3] (L2 resident until the
    // transposition below rewrites the same rows).
    auto chain_done = [&](int j_last) {
      int rec;
      const int pr = pair_of(j_last, rec);
      float* o = row_seg(rec) + (pr / kPpc) * mstride + 4 * t_frame;
#pragma unroll
      for (int g = 0; g < kTMels / 4; ++g) {
        float4 v;
        v.x = fmaf(a.out_gain, acc[4 * g + 0] - csum, a.out_bias);
        v.y = fmaf(a.out_gain, acc[4 * g + 1] - csum, a.out_bias);
        v.z = fmaf(a.out_gain, acc[4 * g + 2] - csum, a.out_bias);
        v.w = fmaf(a.out_gain, acc[4 * g + 3] - csum, a.out_bias);
        t_st_f4_hint(o + 4 * g * mstride, v, pol_keep);
      }
    };
    // A finished record is turned into the final [frame][chain] pixels one group of four mel
    // rows at a time while the next record is computed: the four stash segments of the group
    // are loaded before the fold (whose exchange wait and tensor-memory reads cover the L2
    // latency), the 128 threads of the part meet (every stash value of the group has been
    // read), and each thread writes one float4 pixel per row -- a 4 x 4 transpose in registers.
    float4 tv[4];
    auto transpose_load = [&](int rec, int g) {
      const float* r = row_seg(rec) + 4 * g * mstride + 4 * t_frame;
#pragma unroll
      for (int c = 0; c < 4; ++c) tv[c] = __ldcg(reinterpret_cast<const float4*>(r + c * mstride));
    };
    auto transpose_store = [&](int rec, int g) {
      float* r = row_seg(rec) + 4 * g * mstride + 4 * t_frame;
      // no barrier: the four slots read and the four slots written are this thread's own (chain_done
      // and this function address rows 4 g .. 4 g + 3 at the same 16 bytes, 4 t_frame)
      t_st_f4_hint(r + 0 * mstride, make_float4(tv[0].x, tv[1].x, tv[2].x, tv[3].x), pol_done);
      t_st_f4_hint(r + 1 * mstride, make_float4(tv[0].y, tv[1].y, tv[2].y, tv[3].y), pol_done);
      t_st_f4_hint(r + 2 * mstride, make_float4(tv[0].z, tv[1].z, tv[2].z, tv[3].z), pol_done);
      t_st_f4_hint(r + 3 * mstride, make_float4(tv[0].w, tv[1].w, tv[2].w, tv[3].w), pol_done);
    };

    // Per tile: main (power, mel, lg2, park) -> contribute the maximum -> wait for the exchange
    // -> fold the parked values and hand the accumulator buffer back.  The fold comes LAST: the
    // MMAs of tile j + 2 wait for this buffer while those of tile j + 1 already run in the other
    // one, so the tensor pipe and the operand build overlap the whole epilogue of tile j
    // (folding at the start of the next tile's epilogue put the exchange latency and the fold on
    // the critical loop: period = fold + MMA chain instead of max(epilogue, MMA chain)).
    for (int j = 0; j < J; ++j) {
      c_mbar_wait(mma_bar0 + 8 * (j & 1), (j >> 1) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::);
      const int e_j = sh.e_tile[j & 3];
      if (warp == 0) TL(8);
      const uint32_t lane_addr = lane_base + (j & 1) * kTAcc;
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
#define CH_LOAD(c) c_tmem_ld8x2(lane_addr + col0 + 8 * (c), lane_addr + kTBins + col0 + 8 * (c), re, im);
#define CH_PART(TAG)                                                \
  {                                                                 \
    constexpr int col0 = MELGEN_COL0_##TAG;                         \
    MELGEN_EPILOGUE_##TAG(CH_LOAD, CH_POW, CH_MEL1, CH_MEL2)        \
  }
      if (part == 0) CH_PART(P4_0) else if (part == 1) CH_PART(P4_1) else if (part == 2) CH_PART(P4_2) else CH_PART(P4_3)
      CH_PARK(kTMels / 8 - 1)
      c_tmem_wait_st();
      if (warp == 0) TL(12);
      if (warp == kTEpiWarps - 1) TL(14);
#undef CH_PART
#undef CH_LOAD
#undef CH_POW
#undef CH_MEL_TAIL
#undef CH_PARK
#undef CH_MEL1
#undef CH_MEL2

      // extra frames: lane = table row (bin r - r / 32), column = frame; the right neighbour's power
      // comes from one shuffle, the taps from xw: no shared-memory round trip, no barrier (warps 0-3)
      if (n_extra > 0 && part == 0) {
        const uint32_t xaddr = tmem_base + ((uint32_t)(32 * warp) << 16) + kTXCol + (j & 1) * 2 * kTXN;
        float xr[4], xi[4];
        c_tmem_ld4(xaddr, xr);
        c_tmem_ld4(xaddr + kTXN, xi);
        const float4 w = sh.xw[tid];
#pragma unroll
        for (int x = 0; x < kCMaxExtra; ++x) {   // unused frames are zero rows of the operand: power 0
          const float pw = fmaf(xr[x], xr[x], xi[x] * xi[x]);
          const float pn = __shfl_down_sync(0xffffffffu, pw, 1);
          vmax = fmaxf(vmax, fmaxf(fmaf(w.y, pn, w.x * pw), fmaf(w.w, pn, w.z * pw)));
        }
      }
      vmax = warp_max_nan(vmax);
      // Exchange of the pair maximum.  Local: atomic max into slot j & 3 + arrive on the local
      // exchange barrier.  Remote: the warp's value goes into the peer's pv[j & 1][warp] with an
      // asynchronous store that completes 4 of the 64 transaction bytes warp 0 of the peer
      // expects on its barrier.  Slot (n + 2) & 3 is reset during fold n (fold_tile): a
      // contribution to it can only come from a warp that has seen exchange n + 1 complete,
      // which needs warp 0's arrival, which follows its reset in program order.  pv[n & 1] is
      // rewritten by the peer for tile n + 2 only after it has seen exchange n + 1 complete,
      // i.e. after every warp here has finished fold n.
      if (lane == 0) {
        const float v = c_lg2(vmax) - 2.f * (float)e_j;
        const uint32_t xb = xch_bar0 + 8 * (j & 1);
        atomicMax(&sh.xmax[j & 3], t_f2ord(v));
        t_st_async_peer(peer_pv0 + (j & 1) * (uint32_t)sizeof(sh.pv[0]), peer_xch0 + 8 * (j & 1), v);
        if (warp == 0) c_mbar_expect_tx(xb, 4 * kTEpiWarps);
        else c_mbar_arrive(xb);
      }
      __syncwarp();
      if (warp == 0) TL(13);
      if (warp == kTEpiWarps - 1) TL(15);
      e_prev = e_j;
      // transposition schedule: record number q (of this cluster) is completely stashed by the
      // end of tile 16 q + 15; its group g is handled in tile 16 (q + 1) + 2 g + 1
      const int q_tr = j / tpi - 1, step = j & (tpi - 1);
      const bool tr_now = q_tr >= 0 && (step & 1);
      const int rec_prev = cluster_id + q_tr * n_clusters;
      if (tr_now) transpose_load(rec_prev, step >> 1);   // L2 latency covered by the exchange wait and the fold
      // The chain that ended with tile j - 1 is stashed HERE, inside the exchange wait of tile j: its
      // 64 KB of stores per CTA run into the SM's store throttle (~1 100 ns) and would sit on the
      // critical path at the end of tile j - 1; the accumulators are not touched before fold j.
      if (j > 0 && j % kPpc == 0) {
        chain_done(j - 1);
#pragma unroll
        for (int i = 0; i < kTMels; ++i) acc[i] = 0.f;
        csum = 0.f;
      }
      if (tr_now) transpose_store(rec_prev, step >> 1);
      fold_tile(j);
      if (warp == 0) TL(10);
      if (warp == 0) TL(11);
    }
    if (J > 0) chain_done(J - 1);
    // ---- the last record of this cluster: all groups at once
    {
      const int rec_last = cluster_id + (n_my_items - 1) * n_clusters;
      for (int g = 0; g < kTMels / 4; ++g) {
        transpose_load(rec_last, g);
        transpose_store(rec_last, g);
      }
    }
  }
#undef TL

  asm volatile("tcgen05.fence::before_thread_sync;" ::);
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  c_cluster_arrive();   // nobody exits while its peer may still write or arrive remotely
  c_cluster_wait();
}

// ------------------------------------------------------------------------------------- host
size_t melspec_ts_smem_bytes() {
  return 4 * (size_t)kCTabBytes + sizeof(float) * (2 * (size_t)kTYFloats + 4 * (size_t)kCStageFloats) + 4 * (size_t)kTXTileBytes +
         2 * 16384;   // remapped hi tables of the extra frames
}

// Launches the TS kernel on `clusters` clusters of two CTAs; `a.n_items` must be the number of
// RECORDS (the kernel walks all chains of a record).  Geometry checks are the caller's
// (melspec_chain_launch, below); this form needs 4 chains of 4 pairs, 128 mels and 256 kept frames.
int melspec_ts_run(const ChainArgs& a, long long clusters, cudaStream_t st) {
  if (a.n_chains != 4 || a.ppc != 4 || a.n_mels != 128 || a.n_time != 2 * kTileFrames) return EEGSPEC_EUNSUPPORTED;
  static std::mutex mu;
  static std::atomic<bool> configured[64];   // idempotent set-up; atomic so that concurrent callers do not race on the flag
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return EEGSPEC_EUNSUPPORTED;
  const size_t smem = melspec_ts_smem_bytes();
  std::lock_guard<std::mutex> lock(mu);
  if (!configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute(melspec_ts_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
#ifdef EEGSPEC_TIMELINE
    e = cudaFuncSetAttribute(melspec_ts_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
#endif
    configured[dev] = true;
  }
#ifdef EEGSPEC_TIMELINE
  // EEGSPEC_TIMELINE=<file>: synchronous launch that dumps clock stamps of the pipeline hand-offs of
  // CTAs 0 and 1 (64 tiles from EEGSPEC_TIMELINE_J0) as raw int64 (%globaltimer nanoseconds).
  if (const char* tl_path = getenv("EEGSPEC_TIMELINE")) {
    ChainArgs b = a;
    const size_t tl_bytes = sizeof(long long) * 2 * kTlTilesT * kTlSlotsT;
    const char* j0 = getenv("EEGSPEC_TIMELINE_J0");
    b.tl_j0 = j0 ? atoi(j0) : 0;
    if (cudaMalloc(&b.tl, tl_bytes) != cudaSuccess) return EEGSPEC_EWORKSPACE;
    cudaMemsetAsync(b.tl, 0, tl_bytes, st);
    melspec_ts_kernel<true><<<(unsigned)(2 * clusters), kTThreads, smem, st>>>(b);
    std::vector<long long> hbuf(2 * kTlTilesT * kTlSlotsT);
    cudaStreamSynchronize(st);
    cudaMemcpy(hbuf.data(), b.tl, tl_bytes, cudaMemcpyDeviceToHost);
    cudaFree(b.tl);
    if (FILE* f = fopen(tl_path, "wb")) { fwrite(hbuf.data(), 1, tl_bytes, f); fclose(f); }
    return (int)cudaGetLastError();
  }
#endif
  melspec_ts_kernel<false><<<(unsigned)(2 * clusters), kTThreads, smem, st>>>(a);
  return (int)cudaGetLastError();
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

  static std::mutex mu;
  static int sm_count[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return EEGSPEC_EUNSUPPORTED;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (!sm_count[dev]) {
      const cudaError_t e = cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
      if (e != cudaSuccess) return (int)e;
    }
  }
  if (a.n_chains != 4 || a.ppc != 4 || a.n_mels != 128) return EEGSPEC_EUNSUPPORTED;
  // work item = record: one cluster of two CTAs per SM pair walks records cluster_id, cluster_id + n, ...
  a.n_items = batch;
  a.stagger_ns = 2500;   // start delay per quarter chain: spreads the clusters' stash bursts
  long long clusters = sm_count[dev] / 2;
  if (clusters < 1) clusters = 1;
  if (clusters > batch) clusters = batch;
  if ((batch + clusters - 1) / clusters * 16 > (1LL << 28)) return EEGSPEC_EUNSUPPORTED;
  const int rc = melspec_ts_run(a, clusters, st);
  if (rc == 0) count_launch();
  return rc;
}

}  // namespace eegspec
