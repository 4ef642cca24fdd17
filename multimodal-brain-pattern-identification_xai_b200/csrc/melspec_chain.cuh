// Shared pieces of the mode-N engine-3 kernel (melspec_ts.cu; eegtransform.cu borrows the PTX
// wrappers): geometry constants, the argument block and the PTX wrappers (tcgen05 / TMEM, mbarrier,
// bulk copy, cluster / DSMEM).
#pragma once
#include <cuda_fp16.h>

#include "melspec_common.cuh"

namespace eegspec {

constexpr int kCBins = 112;                         // MMA N
constexpr int kCTabBytes = kCBins * kFold * 2;      // 14 336
constexpr int kCTileBytes = kTileFrames * kFold * 2;  // 16 384
constexpr int kCMargin = 64;                        // padding samples either side of y
constexpr int kCStageFloats = 5376;                 // one staged electrode row segment
constexpr int kCYFloats = kCStageFloats + 2 * kCMargin + 128;
constexpr int kCMaxExtra = 4;                       // frames beyond the kept width (max only)
constexpr int kCAccCols = 2 * kCBins;               // Re | Im
constexpr int kCXCols = 16;                         // MMA N of the extra-frame product
constexpr int kCXTileBytes = kCXCols * kFold * 2;   // 2 048: 16 rows x 64 fp16
constexpr int kCXTmemCol = 2 * kCAccCols;           // 448: [buffer][Re | Im][16] in the spare columns
constexpr float kLog2To10Log10 = 3.0102999566398120f;

struct ChainArgs {
  const float* eeg;
  float* out;
  const unsigned short* tab;   // window-folded fp16 {cos hi, cos lo, sin hi, sin lo}, UMMA layout
  const int* mel_k0;           // runtime taps, used for the extra frames only
  const float* mel_w0;
  const float* mel_w1;
  long long rec_stride, ch_stride;
  long long n_items;           // batch * n_chains
  int crop, crop_off, hop, n_frames, n_mels, n_time, n_chains, ppc, pad_mode;
  int seg_lo[2], seg_n[2];     // staged sample range of each half (multiples of 4)
  float log2_amin, top_log2, out_bias, out_gain;
  int pair_a[EEGSPEC_MAX_PAIRS], pair_b[EEGSPEC_MAX_PAIRS];
  long long* tl;               // debug timeline (builds with -DEEGSPEC_TIMELINE only), normally null
  int tl_j0;
  unsigned int stagger_ns;     // start delay per quarter-chain of stagger
};

// ------------------------------------------------------------------------------ PTX helpers
__device__ __forceinline__ uint32_t c_smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ uint64_t c_umma_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)(128 >> 4) << 16;   // LBO: next K core matrix
  d |= (uint64_t)(1024 >> 4) << 32;  // SBO: next 8-row group
  d |= (uint64_t)1 << 46;            // sm_100 descriptor version
  return d;
}
__host__ __device__ constexpr int c_umma_offset(int row, int k) {
  return (row >> 3) * 1024 + (k >> 3) * 128 + (row & 7) * 16 + (k & 7) * 2;
}
constexpr uint32_t kCIdesc = (1u << 4) | ((uint32_t)(kCBins >> 3) << 17) | ((uint32_t)(kTileFrames >> 4) << 24);
// extra frames: D'[bin][frame] = table (A, M = 128 rows of which 112 are table) x frames (B, N = 16)
constexpr uint32_t kCIdescX = (1u << 4) | ((uint32_t)(kCXCols >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);

__device__ __forceinline__ void c_umma(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t acc,
                                       uint32_t idesc = kCIdesc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
      :
      : "r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc)
      : "memory");
}
// 4 TMEM columns of this thread's lane.
__device__ __forceinline__ void c_tmem_ld4(uint32_t addr, float (&v)[4]) {
  uint32_t a[4];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];\n"
               : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3])
               : "r"(addr));
  asm volatile("tcgen05.wait::ld.sync.aligned;\n" : "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]) : : "memory");
#pragma unroll
  for (int i = 0; i < 4; ++i) v[i] = __uint_as_float(a[i]);
}
__device__ __forceinline__ void c_mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
// The suspend-time hint lets the hardware park the waiting warp until the phase completes instead
// of spinning: spinning waiters steal issue slots from the single-lane MMA / loader warps that
// share their scheduler (measured: 6 MMAs took 800-1400 cycles to ISSUE next to hot spin loops).
__device__ __forceinline__ void c_mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred P1;\n\t"
      "CW_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, %2;\n\t"
      "@!P1 bra CW_%=;\n\t}\n"
      :
      : "r"(bar), "r"(parity), "r"(0x989680u)
      : "memory");
}
__device__ __forceinline__ void c_mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// 1-D bulk copy global -> shared, completion counted on an mbarrier (TMA engine).
__device__ __forceinline__ void c_bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ uint32_t c_cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void c_cluster_arrive() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
}
__device__ __forceinline__ void c_cluster_arrive_relaxed() {
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
}
__device__ __forceinline__ void c_cluster_wait() {
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void c_st_peer_f32(uint32_t local_addr, uint32_t peer_rank, float v) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(peer_rank));
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote), "f"(v) : "memory");
}
__device__ __forceinline__ float c_lg2(float x) {
  float r;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
// 16 TMEM columns of this thread's lane for Re and for Im, and the wait that makes them valid.
__device__ __forceinline__ void c_tmem_ld16x2(uint32_t addr_re, uint32_t addr_im, float (&re)[16], float (&im)[16]) {
  uint32_t a[16], b[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
      : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3]), "=r"(a[4]), "=r"(a[5]), "=r"(a[6]), "=r"(a[7]), "=r"(a[8]),
        "=r"(a[9]), "=r"(a[10]), "=r"(a[11]), "=r"(a[12]), "=r"(a[13]), "=r"(a[14]), "=r"(a[15])
      : "r"(addr_re));
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
      : "=r"(b[0]), "=r"(b[1]), "=r"(b[2]), "=r"(b[3]), "=r"(b[4]), "=r"(b[5]), "=r"(b[6]), "=r"(b[7]), "=r"(b[8]),
        "=r"(b[9]), "=r"(b[10]), "=r"(b[11]), "=r"(b[12]), "=r"(b[13]), "=r"(b[14]), "=r"(b[15])
      : "r"(addr_im));
  asm volatile("tcgen05.wait::ld.sync.aligned;\n"
               : "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(a[4]), "+r"(a[5]), "+r"(a[6]), "+r"(a[7]),
                 "+r"(a[8]), "+r"(a[9]), "+r"(a[10]), "+r"(a[11]), "+r"(a[12]), "+r"(a[13]), "+r"(a[14]), "+r"(a[15]),
                 "+r"(b[0]), "+r"(b[1]), "+r"(b[2]), "+r"(b[3]), "+r"(b[4]), "+r"(b[5]), "+r"(b[6]), "+r"(b[7]),
                 "+r"(b[8]), "+r"(b[9]), "+r"(b[10]), "+r"(b[11]), "+r"(b[12]), "+r"(b[13]), "+r"(b[14]), "+r"(b[15])
               :
               : "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) { re[i] = __uint_as_float(a[i]); im[i] = __uint_as_float(b[i]); }
}
// 8 TMEM columns of this thread's lane for Re and for Im.
__device__ __forceinline__ void c_tmem_ld8x2(uint32_t addr_re, uint32_t addr_im, float (&re)[8], float (&im)[8]) {
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
// 8 TMEM columns of this thread's lane: store (issue only) / load + wait.
__device__ __forceinline__ void c_tmem_st8(uint32_t addr, float v0, float v1, float v2, float v3, float v4, float v5,
                                           float v6, float v7) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};\n" ::"r"(addr),
               "r"(__float_as_uint(v0)), "r"(__float_as_uint(v1)), "r"(__float_as_uint(v2)), "r"(__float_as_uint(v3)),
               "r"(__float_as_uint(v4)), "r"(__float_as_uint(v5)), "r"(__float_as_uint(v6)), "r"(__float_as_uint(v7))
               : "memory");
}
__device__ __forceinline__ void c_tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory"); }
__device__ __forceinline__ void c_tmem_ld8(uint32_t addr, float (&v)[8]) {
  uint32_t a[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
               : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3]), "=r"(a[4]), "=r"(a[5]), "=r"(a[6]), "=r"(a[7])
               : "r"(addr));
  asm volatile("tcgen05.wait::ld.sync.aligned;\n"
               : "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(a[4]), "+r"(a[5]), "+r"(a[6]), "+r"(a[7])
               :
               : "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(a[i]);
}
// x = hi + lo with hi the 11-bit truncation of x (exact in fp16 for the pre-scaled range).
__device__ __forceinline__ void c_split2(float x0, float x1, uint32_t& hi, uint32_t& lo) {
  const float h0 = __uint_as_float(__float_as_uint(x0) & 0xFFFFE000u);
  const float h1 = __uint_as_float(__float_as_uint(x1) & 0xFFFFE000u);
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(hi) : "f"(h1), "f"(h0));
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(lo) : "f"(x1 - h1), "f"(x0 - h0));
}

// ------------------------------------------------------------------------------ more helpers
__device__ __forceinline__ void c_mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// wait with cluster-scope acquire: the data was written by the peer CTA (st.shared::cluster)
__device__ __forceinline__ void c_mbar_wait_cluster(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred P1;\n\t"
      "CX_%=:\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 P1, [%0], %1;\n\t"
      "@!P1 bra CX_%=;\n\t}\n"
      :
      : "r"(bar), "r"(parity)
      : "memory");
}
// arrive on the mbarrier at the same shared-memory offset in CTA `rank` of the cluster
__device__ __forceinline__ void c_mbar_arrive_cluster(uint32_t local_bar, uint32_t rank) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_bar), "r"(rank));
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}
__device__ __forceinline__ void c_named_bar(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// melspec_ts.cu: frame operand in tensor memory, work item = record (a.n_items = records).
int melspec_ts_run(const ChainArgs& a, long long clusters, cudaStream_t st);

}  // namespace eegspec
