// Row (f)1: the raw-EEG transformer in front of the spectrogram path.
//
// Replaces _EEGTransformer.transform (/root/reference/root/src/data/dataset.py:73-104): channel
// selection (:88-90) or the "magic 8" differences (:106-123), clip to +-1024 / NaN -> 0 / divide
// by 32 (:92-94), a 4th-order Butterworth low-pass applied as a CAUSAL IIR along time with
// SciPy's lfilter (:125-131, direct form II transposed, zero initial state) and decimation
// x[::downsample] (:101-102).
//
// The recurrence is sequential in time, so it is parallelised as a block scan over chunks of
// 16 * downsample samples.  The filter is linear in (state, input):
//   * every thread = (chunk, output channel) runs the recurrence over its chunk from a ZERO state
//     (fp32, samples from a shared-memory tile loaded with coalesced float4s) and keeps the 16
//     decimated outputs and the final state;
//   * the true state at the start of a chunk follows from s_{k+1} = M s_k + z_k (M = the 4 x 4
//     zero-input transition over one chunk, computed on the host in double precision); ten
//     chunks per block, the carry runs through the blocks of a record in double precision;
//   * the zero-input response of that state at the decimated positions (a 4 x 16 table G) is
//     added to the zero-state outputs.
// One CTA per record, 10 warps (chunks) x 32 lanes (channels).  Algorithmic bytes per record:
// 4 * L * n_ch in + 4 * ceil(L / ds) * n_out out (HBM bound; 0.95 MB for (10000, 20) -> (2000, 19)).
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "internal.h"
#include "melspec_chain.cuh"

namespace eegspec {

constexpr int kXfP = 16;        // decimated outputs per chunk
constexpr int kXfChunks = 10;   // chunks per block = warps per CTA
constexpr int kXfThreads = 32 * kXfChunks;

struct XfArgs {
  const float* x;
  float* out;
  long long rec_stride, length, l_out;
  int n_ch, n_out, ds, chunk, normalize, filter, out_channel_major, mu_law;
  float clip_abs, scale, mu, mu_scale;
  float b[EEGSPEC_XF_ORDER + 1], a[EEGSPEC_XF_ORDER + 1];
  float g[EEGSPEC_XF_ORDER][kXfP];                    // zero-input response of unit state i at sample m * ds
  double m[EEGSPEC_XF_ORDER][EEGSPEC_XF_ORDER];       // state after one chunk of zero input: s' = M s
  int sel_a[EEGSPEC_MAX_XF_CH], sel_b[EEGSPEC_MAX_XF_CH];
};

// mu-law companding of the filtered value (dataset.py:139-147): sign(y) log(1 + mu |y|) / log(1 + mu);
// mu_scale = 1 / log2(1 + mu).  sign(0) = 0 and NaN stays NaN, like NumPy's.
__device__ __forceinline__ float xf_mu_law(float y, float mu, float mu_scale) {
  return copysignf(log2f(fmaf(mu, fabsf(y), 1.f)) * mu_scale, y);
}

// DS > 0: compile-time decimation with normalize / filter / difference fixed by NORM / FILT / DIFF
// (straight-line recurrence: ~16 instructions per sample instead of ~42 with run-time switches,
// the kernel was issue bound at 65 % with them); DS == 0: everything read from the arguments.
template <int DS, bool NORM, bool FILT, bool DIFF>
__global__ void __launch_bounds__(kXfThreads) eeg_transform_kernel(const XfArgs a) {
  extern __shared__ __align__(16) float xs[];                       // [block samples][n_ch]
  __shared__ float zs[kXfChunks][EEGSPEC_XF_ORDER][32];             // zero-state final state of chunk k, channel c
  __shared__ double carry[EEGSPEC_XF_ORDER][32];                    // state at the start of the block

  const int tid = threadIdx.x, c = tid & 31, k = tid >> 5;
  const long long rec = blockIdx.x;
  const float* xr = a.x + rec * a.rec_stride;
  const int blk_samples = kXfChunks * a.chunk;
  const bool live = c < a.n_out;
  const int ca = live ? a.sel_a[c] : 0, cb = live ? a.sel_b[c] : -1;
  const int ds = DS ? DS : a.ds;
  const bool norm = DS ? NORM : (a.normalize != 0), filt = DS ? FILT : (a.filter != 0);
  if (tid < 32)
    for (int i = 0; i < EEGSPEC_XF_ORDER; ++i) carry[i][tid] = 0.0;

  const float b0 = a.b[0], b1 = a.b[1], b2 = a.b[2], b3 = a.b[3], b4 = a.b[4];
  const float a1 = a.a[1], a2 = a.a[2], a3 = a.a[3], a4 = a.a[4];

  for (long long t0 = 0; t0 < a.length; t0 += blk_samples) {
    __syncthreads();   // the previous block's tile and states have been consumed
    {   // tile load: blk_samples x n_ch floats, contiguous in the time-major record; zeros beyond the
        // end.  cp.async keeps all of a thread's 16-byte requests in flight at once (a plain
        // load -> store loop exposed one DRAM latency per request: 0.55 ms per 1 024 records).
      const long long n_valid = min((long long)blk_samples, a.length - t0) * a.n_ch;
      const float* src = xr + t0 * a.n_ch;
      const uint32_t dst = static_cast<uint32_t>(__cvta_generic_to_shared(xs));
      const int n4 = blk_samples * a.n_ch / 4;
      for (int i = tid; i < n4; i += kXfThreads) {
        const int bytes = (4LL * i < n_valid) ? 16 : 0;   // src-size 0: the 16 bytes are zero-filled
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst + 16u * i),
                     "l"(bytes ? src + 4LL * i : src), "r"(bytes)
                     : "memory");
      }
      asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();

    // ---- zero-state pass over this thread's chunk
    float yz[kXfP];
    float z0 = 0.f, z1 = 0.f, z2 = 0.f, z3 = 0.f;
    {
      const float* xp = xs + (long long)k * a.chunk * a.n_ch;
      const float* xq = xs + (long long)k * a.chunk * a.n_ch + (cb >= 0 ? cb : ca) - ca;   // subtrahend, same indexing
      const int n_ch = a.n_ch;
      const bool diff = DS ? DIFF : (cb >= 0);
      const float clip = a.clip_abs, scale = a.scale;
#pragma unroll
      for (int m = 0; m < kXfP; ++m) {
#pragma unroll
        for (int r = 0; r < (DS ? DS : 8); ++r) {
          if (!DS && r >= ds) break;
          float v = xp[ca];
          if (diff) v -= xq[ca];
          xp += n_ch;
          xq += n_ch;
          if (norm)   // np.clip propagates NaN, nan_to_num zeroes it (dataset.py:93-94)
            v = (v != v) ? 0.f : fminf(fmaxf(v, -clip), clip) * scale;
          float y = v;
          if (filt) {
            y = fmaf(b0, v, z0);
            z0 = fmaf(b1, v, fmaf(-a1, y, z1));
            z1 = fmaf(b2, v, fmaf(-a2, y, z2));
            z2 = fmaf(b3, v, fmaf(-a3, y, z3));
            z3 = fmaf(b4, v, -a4 * y);
          }
          if (r == 0) yz[m] = y;
        }
      }
    }
    if (filt) {
      zs[k][0][c] = z0; zs[k][1][c] = z1; zs[k][2][c] = z2; zs[k][3][c] = z3;
      __syncthreads();
      // ---- true state at the start of chunk k: s_{j+1} = M s_j + z_j from the block's carry
      double s[EEGSPEC_XF_ORDER];
#pragma unroll
      for (int i = 0; i < EEGSPEC_XF_ORDER; ++i) s[i] = carry[i][c];
      const int upto = (k == kXfChunks - 1) ? kXfChunks : k;   // the last warp also produces the next carry
      double s_k[EEGSPEC_XF_ORDER] = {s[0], s[1], s[2], s[3]};
      for (int j = 0; j < upto; ++j) {
        double nx[EEGSPEC_XF_ORDER];
#pragma unroll
        for (int i = 0; i < EEGSPEC_XF_ORDER; ++i)
          nx[i] = a.m[i][0] * s[0] + a.m[i][1] * s[1] + a.m[i][2] * s[2] + a.m[i][3] * s[3] + (double)zs[j][i][c];
#pragma unroll
        for (int i = 0; i < EEGSPEC_XF_ORDER; ++i) s[i] = nx[i];
        if (j == k - 1) {
#pragma unroll
          for (int i = 0; i < EEGSPEC_XF_ORDER; ++i) s_k[i] = s[i];
        }
      }
      __syncthreads();   // every warp has read the old carry
      if (k == kXfChunks - 1) {
#pragma unroll
        for (int i = 0; i < EEGSPEC_XF_ORDER; ++i) carry[i][c] = s[i];
      }
      const float f0 = (float)s_k[0], f1 = (float)s_k[1], f2 = (float)s_k[2], f3 = (float)s_k[3];
#pragma unroll
      for (int m = 0; m < kXfP; ++m)
        yz[m] += a.g[0][m] * f0 + a.g[1][m] * f1 + a.g[2][m] * f2 + a.g[3][m] * f3;
    }
    if (a.mu_law) {
#pragma unroll
      for (int m = 0; m < kXfP; ++m) yz[m] = xf_mu_law(yz[m], a.mu, a.mu_scale);
    }
    // ---- store the decimated outputs of this chunk
    if (live) {
      const long long o0 = t0 / ds + (long long)k * kXfP;   // t0 and the chunk are multiples of ds
      if (a.out_channel_major) {
        float* o = a.out + (rec * a.n_out + c) * a.l_out + o0;
#pragma unroll
        for (int m = 0; m < kXfP; ++m)
          if (o0 + m < a.l_out) o[m] = yz[m];
      } else {
        float* o = a.out + (rec * a.l_out + o0) * a.n_out + c;
#pragma unroll
        for (int m = 0; m < kXfP; ++m)
          if (o0 + m < a.l_out) o[(long long)m * a.n_out] = yz[m];
      }
    }
  }
}

// ----------------------------------------------------------------- span kernels (downsample 5)
// The default configuration (clip / NaN / scale, Butterworth, downsample 5) when the filter
// forgets quickly: the host takes the lane kernel below only if the state transition over three chunks
// (120 samples) is below 1e-9 (1.9e-11 for the reference's filter, far under fp32 resolution);
// slower filters keep the exact block scan above.  Differences from that kernel:
//   * the state at the start of a chunk is rebuilt from the zero-state ends of the THREE chunks
//     before it, s_k = z_{k-1} + M z_{k-2} + M^2 z_{k-3}: no serial carry, every thread does the same
//     32 FMAs (the serial scan left the other warps at a barrier for ~400 cycles per tile);
//   * so a record can be cut anywhere: one CTA = (record, span of the time axis), a span other
//     than the first starts three chunks early and discards those outputs.  The cut changes no bit
//     of the result (the warm-up chunks provide exactly the z the formula wants); the host picks
//     the number of spans that fills whole waves of resident CTAs;
//   * tiles of chunks x 40 samples arrive by bulk copy (TMA) into a double buffer, the copy of
//     tile i + 1 runs under the arithmetic of tile i (the kernel above waits for every tile);
//   * the recurrence keeps two FMAs on its critical path (the b_i x + z_{i+1} halves do not depend
//     on y); the scale is folded into b.
// (A first version with thread = (chunk, channel) packed flat and a shared-memory state ring reached
// 0.21 ms per 1 024 records; git history keeps it.)
constexpr int kSpP = 8;                      // decimated outputs per chunk
constexpr int kSpDs = 5;
constexpr int kSpChunk = kSpP * kSpDs;       // 40 samples
constexpr int kSpChunks = 10;                // chunks per tile
constexpr int kSpTile = kSpChunk * kSpChunks;
constexpr int kSpWarm = 3;                   // warm-up chunks in front of a span

struct SpArgs {
  const float* x;
  float* out;
  long long rec_stride, length, l_out;
  int n_out, units, span, stride, out_channel_major;   // span in samples (multiple of kSpChunk), stride in floats
  float clip, mu, mu_scale;
  int mu_law;
  float b[EEGSPEC_XF_ORDER + 1], a[EEGSPEC_XF_ORDER + 1];   // b already multiplied by the scale
  float g[EEGSPEC_XF_ORDER][kSpP];
  float m[EEGSPEC_XF_ORDER][EEGSPEC_XF_ORDER], m2[EEGSPEC_XF_ORDER][EEGSPEC_XF_ORDER];   // over one chunk, over two
  int sel_a[EEGSPEC_MAX_XF_CH], sel_b[EEGSPEC_MAX_XF_CH];
};

// ------------------------------------------------------- span kernel, one warp per 4 channels
// A warp owns FOUR output channels for the whole span: lane = (chunk k' of the tile's 8,
// channel cc of the warp's 4).  The three previous chunks a lane needs live in the SAME warp (the
// lanes 4, 8 and 12 below it, or the warp's own registers of the previous tile), so the chunk
// states travel by shuffle: no shared-memory state ring and no CTA barrier in the loop.  Warps
// only meet at the tile pipeline: a `full` mbarrier per stage (TMA bytes) and an `empty` one
// (one arrival per warp) before a stage is refilled.  Chunk rows are padded to a stride of
// 4 mod 32 banks: bank = 4 k' + channel, conflict-free for a channel selection of consecutive rows.
constexpr int kLnChunks = 8;                        // chunks per tile = lanes / 4
constexpr int kLnTile = kSpChunk * kLnChunks;       // 320 samples

// Wait for a tile with back-off: the data is microseconds away, and a tight try_wait loop was 10 %
// of all instructions issued (it competes with the warps that have work).
__device__ __forceinline__ void xf_mbar_wait_sleep(uint32_t bar, uint32_t parity) {
  for (;;) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred P1;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, P1;\n\t}\n"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) break;
    __nanosleep(128);
  }
}

template <int NC, bool DIFF>
__global__ void __launch_bounds__(256) eeg_transform_lane_kernel(const SpArgs a) {
  extern __shared__ __align__(128) float xs[];      // [2][kLnChunks][stride]
  __shared__ __align__(8) uint64_t bars[4];         // full[2], empty[2]

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, n_warps = blockDim.x >> 5;
  const int k = lane >> 2, c = warp * 4 + (lane & 3);
  const bool live = c < a.n_out;
  const int ca = live ? a.sel_a[c] : 0, cb = (DIFF && live) ? a.sel_b[c] : 0;
  const int rec = blockIdx.x / a.units;
  const int unit = blockIdx.x - rec * a.units;
  const float* xr = a.x + (long long)rec * a.rec_stride;
  const int length = (int)a.length;
  const int span_start = unit * a.span;
  const int span_end = min(length, span_start + a.span);
  const int t_begin = span_start - (unit ? kSpWarm * kSpChunk : 0);
  const int n_tiles = (span_end - t_begin + kLnTile - 1) / kLnTile;
  const uint32_t xs_u32 = c_smem_u32(xs), full0 = c_smem_u32(&bars[0]), empty0 = c_smem_u32(&bars[2]);
  const int buf_floats = kLnChunks * a.stride;

  if (tid == 0) {
    c_mbar_init(full0, 1);
    c_mbar_init(full0 + 8, 1);
    c_mbar_init(empty0, n_warps);
    c_mbar_init(empty0 + 8, n_warps);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  auto issue = [&](int tile) {   // one thread: the valid rows of each chunk, one bulk copy per chunk
    const int t0 = t_begin + tile * kLnTile;
    const uint32_t bar = full0 + 8u * (tile & 1);
    const uint32_t dst = xs_u32 + 4u * (uint32_t)((tile & 1) * buf_floats);
    int rows_total = min(kLnTile, span_end - t0);
    c_mbar_expect_tx(bar, (uint32_t)rows_total * NC * 4u);
    for (int kk = 0; kk < kLnChunks && rows_total > 0; ++kk, rows_total -= kSpChunk) {
      const int rows = min(rows_total, kSpChunk);
      c_bulk_g2s(dst + 4u * (uint32_t)(kk * a.stride), xr + (long long)(t0 + kk * kSpChunk) * NC,
                 (uint32_t)rows * NC * 4u, bar);
    }
  };
  if (tid == 0) issue(0);

  const float b0 = a.b[0], b1 = a.b[1], b2 = a.b[2], b3 = a.b[3], b4 = a.b[4];
  const float na1 = -a.a[1], na2 = -a.a[2], na3 = -a.a[3], na4 = -a.a[4];
  const float clip = a.clip;
  const int o_begin = span_start / kSpDs, o_end = (span_end + kSpDs - 1) / kSpDs;
  const int l_out = (int)a.l_out;
  float p0 = 0.f, p1 = 0.f, p2 = 0.f, p3 = 0.f;   // this lane's chunk state of the previous tile
  const float* xbase = xs + k * a.stride;
  int t_chunk = t_begin + k * kSpChunk;

  for (int tile = 0; tile < n_tiles; ++tile, t_chunk += kLnTile) {
    if (tid == 0 && tile + 1 < n_tiles) {
      if (tile >= 1) c_mbar_wait(empty0 + 8u * ((tile + 1) & 1), ((tile - 1) >> 1) & 1);   // every warp left that stage
      issue(tile + 1);
    }
    __syncwarp();
    xf_mbar_wait_sleep(full0 + 8u * (tile & 1), (tile >> 1) & 1);
    const int rows = max(0, min(kSpChunk, span_end - t_chunk));
    const float* xp = xbase + (tile & 1) * buf_floats;

    float yz[kSpP];
    float z0 = 0.f, z1 = 0.f, z2 = 0.f, z3 = 0.f;
#define XF_STEP(n, keep)                                                   \
  {                                                                        \
    float v = xp[(n) * NC + ca];                                           \
    if (DIFF) v -= xp[(n) * NC + cb];                                      \
    v = (v != v) ? 0.f : fminf(fmaxf(v, -clip), clip);                     \
    const float t0 = fmaf(b1, v, z1), t1 = fmaf(b2, v, z2), t2 = fmaf(b3, v, z3), t3 = b4 * v; \
    const float y = fmaf(b0, v, z0);                                       \
    z0 = fmaf(na1, y, t0);                                                 \
    z1 = fmaf(na2, y, t1);                                                 \
    z2 = fmaf(na3, y, t2);                                                 \
    z3 = fmaf(na4, y, t3);                                                 \
    if (keep) yz[(n) / kSpDs] = y;                                         \
  }
    if (__all_sync(0xffffffffu, rows == kSpChunk)) {
#pragma unroll
      for (int n = 0; n < kSpChunk; ++n) XF_STEP(n, (n % kSpDs) == 0)
    } else {   // the span ends inside this tile: chunks past the end are never stored
#pragma unroll
      for (int m = 0; m < kSpP; ++m) {
        yz[m] = 0.f;
        for (int r = 0; r < kSpDs; ++r) {
          const int n = m * kSpDs + r;
          if (n < rows) XF_STEP(n, r == 0)
        }
      }
    }
#undef XF_STEP
    __syncwarp();
    if (lane == 0) c_mbar_arrive(empty0 + 8u * (tile & 1));   // this warp is done with the stage

    // ---- start state of chunk k = z_{k-1} + M z_{k-2} + M^2 z_{k-3}, neighbours by shuffle
    float f[EEGSPEC_XF_ORDER];
    {
      const int s1 = (lane - 4) & 31, s2 = (lane - 8) & 31, s3 = (lane - 12) & 31;
      const float zc[4] = {z0, z1, z2, z3}, pc[4] = {p0, p1, p2, p3};
      float q1[4], q2[4], q3[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float a1 = __shfl_sync(0xffffffffu, zc[i], s1), e1 = __shfl_sync(0xffffffffu, pc[i], s1);
        const float a2 = __shfl_sync(0xffffffffu, zc[i], s2), e2 = __shfl_sync(0xffffffffu, pc[i], s2);
        const float a3 = __shfl_sync(0xffffffffu, zc[i], s3), e3 = __shfl_sync(0xffffffffu, pc[i], s3);
        q1[i] = k >= 1 ? a1 : e1;
        q2[i] = k >= 2 ? a2 : e2;
        q3[i] = k >= 3 ? a3 : e3;
      }
#pragma unroll
      for (int i = 0; i < EEGSPEC_XF_ORDER; ++i)
        f[i] = q1[i] + (a.m[i][0] * q2[0] + a.m[i][1] * q2[1] + a.m[i][2] * q2[2] + a.m[i][3] * q2[3]) +
               (a.m2[i][0] * q3[0] + a.m2[i][1] * q3[1] + a.m2[i][2] * q3[2] + a.m2[i][3] * q3[3]);
      p0 = z0; p1 = z1; p2 = z2; p3 = z3;
    }
#pragma unroll
    for (int m = 0; m < kSpP; ++m)
      yz[m] += a.g[0][m] * f[0] + a.g[1][m] * f[1] + a.g[2][m] * f[2] + a.g[3][m] * f[3];
    if (a.mu_law) {
#pragma unroll
      for (int m = 0; m < kSpP; ++m) yz[m] = xf_mu_law(yz[m], a.mu, a.mu_scale);
    }

    if (live) {
      const int o0 = t_chunk / kSpDs;
      const bool whole = o0 >= o_begin && o0 + kSpP <= o_end;
      if (a.out_channel_major) {
        float* o = a.out + ((long long)rec * a.n_out + c) * l_out + o0;
        if (whole && !((l_out | (int)(reinterpret_cast<uintptr_t>(a.out) >> 2)) & 3)) {   // one full sector per thread
          reinterpret_cast<float4*>(o)[0] = make_float4(yz[0], yz[1], yz[2], yz[3]);
          reinterpret_cast<float4*>(o)[1] = make_float4(yz[4], yz[5], yz[6], yz[7]);
        } else {
#pragma unroll
          for (int m = 0; m < kSpP; ++m)
            if (whole || (o0 + m >= o_begin && o0 + m < o_end)) o[m] = yz[m];
        }
      } else if (whole && a.n_out == 19) {   // constant offsets
        float* o = a.out + ((long long)rec * l_out + o0) * 19 + c;
#pragma unroll
        for (int m = 0; m < kSpP; ++m) o[m * 19] = yz[m];
      } else {
        float* o = a.out + ((long long)rec * l_out + o0) * a.n_out + c;
#pragma unroll
        for (int m = 0; m < kSpP; ++m)
          if (whole || (o0 + m >= o_begin && o0 + m < o_end)) o[m * a.n_out] = yz[m];
      }
    }
  }
}

// ------------------------------------------------------------------------------------- host
// Direct form II transposed, one step (double precision): the recurrence SciPy's lfilter runs.
static double df2t_step(const double* b, const double* a, double* z, double v) {
  const double y = b[0] * v + z[0];
  for (int i = 0; i < EEGSPEC_XF_ORDER - 1; ++i) z[i] = b[i + 1] * v - a[i + 1] * y + z[i + 1];
  z[EEGSPEC_XF_ORDER - 1] = b[EEGSPEC_XF_ORDER] * v - a[EEGSPEC_XF_ORDER] * y;
  return y;
}

void xform_params_default(eegspec_xform_params* p) {
  memset(p, 0, sizeof(*p));
  p->n_out = 19;
  for (int c = 0; c < EEGSPEC_MAX_XF_CH; ++c) { p->sel_a[c] = c < 19 ? c : 0; p->sel_b[c] = -1; }
  p->normalize = 1;
  p->clip_abs = 1024.f;
  p->scale = 1.f / 32.f;
  p->filter = 1;
  // scipy.signal.butter(4, 20 / 100, btype="low"): bilinear transform of the 4th-order Butterworth
  // prototype, evaluated here in double precision (two biquads multiplied out).
  {
    const double wc = tan(M_PI * 0.2 / 2.0);   // pre-warped cutoff, fs = 2
    double bb[5] = {1, 0, 0, 0, 0}, aa[5] = {1, 0, 0, 0, 0};
    int nb = 1;
    for (int q = 0; q < 2; ++q) {
      const double th = M_PI * (2.0 * q + 1.0) / 8.0;          // pole angles of the prototype
      const double k1 = 2.0 * sin(th);                          // s^2 + k1 s + 1
      const double w2 = wc * wc, d = 1.0 + k1 * wc + w2;
      const double sb[3] = {w2 / d, 2.0 * w2 / d, w2 / d};
      const double sa[3] = {1.0, 2.0 * (w2 - 1.0) / d, (1.0 - k1 * wc + w2) / d};
      double tb[5] = {0, 0, 0, 0, 0}, ta[5] = {0, 0, 0, 0, 0};
      for (int i = 0; i < nb; ++i)
        for (int j = 0; j < 3; ++j) { tb[i + j] += bb[i] * sb[j]; ta[i + j] += aa[i] * sa[j]; }
      memcpy(bb, tb, sizeof(bb));
      memcpy(aa, ta, sizeof(aa));
      nb += 2;
    }
    for (int i = 0; i < 5; ++i) { p->b[i] = bb[i]; p->a[i] = aa[i]; }
  }
  p->downsample = 5;
  p->out_channel_major = 0;
  p->mu_law = 0;
  p->mu = 1.f;
}

long long xform_out_length(long long length, const eegspec_xform_params* p) {
  const int ds = p->downsample < 1 ? 1 : p->downsample;
  return (length + ds - 1) / ds;
}

// The span kernel's launch: tables for 40-sample chunks, and how many spans to cut a record into.
constexpr int kSpanNotApplicable = -12345;
static int span_launch(const float* x, long long batch, long long length, int n_ch, long long rec_stride,
                       const eegspec_xform_params* p, float* out, bool diff, cudaStream_t st) {
  SpArgs a;
  a.x = x;
  a.out = out;
  a.rec_stride = rec_stride;
  a.length = length;
  a.l_out = xform_out_length(length, p);
  a.n_out = p->n_out;
  a.out_channel_major = p->out_channel_major;
  a.clip = p->clip_abs;
  a.mu_law = p->mu_law;
  a.mu = p->mu;
  a.mu_scale = 1.f / log2f(1.f + p->mu);
  for (int i = 0; i <= EEGSPEC_XF_ORDER; ++i) { a.b[i] = (float)(p->b[i] * (double)p->scale); a.a[i] = (float)p->a[i]; }
  for (int c = 0; c < EEGSPEC_MAX_XF_CH; ++c) { a.sel_a[c] = p->sel_a[c]; a.sel_b[c] = p->sel_b[c] < 0 ? 0 : p->sel_b[c]; }
  double decay = 0.0;   // largest entry of the state transition over the warm-up (three chunks)
  for (int i = 0; i < EEGSPEC_XF_ORDER; ++i) {
    double z[EEGSPEC_XF_ORDER] = {0, 0, 0, 0};
    z[i] = 1.0;
    for (int n = 0; n < kSpWarm * kSpChunk; ++n) {
      const double y = df2t_step(p->b, p->a, z, 0.0);
      if (n < kSpChunk && n % kSpDs == 0) a.g[i][n / kSpDs] = (float)y;
      if (n == kSpChunk - 1)
        for (int r = 0; r < EEGSPEC_XF_ORDER; ++r) a.m[r][i] = (float)z[r];
      if (n == 2 * kSpChunk - 1)
        for (int r = 0; r < EEGSPEC_XF_ORDER; ++r) a.m2[r][i] = (float)z[r];
    }
    for (int r = 0; r < EEGSPEC_XF_ORDER; ++r) decay = fmax(decay, fabs(z[r]));
  }
  if (!(decay < 1e-9)) return kSpanNotApplicable;   // a slowly decaying filter: the exact block scan above
  if (length >= (1LL << 30)) return kSpanNotApplicable;
  a.stride = kSpChunk * n_ch + 4;   // chunk rows 4 (n_ch = 20, 8) or 28 (n_ch = 19) banks apart: conflict-free
  const size_t smem = sizeof(float) * 2 * (size_t)kLnChunks * a.stride;
  const int threads = 32 * ((p->n_out + 3) / 4);
  void (*kernel)(const SpArgs) = diff ? eeg_transform_lane_kernel<8, true>
                                 : n_ch == 19 ? eeg_transform_lane_kernel<19, false>
                                              : eeg_transform_lane_kernel<20, false>;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  int dev = 0, sms = 0, occ = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, threads, smem);
  const long long slots = (long long)(sms > 0 ? sms : 148) * (occ > 0 ? occ : 1);
  // spans per record: whole waves of resident CTAs against the warm-up each extra span re-reads
  const long long n_chunks = (length + kSpChunk - 1) / kSpChunk;
  int best_units = 1;
  long long best_span = n_chunks * kSpChunk;
  {
    double best = 0.0;
    for (int u = 1; u <= 32; ++u) {
      const long long span = (n_chunks + u - 1) / u * kSpChunk;
      if (span < 8 * kSpWarm * kSpChunk && u > 1) break;
      const long long units = (length + span - 1) / span;
      const long long ctas = batch * units, waves = (ctas + slots - 1) / slots;
      const double work = (double)length + (double)(units - 1) * kSpWarm * kSpChunk;
      const double eff = (double)ctas / (double)(waves * slots) * (double)length / work;
      if (eff > best * 1.02) { best = eff; best_units = (int)units; best_span = span; }
    }
  }
  a.units = best_units;
  a.span = (int)best_span;
  if (batch * best_units > 0x7fffffffLL) return EEGSPEC_EUNSUPPORTED;
  kernel<<<(unsigned)(batch * best_units), threads, smem, st>>>(a);
  count_launch();
  return (int)cudaGetLastError();
}

int eeg_transform_launch(const float* x, long long batch, long long length, int n_ch, long long rec_stride,
                         const eegspec_xform_params* p, float* out, cudaStream_t st) {
  if (!x || !out || !p || batch <= 0 || length <= 0 || n_ch <= 0) return EEGSPEC_EINVAL;
  if (p->n_out < 1 || p->n_out > EEGSPEC_MAX_XF_CH) return EEGSPEC_EINVAL;
  const int ds = p->downsample < 1 ? 1 : p->downsample;
  if (ds > 8) return EEGSPEC_EUNSUPPORTED;
  // 16-byte copies (cp.async / TMA) of whole sample rows: every record and every tile of 40 rows must start
  // on a 16-byte boundary and hold a whole number of 16-byte words -- any n_ch when length % 4 == 0
  // (the reference's (10000, 19) windows), n_ch % 4 == 0 otherwise.
  if ((rec_stride & 3) || (reinterpret_cast<uintptr_t>(x) & 15)) return EEGSPEC_EALIGN;
  if ((n_ch & 3) && (length & 3)) return EEGSPEC_EALIGN;
  for (int c = 0; c < p->n_out; ++c)
    if (p->sel_a[c] < 0 || p->sel_a[c] >= n_ch || p->sel_b[c] >= n_ch) return EEGSPEC_EINVAL;
  if (p->filter && fabs(p->a[0] - 1.0) > 1e-12) return EEGSPEC_EINVAL;
  if (p->mu_law && !(p->mu > 0.f)) return EEGSPEC_EINVAL;

  XfArgs a;
  a.x = x;
  a.out = out;
  a.rec_stride = rec_stride;
  a.length = length;
  a.l_out = xform_out_length(length, p);
  a.n_ch = n_ch;
  a.n_out = p->n_out;
  a.ds = ds;
  a.chunk = kXfP * ds;
  a.normalize = p->normalize;
  a.filter = p->filter;
  a.out_channel_major = p->out_channel_major;
  a.clip_abs = p->clip_abs;
  a.scale = p->scale;
  a.mu_law = p->mu_law;
  a.mu = p->mu;
  a.mu_scale = 1.f / log2f(1.f + p->mu);
  for (int i = 0; i <= EEGSPEC_XF_ORDER; ++i) { a.b[i] = (float)p->b[i]; a.a[i] = (float)p->a[i]; }
  for (int c = 0; c < EEGSPEC_MAX_XF_CH; ++c) { a.sel_a[c] = p->sel_a[c]; a.sel_b[c] = p->sel_b[c]; }
  for (int i = 0; i < EEGSPEC_XF_ORDER; ++i) {   // unit state i, zero input, one chunk
    double z[EEGSPEC_XF_ORDER] = {0, 0, 0, 0};
    z[i] = 1.0;
    for (int n = 0; n < a.chunk; ++n) {
      const double y = p->filter ? df2t_step(p->b, p->a, z, 0.0) : 0.0;
      if (n % ds == 0) a.g[i][n / ds] = (float)y;
    }
    for (int r = 0; r < EEGSPEC_XF_ORDER; ++r) a.m[r][i] = z[r];
  }
  const size_t smem = sizeof(float) * (size_t)kXfChunks * a.chunk * n_ch;
  if (smem > 200 * 1024) return EEGSPEC_EUNSUPPORTED;
  bool any_diff = false, all_diff = true;
  for (int c = 0; c < p->n_out; ++c) { any_diff |= p->sel_b[c] >= 0; all_diff &= p->sel_b[c] >= 0; }
  const bool fast = ds == 5 && p->normalize && p->filter && (any_diff == all_diff);
  if (batch > 0x7fffffffLL) return EEGSPEC_EUNSUPPORTED;
  if (fast && (((n_ch == 20 || n_ch == 19) && !all_diff) || (n_ch == 8 && all_diff))) {
    const int rc = span_launch(x, batch, length, n_ch, rec_stride, p, out, all_diff, st);
    if (rc != kSpanNotApplicable) return rc;
  }
  void (*kernel)(const XfArgs) = !fast ? eeg_transform_kernel<0, false, false, false>
                                 : all_diff ? eeg_transform_kernel<5, true, true, true>
                                            : eeg_transform_kernel<5, true, true, false>;
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return EEGSPEC_EUNSUPPORTED;
  if (smem > 48 * 1024) {   // cheap and idempotent; three instantiations, so no per-device cache
    const cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  kernel<<<(unsigned)batch, kXfThreads, smem, st>>>(a);
  count_launch();
  return (int)cudaGetLastError();
}

}  // namespace eegspec
