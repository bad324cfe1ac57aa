// Stem convolution (1 -> 16 channels, 7x7, stride 1, pad 3: reference models/ub_uresnet.py:41,94) on tcgen05.
//
// With one input channel the conv is a GEMM over the 49 filter taps:  out[pixel][co] = sum_k P[pixel][k] * W[co][k],
// k = r*8 + s (7 filter rows x 8 slots; slot 7 and k >= 56 are padding -> K = 64).  P is an im2col tile of 128
// consecutive pixels of one image row, built in shared memory by the CTA's threads from a rolling window of 7 staged
// input rows: pixel p, filter row r is ONE 16-byte chunk (the 8 consecutive inputs x[h+r-3][w-3 .. w+4]) of the
// pixel's 128-byte row, written at the SW128-swizzled position -- the byte layout TMA would produce for a
// [128 pixels][64 channels] box -- so the same tile is
//   * the K-major A operand of the forward MMA        D[128 px][16 co] = P x W^T,                    K = 64 taps
//   * the MN-major A operand of the weight gradient   D[k][16 co]     += P^T x dy,                   K = 128 pixels
// (chunk 7 of every pixel row is {1, 0, ...}: weight-gradient row k = 56 accumulates sum(dy) = the bias gradient).
//
// Precision: the stem is the one conv whose operands the reference-parity budget does not allow in plain bf16 (the
// loss of an untrained net moves 3e-3 under bf16 rounding of conv1.weight alone).  x and W are therefore split into
// two bf16 terms, v = hi + lo with lo = bf16(v - hi) (16-17 significant bits), and the products hi*hi + lo*hi + hi*lo
// are accumulated in the fp32 TMEM accumulator: two im2col tiles (P_hi, P_lo) per pixel tile, two weight tiles.
// The CUDA-core kernels in stem.cu remain for the shapes this file does not take.
#include "common.cuh"
#include "conv_internal.cuh"

namespace ubr {

constexpr int kSTThreads = 256;
// staged input row: 136 bf16 columns (w0-3 .. w0+132).  The 272-byte pitch puts the 8 ring rows 4 banks apart, so the
// 8 lanes of a quarter-warp that read 16 bytes of 8 different rows touch all 32 banks once.
constexpr int kXsCols = 136;
constexpr uint32_t kXsPitch = kXsCols * 2u;
constexpr uint32_t kXsBytes = 8u * kXsPitch;        // one staging array: ring of 8 rows

struct StemTcParams {
  const float* x;                            // [n][h][w] fp32
  int n, h, w;
  int bh, bands, wblocks, nstrips;
  const float* wgt;                          // fwd: [16][49] fp32
  const float* bias;                         // fwd: [16] or null
  __nv_bfloat16* y;                          // fwd: NHWC view
  int yld;
  int relu;                                  // fwd: y = max(y, 0) (inference with BatchNorm folded into wgt / bias)
  const __nv_bfloat16* dy;                   // wgrad: NHWC view (16 channels)
  int dyld;
  float* partials;                           // wgrad: [gridDim.x][128][16]
};

// shared memory (offsets from a 1024-byte aligned base)
//   [0, 16K)      P_hi    [16K, 32K)  P_lo          im2col tiles, 128 rows x 128 B, SW128
//   [32K, 40K)    fwd: W_hi (2 KB) at +0, W_lo at +2048;   wgrad: dy tile (4 KB, SW32)
//   [40K, ...)    2 staging arrays (hi, lo terms of x), each a ring of 8 rows
struct StemSmem {
  uint32_t a_hi, a_lo, aux, xs, bar, tslot;
};
__device__ __forceinline__ StemSmem stem_smem_layout(uint32_t base) {
  StemSmem s;
  s.a_hi = base;
  s.a_lo = base + 16384u;
  s.aux = base + 32768u;
  s.xs = s.aux + 8192u;
  s.bar = s.xs + 2u * kXsBytes;
  s.tslot = s.bar + 16u;
  return s;
}
constexpr uint32_t kStemSmemBytes = 32768u + 8192u + 2u * kXsBytes + 32u + 1024u;

__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void lds128(uint32_t addr, uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d) {
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr));
}
__device__ __forceinline__ void sts16(uint32_t addr, uint16_t v) {
  asm volatile("st.shared.b16 [%0], %1;" ::"r"(addr), "h"(v) : "memory");
}
__device__ __forceinline__ void split_bf16(float v, uint16_t& hi, uint16_t& lo) {
  const __nv_bfloat16 h = __float2bfloat16_rn(v);
  hi = __bfloat16_as_ushort(h);
  lo = __bfloat16_as_ushort(__float2bfloat16_rn(v - __bfloat162float(h)));
}

// input row hh (columns w0-3 .. w0+132) of image n: thread i < 136 fetches one element (out-of-image -> 0) ...
__device__ __forceinline__ float stem_fetch_row(const StemTcParams& p, int n, int hh, int w0) {
  const int i = threadIdx.x;
  const int c = w0 - 3 + i;
  float v = 0.f;
  if (i < 136 && hh >= 0 && hh < p.h && c >= 0 && c < p.w) v = __ldg(p.x + ((size_t)n * p.h + hh) * (size_t)p.w + c);
  return v;
}
// ... and puts its two bf16 terms into ring slot (hh & 7) of the staging arrays
__device__ __forceinline__ void stem_put_row(const StemSmem& s, int hh, float v) {
  const uint32_t slot = (uint32_t)(hh + 8) & 7u;
  const uint32_t i = threadIdx.x;
  if (i < 136u) {
    uint16_t hi, lo;
    split_bf16(v, hi, lo);
    const uint32_t e = slot * kXsPitch + i * 2u;
    sts16(s.xs + e, hi);
    sts16(s.xs + kXsBytes + e, lo);
  }
}

// im2col tiles of the 128 pixels (h, w0 .. w0+127).  Warps 0-3 build P_hi, warps 4-7 P_lo; within a warp the lane is
// (filter row r = lane & 7, run = lane >> 3): the thread loads the 16 staged inputs under a run of 8 consecutive
// pixels of its filter row (2 x 16 bytes) and writes the 8 pixels' chunks, each the 8-input window starting one
// element further (odd starts by funnel shift).  A quarter-warp = the 8 chunks of ONE pixel row of the tile = 128
// contiguous bytes (swizzled order): loads and stores are bank-conflict free.  r = 7 is the constant chunk.
__device__ __forceinline__ void stem_build_tiles(const StemSmem& s, int h) {
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  const uint32_t r = lane & 7u;
  const uint32_t q = (lane >> 3) + 4u * (warp & 3u);                     // run of pixels 8q .. 8q+7
  const bool lo = warp >= 4u;
  const uint32_t tile = (lo ? s.a_lo : s.a_hi) + q * 1024u;
  uint32_t w[8];
  {
    const uint32_t slot = (uint32_t)(h + (int)r - 3 + 8) & 7u;           // (r = 7 reads a valid but unused row)
    const uint32_t src = s.xs + (lo ? kXsBytes : 0u) + slot * kXsPitch + q * 16u;
    lds128(src, w[0], w[1], w[2], w[3]);
    lds128(src + 16u, w[4], w[5], w[6], w[7]);
  }
  const bool konst = r == 7u;
  const uint32_t one = lo ? 0u : 0x00003F80u;                            // k = 56: constant 1 (bias-gradient row)
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    uint32_t c[4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
      c[j] = (i & 1) ? __funnelshift_r(w[(i >> 1) + j], w[(i >> 1) + j + 1 < 8 ? (i >> 1) + j + 1 : 7], 16)
                     : w[(i >> 1) + j];
    if (konst) { c[0] = one; c[1] = 0u; c[2] = 0u; c[3] = 0u; }
    sts128(tile + (uint32_t)i * 128u + ((r ^ (uint32_t)i) << 4), c[0], c[1], c[2], c[3]);
  }
}

struct StemStrip {
  int n, w0, h0;
};
__device__ __forceinline__ StemStrip stem_strip(const StemTcParams& p, int strip) {
  StemStrip st;
  st.w0 = (strip % p.wblocks) * 128;
  st.h0 = ((strip / p.wblocks) % p.bands) * p.bh;
  st.n = strip / (p.wblocks * p.bands);
  return st;
}

// ------------------------------------------------------------------------------------------------ forward
// accumulator columns 0..15 = P_hi W_hi + P_lo W_hi, 16..31 = P_hi W_lo
__device__ __forceinline__ void stem_fwd_store(const StemTcParams& p, uint32_t taddr, const float (&bias_r)[16],
                                               size_t pixel) {
  uint32_t v[16], u[16];
  tmem_ld16(taddr, v);
  tmem_ld16(taddr + 16u, u);
  tmem_ld_wait();
  tc_fence_before();
  uint32_t pk[8];
#pragma unroll
  for (int j = 0; j < 8; ++j)
  {
    float a = __uint_as_float(v[2 * j]) + __uint_as_float(u[2 * j]) + bias_r[2 * j];
    float b = __uint_as_float(v[2 * j + 1]) + __uint_as_float(u[2 * j + 1]) + bias_r[2 * j + 1];
    if (p.relu) { a = fmaxf(a, 0.f); b = fmaxf(b, 0.f); }
    pk[j] = pack_bf16x2(a, b);
  }
  __nv_bfloat16* dst = p.y + pixel * (size_t)p.yld;
  *reinterpret_cast<uint4*>(dst) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
  *reinterpret_cast<uint4*>(dst + 8) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
}

__global__ void __launch_bounds__(kSTThreads, 4) stem_tc_fwd_kernel(const StemTcParams p) {
  pdl_launch_dependents();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const StemSmem s = stem_smem_layout(base);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    mbar_init(s.bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc_dyn(s.tslot, 64);
    tmem_relinquish();
  }
  pdl_wait();                     // the weights below are written by the optimizer step of the previous kernel chain
  // weight tiles W_hi / W_lo [16 co][64 k], K-major SW128 (slot 7 of every filter row and k >= 56 are zero)
  if (threadIdx.x < 128) {
    const int co = threadIdx.x >> 3, j = threadIdx.x & 7;
    uint32_t wh[4] = {0u, 0u, 0u, 0u}, wl[4] = {0u, 0u, 0u, 0u};
    if (j < 7) {
      uint16_t hi[8], lo[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) split_bf16(q < 7 ? __ldg(p.wgt + co * 49 + j * 7 + q) : 0.f, hi[q], lo[q]);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        wh[q] = (uint32_t)hi[2 * q] | ((uint32_t)hi[2 * q + 1] << 16);
        wl[q] = (uint32_t)lo[2 * q] | ((uint32_t)lo[2 * q + 1] << 16);
      }
    }
    const uint32_t off = (uint32_t)co * 128u + (((uint32_t)j ^ ((uint32_t)co & 7u)) << 4);
    sts128(s.aux + off, wh[0], wh[1], wh[2], wh[3]);
    sts128(s.aux + 2048u + off, wl[0], wl[1], wl[2], wl[3]);
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(s.tslot));

  // B operands: [W_hi ; W_lo] as one 32-row tile (N = 32) for P_hi, its first 16 rows (N = 16) for P_lo
  const uint32_t idesc32 = make_idesc_bf16(128, 32, 0, 0), idesc16 = make_idesc_bf16(128, 16, 0, 0);
  float bias_r[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) bias_r[j] = p.bias ? __ldg(p.bias + j) : 0.f;
  const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;

  uint32_t t = 0;                    // tiles issued by this CTA; tile t accumulates in TMEM columns (t & 1) * 32
  size_t prev_pix = 0;               // first output pixel of tile t-1
  for (int strip = blockIdx.x; strip < p.nstrips; strip += gridDim.x) {
    const StemStrip st = stem_strip(p, strip);
    {                                // (the previous strip's last tile was built before its MMA-issue barrier)
      float v[7];
#pragma unroll
      for (int i = 0; i < 7; ++i) v[i] = stem_fetch_row(p, st.n, st.h0 - 3 + i, st.w0);
#pragma unroll
      for (int i = 0; i < 7; ++i) stem_put_row(s, st.h0 - 3 + i, v[i]);
    }
    __syncthreads();                 // the strip's first rows are staged
    float pend = stem_fetch_row(p, st.n, st.h0 + 4, st.w0);             // rows are fetched two tiles ahead of their use
    for (int h = st.h0; h < st.h0 + p.bh; ++h, ++t) {
      const float nxt = stem_fetch_row(p, st.n, h + 5, st.w0);
      // MMA(t-1) has consumed the tile buffers.  No CTA barrier here: staged rows, TMEM reuse and the ring slot
      // written below are all ordered by the barrier that precedes each MMA issue.
      if (t >= 1) mbar_wait(s.bar, (t - 1u) & 1u);
      stem_build_tiles(s, h);
      stem_put_row(s, h + 4, pend);                                     // ring slot of row h-4: not part of this tile
      pend = nxt;
      fence_proxy_async_smem();
      tc_fence_before();
      __syncthreads();
      if (threadIdx.x == 0) {
        tc_fence_after();
        const uint64_t a_hi = make_smem_desc(s.a_hi, 0, 1024u, 2u), a_lo = make_smem_desc(s.a_lo, 0, 1024u, 2u);
        const uint64_t wdesc = make_smem_desc(s.aux, 0, 1024u, 2u);
        const uint32_t d = tmem_base + (t & 1u) * 32u;
#pragma unroll
        for (int k = 0; k < 4; ++k) tc_mma_bf16(d, a_hi + (uint64_t)(2 * k), wdesc + (uint64_t)(2 * k), idesc32, k > 0 ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < 4; ++k) tc_mma_bf16(d, a_lo + (uint64_t)(2 * k), wdesc + (uint64_t)(2 * k), idesc16, 1u);
        tc_commit(s.bar);
      }
      // epilogue of tile t-1 (its MMAs were waited for above): overlaps MMA(t) and the other warps' next fetch
      if (t >= 1 && warp < 4) {
        tc_fence_after();
        stem_fwd_store(p, tmem_base + ((t - 1u) & 1u) * 32u + lane_off, bias_r, prev_pix + (size_t)(warp * 32 + lane));
      }
      prev_pix = ((size_t)st.n * p.h + h) * (size_t)p.w + st.w0;
    }
  }
  if (t >= 1 && warp < 4) {
    mbar_wait(s.bar, (t - 1u) & 1u);
    tc_fence_after();
    stem_fwd_store(p, tmem_base + ((t - 1u) & 1u) * 32u + lane_off, bias_r, prev_pix + (size_t)(warp * 32 + lane));
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc_dyn(tmem_base, 64);
  }
}

// ------------------------------------------------------------------------------------------------ weight gradient
// D[m][co] += sum over the tile's 128 pixels P[px][m] * dy[px][co], m = 0..63 the hi taps, 64..127 the lo taps (the
// two im2col tiles are the two 64-channel blocks of one MN-major A operand, LBO = 16 KB apart).  One fp32
// accumulator [128][16] per CTA lives in TMEM over all its tiles; hi + lo rows are added by the reduce kernel.
__global__ void __launch_bounds__(kSTThreads, 4) stem_tc_wgrad_kernel(const StemTcParams p) {
  pdl_launch_dependents();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const StemSmem s = stem_smem_layout(base);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    mbar_init(s.bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc_dyn(s.tslot, 32);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(s.tslot));
  pdl_wait();

  const uint32_t idesc = make_idesc_bf16(128, 16, 1, 1);
  const int dpx = threadIdx.x >> 1, dj = threadIdx.x & 1;                 // dy tile: thread <-> (pixel, 16-byte half)
  const uint32_t dy_off = (uint32_t)dpx * 32u + (((uint32_t)dj ^ (((uint32_t)dpx >> 2) & 1u)) << 4);   // SW32

  uint32_t t = 0;
  for (int strip = blockIdx.x; strip < p.nstrips; strip += gridDim.x) {
    const StemStrip st = stem_strip(p, strip);
    {
      float v[7];
#pragma unroll
      for (int i = 0; i < 7; ++i) v[i] = stem_fetch_row(p, st.n, st.h0 - 3 + i, st.w0);
#pragma unroll
      for (int i = 0; i < 7; ++i) stem_put_row(s, st.h0 - 3 + i, v[i]);
    }
    __syncthreads();                 // the strip's first rows are staged
    float pend = stem_fetch_row(p, st.n, st.h0 + 4, st.w0);               // x rows and dy tiles are fetched ahead
    const __nv_bfloat16* dyp = p.dy + (((size_t)st.n * p.h + st.h0) * (size_t)p.w + st.w0 + dpx) * (size_t)p.dyld + dj * 8;
    const size_t dy_row = (size_t)p.w * (size_t)p.dyld;
    uint4 dv = __ldg(reinterpret_cast<const uint4*>(dyp));
    for (int h = st.h0; h < st.h0 + p.bh; ++h, ++t) {
      const float nxt = stem_fetch_row(p, st.n, h + 5, st.w0);
      uint4 dn = dv;
      if (h + 1 < st.h0 + p.bh) dn = __ldg(reinterpret_cast<const uint4*>(dyp + (size_t)(h + 1 - st.h0) * dy_row));
      if (t >= 1) mbar_wait(s.bar, (t - 1u) & 1u);                        // MMA(t-1) has consumed the tile buffers
      stem_build_tiles(s, h);
      sts128(s.aux + dy_off, dv.x, dv.y, dv.z, dv.w);
      stem_put_row(s, h + 4, pend);
      pend = nxt;
      dv = dn;
      fence_proxy_async_smem();
      tc_fence_before();
      __syncthreads();
      if (threadIdx.x == 0) {
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < 8; ++k) {                                     // 128 pixels = 8 x K16
          const uint64_t adesc = make_smem_desc(s.a_hi + (uint32_t)k * 16u * 128u, 16384u, 8u * 128u, 2u);
          const uint64_t bdesc = make_smem_desc(s.aux + (uint32_t)k * 16u * 32u, 0u, 8u * 32u, 6u);
          tc_mma_bf16(tmem_base, adesc, bdesc, idesc, (t > 0 || k > 0) ? 1u : 0u);
        }
        tc_commit(s.bar);
      }
    }
  }
  if (warp < 4) {
    float* out = p.partials + ((size_t)blockIdx.x * 128 + (size_t)(warp * 32 + lane)) * 16;
    uint32_t v[16];
    if (t >= 1) {
      mbar_wait(s.bar, (t - 1u) & 1u);                                    // MMAs retire in order: all tiles are in
      tc_fence_after();
      tmem_ld16(tmem_base + ((uint32_t)(warp * 32) << 16), v);
      tmem_ld_wait();
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = 0u;
    }
    float4* d4 = reinterpret_cast<float4*>(out);
#pragma unroll
    for (int j = 0; j < 4; ++j)
      d4[j] = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]),
                          __uint_as_float(v[4 * j + 3]));
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc_dyn(tmem_base, 32);
  }
}

// partials [nblocks][128 m][16 co] -> dw[co][r*7+s] = sum_b part[b][k][co] + part[b][64+k][co]  (k = r*8+s, s < 7),
// dbias[co] = the same for k = 56.  One block per k; 16 slices of the CTA partials per output, combined in fixed order.
__global__ void __launch_bounds__(256) stem_tc_wgrad_reduce_kernel(const float* __restrict__ partials, int nblocks,
                                                                   float* dw, float* db, int accumulate) {
  pdl_prologue();
  __shared__ double red[16][16];
  const int k = blockIdx.x, r = k >> 3, sl = k & 7;
  const bool is_w = r < 7 && sl < 7, is_b = k == 56;
  if (!is_w && !is_b) return;
  const int co = threadIdx.x & 15, slice = threadIdx.x >> 4;
  double acc = 0.0;
  for (int b = slice; b < nblocks; b += 16) {
    const float* pb = partials + (size_t)b * 2048;
    acc += (double)pb[k * 16 + co] + (double)pb[(64 + k) * 16 + co];
  }
  red[slice][co] = acc;
  __syncthreads();
  if (threadIdx.x < 16) {
    double t = 0.0;
#pragma unroll
    for (int i = 0; i < 16; ++i) t += red[i][co];
    if (is_w) {
      if (dw) dw[co * 49 + r * 7 + sl] = (accumulate ? dw[co * 49 + r * 7 + sl] : 0.f) + (float)t;
    } else if (db) {
      db[co] = (accumulate ? db[co] : 0.f) + (float)t;
    }
  }
}

// ------------------------------------------------------------------------------------------------ host side
static bool stem_tc_plan(int n, int h, int w, StemTcParams& p) {
  if (w % 128 != 0 || h % 8 != 0) return false;
  p.n = n; p.h = h; p.w = w;
  p.wblocks = w / 128;
  // strip height: as tall as possible (6 halo rows are re-staged per strip) while the persistent CTAs still get
  // >= 3 strips each, so that the round-robin tail stays small
  p.bh = 8;
  for (int bh = 64; bh > 8; bh >>= 1)
    if (h % bh == 0 && (long)n * (h / bh) * p.wblocks >= 3L * kStemTcMaxBlocks) { p.bh = bh; break; }
  p.bands = h / p.bh;
  p.nstrips = n * p.bands * p.wblocks;
  return true;
}

bool stem_tc_eligible(int n, int h, int w, int out_ch) {
  StemTcParams p{};
  return out_ch == 16 && stem_tc_plan(n, h, w, p);
}

int stem_tc_fwd(const float* x, int n, int h, int w, const float* wgt, const float* bias, __nv_bfloat16* y, int yld,
                cudaStream_t st, int relu) {
  StemTcParams p{};
  UBR_REQUIRE(stem_tc_plan(n, h, w, p), "stem_tc_fwd: shape %dx%dx%d not eligible", n, h, w);
  p.x = x; p.wgt = wgt; p.bias = bias; p.y = y; p.yld = yld; p.relu = relu;
  UBR_CUDA_OK(cudaFuncSetAttribute(stem_tc_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kStemSmemBytes));
  const int grid = p.nstrips < kStemTcMaxBlocks ? p.nstrips : kStemTcMaxBlocks;
  ::ubr::Launch(grid, kSTThreads, kStemSmemBytes, st)(stem_tc_fwd_kernel, p);
  UBR_LAUNCH_CHECK("stem_tc_fwd_kernel");
  return 0;
}

int stem_tc_wgrad(const float* x, int n, int h, int w, const __nv_bfloat16* dy, int dyld, float* dw, float* dbias,
                  int accumulate, float* workspace, cudaStream_t st) {
  StemTcParams p{};
  UBR_REQUIRE(stem_tc_plan(n, h, w, p), "stem_tc_wgrad: shape %dx%dx%d not eligible", n, h, w);
  p.x = x; p.dy = dy; p.dyld = dyld; p.partials = workspace;
  UBR_CUDA_OK(cudaFuncSetAttribute(stem_tc_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kStemSmemBytes));
  const int grid = p.nstrips < kStemTcMaxBlocks ? p.nstrips : kStemTcMaxBlocks;
  ::ubr::Launch(grid, kSTThreads, kStemSmemBytes, st)(stem_tc_wgrad_kernel, p);
  UBR_LAUNCH_CHECK("stem_tc_wgrad_kernel");
  ::ubr::Launch(64, 256, 0, st)(stem_tc_wgrad_reduce_kernel, workspace, grid, dw, dbias, accumulate);
  UBR_LAUNCH_CHECK("stem_tc_wgrad_reduce_kernel");
  return 0;
}

}  // namespace ubr
