// LogSoftmax + PixelWiseNLLLoss (reference models/ub_uresnet.py:70,143; training/pixelwise_nllloss.py:41-61),
// whole-view tile gather / stitch scatter (deploy/run_ubresnet_wholeview.py:238-239; larcv Image2D.cxx:362-367),
// and weight packing.  All coalesced, fp32 math, deterministic reductions.
#include "common.cuh"

namespace ubr {

constexpr int kMaxCls = 16;

// logits [pixel][ld] fp32 -> logp NCHW fp32
__global__ void __launch_bounds__(256) logsoftmax_fwd_kernel(const float* __restrict__ logits, int ld, int n, int ncls,
                                                             int hw, float* __restrict__ logp) {
  pdl_prologue();
  const long total = (long)n * hw;
  for (long p = (long)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += (long)gridDim.x * blockDim.x) {
    const float* src = logits + p * ld;
    float v[kMaxCls];
    float m = -INFINITY;
    if (ld == 4 && ncls <= 4) {
      const float4 q = *reinterpret_cast<const float4*>(src);
      v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
    } else {
      for (int k = 0; k < ncls; ++k) v[k] = src[k];
    }
    for (int k = 0; k < ncls; ++k) m = fmaxf(m, v[k]);
    float s = 0.f;
    for (int k = 0; k < ncls; ++k) s += expf(v[k] - m);
    const float lse = m + logf(s);
    const long b = p / hw, i = p % hw;
    for (int k = 0; k < ncls; ++k) logp[(b * ncls + k) * hw + i] = v[k] - lse;
  }
}

// dlogits[p][k] = dlogp[k] - exp(logp[k]) * sum_j dlogp[j]   (bf16, channels >= ncls zeroed)
__global__ void __launch_bounds__(256) logsoftmax_bwd_kernel(const float* __restrict__ logp,
                                                             const float* __restrict__ dlogp, int n, int ncls, int hw,
                                                             __nv_bfloat16* __restrict__ dl, int dl_c, int dl_ld) {
  pdl_prologue();
  const long total = (long)n * hw;
  for (long p = (long)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += (long)gridDim.x * blockDim.x) {
    const long b = p / hw, i = p % hw;
    float g[kMaxCls], s = 0.f;
    for (int k = 0; k < ncls; ++k) { g[k] = dlogp[(b * ncls + k) * hw + i]; s += g[k]; }
    for (int k = 0; k < ncls; ++k) g[k] -= expf(logp[(b * ncls + k) * hw + i]) * s;
    __nv_bfloat16* dst = dl + p * dl_ld;
    for (int c0 = 0; c0 < dl_c; c0 += 8) {
      float f[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = (c0 + j < ncls) ? g[c0 + j] : 0.f;
      uint4 u;
      u.x = pack_bf16x2(f[0], f[1]); u.y = pack_bf16x2(f[2], f[3]);
      u.z = pack_bf16x2(f[4], f[5]); u.w = pack_bf16x2(f[6], f[7]);
      *reinterpret_cast<uint4*>(dst + c0) = u;
    }
  }
}

constexpr int kNllBlocks = 592;
__global__ void __launch_bounds__(256) nll_fwd_kernel(const float* __restrict__ logp, const long long* __restrict__ target,
                                                      const float* __restrict__ pixw, const float* __restrict__ classw,
                                                      int n, int ncls, int hw, long long ignore_index,
                                                      float* __restrict__ partials) {
  pdl_prologue();
  const long total = (long)n * hw;
  float acc = 0.f;
  for (long p = (long)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += (long)gridDim.x * blockDim.x) {
    const long long t = target[p];
    if (t != ignore_index && t >= 0 && t < ncls) {
      const long b = p / hw, i = p % hw;
      float v = -logp[(b * ncls + t) * hw + i];
      if (classw) v *= classw[t];
      acc = fmaf(v, pixw[p], acc);
    }
  }
  __shared__ float red[8];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int i = 0; i < 8; ++i) s += red[i];
    partials[blockIdx.x] = s;
  }
}
__global__ void nll_finalize_kernel(const float* partials, int nb, double inv_count, float* loss) {
  pdl_prologue();
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    double s = 0.0;
    for (int i = 0; i < nb; ++i) s += (double)partials[i];
    *loss = (float)(s * inv_count);
  }
}
__global__ void __launch_bounds__(256) nll_bwd_kernel(const long long* __restrict__ target, const float* __restrict__ pixw,
                                                      const float* __restrict__ classw, const float* __restrict__ gout,
                                                      int n, int ncls, int hw, long long ignore_index, float inv_count,
                                                      float* __restrict__ dlogp) {
  pdl_prologue();
  const long total = (long)n * hw;
  const float go = gout ? *gout : 1.f;
  for (long p = (long)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += (long)gridDim.x * blockDim.x) {
    const long long t = target[p];
    const long b = p / hw, i = p % hw;
    float g = 0.f;
    if (t != ignore_index && t >= 0 && t < ncls) g = -go * pixw[p] * (classw ? classw[t] : 1.f) * inv_count;
    for (int k = 0; k < ncls; ++k) dlogp[(b * ncls + k) * hw + i] = (k == t) ? g : 0.f;
  }
}

// ------------------------------------------------------------------ tiling
__global__ void __launch_bounds__(256) tile_gather_kernel(const float* __restrict__ plane, int rows, int cols,
                                                          const int* __restrict__ origins, int ntiles, int crop_h,
                                                          int crop, float* __restrict__ crops) {
  pdl_prologue();
  const long per = (long)crop_h * crop;
  const long total = per * ntiles;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int t = (int)(i / per);
    const int r = (int)((i % per) / crop), c = (int)(i % crop);
    const int sr = origins[2 * t] + r, sc = origins[2 * t + 1] + c;
    crops[i] = (sr >= 0 && sr < rows && sc >= 0 && sc < cols) ? plane[(long)sr * cols + sc] : 0.f;
  }
}
// every output pixel takes the value of the highest-index tile that covers it (== sequential overwrite)
__global__ void __launch_bounds__(256) stitch_scatter_kernel(const float* __restrict__ tiles, int nch,
                                                             const int* __restrict__ origins, int ntiles, int crop_h,
                                                             int crop, int rows, int cols, float* __restrict__ out) {
  pdl_prologue();
  const long total = (long)rows * cols;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int r = (int)(i / cols), c = (int)(i % cols);
    int owner = -1;
    for (int t = ntiles - 1; t >= 0; --t) {
      const int r0 = origins[2 * t], c0 = origins[2 * t + 1];
      if (r >= r0 && r < r0 + crop_h && c >= c0 && c < c0 + crop) { owner = t; break; }
    }
    for (int k = 0; k < nch; ++k) {
      float v = 0.f;
      if (owner >= 0) {
        const int r0 = origins[2 * owner], c0 = origins[2 * owner + 1];
        v = tiles[(((long)owner * nch + k) * crop_h + (r - r0)) * crop + (c - c0)];
      }
      out[(long)k * total + i] = v;
    }
  }
}

// per-pixel class = argmax over the C log-probabilities (first maximum, like torch.argmax / `output.max(1)`)
__global__ void __launch_bounds__(256) argmax_classes_kernel(const float* __restrict__ logp, int n, int ncls, long hw,
                                                             uint8_t* __restrict__ cls) {
  pdl_prologue();
  const long total = (long)n * hw;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const long b = i / hw, px = i - b * hw;
    const float* p = logp + b * ncls * hw + px;
    float best = p[0];
    int bi = 0;
    for (int k = 1; k < ncls; ++k) {
      const float v = p[k * hw];
      if (v > best) { best = v; bi = k; }
    }
    cls[i] = (uint8_t)bi;
  }
}
// uint8 class tiles -> stitched class map; same ownership rule as stitch_scatter_kernel; uncovered pixels get `fill`
__global__ void __launch_bounds__(256) stitch_scatter_u8_kernel(const uint8_t* __restrict__ tiles,
                                                                const int* __restrict__ origins, int ntiles, int crop_h,
                                                                int crop, int rows, int cols, int fill,
                                                                uint8_t* __restrict__ out) {
  pdl_prologue();
  const long total = (long)rows * cols;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int r = (int)(i / cols), c = (int)(i % cols);
    uint8_t v = (uint8_t)fill;
    for (int t = ntiles - 1; t >= 0; --t) {
      const int r0 = origins[2 * t], c0 = origins[2 * t + 1];
      if (r >= r0 && r < r0 + crop_h && c >= c0 && c < c0 + crop) {
        v = tiles[((long)t * crop_h + (r - r0)) * crop + (c - c0)];
        break;
      }
    }
    out[i] = v;
  }
}

// ------------------------------------------------------------------ weight packing
// dst[tap][n][k] (bf16, zero padded); one extra all-zero slab at index taps (used by phases no tap reaches)
// one element of the packed operand.  src_layout 0: OIHW conv, 1: IOHW deconv (16 taps),
// 2: IOHW ConvTranspose(k4,s2,p1) re-expressed as a 3x3 stride-1 "sub-pixel" conv with 4*out_ch phase channels
//    (phase channel = (ph*2+pw)*out_ch + co): n_is_out=1 forward  (y4 = conv3x3(x)),  n_is_out=0 data gradient
//    (dx = conv3x3(dy4)); taps that a phase does not use are zero.
__device__ __forceinline__ float pack_value(const float* __restrict__ src, int src_layout, int n_is_out, int out_ch,
                                            int in_ch, int taps, int t, int nn, int k) {
  if (src_layout == 2) {
    const int pc = n_is_out ? nn : k, ci = n_is_out ? k : nn;
    if (pc >= 4 * out_ch || ci >= in_ch) return 0.f;
    const int ph = pc / (2 * out_ch), pw = (pc / out_ch) & 1, co = pc % out_ch;
    const int dh = t / 3 - 1, dw = t % 3 - 1;
    const int kh4 = n_is_out ? (ph + 1 - 2 * dh) : (2 * dh + ph + 1);
    const int kw4 = n_is_out ? (pw + 1 - 2 * dw) : (2 * dw + pw + 1);
    if (kh4 < 0 || kh4 > 3 || kw4 < 0 || kw4 > 3) return 0.f;
    return src[(((long)ci * out_ch + co) * 4 + kh4) * 4 + kw4];
  }
  const int o = n_is_out ? nn : k, ci = n_is_out ? k : nn;
  if (o >= out_ch || ci >= in_ch) return 0.f;
  const long idx = src_layout == 0 ? (((long)o * in_ch + ci) * taps + t) : (((long)ci * out_ch + o) * taps + t);
  return src[idx];
}

__global__ void __launch_bounds__(256) pack_weights_kernel(const float* __restrict__ src, int src_layout, int n_is_out,
                                                           int out_ch, int in_ch, int kh, int kw, int n_pad, int k_pad,
                                                           __nv_bfloat16* __restrict__ dst) {
  pdl_prologue();
  const int taps = kh * kw;
  // 32-bit index arithmetic (the host checks total < 2^31); the grid is wide enough that a thread of the largest
  // conv (512x512x3x3) handles a few tens of elements -- the strided source reads are latency bound
  const unsigned total = (unsigned)(taps + 1) * (unsigned)n_pad * (unsigned)k_pad;
  const unsigned plane = (unsigned)k_pad * (unsigned)n_pad;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned t = i / plane, rem = i - t * plane;
    const unsigned nn = rem / (unsigned)k_pad, k = rem - nn * (unsigned)k_pad;
    const float v = (int)t < taps ? pack_value(src, src_layout, n_is_out, out_ch, in_ch, taps, (int)t, (int)nn, (int)k) : 0.f;
    dst[i] = __float2bfloat16_rn(v);
  }
}

// batched variant: one launch packs every job of a device-resident table (blockIdx.y = job)
constexpr unsigned kPackBlocksPerJob = 384;
__global__ void __launch_bounds__(256) pack_weights_batch_kernel(const long long* __restrict__ jobs) {
  pdl_prologue();
  const long long* j = jobs + (size_t)blockIdx.y * 10;
  const float* src = reinterpret_cast<const float*>(j[0]);
  __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(j[1]);
  const int src_layout = (int)j[2], n_is_out = (int)j[3], out_ch = (int)j[4], in_ch = (int)j[5];
  const int taps = (int)(j[6] * j[7]), n_pad = (int)j[8], k_pad = (int)j[9];
  // 32-bit index arithmetic (the host checks total < 2^31); the grid is wide enough that a thread of the largest
  // conv (512x512x3x3) handles a few tens of elements -- the strided source reads are latency bound
  const unsigned total = (unsigned)(taps + 1) * (unsigned)n_pad * (unsigned)k_pad;
  const unsigned plane = (unsigned)k_pad * (unsigned)n_pad;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned t = i / plane, rem = i - t * plane;
    const unsigned nn = rem / (unsigned)k_pad, k = rem - nn * (unsigned)k_pad;
    const float v = (int)t < taps ? pack_value(src, src_layout, n_is_out, out_ch, in_ch, taps, (int)t, (int)nn, (int)k) : 0.f;
    dst[i] = __float2bfloat16_rn(v);
  }
}

static inline unsigned blocks_for(long total) {
  long nb = (total + 255) / 256;
  const long cap = (long)sm_count() * 16;
  if (nb > cap) nb = cap;
  if (nb < 1) nb = 1;
  return (unsigned)nb;
}

// ------------------------------------------------------------------ accuracy counters
// pred = argmax_c logp[b,c,pixel] (first maximum, as torch.max); counts[c] += (target == c),
// counts[ncls + c] += (target == c && pred == c).  Integer counters: exact and order independent.
constexpr int kAccMaxCls = 16;
__global__ void __launch_bounds__(256) accuracy_counts_kernel(const float* __restrict__ logp,
                                                              const long long* __restrict__ target, int n, int ncls,
                                                              int hw, unsigned long long* __restrict__ counts) {
  pdl_prologue();
  __shared__ unsigned int s_cnt[2 * kAccMaxCls];
  if (threadIdx.x < 2 * kAccMaxCls) s_cnt[threadIdx.x] = 0u;
  __syncthreads();
  const long total = (long)n * hw;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const long b = i / hw, px = i % hw;
    const float* lp = logp + b * (long)ncls * hw + px;
    float best = lp[0];
    int arg = 0;
    for (int c = 1; c < ncls; ++c) {
      const float v = lp[(long)c * hw];
      if (v > best) { best = v; arg = c; }
    }
    const long long t = target[i];
    if (t >= 0 && t < ncls) {
      atomicAdd(&s_cnt[(int)t], 1u);
      if (arg == (int)t) atomicAdd(&s_cnt[ncls + (int)t], 1u);
    }
  }
  __syncthreads();
  if (threadIdx.x < 2 * ncls && s_cnt[threadIdx.x]) atomicAdd(&counts[threadIdx.x], (unsigned long long)s_cnt[threadIdx.x]);
}

}  // namespace ubr

using namespace ubr;

extern "C" {

int ubr_pack_weights(const float* src, int src_layout, int n_is_out, int out_ch, int in_ch, int kh, int kw, int n_pad,
                     int k_pad, void* dst_bf16, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(src && dst_bf16, "pack_weights: null pointer");
  UBR_REQUIRE(src_layout >= 0 && src_layout <= 2, "pack_weights: bad src_layout");
  UBR_REQUIRE(n_pad % 16 == 0 && k_pad % 16 == 0, "pack_weights: pads must be multiples of 16");
  const int oc_eff = src_layout == 2 ? 4 * out_ch : out_ch;
  const int nn = n_is_out ? oc_eff : in_ch, kk = n_is_out ? in_ch : oc_eff;
  UBR_REQUIRE(n_pad >= nn && k_pad >= kk, "pack_weights: pads too small");
  UBR_REQUIRE(src_layout != 2 || (kh == 3 && kw == 3), "pack_weights: sub-pixel deconv pack is 3x3");
  const long total = (long)(kh * kw + 1) * n_pad * k_pad;
  ::ubr::Launch(blocks_for(total), 256, 0, (cudaStream_t)stream)(pack_weights_kernel, src, src_layout, n_is_out, out_ch, in_ch, kh,
                                                                           kw, n_pad, k_pad, (__nv_bfloat16*)dst_bf16);
  UBR_LAUNCH_CHECK("pack_weights_kernel");
  return 0;
}

int ubr_pack_weights_batch(const int64_t* jobs_device, int njobs, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(jobs_device != nullptr && njobs >= 1 && njobs <= 65535, "pack_weights_batch: bad job table");
  ::ubr::Launch(dim3(kPackBlocksPerJob, (unsigned)njobs), 256, 0, (cudaStream_t)stream)(pack_weights_batch_kernel, (const long long*)jobs_device);
  UBR_LAUNCH_CHECK("pack_weights_batch_kernel");
  return 0;
}

int ubr_logsoftmax_fwd(const float* logits, int ld, int n, int ncls, int hw, float* logp, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(logits && logp, "logsoftmax_fwd: null pointer");
  UBR_REQUIRE(ncls >= 1 && ncls <= kMaxCls && ld >= ncls, "logsoftmax_fwd: ncls=%d ld=%d unsupported", ncls, ld);
  ::ubr::Launch(blocks_for((long)n * hw), 256, 0, (cudaStream_t)stream)(logsoftmax_fwd_kernel, logits, ld, n, ncls, hw, logp);
  UBR_LAUNCH_CHECK("logsoftmax_fwd_kernel");
  return 0;
}

int ubr_logsoftmax_bwd(const float* logp, const float* dlogp, int n, int ncls, int hw, ubr_view dlogits, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(logp && dlogp && dlogits.ptr, "logsoftmax_bwd: null pointer");
  UBR_REQUIRE(ncls >= 1 && ncls <= kMaxCls && dlogits.c >= ncls && dlogits.c % 8 == 0 && dlogits.ld % 8 == 0,
              "logsoftmax_bwd: bad dlogits view");
  UBR_REQUIRE((long)dlogits.n * dlogits.h * dlogits.w == (long)n * hw, "logsoftmax_bwd: pixel count mismatch");
  ::ubr::Launch(blocks_for((long)n * hw), 256, 0, (cudaStream_t)stream)(logsoftmax_bwd_kernel, logp, dlogp, n, ncls, hw, (__nv_bfloat16*)dlogits.ptr, dlogits.c, dlogits.ld);
  UBR_LAUNCH_CHECK("logsoftmax_bwd_kernel");
  return 0;
}

uint64_t ubr_nll_workspace_bytes(int n, int hw) {
  (void)n; (void)hw;
  return (uint64_t)kNllBlocks * sizeof(float);
}

int ubr_pixelwise_nll_fwd(const float* logp, const int64_t* target, const float* pixw, const float* classw, int n,
                          int ncls, int hw, int64_t ignore_index, void* workspace, float* loss, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(logp && target && pixw && workspace && loss, "pixelwise_nll_fwd: null pointer");
  UBR_REQUIRE(n > 0 && hw > 0 && ncls >= 1, "pixelwise_nll_fwd: bad sizes");
  cudaStream_t st = (cudaStream_t)stream;
  long nb = ((long)n * hw + 255) / 256;
  if (nb > kNllBlocks) nb = kNllBlocks;
  ::ubr::Launch((unsigned)nb, 256, 0, st)(nll_fwd_kernel, logp, (const long long*)target, pixw, classw, n, ncls, hw,
                                               (long long)ignore_index, (float*)workspace);
  UBR_LAUNCH_CHECK("nll_fwd_kernel");
  ::ubr::Launch(1, 32, 0, st)(nll_finalize_kernel, (const float*)workspace, (int)nb, 1.0 / ((double)n * hw), loss);
  UBR_LAUNCH_CHECK("nll_finalize_kernel");
  return 0;
}

int ubr_pixelwise_nll_bwd(const int64_t* target, const float* pixw, const float* classw, const float* gout, int n,
                          int ncls, int hw, int64_t ignore_index, float* dlogp, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(target && pixw && dlogp, "pixelwise_nll_bwd: null pointer");
  ::ubr::Launch(blocks_for((long)n * hw), 256, 0, (cudaStream_t)stream)(nll_bwd_kernel, (const long long*)target, pixw, classw, gout, n, ncls, hw, (long long)ignore_index,
      (float)(1.0 / ((double)n * hw)), dlogp);
  UBR_LAUNCH_CHECK("nll_bwd_kernel");
  return 0;
}

int ubr_tile_gather_rect(const float* plane, int rows, int cols, const int32_t* origins, int ntiles, int crop_h, int crop_w,
                         float* crops, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(plane && origins && crops, "tile_gather: null pointer");
  UBR_REQUIRE(rows > 0 && cols > 0 && ntiles > 0 && crop_h > 0 && crop_w > 0, "tile_gather: bad sizes");
  ::ubr::Launch(blocks_for((long)ntiles * crop_h * crop_w), 256, 0, (cudaStream_t)stream)(tile_gather_kernel, plane, rows, cols, origins, ntiles, crop_h, crop_w, crops);
  UBR_LAUNCH_CHECK("tile_gather_kernel");
  return 0;
}
int ubr_tile_gather(const float* plane, int rows, int cols, const int32_t* origins, int ntiles, int crop, float* crops,
                    void* stream) {
  return ubr_tile_gather_rect(plane, rows, cols, origins, ntiles, crop, crop, crops, stream);
}

int ubr_stitch_scatter_rect(const float* tiles, int nch, const int32_t* origins, int ntiles, int crop_h, int crop_w, int rows,
                            int cols, float* out, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(tiles && origins && out, "stitch_scatter: null pointer");
  UBR_REQUIRE(rows > 0 && cols > 0 && ntiles > 0 && crop_h > 0 && crop_w > 0 && nch > 0, "stitch_scatter: bad sizes");
  ::ubr::Launch(blocks_for((long)rows * cols), 256, 0, (cudaStream_t)stream)(stitch_scatter_kernel, tiles, nch, origins, ntiles,
                                                                                         crop_h, crop_w, rows, cols, out);
  UBR_LAUNCH_CHECK("stitch_scatter_kernel");
  return 0;
}
int ubr_stitch_scatter(const float* tiles, int nch, const int32_t* origins, int ntiles, int crop, int rows, int cols,
                       float* out, void* stream) {
  return ubr_stitch_scatter_rect(tiles, nch, origins, ntiles, crop, crop, rows, cols, out, stream);
}

int ubr_argmax_classes(const float* logp, int n, int ncls, int hw, uint8_t* classes, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(logp && classes && n > 0 && ncls >= 1 && ncls <= 255 && hw > 0, "argmax_classes: bad arguments");
  ::ubr::Launch(blocks_for((long)n * hw), 256, 0, (cudaStream_t)stream)(argmax_classes_kernel, logp, n, ncls, (long)hw, classes);
  UBR_LAUNCH_CHECK("argmax_classes_kernel");
  return 0;
}

int ubr_stitch_scatter_u8(const uint8_t* tiles, const int32_t* origins, int ntiles, int crop_h, int crop_w, int rows, int cols,
                          int fill, uint8_t* out, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(tiles && origins && out, "stitch_scatter_u8: null pointer");
  UBR_REQUIRE(rows > 0 && cols > 0 && ntiles > 0 && crop_h > 0 && crop_w > 0, "stitch_scatter_u8: bad sizes");
  ::ubr::Launch(blocks_for((long)rows * cols), 256, 0, (cudaStream_t)stream)(stitch_scatter_u8_kernel, tiles, origins, ntiles, crop_h,
                                                                                            crop_w, rows, cols, fill, out);
  UBR_LAUNCH_CHECK("stitch_scatter_u8_kernel");
  return 0;
}

int ubr_accuracy_counts(const float* logp, const int64_t* target, int n, int ncls, int hw, uint64_t* counts, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(logp != nullptr && target != nullptr && counts != nullptr, "accuracy_counts: null pointer");
  UBR_REQUIRE(n > 0 && hw > 0 && ncls >= 1 && ncls <= kAccMaxCls, "accuracy_counts: bad extents n=%d ncls=%d hw=%d", n, ncls, hw);
  UBR_CUDA_OK(cudaMemsetAsync(counts, 0, 2 * (size_t)ncls * sizeof(uint64_t), (cudaStream_t)stream));
  const long total = (long)n * hw;
  long nb = (total + 255) / 256;
  if (nb > 148 * 8) nb = 148 * 8;
  ::ubr::Launch((unsigned)nb, 256, 0, (cudaStream_t)stream)(accuracy_counts_kernel, logp, (const long long*)target, n, ncls, hw,
                                                                       (unsigned long long*)counts);
  UBR_LAUNCH_CHECK("accuracy_counts_kernel");
  return 0;
}

}  // extern "C"
