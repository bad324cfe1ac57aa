// Memory-bound NHWC bf16 kernels: BatchNorm statistics / affine / backward, residual block tail,
// max-pool, pointwise sums.  All are HBM-roofline kernels: 16-byte vector accesses (8 bf16 channels per
// thread), fp32 math, deterministic two-level reductions (per-block partials, fixed-order finalize).
#include "common.cuh"
#include "tree_reduce.cuh"

namespace ubr {

constexpr int kRedThreads = 256;
constexpr int kBwdRedBlocks = 296;   // 2 x 148: the backward reduce kernels keep 2 CTAs (x4 / x2 pixels in flight) per SM

struct DView {
  const __nv_bfloat16* ptr;
  int c, ld;
  long npix;
};
struct DViewW {
  __nv_bfloat16* ptr;
  int c, ld;
  long npix;
};

static inline DView dv(const ubr_view& v) { return DView{(const __nv_bfloat16*)v.ptr, v.c, v.ld, (long)v.n * v.h * v.w}; }
static inline DViewW dvw(const ubr_view& v) { return DViewW{(__nv_bfloat16*)v.ptr, v.c, v.ld, (long)v.n * v.h * v.w}; }

__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&f)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
// raw 16-byte load kept packed while in flight (4 registers instead of 8), unpacked at the point of use
__device__ __forceinline__ uint4 ldraw(const __nv_bfloat16* p) { return *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float (&f)[8]) {
  uint4 u;
  u.x = pack_bf16x2(f[0], f[1]); u.y = pack_bf16x2(f[2], f[3]);
  u.z = pack_bf16x2(f[4], f[5]); u.w = pack_bf16x2(f[6], f[7]);
  *reinterpret_cast<uint4*>(p) = u;
}
__device__ __forceinline__ void ldparam8(const float* p, int c0, float (&f)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p + c0));
  const float4 b = __ldg(reinterpret_cast<const float4*>(p + c0 + 4));
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
// round an fp32 value through bf16 (what a consumer reading the stored tensor would see)
__device__ __forceinline__ float rbf(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

// ---------------------------------------------------------------------------------------------
// block-level reduction of NS x 8 per-thread accumulators -> partials[block][NS][C]
// thread layout: cg = tid % (C/8) (channel group), pixel lane = tid / (C/8)
// ---------------------------------------------------------------------------------------------
template <int NS>
__device__ __forceinline__ void block_reduce_store(float (&acc)[NS][8], int c, float* partials) {
  extern __shared__ float red[];            // [blockDim.x][NS*8]
  const int cgs = c >> 3;
  const int lanes = blockDim.x / cgs;
  float* mine = red + (size_t)threadIdx.x * (NS * 8);
#pragma unroll
  for (int s = 0; s < NS; ++s)
#pragma unroll
    for (int j = 0; j < 8; ++j) mine[s * 8 + j] = acc[s][j];
  __syncthreads();
  float* out = partials + (size_t)blockIdx.x * NS * c;
  for (int o = threadIdx.x; o < cgs * NS * 8; o += blockDim.x) {
    const int cg = o / (NS * 8), r = o % (NS * 8);
    float sum = 0.f;
    for (int l = 0; l < lanes; ++l) sum += red[(size_t)(l * cgs + cg) * (NS * 8) + r];
    const int s = r >> 3, j = r & 7;
    out[s * c + cg * 8 + j] = sum;
  }
}

// Streaming kernels: a thread owns ONE channel group (8 channels; per-channel parameters live in registers) and walks
// pixels p0, p0 + ps, ... with U independent 16-byte loads per tensor in flight.  HBM latency under load is ~1.5 us
// (profiles/r01_ubench_tma_mma_desc.txt T1), so ~64 KB must be in flight per SM to reach the copy bandwidth; the grid
// is sized so that (threads x cgs) tiles evenly (gridDim * blockDim is a multiple of cgs).
struct EwIdx { int cg; long p0, ps; };
__device__ __forceinline__ EwIdx ew_index(int cgs) {
  const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long t = (long)gridDim.x * blockDim.x;
  return EwIdx{(int)(g % cgs), g / cgs, t / cgs};
}

// dynamic shared memory of a reduce kernel: the block_reduce_store scratch, reused by the in-launch finalize tail
static inline size_t red_smem(int ns, int c, bool tree) {
  const size_t a = (size_t)kRedThreads * ns * 8 * sizeof(float);
  const size_t b = tree ? tree_smem_bytes(ns * c, kRedThreads) : 0;
  return a > b ? a : b;
}

static inline int red_blocks(long npix, int c, int cap = kRedBlocks) {
  const int lanes = kRedThreads / (c >> 3);
  long nb = (npix + (long)lanes * 8 - 1) / ((long)lanes * 8);
  if (nb > cap) nb = cap;
  if (nb < 1) nb = 1;
  return (int)nb;
}

// ------------------------------------------------------------------ BN forward statistics
// tr.tickets != null: the last CTA to arrive finishes the reduction and the BatchNorm finalize (tree_reduce.cuh)
__global__ void __launch_bounds__(kRedThreads) bn_stats_kernel(DView y, float* partials, TreeRed tr, BnFinal fin) {
  pdl_prologue();
  const int cgs = y.c >> 3, lanes = blockDim.x / cgs;
  const int cg = threadIdx.x % cgs, pl = threadIdx.x / cgs;
  const long ps = (long)gridDim.x * lanes;
  float acc[2][8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { acc[0][j] = 0.f; acc[1][j] = 0.f; }
  for (long p = (long)blockIdx.x * lanes + pl; p < y.npix; p += 4 * ps) {
    uint4 rf[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long q = p + u * ps;
      rf[u] = q < y.npix ? ldraw(y.ptr + q * y.ld + cg * 8) : make_uint4(0u, 0u, 0u, 0u);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float f[8];
      unpack8(rf[u], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) { acc[0][j] += f[j]; acc[1][j] = fmaf(f[j], f[j], acc[1][j]); }
    }
  }
  block_reduce_store<2>(acc, y.c, partials);
  if (tr.tickets) {
    extern __shared__ __align__(16) double tree_smem[];       // the reduction scratch of block_reduce_store, now idle
    if (tree_arrive(tr, blockIdx.x, 1, 1, tree_smem)) tree_finish_bn(tree_smem, fin);
  }
}

// column sums of partials[nb][stride] for 32 consecutive columns starting at col0, by a (32 x 32) thread block:
// threadIdx.x = column, threadIdx.y = row lane (coalesced 128-byte row reads, fixed summation order; the <= 19
// loads of a thread are independent, so the kernel costs a few memory latencies instead of nb/8 of them).
constexpr int kColLanes = 32;
__device__ __forceinline__ double col_sum_32x8(const float* partials, int nb, int stride, int col, bool active) {
  __shared__ double sred[kColLanes][33];
  double s = 0.0;
  if (active) {
    float v[20];
#pragma unroll
    for (int i = 0; i < 20; ++i) {
      const int b = threadIdx.y + i * kColLanes;
      v[i] = b < nb ? partials[(size_t)b * stride + col] : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 20; ++i) s += (double)v[i];
    for (int b = threadIdx.y + 20 * kColLanes; b < nb; b += kColLanes) s += (double)partials[(size_t)b * stride + col];
  }
  sred[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  double t = 0.0;
#pragma unroll
  for (int l = 0; l < kColLanes; ++l) t += sred[l][threadIdx.x];
  __syncthreads();
  return t;
}

// grid = ceil(c/32), block = (32, 8)
__global__ void bn_finalize_kernel(const float* partials, int nb, int c, double count, const float* gamma,
                                   const float* beta, float* running_mean, float* running_var, float momentum,
                                   float eps, float* scale, float* shift, float* mean_out, float* invstd_out) {
  pdl_prologue();
  const int ch = blockIdx.x * 32 + threadIdx.x;
  const bool act = ch < c;
  const double s = col_sum_32x8(partials, nb, 2 * c, ch, act);
  const double q = col_sum_32x8(partials, nb, 2 * c, c + ch, act);
  if (!act || threadIdx.y != 0) return;
  const double mean = s / count;
  double var = q / count - mean * mean;
  if (var < 0.0) var = 0.0;
  const float invstd = (float)(1.0 / sqrt(var + (double)eps));
  const float g = gamma ? gamma[ch] : 1.f, b = beta ? beta[ch] : 0.f;
  scale[ch] = g * invstd;
  shift[ch] = b - (float)mean * g * invstd;
  if (mean_out) mean_out[ch] = (float)mean;
  if (invstd_out) invstd_out[ch] = invstd;
  if (running_mean) running_mean[ch] = (1.f - momentum) * running_mean[ch] + momentum * (float)mean;
  if (running_var) {
    const double unbiased = count > 1.0 ? var * count / (count - 1.0) : var;
    running_var[ch] = (1.f - momentum) * running_var[ch] + momentum * (float)unbiased;
  }
}

__global__ void bn_eval_affine_kernel(int c, const float* gamma, const float* beta, const float* rm, const float* rv,
                                      float eps, float* scale, float* shift) {
  pdl_prologue();
  const int ch = blockIdx.x * blockDim.x + threadIdx.x;
  if (ch >= c) return;
  const float invstd = rsqrtf(rv[ch] + eps);
  const float g = gamma ? gamma[ch] : 1.f, b = beta ? beta[ch] : 0.f;
  scale[ch] = g * invstd;
  shift[ch] = b - rm[ch] * g * invstd;
}

// ------------------------------------------------------------------ BN apply (+ReLU), block tail
__global__ void __launch_bounds__(256, 2) bn_act_kernel(DView y, const float* scale, const float* shift, int relu,
                                                        DViewW out) {
  pdl_prologue();
  const EwIdx ix = ew_index(y.c >> 3);
  float sc[8], sh[8];
  ldparam8(scale, ix.cg * 8, sc);
  ldparam8(shift, ix.cg * 8, sh);
  const float lo = relu ? 0.f : -INFINITY;
  for (long p = ix.p0; p < y.npix; p += 4 * ix.ps) {
    uint4 rf[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long q = p + u * ix.ps;
      if (q < y.npix) rf[u] = ldraw(y.ptr + q * y.ld + ix.cg * 8);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long q = p + u * ix.ps;
      if (q < y.npix) {
        float f[8];
        unpack8(rf[u], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = fmaxf(fmaf(f[j], sc[j], sh[j]), lo);
        store8(out.ptr + q * out.ld + ix.cg * 8, f);
      }
    }
  }
}

// out = relu( relu(y2*s2+b2) + (res*sr+br | res) )
__global__ void __launch_bounds__(256, 2) block_tail_kernel(DView y2, const float* s2, const float* b2, DView res,
                                                            const float* sr, const float* br, DViewW out) {
  pdl_prologue();
  const EwIdx ix = ew_index(y2.c >> 3);
  float a2[8], c2[8], ar[8], cr[8];
  ldparam8(s2, ix.cg * 8, a2);
  ldparam8(b2, ix.cg * 8, c2);
  if (sr) { ldparam8(sr, ix.cg * 8, ar); ldparam8(br, ix.cg * 8, cr); }
  else {
#pragma unroll
    for (int j = 0; j < 8; ++j) { ar[j] = 1.f; cr[j] = 0.f; }
  }
  for (long p = ix.p0; p < y2.npix; p += 4 * ix.ps) {
    uint4 rf[4], rr4[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long q = p + u * ix.ps;
      if (q < y2.npix) {
        rf[u] = ldraw(y2.ptr + q * y2.ld + ix.cg * 8);
        rr4[u] = ldraw(res.ptr + q * res.ld + ix.cg * 8);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long q = p + u * ix.ps;
      if (q < y2.npix) {
        float f[8], r[8];
        unpack8(rf[u], f);
        unpack8(rr4[u], r);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float z = fmaxf(fmaf(f[j], a2[j], c2[j]), 0.f);
          const float rr = sr ? fmaf(r[j], ar[j], cr[j]) : r[j];
          f[j] = fmaxf(z + rr, 0.f);
        }
        store8(out.ptr + q * out.ld + ix.cg * 8, f);
      }
    }
  }
}

// ------------------------------------------------------------------ BN(+ReLU) backward
// reduce: sums[0] = sum dz, sums[1] = sum dz*xhat
__global__ void __launch_bounds__(kRedThreads, 2) bn_act_bwd_reduce_kernel(DView dout, DView y, const float* scale,
                                                                           const float* shift, const float* mean,
                                                                           const float* invstd, int relu, float* partials,
                                                                           TreeRed tr, SumsFinal fin) {
  pdl_prologue();
  const int cgs = y.c >> 3, lanes = blockDim.x / cgs;
  const int cg = threadIdx.x % cgs, pl = threadIdx.x / cgs;
  const long ps = (long)gridDim.x * lanes;
  float sc[8], sh[8], mu[8], is[8];
  ldparam8(scale, cg * 8, sc); ldparam8(shift, cg * 8, sh);
  ldparam8(mean, cg * 8, mu); ldparam8(invstd, cg * 8, is);
  float acc[2][8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { acc[0][j] = 0.f; acc[1][j] = 0.f; }
  for (long p = (long)blockIdx.x * lanes + pl; p < y.npix; p += 4 * ps) {
    uint4 rg[4], rv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long q = p + u * ps;
      if (q < y.npix) {
        rg[u] = ldraw(dout.ptr + q * dout.ld + cg * 8);
        rv[u] = ldraw(y.ptr + q * y.ld + cg * 8);
      } else {
        rg[u] = make_uint4(0u, 0u, 0u, 0u);
        rv[u] = make_uint4(0u, 0u, 0u, 0u);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float g[8], v[8];
      unpack8(rg[u], g);
      unpack8(rv[u], v);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float z = fmaf(v[j], sc[j], sh[j]);
        const float dz = (!relu || z > 0.f) ? g[j] : 0.f;
        acc[0][j] += dz;
        acc[1][j] = fmaf(dz, (v[j] - mu[j]) * is[j], acc[1][j]);
      }
    }
  }
  block_reduce_store<2>(acc, y.c, partials);
  if (tr.tickets) {
    extern __shared__ __align__(16) double tree_smem[];       // the reduction scratch of block_reduce_store, now idle
    if (tree_arrive(tr, blockIdx.x, 1, 1, tree_smem)) tree_finish_sums(tree_smem, fin);
  }
}

// per-channel pixel sums (conv bias gradients): partials[nb][2][c] of bn_stats -> out[0:nvalid) (fixed order, double)
__global__ void channel_sum_finish_kernel(const float* partials, int nb, int c, int nvalid, float* out, int accumulate) {
  pdl_prologue();
  const int i = blockIdx.x * 32 + threadIdx.x;
  const double s = col_sum_32x8(partials, nb, 2 * c, i, i < c);
  if (i < nvalid && threadIdx.y == 0) out[i] = (accumulate ? out[i] : 0.f) + (float)s;
}

// bumps a list of int64 scalars (BatchNorm2d.num_batches_tracked of every layer) by one
__global__ void increment_i64_kernel(long long* const* ptrs, int n) {
  pdl_prologue();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) *ptrs[i] += 1;
}

// sums partials[nb][NS][C] -> sums[NS][C] (fixed order, double); grid = ceil(nsc/32), block = (32, 8)
struct ParamGradOut { float* p[4]; };    // per sum row: optional fp32 parameter-gradient vector of c entries (dbeta / dgamma)
__global__ void sums_finalize_kernel(const float* partials, int nb, int nsc, float* sums, int c, ParamGradOut out,
                                     int accumulate) {
  pdl_prologue();
  const int i = blockIdx.x * 32 + threadIdx.x;
  const double s = col_sum_32x8(partials, nb, nsc, i, i < nsc);
  if (i < nsc && threadIdx.y == 0) {
    sums[i] = (float)s;
    float* dst = out.p[i / c];
    if (dst) dst[i % c] = (accumulate ? dst[i % c] : 0.f) + (float)s;
  }
}

// dy = scale * (dz - S0/N - xhat * S1/N) with xhat = (y - mean) * invstd, regrouped per channel as
// dy = scale * dz + A * y + B,  A = -scale * invstd * S1/N,  B = -scale * S0/N - A * mean   (gamma*invstd == scale)
__global__ void __launch_bounds__(256, 2) bn_act_bwd_apply_kernel(DView dout, DView y, const float* scale,
                                                                  const float* shift, const float* mean,
                                                                  const float* invstd, const float* gamma, int relu,
                                                                  const float* sums, float inv_count, DViewW dy) {
  pdl_prologue();
  const EwIdx ix = ew_index(y.c >> 3);
  float sc[8], sh[8], ca[8], cb[8];
  {
    float mu[8], is[8], s0[8], s1[8];
    ldparam8(scale, ix.cg * 8, sc); ldparam8(shift, ix.cg * 8, sh);
    ldparam8(mean, ix.cg * 8, mu); ldparam8(invstd, ix.cg * 8, is);
    ldparam8(sums, ix.cg * 8, s0); ldparam8(sums + y.c, ix.cg * 8, s1);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      ca[j] = -sc[j] * is[j] * s1[j] * inv_count;
      cb[j] = -sc[j] * s0[j] * inv_count - ca[j] * mu[j];
    }
  }
  (void)gamma;
  for (long p = ix.p0; p < y.npix; p += 4 * ix.ps) {
    uint4 rg[4], rv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long q = p + u * ix.ps;
      if (q < y.npix) {
        rg[u] = ldraw(dout.ptr + q * dout.ld + ix.cg * 8);
        rv[u] = ldraw(y.ptr + q * y.ld + ix.cg * 8);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long q = p + u * ix.ps;
      if (q < y.npix) {
        float o[8], g[8], v[8];
        unpack8(rg[u], g);
        unpack8(rv[u], v);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float z = fmaf(v[j], sc[j], sh[j]);
          const float dz = (!relu || z > 0.f) ? g[j] : 0.f;
          o[j] = fmaf(sc[j], dz, fmaf(ca[j], v[j], cb[j]));
        }
        store8(dy.ptr + q * dy.ld + ix.cg * 8, o);
      }
    }
  }
}

// ------------------------------------------------------------------ block tail backward
// du = dout*(out>0);  dz2 = du*(bn2(y2)>0)
// sums: [0] sum dz2, [1] sum dz2*xhat2, [2] sum du, [3] sum du*xhat_r   (2,3 only if res_is_bn)
template <bool RES_BN>
__global__ void __launch_bounds__(kRedThreads, 2) tail_bwd_reduce_kernel(DView dout, DView out, DView y2, const float* s2,
                                                                      const float* b2, const float* mean2,
                                                                      const float* invstd2, DView res,
                                                                      const float* meanr, const float* invstdr,
                                                                      const float* sr, const float* br,
                                                                      float* partials, TreeRed tr, SumsFinal fin) {
  pdl_prologue();
  // RES_BN: the mask of the block's final ReLU is recomputed from y2 and the bypass output with the forward pass's own
  // arithmetic (block_tail_kernel) instead of reading `out` back
  constexpr int NS = RES_BN ? 4 : 2;
  constexpr int U = RES_BN ? 1 : 2;      // 4 tensors in flight per pixel group with the residual BN: no unroll (registers)
  const int cgs = y2.c >> 3, lanes = blockDim.x / cgs;
  const int cg = threadIdx.x % cgs, pl = threadIdx.x / cgs;
  float sc[8], sh[8], mu[8], is[8], mur[8], isr[8];
  ldparam8(s2, cg * 8, sc); ldparam8(b2, cg * 8, sh);
  ldparam8(mean2, cg * 8, mu); ldparam8(invstd2, cg * 8, is);
  float ar[8], cr[8];
  if (RES_BN) { ldparam8(meanr, cg * 8, mur); ldparam8(invstdr, cg * 8, isr); ldparam8(sr, cg * 8, ar); ldparam8(br, cg * 8, cr); }
  float acc[NS][8];
#pragma unroll
  for (int s = 0; s < NS; ++s)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[s][j] = 0.f;
  const long ps = (long)gridDim.x * lanes;
  for (long p = (long)blockIdx.x * lanes + pl; p < y2.npix; p += U * ps) {
    uint4 rg[U], ro[U], rv[U], rr4[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long q = p + u * ps;
      if (q < y2.npix) {
        rg[u] = ldraw(dout.ptr + q * dout.ld + cg * 8);
        ro[u] = RES_BN ? make_uint4(0u, 0u, 0u, 0u) : ldraw(out.ptr + q * out.ld + cg * 8);
        rv[u] = ldraw(y2.ptr + q * y2.ld + cg * 8);
        rr4[u] = RES_BN ? ldraw(res.ptr + q * res.ld + cg * 8) : make_uint4(0u, 0u, 0u, 0u);
      } else {
        rg[u] = make_uint4(0u, 0u, 0u, 0u); ro[u] = make_uint4(0u, 0u, 0u, 0u); rv[u] = make_uint4(0u, 0u, 0u, 0u); rr4[u] = make_uint4(0u, 0u, 0u, 0u);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float g[8], o[8], v[8], r[8];
      unpack8(rg[u], g);
      unpack8(ro[u], o);
      unpack8(rv[u], v);
      unpack8(rr4[u], r);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float z = fmaf(v[j], sc[j], sh[j]);
        const bool on = RES_BN ? (fmaxf(z, 0.f) + fmaf(r[j], ar[j], cr[j]) > 0.f) : (o[j] > 0.f);
        const float du = on ? g[j] : 0.f;
        const float dz = z > 0.f ? du : 0.f;
        acc[0][j] += dz;
        acc[1][j] = fmaf(dz, (v[j] - mu[j]) * is[j], acc[1][j]);
        if (RES_BN) {
          acc[2][j] += du;
          acc[3][j] = fmaf(du, (r[j] - mur[j]) * isr[j], acc[3][j]);
        }
      }
    }
  }
  block_reduce_store<NS>(acc, y2.c, partials);
  if (tr.tickets) {
    extern __shared__ __align__(16) double tree_smem[];       // the reduction scratch of block_reduce_store, now idle
    if (tree_arrive(tr, blockIdx.x, 1, 1, tree_smem)) tree_finish_sums(tree_smem, fin);
  }
}

// dy2 = s2 * dz2 + A2 * y2 + B2 (see bn_act_bwd_apply_kernel);  dres = du | Gr * du + Ar * res + Br with
// Gr = gamma_r * invstd_r, Ar = -Gr * invstd_r * S3/N, Br = -Gr * S2/N - Ar * mean_r
template <bool RES_BN>
__global__ void __launch_bounds__(256, 2) tail_bwd_apply_kernel(DView dout, DView out, DView y2, const float* s2,
                                                                const float* b2, const float* mean2,
                                                                const float* invstd2, DView res, const float* meanr,
                                                                const float* invstdr, const float* gammar,
                                                                const float* sr, const float* br,
                                                                const float* sums, float inv_count, DViewW dy2,
                                                                DViewW dres) {
  pdl_prologue();
  const int c = y2.c;
  const EwIdx ix = ew_index(c >> 3);
  float sc[8], sh[8], ca[8], cb[8], gr[8], ra[8], rb[8], ar[8], cr[8];
  if (RES_BN) { ldparam8(sr, ix.cg * 8, ar); ldparam8(br, ix.cg * 8, cr); }
  {
    float mu[8], is[8], t0[8], t1[8];
    ldparam8(s2, ix.cg * 8, sc); ldparam8(b2, ix.cg * 8, sh);
    ldparam8(mean2, ix.cg * 8, mu); ldparam8(invstd2, ix.cg * 8, is);
    ldparam8(sums, ix.cg * 8, t0); ldparam8(sums + c, ix.cg * 8, t1);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      ca[j] = -sc[j] * is[j] * t1[j] * inv_count;
      cb[j] = -sc[j] * t0[j] * inv_count - ca[j] * mu[j];
    }
    if (RES_BN) {
      ldparam8(meanr, ix.cg * 8, mu); ldparam8(invstdr, ix.cg * 8, is); ldparam8(gammar, ix.cg * 8, gr);
      ldparam8(sums + 2 * c, ix.cg * 8, t0); ldparam8(sums + 3 * c, ix.cg * 8, t1);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        gr[j] *= is[j];
        ra[j] = -gr[j] * is[j] * t1[j] * inv_count;
        rb[j] = -gr[j] * t0[j] * inv_count - ra[j] * mu[j];
      }
    }
  }
  for (long p = ix.p0; p < y2.npix; p += 2 * ix.ps) {
    uint4 rg[2], ro[2], rv[2], rr4[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const long q = p + u * ix.ps;
      if (q < y2.npix) {
        rg[u] = ldraw(dout.ptr + q * dout.ld + ix.cg * 8);
        if (!RES_BN) ro[u] = ldraw(out.ptr + q * out.ld + ix.cg * 8);
        rv[u] = ldraw(y2.ptr + q * y2.ld + ix.cg * 8);
        if (RES_BN) rr4[u] = ldraw(res.ptr + q * res.ld + ix.cg * 8);
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const long q = p + u * ix.ps;
      if (q < y2.npix) {
        float du[8], d2[8], g[8], o[8], v[8], r[8];
        unpack8(rg[u], g);
        if (!RES_BN) unpack8(ro[u], o);
        unpack8(rv[u], v);
        if (RES_BN) unpack8(rr4[u], r);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float z = fmaf(v[j], sc[j], sh[j]);
          const bool on = RES_BN ? (fmaxf(z, 0.f) + fmaf(r[j], ar[j], cr[j]) > 0.f) : (o[j] > 0.f);
          du[j] = on ? g[j] : 0.f;
          const float dz = z > 0.f ? du[j] : 0.f;
          d2[j] = fmaf(sc[j], dz, fmaf(ca[j], v[j], cb[j]));
        }
        store8(dy2.ptr + q * dy2.ld + ix.cg * 8, d2);
        if (RES_BN) {
#pragma unroll
          for (int j = 0; j < 8; ++j) du[j] = fmaf(gr[j], du[j], fmaf(ra[j], r[j], rb[j]));
        }
        store8(dres.ptr + q * dres.ld + ix.cg * 8, du);
      }
    }
  }
}

// ------------------------------------------------------------------ pointwise sum
__global__ void __launch_bounds__(256, 2) add_kernel(DView a, DView b, DView c, DViewW out) {
  pdl_prologue();
  const EwIdx ix = ew_index(a.c >> 3);
  for (long p = ix.p0; p < a.npix; p += 4 * ix.ps) {
    uint4 rx[4], ry[4], rz[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long q = p + u * ix.ps;
      if (q < a.npix) {
        rx[u] = ldraw(a.ptr + q * a.ld + ix.cg * 8);
        if (b.ptr) ry[u] = ldraw(b.ptr + q * b.ld + ix.cg * 8);
        if (c.ptr) rz[u] = ldraw(c.ptr + q * c.ld + ix.cg * 8);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long q = p + u * ix.ps;
      if (q < a.npix) {
        float x[8], y[8];
        unpack8(rx[u], x);
        if (b.ptr) {
          unpack8(ry[u], y);
#pragma unroll
          for (int j = 0; j < 8; ++j) x[j] += y[j];
        }
        if (c.ptr) {
          unpack8(rz[u], y);
#pragma unroll
          for (int j = 0; j < 8; ++j) x[j] += y[j];
        }
        store8(out.ptr + q * out.ld + ix.cg * 8, x);
      }
    }
  }
}

// ------------------------------------------------------------------ MaxPool 3x3 / stride 2 / pad 1
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const __nv_bfloat16* x, int xld, int n, int h, int w, int c,
                                                          __nv_bfloat16* y, int yld) {
  pdl_prologue();
  const int cgs = c >> 3, ho = h / 2, wo = w / 2;
  const long total = (long)n * ho * wo * cgs;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int cg = (int)(i % cgs);
    long p = i / cgs;
    const int ow = (int)(p % wo); p /= wo;
    const int oh = (int)(p % ho);
    const int b = (int)(p / ho);
    float m[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) m[j] = -INFINITY;
    for (int kh = 0; kh < 3; ++kh) {
      const int ih = 2 * oh - 1 + kh;
      if (ih < 0 || ih >= h) continue;
      for (int kw = 0; kw < 3; ++kw) {
        const int iw = 2 * ow - 1 + kw;
        if (iw < 0 || iw >= w) continue;
        float f[8];
        load8(x + (((long)b * h + ih) * w + iw) * xld + cg * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) m[j] = fmaxf(m[j], f[j]);
      }
    }
    store8(y + (((long)b * ho + oh) * wo + ow) * yld + cg * 8, m);
  }
}

// forward that also records the argmax tap (kh*3+kw of the FIRST maximum in scan order, as ATen) per element
template <typename I>                  // I: index type of the flat element loop (see maxpool_bwd_idx_kernel)
__global__ void __launch_bounds__(256) maxpool_fwd_idx_kernel(const __nv_bfloat16* x, int xld, int n, int h, int w, int c,
                                                              __nv_bfloat16* y, int yld, uint8_t* __restrict__ idx) {
  pdl_prologue();
  const int cgs = c >> 3, ho = h / 2, wo = w / 2;
  const I total = (I)n * (I)ho * (I)wo * (I)cgs;
  for (I i = (I)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (I)gridDim.x * blockDim.x) {
    const int cg = (int)(i % (I)cgs);
    I p = i / (I)cgs;
    const int ow = (int)(p % (I)wo); p /= (I)wo;
    const int oh = (int)(p % (I)ho);
    const int b = (int)(p / (I)ho);
    float m[8];
    uint32_t arg[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { m[j] = -INFINITY; arg[j] = 255u; }
    for (int kh = 0; kh < 3; ++kh) {
      const int ih = 2 * oh - 1 + kh;
      if (ih < 0 || ih >= h) continue;
      for (int kw = 0; kw < 3; ++kw) {
        const int iw = 2 * ow - 1 + kw;
        if (iw < 0 || iw >= w) continue;
        float f[8];
        load8(x + (((long)b * h + ih) * w + iw) * xld + cg * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (f[j] > m[j] || arg[j] == 255u) { m[j] = f[j]; arg[j] = (uint32_t)(kh * 3 + kw); }
      }
    }
    store8(y + (((long)b * ho + oh) * wo + ow) * yld + cg * 8, m);
    uint2 u;
    u.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
    u.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
    *reinterpret_cast<uint2*>(idx + (size_t)i * 8) = u;  // idx layout [n][ho][wo][c], c contiguous
  }
}

// backward from the recorded argmax: each input pixel gathers from the <= 4 windows covering it (no atomics)
// I: index type of the flat element loop (unsigned 32-bit where the tensor allows it: the three div/mod pairs per element
// are a third of the cost of their 64-bit forms)
template <typename I>
__global__ void __launch_bounds__(256) maxpool_bwd_idx_kernel(const uint8_t* __restrict__ idx, int n, int h, int w, int c,
                                                              const __nv_bfloat16* dy, int dyld, __nv_bfloat16* dx,
                                                              int dxld, int accumulate) {
  pdl_prologue();
  const int cgs = c >> 3, ho = h / 2, wo = w / 2;
  const I total = (I)n * (I)h * (I)w * (I)cgs;
  for (I i = (I)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (I)gridDim.x * blockDim.x) {
    const int cg = (int)(i % (I)cgs);
    I p = i / (I)cgs;
    const int iw = (int)(p % (I)w); p /= (I)w;
    const int ih = (int)(p % (I)h);
    const int b = (int)(p / (I)h);
    // the <= 4 covering windows: every index load is issued before any of them is inspected, then the dy loads of
    // the windows that hit -- independent loads instead of a load -> test -> load chain per window
    const int oh_lo = ih / 2, ow_lo = iw / 2;
    long q[4];
    uint2 u[4];
    uint32_t hit[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int oh = oh_lo + (k >> 1), ow = ow_lo + (k & 1);
      const bool valid = oh <= (ih + 1) / 2 && ow <= (iw + 1) / 2 && oh < ho && ow < wo;
      q[k] = valid ? ((long)b * ho + oh) * wo + ow : -1;
      u[k] = valid ? *reinterpret_cast<const uint2*>(idx + (q[k] * cgs + cg) * 8) : make_uint2(0xffffffffu, 0xffffffffu);
    }
    float acc[8];
    if (accumulate) load8(dx + (((long)b * h + ih) * w + iw) * dxld + cg * 8, acc);
    else {
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    }
    uint32_t e0[4], e1[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int oh = oh_lo + (k >> 1), ow = ow_lo + (k & 1);
      const uint32_t me = (uint32_t)((ih - (2 * oh - 1)) * 3 + (iw - (2 * ow - 1))) & 0xffu;
      const uint32_t m4 = me * 0x01010101u;
      e0[k] = u[k].x ^ m4; e1[k] = u[k].y ^ m4;            // a zero byte <=> this input is that channel's argmax
      hit[k] = q[k] >= 0 ? (((e0[k] - 0x01010101u) & ~e0[k] & 0x80808080u) | ((e1[k] - 0x01010101u) & ~e1[k] & 0x80808080u)) : 0u;
    }
    uint4 graw[4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
      graw[k] = hit[k] ? ldraw(dy + q[k] * dyld + cg * 8) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (hit[k]) {
        float g[8];
        unpack8(graw[k], g);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (((e0[k] >> (8 * j)) & 0xffu) == 0u) acc[j] += g[j];
          if (((e1[k] >> (8 * j)) & 0xffu) == 0u) acc[4 + j] += g[4 + j];
        }
      }
    }
    store8(dx + (((long)b * h + ih) * w + iw) * dxld + cg * 8, acc);
  }
}

// ------------------------------------------------------------------ MaxPool 3x3 / stride 1 / pad 1
// ASPP block 5 (reference models/ASPP_ResNet.py:217-219, 239).  idx (optional) records the argmax tap like the
// stride-2 kernel so the backward is a gather over the <= 9 windows that cover an input pixel.
__global__ void __launch_bounds__(256) maxpool3s1_fwd_kernel(const __nv_bfloat16* x, int xld, int n, int h, int w, int c,
                                                             __nv_bfloat16* y, int yld, uint8_t* __restrict__ idx) {
  pdl_prologue();
  const int cgs = c >> 3;
  const long total = (long)n * h * w * cgs;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int cg = (int)(i % cgs);
    long p = i / cgs;
    const int ow = (int)(p % w); p /= w;
    const int oh = (int)(p % h);
    const int b = (int)(p / h);
    float m[8];
    uint32_t arg[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { m[j] = -INFINITY; arg[j] = 255u; }
    for (int kh = 0; kh < 3; ++kh) {
      const int ih = oh - 1 + kh;
      if (ih < 0 || ih >= h) continue;
      for (int kw = 0; kw < 3; ++kw) {
        const int iw = ow - 1 + kw;
        if (iw < 0 || iw >= w) continue;
        float f[8];
        load8(x + (((long)b * h + ih) * w + iw) * xld + cg * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (f[j] > m[j] || arg[j] == 255u) { m[j] = f[j]; arg[j] = (uint32_t)(kh * 3 + kw); }
      }
    }
    store8(y + (((long)b * h + oh) * w + ow) * yld + cg * 8, m);
    if (idx) {
      uint2 u;
      u.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
      u.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
      *reinterpret_cast<uint2*>(idx + i * 8) = u;          // idx layout [n][h][w][c], c contiguous
    }
  }
}

__global__ void __launch_bounds__(256) maxpool3s1_bwd_idx_kernel(const uint8_t* __restrict__ idx, int n, int h, int w, int c,
                                                                 const __nv_bfloat16* dy, int dyld, __nv_bfloat16* dx,
                                                                 int dxld, int accumulate) {
  pdl_prologue();
  const int cgs = c >> 3;
  const long total = (long)n * h * w * cgs;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int cg = (int)(i % cgs);
    long p = i / cgs;
    const int iw = (int)(p % w); p /= w;
    const int ih = (int)(p % h);
    const int b = (int)(p / h);
    float acc[8];
    if (accumulate) load8(dx + (((long)b * h + ih) * w + iw) * dxld + cg * 8, acc);
    else {
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    }
    for (int oh = ih - 1; oh <= ih + 1; ++oh) {
      if (oh < 0 || oh >= h) continue;
      for (int ow = iw - 1; ow <= iw + 1; ++ow) {
        if (ow < 0 || ow >= w) continue;
        const long q = ((long)b * h + oh) * w + ow;
        const uint2 u = *reinterpret_cast<const uint2*>(idx + (q * cgs + cg) * 8);
        const uint32_t me = (uint32_t)((ih - oh + 1) * 3 + (iw - ow + 1));
        const uint32_t m4 = me * 0x01010101u;
        const uint32_t e0 = u.x ^ m4, e1 = u.y ^ m4;       // a zero byte <=> this input is that channel's argmax
        if (((e0 - 0x01010101u) & ~e0 & 0x80808080u) | ((e1 - 0x01010101u) & ~e1 & 0x80808080u)) {
          float g[8];
          load8(dy + q * dyld + cg * 8, g);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (((e0 >> (8 * j)) & 0xffu) == 0u) acc[j] += g[j];
            if (((e1 >> (8 * j)) & 0xffu) == 0u) acc[4 + j] += g[4 + j];
          }
        }
      }
    }
    store8(dx + (((long)b * h + ih) * w + iw) * dxld + cg * 8, acc);
  }
}

// ------------------------------------------------------------------ pixel unshuffle (ConvTranspose backward)
// y4[n, h, w, (ph*2+pw)*c + co] = y[n, 2h+ph, 2w+pw, co]: the gradient of a ConvTranspose2d(k4,s2,p1) output regrouped
// on the INPUT grid, so that its data / weight gradients become a 3x3 stride-1 conv / wgrad with 4*c channels
// (the transpose of the sub-pixel forward, ubr_deconv4s2_fwd_subpixel).
__global__ void __launch_bounds__(256) pixel_unshuffle2_kernel(const __nv_bfloat16* __restrict__ y, int yld, int n, int h,
                                                               int w, int c, __nv_bfloat16* __restrict__ y4, int y4ld) {
  pdl_prologue();
  const int cgs = c >> 3;                       // 16-byte groups per source pixel
  const long total = (long)n * (2 * h) * (2 * w) * cgs;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int cg = (int)(i % cgs);
    long p = i / cgs;
    const int ow = (int)(p % (2 * w)); p /= (2 * w);
    const int oh = (int)(p % (2 * h));
    const int b = (int)(p / (2 * h));
    const uint4 v = *reinterpret_cast<const uint4*>(y + (((long)b * 2 * h + oh) * (2 * w) + ow) * yld + cg * 8);
    const int ph = oh & 1, pw = ow & 1;
    *reinterpret_cast<uint4*>(y4 + (((long)b * h + (oh >> 1)) * w + (ow >> 1)) * y4ld + (ph * 2 + pw) * c + cg * 8) = v;
  }
}

// dw[ci][co][kh][kw] (+)= dw3[(ph*2+pw)*out_ch + co][ci][r][s] with kh = ph + 1 - 2*(r-1), kw = pw + 1 - 2*(s-1)
// (the inverse of pack src_layout 2: every 4x4 tap is exactly one (phase, 3x3 tap) pair)
__global__ void deconv_wgrad_gather_kernel(const float* __restrict__ dw3, int in_ch, int out_ch, float* __restrict__ dw,
                                           int accumulate) {
  pdl_prologue();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= in_ch * out_ch * 16) return;
  const int kw = i & 3, kh = (i >> 2) & 3, co = (i >> 4) % out_ch, ci = (i >> 4) / out_ch;
  const int ph = (kh + 1) & 1, pw = (kw + 1) & 1;
  const int r = (ph + 1 - kh) / 2 + 1, s = (pw + 1 - kw) / 2 + 1;
  const float v = dw3[(((long)((ph * 2 + pw) * out_ch + co) * in_ch + ci) * 3 + r) * 3 + s];
  dw[i] = accumulate ? dw[i] + v : v;
}

// gather form of the backward: each input pixel re-derives the argmax (first maximum in (kh,kw) scan
// order, as ATen's max_pool2d) of every window that covers it.  Deterministic, no atomics.
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const __nv_bfloat16* x, int xld, int n, int h, int w, int c,
                                                          const __nv_bfloat16* dy, int dyld, __nv_bfloat16* dx,
                                                          int dxld, int accumulate) {
  pdl_prologue();
  const int cgs = c >> 3, ho = h / 2, wo = w / 2;
  const long total = (long)n * h * w * cgs;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int cg = (int)(i % cgs);
    long p = i / cgs;
    const int iw = (int)(p % w); p /= w;
    const int ih = (int)(p % h);
    const int b = (int)(p / h);
    float acc[8];
    if (accumulate) load8(dx + (((long)b * h + ih) * w + iw) * dxld + cg * 8, acc);
    else {
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    }
    // windows oh with 2*oh-1 <= ih <= 2*oh+1
    const int oh_lo = ih / 2, oh_hi = (ih + 1) / 2;
    const int ow_lo = iw / 2, ow_hi = (iw + 1) / 2;
    for (int oh = oh_lo; oh <= oh_hi; ++oh) {
      if (oh >= ho) continue;
      for (int ow = ow_lo; ow <= ow_hi; ++ow) {
        if (ow >= wo) continue;
        float best[8];
        int arg[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; arg[j] = -1; }
        for (int kh = 0; kh < 3; ++kh) {
          const int yh = 2 * oh - 1 + kh;
          if (yh < 0 || yh >= h) continue;
          for (int kw = 0; kw < 3; ++kw) {
            const int yw = 2 * ow - 1 + kw;
            if (yw < 0 || yw >= w) continue;
            float f[8];
            load8(x + (((long)b * h + yh) * w + yw) * xld + cg * 8, f);
#pragma unroll
            for (int j = 0; j < 8; ++j)
              if (f[j] > best[j] || arg[j] < 0) { best[j] = f[j]; arg[j] = kh * 3 + kw; }
          }
        }
        const int me = (ih - (2 * oh - 1)) * 3 + (iw - (2 * ow - 1));
        float g[8];
        load8(dy + (((long)b * ho + oh) * wo + ow) * dyld + cg * 8, g);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (arg[j] == me) acc[j] += g[j];
      }
    }
    store8(dx + (((long)b * h + ih) * w + iw) * dxld + cg * 8, acc);
  }
}

__global__ void scale_f32_kernel(float* buf, unsigned long long count, float alpha) {
  pdl_prologue();
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < count;
       i += (unsigned long long)gridDim.x * blockDim.x)
    buf[i] *= alpha;
}

// ------------------------------------------------------------------ host helpers
static int chk(const ubr_view& v, const char* name) {
  UBR_REQUIRE(v.ptr != nullptr, "%s: null pointer", name);
  UBR_REQUIRE(v.n > 0 && v.h > 0 && v.w > 0 && v.c > 0 && v.c % 8 == 0, "%s: bad extents (c=%d)", name, v.c);
  UBR_REQUIRE(v.ld >= v.c && v.ld % 8 == 0, "%s: bad ld=%d (c=%d)", name, v.ld, v.c);
  UBR_REQUIRE(((uintptr_t)v.ptr & 15) == 0, "%s: pointer must be 16-byte aligned", name);
  return 0;
}
static int same_shape(const ubr_view& a, const ubr_view& b, const char* what) {
  UBR_REQUIRE(a.n == b.n && a.h == b.h && a.w == b.w && a.c == b.c, "%s: shape mismatch (%d,%d,%d,%d) vs (%d,%d,%d,%d)",
              what, a.n, a.h, a.w, a.c, b.n, b.h, b.w, b.c);
  return 0;
}
static int chk_red(const ubr_view& v, const char* name) {
  if (int rc = chk(v, name)) return rc;
  const int cgs = v.c >> 3;
  UBR_REQUIRE(cgs <= kRedThreads && kRedThreads % cgs == 0, "%s: c=%d must divide %d (power-of-two multiples of 8)",
              name, v.c, kRedThreads * 8);
  return 0;
}
// grid of the fixed-channel-group streaming kernels (ew_index): `unroll` pixels per thread per trip, at most 4 CTAs
// per SM, and gridDim * 256 a multiple of cgs so that every thread keeps one channel group
static inline unsigned ew_grid(long npix, int cgs, int unroll) {
  long nb = (npix * cgs + 256L * unroll - 1) / (256L * unroll);
  const long cap = (long)sm_count() * 4;
  if (nb > cap) nb = cap;
  if (nb < 1) nb = 1;
  int g = cgs, r = 256;
  while (r) { const int t = g % r; g = r; r = t; }          // g = gcd(cgs, 256)
  const long m = cgs / g;
  nb = (nb + m - 1) / m * m;
  return (unsigned)nb;
}
static inline unsigned ew_blocks(long total) {
  long nb = (total + 255) / 256;
  const long cap = (long)sm_count() * 16;
  if (nb > cap) nb = cap;
  if (nb < 1) nb = 1;
  return (unsigned)nb;
}

}  // namespace ubr

using namespace ubr;

extern "C" {

uint64_t ubr_bn_workspace_bytes(int c) {
  // partial rows [kRedBlocks][4][c] + finalised sums [4][c]
  return (uint64_t)(kRedBlocks + 1) * 4 * (uint64_t)c * sizeof(float);
}

int ubr_bn_stats(ubr_view y, void* workspace, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk_red(y, "bn_stats.y")) return rc;
  UBR_REQUIRE(workspace != nullptr, "bn_stats: null workspace");
  DView v = dv(y);
  const int nb = red_blocks(v.npix, v.c);
  ::ubr::Launch(nb, kRedThreads, kRedThreads * 16 * sizeof(float), (cudaStream_t)stream)(bn_stats_kernel, v, (float*)workspace, TreeRed{}, BnFinal{});
  UBR_LAUNCH_CHECK("bn_stats_kernel");
  return 0;
}

// bn_stats + finalize in ONE launch: the last CTA to arrive finishes the reduction (tree_reduce.cuh)
int ubr_bn_stats_final(ubr_view y, void* workspace, void* tree, const float* gamma, const float* beta, float* running_mean,
                       float* running_var, float momentum, float eps, float* scale, float* shift, float* mean,
                       float* invstd, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk_red(y, "bn_stats_final.y")) return rc;
  UBR_REQUIRE(workspace && tree && scale && shift, "bn_stats_final: null pointer");
  if (red_smem(2, y.c, true) > 48 * 1024) return 1;      // too wide for the in-launch tail: ubr_bn_stats + ubr_bn_finalize
  DView v = dv(y);
  const int nb = red_blocks(v.npix, v.c);
  BnFinal f{gamma, beta, running_mean, running_var, momentum, eps, (double)v.npix, scale, shift, mean, invstd, v.c, v.c};
  ::ubr::Launch(nb, kRedThreads, red_smem(2, v.c, true), (cudaStream_t)stream)(
      bn_stats_kernel, v, (float*)workspace, make_tree(tree, (const float*)workspace, nb, 2 * v.c), f);
  UBR_LAUNCH_CHECK("bn_stats_kernel");
  return 0;
}

uint64_t ubr_bn_tree_bytes(int c) { return (uint64_t)tree_bytes(c); }

int ubr_bn_finalize(const void* workspace, int c, double count, const float* gamma, const float* beta,
                    float* running_mean, float* running_var, float momentum, float eps, float* scale, float* shift,
                    float* mean, float* invstd, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(workspace && scale && shift, "bn_finalize: null pointer");
  UBR_REQUIRE(c > 0 && c % 8 == 0 && count >= 1.0, "bn_finalize: bad c/count");
  const int nb = red_blocks((long)count, c);
  ::ubr::Launch(ceil_div(c, 32), dim3(32, kColLanes), 0, (cudaStream_t)stream)(bn_finalize_kernel, (const float*)workspace, nb, c, count, gamma,
                                                                         beta, running_mean, running_var, momentum,
                                                                         eps, scale, shift, mean, invstd);
  UBR_LAUNCH_CHECK("bn_finalize_kernel");
  return 0;
}

int ubr_bn_finalize_rows(const void* workspace, int rows, int c, double count, const float* gamma, const float* beta,
                         float* running_mean, float* running_var, float momentum, float eps, float* scale, float* shift,
                         float* mean, float* invstd, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(workspace && scale && shift, "bn_finalize_rows: null pointer");
  UBR_REQUIRE(c > 0 && c % 8 == 0 && count >= 1.0 && rows >= 1 && rows <= 2 * kRedBlocks, "bn_finalize_rows: bad c/count/rows");
  ::ubr::Launch(ceil_div(c, 32), dim3(32, kColLanes), 0, (cudaStream_t)stream)(bn_finalize_kernel, (const float*)workspace, rows, c, count,
                                                                                      gamma, beta, running_mean, running_var,
                                                                                      momentum, eps, scale, shift, mean, invstd);
  UBR_LAUNCH_CHECK("bn_finalize_kernel");
  return 0;
}

int ubr_bn_eval_affine(int c, const float* gamma, const float* beta, const float* running_mean,
                       const float* running_var, float eps, float* scale, float* shift, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(running_mean && running_var && scale && shift && c > 0, "bn_eval_affine: bad args");
  ::ubr::Launch(ceil_div(c, 128), 128, 0, (cudaStream_t)stream)(bn_eval_affine_kernel, c, gamma, beta, running_mean, running_var,
                                                                            eps, scale, shift);
  UBR_LAUNCH_CHECK("bn_eval_affine_kernel");
  return 0;
}

int ubr_bn_act(ubr_view y, const float* scale, const float* shift, int relu, ubr_view out, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk(y, "bn_act.y")) return rc;
  if (int rc = chk(out, "bn_act.out")) return rc;
  if (int rc = same_shape(y, out, "bn_act")) return rc;
  UBR_REQUIRE(scale && shift, "bn_act: null affine");
  DView v = dv(y);
  ::ubr::Launch(ew_grid(v.npix, v.c >> 3, 4), 256, 0, (cudaStream_t)stream)(bn_act_kernel, v, scale, shift, relu, dvw(out));
  UBR_LAUNCH_CHECK("bn_act_kernel");
  return 0;
}

int ubr_block_tail(ubr_view y2, const float* s2, const float* b2, ubr_view res, const float* sr, const float* br,
                   ubr_view out, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk(y2, "block_tail.y2")) return rc;
  if (int rc = chk(res, "block_tail.res")) return rc;
  if (int rc = chk(out, "block_tail.out")) return rc;
  if (int rc = same_shape(y2, res, "block_tail(res)")) return rc;
  if (int rc = same_shape(y2, out, "block_tail(out)")) return rc;
  UBR_REQUIRE(s2 && b2 && ((sr == nullptr) == (br == nullptr)), "block_tail: bad affine pointers");
  DView v = dv(y2);
  ::ubr::Launch(ew_grid(v.npix, v.c >> 3, 4), 256, 0, (cudaStream_t)stream)(block_tail_kernel, v, s2, b2, dv(res), sr, br,
                                                                                      dvw(out));
  UBR_LAUNCH_CHECK("block_tail_kernel");
  return 0;
}

int ubr_bn_act_bwd_reduce(ubr_view dout, ubr_view y, const float* scale, const float* shift, const float* mean,
                          const float* invstd, int relu, void* workspace, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk_red(y, "bn_act_bwd_reduce.y")) return rc;
  if (int rc = chk(dout, "bn_act_bwd_reduce.dout")) return rc;
  if (int rc = same_shape(y, dout, "bn_act_bwd_reduce")) return rc;
  UBR_REQUIRE(scale && shift && mean && invstd && workspace, "bn_act_bwd_reduce: null pointer");
  DView v = dv(y);
  const int nb = red_blocks(v.npix, v.c, kBwdRedBlocks);
  ::ubr::Launch(nb, kRedThreads, kRedThreads * 16 * sizeof(float), (cudaStream_t)stream)(bn_act_bwd_reduce_kernel, dv(dout), v, scale, shift, mean, invstd, relu, (float*)workspace, TreeRed{}, SumsFinal{});
  UBR_LAUNCH_CHECK("bn_act_bwd_reduce_kernel");
  return 0;
}

// reduce pass that also finishes its sums (sums[2][c] behind the partial rows, dbeta / dgamma) in the same launch
int ubr_bn_act_bwd_reduce_final(ubr_view dout, ubr_view y, const float* scale, const float* shift, const float* mean,
                                const float* invstd, int relu, void* workspace, void* tree, float* dgamma, float* dbeta,
                                int accumulate, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk_red(y, "bn_act_bwd_reduce_final.y")) return rc;
  if (int rc = chk(dout, "bn_act_bwd_reduce_final.dout")) return rc;
  if (int rc = same_shape(y, dout, "bn_act_bwd_reduce_final")) return rc;
  UBR_REQUIRE(scale && shift && mean && invstd && workspace && tree, "bn_act_bwd_reduce_final: null pointer");
  if (red_smem(2, y.c, true) > 48 * 1024) return 1;      // too wide: ubr_bn_act_bwd_reduce + ubr_bn_act_bwd_apply
  DView v = dv(y);
  const int nb = red_blocks(v.npix, v.c, kBwdRedBlocks);
  float* sums = (float*)workspace + (size_t)kRedBlocks * 4 * v.c;
  SumsFinal f{sums, {dbeta, dgamma, nullptr, nullptr}, v.c, 2, v.c, accumulate};
  ::ubr::Launch(nb, kRedThreads, red_smem(2, v.c, true), (cudaStream_t)stream)(
      bn_act_bwd_reduce_kernel, dv(dout), v, scale, shift, mean, invstd, relu, (float*)workspace,
      make_tree(tree, (const float*)workspace, nb, 2 * v.c), f);
  UBR_LAUNCH_CHECK("bn_act_bwd_reduce_kernel");
  return 0;
}

// apply pass for sums that are already final (ubr_bn_act_bwd_reduce_final / ubr_conv2d_dgrad_bnreduce_final)
int ubr_bn_act_bwd_apply_final(ubr_view dout, ubr_view y, const float* scale, const float* shift, const float* mean,
                               const float* invstd, int relu, const void* workspace, double count, ubr_view dy,
                               void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk_red(y, "bn_act_bwd_apply_final.y")) return rc;
  if (int rc = chk(dout, "bn_act_bwd_apply_final.dout")) return rc;
  if (int rc = chk(dy, "bn_act_bwd_apply_final.dy")) return rc;
  if (int rc = same_shape(y, dout, "bn_act_bwd_apply_final(dout)")) return rc;
  if (int rc = same_shape(y, dy, "bn_act_bwd_apply_final(dy)")) return rc;
  UBR_REQUIRE(scale && shift && mean && invstd && workspace && count >= 1.0, "bn_act_bwd_apply_final: null pointer");
  DView v = dv(y);
  const float* sums = (const float*)workspace + (size_t)kRedBlocks * 4 * v.c;
  ::ubr::Launch(ew_grid(v.npix, v.c >> 3, 4), 256, 0, (cudaStream_t)stream)(bn_act_bwd_apply_kernel, dv(dout), v, scale, shift, mean,
                                                                             invstd, (const float*)nullptr, relu, sums,
                                                                             (float)(1.0 / count), dvw(dy));
  UBR_LAUNCH_CHECK("bn_act_bwd_apply_kernel");
  return 0;
}

int ubr_bn_act_bwd_apply(ubr_view dout, ubr_view y, const float* scale, const float* shift, const float* mean,
                         const float* invstd, const float* gamma, int relu, const void* workspace, double count,
                         ubr_view dy, float* dgamma, float* dbeta, int accumulate, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk_red(y, "bn_act_bwd_apply.y")) return rc;
  if (int rc = chk(dout, "bn_act_bwd_apply.dout")) return rc;
  if (int rc = chk(dy, "bn_act_bwd_apply.dy")) return rc;
  if (int rc = same_shape(y, dout, "bn_act_bwd_apply(dout)")) return rc;
  if (int rc = same_shape(y, dy, "bn_act_bwd_apply(dy)")) return rc;
  UBR_REQUIRE(scale && shift && mean && invstd && workspace, "bn_act_bwd_apply: null pointer");
  DView v = dv(y);
  const int nb = red_blocks(v.npix, v.c, kBwdRedBlocks);
  const float* partials = (const float*)workspace;
  float* sums = (float*)workspace + (size_t)kRedBlocks * 4 * v.c;
  cudaStream_t st = (cudaStream_t)stream;
  // sums row 0 = sum dz (-> dbeta), row 1 = sum dz*xhat (-> dgamma): parameter gradients written by the finalize
  ::ubr::Launch(ceil_div(2 * v.c, 32), dim3(32, kColLanes), 0, st)(sums_finalize_kernel, partials, nb, 2 * v.c, sums, v.c, ParamGradOut{{dbeta, dgamma, nullptr, nullptr}}, accumulate);
  UBR_LAUNCH_CHECK("sums_finalize_kernel");
  ::ubr::Launch(ew_grid(v.npix, v.c >> 3, 4), 256, 0, st)(bn_act_bwd_apply_kernel, dv(dout), v, scale, shift, mean, invstd,
                                                                          gamma, relu, sums, (float)(1.0 / count),
                                                                          dvw(dy));
  UBR_LAUNCH_CHECK("bn_act_bwd_apply_kernel");
  return 0;
}

int ubr_bn_act_bwd_apply_rows(ubr_view dout, ubr_view y, const float* scale, const float* shift, const float* mean,
                              const float* invstd, const float* gamma, int relu, const void* workspace, int rows,
                              double count, ubr_view dy, float* dgamma, float* dbeta, int accumulate, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk_red(y, "bn_act_bwd_apply_rows.y")) return rc;
  if (int rc = chk(dout, "bn_act_bwd_apply_rows.dout")) return rc;
  if (int rc = chk(dy, "bn_act_bwd_apply_rows.dy")) return rc;
  if (int rc = same_shape(y, dout, "bn_act_bwd_apply_rows(dout)")) return rc;
  if (int rc = same_shape(y, dy, "bn_act_bwd_apply_rows(dy)")) return rc;
  UBR_REQUIRE(scale && shift && mean && invstd && workspace && rows >= 1 && rows <= kRedBlocks,
              "bn_act_bwd_apply_rows: null pointer / bad row count %d", rows);
  DView v = dv(y);
  const float* partials = (const float*)workspace;
  float* sums = (float*)workspace + (size_t)kRedBlocks * 4 * v.c;
  cudaStream_t st = (cudaStream_t)stream;
  ::ubr::Launch(ceil_div(2 * v.c, 32), dim3(32, kColLanes), 0, st)(sums_finalize_kernel, partials, rows, 2 * v.c, sums, v.c, ParamGradOut{{dbeta, dgamma, nullptr, nullptr}}, accumulate);
  UBR_LAUNCH_CHECK("sums_finalize_kernel");
  ::ubr::Launch(ew_grid(v.npix, v.c >> 3, 4), 256, 0, st)(bn_act_bwd_apply_kernel, dv(dout), v, scale, shift, mean, invstd,
                                                                          gamma, relu, sums, (float)(1.0 / count),
                                                                          dvw(dy));
  UBR_LAUNCH_CHECK("bn_act_bwd_apply_kernel");
  return 0;
}

int ubr_block_tail_bwd_reduce(ubr_view dout, ubr_view out, ubr_view y2, const float* s2, const float* b2,
                              const float* mean2, const float* invstd2, ubr_view res, const float* meanr,
                              const float* invstdr, const float* scaler, const float* shiftr, int res_is_bn,
                              void* workspace, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk_red(y2, "block_tail_bwd_reduce.y2")) return rc;
  if (int rc = chk(dout, "block_tail_bwd_reduce.dout")) return rc;
  if (!res_is_bn) {                                 // with a BatchNorm'd residual `out` is not read (mask recomputed)
    if (int rc = chk(out, "block_tail_bwd_reduce.out")) return rc;
    if (int rc = same_shape(y2, out, "block_tail_bwd_reduce(out)")) return rc;
  }
  if (int rc = same_shape(y2, dout, "block_tail_bwd_reduce(dout)")) return rc;
  UBR_REQUIRE(s2 && b2 && mean2 && invstd2 && workspace, "block_tail_bwd_reduce: null pointer");
  DView v = dv(y2);
  const int nb = red_blocks(v.npix, v.c, kBwdRedBlocks);
  cudaStream_t st = (cudaStream_t)stream;
  if (res_is_bn) {
    if (int rc = chk(res, "block_tail_bwd_reduce.res")) return rc;
    if (int rc = same_shape(y2, res, "block_tail_bwd_reduce(res)")) return rc;
    UBR_REQUIRE(meanr && invstdr && scaler && shiftr, "block_tail_bwd_reduce: null residual stats / affine");
    ::ubr::Launch(nb, kRedThreads, kRedThreads * 32 * sizeof(float), st)(tail_bwd_reduce_kernel<true>, dv(dout), v, v, s2, b2, mean2, invstd2, dv(res), meanr, invstdr, scaler, shiftr, (float*)workspace, TreeRed{}, SumsFinal{});
  } else {
    ::ubr::Launch(nb, kRedThreads, kRedThreads * 16 * sizeof(float), st)(tail_bwd_reduce_kernel<false>, dv(dout), dv(out), v, s2, b2, mean2, invstd2, v, nullptr, nullptr, nullptr, nullptr, (float*)workspace, TreeRed{}, SumsFinal{});
  }
  UBR_LAUNCH_CHECK("tail_bwd_reduce_kernel");
  return 0;
}

// reduce pass of the block-tail backward that finishes its 2 / 4 sums and the BatchNorm parameter gradients itself
int ubr_block_tail_bwd_reduce_final(ubr_view dout, ubr_view out, ubr_view y2, const float* s2, const float* b2,
                                    const float* mean2, const float* invstd2, ubr_view res, const float* meanr,
                                    const float* invstdr, const float* scaler, const float* shiftr, int res_is_bn,
                                    void* workspace, void* tree, float* dgamma2, float* dbeta2, float* dgammar,
                                    float* dbetar, int accumulate, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk_red(y2, "block_tail_bwd_reduce_final.y2")) return rc;
  if (int rc = chk(dout, "block_tail_bwd_reduce_final.dout")) return rc;
  if (!res_is_bn) {
    if (int rc = chk(out, "block_tail_bwd_reduce_final.out")) return rc;
    if (int rc = same_shape(y2, out, "block_tail_bwd_reduce_final(out)")) return rc;
  }
  if (int rc = same_shape(y2, dout, "block_tail_bwd_reduce_final(dout)")) return rc;
  UBR_REQUIRE(s2 && b2 && mean2 && invstd2 && workspace && tree, "block_tail_bwd_reduce_final: null pointer");
  if (red_smem(4, y2.c, true) > 48 * 1024) return 1;     // too wide: ubr_block_tail_bwd_reduce + ubr_block_tail_bwd_apply
  DView v = dv(y2);
  const int nb = red_blocks(v.npix, v.c, kBwdRedBlocks);
  const int ns = res_is_bn ? 4 : 2;
  float* sums = (float*)workspace + (size_t)kRedBlocks * 4 * v.c;
  SumsFinal f{sums, {dbeta2, dgamma2, res_is_bn ? dbetar : nullptr, res_is_bn ? dgammar : nullptr}, v.c, ns, v.c, accumulate};
  const TreeRed tr = make_tree(tree, (const float*)workspace, nb, ns * v.c);
  cudaStream_t st = (cudaStream_t)stream;
  if (res_is_bn) {
    if (int rc = chk(res, "block_tail_bwd_reduce_final.res")) return rc;
    if (int rc = same_shape(y2, res, "block_tail_bwd_reduce_final(res)")) return rc;
    UBR_REQUIRE(meanr && invstdr && scaler && shiftr, "block_tail_bwd_reduce_final: null residual stats / affine");
    ::ubr::Launch(nb, kRedThreads, red_smem(4, v.c, true), st)(tail_bwd_reduce_kernel<true>, dv(dout), v, v, s2, b2, mean2, invstd2, dv(res), meanr, invstdr, scaler, shiftr, (float*)workspace, tr, f);
  } else {
    ::ubr::Launch(nb, kRedThreads, red_smem(2, v.c, true), st)(tail_bwd_reduce_kernel<false>, dv(dout), dv(out), v, s2, b2, mean2, invstd2, v, nullptr, nullptr, nullptr, nullptr, (float*)workspace, tr, f);
  }
  UBR_LAUNCH_CHECK("tail_bwd_reduce_kernel");
  return 0;
}

int ubr_block_tail_bwd_apply(ubr_view dout, ubr_view out, ubr_view y2, const float* s2, const float* b2,
                             const float* mean2, const float* invstd2, const float* gamma2, ubr_view res,
                             const float* meanr, const float* invstdr, const float* gammar, const float* scaler,
                             const float* shiftr, int res_is_bn,
                             const void* workspace, double count, ubr_view dy2, ubr_view dres, float* dgamma2,
                             float* dbeta2, float* dgammar, float* dbetar, int accumulate, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk_red(y2, "block_tail_bwd_apply.y2")) return rc;
  if (int rc = chk(dout, "block_tail_bwd_apply.dout")) return rc;
  if (!res_is_bn) {
    if (int rc = chk(out, "block_tail_bwd_apply.out")) return rc;
    if (int rc = same_shape(y2, out, "block_tail_bwd_apply(out)")) return rc;
  }
  if (int rc = chk(dy2, "block_tail_bwd_apply.dy2")) return rc;
  if (int rc = chk(dres, "block_tail_bwd_apply.dres")) return rc;
  if (int rc = same_shape(y2, dout, "block_tail_bwd_apply(dout)")) return rc;
  if (int rc = same_shape(y2, dy2, "block_tail_bwd_apply(dy2)")) return rc;
  if (int rc = same_shape(y2, dres, "block_tail_bwd_apply(dres)")) return rc;
  UBR_REQUIRE(s2 && b2 && mean2 && invstd2 && workspace, "block_tail_bwd_apply: null pointer");
  (void)gamma2;
  DView v = dv(y2);
  const int nb = red_blocks(v.npix, v.c, kBwdRedBlocks);
  const int ns = res_is_bn ? 4 : 2;
  const float* partials = (const float*)workspace;
  float* sums = (float*)workspace + (size_t)kRedBlocks * 4 * v.c;
  cudaStream_t st = (cudaStream_t)stream;
  ::ubr::Launch(ceil_div(ns * v.c, 32), dim3(32, kColLanes), 0, st)(sums_finalize_kernel, partials, nb, ns * v.c, sums, v.c,
      ParamGradOut{{dbeta2, dgamma2, res_is_bn ? dbetar : nullptr, res_is_bn ? dgammar : nullptr}}, accumulate);
  UBR_LAUNCH_CHECK("sums_finalize_kernel");
  const unsigned blocks = ew_grid(v.npix, v.c >> 3, 2);
  if (res_is_bn) {
    if (int rc = chk(res, "block_tail_bwd_apply.res")) return rc;
    UBR_REQUIRE(meanr && invstdr && gammar && scaler && shiftr, "block_tail_bwd_apply: null residual stats / affine");
    ::ubr::Launch(blocks, 256, 0, st)(tail_bwd_apply_kernel<true>, dv(dout), v, v, s2, b2, mean2, invstd2, dv(res), meanr,
                                                        invstdr, gammar, scaler, shiftr, sums, (float)(1.0 / count),
                                                        dvw(dy2), dvw(dres));
  } else {
    ::ubr::Launch(blocks, 256, 0, st)(tail_bwd_apply_kernel<false>, dv(dout), dv(out), v, s2, b2, mean2, invstd2, v, nullptr,
                                                         nullptr, nullptr, nullptr, nullptr, sums, (float)(1.0 / count),
                                                         dvw(dy2), dvw(dres));
  }
  UBR_LAUNCH_CHECK("tail_bwd_apply_kernel");
  return 0;
}

// apply pass of the block-tail backward for sums that are already final (ubr_block_tail_bwd_reduce_final)
int ubr_block_tail_bwd_apply_final(ubr_view dout, ubr_view out, ubr_view y2, const float* s2, const float* b2,
                                   const float* mean2, const float* invstd2, ubr_view res, const float* meanr,
                                   const float* invstdr, const float* gammar, const float* scaler, const float* shiftr,
                                   int res_is_bn, const void* workspace, double count, ubr_view dy2, ubr_view dres,
                                   void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk_red(y2, "block_tail_bwd_apply_final.y2")) return rc;
  if (int rc = chk(dout, "block_tail_bwd_apply_final.dout")) return rc;
  if (!res_is_bn) {
    if (int rc = chk(out, "block_tail_bwd_apply_final.out")) return rc;
    if (int rc = same_shape(y2, out, "block_tail_bwd_apply_final(out)")) return rc;
  }
  if (int rc = chk(dy2, "block_tail_bwd_apply_final.dy2")) return rc;
  if (int rc = chk(dres, "block_tail_bwd_apply_final.dres")) return rc;
  if (int rc = same_shape(y2, dout, "block_tail_bwd_apply_final(dout)")) return rc;
  if (int rc = same_shape(y2, dy2, "block_tail_bwd_apply_final(dy2)")) return rc;
  if (int rc = same_shape(y2, dres, "block_tail_bwd_apply_final(dres)")) return rc;
  UBR_REQUIRE(s2 && b2 && mean2 && invstd2 && workspace && count >= 1.0, "block_tail_bwd_apply_final: null pointer");
  DView v = dv(y2);
  const float* sums = (const float*)workspace + (size_t)kRedBlocks * 4 * v.c;
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned blocks = ew_grid(v.npix, v.c >> 3, 2);
  if (res_is_bn) {
    if (int rc = chk(res, "block_tail_bwd_apply_final.res")) return rc;
    UBR_REQUIRE(meanr && invstdr && gammar && scaler && shiftr, "block_tail_bwd_apply_final: null residual stats / affine");
    ::ubr::Launch(blocks, 256, 0, st)(tail_bwd_apply_kernel<true>, dv(dout), v, v, s2, b2, mean2, invstd2, dv(res), meanr,
                                                        invstdr, gammar, scaler, shiftr, sums, (float)(1.0 / count),
                                                        dvw(dy2), dvw(dres));
  } else {
    ::ubr::Launch(blocks, 256, 0, st)(tail_bwd_apply_kernel<false>, dv(dout), dv(out), v, s2, b2, mean2, invstd2, v, nullptr,
                                                         nullptr, nullptr, nullptr, nullptr, sums, (float)(1.0 / count),
                                                         dvw(dy2), dvw(dres));
  }
  UBR_LAUNCH_CHECK("tail_bwd_apply_kernel");
  return 0;
}

int ubr_maxpool3s2_fwd(ubr_view x, ubr_view y, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk(x, "maxpool_fwd.x")) return rc;
  if (int rc = chk(y, "maxpool_fwd.y")) return rc;
  UBR_REQUIRE(x.h % 2 == 0 && x.w % 2 == 0, "maxpool_fwd: H,W must be even");
  UBR_REQUIRE(y.n == x.n && y.c == x.c && y.h == x.h / 2 && y.w == x.w / 2, "maxpool_fwd: output view mismatch");
  const long total = (long)y.n * y.h * y.w * (y.c >> 3);
  ::ubr::Launch(ew_blocks(total), 256, 0, (cudaStream_t)stream)(maxpool_fwd_kernel, (const __nv_bfloat16*)x.ptr, x.ld, x.n, x.h,
                                                                         x.w, x.c, (__nv_bfloat16*)y.ptr, y.ld);
  UBR_LAUNCH_CHECK("maxpool_fwd_kernel");
  return 0;
}

int ubr_maxpool3s2_bwd(ubr_view x, ubr_view dy, ubr_view dx_accum, int accumulate, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk(x, "maxpool_bwd.x")) return rc;
  if (int rc = chk(dy, "maxpool_bwd.dy")) return rc;
  if (int rc = chk(dx_accum, "maxpool_bwd.dx")) return rc;
  if (int rc = same_shape(x, dx_accum, "maxpool_bwd(dx)")) return rc;
  UBR_REQUIRE(dy.n == x.n && dy.c == x.c && dy.h == x.h / 2 && dy.w == x.w / 2, "maxpool_bwd: dy view mismatch");
  const long total = (long)x.n * x.h * x.w * (x.c >> 3);
  ::ubr::Launch(ew_blocks(total), 256, 0, (cudaStream_t)stream)(maxpool_bwd_kernel, (const __nv_bfloat16*)x.ptr, x.ld, x.n, x.h, x.w, x.c, (const __nv_bfloat16*)dy.ptr, dy.ld,
      (__nv_bfloat16*)dx_accum.ptr, dx_accum.ld, accumulate);
  UBR_LAUNCH_CHECK("maxpool_bwd_kernel");
  return 0;
}

int ubr_maxpool3s1_fwd(ubr_view x, ubr_view y, uint8_t* idx, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk(x, "maxpool3s1_fwd.x")) return rc;
  if (int rc = chk(y, "maxpool3s1_fwd.y")) return rc;
  UBR_REQUIRE(idx == nullptr || ((uintptr_t)idx & 7) == 0, "maxpool3s1_fwd: idx must be 8-byte aligned");
  if (int rc = same_shape(x, y, "maxpool3s1_fwd(y)")) return rc;
  const long total = (long)x.n * x.h * x.w * (x.c / 8);
  ::ubr::Launch(ew_blocks(total), 256, 0, (cudaStream_t)stream)(maxpool3s1_fwd_kernel, (const __nv_bfloat16*)x.ptr, x.ld, x.n, x.h, x.w,
                                                                            x.c, (__nv_bfloat16*)y.ptr, y.ld, idx);
  UBR_LAUNCH_CHECK("maxpool3s1_fwd_kernel");
  return 0;
}

int ubr_maxpool3s1_bwd_idx(const uint8_t* idx, ubr_view dy, ubr_view dx_accum, int accumulate, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk(dy, "maxpool3s1_bwd_idx.dy")) return rc;
  if (int rc = chk(dx_accum, "maxpool3s1_bwd_idx.dx")) return rc;
  UBR_REQUIRE(idx != nullptr, "maxpool3s1_bwd_idx: null idx");
  if (int rc = same_shape(dy, dx_accum, "maxpool3s1_bwd_idx(dx)")) return rc;
  const long total = (long)dy.n * dy.h * dy.w * (dy.c / 8);
  ::ubr::Launch(ew_blocks(total), 256, 0, (cudaStream_t)stream)(maxpool3s1_bwd_idx_kernel, idx, dy.n, dy.h, dy.w, dy.c, (const __nv_bfloat16*)dy.ptr, dy.ld, (__nv_bfloat16*)dx_accum.ptr, dx_accum.ld, accumulate);
  UBR_LAUNCH_CHECK("maxpool3s1_bwd_idx_kernel");
  return 0;
}

int ubr_pixel_unshuffle2(ubr_view y, ubr_view y4, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk(y, "pixel_unshuffle2.y")) return rc;
  if (int rc = chk(y4, "pixel_unshuffle2.y4")) return rc;
  UBR_REQUIRE(y.h % 2 == 0 && y.w % 2 == 0 && y4.n == y.n && y4.h * 2 == y.h && y4.w * 2 == y.w && y4.c == 4 * y.c,
              "pixel_unshuffle2: shape mismatch");
  const long total = (long)y.n * y.h * y.w * (y.c / 8);
  ::ubr::Launch(ew_blocks(total), 256, 0, (cudaStream_t)stream)(pixel_unshuffle2_kernel, (const __nv_bfloat16*)y.ptr, y.ld, y4.n, y4.h,
                                                                              y4.w, y.c, (__nv_bfloat16*)y4.ptr, y4.ld);
  UBR_LAUNCH_CHECK("pixel_unshuffle2_kernel");
  return 0;
}

int ubr_deconv4s2_wgrad_gather(const float* dw3, int in_ch, int out_ch, float* dw, int accumulate, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(dw3 != nullptr && dw != nullptr && in_ch > 0 && out_ch > 0, "deconv4s2_wgrad_gather: bad arguments");
  const int total = in_ch * out_ch * 16;
  ::ubr::Launch((total + 255) / 256, 256, 0, (cudaStream_t)stream)(deconv_wgrad_gather_kernel, dw3, in_ch, out_ch, dw, accumulate);
  UBR_LAUNCH_CHECK("deconv_wgrad_gather_kernel");
  return 0;
}

int ubr_maxpool3s2_fwd_idx(ubr_view x, ubr_view y, uint8_t* idx, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk(x, "maxpool_fwd_idx.x")) return rc;
  if (int rc = chk(y, "maxpool_fwd_idx.y")) return rc;
  UBR_REQUIRE(idx != nullptr && ((uintptr_t)idx & 7) == 0, "maxpool_fwd_idx: idx must be a non-null 8-byte aligned pointer");
  UBR_REQUIRE(x.h % 2 == 0 && x.w % 2 == 0 && y.n == x.n && y.h == x.h / 2 && y.w == x.w / 2 && y.c == x.c,
              "maxpool_fwd_idx: shape mismatch");
  const long total = (long)y.n * y.h * y.w * (x.c >> 3);
  if (total < (1L << 31) - (1L << 26))
    ::ubr::Launch(ew_blocks(total), 256, 0, (cudaStream_t)stream)(maxpool_fwd_idx_kernel<unsigned>, (const __nv_bfloat16*)x.ptr, x.ld, x.n, x.h,
                                                                               x.w, x.c, (__nv_bfloat16*)y.ptr, y.ld, idx);
  else
    ::ubr::Launch(ew_blocks(total), 256, 0, (cudaStream_t)stream)(maxpool_fwd_idx_kernel<long>, (const __nv_bfloat16*)x.ptr, x.ld, x.n, x.h,
                                                                               x.w, x.c, (__nv_bfloat16*)y.ptr, y.ld, idx);
  UBR_LAUNCH_CHECK("maxpool_fwd_idx_kernel");
  return 0;
}

int ubr_maxpool3s2_bwd_idx(const uint8_t* idx, ubr_view dy, ubr_view dx_accum, int accumulate, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk(dy, "maxpool_bwd_idx.dy")) return rc;
  if (int rc = chk(dx_accum, "maxpool_bwd_idx.dx")) return rc;
  UBR_REQUIRE(idx != nullptr, "maxpool_bwd_idx: null idx");
  UBR_REQUIRE(dx_accum.n == dy.n && dx_accum.h == 2 * dy.h && dx_accum.w == 2 * dy.w && dx_accum.c == dy.c,
              "maxpool_bwd_idx: shape mismatch");
  const long total = (long)dx_accum.n * dx_accum.h * dx_accum.w * (dy.c >> 3);
  // grid-stride loop: i + gridDim * blockDim must not wrap either
  if (total < (1L << 31) - (1L << 26))
    ::ubr::Launch(ew_blocks(total), 256, 0, (cudaStream_t)stream)(maxpool_bwd_idx_kernel<unsigned>, idx, dx_accum.n, dx_accum.h, dx_accum.w, dy.c,
                                                                               (const __nv_bfloat16*)dy.ptr, dy.ld,
                                                                               (__nv_bfloat16*)dx_accum.ptr, dx_accum.ld,
                                                                               accumulate);
  else
    ::ubr::Launch(ew_blocks(total), 256, 0, (cudaStream_t)stream)(maxpool_bwd_idx_kernel<long>, idx, dx_accum.n, dx_accum.h, dx_accum.w, dy.c,
                                                                               (const __nv_bfloat16*)dy.ptr, dy.ld,
                                                                               (__nv_bfloat16*)dx_accum.ptr, dx_accum.ld,
                                                                               accumulate);
  UBR_LAUNCH_CHECK("maxpool_bwd_idx_kernel");
  return 0;
}

int ubr_add(ubr_view a, ubr_view b, ubr_view c, ubr_view out, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk(a, "add.a")) return rc;
  if (int rc = chk(out, "add.out")) return rc;
  if (int rc = same_shape(a, out, "add(out)")) return rc;
  DView vb{nullptr, 0, 0, 0}, vc{nullptr, 0, 0, 0};
  if (b.ptr) {
    if (int rc = chk(b, "add.b")) return rc;
    if (int rc = same_shape(a, b, "add(b)")) return rc;
    vb = dv(b);
  }
  if (c.ptr) {
    if (int rc = chk(c, "add.c")) return rc;
    if (int rc = same_shape(a, c, "add(c)")) return rc;
    vc = dv(c);
  }
  DView va = dv(a);
  ::ubr::Launch(ew_grid(va.npix, va.c >> 3, 4), 256, 0, (cudaStream_t)stream)(add_kernel, va, vb, vc, dvw(out));
  UBR_LAUNCH_CHECK("add_kernel");
  return 0;
}

int ubr_channel_sum(ubr_view a, int nvalid, void* workspace, float* out, int accumulate, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk_red(a, "channel_sum.a")) return rc;
  UBR_REQUIRE(workspace && out && nvalid >= 1 && nvalid <= a.c, "channel_sum: bad arguments");
  DView v = dv(a);
  const int nb = red_blocks(v.npix, v.c);
  cudaStream_t st = (cudaStream_t)stream;
  ::ubr::Launch(nb, kRedThreads, kRedThreads * 16 * sizeof(float), st)(bn_stats_kernel, v, (float*)workspace, TreeRed{}, BnFinal{});
  UBR_LAUNCH_CHECK("bn_stats_kernel");
  ::ubr::Launch(ceil_div(v.c, 32), dim3(32, kColLanes), 0, st)(channel_sum_finish_kernel, (const float*)workspace, nb, v.c,
                                                               nvalid, out, accumulate);
  UBR_LAUNCH_CHECK("channel_sum_finish_kernel");
  return 0;
}

// ---- allocation-free variant (no counterpart in the reference) -----------
// Same two kernels as ubr_channel_sum, but the partial rows live in a file-scope device array instead of a caller
// workspace, for callers that cannot allocate (graph capture, serving loops).  Every GPU that loads the library holds
// the array (kStaticMaxC * kRedBlocks * 2 floats = 2.3 MiB) until exit.  The array's name is only a host-side handle:
// its device address is looked up once per device with cudaGetSymbolAddress (with that device current) and cached.
// bn_stats_kernel overwrites every partial row the finish kernel reads, so nothing has to be cleared between calls;
// all calls share the array, so two calls that overlap on the GPU (different streams) race.
constexpr int kStaticMaxC = 512;
static __device__ __align__(256) float g_static_partials[(size_t)kRedBlocks * 2 * kStaticMaxC];
static float* g_static_partials_ptr[64] = {};

int ubr_channel_sum_static(ubr_view a, int nvalid, float* out, int accumulate, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = chk_red(a, "channel_sum_static.a")) return rc;
  UBR_REQUIRE(out && nvalid >= 1 && nvalid <= a.c, "channel_sum_static: bad arguments");
  UBR_REQUIRE(a.c <= kStaticMaxC, "channel_sum_static: c=%d exceeds the built-in scratch (max %d channels); use "
              "ubr_channel_sum with a caller workspace", a.c, kStaticMaxC);
  int dev = 0;
  UBR_CUDA_OK(cudaGetDevice(&dev));
  if (g_static_partials_ptr[dev] == nullptr) {
    float* p = nullptr;
    UBR_CUDA_OK(cudaGetSymbolAddress((void**)&p, g_static_partials));
    g_static_partials_ptr[dev] = p;
  }
  float* scratch = g_static_partials_ptr[dev];
  DView v = dv(a);
  const int nb = red_blocks(v.npix, v.c);
  cudaStream_t st = (cudaStream_t)stream;
  ::ubr::Launch(nb, kRedThreads, kRedThreads * 16 * sizeof(float), st)(bn_stats_kernel, v, scratch, TreeRed{}, BnFinal{});
  UBR_LAUNCH_CHECK("bn_stats_kernel");
  ::ubr::Launch(ceil_div(v.c, 32), dim3(32, kColLanes), 0, st)(channel_sum_finish_kernel, (const float*)scratch, nb, v.c,
                                                               nvalid, out, accumulate);
  UBR_LAUNCH_CHECK("channel_sum_finish_kernel");
  return 0;
}

int ubr_increment_i64(int64_t* const* ptrs_device, int n, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(ptrs_device != nullptr && n >= 1, "increment_i64: bad arguments");
  ::ubr::Launch(ceil_div(n, 128), 128, 0, (cudaStream_t)stream)(increment_i64_kernel, (long long* const*)ptrs_device, n);
  UBR_LAUNCH_CHECK("increment_i64_kernel");
  return 0;
}

int ubr_scale_f32(float* buf, uint64_t count, float alpha, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(buf != nullptr || count == 0, "scale_f32: null buffer");
  if (count == 0) return 0;
  ::ubr::Launch(ew_blocks((long)count), 256, 0, (cudaStream_t)stream)(scale_f32_kernel, buf, (unsigned long long)count, alpha);
  UBR_LAUNCH_CHECK("scale_f32_kernel");
  return 0;
}

}  // extern "C"
