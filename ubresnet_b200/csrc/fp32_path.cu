// fp32 precision mode (UBR_PRECISION=fp32): the reference's arithmetic type end to end.
//
// The reference runs every layer in fp32 (models/ub_uresnet.py:41-70, models/common_layers.py:13-58); BASELINE.json
// asks for a mode that matches it to 1e-4.  bf16 operands cannot (8 significant bits), tf32 cannot either (11 bits
// through 58 convolutions), so this path keeps fp32 NHWC activations and fp32 FMA accumulation on the CUDA cores.
// It is a PARITY mode: correctness first, one generic implicit-GEMM kernel for every convolution flavour of the
// network (conv / strided conv / dilated conv / ConvTranspose and all their data gradients), one generic weight
// gradient kernel, and fp32 pointwise kernels with float64 reductions.  Views are `ubr_view` with fp32 elements
// (`ld` in elements); weights are read in the reference layouts directly (no packing).
#include "common.cuh"

namespace ubr {

// ------------------------------------------------------------------------------------------------
// generic convolution:  out[n,oh,ow,j] (+)= bias[j] + sum_{r,s,k} in[n,ih,iw,k] * w[j*ws_n + k*ws_k + r*kw+s]
//   direct     : ih = oh*stride + r*dil - pad
//   transposed : ih = (oh + pad - r*dil) / stride   when divisible (else the tap does not touch this output)
// The reduce index is flattened kk = (r*kw+s)*in_c + k so that in_c = 1 (stem) and in_c = 512 run the same loop.
// ------------------------------------------------------------------------------------------------
struct F32Conv {
  const float* in; float* out; const float* w; const float* bias;
  int n, in_h, in_w, in_c, in_ld;
  int out_h, out_w, out_c, out_ld;
  long ws_n, ws_k;
  int kh, kw, stride, pad, dil, transposed, accumulate;
};

constexpr int FBM = 64, FBN = 64, FBK = 16;

__global__ void __launch_bounds__(256) f32_conv_kernel(const F32Conv p) {
  __shared__ float As[FBK][FBM + 4];
  __shared__ float Bs[FBK][FBN + 4];
  const int tid = threadIdx.x;
  const long npix = (long)p.n * p.out_h * p.out_w;
  const long pix0 = (long)blockIdx.x * FBM;
  const int ch0 = blockIdx.y * FBN;
  const int ktot = p.kh * p.kw * p.in_c;
  // A-load role: this thread fetches reduce column (tid % 16) of pixels (tid / 16) + 16*j
  const int a_k = tid & 15;
  int a_n[4], a_h[4], a_w[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    long pix = pix0 + (tid >> 4) + 16 * j;
    if (pix < npix) {
      a_w[j] = (int)(pix % p.out_w);
      long t = pix / p.out_w;
      a_h[j] = (int)(t % p.out_h);
      a_n[j] = (int)(t / p.out_h);
    } else {
      a_n[j] = -1; a_h[j] = 0; a_w[j] = 0;
    }
  }
  // B-load role: output channel (tid % 64), reduce rows (tid / 64) + 4*j
  const int b_n = ch0 + (tid & 63);
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  const int ty = tid >> 4, tx = tid & 15;      // compute role: pixels ty*4.., channels tx*4..

  for (int k0 = 0; k0 < ktot; k0 += FBK) {
    {
      const int kk = k0 + a_k;
      int tap = 0, ci = 0, r = 0, s = 0;
      const bool kval = kk < ktot;
      if (kval) { tap = kk / p.in_c; ci = kk - tap * p.in_c; r = tap / p.kw; s = tap - r * p.kw; }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float v = 0.f;
        if (kval && a_n[j] >= 0) {
          int ih, iw;
          bool ok = true;
          if (!p.transposed) {
            ih = a_h[j] * p.stride + r * p.dil - p.pad;
            iw = a_w[j] * p.stride + s * p.dil - p.pad;
          } else {
            const int th = a_h[j] + p.pad - r * p.dil, tw = a_w[j] + p.pad - s * p.dil;
            ok = th >= 0 && tw >= 0 && (th % p.stride) == 0 && (tw % p.stride) == 0;
            ih = th / p.stride; iw = tw / p.stride;
          }
          if (ok && ih >= 0 && ih < p.in_h && iw >= 0 && iw < p.in_w)
            v = __ldg(p.in + (((long)a_n[j] * p.in_h + ih) * p.in_w + iw) * p.in_ld + ci);
        }
        As[a_k][(tid >> 4) + 16 * j] = v;
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int kl = (tid >> 6) + 4 * j;
      const int kk = k0 + kl;
      float v = 0.f;
      if (kk < ktot && b_n < p.out_c) {
        const int tap = kk / p.in_c, ci = kk - tap * p.in_c;
        v = __ldg(p.w + (long)b_n * p.ws_n + (long)ci * p.ws_k + tap);
      }
      Bs[kl][tid & 63] = v;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < FBK; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long pix = pix0 + ty * 4 + i;
    if (pix >= npix) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int ch = ch0 + tx * 4 + j;
      if (ch >= p.out_c) continue;
      float v = acc[i][j] + (p.bias ? p.bias[ch] : 0.f);
      float* o = p.out + pix * p.out_ld + ch;
      *o = p.accumulate ? (*o + v) : v;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// generic weight gradient:  dw[a*(cf*taps) + b*taps + tap] = sum_{n,h,w} coarse[n,h,w,a] * fine[n, h*stride + r*dil - pad,
// w*stride + s*dil - pad, b].  Conv: coarse = dy, fine = x (dw is OIHW); ConvTranspose: coarse = x, fine = dy (IOHW).
// float64 atomics into `acc` (zeroed by the caller), converted to fp32 by f32_wgrad_finish.
// ------------------------------------------------------------------------------------------------
struct F32Wgrad {
  const float* coarse; const float* fine; double* acc;
  int n, ch, cw, cc, c_ld;
  int fh, fw, cf, f_ld;
  int kh, kw, stride, pad, dil;
  int pix_per_block;
};

constexpr int WBM = 32, WBN = 64, WBK = 16;

__global__ void __launch_bounds__(256) f32_wgrad_kernel(const F32Wgrad p) {
  __shared__ float As[WBK][WBM + 4];
  __shared__ float Bs[WBK][WBN + 4];
  const int tid = threadIdx.x;
  const int taps = p.kh * p.kw;
  const int ntot = taps * p.cf;
  const int m0 = blockIdx.x * WBM, n0 = blockIdx.y * WBN;
  const long npix = (long)p.n * p.ch * p.cw;
  const long pbeg = (long)blockIdx.z * p.pix_per_block;
  const long pend = pbeg + p.pix_per_block < npix ? pbeg + p.pix_per_block : npix;
  // B-load role: column nn = n0 + (tid % 64), pixel rows (tid / 64) + 4*j
  const int nn = n0 + (tid & 63);
  int b_tap = 0, b_cf = 0, b_r = 0, b_s = 0;
  const bool nval = nn < ntot;
  if (nval) { b_tap = nn / p.cf; b_cf = nn - b_tap * p.cf; b_r = b_tap / p.kw; b_s = b_tap - b_r * p.kw; }
  const int ty = tid >> 4, tx = tid & 15;      // compute: rows ty*2.., cols tx*4..
  float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
  for (long q0 = pbeg; q0 < pend; q0 += WBK) {
    {   // A tile: 16 pixels x 32 channels = 512 elements, 2 per thread
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int e = tid + 256 * j;
        const int pl = e >> 5, m = e & 31;
        const long q = q0 + pl;
        float v = 0.f;
        if (q < pend && m0 + m < p.cc) v = __ldg(p.coarse + q * p.c_ld + m0 + m);
        As[pl][m] = v;
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int pl = (tid >> 6) + 4 * j;
      const long q = q0 + pl;
      float v = 0.f;
      if (nval && q < pend) {
        const int w = (int)(q % p.cw);
        const long t = q / p.cw;
        const int h = (int)(t % p.ch), n = (int)(t / p.ch);
        const int ih = h * p.stride + b_r * p.dil - p.pad, iw = w * p.stride + b_s * p.dil - p.pad;
        if (ih >= 0 && ih < p.fh && iw >= 0 && iw < p.fw)
          v = __ldg(p.fine + (((long)n * p.fh + ih) * p.fw + iw) * p.f_ld + b_cf);
      }
      Bs[pl][tid & 63] = v;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < WBK; ++k) {
      const float a0 = As[k][ty * 2], a1 = As[k][ty * 2 + 1];
      const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      acc[0][0] = fmaf(a0, b.x, acc[0][0]); acc[0][1] = fmaf(a0, b.y, acc[0][1]);
      acc[0][2] = fmaf(a0, b.z, acc[0][2]); acc[0][3] = fmaf(a0, b.w, acc[0][3]);
      acc[1][0] = fmaf(a1, b.x, acc[1][0]); acc[1][1] = fmaf(a1, b.y, acc[1][1]);
      acc[1][2] = fmaf(a1, b.z, acc[1][2]); acc[1][3] = fmaf(a1, b.w, acc[1][3]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int m = m0 + ty * 2 + i;
    if (m >= p.cc) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = n0 + tx * 4 + j;
      if (c >= ntot) continue;
      const int tap = c / p.cf, b = c - tap * p.cf;
      atomicAdd(p.acc + ((long)m * p.cf + b) * taps + tap, (double)acc[i][j]);
    }
  }
}

__global__ void f32_from_f64_kernel(const double* __restrict__ src, float* __restrict__ dst, long count, int accumulate) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) dst[i] = accumulate ? dst[i] + (float)src[i] : (float)src[i];
}

}  // namespace ubr

namespace ubr {

// ------------------------------------------------------------------------------------------------
// pointwise fp32 kernels.  A thread walks pixels for one channel lane group; reductions go through warp shuffles
// into float64 atomics (statistics over up to 8.4 M pixels keep all 24 bits).
// ------------------------------------------------------------------------------------------------
struct FView { float* p; long npix; int c, ld; };
static inline FView fview(const ubr_view& v) { return FView{(float*)v.ptr, (long)v.n * v.h * v.w, v.c, v.ld}; }

// sums[0:c) += sum y, sums[c:2c) += sum y^2.   block = 256 threads = 8 pixel rows x 32 channel lanes
__global__ void f32_stats_kernel(FView y, double* sums, int chunk) {
  const int lane = threadIdx.x & 31, row = threadIdx.x >> 5;
  __shared__ double sh[2][8][32];
  for (int c0 = 0; c0 < y.c; c0 += 32) {
    const int c = c0 + lane;
    double s = 0.0, q = 0.0;
    if (c < y.c) {
      const long beg = (long)blockIdx.x * chunk;
      const long end = beg + chunk < y.npix ? beg + chunk : y.npix;
      float fs = 0.f, fq = 0.f;
      int cnt = 0;
      for (long px = beg + row; px < end; px += 8) {
        const float v = y.p[px * y.ld + c];
        fs += v; fq = fmaf(v, v, fq);
        if (++cnt == 64) { s += fs; q += fq; fs = fq = 0.f; cnt = 0; }
      }
      s += fs; q += fq;
    }
    sh[0][row][lane] = s; sh[1][row][lane] = q;
    __syncthreads();
    if (row == 0 && c < y.c) {
      double ts = 0.0, tq = 0.0;
      for (int r = 0; r < 8; ++r) { ts += sh[0][r][lane]; tq += sh[1][r][lane]; }
      atomicAdd(sums + c, ts);
      atomicAdd(sums + y.c + c, tq);
    }
    __syncthreads();
  }
}

__global__ void f32_bn_finalize_kernel(const double* sums, int c, double count, const float* gamma, const float* beta,
                                       float* rmean, float* rvar, float momentum, float eps, float* scale, float* shift,
                                       float* mean, float* invstd) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= c) return;
  const double m = sums[i] / count;
  double var = sums[c + i] / count - m * m;
  if (var < 0.0) var = 0.0;
  const double is = 1.0 / sqrt(var + (double)eps);
  const double g = gamma ? (double)gamma[i] : 1.0, b = beta ? (double)beta[i] : 0.0;
  scale[i] = (float)(g * is);
  shift[i] = (float)(b - m * g * is);
  mean[i] = (float)m;
  invstd[i] = (float)is;
  if (rmean) rmean[i] = (float)((1.0 - momentum) * rmean[i] + momentum * m);
  if (rvar) {
    const double unb = count > 1.0 ? var * count / (count - 1.0) : var;
    rvar[i] = (float)((1.0 - momentum) * rvar[i] + momentum * unb);
  }
}

__global__ void f32_sums_to_float_kernel(const double* sums, int c, float* out, int accumulate) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < c) out[i] = accumulate ? out[i] + (float)sums[i] : (float)sums[i];
}

// out = relu?(y*scale+shift)
__global__ void f32_bn_act_kernel(FView y, const float* scale, const float* shift, int relu, FView out) {
  const long total = y.npix * y.c;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const long px = i / y.c; const int c = (int)(i - px * y.c);
    float v = fmaf(y.p[px * y.ld + c], scale[c], shift[c]);
    if (relu) v = fmaxf(v, 0.f);
    out.p[px * out.ld + c] = v;
  }
}

// out = relu( relu(y2*s2+b2) + (res*sr+br | res) )     (common_layers.py:45-56)
__global__ void f32_block_tail_kernel(FView y2, const float* s2, const float* b2, FView res, const float* sr,
                                      const float* br, FView out) {
  const long total = y2.npix * y2.c;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const long px = i / y2.c; const int c = (int)(i - px * y2.c);
    const float a = fmaxf(fmaf(y2.p[px * y2.ld + c], s2[c], b2[c]), 0.f);
    float r = res.p[px * res.ld + c];
    if (sr) r = fmaf(r, sr[c], br[c]);
    out.p[px * out.ld + c] = fmaxf(a + r, 0.f);
  }
}

// du = dout * (out > 0)
__global__ void f32_relu_mask_kernel(FView dout, FView out, FView du) {
  const long total = dout.npix * dout.c;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const long px = i / dout.c; const int c = (int)(i - px * dout.c);
    du.p[px * du.ld + c] = out.p[px * out.ld + c] > 0.f ? dout.p[px * dout.ld + c] : 0.f;
  }
}

// BN(+ReLU) backward, pass 1: sums[0:c) += sum dz, sums[c:2c) += sum dz*xhat,  dz = dout * (!relu | bn(y) > 0)
__global__ void f32_bn_bwd_reduce_kernel(FView dout, FView y, const float* scale, const float* shift, const float* mean,
                                         const float* invstd, int relu, double* sums, int chunk) {
  const int lane = threadIdx.x & 31, row = threadIdx.x >> 5;
  __shared__ double sh[2][8][32];
  for (int c0 = 0; c0 < y.c; c0 += 32) {
    const int c = c0 + lane;
    double s = 0.0, q = 0.0;
    if (c < y.c) {
      const float sc = scale[c], sf = shift[c], mu = mean[c], is = invstd[c];
      const long beg = (long)blockIdx.x * chunk;
      const long end = beg + chunk < y.npix ? beg + chunk : y.npix;
      float fs = 0.f, fq = 0.f;
      int cnt = 0;
      for (long px = beg + row; px < end; px += 8) {
        const float yv = y.p[px * y.ld + c];
        float dz = dout.p[px * dout.ld + c];
        if (relu && !(fmaf(yv, sc, sf) > 0.f)) dz = 0.f;
        fs += dz; fq = fmaf(dz, (yv - mu) * is, fq);
        if (++cnt == 64) { s += fs; q += fq; fs = fq = 0.f; cnt = 0; }
      }
      s += fs; q += fq;
    }
    sh[0][row][lane] = s; sh[1][row][lane] = q;
    __syncthreads();
    if (row == 0 && c < y.c) {
      double ts = 0.0, tq = 0.0;
      for (int r = 0; r < 8; ++r) { ts += sh[0][r][lane]; tq += sh[1][r][lane]; }
      atomicAdd(sums + c, ts);
      atomicAdd(sums + y.c + c, tq);
    }
    __syncthreads();
  }
}

// pass 2: dy = scale * (dz - sum_dz/count - xhat * sum_dzx/count)
__global__ void f32_bn_bwd_apply_kernel(FView dout, FView y, const float* scale, const float* shift, const float* mean,
                                        const float* invstd, int relu, const double* sums, double count, FView dy) {
  const long total = y.npix * y.c;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const long px = i / y.c; const int c = (int)(i - px * y.c);
    const float yv = y.p[px * y.ld + c];
    float dz = dout.p[px * dout.ld + c];
    if (relu && !(fmaf(yv, scale[c], shift[c]) > 0.f)) dz = 0.f;
    const float xh = (yv - mean[c]) * invstd[c];
    const float m1 = (float)(sums[c] / count), m2 = (float)(sums[y.c + c] / count);
    dy.p[px * dy.ld + c] = scale[c] * (dz - m1 - xh * m2);
  }
}

// MaxPool2d(3,2,1): first maximum in scan order, idx = kh*3+kw
__global__ void f32_maxpool_fwd_kernel(FView x, int n, int h, int w, FView y, uint8_t* idx) {
  const int oh = h / 2, ow = w / 2;
  const long total = (long)n * oh * ow * x.c;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i % x.c); long t = i / x.c;
    const int wo = (int)(t % ow); t /= ow;
    const int ho = (int)(t % oh); const int b = (int)(t / oh);
    float best = -INFINITY; int bi = 0;
    for (int r = 0; r < 3; ++r)
      for (int s = 0; s < 3; ++s) {
        const int ih = 2 * ho - 1 + r, iw = 2 * wo - 1 + s;
        if (ih < 0 || ih >= h || iw < 0 || iw >= w) continue;
        const float v = x.p[(((long)b * h + ih) * w + iw) * x.ld + c];
        if (v > best) { best = v; bi = r * 3 + s; }
      }
    y.p[(i / x.c) * y.ld + c] = best;
    if (idx) idx[i] = (uint8_t)bi;
  }
}

// dx (+)= sum over the (<= 4) windows covering this input pixel whose argmax is this pixel
__global__ void f32_maxpool_bwd_kernel(const uint8_t* idx, FView dy, int n, int h, int w, FView dx, int accumulate) {
  const int oh = h / 2, ow = w / 2;
  const long total = (long)n * h * w * dx.c;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i % dx.c); long t = i / dx.c;
    const int iw = (int)(t % w); t /= w;
    const int ih = (int)(t % h); const int b = (int)(t / h);
    float g = 0.f;
    for (int r = 0; r < 3; ++r) {
      const int th = ih + 1 - r;
      if (th < 0 || (th & 1)) continue;
      const int ho = th >> 1;
      if (ho >= oh) continue;
      for (int s = 0; s < 3; ++s) {
        const int tw = iw + 1 - s;
        if (tw < 0 || (tw & 1)) continue;
        const int wo = tw >> 1;
        if (wo >= ow) continue;
        const long o = (((long)b * oh + ho) * ow + wo);
        if (idx[o * dx.c + c] == r * 3 + s) g += dy.p[o * dy.ld + c];
      }
    }
    float* d = dx.p + (i / dx.c) * dx.ld + c;
    *d = accumulate ? *d + g : g;
  }
}

__global__ void f32_add_kernel(FView a, FView b, FView c3, FView out) {
  const long total = a.npix * a.c;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const long px = i / a.c; const int c = (int)(i - px * a.c);
    float v = a.p[px * a.ld + c];
    if (b.p) v += b.p[px * b.ld + c];
    if (c3.p) v += c3.p[px * c3.ld + c];
    out.p[px * out.ld + c] = v;
  }
}

// dlogits[n,hw,k] = dlogp[n,k,hw] - exp(logp[n,k,hw]) * sum_j dlogp[n,j,hw]
__global__ void f32_logsoftmax_bwd_kernel(const float* logp, const float* dlogp, int n, int ncls, int hw, FView dl) {
  const long total = (long)n * hw;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const long b = i / hw, px = i - b * hw;
    float s = 0.f;
    for (int k = 0; k < ncls; ++k) s += dlogp[(b * ncls + k) * hw + px];
    for (int k = 0; k < ncls; ++k) {
      const long o = (b * ncls + k) * hw + px;
      dl.p[i * dl.ld + k] = dlogp[o] - __expf(logp[o]) * s;
    }
  }
}

static inline int pw_grid(long total) {
  long g = (total + 255) / 256;
  const long cap = (long)sm_count() * 16;
  return (int)(g < cap ? (g > 0 ? g : 1) : cap);
}

}  // namespace ubr

// ================================================================================================
// C ABI
// ================================================================================================
using namespace ubr;

static int f32_check_view(const ubr_view& v, const char* what) {
  UBR_REQUIRE(v.ptr != nullptr && v.n > 0 && v.h > 0 && v.w > 0 && v.c > 0 && v.ld >= v.c, "%s: bad fp32 view", what);
  return 0;
}

static int f32_conv_launch(const ubr_view& in, const ubr_view& out, const float* w, long ws_n, long ws_k, const float* bias,
                           int k, int stride, int pad, int dil, int transposed, int accumulate, cudaStream_t st) {
  int rc = ensure_device();
  if (rc) return rc;
  if ((rc = f32_check_view(in, "f32 conv input"))) return rc;
  if ((rc = f32_check_view(out, "f32 conv output"))) return rc;
  UBR_REQUIRE(w != nullptr && k >= 1 && k <= 7 && stride >= 1 && dil >= 1 && in.n == out.n, "f32 conv: bad arguments");
  F32Conv p;
  p.in = (const float*)in.ptr; p.out = (float*)out.ptr; p.w = w; p.bias = bias;
  p.n = in.n; p.in_h = in.h; p.in_w = in.w; p.in_c = in.c; p.in_ld = in.ld;
  p.out_h = out.h; p.out_w = out.w; p.out_c = out.c; p.out_ld = out.ld;
  p.ws_n = ws_n; p.ws_k = ws_k; p.kh = k; p.kw = k; p.stride = stride; p.pad = pad; p.dil = dil;
  p.transposed = transposed; p.accumulate = accumulate;
  const long npix = (long)out.n * out.h * out.w;
  dim3 grid((unsigned)((npix + FBM - 1) / FBM), (unsigned)((out.c + FBN - 1) / FBN), 1);
  f32_conv_kernel<<<grid, 256, 0, st>>>(p);
  UBR_LAUNCH_CHECK("f32_conv_kernel");
  return 0;
}

static int f32_wgrad_launch(const ubr_view& coarse, const ubr_view& fine, int k, int stride, int pad, int dil, float* dw,
                            int accumulate, void* ws, uint64_t ws_bytes, cudaStream_t st) {
  int rc = ensure_device();
  if (rc) return rc;
  if ((rc = f32_check_view(coarse, "f32 wgrad coarse"))) return rc;
  if ((rc = f32_check_view(fine, "f32 wgrad fine"))) return rc;
  const long count = (long)coarse.c * fine.c * k * k;
  UBR_REQUIRE(dw != nullptr && ws != nullptr && ws_bytes >= (uint64_t)count * 8, "f32 wgrad: workspace of %ld bytes needed",
              count * 8);
  UBR_CUDA_OK(cudaMemsetAsync(ws, 0, (size_t)count * 8, st));
  F32Wgrad p;
  p.coarse = (const float*)coarse.ptr; p.fine = (const float*)fine.ptr; p.acc = (double*)ws;
  p.n = coarse.n; p.ch = coarse.h; p.cw = coarse.w; p.cc = coarse.c; p.c_ld = coarse.ld;
  p.fh = fine.h; p.fw = fine.w; p.cf = fine.c; p.f_ld = fine.ld;
  p.kh = k; p.kw = k; p.stride = stride; p.pad = pad; p.dil = dil;
  const long npix = (long)coarse.n * coarse.h * coarse.w;
  const int mt = (coarse.c + WBM - 1) / WBM, nt = (k * k * fine.c + WBN - 1) / WBN;
  // enough pixel splits to fill the GPU a few times over, at least 256 pixels each
  long splits = (long)sm_count() * 8 / ((long)mt * nt) + 1;
  long ppb = (npix + splits - 1) / splits;
  if (ppb < 256) ppb = 256;
  ppb = (ppb + WBK - 1) / WBK * WBK;
  splits = (npix + ppb - 1) / ppb;
  p.pix_per_block = (int)ppb;
  dim3 grid((unsigned)mt, (unsigned)nt, (unsigned)splits);
  f32_wgrad_kernel<<<grid, 256, 0, st>>>(p);
  UBR_LAUNCH_CHECK("f32_wgrad_kernel");
  f32_from_f64_kernel<<<(unsigned)((count + 255) / 256), 256, 0, st>>>((const double*)ws, dw, count, accumulate);
  UBR_LAUNCH_CHECK("f32_from_f64_kernel");
  return 0;
}

extern "C" {

int ubr_f32_conv2d_fwd(ubr_view x, const float* w, const float* bias, int out_ch, int ksize, int stride, int pad,
                       int dilation, ubr_view y, void* stream) {
  UBR_REQUIRE(y.c == out_ch, "f32 conv fwd: y.c != out_ch");
  return f32_conv_launch(x, y, w, (long)x.c * ksize * ksize, (long)ksize * ksize, bias, ksize, stride, pad, dilation, 0, 0,
                         (cudaStream_t)stream);
}

int ubr_f32_conv2d_dgrad(ubr_view dy, const float* w, int in_ch, int ksize, int stride, int pad, int dilation, ubr_view dx,
                         int accumulate, void* stream) {
  UBR_REQUIRE(dx.c == in_ch, "f32 conv dgrad: dx.c != in_ch");
  return f32_conv_launch(dy, dx, w, (long)ksize * ksize, (long)in_ch * ksize * ksize, nullptr, ksize, stride, pad, dilation,
                         1, accumulate, (cudaStream_t)stream);
}

int ubr_f32_conv2d_wgrad(ubr_view x, ubr_view dy, int ksize, int stride, int pad, int dilation, float* dw, int accumulate,
                         void* workspace, uint64_t workspace_bytes, void* stream) {
  return f32_wgrad_launch(dy, x, ksize, stride, pad, dilation, dw, accumulate, workspace, workspace_bytes,
                          (cudaStream_t)stream);
}

int ubr_f32_deconv_fwd(ubr_view x, const float* w, int out_ch, int ksize, int stride, int pad, ubr_view y, void* stream) {
  UBR_REQUIRE(y.c == out_ch, "f32 deconv fwd: y.c != out_ch");
  return f32_conv_launch(x, y, w, (long)ksize * ksize, (long)out_ch * ksize * ksize, nullptr, ksize, stride, pad, 1, 1, 0,
                         (cudaStream_t)stream);
}

int ubr_f32_deconv_dgrad(ubr_view dy, const float* w, int in_ch, int ksize, int stride, int pad, ubr_view dx, void* stream) {
  UBR_REQUIRE(dx.c == in_ch, "f32 deconv dgrad: dx.c != in_ch");
  return f32_conv_launch(dy, dx, w, (long)dy.c * ksize * ksize, (long)ksize * ksize, nullptr, ksize, stride, pad, 1, 0, 0,
                         (cudaStream_t)stream);
}

int ubr_f32_deconv_wgrad(ubr_view x, ubr_view dy, int ksize, int stride, int pad, float* dw, int accumulate, void* workspace,
                         uint64_t workspace_bytes, void* stream) {
  return f32_wgrad_launch(x, dy, ksize, stride, pad, 1, dw, accumulate, workspace, workspace_bytes, (cudaStream_t)stream);
}

int ubr_f32_bn_stats(ubr_view y, double* sums, void* stream) {
  int rc = ensure_device();
  if (rc) return rc;
  if ((rc = f32_check_view(y, "f32 bn_stats"))) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  UBR_CUDA_OK(cudaMemsetAsync(sums, 0, sizeof(double) * 2 * y.c, st));
  FView v = fview(y);
  const int chunk = 2048;
  f32_stats_kernel<<<(unsigned)((v.npix + chunk - 1) / chunk), 256, 0, st>>>(v, sums, chunk);
  UBR_LAUNCH_CHECK("f32_stats_kernel");
  return 0;
}

int ubr_f32_bn_finalize(const double* sums, int c, double count, const float* gamma, const float* beta, float* running_mean,
                        float* running_var, float momentum, float eps, float* scale, float* shift, float* mean,
                        float* invstd, void* stream) {
  f32_bn_finalize_kernel<<<(c + 127) / 128, 128, 0, (cudaStream_t)stream>>>(sums, c, count, gamma, beta, running_mean,
                                                                             running_var, momentum, eps, scale, shift,
                                                                             mean, invstd);
  UBR_LAUNCH_CHECK("f32_bn_finalize_kernel");
  return 0;
}

int ubr_f32_channel_sum(ubr_view a, double* workspace, float* out, int accumulate, void* stream) {
  int rc = ubr_f32_bn_stats(a, workspace, stream);
  if (rc) return rc;
  f32_sums_to_float_kernel<<<(a.c + 127) / 128, 128, 0, (cudaStream_t)stream>>>(workspace, a.c, out, accumulate);
  UBR_LAUNCH_CHECK("f32_sums_to_float_kernel");
  return 0;
}

int ubr_f32_bn_act(ubr_view y, const float* scale, const float* shift, int relu, ubr_view out, void* stream) {
  FView a = fview(y), o = fview(out);
  f32_bn_act_kernel<<<pw_grid(a.npix * a.c), 256, 0, (cudaStream_t)stream>>>(a, scale, shift, relu, o);
  UBR_LAUNCH_CHECK("f32_bn_act_kernel");
  return 0;
}

int ubr_f32_block_tail(ubr_view y2, const float* s2, const float* b2, ubr_view res, const float* sr, const float* br,
                       ubr_view out, void* stream) {
  FView a = fview(y2);
  f32_block_tail_kernel<<<pw_grid(a.npix * a.c), 256, 0, (cudaStream_t)stream>>>(a, s2, b2, fview(res), sr, br, fview(out));
  UBR_LAUNCH_CHECK("f32_block_tail_kernel");
  return 0;
}

int ubr_f32_relu_mask(ubr_view dout, ubr_view out, ubr_view du, void* stream) {
  FView a = fview(dout);
  f32_relu_mask_kernel<<<pw_grid(a.npix * a.c), 256, 0, (cudaStream_t)stream>>>(a, fview(out), fview(du));
  UBR_LAUNCH_CHECK("f32_relu_mask_kernel");
  return 0;
}

int ubr_f32_bn_act_bwd(ubr_view dout, ubr_view y, const float* scale, const float* shift, const float* mean,
                       const float* invstd, int relu, double* workspace, double count, ubr_view dy, float* dgamma,
                       float* dbeta, int accumulate, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  UBR_CUDA_OK(cudaMemsetAsync(workspace, 0, sizeof(double) * 2 * y.c, st));
  FView d = fview(dout), v = fview(y);
  const int chunk = 2048;
  f32_bn_bwd_reduce_kernel<<<(unsigned)((v.npix + chunk - 1) / chunk), 256, 0, st>>>(d, v, scale, shift, mean, invstd, relu,
                                                                                     workspace, chunk);
  UBR_LAUNCH_CHECK("f32_bn_bwd_reduce_kernel");
  f32_bn_bwd_apply_kernel<<<pw_grid(v.npix * v.c), 256, 0, st>>>(d, v, scale, shift, mean, invstd, relu, workspace, count,
                                                                 fview(dy));
  UBR_LAUNCH_CHECK("f32_bn_bwd_apply_kernel");
  if (dbeta) {
    f32_sums_to_float_kernel<<<(y.c + 127) / 128, 128, 0, st>>>(workspace, y.c, dbeta, accumulate);
    UBR_LAUNCH_CHECK("f32_sums_to_float_kernel");
  }
  if (dgamma) {
    f32_sums_to_float_kernel<<<(y.c + 127) / 128, 128, 0, st>>>(workspace + y.c, y.c, dgamma, accumulate);
    UBR_LAUNCH_CHECK("f32_sums_to_float_kernel");
  }
  return 0;
}

int ubr_f32_maxpool3s2_fwd(ubr_view x, ubr_view y, uint8_t* idx, void* stream) {
  FView a = fview(x);
  f32_maxpool_fwd_kernel<<<pw_grid(a.npix * a.c / 4), 256, 0, (cudaStream_t)stream>>>(a, x.n, x.h, x.w, fview(y), idx);
  UBR_LAUNCH_CHECK("f32_maxpool_fwd_kernel");
  return 0;
}

int ubr_f32_maxpool3s2_bwd(const uint8_t* idx, ubr_view dy, ubr_view dx, int accumulate, void* stream) {
  FView d = fview(dx);
  f32_maxpool_bwd_kernel<<<pw_grid(d.npix * d.c), 256, 0, (cudaStream_t)stream>>>(idx, fview(dy), dx.n, dx.h, dx.w, d,
                                                                                  accumulate);
  UBR_LAUNCH_CHECK("f32_maxpool_bwd_kernel");
  return 0;
}

int ubr_f32_add(ubr_view a, ubr_view b, ubr_view c, ubr_view out, void* stream) {
  FView av = fview(a);
  FView bv = b.ptr ? fview(b) : FView{nullptr, 0, 0, 0};
  FView cv = c.ptr ? fview(c) : FView{nullptr, 0, 0, 0};
  f32_add_kernel<<<pw_grid(av.npix * av.c), 256, 0, (cudaStream_t)stream>>>(av, bv, cv, fview(out));
  UBR_LAUNCH_CHECK("f32_add_kernel");
  return 0;
}

int ubr_f32_logsoftmax_bwd(const float* logp, const float* dlogp, int n, int ncls, int hw, ubr_view dlogits, void* stream) {
  UBR_REQUIRE(dlogits.c == ncls, "f32 logsoftmax_bwd: dlogits.c != ncls");
  f32_logsoftmax_bwd_kernel<<<pw_grid((long)n * hw), 256, 0, (cudaStream_t)stream>>>(logp, dlogp, n, ncls, hw, fview(dlogits));
  UBR_LAUNCH_CHECK("f32_logsoftmax_bwd_kernel");
  return 0;
}

}  // extern "C"
