// Stem convolution: nn.Conv2d(1, OC, 7, stride 1, pad 3, bias) on the fp32 single-channel crop
// (reference models/ub_uresnet.py:41,94).  Cin = 1 means K = 49: not a tensor-core GEMM.  Direct
// CUDA-core kernels; forward is bounded by the bf16 NHWC write (32 B/pixel for OC = 16).
#include "common.cuh"
#include "conv_internal.cuh"

namespace ubr {

constexpr int kTW = 32, kTH = 8;            // output tile per block (256 threads, one pixel each)
constexpr int kHW = kTW + 6, kHH = kTH + 6; // halo tile

// ---- forward: register-tiled direct convolution.  Block = 32 x 8 threads, tile = 128 x 8 output pixels; a thread
// owns 4 consecutive pixels x 16 channels (64 fp32 accumulators), so each smem weight vector (warp broadcast) feeds
// 4 pixels and each input value 16 channels: 16 FMAs per shared-memory load instead of 3.
constexpr int kFTW = 128, kFTH = 8;
constexpr int kFHW = kFTW + 8, kFHH = kFTH + 6;      // tile columns start at w0-4 (16-byte aligned float4 reads)

__global__ void __launch_bounds__(256) stem_fwd_kernel(const float* __restrict__ x, int n, int h, int w,
                                                       const float* __restrict__ wgt, const float* __restrict__ bias,
                                                       int oc, __nv_bfloat16* __restrict__ y, int yld) {
  pdl_prologue();
  __shared__ __align__(16) float tile[kFHH][kFHW];
  __shared__ __align__(16) float wsm[49][16];
  __shared__ float bsm[16];
  const int ocg = oc / 16;
  const int b = blockIdx.z / ocg, g = blockIdx.z % ocg;
  const int w0 = blockIdx.x * kFTW, h0 = blockIdx.y * kFTH;
  const float* img = x + (size_t)b * h * w;
  for (int i = threadIdx.x; i < kFHH * kFHW; i += 256) {
    const int r = i / kFHW, c = i % kFHW;
    const int ih = h0 + r - 3, iw = w0 + c - 4;
    tile[r][c] = (ih >= 0 && ih < h && iw >= 0 && iw < w) ? img[(size_t)ih * w + iw] : 0.f;
  }
  for (int i = threadIdx.x; i < 49 * 16; i += 256) {
    const int t = i / 16, o = i % 16;
    wsm[t][o] = wgt[(size_t)(g * 16 + o) * 49 + t];
  }
  if (threadIdx.x < 16) bsm[threadIdx.x] = bias ? bias[g * 16 + threadIdx.x] : 0.f;
  __syncthreads();
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  float acc[4][16];
#pragma unroll
  for (int q = 0; q < 4; ++q)
#pragma unroll
    for (int o = 0; o < 16; ++o) acc[q][o] = bsm[o];
#pragma unroll
  for (int r = 0; r < 7; ++r) {
    // input columns (w0 + 4*tx - 4) .. (+11): outputs 4*tx+q need columns 4*tx+q-3 .. 4*tx+q+3 = v[q+1 .. q+7]
    float v[12];
    const float4* rp = reinterpret_cast<const float4*>(&tile[ty + r][4 * tx]);
    const float4 a = rp[0], bq = rp[1], c = rp[2];
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = bq.x; v[5] = bq.y; v[6] = bq.z; v[7] = bq.w;
    v[8] = c.x; v[9] = c.y; v[10] = c.z; v[11] = c.w;
#pragma unroll
    for (int s = 0; s < 7; ++s) {
      const float4* wp = reinterpret_cast<const float4*>(wsm[r * 7 + s]);
      const float4 w0v = wp[0], w1v = wp[1], w2v = wp[2], w3v = wp[3];
      const float wv[16] = {w0v.x, w0v.y, w0v.z, w0v.w, w1v.x, w1v.y, w1v.z, w1v.w,
                            w2v.x, w2v.y, w2v.z, w2v.w, w3v.x, w3v.y, w3v.z, w3v.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float xv = v[q + 1 + s];
#pragma unroll
        for (int o = 0; o < 16; ++o) acc[q][o] = fmaf(xv, wv[o], acc[q][o]);
      }
    }
  }
  const int oh = h0 + ty;
  if (oh < h) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int ow = w0 + 4 * tx + q;
      if (ow < w) {
        __nv_bfloat16* dst = y + (((size_t)b * h + oh) * w + ow) * yld + g * 16;
        uint4 u0, u1;
        u0.x = pack_bf16x2(acc[q][0], acc[q][1]); u0.y = pack_bf16x2(acc[q][2], acc[q][3]);
        u0.z = pack_bf16x2(acc[q][4], acc[q][5]); u0.w = pack_bf16x2(acc[q][6], acc[q][7]);
        u1.x = pack_bf16x2(acc[q][8], acc[q][9]); u1.y = pack_bf16x2(acc[q][10], acc[q][11]);
        u1.z = pack_bf16x2(acc[q][12], acc[q][13]); u1.w = pack_bf16x2(acc[q][14], acc[q][15]);
        reinterpret_cast<uint4*>(dst)[0] = u0;
        reinterpret_cast<uint4*>(dst)[1] = u1;
      }
    }
  }
}

// weight gradient: dW[o][t] = sum_p dy[p][o] * x[p + t - 3],  db[o] = sum_p dy[p][o]
// persistent blocks loop over 32x8 tiles.  Thread layout: 16 tap groups x 16 pixel subsets; a thread owns FOUR
// horizontally adjacent taps (one filter row, s0 = 0 or 4) x 16 channels = 64 accumulators, so the 4 broadcast
// 16-byte dy loads of a pixel feed 64 FMAs.  Group 14 accumulates the bias gradient, group 15 idles.
constexpr int kWgBlocks = 296;
__global__ void __launch_bounds__(256) stem_wgrad_kernel(const float* __restrict__ x, int n, int h, int w,
                                                         const __nv_bfloat16* __restrict__ dy, int dyld, int oc,
                                                         float* __restrict__ partials /* [blocks][ocg][50][16] */) {
  pdl_prologue();
  __shared__ float tile[kHH][kHW + 2];
  __shared__ __align__(16) float dys[kTH * kTW][16];
  __shared__ float red[50][16];
  const int ocg = oc / 16;
  const int tiles_w = (w + kTW - 1) / kTW, tiles_h = (h + kTH - 1) / kTH;
  const long ntiles = (long)tiles_w * tiles_h * n;
  const int grp = threadIdx.x & 15, sub = threadIdx.x >> 4;
  const int r = grp >> 1, s0 = (grp & 1) * 4;       // grp 14 -> bias, grp 15 -> idle
  const bool is_tap = grp < 14, is_bias = grp == 14;
  for (int g = 0; g < ocg; ++g) {
    float acc[4][16];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
      for (int o = 0; o < 16; ++o) acc[q][o] = 0.f;
    for (long t = blockIdx.x; t < ntiles; t += gridDim.x) {
      const int tw = (int)(t % tiles_w), th = (int)((t / tiles_w) % tiles_h), b = (int)(t / ((long)tiles_w * tiles_h));
      const int w0 = tw * kTW, h0 = th * kTH;
      const float* img = x + (size_t)b * h * w;
      __syncthreads();
      for (int i = threadIdx.x; i < kHH * (kHW + 1); i += 256) {
        const int rr = i / (kHW + 1), cc = i % (kHW + 1);
        const int ih = h0 + rr - 3, iw = w0 + cc - 3;
        tile[rr][cc] = (ih >= 0 && ih < h && iw >= 0 && iw < w) ? img[(size_t)ih * w + iw] : 0.f;
      }
      {
        const int px = threadIdx.x % kTW, py = threadIdx.x / kTW;
        const int oh = h0 + py, ow = w0 + px;
        float f[16];
        if (oh < h && ow < w) {
          const uint4* src = reinterpret_cast<const uint4*>(dy + (((size_t)b * h + oh) * w + ow) * dyld + g * 16);
          const uint4 u0 = src[0], u1 = src[1];
          float2 p;
          p = unpack_bf16x2(u0.x); f[0] = p.x; f[1] = p.y;  p = unpack_bf16x2(u0.y); f[2] = p.x; f[3] = p.y;
          p = unpack_bf16x2(u0.z); f[4] = p.x; f[5] = p.y;  p = unpack_bf16x2(u0.w); f[6] = p.x; f[7] = p.y;
          p = unpack_bf16x2(u1.x); f[8] = p.x; f[9] = p.y;  p = unpack_bf16x2(u1.y); f[10] = p.x; f[11] = p.y;
          p = unpack_bf16x2(u1.z); f[12] = p.x; f[13] = p.y; p = unpack_bf16x2(u1.w); f[14] = p.x; f[15] = p.y;
        } else {
#pragma unroll
          for (int o = 0; o < 16; ++o) f[o] = 0.f;
        }
#pragma unroll
        for (int o = 0; o < 16; ++o) dys[threadIdx.x][o] = f[o];
      }
      __syncthreads();
      if (is_tap) {
        // this subset handles pixels sub, sub+16, ... of the 256-pixel tile
        for (int pix = sub; pix < kTH * kTW; pix += 16) {
          const int px = pix % kTW, py = pix / kTW;
          const float* xr = &tile[py + r][px + s0];
          const float xv[4] = {xr[0], xr[1], xr[2], xr[3]};      // tap s0+3 = 7 reads the zero pad column; discarded
          const float4* dp = reinterpret_cast<const float4*>(dys[pix]);
          const float4 d0 = dp[0], d1 = dp[1], d2 = dp[2], d3 = dp[3];
          const float dv[16] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w, d2.x, d2.y, d2.z, d2.w, d3.x, d3.y, d3.z, d3.w};
#pragma unroll
          for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int o = 0; o < 16; ++o) acc[q][o] = fmaf(xv[q], dv[o], acc[q][o]);
        }
      } else if (is_bias) {
        for (int pix = sub; pix < kTH * kTW; pix += 16) {
          const float4* dp = reinterpret_cast<const float4*>(dys[pix]);
          const float4 d0 = dp[0], d1 = dp[1], d2 = dp[2], d3 = dp[3];
          acc[0][0] += d0.x; acc[0][1] += d0.y; acc[0][2] += d0.z; acc[0][3] += d0.w;
          acc[0][4] += d1.x; acc[0][5] += d1.y; acc[0][6] += d1.z; acc[0][7] += d1.w;
          acc[0][8] += d2.x; acc[0][9] += d2.y; acc[0][10] += d2.z; acc[0][11] += d2.w;
          acc[0][12] += d3.x; acc[0][13] += d3.y; acc[0][14] += d3.z; acc[0][15] += d3.w;
        }
      }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 50 * 16; i += 256) red[i / 16][i % 16] = 0.f;
    __syncthreads();
    // fixed-order reduction over the 16 pixel subsets: subset k adds in round k (deterministic, no float atomics)
    for (int k = 0; k < 16; ++k) {
      if (sub == k) {
        if (is_tap) {
#pragma unroll
          for (int q = 0; q < 4; ++q)
            if (s0 + q < 7) {
#pragma unroll
              for (int o = 0; o < 16; ++o) red[r * 7 + s0 + q][o] += acc[q][o];
            }
        } else if (is_bias) {
#pragma unroll
          for (int o = 0; o < 16; ++o) red[49][o] += acc[0][o];
        }
      }
      __syncthreads();
    }
    float* out = partials + ((size_t)blockIdx.x * ocg + g) * 50 * 16;
    for (int i = threadIdx.x; i < 50 * 16; i += 256) out[i] = red[i / 16][i % 16];
  }
}

__global__ void stem_wgrad_reduce_kernel(const float* partials, int nblocks, int oc, float* dw, float* db,
                                         int accumulate) {
  pdl_prologue();
  const int ocg = oc / 16;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;      // over ocg*50*16
  if (i >= ocg * 800) return;
  const int g = i / 800, sl = (i % 800) / 16, o = i % 16;
  double s = 0.0;
  for (int b = 0; b < nblocks; ++b) s += (double)partials[((size_t)b * ocg + g) * 800 + sl * 16 + o];
  const int ch = g * 16 + o;
  if (sl < 49) {
    if (dw) dw[(size_t)ch * 49 + sl] = (accumulate ? dw[(size_t)ch * 49 + sl] : 0.f) + (float)s;
  } else if (db) {
    db[ch] = (accumulate ? db[ch] : 0.f) + (float)s;
  }
}

}  // namespace ubr

using namespace ubr;

extern "C" {

int ubr_stem_conv_fwd(const float* x, int n, int h, int w, const float* wgt, const float* bias, int out_ch, ubr_view y,
                      void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(x && wgt && y.ptr, "stem_conv_fwd: null pointer");
  UBR_REQUIRE(out_ch > 0 && out_ch % 16 == 0, "stem_conv_fwd: out_ch=%d must be a multiple of 16", out_ch);
  UBR_REQUIRE(y.n == n && y.h == h && y.w == w && y.c == out_ch && y.ld >= out_ch && y.ld % 8 == 0,
              "stem_conv_fwd: output view mismatch");
  if (stem_tc_eligible(n, h, w, out_ch))
    return stem_tc_fwd(x, n, h, w, wgt, bias, (__nv_bfloat16*)y.ptr, y.ld, (cudaStream_t)stream);
  dim3 grid(ceil_div(w, kFTW), ceil_div(h, kFTH), n * (out_ch / 16));
  ::ubr::Launch(grid, 256, 0, (cudaStream_t)stream)(stem_fwd_kernel, x, n, h, w, wgt, bias, out_ch, (__nv_bfloat16*)y.ptr, y.ld);
  UBR_LAUNCH_CHECK("stem_fwd_kernel");
  return 0;
}

int ubr_stem_conv_fwd_relu(const float* x, int n, int h, int w, const float* wgt, const float* bias, int out_ch,
                           ubr_view y, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(x && wgt && y.ptr, "stem_conv_fwd_relu: null pointer");
  UBR_REQUIRE(y.n == n && y.h == h && y.w == w && y.c == out_ch && y.ld >= out_ch && y.ld % 8 == 0,
              "stem_conv_fwd_relu: output view mismatch");
  if (!stem_tc_eligible(n, h, w, out_ch)) return 1;          // caller runs conv + bn_act instead
  return stem_tc_fwd(x, n, h, w, wgt, bias, (__nv_bfloat16*)y.ptr, y.ld, (cudaStream_t)stream, 1);
}

uint64_t ubr_stem_wgrad_workspace_bytes(int n, int h, int w, int out_ch) {
  (void)n; (void)h; (void)w;
  const uint64_t cc = (uint64_t)kWgBlocks * (out_ch / 16) * 800 * sizeof(float);
  const uint64_t tc = (uint64_t)kStemTcMaxBlocks * 128 * 16 * sizeof(float);
  return cc > tc ? cc : tc;
}

int ubr_stem_conv_wgrad(const float* x, int n, int h, int w, ubr_view dy, float* dw, float* dbias, int accumulate,
                        void* workspace, uint64_t workspace_bytes, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(x && dy.ptr && workspace, "stem_conv_wgrad: null pointer");
  UBR_REQUIRE(dy.c % 16 == 0 && dy.n == n && dy.h == h && dy.w == w, "stem_conv_wgrad: dy view mismatch");
  UBR_REQUIRE(workspace_bytes >= ubr_stem_wgrad_workspace_bytes(n, h, w, dy.c), "stem_conv_wgrad: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  if (stem_tc_eligible(n, h, w, dy.c) && dy.ld % 8 == 0)
    return stem_tc_wgrad(x, n, h, w, (const __nv_bfloat16*)dy.ptr, dy.ld, dw, dbias, accumulate, (float*)workspace, st);
  ::ubr::Launch(kWgBlocks, 256, 0, st)(stem_wgrad_kernel, x, n, h, w, (const __nv_bfloat16*)dy.ptr, dy.ld, dy.c, (float*)workspace);
  UBR_LAUNCH_CHECK("stem_wgrad_kernel");
  const int total = (dy.c / 16) * 800;
  ::ubr::Launch(ceil_div(total, 128), 128, 0, st)(stem_wgrad_reduce_kernel, (const float*)workspace, kWgBlocks, dy.c, dw, dbias,
                                                                 accumulate);
  UBR_LAUNCH_CHECK("stem_wgrad_reduce_kernel");
  return 0;
}

}  // extern "C"
