// Stem convolution: nn.Conv2d(1, OC, 7, stride 1, pad 3, bias) on the fp32 single-channel crop
// (reference models/ub_uresnet.py:41,94).  Cin = 1 means K = 49: not a tensor-core GEMM.  Direct
// CUDA-core kernels; forward is bounded by the bf16 NHWC write (32 B/pixel for OC = 16).
#include "common.cuh"

namespace ubr {

constexpr int kTW = 32, kTH = 8;            // output tile per block (256 threads, one pixel each)
constexpr int kHW = kTW + 6, kHH = kTH + 6; // halo tile

// grid: (tiles_w, tiles_h, n * oc_groups)   each block: 16 output channels
__global__ void __launch_bounds__(256) stem_fwd_kernel(const float* __restrict__ x, int n, int h, int w,
                                                       const float* __restrict__ wgt, const float* __restrict__ bias,
                                                       int oc, __nv_bfloat16* __restrict__ y, int yld) {
  __shared__ float tile[kHH][kHW + 1];
  __shared__ __align__(16) float wsm[49][16];
  __shared__ float bsm[16];
  const int ocg = oc / 16;
  const int b = blockIdx.z / ocg, g = blockIdx.z % ocg;
  const int w0 = blockIdx.x * kTW, h0 = blockIdx.y * kTH;
  const float* img = x + (size_t)b * h * w;
  for (int i = threadIdx.x; i < kHH * kHW; i += 256) {
    const int r = i / kHW, c = i % kHW;
    const int ih = h0 + r - 3, iw = w0 + c - 3;
    tile[r][c] = (ih >= 0 && ih < h && iw >= 0 && iw < w) ? img[(size_t)ih * w + iw] : 0.f;
  }
  for (int i = threadIdx.x; i < 49 * 16; i += 256) {
    const int t = i / 16, o = i % 16;
    wsm[t][o] = wgt[(size_t)(g * 16 + o) * 49 + t];
  }
  if (threadIdx.x < 16) bsm[threadIdx.x] = bias ? bias[g * 16 + threadIdx.x] : 0.f;
  __syncthreads();
  const int tx = threadIdx.x % kTW, ty = threadIdx.x / kTW;
  float acc[16];
#pragma unroll
  for (int o = 0; o < 16; ++o) acc[o] = bsm[o];
#pragma unroll
  for (int r = 0; r < 7; ++r)
#pragma unroll
    for (int s = 0; s < 7; ++s) {
      const float v = tile[ty + r][tx + s];
      const float4* wp = reinterpret_cast<const float4*>(wsm[r * 7 + s]);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 wv = wp[q];
        acc[4 * q + 0] = fmaf(v, wv.x, acc[4 * q + 0]);
        acc[4 * q + 1] = fmaf(v, wv.y, acc[4 * q + 1]);
        acc[4 * q + 2] = fmaf(v, wv.z, acc[4 * q + 2]);
        acc[4 * q + 3] = fmaf(v, wv.w, acc[4 * q + 3]);
      }
    }
  const int oh = h0 + ty, ow = w0 + tx;
  if (oh < h && ow < w) {
    __nv_bfloat16* dst = y + (((size_t)b * h + oh) * w + ow) * yld + g * 16;
    uint4 u0, u1;
    u0.x = pack_bf16x2(acc[0], acc[1]); u0.y = pack_bf16x2(acc[2], acc[3]);
    u0.z = pack_bf16x2(acc[4], acc[5]); u0.w = pack_bf16x2(acc[6], acc[7]);
    u1.x = pack_bf16x2(acc[8], acc[9]); u1.y = pack_bf16x2(acc[10], acc[11]);
    u1.z = pack_bf16x2(acc[12], acc[13]); u1.w = pack_bf16x2(acc[14], acc[15]);
    reinterpret_cast<uint4*>(dst)[0] = u0;
    reinterpret_cast<uint4*>(dst)[1] = u1;
  }
}

// weight gradient: dW[o][t] = sum_p dy[p][o] * x[p + t - 3],  db[o] = sum_p dy[p][o]
// persistent blocks loop over 32x8 tiles.  Thread layout: 4 pixel-subsets x 64 tap slots; a thread owns one
// tap (49 real + 1 bias slot) and all 16 channels (dy[p][0:16] is a warp-broadcast smem read).
constexpr int kWgBlocks = 296;
__global__ void __launch_bounds__(256) stem_wgrad_kernel(const float* __restrict__ x, int n, int h, int w,
                                                         const __nv_bfloat16* __restrict__ dy, int dyld, int oc,
                                                         float* __restrict__ partials /* [blocks][ocg][50][16] */) {
  __shared__ float tile[kHH][kHW + 1];
  __shared__ __align__(16) float dys[kTH * kTW][16];
  __shared__ float red[4][50][16];
  const int ocg = oc / 16;
  const int tiles_w = (w + kTW - 1) / kTW, tiles_h = (h + kTH - 1) / kTH;
  const long ntiles = (long)tiles_w * tiles_h * n;
  const int slot = threadIdx.x & 63, sub = threadIdx.x >> 6;
  const bool active = slot < 50;
  const int r = slot / 7, s = slot % 7;     // slot 49 = bias
  for (int g = 0; g < ocg; ++g) {
    float acc[16];
#pragma unroll
    for (int o = 0; o < 16; ++o) acc[o] = 0.f;
    for (long t = blockIdx.x; t < ntiles; t += gridDim.x) {
      const int tw = (int)(t % tiles_w), th = (int)((t / tiles_w) % tiles_h), b = (int)(t / ((long)tiles_w * tiles_h));
      const int w0 = tw * kTW, h0 = th * kTH;
      const float* img = x + (size_t)b * h * w;
      __syncthreads();
      for (int i = threadIdx.x; i < kHH * kHW; i += 256) {
        const int rr = i / kHW, cc = i % kHW;
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
      if (active) {
        // this subset handles pixels sub, sub+4, ... of the 256-pixel tile
        for (int pix = sub; pix < kTH * kTW; pix += 4) {
          const int px = pix % kTW, py = pix / kTW;
          const float xv = slot < 49 ? tile[py + r][px + s] : 1.f;
          const float4* dp = reinterpret_cast<const float4*>(dys[pix]);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float4 d = dp[q];
            acc[4 * q + 0] = fmaf(xv, d.x, acc[4 * q + 0]);
            acc[4 * q + 1] = fmaf(xv, d.y, acc[4 * q + 1]);
            acc[4 * q + 2] = fmaf(xv, d.z, acc[4 * q + 2]);
            acc[4 * q + 3] = fmaf(xv, d.w, acc[4 * q + 3]);
          }
        }
      }
    }
    __syncthreads();
    if (active) {
#pragma unroll
      for (int o = 0; o < 16; ++o) red[sub][slot][o] = acc[o];
    }
    __syncthreads();
    float* out = partials + ((size_t)blockIdx.x * ocg + g) * 50 * 16;
    for (int i = threadIdx.x; i < 50 * 16; i += 256) {
      const int sl = i / 16, o = i % 16;
      out[i] = red[0][sl][o] + red[1][sl][o] + red[2][sl][o] + red[3][sl][o];
    }
  }
}

__global__ void stem_wgrad_reduce_kernel(const float* partials, int nblocks, int oc, float* dw, float* db,
                                         int accumulate) {
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
  dim3 grid(ceil_div(w, kTW), ceil_div(h, kTH), n * (out_ch / 16));
  stem_fwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, n, h, w, wgt, bias, out_ch, (__nv_bfloat16*)y.ptr, y.ld);
  UBR_LAUNCH_CHECK("stem_fwd_kernel");
  return 0;
}

uint64_t ubr_stem_wgrad_workspace_bytes(int n, int h, int w, int out_ch) {
  (void)n; (void)h; (void)w;
  return (uint64_t)kWgBlocks * (out_ch / 16) * 800 * sizeof(float);
}

int ubr_stem_conv_wgrad(const float* x, int n, int h, int w, ubr_view dy, float* dw, float* dbias, int accumulate,
                        void* workspace, uint64_t workspace_bytes, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(x && dy.ptr && workspace, "stem_conv_wgrad: null pointer");
  UBR_REQUIRE(dy.c % 16 == 0 && dy.n == n && dy.h == h && dy.w == w, "stem_conv_wgrad: dy view mismatch");
  UBR_REQUIRE(workspace_bytes >= ubr_stem_wgrad_workspace_bytes(n, h, w, dy.c), "stem_conv_wgrad: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  stem_wgrad_kernel<<<kWgBlocks, 256, 0, st>>>(x, n, h, w, (const __nv_bfloat16*)dy.ptr, dy.ld, dy.c, (float*)workspace);
  UBR_LAUNCH_CHECK("stem_wgrad_kernel");
  const int total = (dy.c / 16) * 800;
  stem_wgrad_reduce_kernel<<<ceil_div(total, 128), 128, 0, st>>>((const float*)workspace, kWgBlocks, dy.c, dw, dbias,
                                                                 accumulate);
  UBR_LAUNCH_CHECK("stem_wgrad_reduce_kernel");
  return 0;
}

}  // extern "C"
