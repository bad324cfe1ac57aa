// internal (non-ABI) declarations shared by the convolution translation units
#pragma once
#include "common.cuh"
#include "tree_reduce.cuh"

namespace ubr {

struct TapSpec { int dh, dw, amap, slab; };

// Fused inference epilogue (eval-mode residual block with BatchNorm folded into weights and bias):
//   v = acc + bias;  if (relu_pre) v = max(v, 0);  if (res) v += res[pixel][channel];  if (relu_post) v = max(v, 0)
struct FusedEpi {
  int relu_pre, relu_post;
  const void* res;              // bf16 NHWC on the output grid (same n, h, w), or null
  int res_ld;
};

#ifdef __CUDACC__
// Warp reduce-scatter of 16 per-lane values over the 32 lanes: 16 shuffles instead of 80.  On return lane l holds
// in the result the warp total of channel (l >> 1) & 15 (both lanes of a pair hold the same total).
__device__ __forceinline__ float warp_reduce_scatter16(float (&v)[16], int lane) {
#pragma unroll
  for (int h = 8, m = 16; h >= 1; h >>= 1, m >>= 1) {
    const bool up = (lane & m) != 0;
#pragma unroll
    for (int i = 0; i < h; ++i) {
      const float send = up ? v[i] : v[i + h];
      const float keep = up ? v[i + h] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, m);
    }
  }
  return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 1);
}
// FusedEpi on 16 consecutive channels of one pixel; res_px = &res[pixel][first channel] (16-byte aligned) or null
__device__ __forceinline__ void fused_epi16(float (&f)[16], int relu_pre, const __nv_bfloat16* res_px, int relu_post) {
  if (relu_pre) {
#pragma unroll
    for (int j = 0; j < 16; ++j) f[j] = fmaxf(f[j], 0.f);
  }
  if (res_px) {
    const uint4 r0 = __ldg(reinterpret_cast<const uint4*>(res_px));
    const uint4 r1 = __ldg(reinterpret_cast<const uint4*>(res_px + 8));
    const uint32_t rw[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float2 t = unpack_bf16x2(rw[j]);
      f[2 * j] += t.x; f[2 * j + 1] += t.y;
    }
  }
  if (relu_post) {
#pragma unroll
    for (int j = 0; j < 16; ++j) f[j] = fmaxf(f[j], 0.f);
  }
}
#endif

// fused BatchNorm statistics: a conv launch may emit at most this many partial rows [2][c] into the bn workspace
constexpr int kMaxStatsRows = 1184;

// Tensor-core stem (stem_tc.cu): 1 -> 16 channels, 7x7, for w % 128 == 0 and h % 8 == 0; stem.cu keeps the other shapes.
constexpr int kStemTcMaxBlocks = 592;          // persistent CTAs (4 per SM); the weight gradient keeps one partial each
bool stem_tc_eligible(int n, int h, int w, int out_ch);
int stem_tc_fwd(const float* x, int n, int h, int w, const float* wgt, const float* bias, __nv_bfloat16* y, int yld,
                cudaStream_t stream, int relu = 0);
int stem_tc_wgrad(const float* x, int n, int h, int w, const __nv_bfloat16* dy, int dyld, float* dw, float* dbias,
                  int accumulate, float* workspace, cudaStream_t stream);

// Fused BatchNorm(+ReLU)-backward REDUCTION in a data-gradient epilogue: the conv output `da` feeds the backward of
// a = relu?(bn(y)); while the epilogue still holds the (bf16-rounded) da tile it reads the matching y pixels and
// accumulates  sum dz  and  sum dz*xhat  (dz = da * [bn(y) > 0], xhat = (y - mean) * invstd) per channel, replacing the
// stand-alone bn_act_bwd_reduce pass (one full read of da and y).  Partial rows [cta][2][c] land in `stats`.
struct BnReduce {
  const void* y;                // bf16 NHWC pre-BatchNorm conv output on the da grid
  int y_ld, relu;
  const float *scale, *shift, *mean, *invstd;
};

// Finalize folded into the producing conv launch (tree_reduce.cuh): the partial rows the epilogue emits are summed by the
// last CTAs to arrive and turned into the BatchNorm affine (kind 1, forward statistics) or the backward sums (kind 2).
struct FusedFinal {
  void* tree;                   // ubr_bn_tree_bytes() of zero-initialised device memory
  int kind;                     // 1: BnFinal, 2: SumsFinal
  BnFinal bn;
  SumsFinal sums;
};

// Halo-resident kernel (conv_halo.cu).  Returns 0 = launched, 1 = shape not eligible (caller falls back to the
// tap-by-tap kernel), <0 = error.  `y` (bf16 view) or `y_f32` (fp32 [pixel][ld]) selects the output mode.
int conv_halo_try(const ubr_view& x, const void* wpk, int nslabs, int n_pad, const TapSpec* taps, int ntaps,
                  const float* bias, const ubr_view* y, float* y_f32, int y_f32_ld, int out_cols,
                  cudaStream_t stream, int accumulate = 0, float* stats = nullptr,
                  int* stats_rows = nullptr, int shuffle_cout = 0, const FusedEpi* epi = nullptr,
                  const BnReduce* bnr = nullptr, const FusedFinal* fin = nullptr);

// Halo-resident weight gradient (wgrad_halo.cu).  Same return convention.
int wgrad_halo_try(const ubr_view& x, const ubr_view& dy, int out_ch, int ksize, float* dw, int accumulate,
                   void* workspace, uint64_t workspace_bytes, cudaStream_t stream);
uint64_t wgrad_halo_workspace_bytes(int in_ch, int out_ch_pad, int ksize);

}  // namespace ubr
