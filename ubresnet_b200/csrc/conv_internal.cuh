// internal (non-ABI) declarations shared by the convolution translation units
#pragma once
#include "common.cuh"

namespace ubr {

struct TapSpec { int dh, dw, amap, slab; };

// Halo-resident kernel (conv_halo.cu).  Returns 0 = launched, 1 = shape not eligible (caller falls back to the
// tap-by-tap kernel), <0 = error.  `y` (bf16 view) or `y_f32` (fp32 [pixel][ld]) selects the output mode.
int conv_halo_try(const ubr_view& x, const void* wpk, int nslabs, int n_pad, const TapSpec* taps, int ntaps,
                  const float* bias, const ubr_view* y, float* y_f32, int y_f32_ld, int out_cols,
                  cudaStream_t stream, int accumulate = 0, float* stats = nullptr,
                  int* stats_rows = nullptr, int shuffle_cout = 0);

// Halo-resident weight gradient (wgrad_halo.cu).  Same return convention.
int wgrad_halo_try(const ubr_view& x, const ubr_view& dy, int out_ch, int ksize, float* dw, int accumulate,
                   void* workspace, uint64_t workspace_bytes, cudaStream_t stream);
uint64_t wgrad_halo_workspace_bytes(int in_ch, int out_ch_pad, int ksize);

}  // namespace ubr
