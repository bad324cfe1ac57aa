// Device-side batch assembly: what the reference's C++ batch fillers do on host threads, as kernels.
//
//   image : larcv Image2D pixels are COLUMN-major (index = col*rows + row, ImageMeta::index); the filler writes the
//           row-major (caffe/NCHW) order and optionally mirrors the columns
//           (larcv/larcv/app/ThreadIO/BatchFillerImage2D.cxx:102-112,217-256; caffe/larcv/app/APICaffe/SegFiller.cxx:254-303)
//   label : ROI-type pixel -> class through the ClassTypeList/ClassTypeDef table (SegFiller.cxx:49-63,339-363), then the
//           training scripts shift by -1 (grid_scripts/train_ubresnet_wlarcv1_tuftsgrid.py:132); int64 out (nll target)
//   weight: SegWeightTrackShower (caffe/larcv/app/ImageMod/SegWeightTrackShower.cxx:109-272): per image
//           w_track = (int)((float)N / (float)n_track), w_shower likewise, background pixels within DistSurrounding of a
//           labelled pixel get (int)((float)N / (float)n_surrounding); combined with the incoming weight image (1.0
//           when none) by sum / max / average / overwrite pooling and clipped at WeightMax.
// All integer / index work is exact; the float divisions use IEEE round-to-nearest like the host code (this file is
// compiled with --use_fast_math, so they are spelled __fdiv_rn).
#include "common.cuh"
#include <cuda_fp16.h>

namespace ubr {

template <typename T> __device__ __forceinline__ float wire_to_float(T v);
template <> __device__ __forceinline__ float wire_to_float<float>(float v) { return v; }
template <> __device__ __forceinline__ float wire_to_float<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float wire_to_float<uint8_t>(uint8_t v) { return (float)v; }

// out[b][row][col] = in[b][(mirror[b] ? cols-1-col : col) * rows + row]          (32x32 tiles through shared memory)
template <typename T>
__global__ void __launch_bounds__(256) fill_image_kernel(const T* __restrict__ in, int rows, int cols,
                                                         const uint8_t* __restrict__ mirror, float* __restrict__ out) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const bool mir = mirror && mirror[b];
  const int r0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const T* src = in + (size_t)b * rows * cols;
  float* dst = out + (size_t)b * rows * cols;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int j = ty; j < 32; j += 8) {                 // j: column inside the tile, tx: row (fast index of the source)
    const int row = r0 + tx, col = c0 + j;
    if (row < rows && col < cols) {
      const int scol = mir ? cols - 1 - col : col;
      tile[j][tx] = wire_to_float<T>(src[(size_t)scol * rows + row]);
    }
  }
  __syncthreads();
  for (int j = ty; j < 32; j += 8) {                 // j: row inside the tile, tx: column (fast index of the output)
    const int row = r0 + j, col = c0 + tx;
    if (row < rows && col < cols) dst[(size_t)row * cols + col] = tile[tx][j];
  }
}

// label[b][row][col] = lut[(size_t)src[...]] + shift ; ROI types outside the table / mapped to 0xFF raise the error flag
template <typename T>
__global__ void __launch_bounds__(256) fill_label_kernel(const T* __restrict__ in, int rows, int cols,
                                                         const uint8_t* __restrict__ mirror, const uint8_t* __restrict__ lut,
                                                         int lut_size, int shift, long long* __restrict__ out,
                                                         uint8_t* __restrict__ out_cls, int* __restrict__ bad) {
  __shared__ long long tile[32][33];
  const int b = blockIdx.z;
  const bool mir = mirror && mirror[b];
  const int r0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const T* src = in + (size_t)b * rows * cols;
  long long* dst = out + (size_t)b * rows * cols;
  uint8_t* dst8 = out_cls ? out_cls + (size_t)b * rows * cols : nullptr;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int j = ty; j < 32; j += 8) {
    const int row = r0 + tx, col = c0 + j;
    if (row < rows && col < cols) {
      const int scol = mir ? cols - 1 - col : col;
      const float v = wire_to_float<T>(src[(size_t)scol * rows + row]);
      const long long t = (long long)v;                           // (size_t)label_value of SegFiller.cxx:351
      uint8_t cls = 0xFF;
      if (t >= 0 && t < lut_size) cls = lut[t];
      if (cls == 0xFF) { atomicAdd(bad, 1); cls = 0; }
      tile[j][tx] = (long long)cls;
    }
  }
  __syncthreads();
  for (int j = ty; j < 32; j += 8) {
    const int row = r0 + j, col = c0 + tx;
    if (row < rows && col < cols) {
      const long long cls = tile[tx][j];
      dst[(size_t)row * cols + col] = cls + shift;
      if (dst8) dst8[(size_t)row * cols + col] = (uint8_t)cls;    // pre-shift class, 1 byte: input of the weight maker
    }
  }
}

// ---- SegWeightTrackShower ------------------------------------------------------------------------------------------
// pass 1: category per pixel (0 none, 1 shower, 2 track, 3 background within `dist` of a labelled pixel) and the three
// per-image pixel counts.  `lab` holds small integer codes (ROI types or classes); cat_lut maps code -> {0,1,2}.
template <typename T>
__global__ void __launch_bounds__(256) seg_weight_count_kernel(const T* __restrict__ lab, int rows, int cols, long row_stride,
                                                               long col_stride, const uint8_t* __restrict__ cat_lut,
                                                               int lut_size, int dist, uint8_t* __restrict__ cat,
                                                               unsigned int* __restrict__ counts) {
  const int b = blockIdx.y;
  const size_t npx = (size_t)rows * cols;
  const T* src = lab + (size_t)b * npx;
  uint8_t* dst = cat + (size_t)b * npx;
  unsigned int ns = 0, nt = 0, nb = 0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < npx; i += (size_t)gridDim.x * blockDim.x) {
    const int row = (int)(i / cols), col = (int)(i % cols);
    auto code = [&](int r, int c) -> int {
      const long long t = (long long)src[(size_t)r * row_stride + (size_t)c * col_stride];
      return (t >= 0 && t < lut_size) ? cat_lut[t] : 0;
    };
    int k = code(row, col);
    if (k == 0 && dist > 0) {
      bool flag = false;
      for (int r = row - dist; r <= row + dist && !flag; ++r) {
        if (r < 0 || r >= rows) continue;
        for (int c = col - dist; c <= col + dist; ++c) {
          if (c < 0 || c >= cols) continue;
          if (code(r, c) != 0) { flag = true; break; }
        }
      }
      if (flag) k = 3;
    }
    dst[i] = (uint8_t)k;
    ns += (k == 1); nt += (k == 2); nb += (k == 3);
  }
  // warp reduce, then integer atomics (exact)
  for (int o = 16; o > 0; o >>= 1) {
    ns += __shfl_down_sync(0xffffffffu, ns, o);
    nt += __shfl_down_sync(0xffffffffu, nt, o);
    nb += __shfl_down_sync(0xffffffffu, nb, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (ns) atomicAdd(counts + 3 * b + 0, ns);
    if (nt) atomicAdd(counts + 3 * b + 1, nt);
    if (nb) atomicAdd(counts + 3 * b + 2, nb);
  }
}

// fast path of pass 1 for row-major uint8 labels, cols % 4 == 0, dist <= 4: a thread owns 4 adjacent pixels and reads
// the (2*dist+1) x 12-pixel neighbourhood as 32-bit words (3.75 word loads per pixel instead of 25 byte loads)
__global__ void __launch_bounds__(256) seg_weight_count_u8x4_kernel(const uint8_t* __restrict__ lab, int rows, int cols,
                                                                    const uint8_t* __restrict__ cat_lut, int lut_size,
                                                                    int dist, uint8_t* __restrict__ cat,
                                                                    unsigned int* __restrict__ counts) {
  __shared__ uint8_t lut[256];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) lut[i] = i < lut_size ? cat_lut[i] : 0;
  __syncthreads();
  const int b = blockIdx.y;
  const size_t npx = (size_t)rows * cols;
  const uint32_t* src = reinterpret_cast<const uint32_t*>(lab + (size_t)b * npx);
  uint32_t* dst = reinterpret_cast<uint32_t*>(cat + (size_t)b * npx);
  const int wpr = cols >> 2;                                   // words per row
  const size_t nwords = npx >> 2;
  unsigned int ns = 0, nt = 0, nb = 0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nwords; i += (size_t)gridDim.x * blockDim.x) {
    const int row = (int)(i / wpr), wc = (int)(i % wpr);
    const uint32_t own = src[i];
    uint32_t k4 = 0, any = 0;                                   // categories of the 4 pixels / "a neighbour is labelled"
#pragma unroll
    for (int k = 0; k < 4; ++k) k4 |= (uint32_t)lut[(own >> (8 * k)) & 0xFFu] << (8 * k);
    bool need = false;
#pragma unroll
    for (int k = 0; k < 4; ++k) need |= ((k4 >> (8 * k)) & 0xFFu) == 0;
    if (need && dist > 0) {
      for (int r = row - dist; r <= row + dist; ++r) {
        if (r < 0 || r >= rows) continue;
        const uint32_t* rp = src + (size_t)r * wpr;
        const uint32_t wl = wc > 0 ? rp[wc - 1] : 0u, wm = rp[wc], wr = wc + 1 < wpr ? rp[wc + 1] : 0u;
        uint32_t bits = 0;                                      // bit j: pixel (4*wc - 4 + j) of this row is labelled
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          bits |= (uint32_t)(lut[(wl >> (8 * j)) & 0xFFu] != 0) << j;
          bits |= (uint32_t)(lut[(wm >> (8 * j)) & 0xFFu] != 0) << (4 + j);
          bits |= (uint32_t)(lut[(wr >> (8 * j)) & 0xFFu] != 0) << (8 + j);
        }
        const uint32_t win = (1u << (2 * dist + 1)) - 1u;
#pragma unroll
        for (int k = 0; k < 4; ++k) any |= (uint32_t)(((bits >> (4 + k - dist)) & win) != 0) << k;
      }
    }
    uint32_t outw = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      uint32_t c = (k4 >> (8 * k)) & 0xFFu;
      if (c == 0 && ((any >> k) & 1u)) c = 3;
      outw |= c << (8 * k);
      ns += (c == 1); nt += (c == 2); nb += (c == 3);
    }
    dst[i] = outw;
  }
  for (int o = 16; o > 0; o >>= 1) {
    ns += __shfl_down_sync(0xffffffffu, ns, o);
    nt += __shfl_down_sync(0xffffffffu, nt, o);
    nb += __shfl_down_sync(0xffffffffu, nb, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (ns) atomicAdd(counts + 3 * b + 0, ns);
    if (nt) atomicAdd(counts + 3 * b + 1, nt);
    if (nb) atomicAdd(counts + 3 * b + 2, nb);
  }
}

// pass 2: weights from the categories and counts (SegWeightTrackShower.cxx:119-125,180-187,244-272)
__global__ void __launch_bounds__(256) seg_weight_apply_kernel(const uint8_t* __restrict__ cat,
                                                               const unsigned int* __restrict__ counts, long npx,
                                                               const float* __restrict__ win, int pool, float weight_max,
                                                               int weight_surrounding, float* __restrict__ wout) {
  const int b = blockIdx.y;
  const unsigned int ns = counts[3 * b], nt = counts[3 * b + 1], nb = counts[3 * b + 2];
  const float n = (float)npx;
  const float w_shower = ns ? (float)(int)__fdiv_rn(n, (float)ns) : 1.f;
  const float w_track = nt ? (float)(int)__fdiv_rn(n, (float)nt) : 1.f;
  const float w_surr = (weight_surrounding && nb) ? (float)(int)__fdiv_rn(n, (float)nb) : 0.f;
  const uint8_t* c = cat + (size_t)b * npx;
  const float* wi = win ? win + (size_t)b * npx : nullptr;
  float* wo = wout + (size_t)b * npx;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < npx; i += (long)gridDim.x * blockDim.x) {
    const int k = c[i];
    const float t = k == 1 ? w_shower : k == 2 ? w_track : k == 3 ? w_surr : 0.f;
    const float w0 = wi ? wi[i] : 1.f;
    float v;
    switch (pool) {
      case 0: v = w0 + t; break;                      // kSumPool
      case 1: v = fmaxf(w0, t); break;                // kMaxPool
      case 2: v = __fdiv_rn(w0 + t, 2.f); break;      // kAveragePool
      default: v = t; break;                          // kOverwrite
    }
    if (v > weight_max) v = weight_max;
    wo[i] = v;
  }
}

}  // namespace ubr

using namespace ubr;

extern "C" {

int ubr_fill_image_batch(const void* src, int src_dtype, int batch, int rows, int cols, const uint8_t* mirror, float* out,
                         void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(src && out && batch > 0 && rows > 0 && cols > 0, "fill_image_batch: bad arguments");
  dim3 grid((unsigned)ceil_div(rows, 32), (unsigned)ceil_div(cols, 32), (unsigned)batch);
  cudaStream_t st = (cudaStream_t)stream;
  if (src_dtype == 0) fill_image_kernel<float><<<grid, 256, 0, st>>>((const float*)src, rows, cols, mirror, out);
  else if (src_dtype == 1) fill_image_kernel<__half><<<grid, 256, 0, st>>>((const __half*)src, rows, cols, mirror, out);
  else if (src_dtype == 2) fill_image_kernel<uint8_t><<<grid, 256, 0, st>>>((const uint8_t*)src, rows, cols, mirror, out);
  else { set_error("fill_image_batch: src_dtype %d (0 f32, 1 f16, 2 u8)", src_dtype); return UBR_E_BADARG; }
  UBR_LAUNCH_CHECK("fill_image_kernel");
  return 0;
}

int ubr_fill_label_batch(const void* src, int src_dtype, int batch, int rows, int cols, const uint8_t* mirror,
                         const uint8_t* lut, int lut_size, int shift, int64_t* out, uint8_t* out_class_u8, int* invalid_count,
                         void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(src && out && lut && invalid_count && batch > 0 && rows > 0 && cols > 0 && lut_size > 0,
              "fill_label_batch: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  UBR_CUDA_OK(cudaMemsetAsync(invalid_count, 0, sizeof(int), st));
  dim3 grid((unsigned)ceil_div(rows, 32), (unsigned)ceil_div(cols, 32), (unsigned)batch);
  if (src_dtype == 0)
    fill_label_kernel<float><<<grid, 256, 0, st>>>((const float*)src, rows, cols, mirror, lut, lut_size, shift,
                                                   (long long*)out, out_class_u8, invalid_count);
  else if (src_dtype == 2)
    fill_label_kernel<uint8_t><<<grid, 256, 0, st>>>((const uint8_t*)src, rows, cols, mirror, lut, lut_size, shift,
                                                     (long long*)out, out_class_u8, invalid_count);
  else { set_error("fill_label_batch: src_dtype %d (0 f32, 2 u8)", src_dtype); return UBR_E_BADARG; }
  UBR_LAUNCH_CHECK("fill_label_kernel");
  return 0;
}

uint64_t ubr_seg_weights_workspace_bytes(int batch, int rows, int cols) {
  return (uint64_t)batch * rows * cols + (uint64_t)batch * 3 * sizeof(unsigned int) + 64;
}

int ubr_seg_weights(const void* label, int label_dtype, int column_major, int batch, int rows, int cols,
                    const uint8_t* cat_lut, int lut_size, int dist_surrounding, int weight_surrounding, int pool_type,
                    float weight_max, const float* weight_in, float* weight_out, void* workspace, uint64_t workspace_bytes,
                    void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(label && cat_lut && weight_out && workspace && batch > 0 && rows > 0 && cols > 0 && lut_size > 0,
              "seg_weights: bad arguments");
  UBR_REQUIRE(workspace_bytes >= ubr_seg_weights_workspace_bytes(batch, rows, cols), "seg_weights: workspace too small");
  UBR_REQUIRE(pool_type >= 0 && pool_type <= 3 && dist_surrounding >= 0 && dist_surrounding <= 8, "seg_weights: bad options");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t npx = (size_t)rows * cols;
  // counts first (16-byte aligned), categories behind them
  unsigned int* counts = (unsigned int*)workspace;
  uint8_t* cat = (uint8_t*)workspace + (((size_t)batch * 3 * sizeof(unsigned int) + 63) / 64) * 64;
  UBR_CUDA_OK(cudaMemsetAsync(counts, 0, (size_t)batch * 3 * sizeof(unsigned int), st));
  // `column_major` describes the LABEL only (larcv Image2D order); the category scratch, weight_in and weight_out are
  // row-major [batch][rows][cols]
  const long rs = column_major ? 1 : cols, cs = column_major ? rows : 1;
  unsigned gx = (unsigned)((npx + 255) / 256);
  if (gx > 148u * 8u) gx = 148u * 8u;
  dim3 grid(gx, (unsigned)batch, 1);
  if (label_dtype == 0)
    seg_weight_count_kernel<float><<<grid, 256, 0, st>>>((const float*)label, rows, cols, rs, cs, cat_lut, lut_size,
                                                         weight_surrounding ? dist_surrounding : 0, cat, counts);
  else if (label_dtype == 2 && !column_major && cols % 4 == 0 && dist_surrounding <= 4 &&
           (((uintptr_t)label | (uintptr_t)cat) & 3) == 0)
    seg_weight_count_u8x4_kernel<<<grid, 256, 0, st>>>((const uint8_t*)label, rows, cols, cat_lut, lut_size,
                                                       weight_surrounding ? dist_surrounding : 0, cat, counts);
  else if (label_dtype == 2)
    seg_weight_count_kernel<uint8_t><<<grid, 256, 0, st>>>((const uint8_t*)label, rows, cols, rs, cs, cat_lut, lut_size,
                                                           weight_surrounding ? dist_surrounding : 0, cat, counts);
  else if (label_dtype == 3)
    seg_weight_count_kernel<long long><<<grid, 256, 0, st>>>((const long long*)label, rows, cols, rs, cs, cat_lut, lut_size,
                                                             weight_surrounding ? dist_surrounding : 0, cat, counts);
  else { set_error("seg_weights: label_dtype %d (0 f32, 2 u8, 3 i64)", label_dtype); return UBR_E_BADARG; }
  UBR_LAUNCH_CHECK("seg_weight_count_kernel");
  seg_weight_apply_kernel<<<grid, 256, 0, st>>>(cat, counts, (long)npx, weight_in, pool_type, weight_max, weight_surrounding,
                                                weight_out);
  UBR_LAUNCH_CHECK("seg_weight_apply_kernel");
  return 0;
}

}  // extern "C"
