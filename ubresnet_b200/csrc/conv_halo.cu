// Halo-resident implicit-GEMM convolution for the high-resolution / low-channel layers (sm_100a).
//
// The tap-by-tap kernel in conv_gemm.cu re-fetches the activation tile once per filter tap (9x / 49x the
// algorithmic bytes through L2) and pays a CTA launch + TMEM allocation per 128 pixels.  For stride-1 convs with
// Cin <= 64 (every 512^2 / 256^2 / 128^2 layer of the U-ResNet: models/ub_uresnet.py:41-64,
// models/common_layers.py:13-15,34) this kernel instead keeps a RING of input image rows resident in shared
// memory and addresses each tap as a SHIFTED WINDOW of those rows:
//
//   * a pixel row of Cin bf16 channels is exactly one TMA/UMMA swizzle span (32/64/128 B), so a row segment of
//     128+2*pad pixels lands in smem as a linear array of spans; the K-major A operand of tap (dh,dw) is the
//     128-span window starting at  slot(row+dh) + (dw+pad)*span  -- the hardware swizzle is a function of the
//     absolute smem address, so any span-aligned start is legal (profiles/r01_ubench_tma_mma_desc.txt, T3).
//   * TMA zero-fills out-of-image pixels => conv padding for free; every input byte is read from L2/HBM once per
//     row band (+ halo rows), every output byte written once.
//   * persistent CTAs (<= 2 per SM) walk (image, 128-column block, 32-row band) tiles; TMEM holds two
//     accumulators so the epilogue of row i overlaps the MMAs of row i+1; all filter taps stay resident in smem.
//
// Warp roles: warp 0 TMA producer (weights once, then one row segment per ring slot), warp 1 TMEM owner +
// single-thread tcgen05.mma issuer, warps 2-5 epilogue (tcgen05.ld -> +bias -> bf16 -> swizzled staging ->
// TMA store; or fp32 direct store for the logits).
#include "common.cuh"
#include "conv_internal.cuh"

namespace ubr {

constexpr int kHThreads = 192;
constexpr int kHMaxTaps = 49;

struct HaloParams {
  CUtensorMap a_map;            // input view (C, W, H, N), box {cin, 128+2p, 1, 1}
  CUtensorMap w_map;            // packed weights (K, N, slab), box {cin, nt, 1}
  CUtensorMap o_map;            // output view, box {o_box_c, 128, 1, 1}
  int ntaps, pad, ksteps, nt;
  int bands, cblocks, bh, ntiles;
  int nslots;
  uint32_t span, row_bytes, slot_bytes, w_tap_bytes, w_region, stage_bytes;
  uint32_t idesc;
  int out_mode, o_box_c;
  const float* bias;
  float* out_f32;
  int out_ld, out_cols, out_h, out_w;
  int8_t dh[kHMaxTaps], dw[kHMaxTaps];
  uint8_t slab[kHMaxTaps];
};

template <int N>
__device__ __forceinline__ void tma_store_wait_read_n() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}

__global__ void __launch_bounds__(kHThreads, 1) conv_halo_kernel(const __grid_constant__ HaloParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t w_base = smem_base;
  const uint32_t ring_base = w_base + p.w_region;
  const uint32_t stage_base = ring_base + (uint32_t)p.nslots * p.slot_bytes;
  const uint32_t bar_base = stage_base + 2u * p.stage_bytes;
  const uint32_t full0 = bar_base;
  const uint32_t empty0 = full0 + 8u * p.nslots;
  const uint32_t wbar = empty0 + 8u * p.nslots;
  const uint32_t accf0 = wbar + 8u;         // [2]
  const uint32_t acce0 = accf0 + 16u;       // [2]
  const uint32_t tmem_slot = acce0 + 16u;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int co0 = blockIdx.y * p.nt;
  const int S = p.nslots;
  const int rows_in = p.bh + 2 * p.pad;

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(full0 + 8u * s, 1);
      mbar_init(empty0 + 8u * s, 1);
    }
    mbar_init(wbar, 1);
    mbar_init(accf0, 1);
    mbar_init(accf0 + 8u, 1);
    mbar_init(acce0, 4);
    mbar_init(acce0 + 8u, 4);
    fence_mbar_init();
    prefetch_tmap(&p.a_map);
    prefetch_tmap(&p.w_map);
    if (p.out_mode == 0) prefetch_tmap(&p.o_map);
  }
  if (warp == 1) {
    tmem_alloc_dyn(tmem_slot, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp == 0) {
    if (lane == 0) {
      // ================= TMA producer =================
      mbar_arrive_expect_tx(wbar, (uint32_t)p.ntaps * p.w_tap_bytes);
      for (int t = 0; t < p.ntaps; ++t)
        tma_load_3d(w_base + (uint32_t)t * p.w_tap_bytes, &p.w_map, wbar, 0, co0, (int)p.slab[t]);
      uint32_t g = 0;
      for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        const int band = tile % p.bands;
        const int cb = (tile / p.bands) % p.cblocks;
        const int n = tile / (p.bands * p.cblocks);
        const int h0 = band * p.bh - p.pad, c0 = cb * 128 - p.pad;
        for (int j = 0; j < rows_in; ++j, ++g) {
          const uint32_t s = g % (uint32_t)S;
          const uint32_t ph = (g / (uint32_t)S) & 1u;
          mbar_wait(empty0 + 8u * s, ph ^ 1u);
          mbar_arrive_expect_tx(full0 + 8u * s, p.row_bytes);
          tma_load_4d(ring_base + s * p.slot_bytes, &p.a_map, full0 + 8u * s, 0, c0, h0 + j, n);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ================= MMA issuer =================
      const uint32_t layout = swizzle_layout_code((int)p.span);
      const uint32_t sbo = 8u * p.span;
      mbar_wait(wbar, 0);
      uint32_t g0 = 0, r = 0;
      for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        for (int i = 0; i < p.bh; ++i, ++r) {
          // rows g0+i .. g0+i+2p must have landed
          const int jfirst = (i == 0) ? 0 : 2 * p.pad;
          for (int j = jfirst; j <= 2 * p.pad; ++j) {
            const uint32_t gi = g0 + (uint32_t)(i + j);
            mbar_wait(full0 + 8u * (gi % (uint32_t)S), (gi / (uint32_t)S) & 1u);
          }
          const uint32_t b = r & 1u;
          mbar_wait(acce0 + 8u * b, ((r >> 1) & 1u) ^ 1u);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + b * 128u;
          for (int t = 0; t < p.ntaps; ++t) {
            const uint32_t gi = g0 + (uint32_t)(i + p.dh[t] + p.pad);
            const uint32_t a_src = ring_base + (gi % (uint32_t)S) * p.slot_bytes + (uint32_t)(p.dw[t] + p.pad) * p.span;
            const uint64_t adesc = make_smem_desc(a_src, 0, sbo, layout);
            const uint64_t bdesc = make_smem_desc(w_base + (uint32_t)t * p.w_tap_bytes, 0, sbo, layout);
            for (int k = 0; k < p.ksteps; ++k)
              tc_mma_bf16(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), p.idesc, (t > 0 || k > 0) ? 1u : 0u);
          }
          tc_commit(accf0 + 8u * b);
          {
            const uint32_t gi = g0 + (uint32_t)i;                       // oldest row: no later output row needs it
            tc_commit(empty0 + 8u * (gi % (uint32_t)S));
          }
          if (i == p.bh - 1)
            for (int j = 1; j <= 2 * p.pad; ++j) {
              const uint32_t gi = g0 + (uint32_t)(i + j);
              tc_commit(empty0 + 8u * (gi % (uint32_t)S));
            }
        }
        g0 += (uint32_t)rows_in;
      }
    }
  } else {
    // ================= epilogue (warps 2..5) =================
    const int quad = warp & 3;                 // TMEM lane quadrant this warp may access
    const int row = quad * 32 + lane;          // tile-local pixel (column inside the 128-wide block)
    const bool leader = (warp == 2 && lane == 0);
    const int bc = p.o_box_c;
    const uint32_t rb = (uint32_t)bc * 2u;
    const uint32_t xr = rb == 128 ? (uint32_t)(row & 7) : (rb == 64 ? (uint32_t)((row >> 1) & 3) : (uint32_t)((row >> 2) & 1));
    uint32_t r = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      const int band = tile % p.bands;
      const int cb = (tile / p.bands) % p.cblocks;
      const int n = tile / (p.bands * p.cblocks);
      const int c0 = cb * 128;
      for (int i = 0; i < p.bh; ++i, ++r) {
        const int h = band * p.bh + i;
        const uint32_t b = r & 1u;
        mbar_wait(accf0 + 8u * b, (r >> 1) & 1u);
        tc_fence_after();
        const uint32_t t_row = tmem_base + b * 128u + ((uint32_t)(quad * 32) << 16);
        if (p.out_mode == 0) {
          const uint32_t sbase = stage_base + b * p.stage_bytes;
          if (leader) tma_store_wait_read_n<1>();          // the store that last read this staging buffer is done
          asm volatile("bar.sync 1, 128;" ::: "memory");
          for (int c = 0; c < p.nt; c += 16) {
            uint32_t v[16];
            tmem_ld16(t_row + (uint32_t)c, v);
            tmem_ld_wait();
            float f[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) f[j] = __uint_as_float(v[j]);
            if (p.bias) {
#pragma unroll
              for (int j = 0; j < 16; ++j) f[j] += __ldg(p.bias + co0 + c + j);
            }
            uint32_t pk[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) pk[j] = pack_bf16x2(f[2 * j], f[2 * j + 1]);
            const int blk = c / bc;
            const uint32_t q = (uint32_t)((c % bc) >> 3);
            const uint32_t base = sbase + (uint32_t)blk * 128u * rb + (uint32_t)row * rb;
            const uint32_t a0 = base + ((q ^ xr) << 4);
            const uint32_t a1 = base + (((q + 1u) ^ xr) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a0), "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3]) : "memory");
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a1), "r"(pk[4]), "r"(pk[5]), "r"(pk[6]), "r"(pk[7]) : "memory");
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(acce0 + 8u * b);      // accumulator drained: MMA may overwrite it
          fence_proxy_async_smem();
          asm volatile("bar.sync 1, 128;" ::: "memory");
          if (leader) {
            const int nblk = p.nt / bc;
            for (int k = 0; k < nblk; ++k)
              tma_store_4d(&p.o_map, sbase + (uint32_t)k * 128u * rb, co0 + k * bc, c0, h, n);
            tma_store_commit();
          }
        } else {
          // fp32 direct store (network logits): few valid columns, pixel-major rows
          float* dst = p.out_f32 + ((size_t)((size_t)n * p.out_h + h) * p.out_w + (c0 + row)) * (size_t)p.out_ld;
          for (int c = 0; c < p.nt; c += 16) {
            uint32_t v[16];
            tmem_ld16(t_row + (uint32_t)c, v);
            tmem_ld_wait();
            if (p.out_ld == 4 && c == 0 && co0 == 0) {
              float4 o;
              o.x = __uint_as_float(v[0]) + ((p.bias && 0 < p.out_cols) ? __ldg(p.bias + 0) : 0.f);
              o.y = __uint_as_float(v[1]) + ((p.bias && 1 < p.out_cols) ? __ldg(p.bias + 1) : 0.f);
              o.z = __uint_as_float(v[2]) + ((p.bias && 2 < p.out_cols) ? __ldg(p.bias + 2) : 0.f);
              o.w = __uint_as_float(v[3]) + ((p.bias && 3 < p.out_cols) ? __ldg(p.bias + 3) : 0.f);
              if (p.out_cols < 4) o.w = 0.f;
              if (p.out_cols < 3) o.z = 0.f;
              if (p.out_cols < 2) o.y = 0.f;
              *reinterpret_cast<float4*>(dst) = o;
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const int col = co0 + c + j;
                if (col < p.out_cols) dst[col] = __uint_as_float(v[j]) + (p.bias ? __ldg(p.bias + col) : 0.f);
              }
            }
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(acce0 + 8u * b);
        }
      }
    }
    if (p.out_mode == 0 && leader) tma_store_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc_dyn(tmem_base, 256);
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static bool halo_disabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("UBR_NO_HALO"); v = (e && e[0] && e[0] != '0') ? 1 : 0; }
  return v == 1;
}

int conv_halo_try(const ubr_view& x, const void* wpk, int nslabs, int n_pad, const TapSpec* taps, int ntaps,
                  const float* bias, const ubr_view* y, float* y_f32, int y_f32_ld, int out_cols,
                  cudaStream_t stream) {
  if (halo_disabled()) return 1;
  const int cin = x.c;
  if (!(cin == 16 || cin == 32 || cin == 64)) return 1;
  if (x.w % 128 != 0 || x.w < 128) return 1;
  if (ntaps < 1 || ntaps > kHMaxTaps) return 1;
  int pad = 0;
  for (int i = 0; i < ntaps; ++i) {
    if (taps[i].amap != 0) return 1;
    pad = max(pad, max(abs(taps[i].dh), abs(taps[i].dw)));
  }
  if (pad > 3) return 1;
  int bh = 32;
  while (bh > 1 && x.h % bh != 0) bh >>= 1;
  if (bh < 8) return 1;
  const uint32_t span = (uint32_t)cin * 2u;
  const uint32_t row_bytes = (uint32_t)(128 + 2 * pad) * span;
  const uint32_t slot_bytes = (row_bytes + 1023u) & ~1023u;
  // N tile: largest divisor tile whose resident weights + ring + staging fit
  const uint32_t budget = 225u * 1024u - 1024u - 512u;
  int nt = 0, nslots = 0;
  uint32_t w_region = 0, stage_bytes = 0;
  for (int cand = n_pad > 128 ? 128 : n_pad; cand >= 16; cand -= 16) {
    if (n_pad % cand) continue;
    const uint32_t wr = ((uint32_t)ntaps * cand * span + 1023u) & ~1023u;
    const uint32_t st = y_f32 ? 0u : 128u * (uint32_t)cand * 2u;
    if (wr + 2u * st + (uint32_t)(2 * pad + 2) * slot_bytes > budget) continue;
    int s = (int)((budget - wr - 2u * st) / slot_bytes);
    if (s > 16) s = 16;
    nt = cand; nslots = s; w_region = wr; stage_bytes = st;
    break;
  }
  if (nt == 0) return 1;

  HaloParams p;
  memset(&p, 0, sizeof(p));
  {
    uint64_t dims[4] = {(uint64_t)x.c, (uint64_t)x.w, (uint64_t)x.h, (uint64_t)x.n};
    uint64_t str[4] = {2, (uint64_t)x.ld * 2, (uint64_t)x.w * x.ld * 2, (uint64_t)x.h * x.w * x.ld * 2};
    uint32_t box[4] = {(uint32_t)cin, (uint32_t)(128 + 2 * pad), 1, 1};
    if (int rc = make_tmap_bf16(&p.a_map, x.ptr, 4, dims, str, box, (int)span)) return rc;
  }
  {
    uint64_t dims[3] = {(uint64_t)cin, (uint64_t)n_pad, (uint64_t)nslabs};
    uint64_t str[3] = {2, (uint64_t)cin * 2, (uint64_t)cin * n_pad * 2};
    uint32_t box[3] = {(uint32_t)cin, (uint32_t)nt, 1};
    if (int rc = make_tmap_bf16(&p.w_map, wpk, 3, dims, str, box, (int)span)) return rc;
  }
  if (y_f32) {
    p.out_mode = 1; p.out_f32 = y_f32; p.out_ld = y_f32_ld; p.out_cols = out_cols;
    p.out_h = x.h; p.out_w = x.w;
  } else {
    p.out_mode = 0;
    p.o_box_c = nt >= 64 ? 64 : nt;
    uint64_t dims[4] = {(uint64_t)y->c, (uint64_t)y->w, (uint64_t)y->h, (uint64_t)y->n};
    uint64_t str[4] = {2, (uint64_t)y->ld * 2, (uint64_t)y->w * y->ld * 2, (uint64_t)y->h * y->w * y->ld * 2};
    uint32_t box[4] = {(uint32_t)p.o_box_c, 128, 1, 1};
    if (int rc = make_tmap_bf16(&p.o_map, y->ptr, 4, dims, str, box, p.o_box_c * 2)) return rc;
  }
  p.ntaps = ntaps; p.pad = pad; p.ksteps = cin / 16; p.nt = nt;
  p.bh = bh; p.bands = x.h / bh; p.cblocks = x.w / 128; p.ntiles = p.bands * p.cblocks * x.n;
  p.nslots = nslots;
  p.span = span; p.row_bytes = row_bytes; p.slot_bytes = slot_bytes;
  p.w_tap_bytes = (uint32_t)nt * span; p.w_region = w_region; p.stage_bytes = stage_bytes;
  p.idesc = make_idesc_bf16(128, nt, 0, 0);
  p.bias = bias;
  for (int i = 0; i < ntaps; ++i) {
    p.dh[i] = (int8_t)taps[i].dh; p.dw[i] = (int8_t)taps[i].dw; p.slab[i] = (uint8_t)taps[i].slab;
  }
  const uint32_t smem = w_region + (uint32_t)nslots * slot_bytes + 2u * stage_bytes + 16u * nslots + 64u + 1024u;
  static bool attr_set = false;
  if (!attr_set) {
    UBR_CUDA_OK(cudaFuncSetAttribute(conv_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set = true;
  }
  const int n_tiles = n_pad / nt;
  int per_sm = (smem <= 110u * 1024u) ? 2 : 1;
  int ctas = per_sm * sm_count() / n_tiles;
  if (const char* e = getenv("UBR_HALO_MAX_CTAS")) { const int v = atoi(e); if (v > 0 && ctas > v) ctas = v; }
  if (ctas < 1) ctas = 1;
  if (ctas > p.ntiles) ctas = p.ntiles;
  dim3 grid((unsigned)ctas, (unsigned)n_tiles, 1);
  conv_halo_kernel<<<grid, kHThreads, smem, stream>>>(p);
  UBR_LAUNCH_CHECK("conv_halo_kernel");
  return 0;
}

}  // namespace ubr
