// Halo-resident weight-gradient GEMM for the high-resolution / low-channel stride-1 convolutions (sm_100a).
//
//   dW[co][ci][r][s] = sum_{n,ho,wo} dy[n,ho,wo,co] * x[n, ho+r-p, wo+s-p, ci]          (autograd of nn.Conv2d,
//   models/common_layers.py:13-15,34; models/ub_uresnet.py:60,64)
//
// The tap-by-tap kernel (wgrad_gemm.cu) issues one M=128 MMA per (tap, 16 pixels) with only Cout of the 128
// accumulator rows useful, and re-fetches x once per tap.  Here ALL k*k taps of a 16-pixel K step are ONE
// tcgen05.mma, by stacking the tap shifts into the operand descriptors of rows that are resident in smem:
//
//   A (M side, MN-major) = x row h:  M index (a, ci), a = horizontal tap s.  MN atom a starts `span_x` bytes
//       (one pixel) after atom a-1  => descriptor LBO = span_x: overlapping windows of the same smem row.
//   B (N side, MN-major) = dy rows h-p .. h+p:  N index (b, co), b <-> vertical tap r = 2p-b.  N atom b is the dy
//       row one ring slot further => descriptor LBO = slot pitch.  The dy ring keeps 2p mirror slots behind its
//       last slot so that every window of 2p+1 consecutive rows is contiguous and ascending in smem.
//   K = 16 consecutive pixels of the 128-pixel row segment (8 MMAs per segment).
//
//   D[(a,ci)][(b,co)] accumulates in TMEM (fp32) over EVERY pixel the persistent CTA walks; one epilogue per CTA
//   writes an fp32 partial, and a second kernel sums the <=148 partials in a fixed order (deterministic).
//   x is read once (+2p halo columns), dy once (+2p halo rows per 32-row band): HBM-bound by construction.
#include "common.cuh"
#include "conv_internal.cuh"

namespace ubr {

constexpr int kWHThreads = 192;

struct WgradHaloParams {
  CUtensorMap x_map;            // x view (C, W, H, N), box {cin, 128+2p, 1, 1}
  CUtensorMap y_map;            // dy view (C, W, H, N), box {cout, 128, 1, 1}
  int pad, ksize, cin, cout;
  int nmma, natoms, nN;         // MMAs per K step, M atoms per MMA, N = ksize*cout
  int bands, cblocks, bh, ntiles;
  int sx, sy;                   // ring slots (dy ring has sy + 2p physical slots)
  uint32_t span_x, span_y, xrow_bytes, xslot_bytes, yrow_bytes, yslot_bytes;
  uint32_t idesc, tmem_cols;
  float* partials;              // [cta][r][s][cout][cin]
};

template <bool ACC>
__device__ __forceinline__ void mma_wg(uint32_t d_tmem, uint32_t alo, uint32_t ahi, uint32_t blo, uint32_t bhi,
                                       uint32_t idesc) {
  asm volatile(
      "{\n\t.reg .b64 da, db;\n\t.reg .pred p;\n\t"
      "mov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\t"
      "setp.ne.b32 p, %6, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}\n"
      ::"r"(d_tmem), "r"(alo), "r"(ahi), "r"(blo), "r"(bhi), "r"(idesc), "n"(ACC ? 1 : 0)
      : "memory");
}

__global__ void __launch_bounds__(kWHThreads, 1) wgrad_halo_kernel(const __grid_constant__ WgradHaloParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t xr_base = smem_base;
  const uint32_t yr_base = xr_base + (uint32_t)p.sx * p.xslot_bytes;
  const uint32_t bar_base = yr_base + (uint32_t)(p.sy + 2 * p.pad) * p.yslot_bytes;
  const uint32_t xfull0 = bar_base;
  const uint32_t xempty0 = xfull0 + 8u * p.sx;
  const uint32_t yfull0 = xempty0 + 8u * p.sx;
  const uint32_t yempty0 = yfull0 + 8u * p.sy;
  const uint32_t accum_bar = yempty0 + 8u * p.sy;
  const uint32_t tmem_slot = accum_bar + 8u;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t SX = (uint32_t)p.sx, SY = (uint32_t)p.sy;
  const int P2 = 2 * p.pad;
  const int rows_y = p.bh + P2;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.sx; ++s) { mbar_init(xfull0 + 8u * s, 1); mbar_init(xempty0 + 8u * s, 1); }
    for (int s = 0; s < p.sy; ++s) { mbar_init(yfull0 + 8u * s, 1); mbar_init(yempty0 + 8u * s, 1); }
    mbar_init(accum_bar, 1);
    fence_mbar_init();
    prefetch_tmap(&p.x_map);
    prefetch_tmap(&p.y_map);
  }
  if (warp == 1) {
    tmem_alloc_dyn(tmem_slot, p.tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp == 0) {
    // ================= TMA producer (whole warp walks the loop, one elected lane issues) =================
    uint32_t sx = 0, phx = 0, sy = 0, phy = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      const int band = tile % p.bands;
      const int cb = (tile / p.bands) % p.cblocks;
      const int n = tile / (p.bands * p.cblocks);
      const int h0 = band * p.bh, c0 = cb * 128;
      for (int j = 0; j < rows_y; ++j) {
        mbar_wait(yempty0 + 8u * sy, phy ^ 1u);
        if (j >= P2) mbar_wait(xempty0 + 8u * sx, phx ^ 1u);
        if (elect_one()) {
          const bool mirror = (int)sy < P2;
          const uint32_t fb = yfull0 + 8u * sy;
          mbar_arrive_expect_tx(fb, mirror ? 2u * p.yrow_bytes : p.yrow_bytes);
          tma_load_4d(yr_base + sy * p.yslot_bytes, &p.y_map, fb, 0, c0, h0 - p.pad + j, n);
          if (mirror) tma_load_4d(yr_base + (SY + sy) * p.yslot_bytes, &p.y_map, fb, 0, c0, h0 - p.pad + j, n);
          if (j >= P2) {
            const uint32_t fx = xfull0 + 8u * sx;
            mbar_arrive_expect_tx(fx, p.xrow_bytes);
            tma_load_4d(xr_base + sx * p.xslot_bytes, &p.x_map, fx, 0, c0 - p.pad, h0 + (j - P2), n);
          }
        }
        __syncwarp();
        if (++sy == SY) { sy = 0; phy ^= 1u; }
        if (j >= P2) { if (++sx == SX) { sx = 0; phx ^= 1u; } }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    // whole warp, warp-uniform control flow; one elected lane issues tcgen05.mma / tcgen05.commit
    // descriptors are (lo, hi) pairs; per MMA only the 14-bit start-address field of lo changes (one add)
    const uint64_t dx0 = make_smem_desc(0, p.span_x, 8u * p.span_x, swizzle_layout_code((int)p.span_x));
    const uint64_t dy0 = make_smem_desc(0, p.yslot_bytes, 8u * p.span_y, swizzle_layout_code((int)p.span_y));
    const uint32_t x_hi = (uint32_t)(dx0 >> 32), x_lo0 = (uint32_t)dx0 + (xr_base >> 4);
    const uint32_t y_hi = (uint32_t)(dy0 >> 32), y_lo0 = (uint32_t)dy0 + (yr_base >> 4);
    const uint32_t xs16 = p.xslot_bytes >> 4, ys16 = p.yslot_bytes >> 4;
    const uint32_t xk16 = p.span_x, yk16 = p.span_y;            // 16 pixels * span / 16
    const uint32_t xm16 = ((uint32_t)p.natoms * p.span_x) >> 4; // second MMA: M atoms shifted by natoms pixels
    const uint32_t idesc = p.idesc, nN = (uint32_t)p.nN;
    const bool two = p.nmma > 1;
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    uint32_t sx = 0, phx = 0, sy = 0, phy = 0;                  // x slot / dy slot of the current row (+ phases)
    bool first = true;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      for (int i = 0; i < p.bh; ++i) {
        {
          uint32_t s = sy, ph = phy;
          for (int j = 0; j <= P2; ++j) {
            if (i == 0 || j == P2) mbar_wait(yfull0 + 8u * s, ph);
            if (++s == SY) { s = 0; ph ^= 1u; }
          }
        }
        mbar_wait(xfull0 + 8u * sx, phx);
        tc_fence_after();
        const uint32_t a0 = x_lo0 + sx * xs16;
        const uint32_t b0 = y_lo0 + sy * ys16;
        const bool last = (i == p.bh - 1);
        if (elect_one()) {
          if (first) {
            mma_wg<false>(tmem_u, a0, x_hi, b0, y_hi, idesc);
            if (two) mma_wg<false>(tmem_u + nN, a0 + xm16, x_hi, b0, y_hi, idesc);
          } else {
            mma_wg<true>(tmem_u, a0, x_hi, b0, y_hi, idesc);
            if (two) mma_wg<true>(tmem_u + nN, a0 + xm16, x_hi, b0, y_hi, idesc);
          }
#pragma unroll
          for (int ks = 1; ks < 8; ++ks) {
            mma_wg<true>(tmem_u, a0 + ks * xk16, x_hi, b0 + ks * yk16, y_hi, idesc);
            if (two) mma_wg<true>(tmem_u + nN, a0 + ks * xk16 + xm16, x_hi, b0 + ks * yk16, y_hi, idesc);
          }
          tc_commit(xempty0 + 8u * sx);
          tc_commit(yempty0 + 8u * sy);
          if (last) {
            uint32_t s = sy;
            for (int j = 0; j < P2; ++j) {
              if (++s == SY) s = 0;
              tc_commit(yempty0 + 8u * s);
            }
          }
        }
        __syncwarp();
        first = false;
        if (++sx == SX) { sx = 0; phx ^= 1u; }
        const int adv = last ? 1 + P2 : 1;
        for (int j = 0; j < adv; ++j)
          if (++sy == SY) { sy = 0; phy ^= 1u; }
      }
    }
    if (elect_one()) tc_commit(accum_bar);
    __syncwarp();
    mbar_wait(accum_bar, 0);                       // the epilogue warps sleep on a named barrier, not on a spin loop
    asm volatile("bar.sync 1, 160;" ::: "memory");
  } else {
    // ================= epilogue: TMEM -> fp32 partial [r][s][co][ci] =================
    const int quad = warp & 3;
    const int m = quad * 32 + lane;
    const int a_loc = m / p.cin, ci = m % p.cin;
    asm volatile("bar.sync 1, 160;" ::: "memory");
    tc_fence_after();
    const uint32_t t_row = tmem_base + ((uint32_t)(quad * 32) << 16);
    float* base = p.partials + (size_t)blockIdx.x * p.ksize * p.ksize * p.cout * p.cin;
    for (int mm = 0; mm < p.nmma; ++mm) {
      const int a = mm * p.natoms + a_loc;                 // horizontal tap s
      const bool valid = a < p.ksize;
      for (int c = 0; c < p.nN; c += 16) {
        uint32_t v[16];
        tmem_ld16(t_row + (uint32_t)(mm * p.nN + c), v);
        tmem_ld_wait();
        if (valid) {
          const int b = c / p.cout, co0 = c % p.cout;      // cout is a multiple of 16: a 16-column group has one b
          const int r = p.ksize - 1 - b;
          float* dst = base + ((size_t)(r * p.ksize + a) * p.cout + co0) * p.cin + ci;
#pragma unroll
          for (int j = 0; j < 16; ++j) dst[(size_t)j * p.cin] = __uint_as_float(v[j]);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc_dyn(tmem_base, p.tmem_cols);
  }
}

// partials [ncta][r][s][cout][cin] -> dw OIHW [out_ch][cin][k][k]  (fixed summation order)
__global__ void wgrad_halo_reduce_kernel(const float* __restrict__ partials, int ncta, int ksize, int cout, int cin,
                                         int out_ch, float* __restrict__ dw, int accumulate) {
  const int total = ksize * ksize * cout * cin;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int ci = idx % cin;
  const int co = (idx / cin) % cout;
  const int rs = idx / (cin * cout);
  if (co >= out_ch) return;
  float acc = 0.f;
  for (int c = 0; c < ncta; ++c) acc += partials[(size_t)c * total + idx];
  float* d = dw + ((size_t)co * cin + ci) * ksize * ksize + rs;
  *d = accumulate ? (*d + acc) : acc;
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static bool wg_halo_disabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("UBR_NO_HALO_WGRAD"); v = (e && e[0] && e[0] != '0') ? 1 : 0; }
  return v == 1;
}

static bool wg_halo_shape_ok(int cin, int cout, int ksize, int h, int w) {
  if (!(cin == 16 || cin == 32 || cin == 64)) return false;
  if (!(cout == 16 || cout == 32 || cout == 64)) return false;
  if (!(ksize == 1 || ksize == 3 || ksize == 7)) return false;
  if (w < 128 || w % 128 != 0 || h % 8 != 0) return false;
  if (ksize * cout > 256) return false;
  const int natoms = 128 / cin;
  const int nmma = (ksize + natoms - 1) / natoms;
  if (nmma * ksize * cout > 512) return false;
  return true;
}

uint64_t wgrad_halo_workspace_bytes(int in_ch, int out_ch_pad, int ksize) {
  return (uint64_t)296 * ksize * ksize * out_ch_pad * in_ch * sizeof(float);
}

int wgrad_halo_try(const ubr_view& x, const ubr_view& dy, int out_ch, int ksize, float* dw, int accumulate,
                   void* workspace, uint64_t workspace_bytes, cudaStream_t stream) {
  if (wg_halo_disabled()) return 1;
  const int cin = x.c, cout = dy.c;
  if (!wg_halo_shape_ok(cin, cout, ksize, x.h, x.w)) return 1;
  if (dy.n != x.n || dy.h != x.h || dy.w != x.w) return 1;
  const int pad = (ksize - 1) / 2;
  int bh = 32;
  while (bh > 1 && x.h % bh != 0) bh >>= 1;
  if (bh < 8) return 1;
  const int natoms = 128 / cin;
  const int nmma = (ksize + natoms - 1) / natoms;
  const int nN = ksize * cout;
  const uint32_t span_x = (uint32_t)cin * 2u, span_y = (uint32_t)cout * 2u;
  const uint32_t xrow = (uint32_t)(128 + 2 * pad) * span_x;
  // junk M atoms (a >= ksize) read up to natoms*nmma-1 pixels past the window: keep them inside the slot
  const uint32_t xslot = ((uint32_t)(128 + natoms * nmma + 8) * span_x + 1023u) & ~1023u;
  const uint32_t yrow = 128u * span_y;
  const uint32_t yslot = yrow;                                  // 4096 / 8192 / 16384: 1024-aligned
  const uint32_t budget = 225u * 1024u - 2048u;
  // deep, balanced rings: ~1.5 us of HBM latency at 23 B/cycle/SM needs >= 64 KB in flight per SM.  The single
  // issuing thread is the other limiter, so prefer TWO CTAs per SM (half-depth rings each) whenever smem/TMEM allow.
  uint32_t cols = 32;
  while ((int)cols < nmma * nN) cols <<= 1;
  int sy = 0, sx = 0, per_sm = 1;
  {
    const uint32_t half = 110u * 1024u - 2048u;
    int s2 = (int)((half - (uint32_t)(2 * pad) * yslot) / (xslot + yslot));
    if (cols <= 256 && s2 >= 2 * pad + 3 && s2 >= 4) {
      if (s2 > 12) s2 = 12;
      sx = sy = s2; per_sm = 2;
    } else {
      int s = (int)((budget - (uint32_t)(2 * pad) * yslot) / (xslot + yslot));
      if (s > 16) s = 16;
      if (s >= 2 * pad + 2) { sx = s; sy = s; }
      else {
        // not enough room for equal depth: give dy its minimum window + 1 and x the rest
        sy = 2 * pad + 2;
        const uint32_t used = (uint32_t)(sy + 2 * pad) * yslot;
        if (used + 2u * xslot <= budget) { sx = (int)((budget - used) / xslot); if (sx > 16) sx = 16; }
        else sy = 0;
      }
    }
  }
  if (sy == 0) return 1;
  const int ntiles = (x.h / bh) * (x.w / 128) * x.n;
  int ctas = sm_count();
  if (ctas > 148) ctas = 148;
  ctas *= per_sm;
  if (const char* e = getenv("UBR_HALO_MAX_CTAS")) { const int v = atoi(e); if (v > 0 && ctas > v) ctas = v; }
  if (ctas > ntiles) ctas = ntiles;
  const uint64_t need = (uint64_t)ctas * ksize * ksize * cout * cin * sizeof(float);
  UBR_REQUIRE(workspace != nullptr && workspace_bytes >= need, "wgrad(halo): workspace too small (%llu < %llu)",
              (unsigned long long)workspace_bytes, (unsigned long long)need);
  UBR_REQUIRE(((uintptr_t)workspace & 15) == 0, "wgrad(halo): workspace must be 16-byte aligned");

  WgradHaloParams p;
  memset(&p, 0, sizeof(p));
  {
    uint64_t dims[4] = {(uint64_t)x.c, (uint64_t)x.w, (uint64_t)x.h, (uint64_t)x.n};
    uint64_t str[4] = {2, (uint64_t)x.ld * 2, (uint64_t)x.w * x.ld * 2, (uint64_t)x.h * x.w * x.ld * 2};
    uint32_t box[4] = {(uint32_t)cin, (uint32_t)(128 + 2 * pad), 1, 1};
    if (int rc = make_tmap_bf16(&p.x_map, x.ptr, 4, dims, str, box, (int)span_x)) return rc;
  }
  {
    uint64_t dims[4] = {(uint64_t)dy.c, (uint64_t)dy.w, (uint64_t)dy.h, (uint64_t)dy.n};
    uint64_t str[4] = {2, (uint64_t)dy.ld * 2, (uint64_t)dy.w * dy.ld * 2, (uint64_t)dy.h * dy.w * dy.ld * 2};
    uint32_t box[4] = {(uint32_t)cout, 128, 1, 1};
    if (int rc = make_tmap_bf16(&p.y_map, dy.ptr, 4, dims, str, box, (int)span_y)) return rc;
  }
  p.pad = pad; p.ksize = ksize; p.cin = cin; p.cout = cout;
  p.nmma = nmma; p.natoms = natoms; p.nN = nN;
  p.bh = bh; p.bands = x.h / bh; p.cblocks = x.w / 128; p.ntiles = ntiles;
  p.sx = sx; p.sy = sy;
  p.span_x = span_x; p.span_y = span_y; p.xrow_bytes = xrow; p.xslot_bytes = xslot;
  p.yrow_bytes = yrow; p.yslot_bytes = yslot;
  p.idesc = make_idesc_bf16(128, nN, 1, 1);
  p.tmem_cols = cols;
  p.partials = (float*)workspace;
  const uint32_t smem = (uint32_t)sx * xslot + (uint32_t)(sy + 2 * pad) * yslot + 16u * (sx + sy) + 32u + 1024u;
  static bool attr_set = false;
  if (!attr_set) {
    UBR_CUDA_OK(cudaFuncSetAttribute(wgrad_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set = true;
  }
  UBR_REQUIRE(smem <= 227u * 1024u, "wgrad(halo): smem plan %u too large", smem);
  wgrad_halo_kernel<<<ctas, kWHThreads, smem, stream>>>(p);
  UBR_LAUNCH_CHECK("wgrad_halo_kernel");
  const int total = ksize * ksize * cout * cin;
  wgrad_halo_reduce_kernel<<<(total + 255) / 256, 256, 0, stream>>>(p.partials, ctas, ksize, cout, cin, out_ch, dw,
                                                                    accumulate);
  UBR_LAUNCH_CHECK("wgrad_halo_reduce_kernel");
  return 0;
}

}  // namespace ubr
