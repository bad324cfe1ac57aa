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

template <int KS, int NMMA>
__global__ void __launch_bounds__(kWHThreads, 1) wgrad_halo_kernel(const __grid_constant__ WgradHaloParams p) {
  pdl_launch_dependents();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  constexpr int P2 = KS - 1;
  constexpr int PAD = (KS - 1) / 2;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  // unified ring: slot = [dy row (128 px) | x row (128+2p px, padded)]; the last P2 slots mirror slots 0..P2-1 (dy part)
  const uint32_t ring = smem_base;
  const uint32_t slot_bytes = p.yslot_bytes + p.xslot_bytes;
  const uint32_t bar_base = ring + (uint32_t)(p.sy + P2) * slot_bytes;
  const uint32_t full0 = bar_base;
  const uint32_t empty0 = full0 + 8u * p.sy;
  const uint32_t accum_bar = empty0 + 8u * p.sy;
  const uint32_t tmem_slot = accum_bar + 8u;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t SY = (uint32_t)p.sy;
  const int rows_y = p.bh + P2;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.sy; ++s) { mbar_init(full0 + 8u * s, 1); mbar_init(empty0 + 8u * s, 1); }
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
  pdl_wait();
  if (warp >= 2) {                                   // accumulators start at zero: every MMA accumulates
    const uint32_t t_q = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    for (uint32_t c = 0; c < p.tmem_cols; c += 16) tmem_st16_zero(t_q + c);
    tmem_st_wait();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (warp == 0) {
    // ================= TMA producer (whole warp walks the loop, one elected lane issues) =================
    uint32_t s = 0, ph = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      const int band = tile % p.bands;
      const int cb = (tile / p.bands) % p.cblocks;
      const int n = tile / (p.bands * p.cblocks);
      const int h0 = band * p.bh, c0 = cb * 128;
      for (int j = 0; j < rows_y; ++j) {
        mbar_wait(empty0 + 8u * s, ph ^ 1u);
        if (elect_one()) {
          const bool mirror = (int)s < P2;
          const bool has_x = j < p.bh;
          const uint32_t fb = full0 + 8u * s;
          mbar_arrive_expect_tx(fb, (mirror ? 2u * p.yrow_bytes : p.yrow_bytes) + (has_x ? p.xrow_bytes : 0u));
          tma_load_4d(ring + s * slot_bytes, &p.y_map, fb, 0, c0, h0 - PAD + j, n);
          if (mirror) tma_load_4d(ring + (SY + s) * slot_bytes, &p.y_map, fb, 0, c0, h0 - PAD + j, n);
          if (has_x) tma_load_4d(ring + s * slot_bytes + p.yslot_bytes, &p.x_map, fb, 0, c0 - PAD, h0 + j, n);
        }
        __syncwarp();
        if (++s == SY) { s = 0; ph ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    // whole warp, warp-uniform control flow; one elected lane issues tcgen05.mma / tcgen05.commit.
    // Per row: ONE barrier wait (the newest dy row), 8*NMMA MMAs (one integer add each), ONE commit.
    const uint64_t dx0 = make_smem_desc(0, p.span_x, 8u * p.span_x, swizzle_layout_code((int)p.span_x));
    const uint64_t dy0 = make_smem_desc(0, slot_bytes, 8u * p.span_y, swizzle_layout_code((int)p.span_y));
    const uint32_t x_hi = (uint32_t)(dx0 >> 32), x_lo0 = (uint32_t)dx0 + ((ring + p.yslot_bytes) >> 4);
    const uint32_t y_hi = (uint32_t)(dy0 >> 32), y_lo0 = (uint32_t)dy0 + (ring >> 4);
    const uint32_t s16 = slot_bytes >> 4;
    const uint32_t xk16 = p.span_x, yk16 = p.span_y;            // 16 pixels * span / 16
    const uint32_t xm16 = ((uint32_t)p.natoms * p.span_x) >> 4; // second MMA: M atoms shifted by natoms pixels
    const uint32_t idesc = p.idesc, nN = (uint32_t)p.nN;
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    uint32_t s = 0, ph = 0;                                     // slot / phase of the current x row
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      for (int i = 0; i < p.bh; ++i) {
        {
          uint32_t sw = s, pw = ph;
#pragma unroll
          for (int j = 0; j <= P2; ++j) {
            if (i == 0 || j == P2) mbar_wait(full0 + 8u * sw, pw);
            if (++sw == SY) { sw = 0; pw ^= 1u; }
          }
        }
        tc_fence_after();
        const uint32_t a0 = x_lo0 + s * s16;
        const uint32_t b0 = y_lo0 + s * s16;
        const bool last = (i == p.bh - 1);
        if (elect_one()) {
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            mma_wg<true>(tmem_u, a0 + ks * xk16, x_hi, b0 + ks * yk16, y_hi, idesc);
            if (NMMA > 1) mma_wg<true>(tmem_u + nN, a0 + ks * xk16 + xm16, x_hi, b0 + ks * yk16, y_hi, idesc);
          }
          tc_commit(empty0 + 8u * s);
          if (last) {
            uint32_t sw = s;
#pragma unroll
            for (int j = 0; j < P2; ++j) {
              if (++sw == SY) sw = 0;
              tc_commit(empty0 + 8u * sw);
            }
          }
        }
        __syncwarp();
        const int adv = last ? 1 + P2 : 1;
        for (int j = 0; j < adv; ++j)
          if (++s == SY) { s = 0; ph ^= 1u; }
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
    float* base = p.partials + (size_t)blockIdx.x * KS * KS * p.cout * p.cin;
    for (int mm = 0; mm < NMMA; ++mm) {
      const int a = mm * p.natoms + a_loc;                 // horizontal tap s
      const bool valid = a < KS;
      for (int c = 0; c < p.nN; c += 16) {
        uint32_t v[16];
        tmem_ld16(t_row + (uint32_t)(mm * p.nN + c), v);
        tmem_ld_wait();
        if (valid) {
          const int b = c / p.cout, co0 = c % p.cout;      // cout is a multiple of 16: a 16-column group has one b
          const int r = KS - 1 - b;
          float* dst = base + ((size_t)(r * KS + a) * p.cout + co0) * p.cin + ci;
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
  pdl_prologue();
  const int total = ksize * ksize * cout * cin;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int ci = idx % cin;
  const int co = (idx / cin) % cout;
  const int rs = idx / (cin * cout);
  if (co >= out_ch) return;
  // eight independent partial sums (fixed assignment c mod 8, fixed combination order: deterministic) keep eight loads
  // in flight instead of one dependent add chain over up to 296 partials
  float a8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  int c = 0;
  for (; c + 8 <= ncta; c += 8) {
#pragma unroll
    for (int j = 0; j < 8; ++j) a8[j] += partials[(size_t)(c + j) * total + idx];
  }
  for (int j = 0; c < ncta; ++c, ++j) a8[j] += partials[(size_t)c * total + idx];
  const float acc = ((a8[0] + a8[1]) + (a8[2] + a8[3])) + ((a8[4] + a8[5]) + (a8[6] + a8[7]));
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
    // unified ring: every slot holds one dy row and one x row; 2p extra (mirror) slots
    const uint32_t slot = xslot + yslot;
    const uint32_t half = 110u * 1024u - 2048u;
    int s2 = (int)(half / slot) - 2 * pad;
    if (cols <= 256 && s2 >= 2 * pad + 3 && s2 >= 4) {
      if (s2 > 12) s2 = 12;
      sy = s2; per_sm = 2;
    } else {
      int s = (int)(budget / slot) - 2 * pad;
      if (s > 16) s = 16;
      if (s >= 2 * pad + 2) sy = s;
    }
    sx = sy;
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
  const uint32_t smem = (uint32_t)(sy + 2 * pad) * (xslot + yslot) + 16u * sy + 32u + 1024u;
  typedef void (*wg_kernel_t)(const WgradHaloParams);
  wg_kernel_t kern = nullptr;
  switch (ksize * 10 + nmma) {
    case 11: kern = wgrad_halo_kernel<1, 1>; break;
    case 31: kern = wgrad_halo_kernel<3, 1>; break;
    case 32: kern = wgrad_halo_kernel<3, 2>; break;
    case 71: kern = wgrad_halo_kernel<7, 1>; break;
    case 72: kern = wgrad_halo_kernel<7, 2>; break;
    default: return 1;
  }
  UBR_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  UBR_REQUIRE(smem <= 227u * 1024u, "wgrad(halo): smem plan %u too large", smem);
  ::ubr::Launch(ctas, kWHThreads, smem, stream)(kern, p);
  UBR_LAUNCH_CHECK("wgrad_halo_kernel");
  const int total = ksize * ksize * cout * cin;
  ::ubr::Launch((total + 255) / 256, 256, 0, stream)(wgrad_halo_reduce_kernel, p.partials, ctas, ksize, cout, cin, out_ch, dw,
                                                                    accumulate);
  UBR_LAUNCH_CHECK("wgrad_halo_reduce_kernel");
  return 0;
}

}  // namespace ubr
