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
//   * persistent CTAs (<= 2 per SM) walk (image, 128-column block, 32-row band) tiles; TMEM holds FOUR
//     accumulators so the MMA issuer runs two rows ahead of the epilogue; all filter taps stay resident in smem.
//
// Warp roles (320 threads): warp 0 TMA producer (weights once, then one row segment per ring slot), warp 1 TMEM
// owner + single-thread tcgen05.mma issuer (fully unrolled per filter size: descriptor = one integer add per MMA),
// warps 2-5 / 6-9 two independent epilogue groups taking alternate output rows
// (tcgen05.ld -> +bias -> bf16 -> direct 16-byte global stores when Cout <= 32, else swizzled staging + TMA store;
// fp32 direct store for the logits).
#include "common.cuh"
#include "conv_internal.cuh"

namespace ubr {

constexpr int kHThreads = 320;

struct HaloParams {
  CUtensorMap a_map;            // input view (C, W, H, N), box {cin, 128+2p, 1, 1}
  CUtensorMap w_map;            // packed weights (K, N, slab), box {cin, nt, 1}
  CUtensorMap o_map;            // output view, box {o_box_c, 128, 1, 1}  (out_mode 0 only)
  int ksize, pad, nt, flip;     // flip: data-gradient tap order (dh = p-r, dw = p-s)
  int bands, cblocks, bh, ntiles;
  int nslots;
  uint32_t span, row_bytes, slot_bytes, w_tap_bytes, w_region, stage_bytes;
  uint32_t idesc, tmem_cols, acc_stride;
  int nb;                       // TMEM ring: nb output-row accumulators of nt columns (power of two)
  int out_mode;                 // 0: bf16 staging + TMA store, 1: fp32 direct (logits), 2: bf16 direct stores,
                                // 3: bf16 pixel-shuffle stores (sub-pixel ConvTranspose: channel (ph,pw,co) -> pixel (2h+ph,2w+pw))
  int ps_cout;                  // mode 3: channels per output pixel
  int o_box_c;
  const float* bias;
  float* out_f32;
  __nv_bfloat16* out_bf16;
  int out_ld, out_cols, out_h, out_w;
  float* stats;                 // out_mode 2 only: per-CTA BN partials [row][2][stats_c] (sum, sum of squares), or null
  int stats_c, stats_rows;
  int accumulate;               // out_mode 2 only: out += result (fp32 add of the stored bf16 value, one rounding)
  int fused, relu_pre, relu_post;   // out_mode 0 / 2: fused inference epilogue (FusedEpi)
  const __nv_bfloat16* res;
  int res_ld;
  // out_mode 2 + stats: fused BatchNorm-backward reduction (BnReduce) instead of the forward statistics
  const __nv_bfloat16* bnr_y;
  int bnr_ld, bnr_relu;
  const float *bnr_scale, *bnr_shift, *bnr_mean, *bnr_invstd;
  // stats: the reduction over the partial rows and its finalize run inside this launch (tree_reduce.cuh)
  int fin_kind;                 // 0: none (rows are left for a finalize launch), 1: BnFinal, 2: SumsFinal
  TreeRed tr;
  BnFinal bnf;
  SumsFinal sf;
};

template <int N>
__device__ __forceinline__ void tma_store_wait_read_n() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}

// D[tmem] (+)= A * B with descriptors given as (lo, hi) 32-bit halves; ACC is a compile-time flag
template <bool ACC>
__device__ __forceinline__ void mma_lohi(uint32_t d_tmem, uint32_t alo, uint32_t ahi, uint32_t blo, uint32_t bhi,
                                         uint32_t idesc) {
  if (ACC) {
    asm volatile(
        "{\n\t.reg .b64 da, db;\n\t.reg .pred p;\n\t"
        "mov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\t"
        "setp.ne.b32 p, 1, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}\n"
        ::"r"(d_tmem), "r"(alo), "r"(ahi), "r"(blo), "r"(bhi), "r"(idesc)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .b64 da, db;\n\t.reg .pred p;\n\t"
        "mov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\t"
        "setp.ne.b32 p, 0, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}\n"
        ::"r"(d_tmem), "r"(alo), "r"(ahi), "r"(blo), "r"(bhi), "r"(idesc)
        : "memory");
  }
}

template <int KS, int KSTEPS>
__global__ void __launch_bounds__(kHThreads, 2) conv_halo_kernel(const __grid_constant__ HaloParams p) {
  pdl_launch_dependents();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t w_base = smem_base;
  const uint32_t ring_base = w_base + p.w_region;
  const uint32_t stage_base = ring_base + (uint32_t)p.nslots * p.slot_bytes;
  const uint32_t bar_base = stage_base + 2u * p.stage_bytes;
  const uint32_t full0 = bar_base;
  const uint32_t empty0 = full0 + 8u * p.nslots;
  const uint32_t wbar = empty0 + 8u * p.nslots;
  const uint32_t accf0 = wbar + 8u;         // [32]
  const uint32_t acce0 = accf0 + 256u;      // [32]
  const uint32_t tmem_slot = acce0 + 256u;

  constexpr int PAD = (KS - 1) / 2;
  constexpr int ROWS_EXTRA = 2 * PAD;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int co0 = blockIdx.y * p.nt;
  const uint32_t S = (uint32_t)p.nslots;
  const int rows_in = p.bh + ROWS_EXTRA;

  if (threadIdx.x == 0) {
    for (uint32_t s = 0; s < S; ++s) {
      mbar_init(full0 + 8u * s, 1);
      mbar_init(empty0 + 8u * s, 1);
    }
    mbar_init(wbar, 1);
    for (uint32_t b = 0; b < (uint32_t)p.nb; ++b) {
      mbar_init(accf0 + 8u * b, 1);
      mbar_init(acce0 + 8u * b, 4);
    }
    fence_mbar_init();
    prefetch_tmap(&p.a_map);
    prefetch_tmap(&p.w_map);
    if (p.out_mode == 0) prefetch_tmap(&p.o_map);
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
  // every accumulator of the TMEM ring starts at zero: all MMAs accumulate (see the issuer)
  if (warp >= 2 && warp < 6) {
    const uint32_t t_q = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    for (uint32_t c = 0; c < p.tmem_cols; c += 16) tmem_st16_zero(t_q + c);
    tmem_st_wait();
  }
  pdl_wait();                    // everything above touched only this CTA's shared memory / TMEM
  // per-channel constants of the fused BatchNorm-backward reduction: [scale | shift | invstd | -mean*invstd][32]
  float* bnr_cst = reinterpret_cast<float*>(smem_raw + (tmem_slot + 16u + 4096u - smem_u32(smem_raw)));
  if (p.bnr_y && threadIdx.x >= 64 && threadIdx.x < 192) {
    const int k = (threadIdx.x - 64) >> 5, ch = (threadIdx.x - 64) & 31;
    float v = 0.f;
    if (ch < p.nt) {
      const int gc = co0 + ch;
      v = k == 0 ? p.bnr_scale[gc] : k == 1 ? p.bnr_shift[gc] : k == 2 ? p.bnr_invstd[gc] : -p.bnr_mean[gc] * p.bnr_invstd[gc];
    }
    bnr_cst[k * 32 + ch] = v;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (warp == 0) {
    // ================= TMA producer (whole warp walks the loop, one elected lane issues) =================
    if (elect_one()) {
      mbar_arrive_expect_tx(wbar, (uint32_t)(KS * KS) * p.w_tap_bytes);
      // smem weight order [s'][j]: s' = column offset index (dw = s'-PAD), j = accumulator-ring order (dh = PAD-j)
      for (int sp = 0; sp < KS; ++sp)
        for (int j = 0; j < KS; ++j) {
          const int slab = p.flip ? (j * KS + (KS - 1 - sp)) : ((KS - 1 - j) * KS + sp);
          tma_load_3d(w_base + (uint32_t)(sp * KS + j) * p.w_tap_bytes, &p.w_map, wbar, 0, co0, slab);
        }
    }
    __syncwarp();
    uint32_t s = 0, ph = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      const int band = tile % p.bands;
      const int cb = (tile / p.bands) % p.cblocks;
      const int n = tile / (p.bands * p.cblocks);
      const int h0 = band * p.bh - PAD, c0 = cb * 128 - PAD;
      for (int j = 0; j < rows_in; ++j) {
        mbar_wait(empty0 + 8u * s, ph ^ 1u);
        if (elect_one()) {
          mbar_arrive_expect_tx(full0 + 8u * s, p.row_bytes);
          tma_load_4d(ring_base + s * p.slot_bytes, &p.a_map, full0 + 8u * s, 0, c0, h0 + j, n);
        }
        __syncwarp();
        if (++s == S) { s = 0; ph ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (row-stationary, taps stacked along N) =================
    // Input row R contributes to the KS output rows R-PAD..R+PAD.  Their accumulators are CONSECUTIVE column
    // blocks of a TMEM ring, so ONE tcgen05.mma per (column offset s', K step) with N = (#rows x nt) updates all
    // of them:  D[pixel][(j, co)] += A[row R shifted by s'][pixel][ci] * W[s'][j][co][ci].   KS x fewer MMA
    // instructions than tap-by-tap issue (each costs >= ~52 cycles at small N, profiles/r01_ubench T2).
    // The WHOLE warp walks the loop with warp-uniform values; one elected lane issues MMAs / commits.
    const uint32_t layout = swizzle_layout_code((int)p.span);
    const uint32_t desc_hi = (uint32_t)((make_smem_desc(0, 0, 8u * p.span, layout)) >> 32);
    const uint32_t span16 = p.span >> 4, slot16 = p.slot_bytes >> 4, wtap16 = p.w_tap_bytes >> 4;
    const uint32_t ring16 = ring_base >> 4, w16 = w_base >> 4;
    const uint32_t idesc0 = p.idesc;                       // N field cleared; N = rows * nt added per MMA
    const uint32_t nunit = ((uint32_t)p.nt >> 3) << 17;
    const uint32_t NBm = (uint32_t)p.nb - 1u, nbsh = (uint32_t)__ffs(p.nb) - 1u;
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    mbar_wait(wbar, 0);
    uint32_t s = 0, ph = 0;                  // ring slot / phase of the current input row
    uint32_t q0 = 0;                         // CTA-wide output row counter at the start of the tile
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, q0 += (uint32_t)p.bh) {
      for (int j = 0; j < rows_in; ++j) {
        mbar_wait(full0 + 8u * s, ph);
        const int olo = j > ROWS_EXTRA ? j - ROWS_EXTRA : 0;
        const int ohi = j < p.bh ? j : p.bh - 1;
        if (j < p.bh) {                      // first touch of output row j's accumulator: drained + zeroed?
          const uint32_t q = q0 + (uint32_t)j;
          mbar_wait(acce0 + 8u * (q & NBm), ((q >> nbsh) & 1u) ^ 1u);
        }
        tc_fence_after();
        const uint32_t blo = (q0 + (uint32_t)olo) & NBm;
        const uint32_t nrows = (uint32_t)(ohi - olo + 1);
        const uint32_t n1 = min(nrows, (uint32_t)p.nb - blo), n2 = nrows - n1;
        const uint32_t jb0 = (uint32_t)(olo - j + ROWS_EXTRA);     // first W block (ring order) this row touches
        if (elect_one()) {
          const uint32_t a_row = ring16 + s * slot16;
          const uint32_t d1 = tmem_u + blo * (uint32_t)p.nt;
          const uint32_t id1 = idesc0 + n1 * nunit, id2 = idesc0 + n2 * nunit;
#pragma unroll
          for (int sp = 0; sp < KS; ++sp) {
            const uint32_t a_t = a_row + (uint32_t)sp * span16;
            const uint32_t b_t = w16 + ((uint32_t)(sp * KS) + jb0) * wtap16;
#pragma unroll
            for (int k = 0; k < KSTEPS; ++k) {
              mma_lohi<true>(d1, a_t + 2u * k, desc_hi, b_t + 2u * k, desc_hi, id1);
              if (n2) mma_lohi<true>(tmem_u, a_t + 2u * k, desc_hi, b_t + n1 * wtap16 + 2u * k, desc_hi, id2);
            }
          }
          tc_commit(empty0 + 8u * s);                            // the row is consumed exactly once
          if (j >= ROWS_EXTRA) tc_commit(accf0 + 8u * ((q0 + (uint32_t)(j - ROWS_EXTRA)) & NBm));
        }
        __syncwarp();
        if (++s == S) { s = 0; ph ^= 1u; }
      }
    }
  } else {
    // ================= epilogue: group 0 = warps 2..5 (even rows), group 1 = warps 6..9 (odd rows) ============
    const int grp = (warp - 2) >> 2;
    const int quad = warp & 3;                 // TMEM lane quadrant this warp may access
    const int row = quad * 32 + lane;          // pixel inside the 128-wide column block
    const bool leader = (((warp - 2) & 3) == 0 && lane == 0);
    const int bc = p.o_box_c;
    const uint32_t rb = (uint32_t)bc * 2u;
    const uint32_t xr = rb == 128 ? (uint32_t)(row & 7) : (rb == 64 ? (uint32_t)((row >> 1) & 3) : (uint32_t)((row >> 2) & 1));
    const uint32_t sbase = stage_base + (uint32_t)grp * p.stage_bytes;
    const int bar_id = 1 + grp;
    float st_sum0 = 0.f, st_sq0 = 0.f, st_sum1 = 0.f, st_sq1 = 0.f;   // fused BN statistics: this lane's channel, per 16-channel chunk
    float st_sum2 = 0.f, st_sq2 = 0.f, st_sum3 = 0.f, st_sq3 = 0.f;   // chunks 2, 3: the 64-channel tiles of the TMA-store path
    uint32_t r0 = 0;                           // global output-row counter at the start of the tile
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, r0 += (uint32_t)p.bh) {
      const int band = tile % p.bands;
      const int cb = (tile / p.bands) % p.cblocks;
      const int n = tile / (p.bands * p.cblocks);
      const int c0 = cb * 128;
      for (int i = grp; i < p.bh; i += 2) {
        const uint32_t r = r0 + (uint32_t)i;
        const int h = band * p.bh + i;
        const uint32_t b = r & ((uint32_t)p.nb - 1u);
        if (p.out_mode == 2 && (p.bnr_y || p.accumulate) && i + 2 < p.bh) {
          // the epilogue's own global reads (y of the fused BatchNorm-backward reduction, the old value of an
          // accumulating store) sit on its serial per-row chain: pull the NEXT row of this group towards L2 now
          const size_t npx = (size_t)((size_t)n * p.out_h + h + 2) * p.out_w + (c0 + row);
          if (p.bnr_y) {
            const char* a = reinterpret_cast<const char*>(p.bnr_y + npx * (size_t)p.bnr_ld + co0);
            asm volatile("prefetch.global.L2 [%0];" ::"l"(a));
            if (p.nt > 16) asm volatile("prefetch.global.L2 [%0];" ::"l"(a + 32));
          }
          if (p.accumulate) {
            const char* a = reinterpret_cast<const char*>(p.out_bf16 + npx * (size_t)p.out_ld + co0);
            asm volatile("prefetch.global.L2 [%0];" ::"l"(a));
            if (p.nt > 16) asm volatile("prefetch.global.L2 [%0];" ::"l"(a + 32));
          }
        }
        mbar_wait(accf0 + 8u * b, (r >> ((uint32_t)__ffs(p.nb) - 1u)) & 1u);
        tc_fence_after();
        const uint32_t t_row = tmem_base + b * (uint32_t)p.nt + ((uint32_t)(quad * 32) << 16);
        const size_t pix = (size_t)((size_t)n * p.out_h + h) * p.out_w + (c0 + row);      // this thread's output pixel
        if (p.out_mode == 2) {
          // bf16 direct stores: one pixel (nt <= 32 channels) per thread, 16-byte vectors
          __nv_bfloat16* dst = p.out_bf16 + ((size_t)((size_t)n * p.out_h + h) * p.out_w + (c0 + row)) * (size_t)p.out_ld + co0;
          for (int c = 0; c < p.nt; c += 16) {
            uint32_t v[16];
            tmem_ld16(t_row + (uint32_t)c, v);
            tmem_ld_wait();
            tmem_st16_zero(t_row + (uint32_t)c);             // hand the accumulator back zeroed
            float f[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) f[j] = __uint_as_float(v[j]);
            if (p.bias) {
#pragma unroll
              for (int j = 0; j < 16; ++j) f[j] += __ldg(p.bias + co0 + c + j);
            }
            if (p.fused)
              fused_epi16(f, p.relu_pre, p.res ? p.res + pix * (size_t)p.res_ld + co0 + c : nullptr, p.relu_post);
            if (p.accumulate) {
              const uint4 o0 = *reinterpret_cast<const uint4*>(dst + c);
              const uint4 o1 = *reinterpret_cast<const uint4*>(dst + c + 8);
              const uint32_t ow[8] = {o0.x, o0.y, o0.z, o0.w, o1.x, o1.y, o1.z, o1.w};
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float2 t = unpack_bf16x2(ow[j]);
                f[2 * j] += t.x; f[2 * j + 1] += t.y;
              }
            }
            uint4 u0, u1;
            u0.x = pack_bf16x2(f[0], f[1]); u0.y = pack_bf16x2(f[2], f[3]);
            u0.z = pack_bf16x2(f[4], f[5]); u0.w = pack_bf16x2(f[6], f[7]);
            u1.x = pack_bf16x2(f[8], f[9]); u1.y = pack_bf16x2(f[10], f[11]);
            u1.z = pack_bf16x2(f[12], f[13]); u1.w = pack_bf16x2(f[14], f[15]);
            *reinterpret_cast<uint4*>(dst + c) = u0;
            *reinterpret_cast<uint4*>(dst + c + 8) = u1;
            if (p.stats) {
              // statistics of the STORED (bf16-rounded) values, like the stand-alone bn_stats kernel
              const uint32_t uw[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
              float g[16], q[16];
              if (p.bnr_y) {
                // BatchNorm(+ReLU) backward sums of the stored gradient: dz = da * [y*scale+shift > 0], dz * xhat
                const __nv_bfloat16* yp = p.bnr_y + pix * (size_t)p.bnr_ld + co0 + c;
                const uint4 y0 = __ldg(reinterpret_cast<const uint4*>(yp));
                const uint4 y1 = __ldg(reinterpret_cast<const uint4*>(yp + 8));
                const uint32_t yw[8] = {y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w};
                // constants are read pair by pair from shared memory (warp-uniform addresses: broadcast), keeping
                // the register footprint of the epilogue below the two-CTAs-per-SM limit
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  const float2 t = unpack_bf16x2(uw[j]);
                  const float2 v = unpack_bf16x2(yw[j]);
                  const float2 ksc = *reinterpret_cast<const float2*>(bnr_cst + c + 2 * j);
                  const float2 ksh = *reinterpret_cast<const float2*>(bnr_cst + 32 + c + 2 * j);
                  const float2 kis = *reinterpret_cast<const float2*>(bnr_cst + 64 + c + 2 * j);
                  const float2 knb = *reinterpret_cast<const float2*>(bnr_cst + 96 + c + 2 * j);
                  const float d0 = (!p.bnr_relu || fmaf(v.x, ksc.x, ksh.x) > 0.f) ? t.x : 0.f;
                  const float d1 = (!p.bnr_relu || fmaf(v.y, ksc.y, ksh.y) > 0.f) ? t.y : 0.f;
                  g[2 * j] = d0; g[2 * j + 1] = d1;
                  q[2 * j] = d0 * fmaf(v.x, kis.x, knb.x);
                  q[2 * j + 1] = d1 * fmaf(v.y, kis.y, knb.y);
                }
              } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  const float2 t = unpack_bf16x2(uw[j]);
                  g[2 * j] = t.x; g[2 * j + 1] = t.y;
                  q[2 * j] = t.x * t.x; q[2 * j + 1] = t.y * t.y;
                }
              }
              const float rs = warp_reduce_scatter16(g, lane), rq = warp_reduce_scatter16(q, lane);
              if (c == 0) { st_sum0 += rs; st_sq0 += rq; } else { st_sum1 += rs; st_sq1 += rq; }
            }
          }
          tmem_st_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(acce0 + 8u * b);
        } else if (p.out_mode == 3) {
          // sub-pixel ConvTranspose: this thread's input-grid pixel (h, c0+row) owns 4 output pixels
          for (int c = 0; c < p.nt; c += 16) {
            uint32_t v[16];
            tmem_ld16(t_row + (uint32_t)c, v);
            tmem_ld_wait();
            tmem_st16_zero(t_row + (uint32_t)c);
            const int cc = co0 + c;
            const int ph = cc / (2 * p.ps_cout), pw = (cc / p.ps_cout) & 1, co = cc % p.ps_cout;
            __nv_bfloat16* dst = p.out_bf16 + ((size_t)((size_t)n * (2 * p.out_h) + (2 * h + ph)) * (2 * p.out_w) +
                                               (size_t)(2 * (c0 + row) + pw)) * (size_t)p.out_ld + co;
            uint4 u0, u1;
            u0.x = pack_bf16x2(__uint_as_float(v[0]), __uint_as_float(v[1]));
            u0.y = pack_bf16x2(__uint_as_float(v[2]), __uint_as_float(v[3]));
            u0.z = pack_bf16x2(__uint_as_float(v[4]), __uint_as_float(v[5]));
            u0.w = pack_bf16x2(__uint_as_float(v[6]), __uint_as_float(v[7]));
            u1.x = pack_bf16x2(__uint_as_float(v[8]), __uint_as_float(v[9]));
            u1.y = pack_bf16x2(__uint_as_float(v[10]), __uint_as_float(v[11]));
            u1.z = pack_bf16x2(__uint_as_float(v[12]), __uint_as_float(v[13]));
            u1.w = pack_bf16x2(__uint_as_float(v[14]), __uint_as_float(v[15]));
            *reinterpret_cast<uint4*>(dst) = u0;
            *reinterpret_cast<uint4*>(dst + 8) = u1;
          }
          tmem_st_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(acce0 + 8u * b);
        } else if (p.out_mode == 0) {
          if (leader) tma_store_wait_read_n<0>();          // this group's previous store has read the staging tile
          asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
          for (int c = 0; c < p.nt; c += 16) {
            uint32_t v[16];
            tmem_ld16(t_row + (uint32_t)c, v);
            tmem_ld_wait();
            tmem_st16_zero(t_row + (uint32_t)c);             // hand the accumulator back zeroed
            float f[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) f[j] = __uint_as_float(v[j]);
            if (p.bias) {
#pragma unroll
              for (int j = 0; j < 16; ++j) f[j] += __ldg(p.bias + co0 + c + j);
            }
            if (p.fused)
              fused_epi16(f, p.relu_pre, p.res ? p.res + pix * (size_t)p.res_ld + co0 + c : nullptr, p.relu_post);
            uint32_t pk[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) pk[j] = pack_bf16x2(f[2 * j], f[2 * j + 1]);
            if (p.stats) {
              // statistics of the STORED (bf16-rounded) values (TMA-store path: no BatchNorm-backward variant here)
              float g[16], q2[16];
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float2 t = unpack_bf16x2(pk[j]);
                g[2 * j] = t.x; g[2 * j + 1] = t.y;
                q2[2 * j] = t.x * t.x; q2[2 * j + 1] = t.y * t.y;
              }
              const float rs = warp_reduce_scatter16(g, lane), rq = warp_reduce_scatter16(q2, lane);
              const int ci = c >> 4;
              if (ci == 0) { st_sum0 += rs; st_sq0 += rq; }
              else if (ci == 1) { st_sum1 += rs; st_sq1 += rq; }
              else if (ci == 2) { st_sum2 += rs; st_sq2 += rq; }
              else { st_sum3 += rs; st_sq3 += rq; }
            }
            const int blk = c / bc;
            const uint32_t q = (uint32_t)((c % bc) >> 3);
            const uint32_t base = sbase + (uint32_t)blk * 128u * rb + (uint32_t)row * rb;
            const uint32_t a0 = base + ((q ^ xr) << 4);
            const uint32_t a1 = base + (((q + 1u) ^ xr) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a0), "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3]) : "memory");
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a1), "r"(pk[4]), "r"(pk[5]), "r"(pk[6]), "r"(pk[7]) : "memory");
          }
          tmem_st_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(acce0 + 8u * b);      // accumulator drained: the issuer may overwrite it
          fence_proxy_async_smem();
          asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
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
            tmem_st16_zero(t_row + (uint32_t)c);             // hand the accumulator back zeroed
            if (p.out_ld == 4 && c == 0 && co0 == 0) {
              float4 o;
              o.x = (0 < p.out_cols) ? __uint_as_float(v[0]) + (p.bias ? __ldg(p.bias + 0) : 0.f) : 0.f;
              o.y = (1 < p.out_cols) ? __uint_as_float(v[1]) + (p.bias ? __ldg(p.bias + 1) : 0.f) : 0.f;
              o.z = (2 < p.out_cols) ? __uint_as_float(v[2]) + (p.bias ? __ldg(p.bias + 2) : 0.f) : 0.f;
              o.w = (3 < p.out_cols) ? __uint_as_float(v[3]) + (p.bias ? __ldg(p.bias + 3) : 0.f) : 0.f;
              *reinterpret_cast<float4*>(dst) = o;
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const int col = co0 + c + j;
                if (col < p.out_cols) dst[col] = __uint_as_float(v[j]) + (p.bias ? __ldg(p.bias + col) : 0.f);
              }
            }
          }
          tmem_st_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(acce0 + 8u * b);
        }
      }
    }
    if (p.out_mode == 0 && leader) tma_store_wait_all();
    if (p.out_mode == 0 && p.stats) {
      // same fixed-order combine as below, for up to four 16-channel chunks; the scratch is the first staging tile, idle
      // once every TMA store of the CTA has completed
      asm volatile("bar.sync 3, 256;" ::: "memory");
      float* scr8 = reinterpret_cast<float*>(smem_raw + (stage_base - smem_u32(smem_raw)));        // [8 warps][32 lanes][8]
      const int we = warp - 2;
      float* mine = scr8 + (we * 32 + lane) * 8;
      mine[0] = st_sum0; mine[1] = st_sq0; mine[2] = st_sum1; mine[3] = st_sq1;
      mine[4] = st_sum2; mine[5] = st_sq2; mine[6] = st_sum3; mine[7] = st_sq3;
      asm volatile("bar.sync 3, 256;" ::: "memory");
      if (we == 0 && (lane & 1) == 0) {
        const int ch = (lane >> 1) & 15;
        float tot[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int k = 0; k < 8; ++k)
#pragma unroll
          for (int j = 0; j < 8; ++j) tot[j] += scr8[(k * 32 + lane) * 8 + j];
        float* rowp = p.stats + (size_t)blockIdx.x * 2 * p.stats_c;
        for (int cc = 0; cc < p.nt; cc += 16) {
          rowp[co0 + cc + ch] = tot[(cc >> 4) * 2];
          rowp[p.stats_c + co0 + cc + ch] = tot[(cc >> 4) * 2 + 1];
        }
      }
    }
    if (p.out_mode == 2 && p.stats) {
      // combine the 8 epilogue warps in a fixed order (deterministic) and publish this CTA's partial row
      float* scr = reinterpret_cast<float*>(smem_raw + (tmem_slot + 16u - smem_u32(smem_raw)));   // [8 warps][32 lanes][4]
      const int we = warp - 2;
      float* mine = scr + (we * 32 + lane) * 4;
      mine[0] = st_sum0; mine[1] = st_sq0; mine[2] = st_sum1; mine[3] = st_sq1;
      asm volatile("bar.sync 3, 256;" ::: "memory");
      if (we == 0 && (lane & 1) == 0) {
        const int ch = (lane >> 1) & 15;
        float tot[4] = {0.f, 0.f, 0.f, 0.f};
        for (int k = 0; k < 8; ++k)
#pragma unroll
          for (int j = 0; j < 4; ++j) tot[j] += scr[(k * 32 + lane) * 4 + j];
        float* rowp = p.stats + (size_t)blockIdx.x * 2 * p.stats_c;
        for (int cc = 0; cc < p.nt; cc += 16) {
          rowp[co0 + cc + ch] = tot[(cc >> 4) * 2];
          rowp[p.stats_c + co0 + cc + ch] = tot[(cc >> 4) * 2 + 1];
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
  if (p.fin_kind) {
    // partial row blockIdx.x is shared by the gridDim.y channel tiles; the last CTA of the grid finishes the reduction
    double* ts = reinterpret_cast<double*>(smem_raw + (smem_base - smem_u32(smem_raw)));     // the idle weight / ring region
    if (tree_arrive(p.tr, (int)blockIdx.x, 1, (int)gridDim.y, ts)) {
      if (p.fin_kind == 1) tree_finish_bn(ts, p.bnf);
      else tree_finish_sums(ts, p.sf);
    }
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

typedef void (*halo_kernel_t)(const HaloParams);
static halo_kernel_t pick_halo_kernel(int ks, int ksteps) {
  switch (ks * 10 + ksteps) {
    case 11: return conv_halo_kernel<1, 1>;
    case 12: return conv_halo_kernel<1, 2>;
    case 14: return conv_halo_kernel<1, 4>;
    case 31: return conv_halo_kernel<3, 1>;
    case 32: return conv_halo_kernel<3, 2>;
    case 34: return conv_halo_kernel<3, 4>;
    case 71: return conv_halo_kernel<7, 1>;
    case 72: return conv_halo_kernel<7, 2>;
    default: return nullptr;
  }
}

int conv_halo_try(const ubr_view& x, const void* wpk, int nslabs, int n_pad, const TapSpec* taps, int ntaps,
                  const float* bias, const ubr_view* y, float* y_f32, int y_f32_ld, int out_cols,
                  cudaStream_t stream, int accumulate, float* stats, int* stats_rows, int shuffle_cout,
                  const FusedEpi* epi, const BnReduce* bnr, const FusedFinal* fin) {
  if (halo_disabled()) return 1;
  const int cin = x.c;
  if (!(cin == 16 || cin == 32 || cin == 64)) return 1;
  if (x.w % 128 != 0 || x.w < 128) return 1;
  int ks = 0;
  if (ntaps == 1) ks = 1; else if (ntaps == 9) ks = 3; else if (ntaps == 49) ks = 7; else return 1;
  const int pad = (ks - 1) / 2;
  // regular tap grid only: forward (dh = r-p, dw = s-p) or data gradient (dh = p-r, dw = p-s), slab = r*ks+s
  int flip = -1;
  for (int sgn = 1; sgn >= -1 && flip < 0; sgn -= 2) {
    bool ok = true;
    for (int r = 0; r < ks && ok; ++r)
      for (int s = 0; s < ks && ok; ++s) {
        const TapSpec& t = taps[r * ks + s];
        ok = t.amap == 0 && t.slab == r * ks + s && t.dh == sgn * (r - pad) && t.dw == sgn * (s - pad);
      }
    if (ok) flip = sgn > 0 ? 0 : 1;
  }
  if (flip < 0) return 1;
  halo_kernel_t kern = pick_halo_kernel(ks, cin / 16);
  if (!kern) return 1;
  int bh = 32;
  while (bh > 1 && x.h % bh != 0) bh >>= 1;
  if (bh < 8) return 1;
  {
    // small images (128^2 layers at batch 32: 128 tiles of 32 rows for 148-296 resident CTAs) leave SMs idle and every
    // CTA pays its pipeline fill once for a single tile: shorter bands give each CTA >= 2 tiles (UBR_HALO_MIN_TILES per
    // SM; the extra halo rows cost 2*pad/bh more reads)
    static const int min_tiles = [] { const char* e = getenv("UBR_HALO_MIN_TILES"); return e ? atoi(e) : 0; }();
    if (min_tiles > 0)
      while (bh > 8 && (x.h / bh) * (x.w / 128) * x.n < min_tiles * sm_count() && x.h % (bh / 2) == 0) bh >>= 1;
  }
  const uint32_t span = (uint32_t)cin * 2u;
  const uint32_t row_bytes = (uint32_t)(128 + 2 * pad) * span;
  const uint32_t slot_bytes = (row_bytes + 1023u) & ~1023u;
  // N tile: largest divisor tile whose resident weights + ring + staging fit
  const uint32_t budget = 225u * 1024u - 1024u - 512u - 4096u - 512u;
  int nt = 0, nslots = 0, out_mode = 0;
  uint32_t w_region = 0, stage_bytes = 0;
  int nt_max = ks == 1 ? 128 : (ks == 3 ? 64 : 32);           // stacked MMA: ks * nt <= 256 columns
  if (shuffle_cout && nt_max > 64) nt_max = 64;
  for (int cand = n_pad > nt_max ? nt_max : n_pad; cand >= 16; cand -= 16) {
    if (n_pad % cand) continue;
    const uint32_t wr = ((uint32_t)ntaps * cand * span + 1023u) & ~1023u;
    const int mode = y_f32 ? 1 : (shuffle_cout ? 3 : (cand <= 32 ? 2 : 0));
    const uint32_t st = mode == 0 ? 128u * (uint32_t)cand * 2u : 0u;
    if (wr + 2u * st + 4u * slot_bytes > budget) continue;
    int s = (int)((budget - wr - 2u * st) / slot_bytes);
    if (s > 16) s = 16;
    // two CTAs per SM (two issuing threads) beat one deep ring: shrink the ring if that makes a second CTA fit
    {
      const uint32_t half = 110u * 1024u - 1024u - 1024u - 4096u - 512u;
      if (wr + 2u * st + 6u * slot_bytes <= half && (ks + 3 <= 8 ? 8 : 16) * cand <= 256) {
        const int s2 = (int)((half - wr - 2u * st) / slot_bytes);
        if (s2 < s) s = s2;
      }
    }
    nt = cand; nslots = s; w_region = wr; stage_bytes = st; out_mode = mode;
    break;
  }
  if (nt == 0) return 1;
  if (accumulate && out_mode != 2 && out_mode != 0) return 1;  // in-epilogue accumulation: bf16 outputs only
  // fused BN statistics: the direct-store epilogue, and (forward statistics only) the TMA-store epilogue of the 64-channel
  // tiles (UBR_HALO_STATS_M0=0 leaves those to the stand-alone bn_stats pass)
  static const bool stats_m0 = [] { const char* e = getenv("UBR_HALO_STATS_M0"); return !(e && e[0] == '0'); }();
  if (stats && out_mode != 2 && !(stats_m0 && out_mode == 0 && !bnr && !accumulate && !epi && nt <= 64)) return 1;
  if (bnr && (!stats || out_mode != 2 || nt > 32 || accumulate)) return 1;

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
  p.out_mode = out_mode;
  p.out_h = x.h; p.out_w = x.w;
  if (out_mode == 1) {
    p.out_f32 = y_f32; p.out_ld = y_f32_ld; p.out_cols = out_cols;
  } else if (out_mode == 2 || out_mode == 3) {
    p.out_bf16 = (__nv_bfloat16*)y->ptr; p.out_ld = y->ld; p.out_cols = out_cols; p.ps_cout = shuffle_cout;
  } else {
    p.o_box_c = nt >= 64 ? 64 : nt;
    uint64_t dims[4] = {(uint64_t)y->c, (uint64_t)y->w, (uint64_t)y->h, (uint64_t)y->n};
    uint64_t str[4] = {2, (uint64_t)y->ld * 2, (uint64_t)y->w * y->ld * 2, (uint64_t)y->h * y->w * y->ld * 2};
    uint32_t box[4] = {(uint32_t)p.o_box_c, 128, 1, 1};
    if (int rc = make_tmap_bf16(&p.o_map, y->ptr, 4, dims, str, box, p.o_box_c * 2)) return rc;
  }
  if (p.o_box_c == 0) p.o_box_c = 16;
  p.ksize = ks; p.pad = pad; p.nt = nt; p.flip = flip;
  p.bh = bh; p.bands = x.h / bh; p.cblocks = x.w / 128; p.ntiles = p.bands * p.cblocks * x.n;
  p.nslots = nslots;
  p.span = span; p.row_bytes = row_bytes; p.slot_bytes = slot_bytes;
  p.w_tap_bytes = (uint32_t)nt * span; p.w_region = w_region; p.stage_bytes = stage_bytes;
  p.idesc = make_idesc_bf16(128, 0, 0, 0);                   // N = (#stacked rows * nt) is added per MMA
  {
    int nb = 4;
    while (nb < ks + 3) nb <<= 1;                            // rows in flight (ks) + slack for the epilogue
    while (nb * nt > 512 && nb > 1) nb >>= 1;
    if (nb < ks + 1) return 1;
    p.nb = nb;
    uint32_t cols = 32;
    while ((int)cols < nb * nt) cols <<= 1;
    p.tmem_cols = cols;
    p.acc_stride = (uint32_t)nt;
  }
  p.bias = bias;
  p.accumulate = accumulate && out_mode == 2;
  p.stats = stats; p.stats_c = n_pad;
  if (bnr) {
    p.bnr_y = (const __nv_bfloat16*)bnr->y; p.bnr_ld = bnr->y_ld; p.bnr_relu = bnr->relu;
    p.bnr_scale = bnr->scale; p.bnr_shift = bnr->shift; p.bnr_mean = bnr->mean; p.bnr_invstd = bnr->invstd;
  }
  if (accumulate && out_mode == 0) {
    // TMA-store path: the old value of the output is the "residual" of the fused epilogue (read by the thread that
    // owns the pixel before the tile is stored -- in place is safe, tiles are disjoint)
    UBR_REQUIRE(epi == nullptr && y != nullptr, "conv_halo: accumulate and a fused epilogue are exclusive");
    p.fused = 1; p.relu_pre = 0; p.relu_post = 0;
    p.res = (const __nv_bfloat16*)y->ptr; p.res_ld = y->ld;
  }
  if (epi) {
    UBR_REQUIRE(p.out_mode == 0 || p.out_mode == 2, "conv_halo: the fused epilogue needs a bf16 output");
    p.fused = 1; p.relu_pre = epi->relu_pre; p.relu_post = epi->relu_post;
    p.res = (const __nv_bfloat16*)epi->res; p.res_ld = epi->res_ld;
  }
  const uint32_t smem = w_region + (uint32_t)nslots * slot_bytes + 2u * stage_bytes + 16u * nslots + 640u + 4096u + 512u + 1024u;
  UBR_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  const int n_tiles = n_pad / nt;
  int per_sm = (smem <= 110u * 1024u && p.tmem_cols <= 256u) ? 2 : 1;
  int ctas = per_sm * sm_count() / n_tiles;
  if (const char* e = getenv("UBR_HALO_MAX_CTAS")) { const int v = atoi(e); if (v > 0 && ctas > v) ctas = v; }
  if (ctas < 1) ctas = 1;
  if (ctas > p.ntiles) ctas = p.ntiles;
  p.stats_rows = ctas;
  if (stats_rows) *stats_rows = ctas;
  if (fin && stats) {
    UBR_REQUIRE(fin->tree != nullptr && ctas <= kTreeGroupCtas * kTreeMaxGroups, "conv_halo: bad fused-finalize state");
    p.fin_kind = fin->kind;
    p.tr = make_tree(fin->tree, stats, ctas, 2 * n_pad);
    p.bnf = fin->bn; p.bnf.sq_off = n_pad;
    p.sf = fin->sums; p.sf.row_stride = n_pad;
  }
  dim3 grid((unsigned)ctas, (unsigned)n_tiles, 1);
  ::ubr::Launch(grid, kHThreads, smem, stream)(kern, p);
  UBR_LAUNCH_CHECK("conv_halo_kernel");
  return 0;
}

}  // namespace ubr
