// Implicit-GEMM "gather" convolution on tcgen05 tensor cores (sm_100a).
//
//   out[pixel, co] = sum_{tap} sum_{ci} A_view[tap.map][pixel + (tap.dh, tap.dw), ci] * Wpk[tap.slab][co][ci]
//
// One kernel covers every dense contraction of the forward and data-gradient passes:
//   conv kxk stride 1 fwd/dgrad   : 1 input view, k*k taps
//   conv kxk stride 2 fwd         : 4 parity views of the input (strided TMA maps), k*k taps
//   conv kxk stride 2 dgrad       : 4 launches, one per output parity phase (strided TMA store map)
//   ConvTranspose 4x4 s2 fwd      : 4 launches (sub-pixel phases, 2x2 taps each)
//   ConvTranspose 4x4 s2 dgrad    : == 4x4 stride-2 conv of dy
//
// Structure (one 128-pixel x NT-channel output tile per CTA, several CTAs resident per SM):
//   warp 0    : TMA producer.  A tile = box {KC ch, bw, bh, bn} of the NHWC activation at the tap's
//               shifted coordinates (TMA zero-fills the halo => conv padding for free); B tile = weight slab.
//   warp 1    : TMEM allocator + single-thread tcgen05.mma issuer (K-major A and B, fp32 accum in TMEM).
//   warps 2-5 : epilogue.  tcgen05.ld -> (+bias) -> bf16 -> swizzled smem staging -> TMA store
//               (channel-slice / parity-phase aware through the tensor map), or fp32 direct store (logits).
#include "common.cuh"
#include "conv_internal.cuh"

namespace ubr {

constexpr int kMaxTaps = 49;
constexpr int kThreads = 192;

struct ConvGemmParams {
  CUtensorMap a_map[4];
  CUtensorMap b_map;
  CUtensorMap o_map[4];         // one store map per output phase (blockIdx.z): parity-phase launches are merged
  int nphase, tap_begin[5];     // phase z owns taps [tap_begin[z], tap_begin[z+1]) (nphase == 1: all ntaps)
  int zin;                      // 1: ONE CTA walks all nphase phases of its tile (own TMEM accumulator per phase, the
                                // operand tile stays in L2, set-up paid once); 0: one CTA per phase (blockIdx.z)
  int ntaps, nkc, kc, nt, stages;
  int mt, ntiles;               // pixel sub-tiles per CTA (1 or 2, sharing every weight stage) / total 128-pixel tiles
  int tiles_w, tiles_h, bw, bh, bn;
  uint32_t a_bytes, b_bytes, a_stride, b_stride;
  uint32_t idesc, tmem_cols;
  int out_mode;                 // 0: bf16 via TMA store, 1: fp32 direct (first out_cols columns)
  int o_box_c;                  // channels per TMA store box
  const float* bias;
  float* out_f32;
  float* stats;                 // out_mode 0: per-tile BN partials [tile][2][stats_c] (sum, sum of squares), or null
  int stats_c;
  int out_ld, out_cols, out_n, out_h, out_w;
  int fused, relu_pre, relu_post;   // out_mode 0: fused inference epilogue (FusedEpi); needs out_n/out_h/out_w
  const __nv_bfloat16* res;
  int res_ld;
  int8_t dh[kMaxTaps], dw[kMaxTaps];
  uint8_t amap[kMaxTaps], slab[kMaxTaps];
  // stats: reduction over the per-tile partial rows + BatchNorm finalize inside this launch (tree_reduce.cuh)
  int fin_kind;
  TreeRed tr;
  BnFinal bnf;
};

__global__ void __launch_bounds__(kThreads, 1) conv_gemm_kernel(const __grid_constant__ ConvGemmParams p) {
  pdl_launch_dependents();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // dynamic smem base is not guaranteed 1024-aligned: align by hand (host adds 1024 slack)
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t stage_bytes = (uint32_t)p.mt * p.a_stride + p.b_stride;
  const uint32_t staging_bytes = (uint32_t)p.mt * 128u * (uint32_t)p.nt * 2u;
  uint32_t data_bytes = (uint32_t)p.stages * stage_bytes;
  if (data_bytes < staging_bytes) data_bytes = staging_bytes;
  const uint32_t bar_base = smem_base + data_bytes;        // full[stages], empty[stages], accum, tmem ptr
  const uint32_t full0 = bar_base;
  const uint32_t empty0 = bar_base + 8u * p.stages;
  const uint32_t accum_bar = bar_base + 16u * p.stages;
  const uint32_t tmem_slot = accum_bar + 8u;
  float* s_part = reinterpret_cast<float*>(smem_raw + (tmem_slot + 8u - smem_u32(smem_raw)));   // [4 warps][2][nt]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // tile coordinates: this CTA owns the 128-pixel tiles blockIdx.x*mt .. +mt-1 (the last CTA may own one only)
  const int tile0 = blockIdx.x * p.mt;
  const int nsub = (tile0 + p.mt <= p.ntiles) ? p.mt : (p.ntiles - tile0);
  int w0s[2], h0s[2], n0s[2];
#pragma unroll
  for (int s = 0; s < 2; ++s) {
    const int tile = tile0 + (s < nsub ? s : 0);
    w0s[s] = (tile % p.tiles_w) * p.bw;
    h0s[s] = ((tile / p.tiles_w) % p.tiles_h) * p.bh;
    n0s[s] = (tile / (p.tiles_w * p.tiles_h)) * p.bn;
  }
  const int co0 = blockIdx.y * p.nt;
  const int z_lo_early = p.zin ? 0 : (int)blockIdx.z;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(full0 + 8u * s, 1);
      mbar_init(empty0 + 8u * s, 1);
    }
    mbar_init(accum_bar, 1);
    fence_mbar_init();
    prefetch_tmap(&p.a_map[0]);
    prefetch_tmap(&p.b_map);
    if (p.out_mode == 0) prefetch_tmap(&p.o_map[z_lo_early]);
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
  pdl_wait();                    // barriers, TMEM and tensor maps are set up; the operands belong to earlier kernels

  const int z_lo = p.zin ? 0 : (int)blockIdx.z, z_hi = p.zin ? p.nphase : (int)blockIdx.z + 1;
  const int tap_lo = p.tap_begin[z_lo], tap_hi = p.tap_begin[z_hi];

  if (warp == 0) {
    if (lane == 0) {
      // ================= TMA producer =================
      int it = 0;
      for (int t = tap_lo; t < tap_hi; ++t) {
        const CUtensorMap* am = &p.a_map[p.amap[t]];
        for (int kc = 0; kc < p.nkc; ++kc, ++it) {
          const int s = it % p.stages;
          const uint32_t ph = (uint32_t)(it / p.stages) & 1u;
          mbar_wait(empty0 + 8u * s, ph ^ 1u);
          const uint32_t a_dst = smem_base + (uint32_t)s * stage_bytes;
          const uint32_t b_dst = a_dst + (uint32_t)p.mt * p.a_stride;
          mbar_arrive_expect_tx(full0 + 8u * s, (uint32_t)nsub * p.a_bytes + p.b_bytes);
          for (int u = 0; u < nsub; ++u)
            tma_load_4d(a_dst + (uint32_t)u * p.a_stride, am, full0 + 8u * s, kc * p.kc, w0s[u] + p.dw[t],
                        h0s[u] + p.dh[t], n0s[u]);
          tma_load_3d(b_dst, &p.b_map, full0 + 8u * s, kc * p.kc, co0, (int)p.slab[t]);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ================= MMA issuer =================
      const uint32_t row_bytes = (uint32_t)p.kc * 2u;
      const uint32_t layout = swizzle_layout_code((int)row_bytes);
      const uint32_t sbo = 8u * row_bytes;
      const int ksteps = p.kc / 16;
      int it = 0;
      for (int z = z_lo; z < z_hi; ++z) {
        const uint32_t acc_z = (uint32_t)((z - z_lo) * p.mt * p.nt);      // this phase's accumulator columns
        const int iters_z = (p.tap_begin[z + 1] - p.tap_begin[z]) * p.nkc;
        for (int iz = 0; iz < iters_z; ++iz, ++it) {
          const int s = it % p.stages;
          const uint32_t ph = (uint32_t)(it / p.stages) & 1u;
          mbar_wait(full0 + 8u * s, ph);
          tc_fence_after();
          const uint32_t a_src = smem_base + (uint32_t)s * stage_bytes;
          const uint32_t b_src = a_src + (uint32_t)p.mt * p.a_stride;
          const uint64_t bdesc = make_smem_desc(b_src, 0, sbo, layout);
          for (int u = 0; u < nsub; ++u) {
            // every pixel sub-tile multiplies the SAME weight stage (halves the weight bytes pulled through L2 per FLOP)
            const uint64_t adesc = make_smem_desc(a_src + (uint32_t)u * p.a_stride, 0, sbo, layout);
            for (int k = 0; k < ksteps; ++k) {
              // advance 16 bf16 = 32 bytes along K inside the swizzle row: +2 in the (addr>>4) field
              tc_mma_bf16(tmem_base + acc_z + (uint32_t)(u * p.nt), adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k),
                          p.idesc, (iz > 0 || k > 0) ? 1u : 0u);
            }
          }
          tc_commit(empty0 + 8u * s);          // frees the smem stage when these MMAs retire
        }
      }
      tc_commit(accum_bar);                  // accumulator complete
    }
  } else {
    // ================= epilogue (warps 2..5) =================
    const int quad = warp & 3;               // TMEM lane quadrant this warp may access
    const int row = quad * 32 + lane;        // tile-local pixel
    mbar_wait(accum_bar, 0);
    tc_fence_after();
    for (int z = z_lo; z < z_hi; ++z) {
    if (z > z_lo) {
      // the staging tiles are reused by the next phase: its TMA stores must have read them first
      if (warp == 2 && lane == 0) tma_store_wait_read();
      asm volatile("bar.sync 1, 128;" ::: "memory");
    }
    for (int u = 0; u < nsub; ++u) {
    const int w0 = w0s[u], h0 = h0s[u], n0 = n0s[u];
    const uint32_t t_row = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(((z - z_lo) * p.mt + u) * p.nt);
    const uint32_t stage_base = smem_base + (uint32_t)u * 128u * (uint32_t)p.nt * 2u;   // own staging per sub-tile
    if (p.out_mode == 0) {
      const int bc = p.o_box_c;                         // channels per staging block
      const uint32_t rb = (uint32_t)bc * 2u;            // staging row bytes (32/64/128)
      const uint32_t xr = rb == 128 ? (uint32_t)(row & 7) : (rb == 64 ? (uint32_t)((row >> 1) & 3)
                                                                       : (uint32_t)((row >> 2) & 1));
      const __nv_bfloat16* res_row = nullptr;           // fused epilogue: this thread's pixel of the residual
      if (p.fused && p.res) {
        const int mw = row % p.bw, mh = (row / p.bw) % p.bh, mn = row / (p.bw * p.bh);
        const int pw_ = w0 + mw, ph_ = h0 + mh, pn_ = n0 + mn;
        if (pw_ < p.out_w && ph_ < p.out_h && pn_ < p.out_n)
          res_row = p.res + ((size_t)((size_t)pn_ * p.out_h + ph_) * p.out_w + pw_) * (size_t)p.res_ld + co0;
      }
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
        if (p.fused) fused_epi16(f, p.relu_pre, res_row ? res_row + c : nullptr, p.relu_post);
        uint32_t pk[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) pk[j] = pack_bf16x2(f[2 * j], f[2 * j + 1]);
        const int blk = c / bc;
        const uint32_t q = (uint32_t)((c % bc) >> 3);   // 16-byte chunk index inside the row
        const uint32_t base = stage_base + (uint32_t)blk * 128u * rb + (uint32_t)row * rb;
        const uint32_t a0 = base + ((q ^ xr) << 4);
        const uint32_t a1 = base + (((q + 1u) ^ xr) << 4);
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a0), "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3]) : "memory");
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a1), "r"(pk[4]), "r"(pk[5]), "r"(pk[6]), "r"(pk[7]) : "memory");
        if (p.stats) {
          // BatchNorm statistics of the STORED (bf16-rounded) values, like the stand-alone bn_stats kernel
          float g[16], q[16];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float2 t = unpack_bf16x2(pk[j]);
            g[2 * j] = t.x; g[2 * j + 1] = t.y;
            q[2 * j] = t.x * t.x; q[2 * j + 1] = t.y * t.y;
          }
          const float rs = warp_reduce_scatter16(g, lane), rq = warp_reduce_scatter16(q, lane);
          if ((lane & 1) == 0) {
            s_part[(quad * 2 + 0) * p.nt + c + (lane >> 1)] = rs;
            s_part[(quad * 2 + 1) * p.nt + c + (lane >> 1)] = rq;
          }
        }
      }
      fence_proxy_async_smem();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (p.stats) {
        // fixed-order sum over the four epilogue warps -> one partial row per tile (deterministic)
        const int e = (int)threadIdx.x - 64;
        for (int i = e; i < 2 * p.nt; i += 128) {
          const int k = i / p.nt, ch = i % p.nt;
          const float v = s_part[(0 * 2 + k) * p.nt + ch] + s_part[(1 * 2 + k) * p.nt + ch] +
                          s_part[(2 * 2 + k) * p.nt + ch] + s_part[(3 * 2 + k) * p.nt + ch];
          p.stats[((size_t)(tile0 + u) * 2 + k) * (size_t)p.stats_c + co0 + ch] = v;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");       // s_part is reused by the next sub-tile
      }
      if (warp == 2 && lane == 0) {
        const int nblk = p.nt / bc;
        for (int b = 0; b < nblk; ++b)
          tma_store_4d(&p.o_map[z], stage_base + (uint32_t)b * 128u * rb, co0 + b * bc, w0, h0, n0);
        tma_store_commit();
        if (u == nsub - 1 && z == z_hi - 1) tma_store_wait_all();
      }
    } else {
      // fp32 direct store (network logits): few valid columns, pixel-major rows
      const int mw = row % p.bw, mh = (row / p.bw) % p.bh, mn = row / (p.bw * p.bh);
      const int pw_ = w0 + mw, ph_ = h0 + mh, pn_ = n0 + mn;
      const bool valid = pw_ < p.out_w && ph_ < p.out_h && pn_ < p.out_n;
      float* dst = p.out_f32 + ((size_t)((size_t)pn_ * p.out_h + ph_) * p.out_w + pw_) * (size_t)p.out_ld;
      for (int c = 0; c < p.nt; c += 16) {
        uint32_t v[16];
        tmem_ld16(t_row + (uint32_t)c, v);
        tmem_ld_wait();
        if (valid) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int col = co0 + c + j;
            if (col < p.out_cols) dst[col] = __uint_as_float(v[j]) + (p.bias ? __ldg(p.bias + col) : 0.f);
          }
        }
      }
    }
    }   // sub-tiles
    }   // phases
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc_dyn(tmem_base, p.tmem_cols);
  }
  if (p.fin_kind) {
    // this CTA wrote the partial rows tile0 .. tile0+mt-1 of its channel tile blockIdx.y
    double* ts = reinterpret_cast<double*>(smem_raw + (smem_base - smem_u32(smem_raw)));     // the idle stage buffers
    if (tree_arrive(p.tr, tile0, p.mt, (int)gridDim.y, ts)) tree_finish_bn(ts, p.bnf);
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
struct TileShape { int bw, bh, bn; };
static int out_box_c(int nt) { return nt >= 64 ? 64 : nt; }

static TileShape pick_tile(int n, int h, int w) {
  // 128 pixels per tile = bw x bh x bn (powers of two); maximise useful fraction, prefer wide rows
  TileShape best{128, 1, 1};
  double best_eff = -1.0;
  for (int bw = 128; bw >= 1; bw >>= 1) {
    for (int bh = 128 / bw; bh >= 1; bh >>= 1) {
      const int bn = 128 / (bw * bh);
      const double eff = ((double)w / (ceil_div(w, bw) * bw)) * ((double)h / (ceil_div(h, bh) * bh)) *
                         ((double)n / (ceil_div(n, bn) * bn));
      // prefer larger bw, then larger bh (fewer images per tile) on ties
      if (eff > best_eff + 1e-9) { best_eff = eff; best = TileShape{bw, bh, bn}; }
    }
  }
  return best;
}

static int check_view(const ubr_view& v, const char* name) {
  UBR_REQUIRE(v.ptr != nullptr, "%s: null pointer", name);
  UBR_REQUIRE(v.n > 0 && v.h > 0 && v.w > 0 && v.c > 0, "%s: bad extents n=%d h=%d w=%d c=%d", name, v.n, v.h, v.w, v.c);
  UBR_REQUIRE(v.ld >= v.c && (v.ld % 8) == 0, "%s: ld=%d must be >= c=%d and a multiple of 8", name, v.ld, v.c);
  UBR_REQUIRE(((uintptr_t)v.ptr & 15) == 0, "%s: pointer must be 16-byte aligned", name);
  return 0;
}

// plain NHWC map (C, W, H, N) over a view, box {box_c, bw, bh, bn}
static int make_act_map(CUtensorMap* m, const ubr_view& v, int box_c, TileShape t) {
  uint64_t dims[4] = {(uint64_t)v.c, (uint64_t)v.w, (uint64_t)v.h, (uint64_t)v.n};
  uint64_t str[4] = {2, (uint64_t)v.ld * 2, (uint64_t)v.w * v.ld * 2, (uint64_t)v.h * v.w * v.ld * 2};
  uint32_t box[4] = {(uint32_t)box_c, (uint32_t)t.bw, (uint32_t)t.bh, (uint32_t)t.bn};
  return make_tmap_bf16(m, v.ptr, 4, dims, str, box, box_c * 2);
}
// parity-phase map: pixels (2i+ph, 2j+pw) of the view, as a (C, W/2, H/2, N) tensor
static int make_phase_map(CUtensorMap* m, const ubr_view& v, int ph, int pw, int box_c, TileShape t) {
  uint64_t dims[4] = {(uint64_t)v.c, (uint64_t)(v.w / 2), (uint64_t)(v.h / 2), (uint64_t)v.n};
  uint64_t str[4] = {2, (uint64_t)v.ld * 4, (uint64_t)v.w * v.ld * 4, (uint64_t)v.h * v.w * v.ld * 2};
  uint32_t box[4] = {(uint32_t)box_c, (uint32_t)t.bw, (uint32_t)t.bh, (uint32_t)t.bn};
  const uint8_t* base = (const uint8_t*)v.ptr + ((size_t)ph * v.w + pw) * (size_t)v.ld * 2;
  return make_tmap_bf16(m, base, 4, dims, str, box, box_c * 2);
}

// output-channel tile: 128 where it divides N, else 64 (192/320/576-channel ASPP concat gradients), else all of N
static int pick_nt(int n_pad) { return (n_pad % 128 == 0) ? 128 : ((n_pad % 64 == 0) ? 64 : n_pad); }

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}

// generic launcher. `grid_view` = (n, h, w) of the output tile grid; a_maps built by caller.
static int launch_conv_gemm(ConvGemmParams& p, int grid_n, int grid_h, int grid_w, TileShape t, int k_ch,
                            int n_pad, const void* wpk, int nslabs, const TapSpec* taps, int ntaps,
                            cudaStream_t stream) {
  UBR_REQUIRE(ntaps >= 1 && ntaps <= kMaxTaps, "conv_gemm: ntaps=%d out of range", ntaps);
  UBR_REQUIRE(k_ch % 16 == 0 && n_pad % 16 == 0, "conv_gemm: channels must be multiples of 16 (k=%d n=%d)", k_ch, n_pad);
  p.kc = k_ch >= 64 ? 64 : k_ch;
  UBR_REQUIRE(p.kc == 16 || p.kc == 32 || p.kc == 64, "conv_gemm: unsupported K chunk %d", p.kc);
  UBR_REQUIRE(k_ch % p.kc == 0, "conv_gemm: K=%d not a multiple of chunk %d", k_ch, p.kc);
  p.nkc = k_ch / p.kc;
  p.nt = pick_nt(n_pad);
  UBR_REQUIRE(p.nt <= 128 && n_pad % p.nt == 0, "conv_gemm: N=%d not a multiple of tile %d", n_pad, p.nt);
  // Tile shape per CTA: mt pixel sub-tiles of 128 x nt channels.  The kernel is bound by the bytes it pulls through L2
  // per FLOP (a 128x128 tile streams 32 KB per 2.1 MFLOP k-step = 64 FLOP/B); two pixel sub-tiles sharing one weight
  // stage give 87 FLOP/B.  Used when it still leaves a full wave of CTAs.
  const int tiles_total = ceil_div(grid_w, t.bw) * ceil_div(grid_h, t.bh) * ceil_div(grid_n, t.bn);
  const int sms = sm_count();
  static const int force_mt = env_int("UBR_GEMM_MT", 0);
  p.mt = 1;
  if (p.out_mode == 0 && (tiles_total / 2) * (n_pad / p.nt) >= sms) p.mt = 2;
  if (force_mt == 1 || force_mt == 2) p.mt = force_mt;
  // parity-phase launches (stride-2 data gradients, ConvTranspose forwards: 1-4 taps per phase): one CTA walks all
  // phases of its tile -- barrier / TMEM / tensor-map set-up paid once per 9-16 taps instead of once per 1-4 -- when the
  // accumulators of all phases fit beside a second CTA (<= 256 TMEM columns, i.e. N tiles of <= 32 channels) and a full
  // wave of CTAs remains.  Measured (batch 32): enc_layer2.res1 data gradients (64 -> 32 channels, 128^2 -> 256^2) 0.177
  // -> 0.120 ms; with 64-channel N tiles the single pixel sub-tile and the serial four-phase epilogue lose more than
  // the set-up saves (0.065 -> 0.076 ms, ConvTranspose 128 -> 64: 0.085 -> 0.106 ms), so those keep one CTA per phase.
  p.zin = 0;
  static const int zin_on = env_int("UBR_GEMM_ZIN", 1);
  if (p.nphase > 1 && zin_on && p.nphase * p.mt * p.nt <= 256 && ceil_div(tiles_total, p.mt) * (n_pad / p.nt) >= sms)
    p.zin = 1;
  p.ntiles = tiles_total;
  p.ntaps = ntaps;
  if (p.nphase <= 1) { p.nphase = 1; p.tap_begin[0] = 0; p.tap_begin[1] = ntaps; }
  for (int i = 0; i < ntaps; ++i) {
    p.dh[i] = (int8_t)taps[i].dh; p.dw[i] = (int8_t)taps[i].dw;
    p.amap[i] = (uint8_t)taps[i].amap; p.slab[i] = (uint8_t)taps[i].slab;
  }
  p.bw = t.bw; p.bh = t.bh; p.bn = t.bn;
  p.tiles_w = ceil_div(grid_w, t.bw);
  p.tiles_h = ceil_div(grid_h, t.bh);
  const int tiles_n = ceil_div(grid_n, t.bn);
  p.a_bytes = 128u * p.kc * 2u;
  p.b_bytes = (uint32_t)p.nt * p.kc * 2u;
  p.a_stride = (p.a_bytes + 1023u) & ~1023u;
  p.b_stride = (p.b_bytes + 1023u) & ~1023u;
  const uint32_t stage = (uint32_t)p.mt * p.a_stride + p.b_stride;
  static const uint32_t smem_budget = (uint32_t)env_int("UBR_GEMM_SMEM_KB", 96) * 1024u;
  int stages = (int)(smem_budget / stage);
  if (stages > 8) stages = 8;
  if (stages < 2) stages = 2;
  if (stages > ntaps * p.nkc) stages = ntaps * p.nkc;
  p.stages = stages;
  p.idesc = make_idesc_bf16(128, p.nt, 0, 0);
  uint32_t cols = 32;
  while ((int)cols < (p.zin ? p.nphase : 1) * p.mt * p.nt) cols <<= 1;
  p.tmem_cols = cols;
  // weights: (K, N, slab) 3-D map, box {kc, nt, 1}
  {
    uint64_t dims[3] = {(uint64_t)k_ch, (uint64_t)n_pad, (uint64_t)nslabs};
    uint64_t str[3] = {2, (uint64_t)k_ch * 2, (uint64_t)k_ch * n_pad * 2};
    uint32_t box[3] = {(uint32_t)p.kc, (uint32_t)p.nt, 1};
    int rc = make_tmap_bf16(&p.b_map, wpk, 3, dims, str, box, p.kc * 2);
    if (rc) return rc;
  }
  uint32_t data = (uint32_t)stages * stage;
  const uint32_t staging = (uint32_t)p.mt * 128u * p.nt * 2u;
  if (data < staging) data = staging;
  const uint32_t smem = data + 16u * stages + 16u + 1024u + 4096u + 16u;      // + [4][2][nt] fp32 statistics scratch
  static bool attr_set[64] = {};             // the opt-in is per device
  int dev_id = 0;
  UBR_CUDA_OK(cudaGetDevice(&dev_id));
  if (!attr_set[dev_id & 63]) {
    UBR_CUDA_OK(cudaFuncSetAttribute(conv_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_set[dev_id & 63] = true;
  }
  dim3 grid((unsigned)ceil_div(p.tiles_w * p.tiles_h * tiles_n, p.mt), (unsigned)(n_pad / p.nt), (unsigned)(p.zin ? 1 : p.nphase));
  ::ubr::Launch(grid, kThreads, smem, stream)(conv_gemm_kernel, p);
  UBR_LAUNCH_CHECK("conv_gemm_kernel");
  return 0;
}


}  // namespace ubr

using namespace ubr;

extern "C" {

static int conv2d_fwd_impl(ubr_view x, const void* wpk, const float* bias, int out_ch, int ksize, int stride, ubr_view y,
                           float* y_f32, int y_f32_ld, void* stream, float* stats, int* stats_rows,
                           const FusedEpi* epi = nullptr, const FusedFinal* fin = nullptr);

int ubr_conv2d_fwd_fused(ubr_view x, const void* wpk, const float* bias, int out_ch, int ksize, int stride, int relu_pre,
                         ubr_view res, int relu_post, ubr_view y, void* stream) {
  FusedEpi e{relu_pre, relu_post, res.ptr, res.ld};
  if (res.ptr) {
    UBR_REQUIRE(res.n == y.n && res.h == y.h && res.w == y.w && res.c == out_ch && res.ld % 8 == 0 &&
                    ((uintptr_t)res.ptr & 15) == 0,
                "conv2d_fwd_fused: residual view must match the output (n,h,w,c) with 16-byte aligned rows");
  }
  return conv2d_fwd_impl(x, wpk, bias, out_ch, ksize, stride, y, nullptr, 0, stream, nullptr, nullptr, &e);
}

int ubr_conv2d_fwd(ubr_view x, const void* wpk, const float* bias, int out_ch, int ksize, int stride, ubr_view y,
                   float* y_f32, int y_f32_ld, void* stream) {
  return conv2d_fwd_impl(x, wpk, bias, out_ch, ksize, stride, y, y_f32, y_f32_ld, stream, nullptr, nullptr);
}

int ubr_conv2d_fwd_stats(ubr_view x, const void* wpk, const float* bias, int out_ch, int ksize, int stride, ubr_view y,
                         void* bn_workspace, int* stats_rows, void* stream) {
  UBR_REQUIRE(bn_workspace != nullptr && stats_rows != nullptr, "conv2d_fwd_stats: null workspace / rows pointer");
  return conv2d_fwd_impl(x, wpk, bias, out_ch, ksize, stride, y, nullptr, 0, stream, (float*)bn_workspace, stats_rows);
}

// conv + BatchNorm statistics + finalize in ONE launch (the partial rows are reduced by the last CTAs to arrive).
// Returns 1 (nothing launched) where the conv kernel has no fused-statistics epilogue.
int ubr_conv2d_fwd_bn(ubr_view x, const void* wpk, const float* bias, int out_ch, int ksize, int stride, ubr_view y,
                      void* bn_workspace, void* tree, const float* gamma, const float* beta, float* running_mean,
                      float* running_var, float momentum, float eps, float* scale, float* shift, float* mean,
                      float* invstd, void* stream) {
  UBR_REQUIRE(bn_workspace && tree && scale && shift, "conv2d_fwd_bn: null pointer");
  FusedFinal f;
  memset(&f, 0, sizeof(f));
  f.tree = tree;
  f.kind = 1;
  f.bn = BnFinal{gamma, beta, running_mean, running_var, momentum, eps, (double)y.n * y.h * y.w, scale, shift, mean, invstd,
                 out_ch, out_ch};
  int rows = 0;
  return conv2d_fwd_impl(x, wpk, bias, out_ch, ksize, stride, y, nullptr, 0, stream, (float*)bn_workspace, &rows, nullptr, &f);
}

static int conv2d_fwd_impl(ubr_view x, const void* wpk, const float* bias, int out_ch, int ksize, int stride, ubr_view y,
                           float* y_f32, int y_f32_ld, void* stream, float* stats, int* stats_rows, const FusedEpi* epi,
                           const FusedFinal* fin) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(!(epi && (y_f32 || stats)), "conv2d_fwd: the fused epilogue writes bf16 and emits no statistics");
  if (int rc = check_view(x, "conv2d_fwd.x")) return rc;
  UBR_REQUIRE(ksize == 1 || ksize == 3 || ksize == 7, "conv2d_fwd: ksize %d unsupported", ksize);
  UBR_REQUIRE(stride == 1 || stride == 2, "conv2d_fwd: stride %d unsupported", stride);
  UBR_REQUIRE(x.c % 16 == 0, "conv2d_fwd: in_ch=%d must be a multiple of 16", x.c);
  UBR_REQUIRE(wpk != nullptr, "conv2d_fwd: null weights");
  const int pad = (ksize - 1) / 2;
  const int n_pad = (out_ch + 15) / 16 * 16;
  int ho, wo;
  if (stride == 1) { ho = x.h; wo = x.w; }
  else {
    UBR_REQUIRE(x.h % 2 == 0 && x.w % 2 == 0, "conv2d_fwd: stride 2 needs even H,W (got %dx%d)", x.h, x.w);
    ho = x.h / 2; wo = x.w / 2;
  }
  ConvGemmParams p;
  memset(&p, 0, sizeof(p));
  TileShape t = pick_tile(x.n, ho, wo);
  const int kc = x.c >= 64 ? 64 : x.c;
  TapSpec taps[kMaxTaps];
  int nt = 0;
  if (stride == 1) {
    if (int rc = make_act_map(&p.a_map[0], x, kc, t)) return rc;
    for (int r = 0; r < ksize; ++r)
      for (int s = 0; s < ksize; ++s) taps[nt++] = TapSpec{r - pad, s - pad, 0, r * ksize + s};
  } else {
    for (int ph = 0; ph < 2; ++ph)
      for (int pw = 0; pw < 2; ++pw)
        if (int rc = make_phase_map(&p.a_map[ph * 2 + pw], x, ph, pw, kc, t)) return rc;
    for (int r = 0; r < ksize; ++r)
      for (int s = 0; s < ksize; ++s) {
        const int oh = r - pad, ow = s - pad;                 // input row = 2*ho + oh
        const int ph = oh & 1, pw = ow & 1;                   // parity (two's complement & works for -1)
        const int dh = (oh - ph) / 2, dw = (ow - pw) / 2;     // floor division
        taps[nt++] = TapSpec{dh, dw, ph * 2 + pw, r * ksize + s};
      }
  }
  if (stride == 1) {
    if (!y_f32) {
      if (int rc = check_view(y, "conv2d_fwd.y")) return rc;
      UBR_REQUIRE(y.n == x.n && y.h == ho && y.w == wo && y.c == out_ch, "conv2d_fwd: output view mismatch");
      UBR_REQUIRE(out_ch % 16 == 0, "conv2d_fwd: bf16 output needs out_ch %% 16 == 0");
    } else {
      UBR_REQUIRE(y_f32_ld >= out_ch, "conv2d_fwd: y_f32_ld too small");
    }
    const int hrc = conv_halo_try(x, wpk, ksize * ksize, n_pad, taps, nt, bias, y_f32 ? nullptr : &y, y_f32, y_f32_ld,
                                  out_ch, (cudaStream_t)stream, 0, stats, stats_rows, 0, epi, nullptr, fin);
    if (hrc <= 0) return hrc;
  }
  if (stats) {
    // fused statistics in the implicit-GEMM epilogue: full tiles only (a zero-filled pixel would still add the bias)
    const int tiles = ceil_div(wo, t.bw) * ceil_div(ho, t.bh) * ceil_div(x.n, t.bn);
    if (y_f32 || wo % t.bw || ho % t.bh || x.n % t.bn || tiles > kMaxStatsRows || out_ch % 16) return 1;
    p.stats = stats;
    p.stats_c = out_ch;
    *stats_rows = tiles;
    if (fin) {
      p.fin_kind = 1;
      p.tr = make_tree(fin->tree, stats, tiles, 2 * out_ch);
      p.bnf = fin->bn;
    }
  }
  p.bias = bias;
  if (y_f32) {
    p.out_mode = 1; p.out_f32 = y_f32; p.out_ld = y_f32_ld; p.out_cols = out_ch;
    p.out_n = x.n; p.out_h = ho; p.out_w = wo;
    UBR_REQUIRE(y_f32_ld >= out_ch, "conv2d_fwd: y_f32_ld too small");
  } else {
    if (int rc = check_view(y, "conv2d_fwd.y")) return rc;
    UBR_REQUIRE(y.n == x.n && y.h == ho && y.w == wo && y.c == out_ch, "conv2d_fwd: output view mismatch");
    UBR_REQUIRE(out_ch % 16 == 0, "conv2d_fwd: bf16 output needs out_ch %% 16 == 0");
    p.out_mode = 0;
    p.o_box_c = out_box_c(pick_nt(n_pad));
    if (int rc = make_act_map(&p.o_map[0], y, p.o_box_c, t)) return rc;
    if (epi) {
      p.fused = 1; p.relu_pre = epi->relu_pre; p.relu_post = epi->relu_post;
      p.res = (const __nv_bfloat16*)epi->res; p.res_ld = epi->res_ld;
      p.out_n = x.n; p.out_h = ho; p.out_w = wo;
    }
  }
  return launch_conv_gemm(p, x.n, ho, wo, t, x.c, n_pad, wpk, ksize * ksize, taps, nt, (cudaStream_t)stream);
}

static int conv2d_dgrad_impl(ubr_view dy, const void* wpk, int in_ch, int ksize, int stride, ubr_view dx, int accumulate,
                             void* stream, const BnReduce* bnr = nullptr, float* stats = nullptr, int* stats_rows = nullptr,
                             const FusedFinal* fin = nullptr);

// ubr_conv2d_dgrad_bnreduce whose launch also finishes the two sums (sums[2][in_ch] behind the partial rows of
// bn_workspace, dbeta / dgamma): follow it with ubr_bn_act_bwd_apply_final
int ubr_conv2d_dgrad_bnreduce_final(ubr_view dy, const void* wpk, int in_ch, int ksize, int stride, ubr_view dx,
                                    ubr_view y_pre, const float* scale, const float* shift, const float* mean,
                                    const float* invstd, int relu, void* bn_workspace, void* tree, float* dgamma,
                                    float* dbeta, int accumulate, void* stream) {
  UBR_REQUIRE(bn_workspace && tree && scale && shift && mean && invstd, "conv2d_dgrad_bnreduce_final: null pointer");
  if (stride != 1) return 1;
  UBR_REQUIRE(y_pre.ptr && y_pre.n == dx.n && y_pre.h == dx.h && y_pre.w == dx.w && y_pre.c == in_ch && y_pre.ld % 8 == 0 &&
                  ((uintptr_t)y_pre.ptr & 15) == 0,
              "conv2d_dgrad_bnreduce_final: y_pre must match dx (n,h,w,c) with 16-byte aligned pixels");
  BnReduce b{y_pre.ptr, y_pre.ld, relu, scale, shift, mean, invstd};
  FusedFinal f;
  memset(&f, 0, sizeof(f));
  f.tree = tree;
  f.kind = 2;
  // same place as the stand-alone reduce passes keep their final sums
  f.sums = SumsFinal{(float*)bn_workspace + (size_t)kRedBlocks * 4 * in_ch, {dbeta, dgamma, nullptr, nullptr}, in_ch, 2, in_ch, accumulate};
  int rows = 0;
  return conv2d_dgrad_impl(dy, wpk, in_ch, ksize, stride, dx, 0, stream, &b, (float*)bn_workspace, &rows, &f);
}

int ubr_conv2d_dgrad_bnreduce(ubr_view dy, const void* wpk, int in_ch, int ksize, int stride, ubr_view dx, ubr_view y_pre,
                              const float* scale, const float* shift, const float* mean, const float* invstd, int relu,
                              void* bn_workspace, int* rows, void* stream) {
  UBR_REQUIRE(bn_workspace && rows && scale && shift && mean && invstd, "conv2d_dgrad_bnreduce: null pointer");
  if (stride != 1) return 1;
  UBR_REQUIRE(y_pre.ptr && y_pre.n == dx.n && y_pre.h == dx.h && y_pre.w == dx.w && y_pre.c == in_ch && y_pre.ld % 8 == 0 &&
                  ((uintptr_t)y_pre.ptr & 15) == 0,
              "conv2d_dgrad_bnreduce: y_pre must match dx (n,h,w,c) with 16-byte aligned pixels");
  BnReduce b{y_pre.ptr, y_pre.ld, relu, scale, shift, mean, invstd};
  return conv2d_dgrad_impl(dy, wpk, in_ch, ksize, stride, dx, 0, stream, &b, (float*)bn_workspace, rows);
}

int ubr_conv2d_dgrad(ubr_view dy, const void* wpk, int in_ch, int ksize, int stride, ubr_view dx, void* stream) {
  return conv2d_dgrad_impl(dy, wpk, in_ch, ksize, stride, dx, 0, stream);
}

int ubr_conv2d_dgrad_acc(ubr_view dy, const void* wpk, int in_ch, int ksize, int stride, ubr_view dx, void* stream) {
  return conv2d_dgrad_impl(dy, wpk, in_ch, ksize, stride, dx, 1, stream);
}

static int conv2d_dgrad_impl(ubr_view dy, const void* wpk, int in_ch, int ksize, int stride, ubr_view dx, int accumulate,
                             void* stream, const BnReduce* bnr, float* stats, int* stats_rows, const FusedFinal* fin) {
  if (int rc = ensure_device()) return rc;
  if (int rc = check_view(dy, "conv2d_dgrad.dy")) return rc;
  if (int rc = check_view(dx, "conv2d_dgrad.dx")) return rc;
  UBR_REQUIRE(ksize == 1 || ksize == 3 || ksize == 7, "conv2d_dgrad: ksize %d unsupported", ksize);
  UBR_REQUIRE(stride == 1 || stride == 2, "conv2d_dgrad: stride %d unsupported", stride);
  UBR_REQUIRE(dy.c % 16 == 0 && in_ch % 16 == 0, "conv2d_dgrad: channels must be multiples of 16");
  UBR_REQUIRE(dx.c == in_ch && dx.n == dy.n && dx.h == dy.h * stride && dx.w == dy.w * stride,
              "conv2d_dgrad: dx view mismatch");
  const int pad = (ksize - 1) / 2;
  const int kc = dy.c >= 64 ? 64 : dy.c;
  TileShape t = pick_tile(dy.n, dy.h, dy.w);
  if (stride == 1) {
    ConvGemmParams p;
    memset(&p, 0, sizeof(p));
    if (int rc = make_act_map(&p.a_map[0], dy, kc, t)) return rc;
    TapSpec taps[kMaxTaps];
    int nt = 0;
    // dx[p] = sum_{r,s} dy[p - (r-pad, s-pad)] * W[:, :, r, s]
    for (int r = 0; r < ksize; ++r)
      for (int s = 0; s < ksize; ++s) taps[nt++] = TapSpec{pad - r, pad - s, 0, r * ksize + s};
    {
      const int hrc = conv_halo_try(dy, wpk, ksize * ksize, in_ch, taps, nt, nullptr, &dx, nullptr, 0, in_ch,
                                    (cudaStream_t)stream, accumulate, stats, stats_rows, 0, nullptr, bnr, fin);
      if (hrc <= 0) return hrc;
      if (bnr) return 1;                     // the fused reduction lives in the halo-resident epilogue only
    }
    p.out_mode = 0;
    p.o_box_c = out_box_c(pick_nt(in_ch));
    if (int rc = make_act_map(&p.o_map[0], dx, p.o_box_c, t)) return rc;
    if (accumulate) {
      // dx += dgrad: the old dx is the residual of the fused epilogue (in place: every tile reads its own pixels
      // before its TMA store, tiles are disjoint)
      p.fused = 1; p.relu_pre = 0; p.relu_post = 0;
      p.res = (const __nv_bfloat16*)dx.ptr; p.res_ld = dx.ld;
      p.out_n = dx.n; p.out_h = dx.h; p.out_w = dx.w;
    }
    return launch_conv_gemm(p, dy.n, dy.h, dy.w, t, dy.c, in_ch, wpk, ksize * ksize, taps, nt, (cudaStream_t)stream);
  }
  if (accumulate) return 1;
  // stride 2: input row x = 2*o + r - pad.  For output parity (ph,pw) only taps with (r-pad)&1 == ph contribute,
  // from dy row o = (x - (r-pad))/2 = i + (ph - (r-pad))/2.  The four parity phases are ONE launch (blockIdx.z = phase,
  // each with its own tap range and strided store map): 4x the CTAs for the small deep layers, 3 launches less.
  ConvGemmParams p;
  memset(&p, 0, sizeof(p));
  if (int rc = make_act_map(&p.a_map[0], dy, kc, t)) return rc;
  TapSpec taps[kMaxTaps];
  int nt = 0;
  p.out_mode = 0;
  p.o_box_c = out_box_c(pick_nt(in_ch));
  p.nphase = 4;
  for (int ph = 0; ph < 2; ++ph)
    for (int pw = 0; pw < 2; ++pw) {
      const int z = ph * 2 + pw;
      p.tap_begin[z] = nt;
      for (int r = 0; r < ksize; ++r)
        for (int s = 0; s < ksize; ++s) {
          const int oh = r - pad, ow = s - pad;
          if ((oh & 1) != ph || (ow & 1) != pw) continue;
          taps[nt++] = TapSpec{(ph - oh) / 2, (pw - ow) / 2, 0, r * ksize + s};
        }
      if (nt == p.tap_begin[z]) {
        // no tap reaches this phase (1x1 stride 2, odd pixels): gradient is zero there.
        // Handled by a zero weight slab: use tap 0 with slab index == ksize*ksize (caller packs +1 zero slab)
        taps[nt++] = TapSpec{0, 0, 0, ksize * ksize};
      }
      if (int rc = make_phase_map(&p.o_map[z], dx, ph, pw, p.o_box_c, t)) return rc;
    }
  p.tap_begin[4] = nt;
  return launch_conv_gemm(p, dy.n, dy.h, dy.w, t, dy.c, in_ch, wpk, ksize * ksize + 1, taps, nt, (cudaStream_t)stream);
}

// 3x3 stride-1 DILATED conv (padding == dilation): the ASPP branches of models/ASPP_ResNet.py:207-214.  Taps are the
// dense 3x3 taps scaled by the dilation; the TMA zero-fill provides the (dilation-wide) padding.
static int conv2d_dilated_fwd_impl(ubr_view x, const void* wpk, const float* bias, int out_ch, int dilation, ubr_view y,
                                   int relu, void* stream);

int ubr_conv2d_dilated_fwd(ubr_view x, const void* wpk, const float* bias, int out_ch, int dilation, ubr_view y,
                           void* stream) {
  return conv2d_dilated_fwd_impl(x, wpk, bias, out_ch, dilation, y, 0, stream);
}

int ubr_conv2d_dilated_fwd_relu(ubr_view x, const void* wpk, const float* bias, int out_ch, int dilation, ubr_view y,
                                void* stream) {
  return conv2d_dilated_fwd_impl(x, wpk, bias, out_ch, dilation, y, 1, stream);
}

static int conv2d_dilated_fwd_impl(ubr_view x, const void* wpk, const float* bias, int out_ch, int dilation, ubr_view y,
                                   int relu, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = check_view(x, "conv2d_dilated_fwd.x")) return rc;
  if (int rc = check_view(y, "conv2d_dilated_fwd.y")) return rc;
  UBR_REQUIRE(dilation >= 1 && dilation <= 16, "conv2d_dilated_fwd: dilation %d unsupported", dilation);
  UBR_REQUIRE(x.c % 16 == 0 && out_ch % 16 == 0, "conv2d_dilated_fwd: channels must be multiples of 16");
  UBR_REQUIRE(wpk != nullptr, "conv2d_dilated_fwd: null weights");
  UBR_REQUIRE(y.n == x.n && y.h == x.h && y.w == x.w && y.c == out_ch, "conv2d_dilated_fwd: output view mismatch");
  ConvGemmParams p;
  memset(&p, 0, sizeof(p));
  TileShape t = pick_tile(x.n, x.h, x.w);
  const int kc = x.c >= 64 ? 64 : x.c;
  if (int rc = make_act_map(&p.a_map[0], x, kc, t)) return rc;
  TapSpec taps[9];
  int nt = 0;
  for (int r = 0; r < 3; ++r)
    for (int s = 0; s < 3; ++s) taps[nt++] = TapSpec{(r - 1) * dilation, (s - 1) * dilation, 0, r * 3 + s};
  p.bias = bias;
  p.out_mode = 0;
  p.o_box_c = out_box_c(pick_nt(out_ch));
  if (int rc = make_act_map(&p.o_map[0], y, p.o_box_c, t)) return rc;
  if (relu) { p.fused = 1; p.relu_pre = 1; }               // inference form: BatchNorm folded into (wpk, bias)
  return launch_conv_gemm(p, x.n, x.h, x.w, t, x.c, out_ch, wpk, 9, taps, nt, (cudaStream_t)stream);
}

int ubr_conv2d_dilated_dgrad(ubr_view dy, const void* wpk, int in_ch, int dilation, ubr_view dx, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = check_view(dy, "conv2d_dilated_dgrad.dy")) return rc;
  if (int rc = check_view(dx, "conv2d_dilated_dgrad.dx")) return rc;
  UBR_REQUIRE(dilation >= 1 && dilation <= 16, "conv2d_dilated_dgrad: dilation %d unsupported", dilation);
  UBR_REQUIRE(dy.c % 16 == 0 && in_ch % 16 == 0, "conv2d_dilated_dgrad: channels must be multiples of 16");
  UBR_REQUIRE(dx.c == in_ch && dx.n == dy.n && dx.h == dy.h && dx.w == dy.w, "conv2d_dilated_dgrad: dx view mismatch");
  ConvGemmParams p;
  memset(&p, 0, sizeof(p));
  TileShape t = pick_tile(dy.n, dy.h, dy.w);
  const int kc = dy.c >= 64 ? 64 : dy.c;
  if (int rc = make_act_map(&p.a_map[0], dy, kc, t)) return rc;
  TapSpec taps[9];
  int nt = 0;
  // dx[p] = sum_{r,s} dy[p - d*(r-1, s-1)] * W[:, :, r, s]
  for (int r = 0; r < 3; ++r)
    for (int s = 0; s < 3; ++s) taps[nt++] = TapSpec{(1 - r) * dilation, (1 - s) * dilation, 0, r * 3 + s};
  p.out_mode = 0;
  p.o_box_c = out_box_c(pick_nt(in_ch));
  if (int rc = make_act_map(&p.o_map[0], dx, p.o_box_c, t)) return rc;
  return launch_conv_gemm(p, dy.n, dy.h, dy.w, t, dy.c, in_ch, wpk, 9, taps, nt, (cudaStream_t)stream);
}

int ubr_deconv4s2_fwd(ubr_view x, const void* wpk, int out_ch, ubr_view y, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = check_view(x, "deconv_fwd.x")) return rc;
  if (int rc = check_view(y, "deconv_fwd.y")) return rc;
  UBR_REQUIRE(x.c % 16 == 0 && out_ch % 16 == 0, "deconv_fwd: channels must be multiples of 16");
  UBR_REQUIRE(y.c == out_ch && y.n == x.n && y.h == 2 * x.h && y.w == 2 * x.w, "deconv_fwd: output view mismatch");
  const int kc = x.c >= 64 ? 64 : x.c;
  TileShape t = pick_tile(x.n, x.h, x.w);
  // out row o = 2*i - 1 + kh  =>  for o = 2*j + ph: i = j + (ph + 1 - kh)/2, kh == (ph+1) mod 2.
  // The four sub-pixel phases (2x2 taps each) are ONE launch: blockIdx.z = phase.
  ConvGemmParams p;
  memset(&p, 0, sizeof(p));
  if (int rc = make_act_map(&p.a_map[0], x, kc, t)) return rc;
  TapSpec taps[16];
  int nt = 0;
  p.out_mode = 0;
  p.o_box_c = out_box_c(pick_nt(out_ch));
  p.nphase = 4;
  for (int ph = 0; ph < 2; ++ph)
    for (int pw = 0; pw < 2; ++pw) {
      const int z = ph * 2 + pw;
      p.tap_begin[z] = nt;
      for (int kh = 0; kh < 4; ++kh)
        for (int kw = 0; kw < 4; ++kw) {
          if (((ph + 1 - kh) & 1) || ((pw + 1 - kw) & 1)) continue;
          taps[nt++] = TapSpec{(ph + 1 - kh) / 2, (pw + 1 - kw) / 2, 0, kh * 4 + kw};
        }
      if (int rc = make_phase_map(&p.o_map[z], y, ph, pw, p.o_box_c, t)) return rc;
    }
  p.tap_begin[4] = nt;
  return launch_conv_gemm(p, x.n, x.h, x.w, t, x.c, out_ch, wpk, 16, taps, nt, (cudaStream_t)stream);
}

int ubr_deconv4s2_fwd_subpixel(ubr_view x, const void* wpk3, int out_ch, ubr_view y, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = check_view(x, "deconv_fwd_subpixel.x")) return rc;
  if (int rc = check_view(y, "deconv_fwd_subpixel.y")) return rc;
  UBR_REQUIRE(x.c % 16 == 0 && out_ch % 16 == 0, "deconv_fwd_subpixel: channels must be multiples of 16");
  UBR_REQUIRE(y.c == out_ch && y.n == x.n && y.h == 2 * x.h && y.w == 2 * x.w, "deconv_fwd_subpixel: output view mismatch");
  UBR_REQUIRE(wpk3 != nullptr, "deconv_fwd_subpixel: null weights");
  TapSpec taps[9];
  int nt = 0;
  for (int r = 0; r < 3; ++r)
    for (int s = 0; s < 3; ++s) taps[nt++] = TapSpec{r - 1, s - 1, 0, r * 3 + s};
  return conv_halo_try(x, wpk3, 9, 4 * out_ch, taps, 9, nullptr, &y, nullptr, 0, 4 * out_ch, (cudaStream_t)stream, 0, nullptr,
                       nullptr, out_ch);
}

int ubr_deconv4s2_dgrad(ubr_view dy, const void* wpk, int in_ch, ubr_view dx, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = check_view(dy, "deconv_dgrad.dy")) return rc;
  if (int rc = check_view(dx, "deconv_dgrad.dx")) return rc;
  UBR_REQUIRE(dy.c % 16 == 0 && in_ch % 16 == 0, "deconv_dgrad: channels must be multiples of 16");
  UBR_REQUIRE(dy.h % 2 == 0 && dy.w % 2 == 0, "deconv_dgrad: dy must have even H,W");
  UBR_REQUIRE(dx.c == in_ch && dx.n == dy.n && dx.h * 2 == dy.h && dx.w * 2 == dy.w, "deconv_dgrad: dx view mismatch");
  // dx[i] = sum_{kh,kw} dy[2*i - 1 + kh] * W[ci, co, kh, kw]   -> a 4x4 stride-2 pad-1 conv of dy
  const int kc = dy.c >= 64 ? 64 : dy.c;
  TileShape t = pick_tile(dx.n, dx.h, dx.w);
  ConvGemmParams p;
  memset(&p, 0, sizeof(p));
  for (int ph = 0; ph < 2; ++ph)
    for (int pw = 0; pw < 2; ++pw)
      if (int rc = make_phase_map(&p.a_map[ph * 2 + pw], dy, ph, pw, kc, t)) return rc;
  TapSpec taps[16];
  int nt = 0;
  for (int kh = 0; kh < 4; ++kh)
    for (int kw = 0; kw < 4; ++kw) {
      const int oh = kh - 1, ow = kw - 1;
      const int ph = oh & 1, pw = ow & 1;
      taps[nt++] = TapSpec{(oh - ph) / 2, (ow - pw) / 2, ph * 2 + pw, kh * 4 + kw};
    }
  p.out_mode = 0;
  p.o_box_c = out_box_c(pick_nt(in_ch));
  if (int rc = make_act_map(&p.o_map[0], dx, p.o_box_c, t)) return rc;
  return launch_conv_gemm(p, dx.n, dx.h, dx.w, t, dy.c, in_ch, wpk, 16, taps, nt, (cudaStream_t)stream);
}

}  // extern "C"
