// conv3x3 -> ReLU -> conv3x3 chain of an eval-mode residual block in ONE kernel (sm_100a).
//
//   a1  = relu(conv1(x) + b1)                                   (BatchNorm folded into the weights / bias)
//   out = relu(relu(conv2(a1) + b2) + res)                      models/common_layers.py:39-58 in eval mode
//
// The unfused path (conv_halo.cu, two launches) writes a1 to HBM and reads it back.  Here a1 never leaves the SM:
//   * a persistent CTA owns FULL-WIDTH row bands (bh output rows x W columns) of one image.  Input rows stream
//     through a TMA ring exactly like conv_halo.cu (row-stationary issue, taps stacked along N);
//   * conv1's accumulators (TMEM) are drained by epilogue group 1, which applies bias + ReLU, rounds to bf16 and
//     writes the row into a SECOND shared-memory ring in the K-major swizzled layout the tensor core reads
//     (the same address-XOR the TMA applies: chunk ^= (addr >> 7) & mask), then fence.proxy.async + mbarrier;
//   * a second issuer warp runs conv2 over that ring; epilogue group 2 adds bias, ReLU, the residual, ReLU and stores.
// Halo recompute: a band of bh output rows needs a1 rows -1..bh (2 recomputed per band) and x rows -2..bh+1; a1 rows
// outside the image are conv2's zero padding and are written as zeros, the left/right padding columns of both rings
// are zeroed once.  Full-width rows mean there is no column halo.
// Shapes: W % 128 == 0, Cin in {16,32,64}, planes in {16,32,64}, (W/128) * (planes/16) <= 4 (TMEM: two rings of four
// row accumulators per 128-column block).  Everything else returns 1 and the caller uses the two-launch path.
#include "common.cuh"
#include "conv_internal.cuh"

namespace ubr {
namespace {

constexpr int kCThreads = 352;   // warp 0 TMA | 1 issuer conv1 | 2-5 epilogue a1 | 6-9 epilogue out | 10 issuer conv2
constexpr uint32_t kLead = 1024; // bytes before pixel 0 of a ring slot (pixel -1 = the last span of it)

struct ChainParams {
  CUtensorMap x_map;            // input view (C, W, H, N), box {cin, 128, 1, 1}
  CUtensorMap w1_map, w2_map;   // packed weights (K, N, slab), box {K, planes, 1}
  int cb, bh, bands, ntiles, H, W, planes;
  int s1, s2;                   // ring depths (x rows, a1 rows)
  uint32_t span1, span2, slot1, slot2, w1_tap, w2_tap, w1_region, w2_region;
  uint32_t idesc, tmem_cols, acc2_col0;
  const float *b1, *b2;
  const __nv_bfloat16* res;
  int res_ld;
  __nv_bfloat16* out;
  int out_ld;
};

__device__ __forceinline__ void mma_acc(uint32_t d_tmem, uint32_t alo, uint32_t ahi, uint32_t blo, uint32_t bhi,
                                        uint32_t idesc) {
  asm volatile(
      "{\n\t.reg .b64 da, db;\n\t.reg .pred p;\n\t"
      "mov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\t"
      "setp.ne.b32 p, 1, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}\n"
      ::"r"(d_tmem), "r"(alo), "r"(ahi), "r"(blo), "r"(bhi), "r"(idesc)
      : "memory");
}

// One conv stage of the row-stationary issuer: ring row `j` of `rows_in` (= rows_out + 2) feeds the output rows
// j-2..j (clipped), whose accumulators are consecutive blocks of a 4-deep TMEM ring per 128-column block.
template <int KSTEPS>
__device__ __forceinline__ void issue_stage(const ChainParams& p, uint32_t tmem_col0, uint32_t ring_base, uint32_t slot_bytes,
                                            uint32_t span, uint32_t w_base, uint32_t w_tap, int nslots, int rows_out,
                                            uint32_t full0, uint32_t empty0, uint32_t accf0, uint32_t acce0) {
  const uint32_t layout = swizzle_layout_code((int)span);
  const uint32_t desc_hi = (uint32_t)((make_smem_desc(0, 0, 8u * span, layout)) >> 32);
  const uint32_t span16 = span >> 4, slot16 = slot_bytes >> 4, wtap16 = w_tap >> 4;
  const uint32_t ring16 = (ring_base + kLead - span) >> 4, w16 = w_base >> 4;   // window of tap sp starts at pixel sp-1
  const uint32_t nt = (uint32_t)p.planes;
  const uint32_t nunit = (nt >> 3) << 17;
  const uint32_t cb16 = (128u * span) >> 4;
  const int rows_in = rows_out + 2;
  uint32_t s = 0, ph = 0, q0 = 0;
  for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, q0 += (uint32_t)rows_out) {
    for (int j = 0; j < rows_in; ++j) {
      mbar_wait(full0 + 8u * s, ph);
      const int olo = j > 2 ? j - 2 : 0;
      const int ohi = j < rows_out ? j : rows_out - 1;
      if (j < rows_out) {
        const uint32_t q = q0 + (uint32_t)j;
        mbar_wait(acce0 + 8u * (q & 3u), ((q >> 2) & 1u) ^ 1u);
      }
      tc_fence_after();
      const uint32_t blo = (q0 + (uint32_t)olo) & 3u;
      const uint32_t nrows = (uint32_t)(ohi - olo + 1);
      const uint32_t n1 = min(nrows, 4u - blo), n2 = nrows - n1;
      const uint32_t jb0 = (uint32_t)(olo - j + 2);
      if (elect_one()) {
        const uint32_t id1 = p.idesc + n1 * nunit, id2 = p.idesc + n2 * nunit;
        for (int cb = 0; cb < p.cb; ++cb) {
          const uint32_t a_row = ring16 + s * slot16 + (uint32_t)cb * cb16;
          const uint32_t dcb = tmem_col0 + (uint32_t)cb * 4u * nt;
#pragma unroll
          for (int sp = 0; sp < 3; ++sp) {
            const uint32_t a_t = a_row + (uint32_t)sp * span16;
            const uint32_t b_t = w16 + ((uint32_t)(sp * 3) + jb0) * wtap16;
#pragma unroll
            for (int k = 0; k < KSTEPS; ++k) {
              mma_acc(dcb + blo * nt, a_t + 2u * k, desc_hi, b_t + 2u * k, desc_hi, id1);
              if (n2) mma_acc(dcb, a_t + 2u * k, desc_hi, b_t + n1 * wtap16 + 2u * k, desc_hi, id2);
            }
          }
        }
        tc_commit(empty0 + 8u * s);
        if (j >= 2) tc_commit(accf0 + 8u * ((q0 + (uint32_t)(j - 2)) & 3u));
      }
      __syncwarp();
      if (++s == (uint32_t)nslots) { s = 0; ph ^= 1u; }
    }
  }
}

template <int K1, int K2>
__global__ void __launch_bounds__(kCThreads, 1) chain_halo_kernel(const __grid_constant__ ChainParams p) {
  pdl_launch_dependents();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t w1_base = smem_base;
  const uint32_t w2_base = w1_base + p.w1_region;
  const uint32_t ring1 = w2_base + p.w2_region;
  const uint32_t ring2 = ring1 + (uint32_t)p.s1 * p.slot1;
  const uint32_t bar_base = ring2 + (uint32_t)p.s2 * p.slot2;
  const uint32_t S1 = (uint32_t)p.s1, S2 = (uint32_t)p.s2;
  const uint32_t xfull0 = bar_base;
  const uint32_t xempty0 = xfull0 + 8u * S1;
  const uint32_t afull0 = xempty0 + 8u * S1;
  const uint32_t aempty0 = afull0 + 8u * S2;
  const uint32_t wbar = aempty0 + 8u * S2;
  const uint32_t acc1f0 = wbar + 8u, acc1e0 = acc1f0 + 32u, acc2f0 = acc1e0 + 32u, acc2e0 = acc2f0 + 32u;
  const uint32_t tmem_slot = acc2e0 + 32u;

  constexpr int NT = 16 * K2;                   // planes: N of both convs, K of the second
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (uint32_t s = 0; s < S1; ++s) { mbar_init(xfull0 + 8u * s, 1); mbar_init(xempty0 + 8u * s, 1); }
    for (uint32_t s = 0; s < S2; ++s) { mbar_init(afull0 + 8u * s, 4); mbar_init(aempty0 + 8u * s, 1); }
    mbar_init(wbar, 1);
    for (uint32_t b = 0; b < 4; ++b) {
      mbar_init(acc1f0 + 8u * b, 1); mbar_init(acc1e0 + 8u * b, 4);
      mbar_init(acc2f0 + 8u * b, 1); mbar_init(acc2e0 + 8u * b, 4);
    }
    fence_mbar_init();
    prefetch_tmap(&p.x_map);
    prefetch_tmap(&p.w1_map);
    prefetch_tmap(&p.w2_map);
  }
  if (warp == 1) {
    tmem_alloc_dyn(tmem_slot, p.tmem_cols);
    tmem_relinquish();
  }
  // left / right padding columns of every ring slot: zero once, nothing ever writes them again
  {
    const uint32_t data1 = (uint32_t)p.W * p.span1, data2 = (uint32_t)p.W * p.span2;
    for (uint32_t s = 0; s < S1 + S2; ++s) {
      const bool first = s < S1;
      const uint32_t base = first ? ring1 + s * p.slot1 : ring2 + (s - S1) * p.slot2;
      const uint32_t slot = first ? p.slot1 : p.slot2, data = first ? data1 : data2;
      for (uint32_t o = threadIdx.x * 16u; o < kLead; o += kCThreads * 16u)
        asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(base + o), "r"(0u) : "memory");
      for (uint32_t o = kLead + data + threadIdx.x * 16u; o < slot; o += kCThreads * 16u)
        asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(base + o), "r"(0u) : "memory");
    }
    fence_proxy_async_smem();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
  if (warp >= 2 && warp < 6) {                  // every accumulator starts at zero: all MMAs accumulate
    const uint32_t t_q = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    for (uint32_t c = 0; c < p.tmem_cols; c += 16) tmem_st16_zero(t_q + c);
    tmem_st_wait();
  }
  pdl_wait();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  const int rows_a = p.bh + 2;                  // a1 rows per band (one recomputed halo row on either side)
  if (warp == 0) {
    // ================= TMA producer: weights once, then bh+4 input rows per band =================
    if (elect_one()) {
      mbar_arrive_expect_tx(wbar, 9u * (p.w1_tap + p.w2_tap));
      for (int sp = 0; sp < 3; ++sp)
        for (int j = 0; j < 3; ++j) {
          const int slab = (2 - j) * 3 + sp;     // ring order j <-> filter row r = 2-j (dh = r-1), column s = sp
          tma_load_3d(w1_base + (uint32_t)(sp * 3 + j) * p.w1_tap, &p.w1_map, wbar, 0, 0, slab);
          tma_load_3d(w2_base + (uint32_t)(sp * 3 + j) * p.w2_tap, &p.w2_map, wbar, 0, 0, slab);
        }
    }
    __syncwarp();
    uint32_t s = 0, ph = 0;
    const uint32_t row_bytes = (uint32_t)p.W * p.span1, blk = 128u * p.span1;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      const int band = tile % p.bands, n = tile / p.bands;
      const int h0 = band * p.bh - 2;
      for (int j = 0; j < rows_a + 2; ++j) {
        mbar_wait(xempty0 + 8u * s, ph ^ 1u);
        if (elect_one()) {
          mbar_arrive_expect_tx(xfull0 + 8u * s, row_bytes);
          for (int cb = 0; cb < p.cb; ++cb)
            tma_load_4d(ring1 + s * p.slot1 + kLead + (uint32_t)cb * blk, &p.x_map, xfull0 + 8u * s, 0, cb * 128, h0 + j, n);
        }
        __syncwarp();
        if (++s == S1) { s = 0; ph ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ================= issuer of conv1: x ring -> a1 accumulators =================
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    mbar_wait(wbar, 0);
    issue_stage<K1>(p, tmem_u, ring1, p.slot1, p.span1, w1_base, p.w1_tap, p.s1, rows_a, xfull0, xempty0, acc1f0, acc1e0);
  } else if (warp == 10) {
    // ================= issuer of conv2: a1 ring -> output accumulators =================
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    mbar_wait(wbar, 0);
    issue_stage<K2>(p, tmem_u + p.acc2_col0, ring2, p.slot2, p.span2, w2_base, p.w2_tap, p.s2, p.bh, afull0, aempty0, acc2f0,
                    acc2e0);
  } else if (warp < 6) {
    // ================= epilogue 1 (warps 2..5): a1 = relu(acc + b1) -> bf16 -> swizzled smem row =================
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const uint32_t swz_mask = p.span2 == 128u ? 7u : (p.span2 == 64u ? 3u : 1u);
    uint32_t r = 0, s = 0, ph = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      const int band = tile % p.bands;
      for (int ja = 0; ja < rows_a; ++ja, ++r) {
        const uint32_t b = r & 3u;
        const int hr = band * p.bh - 1 + ja;
        const bool valid = hr >= 0 && hr < p.H;
        mbar_wait(acc1f0 + 8u * b, (r >> 2) & 1u);
        mbar_wait(aempty0 + 8u * s, ph ^ 1u);
        tc_fence_after();
        const uint32_t slot = ring2 + s * p.slot2;
        for (int cb = 0; cb < p.cb; ++cb) {
          const uint32_t t_row = tmem_base + ((uint32_t)(quad * 32) << 16) + ((uint32_t)cb * 4u + b) * (uint32_t)NT;
          const uint32_t off = kLead + (uint32_t)(cb * 128 + row) * p.span2;
#pragma unroll
          for (int c = 0; c < NT; c += 16) {
            uint32_t v[16];
            tmem_ld16(t_row + (uint32_t)c, v);
            tmem_ld_wait();
            tmem_st16_zero(t_row + (uint32_t)c);
            uint32_t pk[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float f0 = fmaxf(__uint_as_float(v[2 * j]) + __ldg(p.b1 + c + 2 * j), 0.f);
              const float f1 = fmaxf(__uint_as_float(v[2 * j + 1]) + __ldg(p.b1 + c + 2 * j + 1), 0.f);
              pk[j] = valid ? pack_bf16x2(f0, f1) : 0u;
            }
            const uint32_t a0 = off + (uint32_t)c * 2u, a1 = a0 + 16u;
            const uint32_t x0 = a0 ^ (((a0 >> 7) & swz_mask) << 4), x1 = a1 ^ (((a1 >> 7) & swz_mask) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(slot + x0), "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3]) : "memory");
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(slot + x1), "r"(pk[4]), "r"(pk[5]), "r"(pk[6]), "r"(pk[7]) : "memory");
          }
        }
        tmem_st_wait();
        tc_fence_before();
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) { mbar_arrive(acc1e0 + 8u * b); mbar_arrive(afull0 + 8u * s); }
        if (++s == S2) { s = 0; ph ^= 1u; }
      }
    }
  } else {
    // ================= epilogue 2 (warps 6..9): out = relu(relu(acc + b2) + res) =================
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    constexpr int CBMAX = 4 / K2;               // (W/128) * K2 <= 4
    constexpr int V = 2 * K2;                   // 16-byte vectors per pixel
    uint32_t r = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      const int band = tile % p.bands, n = tile / p.bands;
      for (int i = 0; i < p.bh; ++i, ++r) {
        const uint32_t b = r & 3u;
        const int h = band * p.bh + i;
        const size_t pix0 = ((size_t)n * p.H + h) * (size_t)p.W + row;
        uint4 rr[CBMAX * V];
        // the residual is read before the accumulator is waited for: its latency hides behind the wait
#pragma unroll
        for (int cb = 0; cb < CBMAX; ++cb)
          if (cb < p.cb) {
            const uint4* src = reinterpret_cast<const uint4*>(p.res + (pix0 + (size_t)cb * 128) * (size_t)p.res_ld);
#pragma unroll
            for (int k = 0; k < V; ++k) rr[cb * V + k] = __ldg(src + k);
          }
        mbar_wait(acc2f0 + 8u * b, (r >> 2) & 1u);
        tc_fence_after();
#pragma unroll
        for (int cb = 0; cb < CBMAX; ++cb)
          if (cb < p.cb) {
            const uint32_t t_row = tmem_base + p.acc2_col0 + ((uint32_t)(quad * 32) << 16) + ((uint32_t)cb * 4u + b) * (uint32_t)NT;
            uint4* dst = reinterpret_cast<uint4*>(p.out + (pix0 + (size_t)cb * 128) * (size_t)p.out_ld);
#pragma unroll
            for (int c = 0; c < NT; c += 16) {
              uint32_t v[16];
              tmem_ld16(t_row + (uint32_t)c, v);
              tmem_ld_wait();
              tmem_st16_zero(t_row + (uint32_t)c);
              const uint4 q0 = rr[cb * V + c / 8], q1 = rr[cb * V + c / 8 + 1];
              const uint32_t rw[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
              uint32_t pk[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float2 t = unpack_bf16x2(rw[j]);
                const float f0 = fmaxf(__uint_as_float(v[2 * j]) + __ldg(p.b2 + c + 2 * j), 0.f) + t.x;
                const float f1 = fmaxf(__uint_as_float(v[2 * j + 1]) + __ldg(p.b2 + c + 2 * j + 1), 0.f) + t.y;
                pk[j] = pack_bf16x2(fmaxf(f0, 0.f), fmaxf(f1, 0.f));
              }
              dst[c / 8] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
              dst[c / 8 + 1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
            }
          }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc2e0 + 8u * b);
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

typedef void (*chain_kernel_t)(const ChainParams);
chain_kernel_t pick_chain_kernel(int k1, int k2) {
  switch (k1 * 10 + k2) {
    case 11: return chain_halo_kernel<1, 1>;
    case 21: return chain_halo_kernel<2, 1>;
    case 41: return chain_halo_kernel<4, 1>;
    case 12: return chain_halo_kernel<1, 2>;
    case 22: return chain_halo_kernel<2, 2>;
    case 42: return chain_halo_kernel<4, 2>;
    case 14: return chain_halo_kernel<1, 4>;
    case 24: return chain_halo_kernel<2, 4>;
    case 44: return chain_halo_kernel<4, 4>;
    default: return nullptr;
  }
}

bool chain_disabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("UBR_NO_CHAIN"); v = (e && e[0] && e[0] != '0') ? 1 : 0; }
  return v == 1;
}

}  // namespace

// 0 = launched, 1 = shape not eligible, <0 = error
static int chain_halo_try(const ubr_view& x, const void* w1pk, const float* b1, const void* w2pk, const float* b2, int planes,
                          const ubr_view& res, const ubr_view& y, cudaStream_t stream) {
  if (chain_disabled()) return 1;
  const int cin = x.c;
  if (!(cin == 16 || cin == 32 || cin == 64)) return 1;
  if (!(planes == 16 || planes == 32 || planes == 64)) return 1;
  if (x.w % 128 != 0 || x.w < 128) return 1;
  const int cb = x.w / 128, k1 = cin / 16, k2 = planes / 16;
  if (cb * k2 > 4) return 1;
  int bh = 32;
  while (bh > 1 && x.h % bh != 0) bh >>= 1;
  if (bh < 8) return 1;
  chain_kernel_t kern = pick_chain_kernel(k1, k2);
  if (!kern) return 1;

  ChainParams p;
  memset(&p, 0, sizeof(p));
  p.span1 = (uint32_t)cin * 2u;
  p.span2 = (uint32_t)planes * 2u;
  p.slot1 = (kLead + (uint32_t)(x.w + 1) * p.span1 + 1023u) & ~1023u;
  p.slot2 = (kLead + (uint32_t)(x.w + 1) * p.span2 + 1023u) & ~1023u;
  p.w1_tap = (uint32_t)planes * p.span1;
  p.w2_tap = (uint32_t)planes * p.span2;
  p.w1_region = (9u * p.w1_tap + 1023u) & ~1023u;
  p.w2_region = (9u * p.w2_tap + 1023u) & ~1023u;
  const uint32_t budget = 225u * 1024u - 1024u - 1024u;            // alignment slack + barriers
  if (p.w1_region + p.w2_region + 2u * p.slot1 + 2u * p.slot2 > budget) return 1;
  int s2 = 3, s1;
  if (p.w1_region + p.w2_region + 3u * p.slot1 + 3u * p.slot2 > budget) s2 = 2;
  s1 = (int)((budget - p.w1_region - p.w2_region - (uint32_t)s2 * p.slot2) / p.slot1);
  if (s1 > 8) {
    s1 = 8;
    const int more = (int)((budget - p.w1_region - p.w2_region - 8u * p.slot1) / p.slot2);
    s2 = more > 6 ? 6 : more;
  }
  if (s1 < 2 || s2 < 2) return 1;
  p.s1 = s1; p.s2 = s2;
  {
    uint64_t dims[4] = {(uint64_t)x.c, (uint64_t)x.w, (uint64_t)x.h, (uint64_t)x.n};
    uint64_t str[4] = {2, (uint64_t)x.ld * 2, (uint64_t)x.w * x.ld * 2, (uint64_t)x.h * x.w * x.ld * 2};
    uint32_t box[4] = {(uint32_t)cin, 128, 1, 1};
    if (int rc = make_tmap_bf16(&p.x_map, x.ptr, 4, dims, str, box, (int)p.span1)) return rc;
  }
  {
    uint64_t dims[3] = {(uint64_t)cin, (uint64_t)planes, 9};
    uint64_t str[3] = {2, (uint64_t)cin * 2, (uint64_t)cin * planes * 2};
    uint32_t box[3] = {(uint32_t)cin, (uint32_t)planes, 1};
    if (int rc = make_tmap_bf16(&p.w1_map, w1pk, 3, dims, str, box, (int)p.span1)) return rc;
  }
  {
    uint64_t dims[3] = {(uint64_t)planes, (uint64_t)planes, 9};
    uint64_t str[3] = {2, (uint64_t)planes * 2, (uint64_t)planes * planes * 2};
    uint32_t box[3] = {(uint32_t)planes, (uint32_t)planes, 1};
    if (int rc = make_tmap_bf16(&p.w2_map, w2pk, 3, dims, str, box, (int)p.span2)) return rc;
  }
  p.cb = cb; p.bh = bh; p.bands = x.h / bh; p.ntiles = p.bands * x.n; p.H = x.h; p.W = x.w; p.planes = planes;
  p.idesc = make_idesc_bf16(128, 0, 0, 0);
  p.acc2_col0 = (uint32_t)(cb * 4 * planes);
  uint32_t cols = 32;
  while (cols < 2u * p.acc2_col0) cols <<= 1;
  p.tmem_cols = cols;
  p.b1 = b1; p.b2 = b2;
  p.res = (const __nv_bfloat16*)res.ptr; p.res_ld = res.ld;
  p.out = (__nv_bfloat16*)y.ptr; p.out_ld = y.ld;
  const uint32_t smem = p.w1_region + p.w2_region + (uint32_t)s1 * p.slot1 + (uint32_t)s2 * p.slot2 + 16u * (uint32_t)(s1 + s2) +
                        256u + 1024u;
  UBR_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  int ctas = sm_count();
  if (ctas > p.ntiles) ctas = p.ntiles;
  ::ubr::Launch(dim3((unsigned)ctas), kCThreads, smem, stream)(kern, p);
  UBR_LAUNCH_CHECK("chain_halo_kernel");
  return 0;
}

}  // namespace ubr

extern "C" {
using namespace ubr;

int ubr_block_chain_fwd(ubr_view x, const void* w1pk, const float* b1, const void* w2pk, const float* b2, int planes,
                        ubr_view res, ubr_view y, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(x.ptr && w1pk && w2pk && b1 && b2 && res.ptr && y.ptr, "block_chain_fwd: null pointer");
  UBR_REQUIRE(x.c % 16 == 0 && planes % 16 == 0, "block_chain_fwd: channel counts must be multiples of 16");
  UBR_REQUIRE(y.n == x.n && y.h == x.h && y.w == x.w && y.c == planes, "block_chain_fwd: the output must be (n,h,w,planes) of x");
  UBR_REQUIRE(res.n == y.n && res.h == y.h && res.w == y.w && res.c == planes, "block_chain_fwd: the residual must match the output");
  UBR_REQUIRE(x.ld % 8 == 0 && y.ld % 8 == 0 && res.ld % 8 == 0 && ((uintptr_t)x.ptr & 15) == 0 && ((uintptr_t)y.ptr & 15) == 0 &&
                  ((uintptr_t)res.ptr & 15) == 0, "block_chain_fwd: views need 16-byte aligned pixels");
  return chain_halo_try(x, w1pk, b1, w2pk, b2, planes, res, y, (cudaStream_t)stream);
}

}  // extern "C"
