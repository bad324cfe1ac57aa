// Weight-gradient GEMM on tcgen05 tensor cores (sm_100a).
//
//   D[m][tap][n] = sum_{pixels p} MS[p][m] * NS_view[tap.map][p + (tap.dh, tap.dw)][n]
//
// The reduction dimension is the PIXEL axis, so both operands are read exactly as they lie in HBM
// (NHWC: channels contiguous) as MN-major UMMA operands: a TMA box of 128 pixels x C channels IS the
// canonical MN-major swizzled tile (8 pixel-rows per swizzle atom).  No transposes anywhere.
//   conv kxk (stride 1|2) : MS = dy (out-channel rows), NS = x shifted per tap (parity views if stride 2)
//   ConvTranspose 4x4 s2  : MS = x  (in-channel rows),  NS = dy parity views shifted per tap
// Split-K over pixel chunks: each CTA writes an fp32 partial [m][tap][n]; a second kernel sums the
// partials in a FIXED order (deterministic, no atomics) into the fp32 parameter-gradient layout.
#include "common.cuh"
#include "conv_internal.cuh"

namespace ubr {

constexpr int kWMaxTaps = 49;
constexpr int kWThreads = 192;

struct WgradParams {
  CUtensorMap m_map;
  CUtensorMap n_map[4];
  int ntaps, G, ngroups;
  int m_ch, m_blocks, mbox_c;        // M side: total channels, boxes per CTA tile, channels per box
  int n_ch, n_blocks, nbox_c, nt, n_tiles;
  int stages;
  int tiles_w, tiles_h, bw, bh, bn;
  int total_chunks, chunks_per_split;
  uint32_t m_bytes, n_tap_bytes, stage_bytes;
  uint32_t idesc, tmem_cols;
  float* partials;                   // [split][m_ch][ntaps][n_ch]
  int8_t dh[kWMaxTaps], dw[kWMaxTaps];
  uint8_t nmap[kWMaxTaps];
};

__global__ void __launch_bounds__(kWThreads, 1) wgrad_gemm_kernel(const __grid_constant__ WgradParams p) {
  pdl_launch_dependents();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_base = smem_base + (uint32_t)p.stages * p.stage_bytes;
  const uint32_t full0 = bar_base;
  const uint32_t empty0 = bar_base + 8u * p.stages;
  const uint32_t accum_bar = bar_base + 16u * p.stages;
  const uint32_t tmem_slot = accum_bar + 8u;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int split = blockIdx.x;
  const int group = blockIdx.y / p.n_tiles;
  const int n_tile = blockIdx.y % p.n_tiles;
  const int m_tile = blockIdx.z;
  const int tap0 = group * p.G;
  const int gcur = min(p.G, p.ntaps - tap0);
  const int m0 = m_tile * 128;
  const int nn0 = n_tile * p.nt;
  const int chunk_begin = split * p.chunks_per_split;
  const int chunk_end = min(p.total_chunks, chunk_begin + p.chunks_per_split);
  const int nchunks = chunk_end - chunk_begin;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(full0 + 8u * s, 1);
      mbar_init(empty0 + 8u * s, 1);
    }
    mbar_init(accum_bar, 1);
    fence_mbar_init();
    prefetch_tmap(&p.m_map);
    prefetch_tmap(&p.n_map[0]);
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

  const uint32_t rbm = (uint32_t)p.mbox_c * 2u;     // bytes per pixel-row of an M box
  const uint32_t rbn = (uint32_t)p.nbox_c * 2u;

  if (warp == 0) {
    if (lane == 0) {
      // ================= TMA producer =================
      for (int i = 0; i < nchunks; ++i) {
        const int chunk = chunk_begin + i;
        const int tw = chunk % p.tiles_w;
        const int th = (chunk / p.tiles_w) % p.tiles_h;
        const int tn = chunk / (p.tiles_w * p.tiles_h);
        const int w0 = tw * p.bw, h0 = th * p.bh, n0 = tn * p.bn;
        const int s = i % p.stages;
        const uint32_t ph = (uint32_t)(i / p.stages) & 1u;
        mbar_wait(empty0 + 8u * s, ph ^ 1u);
        const uint32_t m_dst = smem_base + (uint32_t)s * p.stage_bytes;
        const uint32_t fb = full0 + 8u * s;
        mbar_arrive_expect_tx(fb, p.m_bytes + (uint32_t)gcur * p.n_tap_bytes);
        for (int b = 0; b < p.m_blocks; ++b)
          tma_load_4d(m_dst + (uint32_t)b * 128u * rbm, &p.m_map, fb, m0 + b * p.mbox_c, w0, h0, n0);
        for (int g = 0; g < gcur; ++g) {
          const int t = tap0 + g;
          const uint32_t n_dst = m_dst + p.m_bytes + (uint32_t)g * p.n_tap_bytes;
          for (int b = 0; b < p.n_blocks; ++b)
            tma_load_4d(n_dst + (uint32_t)b * 128u * rbn, &p.n_map[p.nmap[t]], fb, nn0 + b * p.nbox_c,
                        w0 + p.dw[t], h0 + p.dh[t], n0);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ================= MMA issuer =================
      const uint32_t lay_m = swizzle_layout_code((int)rbm);
      const uint32_t lay_n = swizzle_layout_code((int)rbn);
      // MN-major: SBO = stride between 8-row K groups, LBO = stride between MN atoms (channel blocks).
      // When the tile has a single channel block, every further atom ALIASES block 0 (LBO = 0): the
      // duplicated accumulator rows/cols are simply never read back.
      const uint32_t lbo_m = p.m_blocks > 1 ? 128u * rbm : 0u;
      const uint32_t lbo_n = p.n_blocks > 1 ? 128u * rbn : 0u;
      for (int i = 0; i < nchunks; ++i) {
        const int s = i % p.stages;
        const uint32_t ph = (uint32_t)(i / p.stages) & 1u;
        mbar_wait(full0 + 8u * s, ph);
        tc_fence_after();
        const uint32_t m_src = smem_base + (uint32_t)s * p.stage_bytes;
        for (int g = 0; g < gcur; ++g) {
          const uint32_t n_src = m_src + p.m_bytes + (uint32_t)g * p.n_tap_bytes;
#pragma unroll
          for (int k = 0; k < 8; ++k) {                  // 128 pixels = 8 x K16
            const uint64_t adesc = make_smem_desc(m_src + (uint32_t)k * 16u * rbm, lbo_m, 8u * rbm, lay_m);
            const uint64_t bdesc = make_smem_desc(n_src + (uint32_t)k * 16u * rbn, lbo_n, 8u * rbn, lay_n);
            tc_mma_bf16(tmem_base + (uint32_t)(g * p.nt), adesc, bdesc, p.idesc, (i > 0 || k > 0) ? 1u : 0u);
          }
        }
        tc_commit(empty0 + 8u * s);
      }
      tc_commit(accum_bar);
    }
  } else {
    // ================= epilogue: TMEM -> fp32 partials =================
    const int quad = warp & 3;
    const int m = quad * 32 + lane;
    const int mg = m0 + m;
    mbar_wait(accum_bar, 0);
    tc_fence_after();
    const uint32_t t_row = tmem_base + ((uint32_t)(quad * 32) << 16);
    const bool valid = (m < p.m_blocks * p.mbox_c) && (mg < p.m_ch);
    float* rowp = p.partials + ((size_t)split * p.m_ch + (size_t)(valid ? mg : 0)) * (size_t)p.ntaps * p.n_ch;
    for (int g = 0; g < gcur; ++g) {
      float* dst = rowp + (size_t)(tap0 + g) * p.n_ch + nn0;
      for (int c = 0; c < p.nt; c += 16) {
        uint32_t v[16];
        tmem_ld16(t_row + (uint32_t)(g * p.nt + c), v);
        tmem_ld_wait();
        if (valid) {
          float4* d4 = reinterpret_cast<float4*>(dst + c);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            d4[j] = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]),
                                __uint_as_float(v[4 * j + 3]));
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

// partials [splits][m_ch][ntaps][n_ch] -> dw[m*sm + tap*st + n*sn] (fixed summation order)
__global__ void wgrad_reduce_kernel(const float* __restrict__ partials, int splits, int m_ch, int m_valid, int ntaps,
                                    int n_ch, int n_valid, long sm, long st, long sn, float* __restrict__ dw,
                                    int accumulate) {
  pdl_prologue();
  const long total = (long)m_ch * ntaps * n_ch;
  const long idx = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int n = (int)(idx % n_ch);
  const int t = (int)((idx / n_ch) % ntaps);
  const int m = (int)(idx / ((long)n_ch * ntaps));
  if (m >= m_valid || n >= n_valid) return;
  float a8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};      // see wgrad_halo_reduce_kernel: same fixed-order scheme
  int s = 0;
  for (; s + 8 <= splits; s += 8) {
#pragma unroll
    for (int j = 0; j < 8; ++j) a8[j] += partials[(long)(s + j) * total + idx];
  }
  for (int j = 0; s < splits; ++s, ++j) a8[j] += partials[(long)s * total + idx];
  const float acc = ((a8[0] + a8[1]) + (a8[2] + a8[3])) + ((a8[4] + a8[5]) + (a8[6] + a8[7]));
  float* d = dw + m * sm + t * st + n * sn;
  *d = accumulate ? (*d + acc) : acc;
}

struct WTile { int bw, bh, bn; };
static WTile pick_wtile(int n, int h, int w) {
  WTile best{128, 1, 1};
  double best_eff = -1.0;
  for (int bw = 128; bw >= 1; bw >>= 1)
    for (int bh = 128 / bw; bh >= 1; bh >>= 1) {
      const int bn = 128 / (bw * bh);
      const double eff = ((double)w / (ceil_div(w, bw) * bw)) * ((double)h / (ceil_div(h, bh) * bh)) *
                         ((double)n / (ceil_div(n, bn) * bn));
      if (eff > best_eff + 1e-9) { best_eff = eff; best = WTile{bw, bh, bn}; }
    }
  return best;
}

struct WPlan {
  WTile t;
  int mbox_c, m_blocks, m_tiles, nbox_c, n_blocks, nt, n_tiles, G, ngroups, stages;
  int tiles_w, tiles_h, tiles_n, total_chunks, chunks_per_split, splits;
  uint32_t m_bytes, n_tap_bytes, stage_bytes, tmem_cols;
};

static WPlan plan_wgrad(int grid_n, int grid_h, int grid_w, int m_ch, int n_ch, int ntaps) {
  WPlan pl;
  pl.t = pick_wtile(grid_n, grid_h, grid_w);
  pl.mbox_c = m_ch >= 64 ? 64 : m_ch;
  const int m_tile_ch = m_ch >= 128 ? 128 : m_ch;
  pl.m_blocks = m_tile_ch / pl.mbox_c;
  pl.m_tiles = ceil_div(m_ch, 128);
  pl.nbox_c = n_ch >= 64 ? 64 : n_ch;
  // N tile: 128 channels where they divide the operand, else 64 (192/320/576-channel ASPP concat inputs)
  pl.nt = (n_ch % 128 == 0) ? 128 : ((n_ch % 64 == 0) ? 64 : n_ch);
  pl.n_blocks = pl.nt / pl.nbox_c;
  pl.n_tiles = n_ch / pl.nt;
  pl.m_bytes = (uint32_t)pl.m_blocks * 128u * pl.mbox_c * 2u;
  pl.n_tap_bytes = (uint32_t)pl.n_blocks * 128u * pl.nbox_c * 2u;
  int G = ntaps;
  if (G > 512 / pl.nt) G = 512 / pl.nt;
  int gs = (int)((98304u - pl.m_bytes) / pl.n_tap_bytes);
  if (gs < 1) gs = 1;
  if (G > gs) G = gs;
  pl.G = G;
  pl.ngroups = ceil_div(ntaps, G);
  // balance group sizes (e.g. 9 taps, G=5 -> 5+4)
  pl.G = ceil_div(ntaps, pl.ngroups);
  pl.stage_bytes = pl.m_bytes + (uint32_t)pl.G * pl.n_tap_bytes;
  int stages = (int)(196608u / pl.stage_bytes);
  if (stages > 4) stages = 4;
  if (stages < 2) stages = 2;
  pl.stages = stages;
  uint32_t cols = 32;
  while ((int)cols < pl.G * pl.nt) cols <<= 1;
  pl.tmem_cols = cols;
  pl.tiles_w = ceil_div(grid_w, pl.t.bw);
  pl.tiles_h = ceil_div(grid_h, pl.t.bh);
  pl.tiles_n = ceil_div(grid_n, pl.t.bn);
  pl.total_chunks = pl.tiles_w * pl.tiles_h * pl.tiles_n;
  // split-K over pixel chunks.  One CTA per SM is resident (up to 192 KB of operand stages), so the grid runs in
  // waves of 148 CTAs: pick the split count that minimises  waves x (chunks per CTA x t_chunk + fixed CTA cost)
  // + the cost of writing and re-reading one fp32 partial per split (measured shape: profiles/r01f, wgrad_gemm at
  // 21 % of the tensor peak with 7 splits = 336 CTAs = 2.3 waves on the 16^2 layers).
  const int base = pl.m_tiles * pl.ngroups * pl.n_tiles;
  const double t_chunk = pl.G * 8.0 * (pl.nt / 2 > 52 ? pl.nt / 2 : 52) / 1900.0;            // us of MMA issue per chunk
  const double t_split = (double)m_ch * ntaps * n_ch * 8.0 / 5.0e6;                            // us: partial write + read
  int best_s = 1;
  double best_cost = 1e30;
  const int max_s = pl.total_chunks < 296 ? pl.total_chunks : 296;
  const uint32_t smem_cta = (uint32_t)pl.stages * pl.stage_bytes + 2048u;
  int resident = (int)(232448u / smem_cta);                            // CTAs per SM by shared memory (TMEM: <= 512 columns)
  if (resident * (int)pl.tmem_cols > 512) resident = 512 / (int)pl.tmem_cols;
  if (resident < 1) resident = 1;
  for (int s = 1; s <= max_s; ++s) {
    const int cps = ceil_div(pl.total_chunks, s);
    if (ceil_div(pl.total_chunks, cps) != s) continue;                 // not a distinct split count
    const int waves = ceil_div(base * s, 148 * resident);
    const double cost = waves * (cps * t_chunk + 3.0) + t_split * s;
    if (cost < best_cost - 1e-9) { best_cost = cost; best_s = s; }
  }
  pl.chunks_per_split = ceil_div(pl.total_chunks, best_s);
  pl.splits = ceil_div(pl.total_chunks, pl.chunks_per_split);
  return pl;
}

static uint64_t plan_ws_bytes(const WPlan& pl, int m_ch, int n_ch, int ntaps) {
  return (uint64_t)pl.splits * m_ch * ntaps * n_ch * sizeof(float);
}

static int check_view_w(const ubr_view& v, const char* name) {
  UBR_REQUIRE(v.ptr != nullptr, "%s: null pointer", name);
  UBR_REQUIRE(v.n > 0 && v.h > 0 && v.w > 0 && v.c > 0, "%s: bad extents", name);
  UBR_REQUIRE(v.ld >= v.c && (v.ld % 8) == 0 && v.c % 16 == 0, "%s: c=%d ld=%d unsupported", name, v.c, v.ld);
  UBR_REQUIRE(((uintptr_t)v.ptr & 15) == 0, "%s: pointer must be 16-byte aligned", name);
  return 0;
}

static int act_map(CUtensorMap* m, const ubr_view& v, int box_c, WTile t) {
  uint64_t dims[4] = {(uint64_t)v.c, (uint64_t)v.w, (uint64_t)v.h, (uint64_t)v.n};
  uint64_t str[4] = {2, (uint64_t)v.ld * 2, (uint64_t)v.w * v.ld * 2, (uint64_t)v.h * v.w * v.ld * 2};
  uint32_t box[4] = {(uint32_t)box_c, (uint32_t)t.bw, (uint32_t)t.bh, (uint32_t)t.bn};
  return make_tmap_bf16(m, v.ptr, 4, dims, str, box, box_c * 2);
}
static int phase_map(CUtensorMap* m, const ubr_view& v, int ph, int pw, int box_c, WTile t) {
  uint64_t dims[4] = {(uint64_t)v.c, (uint64_t)(v.w / 2), (uint64_t)(v.h / 2), (uint64_t)v.n};
  uint64_t str[4] = {2, (uint64_t)v.ld * 4, (uint64_t)v.w * v.ld * 4, (uint64_t)v.h * v.w * v.ld * 2};
  uint32_t box[4] = {(uint32_t)box_c, (uint32_t)t.bw, (uint32_t)t.bh, (uint32_t)t.bn};
  const uint8_t* base = (const uint8_t*)v.ptr + ((size_t)ph * v.w + pw) * (size_t)v.ld * 2;
  return make_tmap_bf16(m, base, 4, dims, str, box, box_c * 2);
}

struct WTap { int dh, dw, nmap; };

// ms: M-side view (defines the pixel grid); ns: N-side view; strided = N side uses 4 parity views
static int run_wgrad(const ubr_view& ms, const ubr_view& ns, bool n_strided, const WTap* taps, int ntaps, int m_valid,
                     int n_valid, long sm, long st, long sn, float* dw, int accumulate, void* workspace,
                     uint64_t workspace_bytes, cudaStream_t stream) {
  UBR_REQUIRE(ntaps >= 1 && ntaps <= kWMaxTaps, "wgrad: ntaps=%d out of range", ntaps);
  UBR_REQUIRE(ms.c <= 128 || ms.c % 128 == 0, "wgrad: M-side channels %d must be <= 128 or a multiple of 128", ms.c);
  UBR_REQUIRE(ns.c <= 128 || ns.c % 64 == 0, "wgrad: N-side channels %d must be <= 128 or a multiple of 64", ns.c);
  WPlan pl = plan_wgrad(ms.n, ms.h, ms.w, ms.c, ns.c, ntaps);
  const uint64_t need = plan_ws_bytes(pl, ms.c, ns.c, ntaps);
  UBR_REQUIRE(workspace != nullptr && workspace_bytes >= need, "wgrad: workspace too small (%llu < %llu)",
              (unsigned long long)workspace_bytes, (unsigned long long)need);
  UBR_REQUIRE(((uintptr_t)workspace & 15) == 0, "wgrad: workspace must be 16-byte aligned");
  WgradParams p;
  memset(&p, 0, sizeof(p));
  if (int rc = act_map(&p.m_map, ms, pl.mbox_c, pl.t)) return rc;
  if (n_strided) {
    for (int ph = 0; ph < 2; ++ph)
      for (int pw = 0; pw < 2; ++pw)
        if (int rc = phase_map(&p.n_map[ph * 2 + pw], ns, ph, pw, pl.nbox_c, pl.t)) return rc;
  } else {
    if (int rc = act_map(&p.n_map[0], ns, pl.nbox_c, pl.t)) return rc;
  }
  p.ntaps = ntaps; p.G = pl.G; p.ngroups = pl.ngroups;
  p.m_ch = ms.c; p.m_blocks = pl.m_blocks; p.mbox_c = pl.mbox_c;
  p.n_ch = ns.c; p.n_blocks = pl.n_blocks; p.nbox_c = pl.nbox_c; p.nt = pl.nt; p.n_tiles = pl.n_tiles;
  p.stages = pl.stages;
  p.tiles_w = pl.tiles_w; p.tiles_h = pl.tiles_h; p.bw = pl.t.bw; p.bh = pl.t.bh; p.bn = pl.t.bn;
  p.total_chunks = pl.total_chunks; p.chunks_per_split = pl.chunks_per_split;
  p.m_bytes = pl.m_bytes; p.n_tap_bytes = pl.n_tap_bytes; p.stage_bytes = pl.stage_bytes;
  p.idesc = make_idesc_bf16(128, pl.nt, 1, 1);
  p.tmem_cols = pl.tmem_cols;
  p.partials = (float*)workspace;
  for (int i = 0; i < ntaps; ++i) {
    p.dh[i] = (int8_t)taps[i].dh; p.dw[i] = (int8_t)taps[i].dw; p.nmap[i] = (uint8_t)taps[i].nmap;
  }
  const uint32_t smem = (uint32_t)pl.stages * pl.stage_bytes + 16u * pl.stages + 16u + 1024u;
  static bool attr_set[64] = {};             // the opt-in is per device
  int dev_id = 0;
  UBR_CUDA_OK(cudaGetDevice(&dev_id));
  if (!attr_set[dev_id & 63]) {
    UBR_CUDA_OK(cudaFuncSetAttribute(wgrad_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev_id & 63] = true;
  }
  UBR_REQUIRE(smem <= 227u * 1024u, "wgrad: smem plan %u too large", smem);
  dim3 grid((unsigned)pl.splits, (unsigned)(pl.ngroups * pl.n_tiles), (unsigned)pl.m_tiles);
  ::ubr::Launch(grid, kWThreads, smem, stream)(wgrad_gemm_kernel, p);
  UBR_LAUNCH_CHECK("wgrad_gemm_kernel");
  const long total = (long)ms.c * ntaps * ns.c;
  ::ubr::Launch((unsigned)((total + 255) / 256), 256, 0, stream)(wgrad_reduce_kernel, p.partials, pl.splits, ms.c, m_valid, ntaps,
                                                                           ns.c, n_valid, sm, st, sn, dw, accumulate);
  UBR_LAUNCH_CHECK("wgrad_reduce_kernel");
  return 0;
}

}  // namespace ubr

using namespace ubr;

extern "C" {

uint64_t ubr_wgrad_workspace_bytes(int grid_n, int grid_h, int grid_w, int m_ch, int n_ch, int ntaps) {
  if (grid_n <= 0 || grid_h <= 0 || grid_w <= 0 || m_ch <= 0 || n_ch <= 0 || ntaps <= 0) return 0;
  WPlan pl = plan_wgrad(grid_n, grid_h, grid_w, m_ch, n_ch, ntaps);
  uint64_t need = plan_ws_bytes(pl, m_ch, n_ch, ntaps);
  // the halo-resident kernel (stride-1 convs, <= 64 channels) keeps one partial per persistent CTA
  if (m_ch <= 64 && n_ch <= 64 && (ntaps == 1 || ntaps == 9 || ntaps == 49)) {
    const int k = ntaps == 1 ? 1 : (ntaps == 9 ? 3 : 7);
    const uint64_t h = wgrad_halo_workspace_bytes(n_ch, m_ch, k);
    if (h > need) need = h;
  }
  return need;
}

int ubr_conv2d_wgrad(ubr_view x, ubr_view dy, int out_ch, int ksize, int stride, float* dw, int accumulate,
                     void* workspace, uint64_t workspace_bytes, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = check_view_w(x, "conv2d_wgrad.x")) return rc;
  if (int rc = check_view_w(dy, "conv2d_wgrad.dy")) return rc;
  UBR_REQUIRE(ksize == 1 || ksize == 3 || ksize == 7, "conv2d_wgrad: ksize %d unsupported", ksize);
  UBR_REQUIRE(stride == 1 || stride == 2, "conv2d_wgrad: stride %d unsupported", stride);
  UBR_REQUIRE(dy.n == x.n && dy.h * stride == x.h && dy.w * stride == x.w, "conv2d_wgrad: shape mismatch");
  UBR_REQUIRE(dw != nullptr, "conv2d_wgrad: null dw");
  const int pad = (ksize - 1) / 2, T = ksize * ksize;
  WTap taps[kWMaxTaps];
  int nt = 0;
  for (int r = 0; r < ksize; ++r)
    for (int s = 0; s < ksize; ++s) {
      const int oh = r - pad, ow = s - pad;
      if (stride == 1) taps[nt++] = WTap{oh, ow, 0};
      else {
        const int ph = oh & 1, pw = ow & 1;
        taps[nt++] = WTap{(oh - ph) / 2, (ow - pw) / 2, ph * 2 + pw};
      }
    }
  UBR_REQUIRE(out_ch >= 1 && out_ch <= dy.c, "conv2d_wgrad: out_ch=%d vs dy.c=%d", out_ch, dy.c);
  if (stride == 1) {
    const int hrc = wgrad_halo_try(x, dy, out_ch, ksize, dw, accumulate, workspace, workspace_bytes, (cudaStream_t)stream);
    if (hrc <= 0) return hrc;
  }
  // dw is OIHW with the TRUE out_ch rows; dy.c may be padded (logits): rows >= out_ch are skipped
  return run_wgrad(dy, x, stride == 2, taps, nt, out_ch, x.c, (long)x.c * T, 1, (long)T, dw, accumulate, workspace,
                   workspace_bytes, (cudaStream_t)stream);
}

int ubr_conv2d_dilated_wgrad(ubr_view x, ubr_view dy, int out_ch, int dilation, float* dw, int accumulate,
                             void* workspace, uint64_t workspace_bytes, void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = check_view_w(x, "conv2d_dilated_wgrad.x")) return rc;
  if (int rc = check_view_w(dy, "conv2d_dilated_wgrad.dy")) return rc;
  UBR_REQUIRE(dilation >= 1 && dilation <= 16, "conv2d_dilated_wgrad: dilation %d unsupported", dilation);
  UBR_REQUIRE(dy.n == x.n && dy.h == x.h && dy.w == x.w, "conv2d_dilated_wgrad: shape mismatch");
  UBR_REQUIRE(dw != nullptr, "conv2d_dilated_wgrad: null dw");
  UBR_REQUIRE(out_ch >= 1 && out_ch <= dy.c, "conv2d_dilated_wgrad: out_ch=%d vs dy.c=%d", out_ch, dy.c);
  // dw[co,ci,r,s] = sum_p dy[p,co] * x[p + d*(r-1, s-1), ci]   (3x3, stride 1, padding == dilation)
  WTap taps[9];
  int nt = 0;
  for (int r = 0; r < 3; ++r)
    for (int s = 0; s < 3; ++s) taps[nt++] = WTap{(r - 1) * dilation, (s - 1) * dilation, 0};
  return run_wgrad(dy, x, false, taps, nt, out_ch, x.c, (long)x.c * 9, 1, 9L, dw, accumulate, workspace, workspace_bytes,
                   (cudaStream_t)stream);
}

int ubr_deconv4s2_wgrad(ubr_view x, ubr_view dy, float* dw, int accumulate, void* workspace, uint64_t workspace_bytes,
                        void* stream) {
  if (int rc = ensure_device()) return rc;
  if (int rc = check_view_w(x, "deconv_wgrad.x")) return rc;
  if (int rc = check_view_w(dy, "deconv_wgrad.dy")) return rc;
  UBR_REQUIRE(dy.n == x.n && dy.h == 2 * x.h && dy.w == 2 * x.w, "deconv_wgrad: shape mismatch");
  UBR_REQUIRE(dw != nullptr, "deconv_wgrad: null dw");
  // dw[ci,co,kh,kw] = sum_i x[i,ci] * dy[2i-1+kh, co]
  WTap taps[16];
  int nt = 0;
  for (int kh = 0; kh < 4; ++kh)
    for (int kw = 0; kw < 4; ++kw) {
      const int oh = kh - 1, ow = kw - 1;
      const int ph = oh & 1, pw = ow & 1;
      taps[nt++] = WTap{(oh - ph) / 2, (ow - pw) / 2, ph * 2 + pw};
    }
  return run_wgrad(x, dy, true, taps, nt, x.c, dy.c, (long)dy.c * 16, 1, 16, dw, accumulate, workspace, workspace_bytes,
                   (cudaStream_t)stream);
}

}  // extern "C"
