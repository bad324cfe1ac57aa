// Micro-benchmarks that pin the sm_100a facts the conv kernels are designed on (run on the B200 box):
//   T1  TMA tensor-load throughput per SM vs inner box bytes (32/64/128 B pixel rows)
//   T2  tcgen05.mma issue rate vs operand layout (K-major SW32/64/128, MN-major as used by wgrad), N = 16..256
//   T3  correctness of a K-major A descriptor whose start address is shifted by whole pixel rows (the
//       "shifted window into one resident halo tile" trick), with and without the descriptor base_offset field
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/ubench tools/ubench.cu ubresnet_b200/csrc/runtime.cu
#include <vector>
#include <stdlib.h>
#include <string.h>
#include "../ubresnet_b200/csrc/common.cuh"

using namespace ubr;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

// ------------------------------------------------------------------------------------------- T1
struct T1Params {
  CUtensorMap map;
  int boxes_per_cta, box_bytes, stages, tiles_w, tiles_h, nimg;
  int bw;
};
__global__ void __launch_bounds__(64, 1) t1_kernel(const __grid_constant__ T1Params p, unsigned long long* cycles) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar0 = base + (uint32_t)p.stages * (uint32_t)p.box_bytes;
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(bar0 + 8u * s, 1); mbar_init(bar0 + 8u * (p.stages + s), 1); }
    fence_mbar_init();
    prefetch_tmap(&p.map);
  }
  __syncthreads();
  const long long t0 = clock64();
  if (threadIdx.x == 0) {
    long long box = (long long)blockIdx.x * p.boxes_per_cta;
    int tw = (int)(box % p.tiles_w); box /= p.tiles_w;
    int th = (int)(box % p.tiles_h); box /= p.tiles_h;
    int tn = (int)(box % p.nimg);
    int s = 0; uint32_t ph = 0;
    for (int i = 0; i < p.boxes_per_cta; ++i) {
      mbar_wait(bar0 + 8u * (p.stages + s), ph ^ 1u);
      mbar_arrive_expect_tx(bar0 + 8u * s, (uint32_t)p.box_bytes);
      tma_load_4d(base + (uint32_t)s * p.box_bytes, &p.map, bar0 + 8u * s, 0, tw * p.bw, th, tn);
      if (++s == p.stages) { s = 0; ph ^= 1u; }
      if (++tw == p.tiles_w) { tw = 0; if (++th == p.tiles_h) { th = 0; if (++tn == p.nimg) tn = 0; } }
    }
  } else if (threadIdx.x == 32) {
    int s = 0; uint32_t ph = 0;
    for (int i = 0; i < p.boxes_per_cta; ++i) {
      mbar_wait(bar0 + 8u * s, ph);
      mbar_arrive(bar0 + 8u * (p.stages + s));
      if (++s == p.stages) { s = 0; ph ^= 1u; }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
}

static void run_t1(int c, int bw, int n, int h, int w, int grid, const char* note, int stages = 8) {
  const size_t elems = (size_t)n * h * w * c;
  __nv_bfloat16* buf;
  CK(cudaMalloc(&buf, elems * 2));
  CK(cudaMemset(buf, 0, elems * 2));
  T1Params p;
  memset(&p, 0, sizeof(p));
  uint64_t dims[4] = {(uint64_t)c, (uint64_t)w, (uint64_t)h, (uint64_t)n};
  uint64_t str[4] = {2, (uint64_t)c * 2, (uint64_t)w * c * 2, (uint64_t)h * w * c * 2};
  uint32_t box[4] = {(uint32_t)c, (uint32_t)bw, 1, 1};
  if (make_tmap_bf16(&p.map, buf, 4, dims, str, box, c * 2 > 128 ? 128 : c * 2)) { printf("tmap failed: %s\n", ubr_last_error()); exit(1); }
  p.box_bytes = bw * c * 2;
  p.stages = stages;
  while (p.stages * p.box_bytes > 192 * 1024) p.stages /= 2;
  p.bw = bw; p.tiles_w = w / bw; p.tiles_h = h; p.nimg = n;
  const long long total_boxes = (long long)p.tiles_w * h * n;
  p.boxes_per_cta = (int)(total_boxes / grid);
  unsigned long long* cyc;
  CK(cudaMalloc(&cyc, grid * 8));
  const int smem = p.stages * p.box_bytes + 16 * p.stages + 1024 + 64;
  CK(cudaFuncSetAttribute(t1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 3; ++rep) {
    CK(cudaEventRecord(e0));
    t1_kernel<<<grid, 64, smem>>>(p, cyc);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
  }
  CK(cudaGetLastError());
  const double bytes = (double)p.boxes_per_cta * grid * p.box_bytes;
  printf("T1 %-4s st=%2d C=%3d (row %3d B) box_w=%3d box=%6d B grid=%4d tensor=%7.1f MB : %8.3f ms  %8.1f GB/s  (%.1f B/cyc/SM @1.9GHz)\n",
         note, p.stages, c, c * 2, bw, p.box_bytes, grid, elems * 2 / 1e6, best, bytes / best / 1e6, bytes / best / 1e6 / 148 / 1.9);
  CK(cudaFree(buf)); CK(cudaFree(cyc));
}

// ------------------------------------------------------------------------------------------- T2
struct T2Params {
  int n, iters, ksteps, groups;
  uint32_t a_lbo, a_sbo, a_layout, a_kadv;     // kadv: bytes added to the start address per K step
  uint32_t b_lbo, b_sbo, b_layout, b_kadv;
  uint32_t a_stage_bytes, b_stage_bytes, nstage;
  uint32_t idesc;
  int commit_every;
};
__global__ void __launch_bounds__(128, 1) t2_kernel(const T2Params p, unsigned long long* cycles) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(8) unsigned long long bar;
  __shared__ __align__(8) unsigned long long bar2[2];
  const uint32_t bar_a = smem_u32(&bar);
  const uint32_t bar_b = smem_u32(&bar2[0]);
  // fill smem with small finite bf16 values
  for (uint32_t i = threadIdx.x; i < (p.a_stage_bytes + p.b_stage_bytes) * p.nstage / 4; i += blockDim.x)
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(base + i * 4), "r"(0x3c003c00u));
  if (threadIdx.x == 0) { mbar_init(bar_a, 1); mbar_init(bar_b, 1); mbar_init(bar_b + 8, 1); fence_mbar_init(); }
  if (threadIdx.x < 32) { tmem_alloc_dyn(smem_u32(&tmem_slot), 512); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (threadIdx.x == 0) {
    // descriptors precomputed; the timed loop is 8 x ksteps MMAs per trip with no integer division
    uint64_t ad[8], bd[8];
    for (int k = 0; k < 8; ++k) {
      const int kk = k % p.ksteps;
      ad[k] = make_smem_desc(base + kk * p.a_kadv, p.a_lbo, p.a_sbo, p.a_layout);
      bd[k] = make_smem_desc(base + p.a_stage_bytes + kk * p.b_kadv, p.b_lbo, p.b_sbo, p.b_layout);
    }
    const uint32_t idesc = p.idesc;
    const uint32_t gstep = (uint32_t)p.n, gmask = (uint32_t)p.groups - 1u;   // groups is a power of two
    const int trips = p.iters * p.ksteps / 8;
    tc_mma_bf16(tmem, ad[0], bd[0], idesc, 0u);
    for (uint32_t g = 1; g <= gmask; ++g) tc_mma_bf16(tmem + g * gstep, ad[0], bd[0], idesc, 0u);
    const long long t0 = clock64();
    uint32_t g = 0;
    for (int it = 0; it < trips; ++it) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        tc_mma_bf16(tmem + g * gstep, ad[k], bd[k], idesc, 1u);
        g = (g + 1u) & gmask;
        if (p.commit_every == 1) { tc_commit(bar_b); }
      }
      if (p.commit_every == 8) { tc_commit(bar_b); }
      if (p.commit_every == 82) { tc_commit(bar_b); tc_commit(bar_b + 8); }
      if (p.commit_every == 16 && (it & 1)) { tc_commit(bar_b); }
    }
    tc_commit(bar_a);
    mbar_wait(bar_a, 0);
    cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc_dyn(tmem, 512); }
}

static int g_commit_every = 0;
static void run_t2(const char* name, int n, int row_bytes_a, int row_bytes_b, bool mn_major, int m_blocks, int n_blocks, int groups = 1) {
  T2Params p;
  memset(&p, 0, sizeof(p));
  p.n = n;
  p.groups = groups;
  p.iters = 512;
  if (!mn_major) {
    // K-major: rows of row_bytes (one swizzle atom wide), 128 (A) / n (B) rows, K chunk = row_bytes/2 elements
    p.ksteps = (row_bytes_a < row_bytes_b ? row_bytes_a : row_bytes_b) / 32;
    p.a_lbo = 0; p.a_sbo = 8 * row_bytes_a; p.a_layout = swizzle_layout_code(row_bytes_a); p.a_kadv = 32;
    p.b_lbo = 0; p.b_sbo = 8 * row_bytes_b; p.b_layout = swizzle_layout_code(row_bytes_b); p.b_kadv = 32;
    p.a_stage_bytes = 128 * row_bytes_a;
    p.b_stage_bytes = ((n * row_bytes_b) + 1023) & ~1023;
    p.idesc = make_idesc_bf16(128, n, 0, 0);
  } else {
    // MN-major (wgrad): box = 128 pixels x C channels, rows = pixels (K), row_bytes = C*2
    p.ksteps = 8;
    p.a_lbo = m_blocks > 1 ? 128 * row_bytes_a : 0; p.a_sbo = 8 * row_bytes_a; p.a_layout = swizzle_layout_code(row_bytes_a);
    p.a_kadv = 16 * row_bytes_a;
    p.b_lbo = n_blocks > 1 ? 128 * row_bytes_b : 0; p.b_sbo = 8 * row_bytes_b; p.b_layout = swizzle_layout_code(row_bytes_b);
    p.b_kadv = 16 * row_bytes_b;
    p.a_stage_bytes = m_blocks * 128 * row_bytes_a;
    p.b_stage_bytes = n_blocks * 128 * row_bytes_b;
    p.idesc = make_idesc_bf16(128, n, 1, 1);
  }
  p.nstage = (p.a_stage_bytes + p.b_stage_bytes) > 49152 ? 2 : 4;
  p.commit_every = g_commit_every;
  unsigned long long* cyc;
  CK(cudaMalloc(&cyc, 148 * 8));
  const int smem = (p.a_stage_bytes + p.b_stage_bytes) * p.nstage + 2048;
  CK(cudaFuncSetAttribute(t2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
  t2_kernel<<<148, 128, smem>>>(p, cyc);
  CK(cudaDeviceSynchronize());
  unsigned long long h[148];
  CK(cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost));
  double avg = 0;
  for (int i = 0; i < 148; ++i) avg += (double)h[i];
  avg /= 148;
  const double per = avg / ((double)(p.iters * p.ksteps / 8) * 8);
  const double macs = 128.0 * n * 16;
  printf("T2 %-34s G=%d N=%3d : %7.1f cyc/MMA  (%6.0f MAC/cyc/SM = %5.1f%% of 4096; floor %d cyc)\n", name, groups, n, per, macs / per,
         100.0 * macs / per / 4096.0, n / 2);
  CK(cudaFree(cyc));
}

// T2w: the stacked-tap weight-gradient MMA (wgrad_halo.cu): A MN-major with overlapping atoms (LBO = 1 pixel),
// B MN-major with one atom per dy row (LBO = row pitch), N = ks*cout
static void run_t2w(int cin, int cout, int ks, int groups = 1) {
  T2Params p;
  memset(&p, 0, sizeof(p));
  const int ra = cin * 2, rbb = cout * 2;
  p.n = ks * cout;
  p.groups = groups;
  p.iters = 512;
  p.ksteps = 8;
  p.a_lbo = ra; p.a_sbo = 8 * ra; p.a_layout = swizzle_layout_code(ra); p.a_kadv = 16 * ra;
  p.b_lbo = 128 * rbb; p.b_sbo = 8 * rbb; p.b_layout = swizzle_layout_code(rbb); p.b_kadv = 16 * rbb;
  p.a_stage_bytes = ((160 * ra) + 1023) & ~1023;
  p.b_stage_bytes = ks * 128 * rbb;
  p.idesc = make_idesc_bf16(128, p.n, 1, 1);
  p.nstage = 1;
  unsigned long long* cyc;
  CK(cudaMalloc(&cyc, 148 * 8));
  const int smem = (p.a_stage_bytes + p.b_stage_bytes) * p.nstage + 2048;
  CK(cudaFuncSetAttribute(t2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
  t2_kernel<<<148, 128, smem>>>(p, cyc);
  CK(cudaDeviceSynchronize());
  unsigned long long h[148];
  CK(cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost));
  double avg = 0;
  for (int i = 0; i < 148; ++i) avg += (double)h[i];
  avg /= 148;
  const double per = avg / ((double)(p.iters * p.ksteps / 8) * 8);
  printf("T2w stacked wgrad MMA cin=%2d cout=%2d ks=%d (M=128 overlap LBO=%dB, N=%3d) G=%d : %7.1f cyc/MMA\n", cin, cout, ks, ra,
         p.n, groups, per);
  CK(cudaFree(cyc));
}

// ------------------------------------------------------------------------------------------- T3
struct T3Params {
  CUtensorMap a_map, b_map;
  int c, n, p0, use_base_offset, npix;
};
__global__ void __launch_bounds__(128, 1) t3_kernel(const __grid_constant__ T3Params p, float* out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(8) unsigned long long bars[2];
  const uint32_t bar_ld = smem_u32(&bars[0]), bar_mma = smem_u32(&bars[1]);
  const uint32_t S = (uint32_t)p.c * 2u;
  const uint32_t a_bytes = (uint32_t)p.npix * S, b_bytes = (uint32_t)p.n * S;
  const uint32_t a0 = base, b0 = base + ((a_bytes + 1023u) & ~1023u);
  if (threadIdx.x == 0) { mbar_init(bar_ld, 1); mbar_init(bar_mma, 1); fence_mbar_init(); }
  if (threadIdx.x < 32) { tmem_alloc_dyn(smem_u32(&tmem_slot), 32); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(bar_ld, a_bytes + b_bytes);
    for (int q = 0; q < p.npix; q += 256)
      tma_load_3d(a0 + (uint32_t)q * S, &p.a_map, bar_ld, 0, q, 0);
    tma_load_3d(b0, &p.b_map, bar_ld, 0, 0, 0);
    mbar_wait(bar_ld, 0);
    tc_fence_after();
    const uint32_t layout = swizzle_layout_code((int)S);
    const uint32_t start = a0 + (uint32_t)p.p0 * S;
    for (int k = 0; k < p.c / 16; ++k) {
      uint64_t ad = make_smem_desc(start + 32u * k, 0, 8u * S, layout);
      if (p.use_base_offset) ad |= (uint64_t)((start >> 7) & 7u) << 49;
      const uint64_t bd = make_smem_desc(b0 + 32u * k, 0, 8u * S, layout);
      tc_mma_bf16(tmem, ad, bd, make_idesc_bf16(128, p.n, 0, 0), k ? 1u : 0u);
    }
    tc_commit(bar_mma);
  }
  mbar_wait(bar_mma, 0);
  tc_fence_after();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int c = 0; c < p.n; c += 16) {
    uint32_t v[16];
    tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c, v);
    tmem_ld_wait();
    for (int j = 0; j < 16; ++j) out[(warp * 32 + lane) * p.n + c + j] = __uint_as_float(v[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc_dyn(tmem, 32); }
}

static void run_t3(int c, int p0, int use_bo) {
  const int npix = 512, n = 16;
  std::vector<__nv_bfloat16> ha((size_t)npix * c), hb((size_t)n * c);
  std::vector<float> fa((size_t)npix * c), fb((size_t)n * c);
  for (int p = 0; p < npix; ++p)
    for (int k = 0; k < c; ++k) { float v = (float)((p * 7 + k * 3) % 13 - 6); fa[p * c + k] = v; ha[p * c + k] = __float2bfloat16(v); }
  for (int j = 0; j < n; ++j)
    for (int k = 0; k < c; ++k) { float v = (float)((j * 5 + k * 11) % 7 - 3); fb[j * c + k] = v; hb[j * c + k] = __float2bfloat16(v); }
  __nv_bfloat16 *da, *db; float* dout;
  CK(cudaMalloc(&da, ha.size() * 2)); CK(cudaMalloc(&db, hb.size() * 2)); CK(cudaMalloc(&dout, 128 * n * 4));
  CK(cudaMemcpy(da, ha.data(), ha.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(db, hb.data(), hb.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemset(dout, 0, 128 * n * 4));
  T3Params p;
  memset(&p, 0, sizeof(p));
  p.c = c; p.n = n; p.p0 = p0; p.use_base_offset = use_bo; p.npix = npix;
  {
    uint64_t dims[3] = {(uint64_t)c, (uint64_t)npix, 1};
    uint64_t str[3] = {2, (uint64_t)c * 2, (uint64_t)npix * c * 2};
    uint32_t box[3] = {(uint32_t)c, 256, 1};
    if (make_tmap_bf16(&p.a_map, da, 3, dims, str, box, c * 2)) { printf("tmap a failed: %s\n", ubr_last_error()); exit(1); }
    uint64_t dimb[3] = {(uint64_t)c, (uint64_t)n, 1};
    uint64_t strb[3] = {2, (uint64_t)c * 2, (uint64_t)n * c * 2};
    uint32_t boxb[3] = {(uint32_t)c, (uint32_t)n, 1};
    if (make_tmap_bf16(&p.b_map, db, 3, dimb, strb, boxb, c * 2)) { printf("tmap b failed: %s\n", ubr_last_error()); exit(1); }
  }
  CK(cudaFuncSetAttribute(t3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  t3_kernel<<<1, 128, npix * c * 2 + 4096 + 2048>>>(p, dout);
  CK(cudaDeviceSynchronize());
  std::vector<float> ho(128 * n);
  CK(cudaMemcpy(ho.data(), dout, ho.size() * 4, cudaMemcpyDeviceToHost));
  int bad = 0;
  for (int m = 0; m < 128; ++m)
    for (int j = 0; j < n; ++j) {
      float ref = 0;
      for (int k = 0; k < c; ++k) ref += fa[(p0 + m) * c + k] * fb[j * c + k];
      if (ref != ho[m * n + j]) ++bad;
    }
  printf("T3 C=%3d (SW%3d) pixel shift p0=%3d base_offset=%d : %s (%d / %d mismatches)\n", c, c * 2, p0, use_bo,
         bad ? "WRONG" : "exact", bad, 128 * n);
  CK(cudaFree(da)); CK(cudaFree(db)); CK(cudaFree(dout));
}

int main(int argc, char** argv) {
  const char* which = argc > 1 ? argv[1] : "all";
  const bool all = !strcmp(which, "all");
  if (all || !strcmp(which, "t3")) {
    const int shifts[] = {0, 1, 2, 3, 4, 5, 7, 8, 9, 17, 66, 130, 257};
    for (int c : {16, 32, 64})
      for (int bo = 0; bo < 2; ++bo)
        for (int s : shifts) run_t3(c, s, bo);
  }
  if (all || !strcmp(which, "t2")) {
    for (int g : {1, 2, 4, 8})
      for (int n : {16, 32, 64}) run_t2("K-major SW32  (C=16 rows)", n, 32, 32, false, 1, 1, g);
    for (int g : {1, 2, 4, 8})
      for (int n : {16, 32, 64}) run_t2("K-major SW64  (C=32 rows)", n, 64, 64, false, 1, 1, g);
    for (int g : {1, 2, 4})
      for (int n : {16, 32, 64, 128}) run_t2("K-major SW128 (C=64 rows)", n, 128, 128, false, 1, 1, g);
    run_t2("K-major SW128 (C=64 rows)", 256, 128, 128, false, 1, 1, 1);
    run_t2("K-major SW128 (C=64 rows)", 256, 128, 128, false, 1, 1, 2);
    for (int g : {1, 2, 4, 8}) {
      run_t2("MN-major A C=16 alias, B C=16", 16, 32, 32, true, 1, 1, g);
      run_t2("MN-major A C=32 alias, B C=32", 32, 64, 64, true, 1, 1, g);
      run_t2("MN-major A C=64 alias, B C=64", 64, 128, 128, true, 1, 1, g);
    }
    run_t2("MN-major A C=128,     B C=64", 64, 128, 128, true, 2, 1, 1);
    run_t2("MN-major A C=128,     B C=64", 64, 128, 128, true, 2, 1, 4);
    run_t2("MN-major A C=128,     B C=128", 128, 128, 128, true, 2, 2, 1);
    run_t2("MN-major A C=128,     B C=128", 128, 128, 128, true, 2, 2, 4);
  }
  if (!strcmp(which, "t2c")) {
    // cost of tcgen05.commit in the MMA stream: none / one per 16 / one per 8 / two per 8 / one per MMA
    for (int ce : {0, 16, 8, 82, 1}) {
      g_commit_every = ce;
      printf("commit_every=%d: ", ce);
      run_t2("K-major SW32  (C=16 rows)", 16, 32, 32, false, 1, 1, 1);
    }
    g_commit_every = 0;
  }
  if (!strcmp(which, "t2w")) {
    for (int g : {1, 2}) {
      run_t2w(16, 16, 1, g); run_t2w(16, 16, 3, g); run_t2w(16, 16, 7, g);
      run_t2w(32, 32, 3, g); run_t2w(64, 64, 3, g); run_t2w(64, 32, 1, g);
    }
  }
  if (!strcmp(which, "t2b")) {
    // swapped orientation: A = weights (M rows, wide SW128 rows), B = pixels (N rows of C*2 bytes)
    for (int rb : {32, 64, 128})
      for (int n : {64, 128, 256}) run_t2(rb == 32 ? "A SW128 weights, B SW32 pixels" : rb == 64 ? "A SW128 weights, B SW64 pixels" : "A SW128 weights, B SW128 pixels", n, 128, rb, false, 1, 1, 1);
    for (int rb : {32, 64})
      for (int n : {128, 256}) run_t2(rb == 32 ? "A SW32 weights, B SW32 pixels" : "A SW64 weights, B SW64 pixels", n, rb, rb, false, 1, 1, 1);
    // pixels as A (current orientation) with wide weights rows
    for (int rb : {32, 64}) run_t2(rb == 32 ? "A SW32 pixels, B SW128 weights" : "A SW64 pixels, B SW128 weights", 64, rb, 128, false, 1, 1, 1);
  }
  if (all || !strcmp(which, "t1")) {
    for (int grid : {148, 296})
      for (int st : {4, 8, 16}) {
        run_t1(16, 128, 32, 512, 512, grid, "HBM", st);
        run_t1(16, 256, 32, 512, 512, grid, "HBM", st);
        run_t1(32, 256, 32, 256, 256, grid, "HBM", st);
        run_t1(64, 128, 32, 128, 128, grid, "HBM", st);
        run_t1(64, 128, 32, 256, 256, grid, "HBM", st);
        run_t1(16, 128, 2, 512, 512, grid, "L2", st);
        run_t1(64, 128, 8, 128, 128, grid, "L2", st);
      }
  }
  return 0;
}
