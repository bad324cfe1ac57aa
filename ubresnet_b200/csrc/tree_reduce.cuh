// Two-level "last CTA to arrive finishes the job" reduction of per-CTA partial rows.
//
// Every reduction of the training step (BatchNorm statistics, the sum dz / sum dz*xhat pairs of the BatchNorm backward)
// is produced as one fp32 partial row per CTA and used to be summed by a separate finalize launch (96 launches of a few
// microseconds per step).  Here the producing kernel finishes the reduction itself, still in a FIXED order:
//   level 1: the partial rows are grouped 16 CTAs at a time; the CTA that arrives last in its group (ticket counter)
//            sums the group's rows in row order into one fp64 row of `lvl2`;
//   level 2: the group-finisher that arrives last at the root ticket sums the lvl2 rows in group order and runs the
//            finalize arithmetic (tree_finish_bn / tree_finish_sums) with all its threads.
// Both sums use the whole CTA (tree_colsum: a few independent 16-byte loads per thread, row lanes combined in lane order).
// Which CTA does the work depends on timing, WHAT is summed in which order does not: results are bit-reproducible.
// A single CTA never reads more than 16 rows + ngroups fp64 rows, so the tail stays a few microseconds even for the
// 1184-row x 1024-column case, and only the last group's tail is exposed (the others overlap CTAs still running).
//
// State: `tree` points at caller-owned device memory of ubr_bn_tree_bytes(c) bytes, ZERO-FILLED once after allocation:
//   [ (1 + kTreeMaxGroups) x uint32 tickets, padded to kTreeTicketBytes ][ kTreeMaxGroups x 4c fp64 lvl2 rows ]
// The tickets are back at zero when the launch completes.  Launches that share a `tree` must be stream-ordered (the
// same rule as for the partial-row workspace they already share).
#pragma once
#include "common.cuh"

namespace ubr {

constexpr int kRedBlocks = 592;                // 4 x 148 SMs: partial rows [kRedBlocks][4][c] of a bn workspace; the final
                                               // sums [4][c] sit right behind them (ubr_bn_workspace_bytes)
constexpr int kTreeGroupCtas = 16;
constexpr int kTreeMaxGroups = 80;             // >= ceil(kMaxStatsRows / 16) = 74
constexpr size_t kTreeTicketBytes = 1024;

static inline size_t tree_bytes(int c) { return kTreeTicketBytes + (size_t)kTreeMaxGroups * 4u * (size_t)c * sizeof(double); }

struct TreeRed {
  const float* rows;      // [nrows][width] fp32 partial rows written by THIS launch
  double* lvl2;           // [ngroups][width]
  unsigned* tickets;      // [0] root, [1 + g] group g
  int nrows, width;       // width % 4 == 0
};

static inline TreeRed make_tree(void* tree, const float* rows, int nrows, int width) {
  TreeRed t;
  t.rows = rows;
  t.tickets = reinterpret_cast<unsigned*>(tree);
  t.lvl2 = reinterpret_cast<double*>(reinterpret_cast<char*>(tree) + kTreeTicketBytes);
  t.nrows = nrows;
  t.width = width;
  return t;
}

// training-mode BatchNorm finalize (the arithmetic of bn_finalize_kernel); columns [0,c) = sum, [sq_off, sq_off+c) = sum of squares
struct BnFinal {
  const float *gamma, *beta;
  float *running_mean, *running_var;
  float momentum, eps;
  double count;
  float *scale, *shift, *mean, *invstd;
  int c, sq_off;
};

// plain sums: sums[i] = total of column col_of(i), i < ns*c, optional parameter-gradient vectors per sum row
struct SumsFinal {
  float* sums;            // [ns][c]
  float* grad[4];         // per sum row: fp32 [c] parameter gradient (dbeta / dgamma) or null
  int c, ns, row_stride;  // row_stride: column offset between sum rows inside a partial row (>= c)
  int accumulate;
};

#ifdef __CUDACC__

// shared-memory scratch of the tail, carved out of the kernel's (by then idle) dynamic shared memory:
//   tot[width] fp64 column totals + scr[4 * blockDim.x] fp64 lane partials
__host__ __device__ constexpr size_t tree_smem_bytes(int width, int threads) { return ((size_t)width + 4u * (size_t)threads) * 8u; }

__device__ __forceinline__ void tree_ld4(const float* p, double (&v)[4]) {
  const float4 t = __ldcg(reinterpret_cast<const float4*>(p));
  v[0] = (double)t.x; v[1] = (double)t.y; v[2] = (double)t.z; v[3] = (double)t.w;
}
__device__ __forceinline__ void tree_ld4(const double* p, double (&v)[4]) {
  const double2 a = __ldcg(reinterpret_cast<const double2*>(p)), b = __ldcg(reinterpret_cast<const double2*>(p) + 1);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}

// dst[col] = sum over rows r < nr of src[r * width + col] for every column, by the whole CTA, in a fixed order:
// a thread owns 4 adjacent columns and every (blockDim / columns-in-flight)-th row (independent 16/32-byte loads, a
// few per thread), the row lanes are then combined through `scr` in lane order.  dst: shared or global memory.
template <typename T>
__device__ __forceinline__ void tree_colsum(const T* src, int nr, int width, double* dst, double* scr) {
  const int w4 = width >> 2;
  const int nth = (int)blockDim.x, tid = (int)threadIdx.x;
  for (int base = 0; base < w4; base += nth) {
    const int ncol = min(nth, w4 - base);
    const int nl = min(nth / ncol, nr);                // row lanes
    const int lane = tid / ncol, cc = tid - lane * ncol;
    if (lane < nl) {
      double a[4] = {0.0, 0.0, 0.0, 0.0};
      const T* p = src + 4 * (size_t)(base + cc);
#pragma unroll 8
      for (int r = lane; r < nr; r += nl) {
        double v[4];
        tree_ld4(p + (size_t)r * width, v);
        a[0] += v[0]; a[1] += v[1]; a[2] += v[2]; a[3] += v[3];
      }
      double* o = scr + 4 * (size_t)(lane * ncol + cc);
      o[0] = a[0]; o[1] = a[1]; o[2] = a[2]; o[3] = a[3];
    }
    __syncthreads();
    if (tid < ncol) {
      double a[4] = {0.0, 0.0, 0.0, 0.0};
      for (int l = 0; l < nl; ++l) {
        const double* o = scr + 4 * (size_t)(l * ncol + tid);
        a[0] += o[0]; a[1] += o[1]; a[2] += o[2]; a[3] += o[3];
      }
      double* d = dst + 4 * (size_t)(base + tid);
      d[0] = a[0]; d[1] = a[1]; d[2] = a[2]; d[3] = a[3];
    }
    __syncthreads();
  }
}

// Called by ALL threads of every CTA after its partial-row stores.  first_row: first partial row this CTA wrote,
// rows_per_cta: rows a CTA owns (consecutive), ctas_per_row: CTAs that share one row (channel-split grids).
// smem: tree_smem_bytes(width, blockDim.x) bytes of shared memory nobody else uses any more (16-byte aligned).
// Returns true in exactly one CTA of the grid; its tot[0 .. width) (= smem) then holds the column totals.
__device__ __forceinline__ bool tree_arrive(const TreeRed& t, int first_row, int rows_per_cta, int ctas_per_row,
                                            void* smem) {
  double* tot = reinterpret_cast<double*>(smem);
  double* scr = tot + t.width;
  const int rpg = kTreeGroupCtas * rows_per_cta;
  const int g = first_row / rpg;
  const int ngroups = (t.nrows + rpg - 1) / rpg;
  const int r0 = g * rpg;
  const int nr = min(rpg, t.nrows - r0);
  __threadfence();                 // this thread's partial-row stores are visible device-wide before the ticket is taken
  __syncthreads();
  int last = 0;                    // broadcast with __syncthreads_or: no static shared memory (the conv kernels use all 227 KB)
  if (threadIdx.x == 0) {
    const unsigned need = (unsigned)(((nr + rows_per_cta - 1) / rows_per_cta) * ctas_per_row);
    last = atomicAdd(t.tickets + 1 + g, 1u) == need - 1u;
    if (last) t.tickets[1 + g] = 0u;      // nobody else touches this group's ticket any more: zero for the next launch
  }
  if (!__syncthreads_or(last)) return false;
  __threadfence();
  const float* grows = t.rows + (size_t)r0 * t.width;
  if (ngroups == 1) {              // one group: its finisher is the root
    tree_colsum(grows, nr, t.width, tot, scr);
    return true;
  }
  tree_colsum(grows, nr, t.width, t.lvl2 + (size_t)g * t.width, scr);
  __threadfence();
  __syncthreads();
  last = 0;
  if (threadIdx.x == 0) {
    last = atomicAdd(t.tickets, 1u) == (unsigned)(ngroups - 1);
    if (last) t.tickets[0] = 0u;
  }
  if (!__syncthreads_or(last)) return false;
  __threadfence();
  tree_colsum(t.lvl2, ngroups, t.width, tot, scr);
  return true;
}

// tot: the column totals tree_arrive left in shared memory
__device__ __forceinline__ void tree_finish_bn(const double* tot, const BnFinal& f) {
  for (int ch = threadIdx.x; ch < f.c; ch += blockDim.x) {
    const double s = tot[ch];
    const double q = tot[f.sq_off + ch];
    const double mean = s / f.count;
    double var = q / f.count - mean * mean;
    if (var < 0.0) var = 0.0;
    const float invstd = (float)(1.0 / sqrt(var + (double)f.eps));
    const float g = f.gamma ? f.gamma[ch] : 1.f, b = f.beta ? f.beta[ch] : 0.f;
    f.scale[ch] = g * invstd;
    f.shift[ch] = b - (float)mean * g * invstd;
    if (f.mean) f.mean[ch] = (float)mean;
    if (f.invstd) f.invstd[ch] = invstd;
    if (f.running_mean) f.running_mean[ch] = (1.f - f.momentum) * f.running_mean[ch] + f.momentum * (float)mean;
    if (f.running_var) {
      const double unbiased = f.count > 1.0 ? var * f.count / (f.count - 1.0) : var;
      f.running_var[ch] = (1.f - f.momentum) * f.running_var[ch] + f.momentum * (float)unbiased;
    }
  }
}

__device__ __forceinline__ void tree_finish_sums(const double* tot, const SumsFinal& f) {
  for (int i = threadIdx.x; i < f.ns * f.c; i += blockDim.x) {
    const int k = i / f.c, ch = i - k * f.c;
    const float s = (float)tot[k * f.row_stride + ch];
    f.sums[i] = s;
    float* dst = f.grad[k];
    if (dst) dst[ch] = (f.accumulate ? dst[ch] : 0.f) + s;
  }
}

#endif  // __CUDACC__

}  // namespace ubr
