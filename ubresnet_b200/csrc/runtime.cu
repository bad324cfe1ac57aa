// Library plumbing: error reporting, device gating, TMA descriptor encoding.
#include <stdarg.h>
#include <string.h>
#include <stdlib.h>
#include <atomic>
#include <mutex>
#include "common.cuh"

namespace ubr {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return UBR_E_CUDA;
}

static std::atomic<unsigned long long> g_launches{0};
static std::atomic<int> g_trace{0};
static thread_local char g_names[512];
static thread_local char g_names_out[512];
static thread_local unsigned g_names_len = 0;
void count_launch(const char* name) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  if (g_trace.load(std::memory_order_relaxed) && name) {
    // kernel-name trace for the bench profiler: "a+b" = the kernels this thread launched since the last take
    const size_t n = strlen(name);
    if (g_names_len + n + 2 < sizeof(g_names)) {
      if (g_names_len) g_names[g_names_len++] = '+';
      memcpy(g_names + g_names_len, name, n);
      g_names_len += (unsigned)n;
      g_names[g_names_len] = 0;
    }
  }
}

static thread_local cudaError_t g_launch_err = cudaSuccess;
void note_launch_error(cudaError_t e) { g_launch_err = e; }
cudaError_t take_launch_error() { const cudaError_t e = g_launch_err; g_launch_err = cudaSuccess; return e; }
bool pdl_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("UBR_PDL"); v = (e && e[0] == '0') ? 0 : 1; }
  return v == 1;
}

static int g_dev_ok[64];       // 0 unknown, 1 ok, -1 unsupported
static int g_sm_count[64];

int ensure_device() {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    set_error("no CUDA device (%s): ubresnet_b200 has no CPU fallback", cudaGetErrorString(e));
    cudaGetLastError();
    return UBR_E_UNSUPPORTED;
  }
  if (dev < 0 || dev >= 64) { set_error("device index %d out of range", dev); return UBR_E_UNSUPPORTED; }
  if (g_dev_ok[dev] == 0) {
    cudaDeviceProp p;
    e = cudaGetDeviceProperties(&p, dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceProperties");
    g_sm_count[dev] = p.multiProcessorCount;
    g_dev_ok[dev] = (p.major == 10) ? 1 : -1;
    if (g_dev_ok[dev] < 0)
      set_error("device %d is sm_%d%d; ubresnet_b200 kernels are sm_100a only (no fallback)", dev, p.major, p.minor);
  }
  if (g_dev_ok[dev] < 0) {
    set_error("current device is not sm_100 (B200); ubresnet_b200 kernels are sm_100a only (no fallback)");
    return UBR_E_UNSUPPORTED;
  }
  return 0;
}

int sm_count() {
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || g_sm_count[dev] == 0) return 148;
  return g_sm_count[dev];
}

typedef CUresult (*encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static encode_tiled_fn g_encode = nullptr;
static std::once_flag g_encode_once;

static void resolve_encode() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) g_encode = (encode_tiled_fn)fn;
}

int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                   const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes) {
  std::call_once(g_encode_once, resolve_encode);
  if (!g_encode) { set_error("cuTensorMapEncodeTiled entry point unavailable"); return UBR_E_UNSUPPORTED; }
  cuuint64_t gdim[5], gstr[5];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gdim[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i + 1];
  CUtensorMapSwizzle sw = CU_TENSOR_MAP_SWIZZLE_NONE;
  if (swizzle_bytes == 32) sw = CU_TENSOR_MAP_SWIZZLE_32B;
  else if (swizzle_bytes == 64) sw = CU_TENSOR_MAP_SWIZZLE_64B;
  else if (swizzle_bytes == 128) sw = CU_TENSOR_MAP_SWIZZLE_128B;
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr,
                        bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d): rank %d dims [%llu,%llu,%llu,%llu] box [%u,%u,%u,%u] stride1 %llu sw %d base %p",
              (int)r, rank, (unsigned long long)gdim[0], (unsigned long long)(rank > 1 ? gdim[1] : 0),
              (unsigned long long)(rank > 2 ? gdim[2] : 0), (unsigned long long)(rank > 3 ? gdim[3] : 0), bx[0],
              rank > 1 ? bx[1] : 0, rank > 2 ? bx[2] : 0, rank > 3 ? bx[3] : 0,
              (unsigned long long)(rank > 1 ? gstr[0] : 0), swizzle_bytes, base);
    return UBR_E_CUDA;
  }
  return 0;
}

}  // namespace ubr

extern "C" {

int ubr_abi_version(void) { return UBR_ABI_VERSION; }
const char* ubr_last_error(void) { return ubr::g_err; }
int ubr_device_check(void) { return ubr::ensure_device(); }

uint64_t ubr_launch_count(void) { return (uint64_t)ubr::g_launches.load(); }
void ubr_trace_kernels(int enable) { ubr::g_trace.store(enable ? 1 : 0); ubr::g_names_len = 0; ubr::g_names[0] = 0; }
const char* ubr_take_kernel_names(void) {
  memcpy(ubr::g_names_out, ubr::g_names, ubr::g_names_len + 1);
  ubr::g_names_out[ubr::g_names_len] = 0;
  ubr::g_names_len = 0;
  ubr::g_names[0] = 0;
  return ubr::g_names_out;
}

/* The library keeps no device allocations and no cached tensor maps (every descriptor is encoded per call): shutdown
 * only clears the per-thread trace / error state.  Present so that a binding has a symmetric teardown call. */
int ubr_shutdown(void) {
  ubr::g_trace.store(0);
  ubr::g_names_len = 0;
  ubr::g_names[0] = 0;
  ubr::g_err[0] = 0;
  ubr::take_launch_error();
  return 0;
}

int ubr_sync_check(void* stream) {
  cudaError_t e = cudaStreamSynchronize((cudaStream_t)stream);
  if (e != cudaSuccess) return ubr::cuda_fail(e, "cudaStreamSynchronize");
  e = cudaGetLastError();
  if (e != cudaSuccess) return ubr::cuda_fail(e, "async error");
  return 0;
}

}  // extern "C"
