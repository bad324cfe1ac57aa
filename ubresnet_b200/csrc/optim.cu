// Multi-tensor optimizer updates: one launch updates every parameter of the model.
//   SGD  (train_ubresnet2018_wlarcv1.py:127-129: momentum 0.9, weight decay 1e-3)
//   Adam (train_ubresnet2018_wlarcv2.py:155-157: lr 1e-5, weight decay 1e-4)
// with torch.optim's arithmetic (L2 weight decay added to the gradient; Adam bias correction, no amsgrad).
// The caller uploads a table of {param, grad, state pointers, numel} and a chunk list (slot, element offset); CTA i
// processes chunk i: 256 threads x 16 fp32.  HBM-bound, all fp32: 20 (SGD with momentum) / 28 (Adam) bytes per element.
#include "common.cuh"

namespace ubr {

constexpr int kOptThreads = 256;
constexpr int kOptChunk = 4096;              // elements per CTA (must match ubresnet_b200/training/optim.py)

struct OptHyper {
  float lr, momentum, dampening, weight_decay;      // SGD
  float beta2, eps, step_size, bc2_sqrt;            // Adam: step_size = lr / (1 - beta1^t), bc2_sqrt = sqrt(1 - beta2^t)
  float one_m_beta1, one_m_beta2, one_m_damp;       // rounded from double on the host, as torch's Python scalars are
  int nesterov, first_step;
};

template <bool ADAM>
__global__ void __launch_bounds__(kOptThreads) opt_step_kernel(const ubr_opt_slot* __restrict__ slots,
                                                               const ubr_opt_chunk* __restrict__ chunks,
                                                               const OptHyper hp) {
  pdl_prologue();
  const ubr_opt_chunk ck = chunks[blockIdx.x];
  const ubr_opt_slot sl = slots[ck.slot];
  const long long end = (sl.numel - ck.offset) < kOptChunk ? sl.numel : ck.offset + kOptChunk;
  float* __restrict__ p = sl.param;
  const float* __restrict__ g = sl.grad;
  float* __restrict__ s1 = sl.state1;
  float* __restrict__ s2 = sl.state2;
  const bool vec = sl.vec_ok != 0;
  auto update = [&](float pv, float gv, float& a, float& b) -> float {
    gv = fmaf(hp.weight_decay, pv, gv);
    if (ADAM) {
      a = fmaf(hp.one_m_beta1, gv - a, a);                       // lerp, as torch's multi-tensor Adam
      b = fmaf(hp.one_m_beta2 * gv, gv, hp.beta2 * b);
      const float denom = sqrtf(b) / hp.bc2_sqrt + hp.eps;
      return fmaf(-hp.step_size, a / denom, pv);
    } else {
      float d = gv;
      if (hp.momentum != 0.f) {
        a = hp.first_step ? gv : fmaf(hp.momentum, a, hp.one_m_damp * gv);
        d = hp.nesterov ? fmaf(hp.momentum, a, gv) : a;
      }
      return fmaf(-hp.lr, d, pv);
    }
  };
  if (vec) {
    for (long long i = ck.offset + 4LL * threadIdx.x; i < end; i += 4LL * kOptThreads) {
      if (i + 4 <= end) {
        float4 pv = *reinterpret_cast<const float4*>(p + i);
        const float4 gv = *reinterpret_cast<const float4*>(g + i);
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
        if (s1 && !(hp.first_step && !ADAM)) a = *reinterpret_cast<const float4*>(s1 + i);
        if (ADAM) b = *reinterpret_cast<const float4*>(s2 + i);
        pv.x = update(pv.x, gv.x, a.x, b.x);
        pv.y = update(pv.y, gv.y, a.y, b.y);
        pv.z = update(pv.z, gv.z, a.z, b.z);
        pv.w = update(pv.w, gv.w, a.w, b.w);
        *reinterpret_cast<float4*>(p + i) = pv;
        if (s1) *reinterpret_cast<float4*>(s1 + i) = a;
        if (ADAM) *reinterpret_cast<float4*>(s2 + i) = b;
      } else {
        for (long long j = i; j < end; ++j) {
          float a = (s1 && !(hp.first_step && !ADAM)) ? s1[j] : 0.f, b = ADAM ? s2[j] : 0.f;
          p[j] = update(p[j], g[j], a, b);
          if (s1) s1[j] = a;
          if (ADAM) s2[j] = b;
        }
      }
    }
  } else {
    for (long long j = ck.offset + threadIdx.x; j < end; j += kOptThreads) {
      float a = (s1 && !(hp.first_step && !ADAM)) ? s1[j] : 0.f, b = ADAM ? s2[j] : 0.f;
      p[j] = update(p[j], g[j], a, b);
      if (s1) s1[j] = a;
      if (ADAM) s2[j] = b;
    }
  }
}

}  // namespace ubr

using namespace ubr;

extern "C" {

int ubr_opt_chunk_elems(void) { return kOptChunk; }

int ubr_sgd_step(const ubr_opt_slot* slots, const ubr_opt_chunk* chunks, int nchunks, float lr, float momentum,
                 float dampening, float weight_decay, int nesterov, int first_step, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(slots && chunks && nchunks > 0, "sgd_step: empty tables");
  OptHyper hp{};
  hp.lr = lr; hp.momentum = momentum; hp.dampening = dampening; hp.weight_decay = weight_decay;
  hp.one_m_damp = (float)(1.0 - (double)dampening);
  hp.nesterov = nesterov; hp.first_step = first_step;
  ::ubr::Launch(nchunks, kOptThreads, 0, (cudaStream_t)stream)(opt_step_kernel<false>, slots, chunks, hp);
  UBR_LAUNCH_CHECK("opt_step_kernel<sgd>");
  return 0;
}

int ubr_adam_step(const ubr_opt_slot* slots, const ubr_opt_chunk* chunks, int nchunks, double lr, double beta1,
                  double beta2, double eps, double weight_decay, int step, void* stream) {
  if (int rc = ensure_device()) return rc;
  UBR_REQUIRE(slots && chunks && nchunks > 0 && step >= 1, "adam_step: empty tables or step < 1");
  OptHyper hp{};
  hp.lr = (float)lr; hp.beta2 = (float)beta2; hp.eps = (float)eps; hp.weight_decay = (float)weight_decay;
  hp.one_m_beta1 = (float)(1.0 - beta1);
  hp.one_m_beta2 = (float)(1.0 - beta2);
  hp.step_size = (float)(lr / (1.0 - pow(beta1, (double)step)));
  hp.bc2_sqrt = (float)sqrt(1.0 - pow(beta2, (double)step));
  ::ubr::Launch(nchunks, kOptThreads, 0, (cudaStream_t)stream)(opt_step_kernel<true>, slots, chunks, hp);
  UBR_LAUNCH_CHECK("opt_step_kernel<adam>");
  return 0;
}

}  // extern "C"
