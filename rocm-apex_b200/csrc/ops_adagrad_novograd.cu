// ops_adagrad_novograd.cu -- the two remaining amp_C optimizers on the same harness.
//   b200mt_adagrad  <- amp_C.multi_tensor_adagrad  (csrc/multi_tensor_adagrad.cu:24-100)
//   b200mt_novograd <- amp_C.multi_tensor_novograd (csrc/multi_tensor_novograd.cu:32-188), which first blends
//                      the per-tensor gradient norm into grad_norms (multi_tensor_norm_out_cuda,
//                      csrc/multi_tensor_l2norm_kernel.cu:371-456) and then applies the update.
#include <cmath>
#include "mt_dispatch.h"
#include "mt_norm.cuh"

namespace b200mt {

// lists [g, p, h]; mode 0: L2 (decay folded into g), 1: decoupled decay (csrc/multi_tensor_adagrad.cu:65-72)
template <typename T>
struct AdagradOp : OpBase {
  using Types = TypeList<T, T, T>;
  static constexpr unsigned STORES = 0b110;
  static constexpr int UNROLL = 2;
  float eps, lr, decay;
  int mode;
  __device__ __forceinline__ unsigned load_mask() const { return 0b111; }
  __device__ __forceinline__ void elem(State&, float (&x)[3]) const {
    float g = x[0], p = x[1], h = x[2];
    if (mode == 0) {
      g = g + decay * p;
      h = h + g * g;
      p = p - lr * fast_div(g, fast_sqrt(h) + eps);
    } else {
      h = h + g * g;
      p = p - lr * (fast_div(g, fast_sqrt(h) + eps) + decay * p);
    }
    x[1] = p; x[2] = h;
  }
};

// per-tile max|x| for the L-infinity norm (fmaxf: NaN is dropped, as in the reference's MaxNormFunctor)
template <typename Tx>
struct MaxNormOp : OpBase {
  using Types = TypeList<Tx>;
  static constexpr unsigned STORES = 0;
  static constexpr int VEC = (Elem<Tx>::SIZE == 2) ? 8 : 4;
  static constexpr int UNROLL = (Elem<Tx>::SIZE == 2) ? 1 : 4;   // as L2NormOp
  static constexpr int THREADS = 256;
  struct State { float acc = 0.f; };
  float* partials;
  __device__ __forceinline__ unsigned load_mask() const { return 1u; }
  __device__ __forceinline__ void elem(State& s, float (&x)[1]) const { s.acc = fmaxf(s.acc, fabsf(x[0])); }
  __device__ __forceinline__ void tile_end(State& s, int tile, const TileLoc&, float* smem) const {
    const float t = cta_max<THREADS>(s.acc, smem);
    if (threadIdx.x == 0) partials[tile] = t;
    s.acc = 0.f;
  }
};

// per-tile sum of squares (same arithmetic as the l2norm first pass)
template <typename Tx>
struct SumSqOp : OpBase {
  using Types = TypeList<Tx>;
  static constexpr unsigned STORES = 0;
  static constexpr int VEC = (Elem<Tx>::SIZE == 2) ? 8 : 4;
  static constexpr int UNROLL = (Elem<Tx>::SIZE == 2) ? 1 : 4;   // as L2NormOp
  static constexpr int THREADS = 256;
  struct State { float acc = 0.f; };
  float* partials;
  __device__ __forceinline__ unsigned load_mask() const { return 1u; }
  __device__ __forceinline__ void elem(State& s, float (&x)[1]) const { s.acc += x[0] * x[0]; }
  __device__ __forceinline__ void tile_end(State& s, int tile, const TileLoc&, float* smem) const {
    const float t = cta_sum<THREADS>(s.acc, smem);
    if (threadIdx.x == 0) partials[tile] = t;
    s.acc = 0.f;
  }
};

// lists [g, p, m]; the second moment is one scalar per tensor (grad_norms), csrc/multi_tensor_novograd.cu:98-113
template <typename T>
struct NovoGradOp : OpBase {
  using Types = TypeList<T, T, T>;
  static constexpr unsigned STORES = 0b110;
  static constexpr int UNROLL = 2;
  struct State { float denom, rdenom; };
  float beta1, beta3, bc1, bc2, eps, lr, decay;
  int moment_mode;
  const float* grad_norms;
  __device__ __forceinline__ unsigned load_mask() const { return 0b111; }
  __device__ __forceinline__ void tile_begin(State& s, const TileLoc& L) const {
    s.denom = grad_norms[L.tensor] / bc2 + eps;   // plain load: written by the kernel just before
    s.rdenom = 1.f / s.denom;
  }
  __device__ __forceinline__ void elem(State& s, float (&x)[3]) const {
    float g = x[0], p = x[1], m = x[2];
    const float rbc1 = 1.f / bc1;   // loop-invariant
    if (moment_mode == 0) {   // regularisation inside the moment
      g = div_const(g, s.denom, s.rdenom) + (decay * p);
      m = beta1 * m + beta3 * g;
      p = p - (lr * div_const(m, bc1, rbc1));
    } else {
      m = beta1 * m + beta3 * g;
      const float update = div_const(div_const(m, bc1, rbc1), s.denom, s.rdenom) + (decay * p);
      p = p - (lr * update);
    }
    x[1] = p; x[2] = m;
  }
};

}  // namespace b200mt

using namespace b200mt;

extern "C" int b200mt_adagrad(const b200mt_lists* L, int chunk_size, int32_t* noop, float lr, float eps,
                              int mode, float weight_decay, void* stream_) {
  (void)noop;
  if (check_common(L, chunk_size)) return 1;
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  if (acquire_table(L, 3, 3, elementwise_tiling(), stream, &tv)) return 1;
  B200MT_CHECK(L->dtype[1] == L->dtype[0] && L->dtype[2] == L->dtype[0], "adagrad: g, p, h must share one dtype");
  return dispatch_f64_f32_f16_bf16(L->dtype[0], "adagrad", [&](auto t) {
    AdagradOp<typename decltype(t)::type> op;
    op.eps = eps; op.lr = lr; op.decay = weight_decay; op.mode = mode;
    return launch_stream(tv, op, stream);
  });
}

extern "C" int b200mt_novograd(const b200mt_lists* L, int chunk_size, int32_t* noop, float* grad_norms, float lr,
                               float beta1, float beta2, float eps, int step, int bias_correction,
                               float weight_decay, int grad_averaging, int moment_mode, int norm_type,
                               void* stream_) {
  if (check_common(L, chunk_size)) return 1;
  B200MT_CHECK(noop != nullptr && grad_norms != nullptr, "noop / grad_norms pointer is NULL");
  B200MT_CHECK(L->depth == 3, "tensor_lists.size() != depth");
  B200MT_CHECK(L->dtype[1] == L->dtype[0] && L->dtype[2] == L->dtype[0], "novograd: g, p, m must share one dtype");
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  // csrc/multi_tensor_novograd.cu:147-151
  float bc1 = 1.f, bc2 = 1.f;
  if (bias_correction == 1) {
    bc1 = (float)(1.0 - std::pow((double)beta1, (double)step));
    bc2 = (float)std::sqrt(1.0 - std::pow((double)beta2, (double)step));
  }
  const float beta3 = grad_averaging == 1 ? 1.f - beta1 : 1.f;

  // ---- 1. grad_norms <- blend(grad_norms, per-tensor norm of g): a one-list pass over g + second pass
  b200mt_lists G = *L;
  G.prepared = nullptr;   // a handle describes the whole 3-list set; this one-list view goes through the table cache
  G.depth = 1;   // list 0 (g) only: list-major layout means its pointers are the first ntensors entries
  const int tile_scale = (L->dtype[0] == B200MT_F16 || L->dtype[0] == B200MT_BF16) ? 2 : 1;
  TableView tg;
  if (acquire_table(&G, 1, 1, chunked_tiling(chunk_size, tile_scale), stream, &tg)) return 1;
  void* raw;
  if (acquire_scratch(stream, norm_scratch_bytes(1, tg.ntiles, tg.ntensors), &raw)) return 1;
  NormScratch ns = carve_norm_scratch(raw, 1, tg.ntiles, tg.ntensors);
  const int max_mode = (norm_type == 0) ? 1 : 0;
  int rc = dispatch_f32_f16_bf16(L->dtype[0], max_mode ? "multi_tensor_maxnorm_cuda" : "multi_tensor_l2norm_cuda", [&](auto t) {
    using T = typename decltype(t)::type;
    if (max_mode) { MaxNormOp<T> op; op.partials = ns.partials; return launch_stream(tg, op, stream); }
    SumSqOp<T> op; op.partials = ns.partials;
    return launch_stream(tg, op, stream);
  });
  if (rc) return 1;
  if (launch_norm_cleanup(tg, ns, 1, noop, 0, nullptr, grad_norms, nullptr, nullptr, stream, max_mode, 1, beta2,
                          1.0f - beta2)) return 1;

  // ---- 2. the update
  TableView tv;
  if (acquire_table(L, 3, 3, elementwise_tiling(), stream, &tv)) return 1;
  return dispatch_f64_f32_f16_bf16(L->dtype[0], "novograd", [&](auto t) {
    NovoGradOp<typename decltype(t)::type> op;
    op.beta1 = beta1; op.beta3 = beta3; op.bc1 = bc1; op.bc2 = bc2; op.eps = eps; op.lr = lr; op.decay = weight_decay;
    op.moment_mode = moment_mode; op.grad_norms = grad_norms;
    return launch_stream(tv, op, stream);
  });
}
