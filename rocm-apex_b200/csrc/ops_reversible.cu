// ops_reversible.cu -- the overflow-tolerant Adam pair of fused_adam_cuda.
//   b200mt_reversible_adam <- fused_adam_cuda.reversible_adam
//                             (apex/contrib/csrc/optimizers/fused_adam_cuda_kernel.cu:570-653,814-912)
//   b200mt_maybe_adam_undo <- fused_adam_cuda.maybe_adam_undo (:655-732,961-1040)
// reversible_adam is adam_mt's update applied only where the unscaled gradient is finite; a non-finite
// gradient leaves its element alone and plants an infinity in p_copy[0], which the caller later finds
// with strided_check_finite and answers with maybe_adam_undo: the algebraic inverse of the step.
//
// One conscious divergence.  The reference's undo kernel indexes `j = j_start + i*ILP` inside its ILP
// loop (:687-690,722-724), so every thread reverts the same element four times and three elements
// out of four are never reverted.  This implementation reverts every element -- what the kernel is for --
// and tests/test_reversible_gpu.py shows both facts: elements 0 mod 4 agree with the reference kernel,
// the others are restored here and left at their stepped values there.
#include <cmath>
#include "mt_dispatch.h"

namespace b200mt {

// lists [p, m, v, g] (+ p_copy)
template <typename Tg, typename Tc, bool COPY>
struct ReversibleAdamOp : OpBase {
  using Types = typename std::conditional<COPY, TypeList<float, float, float, Tg, Tc>, TypeList<float, float, float, Tg>>::type;
  static constexpr int NL = COPY ? 5 : 4;
  static constexpr unsigned STORES = COPY ? 0b10111 : 0b00111;
  static constexpr int UNROLL = 2;
  struct State { bool overflow = false; };
  float b1, b2, eps, grad_scale, step_size, decay;
  int mode;
  int* overflow_word;
  __device__ __forceinline__ unsigned load_mask() const { return 0b01111; }
  __device__ __forceinline__ void elem(State& s, float (&x)[NL]) const {
    float p = x[0], m = x[1], v = x[2];
    const float sg = x[3] / grad_scale;   // IEEE: the finite test below must see what the reference sees
    if (isfinite(sg)) {
      m = b1 * m + (1.f - b1) * sg;
      v = b2 * v + (1.f - b2) * sg * sg;
      const float denom = (mode == 0) ? sqrtf(v + eps) : sqrtf(v) + eps;
      p = p - step_size * ((m / denom) + decay * p);
    } else {
      s.overflow = true;
    }
    x[0] = p; x[1] = m; x[2] = v;
    if constexpr (COPY) x[4] = p;
  }
  __device__ __forceinline__ void end(State& s) const {
    if (s.overflow) *overflow_word = 1;
  }
};

// Runs after the step kernel on the same stream: the marker must land after every regular store to
// p_copy[0] (the reference writes it from whichever block saw the overflow, unordered against the
// block that owns element 0).  Resets the word for the next call.
template <typename Tc>
__global__ void overflow_marker_kernel(int* word, void* p_copy0) {
  if (*word != 0) {
    Elem<Tc>::st1(p_copy0, __int_as_float(0x7f800000));
    *word = 0;
  }
}

// lists [p, m, v, g]
template <typename Tg>
struct AdamUndoOp : OpBase {
  using Types = TypeList<float, float, float, Tg>;
  static constexpr int NL = 4;
  static constexpr unsigned STORES = 0b0111;
  static constexpr int UNROLL = 2;
  float b1, b2, eps, grad_scale, step_size, decay;
  int mode;
  const int* flag;   // NULL: always undo
  __device__ __forceinline__ bool begin(State&) const {
    return !(flag && *reinterpret_cast<const volatile int*>(flag) == 0);
  }
  __device__ __forceinline__ unsigned load_mask() const { return 0b1111; }
  __device__ __forceinline__ void elem(State&, float (&x)[NL]) const {
    float p = x[0], m = x[1], v = x[2];
    const float sg = x[3] / grad_scale;
    if (isfinite(sg)) {
      const float denom = (mode == 0) ? sqrtf(v + eps) : sqrtf(v) + eps;
      p = (p + step_size * (m / denom)) / (1.0f - step_size * decay);
      m = (m - (1.f - b1) * sg) / b1;
      v = (v - (1.f - b2) * sg * sg) / b2;
      v = v >= 0.0f ? v : 0.0f;   // round-off may leave a small negative value when the first step is undone
    }
    x[0] = p; x[1] = m; x[2] = v;
  }
};

static float contrib_step_size(float lr, float beta1, float beta2, int step, int bias_correction) {
  if (bias_correction != 1) return lr;
  const float bc1 = 1.f - std::pow(beta1, step);
  const float bc2 = 1.f - std::pow(beta2, step);
  return lr * std::sqrt(bc2) / bc1;
}

}  // namespace b200mt

using namespace b200mt;

extern "C" int b200mt_reversible_adam(const b200mt_lists* L, int chunk_size, float lr, float beta1, float beta2,
                                      float eps, float grad_scale, int step, int mode, int bias_correction,
                                      float decay, void* stream_) {
  if (check_common(L, chunk_size)) return 1;
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  if (acquire_table(L, 4, 5, elementwise_tiling(), stream, &tv)) return 1;
  B200MT_CHECK(L->dtype[0] == B200MT_F32 && L->dtype[1] == B200MT_F32 && L->dtype[2] == B200MT_F32,
               "reversible_adam: expected parameter to be of float type");
  const int dg = L->dtype[3];
  B200MT_CHECK(L->depth == 4 || L->dtype[4] == dg || (L->dtype[4] == B200MT_U8 && dg != B200MT_F32),
               "reversible_adam: p_copy must have the gradient dtype or, for 16-bit gradients, be e5m2 bytes");
  void* scratch = nullptr;
  if (acquire_scratch(stream, 256, &scratch)) return 1;
  int* word = reinterpret_cast<int*>(reinterpret_cast<char*>(scratch) + 128);   // counter zone: zero between calls
  const float step_size = contrib_step_size(lr, beta1, beta2, step, bias_correction);
  auto fill = [&](auto& op) {
    op.b1 = beta1; op.b2 = beta2; op.eps = eps; op.grad_scale = grad_scale; op.step_size = step_size;
    op.decay = decay; op.mode = mode; op.overflow_word = word;
  };
  if (tv.ntiles <= 0) return 0;
  void* p_copy0 = nullptr;   // first element of the first non-empty p_copy
  for (int t = 0; L->depth == 5 && t < L->ntensors && !p_copy0; t++)
    if (L->numel[t] > 0) p_copy0 = const_cast<void*>(L->ptrs[(size_t)4 * L->ntensors + t]);
  return dispatch_f32_f16_bf16(dg, "adam_cuda_kernel", [&](auto tg) {
    using Tg = typename decltype(tg)::type;
    if (L->depth == 4) {
      // no copy, no marker: the word is still cleared so that it stays zero between calls
      ReversibleAdamOp<Tg, Tg, false> op; fill(op);
      if (launch_stream(tv, op, stream)) return 1;
      B200MT_CUDA(cudaMemsetAsync(word, 0, sizeof(int), stream));
      return 0;
    }
    if (L->dtype[4] == B200MT_U8) {
      ReversibleAdamOp<Tg, e5m2, true> op; fill(op);
      if (launch_stream(tv, op, stream)) return 1;
      overflow_marker_kernel<e5m2><<<1, 1, 0, stream>>>(word, p_copy0);
    } else {
      ReversibleAdamOp<Tg, Tg, true> op; fill(op);
      if (launch_stream(tv, op, stream)) return 1;
      overflow_marker_kernel<Tg><<<1, 1, 0, stream>>>(word, p_copy0);
    }
    count_launch();
    B200MT_CUDA(cudaGetLastError());
    return 0;
  });
}

extern "C" int b200mt_maybe_adam_undo(const b200mt_lists* L, int chunk_size, const int32_t* overflow_flag, float lr,
                                      float beta1, float beta2, float eps, float grad_scale, int step, int mode,
                                      int bias_correction, float decay, void* stream_) {
  if (check_common(L, chunk_size)) return 1;
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  if (acquire_table(L, 4, 4, elementwise_tiling(), stream, &tv)) return 1;
  B200MT_CHECK(L->dtype[0] == B200MT_F32 && L->dtype[1] == B200MT_F32 && L->dtype[2] == B200MT_F32,
               "maybe_adam_undo: expected parameter to be of float type");
  const float step_size = contrib_step_size(lr, beta1, beta2, step, bias_correction);
  return dispatch_f32_f16_bf16(L->dtype[3], "adam_cuda_kernel", [&](auto tg) {
    using Tg = typename decltype(tg)::type;
    AdamUndoOp<Tg> op;
    op.b1 = beta1; op.b2 = beta2; op.eps = eps; op.grad_scale = grad_scale; op.step_size = step_size;
    op.decay = decay; op.mode = mode; op.flag = overflow_flag;
    return launch_stream(tv, op, stream);
  });
}
