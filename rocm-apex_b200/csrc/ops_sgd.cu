// ops_sgd.cu -- b200mt_sgd <- amp_C.multi_tensor_sgd (csrc/multi_tensor_sgd_kernel.cu:29-322)
// SGD with momentum / nesterov / dampening / weight decay before or after momentum, the amp
// unscale (grad*scale) fused into the gradient read, optional fp16 model-weight write-out.
// The whole kernel is a no-op on the device when *noop != 0 (:46).
#include "mt_dispatch.h"

namespace b200mt {

template <typename Tg, typename Tw, bool COPY>
struct SgdOp : OpBase {
  // the extra output list is always Half in the reference, whatever the grad dtype (:61-65,160-161)
  using Types = typename std::conditional<COPY, TypeList<Tg, Tw, Tw, __half>, TypeList<Tg, Tw, Tw>>::type;
  static constexpr int NL = COPY ? 4 : 3;
  static constexpr unsigned STORES = COPY ? 0b1110 : 0b0110;
  static constexpr int UNROLL = 2;
  static constexpr int THREADS = 256;
  static constexpr int MIN_CTAS = 4;
  float wd, momentum, dampening, lr, scale;
  int nesterov, first_run, wd_after;
  const int* noop;
  __device__ __forceinline__ bool begin(State&) const {
    return *reinterpret_cast<const volatile int*>(noop) == 0;
  }
  // momentum buffer is neither read nor written when momentum == 0 (:99-120,131-132)
  __device__ __forceinline__ unsigned load_mask() const { return momentum != 0.f ? 0b0111u : 0b0011u; }
  __device__ __forceinline__ unsigned store_mask() const { return momentum != 0.f ? 0b1110u : 0b1010u; }
  __device__ __forceinline__ void elem(State&, float (&x)[NL]) const {
    float g = x[0] * scale;
    const float w = x[1];
    float mo = x[2];
    if (wd != 0.f && !wd_after) g += wd * w;
    if (momentum != 0.f) {
      if (!first_run) mo = mo * momentum + (1.f - dampening) * g;
      else mo = g;
      if (nesterov) g += momentum * mo;
      else g = mo;
    }
    if (wd != 0.f && wd_after) g += wd * w;
    // `weight_in[i] += (-lr * g)` on a Half/BFloat16 weight first converts the increment to the
    // weight's type (c10's operator+= takes a Half), then adds in fp32 and rounds again; for fp32
    // weights both conversions are the identity (csrc/multi_tensor_sgd_kernel.cu:124).
    const float wn = Elem<Tw>::round_trip(w + Elem<Tw>::round_trip(-lr * g));
    x[1] = wn;
    x[2] = mo;
    if constexpr (COPY) x[3] = wn;  // the reference re-reads the stored weight for the fp16 copy
  }
};

}  // namespace b200mt

using namespace b200mt;

extern "C" int b200mt_sgd(const b200mt_lists* L, int chunk_size, int32_t* noop, float wd, float momentum,
                          float dampening, float lr, int nesterov, int first_run, int wd_after_momentum,
                          float scale, void* stream_) {
  if (check_common(L, chunk_size)) return 1;
  B200MT_CHECK(noop != nullptr, "noop flag pointer is NULL");
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  B200MT_CHECK(L->depth == 3 || L->depth == 4, "tensor_lists.size() != depth");
  if (L->depth == 4)
    B200MT_CHECK(L->dtype[3] == B200MT_F16, "Additional output tensors should always be fp16.");
  TableView tv;
  if (acquire_table(L, 3, 4, elementwise_tiling(), stream, &tv)) return 1;
  const int dg = L->dtype[0], dw = L->dtype[1], n = L->depth;
  B200MT_CHECK(L->dtype[2] == dw, "multi_tensor_sgd: momentum buffers must have the weight dtype");
  auto run = [&](auto op) {
    op.wd = wd; op.momentum = momentum; op.dampening = dampening; op.lr = lr; op.scale = scale;
    op.nesterov = nesterov; op.first_run = first_run; op.wd_after = wd_after_momentum; op.noop = noop;
    return launch_stream(tv, op, stream);
  };
  // the six (grad, weight, lists) combinations of the reference (:150-314)
  if (dg == B200MT_F16 && dw == B200MT_F16 && n == 3) return run(SgdOp<__half, __half, false>());
  if (dg == B200MT_F32 && dw == B200MT_F32 && n == 3) return run(SgdOp<float, float, false>());
  if (dg == B200MT_F16 && dw == B200MT_F32 && n == 4) return run(SgdOp<__half, float, true>());
  if (dg == B200MT_F32 && dw == B200MT_F32 && n == 4) return run(SgdOp<float, float, true>());
  if (dg == B200MT_BF16 && dw == B200MT_BF16 && n == 3) return run(SgdOp<__nv_bfloat16, __nv_bfloat16, false>());
  if (dg == B200MT_BF16 && dw == B200MT_F32 && n == 4) return run(SgdOp<__nv_bfloat16, float, true>());
  return set_error("multi_tensor_sgd only supports some combinations of gradient & weight types. Given: "
                   "gradient: %s, weight: %s, num_lists: %d", dtype_name(dg), dtype_name(dw), n);
}
