// ops_adam.cu -- Adam / AdamW updates.
//   b200mt_adam         <- amp_C.multi_tensor_adam (csrc/multi_tensor_adam.cu:23-171)
//   b200mt_dist_adam    <- distributed_adam_cuda.multi_tensor_fused_adam
//                          (apex/contrib/csrc/optimizers/multi_tensor_distopt_adam_kernel.cu:32-228)
//   b200mt_contrib_adam <- fused_adam_cuda.adam_mt
//                          (apex/contrib/csrc/optimizers/fused_adam_cuda_kernel.cu:74-200,283-387)
// All three are one fused read-modify-write pass: every list is read once and written once with
// 128-bit accesses (the reference's multi_tensor_adam has no vector path at all, :66-125).
#include <cmath>
#include "mt_dispatch.h"

// Vector width (elements per access) and vectors in flight per thread; swept on B200
// (profiles/sweep_vec_r01.txt).
#ifndef B200MT_ADAM_VEC
#define B200MT_ADAM_VEC 4
#endif
#ifndef B200MT_ADAM_UNROLL
#define B200MT_ADAM_UNROLL 2
#endif
#ifndef B200MT_NVLS_VEC
#define B200MT_NVLS_VEC 8
#endif
#ifndef B200MT_NVLS_UNROLL
#define B200MT_NVLS_UNROLL 2
#endif

namespace b200mt {

// ops_adam_tma.cu: bulk-copy (TMA) staging for the all-fp32, 16-byte-aligned case
bool adam_tma_wanted(const b200mt_lists* L);
int adam_tma_tile();
int adam_tma_launch(const TableView& tv, float beta1, float beta2, float bc1, float bc2, float eps, float lr,
                    float decay, float grad_scale, int mode, cudaStream_t stream);

// L2 eviction priority per list [g, p, m, v(, copy)], two bits each (mt_device.cuh: l2_policy)
#ifndef B200MT_ADAM_L2_LOADS
#define B200MT_ADAM_L2_LOADS 0
#endif
#ifndef B200MT_ADAM_L2_STORES
#define B200MT_ADAM_L2_STORES 0
#endif

template <typename T> struct is_mc { static constexpr bool value = false; };
template <> struct is_mc<mc_bf16> { static constexpr bool value = true; };
template <> struct is_mc<mc_f16> { static constexpr bool value = true; };
template <> struct is_mc<pair_bf16> { static constexpr bool value = true; };   // same vector shape: whole 16-byte vectors only

// ---- amp_C.multi_tensor_adam ------------------------------------------------------------------
// lists [g, p, m, v] (+ optional p_copy).  Per-element math follows
// csrc/multi_tensor_adam.cu:94-112 in fp32 (division / sqrt: see fast_div in mt_device.cuh).
template <typename Tg, typename Tp, typename Tc, bool COPY>
struct AdamOp : OpBase {
  using Types = typename std::conditional<COPY, TypeList<Tg, Tp, Tp, Tp, Tc>, TypeList<Tg, Tp, Tp, Tp>>::type;
  static constexpr int NL = COPY ? 5 : 4;
  static constexpr unsigned STORES = COPY ? 0b11110 : 0b01110;
  static constexpr int UNROLL = B200MT_ADAM_UNROLL;
  static constexpr int VEC = B200MT_ADAM_VEC;
  static constexpr int THREADS = 256;
#ifdef B200MT_ADAM_MIN_CTAS
  static constexpr int MIN_CTAS = B200MT_ADAM_MIN_CTAS;
#else
  static constexpr int MIN_CTAS = (VEC * UNROLL > 8) ? 2 : 4;
#endif
  static constexpr unsigned L2_LOADS = B200MT_ADAM_L2_LOADS, L2_STORES = B200MT_ADAM_L2_STORES;
  float beta1, beta2, bc1, bc2, eps, lr, decay, grad_scale;
  int mode;
  __device__ __forceinline__ unsigned load_mask() const { return 0b01111; }
  __device__ __forceinline__ void elem(State&, float (&x)[NL]) const {
    float p = x[1], m = x[2], v = x[3];
    adam_rule(x[0], p, m, v, beta1, beta2, bc1, bc2, eps, lr, decay, grad_scale, mode);
    x[1] = p; x[2] = m; x[3] = v;
    if constexpr (COPY) x[4] = p;
  }
};

// ---- distributed_adam_cuda.multi_tensor_fused_adam ---------------------------------------------
// lists [p32, m32, v32, g] (+ p_copy in g's dtype, or e5m2 bytes for the compressed all-gather:
// apex/contrib/optimizers/distributed_fused_adam.py:198,206); per-tensor hyper-parameters in device
// arrays; bias correction evaluated on the device per tile from `step`
// (multi_tensor_distopt_adam_kernel.cu:54-63,103-116).
template <typename Tg, typename Tc, bool COPY>
struct DistAdamOp : OpBase {
  using Types = typename std::conditional<COPY, TypeList<float, float, float, Tg, Tc>, TypeList<float, float, float, Tg>>::type;
  static constexpr int NL = COPY ? 5 : 4;
  static constexpr unsigned STORES = COPY ? 0b10111 : 0b00111;
  static constexpr bool MC = is_mc<Tg>::value || is_mc<Tc>::value;
  // peer-memory form: ONE vector per thread and 4 CTAs/SM -- 32.8 -> 23.0 ms at W = 2 against two vectors at 2 CTAs/SM (the
  // remote loads want resident warps, not registers); the multicast form is indifferent to it (profiles/nvls_probe_r01.txt)
  static constexpr bool PAIR = std::is_same<Tg, pair_bf16>::value;
  static constexpr int UNROLL = PAIR ? 1 : (MC ? B200MT_NVLS_UNROLL : B200MT_ADAM_UNROLL);
  static constexpr int VEC = MC ? B200MT_NVLS_VEC : B200MT_ADAM_VEC;
  static constexpr int THREADS = 256;
#ifdef B200MT_NVLS_MIN_CTAS
  static constexpr int MIN_CTAS = MC ? B200MT_NVLS_MIN_CTAS : ((VEC * UNROLL > 8) ? 2 : 4);
#else
  static constexpr int MIN_CTAS = (VEC * UNROLL > 8) ? 2 : 4;
#endif
  struct State { float b1, b2, eps, decay, bc1, bc2, rbc1, rbc2, rgs; };
  const float *pt_beta1, *pt_beta2, *pt_eps, *pt_decay;
  const int* pt_bias_correction;
  float lr, grad_scale;
  int step, mode;
  __device__ __forceinline__ unsigned load_mask() const { return 0b01111; }
  __device__ __forceinline__ void tile_begin(State& s, const TileLoc& L) const {
    const int t = L.tensor;
    s.b1 = __ldg(pt_beta1 + t); s.b2 = __ldg(pt_beta2 + t);
    s.eps = __ldg(pt_eps + t); s.decay = __ldg(pt_decay + t);
    if (__ldg(pt_bias_correction + t) == 1) {
      s.bc1 = bias_correction_dev(s.b1, step);
      s.bc2 = bias_correction_dev(s.b2, step);
    } else {
      s.bc1 = 1.f; s.bc2 = 1.f;
    }
    s.rbc1 = 1.f / s.bc1; s.rbc2 = 1.f / s.bc2; s.rgs = 1.f / grad_scale;
  }
  __device__ __forceinline__ void elem(State& s, float (&x)[NL]) const {
    float p = x[0], m = x[1], v = x[2];
    const float g = div_const(x[3], grad_scale, s.rgs);
    m = s.b1 * m + (1.f - s.b1) * g;
    v = s.b2 * v + (1.f - s.b2) * g * g;
    const float mu = div_const(m, s.bc1, s.rbc1), vu = div_const(v, s.bc2, s.rbc2);
    const float denom = (mode == 0) ? fast_sqrt(vu + s.eps) : fast_sqrt(vu) + s.eps;
    p = p - lr * (fast_div(mu, denom) + s.decay * p);
    x[0] = p; x[1] = m; x[2] = v;
    if constexpr (COPY) x[4] = p;
  }
};

// ---- fused_adam_cuda.adam_mt --------------------------------------------------------------------
// lists [p, m, v, g] (+ p_copy); step_size = lr*sqrt(bc2)/bc1 on the host, eps not rescaled
// (fused_adam_cuda_kernel.cu:131-146,298-306).
template <typename Tp, typename Tg, bool COPY>
struct ContribAdamOp : OpBase {
  using Types = typename std::conditional<COPY, TypeList<Tp, Tp, Tp, Tg, Tg>, TypeList<Tp, Tp, Tp, Tg>>::type;
  static constexpr int NL = COPY ? 5 : 4;
  static constexpr unsigned STORES = COPY ? 0b10111 : 0b00111;
  static constexpr int UNROLL = B200MT_ADAM_UNROLL;
  static constexpr int VEC = B200MT_ADAM_VEC;
  static constexpr int THREADS = 256;
  static constexpr int MIN_CTAS = (VEC * UNROLL > 8) ? 2 : 4;
  float b1, b2, eps, grad_scale, step_size, decay;
  int mode;
  __device__ __forceinline__ unsigned load_mask() const { return 0b01111; }
  __device__ __forceinline__ void elem(State&, float (&x)[NL]) const {
    float p = x[0], m = x[1], v = x[2];
    const float rgs = 1.f / grad_scale;   // loop-invariant
    const float g = div_const(x[3], grad_scale, rgs);
    m = b1 * m + (1.f - b1) * g;
    v = b2 * v + (1.f - b2) * g * g;
    const float denom = (mode == 0) ? fast_sqrt(v + eps) : fast_sqrt(v) + eps;
    const float update = fast_div(m, denom) + (decay * p);
    p = p - (step_size * update);
    x[0] = p; x[1] = m; x[2] = v;
    if constexpr (COPY) x[4] = p;
  }
};

}  // namespace b200mt

using namespace b200mt;

extern "C" int b200mt_adam(const b200mt_lists* L, int chunk_size, int32_t* noop, float lr, float beta1,
                           float beta2, float eps, int step, int mode, int bias_correction,
                           float weight_decay, float grad_scale, void* stream_) {
  (void)noop;  // multi_tensor_adam never touches the flag
  if (check_common(L, chunk_size)) return 1;
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  B200MT_CHECK(L->depth == 4 || L->depth == 5, "tensor_lists.size() != depth");
  const bool tma = adam_tma_wanted(L);     // all fp32 and 16-byte aligned: bulk-copy staging with its own tile length
  Tiling tiling = elementwise_tiling();
  if (tma) { tiling.tile_elems = adam_tma_tile(); tiling.chunk_elems = tiling.tile_elems; tiling.ppc = 1; }
  if (acquire_table(L, 4, 5, tiling, stream, &tv)) return 1;

  // host-side bias correction, evaluated in double and narrowed (csrc/multi_tensor_adam.cu:145-149)
  float bc1 = 1.f, bc2 = 1.f;
  if (bias_correction == 1) {
    bc1 = (float)(1.0 - std::pow((double)beta1, (double)step));
    bc2 = (float)(1.0 - std::pow((double)beta2, (double)step));
  }
  auto fill = [&](auto& op) {
    op.beta1 = beta1; op.beta2 = beta2; op.bc1 = bc1; op.bc2 = bc2; op.eps = eps; op.lr = lr;
    op.decay = weight_decay; op.grad_scale = grad_scale; op.mode = mode;
  };
  const int dg = L->dtype[0], dp = L->dtype[1];
  B200MT_CHECK(L->dtype[2] == dp && L->dtype[3] == dp, "adam: p, m and v must share one dtype");
  if (L->depth == 4) {
    if (dg == dp) {  // the reference's case: one dtype for all four lists
      if (tma) return adam_tma_launch(tv, beta1, beta2, bc1, bc2, eps, lr, weight_decay, grad_scale, mode, stream);
      return dispatch_f64_f32_f16_bf16(dg, "adam", [&](auto t) {
        using T = typename decltype(t)::type;
        AdamOp<T, T, T, false> op; fill(op);
        return launch_stream(tv, op, stream);
      });
    }
    B200MT_CHECK(dp == B200MT_F32, "adam: mixed-precision lists need fp32 p/m/v");
    return dispatch_f32_f16_bf16(dg, "adam", [&](auto t) {
      using T = typename decltype(t)::type;
      AdamOp<T, float, T, false> op; fill(op);
      return launch_stream(tv, op, stream);
    });
  }
  // depth 5: [g, p32, m32, v32, p_copy] single-pass mixed precision
  B200MT_CHECK(dp == B200MT_F32, "adam: the 5-list form needs fp32 p/m/v");
  const int dc = L->dtype[4];
  B200MT_CHECK(dc == B200MT_F16 || dc == B200MT_BF16, "adam: p_copy must be Half or BFloat16");
  return dispatch_f32_f16_bf16(dg, "adam", [&](auto tg) {
    using Tg = typename decltype(tg)::type;
    if (dc == B200MT_F16) { AdamOp<Tg, float, __half, true> op; fill(op); return launch_stream(tv, op, stream); }
    AdamOp<Tg, float, __nv_bfloat16, true> op; fill(op);
    return launch_stream(tv, op, stream);
  });
}

extern "C" int b200mt_dist_adam(const b200mt_lists* L, int chunk_size, int32_t* noop,
                                const float* pt_beta1, const float* pt_beta2,
                                const int32_t* pt_bias_correction, const float* pt_eps,
                                const float* pt_weight_decay, float lr, float grad_scale, int step,
                                int mode, void* stream_) {
  (void)noop;
  if (check_common(L, chunk_size)) return 1;
  B200MT_CHECK(pt_beta1 && pt_beta2 && pt_bias_correction && pt_eps && pt_weight_decay,
               "per-tensor hyper-parameter array is NULL");
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  if (acquire_table(L, 4, 5, elementwise_tiling(), stream, &tv)) return 1;
  B200MT_CHECK(L->dtype[0] == B200MT_F32 && L->dtype[1] == B200MT_F32 && L->dtype[2] == B200MT_F32,
               "dist_adam: p, m, v must be fp32");
  B200MT_CHECK(L->depth == 4 || L->dtype[4] == L->dtype[3] || L->dtype[4] == B200MT_U8,
               "dist_adam: p_copy must have the gradient dtype or be e5m2 bytes");
  auto fill = [&](auto& op) {
    op.pt_beta1 = pt_beta1; op.pt_beta2 = pt_beta2; op.pt_eps = pt_eps; op.pt_decay = pt_weight_decay;
    op.pt_bias_correction = pt_bias_correction; op.lr = lr; op.grad_scale = grad_scale;
    op.step = step; op.mode = mode;
  };
  return dispatch_f32_f16_bf16(L->dtype[3], "dist_adam_cuda_kernel", [&](auto tg) {
    using Tg = typename decltype(tg)::type;
    if (L->depth == 5 && L->dtype[4] == B200MT_U8) { DistAdamOp<Tg, e5m2, true> op; fill(op); return launch_stream(tv, op, stream); }
    if (L->depth == 5) { DistAdamOp<Tg, Tg, true> op; fill(op); return launch_stream(tv, op, stream); }
    DistAdamOp<Tg, Tg, false> op; fill(op);
    return launch_stream(tv, op, stream);
  });
}

// The sharded step as ONE kernel over NVSwitch multicast memory: list 3 (g) and list 4 (p_copy) are
// addresses inside multicast objects; the gradient load reduces over all ranks in the switch and the
// parameter store is replicated to all ranks by the switch.  Same math as b200mt_dist_adam.
extern "C" int b200mt_dist_adam_nvls(const b200mt_lists* L, int chunk_size, int32_t* noop,
                                     const float* pt_beta1, const float* pt_beta2,
                                     const int32_t* pt_bias_correction, const float* pt_eps,
                                     const float* pt_weight_decay, float lr, float grad_scale, int step,
                                     int mode, void* stream_) {
  (void)noop;
  if (check_common(L, chunk_size)) return 1;
  B200MT_CHECK(pt_beta1 && pt_beta2 && pt_bias_correction && pt_eps && pt_weight_decay,
               "per-tensor hyper-parameter array is NULL");
  B200MT_CHECK(L->depth == 5, "dist_adam_nvls: expected [p32, m32, v32, g_multicast, p_copy_multicast]");
  B200MT_CHECK(L->dtype[0] == B200MT_F32 && L->dtype[1] == B200MT_F32 && L->dtype[2] == B200MT_F32,
               "dist_adam_nvls: p, m, v must be fp32");
  B200MT_CHECK((L->dtype[3] == B200MT_BF16 || L->dtype[3] == B200MT_F16) &&
               (L->dtype[4] == L->dtype[3] || L->dtype[4] == B200MT_U8),
               "dist_adam_nvls: g must be Half or BFloat16 and p_copy the same type or e5m2 bytes");
  const bool packed = L->dtype[4] == B200MT_U8;
  // Measurement switches (tools/nvls_probe.py), carried in the upper bits of `mode`: list 3 / list 4 hold ORDINARY device
  // addresses instead of multicast ones, i.e. the kernel keeps only one of its two collectives --
  //   B200MT_NVLS_LOCAL_COPY:  in-switch reduction of the gradients + Adam, parameters stored locally (no broadcast)
  //   B200MT_NVLS_LOCAL_GRADS: gradients loaded locally (no reduction) + Adam + in-switch broadcast of the parameters
  // Three single launches whose times against the byte model say which direction of the fabric paces the full kernel.
  const bool local_grads = (mode & B200MT_NVLS_LOCAL_GRADS) != 0, local_copy = (mode & B200MT_NVLS_LOCAL_COPY) != 0;
  mode &= 0xff;
  B200MT_CHECK(!(local_grads && local_copy), "dist_adam_nvls: both lists local is b200mt_dist_adam");
  B200MT_CHECK(!(packed && (local_grads || local_copy)), "dist_adam_nvls: the measurement switches take 16-bit copies only");
  for (int t = 0; t < L->ntensors; t++) {
    B200MT_CHECK(L->numel[t] % B200MT_NVLS_VEC == 0, "dist_adam_nvls: fragment %d has %lld elements; multicast fragments "
                 "must be multiples of %d elements", t, (long long)L->numel[t], B200MT_NVLS_VEC);
    for (int d = 0; d < 5; d++) {
      const int al = (d < 3 ? 4 : (d == 4 && packed) ? 1 : 2) * B200MT_NVLS_VEC;
      B200MT_CHECK(((uintptr_t)L->ptrs[(size_t)d * L->ntensors + t]) % al == 0,
                   "dist_adam_nvls: list %d fragment %d is not %d-byte aligned", d, t, al);
    }
  }
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  if (acquire_table(L, 5, 5, elementwise_tiling(), stream, &tv)) return 1;
  auto run = [&](auto op) {
    op.pt_beta1 = pt_beta1; op.pt_beta2 = pt_beta2; op.pt_eps = pt_eps; op.pt_decay = pt_weight_decay;
    op.pt_bias_correction = pt_bias_correction; op.lr = lr; op.grad_scale = grad_scale;
    op.step = step; op.mode = mode;
    return launch_stream(tv, op, stream);
  };
  if (L->dtype[3] == B200MT_BF16) {
    if (local_copy) return run(DistAdamOp<mc_bf16, __nv_bfloat16, true>());
    if (local_grads) return run(DistAdamOp<__nv_bfloat16, mc_bf16, true>());
    return packed ? run(DistAdamOp<mc_bf16, mc_e5m2, true>()) : run(DistAdamOp<mc_bf16, mc_bf16, true>());
  }
  B200MT_CHECK(!(local_grads || local_copy), "dist_adam_nvls: the measurement switches are built for BFloat16 only");
  return packed ? run(DistAdamOp<mc_f16, mc_e5m2, true>()) : run(DistAdamOp<mc_f16, mc_f16, true>());
}

// ---- b200mt_dist_adam_p2p: the sharded step of TWO ranks as one kernel over peer-mapped memory (extension).
// lists = [p32, m32, v32, g, p_copy] with g and p_copy THIS rank's addresses inside buffers that have a twin in the peer's
// memory at the given byte distances.  Gradients: own + peer's, summed on load; parameters: stored here and in the peer.
extern "C" int b200mt_dist_adam_p2p(const b200mt_lists* L, int chunk_size, int32_t* noop,
                                    const float* pt_beta1, const float* pt_beta2,
                                    const int32_t* pt_bias_correction, const float* pt_eps,
                                    const float* pt_weight_decay, float lr, float grad_scale, int step,
                                    int mode, int64_t grad_peer_delta, int64_t copy_peer_delta, void* stream_) {
  (void)noop;
  if (check_common(L, chunk_size)) return 1;
  B200MT_CHECK(pt_beta1 && pt_beta2 && pt_bias_correction && pt_eps && pt_weight_decay,
               "per-tensor hyper-parameter array is NULL");
  B200MT_CHECK(L->depth == 5, "dist_adam_p2p: expected [p32, m32, v32, g, p_copy]");
  B200MT_CHECK(L->dtype[0] == B200MT_F32 && L->dtype[1] == B200MT_F32 && L->dtype[2] == B200MT_F32,
               "dist_adam_p2p: p, m, v must be fp32");
  B200MT_CHECK(L->dtype[3] == B200MT_BF16 && L->dtype[4] == B200MT_BF16, "dist_adam_p2p: g and p_copy must be BFloat16");
  B200MT_CHECK(grad_peer_delta % 16 == 0 && copy_peer_delta % 16 == 0 && grad_peer_delta != 0 && copy_peer_delta != 0,
               "dist_adam_p2p: the peer distances must be non-zero multiples of 16 bytes");
  for (int t = 0; t < L->ntensors; t++) {
    B200MT_CHECK(L->numel[t] % B200MT_NVLS_VEC == 0, "dist_adam_p2p: fragment %d has %lld elements; fragments must be "
                 "multiples of %d elements", t, (long long)L->numel[t], B200MT_NVLS_VEC);
    for (int d = 0; d < 5; d++) {
      const int al = (d < 3 ? 4 : 2) * B200MT_NVLS_VEC;
      B200MT_CHECK(((uintptr_t)L->ptrs[(size_t)d * L->ntensors + t]) % al == 0,
                   "dist_adam_p2p: list %d fragment %d is not %d-byte aligned", d, t, al);
    }
  }
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  if (acquire_table(L, 5, 5, elementwise_tiling(), stream, &tv)) return 1;
  const long long delta[2] = {(long long)grad_peer_delta, (long long)copy_peer_delta};
  // stream-ordered: kernels launched earlier on this stream keep the distances they were launched with
  B200MT_CUDA(cudaMemcpyToSymbolAsync(c_pair_delta, delta, sizeof(delta), 0, cudaMemcpyHostToDevice, stream));
  DistAdamOp<pair_bf16, pair_bf16, true> op;
  op.pt_beta1 = pt_beta1; op.pt_beta2 = pt_beta2; op.pt_eps = pt_eps; op.pt_decay = pt_weight_decay;
  op.pt_bias_correction = pt_bias_correction; op.lr = lr; op.grad_scale = grad_scale;
  op.step = step; op.mode = mode;
  return launch_stream(tv, op, stream, /*allow_pdl=*/false);   // the constant write above must be complete before the kernel starts
}

extern "C" int b200mt_contrib_adam(const b200mt_lists* L, int chunk_size, int32_t* noop, float lr,
                                   float beta1, float beta2, float eps, float grad_scale, int step,
                                   int mode, int bias_correction, float decay, void* stream_) {
  (void)noop;
  if (check_common(L, chunk_size)) return 1;
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  if (acquire_table(L, 4, 5, elementwise_tiling(), stream, &tv)) return 1;
  // fused_adam_cuda_kernel.cu:298-306
  float step_size = lr;
  if (bias_correction == 1) {
    const float bc1 = 1.f - std::pow(beta1, step);
    const float bc2 = 1.f - std::pow(beta2, step);
    step_size = lr * std::sqrt(bc2) / bc1;
  }
  const int dp = L->dtype[0], dg = L->dtype[3];
  B200MT_CHECK(L->dtype[1] == dp && L->dtype[2] == dp, "adam_mt: p, m, v must share one dtype");
  B200MT_CHECK(L->depth == 4 || L->dtype[4] == dg, "adam_mt: p_copy must have the gradient dtype");
  auto fill = [&](auto& op) {
    op.b1 = beta1; op.b2 = beta2; op.eps = eps; op.grad_scale = grad_scale; op.step_size = step_size;
    op.decay = decay; op.mode = mode;
  };
  B200MT_CHECK(dp == B200MT_F32, "adam_mt: p, m, v must be fp32");
  return dispatch_f32_f16_bf16(dg, "adam_cuda_mt_kernel", [&](auto tg) {
    using Tg = typename decltype(tg)::type;
    if (L->depth == 5) { ContribAdamOp<float, Tg, true> op; fill(op); return launch_stream(tv, op, stream); }
    ContribAdamOp<float, Tg, false> op; fill(op);
    return launch_stream(tv, op, stream);
  });
}
