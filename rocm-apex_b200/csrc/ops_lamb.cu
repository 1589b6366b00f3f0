// ops_lamb.cu -- LAMB: global-norm clip, Adam-style update, per-tensor trust ratio.
//   b200mt_lamb        <- amp_C.multi_tensor_lamb     (csrc/multi_tensor_lamb.cu:40-413)
//   b200mt_lamb_mp     <- amp_C.multi_tensor_lamb_mp  (csrc/multi_tensor_lamb_mp.cu:40-496)
//   b200mt_lamb_stage1 <- amp_C.multi_tensor_lamb_stage1_cuda (csrc/multi_tensor_lamb_stage_1.cu:17-151)
//   b200mt_lamb_stage2 <- amp_C.multi_tensor_lamb_stage2_cuda (csrc/multi_tensor_lamb_stage_2.cu:20-125)
//
// The reference runs four passes per call: l2norm(p), stage 1, l2norm(update), stage 2, each of
// them many launches.  Here stage 1 also produces the per-tile sums of squares of p and of the
// update it writes, so a call is three launches (stage 1, norm second pass, stage 2) and two full
// reads fewer (40 instead of 52 B/param in fp32).
#include <cmath>
#include "mt_dispatch.h"
#include "mt_norm.cuh"

// stage 1: 8 elements per access x 1 in flight beats 4 x 2 by 7 % on the mixed-precision lists and
// 2 % on fp32 (profiles/sweep_variants_r01.txt)
#ifndef B200MT_LAMB_VEC
#define B200MT_LAMB_VEC 8
#endif
#ifndef B200MT_LAMB_UNROLL
#define B200MT_LAMB_UNROLL 1
#endif

namespace b200mt {

// ---- stage 1 (fused with both norm partials) ----------------------------------------------------
// lists [g, p, m, v] (a 5th list, if present, is ignored here).  MP: grads are unscaled by
// *inv_scale, clip threshold and step live on the device, and everything is gated on *noop.
template <typename Tg, typename Tp, int DEPTH, bool MP>
struct LambStage1Op : OpBase {
  using Types = typename std::conditional<DEPTH == 5, TypeList<Tg, Tp, Tp, Tp, Tg>, TypeList<Tg, Tp, Tp, Tp>>::type;
  static constexpr int NL = DEPTH;
  static constexpr unsigned STORES = 0b01101;  // g <- update, m, v
  static constexpr int UNROLL = B200MT_LAMB_UNROLL;
  static constexpr int VEC = B200MT_LAMB_VEC;
  static constexpr int THREADS = 256;
  static constexpr int MIN_CTAS = (VEC * UNROLL > 8) ? 2 : 4;
  struct State { float clip, inv_scale, bc1, bc2, rclip, rbc1, rbc2, acc_p, acc_u; };
  float beta1, beta2, beta3, eps, decay;
  float bc1_host, bc2_host, max_norm_host;
  int mode, bias_correction, want_norms;
  const float* global_grad_norm;
  const float* max_norm_dev;   // MP
  const float* inv_scale_dev;  // MP
  const int* step_dev;         // MP
  const int* noop;
  NormFinish fin;              // partial slots [2][ntiles] (set 0 = p, set 1 = update) + the folded second pass

  __device__ __forceinline__ bool begin(State& s) const {
    s.acc_p = 0.f; s.acc_u = 0.f;
    if constexpr (MP) {
      if (*reinterpret_cast<const volatile int*>(noop) != 0) return false;
      const float gn = __ldg(global_grad_norm), mx = __ldg(max_norm_dev);
      s.clip = gn > mx ? gn / mx : 1.f;
      s.inv_scale = __ldg(inv_scale_dev);
      if (bias_correction == 1) {
        const int step = __ldg(step_dev);
        s.bc1 = bias_correction_dev(beta1, step);
        s.bc2 = bias_correction_dev(beta2, step);
      } else {
        s.bc1 = 1.f; s.bc2 = 1.f;
      }
    } else {
      const float gn = __ldg(global_grad_norm);
      s.clip = gn > max_norm_host ? gn / max_norm_host : 1.f;
      s.inv_scale = 1.f;
      s.bc1 = bc1_host; s.bc2 = bc2_host;
    }
    s.rclip = 1.f / s.clip; s.rbc1 = 1.f / s.bc1; s.rbc2 = 1.f / s.bc2;
    return true;
  }
  // p is needed by the math only when decay != 0 (csrc/multi_tensor_lamb.cu:101-102) and by the
  // trust ratio only when norms are wanted; otherwise it is not read at all.
  __device__ __forceinline__ unsigned load_mask() const {
    return (decay != 0.f || want_norms) ? 0b01111u : 0b01101u;
  }
  __device__ __forceinline__ void elem(State& s, float (&x)[NL]) const {
    float g = x[0], m = x[2], v = x[3];
    const float p = x[1];
    if constexpr (MP) g = g * s.inv_scale;
    float u;
    if (mode == 0) {  // L2 on the clipped gradient
      float sg = div_const(g, s.clip, s.rclip);
      sg = sg + decay * p;
      m = m * beta1 + beta3 * sg;
      v = v * beta2 + (1.f - beta2) * sg * sg;
      const float mu = div_const(m, s.bc1, s.rbc1), vu = div_const(v, s.bc2, s.rbc2);
      u = fast_div(mu, fast_sqrt(vu) + eps);
    } else {          // decoupled decay
      const float sg = div_const(g, s.clip, s.rclip);
      m = m * beta1 + beta3 * sg;
      v = v * beta2 + (1.f - beta2) * sg * sg;
      const float mu = div_const(m, s.bc1, s.rbc1), vu = div_const(v, s.bc2, s.rbc2);
      u = fast_div(mu, fast_sqrt(vu) + eps) + decay * p;
    }
    // the reference takes ||update|| over the values as stored in g's dtype
    const float ur = Elem<Tg>::round_trip(u);
    s.acc_p += p * p;
    s.acc_u += ur * ur;
    x[0] = u; x[2] = m; x[3] = v;
  }
  __device__ __forceinline__ void tile_end(State& s, int tile, const TileLoc& L, float* smem) const {
    if (!want_norms) return;
    const float tp = cta_sum<THREADS>(s.acc_p, smem);
    const float tu = cta_sum<THREADS>(s.acc_u, smem);
    const float part[2] = {tp, tu};
    norm_tile_done<THREADS, 2>(fin, tile, L.tensor, part, smem);
    s.acc_p = 0.f; s.acc_u = 0.f;
  }
};

// ---- stage 2 -------------------------------------------------------------------------------------
// Same table as stage 1: reads update (list 0) and p (list 1), writes p and, with 5 lists, the
// low-precision model copy (list 4).  ratio per csrc/multi_tensor_lamb.cu:255-263.
template <typename Tg, typename Tp, int DEPTH, bool MP>
struct LambStage2Op : OpBase {
  using Types = typename std::conditional<DEPTH == 5, TypeList<Tg, Tp, Tp, Tp, Tg>, TypeList<Tg, Tp, Tp, Tp>>::type;
  static constexpr int NL = DEPTH;
  static constexpr unsigned STORES = DEPTH == 5 ? 0b10010 : 0b00010;
  static constexpr int UNROLL = 4;
  static constexpr int THREADS = 256;
  static constexpr int MIN_CTAS = 4;
  struct State { float ratio; };
  const float* param_norm;
  const float* update_norm;
  float lr_host, decay;
  const float* lr_dev;  // MP
  int use_nvlamb;
  const int* noop;
  __device__ __forceinline__ bool begin(State&) const {
    if constexpr (MP) return *reinterpret_cast<const volatile int*>(noop) == 0;
    return true;
  }
  __device__ __forceinline__ unsigned load_mask() const { return 0b00011u; }
  __device__ __forceinline__ void tile_begin(State& s, const TileLoc& L) const {
    const float lr = MP ? __ldg(lr_dev) : lr_host;
    float ratio = lr;
    if (use_nvlamb || decay != 0.f) {
      const float pn = param_norm[L.tensor], un = update_norm[L.tensor];
      ratio = (un != 0.f && pn != 0.f) ? lr * (pn / un) : lr;
    }
    s.ratio = ratio;
  }
  __device__ __forceinline__ void elem(State& s, float (&x)[NL]) const {
    const float p = x[1] - s.ratio * x[0];
    x[1] = p;
    if constexpr (DEPTH == 5) x[4] = p;
  }
};

// ---- split API: stage 1 with explicit update buffer and per-tensor decay --------------------------
// lists [g, p, m, v, update].  The reference evaluates this in the parameter dtype; here it is fp32
// for every dtype (identical for fp32 tensors, more accurate for 16-bit ones).
template <typename Tg, typename Tp, typename Tu>
struct LambSplitStage1Op : OpBase {
  using Types = TypeList<Tg, Tp, Tp, Tp, Tu>;
  static constexpr int NL = 5;
  static constexpr unsigned STORES = 0b11100;
  static constexpr int UNROLL = 2;
  static constexpr int THREADS = 256;
  static constexpr int MIN_CTAS = 4;
  struct State { float decay; };
  const float* per_tensor_decay;
  float beta1, beta2, bc1, bc2, eps, clip;
  __device__ __forceinline__ unsigned load_mask() const { return 0b01111u; }
  __device__ __forceinline__ void tile_begin(State& s, const TileLoc& L) const { s.decay = __ldg(per_tensor_decay + L.tensor); }
  __device__ __forceinline__ void elem(State& s, float (&x)[NL]) const {
    const float rclip = 1.f / clip, rbc1 = 1.f / bc1, rbc2 = 1.f / bc2;   // loop-invariant
    const float sg = div_const(x[0], clip, rclip);
    float m = x[2], v = x[3];
    m = m * beta1 + (1.f - beta1) * sg;
    v = v * beta2 + (1.f - beta2) * sg * sg;
    const float mu = div_const(m, bc1, rbc1), vu = div_const(v, bc2, rbc2);
    x[4] = fast_div(mu, fast_sqrt(vu) + eps) + s.decay * x[1];
    x[2] = m; x[3] = v;
  }
};

// lists [p, update]
template <typename Tp, typename Tu>
struct LambSplitStage2Op : OpBase {
  using Types = TypeList<Tp, Tu>;
  static constexpr int NL = 2;
  static constexpr unsigned STORES = 0b01;
  static constexpr int UNROLL = 4;
  struct State { float ratio; };
  const float* param_norm;
  const float* update_norm;
  float lr, decay;
  int use_nvlamb;
  __device__ __forceinline__ unsigned load_mask() const { return 0b11u; }
  __device__ __forceinline__ void tile_begin(State& s, const TileLoc& L) const {
    float ratio = lr;
    if (use_nvlamb || decay != 0.f) {
      const float pn = param_norm[L.tensor], un = update_norm[L.tensor];
      ratio = (un != 0.f && pn != 0.f) ? lr * (pn / un) : lr;
    }
    s.ratio = ratio;
  }
  __device__ __forceinline__ void elem(State& s, float (&x)[NL]) const { x[0] = x[0] - s.ratio * x[1]; }
};

struct LambArgs {
  float lr_host; const float* lr_dev;
  float beta1, beta2, beta3, eps, decay;
  float bc1, bc2; int bias_correction; const int* step_dev;
  int mode, use_nvlamb;
  const float* global_grad_norm; float max_norm_host; const float* max_norm_dev;
  const float* inv_scale_dev;
  int* noop;
};

template <typename Tg, typename Tp, int DEPTH, bool MP>
int run_lamb(const TableView& tv, const LambArgs& a, cudaStream_t stream) {
  // Both per-tensor norms are always taken, as in the reference: even when the trust ratio does
  // not use them they can raise the overflow flag (csrc/multi_tensor_lamb.cu:370,394).
  const int want_norms = 1;
  void* raw;
  if (acquire_scratch(stream, norm_scratch_bytes(2, tv.ntiles, tv.ntensors), &raw)) return 1;
  NormScratch ns = carve_norm_scratch(raw, 2, tv.ntiles, tv.ntensors);

  LambStage1Op<Tg, Tp, DEPTH, MP> s1;
  s1.beta1 = a.beta1; s1.beta2 = a.beta2; s1.beta3 = a.beta3; s1.eps = a.eps; s1.decay = a.decay;
  s1.bc1_host = a.bc1; s1.bc2_host = a.bc2; s1.max_norm_host = a.max_norm_host;
  s1.mode = a.mode; s1.bias_correction = a.bias_correction; s1.want_norms = want_norms;
  s1.global_grad_norm = a.global_grad_norm; s1.max_norm_dev = a.max_norm_dev;
  s1.inv_scale_dev = a.inv_scale_dev; s1.step_dev = a.step_dev; s1.noop = a.noop;
  float* pn = ns.norms;
  float* un = ns.norms + tv.ntensors;
  // per-tensor ||p|| and ||update||: reduced by the last CTA of each tensor inside stage 1 (two launches per LAMB step)
  const bool folded = make_norm_finish(tv, ns, 2, a.noop, MP ? 1 : 0, nullptr, pn, nullptr, un, tv.nonempty, &s1.fin);
  if (launch_stream(tv, s1, stream)) return 1;

  if (want_norms && !folded) {
    // per-tensor ||p|| and ||update||; a non-finite chunk sum raises the flag, as the reference's
    // two per-tensor l2norm passes would (csrc/multi_tensor_lamb.cu:370,394)
    if (launch_norm_cleanup(tv, ns, 2, a.noop, MP ? 1 : 0, nullptr, pn, nullptr, un, stream)) return 1;
  }

  LambStage2Op<Tg, Tp, DEPTH, MP> s2;
  s2.param_norm = pn; s2.update_norm = un; s2.lr_host = a.lr_host; s2.lr_dev = a.lr_dev;
  s2.decay = a.decay; s2.use_nvlamb = a.use_nvlamb; s2.noop = a.noop;
  return launch_stream(tv, s2, stream);
}

}  // namespace b200mt

using namespace b200mt;

extern "C" int b200mt_lamb(const b200mt_lists* L, int chunk_size, int32_t* noop, float lr, float beta1,
                           float beta2, float eps, int step, int bias_correction, float weight_decay,
                           int grad_averaging, int mode, const float* global_grad_norm,
                           float max_grad_norm, int use_nvlamb, void* stream_) {
  if (check_common(L, chunk_size)) return 1;
  B200MT_CHECK(noop != nullptr && global_grad_norm != nullptr, "noop / global_grad_norm pointer is NULL");
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  if (acquire_table(L, 4, 4, chunked_tiling(chunk_size), stream, &tv)) return 1;
  const int d = L->dtype[0];
  B200MT_CHECK(L->dtype[1] == d && L->dtype[2] == d && L->dtype[3] == d,
               "multi_tensor_lamb: all four lists must share one dtype");
  LambArgs a{};
  a.lr_host = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.decay = weight_decay;
  a.beta3 = grad_averaging == 1 ? 1.f - beta1 : 1.f;
  a.bc1 = 1.f; a.bc2 = 1.f;
  if (bias_correction == 1) {  // csrc/multi_tensor_lamb.cu:356-360
    a.bc1 = (float)(1.0 - std::pow((double)beta1, (double)step));
    a.bc2 = (float)(1.0 - std::pow((double)beta2, (double)step));
  }
  a.bias_correction = bias_correction; a.mode = mode; a.use_nvlamb = use_nvlamb;
  a.global_grad_norm = global_grad_norm; a.max_norm_host = max_grad_norm; a.noop = noop;
  return dispatch_f32_f16_bf16(d, "lamb_stage_1", [&](auto t) {
    using T = typename decltype(t)::type;
    return run_lamb<T, T, 4, false>(tv, a, stream);
  });
}

extern "C" int b200mt_lamb_mp(const b200mt_lists* L, int chunk_size, int32_t* noop, const float* lr,
                              float beta1, float beta2, float eps, const int32_t* step,
                              int bias_correction, float weight_decay, int grad_averaging, int mode,
                              const float* global_grad_norm, const float* max_grad_norm,
                              int use_nvlamb, const float* found_inf, const float* inv_scale,
                              void* stream_) {
  (void)found_inf;  // passed through the reference API but unused by its kernels as well
  if (check_common(L, chunk_size)) return 1;
  B200MT_CHECK(noop && lr && step && global_grad_norm && max_grad_norm && inv_scale,
               "multi_tensor_lamb_mp: a device scalar pointer is NULL");
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  if (acquire_table(L, 4, 5, chunked_tiling(chunk_size), stream, &tv)) return 1;
  const int dg = L->dtype[0], dp = L->dtype[1];
  B200MT_CHECK(L->dtype[2] == dp && L->dtype[3] == dp, "multi_tensor_lamb_mp: p, m, v must share one dtype");
  LambArgs a{};
  a.lr_dev = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.decay = weight_decay;
  a.beta3 = grad_averaging == 1 ? 1.f - beta1 : 1.f;
  a.bias_correction = bias_correction; a.step_dev = step; a.mode = mode; a.use_nvlamb = use_nvlamb;
  a.global_grad_norm = global_grad_norm; a.max_norm_dev = max_grad_norm; a.inv_scale_dev = inv_scale;
  a.noop = noop;
  if (L->depth == 4) {
    // FP32 (or all-fp16) model params, no master copy: g and p share a dtype (:404-426)
    B200MT_CHECK(dg == dp, "multi_tensor_lamb_mp: the 4-list form needs g and p of one dtype");
    return dispatch_f32_f16_bf16(dg, "lamb_stage_1", [&](auto t) {
      using T = typename decltype(t)::type;
      return run_lamb<T, T, 4, true>(tv, a, stream);
    });
  }
  B200MT_CHECK(dp == B200MT_F32, "multi_tensor_lamb_mp: the 5-list form needs fp32 master p/m/v");
  B200MT_CHECK(L->dtype[4] == dg, "multi_tensor_lamb_mp: the model-param copy must have the gradient dtype");
  return dispatch_f32_f16_bf16(dg, "lamb_stage_1", [&](auto t) {
    using T = typename decltype(t)::type;
    return run_lamb<T, float, 5, true>(tv, a, stream);
  });
}

extern "C" int b200mt_lamb_stage1(const b200mt_lists* L, int chunk_size, int32_t* noop,
                                  const float* per_tensor_decay, int step, float beta1, float beta2,
                                  float eps, float clipped_global_grad_norm, void* stream_) {
  (void)noop;
  if (check_common(L, chunk_size)) return 1;
  B200MT_CHECK(per_tensor_decay != nullptr, "per_tensor_decay is NULL");
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  if (acquire_table(L, 5, 5, elementwise_tiling(), stream, &tv)) return 1;
  B200MT_CHECK(L->dtype[2] == L->dtype[1] && L->dtype[3] == L->dtype[1], "lamb_stage_1: p, m, v must share one dtype");
  // csrc/multi_tensor_lamb_stage_1.cu:128-130: float pow with step+1
  const float next_step = float(step + 1);
  const float bc1 = 1.0f - std::pow(beta1, next_step);
  const float bc2 = 1.0f - std::pow(beta2, next_step);
  return dispatch_f32_f16_bf16(L->dtype[0], "lamb_stage_1", [&](auto tg) {
    return dispatch_f32_f16_bf16(L->dtype[1], "lamb_stage_1", [&](auto tp) {
      return dispatch_f32_f16_bf16(L->dtype[4], "lamb_stage_1", [&](auto tu) {
        LambSplitStage1Op<typename decltype(tg)::type, typename decltype(tp)::type, typename decltype(tu)::type> op;
        op.per_tensor_decay = per_tensor_decay; op.beta1 = beta1; op.beta2 = beta2; op.bc1 = bc1; op.bc2 = bc2;
        op.eps = eps; op.clip = clipped_global_grad_norm;
        return launch_stream(tv, op, stream);
      });
    });
  });
}

extern "C" int b200mt_lamb_stage2(const b200mt_lists* L, int chunk_size, int32_t* noop,
                                  const float* per_tensor_param_norm, const float* per_tensor_update_norm,
                                  float lr, float weight_decay, int use_nvlamb, void* stream_) {
  (void)noop;
  if (check_common(L, chunk_size)) return 1;
  B200MT_CHECK(per_tensor_param_norm && per_tensor_update_norm, "per-tensor norm array is NULL");
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  if (acquire_table(L, 2, 2, elementwise_tiling(), stream, &tv)) return 1;
  return dispatch_f32_f16_bf16(L->dtype[0], "lamb_stage_2", [&](auto tp) {
    return dispatch_f32_f16_bf16(L->dtype[1], "lamb_stage_2", [&](auto tu) {
      LambSplitStage2Op<typename decltype(tp)::type, typename decltype(tu)::type> op;
      op.param_norm = per_tensor_param_norm; op.update_norm = per_tensor_update_norm;
      op.lr = lr; op.decay = weight_decay; op.use_nvlamb = use_nvlamb;
      return launch_stream(tv, op, stream);
    });
  });
}
