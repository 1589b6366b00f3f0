// ops_dist_lamb.cu -- the two kernels of the sharded LAMB step (DistributedFusedLAMB).
//   b200mt_dist_lamb_stage1 <- distributed_lamb_cuda.multi_tensor_lamb_compute_update_term
//                              (apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb_kernel.cu:108-316,432-470)
//   b200mt_dist_lamb_stage2 <- distributed_lamb_cuda.multi_tensor_lamb_update_weights (:318-430,472-506)
// Everything the host would have to synchronise for lives on the device: step counter, loss scale, global
// gradient norm and learning rate are device scalars, and both kernels return at once when *noop == 1.
// Stage 1 can also leave the per-fragment sum of squares of the update it writes (the reference runs a
// separate per-tensor l2norm over the update fragments for that, distributed_fused_lamb.py:686-692).
#include "mt_dispatch.h"
#include "mt_norm.cuh"

namespace b200mt {

// lists [g, p, m, v, u]
template <typename Tg, typename Tp, typename Tu>
struct DistLambStage1Op : OpBase {
  using Types = TypeList<Tg, Tp, Tp, Tp, Tu>;
  static constexpr int NL = 5;
  static constexpr unsigned STORES = 0b11100;
  static constexpr int VEC = 8;
  static constexpr int UNROLL = 1;
  static constexpr int THREADS = 256;
  static constexpr int MIN_CTAS = 4;
  struct State { float combined_scale, rcs, b1, b2, b3, bc1, bc2, rbc1, rbc2, eps, decay, acc_u; int step; };
  const float *pt_beta1, *pt_beta2, *pt_eps, *pt_decay;
  const int* pt_bias_correction;
  const int* step_dev;
  const float* global_scale;
  const float* global_grad_norm;
  float max_grad_norm;
  int mode;
  const int* noop;
  float* partials;   // per-tile sum of squares of u, or NULL

  __device__ __forceinline__ bool begin(State& s) const {
    if (*reinterpret_cast<const volatile int*>(noop) == 1) return false;
    const float gs = *global_scale;
    float cs = gs;
    if (max_grad_norm > 0.f) {   // multi_tensor_distopt_lamb_kernel.cu:136-140
      cs = max_grad_norm / (*global_grad_norm / gs + 1e-6f);
      cs = gs / fminf(1.0f, cs);
    }
    s.combined_scale = cs; s.rcs = 1.f / cs;
    s.step = *step_dev;
    s.acc_u = 0.f;
    return true;
  }
  __device__ __forceinline__ unsigned load_mask() const { return 0b01111u; }
  __device__ __forceinline__ void tile_begin(State& s, const TileLoc& L) const {
    const int t = L.tensor;
    s.b1 = __ldg(pt_beta1 + t); s.b2 = __ldg(pt_beta2 + t);
    s.b3 = 1.f - s.b1;                       // the reference recomputes beta3 and ignores the array (:144)
    s.eps = __ldg(pt_eps + t); s.decay = __ldg(pt_decay + t);
    if (__ldg(pt_bias_correction + t) == 1) {
      s.bc1 = bias_correction_dev(s.b1, s.step);
      s.bc2 = bias_correction_dev(s.b2, s.step);
    } else {
      s.bc1 = 1.f; s.bc2 = 1.f;
    }
    s.rbc1 = 1.f / s.bc1; s.rbc2 = 1.f / s.bc2;
  }
  __device__ __forceinline__ void elem(State& s, float (&x)[NL]) const {
    const float p = (s.decay != 0.f) ? x[1] : 0.f;   // p is only used when decay != 0 (:192-203)
    float m = x[2], v = x[3], u;
    float sg = div_const(x[0], s.combined_scale, s.rcs);
    if (mode == 0) {
      sg = sg + s.decay * p;
      m = m * s.b1 + s.b3 * sg;
      v = v * s.b2 + (1.f - s.b2) * sg * sg;
      u = fast_div(div_const(m, s.bc1, s.rbc1), fast_sqrt(div_const(v, s.bc2, s.rbc2)) + s.eps);
    } else {
      m = m * s.b1 + s.b3 * sg;
      v = v * s.b2 + (1.f - s.b2) * sg * sg;
      u = fast_div(div_const(m, s.bc1, s.rbc1), fast_sqrt(div_const(v, s.bc2, s.rbc2)) + s.eps) + s.decay * p;
    }
    const float ur = Elem<Tu>::round_trip(u);
    s.acc_u += ur * ur;
    x[2] = m; x[3] = v; x[4] = u;
  }
  __device__ __forceinline__ void tile_end(State& s, int tile, const TileLoc&, float* smem) const {
    if (!partials) return;
    const float t = cta_sum<THREADS>(s.acc_u, smem);
    if (threadIdx.x == 0) partials[tile] = t;
    s.acc_u = 0.f;
  }
};

// lists [u, p, p_copy]
template <typename Tu, typename Tp, typename Tc>
struct DistLambStage2Op : OpBase {
  using Types = TypeList<Tu, Tp, Tc>;
  static constexpr int NL = 3;
  static constexpr unsigned STORES = 0b110;
  static constexpr int VEC = 8;
  static constexpr int UNROLL = 2;
  struct State { float ratio; };
  const float* param_norm;          // per fragment
  const float* update_norm;         // per parameter, indexed through update_norm_offset
  const long long* update_norm_offset;
  const float* lr;
  const float* pt_decay;
  int use_nvlamb;
  const int* noop;
  __device__ __forceinline__ bool begin(State&) const { return *reinterpret_cast<const volatile int*>(noop) != 1; }
  __device__ __forceinline__ unsigned load_mask() const { return 0b011u; }
  __device__ __forceinline__ void tile_begin(State& s, const TileLoc& L) const {
    const int t = L.tensor;
    const float l = *lr;
    float ratio = l;
    if (use_nvlamb || pt_decay[t] != 0.f) {
      const float pn = param_norm[t], un = update_norm[update_norm_offset[t]];
      ratio = (un != 0.f && pn != 0.f) ? l * (pn / un) : l;
    }
    s.ratio = ratio;
  }
  __device__ __forceinline__ void elem(State& s, float (&x)[NL]) const {
    const float p = x[1] - s.ratio * x[0];
    x[1] = p; x[2] = p;
  }
};

}  // namespace b200mt

using namespace b200mt;

extern "C" int b200mt_dist_lamb_stage1(const b200mt_lists* L, int chunk_size, int32_t* noop,
                                       const float* pt_beta1, const float* pt_beta2, const float* pt_beta3,
                                       const int32_t* pt_bias_correction, const int32_t* step,
                                       const float* pt_eps, int mode, const float* pt_decay,
                                       const float* global_scale, const float* global_grad_norm,
                                       float max_grad_norm, float* update_sumsq, void* stream_) {
  (void)pt_beta3;
  if (check_common(L, chunk_size)) return 1;
  B200MT_CHECK(noop && pt_beta1 && pt_beta2 && pt_bias_correction && step && pt_eps && pt_decay && global_scale &&
               global_grad_norm, "dist_lamb_stage1: a device pointer is NULL");
  B200MT_CHECK(L->depth == 5, "tensor_lists.size() != depth");
  B200MT_CHECK(L->dtype[2] == L->dtype[1] && L->dtype[3] == L->dtype[1], "lamb_stage_1: p, m, v must share one dtype");
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  const Tiling tiling = update_sumsq ? chunked_tiling(chunk_size) : elementwise_tiling();
  if (acquire_table(L, 5, 5, tiling, stream, &tv)) return 1;
  NormScratch ns{};
  if (update_sumsq) {
    void* raw;
    if (acquire_scratch(stream, norm_scratch_bytes(1, tv.ntiles, tv.ntensors), &raw)) return 1;
    ns = carve_norm_scratch(raw, 1, tv.ntiles, tv.ntensors);
  }
  // the reference dispatches float/half for g, p and u (:449-451); bf16 grads are accepted as a superset
  int rc = dispatch_f32_f16_bf16(L->dtype[0], "lamb_stage_1", [&](auto tg) {
    return dispatch_f32_f16(L->dtype[1], "lamb_stage_1", [&](auto tp) {
      return dispatch_f32_f16(L->dtype[4], "lamb_stage_1", [&](auto tu) {
        DistLambStage1Op<typename decltype(tg)::type, typename decltype(tp)::type, typename decltype(tu)::type> op;
        op.pt_beta1 = pt_beta1; op.pt_beta2 = pt_beta2; op.pt_eps = pt_eps; op.pt_decay = pt_decay;
        op.pt_bias_correction = pt_bias_correction; op.step_dev = step; op.global_scale = global_scale;
        op.global_grad_norm = global_grad_norm; op.max_grad_norm = max_grad_norm; op.mode = mode; op.noop = noop;
        op.partials = update_sumsq ? ns.partials : nullptr;
        return launch_stream(tv, op, stream);
      });
    });
  });
  if (rc) return 1;
  if (!update_sumsq) return 0;
  // per-fragment sum of squares of the update (raw: the caller all-reduces over ranks before the sqrt); gated like the kernels
  return launch_norm_cleanup(tv, ns, 1, noop, 1, nullptr, update_sumsq, nullptr, nullptr, stream, 0, 2);
}

extern "C" int b200mt_dist_lamb_stage2(const b200mt_lists* L, int chunk_size, int32_t* noop,
                                       const float* pt_param_norm, const float* pt_update_norm,
                                       const int64_t* update_norm_offset, const float* lr, const float* pt_decay,
                                       const float* global_grad_norm, int use_nvlamb, void* stream_) {
  (void)global_grad_norm;   // passed through the reference API, unused by its kernel as well
  if (check_common(L, chunk_size)) return 1;
  B200MT_CHECK(noop && pt_param_norm && pt_update_norm && update_norm_offset && lr && pt_decay,
               "dist_lamb_stage2: a device pointer is NULL");
  B200MT_CHECK(L->depth == 3, "tensor_lists.size() != depth");
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  if (acquire_table(L, 3, 3, elementwise_tiling(), stream, &tv)) return 1;
  return dispatch_f32_f16(L->dtype[0], "lamb_stage_2", [&](auto tu) {
    return dispatch_f32_f16(L->dtype[1], "lamb_stage_2", [&](auto tp) {
      return dispatch_f32_f16_bf16_u8(L->dtype[2], "lamb_stage_2", [&](auto tc) {
        DistLambStage2Op<typename decltype(tu)::type, typename decltype(tp)::type, typename decltype(tc)::type> op;
        op.param_norm = pt_param_norm; op.update_norm = pt_update_norm;
        op.update_norm_offset = reinterpret_cast<const long long*>(update_norm_offset);
        op.lr = lr; op.pt_decay = pt_decay; op.use_nvlamb = use_nvlamb; op.noop = noop;
        return launch_stream(tv, op, stream);
      });
    });
  });
}
