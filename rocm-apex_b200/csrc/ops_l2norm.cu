// ops_l2norm.cu -- global and per-tensor L2 norm (deterministic two-pass).
//   b200mt_l2norm        <- amp_C.multi_tensor_l2norm     (csrc/multi_tensor_l2norm_kernel.cu:28-110,301-364)
//                           amp_C.multi_tensor_l2norm_mp  (csrc/multi_tensor_l2norm_kernel_mp.cu:28-216), flag GATED
//   b200mt_l2norm_scale  <- amp_C.multi_tensor_l2norm_scale (csrc/multi_tensor_l2norm_scale_kernel.cu:28-324)
#include "mt_dispatch.h"
#include "mt_norm.cuh"

namespace b200mt {

// Pass 1: per-tile sum of squares in fp32 (csrc/multi_tensor_l2norm_kernel.cu:71-75).
#ifndef B200MT_NORM16_VEC
#define B200MT_NORM16_VEC 8   // 16-bit inputs: 8 elements = one 128-bit load per access
#endif
#ifndef B200MT_NORM16_UNROLL
#define B200MT_NORM16_UNROLL 1   // swept 1/2/4/8 on 4x256M fp16: 312 / 325 / 394 / 552 us (profiles/sweep_norm2_r01.txt)
#endif
#ifndef B200MT_NORM32_UNROLL
#define B200MT_NORM32_UNROLL 4
#endif
#ifndef B200MT_NORM32_VEC
#define B200MT_NORM32_VEC 4
#endif
template <typename Tx>
struct L2NormOp : OpBase {
  using Types = TypeList<Tx>;
  static constexpr unsigned STORES = 0;
  static constexpr int MIN_CTAS = (Elem<Tx>::SIZE == 2) ? B200MT_NORM16_MIN_CTAS : B200MT_NORM32_MIN_CTAS;
  static constexpr int VEC = (Elem<Tx>::SIZE == 2) ? B200MT_NORM16_VEC : B200MT_NORM32_VEC;
  static constexpr int UNROLL = (Elem<Tx>::SIZE == 2) ? B200MT_NORM16_UNROLL : B200MT_NORM32_UNROLL;
  static constexpr int THREADS = 256;
  struct State { float acc = 0.f; };
  NormFinish fin;      // partial slots + the folded second pass (mt_norm.cuh)
  const int* noop;
  int gated;
  __device__ __forceinline__ bool begin(State&) const {
    return !(gated && *reinterpret_cast<const volatile int*>(noop) != 0);
  }
  __device__ __forceinline__ unsigned load_mask() const { return 1u; }
  __device__ __forceinline__ void elem(State& s, float (&x)[1]) const { s.acc += x[0] * x[0]; }
  __device__ __forceinline__ void tile_end(State& s, int tile, const TileLoc& L, float* smem) const {
    const float part[1] = {cta_sum<THREADS>(s.acc, smem)};
    norm_tile_done<THREADS, 1>(fin, tile, L.tensor, part, smem);
    s.acc = 0.f;
  }
};

// Pass 1 of l2norm_scale: same partials over the *input*, plus out = in*scale in the same read
// (csrc/multi_tensor_l2norm_scale_kernel.cu:76-84; its per-element finite check is commented out
// in the reference, only the chunk-sum check remains).
template <typename Tin, typename Tout>
struct L2NormScaleOp : OpBase {
  using Types = TypeList<Tin, Tout>;
  static constexpr unsigned STORES = 0b10;
  static constexpr int MIN_CTAS = B200MT_NORM32_MIN_CTAS;
  static constexpr bool ANY16 = Elem<Tin>::SIZE == 2 || Elem<Tout>::SIZE == 2;   // as for ScaleOp
  static constexpr int UNROLL = ANY16 ? 2 : 4;
  static constexpr int VEC = ANY16 ? 8 : 4;
  static constexpr int THREADS = 256;
  struct State { float acc = 0.f; };
  NormFinish fin;
  float scale;
  __device__ __forceinline__ unsigned load_mask() const { return 1u; }
  __device__ __forceinline__ void elem(State& s, float (&x)[2]) const {
    s.acc += x[0] * x[0];
    x[1] = x[0] * scale;
  }
  __device__ __forceinline__ void tile_end(State& s, int tile, const TileLoc& L, float* smem) const {
    const float part[1] = {cta_sum<THREADS>(s.acc, smem)};
    norm_tile_done<THREADS, 1>(fin, tile, L.tensor, part, smem);
    s.acc = 0.f;
  }
};

}  // namespace b200mt

using namespace b200mt;

extern "C" int b200mt_l2norm(const b200mt_lists* L, int chunk_size, int32_t* noop, float* ret,
                             float* ret_per_tensor, int flags, void* stream_) {
  if (check_common(L, chunk_size)) return 1;
  B200MT_CHECK(noop != nullptr && ret != nullptr, "noop / ret pointer is NULL");
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  const int tile_scale = (L->dtype[0] == B200MT_F16 || L->dtype[0] == B200MT_BF16) ? 2 : 1;
  if (acquire_table(L, 1, 1, chunked_tiling(chunk_size, tile_scale), stream, &tv)) return 1;
  void* raw;
  if (acquire_scratch(stream, norm_scratch_bytes(1, tv.ntiles, tv.ntensors), &raw)) return 1;
  NormScratch ns = carve_norm_scratch(raw, 1, tv.ntiles, tv.ntensors);
  const int gated = (flags & B200MT_NORM_NOOP_GATED) ? 1 : 0;
  const char* name = gated ? "multi_tensor_l2norm_mp_cuda" : "multi_tensor_l2norm_cuda";
  NormFinish fin;
  const bool folded = make_norm_finish(tv, ns, 1, noop, gated, ret, ret_per_tensor, nullptr, nullptr, tv.nonempty, &fin);
  auto go = [&](auto tx) {
    L2NormOp<typename decltype(tx)::type> op;
    op.fin = fin; op.noop = noop; op.gated = gated;
    return launch_stream(tv, op, stream);
  };
  // the _mp variant of the reference dispatches fp32/fp16 only (csrc/multi_tensor_l2norm_kernel_mp.cu:186);
  // bf16 is accepted here as a superset.
  if (dispatch_f32_f16_bf16(L->dtype[0], name, go)) return 1;
  if (folded) return 0;      // the last CTA of the streaming kernel has already done the second pass
  return launch_norm_cleanup(tv, ns, 1, noop, gated, ret, ret_per_tensor, nullptr, nullptr, stream);
}

extern "C" int b200mt_l2norm_scale(const b200mt_lists* L, int chunk_size, int32_t* noop, float scale,
                                   float* ret, float* ret_per_tensor, void* stream_) {
  if (check_common(L, chunk_size)) return 1;
  B200MT_CHECK(noop != nullptr && ret != nullptr, "noop / ret pointer is NULL");
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  if (acquire_table(L, 2, 2, chunked_tiling(chunk_size), stream, &tv)) return 1;
  void* raw;
  if (acquire_scratch(stream, norm_scratch_bytes(1, tv.ntiles, tv.ntensors), &raw)) return 1;
  NormScratch ns = carve_norm_scratch(raw, 1, tv.ntiles, tv.ntensors);
  NormFinish fin;
  const bool folded = make_norm_finish(tv, ns, 1, noop, 0, ret, ret_per_tensor, nullptr, nullptr, tv.nonempty, &fin);
  int rc = dispatch_f32_f16_bf16(L->dtype[0], "multi_tensor_l2norm_scale_cuda", [&](auto ti) {
    return dispatch_f32_f16_bf16(L->dtype[1], "multi_tensor_l2norm_scale_cuda", [&](auto to) {
      L2NormScaleOp<typename decltype(ti)::type, typename decltype(to)::type> op;
      op.fin = fin; op.scale = scale;
      return launch_stream(tv, op, stream);
    });
  });
  if (rc) return 1;
  if (folded) return 0;
  return launch_norm_cleanup(tv, ns, 1, noop, 0, ret, ret_per_tensor, nullptr, nullptr, stream);
}
