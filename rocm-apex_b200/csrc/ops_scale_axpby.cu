// ops_scale_axpby.cu -- amp unscale with overflow flag, and axpby.
//   b200mt_scale  <- amp_C.multi_tensor_scale  (csrc/multi_tensor_scale_kernel.cu:29-136)
//   b200mt_axpby  <- amp_C.multi_tensor_axpby  (csrc/multi_tensor_axpby_kernel.cu:27-157)
#include "mt_dispatch.h"

// 8 elements per access (256-bit for fp32, 128-bit for 16-bit types) x 2 in flight: 10 % faster than
// 4 x 4 when a 16-bit list is involved, equal for fp32 (profiles/sweep_variants_r01.txt).
#ifndef B200MT_SCALE_VEC
#define B200MT_SCALE_VEC 8
#endif
#ifndef B200MT_SCALE_UNROLL
#define B200MT_SCALE_UNROLL 2
#endif
// fp32-only lists: vector width and vectors in flight per thread (swept with the occupancy, profiles/sweep_light_r02.txt)
// one 256-bit access per thread at 8 CTAs/SM: equal to 4 x 128-bit on large lists, 3 % faster on the ResNet-50 step
#ifndef B200MT_SCALE32_VEC
#define B200MT_SCALE32_VEC 8
#endif
#ifndef B200MT_SCALE32_UNROLL
#define B200MT_SCALE32_UNROLL 1
#endif
#ifndef B200MT_AXPBY_UNROLL
#define B200MT_AXPBY_UNROLL 2
#endif

// L2 eviction priority of [in, out], two bits per list (mt_device.cuh: l2_policy)
#ifndef B200MT_SCALE_L2_LOADS
#define B200MT_SCALE_L2_LOADS 0
#endif
#ifndef B200MT_SCALE_L2_STORES
#define B200MT_SCALE_L2_STORES 0
#endif

namespace b200mt {

__device__ __forceinline__ bool nonfinite(float v) {
  return (__float_as_uint(v) & 0x7f800000u) == 0x7f800000u;
}

// out = (out_t)((float)in * scale); flag raised iff any *input* is inf/NaN; noop never read
// (csrc/multi_tensor_scale_kernel.cu:66-70,108-109).
template <typename Tin, typename Tout>
struct ScaleOp : OpBase {
  using Types = TypeList<Tin, Tout>;
  static constexpr unsigned STORES = 0b10;
  static constexpr int MIN_CTAS = B200MT_SCALE_MIN_CTAS;
  // a 16-bit list wants 8 elements (one 128-bit access) per vector; fp32-only lists are as fast or
  // faster (small tensor lists) with 4 x 4
  static constexpr bool ANY16 = Elem<Tin>::SIZE == 2 || Elem<Tout>::SIZE == 2;
  static constexpr int UNROLL = ANY16 ? B200MT_SCALE_UNROLL : B200MT_SCALE32_UNROLL;
  static constexpr int VEC = ANY16 ? B200MT_SCALE_VEC : B200MT_SCALE32_VEC;
  static constexpr unsigned L2_LOADS = B200MT_SCALE_L2_LOADS, L2_STORES = B200MT_SCALE_L2_STORES;
  struct State { bool bad = false; };
  float scale;
  int* noop;
  __device__ __forceinline__ unsigned load_mask() const { return 0b01; }
  __device__ __forceinline__ void elem(State& s, float (&x)[2]) const {
    s.bad = s.bad || nonfinite(x[0]);
    x[1] = x[0] * scale;
  }
  __device__ __forceinline__ void end(State& s) const {
    if (s.bad) *noop = 1;  // benign race: every writer stores the same constant
  }
};

// out = a*x + b*y; which operand is checked follows arg_to_check
// (csrc/multi_tensor_axpby_kernel.cu:71-80,123-124).
template <typename Tx, typename Ty, typename Tout>
struct AxpbyOp : OpBase {
  using Types = TypeList<Tx, Ty, Tout>;
  static constexpr unsigned STORES = 0b100;
  static constexpr int MIN_CTAS = B200MT_SCALE_MIN_CTAS;
  static constexpr int UNROLL = B200MT_AXPBY_UNROLL;
  struct State { bool bad = false; };
  float a, b;
  int arg_to_check;
  int* noop;
  __device__ __forceinline__ unsigned load_mask() const { return 0b011; }
  __device__ __forceinline__ void elem(State& s, float (&x)[3]) const {
    if (arg_to_check == -1) s.bad = s.bad || nonfinite(x[0]) || nonfinite(x[1]);
    else if (arg_to_check == 0) s.bad = s.bad || nonfinite(x[0]);
    else if (arg_to_check == 1) s.bad = s.bad || nonfinite(x[1]);
    x[2] = a * x[0] + b * x[1];
  }
  __device__ __forceinline__ void end(State& s) const {
    if (s.bad) *noop = 1;
  }
};

}  // namespace b200mt

using namespace b200mt;

extern "C" int b200mt_scale(const b200mt_lists* L, int chunk_size, int32_t* noop, float scale,
                            void* stream_) {
  if (check_common(L, chunk_size)) return 1;
  B200MT_CHECK(noop != nullptr, "noop flag pointer is NULL");
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  if (acquire_table(L, 2, 2, elementwise_tiling(), stream, &tv)) return 1;
  return dispatch_f32_f16_bf16(L->dtype[0], "multi_tensor_scale_cuda", [&](auto tin) {
    return dispatch_f32_f16_bf16(L->dtype[1], "multi_tensor_scale_cuda", [&](auto tout) {
      ScaleOp<typename decltype(tin)::type, typename decltype(tout)::type> op;
      op.scale = scale; op.noop = noop;
      return launch_stream(tv, op, stream);
    });
  });
}

extern "C" int b200mt_axpby(const b200mt_lists* L, int chunk_size, int32_t* noop, float a, float b,
                            int arg_to_check, void* stream_) {
  if (check_common(L, chunk_size)) return 1;
  B200MT_CHECK(noop != nullptr, "noop flag pointer is NULL");
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  if (acquire_table(L, 3, 3, elementwise_tiling(), stream, &tv)) return 1;
  return dispatch_f32_f16_bf16(L->dtype[0], "multi_tensor_axpby_cuda", [&](auto tx) {
    return dispatch_f32_f16_bf16(L->dtype[1], "multi_tensor_axpby_cuda", [&](auto ty) {
      return dispatch_f32_f16_bf16(L->dtype[2], "multi_tensor_axpby_cuda", [&](auto to) {
        AxpbyOp<typename decltype(tx)::type, typename decltype(ty)::type, typename decltype(to)::type> op;
        op.a = a; op.b = b; op.arg_to_check = arg_to_check; op.noop = noop;
        return launch_stream(tv, op, stream);
      });
    });
  });
}
