// ops_misc.cu -- the small helpers DistributedFusedAdam uses around its step.
//   b200mt_maybe_cast            <- fused_adam_cuda.maybe_cast_mt
//                                   (apex/contrib/csrc/optimizers/fused_adam_cuda_kernel.cu:734-783,939-960)
//   b200mt_strided_check_finite  <- fused_adam_cuda.strided_check_finite (:465-523,785-812)
#include "mt_dispatch.h"

namespace b200mt {

// out = (TO_T)in unless *overflow_flag != 0 (flag may be NULL: always cast).
template <typename Tin, typename Tout>
struct MaybeCastOp : OpBase {
  using Types = TypeList<Tin, Tout>;
  static constexpr unsigned STORES = 0b10;
  static constexpr bool ANY16 = Elem<Tin>::SIZE == 2 || Elem<Tout>::SIZE == 2;   // as for ScaleOp
  static constexpr int UNROLL = ANY16 ? 2 : 4;
  static constexpr int VEC = ANY16 ? 8 : 4;
  const int* flag;
  __device__ __forceinline__ bool begin(State&) const {
    return !(flag && *reinterpret_cast<const volatile int*>(flag) != 0);
  }
  __device__ __forceinline__ unsigned load_mask() const { return 0b01; }
  __device__ __forceinline__ void elem(State&, float (&x)[2]) const { x[1] = x[0]; }
};

// Every stride-th element is tested; one CTA when the flag is cleared first so that the clear is
// ordered before every possible set (the reference makes the same choice, :797-798).
template <typename T>
__global__ void strided_check_finite_kernel(int* noop, const T* __restrict__ p, long long n, int stride,
                                            int clear_first) {
  if (clear_first) {
    if (blockIdx.x == 0 && threadIdx.x == 0) *noop = 0;
    __syncthreads();
  }
  const long long nthreads = (long long)gridDim.x * blockDim.x;
  bool bad = false;
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k * stride < n; k += nthreads) {
    const float v = Elem<T>::ld1(p + k * stride);
    bad = bad || ((__float_as_uint(v) & 0x7f800000u) == 0x7f800000u);
  }
  if (bad) *noop = 1;
}

}  // namespace b200mt

using namespace b200mt;

extern "C" int b200mt_maybe_cast(const b200mt_lists* L, int chunk_size, const int32_t* overflow_flag,
                                 void* stream_) {
  if (check_common(L, chunk_size)) return 1;
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  if (acquire_table(L, 2, 2, elementwise_tiling(), stream, &tv)) return 1;
  // the reference dispatches float / half / byte (e5m2) on both sides (DISPATCH_FLOAT_HALF_AND_BYTE); bf16 is extra
  return dispatch_f32_f16_bf16_u8(L->dtype[0], "maybe_cast_cuda_mt_kernel", [&](auto ti) {
    return dispatch_f32_f16_bf16_u8(L->dtype[1], "maybe_cast_cuda_mt_kernel", [&](auto to) {
      MaybeCastOp<typename decltype(ti)::type, typename decltype(to)::type> op;
      op.flag = overflow_flag;
      return launch_stream(tv, op, stream);
    });
  });
}

// Stand-alone collectives over NVSwitch multicast memory, built from the same streaming kernel:
//   reduce-scatter: lists = [g_mc (multicast address of this rank's shard), out (local)]
//                   out = sum over ranks of g, reduced inside the switch (multimem.ld_reduce)
//   all-gather:     lists = [in (local shard), out_mc (multicast address of the shard's slot)]
//                   one store, replicated into every rank's buffer by the switch (multimem.st)
// Fragments must be multiples of 8 elements and 16-byte aligned.  dtype: Half or BFloat16.
template <typename Tin, typename Tout>
struct McCopyOp : OpBase {
  using Types = TypeList<Tin, Tout>;
  static constexpr unsigned STORES = 0b10;
  static constexpr int UNROLL = 4;
  static constexpr int VEC = 8;
  __device__ __forceinline__ unsigned load_mask() const { return 0b01; }
  __device__ __forceinline__ void elem(State&, float (&x)[2]) const { x[1] = x[0]; }
};

static int nvls_copy(const b200mt_lists* L, int chunk_size, bool reduce, void* stream_) {
  if (check_common(L, chunk_size)) return 1;
  B200MT_CHECK(L->depth == 2, "expected two lists");
  B200MT_CHECK(L->dtype[0] == L->dtype[1] && (L->dtype[0] == B200MT_BF16 || L->dtype[0] == B200MT_F16),
               "nvls collectives: both lists must be Half or both BFloat16");
  for (int t = 0; t < L->ntensors; t++) {
    B200MT_CHECK(L->numel[t] % 8 == 0, "nvls collectives: fragment %d is not a multiple of 8 elements", t);
    for (int d = 0; d < 2; d++)
      B200MT_CHECK(((uintptr_t)L->ptrs[(size_t)d * L->ntensors + t]) % 16 == 0,
                   "nvls collectives: list %d fragment %d is not 16-byte aligned", d, t);
  }
  CallGuard guard(L->device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  TableView tv;
  if (acquire_table(L, 2, 2, elementwise_tiling(), stream, &tv)) return 1;
  const bool bf = L->dtype[0] == B200MT_BF16;
  if (reduce) {
    if (bf) return launch_stream(tv, McCopyOp<mc_bf16, __nv_bfloat16>(), stream);
    return launch_stream(tv, McCopyOp<mc_f16, __half>(), stream);
  }
  if (bf) return launch_stream(tv, McCopyOp<__nv_bfloat16, mc_bf16>(), stream);
  return launch_stream(tv, McCopyOp<__half, mc_f16>(), stream);
}

extern "C" int b200mt_nvls_reduce_scatter(const b200mt_lists* L, int chunk_size, void* stream) {
  return nvls_copy(L, chunk_size, true, stream);
}
extern "C" int b200mt_nvls_all_gather(const b200mt_lists* L, int chunk_size, void* stream) {
  return nvls_copy(L, chunk_size, false, stream);
}

extern "C" int b200mt_strided_check_finite(int32_t* noop, const void* p, int dtype, int64_t numel,
                                           int stride, int clear_overflow_first, int device,
                                           void* stream_) {
  B200MT_CHECK(noop != nullptr, "noop flag pointer is NULL");
  B200MT_CHECK(stride > 0, "stride must be > 0");
  B200MT_CHECK(numel >= 0 && (numel == 0 || p != nullptr), "bad tensor");
  CallGuard guard(device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  const long long niter = (numel + stride - 1) / stride;
  const int threads = 512;
  long long blocks = clear_overflow_first ? 1 : (niter + threads - 1) / threads;
  if (blocks < 1) blocks = 1;
  if (blocks > 148 * 8) blocks = 148 * 8;
  return dispatch_f32_f16_bf16_u8(dtype, "check_finite_cuda_kernel", [&](auto t) {
    using T = typename decltype(t)::type;
    strided_check_finite_kernel<T><<<(int)blocks, threads, 0, stream>>>(noop, (const T*)p, numel, stride,
                                                                     clear_overflow_first);
    count_launch();
    B200MT_CUDA(cudaGetLastError());
    return 0;
  });
}
