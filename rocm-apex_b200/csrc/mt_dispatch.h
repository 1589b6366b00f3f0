// mt_dispatch.h -- runtime dtype code -> compile-time element type.
#pragma once
#include "mt_host.h"

namespace b200mt {

template <typename T> struct Tag { using type = T; };

// f(Tag<T>{}) is called with the element type for `dtype`; returns f's result or an error
// mirroring the reference's AT_ERROR("<name> not implemented for '<dtype>'") (csrc/type_shim.h:32-33).
template <class F>
int dispatch_f32_f16_bf16(int dtype, const char* name, F&& f) {
  switch (dtype) {
    case B200MT_F32: return f(Tag<float>{});
    case B200MT_F16: return f(Tag<__half>{});
    case B200MT_BF16: return f(Tag<__nv_bfloat16>{});
    default: return set_error("%s not implemented for '%s'", name, dtype_name(dtype));
  }
}

template <class F>
int dispatch_f64_f32_f16_bf16(int dtype, const char* name, F&& f) {
  if (dtype == B200MT_F64) return f(Tag<double>{});
  return dispatch_f32_f16_bf16(dtype, name, f);
}

// float / half / bfloat16 / byte, the byte being an e5m2 value (DISPATCH_FLOAT_HALF_AND_BYTE, csrc/type_shim.h:63)
template <class F>
int dispatch_f32_f16_bf16_u8(int dtype, const char* name, F&& f) {
  if (dtype == B200MT_U8) return f(Tag<e5m2>{});
  return dispatch_f32_f16_bf16(dtype, name, f);
}

template <class F>
int dispatch_f32_f16(int dtype, const char* name, F&& f) {
  switch (dtype) {
    case B200MT_F32: return f(Tag<float>{});
    case B200MT_F16: return f(Tag<__half>{});
    default: return set_error("%s not implemented for '%s'", name, dtype_name(dtype));
  }
}

}  // namespace b200mt
