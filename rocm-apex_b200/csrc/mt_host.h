// mt_host.h -- host side shared by every op of libb200mt.so: error reporting, the per-device
// cache of descriptor tables, per-stream scratch, and the launch helper.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdarg.h>
#include <stdio.h>
#include <string>
#include <typeinfo>

#include "../../include/b200mt.h"
#include "mt_device.cuh"

namespace b200mt {

// ---- errors --------------------------------------------------------------------------------
int  set_error(const char* fmt, ...);          // stores message (thread-local), returns 1
void count_launch(int n = 1);

#define B200MT_CUDA(expr)                                                                   \
  do {                                                                                      \
    cudaError_t e__ = (expr);                                                               \
    if (e__ != cudaSuccess)                                                                 \
      return ::b200mt::set_error("CUDA error: %s (%s) at %s:%d", cudaGetErrorString(e__),   \
                                 #expr, __FILE__, __LINE__);                                \
  } while (0)

#define B200MT_CHECK(cond, ...)                                                             \
  do {                                                                                      \
    if (!(cond)) return ::b200mt::set_error(__VA_ARGS__);                                   \
  } while (0)

const char* dtype_name(int dtype);

// ---- per-call guard: takes the library lock, switches to lists->device, restores on exit ----
struct CallGuard {
  explicit CallGuard(int device);
  ~CallGuard();
  int status;  // 0 ok
  int prev_device;
  int device;
};

// ---- tiling policy -------------------------------------------------------------------------
struct Tiling {
  int       tile_elems;
  int       ppc;
  long long chunk_elems;
};
// Element-wise ops: chunk_size has no observable effect, tiles subdivide the tensor directly.
Tiling elementwise_tiling();
// Norm ops: a chunk of `chunk_size` elements is split into ppc tiles so that the per-chunk
// overflow check of the reference (csrc/multi_tensor_l2norm_kernel.cu:103) is preserved.
// tile_scale multiplies the tile length (2 for lists of 16-bit elements only: same bytes per tile);
// fixed_tile_elems > 0 replaces the tunable default (callers whose scratch is sized for one tile length).
Tiling chunked_tiling(int chunk_size, int tile_scale = 1, int fixed_tile_elems = 0);

// Validates the list set (depth, counts, null pointers) and returns a device-resident table for
// it on `stream`, uploading only when no cached copy matches.
int acquire_table(const b200mt_lists* L, int expect_depth_min, int expect_depth_max,
                  const Tiling& tiling, cudaStream_t stream, TableView* out);

// Per-(device, stream) scratch of at least `bytes` bytes, 256-byte aligned, contents undefined
// except for the first kScratchZeroBytes bytes which are zero at creation and which every user must leave
// zero (they hold self-resetting completion counters: 256 bytes of list-level counters, then one tile ticket per
// tensor for the single-launch norms).
constexpr int kNormTicketCap = 65536;
constexpr size_t kScratchZeroBytes = 256 + sizeof(unsigned) * (size_t)kNormTicketCap;
int acquire_scratch(cudaStream_t stream, size_t bytes, void** out);

// ---- launch trace (b200mt_trace_begin / _end): when on, every kernel launch of the library is bracketed by a pair
// of CUDA events on its stream, so a caller can read per-kernel durations of a multi-kernel entry point live
// (bench.py uses it for the roofline of the dominant kernel).  Off by default: one predictable branch per launch.
bool trace_on();
void trace_before(cudaStream_t stream, const char* mangled_name);
void trace_after(cudaStream_t stream);

int grid_for(int ntiles);
bool pdl_enabled();   // env B200MT_PDL (default on): launch streaming kernels with programmatic stream serialization

template <class Op>
int launch_stream(const TableView& tv, const Op& op, cudaStream_t stream, bool allow_pdl = true) {
  if (tv.ntiles <= 0) return 0;
  const bool tracing = trace_on();
  if (tracing) trace_before(stream, typeid(Op).name());
  if (allow_pdl && pdl_enabled() && !tracing) {   // traced launches run serialised: the events must see the whole kernel
    // Programmatic dependent launch: this grid may become resident while the previous kernel of the
    // stream drains its last wave; it touches nothing that kernel produced before griddepcontrol.wait.
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid_for(tv.ntiles));
    cfg.blockDim = dim3(Op::THREADS);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    B200MT_CUDA(cudaLaunchKernelEx(&cfg, mt_stream_kernel<Op>, tv, op));
  } else {
    mt_stream_kernel<Op><<<grid_for(tv.ntiles), Op::THREADS, 0, stream>>>(tv, op);
  }
  if (tracing) trace_after(stream);
  count_launch();
  B200MT_CUDA(cudaGetLastError());
  return 0;
}

// common argument checks shared by all entry points
int check_common(const b200mt_lists* L, int chunk_size);

}  // namespace b200mt
