// mt_norm.cuh -- second pass of the deterministic two-pass norm.
//
// Pass 1 (any streaming Op) leaves one fp32 partial sum of squares per tile in scratch.
// Pass 2 (this kernel): one CTA per tensor adds the tile partials of each chunk in order, checks
// every chunk sum for inf/NaN (the reference raises the noop flag per chunk,
// csrc/multi_tensor_l2norm_kernel.cu:99-108), adds the chunk sums in a fixed thread-strided order
// and a fixed shuffle tree, and writes sqrt(total) per tensor.  The CTA that finishes last sums the
// per-tensor totals in index order for the global norm.  No float atomics anywhere, so results are
// bitwise reproducible run to run.  Replaces cleanup / cleanup_v3 of the reference
// (csrc/multi_tensor_l2norm_kernel.cu:198-237) and its 320-slot accumulate-across-launches scheme.
#pragma once
#include "mt_host.h"

namespace b200mt {

// Scratch layout used by the norm ops (floats after the 256-byte counter zone):
//   counters[0] = ticket, counters[1] = bad-chunk flag   (both self-resetting)
//   partials[nsets][ntiles]      per-tile sums of squares
//   totals[nsets][ntensors]      per-tensor sums of squares
//   norms[nsets][ntensors]       per-tensor norms (used when the caller passes no output array)
struct NormScratch {
  unsigned* counters;
  float* partials;
  float* totals;
  float* norms;
};

inline size_t norm_scratch_bytes(int nsets, int ntiles, int ntensors) {
  return 256 + sizeof(float) * ((size_t)nsets * ntiles + 2 * (size_t)nsets * ntensors) + 64;
}

inline NormScratch carve_norm_scratch(void* base, int nsets, int ntiles, int ntensors) {
  NormScratch s;
  s.counters = reinterpret_cast<unsigned*>(base);
  s.partials = reinterpret_cast<float*>(reinterpret_cast<char*>(base) + 256);
  s.totals = s.partials + (size_t)nsets * ntiles;
  s.norms = s.totals + (size_t)nsets * ntensors;
  return s;
}

constexpr int kCleanupThreads = 256;

struct CleanupArgs {
  const int* tile_prefix;
  int ntensors, ntiles, ppc, nsets;
  const float* partials;
  float* totals;
  unsigned* counters;
  int* noop;
  int gated;                 // 1: skip all output when *noop != 0 on entry (the _mp variants)
  float* ret[2];             // per set: global norm out (float[1]) or NULL
  float* ret_per_tensor[2];  // per set: per-tensor norm out (float[ntensors]) or NULL
  // NovoGrad's norm_out (csrc/multi_tensor_l2norm_kernel.cu:274-297,371-456): partials are per-tile max|x|
  // (max_mode) or sums of squares, and the per-tensor result is blended into ret_per_tensor in place:
  //   L2:   out = sqrt(alpha*out^2 + beta*sum x^2)      L-inf: out = alpha*out + beta*max|x|
  int max_mode;
  int blend;
  float alpha, beta;
};

static __global__ void __launch_bounds__(kCleanupThreads) norm_cleanup_kernel(const CleanupArgs a) {
  __shared__ float red[kCleanupThreads / 32];
  __shared__ int s_last;
  const int tid = threadIdx.x;
  // launched with programmatic stream serialization: this grid may become resident while pass 1 drains its last wave;
  // everything it reads (partials, the flag) is produced by pass 1, so wait for that kernel to complete and flush first
  asm volatile("griddepcontrol.launch_dependents;");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const bool skip = a.gated && (*reinterpret_cast<volatile int*>(a.noop) != 0);
  bool bad = false;

  if (!skip) {
    for (int t = blockIdx.x; t < a.ntensors; t += gridDim.x) {
      const int first = a.tile_prefix[t];
      const int nt = a.tile_prefix[t + 1] - first;
      const int nchunks = (nt + a.ppc - 1) / a.ppc;
      for (int set = 0; set < a.nsets; set++) {
        const float* p = a.partials + (size_t)set * a.ntiles + first;
        float acc = 0.f;
        for (int c = tid; c < nchunks; c += kCleanupThreads) {
          float cs = 0.f;
          const int j0 = c * a.ppc;
          const int j1 = min(j0 + a.ppc, nt);
          if (a.max_mode) {
            for (int j = j0; j < j1; j++) cs = fmaxf(cs, p[j]);
            acc = fmaxf(acc, cs);
          } else {
            for (int j = j0; j < j1; j++) cs += p[j];
            acc += cs;
          }
          bad = bad || !isfinite(cs);
        }
        const float total = a.max_mode ? cta_max<kCleanupThreads>(acc, red) : cta_sum<kCleanupThreads>(acc, red);
        if (tid == 0) {
          a.totals[(size_t)set * a.ntensors + t] = total;
          if (float* out = a.ret_per_tensor[set]) {
            if (!a.blend) out[t] = sqrtf(total);
            else if (a.blend == 2) out[t] = total;   // raw sum of squares: the caller all-reduces before the sqrt
            else if (a.max_mode) out[t] = a.alpha * out[t] + a.beta * total;
            else out[t] = sqrtf(a.alpha * out[t] * out[t] + a.beta * total);
          }
        }
      }
    }
    if (bad) atomicOr(a.counters + 1, 1u);
  }

  // ---- last CTA done: global norm from the per-tensor totals, in index order
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = (atomicAdd(a.counters, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (!skip) {
    for (int set = 0; set < a.nsets; set++) {
      if (!a.ret[set]) continue;
      const volatile float* tot = a.totals + (size_t)set * a.ntensors;
      float acc = 0.f;
      for (int t = tid; t < a.ntensors; t += kCleanupThreads) acc += tot[t];
      const float total = cta_sum<kCleanupThreads>(acc, red);
      if (tid == 0) *a.ret[set] = sqrtf(total);
    }
  }
  if (tid == 0) {
    if (*reinterpret_cast<volatile unsigned*>(a.counters + 1)) *a.noop = 1;
    a.counters[0] = 0;  // self-reset for the next call on this stream
    a.counters[1] = 0;
  }
}

inline int launch_norm_cleanup(const TableView& tv, const NormScratch& s, int nsets, int* noop,
                               int gated, float* ret0, float* per0, float* ret1, float* per1,
                               cudaStream_t stream, int max_mode = 0, int blend = 0, float alpha = 0.f,
                               float beta = 0.f) {
  CleanupArgs a;
  a.max_mode = max_mode; a.blend = blend; a.alpha = alpha; a.beta = beta;
  a.tile_prefix = tv.tile_prefix;
  a.ntensors = tv.ntensors; a.ntiles = tv.ntiles; a.ppc = tv.ppc; a.nsets = nsets;
  a.partials = s.partials; a.totals = s.totals; a.counters = s.counters;
  a.noop = noop; a.gated = gated;
  a.ret[0] = ret0; a.ret[1] = ret1;
  a.ret_per_tensor[0] = per0; a.ret_per_tensor[1] = per1;
  int grid = tv.ntensors < 2048 ? tv.ntensors : 2048;
  const bool tracing = trace_on();
  if (tracing) trace_before(stream, "norm_cleanup_kernel");
  if (pdl_enabled() && !tracing) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kCleanupThreads);
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    B200MT_CUDA(cudaLaunchKernelEx(&cfg, norm_cleanup_kernel, a));
  } else {
    norm_cleanup_kernel<<<grid, kCleanupThreads, 0, stream>>>(a);
  }
  if (tracing) trace_after(stream);
  count_launch();
  B200MT_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace b200mt
