// mt_norm.cuh -- second pass of the deterministic two-pass norm.
//
// Pass 1 (any streaming Op) leaves one fp32 partial sum of squares per tile in scratch.
// Pass 2 (norm_cleanup_kernel): per tensor, the tile partials of each chunk are added in order, every
// chunk sum is checked for inf/NaN (the reference raises the noop flag per chunk,
// csrc/multi_tensor_l2norm_kernel.cu:99-108), the chunk sums are added in a fixed thread-strided order
// and a fixed shuffle tree, and sqrt(total) is written per tensor -- by one thread, one warp or one CTA
// depending on the tensor's size, with identical arithmetic.  The CTA that finishes last (or the only
// one) sums the per-tensor totals in index order for the global norm.  No float atomics anywhere, so results are
// bitwise reproducible run to run.  Replaces cleanup / cleanup_v3 of the reference
// (csrc/multi_tensor_l2norm_kernel.cu:198-237) and its 320-slot accumulate-across-launches scheme.
#pragma once
#include <cstdlib>

#include "mt_host.h"

namespace b200mt {

// Scratch layout used by the norm ops (floats after the zero zone):
//   counters[0] = ticket, counters[1] = bad-chunk flag   (256 bytes, both self-resetting)
//   tickets[kNormTicketCap]      per-tensor tile counters of the single-launch path (self-resetting; the whole zone
//                                [0, kScratchZeroBytes) is zeroed when a scratch buffer is created and every user leaves it zero)
//   partials[nsets][ntiles]      per-tile sums of squares
//   totals[nsets][ntensors]      per-tensor sums of squares
//   norms[nsets][ntensors]       per-tensor norms (used when the caller passes no output array)
struct NormScratch {
  unsigned* counters;
  unsigned* tickets;
  float* partials;
  float* totals;
  float* norms;
};

inline size_t norm_scratch_bytes(int nsets, int ntiles, int ntensors) {
  return kScratchZeroBytes + sizeof(float) * ((size_t)nsets * ntiles + 2 * (size_t)nsets * ntensors) + 64;
}

inline NormScratch carve_norm_scratch(void* base, int nsets, int ntiles, int ntensors) {
  NormScratch s;
  s.counters = reinterpret_cast<unsigned*>(base);
  s.tickets = reinterpret_cast<unsigned*>(reinterpret_cast<char*>(base) + 256);
  s.partials = reinterpret_cast<float*>(reinterpret_cast<char*>(base) + kScratchZeroBytes);
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
  int nbatch;    // norm_cleanup_kernel: CTAs [0, nbatch) handle the one-chunk tensors, 1024 each
  int tickets;   // ... and this many CTAs take part in the last-one-done count: nbatch + the multi-chunk tensors
};

// One tensor's result from its total (sum of squares, or max): per-tensor output rules shared by every path.
__device__ __forceinline__ void norm_store_tensor(const CleanupArgs& a, int set, int t, float total) {
  a.totals[(size_t)set * a.ntensors + t] = total;
  if (float* out = a.ret_per_tensor[set]) {
    if (!a.blend) out[t] = sqrtf(total);
    else if (a.blend == 2) out[t] = total;   // raw sum of squares: the caller all-reduces before the sqrt
    else if (a.max_mode) out[t] = a.alpha * out[t] + a.beta * total;
    else out[t] = sqrtf(a.alpha * out[t] * out[t] + a.beta * total);
  }
}

// Second pass.  Tensors are treated by size, always with the SAME arithmetic -- chunk sums in tile order, chunk c
// accumulated by thread c mod 256, the xor-shuffle tree of cta_sum -- so that the value does not depend on the path
// taken (nor on the folded form of norm_tile_done):
//   one chunk or none  (tensor <= chunk_size)   one THREAD per tensor, 1024 tensors per "batch" CTA (CTAs [0, nbatch))
//   2..32 chunks                                  a CTA per tensor (CTAs [nbatch, nbatch + T)), warp 0 works: lane c = chunk c
//   more                                          a CTA per tensor, thread-strided, cta_sum
// A per-tensor CTA whose tensor has at most one chunk leaves at once (the batch CTAs own it) and takes no ticket; when the
// list has no multi-chunk tensor those CTAs are not launched at all.  Round 1 gave every tensor a CTA: on a list of 1000
// one-chunk tensors that was 1000 CTAs in which one thread worked and 255 went through two barriers, a fence and a ticket --
// 13 us for 32 KB of partials (ncu, profiles/light_r02_summary.md).  Such a list is now ONE CTA: no ticket, no fence.
constexpr int kBatchTensors = 1024;

static __global__ void __launch_bounds__(kCleanupThreads) norm_cleanup_kernel(const CleanupArgs a) {
  __shared__ float red[kCleanupThreads / 32];
  __shared__ int s_last;
  const int tid = threadIdx.x, lane = tid & 31;
  // launched with programmatic stream serialization: this grid may become resident while pass 1 drains its last wave;
  // everything it reads (partials, the flag) is produced by pass 1, so wait for that kernel to complete and flush first
  asm volatile("griddepcontrol.launch_dependents;");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const bool skip = a.gated && (*reinterpret_cast<volatile int*>(a.noop) != 0);
  bool bad = false;

  if ((int)blockIdx.x < a.nbatch) {
    // ---- batch CTA: every tensor of its range that has one chunk or none, one thread each
    const int t1 = min(a.ntensors, ((int)blockIdx.x + 1) * kBatchTensors);
    for (int t = blockIdx.x * kBatchTensors + tid; !skip && t < t1; t += kCleanupThreads) {
      const int first = a.tile_prefix[t];
      const int nt = a.tile_prefix[t + 1] - first;
      if (nt > a.ppc) continue;                    // more than one chunk: its own CTA
      for (int set = 0; set < a.nsets; set++) {
        const float* p = a.partials + (size_t)set * a.ntiles + first;
        float cs = 0.f;
        if (a.max_mode) { for (int j = 0; j < nt; j++) cs = fmaxf(cs, p[j]); }
        else { for (int j = 0; j < nt; j++) cs += p[j]; }
        bad = bad || !isfinite(cs);
        norm_store_tensor(a, set, t, a.max_mode ? fmaxf(0.f, cs) : 0.f + cs);
      }
    }
  } else {
    const int t = blockIdx.x - a.nbatch;
    const int first = a.tile_prefix[t];
    const int nt = a.tile_prefix[t + 1] - first;
    if (nt <= a.ppc) return;                       // CTA-uniform; not counted among the tickets
    const int nchunks = (nt + a.ppc - 1) / a.ppc;
    if (!skip && nchunks <= 32) {
      // ---- warp 0: lane c holds chunk c
      if (tid < 32) {
        for (int set = 0; set < a.nsets; set++) {
          const float* p = a.partials + (size_t)set * a.ntiles + first;
          float acc = 0.f;
          if (lane < nchunks) {
            float cs = 0.f;
            const int j0 = lane * a.ppc, j1 = min(j0 + a.ppc, nt);
            if (a.max_mode) { for (int j = j0; j < j1; j++) cs = fmaxf(cs, p[j]); acc = fmaxf(acc, cs); }
            else { for (int j = j0; j < j1; j++) cs += p[j]; acc += cs; }
            bad = bad || !isfinite(cs);
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1)
            acc = a.max_mode ? fmaxf(acc, __shfl_xor_sync(0xffffffffu, acc, o)) : acc + __shfl_xor_sync(0xffffffffu, acc, o);
          if (lane == 0) norm_store_tensor(a, set, t, acc);
        }
      }
    } else if (!skip) {
      // ---- the whole CTA
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
        if (tid == 0) norm_store_tensor(a, set, t, total);
      }
    }
  }
  if (bad) atomicOr(a.counters + 1, 1u);

  // ---- global norm from the per-tensor totals, in index order: by this CTA if it is the only one with work, else by the
  // last one to finish (a.tickets CTAs take a ticket: the batch CTAs and the CTAs of the multi-chunk tensors)
  if (a.tickets > 1) {
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = (atomicAdd(a.counters, 1u) == (unsigned)(a.tickets - 1));
    __syncthreads();
    if (!s_last) return;
    __threadfence();
  } else {
    __syncthreads();       // one CTA: its own global writes (totals, the bad flag) are visible to all its threads
  }
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

// ------------------------------------------------------------------------------------------------------------------
// Single-launch form (opt-in, see norm_fold_enabled): the second pass folded into the streaming kernel.
//
// The two-kernel form pays a second launch (and its drain) per norm: on a 40 us pass over 1000 x 64K elements that is
// 10-15 % of the op.  Here every CTA, after leaving its tile's partial in the fixed slot, takes a ticket of its tensor;
// the CTA that takes the tensor's LAST ticket re-reads all of that tensor's partials and reduces them -- in exactly the
// order and with exactly the arithmetic of norm_cleanup_kernel, so the result does not depend on which CTA happens to
// finish last -- and then takes a ticket of the list; the CTA that completes the last tensor forms the global norm from
// the per-tensor totals in index order, raises the flag and resets the counters.  Still no float atomics, still
// bitwise reproducible.  Tickets live in a zone of the scratch that is zero between calls (each finisher resets its own).
// ------------------------------------------------------------------------------------------------------------------
struct NormFinish {
  CleanupArgs a;
  unsigned* tickets;   // [ntensors]
  int nonempty;        // tensors that own at least one tile
  int enabled;         // 0: the op only stores partials (two-kernel form)
};

// Called by all threads of the CTA at the end of a tile; part[set] (valid in thread 0) is the tile's partial of set.
template <int THREADS, int NSETS>
__device__ __forceinline__ void norm_tile_done(const NormFinish& nf, int tile, int tensor, const float (&part)[NSETS], float* red) {
  static_assert(THREADS == kCleanupThreads, "the folded second pass keeps the summation order of norm_cleanup_kernel");
  __shared__ int s_last;
  const CleanupArgs& a = nf.a;
  const int tid = threadIdx.x;
  float* partials = const_cast<float*>(a.partials);
  if (tid == 0) {
#pragma unroll
    for (int set = 0; set < NSETS; set++) partials[(size_t)set * a.ntiles + tile] = part[set];
    if (nf.enabled) {
      __threadfence();
      const int nt = a.tile_prefix[tensor + 1] - a.tile_prefix[tensor];
      s_last = (atomicAdd(nf.tickets + tensor, 1u) == (unsigned)(nt - 1));
    }
  }
  if (!nf.enabled) return;
  __syncthreads();
  if (!s_last) return;            // CTA-uniform
  __threadfence();

  // ---- this CTA completes `tensor`: chunk sums in order, per-chunk finiteness, per-tensor total
  const int first = a.tile_prefix[tensor];
  const int nt = a.tile_prefix[tensor + 1] - first;
  const int nchunks = (nt + a.ppc - 1) / a.ppc;
  bool bad = false;
#pragma unroll
  for (int set = 0; set < NSETS; set++) {
    const float* p = partials + (size_t)set * a.ntiles + first;
    float acc = 0.f;
    for (int c = tid; c < nchunks; c += THREADS) {
      float cs = 0.f;
      const int j0 = c * a.ppc;
      const int j1 = min(j0 + a.ppc, nt);
      if (a.max_mode) {
        for (int j = j0; j < j1; j++) cs = fmaxf(cs, __ldcg(p + j));
        acc = fmaxf(acc, cs);
      } else {
        for (int j = j0; j < j1; j++) cs += __ldcg(p + j);
        acc += cs;
      }
      bad = bad || !isfinite(cs);
    }
    const float total = a.max_mode ? cta_max<THREADS>(acc, red) : cta_sum<THREADS>(acc, red);
    if (tid == 0) {
      a.totals[(size_t)set * a.ntensors + tensor] = total;
      if (float* out = a.ret_per_tensor[set]) {
        if (!a.blend) out[tensor] = sqrtf(total);
        else if (a.blend == 2) out[tensor] = total;
        else if (a.max_mode) out[tensor] = a.alpha * out[tensor] + a.beta * total;
        else out[tensor] = sqrtf(a.alpha * out[tensor] * out[tensor] + a.beta * total);
      }
    }
  }
  if (bad) atomicOr(a.counters + 1, 1u);
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    nf.tickets[tensor] = 0u;      // ready for the next call
    s_last = (atomicAdd(a.counters, 1u) == (unsigned)(nf.nonempty - 1));
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();

  // ---- this CTA completes the list: empty tensors, global norm in index order, flag, counter reset
#pragma unroll
  for (int set = 0; set < NSETS; set++) {
    const volatile float* tot = a.totals + (size_t)set * a.ntensors;
    float acc = 0.f;
    for (int t = tid; t < a.ntensors; t += THREADS) {
      const bool empty = a.tile_prefix[t + 1] == a.tile_prefix[t];
      if (empty) {
        if (float* out = a.ret_per_tensor[set]) {         // a tensor without elements: total 0 through the same rules
          if (!a.blend || a.blend == 2) out[t] = 0.f;
          else if (a.max_mode) out[t] = a.alpha * out[t];
          else out[t] = sqrtf(a.alpha * out[t] * out[t]);
        }
      } else {
        acc += tot[t];
      }
    }
    if (a.ret[set]) {
      const float total = cta_sum<THREADS>(acc, red);
      if (tid == 0) *a.ret[set] = sqrtf(total);
    }
  }
  if (tid == 0) {
    if (*reinterpret_cast<volatile unsigned*>(a.counters + 1)) *a.noop = 1;
    a.counters[0] = 0;
    a.counters[1] = 0;
  }
}

// Measured on B200 and NOT the default (profiles/norm_fold_r02.md): the ticket is an atomic whose return value the CTA
// must wait for (~1 us through L2) at the end of every tile, on the critical path of a CTA that lives ~4 us -- 1000 x 64K
// fp32 l2norm 60.5 -> 65.6 us, BERT-large fp16 gradient norm 124 -> 165 us.  The second launch of the two-kernel form,
// started with programmatic dependent launch, costs less than that.  B200MT_NORM_FOLD=1 switches the folded form on
// (it wins only where the launch itself dominates: lists of thousands of tiny tensors).
inline bool norm_fold_enabled() {
  static const bool on = [] { const char* e = getenv("B200MT_NORM_FOLD"); return e && atoi(e) != 0; }();
  return on;
}

// Host side: fills the finish block of an op.  Returns true when the single-launch form applies (then the caller must
// NOT launch norm_cleanup_kernel); false: two-kernel form (lists with more tensors than ticket slots, lists without any
// element -- no streaming kernel is launched for those, the second pass alone writes the zeros --, B200MT_NORM_FOLD=0).
inline bool make_norm_finish(const TableView& tv, const NormScratch& s, int nsets, int* noop, int gated, float* ret0,
                             float* per0, float* ret1, float* per1, int nonempty, NormFinish* nf, int max_mode = 0,
                             int blend = 0, float alpha = 0.f, float beta = 0.f) {
  CleanupArgs& a = nf->a;
  a.max_mode = max_mode; a.blend = blend; a.alpha = alpha; a.beta = beta;
  a.tile_prefix = tv.tile_prefix;
  a.ntensors = tv.ntensors; a.ntiles = tv.ntiles; a.ppc = tv.ppc; a.nsets = nsets;
  a.partials = s.partials; a.totals = s.totals; a.counters = s.counters;
  a.noop = noop; a.gated = gated;
  a.ret[0] = ret0; a.ret[1] = ret1;
  a.ret_per_tensor[0] = per0; a.ret_per_tensor[1] = per1;
  a.nbatch = 0; a.tickets = 0;     // fields of the two-kernel second pass only
  nf->tickets = s.tickets;
  nf->nonempty = nonempty;
  nf->enabled = (norm_fold_enabled() && tv.ntiles > 0 && nonempty > 0 && tv.ntensors <= kNormTicketCap) ? 1 : 0;
  return nf->enabled != 0;
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
  a.nbatch = (tv.ntensors + kBatchTensors - 1) / kBatchTensors;
  a.tickets = a.nbatch + tv.multichunk;
  const int grid = a.nbatch + (tv.multichunk > 0 ? tv.ntensors : 0);
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
