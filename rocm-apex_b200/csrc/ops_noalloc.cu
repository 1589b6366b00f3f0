// ops_noalloc.cu -- b200mt_l2norm_flat_noalloc: the L2 norm of one flat buffer, for hosts that
// cannot let this library allocate.
//
// Who needs it.  The regular entry points create their working memory lazily with cudaMalloc /
// cudaHostAlloc (descriptor-table cache, per-stream partial sums).  That is a problem for a host
// process in which another runtime has already taken essentially all of the GPU's memory for its
// own pool (a late cudaMalloc from a plug-in then simply fails with out-of-memory), and for a step
// loop that must never hit cudaMalloc's implicit device-wide synchronisation.  Such a host still
// wants the gradient norm of its flat gradient shard every step -- the quantity the sharded
// optimizer clips with.  This entry point gives it that with no allocation of any kind: every byte
// of working memory lives in one fixed array that is part of the library image.
//
// The price, which the caller must know about:
//  * the array is B200MT_NOALLOC_SCRATCH_BYTES (about 257 KiB) of device memory on EVERY GPU that
//    loads this library, from load until the process exits, whether or not this function is called;
//  * inputs needing more than B200MT_NOALLOC_MAX_TILES (65536) scheduling tiles are refused with an
//    error.  The limit is in TILES, not elements: a chunk of chunk_size elements is cut into
//    ppc = ceil(chunk_size / 16384) equal tiles (16384 whatever B200MT_NORM_TILE_ELEMS says), so an
//    input of numel elements needs floor(numel / chunk_size) * ppc tiles plus those of the last,
//    partial chunk.  That is 65536 * 16384 = 1,073,741,824 elements only when chunk_size is a
//    multiple of 16384 (65536, the default, is); chunk_size 16400 gives 2 tiles of 8208 per chunk
//    and a limit of 32768 chunks = 537,395,200 elements, chunk_size 20000 about 655 M, and any
//    chunk_size below 16384 gives one tile per chunk, i.e. 65536 * chunk_size elements.  Size your
//    buffer from that formula (or from a trial call: the refusal is decided on the host, before
//    anything is launched);
//  * all calls on a device share the array and it keeps whatever the previous call left in it.  The
//    function re-initialises the few words that must start at zero itself, at the head of its own
//    work on the stream, so back-to-back calls on one stream are fine; two calls whose GPU work can
//    overlap (different streams with no ordering between them, or two host threads on different
//    streams) race on the array and give garbage.  Order them yourself;
//  * the first call on each device looks up the array's device address (a host-side query, no
//    allocation) and keeps it for the life of the process.  Make that first call before you begin
//    any stream capture on that device.  The kept address does not survive cudaDeviceReset() or the
//    destruction of the device's primary context: a host that tears contexts down must not call
//    this entry point afterwards.
#include "mt_dispatch.h"
#include "mt_norm.cuh"

namespace b200mt {

// Layout of the fixed array:
//   [0, 256)        counters of the second pass (ticket, bad-chunk flag)
//   [256, 512)      one-tensor descriptor table: base pointer, numel, tile prefix {0, ntiles}
//   [512, ...)      per-tile partial sums (MAX_TILES floats), then one total and one norm
constexpr int kNoallocMaxTiles = B200MT_NOALLOC_MAX_TILES;
constexpr size_t kNoallocBytes = B200MT_NOALLOC_SCRATCH_BYTES;
static_assert(512 + sizeof(float) * (size_t)(kNoallocMaxTiles + 2) <= kNoallocBytes, "scratch array too small");

// In host code this name is only a handle for the runtime, NOT a device address: the address a
// kernel may use is obtained per device with cudaGetSymbolAddress (noalloc_scratch_for_device).
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoallocBytes];

static void* g_noalloc_addr[64] = {nullptr};

// Once per device, with that device current.
static int noalloc_scratch_for_device(int device, unsigned char** out) {
  if (!g_noalloc_addr[device]) {
    void* p = nullptr;
    B200MT_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    B200MT_CHECK(p != nullptr, "cudaGetSymbolAddress returned a null address for the no-alloc scratch");
    g_noalloc_addr[device] = p;
  }
  *out = static_cast<unsigned char*>(g_noalloc_addr[device]);
  return 0;
}

// Head of every call: writes the one-tensor table and clears the words that must start at zero
// (the array still holds the previous call's state, possibly of a call that was cut short).
static __global__ void noalloc_prepare_kernel(unsigned char* scratch, unsigned long long base, long long numel,
                                              int ntiles) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    unsigned* counters = reinterpret_cast<unsigned*>(scratch);
    counters[0] = 0u;
    counters[1] = 0u;
    *reinterpret_cast<unsigned long long*>(scratch + 256) = base;
    *reinterpret_cast<long long*>(scratch + 264) = numel;
    int* prefix = reinterpret_cast<int*>(scratch + 272);
    prefix[0] = 0;
    prefix[1] = ntiles;
  }
}

// Pass 1, same arithmetic as the regular l2norm (per-tile fp32 sum of squares).
template <typename Tx>
struct FlatL2NormOp : OpBase {
  using Types = TypeList<Tx>;
  static constexpr unsigned STORES = 0;
  static constexpr int UNROLL = 4;
  static constexpr int THREADS = 256;
  struct State { float acc = 0.f; };
  float* partials;
  __device__ __forceinline__ unsigned load_mask() const { return 1u; }
  __device__ __forceinline__ void elem(State& s, float (&x)[1]) const { s.acc += x[0] * x[0]; }
  __device__ __forceinline__ void tile_end(State& s, int tile, const TileLoc&, float* smem) const {
    const float t = cta_sum<THREADS>(s.acc, smem);
    if (threadIdx.x == 0) partials[tile] = t;
    s.acc = 0.f;
  }
};

}  // namespace b200mt

using namespace b200mt;

extern "C" int b200mt_l2norm_flat_noalloc(const void* x, int dtype, int64_t numel, int chunk_size,
                                          int32_t* noop, float* ret, int device, void* stream_) {
  B200MT_CHECK(chunk_size > 0, "chunk_size must be > 0 (got %d)", chunk_size);
  B200MT_CHECK(noop != nullptr && ret != nullptr, "noop / ret pointer is NULL");
  B200MT_CHECK(numel >= 0 && (numel == 0 || x != nullptr), "bad input buffer");
  const Tiling tiling = chunked_tiling(chunk_size, 1, 16384);   // fixed: the array is sized in tiles of this length
  const long long full = numel / tiling.chunk_elems, rem = numel % tiling.chunk_elems;
  const long long ntiles = full * tiling.ppc + (rem + tiling.tile_elems - 1) / tiling.tile_elems;
  B200MT_CHECK(ntiles <= kNoallocMaxTiles,
               "b200mt_l2norm_flat_noalloc: input needs %lld tiles, the fixed scratch holds %d "
               "(numel %lld, chunk_size %d, tile %d elements); use b200mt_l2norm or a larger chunk_size",
               ntiles, kNoallocMaxTiles, (long long)numel, chunk_size, tiling.tile_elems);
  CallGuard guard(device);
  if (guard.status) return 1;
  cudaStream_t stream = (cudaStream_t)stream_;
  unsigned char* scratch = nullptr;
  if (noalloc_scratch_for_device(device, &scratch)) return 1;

  noalloc_prepare_kernel<<<1, 32, 0, stream>>>(scratch, (unsigned long long)x, (long long)numel, (int)ntiles);
  count_launch();
  B200MT_CUDA(cudaGetLastError());

  TableView tv;
  tv.ptrs = reinterpret_cast<const unsigned long long*>(scratch + 256);
  tv.numel = reinterpret_cast<const long long*>(scratch + 264);
  tv.tile_prefix = reinterpret_cast<const int*>(scratch + 272);
  tv.ntensors = 1;
  tv.ntiles = (int)ntiles;
  tv.nonempty = numel > 0 ? 1 : 0;
  tv.multichunk = numel > tiling.chunk_elems ? 1 : 0;
  tv.tile_elems = tiling.tile_elems;
  tv.ppc = tiling.ppc;
  tv.chunk_elems = tiling.chunk_elems;

  NormScratch ns;
  ns.counters = reinterpret_cast<unsigned*>(scratch);
  ns.tickets = nullptr;   // two-kernel form only
  ns.partials = reinterpret_cast<float*>(scratch + 512);
  ns.totals = ns.partials + kNoallocMaxTiles;
  ns.norms = ns.totals + 1;

  int rc = dispatch_f32_f16_bf16(dtype, "b200mt_l2norm_flat_noalloc", [&](auto t) {
    FlatL2NormOp<typename decltype(t)::type> op;
    op.partials = ns.partials;
    // no programmatic early start: the descriptor table itself was written by the kernel just before
    return launch_stream(tv, op, stream, false);
  });
  if (rc) return 1;
  return launch_norm_cleanup(tv, ns, 1, noop, 0, ret, nullptr, nullptr, nullptr, stream);
}
