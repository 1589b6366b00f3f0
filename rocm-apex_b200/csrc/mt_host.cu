// mt_host.cu -- library state of libb200mt.so: errors, descriptor-table cache, scratch.
#include "mt_host.h"

#include <atomic>
#include <climits>
#include <cstdlib>
#include <cstring>
#include <cxxabi.h>
#include <map>
#include <mutex>
#include <vector>

namespace b200mt {

// ---------------------------------------------------------------------------------------------
// errors, counters, tuning
// ---------------------------------------------------------------------------------------------
static thread_local std::string g_err;
static std::atomic<int64_t> g_launches{0}, g_uploads{0}, g_hits{0};
static std::recursive_mutex g_mu;           // library-level state: tuning, trace control, shutdown
// One lock PER DEVICE around a call (table cache, scratch and handles are per device): host threads that drive different
// GPUs of one process do not serialise on each other.
static std::recursive_mutex g_dev_mu[64];
static std::mutex g_trace_mu;               // the trace record list is shared by all devices
static std::once_flag g_tuning_once;
static int g_tile_elems = 0;  // 0 = not yet initialised
static int g_grid_cap = 0;    // 0 = one CTA per tile
static bool g_pdl = true;     // programmatic dependent launch of the streaming kernels (B200MT_PDL=0 disables)
// Kernels that end every tile with a CTA-wide reduction (l2norm, LAMB stage 1) want longer tiles than the
// pure streaming ones: 8192 vs 4096 elements is 1.3x on l2norm and 1.1x on lamb_mp (profiles/sweep_norm_r01.txt);
// with the final kernels 8192 beats 16384 by 1-2 % on l2norm / LAMB and 5 % on 16-bit l2norm
// (profiles/sweep_norm2_r01.txt).  16-bit lists double it (same bytes per tile).
static int g_norm_tile_elems = 8192;

int set_error(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return 1;
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

const char* dtype_name(int d) {
  switch (d) {
    case B200MT_F32: return "Float";
    case B200MT_F16: return "Half";
    case B200MT_BF16: return "BFloat16";
    case B200MT_F64: return "Double";
    case B200MT_U8: return "Byte";
    default: return "Unknown";
  }
}

static void init_tuning_once() {
  g_tile_elems = 4096;  // swept on B200 (profiles/sweep_tiles_r01.txt): one CTA per 4096-element tile is fastest
  if (const char* e = getenv("B200MT_TILE_ELEMS")) { int v = atoi(e); if (v >= 64) g_tile_elems = v; }
  if (const char* e = getenv("B200MT_GRID_CAP")) { int v = atoi(e); if (v > 0) g_grid_cap = v; }
  if (const char* e = getenv("B200MT_PDL")) g_pdl = atoi(e) != 0;
  if (const char* e = getenv("B200MT_NORM_TILE_ELEMS")) { int v = atoi(e); if (v >= 64) g_norm_tile_elems = v; }
}
static void init_tuning() { std::call_once(g_tuning_once, init_tuning_once); }

// ---- launch trace -----------------------------------------------------------------------------------------------
namespace {
struct TraceRec { cudaEvent_t a = nullptr, b = nullptr; std::string name; };
std::vector<TraceRec> g_trace;      // pre-created event pairs
size_t g_trace_used = 0;
std::atomic<bool> g_tracing{false};
}  // namespace

bool trace_on() { return g_tracing.load(std::memory_order_relaxed); }

void trace_before(cudaStream_t stream, const char* mangled) {
  g_trace_mu.lock();                                // released by trace_after: a launch's two records stay adjacent
  if (g_trace_used >= g_trace.size()) return;       // capacity reached: later launches go unrecorded
  TraceRec& r = g_trace[g_trace_used];
  int st = 0;
  char* dem = abi::__cxa_demangle(mangled, nullptr, nullptr, &st);
  r.name = (st == 0 && dem) ? dem : mangled;
  free(dem);
  cudaEventRecord(r.a, stream);
}

void trace_after(cudaStream_t stream) {
  if (g_trace_used < g_trace.size()) {
    cudaEventRecord(g_trace[g_trace_used].b, stream);
    g_trace_used++;
  }
  g_trace_mu.unlock();
}

bool pdl_enabled() {
  init_tuning();
  return g_pdl;
}

Tiling elementwise_tiling() {
  init_tuning();
  Tiling t; t.tile_elems = g_tile_elems; t.ppc = 1; t.chunk_elems = g_tile_elems; return t;
}

Tiling chunked_tiling(int chunk_size, int tile_scale, int fixed_tile_elems) {
  init_tuning();
  Tiling t;
  t.chunk_elems = chunk_size;
  const int target = (fixed_tile_elems > 0 ? fixed_tile_elems : g_norm_tile_elems) * (tile_scale > 0 ? tile_scale : 1);
  if (chunk_size <= target) { t.tile_elems = chunk_size; t.ppc = 1; return t; }
  int ppc = (chunk_size + target - 1) / target;
  int tile = (chunk_size + ppc - 1) / ppc;
  tile = (tile + 15) & ~15;  // keep interior tile boundaries 16-element aligned
  t.ppc = (chunk_size + tile - 1) / tile;
  t.tile_elems = tile;
  return t;
}

int grid_for(int ntiles) {
  if (g_grid_cap > 0 && ntiles > g_grid_cap) return g_grid_cap;
  return ntiles;
}

// ---------------------------------------------------------------------------------------------
// per-device context
// ---------------------------------------------------------------------------------------------
namespace {

constexpr int kSlots = 64;
constexpr size_t kSpareCapBytes = 16u << 20;
constexpr int kMaxDevices = 64;

struct Slot {
  bool valid = false;
  size_t bytes = 0;   // image size
  size_t cap = 0;     // capacity of host/dev buffers
  char* host = nullptr;  // pinned shadow of the image
  char* dev = nullptr;
  cudaEvent_t uploaded = nullptr;
  std::vector<cudaStream_t> users;    // streams already ordered after the upload (emptied once the upload is known complete)
  std::vector<cudaStream_t> readers;  // distinct streams whose kernels may still read `dev` (consulted on eviction; bounded by
                                      // the number of streams the process uses, reset when the slot is refilled)
  uint64_t tick = 0;
  bool upload_done = false;  // the upload has been seen complete from the host: no stream needs to wait for it
  bool pinned = false;       // used while a stream was capturing: a CUDA graph holds `dev`, never evict
};

constexpr uint32_t kMagic = 0xb200;   // first word of every table image
constexpr size_t kAckBytes = 16;
inline volatile uint32_t* slot_ack(Slot* s) {
  return reinterpret_cast<volatile uint32_t*>(s->host + s->cap - kAckBytes);
}

struct StreamScratch { void* ptr = nullptr; size_t cap = 0; };

struct DevCtx {
  Slot slots[kSlots];
  std::map<cudaStream_t, StreamScratch> scratch;
  StreamScratch spare;          // zeroed, unassigned: handed to a stream first met while it is capturing
  std::vector<void*> retired;   // outgrown scratch buffers, possibly still in flight
  cudaEvent_t tmp_ev = nullptr;
  uint64_t tick = 0;
};

// Stream capture (CUDA graphs).  Nothing that allocates, synchronises or crosses the capture boundary may
// run while `stream` is capturing, and everything a captured kernel points at must outlive the graph:
//   * a list must have been seen once before the capture (its table is then resident and the upload is
//     known complete); the slot is pinned so that the graph's table pointer stays valid;
//   * scratch comes from the stream's own buffer if it has one, else from the pre-zeroed spare.
// Violations are reported, never papered over.
static int stream_is_capturing(cudaStream_t stream, bool* capturing) {
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  B200MT_CUDA(cudaStreamIsCapturing(stream, &st));
  *capturing = st != cudaStreamCaptureStatusNone;
  return 0;
}

DevCtx* g_ctx[kMaxDevices] = {nullptr};

struct ImageHeader {
  uint32_t magic, depth, ntensors, tile_elems, ppc, ntiles;
  int64_t chunk_elems;
};

thread_local std::vector<char> g_staging;

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

CallGuard::CallGuard(int dev) : status(0), prev_device(-1), device(dev) {
  if (dev < 0 || dev >= kMaxDevices) { device = -1; status = set_error("invalid device ordinal %d", dev); return; }
  g_dev_mu[dev].lock();
  cudaError_t e = cudaGetDevice(&prev_device);
  if (e != cudaSuccess) { status = set_error("CUDA error: %s (cudaGetDevice)", cudaGetErrorString(e)); return; }
  if (prev_device != dev) {
    e = cudaSetDevice(dev);
    if (e != cudaSuccess) { status = set_error("CUDA error: %s (cudaSetDevice %d)", cudaGetErrorString(e), dev); return; }
  }
}
CallGuard::~CallGuard() {
  if (device < 0) return;
  if (prev_device >= 0 && prev_device != device) cudaSetDevice(prev_device);
  g_dev_mu[device].unlock();
}

static DevCtx* ctx_for_current_device(int device) {
  if (!g_ctx[device]) g_ctx[device] = new DevCtx();
  return g_ctx[device];
}

int check_common(const b200mt_lists* L, int chunk_size) {
  B200MT_CHECK(L != nullptr, "lists is NULL");
  B200MT_CHECK(chunk_size > 0, "chunk_size must be > 0 (got %d)", chunk_size);
  B200MT_CHECK(L->depth >= 1 && L->depth <= B200MT_MAX_DEPTH, "tensor_lists.size() != depth");
  B200MT_CHECK(L->ntensors > 0, "tensor_lists[0].size() is not > 0");
  B200MT_CHECK(L->ptrs != nullptr && L->numel != nullptr, "lists->ptrs / lists->numel is NULL");
  return 0;
}

namespace {

struct ImageLayout { size_t off_ptrs, off_numel, off_prefix, bytes; long long ntiles; int nonempty, multichunk; };

// Builds the table image for (ptrs, numel) under `tiling` in the calling thread's staging buffer.
int build_image(const void* const* ptrs, const int64_t* numel, int D, int T, const Tiling& tiling, ImageLayout* lay,
                char** image) {
  lay->off_ptrs = sizeof(ImageHeader);
  lay->off_numel = lay->off_ptrs + (size_t)D * T * 8;
  lay->off_prefix = lay->off_numel + (size_t)T * 8;
  lay->bytes = align_up(lay->off_prefix + (size_t)(T + 1) * 4, 16);
  if (g_staging.size() < lay->bytes) g_staging.resize(lay->bytes);
  char* img = g_staging.data();
  memset(img + lay->off_prefix, 0, lay->bytes - lay->off_prefix);
  memcpy(img + lay->off_ptrs, ptrs, (size_t)D * T * 8);
  memcpy(img + lay->off_numel, numel, (size_t)T * 8);
  int* prefix = reinterpret_cast<int*>(img + lay->off_prefix);
  long long acc = 0;
  int nonempty = 0, multichunk = 0;
  for (int t = 0; t < T; t++) {
    const long long n = numel[t];
    B200MT_CHECK(n >= 0, "negative numel for tensor %d", t);
    nonempty += n > 0 ? 1 : 0;
    multichunk += n > tiling.chunk_elems ? 1 : 0;
    if (n > 0)
      for (int d = 0; d < D; d++)
        B200MT_CHECK(ptrs[(size_t)d * T + t] != nullptr, "NULL data pointer (list %d, tensor %d)", d, t);
    prefix[t] = (int)acc;
    const long long full = n / tiling.chunk_elems, rem = n % tiling.chunk_elems;
    acc += full * tiling.ppc + (rem + tiling.tile_elems - 1) / tiling.tile_elems;
    B200MT_CHECK(acc < INT_MAX, "tensor list too large: more than 2^31 tiles");
  }
  prefix[T] = (int)acc;
  ImageHeader* h = reinterpret_cast<ImageHeader*>(img);
  h->magic = 0xb200;
  h->depth = D; h->ntensors = T; h->tile_elems = tiling.tile_elems; h->ppc = tiling.ppc;
  h->ntiles = (uint32_t)acc; h->chunk_elems = tiling.chunk_elems;
  lay->ntiles = acc;
  lay->nonempty = nonempty;
  lay->multichunk = multichunk;
  *image = img;
  return 0;
}

// A resident table is about to be used on `stream`: order the stream after the upload if it has not been seen
// complete yet (first use on this stream), pin the slot while a capture refers to it.
int attach_stream(Slot* slot, cudaStream_t stream, bool capturing) {
  // the upload is followed by a 4-byte copy of the table's magic word back into pinned memory: seeing it there
  // proves the upload is over without any CUDA call (none that queries or waits is safe while capturing)
  if (!slot->upload_done && *slot_ack(slot) == kMagic) slot->upload_done = true;
  if (slot->upload_done) {
    slot->users.clear();          // nobody has to be ordered after the upload any more: no need to remember streams
  } else {
    bool known = false;
    for (cudaStream_t u : slot->users) known = known || (u == stream);
    if (!known) {  // first use on this stream: order after the upload
      B200MT_CHECK(!capturing, "stream capture: the upload of this tensor list's descriptor table is still in flight on "
                               "another stream; synchronise before capturing");
      B200MT_CUDA(cudaStreamWaitEvent(stream, slot->uploaded, 0));
      slot->users.push_back(stream);
    }
  }
  if (capturing) slot->pinned = true;
  return 0;
}

// (Re)fills `slot` with `img` and uploads it on `stream`.  The slot must not be in use by kernels of other streams
// (the LRU path orders `stream` after them before calling this; a prepared handle's slot is fresh).
int upload_slot(DevCtx* ctx, Slot* slot, const char* img, size_t bytes, cudaStream_t stream) {
  if (!slot->uploaded) B200MT_CUDA(cudaEventCreateWithFlags(&slot->uploaded, cudaEventDisableTiming));
  slot->valid = false;
  slot->upload_done = false;
  slot->users.clear();
  slot->readers.clear();
  if (slot->cap < bytes + kAckBytes) {
    // Growing: the old device buffer may be read by kernels on this stream -> retire, don't free.
    if (slot->dev) ctx->retired.push_back(slot->dev);
    if (slot->host) cudaFreeHost(slot->host);
    slot->host = slot->dev = nullptr; slot->cap = 0;
    size_t cap = 16384;
    while (cap < bytes + kAckBytes) cap *= 2;     // the last kAckBytes of the pinned shadow hold the upload acknowledgement
    B200MT_CUDA(cudaHostAlloc((void**)&slot->host, cap, cudaHostAllocDefault));
    B200MT_CUDA(cudaMalloc((void**)&slot->dev, cap));
    slot->cap = cap;
  }
  memcpy(slot->host, img, bytes);
  *slot_ack(slot) = 0;
  B200MT_CUDA(cudaMemcpyAsync(slot->dev, slot->host, bytes, cudaMemcpyHostToDevice, stream));
  B200MT_CUDA(cudaMemcpyAsync(const_cast<uint32_t*>(slot_ack(slot)), slot->dev, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
  B200MT_CUDA(cudaEventRecord(slot->uploaded, stream));
  slot->bytes = bytes; slot->valid = true;
  slot->users.push_back(stream);
  g_uploads.fetch_add(1, std::memory_order_relaxed);
  return 0;
}

void fill_view(const Slot* slot, const ImageLayout& lay, int T, const Tiling& tiling, TableView* out) {
  out->ptrs = reinterpret_cast<const unsigned long long*>(slot->dev + lay.off_ptrs);
  out->numel = reinterpret_cast<const long long*>(slot->dev + lay.off_numel);
  out->tile_prefix = reinterpret_cast<const int*>(slot->dev + lay.off_prefix);
  out->ntensors = T;
  out->ntiles = (int)lay.ntiles;
  out->nonempty = lay.nonempty;
  out->multichunk = lay.multichunk;
  out->tile_elems = tiling.tile_elems;
  out->ppc = tiling.ppc;
  out->chunk_elems = tiling.chunk_elems;
}

inline void note_reader(Slot* slot, cudaStream_t stream) {
  for (cudaStream_t u : slot->readers) if (u == stream) return;
  slot->readers.push_back(stream);
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// prepared list sets (b200mt_lists_prepare): caller-owned resident tables, no per-call image / compare
// ---------------------------------------------------------------------------------------------
}  // namespace b200mt

struct b200mt_prepared {
  uint32_t magic = 0;
  int device = 0, depth = 0, ntensors = 0;
  std::vector<const void*> ptrs;
  std::vector<int64_t> numel;
  struct Entry { b200mt::Tiling tiling; b200mt::ImageLayout lay; b200mt::Slot slot; };
  std::vector<Entry*> entries;    // one resident table per tiling the handle has been used with (element-wise, norm ...)
};

namespace b200mt {

constexpr uint32_t kPreparedMagic = 0xb2009e9aU;

static int acquire_prepared(const b200mt_lists* L, const Tiling& tiling, cudaStream_t stream, TableView* out) {
  b200mt_prepared* h = const_cast<b200mt_prepared*>(reinterpret_cast<const b200mt_prepared*>(L->prepared));
  B200MT_CHECK(h->magic == kPreparedMagic, "lists->prepared is not a live handle of b200mt_lists_prepare");
  B200MT_CHECK(h->device == L->device && h->depth == L->depth && h->ntensors == L->ntensors,
               "lists->prepared was made for another list set (device %d depth %d tensors %d, now %d %d %d)", h->device, h->depth,
               h->ntensors, L->device, L->depth, L->ntensors);
  DevCtx* ctx = ctx_for_current_device(L->device);
  bool capturing = false;
  if (stream_is_capturing(stream, &capturing)) return 1;
  b200mt_prepared::Entry* e = nullptr;
  for (b200mt_prepared::Entry* c : h->entries)
    if (c->tiling.tile_elems == tiling.tile_elems && c->tiling.ppc == tiling.ppc && c->tiling.chunk_elems == tiling.chunk_elems) { e = c; break; }
  if (e) {
    g_hits.fetch_add(1, std::memory_order_relaxed);
    if (attach_stream(&e->slot, stream, capturing)) return 1;
  } else {
    B200MT_CHECK(!capturing, "stream capture: this prepared list set has not been used with this kind of op before the capture "
                             "began; run the captured calls once (warm-up) first");
    e = new b200mt_prepared::Entry();
    e->tiling = tiling;
    char* img = nullptr;
    if (build_image(h->ptrs.data(), h->numel.data(), h->depth, h->ntensors, tiling, &e->lay, &img) ||
        upload_slot(ctx, &e->slot, img, e->lay.bytes, stream)) {
      delete e;
      return 1;
    }
    h->entries.push_back(e);
  }
  fill_view(&e->slot, e->lay, h->ntensors, tiling, out);
  return 0;
}

int acquire_table(const b200mt_lists* L, int dmin, int dmax, const Tiling& tiling,
                  cudaStream_t stream, TableView* out) {
  B200MT_CHECK(L->depth >= dmin && L->depth <= dmax, "tensor_lists.size() != depth");
  if (L->prepared) return acquire_prepared(L, tiling, stream, out);
  const int T = L->ntensors, D = L->depth;
  DevCtx* ctx = ctx_for_current_device(L->device);

  // ---- build the image in host staging memory
  ImageLayout lay;
  char* img = nullptr;
  if (build_image(L->ptrs, L->numel, D, T, tiling, &lay, &img)) return 1;
  const size_t bytes = lay.bytes;

  // ---- look it up: images of equal size are compared directly.  memcmp leaves at the first differing word (two list sets
  // differ in their first pointers), so the 64 slots cost a few cache lines each and only the hit is read in full -- a hash of
  // the image (round 1: FNV-1a) read every byte once more for nothing: 30 us per call on a 20,000-pointer list.
  Slot* slot = nullptr;
  for (int i = 0; i < kSlots; i++) {
    Slot& s = ctx->slots[i];
    if (s.valid && s.bytes == bytes && memcmp(s.host, img, bytes) == 0) { slot = &s; break; }
  }
  bool capturing = false;
  if (stream_is_capturing(stream, &capturing)) return 1;
  if (slot) {
    g_hits.fetch_add(1, std::memory_order_relaxed);
    if (attach_stream(slot, stream, capturing)) return 1;
  } else {
    B200MT_CHECK(!capturing, "stream capture: this tensor list has not been seen before the capture began; run the "
                             "captured calls once (warm-up) with the same tensors first");
    // evict the least recently used slot that no CUDA graph refers to
    for (int i = 0; i < kSlots; i++) {
      Slot& s = ctx->slots[i];
      if (!s.valid) { slot = &s; break; }
      if (s.pinned) continue;
      if (!slot || s.tick < slot->tick) slot = &s;
    }
    B200MT_CHECK(slot != nullptr, "all %d descriptor-table slots of device %d are held by captured CUDA graphs", kSlots, L->device);
    if (!ctx->tmp_ev) B200MT_CUDA(cudaEventCreateWithFlags(&ctx->tmp_ev, cudaEventDisableTiming));
    if (slot->valid) {
      // (a) the pinned shadow may still be the source of an in-flight upload
      if (slot->uploaded) B200MT_CUDA(cudaEventSynchronize(slot->uploaded));
      // (b) kernels on other streams may still read the device copy: order this stream after them
      for (cudaStream_t u : slot->readers) {
        if (u == stream) continue;
        B200MT_CUDA(cudaEventRecord(ctx->tmp_ev, u));
        B200MT_CUDA(cudaStreamWaitEvent(stream, ctx->tmp_ev, 0));
      }
    }
    if (upload_slot(ctx, slot, img, bytes, stream)) return 1;
  }
  note_reader(slot, stream);
  slot->tick = ++ctx->tick;
  fill_view(slot, lay, T, tiling, out);
  return 0;
}

int acquire_scratch(cudaStream_t stream, size_t bytes, void** out) {
  int device = 0;
  B200MT_CUDA(cudaGetDevice(&device));
  DevCtx* ctx = ctx_for_current_device(device);
  bool capturing = false;
  if (stream_is_capturing(stream, &capturing)) return 1;
  StreamScratch& s = ctx->scratch[stream];
  if (s.cap < bytes && capturing) {
    // no allocation while capturing: adopt the spare (zeroed with a blocking memset when it was made)
    B200MT_CHECK(ctx->spare.cap >= bytes, "stream capture: %zu bytes of scratch are needed on the capturing stream and none of "
                 "that size exists yet; run the captured calls once before capturing (on the capture stream itself if the "
                 "lists are larger than the warm-up's)", bytes);
    if (s.ptr) ctx->retired.push_back(s.ptr);
    s = ctx->spare;
    ctx->spare = StreamScratch();
  } else if (s.cap < bytes) {
    if (s.ptr) ctx->retired.push_back(s.ptr);  // may still be in flight on this stream
    size_t cap = 1 << 20;
    while (cap < bytes) cap *= 2;
    s.ptr = nullptr; s.cap = 0;
    B200MT_CUDA(cudaMalloc(&s.ptr, cap));
    s.cap = cap;
    B200MT_CUDA(cudaMemsetAsync(s.ptr, 0, kScratchZeroBytes, stream));
  }
  // keep one small unassigned buffer for a stream that shows up already capturing (no allocation is possible then).
  // Capped: mirroring the largest buffer in use would double the scratch of Llama-sized norm calls for a rare case;
  // a capture that needs more is told to warm up on the capture stream (message above).
  const size_t spare_want = s.cap < kSpareCapBytes ? s.cap : kSpareCapBytes;
  if (!capturing && ctx->spare.cap < spare_want) {
    if (ctx->spare.ptr) ctx->retired.push_back(ctx->spare.ptr);
    ctx->spare = StreamScratch();
    B200MT_CUDA(cudaMalloc(&ctx->spare.ptr, spare_want));
    ctx->spare.cap = spare_want;
    B200MT_CUDA(cudaMemset(ctx->spare.ptr, 0, kScratchZeroBytes));
  }
  *out = s.ptr;
  return 0;
}

}  // namespace b200mt

// ---------------------------------------------------------------------------------------------
// C ABI: library-level entry points
// ---------------------------------------------------------------------------------------------
using namespace b200mt;

extern "C" const char* b200mt_last_error(void) { return g_err.c_str(); }
extern "C" int b200mt_version(void) { return 101; }
extern "C" int64_t b200mt_launch_count(void) { return g_launches.load(); }
extern "C" int64_t b200mt_table_uploads(void) { return g_uploads.load(); }
extern "C" int64_t b200mt_table_hits(void) { return g_hits.load(); }

extern "C" int b200mt_lists_prepare(const b200mt_lists* L, b200mt_prepared** out) {
  B200MT_CHECK(out != nullptr, "b200mt_lists_prepare: out is NULL");
  *out = nullptr;
  if (check_common(L, 1)) return 1;
  B200MT_CHECK(L->prepared == nullptr, "b200mt_lists_prepare: lists->prepared must be NULL");
  for (int t = 0; t < L->ntensors; t++) B200MT_CHECK(L->numel[t] >= 0, "negative numel for tensor %d", t);
  b200mt_prepared* h = new b200mt_prepared();
  h->magic = kPreparedMagic;
  h->device = L->device; h->depth = L->depth; h->ntensors = L->ntensors;
  h->ptrs.assign(L->ptrs, L->ptrs + (size_t)L->depth * L->ntensors);
  h->numel.assign(L->numel, L->numel + L->ntensors);
  *out = h;
  return 0;
}

extern "C" int b200mt_lists_release(b200mt_prepared* h) {
  if (!h) return 0;
  B200MT_CHECK(h->magic == kPreparedMagic, "b200mt_lists_release: not a live handle");
  CallGuard guard(h->device);
  if (guard.status) return 1;
  h->magic = 0;
  for (b200mt_prepared::Entry* e : h->entries) {
    // kernels in flight may still read the table: cudaFree waits for the device, which a release may afford
    if (e->slot.uploaded) { cudaEventSynchronize(e->slot.uploaded); cudaEventDestroy(e->slot.uploaded); }
    if (e->slot.dev) cudaFree(e->slot.dev);
    if (e->slot.host) cudaFreeHost(e->slot.host);
    delete e;
  }
  delete h;
  return 0;
}

extern "C" int b200mt_trace_begin(int max_launches) {
  std::lock_guard<std::recursive_mutex> lk(g_mu);
  B200MT_CHECK(max_launches > 0 && max_launches <= 65536, "b200mt_trace_begin: max_launches must be in 1..65536");
  B200MT_CHECK(!g_tracing.load(), "b200mt_trace_begin: a trace is already running");
  while ((int)g_trace.size() < max_launches) {
    TraceRec r;
    B200MT_CUDA(cudaEventCreate(&r.a));
    B200MT_CUDA(cudaEventCreate(&r.b));
    g_trace.push_back(r);
  }
  g_trace_used = 0;
  g_tracing.store(true);
  return 0;
}

extern "C" int b200mt_trace_end(float* ms, int capacity, int* count) {
  std::lock_guard<std::recursive_mutex> lk(g_mu);
  B200MT_CHECK(g_tracing.load(), "b200mt_trace_end: no trace is running");
  g_tracing.store(false);
  const int n = (int)g_trace_used;
  if (count) *count = n;
  for (int i = 0; i < n && i < capacity; i++) {
    B200MT_CUDA(cudaEventSynchronize(g_trace[i].b));
    B200MT_CUDA(cudaEventElapsedTime(&ms[i], g_trace[i].a, g_trace[i].b));
  }
  return 0;
}

extern "C" const char* b200mt_trace_name(int i) {
  std::lock_guard<std::recursive_mutex> lk(g_mu);
  return (i >= 0 && i < (int)g_trace_used) ? g_trace[i].name.c_str() : "";
}

extern "C" int b200mt_set_tuning(int tile_elems, int grid_cap) {
  std::lock_guard<std::recursive_mutex> lk(g_mu);
  init_tuning();
  if (tile_elems > 0) {
    B200MT_CHECK(tile_elems >= 64 && tile_elems % 16 == 0, "tile_elems must be a multiple of 16, >= 64");
    g_tile_elems = tile_elems;
  }
  if (grid_cap >= 0) g_grid_cap = grid_cap;
  return 0;
}

extern "C" int b200mt_shutdown(void) {
  std::lock_guard<std::recursive_mutex> lk(g_mu);
  int prev = -1;
  cudaGetDevice(&prev);
  for (int d = 0; d < kMaxDevices; d++) {
    std::lock_guard<std::recursive_mutex> dlk(g_dev_mu[d]);
    DevCtx* c = g_ctx[d];
    if (!c) continue;
    cudaSetDevice(d);
    for (int i = 0; i < kSlots; i++) {
      Slot& s = c->slots[i];
      if (s.dev) cudaFree(s.dev);
      if (s.host) cudaFreeHost(s.host);
      if (s.uploaded) cudaEventDestroy(s.uploaded);
    }
    for (auto& kv : c->scratch) if (kv.second.ptr) cudaFree(kv.second.ptr);
    if (c->spare.ptr) cudaFree(c->spare.ptr);
    for (void* p : c->retired) cudaFree(p);
    if (c->tmp_ev) cudaEventDestroy(c->tmp_ev);
    delete c;
    g_ctx[d] = nullptr;
  }
  if (prev >= 0) cudaSetDevice(prev);
  return 0;
}
