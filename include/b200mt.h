/*
 * b200mt.h -- C ABI of libb200mt.so: the multi-tensor fused-optimizer hot path for B200 (sm_100a).
 *
 * This is the drop-in boundary.  Every entry point replaces one function of the reference's
 * `amp_C` / `distributed_adam_cuda` / `fused_adam_cuda` pybind modules (cited per function as
 * path:line under the reference tree).  Plain pointers and sizes only -- no torch / ATen types.
 *
 * Conventions
 *  - A "tensor list set" is `depth` lists of `ntensors` device pointers each, passed list-major:
 *    ptrs[d*ntensors + t] is tensor t of list d.  All lists share numel[t] (elements, int64).
 *    The element type of list d is dtype[d] (B200MT_F32 ...).  Tensors must be dense; element 0
 *    of tensor t is at ptrs[..]; there is no alignment requirement (misaligned views take a
 *    peeled/scalar path).
 *  - Every call is asynchronous on `stream` (a cudaStream_t passed as void*; NULL = legacy
 *    default stream) of device `device`.  No call synchronises the host with the GPU.
 *  - `noop` is a device pointer to one int32 (the reference's `noop_flag` / `_overflow_buf`).
 *  - `chunk_size` keeps the reference's meaning where it is observable: the L2-norm kernels raise
 *    `noop` when the sum of squares over one chunk of `chunk_size` elements is non-finite
 *    (csrc/multi_tensor_l2norm_kernel.cu:99-108).  Work is scheduled in tiles that subdivide
 *    chunks; element-wise ops ignore chunk_size apart from validating it (> 0).
 *  - Return value: 0 on success, non-zero on error; b200mt_last_error() returns the message of
 *    the last failing call on the calling thread.
 *  - The library owns two kinds of internal device memory per device: the cache of tensor-list
 *    descriptor tables (uploaded once per distinct list set, reused while pointers are
 *    unchanged) and a per-stream scratch buffer for per-tile partial sums.  Both are created
 *    lazily with cudaMalloc and released by b200mt_shutdown().
 */
#ifndef B200MT_H_
#define B200MT_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { B200MT_F32 = 0, B200MT_F16 = 1, B200MT_BF16 = 2, B200MT_F64 = 3, B200MT_U8 = 4 };

enum { B200MT_MAX_DEPTH = 5 };

typedef struct b200mt_lists {
  const void* const* ptrs;   /* depth*ntensors device pointers, list-major            */
  const int64_t*     numel;  /* ntensors element counts                               */
  int32_t            ntensors;
  int32_t            depth;
  int32_t            dtype[B200MT_MAX_DEPTH];
  int32_t            device; /* CUDA device ordinal the tensors live on               */
  const struct b200mt_prepared* prepared; /* NULL, or a handle of b200mt_lists_prepare for exactly this list set */
} b200mt_lists;

/* ---- library ---------------------------------------------------------------------------- */
const char* b200mt_last_error(void);
int  b200mt_version(void);
/* Number of kernels this library has launched since load (process-wide, all devices). */
int64_t b200mt_launch_count(void);
/* Descriptor-table cache statistics: uploads performed / cache hits. */
int64_t b200mt_table_uploads(void);
int64_t b200mt_table_hits(void);
/* Tuning knobs (0 = keep default): elements per scheduling tile, CTAs per SM cap. */
int  b200mt_set_tuning(int tile_elems, int grid_cap);
/* Frees cached descriptor tables and scratch on every device.  Caller must have synchronised. */
int  b200mt_shutdown(void);

/* ---- amp_C.multi_tensor_scale  (csrc/multi_tensor_scale_kernel.cu:113-136, amp_C_frontend.cpp:3-7)
 * lists = [in, out]; out = (out_t)((float)in * scale); *noop = 1 iff any input is inf/NaN.
 * Never reads noop.  in may alias out. */
int b200mt_scale(const b200mt_lists* lists, int chunk_size, int32_t* noop, float scale, void* stream);

/* ---- amp_C.multi_tensor_axpby  (csrc/multi_tensor_axpby_kernel.cu:128-157)
 * lists = [x, y, out]; out = a*x + b*y; arg_to_check: -1 both, 0 x, 1 y, anything else none. */
int b200mt_axpby(const b200mt_lists* lists, int chunk_size, int32_t* noop, float a, float b,
                 int arg_to_check, void* stream);

/* ---- amp_C.multi_tensor_l2norm / _l2norm_mp / _l2norm_scale
 *      (csrc/multi_tensor_l2norm_kernel.cu:301-364, multi_tensor_l2norm_kernel_mp.cu:153-216,
 *       multi_tensor_l2norm_scale_kernel.cu:258-324)
 * lists = [x]  (l2norm_scale: [in, out], out = in*scale written in the same pass).
 * ret: device float[1] = sqrt(sum x^2);  ret_per_tensor: device float[ntensors] or NULL.
 * Deterministic: per-tile partials are reduced in a fixed order by a second pass; no float atomics.
 * flags: B200MT_NORM_NOOP_GATED makes every kernel return immediately when *noop != 0 and leaves
 * ret / ret_per_tensor untouched (the _mp variant). */
enum { B200MT_NORM_NOOP_GATED = 1 };
int b200mt_l2norm(const b200mt_lists* lists, int chunk_size, int32_t* noop, float* ret,
                  float* ret_per_tensor, int flags, void* stream);
int b200mt_l2norm_scale(const b200mt_lists* lists, int chunk_size, int32_t* noop, float scale,
                        float* ret, float* ret_per_tensor, void* stream);

/* ---- amp_C.multi_tensor_adam  (csrc/multi_tensor_adam.cu:129-171)
 * lists = [g, p, m, v] (one dtype for all four, as in the reference) or, as an extension,
 * [g, p, m, v, p_copy] / mixed dtypes (g 16-bit, p/m/v fp32, p_copy 16-bit): a single fused
 * read-modify-write pass.  mode 0 = L2 (decay added to g), 1 = AdamW.  Bias corrections are
 * computed here on the host exactly as the reference does (1 - pow((double)beta, step) -> float).
 * grad_scale multiplies g on load (1.0f = reference behaviour).  No noop interaction. */
int b200mt_adam(const b200mt_lists* lists, int chunk_size, int32_t* noop, float lr, float beta1,
                float beta2, float eps, int step, int mode, int bias_correction,
                float weight_decay, float grad_scale, void* stream);

/* ---- amp_C.multi_tensor_sgd  (csrc/multi_tensor_sgd_kernel.cu:141-322)
 * lists = [g, p, mom] or [g, p, mom, p_model_fp16].  Returns immediately on the device when *noop. */
int b200mt_sgd(const b200mt_lists* lists, int chunk_size, int32_t* noop, float wd, float momentum,
               float dampening, float lr, int nesterov, int first_run, int wd_after_momentum,
               float scale, void* stream);

/* ---- amp_C.multi_tensor_lamb  (csrc/multi_tensor_lamb.cu:332-413)
 * lists = [g, p, m, v], one dtype.  global_grad_norm: device float[1].  The per-tensor ||p|| and
 * ||update|| partial sums are produced inside the stage-1 pass (two full reads fewer than the
 * reference's four launches); g is overwritten with the update, as in the reference. */
int b200mt_lamb(const b200mt_lists* lists, int chunk_size, int32_t* noop, float lr, float beta1,
                float beta2, float eps, int step, int bias_correction, float weight_decay,
                int grad_averaging, int mode, const float* global_grad_norm, float max_grad_norm,
                int use_nvlamb, void* stream);

/* ---- amp_C.multi_tensor_lamb_mp  (csrc/multi_tensor_lamb_mp.cu:367-496)
 * lists = [g, p32, m32, v32] or [g, p32, m32, v32, p_model(g's dtype)].  lr, step, global_grad_norm,
 * max_grad_norm, found_inf, inv_scale are device scalars (float / int32).  All kernels return
 * immediately when *noop != 0. */
int b200mt_lamb_mp(const b200mt_lists* lists, int chunk_size, int32_t* noop, const float* lr,
                   float beta1, float beta2, float eps, const int32_t* step, int bias_correction,
                   float weight_decay, int grad_averaging, int mode, const float* global_grad_norm,
                   const float* max_grad_norm, int use_nvlamb, const float* found_inf,
                   const float* inv_scale, void* stream);

/* ---- amp_C.multi_tensor_lamb_stage1_cuda / stage2_cuda
 *      (csrc/multi_tensor_lamb_stage_1.cu:112-151, csrc/multi_tensor_lamb_stage_2.cu:94-125)
 * stage1 lists = [g, p, m, v, update]; per_tensor_decay: device float[ntensors];
 * clipped_global_grad_norm is the host-side clip factor (reference dereferences a CPU tensor).
 * stage2 lists = [p, update]; norms: device float[ntensors]. */
int b200mt_lamb_stage1(const b200mt_lists* lists, int chunk_size, int32_t* noop,
                       const float* per_tensor_decay, int step, float beta1, float beta2,
                       float eps, float clipped_global_grad_norm, void* stream);
int b200mt_lamb_stage2(const b200mt_lists* lists, int chunk_size, int32_t* noop,
                       const float* per_tensor_param_norm, const float* per_tensor_update_norm,
                       float lr, float weight_decay, int use_nvlamb, void* stream);

/* ---- amp_C.multi_tensor_adagrad  (csrc/multi_tensor_adagrad.cu:85-100)
 * lists = [g, p, h], one dtype.  mode 0: L2 (decay folded into g), 1: decoupled decay. */
int b200mt_adagrad(const b200mt_lists* lists, int chunk_size, int32_t* noop, float lr, float eps, int mode,
                   float weight_decay, void* stream);

/* ---- amp_C.multi_tensor_novograd  (csrc/multi_tensor_novograd.cu:120-188)
 * lists = [g, p, m], one dtype.  grad_norms: device float[ntensors], updated in place first
 * (multi_tensor_norm_out_cuda, csrc/multi_tensor_l2norm_kernel.cu:371-456): norm_type 2 -> L2 blend
 * sqrt(beta2*gn^2 + (1-beta2)*|g|^2), norm_type 0 -> L-inf blend beta2*gn + (1-beta2)*max|g|; then the update.
 * The norm pass raises *noop on a non-finite chunk. */
int b200mt_novograd(const b200mt_lists* lists, int chunk_size, int32_t* noop, float* grad_norms, float lr,
                    float beta1, float beta2, float eps, int step, int bias_correction, float weight_decay,
                    int grad_averaging, int moment_mode, int norm_type, void* stream);

/* ---- distributed_adam_cuda.multi_tensor_fused_adam
 *      (apex/contrib/csrc/optimizers/multi_tensor_distopt_adam_kernel.cu:169-228)
 * lists = [p32, m32, v32, g] or [p32, m32, v32, g, p_copy(g's dtype)].  Per-tensor device arrays
 * beta1/beta2/eps/weight_decay (float[ntensors]) and bias_correction (int32[ntensors]).
 * mode 0: eps inside sqrt, 1: outside.
 * p_copy may also be B200MT_U8: e5m2 bytes, the payload of the reference's `e5m2_allgather=True`
 * (apex/contrib/optimizers/distributed_fused_adam.py:198,206), packed as fused_adam_cuda_kernel.cu:395-463
 * packs them.  The reference's own kernel writes p_copy through a pointer of g's type whatever the buffer
 * holds (multi_tensor_distopt_adam_kernel.cu:75-76,161), so this is the behaviour its Python side expects,
 * not the one its kernel has. */
int b200mt_dist_adam(const b200mt_lists* lists, int chunk_size, int32_t* noop,
                     const float* per_tensor_beta1, const float* per_tensor_beta2,
                     const int32_t* per_tensor_bias_correction, const float* per_tensor_eps,
                     const float* per_tensor_weight_decay, float lr, float grad_scale, int step,
                     int mode, void* stream);

/* ---- distributed_lamb_cuda.multi_tensor_lamb_compute_update_term / multi_tensor_lamb_update_weights
 *      (apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb.cpp:3-36, ..._lamb_kernel.cu:108-506)
 * The sharded LAMB step of DistributedFusedLAMB.  Per-fragment hyper-parameters are device arrays; step,
 * global_scale, global_grad_norm and lr are device scalars; both kernels return at once when *noop == 1.
 * stage1 lists = [g, p, m, v, u]: u <- Adam-style update of g/combined_scale (clip folded into combined_scale),
 *   m, v updated.  Extension: when update_sumsq != NULL it receives, per fragment, the sum of squares of u
 *   (raw, so that it can be all-reduced over ranks before the square root); pass NULL for reference behaviour.
 * stage2 lists = [u, p, p_copy]: p -= ratio*u with ratio = lr*param_norm[t]/update_norm[update_norm_offset[t]]
 *   when use_nvlamb or decay[t] != 0, else lr; p_copy = (its dtype, Half / BFloat16 / e5m2 byte / Float) p. */
int b200mt_dist_lamb_stage1(const b200mt_lists* lists, int chunk_size, int32_t* noop,
                            const float* per_tensor_beta1, const float* per_tensor_beta2,
                            const float* per_tensor_beta3, const int32_t* per_tensor_bias_correction,
                            const int32_t* step, const float* per_tensor_eps, int mode,
                            const float* per_tensor_decay, const float* global_scale,
                            const float* global_grad_norm, float max_grad_norm, float* update_sumsq,
                            void* stream);
int b200mt_dist_lamb_stage2(const b200mt_lists* lists, int chunk_size, int32_t* noop,
                            const float* per_tensor_param_norm, const float* per_tensor_update_norm,
                            const int64_t* update_norm_offset, const float* lr,
                            const float* per_tensor_decay, const float* global_grad_norm, int use_nvlamb,
                            void* stream);

/* ---- b200mt_dist_adam_nvls -- the sharded step of DistributedFusedAdam as ONE kernel (extension; it
 * replaces the reference's three-step sequence reduce_scatter -> multi_tensor_fused_adam -> all_gather,
 * apex/contrib/optimizers/distributed_fused_adam.py:409,443-467,477).
 * lists = [p32, m32, v32, g_mc, p_copy_mc]: lists 3 and 4 hold addresses inside CUDA multicast objects
 * (NVLS) bound to the same symmetric buffers on every GPU of the NVSwitch domain, at this rank's shard.
 * The gradient load is multimem.ld_reduce (summed over all ranks inside the switch, fp32 accumulation),
 * the parameter store is multimem.st (replicated to all ranks by the switch): no separate collective.
 * dtype[3] in {B200MT_F16, B200MT_BF16}; dtype[4] the same, or B200MT_U8 for an e5m2-compressed
 * parameter broadcast (half the bytes through the switch).  Every fragment must be a multiple of 8
 * elements; lists 0-2 32-byte, list 3 16-byte, list 4 16-byte (8-byte when e5m2) aligned.  The caller brackets the launch with cross-GPU
 * barriers (all ranks' gradients written before; all ranks' kernels finished after). */
/* measurement switches, OR-ed into `mode`: the named list holds ordinary device addresses, so the kernel keeps only one
 * of its two collectives (tools/nvls_probe.py times the three forms) */
#define B200MT_NVLS_LOCAL_GRADS 0x100
#define B200MT_NVLS_LOCAL_COPY  0x200
int b200mt_dist_adam_nvls(const b200mt_lists* lists, int chunk_size, int32_t* noop,
                          const float* per_tensor_beta1, const float* per_tensor_beta2,
                          const int32_t* per_tensor_bias_correction, const float* per_tensor_eps,
                          const float* per_tensor_weight_decay, float lr, float grad_scale, int step,
                          int mode, void* stream);

/* ---- b200mt_dist_adam_p2p -- the same sharded step for TWO ranks over peer-mapped memory (extension, no reference
 * counterpart beyond the three-step sequence above).  lists = [p32, m32, v32, g, p_copy]: g and p_copy are this rank's
 * addresses (BFloat16) inside buffers that have a twin in the peer GPU's memory, mapped into this process at
 * `grad_peer_delta` / `copy_peer_delta` bytes from them (symmetric memory: peer base - own base).  The gradient load adds
 * the peer's value (fp32 sum, one rounding to bf16: the two-rank reduce-scatter), the parameter store also writes the
 * peer's twin (the two-rank all-gather).  Per direction 4 bytes per shard element cross the link; the multicast kernel
 * needs 6 at W = 2 because the switch pulls the requester's own half through the links too.  Fragments: multiples of 8
 * elements, 16-byte aligned (fp32 lists 32-byte).  Barriers around the launch as for b200mt_dist_adam_nvls.  The distances
 * are written to constant memory on `stream` before the launch: concurrent launches with DIFFERENT distances must share a stream. */
int b200mt_dist_adam_p2p(const b200mt_lists* lists, int chunk_size, int32_t* noop,
                         const float* per_tensor_beta1, const float* per_tensor_beta2,
                         const int32_t* per_tensor_bias_correction, const float* per_tensor_eps,
                         const float* per_tensor_weight_decay, float lr, float grad_scale, int step,
                         int mode, int64_t grad_peer_delta, int64_t copy_peer_delta, void* stream);

/* ---- b200mt_nvls_reduce_scatter / b200mt_nvls_all_gather -- the two collectives of the sharded step
 * (torch.distributed.reduce_scatter / all_gather at apex/contrib/optimizers/distributed_fused_adam.py:409,477)
 * as stand-alone kernels over NVSwitch multicast memory, for callers that want them unfused.
 * reduce_scatter: lists = [g_mc, out]: out = sum over ranks of g (multimem.ld_reduce, fp32 accumulation);
 * all_gather:     lists = [in, out_mc]: in is replicated into every rank's buffer (multimem.st).
 * The *_mc lists hold multicast addresses; dtype Half or BFloat16; fragments are multiples of 8 elements,
 * 16-byte aligned.  The caller provides the cross-GPU barriers, as for b200mt_dist_adam_nvls. */
int b200mt_nvls_reduce_scatter(const b200mt_lists* lists, int chunk_size, void* stream);
int b200mt_nvls_all_gather(const b200mt_lists* lists, int chunk_size, void* stream);

/* ---- fused_adam_cuda.adam_mt / maybe_cast_mt / strided_check_finite
 *      (apex/contrib/csrc/optimizers/fused_adam_cuda_kernel.cu:283-387, :939-960, :785-812)
 * adam_mt lists = [p, m, v, g] or [p, m, v, g, p_copy]; step_size form (eps not rescaled). */
int b200mt_contrib_adam(const b200mt_lists* lists, int chunk_size, int32_t* noop, float lr,
                        float beta1, float beta2, float eps, float grad_scale, int step, int mode,
                        int bias_correction, float decay, void* stream);
int b200mt_maybe_cast(const b200mt_lists* lists, int chunk_size, const int32_t* overflow_flag,
                      void* stream);
int b200mt_strided_check_finite(int32_t* noop, const void* p, int dtype, int64_t numel,
                                int stride, int clear_overflow_first, int device, void* stream);

/* ---- fused_adam_cuda.reversible_adam / maybe_adam_undo
 *      (apex/contrib/csrc/optimizers/fused_adam_cuda_kernel.cu:570-653,814-912 and :655-732,961-1040)
 * reversible_adam lists = [p32, m32, v32, g] or [.., p_copy (g's dtype, or e5m2 bytes for 16-bit g)]: adam_mt's
 *   update where g/grad_scale is finite; elsewhere the element is left alone and the first element of p_copy
 *   is overwritten with +inf (found later by strided_check_finite).
 * maybe_adam_undo lists = [p32, m32, v32, g]: the inverse of that step, applied when overflow_flag is NULL or
 *   *overflow_flag != 0.  Every element is reverted; the reference's kernel only reverts elements 0 mod 4
 *   (its index inside the ILP loop is `j_start + i*ILP`, :687-690) -- see csrc/ops_reversible.cu.
 * The reference binds these for flat tensors; lists of one tensor reproduce that, longer lists are an extension. */
int b200mt_reversible_adam(const b200mt_lists* lists, int chunk_size, float lr, float beta1, float beta2,
                           float eps, float grad_scale, int step, int mode, int bias_correction,
                           float decay, void* stream);
int b200mt_maybe_adam_undo(const b200mt_lists* lists, int chunk_size, const int32_t* overflow_flag, float lr,
                           float beta1, float beta2, float eps, float grad_scale, int step, int mode,
                           int bias_correction, float decay, void* stream);

/* ---- prepared list sets -- no counterpart in the reference, whose harness re-packs the tensor addresses into kernel
 * arguments on every call (csrc/multi_tensor_apply.cuh:84-146: the cost this removes).
 * By default every entry point rebuilds the descriptor-table image of its list set and compares it with the
 * cached copies: O(depth * ntensors) host work per call, which for lists of thousands of small tensors is longer than the
 * kernel.  A host that KNOWS its tensors stay where they are (optimizer state, flat gradient views, bucket buffers)
 * can prepare the list set once and pass the handle in lists->prepared on every call: the library then uses the
 * handle's resident table directly -- no image, no compare.  ptrs / numel / ntensors / depth / dtype / device
 * must still describe the same list set (ntensors, depth and device are checked against the handle; the pointers are
 * NOT re-read: moving a tensor without preparing again is the caller's bug).  One handle serves every op: the table
 * for a kind of op (element-wise, norm-type per chunk size) is uploaded on the handle's first use with it, on that
 * call's stream.  b200mt_lists_release frees the tables (it may wait for the device). */
typedef struct b200mt_prepared b200mt_prepared;
int b200mt_lists_prepare(const b200mt_lists* lists, b200mt_prepared** out);
int b200mt_lists_release(b200mt_prepared* handle);

/* ---- launch trace -- no counterpart in the reference (its profiling hook is apex.pyprof's NVTX wrapper around
 * fused_adam_cuda.adam, apex/pyprof/examples/apex/fused_adam.py:6-7).  Between b200mt_trace_begin(max) and
 * b200mt_trace_end every kernel this library launches is bracketed by two CUDA events on its stream (and launched
 * without programmatic early start, so the pair sees the whole kernel).  trace_end waits for the recorded events and
 * returns the durations in launch order; b200mt_trace_name(i) names launch i (the demangled kernel functor).  Events
 * are created in trace_begin: call it outside stream capture and outside timed regions. */
int b200mt_trace_begin(int max_launches);
int b200mt_trace_end(float* ms, int capacity, int* count);
const char* b200mt_trace_name(int i);

/* ---- b200mt_l2norm_flat_noalloc -- no counterpart in the reference.
 * L2 norm of ONE flat device buffer (x, dtype, numel) with the same two-pass arithmetic and the same
 * per-chunk overflow flag as b200mt_l2norm, for hosts that cannot let this library allocate: a
 * process whose GPU memory has already been handed in full to another runtime's pool (a cudaMalloc
 * from here would fail), or a step loop that must never run into cudaMalloc's device-wide
 * synchronisation.  It calls no cudaMalloc / cudaHostAlloc / cudaFree and creates no events; its
 * working memory is a fixed array inside the library image:
 *   - B200MT_NOALLOC_SCRATCH_BYTES of device memory are held on every GPU that loads the library,
 *     until the process exits;
 *   - inputs that need more than B200MT_NOALLOC_MAX_TILES scheduling tiles are refused with an error.
 *     The limit counts tiles: every chunk of chunk_size elements is cut into ceil(chunk_size / 16384)
 *     equal tiles, so numel elements need floor(numel / chunk_size) * ceil(chunk_size / 16384) tiles
 *     plus those of the last partial chunk.  65536 tiles are 1,073,741,824 elements only when
 *     chunk_size is a multiple of 16384 (the default 65536 is); chunk_size 16400 -> 2 tiles per
 *     chunk -> 537,395,200 elements; chunk_size < 16384 -> one tile per chunk -> 65536 * chunk_size;
 *   - all calls on a device share the array: calls whose GPU work may overlap (unordered streams)
 *     race and must be ordered by the caller; back-to-back calls on one stream are safe, the
 *     function re-zeroes what has to start at zero;
 *   - the first call on each device queries the array's device address (host-side, nothing is
 *     allocated) and caches it: make that call before starting any stream capture on that device;
 *     the cached address does not survive cudaDeviceReset() / a context teardown.
 * ret: device float[1].  *noop is set when a chunk's sum of squares is inf/NaN. */
#define B200MT_NOALLOC_MAX_TILES 65536
#define B200MT_NOALLOC_SCRATCH_BYTES (512 + 4 * (B200MT_NOALLOC_MAX_TILES + 2) + 248)
int b200mt_l2norm_flat_noalloc(const void* x, int dtype, int64_t numel, int chunk_size,
                               int32_t* noop, float* ret, int device, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200MT_H_ */
