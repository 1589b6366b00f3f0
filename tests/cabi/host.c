/* A host written in plain C against include/b200mt.h: proves that the header is C (not C++), that the structs can be
 * initialised the way INTEGRATION.md shows, and that argument checking and error reporting work without a GPU.
 * Built and run by tests/test_cabi_symbols.py::test_plain_c_host_compiles_links_and_gets_errors. */
#include <stdio.h>
#include <string.h>

#include "b200mt.h"

int main(void) {
  const void* ptrs[2] = {0, 0};
  int64_t numel[1] = {16};
  b200mt_lists L = {ptrs, numel, 1, 2, {B200MT_F32, B200MT_F16}, 0, 0};
  b200mt_prepared* handle = 0;
  int failures = 0;

  printf("version %d launches %lld\n", b200mt_version(), (long long)b200mt_launch_count());
  if (b200mt_version() < 101) failures++;

  /* chunk_size <= 0 is refused before anything touches CUDA */
  if (b200mt_scale(&L, 0, 0, 1.0f, 0) == 0 || !strstr(b200mt_last_error(), "chunk_size")) failures++;
  /* an empty list set is refused with the reference's message (csrc/multi_tensor_apply.cuh:53-70) */
  L.ntensors = 0;
  if (b200mt_axpby(&L, 65536, 0, 1.0f, 1.0f, -1, 0) == 0 || !strstr(b200mt_last_error(), "is not > 0")) failures++;
  if (b200mt_lists_prepare(&L, &handle) == 0 || handle != 0) failures++;
  L.ntensors = 1;
  /* preparing a list set needs no GPU either: it records the addresses; releasing a never-used handle frees nothing on a device */
  if (b200mt_lists_prepare(&L, &handle) != 0 || handle == 0) { printf("prepare: %s\n", b200mt_last_error()); failures++; }
  /* the no-allocation entry point validates its size limit on the host */
  if (b200mt_l2norm_flat_noalloc((const void*)ptrs, B200MT_F32, (int64_t)70000, 1, (int32_t*)ptrs, (float*)ptrs, 0, 0) == 0 ||
      !strstr(b200mt_last_error(), "fixed scratch holds")) failures++;
  if (b200mt_trace_end(0, 0, 0) == 0 || !strstr(b200mt_last_error(), "no trace is running")) failures++;
  printf("last error: %s\n", b200mt_last_error());
  printf(failures ? "FAILED %d\n" : "ok\n", failures);
  return failures;
}
