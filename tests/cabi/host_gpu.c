/* A host written in plain C that drives the GPU through include/b200mt.h and the CUDA runtime alone -- no torch, no C++:
 * three tensors of unequal, unaligned-length sizes in cudaMalloc'd memory; unscale with overflow flag (b200mt_scale), global
 * and per-tensor norm (b200mt_l2norm), one AdamW step (b200mt_adam) through a prepared handle; every result is compared with
 * the same arithmetic done here in C.  Built and run by tests/test_cabi_gpu.py. */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <cuda_runtime_api.h>

#include "b200mt.h"

#define NT 3
#define CHECK_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA: %s at line %d\n", cudaGetErrorString(e_), __LINE__); return 100; } } while (0)
#define CHECK_MT(x) do { if ((x) != 0) { printf("b200mt: %s at line %d\n", b200mt_last_error(), __LINE__); return 101; } } while (0)

static float frand(unsigned* s) {   /* xorshift in [-1, 1) */
  *s ^= *s << 13; *s ^= *s >> 17; *s ^= *s << 5;
  return (float)(*s >> 8) / 8388608.0f - 1.0f;
}

int main(void) {
  const int64_t numel[NT] = {5, 4097, 70001};
  const float scale = 1.0f / 1024.0f, lr = 1e-2f, b1 = 0.9f, b2 = 0.999f, eps = 1e-8f, wd = 0.01f;
  const int step = 3;
  float *hg[NT], *hp[NT], *hm[NT], *hv[NT], *dg[NT], *dp[NT], *dm[NT], *dv[NT];
  const void* ptrs[4 * NT];
  unsigned seed = 12345u;
  int failures = 0, t, flag_host = -1;
  int32_t* flag;
  float *norms, norms_host[1 + NT];
  cudaStream_t stream;
  b200mt_prepared* handle = 0;

  CHECK_CUDA(cudaSetDevice(0));
  CHECK_CUDA(cudaStreamCreate(&stream));
  CHECK_CUDA(cudaMalloc((void**)&flag, sizeof(int32_t)));
  CHECK_CUDA(cudaMemset(flag, 0, sizeof(int32_t)));
  CHECK_CUDA(cudaMalloc((void**)&norms, sizeof(float) * (1 + NT)));
  for (t = 0; t < NT; t++) {
    const size_t bytes = sizeof(float) * (size_t)numel[t];
    int64_t i;
    hg[t] = (float*)malloc(bytes); hp[t] = (float*)malloc(bytes); hm[t] = (float*)malloc(bytes); hv[t] = (float*)malloc(bytes);
    for (i = 0; i < numel[t]; i++) {
      hg[t][i] = 1024.0f * frand(&seed); hp[t][i] = frand(&seed); hm[t][i] = 0.1f * frand(&seed); hv[t][i] = 0.01f * fabsf(frand(&seed));
    }
    CHECK_CUDA(cudaMalloc((void**)&dg[t], bytes)); CHECK_CUDA(cudaMalloc((void**)&dp[t], bytes));
    CHECK_CUDA(cudaMalloc((void**)&dm[t], bytes)); CHECK_CUDA(cudaMalloc((void**)&dv[t], bytes));
    CHECK_CUDA(cudaMemcpy(dg[t], hg[t], bytes, cudaMemcpyHostToDevice)); CHECK_CUDA(cudaMemcpy(dp[t], hp[t], bytes, cudaMemcpyHostToDevice));
    CHECK_CUDA(cudaMemcpy(dm[t], hm[t], bytes, cudaMemcpyHostToDevice)); CHECK_CUDA(cudaMemcpy(dv[t], hv[t], bytes, cudaMemcpyHostToDevice));
    ptrs[0 * NT + t] = dg[t]; ptrs[1 * NT + t] = dp[t]; ptrs[2 * NT + t] = dm[t]; ptrs[3 * NT + t] = dv[t];
  }

  { /* 1. unscale in place: lists [g, g] */
    const void* p2[2 * NT];
    b200mt_lists L = {0, 0, NT, 2, {B200MT_F32, B200MT_F32}, 0, 0};
    for (t = 0; t < NT; t++) { p2[t] = dg[t]; p2[NT + t] = dg[t]; }
    L.ptrs = p2; L.numel = numel;
    CHECK_MT(b200mt_scale(&L, 65536, flag, scale, stream));
  }
  { /* 2. global + per-tensor norm of the unscaled gradients */
    b200mt_lists L = {0, 0, NT, 1, {B200MT_F32}, 0, 0};
    L.ptrs = ptrs; L.numel = numel;   /* the first NT pointers are g */
    CHECK_MT(b200mt_l2norm(&L, 65536, flag, norms, norms + 1, 0, stream));
  }
  { /* 3. AdamW step through a prepared handle: lists [g, p, m, v] */
    b200mt_lists L = {0, 0, NT, 4, {B200MT_F32, B200MT_F32, B200MT_F32, B200MT_F32}, 0, 0};
    L.ptrs = ptrs; L.numel = numel;
    CHECK_MT(b200mt_lists_prepare(&L, &handle));
    L.prepared = handle;
    CHECK_MT(b200mt_adam(&L, 65536, flag, lr, b1, b2, eps, step, 1, 1, wd, 1.0f, stream));
  }
  CHECK_CUDA(cudaStreamSynchronize(stream));
  CHECK_CUDA(cudaMemcpy(&flag_host, flag, sizeof(int), cudaMemcpyDeviceToHost));
  CHECK_CUDA(cudaMemcpy(norms_host, norms, sizeof(float) * (1 + NT), cudaMemcpyDeviceToHost));
  if (flag_host != 0) { printf("overflow flag %d on finite input\n", flag_host); failures++; }

  { /* the same arithmetic on the host (csrc/multi_tensor_scale_kernel.cu:66-70, multi_tensor_adam.cu:94-112 mode 1) */
    const double bc1 = 1.0 - pow(b1, step), bc2 = 1.0 - pow(b2, step);
    double total = 0.0;
    for (t = 0; t < NT; t++) {
      const size_t bytes = sizeof(float) * (size_t)numel[t];
      float* got = (float*)malloc(bytes);
      double sq = 0.0, worst;
      int64_t i;
      int k;
      CHECK_CUDA(cudaMemcpy(got, dg[t], bytes, cudaMemcpyDeviceToHost));
      for (i = 0; i < numel[t]; i++) {
        const float u = hg[t][i] * scale;
        if (got[i] != u) { if (failures < 5) printf("scale: tensor %d elem %lld got %g want %g\n", t, (long long)i, got[i], u); failures++; }
        hg[t][i] = u; sq += (double)u * u;
      }
      total += sq;
      if (fabs(norms_host[1 + t] - sqrt(sq)) > 1e-5 * sqrt(sq)) { printf("norm of tensor %d: got %g want %g\n", t, norms_host[1 + t], sqrt(sq)); failures++; }
      for (k = 0; k < 3; k++) {   /* p, m, v */
        float* dev = k == 0 ? dp[t] : k == 1 ? dm[t] : dv[t];
        CHECK_CUDA(cudaMemcpy(got, dev, bytes, cudaMemcpyDeviceToHost));
        worst = 0.0;
        for (i = 0; i < numel[t]; i++) {
          const double g = hg[t][i], m = b1 * (double)hm[t][i] + (1.0 - b1) * g, v = b2 * (double)hv[t][i] + (1.0 - b2) * g * g;
          const double p = (double)hp[t][i] - lr * ((m / bc1) / (sqrt(v / bc2) + eps) + wd * (double)hp[t][i]);
          const double want = k == 0 ? p : k == 1 ? m : v;
          /* p: relative; m = b1*m + (1-b1)*g may cancel, so an ulp of its operands (|g| <= 1) is the measure; v: purely relative */
          const double err = fabs(got[i] - want) / (fabs(want) + (k == 2 ? 0.0 : k == 1 ? 0.1 : 1e-3));
          if (err > worst) worst = err;
        }
        if (worst > 2e-6) { printf("adam: tensor %d list %d worst relative error %g\n", t, k, worst); failures++; }
      }
      free(got);
    }
    if (fabs(norms_host[0] - sqrt(total)) > 1e-5 * sqrt(total)) { printf("global norm: got %g want %g\n", norms_host[0], sqrt(total)); failures++; }
  }

  { /* 4. an inf anywhere raises the flag, output still written */
    const float inf = INFINITY;
    const void* p2[2 * NT];
    b200mt_lists L = {0, 0, NT, 2, {B200MT_F32, B200MT_F32}, 0, 0};
    for (t = 0; t < NT; t++) { p2[t] = dg[t]; p2[NT + t] = dg[t]; }
    L.ptrs = p2; L.numel = numel;
    CHECK_CUDA(cudaMemcpy(dg[2] + 69999, &inf, sizeof(float), cudaMemcpyHostToDevice));
    CHECK_MT(b200mt_scale(&L, 65536, flag, 1.0f, stream));
    CHECK_CUDA(cudaStreamSynchronize(stream));
    CHECK_CUDA(cudaMemcpy(&flag_host, flag, sizeof(int), cudaMemcpyDeviceToHost));
    if (flag_host != 1) { printf("overflow flag %d after an inf\n", flag_host); failures++; }
  }
  CHECK_MT(b200mt_lists_release(handle));
  printf("launches %lld, table uploads %lld\n", (long long)b200mt_launch_count(), (long long)b200mt_table_uploads());
  if (b200mt_launch_count() < 5) failures++;
  b200mt_shutdown();
  printf(failures ? "FAILED %d\n" : "ok\n", failures);
  return failures ? 1 : 0;
}
