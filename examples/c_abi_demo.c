/* A plain C host using libfedrec_b200.so: one bf16 Linear (bias epilogue) on the tcgen05 GEMM and the device-side
 * de-dup, both checked against the CPU. Shows that the library needs nothing from Python or PyTorch.
 *
 *   gcc -std=c99 -Iinclude -I/usr/local/cuda/include examples/c_abi_demo.c -o build/c_abi_demo \
 *       -L fedrec-with-pytorchdistributed_b200/ops -lfedrec_b200 -L/usr/local/cuda/lib64 -lcudart -lm \
 *       -Wl,-rpath,$PWD/fedrec-with-pytorchdistributed_b200/ops -Wl,-rpath,/usr/local/cuda/lib64
 *
 * tests/test_build.py compiles and links it on the CPU box; tests/test_gpu_small_ops.py runs it on a B200.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <cuda_runtime_api.h>

#include "fedrec_b200.h"

static uint16_t f2bf(float f) { /* round to nearest even */
  uint32_t u;
  memcpy(&u, &f, 4);
  u += 0x7FFFu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
static float bf2f(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

#define CU(x)                                                                         \
  do {                                                                                \
    cudaError_t e_ = (x);                                                             \
    if (e_ != cudaSuccess) {                                                          \
      fprintf(stderr, "%s -> %s\n", #x, cudaGetErrorString(e_));                      \
      return 2;                                                                       \
    }                                                                                 \
  } while (0)
#define FR(x)                                                                         \
  do {                                                                                \
    int r_ = (x);                                                                     \
    if (r_ != 0) {                                                                    \
      fprintf(stderr, "%s -> %d: %s\n", #x, r_, fr_last_error());                     \
      return 3;                                                                       \
    }                                                                                 \
  } while (0)

static int demo_gemm(void) {
  const int M = 300, N = 256, K = 128;
  uint16_t* A = (uint16_t*)malloc(sizeof(uint16_t) * M * K);
  uint16_t* W = (uint16_t*)malloc(sizeof(uint16_t) * N * K);
  float* bias = (float*)malloc(sizeof(float) * N);
  uint16_t* out = (uint16_t*)malloc(sizeof(uint16_t) * M * N);
  srand(1);
  for (int i = 0; i < M * K; ++i) A[i] = f2bf((float)(rand() % 17 - 8) / 8.f);
  for (int i = 0; i < N * K; ++i) W[i] = f2bf((float)(rand() % 17 - 8) / 16.f);
  for (int i = 0; i < N; ++i) bias[i] = (float)(rand() % 9 - 4) / 4.f;
  void *dA, *dW, *dO;
  float* dB;
  CU(cudaMalloc(&dA, sizeof(uint16_t) * M * K));
  CU(cudaMalloc(&dW, sizeof(uint16_t) * N * K));
  CU(cudaMalloc(&dO, sizeof(uint16_t) * M * N));
  CU(cudaMalloc((void**)&dB, sizeof(float) * N));
  CU(cudaMemcpy(dA, A, sizeof(uint16_t) * M * K, cudaMemcpyHostToDevice));
  CU(cudaMemcpy(dW, W, sizeof(uint16_t) * N * K, cudaMemcpyHostToDevice));
  CU(cudaMemcpy(dB, bias, sizeof(float) * N, cudaMemcpyHostToDevice));
  /* out = A · W^T + bias ; epi 0 = bias, block_n 0 = choose, no device-side row count, default stream */
  FR(fr_gemm_bf16(dA, dW, dB, NULL, dO, 0, M, N, K, K, K, N, 0, 0, 0, NULL, 1, NULL));
  CU(cudaDeviceSynchronize());
  CU(cudaMemcpy(out, dO, sizeof(uint16_t) * M * N, cudaMemcpyDeviceToHost));
  double max_err = 0.0;
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N; ++n) {
      float acc = bias[n];
      for (int k = 0; k < K; ++k) acc += bf2f(A[m * K + k]) * bf2f(W[n * K + k]);
      const double err = fabs((double)bf2f(out[m * N + n]) - acc) / (1.0 + fabs(acc));
      if (err > max_err) max_err = err;
    }
  printf("c_abi_demo: gemm %dx%dx%d max_rel_err=%.3g\n", M, N, K, max_err);
  cudaFree(dA); cudaFree(dW); cudaFree(dO); cudaFree(dB);
  free(A); free(W); free(bias); free(out);
  return max_err < 1e-2 ? 0 : 4;
}

static int demo_dedup(void) {
  enum { K = 1000, NTAB = 4096 };
  int ids[K], uniq[K], inv[K], count = 0;
  srand(2);
  for (int i = 0; i < K; ++i) ids[i] = rand() % 300 * 7 % NTAB;
  int *d_ids, *d_uniq, *d_inv, *d_count;
  CU(cudaMalloc((void**)&d_ids, sizeof(int) * K));
  CU(cudaMalloc((void**)&d_uniq, sizeof(int) * K));
  CU(cudaMalloc((void**)&d_inv, sizeof(int) * K));
  CU(cudaMalloc((void**)&d_count, sizeof(int)));
  CU(cudaMemcpy(d_ids, ids, sizeof(int) * K, cudaMemcpyHostToDevice));
  /* the variant with scratch built into the library: nothing to allocate for it */
  FR(fr_dedup_static(d_ids, K, NTAB, d_uniq, d_count, d_inv, NULL));
  CU(cudaDeviceSynchronize());
  CU(cudaMemcpy(uniq, d_uniq, sizeof(int) * K, cudaMemcpyDeviceToHost));
  CU(cudaMemcpy(inv, d_inv, sizeof(int) * K, cudaMemcpyDeviceToHost));
  CU(cudaMemcpy(&count, d_count, sizeof(int), cudaMemcpyDeviceToHost));
  int bad = 0;
  for (int i = 0; i < K; ++i) bad += (inv[i] < 0 || inv[i] >= count || uniq[inv[i]] != ids[i]);
  for (int i = 1; i < count; ++i) bad += (uniq[i] <= uniq[i - 1]);
  printf("c_abi_demo: dedup %d ids -> %d unique, %d mismatches\n", K, count, bad);
  cudaFree(d_ids); cudaFree(d_uniq); cudaFree(d_inv); cudaFree(d_count);
  return bad == 0 ? 0 : 5;
}

int main(void) {
  int info[4];
  FR(fr_device_info(info));
  printf("c_abi_demo: library version %d, %d SMs, sm_%d%d\n", fr_version(), info[0], info[1], info[2]);
  int rc = demo_gemm();
  if (rc) return rc;
  rc = demo_dedup();
  if (rc) return rc;
  printf("c_abi_demo: OK\n");
  return 0;
}
