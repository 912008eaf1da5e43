/* One dense layer trained for a few steps through the C ABI alone (no Python, no PyTorch): Philox init, the
 * tcgen05 GEMM forward with the fused bias+ReLU epilogue, the weight-gradient GEMM, and the TF-form Adam kernel
 * that also republishes the bf16 shadow the next forward reads.
 *
 *   gcc examples/c_api_dense_step.c -Iinclude -I/usr/local/cuda/include \
 *       -Ltensorflow_distributed_b200/_native -ltfd_b200 -L/usr/local/cuda/lib64 -lcudart -lm -o dense_step
 *   LD_LIBRARY_PATH=tensorflow_distributed_b200/_native ./dense_step          (needs a B200)
 *
 * The layer is y = relu(x W + b) with x [128, 512], W [512, 256].  The objective driving the update is
 * L = sum((x W) * r) for a fixed random r, so dL/dW = x^T r exactly: enough to exercise wgrad and Adam with real
 * data while the forward call shows the fused epilogue. */
#include <cuda_runtime_api.h>
#include <stdio.h>
#include <stdlib.h>

#include "tfd_b200.h"

#define CK(call)                                                                          \
  do {                                                                                    \
    int rc_ = (call);                                                                     \
    if (rc_ != 0) {                                                                       \
      fprintf(stderr, "%s failed (%d): %s\n", #call, rc_, tfd_last_error());              \
      return 1;                                                                           \
    }                                                                                     \
  } while (0)
#define CU(call)                                                                          \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess) {                                                              \
      fprintf(stderr, "%s: %s\n", #call, cudaGetErrorString(e_));                         \
      return 1;                                                                           \
    }                                                                                     \
  } while (0)

enum { B = 128, IN = 512, OUT = 256 };

int main(void) {
  CK(tfd_set_device(0));
  cudaStream_t st;
  CU(cudaStreamCreate(&st));
  const long long nW = (long long)IN * OUT;
  float *x, *r, *w, *m, *v, *g, *bias;
  void *x16, *r16, *w16, *y16;
  unsigned* ctrl; /* [0] Adam step t, [1] block counter, [2] version, [3] global step */
  CU(cudaMalloc((void**)&x, sizeof(float) * B * IN));
  CU(cudaMalloc((void**)&r, sizeof(float) * B * OUT));
  CU(cudaMalloc((void**)&w, sizeof(float) * nW));
  CU(cudaMalloc((void**)&m, sizeof(float) * nW));
  CU(cudaMalloc((void**)&v, sizeof(float) * nW));
  CU(cudaMalloc((void**)&g, sizeof(float) * nW));
  CU(cudaMalloc((void**)&bias, sizeof(float) * OUT));
  CU(cudaMalloc(&x16, 2 * B * IN));
  CU(cudaMalloc(&r16, 2 * B * OUT));
  CU(cudaMalloc(&w16, 2 * nW));
  CU(cudaMalloc(&y16, 2 * B * OUT));
  CU(cudaMalloc((void**)&ctrl, 64));
  CU(cudaMemsetAsync(m, 0, sizeof(float) * nW, st));
  CU(cudaMemsetAsync(v, 0, sizeof(float) * nW, st));
  CU(cudaMemsetAsync(bias, 0, sizeof(float) * OUT, st));
  CU(cudaMemsetAsync(ctrl, 0, 64, st));

  CK(tfd_random_normal(x, (long long)B * IN, 1, 1.0f, 0, st));
  CK(tfd_random_normal(r, (long long)B * OUT, 2, 1.0f, 0, st));
  CK(tfd_random_normal(w, nW, 3, 0.05f, 0, st));
  CK(tfd_cast_bf16(x, IN, x16, IN, B, IN, st));
  CK(tfd_cast_bf16(r, OUT, r16, OUT, B, OUT, st));
  CK(tfd_cast_bf16(w, OUT, w16, OUT, IN, OUT, st)); /* first shadow; Adam republishes it from then on */

  const tfd_piece_t piece = {0, nW, OUT, OUT, 0};
  const float* slots[1];
  slots[0] = g;
  for (int step = 0; step < 5; ++step) {
    /* forward: y = relu(x W + b), W read in place as the MN-major operand ([in, out] layout of the reference) */
    CK(tfd_gemm_bf16(x16, IN, 0, w16, OUT, 1, B, OUT, IN, 128, 1, TFD_EPI_BIAS_RELU_BF16, 1.0f, y16, OUT, bias, NULL, 0,
                     0, NULL, 0, st));
    /* wgrad: dW [IN, OUT] = x^T r, contraction over the batch, both operands read MN-major in place */
    CK(tfd_gemm_bf16(x16, IN, 1, r16, OUT, 1, IN, OUT, B, 128, 1, TFD_EPI_STORE_F32, 1.0f, g, OUT, NULL, NULL, 0, 0, NULL,
                     0, st));
    /* Adam (TF form) over the arena; publishes the bf16 shadow and bumps ctrl[2] (version) and ctrl[3] (step) */
    CK(tfd_adam_apply(w, m, v, w16, slots, 1, 1.0f, nW, &piece, 1, 0.01f, 0.9f, 0.999f, 1e-8f, ctrl, NULL, ctrl + 2,
                      ctrl + 3, st));
  }
  unsigned host_ctrl[4];
  float w0[4];
  CU(cudaMemcpyAsync(host_ctrl, ctrl, sizeof(host_ctrl), cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(w0, w, sizeof(w0), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  printf("adam t=%u version=%u global_step=%u  w[0..3] = %g %g %g %g\n", host_ctrl[0], host_ctrl[2], host_ctrl[3], w0[0],
         w0[1], w0[2], w0[3]);
  return (host_ctrl[0] == 5 && host_ctrl[3] == 5) ? 0 : 2;
}
