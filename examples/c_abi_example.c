/* Using the CUDA library through its C ABI, without Python or torch ("C ABI").
 *
 *   nvcc -o c_abi_example examples/c_abi_example.c -Iinclude \
 *        -L deepfm-tensorflow-distributed-training-on-sagemaker_b200/ops/_lib -ldeepfm_b200 -lcudart
 *
 * One MLP layer forward: Y[B,400] = relu(X[B,1248] W[400,1248]^T + b) in bf16 with fp32 accumulation on the tcgen05
 * tensor cores, then one flat Adam step over a parameter buffer, both asynchronous on a caller-owned stream. */
#include <cuda_runtime.h>
#include <stdio.h>

#include "deepfm_b200.h"

#define CHECK(call)                                                      \
  do {                                                                   \
    if ((call) != 0) {                                                   \
      fprintf(stderr, "%s failed: %s\n", #call, dfm_last_error());       \
      return 1;                                                          \
    }                                                                    \
  } while (0)

int main(void) {
  const int B = 1024, IN = 1248, OUT = 400;
  const long long n = 4096;
  void *X, *W, *Y, *plan;
  float *bias, *w, *g, *m, *v, *reg;
  unsigned long long* step;
  unsigned int* done;
  cudaStream_t st;
  if (dfm_device_is_sm100() != 1) {
    fprintf(stderr, "needs a Blackwell (sm_100) GPU\n");
    return 1;
  }
  cudaStreamCreate(&st);
  cudaMalloc(&X, (size_t)B * IN * 2);
  cudaMalloc(&W, (size_t)OUT * IN * 2);
  cudaMalloc(&Y, (size_t)B * OUT * 2);
  cudaMalloc((void**)&bias, OUT * sizeof(float));
  cudaMalloc((void**)&w, n * sizeof(float));
  cudaMalloc((void**)&g, n * sizeof(float));
  cudaMalloc((void**)&m, n * sizeof(float));
  cudaMalloc((void**)&v, n * sizeof(float));
  cudaMalloc((void**)&reg, sizeof(float));
  cudaMalloc((void**)&step, sizeof(unsigned long long));
  cudaMalloc((void**)&done, sizeof(unsigned int));
  cudaMemset(X, 0, (size_t)B * IN * 2);
  cudaMemset(W, 0, (size_t)OUT * IN * 2);
  cudaMemset(bias, 0, OUT * sizeof(float));
  cudaMemset(w, 0, n * sizeof(float));
  cudaMemset(g, 0, n * sizeof(float));
  cudaMemset(m, 0, n * sizeof(float));
  cudaMemset(v, 0, n * sizeof(float));
  cudaMemset(reg, 0, sizeof(float));
  cudaMemset(step, 0, sizeof(unsigned long long));
  cudaMemset(done, 0, sizeof(unsigned int));

  /* mode 0 = forward; bias + ReLU in the epilogue, keep probability 1 (no dropout) */
  CHECK(dfm_gemm_plan_create(&plan, 0, X, IN, W, IN, B, OUT, IN, Y, OUT, bias, /*relu*/ 1, /*keep_prob*/ 1.0f,
                             /*seed*/ 0ull, /*layer*/ 0, /*row_offset*/ 0, step, NULL, 0, 1.0f, /*split_k*/ 0));
  CHECK(dfm_gemm_run(plan, st));
  /* optimizer id 0 = Adam; no L2 range, no bf16 shadow range */
  CHECK(dfm_flat_opt_step(w, g, m, v, n, 0, 5e-4f, 0.9f, 0.999f, 1e-8f, 0.95f, 1.0f, 0, 0, 0.0f, 0, 0, NULL, step, reg, done,
                          st));
  if (cudaStreamSynchronize(st) != cudaSuccess) {
    fprintf(stderr, "stream failed: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 1;
  }
  dfm_gemm_plan_destroy(plan);
  printf("forward GEMM and one Adam step ran on the GPU\n");
  return 0;
}
