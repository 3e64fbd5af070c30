// fp32 SIMT building blocks for the agent (exact-arithmetic path: the reference runs fp32, AMP off,
// src/my_agents/ddpg.py:163).  Declarations shared between mlp_fp32.cu and agent.cu.
#pragma once
#include "ois_common.cuh"

namespace ois {

enum Epilogue : int {
  EPI_NONE = 0,       // C = acc
  EPI_BIAS = 1,       // C = acc + bias[n]
  EPI_BIAS_ELU = 2,   // C = elu(acc + bias[n])
  EPI_BIAS_SIGMOID = 3,
  EPI_BIAS_TANH = 4,
  EPI_ELU_GRAD = 5,   // C = acc * elu'(z) with elu'(z) recovered from the saved output H: H > 0 ? 1 : H + 1
};

// C[M,N] = epi(op(A) * op(B)); row-major everywhere.
//   NT: A[M,K], B[N,K]  (forward  X * W^T)
//   NN: A[M,K], B[K,N]  (dgrad    dY * W)
//   TN: A[K,M], B[K,N]  (wgrad    dY^T * X)
int gemm_nt(const float* A, const float* B, float* C, int M, int N, int K, int lda, int ldb, int ldc, int epi,
            const float* bias, const float* H, int ldh, cudaStream_t s);
int gemm_nn(const float* A, const float* B, float* C, int M, int N, int K, int lda, int ldb, int ldc, int epi,
            const float* bias, const float* H, int ldh, cudaStream_t s);
int gemm_tn(const float* A, const float* B, float* C, int M, int N, int K, int lda, int ldb, int ldc, cudaStream_t s);

// column sums: out[n] = sum_m X[m, n]
int col_sum(const float* X, float* out, int M, int N, int ldx, cudaStream_t s);

}  // namespace ois
