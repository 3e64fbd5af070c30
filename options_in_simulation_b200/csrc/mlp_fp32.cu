// fp32 SIMT GEMM with fused epilogues + small reductions: the exact-arithmetic path of the agent
// (DDPG._update at B = 512 is launch/latency bound, not tensor bound: 3.4 GFLOP per update; the bf16
// tcgen05 path in mlp_tc.cu serves the large-M actor inference).
#include "mlp_fp32.cuh"

namespace ois {

constexpr int BM = 64, BN = 64, BK = 16, GT = 256;

__device__ __forceinline__ float elu1(float x) { return x > 0.f ? x : expm1f(x); }

template <bool TA, bool TB_KN>
__global__ void __launch_bounds__(GT)
gemm_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ C, int M, int N, int K,
            int lda, int ldb, int ldc, int epi, const float* __restrict__ bias, const float* __restrict__ H, int ldh) {
  __shared__ __align__(16) float As[BK][BM + 4];
  __shared__ __align__(16) float Bs[BK][BN + 4];
  const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < K; k0 += BK) {
    // ---- stage A tile -> As[k][m]
    if (!TA) {   // A[m, k], k contiguous
      const int m = t >> 2, kq = (t & 3) * 4;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int gm = m0 + m, gk = k0 + kq + q;
        As[kq + q][m] = (gm < M && gk < K) ? A[(size_t)gm * lda + gk] : 0.f;
      }
    } else {     // A[k, m], m contiguous
      const int k = t >> 4, mq = (t & 15) * 4;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int gm = m0 + mq + q, gk = k0 + k;
        As[k][mq + q] = (gm < M && gk < K) ? A[(size_t)gk * lda + gm] : 0.f;
      }
    }
    // ---- stage B tile -> Bs[k][n]
    if (!TB_KN) {  // B[n, k], k contiguous
      const int n = t >> 2, kq = (t & 3) * 4;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int gn = n0 + n, gk = k0 + kq + q;
        Bs[kq + q][n] = (gn < N && gk < K) ? B[(size_t)gn * ldb + gk] : 0.f;
      }
    } else {       // B[k, n], n contiguous
      const int k = t >> 4, nq = (t & 15) * 4;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int gn = n0 + nq + q, gk = k0 + k;
        Bs[k][nq + q] = (gn < N && gk < K) ? B[(size_t)gk * ldb + gn] : 0.f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gm = m0 + ty * 4 + i;
    if (gm >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gn = n0 + tx * 4 + j;
      if (gn >= N) continue;
      float v = acc[i][j];
      switch (epi) {
        case EPI_BIAS: v += bias[gn]; break;
        case EPI_BIAS_ELU: v = elu1(v + bias[gn]); break;
        case EPI_BIAS_SIGMOID: v = 1.f / (1.f + expf(-(v + bias[gn]))); break;
        case EPI_BIAS_TANH: v = tanhf(v + bias[gn]); break;
        case EPI_ELU_GRAD: { const float h = H[(size_t)gm * ldh + gn]; v *= (h > 0.f ? 1.f : h + 1.f); } break;
        default: break;
      }
      C[(size_t)gm * ldc + gn] = v;
    }
  }
}

template <bool TA, bool TB_KN>
static int launch_gemm(const float* A, const float* B, float* C, int M, int N, int K, int lda, int ldb, int ldc, int epi,
                       const float* bias, const float* H, int ldh, cudaStream_t s) {
  if (M <= 0 || N <= 0) return OIS_OK;
  dim3 grid((N + BN - 1) / BN, (M + BM - 1) / BM);
  gemm_kernel<TA, TB_KN><<<grid, GT, 0, s>>>(A, B, C, M, N, K, lda, ldb, ldc, epi, bias, H, ldh);
  OIS_CHECK_LAUNCH();
  return OIS_OK;
}

int gemm_nt(const float* A, const float* B, float* C, int M, int N, int K, int lda, int ldb, int ldc, int epi,
            const float* bias, const float* H, int ldh, cudaStream_t s) {
  return launch_gemm<false, false>(A, B, C, M, N, K, lda, ldb, ldc, epi, bias, H, ldh, s);
}
int gemm_nn(const float* A, const float* B, float* C, int M, int N, int K, int lda, int ldb, int ldc, int epi,
            const float* bias, const float* H, int ldh, cudaStream_t s) {
  return launch_gemm<false, true>(A, B, C, M, N, K, lda, ldb, ldc, epi, bias, H, ldh, s);
}
int gemm_tn(const float* A, const float* B, float* C, int M, int N, int K, int lda, int ldb, int ldc, cudaStream_t s) {
  return launch_gemm<true, true>(A, B, C, M, N, K, lda, ldb, ldc, EPI_NONE, nullptr, nullptr, 0, s);
}

__global__ void col_sum_kernel(const float* __restrict__ X, float* __restrict__ out, int M, int N, int ldx) {
  // one block per 32 columns; 8 row-groups of 32 lanes, fixed summation order (deterministic)
  __shared__ float part[8][33];
  const int col = blockIdx.x * 32 + (threadIdx.x & 31), rg = threadIdx.x >> 5;
  float acc = 0.f;
  if (col < N)
    for (int m = rg; m < M; m += 8) acc += X[(size_t)m * ldx + col];
  part[rg][threadIdx.x & 31] = acc;
  __syncthreads();
  if (rg == 0 && col < N) {
    float v = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) v += part[r][threadIdx.x];
    out[col] = v;
  }
}

int col_sum(const float* X, float* out, int M, int N, int ldx, cudaStream_t s) {
  if (N <= 0) return OIS_OK;
  col_sum_kernel<<<(N + 31) / 32, 256, 0, s>>>(X, out, M, N, ldx);
  OIS_CHECK_LAUNCH();
  return OIS_OK;
}

}  // namespace ois
