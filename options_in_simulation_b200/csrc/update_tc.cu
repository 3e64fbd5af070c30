// bf16 tcgen05 DDPG._update (reference: src/my_agents/ddpg.py:357-444, models
// train/train_kamma_4d_nofifo_random_RLIL.py:21-52): every dense contraction of the update — the forward passes of
// the four networks, the data gradients and the weight gradients — runs on the 5th-generation tensor cores as a
// GROUPED GEMM launch (several independent problems share one grid), fp32 accumulation in tensor memory, fp32 master
// weights / Adam / polyak untouched.  The fp32 SIMT path (agent.cu) stays the exact-arithmetic path.
//
// Operand storage ("tile-blocked" bf16): a matrix [R, C] is kept as 16-byte chunks of 8 consecutive columns,
//     chunk(r, c8)  at  (((r / RB) * C/8 + c8) * RB + r % RB) * 16   bytes          (RB rows per block: 128, or 16)
// i.e. [row block][column chunk][row in block][8 elements].  One stored copy serves BOTH contraction directions:
//   * contraction over the columns (forward, data gradient: rows are the M/N index)  -> UMMA K-major canonical
//     no-swizzle layout, SBO = 128 B (next 8 rows), LBO = RB*16 B (next column chunk);
//   * contraction over the rows (weight gradient dY^T X, and W read as [k_in, n_out] for dY W: columns are the
//     M/N index) -> UMMA MN-major canonical layout, LBO = 128 B (next 8 rows = next K group), SBO = RB*16 B.
// so no transposed copies of activations, gradients or weights are ever written.  In both cases the slice one CTA
// needs for one K stage is CONTIGUOUS in global memory: one bulk async copy (TMA engine, UBLKCP) per operand per stage.
//
// Kernel: 128 x tile_n output tile per CTA (tile_n = 128 or 16), K stages of 128 (or the whole K when K = 16) through
// a 3-deep shared-memory ring on mbarriers; warp 0 = bulk-copy producer, warp 1 = tcgen05.mma issuer (one elected
// lane), warps 2-5 = epilogue (tcgen05.ld -> bias / ELU / ELU' / output activation -> bf16 tile-blocked or fp32).
#include <new>
#include <string.h>
#include "agent.cuh"
#include "tc_ptx.cuh"
#include "update_tc.cuh"

namespace ois {
namespace tcu {

constexpr int kThreads = 320;   // warp 0 producer, warp 1 MMA issuer, warps 2-9 epilogue (two per TMEM lane quadrant)
constexpr int kStages = 3;
constexpr int kHalfStage = 32768;
constexpr int kStageBytes = 2 * kHalfStage;           // A (128 x 128 bf16) + B (<= 128 x 128 bf16)
constexpr int kSmemBytes = kStages * kStageBytes + 128 + 512 + 512;
constexpr int kTmemCols = 128;
constexpr int kMaxProblems = 12;

enum Epi : int {
  EPI_HIDDEN = 0,   // out_bf16 = bf16(elu(acc + bias[n]))
  EPI_OUT = 1,      // out_f32[m, n] = act(acc + bias[n]), n < n_valid
  EPI_DGRAD = 2,    // out_bf16 = bf16(acc * elu'(z)), elu' recovered from the saved bf16 output h: h > 0 ? 1 : h + 1
  EPI_F32 = 3,      // out_f32[m, n] (or [n, m] when transposed) = acc, m < m_valid, n < n_valid
  EPI_BIAS = 4,     // no GEMM: out_f32[n] = sum over the K batch rows of the bf16 matrix `a` [K, a_c8 * 8] (a bias gradient);
                    // one CTA per 64 columns, riding in the weight-gradient launch instead of a launch of its own
};

struct Problem {
  const uint8_t* a;
  const uint8_t* b;
  int a_c8, a_rb, a_mn;     // column chunks per row block, rows per block, 1 = contraction over the rows (MN-major)
  int b_c8, b_rb, b_mn;
  int M, N, K, tile_n;
  int epi, out_act;
  const float* bias;
  const uint8_t* h;
  uint8_t* out_bf16;        // tile-blocked [M, N], RB = 128
  float* out_f32;
  int ld, m_valid, n_valid, transposed;
  uint8_t* x_out;           // EPI_OUT: also bf16 into columns x_col.. of a tile-blocked [M, 16] buffer
  int x_col;
  // EPI_HIDDEN with d_out != null (layer 3 of the policy-loss critic pass): the data gradient of the NEXT layer's single
  // output, d_out[m, n] = bf16(d_coef * d_w[n] * elu'(h[m, n])), is emitted with the activation — exactly what the K = 16
  // GEMM dz W4 would produce from dz = (d_coef, 0, ...), so that launch (and the wait for layer 4) leaves the chain
  uint8_t* d_out;
  const uint8_t* d_w;       // row 0 of the next layer's W in its tile-blocked image ([k chunk][16 rows][8])
  float d_coef;
  int tile_begin;
};

struct Group {
  Problem p[kMaxProblems];
  int count;
};

__host__ __device__ inline size_t chunk_off(int r, int c8, int C8, int rb) { return tcu_chunk_off(r, c8, C8, rb); }

__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); }

// Diagnostics (not part of the public ABI): %globaltimer stamps of CTA 0, read by profiles/tcu_trace.py
__device__ unsigned long long g_tcu_trace[16];
__device__ __forceinline__ void tcu_stamp(int slot) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  g_tcu_trace[slot] = t;
}
#define TCU_STAMP(cond, slot) do { if (cond) tcu_stamp(slot); } while (0)

__global__ void __launch_bounds__(kThreads, 1)
tc_gemm_kernel(const __grid_constant__ Group g) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
  uint64_t* empty = full + kStages;
  uint64_t* d_full = empty + kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(d_full + 1);
  float* s_bias = reinterpret_cast<float*>(smem + kStages * kStageBytes + 128);   // [128]
  float* s_dw = s_bias + 128;                                                       // [128]
  const bool tr = blockIdx.x == 0 && (threadIdx.x & 31) == 0;
  TCU_STAMP(tr && threadIdx.x == 0, 0);

  int pi = 0;
#pragma unroll 1
  for (int i = 1; i < g.count; ++i)
    if ((int)blockIdx.x >= g.p[i].tile_begin) pi = i;
  const Problem P = g.p[pi];   // by value: registers, immune to the memory clobbers of the inline-PTX wrappers
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t = blockIdx.x - P.tile_begin;
  if (P.epi == EPI_BIAS) {
    // ===================================================== bias gradient: warp w of CTA t sums 8-column chunk 8 t + w of `a`
    // over the K batch rows (eight independent 16-byte loads in flight per lane; fixed summation order)
    pdl_wait();
    if (threadIdx.x == 0) pdl_trigger();
    const int w = t * 8 + warp;
    if (warp >= 8 || w >= P.a_c8) return;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int r0 = 0; r0 < P.K; r0 += 256) {
      uint4 q[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int r = r0 + u * 32 + lane;
        q[u] = r < P.K ? *reinterpret_cast<const uint4*>(P.a + chunk_off(r, w, P.a_c8, 128)) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const uint32_t ws[4] = {q[u].x, q[u].y, q[u].z, q[u].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) { acc[2 * j] += bf16_lo(ws[j]); acc[2 * j + 1] += bf16_hi(ws[j]); }
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j)
      for (int o = 16; o > 0; o >>= 1) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], o);
    if (lane < 8 && w * 8 + lane < P.n_valid) {
      float sacc = acc[0];
#pragma unroll
      for (int j = 1; j < 8; ++j) sacc = lane == j ? acc[j] : sacc;
      P.out_f32[w * 8 + lane] = sacc;
    }
    return;
  }
  const int tile_n = P.tile_n;
  const int tiles_n = P.N / tile_n;
  const int tm = t / tiles_n, tn = t - tm * tiles_n;
  const int k_stage = P.K < 128 ? P.K : 128;
  const int nst = P.K / k_stage;
  const uint32_t a_bytes = 128u * k_stage * 2u, b_bytes = (uint32_t)tile_n * k_stage * 2u;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kStages; ++i) { ptx::mbar_init(&full[i], 1); ptx::mbar_init(&empty[i], 1); }
    ptx::mbar_init(d_full, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 1) ptx::tmem_alloc<kTmemCols>(tmem_slot);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  // programmatic dependent launch: the set-up above overlapped the predecessor's tail (ois_common.cuh states the rule)
  TCU_STAMP(tr && threadIdx.x == 0, 1);
  pdl_wait();
  if (threadIdx.x == 0) pdl_trigger();
  TCU_STAMP(tr && threadIdx.x == 0, 2);

  if (warp == 0) {
    // ===================================================== producer: one contiguous bulk copy per operand per K stage
    if (lane == 0) {
      for (int s = 0; s < nst; ++s) {
        const int slot = s % kStages;
        if (s >= kStages) ptx::mbar_wait(&empty[slot], ((s / kStages) - 1) & 1);
        uint8_t* sa = smem + slot * kStageBytes;
        uint8_t* sb = sa + kHalfStage;
        const uint8_t* ga = P.a_mn ? P.a + ((size_t)(s * P.a_c8 + tm * 16) * P.a_rb) * 16
                                   : P.a + ((size_t)(tm * P.a_c8 + s * (k_stage >> 3)) * P.a_rb) * 16;
        ptx::mbar_expect_tx(&full[slot], a_bytes + b_bytes);
        ptx::bulk_g2s(sa, ga, a_bytes, &full[slot]);
        if (P.b_mn || P.b_rb == tile_n) {
          const uint8_t* gb = P.b_mn ? P.b + ((size_t)(s * P.b_c8 + tn * (tile_n >> 3)) * P.b_rb) * 16
                                     : P.b + ((size_t)(tn * P.b_c8 + s * (k_stage >> 3)) * P.b_rb) * 16;
          ptx::bulk_g2s(sb, gb, b_bytes, &full[slot]);
        } else {
          // K-major B whose stored row block (128) is wider than the tile: the tile's rows are a contiguous run inside
          // every column chunk of the block, so the stage is one copy per chunk (tile_n * 16 bytes each)
          const int row0 = tn * tile_n;
          const uint32_t piece = (uint32_t)tile_n * 16u;
          for (int j = 0; j < (k_stage >> 3); ++j)
            ptx::bulk_g2s(sb + j * piece, P.b + chunk_off(row0, s * (k_stage >> 3) + j, P.b_c8, P.b_rb), piece, &full[slot]);
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    const bool leader = ptx::elect_one();
    const uint32_t idesc = ptx::idesc_bf16(128, tile_n) | ((uint32_t)P.a_mn << 15) | ((uint32_t)P.b_mn << 16);
    // K-major: LBO = rows*16 (next column chunk), SBO = 128 (next 8 rows), 2 chunks per MMA
    // MN-major: LBO = 128 (next 8 K rows), SBO = k_stage*16 (next chunk of 8 M/N elements), 16 K rows per MMA
    const uint32_t a_lbo = P.a_mn ? 128u : 2048u, a_sbo = P.a_mn ? (uint32_t)k_stage * 16u : 128u;
    const uint32_t a_step = P.a_mn ? 256u : 4096u;
    const uint32_t b_lbo = P.b_mn ? 128u : (uint32_t)tile_n * 16u, b_sbo = P.b_mn ? (uint32_t)k_stage * 16u : 128u;
    const uint32_t b_step = P.b_mn ? 256u : (uint32_t)tile_n * 32u;
    const uint32_t a_hi = ptx::smem_desc_hi(a_sbo), b_hi = ptx::smem_desc_hi(b_sbo);
    // ONE elected thread runs the whole issue loop: what the issuing thread does between two stages (a warp-wide wait, a
    // re-convergence, rebuilding every descriptor) is not hidden by the tensor pipe's queue (measured in the actor kernel,
    // mlp_tct.cu: 85 -> 47 cycles per MMA).  Descriptors of a stage by one addition from its first: the address field is
    // the low 14 bits of `lo` and the ring lies below 256 KB, so no carry.  (Bulk-copy data -> MMA: both the async proxy.)
    if (leader) {
      for (int s = 0; s < nst; ++s) {
        const int slot = s % kStages;
        ptx::mbar_wait(&full[slot], (s / kStages) & 1);
        TCU_STAMP(blockIdx.x == 0 && s < 4, 3 + s);
        const uint32_t sa = ptx::smem_u32(smem + slot * kStageBytes);
        const uint32_t a_lo = ptx::smem_desc_lo(sa, a_lbo), b_lo = ptx::smem_desc_lo(sa + kHalfStage, b_lbo);
        for (int kk = 0; kk < (k_stage >> 4); ++kk)
          ptx::mma_ss(tmem, ptx::make_desc(a_lo + kk * (a_step >> 4), a_hi), ptx::make_desc(b_lo + kk * (b_step >> 4), b_hi), idesc,
                      (s | kk) != 0);
        ptx::mma_commit(&empty[slot]);
      }
      ptx::mma_commit(d_full);
    }
    __syncwarp();
  } else {
    // ===================================================== epilogue: two warps per TMEM lane quadrant, each takes half of
    // the tile's 16-column chunks (the epilogue is the serial tail of every launch of the update's dependent chain)
    // (parameter fields are copied to registers once: the inline-PTX wrappers carry memory clobbers, which would
    //  otherwise force a re-load of every P.field after each of them)
    const int epi = P.epi, Ncols = P.N, ld = P.ld, m_valid = P.m_valid, n_valid = P.n_valid, transposed = P.transposed;
    const int out_act = P.out_act, x_col = P.x_col;
    const float* const bias = P.bias;
    const uint8_t* const hsrc = P.h;
    uint8_t* const out_bf16 = P.out_bf16;
    uint8_t* const x_out = P.x_out;
    float* const out_f32 = P.out_f32;
    const int quad = warp & 3;
    const int half = (warp - 2) >> 2;       // takes the 16-column chunks c with (c & 1) == half
    const int m = quad * 32 + lane;
    const int gm = tm * 128 + m;
    const int n0 = tn * tile_n;
    const uint32_t t_lane = tmem + ((uint32_t)(quad * 32) << 16);
    // everything the epilogue needs from global memory is fetched while the MMAs run: the bias slice goes to shared
    // memory, the saved activations of this thread's row (ELU') to registers
    const int et = threadIdx.x - 64;
    if (epi == EPI_HIDDEN || epi == EPI_OUT)
      if (et < tile_n) s_bias[et] = (n0 + et < n_valid || epi == EPI_HIDDEN) ? bias[n0 + et] : 0.f;
    uint8_t* const d_out = P.d_out;
    if (d_out && et < tile_n) {   // d_coef * W[0][n] for this tile's columns (both factors are bf16: the product is exact)
      const int n = n0 + et;
      const uint32_t wb = *reinterpret_cast<const uint16_t*>(P.d_w + (n >> 3) * 256 + (n & 7) * 2);
      s_dw[et] = P.d_coef * __uint_as_float(wb << 16);
    }
    uint4 hreg[8];
    if (epi == EPI_DGRAD) {
      const uint4* hp = reinterpret_cast<const uint4*>(hsrc + chunk_off(gm, n0 >> 3, Ncols >> 3, 128));
#pragma unroll
      for (int c = 0; c < 8; ++c) {       // 8-column chunk pairs (2c', 2c'+1) of the 16-column chunks c' = 2 * (c >> 1) + half
        const int k8 = (2 * (c >> 1) + half) * 2 + (c & 1);
        hreg[c] = (k8 * 8 < tile_n) ? hp[k8 * 128] : make_uint4(0, 0, 0, 0);
      }
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");      // the eight epilogue warps only: s_bias is complete
    ptx::mbar_wait(d_full, 0);
    ptx::tc_fence_after();
    TCU_STAMP(tr && warp == 2, 8);
    uint32_t v4[4][16];                     // all of this warp's accumulator columns in flight at once, ONE wait
#pragma unroll
    for (int cc = 0; cc < 4; ++cc)
      if ((2 * cc + half) * 16 < tile_n) ptx::tmem_ld16(t_lane + (2 * cc + half) * 16, v4[cc]);
    ptx::tmem_ld_wait();
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) {
      const int c = 2 * cc + half;
      const int c0 = c * 16;
      if (c0 >= tile_n) break;
      uint32_t (&v)[16] = v4[cc];
      const int n = n0 + c0;
      if (epi == EPI_HIDDEN || epi == EPI_DGRAD) {
        float x[16];
        if (epi == EPI_HIDDEN) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            // ELU: exp(z) - 1 with ONE MUFU.EX2 (ex2.approx.ftz; abs. error ~1e-7, far below the bf16 rounding that follows)
            const float z = __uint_as_float(v[j]) + s_bias[c0 + j];
            float ez;
            asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ez) : "f"(z * 1.4426950408889634f));
            x[j] = z > 0.f ? z : ez - 1.0f;
          }
        } else {
          const uint4 h0 = hreg[2 * cc], h1 = hreg[2 * cc + 1];
          const uint32_t hw[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float ha = bf16_lo(hw[j]), hb = bf16_hi(hw[j]);
            x[2 * j] = __uint_as_float(v[2 * j]) * (ha > 0.f ? 1.f : ha + 1.f);
            x[2 * j + 1] = __uint_as_float(v[2 * j + 1]) * (hb > 0.f ? 1.f : hb + 1.f);
          }
        }
        uint4* op = reinterpret_cast<uint4*>(out_bf16 + chunk_off(gm, n >> 3, Ncols >> 3, 128));
        const uint4 o0 = make_uint4(ptx::pack_bf16(x[0], x[1]), ptx::pack_bf16(x[2], x[3]), ptx::pack_bf16(x[4], x[5]), ptx::pack_bf16(x[6], x[7]));
        const uint4 o1 = make_uint4(ptx::pack_bf16(x[8], x[9]), ptx::pack_bf16(x[10], x[11]), ptx::pack_bf16(x[12], x[13]), ptx::pack_bf16(x[14], x[15]));
        op[0] = o0;
        op[128] = o1;
        if (epi == EPI_HIDDEN && d_out) {   // next layer's data gradient from the bf16 activation just produced (see Problem::d_out)
          const uint32_t hw[8] = {o0.x, o0.y, o0.z, o0.w, o1.x, o1.y, o1.z, o1.w};
          float dd[16];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float ha = bf16_lo(hw[j]), hb = bf16_hi(hw[j]);
            dd[2 * j] = s_dw[c0 + 2 * j] * (ha > 0.f ? 1.f : ha + 1.f);
            dd[2 * j + 1] = s_dw[c0 + 2 * j + 1] * (hb > 0.f ? 1.f : hb + 1.f);
          }
          uint4* dp = reinterpret_cast<uint4*>(d_out + chunk_off(gm, n >> 3, Ncols >> 3, 128));
          dp[0] = make_uint4(ptx::pack_bf16(dd[0], dd[1]), ptx::pack_bf16(dd[2], dd[3]), ptx::pack_bf16(dd[4], dd[5]), ptx::pack_bf16(dd[6], dd[7]));
          dp[128] = make_uint4(ptx::pack_bf16(dd[8], dd[9]), ptx::pack_bf16(dd[10], dd[11]), ptx::pack_bf16(dd[12], dd[13]), ptx::pack_bf16(dd[14], dd[15]));
        }
      } else if (epi == EPI_OUT) {
        if (gm < m_valid) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            if (n + j < n_valid) {
              const float z = __uint_as_float(v[j]) + s_bias[c0 + j];
              const float o = out_act == OIS_ACT_SIGMOID ? 1.0f / (1.0f + expf(-z)) : out_act == OIS_ACT_TANH ? tanhf(z) : z;
              out_f32[(size_t)gm * ld + n + j] = o;
              if (x_out) {
                const int xc = x_col + n + j;
                *reinterpret_cast<uint16_t*>(x_out + chunk_off(gm, xc >> 3, 2, 128) + (xc & 7) * 2) = tcu_bf16_bits(o);
              }
            }
          }
        }
      } else {
        if (gm < m_valid) {
          if (transposed) {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (n + j < n_valid) out_f32[(size_t)(n + j) * ld + gm] = __uint_as_float(v[j]);
          } else if (n + 16 <= n_valid && (ld & 3) == 0) {
            float4* o = reinterpret_cast<float4*>(out_f32 + (size_t)gm * ld + n);
#pragma unroll
            for (int j = 0; j < 4; ++j)
              o[j] = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (n + j < n_valid) out_f32[(size_t)gm * ld + n + j] = __uint_as_float(v[j]);
          }
        }
      }
    }
    TCU_STAMP(tr && warp == 2, 9);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc<kTmemCols>(tmem);
}

// ----------------------------------------------------------------------------------------------------------
// small helper kernels
// ----------------------------------------------------------------------------------------------------------
// fp32 master weights (nn.Sequential order) -> four tile-blocked bf16 matrices [n_out_pad, k_in_pad]
__global__ void __launch_bounds__(256)
pack_weights_kernel(const float* __restrict__ P, uint8_t* __restrict__ img, int in_dim, int out_dim) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  // chunk counts: W1 512 x 2, W2 512 x 64, W3 128 x 64, W4 16 x 16
  const int n1 = 1024, n2 = n1 + 32768, n3 = n2 + 8192, n4 = n3 + 256;
  if (c >= n4) return;
  const int64_t w0 = 0, b0 = (int64_t)512 * in_dim, w1 = b0 + 512, b1 = w1 + 512 * 512, w2 = b1 + 512, b2 = w2 + 128 * 512;
  const int64_t w3 = b2 + 128;
  const float* W;
  int r, c8, C8, rb, N, K;
  size_t base;
  if (c < n1) { W = P + w0; r = c >> 1; c8 = c & 1; C8 = 2; rb = 128; N = 512; K = in_dim; base = kImgW1; }
  else if (c < n2) { const int i = c - n1; W = P + w1; r = i >> 6; c8 = i & 63; C8 = 64; rb = 128; N = 512; K = 512; base = kImgW2; }
  else if (c < n3) { const int i = c - n2; W = P + w2; r = i >> 6; c8 = i & 63; C8 = 64; rb = 128; N = 128; K = 512; base = kImgW3; }
  else { const int i = c - n3; W = P + w3; r = i >> 4; c8 = i & 15; C8 = 16; rb = 16; N = out_dim; K = 128; base = kImgW4; }
  float v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = (r < N && c8 * 8 + j < K) ? W[(int64_t)r * K + c8 * 8 + j] : 0.f;
  *reinterpret_cast<uint4*>(img + base + chunk_off(r, c8, C8, rb)) =
      make_uint4(ptx::pack_bf16(v[0], v[1]), ptx::pack_bf16(v[2], v[3]), ptx::pack_bf16(v[4], v[5]), ptx::pack_bf16(v[6], v[7]));
}

// [B, w1] ++ [B, w2] (fp32 row-major, w1 + w2 <= 16) -> tile-blocked bf16 [B, 16] (zero padded);
// src2 may carry a per-element factor: out_act_grad (d(pre-activation) = d(action) * act'(a)) is folded in here
__global__ void __launch_bounds__(256)
pack_rows_kernel(const float* __restrict__ s1, int ld1, int col1, int w1, const float* __restrict__ s2, int ld2, int w2,
                 const float* __restrict__ act_out, int act, uint8_t* __restrict__ dst, int B,
                 const float* __restrict__ q, float* __restrict__ stats) {
  __shared__ float s_sum[8], s_max[8], s_min[8];
  pdl_wait();
  if (threadIdx.x == 0) pdl_trigger();
  if (q && blockIdx.x == 0) {
    // ddpg.py:423 policy loss = -mean(Q(s, pi(s))) and the Q statistics, in policy_loss_kernel's summation order
    // (agent.cu).  They ride here because nothing on the gradient chain needs them: d(loss)/dq is the constant -1/B
    const int t = threadIdx.x, lane = t & 31, wp = t >> 5;
    float a = 0.f, mx = -INFINITY, mn = INFINITY;
    for (int k = t; k < B; k += 256) {
      a += q[k];
      mx = fmaxf(mx, q[k]);
      mn = fminf(mn, q[k]);
    }
    for (int o = 16; o > 0; o >>= 1) {
      a += __shfl_xor_sync(0xffffffffu, a, o);
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    }
    if (lane == 0) { s_sum[wp] = a; s_max[wp] = mx; s_min[wp] = mn; }
    __syncthreads();
    if (t == 0 && stats) {
      float A = 0.f, MX = -INFINITY, MN = INFINITY;
      for (int w = 0; w < 8; ++w) { A += s_sum[w]; MX = fmaxf(MX, s_max[w]); MN = fminf(MN, s_min[w]); }
      stats[1] = -A / (float)B;
      stats[2] = MX; stats[3] = MN; stats[4] = A / (float)B;
    }
  }
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * 2) return;
  const int r = i >> 1, c8 = i & 1;
  float v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = c8 * 8 + j;
    float x = 0.f;
    if (c < w1) {
      x = s1[(size_t)r * ld1 + col1 + c];
      if (act_out) {   // sigmoid' = a (1 - a), tanh' = 1 - a^2  (ddpg.py:424-429 through the output activation)
        const float a = act_out[(size_t)r * w1 + c];
        x = act == OIS_ACT_SIGMOID ? x * a * (1.f - a) : act == OIS_ACT_TANH ? x * (1.f - a * a) : x;
      }
    } else if (c < w1 + w2) {
      x = s2[(size_t)r * ld2 + (c - w1)];
    }
    v[j] = x;
  }
  *reinterpret_cast<uint4*>(dst + chunk_off(r, c8, 2, 128)) =
      make_uint4(ptx::pack_bf16(v[0], v[1]), ptx::pack_bf16(v[2], v[3]), ptx::pack_bf16(v[4], v[5]), ptx::pack_bf16(v[6], v[7]));
}

// d(policy loss)/dq = -1/B for every sample (ddpg.py:423): the critic's dz input of the policy pass never changes
__global__ void fill_policy_dz_kernel(uint8_t* __restrict__ dz, int B) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  *reinterpret_cast<uint4*>(dz + chunk_off(i, 0, 2, 128)) = make_uint4(ptx::pack_bf16(-1.0f / (float)B, 0.f), 0u, 0u, 0u);
  *reinterpret_cast<uint4*>(dz + chunk_off(i, 1, 2, 128)) = make_uint4(0u, 0u, 0u, 0u);
}

// ----------------------------------------------------------------------------------------------------------
// host side
// ----------------------------------------------------------------------------------------------------------
static int validate(const Problem& p) {
  if (p.epi == EPI_BIAS) {
    OIS_CHECK_ARG(p.a && p.out_f32 && p.a_c8 > 0 && p.K > 0, "bad bias-gradient request");
    return OIS_OK;
  }
  const int k_stage = p.K < 128 ? p.K : 128;
  OIS_CHECK_ARG(p.M % 128 == 0 && p.N % p.tile_n == 0 && p.K % k_stage == 0 && p.K % 16 == 0, "bad GEMM extents");
  OIS_CHECK_ARG(p.tile_n % 16 == 0 && p.tile_n >= 16 && p.tile_n <= 128, "bad tile_n");
  OIS_CHECK_ARG(p.a_mn ? p.a_rb == k_stage : p.a_rb == 128, "A operand row block does not match its use");
  OIS_CHECK_ARG(p.b_mn ? p.b_rb == k_stage : p.b_rb % p.tile_n == 0, "B operand row block does not match its use");
  return OIS_OK;
}

int launch_group(Group& g, cudaStream_t s) {
  int tiles = 0;
  for (int i = 0; i < g.count; ++i) {
    if (int e = validate(g.p[i])) return e;
    g.p[i].tile_begin = tiles;
    tiles += g.p[i].epi == EPI_BIAS ? (g.p[i].a_c8 + 7) / 8 : (g.p[i].M / 128) * (g.p[i].N / g.p[i].tile_n);
  }
  OIS_CHECK_CUDA(launch_pdl(tc_gemm_kernel, tiles, kThreads, kSmemBytes, s, g));
  return OIS_OK;
}

// Column width of the output tiles of the 512- and 128-wide problems.  Narrow tiles cut the B bytes, the MMA work and the
// epilogue of every CTA and multiply the CTAs per launch (B = 512 never fills 148 SMs): measured per update on B200
// 159.7 us (128), 155.6 us (64), 149.3 us (32).  OIS_TCU_TILE_N overrides the default for experiments.
static int wide_tile_n() {
  static const int v = [] {
    const char* e = getenv("OIS_TCU_TILE_N");
    const int t = e ? atoi(e) : 32;
    return (t == 128 || t == 64 || t == 32) ? t : 32;
  }();
  return v;
}

static Problem base_problem() {
  Problem p = {};
  p.tile_n = wide_tile_n();
  p.m_valid = p.n_valid = 1 << 30;
  return p;
}

// layer l of y = mlp(x): A = the layer input (rows = batch, contraction over its columns), B = W_l (rows = n_out)
static Problem fwd_problem(const OisMlpShape& sh, const float* P32, const uint8_t* img, int l, const uint8_t* in,
                           uint8_t* out_bf16, float* y, uint8_t* x_out, int x_col, int B) {
  const MlpOffsets o = mlp_offsets(sh);
  const int kin[4] = {16, 512, 512, 128}, nout[4] = {512, 512, 128, 16};
  const size_t wo[4] = {kImgW1, kImgW2, kImgW3, kImgW4};
  Problem p = base_problem();
  p.a = in; p.a_c8 = kin[l] / 8; p.a_rb = 128; p.a_mn = 0;
  p.b = img + wo[l]; p.b_c8 = kin[l] / 8; p.b_rb = l == 3 ? 16 : 128; p.b_mn = 0;
  p.M = B; p.N = nout[l]; p.K = kin[l]; p.tile_n = l == 3 ? 16 : wide_tile_n();
  p.bias = P32 + o.b[l];
  if (l < 3) { p.epi = EPI_HIDDEN; p.out_bf16 = out_bf16; }
  else { p.epi = EPI_OUT; p.out_f32 = y; p.ld = sh.out_dim; p.n_valid = sh.out_dim; p.m_valid = B; p.out_act = sh.out_act; p.x_out = x_out; p.x_col = x_col; }
  return p;
}

struct NetActs { uint8_t *x, *h0, *h1, *h2; };

// up to two independent networks advance layer by layer in the same launches
// d2_out (single network only): layer 3 also emits the next layer's data gradient for dz = (d_coef, 0, ...) into d2_out
// (Problem::d_out), and layer 4 is NOT launched: its problem is returned in *last for the caller to group with its
// next launch (nothing on the gradient chain waits for it).
static int forward(int count, const OisMlpShape* const* sh, const float* const* P32, const uint8_t* const* img,
                   const NetActs* acts, float* const* y, uint8_t* const* x_out, int x_col, int B, cudaStream_t s,
                   uint8_t* d2_out = nullptr, float d_coef = 0.f, Problem* last = nullptr) {
  for (int l = 0; l < 4; ++l) {
    Group g;
    g.count = count;
    for (int i = 0; i < count; ++i) {
      const uint8_t* in = l == 0 ? acts[i].x : l == 1 ? acts[i].h0 : l == 2 ? acts[i].h1 : acts[i].h2;
      uint8_t* out = l == 0 ? acts[i].h0 : l == 1 ? acts[i].h1 : acts[i].h2;
      g.p[i] = fwd_problem(*sh[i], P32[i], img[i], l, in, out, y[i], x_out ? x_out[i] : nullptr, x_col, B);
    }
    if (l == 2 && d2_out) {
      g.p[0].d_out = d2_out;
      g.p[0].d_w = img[0] + kImgW4;
      g.p[0].d_coef = d_coef;
    }
    if (l == 3 && last) {
      *last = g.p[0];
      return OIS_OK;
    }
    if (int e = launch_group(g, s)) return e;
  }
  return OIS_OK;
}

// backward given dz = d(loss)/d(pre-activation of the last layer) as tile-blocked bf16 [B, 16].
// G != null: parameter gradients (flat fp32).  dx != null: d(loss)/d(input) [B, in_dim] fp32.
// have_d2: d2 = (dz W4) * elu'(h2) was already written by the producer of dz (critic_loss_tc_kernel, actor_dz_kernel,
// or layer 3 of the forward: Problem::d_out), so the chain starts at layer 3.  extra: a problem that is independent of
// this backward (the deferred layer 4 of a forward) and rides in its first launch.
static int backward(const OisMlpShape& sh, const uint8_t* img, float* G, const NetActs& a, const uint8_t* dz, uint8_t* d2,
                    uint8_t* d1, uint8_t* d0, float* dx, int B, cudaStream_t s, bool have_d2 = false, const Problem* extra = nullptr) {
  const MlpOffsets o = mlp_offsets(sh);
  // ---- data gradients: d_l = (d_{l+1} W_{l+1}) * elu'(h_l); W read as [k_in, n_out] from the same stored copy
  const uint8_t* din[3] = {dz, d2, d1};
  uint8_t* dout[3] = {d2, d1, d0};
  const uint8_t* hs[3] = {a.h2, a.h1, a.h0};
  const size_t wo[3] = {kImgW4, kImgW3, kImgW2};
  const int K[3] = {16, 128, 512}, N[3] = {128, 512, 512};
  for (int i = have_d2 ? 1 : 0; i < 3; ++i) {
    Group g;
    g.count = 1;
    if (extra) {
      g.p[1] = *extra;
      g.count = 2;
      extra = nullptr;
    }
    Problem p = base_problem();
    p.a = din[i]; p.a_c8 = K[i] / 8; p.a_rb = 128; p.a_mn = 0;
    p.b = img + wo[i]; p.b_c8 = N[i] / 8; p.b_rb = i == 0 ? 16 : 128; p.b_mn = 1;
    p.M = B; p.N = N[i]; p.K = K[i];
    p.epi = EPI_DGRAD; p.h = hs[i]; p.out_bf16 = dout[i];
    g.p[0] = p;
    if (int e = launch_group(g, s)) return e;
  }
  Group g;
  g.count = 0;
  if (dx) {   // d(loss)/d(input) = d0 W1   (N = in_dim padded to 16)
    Problem p = base_problem();
    p.a = d0; p.a_c8 = 64; p.a_rb = 128; p.a_mn = 0;
    p.b = img + kImgW1; p.b_c8 = 2; p.b_rb = 128; p.b_mn = 1;
    p.M = B; p.N = 16; p.K = 512; p.tile_n = 16;
    p.epi = EPI_F32; p.out_f32 = dx; p.ld = sh.in_dim; p.n_valid = sh.in_dim; p.m_valid = B;
    g.p[g.count++] = p;
  }
  if (G) {
    // ---- weight gradients dW_l = d_l^T h_{l-1}: contraction over the batch rows of both stored matrices
    {   // layer 4, computed transposed (dW4^T = h2^T dz) so that the 128-wide side is M
      Problem p = base_problem();
      p.a = a.h2; p.a_c8 = 16; p.a_rb = 128; p.a_mn = 1;
      p.b = dz; p.b_c8 = 2; p.b_rb = 128; p.b_mn = 1;
      p.M = 128; p.N = 16; p.K = B; p.tile_n = 16;
      p.epi = EPI_F32; p.out_f32 = G + o.w[3]; p.ld = 128; p.transposed = 1; p.m_valid = 128; p.n_valid = sh.out_dim;
      g.p[g.count++] = p;
    }
    const uint8_t* ds[3] = {d2, d1, d0};
    const uint8_t* xs[3] = {a.h1, a.h0, a.x};
    const int Mw[3] = {128, 512, 512}, Nw[3] = {512, 512, 16}, nv[3] = {512, 512, sh.in_dim};
    const int64_t go[3] = {o.w[2], o.w[1], o.w[0]};
    for (int i = 0; i < 3; ++i) {
      Problem p = base_problem();
      p.a = ds[i]; p.a_c8 = Mw[i] / 8; p.a_rb = 128; p.a_mn = 1;
      p.b = xs[i]; p.b_c8 = Nw[i] / 8; p.b_rb = 128; p.b_mn = 1;
      p.M = Mw[i]; p.N = Nw[i]; p.K = B; p.tile_n = Nw[i] == 16 ? 16 : wide_tile_n();
      p.epi = EPI_F32; p.out_f32 = G + go[i]; p.ld = nv[i]; p.m_valid = Mw[i]; p.n_valid = nv[i];
      g.p[g.count++] = p;
    }
  }
  if (G) {   // bias gradients: column sums of the four pre-activation gradients, in the same launch (EPI_BIAS)
    const uint8_t* dd[4] = {d0, d1, d2, dz};
    const int c8[4] = {64, 64, 16, 2}, nv[4] = {512, 512, 128, sh.out_dim};
    for (int l = 0; l < 4; ++l) {
      Problem p = base_problem();
      p.epi = EPI_BIAS; p.a = dd[l]; p.a_c8 = c8[l]; p.K = B; p.out_f32 = G + o.b[l]; p.n_valid = nv[l];
      g.p[g.count++] = p;
    }
  }
  if (g.count)
    if (int e = launch_group(g, s)) return e;
  return OIS_OK;
}

struct State {
  uint8_t* slab = nullptr;
  uint8_t* img[4] = {nullptr, nullptr, nullptr, nullptr};      // TCU_ACTOR, TCU_ACTOR_T, TCU_CRITIC, TCU_CRITIC_T
  const float* packed_from[4] = {nullptr, nullptr, nullptr, nullptr};
  bool valid[4] = {false, false, false, false};
  NetActs acts[4];             // TCU_ACTS_ACTOR, TCU_ACTS_CRITIC (saved for the backward), TCU_ACTS_SCRATCH (targets),
                               // TCU_ACTS_POLICY (own input buffer, hidden buffers shared with TCU_ACTS_CRITIC)
  uint8_t *dz, *d2, *d1, *d0;
  uint8_t* dz_policy;          // constant: -1/B in column 0 (the policy loss's gradient w.r.t. q)
};

}  // namespace tcu

using namespace tcu;

int tcu_init(OisAgent* a) {
  const OisMlpShape* shapes[2] = {&a->actor, &a->critic};
  for (const OisMlpShape* sh : shapes) {
    OIS_CHECK_ARG(sh->h0 == 512 && sh->h1 == 512 && sh->h2 == 128, "the tcgen05 update is built for 512-512-128 hidden layers");
    OIS_CHECK_ARG(sh->in_dim <= 16 && sh->out_dim <= 16, "in_dim and out_dim must be <= 16");
  }
  if (a->batch % 128 != 0) return OIS_OK;   // inference-only use stays possible; ois_ddpg_update_phase reports the reason
  State* st = new (std::nothrow) State();
  OIS_CHECK_ARG(st, "out of host memory");
  const size_t B = (size_t)a->batch;
  const size_t wide = B * 512 * 2, mid = B * 128 * 2, thin = B * 16 * 2;
  const size_t acts = thin + 2 * wide + mid;
  const size_t total = 4 * kImgBytes + 3 * acts + thin + (thin + mid + 2 * wide) + thin;
  cudaError_t e = cudaMalloc(&st->slab, total);
  if (e != cudaSuccess) {
    delete st;
    set_error("tcu_init: cudaMalloc(%zu) -> %s", total, cudaGetErrorString(e));
    return OIS_ERR_CUDA;
  }
  e = cudaFuncSetAttribute(tc_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  if (e != cudaSuccess) {
    cudaFree(st->slab);
    delete st;
    set_error("tcu_init: cudaFuncSetAttribute(%d B smem) -> %s", kSmemBytes, cudaGetErrorString(e));
    return OIS_ERR_CUDA;
  }
  uint8_t* p = st->slab;
  for (int i = 0; i < 4; ++i) { st->img[i] = p; p += kImgBytes; }
  for (int i = 0; i < 3; ++i) {
    NetActs& n = st->acts[i];
    n.x = p; p += thin; n.h0 = p; p += wide; n.h1 = p; p += wide; n.h2 = p; p += mid;
  }
  // the actor reads the first in_dim columns of the CRITIC's input buffer (its image has zero weights beyond them):
  // one buffer, no concatenation kernel.  Its action goes into columns in_dim.. of the POLICY set's input, whose
  // first columns hold the same scaled states; that set reuses the critic's hidden buffers (the two critic passes of
  // an update are sequential: ddpg.py:399-407, then :420-429)
  st->acts[TCU_ACTS_ACTOR].x = st->acts[TCU_ACTS_CRITIC].x;
  st->acts[TCU_ACTS_POLICY] = st->acts[TCU_ACTS_CRITIC];
  st->acts[TCU_ACTS_POLICY].x = p; p += thin;
  st->dz = p; p += thin; st->d2 = p; p += mid; st->d1 = p; p += wide; st->d0 = p; p += wide;
  st->dz_policy = p; p += thin;
  fill_policy_dz_kernel<<<(a->batch + 255) / 256, 256>>>(st->dz_policy, a->batch);
  e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    cudaFree(st->slab);
    delete st;
    set_error("tcu_init: fill_policy_dz_kernel -> %s", cudaGetErrorString(e));
    return OIS_ERR_CUDA;
  }
  a->tcu = st;
  return OIS_OK;
}

void tcu_free(OisAgent* a) {
  if (!a->tcu) return;
  State* st = (State*)a->tcu;
  if (st->slab) cudaFree(st->slab);
  delete st;
  a->tcu = nullptr;
}

void tcu_invalidate(OisAgent* a, int which) {
  if (!a->tcu) return;
  State* st = (State*)a->tcu;
  for (int i = 0; i < 4; ++i)
    if (which < 0 || which == i) st->valid[i] = false;
}

static const OisMlpShape& shape_of(const OisAgent* a, int net) { return net >= TCU_CRITIC ? a->critic : a->actor; }

int tcu_image(OisAgent* a, int net, const float* P32, cudaStream_t s) {
  State* st = (State*)a->tcu;
  OIS_CHECK_ARG(st, "agent was not created with OIS_PRECISION_BF16");
  if (st->valid[net] && st->packed_from[net] == P32) return OIS_OK;
  const OisMlpShape& sh = shape_of(a, net);
  const int chunks = 1024 + 32768 + 8192 + 256;
  pack_weights_kernel<<<(chunks + 255) / 256, 256, 0, s>>>(P32, st->img[net], sh.in_dim, sh.out_dim);
  OIS_CHECK_LAUNCH();
  st->valid[net] = true;
  st->packed_from[net] = P32;
  return OIS_OK;
}

uint8_t* tcu_image_if_valid(OisAgent* a, int net, const float* P32) {
  State* st = (State*)a->tcu;
  return (st && st->valid[net] && st->packed_from[net] == P32) ? st->img[net] : nullptr;
}

uint8_t* tcu_x_buffer(OisAgent* a, int which) {
  State* st = (State*)a->tcu;
  return !st ? nullptr : which == TCU_DZ ? st->dz : st->acts[which].x;
}

int tcu_pack_x(OisAgent* a, int dst, const float* s1, int ld1, int col1, int w1, const float* s2, int ld2, int w2,
               const float* act_out, int act, const float* q, float* stats, cudaStream_t s) {
  State* st = (State*)a->tcu;
  OIS_CHECK_ARG(st && w1 + w2 <= 16, "bad pack request");
  uint8_t* out = dst == TCU_DZ ? st->dz : st->acts[dst].x;
  OIS_CHECK_CUDA(launch_pdl(pack_rows_kernel, (a->batch * 2 + 255) / 256, 256, 0, s, s1, ld1, col1, w1, s2, ld2, w2, act_out, act, out, a->batch, q, stats));
  return OIS_OK;
}

int tcu_forward(OisAgent* a, int count, const int* nets, const float* const* P32, const int* acts, float* const* y,
                uint8_t* const* x_out, int x_col, cudaStream_t s) {
  State* st = (State*)a->tcu;
  OIS_CHECK_ARG(st && count >= 1 && count <= 2, "bad forward request");
  const OisMlpShape* sh[2];
  const uint8_t* img[2];
  NetActs na[2];
  for (int i = 0; i < count; ++i) {
    if (int e = tcu_image(a, nets[i], P32[i], s)) return e;
    sh[i] = &shape_of(a, nets[i]);
    img[i] = st->img[nets[i]];
    na[i] = st->acts[acts[i]];
  }
  return forward(count, sh, P32, img, na, y, x_out, x_col, a->batch, s);
}

int tcu_backward(OisAgent* a, int net, float* G, int acts, float* dx, bool policy_dz, bool have_d2, cudaStream_t s) {
  State* st = (State*)a->tcu;
  OIS_CHECK_ARG(st && st->valid[net], "backward needs the weight image its forward used");
  return backward(shape_of(a, net), st->img[net], G, st->acts[acts], policy_dz ? st->dz_policy : st->dz, st->d2, st->d1,
                  st->d0, dx, a->batch, s, have_d2);
}

// ddpg.py:420-429 through the critic in ONE chain of launches: q = critic(x) for activation set `acts` with layer 3 also
// emitting d2 = d(-mean q)/d(pre-activation 3) (the loss gradient w.r.t. q is the constant -1/B, so d2 needs no q), then
// the data gradients down to dx [B, in_dim]; layer 4 of the forward (q itself, read only by the loss statistics) rides
// in the first data-gradient launch.  6 launches instead of 4 + 4.
int tcu_policy_pass(OisAgent* a, int net, const float* P32, int acts, float* q, float* dx, cudaStream_t s) {
  State* st = (State*)a->tcu;
  OIS_CHECK_ARG(st, "agent was not created with OIS_PRECISION_BF16");
  if (int e = tcu_image(a, net, P32, s)) return e;
  const OisMlpShape* sh[1] = {&shape_of(a, net)};
  const uint8_t* img[1] = {st->img[net]};
  const float* P[1] = {P32};
  NetActs na[1] = {st->acts[acts]};
  float* y[1] = {q};
  Problem last;
  // the coefficient the K = 16 GEMM would have read from dz_policy: bf16(-1 / B)
  uint16_t cb;
  {
    const float c = -1.0f / (float)a->batch;
    uint32_t u;
    memcpy(&u, &c, 4);
    u += 0x7FFFu + ((u >> 16) & 1u);      // round to nearest even
    cb = (uint16_t)(u >> 16);
  }
  const uint32_t cu = (uint32_t)cb << 16;
  float coef;
  memcpy(&coef, &cu, 4);
  if (int e = forward(1, sh, P, img, na, y, nullptr, 0, a->batch, s, st->d2, coef, &last)) return e;
  return backward(*sh[0], st->img[net], nullptr, st->acts[acts], st->dz_policy, st->d2, st->d1, st->d0, dx, a->batch, s, true, &last);
}

// buffers the fused loss / output-gradient kernels of agent.cu write: the dz and d2 inputs of a backward, the saved
// layer-3 activations of activation set `acts`, and row block 0 of W4 inside the image of `net`
uint8_t* tcu_d2_buffer(OisAgent* a) { return a->tcu ? ((State*)a->tcu)->d2 : nullptr; }
const uint8_t* tcu_h2_buffer(OisAgent* a, int acts) { return a->tcu ? ((State*)a->tcu)->acts[acts].h2 : nullptr; }
const uint8_t* tcu_w4_image(OisAgent* a, int net) { return a->tcu ? ((State*)a->tcu)->img[net] + kImgW4 : nullptr; }

// diagnostics hook for tests (deliberately not declared in include/ois_b200.h): C = op(A) op(B) on the tensor cores
// from tile-blocked bf16 operands the caller prepared; mode bits: 1 = A contracts over its rows, 2 = B likewise
int tcu_debug_gemm(const void* a, int a_c8, int a_rb, const void* b, int b_c8, int b_rb, int M, int N, int K, int tile_n,
                   int mode, float* out, int ld, int transposed, cudaStream_t s, int epi = EPI_F32, const float* bias = nullptr,
                   const void* h = nullptr, void* out_bf16 = nullptr) {
  static bool attr = false;
  if (!attr) {
    OIS_CHECK_CUDA(cudaFuncSetAttribute(tc_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    attr = true;
  }
  Group g;
  g.count = 1;
  Problem p = base_problem();
  p.a = (const uint8_t*)a; p.a_c8 = a_c8; p.a_rb = a_rb; p.a_mn = mode & 1;
  p.b = (const uint8_t*)b; p.b_c8 = b_c8; p.b_rb = b_rb; p.b_mn = (mode >> 1) & 1;
  p.M = M; p.N = N; p.K = K; p.tile_n = tile_n;
  p.epi = epi; p.out_f32 = out; p.ld = ld; p.transposed = transposed; p.m_valid = M; p.n_valid = N;
  p.bias = bias; p.h = (const uint8_t*)h; p.out_bf16 = (uint8_t*)out_bf16;
  g.p[0] = p;
  return launch_group(g, s);
}

}  // namespace ois

extern "C" int ois_debug_tcu_trace(unsigned long long* out_host, int n) {
  if (n > 16) n = 16;
  return cudaMemcpyFromSymbol(out_host, ois::tcu::g_tcu_trace, sizeof(unsigned long long) * n) == cudaSuccess ? 0 : -2;
}

extern "C" int ois_debug_tc_gemm_epi(const void* a, int a_c8, int a_rb, const void* b, int b_c8, int b_rb, int M, int N, int K,
                                     int tile_n, int mode, int epi, const float* bias, const void* h, void* out_bf16,
                                     void* stream) {
  return ois::tcu_debug_gemm(a, a_c8, a_rb, b, b_c8, b_rb, M, N, K, tile_n, mode, nullptr, N, 0, (cudaStream_t)stream, epi, bias, h,
                             out_bf16);
}

extern "C" int ois_debug_tc_gemm(const void* a, int a_c8, int a_rb, const void* b, int b_c8, int b_rb, int M, int N, int K,
                                 int tile_n, int mode, float* out, int ld, int transposed, void* stream) {
  return ois::tcu_debug_gemm(a, a_c8, a_rb, b, b_c8, b_rb, M, N, K, tile_n, mode, out, ld, transposed, (cudaStream_t)stream);
}
