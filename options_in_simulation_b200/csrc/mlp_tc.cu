// bf16 tcgen05 + TMA(bulk) fused actor: scaler -> 6->512->512->128->A MLP (ELU, Sigmoid/Tanh) -> OU noise -> clamp
// in ONE persistent, warp-specialised kernel (reference: src/my_agents/ddpg.py:215-268 with the
// model of train/train_kamma_4d_nofifo_random_RLIL.py:21-36).
//
// One CTA owns a 128-row tile of envs at a time; all four layers run back to back with the activations never
// leaving the SM:
//   * weights: packed once per parameter change into a bf16 "stream image" (canonical UMMA no-swizzle K-major
//     core-matrix layout, in exactly the order the MMA warp consumes it: 42 stages of 16 KB), streamed from L2
//     through a 3-deep shared-memory ring by 1-D bulk async copies (UBLKCP) completing on mbarriers;
//   * MMA: one elected thread issues tcgen05.mma.cta_group::1.kind::f16 (M=128, N=128/16, K=16), A = activations
//     in shared memory, B = weight stage, D = fp32 accumulators in TMEM (2 x 128 ping-pong columns + L3 + L4);
//   * epilogue: 8 warps (two per TMEM lane quadrant) tcgen05.ld the accumulators, add bias, ELU, convert to
//     bf16 and write the next layer's A operand straight into shared memory; layer-3 partial products are
//     issued as soon as each 128-column slice of layer 2 has been activated, so act2 only ever exists as one
//     32 KB slice;
//   * the last epilogue applies Sigmoid/Tanh, the OU exploration noise (explore.cuh) and the clamp.
// fp32 master weights are untouched; accumulation is fp32; tolerance vs the fp32 oracle: tests/test_agent_gpu.py.
#include "agent.cuh"
#include "explore.cuh"
#include "tc_ptx.cuh"

namespace ois {

namespace tc {
constexpr int kTileM = 128;
constexpr int kThreads = 320;             // warp 0: bulk-copy producer, warp 1: MMA issuer, warps 2-9: epilogue
constexpr int kEpiWarps = 8;
constexpr int kH0 = 512, kH1 = 512, kH2 = 128, kInPad = 16, kOutPad = 16;
constexpr int kStageBytes = 16384;
constexpr int kStages = 3;
constexpr int kStreamStages = 42;         // 1 (W1) + 32 (W2) + 8 (W3) + 1 (W4)
constexpr int kLastStageBytes = kOutPad * kH2 * 2;                       // 4096
constexpr size_t kImageBytes = (size_t)(kStreamStages - 1) * kStageBytes + kLastStageBytes;
// shared memory map (bytes)
constexpr int kOffAct1 = 0;                                // [64 kc][128 rows][8] bf16 = 128 KB
constexpr int kOffRing = kOffAct1 + kTileM * kH0 * 2;      // act2 slice / act3: [16 kc][128][8] = 32 KB
constexpr int kOffX = kOffRing + kTileM * 128 * 2;         // [2 kc][128][8] = 4 KB
constexpr int kOffW = kOffX + kTileM * kInPad * 2;         // 3 x 16 KB
constexpr int kOffBias = kOffW + kStages * kStageBytes;    // b1 512, b2 512, b3 128, b4 16 fp32
constexpr int kOffBar = kOffBias + (kH0 + kH1 + kH2 + kOutPad) * 4;
constexpr int kSmemBytes = kOffBar + 256;
// TMEM columns
constexpr int kColD = 0;        // two ping-pong 128-column accumulators (layers 1 and 2)
constexpr int kColD3 = 256;     // layer 3, 128 columns
constexpr int kColD4 = 384;     // layer 4, 16 columns
constexpr int kTmemCols = 512;

// stream order: W1 | W2 q0 (8 k-blocks) | W2 q1 | W3 p0 (2 halves) | W2 q2 | W3 p1 | W2 q3 | W3 p2 | W3 p3 | W4
__host__ __device__ inline void stage_info(int s, int& layer, int& piece, int& kb) {
  if (s == 0) { layer = 1; piece = 0; kb = 0; }
  else if (s < 17) { layer = 2; piece = (s - 1) >> 3; kb = (s - 1) & 7; }
  else if (s < 19) { layer = 3; piece = 0; kb = s - 17; }
  else if (s < 27) { layer = 2; piece = 2; kb = s - 19; }
  else if (s < 29) { layer = 3; piece = 1; kb = s - 27; }
  else if (s < 37) { layer = 2; piece = 3; kb = s - 29; }
  else if (s < 39) { layer = 3; piece = 2; kb = s - 37; }
  else if (s < 41) { layer = 3; piece = 3; kb = s - 39; }
  else { layer = 4; piece = 0; kb = 0; }
}
}  // namespace tc

struct TcState {
  void* image = nullptr;      // bf16 stream image of the actor weights
  const float* packed_from = nullptr;
  bool valid = false;
  int sms = 148;
};

// ----------------------------------------------------------------------------------------------------------
// fp32 master weights (nn.Sequential order, include/ois_b200.h OisMlpShape) -> bf16 stream image.
// One thread per 16-byte chunk (8 consecutive k of one output row n).
// ----------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
tc_pack_kernel(const float* __restrict__ P, uint4* __restrict__ image, int in_dim, int out_dim) {
  using namespace tc;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int total = (int)(kImageBytes / 16);
  if (c >= total) return;
  const int s = c >> 10, cc = c & 1023;
  int layer, piece, kb;
  stage_info(s, layer, piece, kb);
  // parameter offsets
  const int64_t w0 = 0, b0 = (int64_t)kH0 * in_dim;
  const int64_t w1 = b0 + kH0, b1 = w1 + (int64_t)kH1 * kH0;
  const int64_t w2 = b1 + kH1, b2 = w2 + (int64_t)kH2 * kH1;
  const int64_t w3 = b2 + kH2;
  const float* W;
  int n, k0, N, K;
  if (layer == 1) { W = P + w0; n = cc & 511; k0 = (cc >> 9) * 8; N = kH0; K = in_dim; }
  else if (layer == 2) { W = P + w1; n = piece * 128 + (cc & 127); k0 = kb * 64 + (cc >> 7) * 8; N = kH1; K = kH0; }
  else if (layer == 3) { W = P + w2; n = cc & 127; k0 = piece * 128 + kb * 64 + (cc >> 7) * 8; N = kH2; K = kH1; }
  else { W = P + w3; n = cc & 15; k0 = (cc >> 4) * 8; N = out_dim; K = kH2; }
  float v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = (n < N && k0 + j < K) ? W[(int64_t)n * K + k0 + j] : 0.f;
  image[c] = make_uint4(ptx::pack_bf16(v[0], v[1]), ptx::pack_bf16(v[2], v[3]), ptx::pack_bf16(v[4], v[5]),
                        ptx::pack_bf16(v[6], v[7]));
}


// ----------------------------------------------------------------------------------------------------------
// epilogue helpers (8 warps; warp quadrant = warp_id % 4 is the TMEM lane group the hardware lets it read)
// ----------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float elu_fast(float x) {   // bf16 output: ex2.approx is far inside the tolerance
  return x > 0.f ? x : __expf(x) - 1.0f;
}

// 32 accumulator columns [col0, col0+32) of this thread's row -> +bias -> ELU -> bf16 -> 4 x 16-byte chunks of
// the next layer's A operand ([kc][row][8] layout, kc = k / 8).
__device__ __forceinline__ void epi_chunk_to_smem(uint32_t taddr, const float* __restrict__ bias, uint8_t* dst_kc0,
                                                  int row) {
  uint32_t v[32];
  ptx::tmem_ld32(taddr, v);
  ptx::tmem_ld_wait();
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint32_t w[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int c = j * 8 + e * 2;
      const float a = elu_fast(__uint_as_float(v[c]) + bias[c]);
      const float b = elu_fast(__uint_as_float(v[c + 1]) + bias[c + 1]);
      w[e] = ptx::pack_bf16(a, b);
    }
    *reinterpret_cast<uint4*>(dst_kc0 + j * 2048 + row * 16) = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

struct TcActArgs {
  const float* params;        // fp32 master (biases are read from it)
  const uint8_t* image;       // bf16 stream image
  const double* mean;         // scaler (nullable)
  const double* var;
  int in_dim, out_dim, out_act;
  int64_t rows;
  OisActIO io;
};

__global__ void __launch_bounds__(tc::kThreads, 1)
tc_actor_kernel(const __grid_constant__ TcActArgs args) {
  using namespace tc;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* s_act1 = smem + kOffAct1;
  uint8_t* s_ring = smem + kOffRing;
  uint8_t* s_x = smem + kOffX;
  uint8_t* s_w = smem + kOffW;
  float* s_bias = reinterpret_cast<float*>(smem + kOffBias);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* w_full = bars;            // [3]
  uint64_t* w_empty = bars + 3;       // [3]
  uint64_t* x_ready = bars + 6;
  uint64_t* d_full = bars + 7;        // [2]
  uint64_t* d_empty = bars + 9;       // [2]
  uint64_t* act1_ready = bars + 11;   // [4]
  uint64_t* act2_ready = bars + 15;
  uint64_t* act2_free = bars + 16;
  uint64_t* d3_full = bars + 17;
  uint64_t* act3_ready = bars + 18;
  uint64_t* d4_full = bars + 19;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 20);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_tiles = (int)((args.rows + kTileM - 1) / kTileM);

  // ---- one-time setup
  for (int i = threadIdx.x; i < kH0 + kH1 + kH2 + kOutPad; i += kThreads) {
    const int64_t b0 = (int64_t)kH0 * args.in_dim, w1 = b0 + kH0, b1 = w1 + (int64_t)kH1 * kH0, w2 = b1 + kH1;
    const int64_t b2 = w2 + (int64_t)kH2 * kH1, w3 = b2 + kH2, b3 = w3 + (int64_t)args.out_dim * kH2;
    float v;
    if (i < kH0) v = args.params[b0 + i];
    else if (i < kH0 + kH1) v = args.params[b1 + i - kH0];
    else if (i < kH0 + kH1 + kH2) v = args.params[b2 + i - kH0 - kH1];
    else v = (i - kH0 - kH1 - kH2) < args.out_dim ? args.params[b3 + i - kH0 - kH1 - kH2] : 0.f;
    s_bias[i] = v;
  }
  if (threadIdx.x == 0) {
    for (int i = 0; i < kStages; ++i) { ptx::mbar_init(&w_full[i], 1); ptx::mbar_init(&w_empty[i], 1); }
    ptx::mbar_init(x_ready, kEpiWarps);
    for (int i = 0; i < 2; ++i) { ptx::mbar_init(&d_full[i], 1); ptx::mbar_init(&d_empty[i], kEpiWarps); }
    for (int i = 0; i < 4; ++i) ptx::mbar_init(&act1_ready[i], kEpiWarps);
    ptx::mbar_init(act2_ready, kEpiWarps);
    ptx::mbar_init(act2_free, 1);
    ptx::mbar_init(d3_full, 1);
    ptx::mbar_init(act3_ready, kEpiWarps);
    ptx::mbar_init(d4_full, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 1) ptx::tmem_alloc<kTmemCols>(tmem_slot);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ===================================================== producer: stream the weight image, tile after tile
    if (lane == 0) {
      uint32_t n = 0;   // stages issued so far
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        for (int s = 0; s < kStreamStages; ++s, ++n) {
          const int slot = n % kStages;
          ptx::mbar_wait(&w_empty[slot], ((n / kStages) & 1) ^ 1);
          const uint32_t bytes = s == kStreamStages - 1 ? kLastStageBytes : kStageBytes;
          ptx::mbar_expect_tx(&w_full[slot], bytes);
          ptx::bulk_g2s(s_w + slot * kStageBytes, args.image + (size_t)s * kStageBytes, bytes, &w_full[slot]);
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer (one thread)
    if (lane == 0) {
      const uint32_t id128 = ptx::idesc_bf16(128, 128), id16 = ptx::idesc_bf16(128, 16);
      const uint32_t a1 = ptx::smem_u32(s_act1), ring = ptx::smem_u32(s_ring), xs = ptx::smem_u32(s_x), ws = ptx::smem_u32(s_w);
      uint32_t n = 0;            // weight stages consumed
      uint32_t d_uses[2] = {0, 0};
      uint32_t tiles_done = 0, a2_cnt = 0;
      auto wait_stage = [&]() -> uint32_t {
        const int slot = n % kStages;
        ptx::mbar_wait(&w_full[slot], (n / kStages) & 1);
        ptx::tc_fence_after();
        return ws + slot * kStageBytes;
      };
      auto release_stage = [&]() { ptx::mma_commit(&w_empty[n % kStages]); ++n; };
      auto acquire_d = [&](int b) {
        ptx::mbar_wait(&d_empty[b], (d_uses[b] & 1) ^ 1);
        ptx::tc_fence_after();
        ++d_uses[b];
      };
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tiles_done) {
        const uint32_t tpar = tiles_done & 1;
        // ---- layer 1: X[128,16] x W1^T, four 128-column pieces
        ptx::mbar_wait(x_ready, tpar);
        ptx::tc_fence_after();
        {
          const uint32_t w = wait_stage();
          for (int q = 0; q < 4; ++q) {
            acquire_d(q & 1);
            ptx::mma_ss(tmem + kColD + (q & 1) * 128, ptx::smem_desc(xs, 2048, 128),
                        ptx::smem_desc(w + q * 2048, 8192, 128), id128, false);
            ptx::mma_commit(&d_full[q & 1]);
          }
          release_stage();
        }
        auto l2_piece = [&](int q) {
          acquire_d(q & 1);
          const uint32_t d = tmem + kColD + (q & 1) * 128;
          for (int kb = 0; kb < 8; ++kb) {
            const uint32_t w = wait_stage();
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)
              ptx::mma_ss(d, ptx::smem_desc(a1 + (kb * 4 + ks) * 4096, 2048, 128), ptx::smem_desc(w + ks * 4096, 2048, 128),
                          id128, (kb | ks) != 0);
            release_stage();
          }
          ptx::mma_commit(&d_full[q & 1]);
        };
        auto l3_partial = [&](int p) {
          ptx::mbar_wait(act2_ready, a2_cnt & 1);
          ++a2_cnt;
          ptx::tc_fence_after();
          for (int h = 0; h < 2; ++h) {
            const uint32_t w = wait_stage();
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)
              ptx::mma_ss(tmem + kColD3, ptx::smem_desc(ring + (h * 4 + ks) * 4096, 2048, 128),
                          ptx::smem_desc(w + ks * 4096, 2048, 128), id128, (p | h | ks) != 0);
            release_stage();
          }
          ptx::mma_commit(act2_free);
          if (p == 3) ptx::mma_commit(d3_full);
        };
        // ---- layer 2 (+ layer-3 partial products one slice behind)
        for (int q = 0; q < 4; ++q) { ptx::mbar_wait(&act1_ready[q], tpar); }
        ptx::tc_fence_after();
        l2_piece(0);
        l2_piece(1);
        l3_partial(0);
        l2_piece(2);
        l3_partial(1);
        l2_piece(3);
        l3_partial(2);
        l3_partial(3);
        // ---- layer 4: act3[128,128] x W4^T (N padded to 16)
        ptx::mbar_wait(act3_ready, tpar);
        ptx::tc_fence_after();
        {
          const uint32_t w = wait_stage();
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            ptx::mma_ss(tmem + kColD4, ptx::smem_desc(ring + ks * 4096, 2048, 128), ptx::smem_desc(w + ks * 512, 256, 128),
                        id16, ks != 0);
          release_stage();
          ptx::mma_commit(d4_full);
        }
      }
    }
  } else {
    // ===================================================== epilogue warps
    const int e = warp - 2;                 // 0..7
    const int quad = warp & 3;              // TMEM lane quadrant this warp may access
    const int half = e >> 2;                // which 64 of the piece's 128 columns
    const int row = quad * 32 + lane;       // row of the tile == TMEM lane
    const uint32_t t_lane = tmem + ((uint32_t)(quad * 32) << 16);
    float mu[16], inv[16];
    const bool has_scaler = args.mean != nullptr;
    for (int c = 0; c < args.in_dim && c < 16; ++c) {
      mu[c] = has_scaler ? (float)args.mean[c] : 0.f;
      inv[c] = has_scaler ? __fadd_rn(sqrtf((float)args.var[c]), 1e-8f) : 1.f;
    }
    uint32_t d_cnt[2] = {0, 0}, tiles_done = 0, free_cnt = 0;
    auto warp_arrive = [&](uint64_t* bar) {
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(bar);
    };
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tiles_done) {
      const uint32_t tpar = tiles_done & 1;
      const int64_t grow = (int64_t)tile * kTileM + row;
      // ---- X: scaler + bf16, K padded to 16 (half 0 warps write kc 0, half 1 warps kc 1)
      {
        float x[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int c = half * 8 + j;
          float v = 0.f;
          if (c < args.in_dim && grow < args.rows) {
            v = args.io.states[grow * args.in_dim + c];
            if (has_scaler) v = fminf(fmaxf(__fdiv_rn(__fsub_rn(v, mu[c]), inv[c]), -5.0f), 5.0f);
          }
          x[j] = v;
        }
        *reinterpret_cast<uint4*>(s_x + half * 2048 + row * 16) =
            make_uint4(ptx::pack_bf16(x[0], x[1]), ptx::pack_bf16(x[2], x[3]), ptx::pack_bf16(x[4], x[5]), ptx::pack_bf16(x[6], x[7]));
        ptx::fence_async_smem();
        warp_arrive(x_ready);
      }
      // ---- layer 1 pieces -> act1
      for (int q = 0; q < 4; ++q) {
        const int b = q & 1;
        ptx::mbar_wait(&d_full[b], d_cnt[b] & 1);
        ++d_cnt[b];
        ptx::tc_fence_after();
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          const int col = half * 64 + c * 32;                 // column inside the piece
          epi_chunk_to_smem(t_lane + kColD + b * 128 + col, s_bias + q * 128 + col, s_act1 + ((q * 128 + col) >> 3) * 2048, row);
        }
        ptx::tc_fence_before();
        ptx::fence_async_smem();
        warp_arrive(&d_empty[b]);
        warp_arrive(&act1_ready[q]);
      }
      // ---- layer 2 pieces -> ring (one 128-column slice of act2 at a time)
      for (int q = 0; q < 4; ++q) {
        const int b = q & 1;
        ptx::mbar_wait(&d_full[b], d_cnt[b] & 1);
        ++d_cnt[b];
        ptx::tc_fence_after();
        if (q > 0) {   // the previous slice has been consumed by its layer-3 partial product
          ptx::mbar_wait(act2_free, free_cnt & 1);
          ++free_cnt;
        }
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          const int col = half * 64 + c * 32;
          epi_chunk_to_smem(t_lane + kColD + b * 128 + col, s_bias + kH0 + q * 128 + col, s_ring + (col >> 3) * 2048, row);
        }
        ptx::tc_fence_before();
        ptx::fence_async_smem();
        warp_arrive(&d_empty[b]);
        warp_arrive(act2_ready);
      }
      // ---- layer 3 -> ring (act3)
      ptx::mbar_wait(act2_free, free_cnt & 1);   // completion of the 4th partial product (keeps the phase count in step)
      ++free_cnt;
      ptx::mbar_wait(d3_full, tpar);
      ptx::tc_fence_after();
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const int col = half * 64 + c * 32;
        epi_chunk_to_smem(t_lane + kColD3 + col, s_bias + kH0 + kH1 + col, s_ring + (col >> 3) * 2048, row);
      }
      ptx::tc_fence_before();
      ptx::fence_async_smem();
      warp_arrive(act3_ready);
      // ---- layer 4 -> output activation -> exploration -> actions
      ptx::mbar_wait(d4_full, tpar);
      ptx::tc_fence_after();
      if (half == 0) {
        uint32_t v[16];
        ptx::tmem_ld16(t_lane + kColD4, v);
        ptx::tmem_ld_wait();
        float mx = -INFINITY, mn = INFINITY, sm = 0.f;
        if (grow < args.rows) {
          const float* b4 = s_bias + kH0 + kH1 + kH2;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if (j < args.out_dim) {
              const float z = __uint_as_float(v[j]) + b4[j];
              float a = args.out_act == OIS_ACT_SIGMOID ? 1.0f / (1.0f + __expf(-z)) : args.out_act == OIS_ACT_TANH ? tanhf(z) : z;
              const int64_t i = grow * args.out_dim + j;
              if (args.io.ou_state) {
                float nz = 0.f;
                bool have = false;
                a = explore_elem(args.io, i, a, nz, have);
                if (have) { mx = fmaxf(mx, nz); mn = fminf(mn, nz); sm += nz; }
              }
              args.io.actions[i] = a;
            }
          }
        }
        if (args.io.noise_stats && args.io.ou_state && args.io.noise_scale >= 0.f) noise_stats_warp(args.io, mx, mn, sm);
      }
      ptx::tc_fence_before();
    }
  }
  // ---- teardown
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc<tc::kTmemCols>(tmem);
}

// ----------------------------------------------------------------------------------------------------------
// host side
// ----------------------------------------------------------------------------------------------------------
int tc_agent_init(OisAgent* a, cudaStream_t) {
  using namespace tc;
  OIS_CHECK_ARG(a->actor.h0 == kH0 && a->actor.h1 == kH1 && a->actor.h2 == kH2, "the tcgen05 actor is built for 512-512-128 hidden layers");
  OIS_CHECK_ARG(a->actor.in_dim <= kInPad && a->actor.out_dim <= 8, "actor in_dim <= 16 and out_dim <= 8");
  TcState* st = new TcState();
  cudaError_t e = cudaMalloc(&st->image, kImageBytes);
  if (e != cudaSuccess) {
    delete st;
    set_error("tc_agent_init: cudaMalloc -> %s", cudaGetErrorString(e));
    return OIS_ERR_CUDA;
  }
  e = cudaFuncSetAttribute(tc_actor_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  if (e != cudaSuccess) {
    cudaFree(st->image);
    delete st;
    set_error("tc_agent_init: cudaFuncSetAttribute(%d B smem) -> %s", kSmemBytes, cudaGetErrorString(e));
    return OIS_ERR_CUDA;
  }
  st->sms = num_sms();
  a->tc = st;
  return OIS_OK;
}

void tc_agent_free(OisAgent* a) {
  if (!a->tc) return;
  TcState* st = (TcState*)a->tc;
  if (st->image) cudaFree(st->image);
  delete st;
  a->tc = nullptr;
}

void tc_invalidate_weights(OisAgent* a) {
  if (a->tc) ((TcState*)a->tc)->valid = false;
}

int tc_actor_forward(OisAgent* a, const float* actor_params, const OisScaler* scaler, const OisActIO* io, int64_t rows,
                     cudaStream_t s) {
  using namespace tc;
  TcState* st = (TcState*)a->tc;
  OIS_CHECK_ARG(st, "agent was not created with OIS_PRECISION_BF16");
  if (!st->valid || st->packed_from != actor_params) {
    const int chunks = (int)(kImageBytes / 16);
    tc_pack_kernel<<<(chunks + 255) / 256, 256, 0, s>>>(actor_params, (uint4*)st->image, a->actor.in_dim, a->actor.out_dim);
    OIS_CHECK_LAUNCH();
    st->valid = true;
    st->packed_from = actor_params;
  }
  TcActArgs args;
  args.params = actor_params;
  args.image = (const uint8_t*)st->image;
  args.mean = scaler ? scaler->running_mean : nullptr;
  args.var = scaler ? scaler->running_variance : nullptr;
  args.in_dim = a->actor.in_dim;
  args.out_dim = a->actor.out_dim;
  args.out_act = a->actor.out_act;
  args.rows = rows;
  args.io = *io;
  const int tiles = (int)((rows + kTileM - 1) / kTileM);
  const int grid = tiles < st->sms ? tiles : st->sms;
  tc_actor_kernel<<<grid, kThreads, kSmemBytes, s>>>(args);
  OIS_CHECK_LAUNCH();
  return OIS_OK;
}

}  // namespace ois
