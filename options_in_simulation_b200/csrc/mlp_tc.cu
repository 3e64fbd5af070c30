// bf16 tcgen05 + TMA(bulk) fused actor: scaler -> 6->512->512->128->A MLP (ELU, Sigmoid/Tanh) -> OU noise -> clamp
// in ONE persistent, warp-specialised kernel (reference: src/my_agents/ddpg.py:215-268 with the
// model of train/train_kamma_4d_nofifo_random_RLIL.py:21-36).
//
// One CTA owns a 128-row tile of envs at a time; all four layers run back to back and the activations never
// leave the SM:
//   * weights: packed once per parameter change into a bf16 "stream image" (canonical UMMA no-swizzle K-major
//     core-matrix layout, in exactly the order the MMA warp consumes it: 42 stages of 16 KB), streamed from L2
//     through a 10-deep shared-memory ring (160 KB in flight) by 1-D bulk async copies (UBLKCP) on mbarriers;
//   * layer-1 / layer-3 outputs (act1 128x512, act3 128x128, bf16) live in TENSOR MEMORY and feed the next
//     layer as the A operand of tcgen05.mma (A-from-TMEM form): no shared-memory traffic for A, and shared
//     memory is left to the weight ring;
//   * MMA: one elected thread issues tcgen05.mma.cta_group::1.kind::f16, M=128, N=64 pieces with fp32
//     accumulators ping-ponging between two 64-column TMEM buffers; layer-3 partial products (A = the 64-column
//     slice of act2 just activated, in shared memory) are issued one piece behind layer 2;
//   * epilogue: two groups of 4 warps (one warp per TMEM lane quadrant in each group) take ALTERNATE 64-column
//     pieces, so one group's tcgen05.ld / mbarrier latencies hide behind the other group's arithmetic; layer 1
//     runs on four accumulator buffers (the layer-3 region is free then), layers 2 on two; each warp
//     tcgen05.ld's the accumulators, adds bias, ELU, converts to bf16 and tcgen05.st / st.shared's the operand;
//   * the last epilogue applies Sigmoid/Tanh, the OU exploration noise (explore.cuh) and the clamp.
// fp32 master weights are untouched; accumulation is fp32; tolerance vs the fp32 oracle: tests/test_agent_gpu.py.
#include "agent.cuh"
#include "explore.cuh"
#include "tc_ptx.cuh"
#include "update_tc.cuh"

namespace ois {

namespace tc {
constexpr int kTileM = 128;
// warp 0 and warps 10-12: bulk-copy producers; warp 1: MMA issuer A1 (layer 1, even pieces of layer 2); warps 2-9:
// epilogue; warp 13: MMA issuer B (layers 3, 4); warp 14: MMA issuer A2 (odd pieces of layer 2)
constexpr int kThreads = 480;
constexpr int kEpiWarps = 8;
constexpr int kWarpB = 2 + kEpiWarps + 3;  // 13
constexpr int kWarpA2 = kWarpB + 1;        // 14
constexpr int kH0 = 512, kH1 = 512, kH2 = 128, kInPad = 16, kOutPad = 16;
constexpr int kStageBytes = 16384;         // unit of the weight image
constexpr int kW2SlotBytes = 32768;        // layer-2 ring slot: half a 64-column piece (k-blocks 0-1 or 2-3)
constexpr int kW2Slots = 4;                // slots 0, 1 serve the even pieces (issuer A1), 2, 3 the odd ones (A2)
constexpr int kW3Slots = 2;                // layer-3 ring slot: one 64-wide K slice of W3, 16 KB
constexpr int kRingSlots = 2;              // act2 slices, one per epilogue group: [8 kc][128 rows][8] bf16 = 16 KB each
constexpr int kStreamStages = 42;          // 1 (W1) + 32 (W2) + 8 (W3) + 1 (W4)
constexpr int kLastStageBytes = kOutPad * kH2 * 2;                       // 4096
constexpr size_t kImageBytes = (size_t)(kStreamStages - 1) * kStageBytes + kLastStageBytes;
constexpr size_t kImgW1 = 0, kImgW2 = kStageBytes, kImgW3 = 33 * (size_t)kStageBytes, kImgW4 = 41 * (size_t)kStageBytes;
// shared memory map (bytes)
constexpr int kOffW2 = 0;                                      // 4 x 32 KB layer-2 weight ring
constexpr int kOffW3 = kOffW2 + kW2Slots * kW2SlotBytes;       // 2 x 16 KB layer-3 weight ring
constexpr int kOffW1 = kOffW3 + kW3Slots * kStageBytes;        // 16 KB, resident for the whole launch
constexpr int kOffW4 = kOffW1 + kStageBytes;                   // 4 KB, resident
constexpr int kOffRing = kOffW4 + kLastStageBytes;             // 2 x 16 KB act2 slices
constexpr int kOffX = kOffRing + kRingSlots * 16384;           // [2 kc][128][8] = 4 KB
constexpr int kOffBias = kOffX + kTileM * kInPad * 2;          // b1 512, b2 512, b3 128, b4 16 fp32
constexpr int kOffScale = kOffBias + (kH0 + kH1 + kH2 + kOutPad) * 4;   // scaler: mean[16], sqrt(var)+eps[16] fp32
constexpr int kOffBar = kOffScale + 2 * kInPad * 4;
constexpr int kSmemBytes = kOffBar + 512;
// TMEM columns
constexpr int kColAct = 0;      // act1 as bf16 pairs: 256 columns (act3 later reuses columns 0..63)
constexpr int kColD = 256;      // 64-column fp32 accumulator buffers: 0, 1 (layers 1, 2; layer 4 uses buffer 0) ...
constexpr int kColD3 = 384;     // ... 2, 3 during layer 1 only; then the 128 columns of layer 3
constexpr int kDBufs = 4;
constexpr int kBuf0UsesPerTile = 7;   // layer 1 pieces 0, 4; layer 2 pieces 0, 2, 4, 6; layer 4
constexpr int kBuf1UsesPerTile = 6;   // layer 1 pieces 1, 5; layer 2 pieces 1, 3, 5, 7
constexpr int kBuf23UsesPerTile = 2;  // layer 1 pieces 2, 6 / 3, 7
constexpr int kTmemCols = 512;

// image order (16 KB stages): W1 | W2 piece 0 (4 k-blocks of 128) | ... | W2 piece 7 | W3 slice 0 .. 7 | W4 (4 KB)
__host__ __device__ inline void stage_info(int s, int& layer, int& piece, int& kb) {
  if (s == 0) { layer = 1; piece = 0; kb = 0; }
  else if (s < 33) { layer = 2; piece = (s - 1) >> 2; kb = (s - 1) & 3; }
  else if (s < 41) { layer = 3; piece = s - 33; kb = 0; }
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
  else if (layer == 2) { W = P + w1; n = piece * 64 + (cc & 63); k0 = kb * 128 + (cc >> 6) * 8; N = kH1; K = kH0; }
  else if (layer == 3) { W = P + w2; n = cc & 127; k0 = piece * 64 + (cc >> 7) * 8; N = kH2; K = kH1; }
  else { W = P + w3; n = cc & 15; k0 = (cc >> 4) * 8; N = out_dim; K = kH2; }
  float v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = (n < N && k0 + j < K) ? W[(int64_t)n * K + k0 + j] : 0.f;
  image[c] = make_uint4(ptx::pack_bf16(v[0], v[1]), ptx::pack_bf16(v[2], v[3]), ptx::pack_bf16(v[4], v[5]),
                        ptx::pack_bf16(v[6], v[7]));
}




// ----------------------------------------------------------------------------------------------------------
// epilogue helpers (8 warps; warp quadrant = warp_id % 4 is the TMEM lane group the hardware lets it access)
// ----------------------------------------------------------------------------------------------------------
// ELU on a pair of pre-activations.  exp(x) = ex2(x log2 e) is ONE MUFU.EX2 (ex2.approx.ftz: no range fix-up — the result
// is rounded to bf16, and for x < -126/log2(e) it flushes to 0, i.e. ELU = -1, which is the bf16 answer anyway); the
// bias add, the log2(e) scale and the "- 1" run as packed FADD2 / FMUL2 (two elements per instruction on sm_100).
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t bias_elu_pack(float2 acc, float2 bias) {
  const float2 x = __fadd2_rn(acc, bias);
  const float2 t = __fmul2_rn(x, make_float2(1.4426950408889634f, 1.4426950408889634f));
  const float2 em1 = __fadd2_rn(make_float2(ex2_approx(t.x), ex2_approx(t.y)), make_float2(-1.0f, -1.0f));
  return ptx::pack_bf16(x.x > 0.f ? x.x : em1.x, x.y > 0.f ? x.y : em1.y);
}

// 32 fp32 accumulator columns of this thread's row -> +bias -> ELU -> 16 bf16 pairs.  The TMEM load and the arithmetic
// are separate so that the accumulator buffer can be handed back to the MMA warp as soon as the load has landed.
__device__ __forceinline__ void epi_act(const uint32_t (&v)[32], const float* __restrict__ bias, uint32_t (&w)[16]) {
  const float2* b2 = reinterpret_cast<const float2*>(bias);
#pragma unroll
  for (int e = 0; e < 16; ++e)
    w[e] = bias_elu_pack(make_float2(__uint_as_float(v[2 * e]), __uint_as_float(v[2 * e + 1])), b2[e]);
}
__device__ __forceinline__ void epi_load_act(uint32_t taddr, const float* __restrict__ bias, uint32_t (&w)[16]) {
  uint32_t v[32];
  ptx::tmem_ld32(taddr, v);
  ptx::tmem_ld_wait();
  epi_act(v, bias, w);
}
__device__ __forceinline__ float sigmoid_fast(float z) { return __fdividef(1.0f, 1.0f + ex2_approx(-1.4426950408889634f * z)); }
__device__ __forceinline__ float tanh_fast(float z) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(z));
  return y;
}

// Diagnostics (not part of the public ABI): per-phase %globaltimer stamps of CTA 0, read by profiles/tc_trace.py.
__device__ unsigned long long g_tc_trace[64];
__device__ __forceinline__ void tc_stamp(int slot) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  g_tc_trace[slot] = t;
}
#define TC_STAMP(cond, slot) do { if (cond) tc_stamp(slot); } while (0)
#ifdef OIS_TC_DIAG      // per-wait cycle counters of the MMA warp (profiles/tc_trace.py); off in the product build
#define TC_CLOCK() clock64()
#else
#define TC_CLOCK() 0ll
#endif

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
  uint8_t* s_w2 = smem + kOffW2;
  uint8_t* s_w3 = smem + kOffW3;
  uint8_t* s_w1 = smem + kOffW1;
  uint8_t* s_w4 = smem + kOffW4;
  uint8_t* s_ring = smem + kOffRing;
  uint8_t* s_x = smem + kOffX;
  float* s_bias = reinterpret_cast<float*>(smem + kOffBias);
  float* s_scale = reinterpret_cast<float*>(smem + kOffScale);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* w2_full = bars;                   // [4]
  uint64_t* w2_empty = w2_full + kW2Slots;    // [4]
  uint64_t* w3_full = w2_empty + kW2Slots;    // [2]
  uint64_t* w3_empty = w3_full + kW3Slots;    // [2]
  uint64_t* wconst_full = w3_empty + kW3Slots;
  uint64_t* x_ready = wconst_full + 1;
  uint64_t* d_full = x_ready + 1;             // [4]
  uint64_t* d_empty = d_full + kDBufs;        // [4]
  uint64_t* act1_ready = d_empty + kDBufs;
  uint64_t* slot_ready = act1_ready + 1;      // [2]
  uint64_t* slot_free = slot_ready + kRingSlots;   // [2]
  uint64_t* d3_full = slot_free + kRingSlots;
  uint64_t* act3_ready = d3_full + 1;
  uint64_t* stagger = act3_ready + 1;         // issuer A1 is half a piece into layer 2: issuer A2 may start
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(stagger + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_tiles = (int)((args.rows + kTileM - 1) / kTileM);
  const bool tr = blockIdx.x == 0 && lane == 0;      // trace CTA 0
  TC_STAMP(tr && warp == 2, 0);

  // ---- one-time setup
  if (threadIdx.x == 0) {
    for (int i = 0; i < kW2Slots; ++i) { ptx::mbar_init(&w2_full[i], 1); ptx::mbar_init(&w2_empty[i], 1); }
    for (int i = 0; i < kW3Slots; ++i) { ptx::mbar_init(&w3_full[i], 1); ptx::mbar_init(&w3_empty[i], 1); }
    ptx::mbar_init(wconst_full, 1);
    ptx::mbar_init(x_ready, kEpiWarps);
    for (int i = 0; i < kDBufs; ++i) { ptx::mbar_init(&d_full[i], 1); ptx::mbar_init(&d_empty[i], kEpiWarps / 2); }
    ptx::mbar_init(act1_ready, (kEpiWarps / 2) * 8);       // 8 pieces, each finished by the 4 warps of one group
    for (int i = 0; i < kRingSlots; ++i) { ptx::mbar_init(&slot_ready[i], kEpiWarps / 2); ptx::mbar_init(&slot_free[i], 1); }
    ptx::mbar_init(d3_full, 1);
    ptx::mbar_init(act3_ready, kEpiWarps);
    ptx::mbar_init(stagger, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 1) ptx::tmem_alloc<kTmemCols>(tmem_slot);
  if (warp >= 2 && warp < 2 + kEpiWarps) {   // biases: fp32, straight from the master parameters
    const int64_t b0 = (int64_t)kH0 * args.in_dim, w1 = b0 + kH0, b1 = w1 + (int64_t)kH1 * kH0, w2 = b1 + kH1;
    const int64_t b2 = w2 + (int64_t)kH2 * kH1, w3 = b2 + kH2, b3 = w3 + (int64_t)args.out_dim * kH2;
    // all of a thread's loads are issued before any is consumed: the set-up is on the critical path whenever the kernel
    // is not overlapping a predecessor (first kernel of a graph replay, cold start), and five dependent global-memory
    // round trips were most of its 3 us
    constexpr int kBiasTotal = kH0 + kH1 + kH2 + kOutPad;
    constexpr int kBiasIters = (kBiasTotal + kEpiWarps * 32 - 1) / (kEpiWarps * 32);
    float bv[kBiasIters];
#pragma unroll
    for (int k = 0; k < kBiasIters; ++k) {
      const int i = threadIdx.x - 64 + k * kEpiWarps * 32;
      float v = 0.f;
      if (i < kH0) v = __ldg(args.params + b0 + i);
      else if (i < kH0 + kH1) v = __ldg(args.params + b1 + i - kH0);
      else if (i < kH0 + kH1 + kH2) v = __ldg(args.params + b2 + i - kH0 - kH1);
      else if (i < kBiasTotal && (i - kH0 - kH1 - kH2) < args.out_dim) v = __ldg(args.params + b3 + i - kH0 - kH1 - kH2);
      bv[k] = v;
    }
#pragma unroll
    for (int k = 0; k < kBiasIters; ++k) {
      const int i = threadIdx.x - 64 + k * kEpiWarps * 32;
      if (i < kBiasTotal) s_bias[i] = bv[k];
    }
    if (threadIdx.x - 64 < 2 * kInPad) {   // RunningStandardScaler constants: mean.float(), sqrt(var.float()) + eps
      const int c = (threadIdx.x - 64) & (kInPad - 1);
      const bool is_mean = threadIdx.x - 64 < kInPad;
      float v = is_mean ? 0.f : 1.f;
      if (args.mean != nullptr && c < args.in_dim) v = is_mean ? (float)args.mean[c] : __fadd_rn(sqrtf((float)args.var[c]), 1e-8f);
      s_scale[threadIdx.x - 64] = v;
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  TC_STAMP(tr && warp == 2, 1);

  if (warp == 0 || (warp >= 2 + kEpiWarps && warp < kWarpB)) {
    // ===================================================== producers (one thread each; a thread sustains only one bulk copy
    // per ~520 cycles whatever its size — profiles/mma_rate.cu): 0 and 1 alternate on the 32 KB half-pieces of W2,
    // 2 streams the 16 KB K-slices of W3, 3 loads W1 and W4 once (they stay resident)
    // (programmatic dependent launch: the weight image, like the parameters, was written at least two kernels ago —
    // the repack kernel never triggers early — so the producers start streaming before the predecessor has finished)
    if (lane == 0) {
      const int p = warp == 0 ? 0 : warp - (1 + kEpiWarps);
      if (p < 2) {
        uint32_t n = 0;   // half-pieces issued so far by both W2 producers
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
          for (int h = 0; h < 16; ++h, ++n) {
            if ((int)(n & 1) != p) continue;
            const int slot = n % kW2Slots;
            ptx::mbar_wait(&w2_empty[slot], ((n / kW2Slots) & 1) ^ 1);
            ptx::mbar_expect_tx(&w2_full[slot], kW2SlotBytes);
            ptx::bulk_g2s(s_w2 + slot * kW2SlotBytes, args.image + tc::kImgW2 + (size_t)h * kW2SlotBytes, kW2SlotBytes, &w2_full[slot]);
          }
        }
      } else if (p == 2) {
        uint32_t n = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
          for (int k = 0; k < 8; ++k, ++n) {
            const int slot = n % kW3Slots;
            ptx::mbar_wait(&w3_empty[slot], ((n / kW3Slots) & 1) ^ 1);
            ptx::mbar_expect_tx(&w3_full[slot], kStageBytes);
            ptx::bulk_g2s(s_w3 + slot * kStageBytes, args.image + tc::kImgW3 + (size_t)k * kStageBytes, kStageBytes, &w3_full[slot]);
          }
        }
      } else {
        ptx::mbar_expect_tx(wconst_full, kStageBytes + kLastStageBytes);
        ptx::bulk_g2s(s_w1, args.image + tc::kImgW1, kStageBytes, wconst_full);
        ptx::bulk_g2s(s_w4, args.image + tc::kImgW4, kLastStageBytes, wconst_full);
      }
    }
  } else if (warp == 1 || warp == kWarpA2) {
    // ===================================================== MMA issuers A1 (warp 1: layer 1, even pieces of layer 2) and
    // A2 (odd pieces of layer 2).  A whole warp runs the (warp-uniform) control flow and the mbarrier waits; one elected
    // lane issues tcgen05.mma / tcgen05.commit.  The tensor pipe's issue queue is shallow (profiles/mma_rate.cu: a wait +
    // fence + commit between groups of 16 MMAs costs ~13 cycles per 32-cycle MMA), so two issuers work on two pieces —
    // two accumulator buffers — at once, A2 half a piece behind A1: whenever one of them is between groups or waiting for
    // its accumulator to be drained, the other one's MMAs keep the pipe busy.
    const bool a1 = warp == 1;
    const int odd = a1 ? 0 : 1;
    const bool leader = ptx::elect_one();
    const uint32_t id64 = ptx::idesc_bf16(128, 64);
    constexpr uint32_t hi = ptx::smem_desc_hi(128);
    const uint32_t xs = ptx::smem_u32(s_x), ws2 = ptx::smem_u32(s_w2);
    const uint32_t x_lo = ptx::smem_desc_lo(xs, 2048);
    const uint32_t w1_lo = ptx::smem_desc_lo(ptx::smem_u32(s_w1), 8192);
    uint32_t tiles_done = 0;
    long long t_w = 0, t_d = 0, t_i = 0, t_c = 0, t_l2 = 0;   // diagnostics (OIS_TC_DIAG): cycles waiting for weights /
                                                              // accumulators, issuing MMAs, committing, whole layer 2
    // accumulator buffer b, its use number `idx` within the tile (kBuf*UsesPerTile uses per tile, in a fixed order)
    auto acquire_d = [&](int b, int idx) {
      const uint32_t per_tile = b == 0 ? kBuf0UsesPerTile : b == 1 ? kBuf1UsesPerTile : kBuf23UsesPerTile;
      const long long c0 = TC_CLOCK();
      ptx::mbar_wait(&d_empty[b], ((tiles_done * per_tile + (uint32_t)idx) & 1) ^ 1);
      t_d += TC_CLOCK() - c0;
      ptx::tc_fence_after();
    };
    if (a1) ptx::mbar_wait(wconst_full, 0);
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tiles_done) {
      const uint32_t tpar = tiles_done & 1;
      if (a1) {
        // ---- layer 1: X[128,16] (smem) x W1^T, eight 64-column pieces on four accumulator buffers
        ptx::mbar_wait(x_ready, tpar);
        ptx::tc_fence_after();
        TC_STAMP(tr && tiles_done == 0, 32);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          acquire_d(q & 3, q >> 2);
          if (leader) {
            ptx::mma_ss(tmem + kColD + (q & 3) * 64, ptx::make_desc(x_lo, hi), ptx::make_desc(w1_lo + q * (1024 >> 4), hi), id64, false);
            ptx::mma_commit(&d_full[q & 3]);
          }
          __syncwarp();
        }
        TC_STAMP(tr && tiles_done == 0, 34);
      }
      // ---- layer 2: act1 (TMEM) x W2[64q .. 64q+64, :]^T, two 32 KB half-pieces (K = 256 each) per piece
      ptx::mbar_wait(act1_ready, tpar);
      if (!a1) ptx::mbar_wait(stagger, tpar);
      ptx::tc_fence_after();
      TC_STAMP(tr && a1 && tiles_done == 0, 35);
      const long long l2_0 = TC_CLOCK();
#pragma unroll 1
      for (int q = odd; q < 8; q += 2) {
        acquire_d(odd, 2 + (q >> 1));
        const uint32_t d = tmem + kColD + odd * 64;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const uint32_t n2 = tiles_done * 16u + (uint32_t)(q * 2 + half);   // half-piece number in stream order
          const int slot = n2 % kW2Slots;
          const long long c0 = TC_CLOCK();
          ptx::mbar_wait(&w2_full[slot], (n2 / kW2Slots) & 1);
          t_w += TC_CLOCK() - c0;
          ptx::tc_fence_after();
          const long long c1 = TC_CLOCK();
          if (leader) {
            const uint32_t w_lo = ptx::smem_desc_lo(ws2 + slot * kW2SlotBytes, 1024);
#pragma unroll
            for (int j = 0; j < 16; ++j)
              ptx::mma_ts(d, tmem + kColAct + (half * 16 + j) * 8, ptx::make_desc(w_lo + j * (2048 >> 4), hi), id64, (half | j) != 0);
          }
          __syncwarp();
          const long long c2 = TC_CLOCK();
          if (leader) {
            ptx::mma_commit(&w2_empty[slot]);
            if (a1 && q == 0 && half == 0) ptx::mma_commit(stagger);
            if (half == 1) ptx::mma_commit(&d_full[odd]);
          }
          __syncwarp();
          t_i += c2 - c1;
          t_c += TC_CLOCK() - c2;
        }
        TC_STAMP(tr && a1 && tiles_done == 0 && q == 0, 36);
      }
      TC_STAMP(tr && a1 && tiles_done == 0, 37);
      t_l2 = TC_CLOCK() - l2_0;
      if (tr && tiles_done == 0) {
        const int o = a1 ? 40 : 50;
        g_tc_trace[o] = t_w; g_tc_trace[o + 1] = t_d; g_tc_trace[o + 3] = t_i; g_tc_trace[o + 4] = t_c; g_tc_trace[o + 5] = t_l2;
      }
    }
  } else if (warp == kWarpB) {
    // ===================================================== MMA issuer B: layer-3 partial products (A = the act2 slice the
    // epilogue just wrote to shared memory, accumulating over the eight slices) and layer 4 (A = act3 in TMEM)
    const bool leader = ptx::elect_one();
    const uint32_t id128 = ptx::idesc_bf16(128, 128), id16 = ptx::idesc_bf16(128, 16);
    constexpr uint32_t hi = ptx::smem_desc_hi(128);
    const uint32_t ring = ptx::smem_u32(s_ring), ws3 = ptx::smem_u32(s_w3);
    const uint32_t w4_lo = ptx::smem_desc_lo(ptx::smem_u32(s_w4), 256);
    uint32_t n3 = 0, slot_uses[kRingSlots] = {0, 0}, tiles_done = 0;
    long long t_s = 0;
    ptx::mbar_wait(wconst_full, 0);
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tiles_done) {
      const uint32_t tpar = tiles_done & 1;
#pragma unroll 1
      for (int p = 0; p < 8; ++p) {
        const int slot = p & 1, wslot = n3 % kW3Slots;
        const long long c0 = TC_CLOCK();
        ptx::mbar_wait(&slot_ready[slot], slot_uses[slot] & 1);
        t_s += TC_CLOCK() - c0;
        ++slot_uses[slot];
        ptx::mbar_wait(&w3_full[wslot], (n3 / kW3Slots) & 1);
        ptx::tc_fence_after();
        if (leader) {
          const uint32_t w_lo = ptx::smem_desc_lo(ws3 + wslot * kStageBytes, 2048);
          const uint32_t a_lo = ptx::smem_desc_lo(ring + slot * 16384, 2048);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            ptx::mma_ss(tmem + kColD3, ptx::make_desc(a_lo + j * (4096 >> 4), hi), ptx::make_desc(w_lo + j * (4096 >> 4), hi), id128,
                        (p | j) != 0);
          ptx::mma_commit(&w3_empty[wslot]);
          ptx::mma_commit(&slot_free[slot]);
          if (p == 7) ptx::mma_commit(d3_full);
        }
        __syncwarp();
        ++n3;
      }
      TC_STAMP(tr && tiles_done == 0, 38);
      // ---- layer 4: act3 (TMEM) x W4^T (N padded to 16) into accumulator buffer 0 (its 7th use in this tile)
      ptx::mbar_wait(act3_ready, tpar);
      ptx::mbar_wait(&d_empty[0], ((tiles_done * kBuf0UsesPerTile + (kBuf0UsesPerTile - 1)) & 1) ^ 1);
      ptx::tc_fence_after();
      if (leader) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          ptx::mma_ts(tmem + kColD, tmem + kColAct + j * 8, ptx::make_desc(w4_lo + j * (512 >> 4), hi), id16, j != 0);
        ptx::mma_commit(&d_full[0]);
      }
      __syncwarp();
      TC_STAMP(tr && tiles_done == 0, 39);
      if (tr && tiles_done == 0) g_tc_trace[42] = t_s;
    }
  } else {
    // ===================================================== epilogue warps: two groups of four (one per TMEM lane quadrant)
    const int e = warp - 2;                 // 0..7
    const int quad = warp & 3;              // TMEM lane quadrant this warp may access
    const int g = e >> 2;                   // group: takes the pieces q with (q & 1) == g
    const int row = quad * 32 + lane;       // row of the tile == TMEM lane
    const uint32_t t_lane = tmem + ((uint32_t)(quad * 32) << 16);
    uint32_t d_cnt[kDBufs] = {0, 0, 0, 0}, tiles_done = 0;
    auto warp_arrive = [&](uint64_t* bar) {
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(bar);
    };
    auto wait_d = [&](int b) {
      ptx::mbar_wait(&d_full[b], d_cnt[b] & 1);
      ++d_cnt[b];
      ptx::tc_fence_after();
    };
    const int A = args.out_dim;
    const bool row_noise = args.io.ou_state != nullptr && (A == 1 || A == 2 || A == 4);
    // X of one tile: raw loads into registers (issued early, consumed a layer later), then scaler + bf16 + smem
    auto load_x = [&](int tile, float (&x)[8]) {
      const int64_t r = (int64_t)tile * kTileM + row;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int c = g * 8 + j;
        x[j] = (c < args.in_dim && r < args.rows) ? args.io.states[r * args.in_dim + c] : 0.f;
      }
    };
    auto write_x = [&](float (&x)[8]) {     // K padded to 16: group 0 writes k 0..7, group 1 k 8..15
      if (args.mean != nullptr) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int c = g * 8 + j;
          if (c < args.in_dim) x[j] = fminf(fmaxf(__fdiv_rn(__fsub_rn(x[j], s_scale[c]), s_scale[kInPad + c]), -5.0f), 5.0f);
        }
      }
      *reinterpret_cast<uint4*>(s_x + g * 2048 + row * 16) =
          make_uint4(ptx::pack_bf16(x[0], x[1]), ptx::pack_bf16(x[2], x[3]), ptx::pack_bf16(x[4], x[5]), ptx::pack_bf16(x[6], x[7]));
      ptx::fence_async_smem();
      warp_arrive(x_ready);
    };
    // Programmatic dependent launch: everything above (barriers, TMEM, biases, the weight stream) overlapped the tail of
    // the previous kernel; the observations are the first thing that kernel (the env step) produced.  One thread per
    // CTA lets the NEXT kernel start, after its own wait (ois_common.cuh states the rule).
    pdl_wait();
    if (warp == 2 && lane == 0) pdl_trigger();
    float xr[8];
    if ((int)blockIdx.x < num_tiles) {
      load_x(blockIdx.x, xr);
      write_x(xr);
      TC_STAMP(tr && warp == 2, 2);
    }
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tiles_done) {
      const uint32_t tpar = tiles_done & 1;
      const int64_t grow = (int64_t)tile * kTileM + row;
      const int next_tile = tile + gridDim.x;
      // ---- layer 1: this group's pieces (64 columns each) -> act1 (TMEM, bf16 pairs)
      if (g == 1 && tiles_done > 0) {
        // act1 overwrites the columns the previous tile's act3 lives in, which layer 4 (issuer B) reads.  Group 0 has
        // waited for that layer 4 itself (it owns the output); group 1 checks the same phase of d_full[0] here.
        ptx::mbar_wait(&d_full[0], ((tiles_done - 1) * kBuf0UsesPerTile + (kBuf0UsesPerTile - 1)) & 1);
        ptx::tc_fence_after();
      }
      for (int q = g; q < 8; q += 2) {
        const int b = q & 3;
        wait_d(b);
        uint32_t v0[32], v1[32];
        ptx::tmem_ld32(t_lane + kColD + b * 64, v0);
        ptx::tmem_ld32(t_lane + kColD + b * 64 + 32, v1);
        ptx::tmem_ld_wait();
        ptx::tc_fence_before();
        warp_arrive(&d_empty[b]);                                           // accumulator drained: the MMA warp may refill it
        uint32_t w[16];
        epi_act(v0, s_bias + q * 64, w);
        ptx::tmem_st16(t_lane + kColAct + q * 32, w);
        epi_act(v1, s_bias + q * 64 + 32, w);
        ptx::tmem_st16(t_lane + kColAct + q * 32 + 16, w);
        ptx::tmem_st_wait();
        ptx::tc_fence_before();
        warp_arrive(act1_ready);
        TC_STAMP(tr && (warp == 2 || warp == 6) && tiles_done == 0, 3 + q);
      }
      long long e_d = 0, e_s = 0;     // diagnostics (OIS_TC_DIAG): layer-2 waits of this warp for accumulators / ring slots
      // the next tile's observations: loads in flight during layer 2
      if (next_tile < num_tiles) load_x(next_tile, xr);
      // ---- layer 2: this group's pieces -> one 64-column slice of act2 each, in the shared-memory ring
      for (int q = g; q < 8; q += 2) {
        const int slot = g;              // each group owns one act2 slice buffer
        const long long c0 = TC_CLOCK();
        wait_d(g);
        e_d += TC_CLOCK() - c0;
        uint32_t v0[32], v1[32];
        ptx::tmem_ld32(t_lane + kColD + g * 64, v0);
        ptx::tmem_ld32(t_lane + kColD + g * 64 + 32, v1);
        ptx::tmem_ld_wait();
        ptx::tc_fence_before();
        warp_arrive(&d_empty[g]);                                           // accumulator drained
        uint32_t w0[16], w1[16];
        epi_act(v0, s_bias + kH0 + q * 64, w0);
        epi_act(v1, s_bias + kH0 + q * 64 + 32, w1);
        // this group's slot is used four times per tile (pieces g, g + 2, g + 4, g + 6)
        const uint32_t use = tiles_done * 4u + (uint32_t)(q >> 1);
        const long long c1 = TC_CLOCK();
        ptx::mbar_wait(&slot_free[slot], (use & 1) ^ 1);                    // previous user of the slot consumed
        e_s += TC_CLOCK() - c1;
        uint8_t* dst = s_ring + slot * 16384 + row * 16;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          *reinterpret_cast<uint4*>(dst + j * 2048) = make_uint4(w0[4 * j], w0[4 * j + 1], w0[4 * j + 2], w0[4 * j + 3]);
          *reinterpret_cast<uint4*>(dst + (4 + j) * 2048) = make_uint4(w1[4 * j], w1[4 * j + 1], w1[4 * j + 2], w1[4 * j + 3]);
        }
        ptx::fence_async_smem();
        warp_arrive(&slot_ready[slot]);
        TC_STAMP(tr && (warp == 2 || warp == 6) && tiles_done == 0, 11 + q);
      }
      // every layer-1 MMA of this tile has completed (a layer-2 accumulator was seen full): the X buffer is free, so the
      // next tile's layer 1 can start the moment the MMA warp gets there
      if (next_tile < num_tiles) write_x(xr);
      if (tr && tiles_done == 0 && (warp == 2 || warp == 6)) { g_tc_trace[46 + 2 * g] = e_d; g_tc_trace[47 + 2 * g] = e_s; }
      // ---- exploration noise of this row, prepared while the last layer-2 / layer-3 MMAs run (group 0 owns the output)
      float nz[4] = {0.f, 0.f, 0.f, 0.f};
      bool applied = false;
      if (g == 0 && row_noise && grow < args.rows) explore_row_prepare(args.io, grow, A, nz, applied);
      // ---- layer 3 -> act3 (TMEM columns 0..63; act1 is dead: every layer-2 MMA has completed)
      ptx::mbar_wait(d3_full, tpar);
      ptx::tc_fence_after();
      TC_STAMP(tr && warp == 2 && tiles_done == 0, 19);
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t w[16];
        epi_load_act(t_lane + kColD3 + g * 64 + c * 32, s_bias + kH0 + kH1 + g * 64 + c * 32, w);
        ptx::tmem_st16(t_lane + kColAct + g * 32 + c * 16, w);
      }
      ptx::tmem_st_wait();
      ptx::tc_fence_before();
      warp_arrive(act3_ready);
      TC_STAMP(tr && warp == 2 && tiles_done == 0, 20);
      // ---- layer 4 -> output activation -> exploration -> actions (group 0; accumulator buffer 0)
      if (g == 0) {
        wait_d(0);
        TC_STAMP(tr && warp == 2 && tiles_done == 0, 21);
        uint32_t v[16];
        ptx::tmem_ld16(t_lane + kColD, v);
        ptx::tmem_ld_wait();
        ptx::tc_fence_before();
        warp_arrive(&d_empty[0]);
        float mx = -INFINITY, mn = INFINITY, sm = 0.f;
        if (grow < args.rows) {
          const float* b4 = s_bias + kH0 + kH1 + kH2;
          float a[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float z = __uint_as_float(v[j]) + b4[j];
            a[j] = args.out_act == OIS_ACT_SIGMOID ? sigmoid_fast(z) : args.out_act == OIS_ACT_TANH ? tanh_fast(z) : z;
          }
          if (row_noise) {
            if (applied) {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                if (j < A) {
                  a[j] = fminf(fmaxf(__fadd_rn(a[j], nz[j]), args.io.clip_low), args.io.clip_high);
                  mx = fmaxf(mx, nz[j]); mn = fminf(mn, nz[j]); sm += nz[j];
                }
              }
            }
          } else if (args.io.ou_state) {   // act_dim not in {1, 2, 4}: element-wise path
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              if (j < A) {
                float n1 = 0.f;
                bool have = false;
                a[j] = explore_elem(args.io, grow * A + j, a[j], n1, have);
                if (have) { mx = fmaxf(mx, n1); mn = fminf(mn, n1); sm += n1; }
              }
            }
          }
          float* out = args.io.actions + grow * A;
          if (A == 4) *reinterpret_cast<float4*>(out) = make_float4(a[0], a[1], a[2], a[3]);
          else if (A == 2) *reinterpret_cast<float2*>(out) = make_float2(a[0], a[1]);
          else {
#pragma unroll
            for (int j = 0; j < 8; ++j) if (j < A) out[j] = a[j];
          }
        }
        if (args.io.noise_stats && args.io.ou_state && act_noise_scale(args.io) >= 0.f) noise_stats_warp(args.io, mx, mn, sm);
        TC_STAMP(tr && warp == 2 && tiles_done == 0, 22);
      }
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

// Which kernel serves `rows` envs: the transposed small-tile kernel (mlp_tct.cu: 64 envs per CTA, reads the update's
// tile-blocked image) up to OIS_TCT_MAX_ROWS rows (default 64 x SMs), the 128-env-row-tile kernel of this file beyond.
static int64_t g_tct_max_rows = -1;     // < 0: OIS_TCT_MAX_ROWS or the default; set by ois_debug_set_tct_max_rows (tests)
bool tc_use_transposed(const OisAgent* a, int64_t rows) {
  // default: while the 64-env tiles still fit ONE wave of the SMs (9472 rows on 148 SMs).  Beyond that a CTA streams the
  // 657 KB weight image once per 64 envs instead of once per 128: measured at 16 384 envs 39.9 us against 24.6 us
  static const int64_t dflt = [] {
    const char* e = getenv("OIS_TCT_MAX_ROWS");
    return e ? (int64_t)atoll(e) : (int64_t)64 * num_sms();
  }();
  const int64_t max_rows = g_tct_max_rows >= 0 ? g_tct_max_rows : dflt;
  return a->tcu != nullptr && a->actor.in_dim <= 8 && rows <= max_rows;
}

// (Re)build the bf16 weight image the inference kernel for `rows` envs reads, if the fp32 master changed since.
int tc_sync_image(OisAgent* a, const float* actor_params, int64_t rows, cudaStream_t s) {
  using namespace tc;
  if (tc_use_transposed(a, rows)) return tcu_image(a, TCU_ACTOR, actor_params, s);   // kept current by the update's Adam
  TcState* st = (TcState*)a->tc;
  OIS_CHECK_ARG(st, "agent was not created with OIS_PRECISION_BF16");
  if (!st->valid || st->packed_from != actor_params) {
    const int chunks = (int)(kImageBytes / 16);
    tc_pack_kernel<<<(chunks + 255) / 256, 256, 0, s>>>(actor_params, (uint4*)st->image, a->actor.in_dim, a->actor.out_dim);
    OIS_CHECK_LAUNCH();
    st->valid = true;
    st->packed_from = actor_params;
  }
  return OIS_OK;
}

int tc_actor_forward(OisAgent* a, const float* actor_params, const OisScaler* scaler, const OisActIO* io, int64_t rows,
                     cudaStream_t s) {
  using namespace tc;
  if (int e = tc_sync_image(a, actor_params, rows, s)) return e;
  if (tc_use_transposed(a, rows))
    return tct_actor_forward(a, tcu_image_if_valid(a, TCU_ACTOR, actor_params), actor_params, scaler, io, rows, s);
  TcState* st = (TcState*)a->tc;
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
  OIS_CHECK_CUDA(launch_pdl(tc_actor_kernel, grid, kThreads, kSmemBytes, s, args));
  return OIS_OK;
}

}  // namespace ois

// diagnostics hook for tests and profiles (deliberately not declared in include/ois_b200.h): row count up to which
// ois_agent_act uses the transposed small-tile kernel; 0 = always the 128-env-row-tile kernel, < 0 = the default
extern "C" int ois_debug_set_tct_max_rows(long long rows) {
  ois::g_tct_max_rows = rows;
  return 0;
}

// diagnostics hook for profiles/tc_trace.py (deliberately not declared in include/ois_b200.h)
extern "C" int ois_debug_tc_trace(unsigned long long* out_host, int n) {
  if (n > 64) n = 64;
  return cudaMemcpyFromSymbol(out_host, ois::g_tc_trace, sizeof(unsigned long long) * n) == cudaSuccess ? 0 : -2;
}
