// bf16 tcgen05 fused actor for SMALL env counts per GPU (the configuration BASELINE.json names: 4096 envs): the same
// scaler -> 6->512->512->128->A MLP (ELU, Sigmoid/Tanh) -> OU noise -> clamp as mlp_tc.cu (reference:
// src/my_agents/ddpg.py:215-268, model of train/train_kamma_4d_nofifo_random_RLIL.py:21-36), computed TRANSPOSED:
//
//      H_l^T [features, envs] = W_l [features_out, features_in] . H_{l-1}^T [features_in, envs]
//
// The weights are the tcgen05 A operand (M = 128 output features per MMA, straight from the tile-blocked bf16 image the
// update keeps current — update_tc.cuh — so nothing is re-packed after an optimizer step), the envs are the N dimension.
// Why: with the envs on M (mlp_tc.cu) a tile is 128 envs, 4096 envs are 32 CTAs on 148 SMs and every epilogue thread owns
// one env row and walks all 512 + 512 + 128 columns of it — 22 us of mostly serial epilogue.  With the envs on N a tile is
// 64 envs (64 CTAs), an epilogue thread owns one FEATURE (one TMEM lane) and 64 envs of it, sixteen epilogue warps cover
// the four 128-feature blocks of a layer at once, and the bias is a per-thread scalar.  What is left on the critical path
// is the weight stream (657 KB per tile from L2 through a 4 x 32 KB ring) and the MMAs themselves.
//
//   * activations never leave the SM: the epilogue writes H_l^T as bf16 into shared memory in the canonical UMMA
//     MN-major (here: env-major) no-swizzle layout [env chunk of 8][feature][8 envs] — a thread's eight consecutive envs
//     are one 16-byte store, a warp's 32 features are 512 contiguous bytes — and that buffer is the B operand of the next
//     layer (LBO 128 B = next 8 features, SBO = features * 16 B = next 8 envs);
//   * layer 2 runs k-block-outer (block kb of H1 = the 128 features epilogue group kb produced), so its MMAs start as
//     soon as the first group is done; all four accumulator blocks of a layer complete together and H2 overwrites H1;
//   * accumulators: D1 in TMEM columns 0..255, D2 in 256..511, D3 / D4 (layers 3, 4) reuse 0..127;
//   * layer 4 (A outputs) is an M = 128 MMA over the 16-row W4 image (rows >= 16 read whatever follows it in shared
//     memory and land in TMEM lanes nobody reads); lanes 0..A-1 are transposed through shared memory so that one warp
//     finishes two envs per lane: output activation, OU noise (prepared by an auxiliary warp while the MLP runs), clamp.
// Numerics: identical contract to mlp_tc.cu (bf16 operands, fp32 accumulation, fp32 bias / activation); tests compare
// both kernels with the fp32 oracle (tests/test_agent_gpu.py).
#include "agent.cuh"
namespace ois { namespace tct {
__device__ unsigned long long g_tct_trace[64];
// latest warp exit of the launch (reported as slot 61).  A symbol of its own: with the exit stamp in g_tct_trace the
// compiler kept &g_tct_trace[60] from the entry stamps alive to the end of the kernel, spilled it in the kEnv build,
// and that spill store — the first local-memory access of every thread, a cold translation after an L2 flush — delayed
// the bias loads queued behind it by ~2 us.
__device__ unsigned long long g_tct_last_end;
} }
#define ENV_ONE_STAMP(slot)                                              \
  do {                                                                   \
    if (blockIdx.x == 0 && (threadIdx.x & 31) == 0) {                    \
      unsigned long long t_;                                             \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));             \
      ::ois::tct::g_tct_trace[slot] = t_;                                \
    }                                                                    \
  } while (0)
#include "env_one.cuh"
#include "explore.cuh"
#include "tc_ptx.cuh"
#include "update_tc.cuh"

namespace ois {
namespace tct {

constexpr int kEpiWarps = 16;                // warps 0..15: group g = warp >> 2 owns feature block g, quadrant = warp & 3
constexpr int kWarpMma = 16;                 // tcgen05.mma issuer (one elected lane), TMEM allocation
constexpr int kWarpProd = 17;                // 17, 18: bulk-copy producers (alternate ring slots)
constexpr int kWarpAux = 19;                 // observations -> scaler -> X operand; OU noise of the tile
constexpr int kThreads = 20 * 32;
constexpr int kH0 = 512, kH1 = 512, kH2 = 128, kInPad = 16, kOutPad = 16;
constexpr int kSlotBytes = 32768;            // one [128 features, 128 k] bf16 weight tile
constexpr int kStagesPerTile = 20;           // 16 tiles of W2 (kb-major), 4 of W3
constexpr int kBiasTotal = kH0 + kH1 + kH2 + kOutPad;
// NT = envs per tile = MMA N: 32 (up to ~4.7k envs: twice the CTAs, half the epilogue per thread, a 5-slot ring) or 64
template <int NT>
struct Cfg {
  static constexpr int kSlots = NT == 32 ? 5 : 4;
  // shared memory map (bytes)
  static constexpr int kOffRing = 0;
  static constexpr int kOffH = kOffRing + kSlots * kSlotBytes;   // [NT/8 env chunks][512 features][8] bf16 (H1, then H2, then H3)
  static constexpr int kOffW1 = kOffH + NT * kH0 * 2;            // 4 blocks x [2 k chunks][128][8] = 16 KB, resident
  static constexpr int kOffW4 = kOffW1 + kH0 * kInPad * 2;       // [16 k chunks][16][8] = 4 KB, resident (over-read into X / bias: finite)
  static constexpr int kOffX = kOffW4 + kOutPad * kH2 * 2;       // [2 k chunks][NT envs][8]
  static constexpr int kOffBias = kOffX + NT * kInPad * 2;       // b1 512, b2 512, b3 128, b4 16 fp32
  static constexpr int kOffScale = kOffBias + kBiasTotal * 4;    // mean[16], 1 / (sqrt(var)+eps) [16]
  static constexpr int kOffOut = kOffScale + 2 * kInPad * 4;     // 2 x [NT envs][8] fp32 (by tile parity)
  static constexpr int kOffNoise = kOffOut + 2 * NT * 8 * 4;     // 2 x [NT envs][4] fp32
  static constexpr int kOffStage = kOffNoise + 2 * NT * 4 * 4;   // [NT envs][8] fp32: raw observations of the next tile (LDGSTS)
  static constexpr int kOffBar = kOffStage + NT * 8 * 4;
  static constexpr int kSmemBytes = kOffBar + 320;
  static_assert(kSmemBytes <= 232448, "shared memory budget");
  // TMEM columns: D1 blocks at 0, D2 blocks at 4 NT, D3 / D4 (layers 3, 4) reuse D1's
  static constexpr int kColD1 = 0, kColD2 = 4 * NT, kColD3 = 0, kColD4 = NT;
  static constexpr int kTmemCols = 8 * NT;
};

__device__ __forceinline__ void stamp(int slot) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  g_tct_trace[slot] = t;
}
#define TCT_STAMP(cond, slot) do { if (cond) stamp(slot); } while (0)

// Suspend-time hint of every mbarrier wait below.  Measured (OIS_TCT_DEBUG_MODE=4): layer 3's MMAs take 48 cycles each with
// the other warps parked and 108 with 16 epilogue warps re-polling d3_full at the default (short) time limit.
#ifndef OIS_TCT_WAIT_HINT
#define OIS_TCT_WAIT_HINT 20000
#endif
constexpr uint32_t kWaitHintNs = OIS_TCT_WAIT_HINT;
// mbarrier wait that cannot hang the GPU: a protocol bug traps after 2^20 failed waits instead of spinning forever.
// (No printf in the product build: a kernel that contains one pays several microseconds per launch for the print FIFO.)
__device__ __forceinline__ void wait(uint64_t* bar, uint32_t parity) {
  uint32_t polls = 0;
  while (!(kWaitHintNs ? ptx::mbar_try_wait_hint(bar, parity, kWaitHintNs) : ptx::mbar_try_wait(bar, parity))) {
    if (++polls > (1u << 20)) {   // >= 1 s of failed waits (each up to the 20 us hint): a protocol bug, not a slow peer
#ifdef OIS_TCT_DEBUG
      printf("tct_actor_kernel: barrier at shared 0x%x of CTA %d never completed (warp %d)\n", ptx::smem_u32(bar), blockIdx.x,
             threadIdx.x >> 5);
#endif
      __trap();
    }
  }
}

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// ELU(acc + bias) for two envs of one feature -> packed bf16 pair (same arithmetic as mlp_tc.cu's bias_elu_pack)
__device__ __forceinline__ uint32_t bias_elu_pack(float a0, float a1, float2 bias) {
  const float2 x = __fadd2_rn(make_float2(a0, a1), bias);
  const float2 t = __fmul2_rn(x, make_float2(1.4426950408889634f, 1.4426950408889634f));
  const float2 em1 = __fadd2_rn(make_float2(ex2_approx(t.x), ex2_approx(t.y)), make_float2(-1.0f, -1.0f));
  return ptx::pack_bf16(x.x > 0.f ? x.x : em1.x, x.y > 0.f ? x.y : em1.y);
}
__device__ __forceinline__ float sigmoid_fast(float z) { return __fdividef(1.0f, 1.0f + ex2_approx(-1.4426950408889634f * z)); }
__device__ __forceinline__ float tanh_fast(float z) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(z));
  return y;
}

// The env step fused behind the actor (kEnv): the auxiliary warp, which has nothing to do once the tile's observations
// and noise are prepared, loads the state of the tile's envs while the MLP runs and steps them the moment warp 0 has the
// actions — no second kernel, no hand-over through a grid boundary, and no wait for the slowest actor CTA.
struct EnvFuse {
  OisEnvParams p;
  OisEnvState st;
  OisEnvIO io;
  EnvDerived dv;
  int64_t n;
};

struct Args {
  const float* params;        // fp32 master (biases)
  const uint8_t* image;       // tile-blocked bf16 image of the actor (update_tc.cuh: kImgW1..kImgW4)
  const double* mean;         // scaler (nullable)
  const double* var;
  int in_dim, out_dim, out_act;
  int64_t rows;
  OisActIO io;
  int debug;                  // diagnostics (OIS_TCT_DEBUG_MODE): 1 = return right after the set-up
};

// 32 envs of this thread's feature: +bias, ELU, bf16 -> four 16-byte chunks of the env-major operand
__device__ __forceinline__ void act_store32(const uint32_t (&v)[32], float bias, uint8_t* dst, int chunk_stride) {
  const float2 b2 = make_float2(bias, bias);
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    uint32_t w[4];
#pragma unroll
    for (int e = 0; e < 4; ++e)
      w[e] = bias_elu_pack(__uint_as_float(v[8 * c + 2 * e]), __uint_as_float(v[8 * c + 2 * e + 1]), b2);
    *reinterpret_cast<uint4*>(dst + c * chunk_stride) = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

template <int NT, bool kEnv>
__global__ void __launch_bounds__(kThreads, 1)
tct_actor_kernel(const __grid_constant__ Args args, const __grid_constant__ EnvFuse ef) {
  using C = Cfg<NT>;
  constexpr int kSlots = C::kSlots;
  constexpr int kRowsPerLane = NT / 32;        // aux / final warp: envs per lane
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* s_ring = smem + C::kOffRing;
  uint8_t* s_h = smem + C::kOffH;
  uint8_t* s_w1 = smem + C::kOffW1;
  uint8_t* s_w4 = smem + C::kOffW4;
  uint8_t* s_x = smem + C::kOffX;
  float* s_bias = reinterpret_cast<float*>(smem + C::kOffBias);
  float* s_scale = reinterpret_cast<float*>(smem + C::kOffScale);
  float* s_out = reinterpret_cast<float*>(smem + C::kOffOut);
  float* s_noise = reinterpret_cast<float*>(smem + C::kOffNoise);
  float* s_stage = reinterpret_cast<float*>(smem + C::kOffStage);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kOffBar);
  uint64_t* full = bars;                    // [kSlots] ring slot filled (producer tx bytes)
  uint64_t* empty = full + kSlots;          // [kSlots] ring slot consumed (MMA commit)
  uint64_t* wconst_full = empty + kSlots;   // W1 + W4 resident
  uint64_t* x_ready = wconst_full + 1;      // X operand of the tile written (aux warp)
  uint64_t* d1_full = x_ready + 1;          // [4] layer-1 accumulator block complete
  uint64_t* h1_ready = d1_full + 4;         // [4] H1 block written (4 warps)
  uint64_t* d2_full = h1_ready + 4;         // all of layer 2 complete
  uint64_t* h2_ready = d2_full + 1;         // [4]
  uint64_t* d3_full = h2_ready + 4;
  uint64_t* h3_ready = d3_full + 1;         // 16 warps
  uint64_t* d4_full = h3_ready + 1;
  // aux warp: the tile's OU noise is in s_noise[tile parity].  TWO barriers, used alternately: the auxiliary warp runs up
  // to one tile ahead of the warp that consumes the noise at the END of a tile, and a single barrier would then be two
  // phase completions ahead of its waiter — which a parity wait cannot tell from "not yet" (it would block for ever)
  uint64_t* noise_ready = d4_full + 1;      // [2]
  uint64_t* tile_done = noise_ready + 2;    // 16 warps: every TMEM read of the tile is done
  uint64_t* act_ready = tile_done + 1;      // kEnv: warp 0 has left the tile's final actions in s_out
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(act_ready + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_tiles = (int)((args.rows + NT - 1) / NT);
  const bool tr = blockIdx.x == 0 && lane == 0;
  TCT_STAMP(tr && warp == 0, 0);
  // diagnostics: earliest / latest CTA start and latest warp end of the launch (slots 60, 62, 61).  Reductions without a
  // return value, spelled out: as atomicMin / atomicMax the kEnv build compiled them to ATOMG (with a return the warp
  // later waits for), 2560 of them per launch on one address — 2.4 us before warps 0..15 reached the set-up barrier.
  if (threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    asm volatile("red.relaxed.gpu.global.min.u64 [%0], %1;" ::"l"(&g_tct_trace[60]), "l"(t) : "memory");
    asm volatile("red.relaxed.gpu.global.max.u64 [%0], %1;" ::"l"(&g_tct_trace[62]), "l"(t) : "memory");
  }

  // Set-up.  The weight stream starts as early as possible: thread 0 initialises the barriers, warp 0 signals the two
  // producer warps (named barrier 2) and they go; they never join another CTA-wide barrier.  Everybody else meets at
  // named barrier 1 below — including the auxiliary warp, which by then has sat out griddepcontrol.wait (the predecessor
  // has completed) and has the first tile's observations in registers, so the set-up hides the observation load.
  if (threadIdx.x == 0) {
    for (int i = 0; i < kSlots; ++i) { ptx::mbar_init(&full[i], 1); ptx::mbar_init(&empty[i], 1); }
    ptx::mbar_init(wconst_full, 1);
    ptx::mbar_init(x_ready, 1);
    for (int i = 0; i < 4; ++i) { ptx::mbar_init(&d1_full[i], 1); ptx::mbar_init(&h1_ready[i], 4); ptx::mbar_init(&h2_ready[i], 4); }
    ptx::mbar_init(d2_full, 1);
    ptx::mbar_init(d3_full, 1);
    ptx::mbar_init(h3_ready, kEpiWarps);
    ptx::mbar_init(d4_full, 1);
    ptx::mbar_init(&noise_ready[0], 1);
    ptx::mbar_init(&noise_ready[1], 1);
    ptx::mbar_init(tile_done, kEpiWarps);
    ptx::mbar_init(act_ready, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 0) {
    __syncwarp();
    asm volatile("bar.arrive 2, 96;" ::: "memory");
    asm volatile("bar.arrive 3, 64;" ::: "memory");     // ... and the auxiliary warp (its x_ready / noise_ready arrivals)
  }
  if (warp == kWarpProd || warp == kWarpProd + 1) {
    // ===================================================== producers: one 32 KB weight tile per ring slot, the two threads
    // alternate (a thread sustains about one bulk copy per 520 cycles whatever its size: profiles/mma_rate.cu).  The
    // image is older than the previous kernel in the stream (the update's last kernel never triggers its dependents
    // early, agent.cu), so the stream starts before griddepcontrol.wait.
    asm volatile("bar.sync 2, 96;" ::: "memory");
    if (lane == 0 && args.debug != 1) {
      const int p = warp - kWarpProd;
      if (p == 0) {
        ptx::mbar_expect_tx(wconst_full, kH0 * kInPad * 2 + kOutPad * kH2 * 2);
        ptx::bulk_g2s(s_w1, args.image + kImgW1, kH0 * kInPad * 2, wconst_full);
        ptx::bulk_g2s(s_w4, args.image + kImgW4, kOutPad * kH2 * 2, wconst_full);
      }
      uint32_t n = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        for (int s = 0; s < kStagesPerTile; ++s, ++n) {
          if ((int)(n & 1) != p) continue;
          const int slot = n % kSlots;
          // The first two stages go out at once; the rest of the first ring fill waits until the first tile's X operand is
          // written.  All CTAs filling their rings at the same moment is 20 MB through the L2 -> SM fabric (~1.7 us at
          // ~6300 B/clk chip-wide), and the tile's few hundred bytes of observations queue behind it: they are the head of
          // the whole dependent chain, the weights of stage k are not needed before ~0.6 + 0.4 k us after X.
          if (n == 2 || n == 3) wait(x_ready, 0);
          wait(&empty[slot], ((n / kSlots) & 1) ^ 1);
          // W2 tile (mb, kb) in kb-major order; W3 tile kb.  Tile-blocked image: [row block][k chunk][128 rows][8]
          const size_t src = s < 16 ? kImgW2 + (size_t)((s & 3) * 64 + (s >> 2) * 16) * 2048 : kImgW3 + (size_t)((s - 16) * 16) * 2048;
          ptx::mbar_expect_tx(&full[slot], kSlotBytes);
          ptx::bulk_g2s(s_ring + slot * kSlotBytes, args.image + src, kSlotBytes, &full[slot]);
          TCT_STAMP(blockIdx.x == 0 && n >= 12 && n < 20, 40 + n);     // 52..59: stages 12..19 requested
        }
      }
    }
    return;   // every copy issued is consumed by an MMA of this CTA before the other warps leave
  }
  TCT_STAMP(tr && warp == kWarpMma, 37);
  if (warp == kWarpMma) ptx::tmem_alloc<C::kTmemCols>(tmem_slot);
  // Biases, fp32, from the master parameters: REQUESTED here (all loads of a thread in flight at once), consumed after the
  // set-up barrier — nobody waits at the barrier for a cold load.  Thread t's first two values are the layer-1 / layer-2
  // biases of the feature it owns in the epilogue (bv[0], bv[1] stay in registers); layers 3 / 4 read the third through
  // shared memory (written after the barrier; its readers come through later mbarrier waits).
  constexpr int kBiasIters = (kBiasTotal + kEpiWarps * 32 - 1) / (kEpiWarps * 32);
  float bv[kBiasIters] = {};
  if (warp < kEpiWarps) {
    const int64_t b0 = (int64_t)kH0 * args.in_dim, w1 = b0 + kH0, b1 = w1 + (int64_t)kH1 * kH0, w2 = b1 + kH1;
    const int64_t b2 = w2 + (int64_t)kH2 * kH1, w3 = b2 + kH2, b3 = w3 + (int64_t)args.out_dim * kH2;
#pragma unroll
    for (int k = 0; k < kBiasIters; ++k) {
      const int i = threadIdx.x + k * kEpiWarps * 32;
      float v = 0.f;
      if (i < kH0) v = __ldg(args.params + b0 + i);
      else if (i < kH0 + kH1) v = __ldg(args.params + b1 + i - kH0);
      else if (i < kH0 + kH1 + kH2) v = __ldg(args.params + b2 + i - kH0 - kH1);
      else if (i < kBiasTotal && (i - kH0 - kH1 - kH2) < args.out_dim) v = __ldg(args.params + b3 + i - kH0 - kH1 - kH2);
      bv[k] = v;
    }
  }
  // RunningStandardScaler constants (in_dim <= 8): s_scale[0..7] = mean.float(), [8..15] = d = sqrt(var.float()) + eps,
  // [16..23] = RN(1 / d).  The auxiliary warp REQUESTS them here and consumes them only after it has issued the
  // observation copies below: a cold load (0.7 us after an L2 flush) between the kernel's start and the first observation
  // request would sit on the critical path of the whole tile.
  const int sc_kind = lane >> 3, sc_c = lane & 7;
  const bool sc_live = warp == kWarpAux && args.mean != nullptr && sc_kind < 3 && sc_c < args.in_dim;
  double sc_raw = 0.0;
  if (sc_live) sc_raw = sc_kind == 0 ? args.mean[sc_c] : args.var[sc_c];
  // auxiliary warp: the raw observations of its next tile travel global -> shared asynchronously (LDGSTS), each lane its
  // own rows.  (Held in registers across the set-up barrier and the MLP they were spilled by ptxas in the kEnv build, and
  // the spill store waits for the load: 2.4 us of exposed HBM latency before the first MMA.)
  // Layout of the staging buffer: the tile's rows packed, [NT][in_dim].  A full tile is one contiguous, 16-byte aligned
  // block of the observation matrix: it is copied in 16-byte pieces (whole sectors, each read once — it matters when the
  // observations sit in pinned HOST memory and every piece is a PCIe read); a ragged last tile goes element by element.
  const bool x_wide = (reinterpret_cast<uintptr_t>(args.io.states) & 15) == 0;
  auto load_x = [&](int tile) {
    const int64_t r0 = (int64_t)tile * NT;
    if (x_wide && r0 + NT <= args.rows) {
      const float* src = args.io.states + r0 * args.in_dim;
      const int chunks = NT * args.in_dim / 4;
      for (int c = lane; c < chunks; c += 32) ptx::cp_async16(s_stage + 4 * c, src + 4 * c);
    } else {
#pragma unroll
      for (int h = 0; h < kRowsPerLane; ++h) {
        const int64_t r = r0 + lane + 32 * h;
        if (r < args.rows) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (j < args.in_dim) ptx::cp_async4(s_stage + (lane + 32 * h) * args.in_dim + j, args.io.states + r * args.in_dim + j);
        }
      }
    }
  };
  if (warp == kWarpAux) {
    // Programmatic dependent launch: everything else in this set-up overlaps the predecessor's tail; the observations are
    // what the predecessor (the env step) produced.  The next kernel is let go at once: it reads nothing of ours before
    // its own wait.
    TCT_STAMP(tr, 24);
    pdl_wait();
    TCT_STAMP(tr, 25);
    if (lane == 0) pdl_trigger();
    if ((int)blockIdx.x < num_tiles) load_x(blockIdx.x);
    asm volatile("" : "+d"(sc_raw));      // the scaler constants are consumed from here on (after the copies were issued)
    float v = sc_kind == 0 ? 0.f : 1.f;
    if (sc_live && sc_kind == 0) v = (float)sc_raw;
    if (sc_live && sc_kind != 0) {
      v = __fadd_rn(sqrtf((float)sc_raw), 1e-8f);
      if (sc_kind == 2) v = __frcp_rn(v);
    }
    s_scale[lane] = v;
    // this warp does not join the set-up barrier below: it needs neither the TMEM address nor the biases, only the
    // initialised mbarriers (named barrier 3 with warp 0) — X is then written as soon as the observations have landed
    asm volatile("bar.sync 3, 64;" ::: "memory");
  }
  TCT_STAMP(tr && (warp == 0 || warp >= kWarpMma), warp == 0 ? 34 : warp == kWarpMma ? 35 : 36);
  uint32_t tmem = 0;
  if (warp != kWarpAux) {
    ptx::tc_fence_before();
    asm volatile("bar.sync 1, 544;" ::: "memory");      // epilogue warps + the MMA warp
    ptx::tc_fence_after();
    tmem = *tmem_slot;
  }
  TCT_STAMP(tr && warp == 0, 1);
  if (args.debug == 1) {
    if (warp == kWarpMma) ptx::tmem_dealloc<C::kTmemCols>(tmem);
    return;
  }
  if (warp < kEpiWarps) {
#pragma unroll
    for (int k = 2; k < kBiasIters; ++k) {
      const int i = threadIdx.x + k * kEpiWarps * 32;
      if (i < kBiasTotal) s_bias[i] = bv[k];
    }
  }

  if (warp == kWarpMma) {
    // ===================================================== MMA issuer: ONE elected thread runs the whole issue loop.
    // What the issuing thread does BETWEEN two tiles is not hidden by the tensor pipe's queue — measured with the other
    // warps parked (OIS_TCT_DEBUG_MODE, since removed): layer 3's 32 MMAs take 47 cycles each issued back to back and 85 in
    // a loop that per tile re-converged the warp, waited with all 32 lanes and rebuilt sixteen descriptors.  So: no
    // warp-wide step inside a layer (the other 31 lanes wait at a __syncwarp per layer), descriptors by one addition from
    // the tile's first (the address field is the low 14 bits of `lo`; ring and H lie below 256 KB, so no carry).
    const bool leader = ptx::elect_one();
    const bool trl = blockIdx.x == 0;
    const uint32_t id_k = ptx::idesc_bf16(128, NT);                 // A, B K-major (layer 1: B = X)
    const uint32_t id_n = id_k | (1u << 16);                        // B env-major (MN-major)
    constexpr uint32_t hi_a = ptx::smem_desc_hi(128);               // K-major: SBO = next 8 rows
    constexpr uint32_t hi_b512 = ptx::smem_desc_hi(kH0 * 16);       // env-major over 512 features: SBO = next 8 envs
    constexpr uint32_t hi_b128 = ptx::smem_desc_hi(kH2 * 16);
    const uint32_t ring = ptx::smem_u32(s_ring), hb = ptx::smem_u32(s_h);
    const uint32_t x_lo = ptx::smem_desc_lo(ptx::smem_u32(s_x), NT * 16);
    const uint32_t w1a = ptx::smem_u32(s_w1), w4a = ptx::smem_u32(s_w4);
    uint32_t tiles_done = 0, n = 0;
    if (leader) wait(wconst_full, 0);
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tiles_done) {
      const uint32_t par = tiles_done & 1;
      if (leader) {
        if (tiles_done > 0) wait(tile_done, par ^ 1);
        // ---- layer 1: four 128-feature blocks, K = 16 (one MMA each)
        wait(x_ready, par);
        ptx::tc_fence_after();
        TCT_STAMP(trl && tiles_done == 0, 3);
#pragma unroll
        for (int mb = 0; mb < 4; ++mb) {
          ptx::mma_ss(tmem + C::kColD1 + mb * NT, ptx::make_desc(ptx::smem_desc_lo(w1a + mb * 4096, 2048), hi_a),
                      ptx::make_desc(x_lo, hi_a), id_k, false);
          ptx::mma_commit(&d1_full[mb]);
        }
        // ---- layer 2: kb-outer; tile (mb, kb) of W2 from the ring, B = block kb of H1
#pragma unroll 1
        for (int kb = 0; kb < 4; ++kb) {
          wait(&h1_ready[kb], par);
          ptx::tc_fence_after();
          TCT_STAMP(trl && tiles_done == 0, 4 + kb);
          const uint32_t b_lo = ptx::smem_desc_lo(hb + kb * 128 * 16, 128);
#pragma unroll 1
          for (int mb = 0; mb < 4; ++mb, ++n) {
            const int slot = n % kSlots;
            wait(&full[slot], (n / kSlots) & 1);      // (bulk-copy data -> MMA: both the async proxy, no tcgen05 fence)
            const uint32_t a_lo = ptx::smem_desc_lo(ring + slot * kSlotBytes, 2048);
#pragma unroll
            for (int kk = 0; kk < 8; ++kk)
              ptx::mma_ss(tmem + C::kColD2 + mb * NT, ptx::make_desc(a_lo + kk * (4096 >> 4), hi_a),
                          ptx::make_desc(b_lo + kk * (256 >> 4), hi_b512), id_n, (kb | kk) != 0);
            ptx::mma_commit(&empty[slot]);
          }
        }
        ptx::mma_commit(d2_full);
        TCT_STAMP(trl && tiles_done == 0, 8);
        wait(d2_full, par);
      }
      // The epilogue warps do not poll d2_full / d3_full themselves: they sit in a hardware named barrier (5 / 6) and this
      // warp — idle until the next layer's operand is written anyway — waits for the commit and lets them go.
      __syncwarp();
      asm volatile("bar.arrive 5, 544;" ::: "memory");
      if (leader) {
        // ---- layer 3: 128 features, K = 512 in four ring tiles, B = block kb of H2
#pragma unroll 1
        for (int kb = 0; kb < 4; ++kb, ++n) {
          const int slot = n % kSlots;
          wait(&h2_ready[kb], par);
          ptx::tc_fence_after();
          wait(&full[slot], (n / kSlots) & 1);
          const uint32_t a_lo = ptx::smem_desc_lo(ring + slot * kSlotBytes, 2048);
          const uint32_t b_lo = ptx::smem_desc_lo(hb + kb * 128 * 16, 128);
#pragma unroll
          for (int kk = 0; kk < 8; ++kk)
            ptx::mma_ss(tmem + C::kColD3, ptx::make_desc(a_lo + kk * (4096 >> 4), hi_a),
                        ptx::make_desc(b_lo + kk * (256 >> 4), hi_b512), id_n, (kb | kk) != 0);
          ptx::mma_commit(&empty[slot]);
        }
        ptx::mma_commit(d3_full);
        TCT_STAMP(trl && tiles_done == 0, 9);
        wait(d3_full, par);
      }
      __syncwarp();
      asm volatile("bar.arrive 6, 544;" ::: "memory");
      if (leader) {
        // ---- layer 4: W4 [16 (read as 128) rows, 128 k], B = H3 (env-major over 128 features)
        wait(h3_ready, par);
        ptx::tc_fence_after();
        const uint32_t a_lo = ptx::smem_desc_lo(w4a, 256), b_lo = ptx::smem_desc_lo(hb, 128);
#pragma unroll
        for (int kk = 0; kk < 8; ++kk)
          ptx::mma_ss(tmem + C::kColD4, ptx::make_desc(a_lo + kk * (512 >> 4), hi_a),
                      ptx::make_desc(b_lo + kk * (256 >> 4), hi_b128), id_n, kk != 0);
        ptx::mma_commit(d4_full);
        TCT_STAMP(trl && tiles_done == 0, 10);
      }
      __syncwarp();
    }
  } else if (warp == kWarpAux) {
    // ===================================================== auxiliary warp: observations -> X operand, then the tile's OU noise
    const int A = args.out_dim;
    const bool row_noise = args.io.ou_state != nullptr && (A == 1 || A == 2 || A == 4);
    float x[kRowsPerLane][8];
    uint32_t tiles_done = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tiles_done) {
      const uint32_t par = tiles_done & 1;
      TCT_STAMP(tr && tiles_done == 0, 42);
      ptx::cp_async_wait_all();                                // every lane's copies have landed ...
      __syncwarp();                                            // ... before any lane reads the rows others fetched
      TCT_STAMP(tr && tiles_done == 0, 43);
#pragma unroll
      for (int h = 0; h < kRowsPerLane; ++h) {
        const bool live = (int64_t)tile * NT + lane + 32 * h < args.rows;
#pragma unroll
        for (int j = 0; j < 8; ++j) x[h][j] = (live && j < args.in_dim) ? s_stage[(lane + 32 * h) * args.in_dim + j] : 0.f;
      }
      __syncwarp();                                            // all reads done before the next tile's copies are issued
      if (tiles_done > 0) wait(&d1_full[3], par ^ 1);          // the previous tile's layer 1 has read X
#pragma unroll
      for (int h = 0; h < kRowsPerLane; ++h) {
        const int r = lane + 32 * h;
        if (args.mean != nullptr) {
          // (x - mean) / d as q0 = t rc, q = q0 + (t - d q0) rc: the correctly rounded quotient for all but rare operands
          // (<= 1 ulp otherwise), three dependent FMA-pipe instructions instead of the IEEE division's ~12 with a
          // slow-path call — this warp's few hundred instructions run once per launch from a cold instruction cache and
          // head the tile's whole dependent chain.  The result is rounded to bf16 right below.  (Padding columns: mean 0,
          // d = rc = 1.)
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float t = __fsub_rn(x[h][j], s_scale[j]), d = s_scale[8 + j], rc = s_scale[16 + j];
            const float q0 = __fmul_rn(t, rc);
            const float q = fabsf(q0) < 1e30f ? __fmaf_rn(__fmaf_rn(-d, q0, t), rc, q0) : q0;
            x[h][j] = fminf(fmaxf(q, -5.0f), 5.0f);
          }
        }
        *reinterpret_cast<uint4*>(s_x + r * 16) = make_uint4(ptx::pack_bf16(x[h][0], x[h][1]), ptx::pack_bf16(x[h][2], x[h][3]),
                                                             ptx::pack_bf16(x[h][4], x[h][5]), ptx::pack_bf16(x[h][6], x[h][7]));
        *reinterpret_cast<uint4*>(s_x + NT * 16 + r * 16) = make_uint4(0u, 0u, 0u, 0u);     // k 8..15 (in_dim <= 8)
      }
      TCT_STAMP(tr && tiles_done == 0, 44);
      ptx::fence_async_smem();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(x_ready);
      TCT_STAMP(tr && tiles_done == 0, 2);
      if (row_noise) {
        float* sn = s_noise + par * (NT * 4);
#pragma unroll
        for (int h = 0; h < kRowsPerLane; ++h) {
          const int r = lane + 32 * h;
          const int64_t grow = (int64_t)tile * NT + r;
          float nz[4] = {0.f, 0.f, 0.f, 0.f};
          bool applied = false;
          if (grow < args.rows) explore_row_prepare(args.io, grow, A, nz, applied);
          *reinterpret_cast<float4*>(sn + r * 4) = make_float4(nz[0], nz[1], nz[2], nz[3]);
        }
      }
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&noise_ready[par]);
      // next tile's observations: in flight while this tile's MLP runs (kEnv: one tile per CTA — and the observations must
      // not stay live across the env step, where they would be spilled and the spill store would wait for the load)
      if (!kEnv && tile + (int)gridDim.x < num_tiles) load_x(tile + gridDim.x);
      if (kEnv) {
        // ---- the env step of this tile's envs (one tile per CTA: the host guarantees it).  State loads are in flight
        // while the MLP runs (this warp sat out griddepcontrol.wait at the top: the previous step's stores are visible)
        EnvLoaded el[kRowsPerLane];
#pragma unroll
        for (int h = 0; h < kRowsPerLane; ++h) {
          const int64_t i = (int64_t)tile * NT + lane + 32 * h;
          if (i < ef.n) env_load(ef.st, ef.io, i, el[h]);
        }
        // the step-control block of the next step, evaluated while the MLP runs (lane 0; parked in the scaler constants'
        // shared memory, which is dead once X is written: one tile per CTA) — the last CTA only stores it
        const bool advance = ef.io.ctl != nullptr && ef.io.ctl->self_advance;
        StepCtlNext* s_next = reinterpret_cast<StepCtlNext*>(s_scale);
        static_assert(sizeof(StepCtlNext) <= 2 * kInPad * 4, "StepCtlNext is parked in the scaler constants");
        if (advance && lane == 0) *s_next = step_ctl_next(ef.io.ctl);
        TCT_STAMP(tr && tiles_done == 0, 28);
        const float* sa_ = s_out + par * (NT * 8);
        // The wrapper's observations [n, 6] (what the host reads back — pinned HOST memory in the host-in-the-loop mode) of a
        // full tile leave as whole 16-byte pieces of one contiguous block, staged through shared memory, instead of three
        // 8-byte stores per lane at a 24-byte stride (partial sectors: several PCIe writes per sector).
        const int64_t t0 = (int64_t)tile * NT;
        const bool oc_wide = ef.io.obs_clipped != nullptr && t0 + NT <= ef.n && (reinterpret_cast<uintptr_t>(ef.io.obs_clipped) & 15) == 0;
#pragma unroll
        for (int h = 0; h < kRowsPerLane; ++h) {
          const int r = lane + 32 * h;
          const int64_t i = t0 + r;
          if (i < ef.n) {
            float oc[6];
            // the env's arithmetic up to the first use of its actions runs while the MLP is still busy; act_ready is waited
            // for inside (a completed phase: a second wait of the same parity returns at once).  (A dry pass of the same
            // code on made-up actions beforehand, to warm the instruction cache, was measured: no gain — this part is bound
            // by its dependent chains.)
            auto actions = [&]() {
              wait(act_ready, par);
              TCT_STAMP(tr && tiles_done == 0 && h == 0, 29);
              return *reinterpret_cast<const float4*>(sa_ + r * 8);
            };
            env_one_direct(ef.p, ef.st, ef.io, ef.dv, ef.n, i, el[h], actions, oc_wide, oc);
            if (oc_wide) {
#pragma unroll
              for (int k = 0; k < 6; k += 2) *reinterpret_cast<float2*>(s_stage + r * 6 + k) = make_float2(oc[k], oc[k + 1]);
            }
          }
        }
        if (oc_wide) {
          __syncwarp();
          float4* dst = reinterpret_cast<float4*>(ef.io.obs_clipped + t0 * 6);
          for (int c = lane; c < NT * 6 / 4; c += 32) dst[c] = *reinterpret_cast<const float4*>(s_stage + 4 * c);
        }
        TCT_STAMP(tr && tiles_done == 0, 30);
        // The last CTA to finish its envs moves the step-control block on (as the tail of env_step_kernel does).  Every
        // reader of the block in this CTA is done: warp 0 read it before it signalled act_ready, this warp's reads fed the
        // addresses and values of stores it has issued.  With a host waiting (done_flag) the CTA's results — stores to
        // pinned host memory — are made visible system-wide before the ticket; the last CTA, having seen every ticket, then
        // only has to keep its flag store behind that observation (a GPU-scope fence), and stores the block afterwards:
        // the host does not wait for it, the next launch is ordered behind this kernel anyway.
        if (advance) {
          __syncwarp();
          if (lane == 0) {
            OisStepCtl* c = const_cast<OisStepCtl*>(ef.io.ctl);
            if (ef.io.done_flag) __threadfence_system();
            TCT_STAMP(tr, 38);
            if (atomicAdd(&c->ticket, 1u) == gridDim.x - 1) {
              stamp(40);
              const StepCtlNext nx = *s_next;
              if (ef.io.done_flag) {
                __threadfence();
                *reinterpret_cast<volatile uint32_t*>(ef.io.done_flag) = (uint32_t)nx.step;
              }
              c->ticket = 0;
              step_ctl_store(c, nx);
              stamp(41);
            }
            TCT_STAMP(tr, 39);
          }
        }
      }
    }
  } else {
    // ===================================================== epilogue warps: group g = feature block, quadrant = TMEM lane group
    const int g = warp >> 2, q = warp & 3;
    const int f = q * 32 + lane;                       // feature within the 128-block == TMEM lane
    const uint32_t t_lane = tmem + ((uint32_t)(q * 32) << 16);
    auto warp_arrive = [&](uint64_t* bar) {
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(bar);
    };
    // a [128 features, NT envs] accumulator block at TMEM column `col` -> bias, ELU, bf16 -> this thread's chunks of H
    auto block_to_h = [&](int col, float b, uint8_t* hdst) {
      uint32_t v[NT / 32][32];
#pragma unroll
      for (int i = 0; i < NT / 32; ++i) ptx::tmem_ld32(t_lane + col + 32 * i, v[i]);
      ptx::tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < NT / 32; ++i) act_store32(v[i], b, hdst + 4 * i * kH0 * 16, kH0 * 16);
    };
    const int A = args.out_dim;
    const bool row_noise = args.io.ou_state != nullptr && (A == 1 || A == 2 || A == 4);
    uint32_t tiles_done = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tiles_done) {
      const uint32_t par = tiles_done & 1;
      uint8_t* hdst = s_h + (g * 128 + f) * 16;        // chunk n8 of feature (g, f) at + n8 * 512 * 16
      // ---- layer 1 block g -> H1 block g
      wait(&d1_full[g], par);
      ptx::tc_fence_after();
      block_to_h(C::kColD1 + g * NT, bv[0], hdst);
      ptx::fence_async_smem();
      warp_arrive(&h1_ready[g]);
      TCT_STAMP(tr && q == 0 && tiles_done == 0, 12 + g);
      // ---- layer 2 block g -> H2 block g (H1 is dead: every layer-2 MMA has completed)
      asm volatile("bar.sync 5, 544;" ::: "memory");      // d2_full has completed (the MMA warp saw it)
      ptx::tc_fence_after();
      TCT_STAMP(tr && q == 0 && g == 0 && tiles_done == 0, 16);
      block_to_h(C::kColD2 + g * NT, bv[1], hdst);
      ptx::fence_async_smem();
      warp_arrive(&h2_ready[g]);
      TCT_STAMP(tr && q == 0 && tiles_done == 0, 17 + g);
      // ---- layer 3: group g takes envs g NT/4 .. of the 128 features -> H3 (env-major over 128 features)
      {
        asm volatile("bar.sync 6, 544;" ::: "memory");    // d3_full has completed
        ptx::tc_fence_after();
        TCT_STAMP(tr && q == 0 && g == 0 && tiles_done == 0, 21);
        constexpr int kE = NT / 4;                     // envs per group: 16 or 8
        uint32_t v[kE];
        if constexpr (kE == 16) ptx::tmem_ld16(t_lane + C::kColD3 + g * kE, v);
        else ptx::tmem_ld8(t_lane + C::kColD3 + g * kE, v);
        ptx::tmem_ld_wait();
        const float b = s_bias[kH0 + kH1 + f];
        const float2 b2 = make_float2(b, b);
#pragma unroll
        for (int c = 0; c < kE / 8; ++c) {
          uint32_t w[4];
#pragma unroll
          for (int e = 0; e < 4; ++e)
            w[e] = bias_elu_pack(__uint_as_float(v[8 * c + 2 * e]), __uint_as_float(v[8 * c + 2 * e + 1]), b2);
          *reinterpret_cast<uint4*>(s_h + ((g * (kE / 8) + c) * kH2 + f) * 16) = make_uint4(w[0], w[1], w[2], w[3]);
        }
        ptx::fence_async_smem();
        warp_arrive(h3_ready);
      }
      // ---- layer 4 -> output activation -> exploration -> actions: warp 0 (TMEM lanes 0..A-1 hold the outputs)
      if (warp == 0) {
        wait(d4_full, par);
        ptx::tc_fence_after();
        TCT_STAMP(tr && tiles_done == 0, 22);
        float* so = s_out + par * (NT * 8);
        {
          uint32_t y[NT / 32][32];
#pragma unroll
          for (int i = 0; i < NT / 32; ++i) ptx::tmem_ld32(t_lane + C::kColD4 + 32 * i, y[i]);
          ptx::tmem_ld_wait();
          if (lane < 8) {
#pragma unroll
            for (int i = 0; i < NT / 32; ++i)
#pragma unroll
              for (int e = 0; e < 32; ++e) so[(e + 32 * i) * 8 + lane] = __uint_as_float(y[i][e]);
          }
        }
        __syncwarp();
        wait(&noise_ready[par], (tiles_done >> 1) & 1);
        const float* b4 = s_bias + kH0 + kH1 + kH2;
        const float* sn = s_noise + par * (NT * 4);
        const bool applied = act_noise_scale(args.io) >= 0.f;
        float mx = -INFINITY, mn = INFINITY, sm = 0.f;
#pragma unroll
        for (int h = 0; h < kRowsPerLane; ++h) {
          const int r = lane + 32 * h;
          const int64_t grow = (int64_t)tile * NT + r;
          if (grow < args.rows) {
            float a[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              a[j] = 0.f;
              if (j < A) {      // (warp-uniform: the outputs beyond act_dim are never stored)
                const float z = so[r * 8 + j] + b4[j];
                a[j] = args.out_act == OIS_ACT_SIGMOID ? sigmoid_fast(z) : args.out_act == OIS_ACT_TANH ? tanh_fast(z) : z;
              }
            }
            if (row_noise) {
              if (applied) {
                const float4 n4 = *reinterpret_cast<const float4*>(sn + r * 4);
                const float nz[4] = {n4.x, n4.y, n4.z, n4.w};
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
            if (kEnv) *reinterpret_cast<float4*>(so + r * 8) = make_float4(a[0], A > 1 ? a[1] : 0.f, A > 2 ? a[2] : 0.f, A > 3 ? a[3] : 0.f);
            float* out = args.io.actions + grow * A;
            if (A == 4) *reinterpret_cast<float4*>(out) = make_float4(a[0], a[1], a[2], a[3]);
            else if (A == 2) *reinterpret_cast<float2*>(out) = make_float2(a[0], a[1]);
            else {
#pragma unroll
              for (int j = 0; j < 8; ++j) if (j < A) out[j] = a[j];
            }
          }
        }
        if (kEnv) {
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(act_ready);
        }
        if (args.io.noise_stats && args.io.ou_state && applied) noise_stats_warp(args.io, mx, mn, sm);
        TCT_STAMP(tr && tiles_done == 0, 23);
      }
      ptx::tc_fence_before();
      warp_arrive(tile_done);
    }
  }
  // ---- teardown (the producers have left; their last copy was consumed by the last layer-3 MMA)
  ptx::tc_fence_before();
  asm volatile("bar.sync 4, 576;" ::: "memory");
  if (warp == kWarpMma) ptx::tmem_dealloc<C::kTmemCols>(tmem);
  if (lane == 0) {     // diagnostics: every warp's exit time
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    asm volatile("red.relaxed.gpu.global.max.u64 [%0], %1;" ::"l"(&g_tct_last_end), "l"(t) : "memory");
  }
}

}  // namespace tct

template <int NT, bool kEnv>
static cudaError_t launch(const tct::Args& args, const tct::EnvFuse& ef, int64_t rows, cudaStream_t s) {
  using namespace tct;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(tct_actor_kernel<NT, kEnv>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<NT>::kSmemBytes);
    if (e != cudaSuccess) return e;
    attr = true;
  }
  const int tiles = (int)((rows + NT - 1) / NT);
  const int sms = num_sms();
  return launch_pdl(tct_actor_kernel<NT, kEnv>, tiles < sms ? tiles : sms, kThreads, Cfg<NT>::kSmemBytes, s, args, ef);
}

static long long g_tct_launches = 0;     // diagnostics: launches of the transposed kernel (tests check which kernel served a call)
static long long g_tct_fused_launches = 0;

static int fill_args(OisAgent* a, const uint8_t* image, const float* actor_params, const OisScaler* scaler, const OisActIO* io,
                     int64_t rows, tct::Args& args) {
  using namespace tct;
  OIS_CHECK_ARG(a->actor.h0 == kH0 && a->actor.h1 == kH1 && a->actor.h2 == kH2, "the tcgen05 actor is built for 512-512-128 hidden layers");
  OIS_CHECK_ARG(a->actor.in_dim <= 8 && a->actor.out_dim <= 8, "actor in_dim <= 8 and out_dim <= 8");
  OIS_CHECK_ARG(image, "no bf16 image of the actor weights");
  args.params = actor_params;
  args.image = image;
  args.mean = scaler ? scaler->running_mean : nullptr;
  args.var = scaler ? scaler->running_variance : nullptr;
  args.in_dim = a->actor.in_dim;
  args.out_dim = a->actor.out_dim;
  args.out_act = a->actor.out_act;
  args.rows = rows;
  args.io = *io;
  static const int debug = getenv("OIS_TCT_DEBUG_MODE") ? atoi(getenv("OIS_TCT_DEBUG_MODE")) : 0;
  args.debug = debug;
  return OIS_OK;
}

int tct_actor_forward(OisAgent* a, const uint8_t* image, const float* actor_params, const OisScaler* scaler, const OisActIO* io,
                      int64_t rows, cudaStream_t s) {
  using namespace tct;
  Args args;
  if (int e = fill_args(a, image, actor_params, scaler, io, rows, args)) return e;
  static const EnvFuse none = {};
  // 32-env tiles while they still fit one wave of the SMs (twice the CTAs, half the epilogue per thread), else 64
  static const int force_nt = getenv("OIS_TCT_NT") ? atoi(getenv("OIS_TCT_NT")) : 0;
  const bool nt32 = force_nt ? force_nt == 32 : (rows + 31) / 32 <= num_sms();
  OIS_CHECK_CUDA((nt32 ? launch<32, false>(args, none, rows, s) : launch<64, false>(args, none, rows, s)));
  ++g_tct_launches;
  return OIS_OK;
}

// rows that the fused policy step (actor + env step in one launch) serves: one 32-env tile per CTA, one wave
bool tct_can_fuse_env(int64_t rows) { return rows > 0 && (rows + 31) / 32 <= num_sms(); }

int tct_rollout_forward(OisAgent* a, const uint8_t* image, const float* actor_params, const OisScaler* scaler, const OisActIO* io,
                        const OisEnvParams* ep, const OisEnvState* es, const OisEnvIO* eio, int64_t rows, cudaStream_t s) {
  using namespace tct;
  OIS_CHECK_ARG(tct_can_fuse_env(rows), "the fused policy step needs one 32-env tile per CTA");
  OIS_CHECK_ARG(a->actor.out_dim <= 4 && a->actor.out_dim == ep->act_dim, "actor out_dim must be the env's act_dim (<= 4)");
  Args args;
  if (int e = fill_args(a, image, actor_params, scaler, io, rows, args)) return e;
  EnvFuse ef;
  ef.p = *ep;
  ef.st = *es;
  ef.io = *eio;
  ef.dv = make_env_derived(ep);
  ef.n = rows;
  OIS_CHECK_CUDA((launch<32, true>(args, ef, rows, s)));
  ++g_tct_launches;
  ++g_tct_fused_launches;
  return OIS_OK;
}

}  // namespace ois

// diagnostics hook for profiles/tct_trace.py (deliberately not declared in include/ois_b200.h)
extern "C" long long ois_debug_tct_launches() { return ois::g_tct_launches; }
extern "C" long long ois_debug_tct_fused_launches() { return ois::g_tct_fused_launches; }
extern "C" int64_t ois_rollout_fused_launches(void) { return (int64_t)ois::g_tct_fused_launches; }
extern "C" int ois_debug_tct_trace_reset() {
  unsigned long long z[64] = {};
  z[60] = ~0ull;
  if (cudaMemcpyToSymbol(ois::tct::g_tct_last_end, z, sizeof(unsigned long long)) != cudaSuccess) return -2;
  return cudaMemcpyToSymbol(ois::tct::g_tct_trace, z, sizeof(z)) == cudaSuccess ? 0 : -2;
}
extern "C" int ois_debug_tct_trace(unsigned long long* out_host, int n) {
  if (n > 64) n = 64;
  if (cudaMemcpyFromSymbol(out_host, ois::tct::g_tct_trace, sizeof(unsigned long long) * n) != cudaSuccess) return -2;
  if (n > 61 && cudaMemcpyFromSymbol(out_host + 61, ois::tct::g_tct_last_end, sizeof(unsigned long long)) != cudaSuccess) return -2;
  return 0;
}
