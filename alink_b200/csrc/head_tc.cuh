// head_tc.cuh — BF16_TC siamese head on the 5th-gen tensor cores (tcgen05, accumulators in TMEM).
//
// Replaces, for all K learners of a committee at once, the Keras graph of
// /root/reference/code/siamese.py:24-33 (|l-r| -> Dense(512,relu) -> Dense(64,relu) -> Dense(2)+softmax),
// siamese3.py:18-26 (-> Dense(1,sigmoid)), the per-learner loop + mean of Bagging.predict
// (code/committee.py:13-20) and the utility measures of code/uncertainty.py:63-130, in ONE persistent
// kernel: index-driven gather -> |a-b| (bf16, swizzled shared memory) -> tcgen05 layer 1 -> fused
// bias/relu/bf16 epilogue -> tcgen05 layer 2 -> fused layer 3 + softmax/sigmoid + K-learner
// consensus/vote/entropy epilogue -> probabilities [N][K][C] and the utility [N].
//
// One CTA per SM, one 128-pair tile at a time, 512 threads in five roles:
//   warp 0     : weight producer — streams pre-swizzled bf16 weight images (16 KB per stage) from L2 with
//                cp.async.bulk (TMA bulk copy) into a 4-stage ring; completion on mbarriers.
//   warp 1     : MMA issuer — one thread issues every tcgen05.mma (M=128, N=128 layer 1 / N=H2 layer 2,
//                K=16), commits completion to mbarriers (tcgen05.commit).
//   warp 2     : TMEM allocator.   warp 3 : L2 prefetcher of the NEXT tile's table rows.
//   warps 4-7  : epilogue — tcgen05.ld accumulator -> +bias, relu -> bf16 -> swizzled shared (A operand of
//                layer 2); after a learner's last chunk: layer-2 accumulator -> +bias, relu -> Dense(H2,C)
//                in fp32 -> activation (include/alink_math.h, same arithmetic as the oracle) -> registers;
//                after the last learner: Bagging mean + utility -> global.
//   warps 8-15 : gather producers — rows ia[n] / ib[n] of the embedding tables with 128-bit loads,
//                |a-b| (or a*b for the opt-in Hadamard half), bf16, 128B-swizzled K-major tile X.
// The feature tile X (128 x D_in bf16, <= 128 KB) is learner-independent: it stays resident in shared
// memory while the K learners x (H1/128) weight chunks stream past it. Work items q = (tile, learner,
// chunk) are software-pipelined on the tensor pipe as L1(q), L2(q-1), with two layer-1 accumulators in
// TMEM (2 x 128 columns) + one layer-2 accumulator (H2 columns).
//
// Rooflines: tensor-bound — 2*K*(D_in*H1 + H1*H2) flop per pair on the tensor pipe; the weight stream is
// K*(D_in*H1 + H1*H2)*2 bytes of L2->SMEM traffic per tile; HBM: 2*D*s + 8 bytes gathered per pair +
// 4*K*C + 4 written.
#pragma once
#include "common.cuh"
#include "ptx.cuh"
#include <stdlib.h>

namespace tc {

constexpr int TILE_M = 128;
constexpr int CHUNK_N = 128;
constexpr int KBLK = 64;            // bf16 per 128-byte swizzle row
constexpr int STAGE_BYTES = 16384;  // one [128 x 64] bf16 tile
constexpr int NSTAGE = 4;
constexpr int MAX_KB = 8;           // D_in <= 512
constexpr int NTHREADS = 512;

// dynamic shared memory map (bytes)
constexpr uint32_t XOFF = 0;                                  // feature tile: MAX_KB x 16 KB
constexpr uint32_t HOFF = XOFF + MAX_KB * STAGE_BYTES;        // hidden-1 chunk: 2 x 16 KB
constexpr uint32_t ROFF = HOFF + 2 * STAGE_BYTES;             // weight ring: NSTAGE x 16 KB
constexpr uint32_t POFF = ROFF + NSTAGE * STAGE_BYTES;        // staged biases / last layer
constexpr uint32_t P_B1 = POFF;                               // float [2][128]
constexpr uint32_t P_B2 = P_B1 + 2 * 128 * 4;                 // float [2][64]
constexpr uint32_t P_W3 = P_B2 + 2 * 64 * 4;                  // float [2][128]
constexpr uint32_t P_B3 = P_W3 + 2 * 128 * 4;                 // float [2][2]
constexpr uint32_t BOFF = P_B3 + 16;                          // mbarriers
enum { W_FULL = 0, W_EMPTY = 4, X_FULL = 8, X_EMPTY = 16, A1_FULL = 24, A1_EMPTY = 26, H1_FULL = 28, H1_EMPTY = 30,
       A2_FULL = 32, A2_EMPTY = 34, S_FULL = 36, S_EMPTY = 38, NBARS = 40 };  // H1_*: two in ATM mode (the hidden chunk is double-buffered in TMEM), the first one otherwise  // A2_*: one per layer-2 accumulator; S_*: staged feature tiles (2 buffers)
constexpr uint32_t TOFF = BOFF + NBARS * 8;                   // TMEM base address
constexpr uint32_t SMEM_BYTES = TOFF + 16;
static_assert(SMEM_BYTES <= 232448, "exceeds the 227 KB per-CTA shared memory limit");
static_assert(BOFF % 8 == 0, "mbarriers need 8-byte alignment");

// Pair mode adds the layer-1 bias ON THE TENSOR PIPE: one extra K = 16 MMA per item, A = a constant tile whose
// first three k-columns are 1 (256 bytes of shared memory: every 8-row group aliases the same rows), B = the
// bias split into three operand-format parts (hi + mid + lo = the fp32 value). The epilogue — the critical
// path of the kernel — then only clamps, rounds and packs. The B tile (64 n x 16 k per CTA, no swizzle:
// core matrices 128 B apart along k, 256 B apart along n) travels in the ring slot of the layer-2 slice.
constexpr uint32_t BIAS_TILE = 2048, BIAS_OFF = 8192, BIAS_LBO = 128, BIAS_SBO = 256;
constexpr uint32_t P_ONES = P_B1;  // the constant A tile replaces the staged bias in shared memory

constexpr uint32_t TMEM_COLS = 512;
constexpr uint32_t ACC2_COL = 256;  // two layer-2 accumulators of 64 columns (consecutive learners alternate)
constexpr uint32_t HTM_COL = 384;   // ATM: the hidden chunk as layer 2's A operand IN TENSOR MEMORY: 128 halves = 64 columns, x 2 buffers

struct Args {
  const void *ta, *tb;       // embedding tables (fp32 or bf16), D columns
  const int32_t *ia, *ib;    // pair -> row (nullptr = identity, or the all-pairs enumeration when G > 0)
  uint32_t G;                // all-pairs: gallery rows; pair p = (probe p / G, gallery p % G)
  int64_t pair0;             // all-pairs: enumeration index of pair 0
  int64_t N;
  int D, nkb, hadamard;
  int K, nc, H1, H2, C, act, measure;
  const uint8_t *w1img;      // [K][nc][nkb] blocks of 16 KB: [128 n x 64 k] bf16, SWIZZLE_128B image
  const uint8_t *w2img;      // [K][nc] blocks of 2*H2*128 B: 2 x [H2 n x 64 k]
  const float *b1, *b2, *W3, *b3;  // fp32 [K][H1], [K][H2], [K][H2][C], [K][C]
  float *probs;              // [N][K][C] or nullptr
  float *util;               // [N] or nullptr
  int ntiles;
  const uint8_t *bimg;       // pair mode: [K][nc][2 ranks] layer-1 bias tiles of 2 KB (BIAS_TILE below)
  uint8_t *xscratch;         // STAGED: [gridDim.x][2][MAX_KB] feature-tile images of 16 KB (L2-resident scratch)
  float *vmin;               // [N] or nullptr: min over learners of |p1 - p0| — how far the pair is from a flipped vote
                             // (the vote-entropy filter of FP32_REFINE, refine.cuh)
  // debug build only (-DALINK_TC_DEBUG): bits 0-5 slow down one role per loop iteration by (dbg >> 8) ns —
  // 1 weight producer, 2 MMA issuer, 8 epilogue, 16 gather producers, 32 peer relay, 128 L2 prefetcher;
  // bit 6 publishes per-role progress counters in the watchdog buffer (ptx.cuh)
  uint32_t dbg;
};

// table rows of pair n: explicit indices, the all-pairs enumeration, or identity
__device__ __forceinline__ void pair_rows(const Args &a, int64_t n, int64_t &ra, int64_t &rb) {
  if (a.ia) {
    ra = (int64_t)__ldg(a.ia + n);
    rb = (int64_t)__ldg(a.ib + n);
  } else if (a.G) {
    const int64_t p = a.pair0 + n;
    ra = p / a.G;
    rb = p - ra * a.G;
  } else {
    ra = rb = n;
  }
}

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t *>(&h);
}
__device__ __forceinline__ uint32_t pack_f16(float lo, float hi) {
  __half2 h = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<uint32_t *>(&h);
}
// operand format of the tensor-core path: bf16 (BF16_TC, and the fallback filter of FP32_REFINE) or IEEE
// half (the default FP32_REFINE filter: 8x smaller rounding error, so 8x fewer candidates to re-score)
template <bool F16>
__device__ __forceinline__ uint32_t pack16(float lo, float hi) {
  return F16 ? pack_f16(lo, hi) : pack_bf16(lo, hi);
}
// relu, round to the operand format and pack two values in one instruction
template <bool F16>
__device__ __forceinline__ uint32_t pack16_relu(float lo, float hi) {
  uint32_t d;
  if (F16) asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  else asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}
__device__ __forceinline__ void sts64(uint32_t addr, uint32_t x, uint32_t y) {
  asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ float4 ldg_f4(const float *p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ float4 ldg_bf4(const __nv_bfloat16 *p) {  // 4 bf16 -> 4 fp32
  uint32_t lo, hi;
  asm volatile("ld.global.nc.L1::no_allocate.v2.b32 {%0, %1}, [%2];" : "=r"(lo), "=r"(hi) : "l"(p));
  float4 v;
  v.x = __uint_as_float(lo << 16);
  v.y = __uint_as_float(lo & 0xFFFF0000u);
  v.z = __uint_as_float(hi << 16);
  v.w = __uint_as_float(hi & 0xFFFF0000u);
  return v;
}

// NCTA = 1: every CTA is independent (tcgen05 cta_group::1, M = 128).
// NCTA = 2: CTA pairs (cluster of 2, tcgen05 cta_group::2, M = 256): each CTA keeps its own 128-pair tile,
//           loads HALF of every weight block, and the leader's MMAs read both halves — half the TMA
//           traffic and half the B-operand shared-memory reads per CTA.
// STAGED: the gather producers run up to two tiles AHEAD of the MMAs: they write the converted feature tile
//           (the exact swizzled shared-memory image) to a per-CTA scratch buffer in global memory, which stays in
//           L2, and warp 3 refills the shared-memory tile from it with one 16 KB bulk copy per k-block as soon as
//           the last MMA of the previous tile has read that block. Without it the refill at a tile boundary is
//           four dependent rounds of table loads (only two k-blocks of fp32 rows fit in the producers' registers)
//           during which the tensor pipe idles — 22 % of the kernel in the profile of the unstaged version.
// ATM:    the epilogue writes the relu'd hidden chunk straight into tensor memory (tcgen05.st) and layer 2 reads its A
//           operand from there (tcgen05.mma with a TMEM A address) instead of a swizzled shared-memory tile: the
//           shared-memory data pipe — 85 % busy feeding operands in the profile of round 1 — loses the 32 KB store
//           and the 32 KB re-read per work item.
template <bool TAB_BF16, int NCTA, bool F16, bool STAGED = false, bool ATM = false>
__global__ void __launch_bounds__(NTHREADS, 1) head_kernel(const Args a) {
  constexpr uint32_t KPS = NCTA;                // k-blocks per ring slot (a pair CTA holds half of each block)
  constexpr uint32_t HB = STAGE_BYTES / NCTA;   // bytes of this CTA's part of one [128 n x 64 k] weight block
  constexpr uint32_t NS = NSTAGE;               // ring slots of STAGE_BYTES
  constexpr bool BMMA = NCTA == 2;              // layer-1 bias added by an MMA (see BIAS_TILE)
  const uint32_t rank = NCTA == 2 ? ptx::cluster_ctarank() : 0u;
  const uint32_t unit0 = blockIdx.x / NCTA, nunit = gridDim.x / NCTA;  // tile (pair) scheduling units
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sbase = ptx::smem_u32(smem);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  auto bar = [&](int i) -> uint32_t { return sbase + BOFF + 8u * (uint32_t)i; };
#ifdef ALINK_TC_DEBUG
  // the first two clusters publish each role's loop counter in the host-mapped watchdog buffer
  // (words 128 + 16 * block + role), so a deadlock report also says how far everyone got
  auto progress = [&](uint32_t role, uint32_t v) {
    if ((a.dbg & 64u) && blockIdx.x < 4 && (threadIdx.x & 31) == 0 && ptx::g_alink_wd)
      reinterpret_cast<volatile unsigned long long *>(ptx::g_alink_wd)[128 + 16 * blockIdx.x + role] = v + 1;
  };
#define ALINK_DBG_ROLE(bit, role, v)                      \
  do {                                                    \
    if (a.dbg & (bit)) __nanosleep(a.dbg >> 8);           \
    progress((role), (v));                                \
  } while (0)
#else
#define ALINK_DBG_ROLE(bit, role, v) do { } while (0)
#endif

  if (tid == 0) {
    if (sbase & 1023u) __trap();  // the swizzled operand tiles need a 1024-byte aligned base
    // The leader CTA of a pair also needs the peer's operands: the peer's relay warp forwards one
    // arrive per phase onto the leader's "operand ready" / "buffer free" barriers (count + 1).
    const uint32_t fwd = (NCTA == 2 && rank == 0) ? 1u : 0u;
    for (int s = 0; s < NSTAGE; ++s) {
      ptx::mbar_init(bar(W_FULL + s), 1 + fwd);
      ptx::mbar_init(bar(W_EMPTY + s), 1);
    }
    for (int k = 0; k < MAX_KB; ++k) {
      ptx::mbar_init(bar(X_FULL + k), (STAGED ? 1 : 256) + fwd);
      ptx::mbar_init(bar(X_EMPTY + k), 1);
    }
    for (int b = 0; b < 2; ++b) {
      ptx::mbar_init(bar(A1_FULL + b), 1);
      ptx::mbar_init(bar(A1_EMPTY + b), 128 + fwd);
    }
    for (int b = 0; b < 2; ++b) {
      ptx::mbar_init(bar(H1_FULL + b), 128 + fwd);
      ptx::mbar_init(bar(H1_EMPTY + b), 1);
    }
    ptx::mbar_init(bar(A2_FULL), 1);
    ptx::mbar_init(bar(A2_FULL + 1), 1);
    ptx::mbar_init(bar(A2_EMPTY), 128 + fwd);
    ptx::mbar_init(bar(A2_EMPTY + 1), 128 + fwd);
    for (int b = 0; b < 2; ++b) {
      ptx::mbar_init(bar(S_FULL + b), 256);
      ptx::mbar_init(bar(S_EMPTY + b), 1);
    }
    ptx::fence_barrier_init();
  }
  if (BMMA && warp == 3 && lane < 16) {
    // constant A tile: 8 rows x 16 B with k = 0, 1, 2 set to 1.0, then 8 rows x 16 B of zeros (k = 8..15)
    const uint32_t one = F16 ? 0x3C00u : 0x3F80u;
    if (lane < 8) sts128(sbase + P_ONES + lane * 16u, one | (one << 16), one, 0u, 0u);
    else sts128(sbase + P_ONES + lane * 16u, 0u, 0u, 0u, 0u);
    ptx::fence_proxy_async_smem();
  }
  if (warp == 2) {
    if (NCTA == 2) ptx::tmem_alloc2(sbase + TOFF, TMEM_COLS);
    else ptx::tmem_alloc(sbase + TOFF, TMEM_COLS);
  }
  ptx::tc_fence_before();
  if (NCTA == 2) ptx::cluster_sync();
  else __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = __shfl_sync(0xffffffffu, *reinterpret_cast<volatile uint32_t *>(smem + TOFF), 0);

  const uint32_t G = (uint32_t)(a.K * a.nc);  // work items per tile
  const uint32_t nunits = ((uint32_t)a.ntiles + NCTA - 1) / NCTA;
  const uint32_t T = nunits > unit0 ? (nunits - unit0 + nunit - 1) / nunit : 0u;
  auto tile_of = [&](uint32_t i) -> int64_t { return ((int64_t)unit0 + (int64_t)i * nunit) * NCTA + rank; };
  const uint32_t Q = T * G;
  const uint32_t nkb = (uint32_t)a.nkb;
  const uint32_t l2bytes = 2u * (uint32_t)a.H2 * 128u;

  // Register budget per warpgroup: light control warps and the epilogue give registers to the gather
  // producers, which keep two k-blocks of loads in flight (56*128 + 112*128 + 168*256 <= 64 K).
  // (each role branch starts with its warpgroup's setmaxnreg so that ptxas budgets the branch with it)

  if (warp == 0) {
    // ------------------------------------------------------------------ weight producer (TMA bulk)
    // The whole warp runs the loop converged; one elected lane issues the copies. One ring slot holds
    // KPS k-blocks of this CTA's rows: all 128 of each block, or the half [64 * rank, +64) in pair mode.
    ptx::setmaxnreg_dec<56>();
    const uint32_t leader = ptx::elect_one();
    const uint32_t ring = sbase + ROFF;
    uint32_t s = 0, ph = 0;  // ring slot and its phase parity
    for (uint32_t q = 0; q <= Q; ++q) {
      ALINK_DBG_ROLE(1u, 0, q);
      if (q < Q) {
        const uint8_t *src = a.w1img + (size_t)(q % G) * nkb * STAGE_BYTES + rank * HB;
        for (uint32_t kb = 0; kb < nkb; kb += KPS) {
          ptx::mbar_wait(bar(W_EMPTY + s), ph ^ 1u);
          if (leader) {
            ptx::mbar_expect_tx(bar(W_FULL + s), KPS * HB);
#pragma unroll
            for (uint32_t kk = 0; kk < KPS; ++kk)
              ptx::bulk_g2s(ring + s * STAGE_BYTES + kk * HB, src + (size_t)(kb + kk) * STAGE_BYTES, HB, bar(W_FULL + s));
          }
          __syncwarp();
          if (++s == NS) { s = 0; ph ^= 1u; }
        }
      }
      if (BMMA || q >= 1) {
        // layer-2 slice of item q - 1: 2 k-blocks of [H2 n x 64 k]; a pair CTA takes rows [H2/2 * rank, +H2/2)
        // of each. In pair mode the slot also carries the layer-1 bias tile of item q.
        const uint8_t *src = a.w2img + (size_t)((q ? q - 1 : 0u) % G) * l2bytes;
        ptx::mbar_wait(bar(W_EMPTY + s), ph ^ 1u);
        if (leader) {
          if (NCTA == 1) {
            ptx::mbar_expect_tx(bar(W_FULL + s), l2bytes);
            ptx::bulk_g2s(ring + s * STAGE_BYTES, src, l2bytes, bar(W_FULL + s));
          } else {
            const uint32_t piece = l2bytes / 4;  // H2/2 rows x 128 B
            ptx::mbar_expect_tx(bar(W_FULL + s), (q >= 1 ? 2u * piece : 0u) + (q < Q ? BIAS_TILE : 0u));
            if (q >= 1) {
              ptx::bulk_g2s(ring + s * STAGE_BYTES, src + rank * piece, piece, bar(W_FULL + s));
              ptx::bulk_g2s(ring + s * STAGE_BYTES + piece, src + 2 * piece + rank * piece, piece, bar(W_FULL + s));
            }
            if (q < Q)
              ptx::bulk_g2s(ring + s * STAGE_BYTES + BIAS_OFF, a.bimg + ((size_t)(q % G) * 2u + rank) * BIAS_TILE,
                            BIAS_TILE, bar(W_FULL + s));
          }
        }
        __syncwarp();
        if (++s == NS) { s = 0; ph ^= 1u; }
      }
    }
  } else if (warp == 1 && rank == 0) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA of a pair)
    // Converged warp, uniform control flow; one elected lane issues tcgen05.mma / tcgen05.commit.
    ptx::setmaxnreg_dec<56>();
    const uint32_t leader = ptx::elect_one();
    const uint32_t idesc1 = F16 ? ptx::umma_idesc_f16(TILE_M * NCTA, CHUNK_N) : ptx::umma_idesc_bf16(TILE_M * NCTA, CHUNK_N);
    const uint32_t idesc2 = F16 ? ptx::umma_idesc_f16(TILE_M * NCTA, a.H2) : ptx::umma_idesc_bf16(TILE_M * NCTA, a.H2);
    const uint32_t kb2_step = ((uint32_t)a.H2 * 128u / NCTA) >> 4;  // layer-2 B sub-block stride, 16-byte units
    constexpr uint64_t STAGE_STEP = STAGE_BYTES >> 4;
    constexpr uint64_t HB_STEP = HB >> 4;
    const uint64_t xdesc0 = ptx::umma_desc_sw128(sbase + XOFF);
    const uint64_t hdesc0 = ptx::umma_desc_sw128(sbase + HOFF);
    const uint64_t rdesc0 = ptx::umma_desc_sw128(sbase + ROFF);
    const uint32_t acc2_0 = tmem + ACC2_COL;
    const uint64_t onesdesc = ptx::umma_desc_nosw(sbase + P_ONES, 128u, 0u);
    auto mma = [&](uint32_t d, uint64_t ad, uint64_t bd, uint32_t idesc, uint32_t acc) {
      if (NCTA == 2) ptx::umma2_bf16(d, ad, bd, idesc, acc);
      else ptx::umma_bf16(d, ad, bd, idesc, acc);
    };
    auto commit = [&](int b_) {
      if (NCTA == 2) ptx::umma2_commit_mc(bar(b_));  // arrives in both CTAs of the pair
      else ptx::umma_commit(bar(b_));
    };
    uint32_t s = 0, ph = 0;
    // work-item coordinates as running counters (no integer division in the issue loop — this one warp's
    // instruction stream is the critical path of the kernel): g1 / xpar for layer 1 of item q,
    // c2 / v2 (= tile * K + learner, mod 4: accumulator v2 & 1, its use count parity (v2 >> 1) & 1) for layer 2
    // of item q - 1
    uint32_t g1 = 0, xpar = 0, c2 = 0, v2 = 0;
    for (uint32_t q = 0; q <= Q; ++q) {
      ALINK_DBG_ROLE(2u, 1, q);
      if (q < Q) {
        const uint32_t g = g1, b = q & 1u, u = q >> 1;
        ptx::mbar_wait(bar(A1_EMPTY + b), (u & 1u) ^ 1u);
        const uint32_t d_tmem = tmem + b * CHUNK_N;
        for (uint32_t kb = 0; kb < nkb; kb += KPS) {
          if (g == 0) {
#pragma unroll
            for (uint32_t kk = 0; kk < KPS; ++kk) ptx::mbar_wait(bar(X_FULL + kb + kk), xpar);
          }
          ptx::mbar_wait(bar(W_FULL + s), ph);
          ptx::tc_fence_after();
          if (leader) {
            const uint64_t ad = xdesc0 + kb * STAGE_STEP;
            const uint64_t bd = rdesc0 + s * STAGE_STEP;
#pragma unroll
            for (uint32_t kk = 0; kk < KPS; ++kk) {
#pragma unroll
              for (uint32_t k = 0; k < 4; ++k)
                mma(d_tmem, ad + kk * STAGE_STEP + 2 * k, bd + kk * HB_STEP + 2 * k, idesc1, (kb | kk | k) != 0u);
            }
            commit(W_EMPTY + s);
            if (g == G - 1) {  // last reader of this tile's X blocks
#pragma unroll
              for (uint32_t kk = 0; kk < KPS; ++kk) commit(X_EMPTY + kb + kk);
            }
          }
          __syncwarp();
          if (++s == NS) { s = 0; ph ^= 1u; }
        }
        if (BMMA) {
          // + bias: the tile rides in the next ring slot (the one of the layer-2 slice below)
          ptx::mbar_wait(bar(W_FULL + s), ph);
          ptx::tc_fence_after();
          if (leader)
            mma(d_tmem, onesdesc, ptx::umma_desc_nosw(sbase + ROFF + s * STAGE_BYTES + BIAS_OFF, BIAS_LBO, BIAS_SBO), idesc1, 1u);
        }
        if (leader) commit(A1_FULL + b);
        __syncwarp();
        if (++g1 == G) { g1 = 0; xpar ^= 1u; }
      }
      if (BMMA && q == 0) {  // the first slot of this kind carries a bias tile only
        if (leader) commit(W_EMPTY + s);
        __syncwarp();
        if (++s == NS) { s = 0; ph ^= 1u; }
      }
      if (q >= 1) {
        const uint32_t p = q - 1, c = c2;
        const uint32_t hbuf = ATM ? (p & 1u) : 0u, hpar = ATM ? ((p >> 1) & 1u) : (p & 1u);  // hidden-chunk buffer of item p and its phase
        ptx::mbar_wait(bar(H1_FULL + hbuf), hpar);
        const uint32_t acc2 = acc2_0 + (v2 & 1u) * 64u;
        // the epilogue has read what the learner before last left in this accumulator
        if (c == 0) ptx::mbar_wait(bar(A2_EMPTY + (v2 & 1u)), ((v2 >> 1) & 1u) ^ 1u);
        if (!BMMA || q == Q) ptx::mbar_wait(bar(W_FULL + s), ph);
        ptx::tc_fence_after();
        if (leader) {
          const uint64_t bd = rdesc0 + s * STAGE_STEP;
#pragma unroll
          for (uint32_t kb2 = 0; kb2 < 2; ++kb2) {
            const uint64_t ad = hdesc0 + kb2 * STAGE_STEP;
#pragma unroll
            for (uint32_t k = 0; k < 4; ++k) {
              if (ATM) {
                const uint32_t at = tmem + HTM_COL + hbuf * 64u + (kb2 * 4u + k) * 8u;  // 16 halves of K = 8 columns per MMA
                if (NCTA == 2) ptx::umma2_f16_ts(acc2, at, bd + kb2 * kb2_step + 2 * k, idesc2, (c | kb2 | k) != 0u);
                else ptx::umma_f16_ts(acc2, at, bd + kb2 * kb2_step + 2 * k, idesc2, (c | kb2 | k) != 0u);
              } else {
                mma(acc2, ad + 2 * k, bd + kb2 * kb2_step + 2 * k, idesc2, (c | kb2 | k) != 0u);
              }
            }
          }
          commit(W_EMPTY + s);
          commit(H1_EMPTY + hbuf);
          if (c == (uint32_t)a.nc - 1) commit(A2_FULL + (v2 & 1u));
        }
        __syncwarp();
        if (++s == NS) { s = 0; ph ^= 1u; }
        if (++c2 == (uint32_t)a.nc) { c2 = 0; v2 = (v2 + 1u) & 3u; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ readiness relay (peer CTA of a pair)
    // Mirrors the leader's MMA loop: wherever the leader needs this CTA's operands or buffers, wait for
    // the local barrier and forward ONE arrive to the same barrier of the leader. "Buffer free" barriers
    // pass freely on their first use, so nothing is forwarded for it.
    ptx::setmaxnreg_dec<56>();
    const uint32_t leader = ptx::elect_one();
    auto relay = [&](int bi, uint32_t parity) {
      ptx::mbar_wait(bar(bi), parity);
      if (leader) ptx::mbar_arrive_remote(ptx::mapa(bar(bi), 0));
      __syncwarp();
    };
    uint32_t s = 0, ph = 0;
    uint32_t g1 = 0, xpar = 0, c2 = 0, v2 = 0;  // the leader's running counters (v2 = tile * K + learner)
    for (uint32_t q = 0; q <= Q; ++q) {
      ALINK_DBG_ROLE(32u, 1, q);
      if (q < Q) {
        const uint32_t g = g1, b = q & 1u, u = q >> 1;
        if (u > 0) relay(A1_EMPTY + b, (u & 1u) ^ 1u);
        for (uint32_t kb = 0; kb < nkb; kb += KPS) {
          if (g == 0) {
#pragma unroll
            for (uint32_t kk = 0; kk < KPS; ++kk) relay(X_FULL + kb + kk, xpar);
          }
          relay(W_FULL + s, ph);
          if (++s == NS) { s = 0; ph ^= 1u; }
        }
        if (++g1 == G) { g1 = 0; xpar ^= 1u; }
      }
      // (pair mode: the slot of the layer-2 slice exists for every q — it carries the bias tile of item q — and
      // the leader needs it BEFORE the hidden chunk of item q - 1)
      relay(W_FULL + s, ph);
      if (++s == NS) { s = 0; ph ^= 1u; }
      if (q >= 1) {
        const uint32_t p = q - 1;
        relay(H1_FULL + (ATM ? (p & 1u) : 0u), ATM ? ((p >> 1) & 1u) : (p & 1u));
        if (c2 == 0 && v2 > 1) relay(A2_EMPTY + (v2 & 1u), ((v2 >> 1) & 1u) ^ 1u);
        if (++c2 == (uint32_t)a.nc) { c2 = 0; ++v2; }
      }
    }
  } else if (warp == 2) {
    ptx::setmaxnreg_dec<56>();
  } else if (STAGED && warp == 3) {
    // ------------------------------------------------------------------ staged mode: feature-tile loader
    ptx::setmaxnreg_dec<56>();
    const uint32_t leader = ptx::elect_one();
    const uint8_t *scr = a.xscratch + (size_t)blockIdx.x * (2u * MAX_KB * STAGE_BYTES);
    for (uint32_t i = 0; i < T; ++i) {
      ALINK_DBG_ROLE(128u, 3, i);
      const uint32_t buf = i & 1u;
      ptx::mbar_wait(bar(S_FULL + buf), (i >> 1) & 1u);  // the producers have written tile i to the scratch buffer
      for (uint32_t kb = 0; kb < nkb; ++kb) {
        ptx::mbar_wait(bar(X_EMPTY + kb), (i & 1u) ^ 1u);  // the previous tile's MMAs have read this block
        if (leader) {
          ptx::mbar_expect_tx(bar(X_FULL + kb), STAGE_BYTES);
          ptx::bulk_g2s(sbase + XOFF + kb * STAGE_BYTES, scr + (size_t)(buf * MAX_KB + kb) * STAGE_BYTES, STAGE_BYTES,
                        bar(X_FULL + kb));
        }
        __syncwarp();
      }
      // the copies have landed (this warp arms every phase of X_FULL, so it sees every phase): the scratch
      // buffer may be overwritten with tile i + 2
      for (uint32_t kb = 0; kb < nkb; ++kb) ptx::mbar_wait(bar(X_FULL + kb), i & 1u);
      if (leader) ptx::mbar_arrive(bar(S_EMPTY + buf));
      __syncwarp();
    }
  } else if (warp == 3) {
    ptx::setmaxnreg_dec<56>();
    // ------------------------------------------------------------------ L2 prefetch of the next tile's rows
    const uint32_t rowbytes = (uint32_t)a.D * (TAB_BF16 ? 2u : 4u);
    for (uint32_t i = 0; i + 1 < T; ++i) {
      ALINK_DBG_ROLE(128u, 3, i);
      // tile i is being consumed: fetch tile i + 1. This warp only observes X_FULL (it neither fills nor frees
      // the block), so the wait is bounded: a lagging prefetcher skips a tile instead of waiting on a stale parity.
      if (!ptx::mbar_wait_soft(bar(X_FULL + 0), i & 1u, 2000000)) continue;
      const int64_t tile = tile_of(i + 1);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        int64_t n = tile * TILE_M + lane * 4 + e;
        if (n >= a.N) n = a.N - 1;
        int64_t ra, rb;
        pair_rows(a, n, ra, rb);
        ptx::bulk_prefetch_l2((const uint8_t *)a.ta + (size_t)ra * rowbytes, rowbytes);
        ptx::bulk_prefetch_l2((const uint8_t *)a.tb + (size_t)rb * rowbytes, rowbytes);
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // ------------------------------------------------------------------ epilogue
    // One thread per tile row (= TMEM lane). This role is the critical path of the kernel (layer 2 of an item
    // cannot start before the item's hidden chunk is in shared memory), so it is written for latency: the
    // accumulator is read 16 columns at a time with the next read in flight, the bias loads sit before the
    // TMEM wait, relu + rounding + packing is one cvt per two values, the per-item staging (bias, and the next
    // learner's layer-2 bias / last layer) is fetched one item ahead, and there is no integer division.
    ptx::setmaxnreg_dec<112>();
    const int et = tid - 128;             // 0..127 = row of the tile = TMEM lane
    const uint32_t quad = (uint32_t)(warp & 3);
    const uint32_t trow = tmem + ((quad * 32u) << 16);
    float *b1s = reinterpret_cast<float *>(smem + P_B1);
    float *b2s = reinterpret_cast<float *>(smem + P_B2);
    float *w3s = reinterpret_cast<float *>(smem + P_W3);
    float *b3s = reinterpret_cast<float *>(smem + P_B3);
    float rows[2 * ALINK_MAX_K];
    const uint32_t nc = (uint32_t)a.nc;
    const uint32_t hrow = sbase + HOFF + ((uint32_t)et >> 3) * 1024u + ((uint32_t)et & 7u) * 128u, hx = (uint32_t)et & 7u;
    if (Q > 0) {  // staging of item 0 / learner 0
      if (!BMMA) b1s[et] = __ldg(a.b1 + et);
      if (et < a.H2) b2s[et] = __ldg(a.b2 + et);
      if (et < a.H2 * a.C) w3s[et] = __ldg(a.W3 + et);
      if (et < a.C) b3s[et] = __ldg(a.b3 + et);
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    uint32_t i = 0, l = 0, c = 0, vp = 0;  // tile, learner, chunk of item q; learners done so far, mod 4
    // A finished learner's layer-2 accumulator is read one item LATER (after that item's hidden chunk has been
    // handed to the tensor pipe), so that this work overlaps the MMAs instead of extending the
    // epilogue -> layer 2 -> epilogue chain; consecutive learners use alternate accumulators.
    bool pend = false;
    uint32_t pl = 0, pi = 0, pvp = 0;  // the pending learner, its tile, its running index mod 4
    float vmn = 0.0f;                  // min over the learners so far of |p1 - p0|
    // vote entropy takes K + 1 values: lane v of every epilogue warp holds H(v votes of K) — the very expression of
    // alink_utility2, so the bits are the same — and the row's value is one shuffle instead of two logs in the tail
    const bool vote_lut = a.measure == ALINK_UTIL_VOTE_ENTROPY && a.K < 32;
    const int lane_e = threadIdx.x & 31;
    float vlut = 0.0f;
    if (vote_lut && lane_e <= a.K) {
      const float fk = (float)a.K;
      vlut = alink_entropy2(ALINK_DIV((float)(a.K - lane_e), fk), ALINK_DIV((float)lane_e, fk));
    }
    int nvote = 0;
    for (uint32_t q = 0; q <= Q; ++q) {
      ALINK_DBG_ROLE(8u, 4 + (warp & 3), q);
      const uint32_t b = q & 1u, u = q >> 1;
      uint32_t c1 = 0, l1 = 0, i1 = 0;
      bool more = false;
      float nb1 = 0.0f, nb2 = 0.0f, nw3 = 0.0f, nb3 = 0.0f;
      if (q < Q) {
      // item q + 1: fetch what it needs now, store it at the end of this item
      c1 = c + 1, l1 = l, i1 = i;
      if (c1 == nc) {
        c1 = 0;
        if (++l1 == (uint32_t)a.K) { l1 = 0; ++i1; }
      }
      more = q + 1 < Q;
      if (more) {
        if (!BMMA) nb1 = __ldg(a.b1 + (size_t)l1 * a.H1 + c1 * CHUNK_N + et);
        if (c1 == 0) {
          if (et < a.H2) nb2 = __ldg(a.b2 + (size_t)l1 * a.H2 + et);
          if (et < a.H2 * a.C) nw3 = __ldg(a.W3 + (size_t)l1 * a.H2 * a.C + et);
          if (et < a.C) nb3 = __ldg(a.b3 + (size_t)l1 * a.C + et);
        }
      }
      const uint32_t tacc = trow + b * CHUNK_N;
      const float *bias_q = b1s + b * 128;
      // 16 columns: + bias, relu, round, pack -> two 16-byte chunks of the swizzled hidden tile
      // (shared memory is busy feeding the tensor pipe, so its loads are slow: the bias of a stage is loaded
      // one stage ahead, the first one before the accumulator is even ready)
      auto load_bias = [&](float4(&bz)[4], uint32_t k) __attribute__((always_inline)) {
        if (BMMA) return;  // the accumulator already holds x W1 + b1
        const float4 *bias = reinterpret_cast<const float4 *>(bias_q + k * 16);  // 128-bit broadcast loads
        bz[0] = bias[0]; bz[1] = bias[1]; bz[2] = bias[2]; bz[3] = bias[3];
      };
      auto addb = [&](uint32_t v, float bv) __attribute__((always_inline)) -> float {
        return BMMA ? __uint_as_float(v) : __fadd_rn(__uint_as_float(v), bv);
      };
      auto convert = [&](const uint32_t(&v)[16], uint32_t k, const float4(&bz)[4]) __attribute__((always_inline)) {
        const float4 ba = bz[0], bb = bz[1], bc = bz[2], bd = bz[3];
        const uint32_t w0 = pack16_relu<F16>(addb(v[0], ba.x), addb(v[1], ba.y));
        const uint32_t w1 = pack16_relu<F16>(addb(v[2], ba.z), addb(v[3], ba.w));
        const uint32_t w2 = pack16_relu<F16>(addb(v[4], bb.x), addb(v[5], bb.y));
        const uint32_t w3 = pack16_relu<F16>(addb(v[6], bb.z), addb(v[7], bb.w));
        const uint32_t w4 = pack16_relu<F16>(addb(v[8], bc.x), addb(v[9], bc.y));
        const uint32_t w5 = pack16_relu<F16>(addb(v[10], bc.z), addb(v[11], bc.w));
        const uint32_t w6 = pack16_relu<F16>(addb(v[12], bd.x), addb(v[13], bd.y));
        const uint32_t w7 = pack16_relu<F16>(addb(v[14], bd.z), addb(v[15], bd.w));
        // layer 2 of the item that used this hidden buffer last (q - 1; q - 2 in ATM mode) has read it
        if (k == 0) ptx::mbar_wait(bar(H1_EMPTY + (ATM ? b : 0u)), ((ATM ? u : q) & 1u) ^ 1u);
        if (ATM) {
          if (k == 0) ptx::tc_fence_after();
          ptx::tmem_st8(trow + HTM_COL + b * 64u + k * 8u, w0, w1, w2, w3, w4, w5, w6, w7);
        } else {
          const uint32_t hb = hrow + (k >> 2) * STAGE_BYTES, ch = (k & 3u) * 2u;
          sts128(hb + ((ch ^ hx) << 4), w0, w1, w2, w3);
          sts128(hb + (((ch + 1u) ^ hx) << 4), w4, w5, w6, w7);
        }
      };
      uint32_t va[16], vb[16];
      float4 ba4[4] = {}, bb4[4] = {};
      load_bias(ba4, 0);
      ptx::mbar_wait(bar(A1_FULL + b), u & 1u);
      ptx::tc_fence_after();
      ptx::tmem_ld16(tacc, va);
      ptx::tc_wait_ld();
#pragma unroll
      for (uint32_t kk = 0; kk < 4; ++kk) {
        ptx::tmem_ld16(tacc + (2 * kk + 1) * 16, vb);  // in flight while va is converted
        load_bias(bb4, 2 * kk + 1);
        convert(va, 2 * kk, ba4);
        ptx::tc_wait_ld();
        if (kk < 3) {
          ptx::tmem_ld16(tacc + (2 * kk + 2) * 16, va);
          load_bias(ba4, 2 * kk + 2);
        }
        convert(vb, 2 * kk + 1, bb4);
        if (kk < 3) ptx::tc_wait_ld();
      }
      if (ATM) ptx::tc_wait_st();
      else ptx::fence_proxy_async_smem();
      ptx::tc_fence_before();
      ptx::mbar_arrive(bar(A1_EMPTY + b));
      ptx::mbar_arrive(bar(H1_FULL + (ATM ? b : 0u)));
      }  // q < Q
      if (pend) {
        // ---- learner pl is complete for tile pi: layer-2 accumulator -> Dense(H2, C) -> activation
        pend = false;
        const uint32_t lb = pvp & 1u;  // accumulator and parameter buffer of this learner
        ptx::mbar_wait(bar(A2_FULL + lb), (pvp >> 1) & 1u);
        ptx::tc_fence_after();
        float z0 = 0.0f, z1 = 0.0f;
        // (shared memory is busy feeding the tensor pipe: its loads are slow, so they are issued in batches
        // well ahead of their use, and the C == 2 / C == 1 variants are separate loops without inner branches)
        const float *b2p = b2s + lb * 64, *w3p = w3s + lb * 128;
        const uint32_t tacc2 = trow + ACC2_COL + lb * 64;
        if (a.C == 2) {
          for (int jj = 0; jj < a.H2 / 32; ++jj) {
            uint32_t v[32];
            ptx::tmem_ld32(tacc2 + jj * 32, v);
            const float4 *b2v = reinterpret_cast<const float4 *>(b2p + jj * 32);
            const float4 *wv = reinterpret_cast<const float4 *>(w3p + 2 * jj * 32);
            float4 bq[4], wq[8];
#pragma unroll
            for (int t = 0; t < 4; ++t) bq[t] = b2v[t];
#pragma unroll
            for (int t = 0; t < 8; ++t) wq[t] = wv[t];
            ptx::tc_wait_ld();
#pragma unroll
            for (int hf = 0; hf < 2; ++hf) {
              float4 bn[4], wn[8];
              if (hf == 0) {  // second half's operands, in flight while the first half is consumed
#pragma unroll
                for (int t = 0; t < 4; ++t) bn[t] = b2v[4 + t];
#pragma unroll
                for (int t = 0; t < 8; ++t) wn[t] = wv[8 + t];
              }
#pragma unroll
              for (int t = 0; t < 4; ++t) {
                const int e = hf * 16 + t * 4;
                const float h0 = fmaxf(__fadd_rn(__uint_as_float(v[e + 0]), bq[t].x), 0.0f);
                const float h1 = fmaxf(__fadd_rn(__uint_as_float(v[e + 1]), bq[t].y), 0.0f);
                const float h2 = fmaxf(__fadd_rn(__uint_as_float(v[e + 2]), bq[t].z), 0.0f);
                const float h3 = fmaxf(__fadd_rn(__uint_as_float(v[e + 3]), bq[t].w), 0.0f);
                const float4 wa = wq[2 * t], wb = wq[2 * t + 1];
                z0 = __fmaf_rn(h0, wa.x, z0); z1 = __fmaf_rn(h0, wa.y, z1);
                z0 = __fmaf_rn(h1, wa.z, z0); z1 = __fmaf_rn(h1, wa.w, z1);
                z0 = __fmaf_rn(h2, wb.x, z0); z1 = __fmaf_rn(h2, wb.y, z1);
                z0 = __fmaf_rn(h3, wb.z, z0); z1 = __fmaf_rn(h3, wb.w, z1);
              }
              if (hf == 0) {
#pragma unroll
                for (int t = 0; t < 4; ++t) bq[t] = bn[t];
#pragma unroll
                for (int t = 0; t < 8; ++t) wq[t] = wn[t];
              }
            }
          }
        } else {
          for (int jj = 0; jj < a.H2 / 32; ++jj) {
            uint32_t v[32];
            ptx::tmem_ld32(tacc2 + jj * 32, v);
            const float4 *b2v = reinterpret_cast<const float4 *>(b2p + jj * 32);
            const float4 *wv = reinterpret_cast<const float4 *>(w3p + jj * 32);
            float4 bq[8], wq[8];
#pragma unroll
            for (int t = 0; t < 8; ++t) { bq[t] = b2v[t]; wq[t] = wv[t]; }
            ptx::tc_wait_ld();
#pragma unroll
            for (int t = 0; t < 8; ++t) {
              const int e = t * 4;
              const float h0 = fmaxf(__fadd_rn(__uint_as_float(v[e + 0]), bq[t].x), 0.0f);
              const float h1 = fmaxf(__fadd_rn(__uint_as_float(v[e + 1]), bq[t].y), 0.0f);
              const float h2 = fmaxf(__fadd_rn(__uint_as_float(v[e + 2]), bq[t].z), 0.0f);
              const float h3 = fmaxf(__fadd_rn(__uint_as_float(v[e + 3]), bq[t].w), 0.0f);
              z0 = __fmaf_rn(h0, wq[t].x, z0);
              z0 = __fmaf_rn(h1, wq[t].y, z0);
              z0 = __fmaf_rn(h2, wq[t].z, z0);
              z0 = __fmaf_rn(h3, wq[t].w, z0);
            }
          }
        }
        ptx::tc_fence_before();
        ptx::mbar_arrive(bar(A2_EMPTY + lb));
        const int64_t n = tile_of(pi) * TILE_M + et;
        float p0, p1;
        if (a.C == 2) {
          alink_softmax2(__fadd_rn(z0, b3s[lb * 2 + 0]), __fadd_rn(z1, b3s[lb * 2 + 1]), &p0, &p1);
          if (a.probs && n < a.N)
            *reinterpret_cast<float2 *>(a.probs + ((size_t)n * a.K + pl) * 2) = make_float2(p0, p1);
        } else {
          p1 = alink_sigmoid(__fadd_rn(z0, b3s[lb * 2 + 0]));
          p0 = __fsub_rn(1.0f, p1);  // code/keras_wrapper.py:268
          if (a.probs && n < a.N) a.probs[(size_t)n * a.K + pl] = p1;
        }
        rows[2 * pl] = p0;
        rows[2 * pl + 1] = p1;
        const float dv = fabsf(__fsub_rn(p1, p0));
        vmn = pl == 0 ? dv : fminf(vmn, dv);
        nvote = (pl == 0 ? 0 : nvote) + (p1 > p0 ? 1 : 0);
        if (vote_lut) {
          if (pl == (uint32_t)a.K - 1) {  // (uniform across the warp: every lane shuffles)
            const float ve = __shfl_sync(0xffffffffu, vlut, nvote);
            if (a.util && n < a.N) {
              a.util[n] = ve;
              if (a.vmin) a.vmin[n] = vmn;
            }
          }
        } else if (pl == (uint32_t)a.K - 1 && a.util && n < a.N) {
          float c0, c1u;
          alink_consensus2(rows, a.K, &c0, &c1u);
          a.util[n] = alink_utility2(a.measure, rows, a.K, c0, c1u);
          if (a.vmin) a.vmin[n] = vmn;
        }
      }
      if (q == Q) break;
      // one-chunk learners: the parameters staged below replace those of the learner just finished above
      if (nc == 1) asm volatile("bar.sync 1, 128;" ::: "memory");
      // staging of item q + 1 (the other bias buffer; the other learner buffer: consecutive learners alternate)
      if (more) {
        if (!BMMA) b1s[(b ^ 1u) * 128 + et] = nb1;
        if (c1 == 0) {
          const uint32_t lb1 = (vp & 1u) ^ 1u;
          if (et < a.H2) b2s[lb1 * 64 + et] = nb2;
          if (et < a.H2 * a.C) w3s[lb1 * 128 + et] = nw3;
          if (et < a.C) b3s[lb1 * 2 + et] = nb3;
        }
      }
      if (!BMMA || (more && c1 == 0)) asm volatile("bar.sync 1, 128;" ::: "memory");

      if (c == nc - 1) {  // learner l is complete once layer 2 of this item has run: read it during the next item
        pend = true;
        pl = l; pi = i; pvp = vp;
        vp = (vp + 1u) & 3u;
      }
      c = c1; l = l1; i = i1;
    }
  } else if (warp >= 8) {
    // ------------------------------------------------------------------ gather producers
    // 16 lanes per row segment (64 features = 256 B of fp32), 8 rows per thread and k-block. Loads run
    // two k-blocks deep (the registers come from setmaxnreg): while block kb is converted and stored,
    // the loads of block kb + 1 — possibly of the NEXT tile — are already in flight.
    ptx::setmaxnreg_inc<168>();
    const int pt = tid - 256, sub = pt & 15, rbase = pt >> 4;
    const uint32_t nkb_feat = (uint32_t)a.D / KBLK;  // k-blocks of the |a-b| half
    uint32_t ra[8], rb[8];                           // table rows of this thread's 8 pairs
    auto load_rows = [&](uint32_t i) __attribute__((always_inline)) {
      const int64_t tile = tile_of(i);
#pragma unroll
      for (int ii = 0; ii < 8; ++ii) {
        int64_t n = tile * TILE_M + rbase + 16 * ii;
        if (n >= a.N) n = a.N - 1;
        int64_t r0, r1;
        pair_rows(a, n, r0, r1);
        ra[ii] = (uint32_t)r0;
        rb[ii] = (uint32_t)r1;
      }
    };
    auto issue = [&](float4(&va)[8], float4(&vb)[8], uint32_t kb) __attribute__((always_inline)) {
      const uint32_t col = (kb >= nkb_feat ? kb - nkb_feat : kb) * KBLK + (uint32_t)sub * 4;
#pragma unroll
      for (int ii = 0; ii < 8; ++ii) {
        const size_t oa = (size_t)ra[ii] * (uint32_t)a.D + col, ob = (size_t)rb[ii] * (uint32_t)a.D + col;
        if (TAB_BF16) {
          va[ii] = ldg_bf4((const __nv_bfloat16 *)a.ta + oa);
          vb[ii] = ldg_bf4((const __nv_bfloat16 *)a.tb + ob);
        } else {
          va[ii] = ldg_f4((const float *)a.ta + oa);
          vb[ii] = ldg_f4((const float *)a.tb + ob);
        }
      }
    };
    auto finish = [&](const float4(&va)[8], const float4(&vb)[8], uint32_t i, uint32_t kb) __attribute__((always_inline)) {
      const bool had = kb >= nkb_feat;
      uint8_t *xg = nullptr;
      if (STAGED) {
        if (kb == 0) ptx::mbar_wait(bar(S_EMPTY + (i & 1u)), ((i >> 1) & 1u) ^ 1u);  // tile i - 2 has left the buffer
        xg = a.xscratch + ((size_t)blockIdx.x * 2u + (i & 1u)) * (MAX_KB * STAGE_BYTES) + (size_t)kb * STAGE_BYTES;
      } else {
        ptx::mbar_wait(bar(X_EMPTY + kb), (i & 1u) ^ 1u);  // the previous tile's MMAs have read this block
      }
      const uint32_t xb = sbase + XOFF + kb * STAGE_BYTES;
#pragma unroll
      for (int ii = 0; ii < 8; ++ii) {
        float x0, x1, x2, x3;
        if (had) {
          x0 = __fmul_rn(va[ii].x, vb[ii].x); x1 = __fmul_rn(va[ii].y, vb[ii].y);
          x2 = __fmul_rn(va[ii].z, vb[ii].z); x3 = __fmul_rn(va[ii].w, vb[ii].w);
        } else {  // K.abs(l - r), code/siamese.py:27
          x0 = fabsf(__fsub_rn(va[ii].x, vb[ii].x)); x1 = fabsf(__fsub_rn(va[ii].y, vb[ii].y));
          x2 = fabsf(__fsub_rn(va[ii].z, vb[ii].z)); x3 = fabsf(__fsub_rn(va[ii].w, vb[ii].w));
        }
        const uint32_t r = (uint32_t)(rbase + 16 * ii);
        const uint32_t off = ptx::sw128_offset(r, (uint32_t)sub >> 1) + ((uint32_t)sub & 1u) * 8u;
        if (STAGED) *reinterpret_cast<uint2 *>(xg + off) = make_uint2(pack16<F16>(x0, x1), pack16<F16>(x2, x3));
        else sts64(xb + off, pack16<F16>(x0, x1), pack16<F16>(x2, x3));
      }
      if (STAGED) {
        if (kb + 1 == nkb) {  // tile complete: publish the buffer to the loader's bulk copies (async proxy)
          ptx::fence_proxy_async_global();
          ptx::mbar_arrive(bar(S_FULL + (i & 1u)));
        }
      } else {
        ptx::fence_proxy_async_smem();
        ptx::mbar_arrive(bar(X_FULL + kb));
      }
    };
    float4 va0[8], vb0[8], va1[8], vb1[8];
    if (T > 0) {
      load_rows(0);
      issue(va0, vb0, 0);
    }
    for (uint32_t i = 0; i < T; ++i) {
      ALINK_DBG_ROLE(16u, 8 + ((warp - 8) & 7), i);
      for (uint32_t kb = 0; kb < nkb; kb += 2) {
        const bool has1 = kb + 1 < nkb;
        if (has1) issue(va1, vb1, kb + 1);
        finish(va0, vb0, i, kb);
        if (kb + 2 < nkb) {
          issue(va0, vb0, kb + 2);
        } else if (i + 1 < T) {
          load_rows(i + 1);
          issue(va0, vb0, 0);
        }
        if (has1) finish(va1, vb1, i, kb + 1);
      }
    }
  }

  ptx::tc_fence_before();
  if (NCTA == 2) ptx::cluster_sync();  // the peer may still be signalling / reading this CTA's shared memory
  else __syncthreads();
  if (warp == 2) {
    ptx::tc_fence_after();
    if (NCTA == 2) ptx::tmem_dealloc2(tmem, TMEM_COLS);
    else ptx::tmem_dealloc(tmem, TMEM_COLS);
  }
}

// ---- one-time weight repack: Keras (in,out) fp32 -> bf16 UMMA operand images (N rows x 64 k, swizzled)
template <bool F16>
__global__ void prep_w1_kernel(const float *__restrict__ W1, uint8_t *__restrict__ img, int K, int Din, int H1) {
  const int nc = H1 / CHUNK_N, nkb = Din / KBLK;
  const int64_t total = (int64_t)K * nc * nkb * 1024;  // 16-byte chunks
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int chunk = (int)(idx & 7), r = (int)((idx >> 3) & 127);
    const int64_t blk = idx >> 10;
    const int kb = (int)(blk % nkb), c = (int)((blk / nkb) % nc), l = (int)(blk / ((int64_t)nkb * nc));
    const float *src = W1 + ((size_t)l * Din + (size_t)kb * KBLK + chunk * 8) * H1 + c * CHUNK_N + r;
    uint32_t w[4];
    for (int h = 0; h < 4; ++h) w[h] = pack16<F16>(src[(size_t)(2 * h) * H1], src[(size_t)(2 * h + 1) * H1]);
    *reinterpret_cast<uint4 *>(img + (size_t)blk * STAGE_BYTES + ptx::sw128_offset((uint32_t)r, (uint32_t)chunk)) =
        make_uint4(w[0], w[1], w[2], w[3]);
  }
}
template <bool F16>
__global__ void prep_w2_kernel(const float *__restrict__ W2, uint8_t *__restrict__ img, int K, int H1, int H2) {
  const int nc = H1 / CHUNK_N;
  const int64_t total = (int64_t)K * nc * 2 * H2 * 8;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int chunk = (int)(idx & 7);
    const int n = (int)((idx >> 3) % H2);
    const int64_t sb = (idx >> 3) / H2;  // (l * nc + c) * 2 + kb2
    const int kb2 = (int)(sb & 1);
    const int64_t blk = sb >> 1;
    const int c = (int)(blk % nc), l = (int)(blk / nc);
    const float *src = W2 + ((size_t)l * H1 + (size_t)c * CHUNK_N + kb2 * KBLK + chunk * 8) * H2 + n;
    uint32_t w[4];
    for (int h = 0; h < 4; ++h) w[h] = pack16<F16>(src[(size_t)(2 * h) * H2], src[(size_t)(2 * h + 1) * H2]);
    *reinterpret_cast<uint4 *>(img + (size_t)blk * (2 * H2 * 128) + (size_t)kb2 * (H2 * 128) +
                               ptx::sw128_offset((uint32_t)n, (uint32_t)chunk)) = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

// layer-1 bias tiles for the bias MMA of pair mode: [K][nc][2 ranks][BIAS_TILE]; element (n, k) of a tile at
// (n / 8) * BIAS_SBO + (k / 8) * BIAS_LBO + (n % 8) * 16 + (k % 8) * 2; k = 0, 1, 2 hold hi, mid, lo with
// hi + mid + lo = b1 (each part rounded to the operand format), the other k are zero
template <bool F16>
__global__ void prep_b1_kernel(const float *__restrict__ b1, uint8_t *__restrict__ img, int K, int H1) {
  const int nc = H1 / CHUNK_N;
  const int total = K * nc * 2 * 64;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
    const int n = e & 63, r = (e >> 6) & 1, lc = e >> 7;  // lc = l * nc + c
    const float b = b1[(size_t)(lc / nc) * H1 + (lc % nc) * CHUNK_N + r * 64 + n];
    uint16_t part[3];
    float rest = b;
#pragma unroll
    for (int t = 0; t < 3; ++t) {
      float back;
      if (F16) {
        const __half h = __float2half_rn(rest);
        part[t] = __half_as_ushort(h);
        back = __half2float(h);
      } else {
        const __nv_bfloat16 h = __float2bfloat16_rn(rest);
        part[t] = __bfloat16_as_ushort(h);
        back = __bfloat162float(h);
      }
      rest = __fsub_rn(rest, back);
    }
    uint8_t *tile = img + ((size_t)lc * 2 + r) * BIAS_TILE + (n >> 3) * BIAS_SBO + (n & 7) * 16;
    *reinterpret_cast<uint4 *>(tile) = make_uint4((uint32_t)part[0] | ((uint32_t)part[1] << 16), (uint32_t)part[2], 0u, 0u);
    *reinterpret_cast<uint4 *>(tile + BIAS_LBO) = make_uint4(0u, 0u, 0u, 0u);
  }
}

}  // namespace tc

// formats: bit 0 = bf16 images, bit 1 = half images
static int tc_prepare_heads(alink_ctx *ctx, HeadGroup &g, int formats = 1) {
  if (g.D_in % tc::KBLK || g.D_in > 8192)
    return alink_fail(ctx, ALINK_E_UNSUPPORTED, "BF16_TC head needs fan-in %% 64 == 0 and <= 8192 (got %d)", g.D_in);
  if (g.D_in > tc::MAX_KB * tc::KBLK && g.H1 > 512)
    return alink_fail(ctx, ALINK_E_UNSUPPORTED, "BF16_TC head with fan-in > 512 needs H1 <= 512 (got %d)", g.H1);
  if (g.H1 % tc::CHUNK_N || g.H1 > 4096)
    return alink_fail(ctx, ALINK_E_UNSUPPORTED, "BF16_TC head needs H1 %% 128 == 0 (got %d)", g.H1);
  if (g.H2 != 32 && g.H2 != 64) return alink_fail(ctx, ALINK_E_UNSUPPORTED, "BF16_TC head needs H2 in {32, 64}");
  // the layer-1 bias tiles of pair mode (tc::BIAS_TILE) live behind the layer-2 image
  const size_t w1b = (size_t)g.K * g.D_in * g.H1 * 2, w2b = (size_t)g.K * g.H1 * g.H2 * 2;
  const size_t bb = (size_t)g.K * (g.H1 / tc::CHUNK_N) * 2 * tc::BIAS_TILE;
  if (formats & 1) {
    ALINK_CUDA(ctx, cudaMalloc((void **)&g.W1t, w1b));
    ALINK_CUDA(ctx, cudaMalloc((void **)&g.W2t, w2b + bb));
    tc::prep_w1_kernel<false><<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(g.W1, (uint8_t *)g.W1t, g.K, g.D_in, g.H1);
    ALINK_LAUNCHED(ctx);
    tc::prep_w2_kernel<false><<<ctx->sm_count, 256, 0, ctx->stream>>>(g.W2, (uint8_t *)g.W2t, g.K, g.H1, g.H2);
    ALINK_LAUNCHED(ctx);
    tc::prep_b1_kernel<false><<<ctx->sm_count, 256, 0, ctx->stream>>>(g.b1, (uint8_t *)g.W2t + w2b, g.K, g.H1);
    ALINK_LAUNCHED(ctx);
  }
  if (formats & 2) {
    ALINK_CUDA(ctx, cudaMalloc((void **)&g.W1h, w1b));
    ALINK_CUDA(ctx, cudaMalloc((void **)&g.W2h, w2b + bb));
    tc::prep_w1_kernel<true><<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(g.W1, (uint8_t *)g.W1h, g.K, g.D_in, g.H1);
    ALINK_LAUNCHED(ctx);
    tc::prep_w2_kernel<true><<<ctx->sm_count, 256, 0, ctx->stream>>>(g.W2, (uint8_t *)g.W2h, g.K, g.H1, g.H2);
    ALINK_LAUNCHED(ctx);
    tc::prep_b1_kernel<true><<<ctx->sm_count, 256, 0, ctx->stream>>>(g.b1, (uint8_t *)g.W2h + w2b, g.K, g.H1);
    ALINK_LAUNCHED(ctx);
  }
  return ALINK_OK;
}

static int tc_score_ko(alink_ctx *ctx, const HeadGroup &g, const tc::Args &a0, bool tb16, bool f16);  // head_tc_ko.cuh

// whether tc_score can also return vmin (the k-inner kernel only)
static bool tc_takes_vmin(const HeadGroup &g) {
  const char *e = getenv("ALINK_TC_KO");
  return g.D_in / tc::KBLK <= tc::MAX_KB && !(e && atoi(e) == 1);
}

// Scores N pairs with every learner; probs [N][K][C] and/or util [N] (measure) are written by the
// kernel's epilogue. ia/ib are device pointers or null.
static int tc_score(alink_ctx *ctx, const HeadGroup &g, const Table &A, const Table &B, const int32_t *ia,
                    const int32_t *ib, int64_t N, float *probs, float *util, int measure, bool f16 = false,
                    float *vmin = nullptr) {
  if (f16 ? (!g.W1h || !g.W2h || A.dtype != ALINK_F32) : (!g.W1t || !g.W2t))
    return alink_fail(ctx, ALINK_E_STATE, "tensor-core weight images were not prepared");
  tc::Args a;
  memset(&a, 0, sizeof(a));
  a.ta = A.p; a.tb = B.p; a.ia = ia; a.ib = ib; a.N = N; a.D = A.D;
  a.G = ia ? 0u : ctx->ap_G; a.pair0 = ctx->ap_pair0;
  a.nkb = g.D_in / tc::KBLK; a.hadamard = g.feat_mode == ALINK_FEAT_ABSDIFF_HADAMARD;
  a.K = g.K; a.nc = g.H1 / tc::CHUNK_N; a.H1 = g.H1; a.H2 = g.H2; a.C = g.C; a.act = g.act; a.measure = measure;
  a.w1img = (const uint8_t *)(f16 ? (void *)g.W1h : (void *)g.W1t);
  a.w2img = (const uint8_t *)(f16 ? (void *)g.W2h : (void *)g.W2t);
  a.bimg = a.w2img + (size_t)g.K * g.H1 * g.H2 * 2;
  a.b1 = g.b1; a.b2 = g.b2; a.W3 = g.W3; a.b3 = g.b3;
  a.probs = probs; a.util = util; a.vmin = vmin;
  a.ntiles = (int)((N + tc::TILE_M - 1) / tc::TILE_M);
  {
    const char *e = getenv("ALINK_TC_DBG_DELAY");  // "<role bits>:<ns>", read per launch; acts in the debug build only
    unsigned bits = 0, ns = 0;
    if (e && sscanf(e, "%u:%u", &bits, &ns) == 2) a.dbg = (bits & 255u) | (ns << 8);
  }
  // Wide features (the tile does not fit in shared memory) take the k-outer kernel; ALINK_TC_KO=1 forces it.
  static const int ko_env = [] { const char *e = getenv("ALINK_TC_KO"); return e ? atoi(e) : 0; }();
  if (a.nkb > tc::MAX_KB || (ko_env == 1 && g.H1 <= 512)) {
    if (vmin) return alink_fail(ctx, ALINK_E_UNSUPPORTED, "the vote filter is not in the k-outer kernel");
    return tc_score_ko(ctx, g, a, A.dtype == ALINK_BF16, f16);
  }
  // CTA pairs (cta_group::2) by default; ALINK_TC_NCTA=1 selects the independent-CTA variant.
  static const int ncta_env = [] { const char *e = getenv("ALINK_TC_NCTA"); return e ? atoi(e) : 2; }();
  const int ncta = (ncta_env == 1 || ctx->sm_count < 2 || (a.nkb & 1)) ? 1 : 2;  // a pair slot holds 2 k-blocks
  const bool tb16 = A.dtype == ALINK_BF16;
  // Staged feature tiles (see the kernel) unless ALINK_TC_STAGED=0; the scratch buffer is allocated once per context.
  static const int staged_env = [] { const char *e = getenv("ALINK_TC_STAGED"); return e ? atoi(e) : 1; }();
  const bool staged = staged_env != 0;
  if (staged) {
    const size_t need = (size_t)ctx->sm_count * 2u * tc::MAX_KB * tc::STAGE_BYTES;
    ALINK_TRY(dev_reserve(ctx, ctx->xscratch, need));  // shared with the k-outer kernel (same stream)
    a.xscratch = (uint8_t *)ctx->xscratch.p;
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.blockDim = dim3(tc::NTHREADS);
  cfg.dynamicSmemBytes = tc::SMEM_BYTES;
  cfg.stream = ctx->stream;
  cudaLaunchAttribute attr[1];
  void (*kern)(const tc::Args);
  if (ncta == 2) {
    const int units = (a.ntiles + 1) / 2, max_units = (ctx->sm_count - ctx->reserved_sms) / 2;
    cfg.gridDim = dim3(2u * (unsigned)(units < max_units ? units : max_units));
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    // layer 2's A operand from tensor memory (ATM) for the staged CTA-pair kernels; ALINK_TC_ATM=0 keeps it in shared memory
    static const int atm_env = [] { const char *e = getenv("ALINK_TC_ATM"); return e ? atoi(e) : 1; }();
    if (staged && atm_env)
      kern = f16 ? tc::head_kernel<false, 2, true, true, true>
                 : tb16 ? tc::head_kernel<true, 2, false, true, true> : tc::head_kernel<false, 2, false, true, true>;
    else if (staged)
      kern = f16 ? tc::head_kernel<false, 2, true, true>
                 : tb16 ? tc::head_kernel<true, 2, false, true> : tc::head_kernel<false, 2, false, true>;
    else
      kern = f16 ? tc::head_kernel<false, 2, true> : tb16 ? tc::head_kernel<true, 2, false> : tc::head_kernel<false, 2, false>;
  } else {
    cfg.gridDim = dim3((unsigned)(a.ntiles < ctx->sm_count - ctx->reserved_sms ? a.ntiles : ctx->sm_count - ctx->reserved_sms));
    if (staged)
      kern = f16 ? tc::head_kernel<false, 1, true, true>
                 : tb16 ? tc::head_kernel<true, 1, false, true> : tc::head_kernel<false, 1, false, true>;
    else
      kern = f16 ? tc::head_kernel<false, 1, true> : tb16 ? tc::head_kernel<true, 1, false> : tc::head_kernel<false, 1, false>;
  }
  ALINK_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc::SMEM_BYTES));
  ALINK_CUDA(ctx, cudaLaunchKernelEx(&cfg, kern, a));
  ALINK_LAUNCHED(ctx);
  return ALINK_OK;
}
