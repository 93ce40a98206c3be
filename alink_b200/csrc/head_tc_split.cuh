// head_tc_split.cuh — the siamese head on tcgen05 with SPLIT half-precision operands: ~fp32 accuracy at tensor speed.
//
// Second stage of FP32_REFINE (refine.cuh): the few thousand pairs the half-precision filter left near the B-th best
// utility are re-scored here with every operand carried as hi + lo IEEE halves,
//     x = (x_hi + x_lo) / sx,  W = (W_hi + W_lo) / sw      (power-of-two scales keep the lo parts normal numbers)
//     x W ~= x_hi W_hi + x_hi W_lo + x_lo W_hi             (the dropped x_lo W_lo term is 2^-22 relative)
// three kind::f16 MMAs per product with fp32 accumulation in TMEM, for layer 1 (|a-b| W1) and, with the hidden
// activations split the same way in the epilogue, for layer 2; bias, relu, Dense(H2, C), softmax / sigmoid and the
// K-learner utility are fp32 as in head_tc.cuh. The result differs from the bit-exact FFMA chain
// (head_ffma.cuh; /root/reference/code/siamese.py:24-33) by ~1e-6 instead of the filter's ~1e-4, so only the pairs
// within ~1e-5 of the B-th value still need the exact CUDA-core re-score.
//
// Structure = the k-outer kernel (head_tc_ko.cuh): CTA pairs (cta_group::2, M = 256), gather producers write the
// tile's operand images (hi blocks, then lo blocks) into an L2-resident scratch one tile ahead, a TMA stage ring
// streams [X block | this CTA's half of the W1 blocks of all H1/128 chunks] and the MMAs of a stage accumulate
// into all chunk accumulators (4 x 128 TMEM columns at H1 = 512). The k loop runs over 3 x nkb "virtual" blocks:
// (x_hi, W_hi), (x_hi, W_lo), (x_lo, W_hi). Layer 2 per chunk: 24 MMAs on (h_hi, W2_hi), (h_hi, W2_lo), (h_lo, W2_hi).
// The number of pairs is read from DEVICE memory (n_dev) so that the host never waits for the candidate count.
// Work unit = (pair of tiles, learner), units dealt round-robin to the CTA pairs: the candidate lists are a wave or two
// of tiles, and at tile granularity 1.16 waves of work would cost two full waves; at learner granularity they cost 1.2.
// The kernel therefore writes per-learner probabilities; the K-learner utility is a separate streaming pass.
#pragma once
#include "head_tc_ko.cuh"

namespace tcsp {

using tc::Args;
using tc::CHUNK_N;
using tc::KBLK;
using tc::TILE_M;

constexpr int NTHREADS = 512;
constexpr int MAX_NC = 4;
constexpr uint32_t XBLK = 16384;
constexpr uint32_t WPART = 8192;
constexpr uint32_t STAGE = XBLK + MAX_NC * WPART;            // 48 KB
constexpr int NS = 3;
constexpr uint32_t ROFF = 0;
constexpr uint32_t HOFF = ROFF + NS * STAGE;                 // hidden chunk: hi image 32 KB, lo image 32 KB
constexpr uint32_t W2OFF = HOFF + 2 * 32768;                 // layer-2 slice: hi 8 KB, lo 8 KB (this CTA's halves)
constexpr uint32_t POFF = W2OFF + 2 * 8192;
constexpr uint32_t P_B1 = POFF;                              // float [128]
constexpr uint32_t P_B2 = P_B1 + 2 * 128 * 4;                // float [2][64]
constexpr uint32_t P_W3 = P_B2 + 2 * 64 * 4;                 // float [2][128]
constexpr uint32_t P_B3 = P_W3 + 2 * 128 * 4;                // float [2][2]
constexpr uint32_t BOFF = P_B3 + 16;
enum { W_FULL = 0, W_EMPTY = 3, W2_FULL = 6, W2_EMPTY = 7, A1_FULL = 8, A1_EMPTY = 9, H1_FULL = 10, H1_EMPTY = 11,
       A2_FULL = 12, XS_FULL = 13, XS_EMPTY = 15, NBARS = 17 };
constexpr uint32_t TOFF = BOFF + NBARS * 8;
constexpr uint32_t SMEM_BYTES = TOFF + 16;
static_assert(SMEM_BYTES <= 232448, "exceeds the 227 KB per-CTA shared memory limit");
static_assert(BOFF % 8 == 0 && STAGE % 1024 == 0, "alignment");
constexpr uint32_t TMEM_COLS = 512;

struct SplitArgs {
  Args a;                 // a.w1img / a.w2img = the scaled HI images; a.N = capacity of the pair list
  const uint8_t *w1lo;    // scaled LO images, same block layout
  const uint8_t *w2lo;
  uint8_t *xscratch;      // [gridDim.x][2][2 * nkb] blocks of 16 KB (hi blocks, then lo blocks)
  const unsigned int *n_dev;  // number of pairs (device; nullptr = a.N)
  float sx, inv1;         // x is scaled by sx; layer-1 accumulator * inv1 = x W1   (inv1 = 1 / (sx * sw1))
  float sh, inv2;         // hidden activations are scaled by sh; layer-2 accumulator * inv2 = h1 W2
};

// hi / lo halves of two scaled values, packed (lo = the rounding error of hi, itself rounded to half)
__device__ __forceinline__ void split2(float v0, float v1, uint32_t &hi, uint32_t &lo) {
  const __half2 h = __floats2half2_rn(v0, v1);
  const float2 back = __half22float2(h);
  const __half2 l = __floats2half2_rn(__fsub_rn(v0, back.x), __fsub_rn(v1, back.y));
  hi = *reinterpret_cast<const uint32_t *>(&h);
  lo = *reinterpret_cast<const uint32_t *>(&l);
}

__global__ void __launch_bounds__(NTHREADS, 1) head_kernel_split(const SplitArgs sa) {
  const Args &a = sa.a;
  const uint32_t rank = ptx::cluster_ctarank();
  const uint32_t unit0 = blockIdx.x / 2, nunit = gridDim.x / 2;
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sbase = ptx::smem_u32(smem);
  const int tid = threadIdx.x, warp = tid >> 5;
  auto bar = [&](int i) -> uint32_t { return sbase + BOFF + 8u * (uint32_t)i; };
  const uint32_t nkb = (uint32_t)a.nkb, nc = (uint32_t)a.nc, K = (uint32_t)a.K;
  int64_t N = a.N;
  if (sa.n_dev) {
    const int64_t nd = (int64_t)__ldg(sa.n_dev);
    if (nd < N) N = nd;
  }
  const uint32_t ntiles = (uint32_t)((N + TILE_M - 1) / TILE_M);
  uint8_t *xs_mine = sa.xscratch + (size_t)blockIdx.x * 2 * (2 * nkb) * XBLK;

  if (tid == 0) {
    if (sbase & 1023u) __trap();
    const uint32_t fwd = rank == 0 ? 1u : 0u;  // the peer's relay forwards one arrive per phase (head_tc.cuh)
    for (int s = 0; s < NS; ++s) {
      ptx::mbar_init(bar(W_FULL + s), 1 + fwd);
      ptx::mbar_init(bar(W_EMPTY + s), 1);
    }
    ptx::mbar_init(bar(W2_FULL), 1 + fwd);
    ptx::mbar_init(bar(W2_EMPTY), 1);
    ptx::mbar_init(bar(H1_FULL), 128 + fwd);
    ptx::mbar_init(bar(H1_EMPTY), 1);
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(bar(XS_FULL + s), 256);
      ptx::mbar_init(bar(XS_EMPTY + s), 1);
    }
    ptx::mbar_init(bar(A1_FULL), 1);
    ptx::mbar_init(bar(A1_EMPTY), 128 + fwd);
    ptx::mbar_init(bar(A2_FULL), 1);
    ptx::fence_barrier_init();
  }
  if (warp == 2) ptx::tmem_alloc2(sbase + TOFF, TMEM_COLS);
  ptx::tc_fence_before();
  ptx::cluster_sync();
  ptx::tc_fence_after();
  const uint32_t tmem = __shfl_sync(0xffffffffu, *reinterpret_cast<volatile uint32_t *>(smem + TOFF), 0);

  const uint32_t nunits = ((ntiles + 1) / 2) * K;  // (tile pair, learner) units, learner fastest
  const uint32_t T = nunits > unit0 ? (nunits - unit0 + nunit - 1) / nunit : 0u;  // units of this CTA pair
  auto unit_of = [&](uint32_t i) -> uint32_t { return unit0 + i * nunit; };
  auto tile_of = [&](uint32_t i) -> int64_t { return (int64_t)(unit_of(i) / K) * 2 + rank; };
  auto learner_of = [&](uint32_t i) -> uint32_t { return unit_of(i) % K; };
  const uint32_t l2bytes = 2u * (uint32_t)a.H2 * 128u;  // one layer-2 slice of the pair; this CTA loads half
  const uint32_t stage_tx = XBLK + nc * WPART;

  if (warp == 0) {
    // ------------------------------------------------------------------ stage producer (TMA bulk)
    ptx::setmaxnreg_dec<56>();
    const uint32_t leader = ptx::elect_one();
    const uint32_t ring = sbase + ROFF;
    uint32_t s = 0, ph = 0, w2n = 0;
    auto issue_w2 = [&](uint32_t l, uint32_t c) __attribute__((always_inline)) {
      ptx::mbar_wait(bar(W2_EMPTY), (w2n & 1u) ^ 1u);
      if (leader) {
        const size_t off = (size_t)(l * nc + c) * l2bytes;
        const uint32_t piece = l2bytes / 4;  // H2/2 rows x 128 B of one k-block
        const uint32_t dst = sbase + W2OFF;
        ptx::mbar_expect_tx(bar(W2_FULL), 4 * piece);
        ptx::bulk_g2s(dst, a.w2img + off + rank * piece, piece, bar(W2_FULL));
        ptx::bulk_g2s(dst + piece, a.w2img + off + 2 * piece + rank * piece, piece, bar(W2_FULL));
        ptx::bulk_g2s(dst + 8192u, sa.w2lo + off + rank * piece, piece, bar(W2_FULL));
        ptx::bulk_g2s(dst + 8192u + piece, sa.w2lo + off + 2 * piece + rank * piece, piece, bar(W2_FULL));
      }
      __syncwarp();
      ++w2n;
    };
    for (uint32_t i = 0; i < T; ++i) {
      ptx::mbar_wait(bar(XS_FULL + (i & 1u)), (i >> 1) & 1u);  // this tile's operand images are in the scratch
      const uint8_t *xsrc = xs_mine + (size_t)(i & 1u) * (2 * nkb) * XBLK;
      {
        const uint32_t l = learner_of(i);
        for (uint32_t part = 0; part < 3; ++part) {
          const uint8_t *xp = xsrc + (part == 2 ? (size_t)nkb * XBLK : 0);
          const uint8_t *wp = part == 1 ? sa.w1lo : a.w1img;
          for (uint32_t kb = 0; kb < nkb; ++kb) {
            if (part == 2 && kb == 0) issue_w2(l, 0);  // the first slice, well before layer 2 needs it
            ptx::mbar_wait(bar(W_EMPTY + s), ph ^ 1u);
            if (leader) {
              const uint32_t dst = ring + s * STAGE;
              ptx::mbar_expect_tx(bar(W_FULL + s), stage_tx);
              ptx::bulk_g2s(dst, xp + (size_t)kb * XBLK, XBLK, bar(W_FULL + s));
              for (uint32_t c = 0; c < nc; ++c)
                ptx::bulk_g2s(dst + XBLK + c * WPART,
                              wp + ((size_t)(l * nc + c) * nkb + kb) * tc::STAGE_BYTES + rank * WPART, WPART,
                              bar(W_FULL + s));
            }
            __syncwarp();
            if (++s == NS) { s = 0; ph ^= 1u; }
          }
        }
        for (uint32_t c = 1; c < nc; ++c) issue_w2(l, c);
      }
    }
  } else if (warp == 1 && rank == 0) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA)
    ptx::setmaxnreg_dec<56>();
    const uint32_t leader = ptx::elect_one();
    const uint32_t idesc1 = ptx::umma_idesc_f16(2 * TILE_M, CHUNK_N);
    const uint32_t idesc2 = ptx::umma_idesc_f16(2 * TILE_M, a.H2);
    const uint32_t kb2_step = ((uint32_t)a.H2 * 64u) >> 4;  // this CTA's half of a layer-2 k-block, 16-byte units
    const uint64_t rdesc0 = ptx::umma_desc_sw128(sbase + ROFF);
    const uint64_t hdesc0 = ptx::umma_desc_sw128(sbase + HOFF);
    const uint64_t w2desc0 = ptx::umma_desc_sw128(sbase + W2OFF);
    const uint32_t nkb3 = 3 * nkb;
    uint32_t s = 0, ph = 0, hq = 0, v = 0;
    for (uint32_t i = 0; i < T; ++i) {
      for (uint32_t l = K - 1; l < K; ++l, ++v) {  // one learner per unit (the loop body is kept for its structure)
        ptx::mbar_wait(bar(A1_EMPTY), (v & 1u) ^ 1u);  // the previous unit's accumulators are drained
        ptx::tc_fence_after();
        for (uint32_t kb = 0; kb < nkb3; ++kb) {
          ptx::mbar_wait(bar(W_FULL + s), ph);
          ptx::tc_fence_after();
          if (leader) {
            const uint64_t ad = rdesc0 + (uint64_t)s * (STAGE >> 4);
            for (uint32_t c = 0; c < nc; ++c) {
              const uint64_t bd = ad + ((XBLK + c * WPART) >> 4);
#pragma unroll
              for (uint32_t k = 0; k < 4; ++k)
                ptx::umma2_bf16(tmem + c * CHUNK_N, ad + 2 * k, bd + 2 * k, idesc1, (kb | k) != 0u);
            }
            ptx::umma2_commit_mc(bar(W_EMPTY + s));
            if (l == K - 1 && kb == nkb3 - 1) ptx::mbar_arrive(bar(XS_EMPTY + (i & 1u)));  // last load of the tile landed
          }
          __syncwarp();
          if (++s == NS) { s = 0; ph ^= 1u; }
        }
        if (leader) ptx::umma2_commit_mc(bar(A1_FULL));
        __syncwarp();
        for (uint32_t c = 0; c < nc; ++c, ++hq) {
          ptx::mbar_wait(bar(H1_FULL), hq & 1u);
          ptx::mbar_wait(bar(W2_FULL), hq & 1u);
          ptx::tc_fence_after();
          if (leader) {
#pragma unroll
            for (uint32_t part = 0; part < 3; ++part) {
              const uint64_t abase = hdesc0 + (part == 2 ? (32768u >> 4) : 0u);
              const uint64_t bbase = w2desc0 + (part == 1 ? (8192u >> 4) : 0u);
#pragma unroll
              for (uint32_t kb2 = 0; kb2 < 2; ++kb2) {
#pragma unroll
                for (uint32_t k = 0; k < 4; ++k)  // layer-2 accumulator: the (drained) first columns of chunk 0
                  ptx::umma2_bf16(tmem, abase + kb2 * (16384u >> 4) + 2 * k, bbase + kb2 * kb2_step + 2 * k, idesc2,
                                  (c | part | kb2 | k) != 0u);
              }
            }
            ptx::umma2_commit_mc(bar(W2_EMPTY));
            ptx::umma2_commit_mc(bar(H1_EMPTY));
            if (c == nc - 1) ptx::umma2_commit_mc(bar(A2_FULL));
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ readiness relay (peer CTA)
    ptx::setmaxnreg_dec<56>();
    const uint32_t leader = ptx::elect_one();
    auto relay = [&](int bi, uint32_t parity) __attribute__((always_inline)) {
      ptx::mbar_wait(bar(bi), parity);
      if (leader) ptx::mbar_arrive_remote(ptx::mapa(bar(bi), 0));
      __syncwarp();
    };
    const uint32_t nkb3 = 3 * nkb;
    uint32_t s = 0, ph = 0, hq = 0, v = 0;
    for (uint32_t i = 0; i < T; ++i) {
      for (uint32_t l = K - 1; l < K; ++l, ++v) {
        if (v > 0) relay(A1_EMPTY, (v & 1u) ^ 1u);
        for (uint32_t kb = 0; kb < nkb3; ++kb) {
          relay(W_FULL + s, ph);
          if (leader && l == K - 1 && kb == nkb3 - 1) ptx::mbar_arrive(bar(XS_EMPTY + (i & 1u)));
          __syncwarp();
          if (++s == NS) { s = 0; ph ^= 1u; }
        }
        for (uint32_t c = 0; c < nc; ++c, ++hq) {
          relay(H1_FULL, hq & 1u);
          relay(W2_FULL, hq & 1u);
        }
      }
    }
  } else if (warp == 2 || warp == 3) {
    ptx::setmaxnreg_dec<56>();
  } else if (warp >= 4 && warp < 8) {
    // ------------------------------------------------------------------ epilogue (after each learner's layer 1)
    ptx::setmaxnreg_dec<112>();
    const int et = tid - 128;  // row of the tile = TMEM lane
    const uint32_t quad = (uint32_t)(warp & 3);
    const uint32_t trow = tmem + ((quad * 32u) << 16);
    float *b1s = reinterpret_cast<float *>(smem + P_B1);
    float *b2s = reinterpret_cast<float *>(smem + P_B2);
    float *w3s = reinterpret_cast<float *>(smem + P_W3);
    float *b3s = reinterpret_cast<float *>(smem + P_B3);
    const float inv1 = sa.inv1, sh = sa.sh, inv2 = sa.inv2;
    uint32_t hq = 0, v = 0;
    for (uint32_t i = 0; i < T; ++i) {
      {
        const uint32_t l = learner_of(i);
        const uint32_t lb = v & 1u;
        if (et < a.H2) b2s[lb * 64 + et] = __ldg(a.b2 + (size_t)l * a.H2 + et);
        if (et < a.H2 * a.C) w3s[lb * 128 + et] = __ldg(a.W3 + (size_t)l * a.H2 * a.C + et);
        if (et < a.C) b3s[lb * 2 + et] = __ldg(a.b3 + (size_t)l * a.C + et);
        ptx::mbar_wait(bar(A1_FULL), v & 1u);  // all chunk accumulators of this learner are complete
        ptx::tc_fence_after();
        for (uint32_t c = 0; c < nc; ++c, ++hq) {
          const uint32_t pb = hq & 1u;
          b1s[pb * 128 + et] = __ldg(a.b1 + (size_t)l * a.H1 + c * CHUNK_N + et);
          asm volatile("bar.sync 1, 128;" ::: "memory");
          ptx::mbar_wait(bar(H1_EMPTY), (hq & 1u) ^ 1u);  // layer 2 has read the hidden images
#pragma unroll 1
          for (uint32_t jj = 0; jj < 4; ++jj) {
            uint32_t r[32];
            ptx::tmem_ld32(trow + c * CHUNK_N + jj * 32, r);
            ptx::tc_wait_ld();
            const float4 *bias = reinterpret_cast<const float4 *>(b1s + pb * 128 + jj * 32);
            const uint32_t hbase = sbase + HOFF + (jj >> 1) * 16384u;
#pragma unroll
            for (uint32_t e = 0; e < 4; ++e) {
              const float4 ba = bias[2 * e], bb = bias[2 * e + 1];
              // h1 = relu(x W1 + b1) in fp32, then scaled and split
              const float x0 = __fmul_rn(fmaxf(__fmaf_rn(__uint_as_float(r[e * 8 + 0]), inv1, ba.x), 0.0f), sh);
              const float x1 = __fmul_rn(fmaxf(__fmaf_rn(__uint_as_float(r[e * 8 + 1]), inv1, ba.y), 0.0f), sh);
              const float x2 = __fmul_rn(fmaxf(__fmaf_rn(__uint_as_float(r[e * 8 + 2]), inv1, ba.z), 0.0f), sh);
              const float x3 = __fmul_rn(fmaxf(__fmaf_rn(__uint_as_float(r[e * 8 + 3]), inv1, ba.w), 0.0f), sh);
              const float x4 = __fmul_rn(fmaxf(__fmaf_rn(__uint_as_float(r[e * 8 + 4]), inv1, bb.x), 0.0f), sh);
              const float x5 = __fmul_rn(fmaxf(__fmaf_rn(__uint_as_float(r[e * 8 + 5]), inv1, bb.y), 0.0f), sh);
              const float x6 = __fmul_rn(fmaxf(__fmaf_rn(__uint_as_float(r[e * 8 + 6]), inv1, bb.z), 0.0f), sh);
              const float x7 = __fmul_rn(fmaxf(__fmaf_rn(__uint_as_float(r[e * 8 + 7]), inv1, bb.w), 0.0f), sh);
              uint32_t h0, h1, h2, h3, l0, l1, l2, l3;
              split2(x0, x1, h0, l0);
              split2(x2, x3, h1, l1);
              split2(x4, x5, h2, l2);
              split2(x6, x7, h3, l3);
              const uint32_t off = ptx::sw128_offset((uint32_t)et, (jj & 1u) * 4 + e);
              tc::sts128(hbase + off, h0, h1, h2, h3);
              tc::sts128(hbase + 32768u + off, l0, l1, l2, l3);
            }
          }
          ptx::fence_proxy_async_smem();
          ptx::tc_fence_before();
          ptx::mbar_arrive(bar(H1_FULL));
        }
        // ---- layer-2 accumulator (TMEM columns [0, H2)) -> Dense(H2, C) -> activation
        ptx::mbar_wait(bar(A2_FULL), v & 1u);
        ptx::tc_fence_after();
        float z0 = 0.0f, z1 = 0.0f;
        for (int jj = 0; jj < a.H2 / 32; ++jj) {
          uint32_t r[32];
          ptx::tmem_ld32(trow + jj * 32, r);
          ptx::tc_wait_ld();
#pragma unroll
          for (int e = 0; e < 32; ++e) {
            const int jn = jj * 32 + e;
            const float h = fmaxf(__fmaf_rn(__uint_as_float(r[e]), inv2, b2s[lb * 64 + jn]), 0.0f);
            if (a.C == 2) {
              z0 = __fmaf_rn(h, w3s[lb * 128 + 2 * jn], z0);
              z1 = __fmaf_rn(h, w3s[lb * 128 + 2 * jn + 1], z1);
            } else {
              z0 = __fmaf_rn(h, w3s[lb * 128 + jn], z0);
            }
          }
        }
        ptx::tc_fence_before();
        ptx::mbar_arrive(bar(A1_EMPTY));  // every accumulator column of this learner has been read
        const int64_t n = tile_of(i) * TILE_M + et;
        float p0, p1;
        if (a.C == 2) {
          alink_softmax2(__fadd_rn(z0, b3s[lb * 2 + 0]), __fadd_rn(z1, b3s[lb * 2 + 1]), &p0, &p1);
          if (a.probs && n < N)
            *reinterpret_cast<float2 *>(a.probs + ((size_t)n * a.K + l) * 2) = make_float2(p0, p1);
        } else {
          p1 = alink_sigmoid(__fadd_rn(z0, b3s[lb * 2 + 0]));
          p0 = __fsub_rn(1.0f, p1);  // code/keras_wrapper.py:268
          if (a.probs && n < N) a.probs[(size_t)n * a.K + l] = p1;
        }
        (void)p0;
        ++v;
      }
    }
  } else if (warp >= 8) {
    // ------------------------------------------------------------------ gather producers -> hi / lo images in the scratch
    ptx::setmaxnreg_inc<168>();
    const int pt = tid - 256, sub = pt & 15, rbase = pt >> 4;
    const uint32_t nkb_feat = (uint32_t)a.D / KBLK;
    const float sx = sa.sx;
    uint32_t ra[8], rb[8];
    auto load_rows = [&](uint32_t i) __attribute__((always_inline)) {
      const int64_t tile = tile_of(i);
#pragma unroll
      for (int ii = 0; ii < 8; ++ii) {
        int64_t n = tile * TILE_M + rbase + 16 * ii;
        if (n >= N) n = N - 1;
        int64_t r0, r1;
        tc::pair_rows(a, n, r0, r1);
        ra[ii] = (uint32_t)r0;
        rb[ii] = (uint32_t)r1;
      }
    };
    auto issue = [&](float4(&va)[8], float4(&vb)[8], uint32_t kb) __attribute__((always_inline)) {
      const uint32_t col = (kb >= nkb_feat ? kb - nkb_feat : kb) * KBLK + (uint32_t)sub * 4;
#pragma unroll
      for (int ii = 0; ii < 8; ++ii) {
        va[ii] = tc::ldg_f4((const float *)a.ta + (size_t)ra[ii] * (uint32_t)a.D + col);
        vb[ii] = tc::ldg_f4((const float *)a.tb + (size_t)rb[ii] * (uint32_t)a.D + col);
      }
    };
    auto finish = [&](const float4(&va)[8], const float4(&vb)[8], uint32_t i, uint32_t kb) __attribute__((always_inline)) {
      const bool had = kb >= nkb_feat;
      uint8_t *xb = xs_mine + ((size_t)(i & 1u) * (2 * nkb) + kb) * XBLK;
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
        uint32_t h0, h1, l0, l1;
        split2(__fmul_rn(x0, sx), __fmul_rn(x1, sx), h0, l0);
        split2(__fmul_rn(x2, sx), __fmul_rn(x3, sx), h1, l1);
        const uint32_t r = (uint32_t)(rbase + 16 * ii);
        const uint32_t off = ptx::sw128_offset(r, (uint32_t)sub >> 1) + ((uint32_t)sub & 1u) * 8u;
        tcko::stg64(xb + off, h0, h1);
        tcko::stg64(xb + (size_t)nkb * XBLK + off, l0, l1);
      }
    };
    float4 va0[8], vb0[8], va1[8], vb1[8];
    for (uint32_t i = 0; i < T; ++i) {
      load_rows(i);
      issue(va0, vb0, 0);
      ptx::mbar_wait(bar(XS_EMPTY + (i & 1u)), ((i >> 1) & 1u) ^ 1u);  // tile i - 2 has been streamed out of this buffer
      for (uint32_t kb = 0; kb < nkb; kb += 2) {
        const bool has1 = kb + 1 < nkb;
        if (has1) issue(va1, vb1, kb + 1);
        finish(va0, vb0, i, kb);
        if (kb + 2 < nkb) issue(va0, vb0, kb + 2);
        if (has1) finish(va1, vb1, i, kb + 1);
      }
      __threadfence();                 // the images are in L2 ...
      asm volatile("fence.proxy.async;" ::: "memory");  // ... and ordered before the TMA (async proxy) reads
      ptx::mbar_arrive(bar(XS_FULL + (i & 1u)));
    }
  }

  ptx::tc_fence_before();
  ptx::cluster_sync();  // the peer may still be signalling / reading this CTA's shared memory
  if (warp == 2) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc2(tmem, TMEM_COLS);
  }
}

// scaled hi / lo operand images of W1 and W2 (block layouts of tc::prep_w1_kernel / tc::prep_w2_kernel)
__device__ __forceinline__ uint16_t split_part(float w, float scale, int lo) {
  const float v = __fmul_rn(w, scale);
  const __half h = __float2half_rn(v);
  if (!lo) return __half_as_ushort(h);
  return __half_as_ushort(__float2half_rn(__fsub_rn(v, __half2float(h))));
}
__global__ void prep_w1_split_kernel(const float *__restrict__ W1, uint8_t *__restrict__ img, int K, int Din, int H1, float scale,
                                     int lo) {
  const int nc = H1 / CHUNK_N, nkb = Din / KBLK;
  const int64_t total = (int64_t)K * nc * nkb * 1024;  // 16-byte chunks
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int chunk = (int)(idx & 7), r = (int)((idx >> 3) & 127);
    const int64_t blk = idx >> 10;
    const int kb = (int)(blk % nkb), c = (int)((blk / nkb) % nc), l = (int)(blk / ((int64_t)nkb * nc));
    const float *src = W1 + ((size_t)l * Din + (size_t)kb * KBLK + chunk * 8) * H1 + c * CHUNK_N + r;
    uint32_t w[4];
    for (int h = 0; h < 4; ++h)
      w[h] = (uint32_t)split_part(src[(size_t)(2 * h) * H1], scale, lo) | ((uint32_t)split_part(src[(size_t)(2 * h + 1) * H1], scale, lo) << 16);
    *reinterpret_cast<uint4 *>(img + (size_t)blk * tc::STAGE_BYTES + ptx::sw128_offset((uint32_t)r, (uint32_t)chunk)) =
        make_uint4(w[0], w[1], w[2], w[3]);
  }
}
__global__ void prep_w2_split_kernel(const float *__restrict__ W2, uint8_t *__restrict__ img, int K, int H1, int H2, float scale,
                                     int lo) {
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
    for (int h = 0; h < 4; ++h)
      w[h] = (uint32_t)split_part(src[(size_t)(2 * h) * H2], scale, lo) | ((uint32_t)split_part(src[(size_t)(2 * h + 1) * H2], scale, lo) << 16);
    *reinterpret_cast<uint4 *>(img + (size_t)blk * (2 * H2 * 128) + (size_t)kb2 * (H2 * 128) +
                               ptx::sw128_offset((uint32_t)n, (uint32_t)chunk)) = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

}  // namespace tcsp

// power-of-two scale that puts `absmax` just below 2^14 (half's largest finite value is 65504 = 2^16 - 32)
static inline float split_scale_for(double absmax) {
  if (!(absmax > 0.0) || !(absmax < 1e30)) return 1.0f;
  int e;
  frexp(absmax, &e);  // absmax = m * 2^e, m in [0.5, 1)
  int s = 14 - e;
  if (s > 60) s = 60;
  if (s < -60) s = -60;
  return ldexpf(1.0f, s);
}

// shapes the split kernel takes: CTA pairs, H1 <= 512 (4 chunk accumulators fill TMEM), fp32 tables
static bool split_supported(const alink_ctx *ctx, const HeadGroup &g) {
  return ctx->sm_count >= 2 && g.H1 % tc::CHUNK_N == 0 && g.H1 <= tcsp::MAX_NC * tc::CHUNK_N && g.D_in % tc::KBLK == 0 &&
         g.D_in <= 2048 && (g.H2 == 32 || g.H2 == 64);
}

// (re)build the scaled hi / lo images of a group (alink_set_heads, and after a finetune step changed the weights)
static int split_prepare_heads(alink_ctx *ctx, HeadGroup &g) {
  const size_t w1b = (size_t)g.K * g.D_in * g.H1 * 2, w2b = (size_t)g.K * g.H1 * g.H2 * 2;
  if (!g.S1hi) {
    ALINK_CUDA(ctx, cudaMalloc((void **)&g.S1hi, w1b));
    ALINK_CUDA(ctx, cudaMalloc((void **)&g.S1lo, w1b));
    ALINK_CUDA(ctx, cudaMalloc((void **)&g.S2hi, w2b));
    ALINK_CUDA(ctx, cudaMalloc((void **)&g.S2lo, w2b));
  }
  g.sw1 = split_scale_for(g.w1_absmax);
  g.sw2 = split_scale_for(g.w_absmax);
  for (int lo = 0; lo < 2; ++lo) {
    tcsp::prep_w1_split_kernel<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(g.W1, (uint8_t *)(lo ? g.S1lo : g.S1hi), g.K, g.D_in,
                                                                         g.H1, g.sw1, lo);
    ALINK_LAUNCHED(ctx);
    tcsp::prep_w2_split_kernel<<<ctx->sm_count, 256, 0, ctx->stream>>>(g.W2, (uint8_t *)(lo ? g.S2lo : g.S2hi), g.K, g.H1, g.H2,
                                                                     g.sw2, lo);
    ALINK_LAUNCHED(ctx);
  }
  return ALINK_OK;
}

// Split-operand scoring of the pairs (ia[n], ib[n]), n < *n_dev (<= cap): probs[n][K][C] per-learner probabilities
// (`util` is unused: the unit of work is one learner, see the kernel).
// xmax / h1max: a-priori bounds of |x| and |h1| (they set the power-of-two scales).
static int tc_score_split(alink_ctx *ctx, const HeadGroup &g, const Table &A, const Table &B, const int32_t *ia,
                          const int32_t *ib, int64_t cap, const unsigned int *n_dev, float *probs, float *util, int measure,
                          double xmax, double h1max) {
  if (!g.S1hi) return alink_fail(ctx, ALINK_E_STATE, "split operand images were not prepared");
  if (!probs) return alink_fail(ctx, ALINK_E_INVALID, "the split kernel writes per-learner probabilities");
  (void)util;
  tcsp::SplitArgs sa;
  memset(&sa, 0, sizeof(sa));
  tc::Args &a = sa.a;
  a.ta = A.p; a.tb = B.p; a.ia = ia; a.ib = ib; a.N = cap; a.D = A.D;
  a.G = ia ? 0u : ctx->ap_G; a.pair0 = ctx->ap_pair0;
  a.nkb = g.D_in / tc::KBLK; a.hadamard = g.feat_mode == ALINK_FEAT_ABSDIFF_HADAMARD;
  a.K = g.K; a.nc = g.H1 / tc::CHUNK_N; a.H1 = g.H1; a.H2 = g.H2; a.C = g.C; a.act = g.act; a.measure = measure;
  a.w1img = (const uint8_t *)g.S1hi; a.w2img = (const uint8_t *)g.S2hi;
  sa.w1lo = (const uint8_t *)g.S1lo; sa.w2lo = (const uint8_t *)g.S2lo;
  a.b1 = g.b1; a.b2 = g.b2; a.W3 = g.W3; a.b3 = g.b3;
  a.probs = probs; a.util = util;
  a.ntiles = (int)((cap + tc::TILE_M - 1) / tc::TILE_M);
  sa.n_dev = n_dev;
  sa.sx = split_scale_for(xmax);
  sa.sh = split_scale_for(h1max);
  sa.inv1 = 1.0f / (sa.sx * g.sw1);
  sa.inv2 = 1.0f / (sa.sh * g.sw2);
  const int64_t units = (int64_t)((a.ntiles + 1) / 2) * g.K, max_units = (ctx->sm_count - ctx->reserved_sms) / 2;
  const unsigned grid = 2u * (unsigned)(units < max_units ? units : max_units);
  ALINK_TRY(dev_reserve(ctx, ctx->xscratch, (size_t)grid * 2 * (2 * a.nkb) * tcsp::XBLK));
  sa.xscratch = (uint8_t *)ctx->xscratch.p;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(tcsp::NTHREADS);
  cfg.dynamicSmemBytes = tcsp::SMEM_BYTES;
  cfg.stream = ctx->stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  ALINK_CUDA(ctx, cudaFuncSetAttribute(tcsp::head_kernel_split, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tcsp::SMEM_BYTES));
  ALINK_CUDA(ctx, cudaLaunchKernelEx(&cfg, tcsp::head_kernel_split, sa));
  ALINK_LAUNCHED(ctx);
  return ALINK_OK;
}
