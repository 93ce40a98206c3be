// head_tc_ko.cuh — BF16_TC head for WIDE features (D_in > 512, e.g. the 2048-d VGGFace-ResNet50 pool
// of /root/reference/code/siamese.py:203-216, FEATURERES at code/ALINK.py:30): "k-outer" variant.
//
// The 128 x D_in bf16 feature tile no longer fits in shared memory (512 KB at D_in = 2048), so:
//   * the gather producers write the converted tile ONCE, as UMMA shared-memory images (16 KB blocks
//     [128 pairs x 64 k], SWIZZLE_128B), into a per-CTA double-buffered scratch in global memory (it stays
//     in L2) — one tile ahead of the tensor pipe;
//   * per learner, the k dimension is the OUTER loop: each ring stage carries one X block + the matching
//     [128 n x 64 k] weight blocks of all H1/128 chunks (this CTA's half of each), and the MMAs of a stage
//     accumulate into all chunk accumulators at once — 4 x 128 = all 512 TMEM columns at H1 = 512;
//   * the epilogue of a learner therefore follows its layer-1 MMAs instead of overlapping the next
//     chunk (~10 % of the tensor time is exposed); the layer-2 accumulator reuses the TMEM columns of
//     chunk 0 once that chunk has been drained.
// Always CTA pairs (cluster of 2, tcgen05 cta_group::2, M = 256) with the readiness relay of head_tc.cuh.
// Same epilogue arithmetic, same outputs (probs [N][K][C], util [N]) as tc::head_kernel.
#pragma once
#include "head_tc.cuh"

namespace tcko {

using tc::Args;
using tc::CHUNK_N;
using tc::KBLK;
using tc::TILE_M;

constexpr int NTHREADS = 512;
constexpr int MAX_NC = 4;                                    // H1 <= 512 (TMEM: MAX_NC x 128 columns)
constexpr uint32_t XBLK = 16384;                             // one [128 x 64] bf16 X block
constexpr uint32_t WPART = 8192;                             // this CTA's half of one [128 n x 64 k] W1 block
constexpr uint32_t STAGE = XBLK + MAX_NC * WPART;            // 48 KB
constexpr int NS = 3;
constexpr uint32_t ROFF = 0;
constexpr uint32_t HOFF = ROFF + NS * STAGE;                 // hidden-1 chunk, double buffered: 2 x 32 KB
constexpr uint32_t W2OFF = HOFF + 2 * 32768;                 // layer-2 slices: 2 x 8 KB
constexpr uint32_t POFF = W2OFF + 2 * 8192;
constexpr uint32_t P_B1 = POFF;                              // float [2][128]
constexpr uint32_t P_B2 = P_B1 + 2 * 128 * 4;                // float [2][64]
constexpr uint32_t P_W3 = P_B2 + 2 * 64 * 4;                 // float [2][128]
constexpr uint32_t P_B3 = P_W3 + 2 * 128 * 4;                // float [2][2]
constexpr uint32_t BOFF = P_B3 + 16;
enum { W_FULL = 0, W_EMPTY = 3, W2_FULL = 6, W2_EMPTY = 8, A1_FULL = 10, A1_EMPTY = 11, H1_FULL = 12, H1_EMPTY = 14,
       A2_FULL = 16, XS_FULL = 17, XS_EMPTY = 19, NBARS = 21 };
constexpr uint32_t TOFF = BOFF + NBARS * 8;
constexpr uint32_t SMEM_BYTES = TOFF + 16;
static_assert(SMEM_BYTES <= 232448, "exceeds the 227 KB per-CTA shared memory limit");
static_assert(BOFF % 8 == 0 && STAGE % 1024 == 0, "alignment");
constexpr uint32_t TMEM_COLS = 512;

struct KoArgs {
  Args a;
  uint8_t *xscratch;  // [gridDim.x][2][nkb] blocks of 16 KB
};

__device__ __forceinline__ void stg64(void *p, uint32_t x, uint32_t y) {
  asm volatile("st.global.v2.b32 [%0], {%1, %2};" ::"l"(p), "r"(x), "r"(y) : "memory");
}

template <bool TAB_BF16, bool F16>
__global__ void __launch_bounds__(NTHREADS, 1) head_kernel_ko(const KoArgs ka) {
  const Args &a = ka.a;
  const uint32_t rank = ptx::cluster_ctarank();
  const uint32_t unit0 = blockIdx.x / 2, nunit = gridDim.x / 2;
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sbase = ptx::smem_u32(smem);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  auto bar = [&](int i) -> uint32_t { return sbase + BOFF + 8u * (uint32_t)i; };
  const uint32_t nkb = (uint32_t)a.nkb, nc = (uint32_t)a.nc, K = (uint32_t)a.K;
  uint8_t *xs_mine = ka.xscratch + (size_t)blockIdx.x * 2 * nkb * XBLK;

  if (tid == 0) {
    if (sbase & 1023u) __trap();
    const uint32_t fwd = rank == 0 ? 1u : 0u;  // the peer's relay forwards one arrive per phase (head_tc.cuh)
    for (int s = 0; s < NS; ++s) {
      ptx::mbar_init(bar(W_FULL + s), 1 + fwd);
      ptx::mbar_init(bar(W_EMPTY + s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(bar(W2_FULL + s), 1 + fwd);
      ptx::mbar_init(bar(W2_EMPTY + s), 1);
      ptx::mbar_init(bar(H1_FULL + s), 128 + fwd);
      ptx::mbar_init(bar(H1_EMPTY + s), 1);
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

  const uint32_t nunits = ((uint32_t)a.ntiles + 1) / 2;
  const uint32_t T = nunits > unit0 ? (nunits - unit0 + nunit - 1) / nunit : 0u;  // tiles of this CTA
  auto tile_of = [&](uint32_t i) -> int64_t { return ((int64_t)unit0 + (int64_t)i * nunit) * 2 + rank; };
  const uint32_t l2bytes = 2u * (uint32_t)a.H2 * 128u;  // one layer-2 slice of the pair; this CTA loads half
  const uint32_t stage_tx = XBLK + nc * WPART;

  if (warp == 0) {
    // ------------------------------------------------------------------ stage producer (TMA bulk)
    ptx::setmaxnreg_dec<56>();
    const uint32_t leader = ptx::elect_one();
    const uint32_t ring = sbase + ROFF;
    uint32_t s = 0, ph = 0, w2n = 0;
    auto issue_w2 = [&](uint32_t l, uint32_t c) __attribute__((always_inline)) {
      const uint32_t slot = w2n & 1u;
      ptx::mbar_wait(bar(W2_EMPTY + slot), ((w2n >> 1) & 1u) ^ 1u);
      if (leader) {
        const uint8_t *src = a.w2img + (size_t)(l * nc + c) * l2bytes;
        const uint32_t piece = l2bytes / 4;  // H2/2 rows x 128 B of one k-block
        const uint32_t dst = sbase + W2OFF + slot * 8192u;
        ptx::mbar_expect_tx(bar(W2_FULL + slot), 2 * piece);
        ptx::bulk_g2s(dst, src + rank * piece, piece, bar(W2_FULL + slot));
        ptx::bulk_g2s(dst + piece, src + 2 * piece + rank * piece, piece, bar(W2_FULL + slot));
      }
      __syncwarp();
      ++w2n;
    };
    const uint32_t early = nkb < (uint32_t)NS ? nkb : (uint32_t)NS;  // stages issued before the first W2 slices
    for (uint32_t i = 0; i < T; ++i) {
      ptx::mbar_wait(bar(XS_FULL + (i & 1u)), (i >> 1) & 1u);  // this tile's X images are in the scratch
      const uint8_t *xsrc = xs_mine + (size_t)(i & 1u) * nkb * XBLK;
      for (uint32_t l = 0; l < K; ++l) {
        for (uint32_t kb = 0; kb < nkb; ++kb) {
          if (kb == early) {
            issue_w2(l, 0);
            if (nc > 1) issue_w2(l, 1);
          }
          ptx::mbar_wait(bar(W_EMPTY + s), ph ^ 1u);
          if (leader) {
            const uint32_t dst = ring + s * STAGE;
            ptx::mbar_expect_tx(bar(W_FULL + s), stage_tx);
            ptx::bulk_g2s(dst, xsrc + (size_t)kb * XBLK, XBLK, bar(W_FULL + s));
            for (uint32_t c = 0; c < nc; ++c)
              ptx::bulk_g2s(dst + XBLK + c * WPART,
                            a.w1img + ((size_t)(l * nc + c) * nkb + kb) * tc::STAGE_BYTES + rank * WPART, WPART,
                            bar(W_FULL + s));
          }
          __syncwarp();
          if (++s == NS) { s = 0; ph ^= 1u; }
        }
        if (nkb <= early) {
          issue_w2(l, 0);
          if (nc > 1) issue_w2(l, 1);
        }
        for (uint32_t c = 2; c < nc; ++c) issue_w2(l, c);
      }
    }
  } else if (warp == 1 && rank == 0) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA)
    ptx::setmaxnreg_dec<56>();
    const uint32_t leader = ptx::elect_one();
    const uint32_t idesc1 = F16 ? ptx::umma_idesc_f16(2 * TILE_M, CHUNK_N) : ptx::umma_idesc_bf16(2 * TILE_M, CHUNK_N);
    const uint32_t idesc2 = F16 ? ptx::umma_idesc_f16(2 * TILE_M, a.H2) : ptx::umma_idesc_bf16(2 * TILE_M, a.H2);
    const uint32_t kb2_step = ((uint32_t)a.H2 * 64u) >> 4;  // this CTA's half of a layer-2 k-block, 16-byte units
    const uint64_t rdesc0 = ptx::umma_desc_sw128(sbase + ROFF);
    const uint64_t hdesc0 = ptx::umma_desc_sw128(sbase + HOFF);
    const uint64_t w2desc0 = ptx::umma_desc_sw128(sbase + W2OFF);
    uint32_t s = 0, ph = 0, w2n = 0, hq = 0, v = 0;
    for (uint32_t i = 0; i < T; ++i) {
      for (uint32_t l = 0; l < K; ++l, ++v) {
        ptx::mbar_wait(bar(A1_EMPTY), (v & 1u) ^ 1u);  // the previous learner's accumulators are drained
        ptx::tc_fence_after();
        for (uint32_t kb = 0; kb < nkb; ++kb) {
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
            if (l == K - 1 && kb == nkb - 1) ptx::mbar_arrive(bar(XS_EMPTY + (i & 1u)));  // last load of the tile landed
          }
          __syncwarp();
          if (++s == NS) { s = 0; ph ^= 1u; }
        }
        if (leader) ptx::umma2_commit_mc(bar(A1_FULL));
        __syncwarp();
        for (uint32_t c = 0; c < nc; ++c, ++hq, ++w2n) {
          const uint32_t hb = hq & 1u, slot = w2n & 1u;
          ptx::mbar_wait(bar(H1_FULL + hb), (hq >> 1) & 1u);
          ptx::mbar_wait(bar(W2_FULL + slot), (w2n >> 1) & 1u);
          ptx::tc_fence_after();
          if (leader) {
#pragma unroll
            for (uint32_t kb2 = 0; kb2 < 2; ++kb2) {
              const uint64_t ad = hdesc0 + (uint64_t)hb * (32768u >> 4) + kb2 * (16384u >> 4);
              const uint64_t bd = w2desc0 + (uint64_t)slot * (8192u >> 4) + kb2 * kb2_step;
#pragma unroll
              for (uint32_t k = 0; k < 4; ++k)  // layer-2 accumulator: the (drained) first columns of chunk 0
                ptx::umma2_bf16(tmem, ad + 2 * k, bd + 2 * k, idesc2, (c | kb2 | k) != 0u);
            }
            ptx::umma2_commit_mc(bar(W2_EMPTY + slot));
            ptx::umma2_commit_mc(bar(H1_EMPTY + hb));
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
    uint32_t s = 0, ph = 0, w2n = 0, hq = 0, v = 0;
    for (uint32_t i = 0; i < T; ++i) {
      for (uint32_t l = 0; l < K; ++l, ++v) {
        if (v > 0) relay(A1_EMPTY, (v & 1u) ^ 1u);
        for (uint32_t kb = 0; kb < nkb; ++kb) {
          relay(W_FULL + s, ph);
          if (leader && l == K - 1 && kb == nkb - 1) ptx::mbar_arrive(bar(XS_EMPTY + (i & 1u)));
          __syncwarp();
          if (++s == NS) { s = 0; ph ^= 1u; }
        }
        for (uint32_t c = 0; c < nc; ++c, ++hq, ++w2n) {
          relay(H1_FULL + (hq & 1u), (hq >> 1) & 1u);
          relay(W2_FULL + (w2n & 1u), (w2n >> 1) & 1u);
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
    float rows[2 * ALINK_MAX_K];
    uint32_t hq = 0, v = 0;
    for (uint32_t i = 0; i < T; ++i) {
      for (uint32_t l = 0; l < K; ++l, ++v) {
        const uint32_t lb = l & 1u;
        if (et < a.H2) b2s[lb * 64 + et] = __ldg(a.b2 + (size_t)l * a.H2 + et);
        if (et < a.H2 * a.C) w3s[lb * 128 + et] = __ldg(a.W3 + (size_t)l * a.H2 * a.C + et);
        if (et < a.C) b3s[lb * 2 + et] = __ldg(a.b3 + (size_t)l * a.C + et);
        ptx::mbar_wait(bar(A1_FULL), v & 1u);  // all chunk accumulators of this learner are complete
        ptx::tc_fence_after();
        for (uint32_t c = 0; c < nc; ++c, ++hq) {
          const uint32_t hb = hq & 1u;
          b1s[hb * 128 + et] = __ldg(a.b1 + (size_t)l * a.H1 + c * CHUNK_N + et);
          asm volatile("bar.sync 1, 128;" ::: "memory");
          ptx::mbar_wait(bar(H1_EMPTY + hb), ((hq >> 1) & 1u) ^ 1u);  // layer 2 has read this h1 buffer
#pragma unroll 1
          for (uint32_t jj = 0; jj < 4; ++jj) {
            uint32_t r[32];
            ptx::tmem_ld32(trow + c * CHUNK_N + jj * 32, r);
            ptx::tc_wait_ld();
            const float4 *bias = reinterpret_cast<const float4 *>(b1s + hb * 128 + jj * 32);
            const uint32_t hbase = sbase + HOFF + hb * 32768u + (jj >> 1) * 16384u;
#pragma unroll
            for (uint32_t e = 0; e < 4; ++e) {
              const float4 ba = bias[2 * e], bb = bias[2 * e + 1];
              const float x0 = fmaxf(__fadd_rn(__uint_as_float(r[e * 8 + 0]), ba.x), 0.0f);
              const float x1 = fmaxf(__fadd_rn(__uint_as_float(r[e * 8 + 1]), ba.y), 0.0f);
              const float x2 = fmaxf(__fadd_rn(__uint_as_float(r[e * 8 + 2]), ba.z), 0.0f);
              const float x3 = fmaxf(__fadd_rn(__uint_as_float(r[e * 8 + 3]), ba.w), 0.0f);
              const float x4 = fmaxf(__fadd_rn(__uint_as_float(r[e * 8 + 4]), bb.x), 0.0f);
              const float x5 = fmaxf(__fadd_rn(__uint_as_float(r[e * 8 + 5]), bb.y), 0.0f);
              const float x6 = fmaxf(__fadd_rn(__uint_as_float(r[e * 8 + 6]), bb.z), 0.0f);
              const float x7 = fmaxf(__fadd_rn(__uint_as_float(r[e * 8 + 7]), bb.w), 0.0f);
              tc::sts128(hbase + ptx::sw128_offset((uint32_t)et, (jj & 1u) * 4 + e), tc::pack16<F16>(x0, x1),
                         tc::pack16<F16>(x2, x3), tc::pack16<F16>(x4, x5), tc::pack16<F16>(x6, x7));
            }
          }
          ptx::fence_proxy_async_smem();
          ptx::tc_fence_before();
          ptx::mbar_arrive(bar(H1_FULL + hb));
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
            const float h = fmaxf(__fadd_rn(__uint_as_float(r[e]), b2s[lb * 64 + jn]), 0.0f);
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
          if (a.probs && n < a.N)
            *reinterpret_cast<float2 *>(a.probs + ((size_t)n * a.K + l) * 2) = make_float2(p0, p1);
        } else {
          p1 = alink_sigmoid(__fadd_rn(z0, b3s[lb * 2 + 0]));
          p0 = __fsub_rn(1.0f, p1);  // code/keras_wrapper.py:268
          if (a.probs && n < a.N) a.probs[(size_t)n * a.K + l] = p1;
        }
        rows[2 * l] = p0;
        rows[2 * l + 1] = p1;
        if (l == K - 1 && a.util && n < a.N) {
          float c0, c1;
          alink_consensus2(rows, a.K, &c0, &c1);
          a.util[n] = alink_utility2(a.measure, rows, a.K, c0, c1);
        }
      }
    }
  } else if (warp >= 8) {
    // ------------------------------------------------------------------ gather producers -> X images in the scratch
    // Same mapping as tc::head_kernel (16 lanes per 256-byte row segment, 8 rows per thread and k-block,
    // loads two k-blocks deep), but the destination is this CTA's global scratch, one tile AHEAD of the
    // tensor pipe: block kb of tile i at xs_mine[(i & 1)][kb], already in the swizzled shared-memory image.
    ptx::setmaxnreg_inc<168>();
    const int pt = tid - 256, sub = pt & 15, rbase = pt >> 4;
    const uint32_t nkb_feat = (uint32_t)a.D / KBLK;
    uint32_t ra[8], rb[8];
    auto load_rows = [&](uint32_t i) __attribute__((always_inline)) {
      const int64_t tile = tile_of(i);
#pragma unroll
      for (int ii = 0; ii < 8; ++ii) {
        int64_t n = tile * TILE_M + rbase + 16 * ii;
        if (n >= a.N) n = a.N - 1;
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
        const size_t oa = (size_t)ra[ii] * (uint32_t)a.D + col, ob = (size_t)rb[ii] * (uint32_t)a.D + col;
        if (TAB_BF16) {
          va[ii] = tc::ldg_bf4((const __nv_bfloat16 *)a.ta + oa);
          vb[ii] = tc::ldg_bf4((const __nv_bfloat16 *)a.tb + ob);
        } else {
          va[ii] = tc::ldg_f4((const float *)a.ta + oa);
          vb[ii] = tc::ldg_f4((const float *)a.tb + ob);
        }
      }
    };
    auto finish = [&](const float4(&va)[8], const float4(&vb)[8], uint32_t i, uint32_t kb) __attribute__((always_inline)) {
      const bool had = kb >= nkb_feat;
      uint8_t *xb = xs_mine + ((size_t)(i & 1u) * nkb + kb) * XBLK;
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
        stg64(xb + ptx::sw128_offset(r, (uint32_t)sub >> 1) + ((uint32_t)sub & 1u) * 8u, tc::pack16<F16>(x0, x1),
              tc::pack16<F16>(x2, x3));
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

}  // namespace tcko

// Wide-feature scoring (D_in > 512 or when forced): same contract as tc_score.
static int tc_score_ko(alink_ctx *ctx, const HeadGroup &g, const tc::Args &a0, bool tb16, bool f16) {
  if (ctx->sm_count < 2) return alink_fail(ctx, ALINK_E_UNSUPPORTED, "the wide-feature tensor-core head needs CTA pairs");
  tcko::KoArgs ka;
  ka.a = a0;
  const int units = (a0.ntiles + 1) / 2, max_units = (ctx->sm_count - ctx->reserved_sms) / 2;
  const unsigned grid = 2u * (unsigned)(units < max_units ? units : max_units);
  ALINK_TRY(dev_reserve(ctx, ctx->xscratch, (size_t)grid * 2 * a0.nkb * tcko::XBLK));
  ka.xscratch = (uint8_t *)ctx->xscratch.p;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(tcko::NTHREADS);
  cfg.dynamicSmemBytes = tcko::SMEM_BYTES;
  cfg.stream = ctx->stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  void (*kern)(const tcko::KoArgs) = f16 ? tcko::head_kernel_ko<false, true>
                                     : tb16 ? tcko::head_kernel_ko<true, false> : tcko::head_kernel_ko<false, false>;
  ALINK_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tcko::SMEM_BYTES));
  ALINK_CUDA(ctx, cudaLaunchKernelEx(&cfg, kern, ka));
  ALINK_LAUNCHED(ctx);
  (void)g;
  return ALINK_OK;
}
