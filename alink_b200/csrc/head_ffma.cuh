// head_ffma.cuh — FP32_EXACT siamese head on CUDA cores (FFMA), bit-identical to oracle/alink_oracle.c.
//
// Replaces, per learner, the Keras graph of /root/reference/code/siamese.py:24-33
// (|l-r| -> Dense(512,relu) -> Dense(64,relu) -> Dense(2)+softmax) and siamese3.py:18-26 (-> Dense(1,sigmoid)).
//
// Every dense output is ONE sequential chain acc = fmaf(x[k], W[k][j], acc), k ascending, acc0 = 0,
// then acc + b[j], then relu — the order the oracle defines. The tiling below parallelises over pairs
// (rows) and output neurons (columns) only, never over k, so the result is bit-exact by construction.
//
// Kernel: 128 x BN output tile per CTA, BK = 16 (layer 1) or 8 (layer 2: its epilogue tile needs the shared
// memory), 256 threads, 8 x (BN/16) accumulators per thread, double-buffered shared memory, one __syncthreads
// per k-chunk.
//   A operand: AMODE 0 dense fp32 rows; AMODE 1 index-driven gather of two table rows -> |a-b|;
//              AMODE 2 gather -> [|a-b| ; a*b] (fan-in 2D, north_star's opt-in Hadamard feature).
//   epilogue : EPI 0 bias + relu -> hidden activations (fp32, global);
//              EPI 1 bias + relu -> shared -> last Dense (H2 x C) + softmax/sigmoid -> probabilities.
// Roofline: CUDA-core FFMA (2 * 128 lanes * 148 SMs * clock); this is the correctness mode — the
// tensor-core path is head_tc.cuh.
#pragma once
#include "common.cuh"

struct FfmaArgs {
  // A operand
  const float *A;              // AMODE 0: A + learner * a_learner_stride + row * lda
  int64_t lda, a_learner_stride;
  const float *ta, *tb;        // AMODE 1/2: fp32 tables, D columns
  const int32_t *ia, *ib;      // pair -> row (nullptr = identity, or the all-pairs enumeration when G > 0)
  uint32_t G;                  // all-pairs: gallery rows; pair p = (p / G, p % G)
  int64_t pair0;               // all-pairs: enumeration index of pair 0
  int D;
  int64_t row0;                // first pair of this chunk
  int M;                       // rows (pairs) in this chunk
  const unsigned int *m_dev;   // optional: the row count lives on the device (M is then its capacity: the grid)
  int Kdim;                    // fan-in
  int Ncols;                   // fan-out of this layer
  int ntiles;                  // column tiles
  const float *const *W;       // [learner] -> (Kdim, Ncols) row-major
  const float *const *bias;    // [learner] -> (Ncols)
  // EPI 0
  float *H;                    // H + learner * h_learner_stride + row * Ncols
  int64_t h_learner_stride;
  // EPI 1
  const float *const *W3;      // [learner] -> (Ncols, C)
  const float *const *b3;
  int C, act, K;
  float *probs;                // [(row0 + row) * K + learner] * C + c
};

// BM = 128 (8 rows per thread) for bulk scoring; BM = 64 (4 rows per thread, 3 CTAs per SM) for the short candidate lists of
// FP32_REFINE, where twice as many CTAs fill the 148 SMs better than one wave of large tiles.
template <int BN, int AMODE, int EPI, int BK = (EPI == 0 ? 16 : 8), int BM = 128>
__global__ void __launch_bounds__(256, (BM == 64 ? 3 : 2)) ffma_layer_kernel(const FfmaArgs a) {
  constexpr int TM = BM / 16, TN = BN / 16;
  constexpr int TPR = 256 / BM;                                // threads sharing one A row
  constexpr int NA = BK / (4 * TPR);                           // float4 of A per thread and chunk
  static_assert(BM == 128 || BM == 64, "row tile");
  static_assert(NA >= 1, "BK too small for this row tile");
  constexpr int NB = (BK * BN / 4 + 255) / 256;                // float4 of B per thread and chunk
  __shared__ __align__(16) float As[2][BK][BM];
  __shared__ __align__(16) float Bs[2][BK][BN];

  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int ntile = blockIdx.x % a.ntiles;
  const int learner = blockIdx.x / a.ntiles;
  const int mtile = blockIdx.y;
  const int n0 = ntile * BN;
  int Mrows = a.M;
  if (a.m_dev) {
    const unsigned int md = __ldg(a.m_dev);
    if (md < (unsigned int)Mrows) Mrows = (int)md;
    if (mtile * BM >= Mrows) return;  // the whole CTA: grids are sized for the capacity
  }

  // ---- A loader state: this thread owns row (tid>>1), k-quads (tid&1)*4*NA + 4*i of every chunk
  const int lrow = tid / TPR, kq = (tid % TPR) * 4 * NA;
  int grow = mtile * BM + lrow;
  if (grow >= Mrows) grow = Mrows - 1;
  const float *pa, *pb = nullptr;
  if constexpr (AMODE == 0) {
    pa = a.A + (size_t)learner * a.a_learner_stride + (size_t)grow * a.lda;
  } else {
    const int64_t pair = a.row0 + grow;
    int64_t ra, rb;
    if (a.ia) {
      ra = (int64_t)__ldg(a.ia + pair);
      rb = (int64_t)__ldg(a.ib + pair);
    } else if (a.G) {  // utilities/generateMatrixDFW.py:28-33: row-major (probe i, every gallery j)
      const int64_t p = a.pair0 + pair;
      ra = p / a.G;
      rb = p - ra * a.G;
    } else {
      ra = rb = pair;
    }
    pa = a.ta + (size_t)ra * a.D;
    pb = a.tb + (size_t)rb * a.D;
  }
  // ---- B loader state: float4 number tid + 256 * i of the [BK][BN] chunk
  constexpr int BTOTAL = BK * BN / 4;
  constexpr int BROW = BN / 4;  // float4 per k-row
  const float *pwl = a.W[learner] + n0;

  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.0f;

  float4 ra4[NA], rb4[NB];
  auto load_chunk = [&](int k0) {
#pragma unroll
    for (int i = 0; i < NA; ++i) {
      if constexpr (AMODE == 0) {
        ra4[i] = __ldg(reinterpret_cast<const float4 *>(pa + k0 + kq + 4 * i));
      } else {
        int kk = k0;
        bool had = false;
        if (AMODE == 2 && k0 >= a.D) {
          kk = k0 - a.D;
          had = true;
        }
        const float4 va = __ldg(reinterpret_cast<const float4 *>(pa + kk + kq + 4 * i));
        const float4 vb = __ldg(reinterpret_cast<const float4 *>(pb + kk + kq + 4 * i));
        if (had) {
          ra4[i].x = __fmul_rn(va.x, vb.x);
          ra4[i].y = __fmul_rn(va.y, vb.y);
          ra4[i].z = __fmul_rn(va.z, vb.z);
          ra4[i].w = __fmul_rn(va.w, vb.w);
        } else {
          ra4[i].x = fabsf(__fsub_rn(va.x, vb.x));  // K.abs(l - r), code/siamese.py:27
          ra4[i].y = fabsf(__fsub_rn(va.y, vb.y));
          ra4[i].z = fabsf(__fsub_rn(va.z, vb.z));
          ra4[i].w = fabsf(__fsub_rn(va.w, vb.w));
        }
      }
    }
#pragma unroll
    for (int i = 0; i < NB; ++i) {
      const int f = tid + 256 * i, kb = f / BROW, n4 = (f % BROW) * 4;
      if (f < BTOTAL && n0 + n4 < a.Ncols)
        rb4[i] = __ldg(reinterpret_cast<const float4 *>(pwl + (size_t)(k0 + kb) * a.Ncols + n4));
      else
        rb4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  auto store_chunk = [&](int buf) {
#pragma unroll
    for (int i = 0; i < NA; ++i) {
      As[buf][kq + 4 * i + 0][lrow] = ra4[i].x;
      As[buf][kq + 4 * i + 1][lrow] = ra4[i].y;
      As[buf][kq + 4 * i + 2][lrow] = ra4[i].z;
      As[buf][kq + 4 * i + 3][lrow] = ra4[i].w;
    }
#pragma unroll
    for (int i = 0; i < NB; ++i) {
      const int f = tid + 256 * i, kb = f / BROW, n4 = (f % BROW) * 4;
      if (f < BTOTAL) *reinterpret_cast<float4 *>(&Bs[buf][kb][n4]) = rb4[i];
    }
  };

  const int nk = a.Kdim / BK;
  load_chunk(0);
  store_chunk(0);
  __syncthreads();
  for (int kc = 0; kc < nk; ++kc) {
    const int buf = kc & 1;
    if (kc + 1 < nk) load_chunk((kc + 1) * BK);
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float av[TM], bv[TN];
      const float4 a0 = *reinterpret_cast<const float4 *>(&As[buf][k][ty * 4]);
      av[0] = a0.x; av[1] = a0.y; av[2] = a0.z; av[3] = a0.w;
      if constexpr (TM == 8) {
        const float4 a1 = *reinterpret_cast<const float4 *>(&As[buf][k][64 + ty * 4]);
        av[4] = a1.x; av[5] = a1.y; av[6] = a1.z; av[7] = a1.w;
      }
      if constexpr (TN == 8) {
        const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[buf][k][tx * 4]);
        const float4 b1 = *reinterpret_cast<const float4 *>(&Bs[buf][k][BN / 2 + tx * 4]);
        bv[0] = b0.x; bv[1] = b0.y; bv[2] = b0.z; bv[3] = b0.w;
        bv[4] = b1.x; bv[5] = b1.y; bv[6] = b1.z; bv[7] = b1.w;
      } else if constexpr (TN == 4) {
        const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[buf][k][tx * 4]);
        bv[0] = b0.x; bv[1] = b0.y; bv[2] = b0.z; bv[3] = b0.w;
      } else {
        const float2 b0 = *reinterpret_cast<const float2 *>(&Bs[buf][k][tx * 2]);
        bv[0] = b0.x; bv[1] = b0.y;
      }
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = __fmaf_rn(av[i], bv[j], acc[i][j]);
    }
    if (kc + 1 < nk) store_chunk(buf ^ 1);
    __syncthreads();
  }

  // ---- epilogue
  const float *bias = a.bias[learner];
  auto col_of = [&](int j) -> int {
    if (TN == 8) return (j < 4) ? tx * 4 + j : BN / 2 + tx * 4 + (j - 4);
    if (TN == 4) return tx * 4 + j;
    return tx * 2 + j;
  };
  auto row_of = [&](int i) -> int { return (i < 4) ? ty * 4 + i : 64 + ty * 4 + (i - 4); };

  if constexpr (EPI == 0) {
    float *H = a.H + (size_t)learner * a.h_learner_stride;
#pragma unroll
    for (int i = 0; i < TM; ++i) {
      const int row = mtile * BM + row_of(i);
      if (row >= Mrows) continue;
#pragma unroll
      for (int j0 = 0; j0 < TN; j0 += (TN >= 4 ? 4 : 2)) {
        const int col = n0 + col_of(j0);
        if (col >= a.Ncols) continue;
        float v[4];
#pragma unroll
        for (int j = 0; j < (TN >= 4 ? 4 : 2); ++j) {
          float t = __fadd_rn(acc[i][j0 + j], __ldg(bias + col + j));
          v[j] = t > 0.0f ? t : 0.0f;
        }
        if (TN >= 4)
          *reinterpret_cast<float4 *>(H + (size_t)row * a.Ncols + col) = make_float4(v[0], v[1], v[2], v[3]);
        else
          *reinterpret_cast<float2 *>(H + (size_t)row * a.Ncols + col) = make_float2(v[0], v[1]);
      }
    }
  } else {
    // hidden-2 tile -> shared, then the last Dense + activation, one thread per pair
    __shared__ float h2s[BM][BN + 1];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
      for (int j = 0; j < TN; ++j) {
        const int col = col_of(j);
        float t = __fadd_rn(acc[i][j], __ldg(bias + col));
        h2s[row_of(i)][col] = t > 0.0f ? t : 0.0f;
      }
    __syncthreads();
    if (tid < BM) {
      const int row = mtile * BM + tid;
      if (row < Mrows) {
        const float *w3 = a.W3[learner];
        const float *b3 = a.b3[learner];
        float *out = a.probs + ((size_t)(a.row0 + row) * a.K + learner) * a.C;
        if (a.C == 2) {
          float z0 = 0.0f, z1 = 0.0f;
#pragma unroll 8
          for (int j = 0; j < BN; ++j) {
            const float h = h2s[tid][j];
            z0 = __fmaf_rn(h, __ldg(w3 + 2 * j + 0), z0);
            z1 = __fmaf_rn(h, __ldg(w3 + 2 * j + 1), z1);
          }
          z0 = __fadd_rn(z0, __ldg(b3 + 0));
          z1 = __fadd_rn(z1, __ldg(b3 + 1));
          float p0, p1;
          alink_softmax2(z0, z1, &p0, &p1);
          *reinterpret_cast<float2 *>(out) = make_float2(p0, p1);
        } else {
          float z = 0.0f;
#pragma unroll 8
          for (int j = 0; j < BN; ++j) z = __fmaf_rn(h2s[tid][j], __ldg(w3 + j), z);
          z = __fadd_rn(z, __ldg(b3));
          out[0] = alink_sigmoid(z);
        }
      }
    }
  }
}

// Launch both layers for one chunk of pairs. H1 workspace layout [K][Mc][H1].
static int ffma_head_chunk(alink_ctx *ctx, const HeadGroup &g, const float *ta, const float *tb, int D,
                           const int32_t *ia, const int32_t *ib, int64_t row0, int Mc, float *h1ws,
                           int64_t h1_learner_stride, float *probs, const unsigned int *m_dev = nullptr) {
  FfmaArgs a;
  memset(&a, 0, sizeof(a));
  a.ta = ta; a.tb = tb; a.ia = ia; a.ib = ib; a.D = D;
  a.G = ia ? 0u : ctx->ap_G; a.pair0 = ctx->ap_pair0;
  a.row0 = row0; a.M = Mc; a.m_dev = m_dev; a.Kdim = g.D_in; a.Ncols = g.H1;
  a.ntiles = (g.H1 + 127) / 128;
  a.W = g.pW1; a.bias = g.pb1;
  a.H = h1ws; a.h_learner_stride = h1_learner_stride;
  a.K = g.K; a.C = g.C; a.act = g.act;
  // (measured: the 64-row tile is SLOWER for FP32_REFINE's ~1.5 k exact rows — 151 + 50 us against 125 + 46 us: half the
  // register reuse per shared-memory load outweighs the better fill; kept selectable for other shapes)
  static const int small_env = [] { const char *e = getenv("ALINK_FFMA_SMALL"); return e ? atoi(e) : 0; }();
  const bool small = m_dev != nullptr && small_env == 1;
  if (small) {
    dim3 grid1(a.ntiles * g.K, (Mc + 63) / 64);
    if (g.feat_mode == ALINK_FEAT_ABSDIFF)
      ffma_layer_kernel<128, 1, 0, 16, 64><<<grid1, 256, 0, ctx->stream>>>(a);
    else
      ffma_layer_kernel<128, 2, 0, 16, 64><<<grid1, 256, 0, ctx->stream>>>(a);
  } else {
    dim3 grid1(a.ntiles * g.K, (Mc + 127) / 128);
    if (g.feat_mode == ALINK_FEAT_ABSDIFF)
      ffma_layer_kernel<128, 1, 0><<<grid1, 256, 0, ctx->stream>>>(a);
    else
      ffma_layer_kernel<128, 2, 0><<<grid1, 256, 0, ctx->stream>>>(a);
  }
  ALINK_LAUNCHED(ctx);

  FfmaArgs b;
  memset(&b, 0, sizeof(b));
  b.A = h1ws; b.lda = g.H1; b.a_learner_stride = h1_learner_stride;
  b.row0 = row0; b.M = Mc; b.m_dev = m_dev; b.Kdim = g.H1; b.Ncols = g.H2; b.ntiles = 1;
  b.W = g.pW2; b.bias = g.pb2;
  b.W3 = g.pW3; b.b3 = g.pb3; b.C = g.C; b.act = g.act; b.K = g.K; b.probs = probs;
  dim3 grid2(g.K, small ? (Mc + 63) / 64 : (Mc + 127) / 128);
  if (g.H2 == 64 && small)
    ffma_layer_kernel<64, 0, 1, 16, 64><<<grid2, 256, 0, ctx->stream>>>(b);
  else if (g.H2 == 32 && small)
    ffma_layer_kernel<32, 0, 1, 16, 64><<<grid2, 256, 0, ctx->stream>>>(b);
  else if (g.H2 == 64)
    ffma_layer_kernel<64, 0, 1><<<grid2, 256, 0, ctx->stream>>>(b);
  else if (g.H2 == 32)
    ffma_layer_kernel<32, 0, 1><<<grid2, 256, 0, ctx->stream>>>(b);
  else
    return alink_fail(ctx, ALINK_E_UNSUPPORTED, "FP32_EXACT head needs H2 in {32, 64}, got %d", g.H2);
  ALINK_LAUNCHED(ctx);
  return ALINK_OK;
}
