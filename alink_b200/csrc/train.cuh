// train.cuh — one finetune ("teach") minibatch step of a siamese head on the GPU.
//
// Replaces, for one learner, what Keras runs per batch inside /root/reference/code/siamese.py:52-58
// (`finetune`: fit(batch_size=16) on the pairs the query step selected, code/ALINK.py:250) and :86-112
// (`train_on_batch` / `test_on_batch`): forward of the head (siamese.py:24-33), binary cross-entropy + binary
// accuracy, backward through the three Dense layers, Adadelta update (siamese.py:35). The per-element
// arithmetic is include/alink_train_math.h (shared with the oracle); every reduction here is ONE sequential
// fmaf chain in a fixed order (k ascending for the forward, j ascending for the input gradients, batch
// index ascending for the weight gradients), so the step is bit-identical to oracle/alink_oracle.c.
// A finetune batch is 16 pairs (ALINK.py:250): the step is launch-latency-bound, not a roofline kernel; the
// layout exists so the weights never leave the device between the query step and the next one.
#pragma once
#include "common.cuh"
#include "../../include/alink_train_math.h"

// X[n][k] = |a - b| (and a * b for the Hadamard half), the head's input (code/siamese.py:27)
__global__ void __launch_bounds__(256) train_feat_kernel(const float *__restrict__ ta, const float *__restrict__ tb,
                                                         const int32_t *__restrict__ ia, const int32_t *__restrict__ ib,
                                                         int D, int hadamard, int nb, float *__restrict__ X) {
  const int Din = hadamard ? 2 * D : D;
  for (int e = blockIdx.x * 256 + threadIdx.x; e < nb * Din; e += gridDim.x * 256) {
    const int n = e / Din, k = e - n * Din;
    const int64_t ra = ia ? (int64_t)ia[n] : n, rb = ib ? (int64_t)ib[n] : n;
    const int kk = k < D ? k : k - D;
    const float a = ta[ra * D + kk], b = tb[rb * D + kk];
    X[e] = k < D ? fabsf(__fsub_rn(a, b)) : __fmul_rn(a, b);
  }
}

// out[n][j] = act(sum_k X[n][k] W[k][j] + b[j]); one thread per output, k ascending (the oracle's chain, bit-identical).
// A CTA owns an 8-row x 32-column tile of the output and walks k in slabs of 64 staged in shared memory: the slab's loads
// (8 coalesced weight loads + 2 input loads per thread) are all in flight together, so a slab costs ONE L2 latency instead
// of one per FMA — with a load in front of every FMA of the chain the layer took ~45 us for a batch of 16 (round 2: 5 us).
constexpr int TD_ROWS = 8, TD_COLS = 32, TD_K = 64;
__global__ void __launch_bounds__(256) train_dense_fwd_kernel(const float *__restrict__ X, int Kd, const float *__restrict__ W,
                                                              const float *__restrict__ b, int J, int nb, float *__restrict__ out) {
  __shared__ float sW[TD_K][TD_COLS];
  __shared__ float sX[TD_ROWS][TD_K];
  const int c = threadIdx.x & 31, r = threadIdx.x >> 5;
  const int ntj = (J + TD_COLS - 1) / TD_COLS, ntn = (nb + TD_ROWS - 1) / TD_ROWS;
  for (int tile = blockIdx.x; tile < ntj * ntn; tile += gridDim.x) {
    const int j0 = (tile % ntj) * TD_COLS, n0 = (tile / ntj) * TD_ROWS;
    const int j = j0 + c, n = n0 + r;
    float acc = 0.0f;
    for (int k0 = 0; k0 < Kd; k0 += TD_K) {
      const int kk = Kd - k0 < TD_K ? Kd - k0 : TD_K;
#pragma unroll
      for (int i = 0; i < TD_K / TD_ROWS; ++i) {
        const int k = r + TD_ROWS * i;
        sW[k][c] = (k < kk && j < J) ? __ldg(W + (size_t)(k0 + k) * J + j) : 0.0f;
      }
#pragma unroll
      for (int i = 0; i < TD_ROWS * TD_K / 256; ++i) {
        const int idx = threadIdx.x + 256 * i, rr = idx / TD_K, k = idx - rr * TD_K;
        sX[rr][k] = (k < kk && n0 + rr < nb) ? X[(size_t)(n0 + rr) * Kd + k0 + k] : 0.0f;
      }
      __syncthreads();
      if (kk == TD_K) {
#pragma unroll 16
        for (int k = 0; k < TD_K; ++k) acc = __fmaf_rn(sX[r][k], sW[k][c], acc);
      } else {
        for (int k = 0; k < kk; ++k) acc = __fmaf_rn(sX[r][k], sW[k][c], acc);
      }
      __syncthreads();
    }
    if (n < nb && j < J) {
      acc = __fadd_rn(acc, b[j]);
      out[(size_t)n * J + j] = acc > 0.0f ? acc : 0.0f;
    }
  }
}

// last Dense + activation + loss + metric + d loss / d logits (already divided by the batch size)
__global__ void __launch_bounds__(256) train_head_kernel(const float *__restrict__ h2, const float *__restrict__ W3,
                                                         const float *__restrict__ b3, int H2, int C, const float *__restrict__ y,
                                                         int nb, float *__restrict__ loss, float *__restrict__ acc,
                                                         float *__restrict__ dz3, float *__restrict__ probs,
                                                         float *__restrict__ mean_out = nullptr) {
  const float inv_nb = __fdiv_rn(1.0f, (float)nb);
  for (int n = blockIdx.x * 256 + threadIdx.x; n < nb; n += gridDim.x * 256) {
    const float *h = h2 + (size_t)n * H2;
    float z0 = 0.0f, z1 = 0.0f;
    for (int j = 0; j < H2; ++j) {
      z0 = __fmaf_rn(h[j], W3[j * C + 0], z0);
      if (C == 2) z1 = __fmaf_rn(h[j], W3[j * C + 1], z1);
    }
    z0 = __fadd_rn(z0, b3[0]);
    if (C == 2) z1 = __fadd_rn(z1, b3[1]);
    float p0, p1, a, d0, d1;
    loss[n] = alink_head_loss(C, z0, z1, y[n], &p0, &p1, &a, &d0, &d1);
    acc[n] = a;
    dz3[n * C + 0] = __fmul_rn(d0, inv_nb);
    if (C == 2) dz3[n * C + 1] = __fmul_rn(d1, inv_nb);
    if (probs) {
      if (C == 2) { probs[n * 2] = p0; probs[n * 2 + 1] = p1; }
      else probs[n] = p1;
    }
  }
  if (mean_out) {  // single-CTA launches (nb <= 256): the batch means in the same launch, same sequential sums as train_mean_kernel
    __syncthreads();
    if (threadIdx.x == 0) {
      float sl = 0.0f, sa = 0.0f;
      for (int n = 0; n < nb; ++n) {
        sl = __fadd_rn(sl, loss[n]);
        sa = __fadd_rn(sa, acc[n]);
      }
      mean_out[0] = __fdiv_rn(sl, (float)nb);
      mean_out[1] = __fdiv_rn(sa, (float)nb);
    }
  }
}

// gradient w.r.t. a relu layer's pre-activation: out[n][k] = act[n][k] > 0 ? sum_j dZ[n][j] W[k][j] : 0 (j ascending)
__global__ void __launch_bounds__(256) train_grad_in_kernel(const float *__restrict__ dZ, int J, const float *__restrict__ W, int Kd,
                                                            const float *__restrict__ act, int nb, float *__restrict__ out) {
  for (int e = blockIdx.x * 256 + threadIdx.x; e < nb * Kd; e += gridDim.x * 256) {
    const int n = e / Kd, k = e - n * Kd;
    float acc = 0.0f;
    if (act[e] > 0.0f) {
      const float *dz = dZ + (size_t)n * J, *w = W + (size_t)k * J;
      for (int j = 0; j < J; ++j) acc = __fmaf_rn(dz[j], w[j], acc);
    }
    out[e] = acc;
  }
}

// All six tensors' gradients + Adadelta updates in ONE launch (they are independent of each other once dz1..dz3 exist):
// a flat index over [W | b] of the three layers. Weights: g[k][j] = sum_n A[n][k] dZ[n][j] (n ascending, fma chain), biases:
// g[j] = sum_n dZ[n][j] (n ascending), then alink_adadelta on the fp32 master and its two accumulators — per element exactly
// what the six separate launches of round 1 computed (bit-identical; tests/test_train_step.py against the oracle).
struct TrainUpdArgs {
  const float *A[3];   // layer input  [nb][Kd]
  const float *dZ[3];  // d loss / d pre-activation [nb][J]
  int Kd[3], J[3];
  float *W[3], *b[3], *ag_w[3], *ad_w[3], *ag_b[3], *ad_b[3];
  int nb;
  float lr, rho, eps;
};

__global__ void __launch_bounds__(256) train_update_all_kernel(const TrainUpdArgs a) {
  const long long n0 = (long long)a.Kd[0] * a.J[0] + a.J[0], n1 = (long long)a.Kd[1] * a.J[1] + a.J[1],
                  n2 = (long long)a.Kd[2] * a.J[2] + a.J[2];
  for (long long f = (long long)blockIdx.x * 256 + threadIdx.x; f < n0 + n1 + n2; f += (long long)gridDim.x * 256) {
    const int l = f < n0 ? 0 : (f < n0 + n1 ? 1 : 2);
    const long long e = f - (l == 0 ? 0 : (l == 1 ? n0 : n0 + n1));
    const int Kd = a.Kd[l], J = a.J[l];
    const float *dZ = a.dZ[l];
    if (e < (long long)Kd * J) {
      const int k = (int)(e / J), j = (int)(e - (long long)k * J);
      const float *A = a.A[l];
      float g = 0.0f;
      for (int n = 0; n < a.nb; ++n) g = __fmaf_rn(A[(size_t)n * Kd + k], dZ[(size_t)n * J + j], g);
      float ga = a.ag_w[l][e], d = a.ad_w[l][e];
      a.W[l][e] = alink_adadelta(a.W[l][e], g, &ga, &d, a.lr, a.rho, a.eps);
      a.ag_w[l][e] = ga;
      a.ad_w[l][e] = d;
    } else {
      const int j = (int)(e - (long long)Kd * J);
      float g = 0.0f;
      for (int n = 0; n < a.nb; ++n) g = __fadd_rn(g, dZ[(size_t)n * J + j]);
      float ga = a.ag_b[l][j], d = a.ad_b[l][j];
      a.b[l][j] = alink_adadelta(a.b[l][j], g, &ga, &d, a.lr, a.rho, a.eps);
      a.ag_b[l][j] = ga;
      a.ad_b[l][j] = d;
    }
  }
}

// batch means of loss and accuracy: sequential sums (n ascending) / nb
__global__ void train_mean_kernel(const float *__restrict__ loss, const float *__restrict__ acc, int nb, float *__restrict__ out) {
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    float sl = 0.0f, sa = 0.0f;
    for (int n = 0; n < nb; ++n) {
      sl = __fadd_rn(sl, loss[n]);
      sa = __fadd_rn(sa, acc[n]);
    }
    out[0] = __fdiv_rn(sl, (float)nb);
    out[1] = __fdiv_rn(sa, (float)nb);
  }
}
