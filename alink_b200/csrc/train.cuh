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

// out[n][j] = act(sum_k X[n][k] W[k][j] + b[j]); one thread per output, k ascending (the oracle's chain)
__global__ void __launch_bounds__(256) train_dense_fwd_kernel(const float *__restrict__ X, int Kd, const float *__restrict__ W,
                                                              const float *__restrict__ b, int J, int nb, float *__restrict__ out) {
  for (int e = blockIdx.x * 256 + threadIdx.x; e < nb * J; e += gridDim.x * 256) {
    const int n = e / J, j = e - n * J;
    const float *x = X + (size_t)n * Kd;
    float acc = 0.0f;
    for (int k = 0; k < Kd; ++k) acc = __fmaf_rn(x[k], W[(size_t)k * J + j], acc);
    acc = __fadd_rn(acc, b[j]);
    out[e] = acc > 0.0f ? acc : 0.0f;
  }
}

// last Dense + activation + loss + metric + d loss / d logits (already divided by the batch size)
__global__ void __launch_bounds__(256) train_head_kernel(const float *__restrict__ h2, const float *__restrict__ W3,
                                                         const float *__restrict__ b3, int H2, int C, const float *__restrict__ y,
                                                         int nb, float *__restrict__ loss, float *__restrict__ acc,
                                                         float *__restrict__ dz3, float *__restrict__ probs) {
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

// weight gradient + Adadelta in one pass: g[k][j] = sum_n A[n][k] dZ[n][j] (n ascending), then the update
__global__ void __launch_bounds__(256) train_update_w_kernel(const float *__restrict__ A, int Kd, const float *__restrict__ dZ, int J,
                                                             int nb, float *__restrict__ W, float *__restrict__ acc_g,
                                                             float *__restrict__ acc_d, float lr, float rho, float eps) {
  for (int e = blockIdx.x * 256 + threadIdx.x; e < Kd * J; e += gridDim.x * 256) {
    const int k = e / J, j = e - k * J;
    float g = 0.0f;
    for (int n = 0; n < nb; ++n) g = __fmaf_rn(A[(size_t)n * Kd + k], dZ[(size_t)n * J + j], g);
    float a = acc_g[e], d = acc_d[e];
    W[e] = alink_adadelta(W[e], g, &a, &d, lr, rho, eps);
    acc_g[e] = a;
    acc_d[e] = d;
  }
}

// bias gradient (sum over the batch, n ascending) + Adadelta
__global__ void __launch_bounds__(256) train_update_b_kernel(const float *__restrict__ dZ, int J, int nb, float *__restrict__ b,
                                                             float *__restrict__ acc_g, float *__restrict__ acc_d, float lr,
                                                             float rho, float eps) {
  for (int j = blockIdx.x * 256 + threadIdx.x; j < J; j += gridDim.x * 256) {
    float g = 0.0f;
    for (int n = 0; n < nb; ++n) g = __fadd_rn(g, dZ[(size_t)n * J + j]);
    float a = acc_g[j], d = acc_d[j];
    b[j] = alink_adadelta(b[j], g, &a, &d, lr, rho, eps);
    acc_g[j] = a;
    acc_d[j] = d;
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
