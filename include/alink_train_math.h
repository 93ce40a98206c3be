/* alink_train_math.h — per-element arithmetic of the finetune ("teach") step, shared by gcc (oracle/) and nvcc
 * (csrc/train.cuh) like alink_math.h, so that the CUDA step and its CPU restatement are the same expression trees.
 *
 * What it restates (SURVEY.md 8f.2 — the step right after query selection):
 *   loss ...... /root/reference/code/siamese.py:35 `loss="binary_crossentropy"` on the head's activation output;
 *               [external: Keras 2.1.2 backend binary_crossentropy, from_logits=False] clips the probability to
 *               [eps, 1 - eps] (eps = K.epsilon() = 1e-7) and returns -(y log p + (1 - y) log(1 - p)), averaged
 *               over the outputs of a sample and then over the batch. (TF evaluates the same value through
 *               logits = log(p / (1 - p)); restated here directly on the clipped probability.)
 *   metric .... `metrics=['accuracy']` with that loss = binary accuracy: mean over outputs of round(p) == y.
 *   optimizer . code/siamese.py:35 `Adadelta(self.learningRate)`; [external: Keras 2.1.2 optimizers.Adadelta,
 *               rho = 0.95, epsilon = 1e-7, decay = 0]:
 *                   a <- rho a + (1 - rho) g^2;  u = g sqrt(d + eps) / sqrt(a + eps);  p <- p - lr u;
 *                   d <- rho d + (1 - rho) u^2
 *   labels .... code/siamese.py:56 `to_categorical(Y, num_classes=2)`: y -> (1 - y, y) for the softmax-2 head;
 *               the sigmoid-1 head (siamese3.py) takes y as it is.
 * Every operation below is a single IEEE fp32 operation (no contraction: gcc -ffp-contract=off, nvcc -fmad=false).
 */
#ifndef ALINK_TRAIN_MATH_H
#define ALINK_TRAIN_MATH_H

#include "alink_math.h"

#define ALINK_KERAS_EPS 1e-7f

/* binary cross-entropy of one output; *dldp = d loss / d p (0 where the clip is active, as tf.clip_by_value) */
ALINK_HD float alink_bce(float p, float y, float *dldp) {
  const float lo = ALINK_KERAS_EPS, hi = ALINK_SUB(1.0f, ALINK_KERAS_EPS);
  float pc = p;
  int clipped = 0;
  if (!(pc >= lo)) { pc = lo; clipped = 1; }
  if (pc > hi) { pc = hi; clipped = 1; }
  const float q = ALINK_SUB(1.0f, pc);
  const float ny = ALINK_SUB(1.0f, y);
  const float loss = -ALINK_ADD(ALINK_MUL(y, alink_logf(pc)), ALINK_MUL(ny, alink_logf(q)));
  *dldp = clipped ? 0.0f : ALINK_SUB(ALINK_DIV(ny, q), ALINK_DIV(y, pc));
  return loss;
}

/* Output layer of one sample: logits -> probabilities, loss, binary accuracy, d loss / d logits (NOT yet divided
 * by the batch size). C = 2: softmax + one-hot labels (1 - y, y); C = 1: sigmoid. Returns the sample's loss. */
ALINK_HD float alink_head_loss(int C, float z0, float z1, float y, float *p0, float *p1, float *acc, float *dz0,
                               float *dz1) {
  if (C == 2) {
    alink_softmax2(z0, z1, p0, p1);
    float g0, g1;
    const float y0 = ALINK_SUB(1.0f, y);
    const float l0 = alink_bce(*p0, y0, &g0), l1 = alink_bce(*p1, y, &g1);
    g0 = ALINK_MUL(g0, 0.5f); /* mean over the two outputs */
    g1 = ALINK_MUL(g1, 0.5f);
    const float dot = ALINK_ADD(ALINK_MUL(*p0, g0), ALINK_MUL(*p1, g1)); /* softmax Jacobian: dz_i = p_i (g_i - sum_c p_c g_c) */
    *dz0 = ALINK_MUL(*p0, ALINK_SUB(g0, dot));
    *dz1 = ALINK_MUL(*p1, ALINK_SUB(g1, dot));
    const float r0 = *p0 > 0.5f ? 1.0f : 0.0f, r1 = *p1 > 0.5f ? 1.0f : 0.0f; /* K.round: half to even -> 0.5 rounds to 0 */
    *acc = ALINK_MUL(0.5f, ALINK_ADD(r0 == y0 ? 1.0f : 0.0f, r1 == y ? 1.0f : 0.0f));
    return ALINK_MUL(0.5f, ALINK_ADD(l0, l1));
  }
  *p1 = alink_sigmoid(z0);
  *p0 = ALINK_SUB(1.0f, *p1);
  float g;
  const float l = alink_bce(*p1, y, &g);
  *dz0 = ALINK_MUL(g, ALINK_MUL(*p1, ALINK_SUB(1.0f, *p1)));
  *dz1 = 0.0f;
  *acc = ((*p1 > 0.5f ? 1.0f : 0.0f) == y) ? 1.0f : 0.0f;
  return l;
}

/* one Adadelta update of one parameter; returns the new parameter, updates the two accumulators in place */
ALINK_HD float alink_adadelta(float p, float g, float *a, float *d, float lr, float rho, float eps) {
  const float omr = ALINK_SUB(1.0f, rho);
  const float na = ALINK_ADD(ALINK_MUL(rho, *a), ALINK_MUL(omr, ALINK_MUL(g, g)));
#if defined(__CUDA_ARCH__)
  const float num = __fsqrt_rn(ALINK_ADD(*d, eps)), den = __fsqrt_rn(ALINK_ADD(na, eps));
#else
  const float num = sqrtf(ALINK_ADD(*d, eps)), den = sqrtf(ALINK_ADD(na, eps));
#endif
  const float u = ALINK_DIV(ALINK_MUL(g, num), den);
  *a = na;
  *d = ALINK_ADD(ALINK_MUL(rho, *d), ALINK_MUL(omr, ALINK_MUL(u, u)));
  return ALINK_SUB(p, ALINK_MUL(lr, u));
}

#endif /* ALINK_TRAIN_MATH_H */
