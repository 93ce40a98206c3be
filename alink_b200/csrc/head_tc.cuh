// head_tc.cuh — BF16_TC siamese head on tcgen05 tensor cores (placeholder until the UMMA path lands).
#pragma once
#include "common.cuh"
static int tc_prepare_heads(alink_ctx *ctx, HeadGroup &g) {
  (void)g;
  return alink_fail(ctx, ALINK_E_UNSUPPORTED, "BF16_TC mode not built yet");
}
static int tc_score(alink_ctx *ctx, const HeadGroup &g, const Table &A, const Table &B, const int32_t *ia,
                    const int32_t *ib, int64_t N, float *probs) {
  (void)g; (void)A; (void)B; (void)ia; (void)ib; (void)N; (void)probs;
  return alink_fail(ctx, ALINK_E_UNSUPPORTED, "BF16_TC mode not built yet");
}
