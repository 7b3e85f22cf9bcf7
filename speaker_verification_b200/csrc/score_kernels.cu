// Scoring path (cosine score matrix on tcgen05, trial gather, EER) -- placeholder until
// the tensor-core kernel lands; the entry points fail loudly rather than fall back.
#include "spb_internal.cuh"

namespace spb {
int launch_cosine_scores(spb_context *ctx, const float *, int64_t, const float *, int64_t, int, float *) {
  return fail(ctx, SPB_ERR_UNSUPPORTED, "cosine scores: kernel not built yet");
}
int launch_trial_gather(spb_context *ctx, const float *, int64_t, int64_t, const int32_t *, const int32_t *, int64_t, float *) {
  return fail(ctx, SPB_ERR_UNSUPPORTED, "trial gather: kernel not built yet");
}
int launch_argmax_rows(spb_context *ctx, const float *, int64_t, int64_t, int32_t *) {
  return fail(ctx, SPB_ERR_UNSUPPORTED, "argmax: kernel not built yet");
}
int run_eer(spb_context *ctx, const float *, const int32_t *, int64_t, double *, double *) {
  return fail(ctx, SPB_ERR_UNSUPPORTED, "eer: kernel not built yet");
}
}  // namespace spb

using namespace spb;
extern "C" {
int spb_cosine_scores(spb_context *ctx, const float *, int64_t, const float *, int64_t, int32_t, float *) {
  return fail(ctx, SPB_ERR_UNSUPPORTED, "cosine scores: kernel not built yet");
}
int spb_cosine_scores_dev(spb_context *ctx, const float *E, int64_t m, const float *T, int64_t n, int32_t d, float *S) {
  return launch_cosine_scores(ctx, E, m, T, n, d, S);
}
int spb_trial_scores_dev(spb_context *ctx, const float *S, int64_t m, int64_t n, const int32_t *ti, const int32_t *tj, int64_t nt, float *out) {
  return launch_trial_gather(ctx, S, m, n, ti, tj, nt, out);
}
int spb_argmax_rows_dev(spb_context *ctx, const float *S, int64_t m, int64_t n, int32_t *out) {
  return launch_argmax_rows(ctx, S, m, n, out);
}
int spb_eer(spb_context *ctx, const float *, const int32_t *, int64_t, double *, double *) {
  return fail(ctx, SPB_ERR_UNSUPPORTED, "eer: kernel not built yet");
}
int spb_eer_dev(spb_context *ctx, const float *s, const int32_t *l, int64_t n, double *eer, double *auc) {
  return run_eer(ctx, s, l, n, eer, auc);
}
}
