// Dosage decode, dosage-aware counts and plink_score (filled in below).
#include "ctx.h"
using namespace pdx;
extern "C" {
int pd_dosage_f64(pd_ctx* c, const uint32_t*, uint32_t, double*) {
  return c ? c->SetError(PD_ERR_NOT_SUPPORTED, "pd_dosage_f64: not built yet") : PD_ERR_INVALID_ARG;
}
int pd_dosage_counts(pd_ctx* c, uint32_t, uint32_t, uint32_t*, uint64_t*, double*) {
  return c ? c->SetError(PD_ERR_NOT_SUPPORTED, "pd_dosage_counts: not built yet") : PD_ERR_INVALID_ARG;
}
int pd_score(pd_ctx* c, const uint32_t*, const uint8_t*, const double*, uint32_t, uint32_t, uint32_t, double*, double*,
             uint32_t*) {
  return c ? c->SetError(PD_ERR_NOT_SUPPORTED, "pd_score: not built yet") : PD_ERR_INVALID_ARG;
}
}
