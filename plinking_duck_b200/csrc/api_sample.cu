// Per-sample accumulations and genotype / dosage decode (plink_missing sample mode, read_pfile
// sample-orient counts, read_pgen / read_pfile decode, plink_score).
#include <algorithm>
#include <cstdio>
#include <cstring>

#include "ctx.h"

using namespace pdx;

extern "C" {

int pd_missing_by_sample(pd_ctx* c, uint32_t, uint32_t, uint32_t*) {
  return c ? c->SetError(PD_ERR_NOT_SUPPORTED, "pd_missing_by_sample: not built yet") : PD_ERR_INVALID_ARG;
}
int pd_sample_counts(pd_ctx* c, uint32_t, uint32_t, uint32_t*, uint32_t*, uint32_t*) {
  return c ? c->SetError(PD_ERR_NOT_SUPPORTED, "pd_sample_counts: not built yet") : PD_ERR_INVALID_ARG;
}
int pd_decode_i8(pd_ctx* c, uint32_t, uint32_t, int8_t*) {
  return c ? c->SetError(PD_ERR_NOT_SUPPORTED, "pd_decode_i8: not built yet") : PD_ERR_INVALID_ARG;
}
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

}  // extern "C"
