// stubs.cu -- entry points declared in include/gpt.h whose stage is not built yet.
// They fail loudly (GPT_E_UNSUPPORTED); nothing falls back to a CPU path.
#include "common.cuh"


#define UNSUPPORTED(ctx, name) gpt_fail((ctx), GPT_E_UNSUPPORTED, name " is not implemented yet", __FILE__, __LINE__)

extern "C" {
void gpt_sppm_free(gpt_ctx *) {}
int gpt_sppm_init(gpt_ctx *ctx) { return UNSUPPORTED(ctx, "gpt_sppm_init"); }
int gpt_sppm_pass(gpt_ctx *ctx, int, int64_t, int64_t) { return UNSUPPORTED(ctx, "gpt_sppm_pass"); }
int gpt_sppm_flux_device(gpt_ctx *ctx, void **, int64_t *) { return UNSUPPORTED(ctx, "gpt_sppm_flux_device"); }
int gpt_sppm_update(gpt_ctx *ctx, int) { return UNSUPPORTED(ctx, "gpt_sppm_update"); }
int gpt_sppm_resolve(gpt_ctx *ctx, int, int, float *) { return UNSUPPORTED(ctx, "gpt_sppm_resolve"); }
}
