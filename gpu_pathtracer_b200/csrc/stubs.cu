// stubs.cu -- entry points declared in include/gpt.h whose stage is not built yet.
// They fail loudly (GPT_E_UNSUPPORTED); nothing falls back to a CPU path.
#include "common.cuh"

#define UNSUPPORTED(ctx, name) gpt_fail((ctx), GPT_E_UNSUPPORTED, name " is not implemented yet", __FILE__, __LINE__)

extern "C" {
}
