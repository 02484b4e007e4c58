// placeholder until fast.cu lands
#include "context.h"
namespace wg {
int fast_render(wgpt_ctx* c, int, int, int) { return set_err(c, WGPT_E_STATE, "fast mode not built"); }
int fast_trace_rays(wgpt_ctx* c, int, const float*, const uint32_t*, int*, int*, float*) { return set_err(c, WGPT_E_STATE, "fast mode not built"); }
void fast_destroy(wgpt_ctx*) {}
}
