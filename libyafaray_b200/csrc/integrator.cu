// placeholder, replaced by the wavefront integrator
#include "context.h"
namespace yg {
int renderPass(yafgpu_ctx *ctx, const yafgpu_pass_desc *) { return setErr(ctx, YAFGPU_ERR_UNSUPPORTED, "integrator not built yet"); }
void freeWavefront(yafgpu_ctx *) {}
}
