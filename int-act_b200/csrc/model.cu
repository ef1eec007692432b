// placeholder until the model orchestration lands (next commit)
#include "../../include/pi0b200.h"
#include "common.cuh"
using namespace pi0;
extern "C" {
size_t pi0_workspace_bytes(const pi0_model_cfg*) { return 0; }
int pi0_ctx_create(const pi0_model_cfg*, const pi0_weights*, void*, size_t, pi0_ctx**) { set_error("not implemented"); return -3; }
void pi0_ctx_destroy(pi0_ctx*) {}
int pi0_siglip_forward(pi0_ctx*, int, int, const float*, float, float, void*) { set_error("not implemented"); return -3; }
int pi0_embed_language(pi0_ctx*, int, const int64_t*, void*) { set_error("not implemented"); return -3; }
int pi0_prefill(pi0_ctx*, int, const uint8_t*, const uint8_t*, void*) { set_error("not implemented"); return -3; }
int pi0_denoise_loop(pi0_ctx*, int, const float*, float*, void*) { set_error("not implemented"); return -3; }
int pi0_infer(pi0_ctx*, int, const float* const*, float, float, const uint8_t*, const int64_t*, const uint8_t*, const float*, float*, void*) { set_error("not implemented"); return -3; }
const void* pi0_debug_buffer(pi0_ctx*, const char*, int64_t*) { return nullptr; }
}
