// attention.cu — placeholder while the tcgen05 kernels land; every entry point fails loudly.
#include "common.cuh"

extern "C" int fx_attn_causal_fwd(void*, const void*, void*, float*, int, int, int, int, float, float,
                                  uint64_t, uint64_t) {
  fx::set_error("fx_attn_causal_fwd: not built yet");
  return FX_ERR_UNSUPPORTED;
}
extern "C" size_t fx_attn_causal_bwd_workspace_bytes(int, int, int, int) { return 0; }
extern "C" int fx_attn_causal_bwd(void*, const void*, const void*, const void*, const float*, void*, void*,
                                  int, int, int, int, float, float, uint64_t, uint64_t) {
  fx::set_error("fx_attn_causal_bwd: not built yet");
  return FX_ERR_UNSUPPORTED;
}
