#include "../../include/jumbo_mae_b200.h"
#include "common.cuh"
extern "C" int jm_attention_fwd(const jm_bf16*, long, int, int, int, int, jm_bf16*, long, float*, void*) {
  return jm::fail("jm_attention_fwd: not built yet");
}
extern "C" int jm_attention_bwd(const jm_bf16*, long, const jm_bf16*, long, const jm_bf16*, long, const float*, int,
                                int, int, int, float, jm_bf16*, long, void*) {
  return jm::fail("jm_attention_bwd: not built yet");
}
