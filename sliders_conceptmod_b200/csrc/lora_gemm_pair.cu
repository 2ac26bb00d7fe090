// lora_gemm_kernel in CTA-pair form (cluster of 2, cta_group::2 MMAs): bf16 storage, 128 x 160 tiles, all ranks.
#include "lora_gemm_kernel.cuh"

namespace slider {

int launch_gemm_pair(int r4, const CUtensorMap& a, const CUtensorMap& w, const CUtensorMap& l, const CUtensorMap& xe,
                     const CUtensorMap& we, const CUtensorMap& my, const CUtensorMap& mp, const GemmParams& p,
                     cudaStream_t st) {
  return launch_gemm_t<__nv_bfloat16, 160, true>(r4, a, w, l, xe, we, my, mp, p, st);
}

}  // namespace slider
