// lora_gemm_kernel / splitk_finalize_kernel for fp16 storage (kind::f16 MMAs, fp32 accumulate), both tile widths: the
// reference's inference scripts run the UNet + adapter in fp16 (eval-scripts/generate_images_xl.py:443).
#include "lora_gemm_kernel.cuh"

namespace slider {

int launch_gemm_f16(int bn, int r4, const CUtensorMap& a, const CUtensorMap& w, const CUtensorMap& l, const CUtensorMap& xe,
                    const CUtensorMap& we, const CUtensorMap& my, const CUtensorMap& mp, const GemmParams& p,
                    cudaStream_t st) {
  if (bn == 160) return launch_gemm_t<__half, 160, false>(r4, a, w, l, xe, we, my, mp, p, st);
  return launch_gemm_t<__half, 80, false>(r4, a, w, l, xe, we, my, mp, p, st);
}

}  // namespace slider
