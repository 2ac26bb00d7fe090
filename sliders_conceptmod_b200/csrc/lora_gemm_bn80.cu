// lora_gemm_kernel / splitk_finalize_kernel for the 128 x 80 tile (all ranks, bf16 and fp32 storage).
#include "lora_gemm_kernel.cuh"

namespace slider {

template int launch_gemm_tile<80>(int, int, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                                  const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                                  const GemmParams&, cudaStream_t);

}  // namespace slider
