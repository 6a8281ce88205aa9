// K1 TMA star kernel, COVER epilogue, operand set -1, bf16-input mode (see stencil_tma_impl.cuh).
#define CPPRE_TMA_MODE CPPRE_EPI_COVER
#define CPPRE_TMA_OPER -1
#define CPPRE_TMA_BF16 1
#define CPPRE_TMA_SUFFIX cover_dyn_bf16
#include "stencil_tma_impl.cuh"
