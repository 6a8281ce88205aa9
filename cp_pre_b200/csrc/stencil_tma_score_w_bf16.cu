// K1 TMA star kernel, SCORE epilogue, operand set 2, bf16-input mode (see stencil_tma_impl.cuh).
#define CPPRE_TMA_MODE CPPRE_EPI_SCORE
#define CPPRE_TMA_OPER 2
#define CPPRE_TMA_BF16 1
#define CPPRE_TMA_SUFFIX score_w_bf16
#include "stencil_tma_impl.cuh"
