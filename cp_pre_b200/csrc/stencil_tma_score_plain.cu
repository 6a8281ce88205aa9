// K1 TMA star kernel, SCORE epilogue, operands none (see stencil_tma_impl.cuh).
#define CPPRE_TMA_MODE CPPRE_EPI_SCORE
#define CPPRE_TMA_OPER 0
#define CPPRE_TMA_SUFFIX score_plain
#include "stencil_tma_impl.cuh"
