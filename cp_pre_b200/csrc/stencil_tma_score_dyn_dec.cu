// K1 TMA star kernel, SCORE epilogue, operand set -1, with the normaliser-decode fold (see stencil_tma_impl.cuh).
#define CPPRE_TMA_MODE CPPRE_EPI_SCORE
#define CPPRE_TMA_OPER -1
#define CPPRE_TMA_DECODE 1
#define CPPRE_TMA_SUFFIX score_dyn_dec
#include "stencil_tma_impl.cuh"
