// K1 TMA star kernel, COVER epilogue, operands weight only (see stencil_tma_impl.cuh).
#define CPPRE_TMA_MODE CPPRE_EPI_COVER
#define CPPRE_TMA_OPER 2
#define CPPRE_TMA_SUFFIX cover_w
#include "stencil_tma_impl.cuh"
