// K1 TMA star kernel, STORE epilogue, operands none (see stencil_tma_impl.cuh).
#define CPPRE_TMA_MODE CPPRE_EPI_STORE
#define CPPRE_TMA_OPER 0
#define CPPRE_TMA_SUFFIX store_plain
#include "stencil_tma_impl.cuh"
