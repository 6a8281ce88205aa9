// K1 TMA star kernel, STORE epilogue, operands decided at run time (see stencil_tma_impl.cuh).
#define CPPRE_TMA_MODE CPPRE_EPI_STORE
#define CPPRE_TMA_OPER -1
#define CPPRE_TMA_SUFFIX store_dyn
#include "stencil_tma_impl.cuh"
