// K1 TMA star kernel, COVER epilogue instantiations (see stencil_tma_impl.cuh).
#define CPPRE_TMA_MODE CPPRE_EPI_COVER
#define CPPRE_TMA_SUFFIX cover
#include "stencil_tma_impl.cuh"
