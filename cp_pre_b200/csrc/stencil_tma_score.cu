// K1 TMA star kernel, SCORE epilogue instantiations (see stencil_tma_impl.cuh).
#define CPPRE_TMA_MODE CPPRE_EPI_SCORE
#define CPPRE_TMA_SUFFIX score
#include "stencil_tma_impl.cuh"
