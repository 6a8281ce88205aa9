// K1 TMA star kernel, STORE epilogue instantiations (see stencil_tma_impl.cuh).
#define CPPRE_TMA_MODE CPPRE_EPI_STORE
#define CPPRE_TMA_SUFFIX store
#include "stencil_tma_impl.cuh"
