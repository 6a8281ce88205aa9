// K1t t-innermost star kernel, COVER epilogue (see stencil_tin_impl.cuh).
#define CPPRE_TIN_MODE CPPRE_EPI_COVER
#define CPPRE_TIN_SUFFIX cover
#include "stencil_tin_impl.cuh"
