// K1t t-innermost star kernel, SCORE epilogue (see stencil_tin_impl.cuh).
#define CPPRE_TIN_MODE CPPRE_EPI_SCORE
#define CPPRE_TIN_SUFFIX score
#include "stencil_tin_impl.cuh"
