// K1t t-innermost star kernel, STORE epilogue (see stencil_tin_impl.cuh).
#define CPPRE_TIN_MODE CPPRE_EPI_STORE
#define CPPRE_TIN_SUFFIX store
#include "stencil_tin_impl.cuh"
