// K1t t-innermost star kernel for the FNO views the plain units decline (odd Nt, unaligned bases), STORE epilogue (see stencil_tin_impl.cuh).
#define CPPRE_TIN_MODE CPPRE_EPI_STORE
#define CPPRE_TIN_FLEX 1
#define CPPRE_TIN_SUFFIX store_x
#include "stencil_tin_impl.cuh"
