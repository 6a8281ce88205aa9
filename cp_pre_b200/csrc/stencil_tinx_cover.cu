// K1t t-innermost star kernel for the FNO views the plain units decline (odd Nt, unaligned bases), COVER epilogue (see stencil_tin_impl.cuh).
#define CPPRE_TIN_MODE CPPRE_EPI_COVER
#define CPPRE_TIN_FLEX 1
#define CPPRE_TIN_SUFFIX cover_x
#include "stencil_tin_impl.cuh"
