// K1t t-innermost star kernel for the FNO views the plain units decline (odd Nt, unaligned bases), SCORE epilogue (see stencil_tin_impl.cuh).
#define CPPRE_TIN_MODE CPPRE_EPI_SCORE
#define CPPRE_TIN_FLEX 1
#define CPPRE_TIN_SUFFIX score_x
#include "stencil_tin_impl.cuh"
