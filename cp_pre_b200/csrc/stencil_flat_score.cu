// K1f flat-line star kernel on contiguous rows the TMA kernel cannot stage, SCORE epilogue (see stencil_tin_impl.cuh, role 1).
#define CPPRE_TIN_MODE CPPRE_EPI_SCORE
#define CPPRE_TIN_ROLE 1
#define CPPRE_TIN_SUFFIX score
#include "stencil_tin_impl.cuh"
