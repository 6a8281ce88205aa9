// K1f flat-line star kernel on contiguous rows the TMA kernel cannot stage, STORE epilogue (see stencil_tin_impl.cuh, role 1).
#define CPPRE_TIN_MODE CPPRE_EPI_STORE
#define CPPRE_TIN_ROLE 1
#define CPPRE_TIN_SUFFIX store
#include "stencil_tin_impl.cuh"
