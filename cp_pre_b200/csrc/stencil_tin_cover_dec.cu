// K1t t-innermost star kernel, COVER epilogue, with the normaliser-decode fold (see stencil_tin_impl.cuh).
#define CPPRE_TIN_MODE CPPRE_EPI_COVER
#define CPPRE_TIN_DECODE 1
#define CPPRE_TIN_SUFFIX cover_dec
#include "stencil_tin_impl.cuh"
