// K1t t-innermost star kernel, STORE epilogue, with the normaliser-decode fold (see stencil_tin_impl.cuh).
#define CPPRE_TIN_MODE CPPRE_EPI_STORE
#define CPPRE_TIN_DECODE 1
#define CPPRE_TIN_SUFFIX store_dec
#include "stencil_tin_impl.cuh"
