// K1t t-innermost star kernel, SCORE epilogue, with the normaliser-decode fold (see stencil_tin_impl.cuh).
#define CPPRE_TIN_MODE CPPRE_EPI_SCORE
#define CPPRE_TIN_DECODE 1
#define CPPRE_TIN_SUFFIX score_dec
#include "stencil_tin_impl.cuh"
