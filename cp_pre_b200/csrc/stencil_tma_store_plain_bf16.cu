// K1 TMA star kernel, STORE epilogue, operand set 0, bf16-input mode (see stencil_tma_impl.cuh).
#define CPPRE_TMA_MODE CPPRE_EPI_STORE
#define CPPRE_TMA_OPER 0
#define CPPRE_TMA_BF16 1
#define CPPRE_TMA_SUFFIX store_plain_bf16
#include "stencil_tma_impl.cuh"
