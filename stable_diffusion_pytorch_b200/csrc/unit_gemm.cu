// Translation unit 5: tcgen05 GEMM / implicit-GEMM convolution.
#include "gemm_umma.cu"
