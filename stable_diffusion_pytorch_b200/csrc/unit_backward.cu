// Translation unit 6: HBM-bound backward kernels, optimiser, tcgen05 weight gradients.
#include "backward_vec.cu"
#include "backward.cu"
#include "wgrad_umma.cu"
