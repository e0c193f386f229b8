// Translation unit 4: tcgen05 attention backward.
#include "attention_bwd_umma.cu"
