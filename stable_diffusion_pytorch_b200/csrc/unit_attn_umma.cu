// Translation unit 3: tcgen05 attention forward (both CTA shapes).
#include "attention_umma.cu"
