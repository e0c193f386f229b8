// Unity translation unit of libsdb200.so (one TU keeps the __device__ debug flag and the launch
// counter shared without relocatable device code).
#include "runtime.cu"
#include "sampler.cu"
#include "norm.cu"
#include "conv_small.cu"
#include "precise_f32.cu"
#include "attention.cu"
#include "attention_ctx.cu"
#include "attention_bwd.cu"
#include "attention_umma.cu"
#include "gemm_umma.cu"
#include "backward.cu"
#include "wgrad_umma.cu"
