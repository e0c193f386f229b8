// Translation unit 1 of libsdb200.so: runtime plumbing and the CUDA-core / HBM-bound kernels.
// The library is built from the unit_*.cu files, compiled in parallel (build.py) and linked without
// relocatable device code; kernels that share device helpers stay in one unit.
#include "runtime.cu"
#include "sampler.cu"
#include "norm.cu"
#include "conv_small.cu"
#include "precise_f32.cu"
#include "clip_text.cu"
