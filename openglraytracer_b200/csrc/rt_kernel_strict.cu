// Strict-arithmetic twin: compiled with -fmad=false; mirrors the shader op by op.
#define RT_STRICT 1
#include "rt_kernel.cuh"
#include "rt_kernel_impl.inc"
