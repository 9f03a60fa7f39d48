// Shipped kernel: default FMA contraction (-fmad=true), no --use_fast_math.
#define RT_STRICT 0
#include "rt_kernel.cuh"
#include "rt_kernel_impl.inc"
