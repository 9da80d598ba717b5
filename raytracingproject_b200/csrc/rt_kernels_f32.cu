// fast mode: float arithmetic, FMA contraction on
#define RT_REAL float
#include "rt_kernels_impl.inl"
