// parity mode: IEEE double in the reference operation order; built with --fmad=false
#define RT_REAL double
#include "rt_kernels_impl.inl"
