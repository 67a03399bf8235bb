// parity flavour of the path kernels: compiled with --fmad=false (see Makefile, pt_math.cuh)
#define PT_PARITY 1
#define PT_NS pt_parity
#include "pt_kernels.cuh"
