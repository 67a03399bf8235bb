// fast flavour of the path kernels: FMA contraction and approximate reciprocals (see pt_math.cuh)
#define PT_PARITY 0
#define PT_NS pt_fast
#include "pt_kernels.cuh"
