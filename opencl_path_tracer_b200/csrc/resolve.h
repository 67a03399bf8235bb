// resolve.h — output-side kernels next to the hot path: tone map, median, sum scale.
#ifndef PTB_RESOLVE_H_
#define PTB_RESOLVE_H_
#include <cuda_runtime.h>
#include <stddef.h>
void resolve_scale(float4 *colors, size_t n, float scale, cudaStream_t st);
void resolve_rgba8(const float4 *colors, int w, int h, int mode, unsigned char *dst, cudaStream_t st);
void filter_median(const float4 *colors, float4 *image, int w, int h, cudaStream_t st);
#endif
