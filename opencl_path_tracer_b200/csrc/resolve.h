// resolve.h — output-side kernels next to the hot path (SURVEY §8f rows 1-2, §8e root epilogue): fused sum -> mean -> tone map -> 8 bit, median.
#ifndef PTB_RESOLVE_H_
#define PTB_RESOLVE_H_
#include <cuda_runtime.h>
#include <stddef.h>
#define RESOLVE_MAX_SRC 8
struct ResolveSrc {
    const float4 *p[RESOLVE_MAX_SRC];   // accumulation buffers to add up (device pointers; peers' memory is fine with peer access on)
    int n;
};
// mean = (sum over sources) * scale; any of the three outputs may be null.  mode 0 = reinhard + sRGB, 1 = filmic.
// tex_out keeps the reference's NaN for L == 0 (prog.cl:264-269); rgba_out maps it to black and is written top row first.
void resolve_frame(const ResolveSrc &src, float scale, int w, int h, int mode, float4 *mean_out, float4 *tex_out, unsigned char *rgba_out,
                   cudaStream_t st);
// the same tone map -> RGBA8 -> host copy with the staging buffer in a static __device__ array (no cudaMalloc / cudaFree): resolve.cu
#define RESOLVE_NOALLOC_MAX_PIXELS (2 * 1024 * 1024)      /* 8 MiB of scratch per GPU; a 1920 x 1080 frame fits */
int resolve_rgba8_noalloc(const float4 *colors, int w, int h, int mode, unsigned char *dst_host, cudaStream_t st, char *err, size_t errcap);
void filter_median(const float4 *colors, float4 *image, int w, int h, cudaStream_t st);
#endif
