// resolve.cu — the arithmetic that follows the path kernel in the reference, as standalone
// HBM-bound passes over the linear colors buffer (16 B per pixel in, 4 or 16 B out):
//   resolve_rgba8   reinhard_tone_map + sRGB (prog.cl:247-258, 264-269) or filmic_tone_map
//                   (prog.cl:259-263) → 8-bit RGBA, rows flipped to top-down for image files
//                   (the reference image is bottom-up: prog.cl:86-88, main.cpp:1027-1034)
//   filter_median   filt_im (prog.cl:391-427): 3x3 median by grey value + filmic tone map
//   resolve_scale   sum → mean for the spp-split accumulation
// One thread per pixel, consecutive threads on consecutive pixels: fully coalesced 16-byte loads.
#include "resolve.h"

namespace {

__device__ __forceinline__ float srgb1(float a) { return a <= 0.00304f ? 12.92f * a : 1.055f * powf(a, 0.4167f) - 0.055f; }
__device__ __forceinline__ float filmic1(float c) {
    c = fmaxf(0.0f, c - 0.004f);
    return (c * (c * 6.2f + 0.5f)) / (c * (c * 6.2f + 1.7f) + 0.06f);
}
__device__ __forceinline__ unsigned char to8(float v) {
    if (!(v > 0.0f)) return 0;          // NaN (the reference's L == 0 case) and negatives → black
    if (v >= 1.0f) return 255;
    return (unsigned char)(v * 255.0f + 0.5f);
}

__global__ void k_scale(float4 *c, size_t n, float s) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 v = c[i];
    c[i] = make_float4(v.x * s, v.y * s, v.z * s, v.w);
}

__global__ void k_rgba8(const float4 *colors, int w, int h, int mode, uchar4 *dst) {
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w || y >= h) return;
    float4 c = colors[(size_t)y * w + x];
    float r, g, b;
    if (mode == 1) {
        r = filmic1(c.x); g = filmic1(c.y); b = filmic1(c.z);
    } else {
        float L = 0.2126f * c.x + 0.7152f * c.y + 0.0722f * c.z;
        float k = L > 0.0f ? (L / (1.0f + L)) / L : 0.0f;
        r = srgb1(c.x * k); g = srgb1(c.y * k); b = srgb1(c.z * k);
    }
    dst[(size_t)(h - 1 - y) * w + x] = make_uchar4(to8(r), to8(g), to8(b), 255);
}

// filt_im: interior pixels only (the reference guards x>0 && y>0 && x<width && y<height and then
// reads x+1 / y+1, i.e. one past the edge on the last column/row — here those pixels are skipped
// like the first column/row, which is the in-bounds subset of what the reference computes).
__global__ void k_median(const float4 *colors, float4 *image, int w, int h) {
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x <= 0 || y <= 0 || x >= w - 1 || y >= h - 1) return;
    float4 arr[9];
    float grey[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            float4 v = colors[(size_t)(y - 1 + i) * w + (x - 1 + j)];
            arr[i * 3 + j] = v;
            grey[i * 3 + j] = (v.x + v.y + v.z) / 3.0f;
        }
    // selection sort by grey, exactly the reference's order of swaps (first maximum wins)
    for (int j = 9; j > 1; --j) {
        int maxi = 0;
        for (int i = 1; i < j; ++i)
            if (grey[i] > grey[maxi]) maxi = i;
        float tg = grey[j - 1]; grey[j - 1] = grey[maxi]; grey[maxi] = tg;
        float4 ta = arr[j - 1]; arr[j - 1] = arr[maxi]; arr[maxi] = ta;
    }
    float4 m = arr[4];
    image[(size_t)y * w + x] = make_float4(filmic1(m.x), filmic1(m.y), filmic1(m.z), 1.0f);
}

}  // namespace

void resolve_scale(float4 *colors, size_t n, float scale, cudaStream_t st) {
    if (n) k_scale<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(colors, n, scale);
}
void resolve_rgba8(const float4 *colors, int w, int h, int mode, unsigned char *dst, cudaStream_t st) {
    k_rgba8<<<dim3((w + 255) / 256, h), 256, 0, st>>>(colors, w, h, mode, (uchar4 *)dst);
}
void filter_median(const float4 *colors, float4 *image, int w, int h, cudaStream_t st) {
    k_median<<<dim3((w + 127) / 128, h), 128, 0, st>>>(colors, image, w, h);
}
