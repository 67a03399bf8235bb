// resolve.cu — the arithmetic that follows the path kernel in the reference, as standalone HBM-bound passes over the
// linear colors buffer (compiled with --fmad=false: every expression keeps prog.cl's operations and order, and pow is the
// correctly rounded float of the double function like in the parity path kernels, so these passes reproduce the
// reference's texture values bit for bit; at 16 B per pixel they are bound by HBM, not by the arithmetic):
//   k_resolve       the root's epilogue of a (multi-GPU) render in ONE pass: sum of up to RESOLVE_MAX_SRC accumulation
//                   buffers (local or peer memory over NVLink) -> * 1/total_spp -> linear colors, tone-mapped float
//                   image (prog.cl:380, bottom-up like the reference's texture) and 8-bit RGBA (rows flipped to top-down
//                   for image files; the reference image is bottom-up: prog.cl:86-88, main.cpp:1027-1034)
//   k_median        filt_im (prog.cl:391-427): 3x3 median by grey value + filmic tone map; shared-memory tile (each
//                   pixel is read 1.3x instead of 9x), 19-exchange median network on (grey, slot) pairs, and the
//                   reference's literal selection sort only where its tie handling could matter
#include "resolve.h"
#include <stdio.h>

namespace {

__device__ __forceinline__ float pow_cr(float a, float b) { return (float)pow((double)a, (double)b); }
__device__ __forceinline__ float srgb1(float a) { return a <= 0.00304f ? 12.92f * a : 1.055f * pow_cr(a, 0.4167f) - 0.055f; }   // prog.cl:247-258
__device__ __forceinline__ float filmic1(float c) {                                                                            // prog.cl:259-263
    c = fmaxf(0.0f, c - 0.004f);
    return (c * (c * 6.2f + 0.5f)) / (c * (c * 6.2f + 1.7f) + 0.06f);
}
// reinhard_tone_map (prog.cl:264-269): L == 0 gives NaN there (0 * 0 / 0); nan_black selects the defined value (black)
__device__ __forceinline__ float4 reinhard(float4 c, bool nan_black) {
    const float L = (0.2126f * c.x + 0.7152f * c.y) + 0.0722f * c.z;
    if (nan_black && !(L > 0.0f)) return make_float4(0.0f, 0.0f, 0.0f, 1.0f);
    const float L2 = L / (1.0f + L);
    return make_float4(srgb1((c.x * L2) / L), srgb1((c.y * L2) / L), srgb1((c.z * L2) / L), 1.0f);
}
__device__ __forceinline__ float4 filmic(float4 c) { return make_float4(filmic1(c.x), filmic1(c.y), filmic1(c.z), 1.0f); }
__device__ __forceinline__ unsigned char to8(float v) {
    if (!(v > 0.0f)) return 0;          // NaN (the reference's L == 0 case) and negatives -> black
    if (v >= 1.0f) return 255;
    return (unsigned char)(v * 255.0f + 0.5f);
}

// One thread per pixel, consecutive threads on consecutive pixels: coalesced 16-byte loads from every source.
// scale == 1 and one source: a pure tone map.  mean_out / tex_out / rgba_out may each be null.
__global__ void __launch_bounds__(256) k_resolve(const ResolveSrc src, float scale, int w, int h, int mode,
                                                 float4 *mean_out, float4 *tex_out, uchar4 *rgba_out) {
    const size_t n = (size_t)w * h;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        float4 v = src.p[0][i];
        for (int k = 1; k < src.n; ++k) {                 // rank order: the sum's rounding does not depend on timing
            const float4 o = src.p[k][i];
            v.x += o.x; v.y += o.y; v.z += o.z;
        }
        if (scale != 1.0f) { v.x *= scale; v.y *= scale; v.z *= scale; }
        if (mean_out) mean_out[i] = make_float4(v.x, v.y, v.z, 0.0f);
        if (tex_out) tex_out[i] = mode == 1 ? filmic(v) : reinhard(v, false);
        if (rgba_out) {
            const float4 t = mode == 1 ? filmic(v) : reinhard(v, true);
            const size_t y = i / (size_t)w, x = i - y * (size_t)w;
            rgba_out[(size_t)(h - 1 - (int)y) * w + x] = make_uchar4(to8(t.x), to8(t.y), to8(t.z), 255);
        }
    }
}

// ---------------------------------------------------------------- filt_im
// The reference guards x>0 && y>0 && x<width && y<height and reads its neighbourhood through LINEAR ids
// (y-1+i)*width + (x-1+j): on the last column that wraps into the next scanline, on the last scanline (and for the last
// pixel of the one before) it runs past the end of colors[].  Here: the same linear addressing, every pixel whose nine
// reads stay inside the buffer is computed (wrap included), the ones that would read out of range are left untouched.
#define MED_BX 32
#define MED_BY 8
#define MED_CE(a, b) do { const bool sw_ = g[a] > g[b]; const float tg_ = sw_ ? g[b] : g[a]; const int ti_ = sw_ ? s[b] : s[a]; \
                          g[b] = sw_ ? g[a] : g[b]; s[b] = sw_ ? s[a] : s[b]; g[a] = tg_; s[a] = ti_; } while (0)
__global__ void __launch_bounds__(MED_BX * MED_BY) k_median(const float4 *colors, float4 *image, int w, int h) {
    __shared__ float4 tile[MED_BY + 2][MED_BX + 2];
    const int x0 = blockIdx.x * MED_BX, y0 = blockIdx.y * MED_BY;
    const long long n = (long long)w * h;
    for (int t = threadIdx.y * MED_BX + threadIdx.x; t < (MED_BY + 2) * (MED_BX + 2); t += MED_BX * MED_BY) {
        const int r = t / (MED_BX + 2), c = t - r * (MED_BX + 2);
        const long long id = (long long)(y0 - 1 + r) * w + (x0 - 1 + c);
        float4 v = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (id >= 0 && id < n) { v = colors[id]; v.w = ((v.x + v.y) + v.z) / 3.0f; }       // grey once per pixel (prog.cl:405-408)
        tile[r][c] = v;
    }
    __syncthreads();
    const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
    if (x <= 0 || y <= 0 || x >= w || y >= h || (long long)(y + 1) * w + (x + 1) >= n) return;
    float g[9];
    int s[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) { g[k] = tile[threadIdx.y + k / 3][threadIdx.x + k % 3].w; s[k] = k; }
    // median-of-9 exchange network (19 compare-exchanges), slot index carried along
    MED_CE(1, 2); MED_CE(4, 5); MED_CE(7, 8); MED_CE(0, 1); MED_CE(3, 4); MED_CE(6, 7); MED_CE(1, 2); MED_CE(4, 5); MED_CE(7, 8);
    MED_CE(0, 3); MED_CE(5, 8); MED_CE(4, 7); MED_CE(3, 6); MED_CE(1, 4); MED_CE(2, 5); MED_CE(4, 7); MED_CE(4, 2); MED_CE(6, 4); MED_CE(4, 2);
    const float gm = g[4];
    float4 m = tile[threadIdx.y + s[4] / 3][threadIdx.x + s[4] % 3];
    // The reference's selection sort is not stable; which of several pixels with the SAME grey value ends up in the middle
    // depends on its swap order.  That only matters when such a pixel differs in colour from the one picked here (or a
    // NaN makes the comparisons non-transitive): then run the reference's own sort (first maximum wins, prog.cl:410-424).
    bool exact = false;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        const float4 v = tile[threadIdx.y + k / 3][threadIdx.x + k % 3];
        exact |= !(v.w == v.w);
        exact |= (v.w == gm) & ((__float_as_uint(v.x) != __float_as_uint(m.x)) | (__float_as_uint(v.y) != __float_as_uint(m.y)) |
                                (__float_as_uint(v.z) != __float_as_uint(m.z)));
    }
    if (exact) {
        float gr[9];
        int sl[9];
        for (int k = 0; k < 9; ++k) { gr[k] = tile[threadIdx.y + k / 3][threadIdx.x + k % 3].w; sl[k] = k; }
        for (int j = 9; j > 1; --j) {
            int maxi = 0;
            for (int i = 1; i < j; ++i)
                if (gr[i] > gr[maxi]) maxi = i;
            const float tg = gr[j - 1]; gr[j - 1] = gr[maxi]; gr[maxi] = tg;
            const int ts = sl[j - 1]; sl[j - 1] = sl[maxi]; sl[maxi] = ts;
        }
        m = tile[threadIdx.y + sl[4] / 3][threadIdx.x + sl[4] % 3];
    }
    image[(size_t)y * w + x] = filmic(m);
}

}  // namespace

void resolve_frame(const ResolveSrc &src, float scale, int w, int h, int mode, float4 *mean_out, float4 *tex_out, unsigned char *rgba_out,
                   cudaStream_t st) {
    const size_t n = (size_t)w * h;
    if (!n || src.n < 1) return;
    size_t blocks = (n + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;         // grid-stride over a multiple of the SM count
    k_resolve<<<(unsigned)blocks, 256, 0, st>>>(src, scale, w, h, mode, mean_out, tex_out, (uchar4 *)rgba_out);
}
// ---------------------------------------------------------------- the resolve without any allocation (ptb_resolve_rgba8_noalloc)
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it): the 8-bit
// picture is staged in a file-scope device array instead of a per-call cudaMalloc'ed buffer.  Every GPU that loads the library holds
// the array (RESOLVE_NOALLOC_MAX_PIXELS * 4 bytes = 8 MiB) until exit.  ALL calls share it and it keeps the last call's contents: the
// kernel writes every one of the w * h pixels it is asked for, so nothing has to start at zero; calls that overlap on the GPU race.
// In host code `g_noalloc_rgba` is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress, once per device.
static __device__ __align__(256) unsigned char g_noalloc_rgba[(size_t)RESOLVE_NOALLOC_MAX_PIXELS * 4];

int resolve_rgba8_noalloc(const float4 *colors, int w, int h, int mode, unsigned char *dst_host, cudaStream_t st, char *err, size_t errcap) {
    static unsigned char *dev_ptr[64] = {};          // per device ordinal
    const size_t n = (size_t)w * h;
    if (n == 0 || n > (size_t)RESOLVE_NOALLOC_MAX_PIXELS) {
        snprintf(err, errcap, "resolve without allocation: %d x %d = %zu pixels, the static scratch holds at most %d", w, h, n, RESOLVE_NOALLOC_MAX_PIXELS);
        return -1;
    }
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) { snprintf(err, errcap, "resolve without allocation: cudaGetDevice: %s", cudaGetErrorString(e)); return -1; }
    if (!dev_ptr[dev]) {                              // set-up, once per device, with that device current
        void *ptr = nullptr;
        e = cudaGetSymbolAddress(&ptr, g_noalloc_rgba);
        if (e != cudaSuccess || !ptr) { snprintf(err, errcap, "resolve without allocation: cudaGetSymbolAddress: %s", cudaGetErrorString(e)); return -1; }
        dev_ptr[dev] = (unsigned char *)ptr;
    }
    ResolveSrc src{};
    src.p[0] = colors; src.n = 1;
    size_t blocks = (n + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    k_resolve<<<(unsigned)blocks, 256, 0, st>>>(src, 1.0f, w, h, mode, nullptr, nullptr, (uchar4 *)dev_ptr[dev]);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(dst_host, dev_ptr[dev], n * 4, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { snprintf(err, errcap, "resolve without allocation: %s", cudaGetErrorString(e)); return -1; }
    return 0;
}

void filter_median(const float4 *colors, float4 *image, int w, int h, cudaStream_t st) {
    k_median<<<dim3((w + MED_BX - 1) / MED_BX, (h + MED_BY - 1) / MED_BY), dim3(MED_BX, MED_BY), 0, st>>>(colors, image, w, h);
}
