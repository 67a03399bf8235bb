// oclshim.h — TEST INFRASTRUCTURE (oracle).  Not part of the product path.
//
// Lets the reference's OpenCL C source (/root/reference/prog.cl) compile as C++17 where it
// lies, so that the reference's own kernel code is the CPU oracle.  This header
// supplies only what OpenCL C has built in and C++ has not: the vector types, their operators,
// the geometric/math built-ins prog.cl calls, work-item ids and write_imagef.
//
// Canonical choices where OpenCL only bounds the result in ulps (stated):
//   half_sqrt(x) = sqrtf(x) (correctly rounded)      normalize(v) = v / sqrtf(dot(v,v))
//   dot(a,b) = (a.x*b.x + a.y*b.y) + a.z*b.z
//   pow/cos/sin = the CORRECTLY ROUNDED float result, obtained as (float)f((double)x) with glibc's
//                 double-precision functions (< 1 ulp in double, so the narrowed value is the correctly
//                 rounded float except for arguments within ~2^-29 relative of a rounding boundary).
//                 This is the one choice every implementation can reproduce: the CUDA parity build
//                 evaluates the same expression with CUDA's double functions.
//   fmin/fmax = fminf/fmaxf (NaN → the other operand, as OpenCL specifies)
// and the translation unit is compiled with -ffp-contract=off (no FMA).
#ifndef ORACLE_OCLSHIM_H_
#define ORACLE_OCLSHIM_H_

#include <cmath>
#include <cstddef>
#include <cstdint>

namespace oclshim {

typedef unsigned short ushort;
typedef unsigned long ulong;   // 64-bit on LP64, like OpenCL's ulong

struct alignas(16) float3 {
    float x, y, z, w_;
};
struct alignas(16) float4 {
    float x, y, z, w;
};
struct alignas(8) int2 {
    int x, y;
};

inline float3 float3_(float x, float y, float z) { float3 r; r.x = x; r.y = y; r.z = z; r.w_ = 0.0f; return r; }
inline float4 float4_(float x, float y, float z, float w) { float4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
inline float4 float4_(float3 v, float w) { return float4_(v.x, v.y, v.z, w); }
inline int2 int2_(int x, int y) { int2 r; r.x = x; r.y = y; return r; }

inline float3 operator+(float3 a, float3 b) { return float3_(a.x + b.x, a.y + b.y, a.z + b.z); }
inline float3 operator-(float3 a, float3 b) { return float3_(a.x - b.x, a.y - b.y, a.z - b.z); }
inline float3 operator*(float3 a, float3 b) { return float3_(a.x * b.x, a.y * b.y, a.z * b.z); }
inline float3 operator/(float3 a, float3 b) { return float3_(a.x / b.x, a.y / b.y, a.z / b.z); }
inline float3 operator+(float3 a, float s) { return float3_(a.x + s, a.y + s, a.z + s); }
inline float3 operator-(float3 a, float s) { return float3_(a.x - s, a.y - s, a.z - s); }
inline float3 operator*(float3 a, float s) { return float3_(a.x * s, a.y * s, a.z * s); }
inline float3 operator/(float3 a, float s) { return float3_(a.x / s, a.y / s, a.z / s); }
inline float3 operator+(float s, float3 a) { return float3_(s + a.x, s + a.y, s + a.z); }
inline float3 operator-(float s, float3 a) { return float3_(s - a.x, s - a.y, s - a.z); }
inline float3 operator*(float s, float3 a) { return float3_(s * a.x, s * a.y, s * a.z); }
inline float3 operator/(float s, float3 a) { return float3_(s / a.x, s / a.y, s / a.z); }
inline float3 operator-(float3 a) { return float3_(-a.x, -a.y, -a.z); }

inline float dot(float3 a, float3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline float3 cross(float3 a, float3 b) {
    return float3_(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
inline float3 normalize(float3 v) { return v / sqrtf(dot(v, v)); }
inline float half_sqrt(float x) { return sqrtf(x); }
inline float fabs(float x) { return fabsf(x); }
inline float fmin(float a, float b) { return fminf(a, b); }
inline float fmax(float a, float b) { return fmaxf(a, b); }
inline float pow(float a, float b) { return (float)::pow((double)a, (double)b); }
inline float pow(float a, int b) { return (float)::pow((double)a, (double)b); }
inline float cos(float a) { return (float)::cos((double)a); }
inline float sin(float a) { return (float)::sin((double)a); }
inline float max(float a, float b) { return a < b ? b : a; }
inline float3 max(float s, float3 v) { return float3_(max(s, v.x), max(s, v.y), max(s, v.z)); }
inline float3 max(float3 v, float s) { return float3_(max(v.x, s), max(v.y, s), max(v.z, s)); }

// work-item ids: set by the executor before each call
struct WorkItem {
    size_t gid[3];
    size_t gsz[3];
};
inline WorkItem &work_item() {
    static thread_local WorkItem wi;
    return wi;
}
inline size_t get_global_id(int d) { return work_item().gid[d]; }
inline size_t get_global_size(int d) { return work_item().gsz[d]; }

// image2d_t: a plain W×H float4 array (the GL texture of main.cpp:512-519 is headless here)
struct image2d {
    float4 *texels;
    int width, height;
};
typedef image2d *image2d_t;
inline void write_imagef(image2d_t img, int2 xy, float4 v) {
    if (img && img->texels && xy.x >= 0 && xy.y >= 0 && xy.x < img->width && xy.y < img->height)
        img->texels[(size_t)xy.y * img->width + xy.x] = v;
}

// work counters: rays are always counted, the rest only in the counting namespace
struct Counters {
    uint64_t rays, box, tri, tri1, tri2, tri3;
};
inline Counters &counters() {
    static thread_local Counters c = {0, 0, 0, 0, 0, 0};
    return c;
}

}  // namespace oclshim

#define global
#define kernel
#define write_only
#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

#endif  // ORACLE_OCLSHIM_H_
