// pt_math.cuh — float3 arithmetic, the reference's per-pixel RNG and the OpenCL built-ins the
// path kernels need, in two arithmetic flavours selected by PT_PARITY at compile time.
//
//   PT_PARITY=1  (pt_kernels_parity.cu, nvcc --fmad=false): every expression keeps the operation
//                order and the IEEE-754 single operations of prog.cl as the oracle compiles it
//                (no FMA contraction, correctly rounded / and sqrt, pow/cos/sin as the correctly
//                rounded float of the double-precision function).  Results are meant to be
//                bit-identical to oracle/ (see oracle/oclshim.h for the same canonical choices).
//   PT_PARITY=0  (pt_kernels_fast.cu, --fmad=true): FMA contraction, rsqrt-based normalize,
//                sincospi, exp2/log2 pow.  Checked against the parity build within the
//                tolerance written in tests/.
#ifndef PT_MATH_CUH_
#define PT_MATH_CUH_

#include <cuda_runtime.h>
#include <stdint.h>

#ifndef PT_PARITY
#error "compile with -DPT_PARITY=0 or 1"
#endif

#define PT_DEV __device__ __forceinline__

struct V3 {
    float x, y, z;
};
PT_DEV V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
PT_DEV V3 v3(float4 q) { return v3(q.x, q.y, q.z); }
PT_DEV V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
PT_DEV V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
PT_DEV V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
PT_DEV V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
PT_DEV V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
PT_DEV V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
PT_DEV V3 operator-(float s, V3 a) { return v3(s - a.x, s - a.y, s - a.z); }
PT_DEV V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
// OpenCL dot(): the oracle's association (oclshim.h) — (x*x' + y*y') + z*z'
PT_DEV float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
PT_DEV V3 cross(V3 a, V3 b) {
    return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}

#if PT_PARITY
PT_DEV V3 normalize(V3 v) { return v / sqrtf(dot(v, v)); }
PT_DEV float pt_sqrt(float x) { return sqrtf(x); }                    // half_sqrt := sqrtf
PT_DEV float pt_rcp_sqrt(float x) { return 1.0f / sqrtf(x); }         // 1/half_sqrt(..)  prog.cl:191,196
PT_DEV float pt_div(float a, float b) { return a / b; }
PT_DEV float pt_pow(float a, float b) { return (float)pow((double)a, (double)b); }
// theta = 2*M_PI*rnd2 is a double product narrowed to float (prog.cl:212), then cos/sin of the float
PT_DEV void pt_disk_angle(float rnd2, float &c, float &s) {
    float theta = (float)(2.0 * 3.14159265358979323846 * (double)rnd2);
    c = (float)cos((double)theta);
    s = (float)sin((double)theta);
}
#else
PT_DEV V3 normalize(V3 v) { return v * rsqrtf(dot(v, v)); }
PT_DEV float pt_sqrt(float x) { return sqrtf(x); }
PT_DEV float pt_rcp_sqrt(float x) { return rsqrtf(x); }
PT_DEV float pt_div(float a, float b) { return __fdividef(a, b); }
PT_DEV float pt_pow(float a, float b) {
    // a >= 0 at every call site (fmax(0,..) or 1-|cos|); 0^0 = 1 like powf
    if (b == 0.0f) return 1.0f;
    if (a <= 0.0f) return 0.0f;
    return exp2f(b * __log2f(a));
}
PT_DEV void pt_disk_angle(float rnd2, float &c, float &s) { sincospif(2.0f * rnd2, &s, &c); }
#endif
// one MUFU.RCP, no denormal rescue (a plane-distance denominator that small means a grazing ray: t overflows either way).
// The parity flavour uses it only inside its conservative pre-filter, never in a value that reaches the image.
PT_DEV float pt_rcp_approx(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

// OpenCL fmin/fmax/max on floats (NaN -> the other operand for fmin/fmax; prog.cl uses fmax on scalars)
PT_DEV float pt_fmax(float a, float b) { return fmaxf(a, b); }
PT_DEV float pt_fmin(float a, float b) { return fminf(a, b); }

// ---------------------------------------------------------------- RNG (prog.cl:72-77)
// n = (ulong)seed; n = n*48271 % 2147483647; seed = n; return n/2147483647.0f
// 2147483647.0f == 2^31 in float, so the division is an exact scale of (float)n.
// The modulus is the Mersenne prime 2^31-1: fold instead of dividing — bit-exact for any state
// in [0, 2^31); other (never produced by the generator itself) states take the generic path.
PT_DEV float pt_rand(uint32_t &state) {
    uint32_t r;
    if ((int32_t)state >= 0) {
        uint64_t n = (uint64_t)state * 48271ull;              // < 2^47
        r = (uint32_t)(n & 0x7fffffffull) + (uint32_t)(n >> 31);
        if (r >= 0x7fffffffu) r -= 0x7fffffffu;
    } else {
        uint64_t n = (uint64_t)(int64_t)(int32_t)state;        // sign-extends like (ulong)(int)
        n = (n * 48271ull) % 2147483647ull;
        r = (uint32_t)n;
    }
    state = r;
    return __uint2float_rn(r) * 4.656612873077392578125e-10f;  // * 2^-31
}
// The same step for a state KNOWN to lie in [0, 2^31): every state the generator itself produces, and every seed the
// reference host uploads (main.cpp:522-526).  The C-ABI verifies the uploaded buffer once (ptb_abi.cu, k_any_negative)
// before it launches a kernel built on this variant; the generic function above stays the default.
PT_DEV float pt_rand_nonneg(uint32_t &state) {
    const uint64_t n = (uint64_t)state * 48271ull;
    uint32_t r = (uint32_t)(n & 0x7fffffffull) + (uint32_t)(n >> 31);
    if (r >= 0x7fffffffu) r -= 0x7fffffffu;
    state = r;
    return __uint2float_rn(r) * 4.656612873077392578125e-10f;
}
template <bool NONNEG> PT_DEV float pt_rand_t(uint32_t &state) { return NONNEG ? pt_rand_nonneg(state) : pt_rand(state); }

#endif  // PT_MATH_CUH_
