// pt_kernels.cuh — the path-trace hot path (prog.cl:64-381, 384-389) as hand-written CUDA for
// sm_100a.  Included by exactly two translation units that differ only in arithmetic flavour
// (see pt_math.cuh); PT_NS names the namespace the entry points land in.
//
// Design (B200-first, not a translation of the OpenCL kernels):
//   * ONE fused kernel, k_render: camera-ray generation (gen_ray), the bounce loop (trace_ray) and
//     the running-mean accumulation run in registers for `spp` consecutive samples of a pixel; the
//     reference's rays[] / rnds[] / colors[] global round trips per bounce (prog.cl:319-379)
//     become one 4-byte read + 20-byte write per pixel per LAUNCH.
//   * path regeneration: a lane whose path ended starts its next sample at once, so the warp stays
//     converged in the intersection loop instead of idling until the longest path of the warp ends.
//   * warps own 8x4 pixel tiles (coherent primary rays), blocks 16x8.
//   * closest hit through one of three structures, selected per launch:
//       BRUTE   all triangles from shared memory (warp-uniform LDS.128 broadcasts)  prog.cl:113-122,318
//       REFBVH  the host-built per-object trees as uploaded                         prog.cl:123-184
//       LBVH    the GPU-built Morton LBVH (lbvh.cu), == BRUTE incl. lowest-index tie-break
//   * the drop-in single-sample kernels of the reference surface (gen_ray → rays[], trace_ray
//     reading rays[]) are the same code with spp = 1 and from_rays = 1.
#ifndef PT_NS
#error "define PT_NS"
#endif

#include <stdio.h>
#include <stdlib.h>
#include "pt_math.cuh"
#include "pt_params.h"
#include "lbvh.h"

namespace PT_NS {

// ---------------------------------------------------------------- small helpers
PT_DEV float4 ldg4(const float4 *p) { return __ldg(p); }
// 32 bytes per lane in one instruction (sm_100 LDG.E.256, p 32-byte aligned).  A scattered per-lane fetch costs one L1
// wavefront per lane and INSTRUCTION, and the LBVH traversal is bound by exactly that (l1tex__throughput 90 %,
// profiles/r01_wf_trace_v3.txt): fetching a 64-byte node with two loads instead of four halves the L1 work.
PT_DEV void ldg8(const float4 *p, float4 &a, float4 &b) {
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p));
}

struct Work {              // per-thread work counters
    unsigned rays, box, tri, tri1, tri2, tri3, samples;
};

struct Mat {               // one Material record in registers (prog.cl:1-5)
    V3 kd, ks, emission, F0;
    float n, shininess;
    int type;
};
PT_DEV Mat load_material(const RefMaterial *mats, int i) {
    const float4 *p = reinterpret_cast<const float4 *>(mats + i);
    float4 a = ldg4(p), b = ldg4(p + 1), c = ldg4(p + 2), d = ldg4(p + 3), e = ldg4(p + 4);
    Mat m;
    m.kd = v3(a); m.ks = v3(b); m.emission = v3(c); m.F0 = v3(d);
    m.n = e.x; m.shininess = e.y; m.type = __float_as_int(e.z);
    return m;
}

// ---------------------------------------------------------------- camera (prog.cl:82-92)
// px, py: id % X and id / X — the fused kernel knows them from its thread coordinates when the launch
// width equals the camera's XM (always, for the reference host), and skips the integer division
PT_DEV void camera_get_ray_xy(int px, int py, const RefCamera &cam, float rnd1, float rnd2, V3 &P, V3 &D) {
    int X = (int)cam.XM;
    int Y = (int)cam.YM;
    float x = (float)px + rnd1;
    float y = (float)py + rnd2;
    float sx = pt_div(2.0f * x, (float)X) - 1.0f;
    float sy = pt_div(2.0f * y, (float)Y) - 1.0f;
    V3 right = v3(cam.right.x, cam.right.y, cam.right.z) * sx;
    V3 up = v3(cam.up.x, cam.up.y, cam.up.z) * sy;
    V3 eye = v3(cam.eye.x, cam.eye.y, cam.eye.z);
    V3 p = (v3(cam.lookat.x, cam.lookat.y, cam.lookat.z) + right) + up;
    P = eye;
    D = normalize(p - eye);
}
// Launch-level form: the parity flavour evaluates prog.cl's expressions literally; the fast flavour uses the
// folded affine map of pt_params.h (6 FMAs + normalize instead of two divisions and the basis arithmetic)
PT_DEV void camera_ray(const RenderParams &p, int px, int py, float rnd1, float rnd2, V3 &P, V3 &D) {
#if PT_PARITY
    camera_get_ray_xy(px, py, p.cam, rnd1, rnd2, P, D);
#else
    const float x = (float)px + rnd1, y = (float)py + rnd2;
    P = v3(p.cam.eye.x, p.cam.eye.y, p.cam.eye.z);
    D = normalize(v3(fmaf(p.cam_dy.x, y, fmaf(p.cam_dx.x, x, p.cam_o.x)),
                     fmaf(p.cam_dy.y, y, fmaf(p.cam_dx.y, x, p.cam_o.y)),
                     fmaf(p.cam_dy.z, y, fmaf(p.cam_dx.z, x, p.cam_o.z))));
#endif
}
PT_DEV void camera_get_ray(int id, const RefCamera &cam, float rnd1, float rnd2, V3 &P, V3 &D) {
    int X = (int)cam.XM;
    camera_get_ray_xy(id % X, id / X, cam, rnd1, rnd2, P, D);
}

// ---------------------------------------------------------------- ray–triangle (prog.cl:94-112)
// BRANCH-FREE: every lane evaluates the whole test and the acceptance is a predicate, so a warp
// that scans a triangle list stays converged (the early returns of the reference cost ~3x in
// divergence: measured 11 of 32 lanes active, profiles/r01_*).  The predicate equals the
// reference's nested ifs, NaNs included: accept <=> !(t<0) && e1>=0 && e2>=0 && e3>=0.
//
// The three float4 of a triangle mean different things in the two arithmetic flavours:
//   parity: (r1,N.x) (r2,N.y) (r3,N.z)  — the uploaded record, evaluated with prog.cl's own expressions
//   fast:   (N, r1.x) (r1.y, r1.z, mu.x, mu.y) (mu.z, mv)  — unit normal, anchor vertex and the two
//           barycentric gradients mu = (e2 x N)/(N.(e1 x e2)), mv = (N x e1)/(N.(e1 x e2)) precomputed at
//           upload (ptb_abi.cu): u = mu.w, v = mv.w with w = p - r1; inside <=> u>=0, v>=0, u+v<=1, the
//           same three half-planes as the reference's edge tests (edge1 <=> v>=0, edge3 <=> u>=0, edge2 <=> u+v<=1)
#if PT_PARITY
PT_DEV V3 tri_normal(float4 a, float4 b, float4 c) { return v3(a.w, b.w, c.w); }
template <bool COUNT>
PT_DEV bool tri_accept(float4 q0, float4 q1, float4 q2, V3 P, V3 V, float &t, Work &wk) {
    V3 r1 = v3(q0), r2 = v3(q1), r3 = v3(q2), N = v3(q0.w, q1.w, q2.w);
    t = pt_div(dot(r1 - P, N), dot(V, N));
    V3 p = P + V * t;
    const bool s0 = !(t < 0.0f);
    const bool s1 = dot(cross(r2 - r1, p - r1), N) >= 0.0f;
    const bool s2 = dot(cross(r3 - r2, p - r2), N) >= 0.0f;
    const bool s3 = dot(cross(r1 - r3, p - r3), N) >= 0.0f;
    if (COUNT) { wk.tri++; wk.tri1 += s0; wk.tri2 += s0 & s1; wk.tri3 += s0 & s1 & s2; }
    return s0 & s1 & s2 & s3;
}
#else
PT_DEV V3 tri_normal(float4 a, float4, float4) { return v3(a.x, a.y, a.z); }
template <bool COUNT>
PT_DEV bool tri_accept(float4 A, float4 B, float4 C, V3 P, V3 V, float &t, Work &wk) {
    const V3 N = v3(A.x, A.y, A.z);
    const V3 w0 = v3(A.w - P.x, B.x - P.y, B.y - P.z);          // r1 - P
    t = dot(w0, N) * pt_rcp_approx(dot(V, N));
    const V3 w = V * t - w0;                                     // p - r1
    const float u = dot(v3(B.z, B.w, C.x), w);
    const float v = dot(v3(C.y, C.z, C.w), w);
    const float k = fminf(fminf(u, v), 1.0f - (u + v));
    if (COUNT) { wk.tri++; const bool s0 = !(t < 0.0f); wk.tri1 += s0; wk.tri2 += s0 & (v >= 0.0f); wk.tri3 += s0 & (v >= 0.0f) & (u + v <= 1.0f); }
    return k >= 0.0f;                                            // t > 0 is applied by closest_update (t NaN fails it)
}
#endif

struct Closest {
    float t;     // +inf: none yet (converted to the reference's -1 by the caller)
    int tri;     // index in upload order
};
#define PT_NO_HIT __int_as_float(0x7f800000)

// first_intersect (prog.cl:113-122): accept t > 0 and strictly closer — the lowest index wins ties
PT_DEV void closest_update(Closest &best, bool accept, float t, int i) {
    const bool better = accept & (t > 0.0f) & (t < best.t);
    best.t = better ? t : best.t;
    best.tri = better ? i : best.tri;
}
template <bool COUNT>
PT_DEV void brute_smem(const float4 *s_tris, int n, V3 P, V3 V, Closest &best, Work &wk) {
#pragma unroll 4
    for (int i = 0; i < n; ++i) {
        float t;
        const bool acc = tri_accept<COUNT>(s_tris[3 * i], s_tris[3 * i + 1], s_tris[3 * i + 2], P, V, t, wk);
        closest_update(best, acc, t, i);
    }
}
template <bool COUNT>
PT_DEV void brute_global(const SceneDev &sc, int from, int to, V3 P, V3 V, Closest &best, Work &wk) {
    for (int i = from; i < to; ++i) {
        float t;
        const bool acc = tri_accept<COUNT>(ldg4(sc.q0 + i), ldg4(sc.q1 + i), ldg4(sc.q2 + i), P, V, t, wk);
        closest_update(best, acc, t, i);
    }
}

// ---- Blackwell packed-FP32 scan: two triangles per step with FFMA2 / FADD2 / FMUL2 (sm_100 f32x2)
// The brute-force scan is issue-bound (86 % issue-slot utilisation, profiles/r01_render_v3.txt), so
// the fix is fewer instructions: lane .x of every float2 carries triangle 2j, lane .y triangle 2j+1,
// and the 21 FP32 operations of the test issue once for both.  Shared-memory layout per pair,
// 6 float4 (built by the block at kernel start from the fast per-triangle layout):
//   (N.x N.y) (N.z -r1.x) (-r1.y -r1.z) (mu.x mu.y) (mu.z mv.x) (mv.y mv.z), each component as an (A,B) pair
// With w0 = P - r1:  t = (N.w0) / (N.(-V)),  w = p - r1 = w0 + V t,  u = mu.w, v = mv.w, k = min(u, v, 1-u-v).
struct RayPair {                 // the ray broadcast into both lanes
    float2 Px, Py, Pz, Vx, Vy, Vz, nVx, nVy, nVz;
};
PT_DEV RayPair ray_pair(V3 P, V3 V) {
    RayPair r;
    r.Px = make_float2(P.x, P.x); r.Py = make_float2(P.y, P.y); r.Pz = make_float2(P.z, P.z);
    r.Vx = make_float2(V.x, V.x); r.Vy = make_float2(V.y, V.y); r.Vz = make_float2(V.z, V.z);
    r.nVx = make_float2(-V.x, -V.x); r.nVy = make_float2(-V.y, -V.y); r.nVz = make_float2(-V.z, -V.z);
    return r;
}
PT_DEV float2 dot2(float2 ax, float2 ay, float2 az, float2 bx, float2 by, float2 bz) {
    return __ffma2_rn(az, bz, __ffma2_rn(ay, by, __fmul2_rn(ax, bx)));
}
#if !PT_PARITY
template <int U>
PT_DEV void brute_smem_pairs(const float4 *s_pairs, int n_pairs, V3 P, V3 V, Closest &best) {
    const RayPair r = ray_pair(P, V);
    const float2 one = make_float2(1.0f, 1.0f), neg1 = make_float2(-1.0f, -1.0f);
#pragma unroll U
    for (int j = 0; j < n_pairs; ++j) {
        const float4 L0 = s_pairs[6 * j], L1 = s_pairs[6 * j + 1], L2 = s_pairs[6 * j + 2];
        const float4 L3 = s_pairs[6 * j + 3], L4 = s_pairs[6 * j + 4], L5 = s_pairs[6 * j + 5];
        const float2 Nx = make_float2(L0.x, L0.y), Ny = make_float2(L0.z, L0.w), Nz = make_float2(L1.x, L1.y);
        const float2 wx = __fadd2_rn(make_float2(L1.z, L1.w), r.Px);          // w0 = P - r1
        const float2 wy = __fadd2_rn(make_float2(L2.x, L2.y), r.Py);
        const float2 wz = __fadd2_rn(make_float2(L2.z, L2.w), r.Pz);
        const float2 a = dot2(Nx, Ny, Nz, wx, wy, wz);                         // N.(P - r1)
        const float2 b = dot2(Nx, Ny, Nz, r.nVx, r.nVy, r.nVz);                // N.(-V)
        const float2 t = __fmul2_rn(a, make_float2(pt_rcp_approx(b.x), pt_rcp_approx(b.y)));
        const float2 px = __ffma2_rn(r.Vx, t, wx), py = __ffma2_rn(r.Vy, t, wy), pz = __ffma2_rn(r.Vz, t, wz);   // p - r1
        const float2 u = dot2(make_float2(L3.x, L3.y), make_float2(L3.z, L3.w), make_float2(L4.x, L4.y), px, py, pz);
        const float2 v = dot2(make_float2(L4.z, L4.w), make_float2(L5.x, L5.y), make_float2(L5.z, L5.w), px, py, pz);
        const float2 q = __ffma2_rn(__fadd2_rn(u, v), neg1, one);              // 1 - (u + v)
        closest_update(best, fminf(fminf(u.x, v.x), q.x) >= 0.0f, t.x, 2 * j);
        closest_update(best, fminf(fminf(u.y, v.y), q.y) >= 0.0f, t.y, 2 * j + 1);
    }
}
// The same scan over merged primitives (prims.cu): a convex planar quad (two coplanar triangles sharing a
// diagonal) costs one triangle test: inside iff min(u, v, 1 - u - b*v, 1 - v - a*u) >= 0, with (b, a) = (1, 0)
// for a lone triangle.  7 float4 per pair, the last holds (-b.x, -b.y, -a.x, -a.y).  best.tri = PRIMITIVE index.
// COUNT: work counters of the SHIPPING scan: one test per primitive actually tested, stage exits of the quad
// test in the reference's order (t >= 0, then the half-planes v >= 0 and 1 - u - b*v >= 0); n_valid = primitives without the padding one
template <int U, bool COUNT = false>
PT_DEV void brute_smem_prims(const float4 *s_pairs, int n_pairs, V3 P, V3 V, Closest &best, Work *wk = nullptr, int n_valid = 0) {
    const RayPair r = ray_pair(P, V);
    const float2 one = make_float2(1.0f, 1.0f), neg1 = make_float2(-1.0f, -1.0f);
#pragma unroll U
    for (int j = 0; j < n_pairs; ++j) {
        const float4 L0 = s_pairs[7 * j], L1 = s_pairs[7 * j + 1], L2 = s_pairs[7 * j + 2];
        const float4 L3 = s_pairs[7 * j + 3], L4 = s_pairs[7 * j + 4], L5 = s_pairs[7 * j + 5], L6 = s_pairs[7 * j + 6];
        const float2 Nx = make_float2(L0.x, L0.y), Ny = make_float2(L0.z, L0.w), Nz = make_float2(L1.x, L1.y);
        const float2 wx = __fadd2_rn(make_float2(L1.z, L1.w), r.Px);
        const float2 wy = __fadd2_rn(make_float2(L2.x, L2.y), r.Py);
        const float2 wz = __fadd2_rn(make_float2(L2.z, L2.w), r.Pz);
        const float2 a = dot2(Nx, Ny, Nz, wx, wy, wz);
        const float2 b = dot2(Nx, Ny, Nz, r.nVx, r.nVy, r.nVz);
        const float2 t = __fmul2_rn(a, make_float2(pt_rcp_approx(b.x), pt_rcp_approx(b.y)));
        const float2 px = __ffma2_rn(r.Vx, t, wx), py = __ffma2_rn(r.Vy, t, wy), pz = __ffma2_rn(r.Vz, t, wz);
        const float2 u = dot2(make_float2(L3.x, L3.y), make_float2(L3.z, L3.w), make_float2(L4.x, L4.y), px, py, pz);
        const float2 v = dot2(make_float2(L4.z, L4.w), make_float2(L5.x, L5.y), make_float2(L5.z, L5.w), px, py, pz);
        const float2 q = __ffma2_rn(v, make_float2(L6.x, L6.y), __ffma2_rn(u, neg1, one));     // 1 - u - b*v
        const float2 w = __ffma2_rn(u, make_float2(L6.z, L6.w), __ffma2_rn(v, neg1, one));     // 1 - v - a*u
        if (COUNT) {
            const bool sa = !(t.x < 0.0f), sb = !(t.y < 0.0f) && 2 * j + 1 < n_valid;
            wk->tri += 1 + (2 * j + 1 < n_valid);
            wk->tri1 += sa + sb;
            wk->tri2 += (sa & (v.x >= 0.0f)) + (sb & (v.y >= 0.0f));
            wk->tri3 += (sa & (v.x >= 0.0f) & (q.x >= 0.0f)) + (sb & (v.y >= 0.0f) & (q.y >= 0.0f));
        }
        closest_update(best, fminf(fminf(u.x, v.x), fminf(q.x, w.x)) >= 0.0f, t.x, 2 * j);
        closest_update(best, fminf(fminf(u.y, v.y), fminf(q.y, w.y)) >= 0.0f, t.y, 2 * j + 1);
    }
}
#endif

#if PT_PARITY
// ---- Parity flavour: conservative pre-filter in front of the exact test
// The exact test (tri_accept above: true divisions, three cross products, ~100 instructions) decides every hit, so the
// image stays bit-identical to prog.cl; the filter only removes triangles the exact test is CERTAIN to reject, at the
// packed-FP32 cost of the fast flavour's test.  With e1 = r2-r1, e2 = r3-r1, den = N.(e1 x e2) and the decomposition
// p - r1 = u e1 + v e2 + h N, prog.cl's three edge expressions are, in exact arithmetic,
//     edge1 = v den,   edge2 = (1-u-v) den,   edge3 = u den                 (N.(e x N) = 0 kills the h term)
// and the fast layout's (mu, mv) produce exactly this u, v.  In float arithmetic (unit roundoff eps = 2^-24):
//   * an edge expression that prog.cl computes as >= 0 is >= -16 eps |e_k| |p - v_k| exactly, i.e. its barycentric
//     coordinate is >= -16 eps L (|p - r1| + L) / den, L = the longest edge;
//   * the filter's t' and prog.cl's t both approximate N.(r1-P)/N.V; their difference is below
//     eps (|r1-P| + |t|) (19/|b| + 6) with b = N.V, as long as |b| is not within rounding of 0;
//   * a barycentric coordinate changes by at most L/den per unit of distance, the roundings of P + V t and of the
//     filter's own dot products add about 8 eps (|P| + |t| + |r1-P|) L / den.
// Summing: |filter coordinate - reference coordinate| < eps (L/den) D (40 + 19/|b|) with D = 2|P|_1 + 2(L + |r1|_1) + |t'|
// (an upper bound of every length above).  The filter rejects only when
//     min(u', v', 1-u'-v') < -delta,  delta = 64 eps (L/den) D g,   g = 2 + 1/|b'| + 2^-11/|b'|^2     (> 3x the bound)
//     or t' < -dt,                    dt    = 64 eps D g                                              (prog.cl: t < 0 misses)
// By the same bound a triangle with min(u', v', 1-u'-v') > +delta is CERTAIN to pass prog.cl's three edge comparisons
// whatever the sign of t (each edge expression is positive by more than its own rounding error, and the bound covers
// the whole t' - t difference), so only prog.cl's t is evaluated for it — every bounce ray leaves a surface at
// t ~ +-1e-3, inside dt, so the sign of t is never taken from the filter.  prog.cl's whole test runs only on what
// lies within delta of an edge, a ~1e-4-wide band.
// The quadratic term makes grazing rays (|b'| < 2^-17, where b' is rounding noise) candidates unconditionally:
// g >= 2^23 there, so delta >= 32 D L/den > |any coordinate| and dt >= 32 D > |t'|.  NaN or inf anywhere compares
// false and the triangle stays a candidate; a record whose den is not positive (normal against the winding, or a
// degenerate triangle) gets c = inf and is always a candidate.  Survivors take the exact test in upload order, so ties
// still go to the lower index (prog.cl:118).
// Shared memory per 32 triangles: 16 pair records of 7 float4 — the six of brute_smem_pairs and (c.A c.B Lt.A Lt.B),
// c = 2^-18 L/den, Lt = 2 (L + |r1|_1).
#define PT_FILTER_MAX_TRIS 1024
PT_DEV void filter_record(float4 r1n, float4 r2n, float4 r3n, float &c, float &Lt) {
    const V3 r1 = v3(r1n), r2 = v3(r2n), r3 = v3(r3n), N = v3(r1n.w, r2n.w, r3n.w);
    const V3 e1 = r2 - r1, e2 = r3 - r1, e3 = r3 - r2;
    const float L = sqrtf(fmaxf(fmaxf(dot(e1, e1), dot(e2, e2)), dot(e3, e3))) * 1.0001f;
    const float den = dot(N, cross(e1, e2));
    c = den > 0.0f ? 3.8147e-6f * (L / den) * 1.0001f : PT_NO_HIT;        // 2^-18 = 3.8146973e-6, rounded up
    Lt = 2.0f * (L + (fabsf(r1.x) + fabsf(r1.y) + fabsf(r1.z)));
}
PT_DEV void brute_smem_filtered(const float4 *s_exact, const float4 *s_filt, int n, V3 P, V3 V, Closest &best, Work &wk) {
    const RayPair r = ray_pair(P, V);
    const float Pn = 2.0f * (fabsf(P.x) + fabsf(P.y) + fabsf(P.z));
    const float2 Pn2 = make_float2(Pn, Pn), one = make_float2(1.0f, 1.0f), neg1 = make_float2(-1.0f, -1.0f);
    const float2 two = make_float2(2.0f, 2.0f), q11 = make_float2(4.8828125e-4f, 4.8828125e-4f);   // 2^-11
    for (int base = 0; base < n; base += 32) {
        unsigned mask = 0, sure = 0;
        const float4 *f = s_filt + 7 * (base >> 1);
#pragma unroll 1
        for (int jj = 0; jj < 16; jj += 4) {
        unsigned m8 = 0, s8 = 0;
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) {
            const int j = jj + k4;
            const float4 L0 = f[7 * j], L1 = f[7 * j + 1], L2 = f[7 * j + 2], L3 = f[7 * j + 3];
            const float4 L4 = f[7 * j + 4], L5 = f[7 * j + 5], L6 = f[7 * j + 6];
            const float2 Nx = make_float2(L0.x, L0.y), Ny = make_float2(L0.z, L0.w), Nz = make_float2(L1.x, L1.y);
            const float2 wx = __fadd2_rn(make_float2(L1.z, L1.w), r.Px);
            const float2 wy = __fadd2_rn(make_float2(L2.x, L2.y), r.Py);
            const float2 wz = __fadd2_rn(make_float2(L2.z, L2.w), r.Pz);
            const float2 a = dot2(Nx, Ny, Nz, wx, wy, wz);
            const float2 b = dot2(Nx, Ny, Nz, r.nVx, r.nVy, r.nVz);
            const float2 inv = make_float2(pt_rcp_approx(b.x), pt_rcp_approx(b.y));
            const float2 t = __fmul2_rn(a, inv);
            const float2 px = __ffma2_rn(r.Vx, t, wx), py = __ffma2_rn(r.Vy, t, wy), pz = __ffma2_rn(r.Vz, t, wz);
            const float2 u = dot2(make_float2(L3.x, L3.y), make_float2(L3.z, L3.w), make_float2(L4.x, L4.y), px, py, pz);
            const float2 v = dot2(make_float2(L4.z, L4.w), make_float2(L5.x, L5.y), make_float2(L5.z, L5.w), px, py, pz);
            const float2 q = __ffma2_rn(__fadd2_rn(u, v), neg1, one);
            const float2 ai = make_float2(fabsf(inv.x), fabsf(inv.y));
            const float2 g = __ffma2_rn(__ffma2_rn(ai, q11, one), ai, two);                 // 2 + ai + 2^-11 ai^2
            const float2 D = __fadd2_rn(__fadd2_rn(Pn2, make_float2(L6.z, L6.w)), make_float2(fabsf(t.x), fabsf(t.y)));
            const float2 dg = __fmul2_rn(D, g);
            const float2 delta = __fmul2_rn(make_float2(L6.x, L6.y), dg);
            const float2 dt = __fmul2_rn(make_float2(3.8147e-6f, 3.8147e-6f), dg);
            const bool rej_a = (fminf(fminf(u.x, v.x), q.x) < -delta.x) | (t.x < -dt.x);
            const bool rej_b = (fminf(fminf(u.y, v.y), q.y) < -delta.y) | (t.y < -dt.y);
            m8 |= rej_a ? 0u : 1u << (2 * k4);
            m8 |= rej_b ? 0u : 2u << (2 * k4);
            const bool sure_a = fminf(fminf(u.x, v.x), q.x) > delta.x;
            const bool sure_b = fminf(fminf(u.y, v.y), q.y) > delta.y;
            s8 |= sure_a ? 1u << (2 * k4) : 0u;
            s8 |= sure_b ? 2u << (2 * k4) : 0u;
        }
        mask |= m8 << (2 * jj);
        sure |= s8 << (2 * jj);
        }
        const int left = n - base;
        if (left < 32) mask &= (1u << left) - 1u;
        while (mask) {                                           // survivors in upload order
            const unsigned bit = mask & (0u - mask);
            const int i = base + __ffs(mask) - 1;
            mask ^= bit;
            const float4 q0 = s_exact[3 * i], q1 = s_exact[3 * i + 1], q2 = s_exact[3 * i + 2];
            if (sure & bit) {                                    // certainly inside: prog.cl's t decides alone (t > 0 in closest_update)
                const V3 r1 = v3(q0), N = v3(q0.w, q1.w, q2.w);
                closest_update(best, true, pt_div(dot(r1 - P, N), dot(V, N)), i);
            } else {                                             // within the margin of an edge: prog.cl's whole test
                float t;
                const bool acc = tri_accept<false>(q0, q1, q2, P, V, t, wk);
                closest_update(best, acc, t, i);
            }
        }
    }
}
#endif

// ---------------------------------------------------------------- reference trees (prog.cl:123-184)
// BBox_intersection with true divisions and OpenCL fmin/fmax NaN rules; kd_intersect's control flow:
// per object an explicit DFS, left child first, prune on miss / tmax<0 / entry beyond this object's
// current best (tmin starts at 999999 for every object), best hit kept across objects by strict '<'.
PT_DEV bool bbox_intersection(const RefNode *node, V3 P, V3 D, float &tmin, float &tmax) {
    const float4 *p = reinterpret_cast<const float4 *>(node);
    float4 bl = ldg4(p + 1), tr = ldg4(p + 2);
    float tx1 = pt_div(bl.x - P.x, D.x), tx2 = pt_div(tr.x - P.x, D.x);
    tmin = pt_fmin(tx1, tx2);
    tmax = pt_fmax(tx1, tx2);
    float ty1 = pt_div(bl.y - P.y, D.y), ty2 = pt_div(tr.y - P.y, D.y);
    tmin = pt_fmax(tmin, pt_fmin(ty1, ty2));
    tmax = pt_fmin(tmax, pt_fmax(ty1, ty2));
    float tz1 = pt_div(bl.z - P.z, D.z), tz2 = pt_div(tr.z - P.z, D.z);
    tmin = pt_fmax(tmin, pt_fmin(tz1, tz2));
    tmax = pt_fmin(tmax, pt_fmax(tz1, tz2));
    return tmax >= tmin;
}
template <bool COUNT>
PT_DEV void refbvh_intersect(const SceneDev &sc, V3 P, V3 D, Closest &best, Work &wk) {
    int stack[64];   // the reference reserves 300 ints; a heap-indexed tree cannot be deeper than 31
    for (int o = 0; o < sc.n_obj; ++o) {
        const int shift = __ldg(sc.kd_shift + o);
        int ptr = 1 + shift;
        float tmin = 999999.0f;
        int sp = 0;
        for (;;) {
            const RefNode *node = sc.kd_tree + ptr;
            float dist, tmax;
            if (COUNT) wk.box++;
            bool descend = false;
            if (bbox_intersection(node, P, D, dist, tmax) && tmax >= 0.0f && !(dist > tmin)) {
                int2 trii = __ldg(reinterpret_cast<const int2 *>(node));
                if (trii.x < 0) {
                    if (sp < 64) stack[sp++] = 2 * (ptr - shift) + 1 + shift;
                    ptr = 2 * (ptr - shift) + shift;
                    descend = true;
                } else {
                    // first_intersect over the leaf, then "keep track of the best hit" with strict '<' (prog.cl:168-172):
                    // one running best in visit order is the same thing; tmin follows only real improvements
                    const float before = best.t;
                    brute_global<COUNT>(sc, trii.x, trii.y, P, D, best, wk);
                    if (best.t < before) tmin = best.t;
                }
            }
            if (!descend) {
                if (sp == 0) break;
                ptr = stack[--sp];
            }
        }
    }
}

// ---------------------------------------------------------------- GPU LBVH (lbvh.cu)
// Exactness contract: the result equals brute force over the upload order, i.e. the minimum of
// (t, index) over all triangles the reference test accepts with t > 0.  Boxes are inflated at
// build time and the slab interval is widened here, so a box is never culled while it can hold
// an accepted hit with t <= best.t; equal-t candidates are resolved by index, not visit order.
// Node = 4 float4 holding both children's boxes (see lbvh.cu); near child first, far child pushed.
#define PT_LBVH_STACK 64
PT_DEV float safe_rcp(float d) {
    // zero (or denormal) direction components: a huge finite reciprocal keeps (lo-P)*inv free of 0*inf NaNs
    return fabsf(d) > 1e-30f ? 1.0f / d : copysignf(1e30f, d);
}
// The outlier triangles the builder kept out of the tree (lbvh.cu step 2: room walls, emitter quads — at most
// LBVH_BIG_MAX): tested once per ray BEFORE the tree, which also hands the traversal a finite best.t to cull with.
// Same acceptance and tie rule as a leaf, so closest hit stays the minimum of (t, upload index) over ALL triangles.
template <bool COUNT>
PT_DEV void lbvh_big_scan(const SceneDev &sc, V3 P, V3 D, Closest &best, Work &wk) {
    for (int k = 0; k < sc.big_count; ++k) {
        const int pos = sc.big_first + k;
        float t;
        const bool acc_ = tri_accept<COUNT>(ldg4(sc.sq0 + pos), ldg4(sc.sq1 + pos), ldg4(sc.sq2 + pos), P, D, t, wk);
        if (acc_ & (t > 0.0f) & (t <= best.t)) {
            const int tid = __ldg(sc.sorted_id + pos);
            if (t < best.t || tid < best.tri) { best.t = t; best.tri = tid; }
        }
    }
}
template <bool COUNT>
PT_DEV void lbvh_intersect(const SceneDev &sc, V3 P, V3 D, Closest &best, Work &wk) {
    lbvh_big_scan<COUNT>(sc, P, D, best, wk);
    const float ix = safe_rcp(D.x), iy = safe_rcp(D.y), iz = safe_rcp(D.z);
    int stack[PT_LBVH_STACK];
    int sp = 0;
    int link = sc.bvh_root;
    for (;;) {
        while (link >= 0) {
            const float4 *nd = sc.bvh + 4 * (size_t)link;
            float4 n0 = ldg4(nd), n1 = ldg4(nd + 1), n2 = ldg4(nd + 2), n3 = ldg4(nd + 3);
            if (COUNT) wk.box += 2;
            float a, b, l0, l1, r0, r1;
            a = (n0.x - P.x) * ix; b = (n0.y - P.x) * ix; l0 = fminf(a, b); l1 = fmaxf(a, b);
            a = (n0.z - P.y) * iy; b = (n0.w - P.y) * iy; l0 = fmaxf(l0, fminf(a, b)); l1 = fminf(l1, fmaxf(a, b));
            a = (n2.x - P.z) * iz; b = (n2.y - P.z) * iz; l0 = fmaxf(l0, fminf(a, b)); l1 = fminf(l1, fmaxf(a, b));
            a = (n1.x - P.x) * ix; b = (n1.y - P.x) * ix; r0 = fminf(a, b); r1 = fmaxf(a, b);
            a = (n1.z - P.y) * iy; b = (n1.w - P.y) * iy; r0 = fmaxf(r0, fminf(a, b)); r1 = fminf(r1, fmaxf(a, b));
            a = (n2.z - P.z) * iz; b = (n2.w - P.z) * iz; r0 = fmaxf(r0, fminf(a, b)); r1 = fminf(r1, fmaxf(a, b));
            // widen both ends by 8 ulp (each product above carries <= 1.5 ulp)
            l0 -= fabsf(l0) * 1.0e-6f; l1 += fabsf(l1) * 1.0e-6f;
            r0 -= fabsf(r0) * 1.0e-6f; r1 += fabsf(r1) * 1.0e-6f;
            const bool hl = l0 <= l1 && l1 >= 0.0f && l0 <= best.t;
            const bool hr = r0 <= r1 && r1 >= 0.0f && r0 <= best.t;
            const int ll = __float_as_int(n3.x), lr = __float_as_int(n3.y);
            if (hl && hr) {
                const bool left_first = l0 <= r0;
                if (sp < PT_LBVH_STACK) stack[sp++] = left_first ? lr : ll;
                link = left_first ? ll : lr;
            } else if (hl) {
                link = ll;
            } else if (hr) {
                link = lr;
            } else {
                if (sp == 0) return;
                link = stack[--sp];
            }
        }
        {   // leaf: [first, first+count) in Morton order
            const int enc = ~link;
            const int first = enc >> 3, count = (enc & 7) + 1;
            for (int k = 0; k < count; ++k) {
                const int pos = first + k;
                float t;
                const bool acc = tri_accept<COUNT>(ldg4(sc.sq0 + pos), ldg4(sc.sq1 + pos), ldg4(sc.sq2 + pos), P, D, t, wk);
                if (acc & (t > 0.0f) & (t <= best.t)) {
                    const int id = __ldg(sc.sorted_id + pos);
                    if (t < best.t || id < best.tri) { best.t = t; best.tri = id; }
                }
            }
        }
        if (sp == 0) return;
        link = stack[--sp];
    }
}

// ---------------------------------------------------------------- bounce sampling (prog.cl:186-245)
PT_DEV void orthonormal_base(V3 v1, V3 &v2, V3 &v3o) {
    const float E = 0.001f;
    if (fabsf(v1.x) <= E && fabsf(v1.z) <= E) {
        float rl = pt_rcp_sqrt(v1.y * v1.y + v1.z * v1.z);
        v2 = v3(0.0f, -v1.z * rl, v1.y * rl);
    } else {
        float rl = pt_rcp_sqrt(v1.x * v1.x + v1.z * v1.z);
        v2 = v3(-v1.z * rl, 0.0f, v1.x * rl);
    }
    v3o = cross(v1, v2);
}
// frame: optional precomputed (X, Z) of the UNFLIPPED stored normal (prims.cu); flipped: N is its negation, for which
// orthonormal_base yields (X, -Z)
PT_DEV void new_ray_diffuse(V3 hitP, V3 N, float rnd1, float rnd2, V3 &P, V3 &D, const V3 *frame = nullptr, bool flipped = false) {
    V3 X, Y = N, Z;
    if (frame) { X = frame[0]; Z = flipped ? -frame[1] : frame[1]; }
    else orthonormal_base(Y, Z, X);
    float r = pt_sqrt(rnd1);
    float c, s;
    pt_disk_angle(rnd2, c, s);
    float x = r * c, y = r * s, z = pt_sqrt(1.0f - rnd1);
    D = normalize((X * x + Y * z) + Z * y);
    P = hitP + Y * 0.001f;
}
PT_DEV V3 fresnel(const Mat &m, V3 N, V3 oldD) {
    float cosa = fabsf(dot(N, oldD));
    return m.F0 + (1.0f - m.F0) * pt_pow(1.0f - cosa, 5.0f);
}
PT_DEV void new_ray_specular(V3 hitP, V3 N, V3 oldD, V3 &P, V3 &D) {
    float cosa = dot(N, oldD);
    D = normalize(oldD - (N * cosa) * 2.0f);
    P = hitP + N * 0.001f;
}

// colors = (colors*s + color)/(s+1), prog.cl:379 — the parity flavour divides, the fast one multiplies by 1/(s+1)
PT_DEV V3 pt_mean_step(V3 acc, V3 color, int s) {
#if PT_PARITY
    return (acc * (float)s + color) / (float)(s + 1);
#else
    return (acc * (float)s + color) * pt_rcp_approx((float)(s + 1));
#endif
}

// per-path throughput state of trace_ray (prog.cl:307-316)
struct Path {
    V3 fL, fB, fS, fR, color;
    bool in;
};
PT_DEV void path_reset(Path &p) {
    p.fL = p.fB = p.fS = p.fR = v3(1.0f, 1.0f, 1.0f);
    p.color = v3(0.0f, 0.0f, 0.0f);
    p.in = false;
}

// One iteration of the bounce loop body after a hit (prog.cl:321-366).  hitN is the stored
// triangle normal (unflipped).  Updates the ray and the path state.
template <bool NONNEG = false>
PT_DEV void shade_hit(const Mat &m, V3 hitP, V3 hitN, V3 &rayP, V3 &rayD, uint32_t &rng, Path &ps, V3 eye, int iterations,
                      const V3 *frame = nullptr) {
    if (iterations == 1) ps.color = m.kd + m.emission;                       // flat preview, prog.cl:323-325
    V3 N = hitN;
    const bool flipped = dot(rayD, N) > 0.0f;
    if (flipped) N = -N;                                                    // prog.cl:326-328
    if (m.type == 0 || m.type == 3) {
        // diffuse + Phong lobe (prog.cl:329-340) and emitter (prog.cl:358-366) share the cosine-weighted
        // bounce, so the warp stays converged through the sampling code
        const float intensity = pt_fmax(0.0f, dot(-rayD, N));               // emitter only: uses the OLD ray
        float r1 = pt_rand_t<NONNEG>(rng), r2 = pt_rand_t<NONNEG>(rng);
        V3 nP, nD;
        new_ray_diffuse(hitP, N, r1, r2, nP, nD, frame, flipped);
        if (m.type == 0) {
            float cos_theta = dot(nD, N);
            ps.fL = ps.fL * (m.kd * pt_fmax(0.0f, cos_theta));
            if (m.ks.x == 0.0f && m.ks.y == 0.0f && m.ks.z == 0.0f && m.shininess >= 0.0f) {
                // ks == 0: ks*pow(..) is exactly +0 for any finite pow (its base is in [0,1]), skip the pow
                ps.fB = ps.fB * v3(0.0f, 0.0f, 0.0f);
            } else {
                V3 halfway = normalize(normalize(eye - hitP) + nD);
                float cos_delta = dot(N, halfway);
                ps.fB = ps.fB * (m.ks * pt_pow(pt_fmax(0.0f, cos_delta), m.shininess));
            }
        } else {
            ps.color = ps.color + (((m.emission * (ps.fL + ps.fB)) * ps.fS) * ps.fR) * intensity;
        }
        rayP = nP; rayD = nD;
    } else if (m.type == 1) {                                               // mirror, prog.cl:341-345
        V3 oldD = rayD;
        new_ray_specular(hitP, N, oldD, rayP, rayD);
        ps.fS = ps.fS * fresnel(m, N, oldD);
    } else if (m.type == 2) {                                               // glass, prog.cl:346-357 + 228-245
        V3 oldD = rayD;
        bool before = ps.in;
        float rnd = pt_rand_t<NONNEG>(rng);
        float n = ps.in ? 1.0f / m.n : m.n;
        float cosa = dot(-oldD, N);
        float disc = 1.0f - pt_div(pt_div(1.0f - cosa * cosa, n), n);
        V3 F = fresnel(m, N, oldD);
        float prob = ((F.x + F.y) + F.z) / 3.0f;
        if (disc > 0.0f && rnd > prob) {
            ps.in = !ps.in;
            rayP = hitP - N * 0.001f;
            rayD = normalize(oldD / n + N * (cosa / n - pt_sqrt(disc)));
        } else {
            new_ray_specular(hitP, N, oldD, rayP, rayD);
        }
#if PT_PARITY
        if (before != ps.in) ps.fR = (ps.fR * (1.0f - F)) * (1.0f / (1.0f - prob));
        else ps.fR = (ps.fR * F) * (1.0f / prob);
#else   // fS and fR only ever appear as the product fS*fR (prog.cl:363): the fast flavour keeps one factor, fR stays 1
        if (before != ps.in) ps.fS = (ps.fS * (1.0f - F)) * (1.0f / (1.0f - prob));
        else ps.fS = (ps.fS * F) * (1.0f / prob);
#endif
    }
    // any other type: ray unchanged, nothing added
}

// ---------------------------------------------------------------- tone map (prog.cl:247-269)
PT_DEV float srgb1(float a) { return a <= 0.00304f ? 12.92f * a : 1.055f * pt_pow(a, 0.4167f) - 0.055f; }
PT_DEV float4 reinhard_tone_map(V3 c) {
    float L = (0.2126f * c.x + 0.7152f * c.y) + 0.0722f * c.z;
    float L2 = L / (1.0f + L);
    c = (c * L2) / L;                     // L == 0 → NaN, as in the reference
    return make_float4(srgb1(c.x), srgb1(c.y), srgb1(c.z), 1.0f);
}

// ---------------------------------------------------------------- the fused kernel
#define PT_BLOCK 128
#define PT_SMEM_MAX_TRIS 4096   // 192 KB of triangles at most in shared memory

// NONNEG: the launch is known to start from RNG states in [0, 2^31) (see pt_rand_nonneg)
template <int ACCEL, bool COUNT, int SMEM, int MINB, int U, bool NONNEG>
__global__ void __launch_bounds__(PT_BLOCK, MINB) k_render(const RenderParams p) {
    extern __shared__ float4 s_tris[];
    const SceneDev &sc = p.sc;
    if (ACCEL == PT_ACCEL_BRUTE && SMEM == 1) {
        for (int i = threadIdx.x; i < sc.n_tris; i += PT_BLOCK) {
            s_tris[3 * i] = ldg4(sc.q0 + i);
            s_tris[3 * i + 1] = ldg4(sc.q1 + i);
            s_tris[3 * i + 2] = ldg4(sc.q2 + i);
        }
        __syncthreads();
    }
#if PT_PARITY
    const int n_filt = ((sc.n_tris + 31) >> 5) * 16 * 7;          // float4 of filter records (whole 32-triangle chunks)
    if (ACCEL == PT_ACCEL_BRUTE && SMEM == 4) {      // filter records, then the exact records (brute_smem_filtered)
        const float4 z = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        for (int j = threadIdx.x; j < n_filt / 7; j += PT_BLOCK) {
            const int ia = 2 * j, ib = 2 * j + 1;
            const bool ha = ia < sc.n_tris, hb = ib < sc.n_tris;      // padding pairs are masked off by the scan
            const float4 A0 = ha ? ldg4(sc.fq0 + ia) : z, A1 = ha ? ldg4(sc.fq1 + ia) : z, A2 = ha ? ldg4(sc.fq2 + ia) : z;
            const float4 B0 = hb ? ldg4(sc.fq0 + ib) : z, B1 = hb ? ldg4(sc.fq1 + ib) : z, B2 = hb ? ldg4(sc.fq2 + ib) : z;
            float ca = PT_NO_HIT, cb = PT_NO_HIT, la = 0.0f, lb = 0.0f;
            if (ha) filter_record(ldg4(sc.q0 + ia), ldg4(sc.q1 + ia), ldg4(sc.q2 + ia), ca, la);
            if (hb) filter_record(ldg4(sc.q0 + ib), ldg4(sc.q1 + ib), ldg4(sc.q2 + ib), cb, lb);
            s_tris[7 * j + 0] = make_float4(A0.x, B0.x, A0.y, B0.y);
            s_tris[7 * j + 1] = make_float4(A0.z, B0.z, -A0.w, -B0.w);
            s_tris[7 * j + 2] = make_float4(-A1.x, -B1.x, -A1.y, -B1.y);
            s_tris[7 * j + 3] = make_float4(A1.z, B1.z, A1.w, B1.w);
            s_tris[7 * j + 4] = make_float4(A2.x, B2.x, A2.y, B2.y);
            s_tris[7 * j + 5] = make_float4(A2.z, B2.z, A2.w, B2.w);
            s_tris[7 * j + 6] = make_float4(ca, cb, la, lb);
        }
        for (int i = threadIdx.x; i < sc.n_tris; i += PT_BLOCK) {
            s_tris[n_filt + 3 * i] = ldg4(sc.q0 + i);
            s_tris[n_filt + 3 * i + 1] = ldg4(sc.q1 + i);
            s_tris[n_filt + 3 * i + 2] = ldg4(sc.q2 + i);
        }
        __syncthreads();
    }
#endif
#if !PT_PARITY
    if (ACCEL == PT_ACCEL_BRUTE && SMEM == 2) {      // pair layout for the packed-FP32 scan
        const int n_pairs = (sc.n_tris + 1) >> 1;
        for (int j = threadIdx.x; j < n_pairs; j += PT_BLOCK) {
            const int ia = 2 * j, ib = 2 * j + 1;
            const float4 z = make_float4(0.0f, 0.0f, 0.0f, 0.0f);     // padding triangle: N = 0 -> t = NaN -> never accepted
            const float4 A0 = ldg4(sc.q0 + ia), A1 = ldg4(sc.q1 + ia), A2 = ldg4(sc.q2 + ia);
            const float4 B0 = ib < sc.n_tris ? ldg4(sc.q0 + ib) : z, B1 = ib < sc.n_tris ? ldg4(sc.q1 + ib) : z, B2 = ib < sc.n_tris ? ldg4(sc.q2 + ib) : z;
            s_tris[6 * j + 0] = make_float4(A0.x, B0.x, A0.y, B0.y);          // N.x N.y
            s_tris[6 * j + 1] = make_float4(A0.z, B0.z, -A0.w, -B0.w);        // N.z -r1.x
            s_tris[6 * j + 2] = make_float4(-A1.x, -B1.x, -A1.y, -B1.y);      // -r1.y -r1.z
            s_tris[6 * j + 3] = make_float4(A1.z, B1.z, A1.w, B1.w);          // mu.x mu.y
            s_tris[6 * j + 4] = make_float4(A2.x, B2.x, A2.y, B2.y);          // mu.z mv.x
            s_tris[6 * j + 5] = make_float4(A2.z, B2.z, A2.w, B2.w);          // mv.y mv.z
        }
        __syncthreads();
    }
    const int n_prim_pairs = (sc.n_prims + 1) >> 1;
    if (ACCEL == PT_ACCEL_BRUTE && SMEM == 3) {      // pair layout over the merged primitives + their edge slopes
        for (int j = threadIdx.x; j < n_prim_pairs; j += PT_BLOCK) {
            const int ia = 2 * j, ib = 2 * j + 1;
            const bool hb = ib < sc.n_prims;
            const float4 z = make_float4(0.0f, 0.0f, 0.0f, 0.0f);     // padding primitive: N = 0 -> t = NaN -> never accepted
            const float4 A0 = ldg4(sc.pq0 + ia), A1 = ldg4(sc.pq1 + ia), A2 = ldg4(sc.pq2 + ia);
            const float4 B0 = hb ? ldg4(sc.pq0 + ib) : z, B1 = hb ? ldg4(sc.pq1 + ib) : z, B2 = hb ? ldg4(sc.pq2 + ib) : z;
            const float2 sa = __ldg(sc.prim_ab + ia), sb = hb ? __ldg(sc.prim_ab + ib) : make_float2(1.0f, 0.0f);
            s_tris[7 * j + 0] = make_float4(A0.x, B0.x, A0.y, B0.y);
            s_tris[7 * j + 1] = make_float4(A0.z, B0.z, -A0.w, -B0.w);
            s_tris[7 * j + 2] = make_float4(-A1.x, -B1.x, -A1.y, -B1.y);
            s_tris[7 * j + 3] = make_float4(A1.z, B1.z, A1.w, B1.w);
            s_tris[7 * j + 4] = make_float4(A2.x, B2.x, A2.y, B2.y);
            s_tris[7 * j + 5] = make_float4(A2.z, B2.z, A2.w, B2.w);
            s_tris[7 * j + 6] = make_float4(-sa.x, -sb.x, -sa.y, -sb.y);
        }
        for (int j = threadIdx.x; j < 3 * sc.n_prims; j += PT_BLOCK) s_tris[7 * n_prim_pairs + j] = ldg4(sc.prim_shade + j);
        __syncthreads();
    }
#endif
    // warp → 8x4 pixel tile, block → 16x8
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int x = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const int y = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
    Work wk = {0, 0, 0, 0, 0, 0, 0};
    const int id = y * p.W + x;
    bool mine = x < p.W && y < p.H;
    if (mine && p.shard_n > 1) mine = ((id / p.shard_group) % p.shard_n) == p.shard_rank;
    // No lane leaves before the loop ends: the loop is WARP-CONVERGED.  Every iteration starts with a
    // warp vote, so all 32 lanes walk the intersection code together; a lane without work (pixel outside
    // the image/shard, or all its samples done) rides along predicated off.  (Without the vote the
    // compiler threads the miss -> regenerate path around the loop head and the warp splits for good:
    // measured 13.7 of 32 lanes active, profiles/r01_render_v2.txt.)
    uint32_t rng = mine ? (uint32_t)p.rnds[id] : 1u;
    V3 acc = v3(0.0f, 0.0f, 0.0f);
    if (mine && (p.sample0 != 0 || p.sum_mode)) acc = v3(p.colors[id]);
    const V3 eye = v3(p.cam.eye.x, p.cam.eye.y, p.cam.eye.z);
    const int s_end = p.sample0 + p.spp;
    int s = mine ? p.sample0 : s_end;
    const int camX = (int)p.cam.XM;
    const int cam_x = camX == p.W ? x : id % max(camX, 1), cam_y = camX == p.W ? y : id / max(camX, 1);
    int depth = 0;
    bool alive = false;
    V3 rayP = eye, rayD = v3(0.0f, 0.0f, 1.0f);
    Path ps;
    path_reset(ps);
    for (;;) {
        // path regeneration: the next sample of this pixel starts when at least p.regen_min lanes of the warp are waiting for one (or
        // nobody has a path left) — the camera code then runs at that many lanes instead of at however many paths just ended (9 of 32
        // on Cornell, profiles/r01_render_v7.txt); a waiting lane rides through the scan predicated off meanwhile
        const bool waiting = !alive && s < s_end;
        const int n_wait = __popc(__ballot_sync(0xffffffffu, waiting));
        const bool any_alive = __any_sync(0xffffffffu, alive);     // (both votes by the whole warp, outside the divergent test below)
        if (waiting && (n_wait >= p.regen_min || !any_alive)) {
            if (p.from_rays) {                                   // drop-in trace_ray: gen_ray already ran
                const float4 *r = reinterpret_cast<const float4 *>(p.rays + id);
                rayP = v3(r[0]); rayD = v3(r[1]);
            } else {                                             // gen_ray fused, prog.cl:384-389
                float r1 = pt_rand_t<NONNEG>(rng), r2 = pt_rand_t<NONNEG>(rng);
                camera_ray(p, cam_x, cam_y, r1, r2, rayP, rayD);
            }
            path_reset(ps);
            depth = 0;
            alive = true;
        }
        if (!__any_sync(0xffffffffu, alive)) break;
        const bool trace = alive && depth < p.iterations;        // iterations <= 0: the bounce loop body never runs
        Closest best; best.t = PT_NO_HIT; best.tri = 0x7fffffff;
        const Work wk_before = wk;
        if (ACCEL == PT_ACCEL_BRUTE) {                           // converged scan: idle lanes scan too (free)
#if !PT_PARITY
            if (SMEM == 2) brute_smem_pairs<U>(s_tris, (sc.n_tris + 1) >> 1, rayP, rayD, best); else
            if (SMEM == 3) {
                brute_smem_prims<U, COUNT>(s_tris, n_prim_pairs, rayP, rayD, best, &wk, sc.n_prims);     // best.tri = primitive index
            } else
#endif
#if PT_PARITY
            if (SMEM == 4) brute_smem_filtered(s_tris + n_filt, s_tris, sc.n_tris, rayP, rayD, best, wk); else
#endif
            if (SMEM == 1) brute_smem<COUNT>(s_tris, sc.n_tris, rayP, rayD, best, wk);
            else brute_global<COUNT>(sc, 0, sc.n_tris, rayP, rayD, best, wk);
            if (COUNT && !trace) wk = wk_before;                 // ... but only real rays count as work
        } else if (trace) {
            if (ACCEL == PT_ACCEL_REFBVH) refbvh_intersect<COUNT>(sc, rayP, rayD, best, wk);
            else lbvh_intersect<COUNT>(sc, rayP, rayD, best, wk);
        }
        if (alive) {
            if (trace) wk.rays++;
            if (trace && best.t < PT_NO_HIT) {
                // Hit record of triangle_intersect (prog.cl:107): p = P + V*t, N = tri->N, mati
                V3 hitP = rayP + rayD * best.t;
#if !PT_PARITY
                if (SMEM == 3) {                                 // normal, material and tangent frame of the primitive from shared memory
                    const float4 *sh = s_tris + 7 * n_prim_pairs + 3 * best.tri;
                    const float4 h0 = sh[0];
                    const V3 frame[2] = {v3(sh[1]), v3(sh[2])};
                    Mat m = load_material(sc.mats, __float_as_int(h0.w));
                    shade_hit<NONNEG>(m, hitP, v3(h0), rayP, rayD, rng, ps, eye, p.iterations, frame);
                } else
#endif
                {
                    V3 hitN = tri_normal(ldg4(sc.q0 + best.tri), ldg4(sc.q1 + best.tri), ldg4(sc.q2 + best.tri));
                    Mat m = load_material(sc.mats, (int)__ldg(sc.mati + best.tri));
                    shade_hit<NONNEG>(m, hitP, hitN, rayP, rayD, rng, ps, eye, p.iterations);
                }
                if (++depth >= p.iterations) alive = false;
            } else {
                alive = false;                                   // miss: break, prog.cl:367-376
            }
            if (!alive) {                                        // sample finished, prog.cl:379
                if (p.sum_mode) acc = acc + ps.color;
                else acc = pt_mean_step(acc, ps.color, s);
                ++s;
                wk.samples++;
            }
        }
    }
    if (mine) {
        p.rnds[id] = (int)rng;
        p.colors[id] = make_float4(acc.x, acc.y, acc.z, 0.0f);
        if (p.tex) p.tex[id] = reinhard_tone_map(acc);           // prog.cl:380
    }
    // work counters: one atomic per warp and counter
    if (p.counters) {
        unsigned v[7] = {wk.rays, wk.box, wk.tri, wk.tri1, wk.tri2, wk.tri3, wk.samples};
        const int idx[7] = {PT_CNT_RAYS, PT_CNT_BOX, PT_CNT_TRI, PT_CNT_TRI1, PT_CNT_TRI2, PT_CNT_TRI3, PT_CNT_PIXEL_SAMPLES};
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            if (!COUNT && k >= 1 && k <= 5) continue;
            unsigned long long tot = v[k];
            for (int o = 16; o; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
            if (lane == 0 && tot) atomicAdd(p.counters + idx[k], tot);
        }
    }
}

#include "pt_sched.cuh"
#include "pt_wavefront.cuh"
#include "pt_pool.cuh"

// gen_ray as its own launch (prog.cl:384-389): only for hosts that drive the reference's two-kernel sequence
__global__ void __launch_bounds__(256) k_gen_ray(const RenderParams p, size_t n) {
    size_t id = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n) return;
    if (p.shard_n > 1 && (int)((id / (size_t)p.shard_group) % (size_t)p.shard_n) != p.shard_rank) return;
    uint32_t rng = (uint32_t)p.rnds[id];
    float r1 = pt_rand(rng), r2 = pt_rand(rng);
    V3 P, D;
    const int camX = max((int)p.cam.XM, 1);
    camera_ray(p, (int)id % camX, (int)id / camX, r1, r2, P, D);
    p.rnds[id] = (int)rng;
    float4 *r = reinterpret_cast<float4 *>(p.rays + id);
    r[0] = make_float4(P.x, P.y, P.z, 0.0f);
    r[1] = make_float4(D.x, D.y, D.z, 0.0f);
}

// ---------------------------------------------------------------- closest hit only (ptb_closest_hit)
// One thread per ray through the same closest-hit code the path kernels call: the test surface for
// "LBVH == brute force on arbitrary rays, ties included" (SURVEY §7 step 5).  out = (t, bits(triangle)), (-1, -1) on a miss.
template <int ACCEL>
__global__ void __launch_bounds__(128) k_closest_hit(const RenderParams p, const RefRay *rays, size_t n, float2 *out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 *r = reinterpret_cast<const float4 *>(rays + i);
    const V3 P = v3(r[0]), D = v3(r[1]);
    Closest best; best.t = PT_NO_HIT; best.tri = 0x7fffffff;
    Work wk = {0, 0, 0, 0, 0, 0, 0};
    if (ACCEL == PT_ACCEL_BRUTE) brute_global<false>(p.sc, 0, p.sc.n_tris, P, D, best, wk);
    else if (ACCEL == PT_ACCEL_REFBVH) refbvh_intersect<false>(p.sc, P, D, best, wk);
    else lbvh_intersect<false>(p.sc, P, D, best, wk);
    const bool hit = best.t < PT_NO_HIT;
    out[i] = make_float2(hit ? best.t : -1.0f, __int_as_float(hit ? best.tri : -1));
}
// the wavefront trace kernel on an arbitrary ray list: rays -> its ray records + an identity queue, and back
__global__ void k_ch_pack(const RenderParams p, const RefRay *rays, size_t n, float4 *ray0, float4 *ray1, int *queue, int *ctrl) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { ctrl[0] = (int)n; ctrl[1] = 0; ctrl[2] = 0; ctrl[3] = 0; }
    if (i >= n) return;
    const float4 *r = reinterpret_cast<const float4 *>(rays + i);
    const float4 a = r[0], b = r[1];
    Closest best; best.t = PT_NO_HIT; best.tri = 0x7fffffff;
    Work wk = {0, 0, 0, 0, 0, 0, 0};
    lbvh_big_scan<false>(p.sc, v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), best, wk);     // the outlier list, as the path kernels do where a ray is created
    ray0[i] = make_float4(a.x, a.y, a.z, b.x);
    ray1[i] = make_float4(b.y, b.z, best.t, __int_as_float(best.tri));
    queue[i] = (int)i;
}
__global__ void k_ch_unpack(const float4 *ray1, size_t n, float2 *out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 b = ray1[i];
    const bool hit = b.z < PT_NO_HIT;
    out[i] = make_float2(hit ? b.z : -1.0f, hit ? b.w : __int_as_float(-1));
}

// ---------------------------------------------------------------- host-side dispatch
template <int ACCEL, bool COUNT, int SMEM, int MINB = 1, int U = 2, bool NONNEG = false>
static void launch_one(const RenderParams &p, cudaStream_t st) {
    dim3 grid((p.W + 15) / 16, (p.H + 7) / 8);
    size_t smem = 0;
    if (ACCEL == PT_ACCEL_BRUTE && SMEM == 1) smem = (size_t)p.sc.n_tris * 3 * sizeof(float4);
    if (ACCEL == PT_ACCEL_BRUTE && SMEM == 2) smem = (size_t)((p.sc.n_tris + 1) / 2) * 6 * sizeof(float4);
    if (ACCEL == PT_ACCEL_BRUTE && SMEM == 3) smem = ((size_t)((p.sc.n_prims + 1) / 2) * 7 + 3 * (size_t)p.sc.n_prims) * sizeof(float4);
    if (ACCEL == PT_ACCEL_BRUTE && SMEM == 4) smem = ((size_t)((p.sc.n_tris + 31) / 32) * 16 * 7 + 3 * (size_t)p.sc.n_tris) * sizeof(float4);
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(k_render<ACCEL, COUNT, SMEM, MINB, U, NONNEG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_render<ACCEL, COUNT, SMEM, MINB, U, NONNEG><<<grid, PT_BLOCK, smem, st>>>(p);
}

void launch_render(const RenderParams &p, cudaStream_t st) {
    const bool c = p.count_work != 0;
    if (p.accel == PT_ACCEL_BRUTE) {
        const bool smem = p.sc.n_tris <= PT_SMEM_MAX_TRIS;
        if (smem) {
#if !PT_PARITY
            if (c && p.sc.n_prims > 0) launch_one<PT_ACCEL_BRUTE, true, 3, 1, 2>(p, st);     // the counting build of the SHIPPING scan (merged quads)
            else
#endif
            if (c) launch_one<PT_ACCEL_BRUTE, true, 1>(p, st);
#if !PT_PARITY
            else if (p.sc.n_prims > 0) {                              // same scan over merged convex quads (prims.cu)
                static const bool verbose = getenv("PTB_VERBOSE") != nullptr;
                if (verbose) fprintf(stderr, "ptb: k_render scan over %d primitives, rng_nonneg=%d\n", p.sc.n_prims, p.rng_nonneg);
                if (p.rng_nonneg) launch_one<PT_ACCEL_BRUTE, false, 3, 8, 4, true>(p, st);
                else launch_one<PT_ACCEL_BRUTE, false, 3, 8, 4>(p, st);
            } else if (p.rng_nonneg) launch_one<PT_ACCEL_BRUTE, false, 2, 8, 4, true>(p, st);
            else launch_one<PT_ACCEL_BRUTE, false, 2, 8, 4>(p, st);   // packed-FP32 pair scan; 64 registers -> 8 blocks/SM, 4 pairs in flight
                                                                      // (measured best of 8 (occupancy, unroll) variants: 23.0 vs 20.7 Grays/s)
#else
            else if (p.parity_filter && p.sc.n_tris <= PT_FILTER_MAX_TRIS) launch_one<PT_ACCEL_BRUTE, false, 4, 4>(p, st);
            else launch_one<PT_ACCEL_BRUTE, false, 1>(p, st);
#endif
        } else {
            if (c) launch_one<PT_ACCEL_BRUTE, true, 0>(p, st); else launch_one<PT_ACCEL_BRUTE, false, 0>(p, st);
        }
    } else if (p.accel == PT_ACCEL_REFBVH) {
        if (c) launch_one<PT_ACCEL_REFBVH, true, 0>(p, st); else launch_one<PT_ACCEL_REFBVH, false, 0>(p, st);
    } else {
        static const int sched = getenv("PTB_LBVH_SCHED") ? atoi(getenv("PTB_LBVH_SCHED")) : 1;
        dim3 grid((p.W + 15) / 16, (p.H + 7) / 8);
        if (sched == 0) { if (c) launch_one<PT_ACCEL_LBVH, true, 0>(p, st); else launch_one<PT_ACCEL_LBVH, false, 0>(p, st); }
        else if (c) k_render_sched<true, 1><<<grid, PT_BLOCK, 0, st>>>(p);
        else if (sched == 2) k_render_sched<false, 6><<<grid, PT_BLOCK, 0, st>>>(p);
        else if (sched == 3) k_render_sched<false, 8><<<grid, PT_BLOCK, 0, st>>>(p);
        else k_render_sched<false, 1><<<grid, PT_BLOCK, 0, st>>>(p);
    }
}

// scratch (wavefront only): n * 36 + 16 bytes
int launch_closest_hit(const RenderParams &p0, const RefRay *rays, size_t n, float2 *out, bool wavefront, void *scratch, cudaStream_t st) {
    if (n == 0) return 0;
    const unsigned blocks = (unsigned)((n + 127) / 128);
    if (wavefront && p0.accel == PT_ACCEL_LBVH && scratch) {
        RenderParams p = p0;
        char *b = (char *)scratch;
        p.wf_ray0 = (float4 *)b; b += n * 16;
        p.wf_ray1 = (float4 *)b; b += n * 16;
        p.wf_queue[0] = (int *)b; b += n * 4;
        p.wf_ctrl = (int *)b;
        k_ch_pack<<<blocks, 128, 0, st>>>(p, rays, n, p.wf_ray0, p.wf_ray1, p.wf_queue[0], p.wf_ctrl);
        int sms = 148, dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        const int wide = wf_wide_mode(p.sc);
        const int3 tune = make_int3(WF_FETCH_MIN, 4, PT_SCHED_NODE_STEPS);
        if (wide == 2) k_wf_trace<false, 2><<<sms * 8, PT_BLOCK, 0, st>>>(p, p.wf_queue[0], p.wf_ctrl, p.wf_ctrl + 2, tune);
        else if (wide == 1) k_wf_trace<false, 1><<<sms * 8, PT_BLOCK, 0, st>>>(p, p.wf_queue[0], p.wf_ctrl, p.wf_ctrl + 2, tune);
        else k_wf_trace<false, 0><<<sms * 8, PT_BLOCK, 0, st>>>(p, p.wf_queue[0], p.wf_ctrl, p.wf_ctrl + 2, tune);
        k_ch_unpack<<<blocks, 128, 0, st>>>(p.wf_ray1, n, out);
        return 3;
    }
    if (p0.accel == PT_ACCEL_BRUTE) k_closest_hit<PT_ACCEL_BRUTE><<<blocks, 128, 0, st>>>(p0, rays, n, out);
    else if (p0.accel == PT_ACCEL_REFBVH) k_closest_hit<PT_ACCEL_REFBVH><<<blocks, 128, 0, st>>>(p0, rays, n, out);
    else k_closest_hit<PT_ACCEL_LBVH><<<blocks, 128, 0, st>>>(p0, rays, n, out);
    return 1;
}

void launch_gen_ray(const RenderParams &p, size_t n, cudaStream_t st) {
    if (n == 0) return;
    k_gen_ray<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(p, n);
}

}  // namespace PT_NS
