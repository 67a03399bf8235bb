// prims.cu — primitive merging for the shared-memory brute-force scan (fast flavour only).
//
// Meshes authored with quads reach the reference as fan-triangulated pairs (tiny_obj_loader
// triangulates, main.cpp:558).  Two triangles that share an edge, lie in one plane, carry the same
// material and together form a CONVEX QUAD can be tested as one primitive at the cost of one triangle.
// Anchor the barycentric frame at z, the corner of triangle A opposite the shared edge x–y:
// u runs along y-z, v along x-z, so y = (1,0), x = (0,1) and B's far corner d = (du,dv).  Then
//     triangle A   : u >= 0, v >= 0, 1 - u - v >= 0
//     convex quad  : u >= 0, v >= 0, 1 - u - b*v >= 0 (edge y–d), 1 - v - a*u >= 0 (edge x–d)
// with b = (1-du)/dv and a = (1-dv)/du; a lone triangle is the same test with b = 1, a = 0 (the fourth
// bound is then implied).  Two scalars per primitive select the shape:
//     inside  <=>  min(u, v, 1 - u - b*v, 1 - v - a*u) >= 0.
// The pair's normals agree and so do their materials, so shading never needs to know which half was
// hit: the lower upload index stands for both.  The Cornell box (32 triangles = 16 quads, none of the
// walls an exact parallelogram) halves its scan this way.  Only exact (bitwise) shared vertices are
// matched; coplanarity / normal agreement use a 1e-5 relative tolerance, inside the stated tolerance
// of the fast flavour.  The parity flavour never uses primitives.
#include <cuda_runtime.h>
#include <stdint.h>

#include "prims.h"

namespace {

struct D3 { double x, y, z; };
__device__ D3 d3(float4 q) { D3 r = {q.x, q.y, q.z}; return r; }
__device__ D3 sub(D3 a, D3 b) { D3 r = {a.x - b.x, a.y - b.y, a.z - b.z}; return r; }
__device__ double dotd(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ D3 crossd(D3 a, D3 b) { D3 r = {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; return r; }
__device__ bool same(float4 a, float4 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }

// vertices of triangle i from the parity layout (r1,N.x) (r2,N.y) (r3,N.z)
__device__ void load_tri(const float4 *q0, const float4 *q1, const float4 *q2, int i, float4 v[3], D3 &N) {
    v[0] = q0[i]; v[1] = q1[i]; v[2] = q2[i];
    N.x = v[0].w; N.y = v[1].w; N.z = v[2].w;
}

struct Quad { int rot; double b, a; };   // rot: index of A's vertex opposite the shared edge

// Can triangle j complete triangle i to a convex planar quad?  (rot < 0: no)
__device__ Quad match(const float4 *q0, const float4 *q1, const float4 *q2, const uint16_t *mati, const int *idx, int i, int j) {
    Quad no = {-1, 1.0, 0.0};
    if (mati[idx ? idx[i] : i] != mati[idx ? idx[j] : j]) return no;
    float4 a[3], b[3];
    D3 Na, Nb;
    load_tri(q0, q1, q2, i, a, Na);
    load_tri(q0, q1, q2, j, b, Nb);
    const D3 cr = crossd(Na, Nb);
    if (!(dotd(Na, Nb) > 0.0) || !(dotd(cr, cr) <= 1e-10)) return no;     // normals within 1e-5 rad
    for (int r = 0; r < 3; ++r) {
        const float4 z = a[r], y = a[(r + 1) % 3], x = a[(r + 2) % 3];
        int shared = 0, other = -1;
        for (int k = 0; k < 3; ++k) {
            if (same(b[k], x) || same(b[k], y)) shared++; else other = k;
        }
        if (shared != 2 || other < 0 || same(b[other], z)) continue;
        const D3 e1 = sub(d3(y), d3(z)), e2 = sub(d3(x), d3(z)), w = sub(d3(b[other]), d3(z));
        const double den = dotd(Na, crossd(e1, e2));
        if (den == 0.0) continue;
        const double scale = sqrt(fmax(dotd(e1, e1), dotd(e2, e2))) + 1e-30;
        if (fabs(dotd(w, Na)) > 1e-5 * scale) continue;                     // far corner off the plane
        const double du = dotd(crossd(e2, Na), w) / den, dv = dotd(crossd(Na, e1), w) / den;
        if (!(du > 1e-6 && dv > 1e-6 && du + dv > 1.0 + 1e-6)) continue;    // not strictly convex
        Quad q = {r, (1.0 - du) / dv, (1.0 - dv) / du};
        return q;
    }
    return no;
}

__device__ void emit(const float4 *q0, const float4 *q1, const float4 *q2, int i, Quad q,
                     float4 *p0, float4 *p1, float4 *p2, float2 *ab, int *rep, float4 *shade, const uint16_t *mati, const int *idx, int slot) {
    float4 a[3];
    D3 N;
    load_tri(q0, q1, q2, i, a, N);
    const int rot = q.rot < 0 ? 0 : q.rot;
    const D3 r1 = d3(a[rot]), r2 = d3(a[(rot + 1) % 3]), r3 = d3(a[(rot + 2) % 3]);
    const D3 e1 = sub(r2, r1), e2 = sub(r3, r1);
    const double inv = 1.0 / dotd(N, crossd(e1, e2));
    const D3 mu = crossd(e2, N), mv = crossd(N, e1);
    p0[slot] = make_float4((float)N.x, (float)N.y, (float)N.z, (float)r1.x);
    p1[slot] = make_float4((float)r1.y, (float)r1.z, (float)(mu.x * inv), (float)(mu.y * inv));
    p2[slot] = make_float4((float)(mu.z * inv), (float)(mv.x * inv), (float)(mv.y * inv), (float)(mv.z * inv));
    ab[slot] = make_float2((float)q.b, (float)q.a);
    rep[slot] = idx ? idx[i] : i;
    // shading record: stored normal + material index, and the tangent frame orthonormal_base (prog.cl:186-204) derives
    // from the UNFLIPPED normal: v2 = (0,-N.z,N.y)/|..| when |N.x|,|N.z| <= 1e-3 else (-N.z,0,N.x)/|..|, v3 = N x v2.
    // (For the flipped normal -N the kernel negates v2 and keeps v3: both are linear/bilinear in N.)
    const float nx = (float)N.x, ny = (float)N.y, nz = (float)N.z;
    D3 v2;
    if (fabsf(nx) <= 0.001f && fabsf(nz) <= 0.001f) {
        const double rl = 1.0 / sqrt(N.y * N.y + N.z * N.z);
        v2.x = 0.0; v2.y = -N.z * rl; v2.z = N.y * rl;
    } else {
        const double rl = 1.0 / sqrt(N.x * N.x + N.z * N.z);
        v2.x = -N.z * rl; v2.y = 0.0; v2.z = N.x * rl;
    }
    const D3 v3 = crossd(N, v2);
    shade[3 * slot + 0] = make_float4(nx, ny, nz, __int_as_float((int)mati[idx ? idx[i] : i]));
    shade[3 * slot + 1] = make_float4((float)v3.x, (float)v3.y, (float)v3.z, 0.0f);     // X
    shade[3 * slot + 2] = make_float4((float)v2.x, (float)v2.y, (float)v2.z, 0.0f);     // Z
}

// one block; n <= PRIMS_MAX_TRIS.  Greedy in upload order: triangle i takes the first unpaired j > i that matches.
__global__ void k_build_prims(const float4 *q0, const float4 *q1, const float4 *q2, const uint16_t *mati, const int *idx, int n,
                              float4 *p0, float4 *p1, float4 *p2, float2 *ab, int *rep, float4 *shade, int *n_prims) {
    __shared__ unsigned char used[PRIMS_MAX_TRIS];
    for (int i = threadIdx.x; i < n; i += blockDim.x) used[i] = 0;
    __syncthreads();
    if (threadIdx.x == 0) {
        int slot = 0;
        for (int i = 0; i < n; ++i) {
            if (used[i]) continue;
            Quad q = {-1, 1.0, 0.0};
            int partner = -1;
            for (int j = i + 1; j < n && partner < 0; ++j) {
                if (used[j]) continue;
                const Quad m = match(q0, q1, q2, mati, idx, i, j);
                if (m.rot >= 0) { partner = j; q = m; }
            }
            used[i] = 1;
            if (partner >= 0) used[partner] = 1;
            emit(q0, q1, q2, i, q, p0, p1, p2, ab, rep, shade, mati, idx, slot++);
        }
        *n_prims = slot;
    }
}

}  // namespace

void prims_build(const float4 *q0, const float4 *q1, const float4 *q2, const uint16_t *mati, const int *idx, int n, float4 *p0,
                 float4 *p1, float4 *p2, float2 *ab, int *rep, float4 *shade, int *n_prims_dev, cudaStream_t st) {
    k_build_prims<<<1, 64, 0, st>>>(q0, q1, q2, mati, idx, n, p0, p1, p2, ab, rep, shade, n_prims_dev);
}
