// prims.cu — primitive merging for the shared-memory brute-force scan (fast flavour only).
//
// Meshes authored with quads reach the reference as fan-triangulated pairs (tiny_obj_loader
// triangulates, main.cpp:558).  Two triangles that share an edge, lie in one plane and together
// form a PARALLELOGRAM can be tested as one primitive at the cost of one triangle:
// with the shared edge as the diagonal a1–a3 of triangle A = (a1, a2, a3) and A's barycentric
// coordinates (u along a2-a1, v along a3-a1) of the plane hit point,
//     A      : u >= 0, v >= 0, u+v <= 1
//     B      : u <= 0, u+v >= 0, v <= 1            (B's far corner is a1 + a3 - a2 = (u,v) = (-1,1))
//     A or B : 0 <= u+v <= 1  and  0 <= v <= 1
// Only pairs with the same material are merged, and their normals agree, so the shading never needs to
// know which half was hit: the lower upload index stands for both.
// A lone triangle is the same test with the u+v >= 0 bound replaced by u >= 0, so one scalar
// f in {0,1} per primitive selects between them:  k = min(v, 1-v, 1-(u+v), u + f*v) >= 0.
// The Cornell box (32 triangles = 16 quads) halves its scan this way.  Only exact (bitwise) shared
// vertices are matched; the plane / parallelogram conditions use a 1e-5 relative tolerance, which
// is inside the stated tolerance of the fast flavour.  The parity flavour never uses primitives.
#include <cuda_runtime.h>
#include <stdint.h>

#include "prims.h"

namespace {

struct D3 { double x, y, z; };
__device__ D3 d3(float4 q) { D3 r = {q.x, q.y, q.z}; return r; }
__device__ D3 sub(D3 a, D3 b) { D3 r = {a.x - b.x, a.y - b.y, a.z - b.z}; return r; }
__device__ D3 add(D3 a, D3 b) { D3 r = {a.x + b.x, a.y + b.y, a.z + b.z}; return r; }
__device__ double dotd(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ D3 crossd(D3 a, D3 b) { D3 r = {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; return r; }
__device__ bool same(float4 a, float4 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }

// vertices of triangle i from the parity layout (r1,N.x) (r2,N.y) (r3,N.z)
__device__ void load_tri(const float4 *q0, const float4 *q1, const float4 *q2, int i, float4 v[3], D3 &N) {
    v[0] = q0[i]; v[1] = q1[i]; v[2] = q2[i];
    N.x = v[0].w; N.y = v[1].w; N.z = v[2].w;
}

// If triangle j completes triangle i to a parallelogram, returns the rotation r of i's vertices that puts the
// shared edge on (first, third); otherwise -1.
__device__ int match(const float4 *q0, const float4 *q1, const float4 *q2, const uint16_t *mati, int i, int j) {
    if (mati[i] != mati[j]) return -1;
    float4 a[3], b[3];
    D3 Na, Nb;
    load_tri(q0, q1, q2, i, a, Na);
    load_tri(q0, q1, q2, j, b, Nb);
    if (!(dotd(Na, Nb) > 1.0 - 1e-9)) return -1;                 // same plane orientation (unit normals)
    for (int r = 0; r < 3; ++r) {
        const float4 x = a[r], z = a[(r + 1) % 3], y = a[(r + 2) % 3];   // rotated A = (x, z, y): diagonal x–y
        int shared = 0, other = -1;
        for (int k = 0; k < 3; ++k) {
            if (same(b[k], x) || same(b[k], y)) shared++; else other = k;
        }
        if (shared != 2 || other < 0) continue;
        const D3 want = sub(add(d3(x), d3(y)), d3(z));            // x + y - z
        const D3 got = d3(b[other]);
        const D3 diff = sub(got, want);
        const D3 e = sub(d3(y), d3(x));
        const double scale = sqrt(dotd(e, e)) + 1e-30;
        if (sqrt(dotd(diff, diff)) > 1e-5 * scale) continue;     // not a parallelogram
        if (fabs(dotd(sub(got, d3(x)), Na)) > 1e-5 * scale) continue;   // not coplanar
        return r;
    }
    return -1;
}

__device__ void emit(const float4 *q0, const float4 *q1, const float4 *q2, int i, int rot, int partner,
                     float4 *p0, float4 *p1, float4 *p2, float *f, int *rep, int slot) {
    float4 a[3];
    D3 N;
    load_tri(q0, q1, q2, i, a, N);
    const D3 r1 = d3(a[rot]), r2 = d3(a[(rot + 1) % 3]), r3 = d3(a[(rot + 2) % 3]);
    const D3 e1 = sub(r2, r1), e2 = sub(r3, r1);
    const double inv = 1.0 / dotd(N, crossd(e1, e2));
    const D3 mu = crossd(e2, N), mv = crossd(N, e1);
    p0[slot] = make_float4((float)N.x, (float)N.y, (float)N.z, (float)r1.x);
    p1[slot] = make_float4((float)r1.y, (float)r1.z, (float)(mu.x * inv), (float)(mu.y * inv));
    p2[slot] = make_float4((float)(mu.z * inv), (float)(mv.x * inv), (float)(mv.y * inv), (float)(mv.z * inv));
    rep[slot] = i;
    f[slot] = partner >= 0 ? 1.0f : 0.0f;
}

// one block; n <= PRIMS_MAX_TRIS.  Greedy in upload order: triangle i takes the first unpaired j > i that matches.
__global__ void k_build_prims(const float4 *q0, const float4 *q1, const float4 *q2, const uint16_t *mati, int n,
                              float4 *p0, float4 *p1, float4 *p2, float *f, int *rep, int *n_prims) {
    __shared__ unsigned char used[PRIMS_MAX_TRIS];
    for (int i = threadIdx.x; i < n; i += blockDim.x) used[i] = 0;
    __syncthreads();
    if (threadIdx.x == 0) {
        int slot = 0;
        for (int i = 0; i < n; ++i) {
            if (used[i]) continue;
            int partner = -1, rot = 0;
            for (int j = i + 1; j < n && partner < 0; ++j) {
                if (used[j]) continue;
                const int r = match(q0, q1, q2, mati, i, j);
                if (r >= 0) { partner = j; rot = r; }
            }
            used[i] = 1;
            if (partner >= 0) used[partner] = 1;
            emit(q0, q1, q2, i, rot, partner, p0, p1, p2, f, rep, slot++);
        }
        *n_prims = slot;
    }
}

}  // namespace

void prims_build(const float4 *q0, const float4 *q1, const float4 *q2, const uint16_t *mati, int n, float4 *p0, float4 *p1,
                 float4 *p2, float *f, int *rep, int *n_prims_dev, cudaStream_t st) {
    k_build_prims<<<1, 64, 0, st>>>(q0, q1, q2, mati, n, p0, p1, p2, f, rep, n_prims_dev);
}
