// lbvh.cu — GPU construction of the closest-hit structure used by the LBVH traversal in
// pt_kernels.cuh.  New to this implementation: the reference builds one median-split tree per OBJ
// shape on the host (main.cpp:195-304, 536-551) and walks them one after the other
// (prog.cl:144-184); here ONE tree over all uploaded triangles is built on the device:
//
//   1. k_bounds     centroid box + largest |coordinate| (block reduce + ordered-int atomics)
//   2. k_morton     63-bit Morton code of the triangle's box centre (21 bits per axis).  OUTLIERS: a triangle whose box
//                   spans more than a quarter of the scene (room walls, emitter quads) would inflate every node above
//                   it; up to LBVH_BIG_MAX of them get the key 2^63, sort behind everything else and stay OUT of the
//                   tree — the traversal kernels scan that short list once per ray before they enter the tree
//   3. cub radix sort of (code, upload index) — stable, so equal codes keep upload order
//   4. k_gather     triangles re-laid in Morton order (3 float4 planes) + their boxes
//   5. k_hierarchy  Karras 2012 radix tree over the sorted codes (ties split on the index bits)
//   6. k_refit      bottom-up boxes, second arrival at a node (atomic counter) continues upward
//   7. k_emit       64-byte traversal nodes holding BOTH children's (inflated) boxes and links;
//                   a subtree with <= leaf_size triangles becomes one leaf [first, first+count)
//   8. k_depth      depth of the emitted tree (the traversal stack is bounded; checked by the host)
//   9. k_flag4 / scan / k_emit4   the same tree collapsed to 4-ary nodes for the wavefront traversal: every
//                   emitted binary node at EVEN depth becomes a wide node whose slots are its grandchildren
//                   (a child that is a leaf keeps its slot), compacted by a prefix sum.  Halves the chain of
//                   dependent node fetches, which is what bounds the traversal (profiles/r01_wf_trace_v2.txt).
//
// Node layout (4 x float4, Aila–Laine style):
//   n0 = (L.lo.x, L.hi.x, L.lo.y, L.hi.y)  n1 = (R.lo.x, R.hi.x, R.lo.y, R.hi.y)
//   n2 = (L.lo.z, L.hi.z, R.lo.z, R.hi.z)  n3 = (bits(left link), bits(right link), 0, 0)
// link >= 0: node index.  link < 0: leaf, ~link = (first << 3) | (count-1) in Morton order.
// Wide node layout (8 x float4 = one 128-byte line): lo.x[4] hi.x[4] lo.y[4] hi.y[4] lo.z[4] hi.z[4] links[4] pad;
// an empty slot has an inverted infinite box (lo = +inf, hi = -inf), which no slab test accepts.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "lbvh.h"

namespace {

__device__ __forceinline__ unsigned f2ord(float f) {
    unsigned b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ __forceinline__ float ord2f(unsigned u) {
    unsigned b = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
    return __uint_as_float(b);
#else
    float f; memcpy(&f, &b, 4); return f;
#endif
}

struct Box {
    float lo[3], hi[3];
};
__device__ __forceinline__ Box tri_box(float4 a, float4 b, float4 c) {
    Box r;
    r.lo[0] = fminf(a.x, fminf(b.x, c.x)); r.hi[0] = fmaxf(a.x, fmaxf(b.x, c.x));
    r.lo[1] = fminf(a.y, fminf(b.y, c.y)); r.hi[1] = fmaxf(a.y, fmaxf(b.y, c.y));
    r.lo[2] = fminf(a.z, fminf(b.z, c.z)); r.hi[2] = fmaxf(a.z, fmaxf(b.z, c.z));
    return r;
}

// bounds[0..2] = min centre, [3..5] = max centre (ordered-uint), [6] = max |coordinate|
__global__ void k_bounds(const float4 *q0, const float4 *q1, const float4 *q2, int n, unsigned *bounds) {
    float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f}, mabs = 0.0f;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        Box b = tri_box(q0[i], q1[i], q2[i]);
        for (int k = 0; k < 3; ++k) {
            float c = 0.5f * (b.lo[k] + b.hi[k]);
            if (c == c && fabsf(c) < 1e30f) { lo[k] = fminf(lo[k], c); hi[k] = fmaxf(hi[k], c); }
            float m = fmaxf(fabsf(b.lo[k]), fabsf(b.hi[k]));
            if (m < 1e30f) mabs = fmaxf(mabs, m);
        }
    }
    for (int o = 16; o; o >>= 1) {
        for (int k = 0; k < 3; ++k) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
        mabs = fmaxf(mabs, __shfl_xor_sync(0xffffffffu, mabs, o));
    }
    if ((threadIdx.x & 31) == 0) {
        for (int k = 0; k < 3; ++k) {
            atomicMin(bounds + k, f2ord(lo[k]));
            atomicMax(bounds + 3 + k, f2ord(hi[k]));
        }
        atomicMax(bounds + 6, f2ord(mabs));
    }
}

__device__ __forceinline__ unsigned long long spread21(unsigned v) {
    unsigned long long x = v & 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void k_morton(const float4 *q0, const float4 *q1, const float4 *q2, int n, const unsigned *bounds,
                         unsigned long long *keys, int *vals, float big_frac, int *n_big) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Box b = tri_box(q0[i], q1[i], q2[i]);
    if (big_frac > 0.0f) {
        float scene = 0.0f, mine = 0.0f;
        for (int k = 0; k < 3; ++k) {
            scene = fmaxf(scene, ord2f(bounds[3 + k]) - ord2f(bounds[k]));
            mine = fmaxf(mine, b.hi[k] - b.lo[k]);
        }
        if (mine > big_frac * scene && scene > 0.0f) {
            keys[i] = 0x8000000000000000ull;
            vals[i] = i;
            atomicAdd(n_big, 1);
            return;
        }
    }
    unsigned g[3];
    for (int k = 0; k < 3; ++k) {
        float lo = ord2f(bounds[k]), hi = ord2f(bounds[3 + k]);
        float c = 0.5f * (b.lo[k] + b.hi[k]);
        float ext = hi - lo;
        float u = ext > 0.0f ? (c - lo) / ext : 0.0f;
        if (!(u >= 0.0f)) u = 0.0f;         // NaN / below → 0
        if (u > 1.0f) u = 1.0f;
        double s = (double)u * 2097151.0;    // 2^21 - 1
        g[k] = (unsigned)s;
    }
    keys[i] = (spread21(g[0]) << 2) | (spread21(g[1]) << 1) | spread21(g[2]);
    vals[i] = i;
}

__global__ void k_gather(const float4 *q0, const float4 *q1, const float4 *q2, const float4 *f0, const float4 *f1, const float4 *f2,
                         const int *vals, int n, float4 *sq0, float4 *sq1, float4 *sq2, float4 *sf0, float4 *sf1, float4 *sf2,
                         float4 *leaf_lo, float4 *leaf_hi, float4 *rq, float4 *rf) {
    int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= n) return;
    int i = vals[pos];
    float4 a = q0[i], b = q1[i], c = q2[i];
    sq0[pos] = a; sq1[pos] = b; sq2[pos] = c;
    const float4 fa = f0[i], fb = f1[i], fc = f2[i];
    sf0[pos] = fa; sf1[pos] = fb; sf2[pos] = fc;
    const float4 idrow = make_float4(__int_as_float(i), 0.0f, 0.0f, 0.0f);
    rq[4 * (size_t)pos] = a; rq[4 * (size_t)pos + 1] = b; rq[4 * (size_t)pos + 2] = c; rq[4 * (size_t)pos + 3] = idrow;
    rf[4 * (size_t)pos] = fa; rf[4 * (size_t)pos + 1] = fb; rf[4 * (size_t)pos + 2] = fc; rf[4 * (size_t)pos + 3] = idrow;
    Box bx = tri_box(a, b, c);
    leaf_lo[pos] = make_float4(bx.lo[0], bx.lo[1], bx.lo[2], 0.0f);
    leaf_hi[pos] = make_float4(bx.hi[0], bx.hi[1], bx.hi[2], 0.0f);
}

// common-prefix length of sorted keys i and j; equal codes fall back to the position bits (Karras 2012 §4)
__device__ __forceinline__ int delta(const unsigned long long *keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    unsigned long long a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz((unsigned)i ^ (unsigned)j);
    return __clzll((long long)(a ^ b));
}

// internal node i ∈ [0, n-2]: range, split, children (child >= 0 internal, < 0 → ~leaf position), parents
__global__ void k_hierarchy(const unsigned long long *keys, int n, int2 *range, int2 *child, int *parent_of_node, int *parent_of_leaf) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(keys, n, i, j);
    int s = 0;
    int t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + min(d, 0);
    int first = min(i, j), last = max(i, j);
    int left = (first == gamma) ? ~gamma : gamma;
    int right = (last == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    range[i] = make_int2(first, last);
    child[i] = make_int2(left, right);
    if (left >= 0) parent_of_node[left] = i; else parent_of_leaf[~left] = i;
    if (right >= 0) parent_of_node[right] = i; else parent_of_leaf[~right] = i;
    if (i == 0) parent_of_node[0] = -1;
}

__global__ void k_refit(int n, const int2 *child, const int *parent_of_node, const int *parent_of_leaf,
                        const float4 *leaf_lo, const float4 *leaf_hi, float4 *node_lo, float4 *node_hi, int *arrivals) {
    int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= n) return;
    int node = parent_of_leaf[pos];
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(arrivals + node, 1) == 0) return;   // the sibling subtree is not finished yet
        __threadfence();
        int2 c = child[node];
        // volatile reads: the sibling's box was written by another thread in this launch
        const volatile float4 *llo = c.x >= 0 ? node_lo + c.x : leaf_lo + ~c.x;
        const volatile float4 *lhi = c.x >= 0 ? node_hi + c.x : leaf_hi + ~c.x;
        const volatile float4 *rlo = c.y >= 0 ? node_lo + c.y : leaf_lo + ~c.y;
        const volatile float4 *rhi = c.y >= 0 ? node_hi + c.y : leaf_hi + ~c.y;
        float4 lo = make_float4(fminf(llo->x, rlo->x), fminf(llo->y, rlo->y), fminf(llo->z, rlo->z), 0.0f);
        float4 hi = make_float4(fmaxf(lhi->x, rhi->x), fmaxf(lhi->y, rhi->y), fmaxf(lhi->z, rhi->z), 0.0f);
        volatile float4 *olo = node_lo + node, *ohi = node_hi + node;
        olo->x = lo.x; olo->y = lo.y; olo->z = lo.z; olo->w = 0.0f;
        ohi->x = hi.x; ohi->y = hi.y; ohi->z = hi.z; ohi->w = 0.0f;
        node = parent_of_node[node];
    }
}

// ---------------------------------------------------------------- PLOC: the binary tree by parallel locally-ordered clustering
// (Meister & Bittner 2018) instead of the Karras radix tree.  The radix tree splits where a Morton bit changes, whatever the boxes
// look like; PLOC builds bottom-up: every cluster (at first: one triangle) looks at its R neighbours on either side in Morton order,
// picks the one whose union with it has the smallest surface area, and mutually chosen pairs merge; repeat until one cluster is left.
// Measured on the two LBVH scenes (tests per ray as surface-area cost, leaves of <= 3): box cost 77.4 -> 66.5 / 81.9 -> 68.8, triangle
// cost 7.4 -> 6.6 / 5.9 -> 5.3 (a binned-SAH top-down build: 62.3 / 6.7).  Everything downstream (leaf collapse, 4-ary collapse,
// quantised nodes, traversal) only sees child[], range[], parents and boxes, so the closest-hit contract is untouched.
// Cluster record: lo.xyz | bits(node code: >= 0 internal node id, < 0 ~Morton position of a leaf), hi.xyz | bits(triangles below).
#define PLOC_R 8
__device__ __forceinline__ float ploc_area(float4 alo, float4 ahi, float4 blo, float4 bhi) {
    const float ex = fmaxf(ahi.x, bhi.x) - fminf(alo.x, blo.x), ey = fmaxf(ahi.y, bhi.y) - fminf(alo.y, blo.y), ez = fmaxf(ahi.z, bhi.z) - fminf(alo.z, blo.z);
    float a = ex * ey + ey * ez + ez * ex;
    if (!(a < 3.0e38f)) a = 3.0e38f;                 // NaN / inf boxes (degenerate input) still get a partner
    return a;
}
__global__ void k_ploc_init(int m, const float4 *leaf_lo, const float4 *leaf_hi, float4 *clo, float4 *chi) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= m) return;
    const float4 l = leaf_lo[p], h = leaf_hi[p];
    clo[p] = make_float4(l.x, l.y, l.z, __int_as_float(~p));
    chi[p] = make_float4(h.x, h.y, h.z, __int_as_float(1));
}
// nearest neighbour of every cluster within PLOC_R positions.  Candidate pairs are compared by (area, not the aligned pair i^1, distance,
// lower index): a strict total order over unordered pairs, symmetric in the pair, so the globally smallest pair is chosen by both of its
// clusters — at least one merge per round, also when all areas are equal (a stack of identical triangles).
template <int R>
__global__ void __launch_bounds__(256) k_ploc_nn(int m, const float4 *clo, const float4 *chi, int *nn) {
    __shared__ float4 slo[256 + 2 * R], shi[256 + 2 * R];
    const int base = blockIdx.x * 256 - R;
    for (int t = threadIdx.x; t < 256 + 2 * R; t += 256) {
        const int g = base + t;
        if (g >= 0 && g < m) { slo[t] = clo[g]; shi[t] = chi[g]; }
    }
    __syncthreads();
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= m) return;
    const int li = threadIdx.x + R;
    const float4 mlo = slo[li], mhi = shi[li];
    float best = 4.0e38f;
    int bestj = -1, best_na = 2, best_d = 1 << 30, best_mn = 1 << 30;
#pragma unroll
    for (int o = -R; o <= R; ++o) {
        if (o == 0) continue;
        const int j = i + o;
        if (j < 0 || j >= m) continue;
        const float a = ploc_area(mlo, mhi, slo[li + o], shi[li + o]);
        const int na = ((i ^ 1) == j) ? 0 : 1, d = o < 0 ? -o : o, mn = i < j ? i : j;
        const bool better = a < best || (a == best && (na < best_na || (na == best_na && (d < best_d || (d == best_d && mn < best_mn)))));
        if (better) { best = a; bestj = j; best_na = na; best_d = d; best_mn = mn; }
    }
    nn[i] = bestj;
}
// flags: bit 0 = the cluster survives into the next round, bit 32 = it is the lower half of a mutual pair (creates a node)
__global__ void k_ploc_flag(int m, const int *nn, unsigned long long *flags) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const int j = nn[i];
    const bool mutual = j >= 0 && nn[j] == i;
    flags[i] = ((mutual && i > j) ? 0ull : 1ull) | ((mutual && i < j) ? (1ull << 32) : 0ull);
}
// node ids are handed out downwards from next_id, so that the last merge — the root — is node 0
__global__ void k_ploc_apply(int m, int next_id, const int *nn, const unsigned long long *flags, const unsigned long long *scan,
                             const float4 *clo, const float4 *chi, float4 *olo, float4 *ohi, int2 *child, float4 *node_lo, float4 *node_hi,
                             int *size, int *parent_of_node, int *parent_of_leaf) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const unsigned long long f = flags[i], sc = scan[i];
    if (!(f & 1ull)) return;
    float4 lo = clo[i], hi = chi[i];
    if (f >> 32) {
        const int j = nn[i];
        const float4 lo2 = clo[j], hi2 = chi[j];
        const int id = next_id - (int)(sc >> 32);
        const int ca = __float_as_int(lo.w), cb = __float_as_int(lo2.w), n_below = __float_as_int(hi.w) + __float_as_int(hi2.w);
        child[id] = make_int2(ca, cb);
        lo = make_float4(fminf(lo.x, lo2.x), fminf(lo.y, lo2.y), fminf(lo.z, lo2.z), __int_as_float(id));
        hi = make_float4(fmaxf(hi.x, hi2.x), fmaxf(hi.y, hi2.y), fmaxf(hi.z, hi2.z), __int_as_float(n_below));
        node_lo[id] = make_float4(lo.x, lo.y, lo.z, 0.0f);
        node_hi[id] = make_float4(hi.x, hi.y, hi.z, 0.0f);
        size[id] = n_below;
        if (ca >= 0) parent_of_node[ca] = id; else parent_of_leaf[~ca] = id;
        if (cb >= 0) parent_of_node[cb] = id; else parent_of_leaf[~cb] = id;
    }
    const int pos = (int)(sc & 0xffffffffull);
    olo[pos] = lo; ohi[pos] = hi;
}
// final position of every Morton leaf: the leaves of a subtree must be contiguous (a collapsed leaf is a range [first, first + count)),
// so leaf p lands at the number of triangles to its left in the tree = sum of the left siblings' sizes on its way up
__global__ void k_ploc_newpos(int m, const int2 *child, const int *size, const int *parent_of_node, const int *parent_of_leaf, int *newpos) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= m) return;
    int off = 0, code = ~p;
    for (int node = parent_of_leaf[p]; node >= 0; node = parent_of_node[node]) {
        const int2 c = child[node];
        if (c.y == code) off += c.x >= 0 ? size[c.x] : 1;
        code = node;
    }
    newpos[p] = off;
}
__global__ void k_ploc_range(int m, const int2 *child, const int *size, const int *newpos, int2 *range) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m - 1) return;
    int c = child[i].x;
    while (c >= 0) c = child[c].x;
    const int first = newpos[~c];
    range[i] = make_int2(first, first + size[i] - 1);
}
__global__ void k_ploc_remap(int m, int n, const int *newpos, int2 *child, const int *vals_in, int *vals_out, const int *pleaf_in, int *pleaf_out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m - 1) {
        int2 c = child[i];
        if (c.x < 0) c.x = ~newpos[~c.x];
        if (c.y < 0) c.y = ~newpos[~c.y];
        child[i] = c;
    }
    if (i < m) { vals_out[newpos[i]] = vals_in[i]; pleaf_out[newpos[i]] = pleaf_in[i]; }
    else if (i < n) vals_out[i] = vals_in[i];                    // the outlier list keeps its place behind the tree's triangles
}

__device__ __forceinline__ int leaf_link(int first, int count) { return ~((first << 3) | (count - 1)); }

__global__ void k_emit(int n, int leaf_size, float eps, const int2 *range, const int2 *child,
                       const float4 *leaf_lo, const float4 *leaf_hi, const float4 *node_lo, const float4 *node_hi, float4 *bvh) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int2 rg = range[i];
    if (rg.y - rg.x + 1 <= leaf_size && i != 0) return;   // inside a collapsed leaf: never visited
    int2 c = child[i];
    float4 lo[2], hi[2];
    int link[2];
    int cc[2] = {c.x, c.y};
    for (int k = 0; k < 2; ++k) {
        if (cc[k] >= 0) {
            lo[k] = node_lo[cc[k]]; hi[k] = node_hi[cc[k]];
            int2 cr = range[cc[k]];
            int cnt = cr.y - cr.x + 1;
            link[k] = cnt <= leaf_size ? leaf_link(cr.x, cnt) : cc[k];
        } else {
            lo[k] = leaf_lo[~cc[k]]; hi[k] = leaf_hi[~cc[k]];
            link[k] = leaf_link(~cc[k], 1);
        }
        lo[k].x -= eps; lo[k].y -= eps; lo[k].z -= eps;
        hi[k].x += eps; hi[k].y += eps; hi[k].z += eps;
    }
    bvh[4 * i + 0] = make_float4(lo[0].x, hi[0].x, lo[0].y, hi[0].y);
    bvh[4 * i + 1] = make_float4(lo[1].x, hi[1].x, lo[1].y, hi[1].y);
    bvh[4 * i + 2] = make_float4(lo[0].z, hi[0].z, lo[1].z, hi[1].z);
    bvh[4 * i + 3] = make_float4(__int_as_float(link[0]), __int_as_float(link[1]), 0.0f, 0.0f);
}

// ---- 4-ary collapse
__device__ __forceinline__ bool emitted_node(const int2 *range, int i, int leaf_size) {
    const int2 rg = range[i];
    return rg.y - rg.x + 1 > leaf_size || i == 0;
}
// flag[i] = 1 iff binary node i is emitted and has an even number of ancestors
__global__ void k_flag4(int n, int leaf_size, const int2 *range, const int *parent_of_node, int *flag) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int f = 0;
    if (emitted_node(range, i, leaf_size)) {
        int depth = 0;
        for (int a = parent_of_node[i]; a >= 0; a = parent_of_node[a]) depth++;
        f = (depth & 1) == 0;
    }
    flag[i] = f;
}
struct Slot { float4 lo, hi; int link; };
// the traversal link of binary child c (as stored in child[]): leaf, collapsed leaf, or wide node id
__device__ __forceinline__ bool slot_of(int c, int leaf_size, const int2 *range, const int *id4, const float4 *leaf_lo, const float4 *leaf_hi,
                                        const float4 *node_lo, const float4 *node_hi, Slot &s) {
    if (c < 0) { s.lo = leaf_lo[~c]; s.hi = leaf_hi[~c]; s.link = leaf_link(~c, 1); return true; }
    s.lo = node_lo[c]; s.hi = node_hi[c];
    const int2 cr = range[c];
    const int cnt = cr.y - cr.x + 1;
    if (cnt <= leaf_size) { s.link = leaf_link(cr.x, cnt); return true; }
    s.link = id4[c];
    return false;                                      // an emitted internal node
}
__global__ void k_emit4(int n, int leaf_size, float eps, const int2 *range, const int2 *child, const int *flag, const int *id4,
                        const float4 *leaf_lo, const float4 *leaf_hi, const float4 *node_lo, const float4 *node_hi, float4 *bvh4) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1 || !flag[i]) return;
    Slot s[4];
    int ns = 0;
    const int2 c = child[i];
    const int cc[2] = {c.x, c.y};
    for (int k = 0; k < 2; ++k) {
        Slot t;
        if (slot_of(cc[k], leaf_size, range, id4, leaf_lo, leaf_hi, node_lo, node_hi, t)) { s[ns++] = t; continue; }
        const int2 g = child[cc[k]];                   // odd-depth internal node: absorbed, its children take the slots
        const int gg[2] = {g.x, g.y};
        for (int j = 0; j < 2; ++j) {
            slot_of(gg[j], leaf_size, range, id4, leaf_lo, leaf_hi, node_lo, node_hi, t);
            s[ns++] = t;
        }
    }
    // empty slots: an inverted infinite box (entry +inf, exit -inf for every direction sign: never hit; NaNs would be
    // dropped by the min/max of the slab test)
    const float pinf = __int_as_float(0x7f800000), ninf = __int_as_float(0xff800000);
    float v[6][4];
    int link[4];
    for (int k = 0; k < 4; ++k) {
        const bool on = k < ns;
        v[0][k] = on ? s[k].lo.x - eps : pinf; v[1][k] = on ? s[k].hi.x + eps : ninf;
        v[2][k] = on ? s[k].lo.y - eps : pinf; v[3][k] = on ? s[k].hi.y + eps : ninf;
        v[4][k] = on ? s[k].lo.z - eps : pinf; v[5][k] = on ? s[k].hi.z + eps : ninf;
        link[k] = on ? s[k].link : leaf_link(0, 1);
    }
    float4 *o = bvh4 + 8 * (size_t)id4[i];
    for (int r = 0; r < 6; ++r) o[r] = make_float4(v[r][0], v[r][1], v[r][2], v[r][3]);
    o[6] = make_float4(__int_as_float(link[0]), __int_as_float(link[1]), __int_as_float(link[2]), __int_as_float(link[3]));
    o[7] = make_float4(__int_as_float(ns), 0.0f, 0.0f, 0.0f);
}

// ---- quantised 4-ary nodes: 64 bytes instead of 128, so that a node is TWO 256-bit loads (two L1 wavefronts per lane) instead
// of five fetches — the L1 data pipes, one wavefront per lane and scattered access, are what bounds the traversal
// (profiles/r02_pool_v1.txt).  Layout, 16 words:
//   w0-2  origin p (float): two quantisation steps below the union box of the children
//   w3    biased exponents ex | ey << 8 | ez << 16 of 128 * step (step = 2^k per axis, >= extent / 250)
//   w4-9  per axis the four children's lower bounds (one byte each), then their upper bounds: x lo, x hi, y lo, y hi, z lo, z hi
//   w12-15 links (same encoding as the 128-byte node)
// A child's box is [p + qlo * step, p + qhi * step] with qlo = floor(x - 2^-7) and qhi = ceil(x + 2^-7), x in steps from p: it contains
// the (already eps-inflated) float box with 2^-7 of a step to spare on every side, which covers the rounding of x here and of the decode
// in the traversal (< 2^-8 step, see k_pool).  (A whole step of slack, the first version, cost 7 % more box tests per ray.)
// Empty slot: qlo = 255, qhi = 0 (inverted on every axis: never hit).
__global__ void k_emit4q(int n, int leaf_size, float eps, const int2 *range, const int2 *child, const int *flag, const int *id4,
                         const float4 *leaf_lo, const float4 *leaf_hi, const float4 *node_lo, const float4 *node_hi, uint4 *bvh4q) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1 || !flag[i]) return;
    Slot s[4];
    int ns = 0;
    const int2 c = child[i];
    const int cc[2] = {c.x, c.y};
    for (int k = 0; k < 2; ++k) {
        Slot t;
        if (slot_of(cc[k], leaf_size, range, id4, leaf_lo, leaf_hi, node_lo, node_hi, t)) { s[ns++] = t; continue; }
        const int2 g = child[cc[k]];
        const int gg[2] = {g.x, g.y};
        for (int j = 0; j < 2; ++j) {
            slot_of(gg[j], leaf_size, range, id4, leaf_lo, leaf_hi, node_lo, node_hi, t);
            s[ns++] = t;
        }
    }
    float lo[4][3], hi[4][3], ulo[3] = {3e38f, 3e38f, 3e38f}, uhi[3] = {-3e38f, -3e38f, -3e38f};
    for (int k = 0; k < ns; ++k) {
        lo[k][0] = s[k].lo.x - eps; lo[k][1] = s[k].lo.y - eps; lo[k][2] = s[k].lo.z - eps;
        hi[k][0] = s[k].hi.x + eps; hi[k][1] = s[k].hi.y + eps; hi[k][2] = s[k].hi.z + eps;
        for (int a = 0; a < 3; ++a) { ulo[a] = fminf(ulo[a], lo[k][a]); uhi[a] = fmaxf(uhi[a], hi[k][a]); }
    }
    unsigned w[16];
    for (int k = 0; k < 16; ++k) w[k] = 0u;
    unsigned exps = 0u;
    for (int a = 0; a < 3; ++a) {
        float ext = uhi[a] - ulo[a];
        if (!(ext > 1e-30f)) ext = 1e-30f;
        int e2 = 0;
        frexpf(ext * (1.0f / 250.0f), &e2);                 // ext / 250 = m * 2^e2, m in [0.5, 1): step = 2^e2 >= ext / 250
        e2 = max(-120, min(120, e2));
        const float step = ldexpf(1.0f, e2), inv_step = ldexpf(1.0f, -e2);
        const float p = ulo[a] - step;
        w[a] = __float_as_uint(p);
        exps |= (unsigned)(e2 + 7 + 127) << (8 * a);
        unsigned qlo = 0u, qhi = 0u;
        for (int k = 0; k < 4; ++k) {
            int ql = 255, qh = 0;
            if (k < ns) {
                ql = (int)floorf((lo[k][a] - p) * inv_step - 0.0078125f);      // 2^-7 of a step outward: covers the subtraction's rounding here
                qh = (int)ceilf((hi[k][a] - p) * inv_step + 0.0078125f);       // (<= 2^-16 step) and the decode's in the traversal (< 2^-8 step)
                ql = max(0, min(255, ql)); qh = max(0, min(255, qh));
            }
            qlo |= (unsigned)ql << (8 * k); qhi |= (unsigned)qh << (8 * k);
        }
        w[4 + 2 * a] = qlo; w[5 + 2 * a] = qhi;
    }
    w[3] = exps;
    for (int k = 0; k < 4; ++k) w[12 + k] = (unsigned)(k < ns ? s[k].link : leaf_link(0, 1));
    uint4 *o = bvh4q + 4 * (size_t)id4[i];
    for (int r = 0; r < 4; ++r) o[r] = make_uint4(w[4 * r], w[4 * r + 1], w[4 * r + 2], w[4 * r + 3]);
}

// depth of every triangle's leaf in the EMITTED tree = number of emitted ancestors
__global__ void k_depth(int n, int leaf_size, const int2 *range, const int *parent_of_node, const int *parent_of_leaf, int *max_depth, int *n_emitted) {
    int pos = blockIdx.x * blockDim.x + threadIdx.x;
    int depth = 0;
    if (pos < n) {
        for (int node = parent_of_leaf[pos]; node >= 0; node = parent_of_node[node]) {
            int2 rg = range[node];
            if (rg.y - rg.x + 1 > leaf_size || node == 0) depth++;
        }
    }
    for (int o = 16; o; o >>= 1) depth = max(depth, __shfl_xor_sync(0xffffffffu, depth, o));
    if ((threadIdx.x & 31) == 0) atomicMax(max_depth, depth);
    if (pos < n - 1) {
        int2 rg = range[pos];
        if (rg.y - rg.x + 1 > leaf_size || pos == 0) atomicAdd(n_emitted, 1);
    }
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { snprintf(err, errcap, "lbvh: %s failed: %s", #x, cudaGetErrorString(e_)); return -1; } } while (0)


}  // namespace

static cudaTextureObject_t linear_tex(const float4 *p, size_t count, bool as_uint = false) {
    if (!p || count == 0) return 0;
    cudaResourceDesc rd;
    memset(&rd, 0, sizeof rd);
    rd.resType = cudaResourceTypeLinear;
    rd.res.linear.devPtr = const_cast<float4 *>(p);
    rd.res.linear.desc = as_uint ? cudaCreateChannelDesc<uint4>() : cudaCreateChannelDesc<float4>();     // quantised nodes are bit patterns, not floats
    rd.res.linear.sizeInBytes = count * sizeof(float4);
    cudaTextureDesc td;
    memset(&td, 0, sizeof td);
    td.readMode = cudaReadModeElementType;
    cudaTextureObject_t t = 0;
    if (cudaCreateTextureObject(&t, &rd, &td, nullptr) != cudaSuccess) { cudaGetLastError(); return 0; }
    return t;
}

void lbvh_free(Lbvh *b) {
    if (!b) return;
    if (b->bvh_tex) cudaDestroyTextureObject(b->bvh_tex);
    if (b->bvh4_tex) cudaDestroyTextureObject(b->bvh4_tex);
    if (b->bvh4q_tex) cudaDestroyTextureObject(b->bvh4q_tex);
    cudaFree(b->arena);
    memset(b, 0, sizeof(*b));
}

namespace {
// bump allocator over the arena: every array 256-byte aligned; with base == nullptr it only measures
struct Carver {
    char *base;
    size_t off;
    template <typename T> T *take(size_t count) {
        off = (off + 255) & ~(size_t)255;
        T *p = base ? reinterpret_cast<T *>(base + off) : nullptr;
        off += count * sizeof(T) + 16;
        return p;
    }
};
struct Scratch {
    unsigned *bounds; unsigned long long *keys, *keys2; int *vals, *vals2, *pleaf, *arrivals, *stats;
    void *sort_tmp, *scan_tmp;
    // PLOC: two cluster lists (lo | code, hi | size), the round's flags' prefix sums, subtree sizes, the tree-order permutation
    float4 *clo[2], *chi[2];
    unsigned long long *pscan;
    int *size, *vals3, *pleaf2;
};
// the layout is a function of the capacity alone, so the pointers stay put across rebuilds
size_t lbvh_carve(Lbvh *b, Scratch &t, char *base, size_t cap, size_t sort_bytes, size_t scan_bytes) {
    Carver c{base, 0};
    b->bvh = c.take<float4>(4 * cap); b->bvh4 = c.take<float4>(8 * cap); b->bvh4q = c.take<float4>(4 * cap);
    b->parent = c.take<int>(cap); b->flag4 = c.take<int>(cap); b->id4 = c.take<int>(cap + 1);
    b->sq0 = c.take<float4>(cap); b->sq1 = c.take<float4>(cap); b->sq2 = c.take<float4>(cap);
    b->sf0 = c.take<float4>(cap); b->sf1 = c.take<float4>(cap); b->sf2 = c.take<float4>(cap);
    b->sorted_id = c.take<int>(cap);
    b->rq = c.take<float4>(4 * cap); b->rf = c.take<float4>(4 * cap);
    b->range = c.take<int2>(cap); b->child = c.take<int2>(cap);
    b->leaf_lo = c.take<float4>(cap); b->leaf_hi = c.take<float4>(cap); b->node_lo = c.take<float4>(cap); b->node_hi = c.take<float4>(cap);
    t.bounds = c.take<unsigned>(8); t.keys = c.take<unsigned long long>(cap); t.keys2 = c.take<unsigned long long>(cap);
    t.vals = c.take<int>(cap); t.vals2 = c.take<int>(cap); t.pleaf = c.take<int>(cap); t.arrivals = c.take<int>(cap); t.stats = c.take<int>(4);
    t.sort_tmp = c.take<char>(sort_bytes); t.scan_tmp = c.take<char>(scan_bytes);
    for (int k = 0; k < 2; ++k) { t.clo[k] = c.take<float4>(cap); t.chi[k] = c.take<float4>(cap); }
    t.pscan = c.take<unsigned long long>(cap); t.size = c.take<int>(cap); t.vals3 = c.take<int>(cap); t.pleaf2 = c.take<int>(cap);
    return c.off + 256;
}
}  // namespace

int lbvh_emit(Lbvh *b, float reach, cudaStream_t st, char *err, size_t errcap) {
    // inflate every box by 2^-16 of the largest coordinate magnitude a ray can carry (scene or eye):
    // covers the rounding slop of the reference's edge tests and of p = P + V*t
    float eps = ldexpf(fmaxf(reach, b->max_abs), -16);
    if (!(eps > 0.0f)) eps = 1e-30f;
    b->reach = fmaxf(reach, b->max_abs);
    b->eps = eps;
    if (b->n_tree >= 2) {
        int n = b->n_tree;
        k_emit<<<(n - 1 + 255) / 256, 256, 0, st>>>(n, b->leaf_size, eps, (const int2 *)b->range, (const int2 *)b->child,
                                                   b->leaf_lo, b->leaf_hi, b->node_lo, b->node_hi, b->bvh);
        CK(cudaGetLastError());
        b->launches++;
        if (b->n_nodes4 > 0) {
            k_emit4<<<(n - 1 + 255) / 256, 256, 0, st>>>(n, b->leaf_size, eps, (const int2 *)b->range, (const int2 *)b->child, b->flag4, b->id4,
                                                        b->leaf_lo, b->leaf_hi, b->node_lo, b->node_hi, b->bvh4);
            CK(cudaGetLastError());
            b->launches++;
            k_emit4q<<<(n - 1 + 255) / 256, 256, 0, st>>>(n, b->leaf_size, eps, (const int2 *)b->range, (const int2 *)b->child, b->flag4, b->id4,
                                                         b->leaf_lo, b->leaf_hi, b->node_lo, b->node_hi, (uint4 *)b->bvh4q);
            CK(cudaGetLastError());
            b->launches++;
        }
    }
    return 0;
}

int lbvh_build(Lbvh *b, const float4 *q0, const float4 *q1, const float4 *q2, const float4 *f0, const float4 *f1, const float4 *f2,
               int n, int leaf_size, float big_frac, cudaStream_t st, char *err, size_t errcap) {
    if (b->bvh_tex) cudaDestroyTextureObject(b->bvh_tex);
    if (b->bvh4_tex) cudaDestroyTextureObject(b->bvh4_tex);
    if (b->bvh4q_tex) cudaDestroyTextureObject(b->bvh4q_tex);
    {   // forget the previous tree, keep the arena
        void *arena = b->arena; const size_t arena_bytes = b->arena_bytes; const int cap = b->cap;
        memset(b, 0, sizeof(*b));
        b->arena = arena; b->arena_bytes = arena_bytes; b->cap = cap;
    }
    b->n_tris = n;
    b->n_tree = n;
    b->n_big = 0;
    b->leaf_size = leaf_size < 1 ? 1 : (leaf_size > 8 ? 8 : leaf_size);
    if (n <= 0) { b->root = ~0; b->n_tris = 0; b->n_tree = 0; return 0; }
    if (n >= (1 << 28)) { snprintf(err, errcap, "lbvh: too many triangles (%d)", n); return -1; }
    if (n > b->cap) { cudaFree(b->arena); b->arena = nullptr; b->arena_bytes = 0; b->cap = 0; }
    const int cap = b->cap > 0 ? b->cap : n;
    Scratch t{};
    size_t sort_bytes = 0, scan_bytes = 0;
    CK(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (unsigned long long *)nullptr, (unsigned long long *)nullptr, (int *)nullptr, (int *)nullptr, cap, 0, 64, st));
    CK(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (int *)nullptr, (int *)nullptr, cap, st));
    {
        size_t scan64 = 0;
        CK(cub::DeviceScan::ExclusiveSum(nullptr, scan64, (unsigned long long *)nullptr, (unsigned long long *)nullptr, cap, st));
        if (scan64 > scan_bytes) scan_bytes = scan64;
    }
    if (!b->arena) {
        const size_t bytes = lbvh_carve(b, t, nullptr, (size_t)cap, sort_bytes, scan_bytes);
        CK(cudaMalloc(&b->arena, bytes));
        b->arena_bytes = bytes;
        b->cap = cap;
    }
    lbvh_carve(b, t, (char *)b->arena, (size_t)cap, sort_bytes, scan_bytes);
    unsigned *bounds = t.bounds; unsigned long long *keys = t.keys, *keys2 = t.keys2;
    int *vals = t.vals, *vals2 = t.vals2, *pnode = b->parent, *pleaf = t.pleaf, *arrivals = t.arrivals, *stats = t.stats;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0, st));

    unsigned init[8] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u, 0u, 0u};
    CK(cudaMemcpyAsync(bounds, init, sizeof(init), cudaMemcpyHostToDevice, st));
    int grid = (n + 255) / 256;
    k_bounds<<<grid > 1184 ? 1184 : grid, 256, 0, st>>>(q0, q1, q2, n, bounds);
    CK(cudaMemsetAsync(stats, 0, 16, st));
    k_morton<<<grid, 256, 0, st>>>(q0, q1, q2, n, bounds, keys, vals, big_frac, stats + 2);
    b->launches += 2;
    int n_big = 0;
    if (big_frac > 0.0f) {
        CK(cudaMemcpyAsync(&n_big, stats + 2, sizeof(int), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (n_big > LBVH_BIG_MAX || n - n_big < 2) {        // too many outliers (or nothing else): everything goes into the tree
            k_morton<<<grid, 256, 0, st>>>(q0, q1, q2, n, bounds, keys, vals, 0.0f, stats + 2);
            b->launches += 1;
            n_big = 0;
        }
    }
    b->n_big = n_big;
    b->n_tree = n - n_big;
    const int m = b->n_tree;                                // triangles under the tree: Morton positions [0, m)
    CK(cub::DeviceRadixSort::SortPairs(t.sort_tmp, sort_bytes, keys, keys2, vals, vals2, n, 0, 64, st));
    k_gather<<<grid, 256, 0, st>>>(q0, q1, q2, f0, f1, f2, vals2, n, b->sq0, b->sq1, b->sq2, b->sf0, b->sf1, b->sf2, b->leaf_lo, b->leaf_hi, b->rq, b->rf);
    CK(cudaMemcpyAsync(b->sorted_id, vals2, (size_t)n * sizeof(int), cudaMemcpyDeviceToDevice, st));
    b->launches += 1;
    if (m >= 2) {
        static const int use_ploc = getenv("PTB_LBVH_PLOC") ? atoi(getenv("PTB_LBVH_PLOC")) : 1;
        bool ploc_done = false;
        b->ploc_rounds = 0;
        if (use_ploc) {
            // ---- PLOC rounds over the Morton-ordered triangles [0, m); keys (the unsorted codes) and vals are free after the sort: flags and nn
            unsigned long long *flags = keys;
            int *nn = vals;
            k_ploc_init<<<(m + 255) / 256, 256, 0, st>>>(m, b->leaf_lo, b->leaf_hi, t.clo[0], t.chi[0]);
            int cur = 0, mc = m, next_id = m - 2, rounds = 0;
            unsigned long long h_last[2];
            static const int ploc_r = getenv("PTB_PLOC_R") ? atoi(getenv("PTB_PLOC_R")) : PLOC_R;      // search radius: 8 (default), 16, 32
            while (mc > 1 && rounds < 1024) {
                const int g = (mc + 255) / 256;
                if (ploc_r >= 32) k_ploc_nn<32><<<g, 256, 0, st>>>(mc, t.clo[cur], t.chi[cur], nn);
                else if (ploc_r >= 16) k_ploc_nn<16><<<g, 256, 0, st>>>(mc, t.clo[cur], t.chi[cur], nn);
                else k_ploc_nn<PLOC_R><<<g, 256, 0, st>>>(mc, t.clo[cur], t.chi[cur], nn);
                k_ploc_flag<<<g, 256, 0, st>>>(mc, nn, flags);
                CK(cub::DeviceScan::ExclusiveSum(t.scan_tmp, scan_bytes, flags, t.pscan, mc, st));
                k_ploc_apply<<<g, 256, 0, st>>>(mc, next_id, nn, flags, t.pscan, t.clo[cur], t.chi[cur], t.clo[cur ^ 1], t.chi[cur ^ 1], (int2 *)b->child,
                                                b->node_lo, b->node_hi, t.size, pnode, pleaf);
                CK(cudaMemcpyAsync(&h_last[0], flags + (mc - 1), 8, cudaMemcpyDeviceToHost, st));
                CK(cudaMemcpyAsync(&h_last[1], t.pscan + (mc - 1), 8, cudaMemcpyDeviceToHost, st));
                CK(cudaStreamSynchronize(st));
                const unsigned long long tot = h_last[0] + h_last[1];
                const int kept = (int)(tot & 0xffffffffull), merged = (int)(tot >> 32);
                b->launches += 5;
                ++rounds;
                if (merged < 1 || kept != mc - merged) break;           // cannot happen (see k_ploc_nn); the radix tree takes over
                next_id -= merged;
                mc = kept;
                cur ^= 1;
            }
            if (mc == 1 && next_id == -1) {
                // ---- tree order: positions, ranges, remapped leaf links; the triangles are gathered again in that order below
                const int root_parent = -1;
                CK(cudaMemcpyAsync(pnode, &root_parent, sizeof(int), cudaMemcpyHostToDevice, st));      // node 0 is the root
                k_ploc_newpos<<<grid, 256, 0, st>>>(m, (const int2 *)b->child, t.size, pnode, pleaf, arrivals);
                k_ploc_range<<<grid, 256, 0, st>>>(m, (const int2 *)b->child, t.size, arrivals, (int2 *)b->range);
                k_ploc_remap<<<grid, 256, 0, st>>>(m, n, arrivals, (int2 *)b->child, vals2, t.vals3, pleaf, t.pleaf2);
                k_gather<<<grid, 256, 0, st>>>(q0, q1, q2, f0, f1, f2, t.vals3, n, b->sq0, b->sq1, b->sq2, b->sf0, b->sf1, b->sf2, b->leaf_lo, b->leaf_hi, b->rq, b->rf);
                CK(cudaMemcpyAsync(b->sorted_id, t.vals3, (size_t)n * sizeof(int), cudaMemcpyDeviceToDevice, st));
                CK(cudaMemcpyAsync(pleaf, t.pleaf2, (size_t)m * sizeof(int), cudaMemcpyDeviceToDevice, st));
                CK(cudaGetLastError());
                b->launches += 4;
                b->ploc_rounds = rounds;
                ploc_done = true;
            }
        }
        if (!ploc_done) {
            CK(cudaMemsetAsync(arrivals, 0, (size_t)n * sizeof(int), st));
            k_hierarchy<<<grid, 256, 0, st>>>(keys2, m, (int2 *)b->range, (int2 *)b->child, pnode, pleaf);
            k_refit<<<grid, 256, 0, st>>>(m, (const int2 *)b->child, pnode, pleaf, b->leaf_lo, b->leaf_hi, b->node_lo, b->node_hi, arrivals);
            b->launches += 2;
        }
        k_depth<<<grid, 256, 0, st>>>(m, b->leaf_size, (const int2 *)b->range, pnode, pleaf, stats, stats + 1);
        b->launches += 1;
    }
    unsigned hb[8];
    int hs[4];
    CK(cudaMemcpyAsync(hb, bounds, sizeof(hb), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(hs, stats, sizeof(hs), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    b->max_abs = ord2f(hb[6]);
    b->max_depth = hs[0];
    b->n_nodes = m >= 2 ? hs[1] : 0;
    b->root = (m <= b->leaf_size || m < 2) ? ~(((0) << 3) | (m - 1)) : 0;
    b->n_nodes4 = 0;
    if (b->root >= 0) {                              // 4-ary collapse: flags, compact ids, node storage
        k_flag4<<<grid, 256, 0, st>>>(m, b->leaf_size, (const int2 *)b->range, pnode, b->flag4);
        CK(cudaMemsetAsync(b->flag4 + (m - 1), 0, sizeof(int), st));
        CK(cub::DeviceScan::ExclusiveSum(t.scan_tmp, scan_bytes, b->flag4, b->id4, m, st));
        int total = 0;
        CK(cudaMemcpyAsync(&total, b->id4 + (m - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        b->launches += 2;
        b->n_nodes4 = total;                         // <= m - 1 < cap: the arena's 8 * cap float4 always hold them
    }
    if (lbvh_emit(b, 0.0f, st, err, errcap)) return -1;
    if (m >= 2 && (size_t)4 * m < ((size_t)1 << 27)) b->bvh_tex = linear_tex(b->bvh, (size_t)4 * m);
    if (b->n_nodes4 > 0 && (size_t)8 * b->n_nodes4 < ((size_t)1 << 27)) b->bvh4_tex = linear_tex(b->bvh4, (size_t)8 * b->n_nodes4);
    if (b->n_nodes4 > 0 && (size_t)4 * b->n_nodes4 < ((size_t)1 << 27)) b->bvh4q_tex = linear_tex(b->bvh4q, (size_t)4 * b->n_nodes4, true);
    CK(cudaEventRecord(e1, st));
    CK(cudaEventSynchronize(e1));
    CK(cudaEventElapsedTime(&b->build_ms, e0, e1));
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    CK(cudaGetLastError());
    return 0;
}
