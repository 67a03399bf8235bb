// ptb_abi.cu — the C-ABI of include/ptb200.h on CUDA (libptb200.so).
//
// Replaces the reference host's OpenCL plumbing (main.cpp:389-528, 618-682): buffers are plain
// device allocations behind ref-counted handles, "kernels" are argument tables for the two entry
// points of prog.cl, and a launch resolves the arguments into a RenderParams, stages the uploaded
// triangle records into the SoA float4 layout (once per upload), builds the LBVH (once per upload)
// and enqueues the fused path kernel on one in-order stream.  There is no CPU path: every entry
// point that needs the device fails loudly when CUDA is not available.
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <string>
#include <vector>

#include "comm.h"
#include "ingest.h"
#include "lbvh.h"
#include "prims.h"
#include "pt_params.h"
#include "ptb200.h"
#include "resolve.h"
static_assert(PTB_NOALLOC_MAX_PIXELS == RESOLVE_NOALLOC_MAX_PIXELS, "ptb200.h and resolve.h disagree on the static scratch size");

namespace {

struct MemObj {
    void *dptr = nullptr;
    size_t bytes = 0;
    int refs = 1;
    bool owned = true;
    bool ipc = false;                 // mapped from another process (ptb_ipc_import): closed, not freed, on release
    bool image = false;
    int w = 0, h = 0;
    unsigned long long version = 0;   // bumped by every write: invalidates staged copies
    // RNG-state buffers: version + 1 at which the first `seeds_checked` ints were scanned, and what the scan found
    unsigned long long seeds_version = 0;
    size_t seeds_checked = 0;
    bool seeds_nonneg = false;
};

struct Arg {
    std::vector<unsigned char> value;
    ptb_mem mem = 0;
    bool set = false;
};

struct Staged {                        // device-internal copy of the scene, valid for one upload
    ptb_mem tris_handle = 0;
    unsigned long long tris_version = ~0ull;
    int n_tris = -1;
    unsigned max_mati = 0;             // largest Triangle::mati of the upload (checked against the materials buffer per launch)
    float4 *q0 = nullptr, *q1 = nullptr, *q2 = nullptr;   // parity layout: (r1,N.x) (r2,N.y) (r3,N.z)
    float4 *f0 = nullptr, *f1 = nullptr, *f2 = nullptr;   // fast layout: (N,r1.x) (r1.yz,mu.xy) (mu.z,mv)
    uint16_t *mati = nullptr;
    int capacity = 0;
    Lbvh lbvh;
    bool lbvh_valid = false;
    int lbvh_leaf = 0;
    // merged scan primitives (prims.cu), built on first use by a fast brute-force launch
    float4 *p0 = nullptr, *p1 = nullptr, *p2 = nullptr;
    float2 *prim_ab = nullptr;
    float4 *prim_shade = nullptr;
    int *prim_rep = nullptr, *n_prims_dev = nullptr;
    int n_prims = 0;
    bool prims_valid = false;
    // the LBVH's outlier list as merged primitives (same records, at most LBVH_BIG_MAX), rebuilt with the tree
    float4 *b0 = nullptr, *b1 = nullptr, *b2 = nullptr, *big_shade = nullptr;
    float2 *big_ab = nullptr;
    int *big_rep = nullptr;
    int n_big_prims = 0;
};

struct PendingGen {                    // a deferred gen_ray launch (ptb_set_fuse_gen)
    bool active = false;
    ptb_mem rays = 0, rnds = 0;
    RefCamera cam;
    size_t n = 0;
};

// Everything that lives on ONE GPU.  dev[0] is the primary device (the one ptb_init chose: all ptb_mem buffers live there);
// dev[1..] exist only in the single-process multi-GPU mode (ptb_set_gpus): they keep their own staged scene + LBVH, scratch and
// counters, read the uploaded scene through peer access when they stage it, and reach the frame buffers of the primary over NVLink.
#define PTB_MAX_GPUS 8
struct DevCtx {
    int device = -1;
    cudaStream_t stream = nullptr;     // secondaries only (the primary launches on State::stream)
    cudaEvent_t ev_done = nullptr;
    unsigned long long *counters = nullptr;
    Staged staged;
    void *wf_blob = nullptr;           // wavefront scratch (ray + path state + queues), grown on demand
    size_t wf_pixels = 0;
    void *pool_blob = nullptr;         // scratch of the persistent pool kernel (pt_pool.cuh): cursor, stack overflow, path state
    size_t pool_bytes = 0;
    // secondaries: local copy of the materials (one 80-byte gather per hit must not cross NVLink)
    void *mats = nullptr;
    size_t mats_bytes = 0;
    ptb_mem mats_handle = 0;
    unsigned long long mats_version = ~0ull;
    // spp split: this device's own RNG states and per-pixel sums
    int *rnds = nullptr;
    float4 *sum = nullptr;
    size_t frame_pixels = 0;
    ptb_mem rnds_handle = 0;
    unsigned long long rnds_version = ~0ull;
    long long sum_index = -1;                          // sample index the sum stands for (-1: not valid)
    unsigned long long sum_colors_version = ~0ull;     // version of the colors buffer the last resolve wrote
};

struct State {
    bool inited = false;
    int device = -1;
    DevCtx dev[PTB_MAX_GPUS];
    int n_gpus = 1, gpu_mode = PTB_SPLIT_TILE, cur = 0;
    cudaEvent_t ev_ready = nullptr;
    cudaDeviceProp prop;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    std::map<ptb_mem, MemObj> mems;
    ptb_mem next = 1;
    Arg args[PTB_KERNEL_COUNT][16];
    std::string err;
    int accel = PTB_ACCEL_AUTO, math = PTB_MATH_FAST, pipe = PTB_PIPE_AUTO;
    int fuse_gen = 1;
    int shard_rank = 0, shard_n = 1, shard_group = 1;
    int sum_mode = 0, count_work = 0;
    int lbvh_leaf = 3;
    int merge_prims = 1;
    int parity_filter = 1;
    int wf_subs = 2;                   // PTB_WF_SUBS / ptb_set_wf_subs: sub-wavefronts of the wavefront pipeline (pt_wavefront.cuh)
    int wf_sort = 0;                   // PTB_WF_SORT=1: wavefront queues grouped by direction octant (measured slower)
    float lbvh_big = 0.25f;            // triangles spanning more than this fraction of the scene stay out of the tree (PTB_LBVH_BIG, 0 = off)
    int tex_nodes = 1;                 // node fetches through the texture pipe (PTB_TEX_NODES=0: LSU)
    int wide_bvh = 1;                  // wavefront traversal over the 4-ary collapse of the LBVH (PTB_WIDE_BVH=0: binary)
    int brute_max = 64;                // AUTO: brute force from shared memory up to this many triangles
    unsigned long long launches = 0;
    PendingGen pending;
    int regen_min = 8;                 // k_render: batch path regeneration until this many lanes of a warp wait (PTB_REGEN_MIN)
    size_t quant_min_bytes = 8u << 20;
    int quant_nodes = 2;               // pool kernel: traverse the 64-byte quantised wide nodes (PTB_QUANT_NODES=0: the 128-byte float nodes)
    int auto_pool = 2;                 // PTB_PIPE_AUTO on LBVH scenes: 0 = the wavefront, 1 = the pool kernel, 2 = by owned pixels (PTB_AUTO_POOL)
    long long auto_pool_pixels = 1500000;   // ... the pool kernel up to this many pixels on this GPU (PTB_AUTO_POOL_PIXELS)
};
// immortal: the reference keeps its Scene (and thus its cl::Buffers) in a global that is destroyed
// during static destruction, after which ptb_release still gets called
State &G = *new State();
inline DevCtx &D() { return G.dev[G.cur]; }                                    // the device the launch path is working on
inline cudaStream_t S() { return G.cur == 0 ? G.stream : G.dev[G.cur].stream; }

int fail(int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (G.err.empty()) {
        G.err = buf;
        fprintf(stderr, "ptb(cuda): %s\n", buf);
    }
    return code;
}
#define CUDA_OK(call)                                                                                   \
    do {                                                                                                \
        cudaError_t e_ = (call);                                                                        \
        if (e_ != cudaSuccess) return fail(PTB_ERR_DEVICE, "%s: %s", #call, cudaGetErrorString(e_));   \
    } while (0)
#define NEED_INIT()                                            \
    do {                                                       \
        if (!G.inited) { int rc_ = ptb_init(-1); if (rc_) return rc_; } \
        if (!G.err.empty()) return PTB_ERR_DEVICE;             \
    } while (0)

MemObj *find(ptb_mem h) {
    auto it = G.mems.find(h);
    return it == G.mems.end() ? nullptr : &it->second;
}

template <typename T> T arg_val(int k, int i, bool *ok) {
    T v;
    memset(&v, 0, sizeof v);
    const Arg &a = G.args[k][i];
    if (a.set && a.value.size() >= sizeof(T)) memcpy(&v, a.value.data(), sizeof(T));
    else if (ok) *ok = false;
    return v;
}
MemObj *arg_mem(int k, int i) { return G.args[k][i].set ? find(G.args[k][i].mem) : nullptr; }

// ---------------------------------------------------------------- staging (AoS upload → SoA planes)
__global__ void k_stage_tris(const RefTriangle *tris, int n, float4 *q0, float4 *q1, float4 *q2,
                             float4 *f0, float4 *f1, float4 *f2, uint16_t *mati, unsigned *max_mati) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 *t = reinterpret_cast<const float4 *>(tris + i);   // 5 x 16 B per record
    float4 r1 = t[0], r2 = t[1], r3 = t[2], N = t[3];
    q0[i] = make_float4(r1.x, r1.y, r1.z, N.x);
    q1[i] = make_float4(r2.x, r2.y, r2.z, N.y);
    q2[i] = make_float4(r3.x, r3.y, r3.z, N.z);
    mati[i] = tris[i].mati;
    if (tris[i].mati > *max_mati) atomicMax(max_mati, (unsigned)tris[i].mati);
    // fast layout: barycentric gradients mu = (e2 x N)/d, mv = (N x e1)/d with d = N.(e1 x e2), in double
    double e1[3] = {(double)r2.x - r1.x, (double)r2.y - r1.y, (double)r2.z - r1.z};
    double e2[3] = {(double)r3.x - r1.x, (double)r3.y - r1.y, (double)r3.z - r1.z};
    double nn[3] = {N.x, N.y, N.z};
    double cx = e1[1] * e2[2] - e1[2] * e2[1], cy = e1[2] * e2[0] - e1[0] * e2[2], cz = e1[0] * e2[1] - e1[1] * e2[0];
    double inv = 1.0 / (nn[0] * cx + nn[1] * cy + nn[2] * cz);
    double mu[3] = {(e2[1] * nn[2] - e2[2] * nn[1]) * inv, (e2[2] * nn[0] - e2[0] * nn[2]) * inv, (e2[0] * nn[1] - e2[1] * nn[0]) * inv};
    double mv[3] = {(nn[1] * e1[2] - nn[2] * e1[1]) * inv, (nn[2] * e1[0] - nn[0] * e1[2]) * inv, (nn[0] * e1[1] - nn[1] * e1[0]) * inv};
    f0[i] = make_float4(N.x, N.y, N.z, r1.x);
    f1[i] = make_float4(r1.y, r1.z, (float)mu[0], (float)mu[1]);
    f2[i] = make_float4((float)mu[2], (float)mv[0], (float)mv[1], (float)mv[2]);
}

// 1 into *flag if any of the n ints is negative
__global__ void k_any_negative(const int *v, size_t n, int *flag) {
    bool neg = false;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) neg |= v[i] < 0;
    if (__any_sync(0xffffffffu, neg) && (threadIdx.x & 31) == 0) atomicOr(flag, 1);
}

// Are the first n RNG states of this buffer all in [0, 2^31)?  True for every seed the reference host uploads and for
// every state the kernels write back, so one scan per host write decides whether launches may use the kernels built on
// pt_rand_nonneg (pt_math.cuh).
int seeds_nonneg(MemObj *rnds, size_t n, bool *out) {
    *out = false;
    // one scan per content version: ptb_write / ptb_device_ptr / ptb_invalidate bump it (the owner of wrapped memory calls
    // ptb_invalidate after writing it behind the library's back); the kernels themselves only ever store states in [0, 2^31)
    if (rnds->seeds_version != rnds->version + 1 || rnds->seeds_checked < n) {
        if (!D().counters) return PTB_SUCCESS;
        int *flag = reinterpret_cast<int *>(D().counters + PT_CNT_WORDS);      // spare word behind the work counters
        CUDA_OK(cudaMemsetAsync(flag, 0, sizeof(int), S()));
        k_any_negative<<<592, 256, 0, S()>>>((const int *)rnds->dptr, n, flag);
        CUDA_OK(cudaGetLastError());
        G.launches++;
        int h = 1;
        CUDA_OK(cudaMemcpyAsync(&h, flag, sizeof h, cudaMemcpyDeviceToHost, S()));
        CUDA_OK(cudaStreamSynchronize(S()));
        rnds->seeds_version = rnds->version + 1;
        rnds->seeds_checked = n;
        rnds->seeds_nonneg = h == 0;
    }
    *out = rnds->seeds_nonneg;
    return PTB_SUCCESS;
}

int stage_scene(MemObj *tris, ptb_mem handle, int n_tris) {
    Staged &s = D().staged;
    if (s.tris_handle == handle && s.tris_version == tris->version && s.n_tris == n_tris) return PTB_SUCCESS;
    if ((size_t)n_tris * sizeof(RefTriangle) > tris->bytes)
        return fail(PTB_ERR_INVALID_ARGS, "trace_ray: tris_size %d exceeds the triangle buffer (%zu bytes)", n_tris, tris->bytes);
    if (n_tris > s.capacity) {
        cudaFree(s.q0); cudaFree(s.q1); cudaFree(s.q2); cudaFree(s.f0); cudaFree(s.f1); cudaFree(s.f2); cudaFree(s.mati);
        s.q0 = s.q1 = s.q2 = s.f0 = s.f1 = s.f2 = nullptr; s.mati = nullptr;    // a failing cudaMalloc below must not leave freed pointers behind
        s.capacity = 0;
        s.tris_handle = 0;
        size_t cap = (size_t)n_tris + 16;
        CUDA_OK(cudaMalloc((void **)&s.q0, cap * sizeof(float4)));
        CUDA_OK(cudaMalloc((void **)&s.q1, cap * sizeof(float4)));
        CUDA_OK(cudaMalloc((void **)&s.q2, cap * sizeof(float4)));
        CUDA_OK(cudaMalloc((void **)&s.f0, cap * sizeof(float4)));
        CUDA_OK(cudaMalloc((void **)&s.f1, cap * sizeof(float4)));
        CUDA_OK(cudaMalloc((void **)&s.f2, cap * sizeof(float4)));
        CUDA_OK(cudaMalloc((void **)&s.mati, cap * sizeof(uint16_t)));
        s.capacity = n_tris;
    }
    s.max_mati = 0;
    if (n_tris > 0) {
        // largest material index of the upload, behind the work counters (next to the seed-scan flag)
        unsigned *max_mati = reinterpret_cast<unsigned *>(D().counters + PT_CNT_WORDS + 1);
        CUDA_OK(cudaMemsetAsync(max_mati, 0, sizeof(unsigned), S()));
        k_stage_tris<<<(n_tris + 255) / 256, 256, 0, S()>>>((const RefTriangle *)tris->dptr, n_tris, s.q0, s.q1, s.q2, s.f0, s.f1, s.f2, s.mati, max_mati);
        CUDA_OK(cudaGetLastError());
        G.launches++;
        CUDA_OK(cudaMemcpyAsync(&s.max_mati, max_mati, sizeof(unsigned), cudaMemcpyDeviceToHost, S()));
        CUDA_OK(cudaStreamSynchronize(S()));
    }
    s.tris_handle = handle;
    s.tris_version = tris->version;
    s.n_tris = n_tris;
    s.lbvh_valid = false;
    s.prims_valid = false;
    return PTB_SUCCESS;
}

// Merge coplanar same-material triangle pairs into convex quads for the fast shared-memory scan (prims.cu).
// Leaves n_prims = 0 when nothing merges, so the launch keeps the plain triangle scan.
int ensure_prims() {
    Staged &s = D().staged;
    if (s.prims_valid) return PTB_SUCCESS;
    s.n_prims = 0;
    s.prims_valid = true;
    if (!G.merge_prims || s.n_tris < 2 || s.n_tris > PRIMS_MAX_TRIS) return PTB_SUCCESS;
    if (!s.p0) {
        CUDA_OK(cudaMalloc((void **)&s.p0, PRIMS_MAX_TRIS * sizeof(float4)));
        CUDA_OK(cudaMalloc((void **)&s.p1, PRIMS_MAX_TRIS * sizeof(float4)));
        CUDA_OK(cudaMalloc((void **)&s.p2, PRIMS_MAX_TRIS * sizeof(float4)));
        CUDA_OK(cudaMalloc((void **)&s.prim_ab, PRIMS_MAX_TRIS * sizeof(float2)));
        CUDA_OK(cudaMalloc((void **)&s.prim_shade, 3 * PRIMS_MAX_TRIS * sizeof(float4)));
        CUDA_OK(cudaMalloc((void **)&s.prim_rep, PRIMS_MAX_TRIS * sizeof(int)));
        CUDA_OK(cudaMalloc((void **)&s.n_prims_dev, sizeof(int)));
    }
    prims_build(s.q0, s.q1, s.q2, s.mati, nullptr, s.n_tris, s.p0, s.p1, s.p2, s.prim_ab, s.prim_rep, s.prim_shade, s.n_prims_dev, S());
    CUDA_OK(cudaGetLastError());
    G.launches++;
    int n = 0;
    CUDA_OK(cudaMemcpyAsync(&n, s.n_prims_dev, sizeof n, cudaMemcpyDeviceToHost, S()));
    CUDA_OK(cudaStreamSynchronize(S()));
    if (n > 0 && n < s.n_tris) s.n_prims = n;
    return PTB_SUCCESS;
}

int resolve_accel(int n_tris, bool have_tree) {
    if (G.accel == PTB_ACCEL_BRUTE) return PT_ACCEL_BRUTE;
    if (G.accel == PTB_ACCEL_REFBVH) return have_tree ? PT_ACCEL_REFBVH : PT_ACCEL_BRUTE;
    if (G.accel == PTB_ACCEL_LBVH) return n_tris > 0 ? PT_ACCEL_LBVH : PT_ACCEL_BRUTE;
    return n_tris <= G.brute_max ? PT_ACCEL_BRUTE : PT_ACCEL_LBVH;
}

int ensure_lbvh(float reach) {
    Staged &s = D().staged;
    char err[256];
    err[0] = 0;
    if (!s.lbvh_valid || s.lbvh_leaf != G.lbvh_leaf) {
        unsigned long long before = s.lbvh.launches;
        if (lbvh_build(&s.lbvh, s.q0, s.q1, s.q2, s.f0, s.f1, s.f2, s.n_tris, G.lbvh_leaf, G.lbvh_big, S(), err, sizeof err))
            return fail(PTB_ERR_DEVICE, "%s", err);
        G.launches += s.lbvh.launches - before + 4;   // + the radix-sort passes
        s.lbvh_valid = true;
        s.lbvh_leaf = G.lbvh_leaf;
        s.n_big_prims = 0;
        if (s.lbvh.n_big >= 1 && G.merge_prims) {      // the outlier list as merged quads for the fast flavour's packed scan
            if (!s.b0) {
                CUDA_OK(cudaMalloc((void **)&s.b0, LBVH_BIG_MAX * sizeof(float4)));
                CUDA_OK(cudaMalloc((void **)&s.b1, LBVH_BIG_MAX * sizeof(float4)));
                CUDA_OK(cudaMalloc((void **)&s.b2, LBVH_BIG_MAX * sizeof(float4)));
                CUDA_OK(cudaMalloc((void **)&s.big_ab, LBVH_BIG_MAX * sizeof(float2)));
                CUDA_OK(cudaMalloc((void **)&s.big_shade, 3 * LBVH_BIG_MAX * sizeof(float4)));
                CUDA_OK(cudaMalloc((void **)&s.big_rep, LBVH_BIG_MAX * sizeof(int)));
            }
            if (!s.n_prims_dev) CUDA_OK(cudaMalloc((void **)&s.n_prims_dev, sizeof(int)));
            const int first = s.lbvh.n_tree;
            prims_build(s.lbvh.sq0 + first, s.lbvh.sq1 + first, s.lbvh.sq2 + first, s.mati, s.lbvh.sorted_id + first, s.lbvh.n_big,
                        s.b0, s.b1, s.b2, s.big_ab, s.big_rep, s.big_shade, s.n_prims_dev, S());
            CUDA_OK(cudaGetLastError());
            G.launches++;
            int nb = 0;
            CUDA_OK(cudaMemcpyAsync(&nb, s.n_prims_dev, sizeof nb, cudaMemcpyDeviceToHost, S()));
            CUDA_OK(cudaStreamSynchronize(S()));
            s.n_big_prims = nb > 0 ? nb : 0;
        }
        if (s.lbvh.max_depth > 62)
            return fail(PTB_ERR_DEVICE, "lbvh: tree depth %d exceeds the traversal stack", s.lbvh.max_depth);
    }
    if (reach > 4.0f * s.lbvh.reach) {   // camera far outside the scene: widen the boxes (cheap re-emit)
        if (lbvh_emit(&s.lbvh, reach, S(), err, sizeof err)) return fail(PTB_ERR_DEVICE, "%s", err);
        G.launches++;
    }
    return PTB_SUCCESS;
}

// Build the launch descriptor from the trace_ray argument table (prog.cl:292-304 ↔ main.cpp:648-660)
int make_params(RenderParams &p, size_t W, size_t H, int spp, bool from_rays) {
    const int K = PTB_KERNEL_TRACE_RAY;
    memset(&p, 0, sizeof p);
    MemObj *tex = arg_mem(K, 0), *tris = arg_mem(K, 1), *mats = arg_mem(K, 3), *tree = arg_mem(K, 4), *shift = arg_mem(K, 5),
           *rays = arg_mem(K, 7), *rnds = arg_mem(K, 8), *colors = arg_mem(K, 12);
    bool ok = true;
    int tris_size = arg_val<int>(K, 2, &ok);
    int shift_size = arg_val<int>(K, 6, &ok);
    p.iterations = arg_val<int>(K, 9, &ok);
    p.sample0 = arg_val<int>(K, 10, &ok);
    p.cam = arg_val<RefCamera>(K, 11, &ok);
    if (!ok) return fail(PTB_ERR_INVALID_ARGS, "trace_ray: a by-value argument (2, 6, 9, 10 or 11) is not set");
    if (!tris || !mats || !rnds || !colors) return fail(PTB_ERR_INVALID_ARGS, "trace_ray: buffer argument 1, 3, 8 or 12 is not set");
    if (from_rays && !rays) return fail(PTB_ERR_INVALID_ARGS, "trace_ray: rays (argument 7) is not set");
    size_t npix = W * H;
    if (npix == 0 || W > 65535u * 16u || H > 65535u * 8u) return fail(PTB_ERR_INVALID_VALUE, "trace_ray: bad launch size %zux%zu", W, H);
    if (rnds->bytes < npix * 4 || colors->bytes < npix * 16 || (from_rays && rays->bytes < npix * 32))
        return fail(PTB_ERR_INVALID_ARGS, "trace_ray: rnds/colors/rays smaller than the %zux%zu launch", W, H);
    if (tex && tex->bytes < npix * 16) tex = nullptr;
    if (tris_size < 0) tris_size = 0;
    int rc = stage_scene(tris, G.args[K][1].mem, tris_size);
    if (rc) return rc;
    Staged &s = D().staged;
    bool have_tree = tree && shift && shift_size > 0 && shift->bytes >= (size_t)shift_size * 4;
    p.accel = resolve_accel(tris_size, have_tree);
    const bool fastm = G.math == PTB_MATH_FAST;   // each arithmetic flavour reads its own triangle layout
    p.sc.q0 = fastm ? s.f0 : s.q0; p.sc.q1 = fastm ? s.f1 : s.q1; p.sc.q2 = fastm ? s.f2 : s.q2; p.sc.mati = s.mati; p.sc.n_tris = tris_size;
    p.sc.fq0 = s.f0; p.sc.fq1 = s.f1; p.sc.fq2 = s.f2;
    p.parity_filter = G.parity_filter;
    p.regen_min = G.regen_min;
    p.wf_sort = G.wf_sort;
    p.wf_subs = G.wf_subs;
    if (p.accel == PT_ACCEL_BRUTE && fastm) {
        rc = ensure_prims();
        if (rc) return rc;
        p.sc.pq0 = s.p0; p.sc.pq1 = s.p1; p.sc.pq2 = s.p2; p.sc.prim_ab = s.prim_ab; p.sc.prim_shade = s.prim_shade; p.sc.prim_rep = s.prim_rep; p.sc.n_prims = s.n_prims;
    }
    p.sc.mats = (const RefMaterial *)mats->dptr;
    p.sc.n_mats = (int)(mats->bytes / sizeof(RefMaterial));
    // prog.cl:322 indexes materials[] unchecked; here a record pointing past the buffer is an error, not a stray device read
    if (tris_size > 0 && (size_t)s.max_mati >= mats->bytes / sizeof(RefMaterial))
        return fail(PTB_ERR_INVALID_ARGS, "trace_ray: a triangle uses material %u but the materials buffer holds %zu", s.max_mati, mats->bytes / sizeof(RefMaterial));
    if (have_tree) {
        p.sc.kd_tree = (const RefNode *)tree->dptr;
        p.sc.kd_shift = (const int *)shift->dptr;
        p.sc.n_obj = shift_size;
    }
    if (p.accel == PT_ACCEL_LBVH) {
        float reach = fmaxf(fabsf(p.cam.eye.x), fmaxf(fabsf(p.cam.eye.y), fabsf(p.cam.eye.z)));
        rc = ensure_lbvh(reach);
        if (rc) return rc;
        p.sc.bvh = s.lbvh.bvh;
        p.sc.bvh_tex = G.tex_nodes ? s.lbvh.bvh_tex : 0;
        p.sc.bvh4_tex = G.tex_nodes ? s.lbvh.bvh4_tex : 0;
        // the 4-ary traversal pushes up to 3 entries per level: usable when the stack bound (64) covers the tree
        const int wide_levels = (s.lbvh.max_depth + 1) / 2;
        p.sc.bvh4 = (G.wide_bvh && s.lbvh.n_nodes4 > 0 && 3 * wide_levels + 1 <= 64) ? s.lbvh.bvh4 : nullptr;
        // node format (pool kernel; the wavefront keeps the float nodes unless forced): the 64-byte quantised nodes pay when the float nodes (128 bytes each) do not sit comfortably in L1 / L2 next to
        // the triangles and the path state (1 M triangles: 30 MB of nodes, +20 %); on small trees their slightly wider boxes cost more
        // than the bytes return (104 k triangles: 3 MB, -5 %) — profiles/r02_quant_nodes.txt.  2 = by size, 1 = always, 0 = never
        const bool quant = G.quant_nodes == 1 || (G.quant_nodes == 2 && (size_t)s.lbvh.n_nodes4 * 128 > G.quant_min_bytes);
        p.sc.bvh4q = (p.sc.bvh4 && quant) ? s.lbvh.bvh4q : nullptr;
        p.sc.quant_wf = G.quant_nodes == 1;
        p.sc.bvh4q_tex = G.tex_nodes ? s.lbvh.bvh4q_tex : 0;
        p.sc.n_nodes4 = s.lbvh.n_nodes4;
        p.sc.srec = fastm ? s.lbvh.rf : s.lbvh.rq;
        p.sc.sq0 = fastm ? s.lbvh.sf0 : s.lbvh.sq0; p.sc.sq1 = fastm ? s.lbvh.sf1 : s.lbvh.sq1; p.sc.sq2 = fastm ? s.lbvh.sf2 : s.lbvh.sq2;
        p.sc.sorted_id = s.lbvh.sorted_id; p.sc.bvh_root = s.lbvh.root;
        p.sc.big_first = s.lbvh.n_tree; p.sc.big_count = s.lbvh.n_big;
        if (fastm && s.n_big_prims > 0) {
            p.sc.bq0 = s.b0; p.sc.bq1 = s.b1; p.sc.bq2 = s.b2; p.sc.big_ab = s.big_ab; p.sc.big_rep = s.big_rep; p.sc.n_big_prims = s.n_big_prims;
        }
    }
    p.rnds = (int *)rnds->dptr;
    p.colors = (float4 *)colors->dptr;
    p.tex = tex ? (float4 *)tex->dptr : nullptr;
    p.rays = rays ? (RefRay *)rays->dptr : nullptr;
    p.W = (int)W; p.H = (int)H;
    p.spp = spp;
    p.from_rays = from_rays ? 1 : 0;
    p.shard_rank = G.shard_rank; p.shard_n = G.shard_n; p.shard_group = G.shard_group;
    p.sum_mode = G.sum_mode;
    p.count_work = G.count_work;
    p.counters = D().counters;
    if (fastm && !from_rays) {                       // (a two-kernel host may hand trace_ray states gen_ray never saw)
        bool ok_seeds = false;
        rc = seeds_nonneg(rnds, npix, &ok_seeds);
        if (rc) return rc;
        p.rng_nonneg = ok_seeds ? 1 : 0;
    }
    pt_cam_affine(p);
    return PTB_SUCCESS;
}

int run_gen_ray_one(ptb_mem rays_h, ptb_mem rnds_h, const RefCamera &cam, size_t n, int shard_rank, int shard_n, int shard_group) {
    MemObj *rays = find(rays_h), *rnds = find(rnds_h);
    if (!rays || !rnds) return fail(PTB_ERR_INVALID_ARGS, "gen_ray: buffer argument 0 or 2 is not set");
    if (rays->bytes < n * 32 || rnds->bytes < n * 4) return fail(PTB_ERR_INVALID_ARGS, "gen_ray: rays/rnds smaller than the launch (%zu)", n);
    RenderParams p;
    memset(&p, 0, sizeof p);
    p.cam = cam;
    pt_cam_affine(p);
    p.rays = (RefRay *)rays->dptr;
    p.rnds = (int *)rnds->dptr;
    p.shard_rank = shard_rank; p.shard_n = shard_n; p.shard_group = shard_group;
    if (G.math == PTB_MATH_PARITY) pt_parity::launch_gen_ray(p, n, S());
    else pt_fast::launch_gen_ray(p, n, S());
    CUDA_OK(cudaGetLastError());
    G.launches++;
    return PTB_SUCCESS;
}

// switch the launch path to device index i of the single-process multi-GPU mode (and back when the scope ends)
struct DevScope {
    int prev;
    explicit DevScope(int i) : prev(G.cur) { G.cur = i; cudaSetDevice(G.dev[i].device); }
    ~DevScope() { G.cur = prev; cudaSetDevice(G.dev[prev].device); }
};
int multi_fork() {          // the secondaries' streams continue after everything enqueued on the primary's so far
    if (!G.ev_ready) CUDA_OK(cudaEventCreateWithFlags(&G.ev_ready, cudaEventDisableTiming));
    CUDA_OK(cudaEventRecord(G.ev_ready, G.stream));
    for (int d = 1; d < G.n_gpus; ++d) CUDA_OK(cudaStreamWaitEvent(G.dev[d].stream, G.ev_ready, 0));
    return PTB_SUCCESS;
}
int multi_join() {          // ... and the primary's stream after everything the secondaries were given
    for (int d = 1; d < G.n_gpus; ++d) {
        DevScope scope(d);
        CUDA_OK(cudaEventRecord(D().ev_done, D().stream));
    }
    for (int d = 1; d < G.n_gpus; ++d) CUDA_OK(cudaStreamWaitEvent(G.stream, G.dev[d].ev_done, 0));
    return PTB_SUCCESS;
}
int run_gen_ray(ptb_mem rays_h, ptb_mem rnds_h, const RefCamera &cam, size_t n) {
    if (G.n_gpus <= 1) return run_gen_ray_one(rays_h, rnds_h, cam, n, G.shard_rank, G.shard_n, G.shard_group);
    if (G.gpu_mode != PTB_SPLIT_TILE) return fail(PTB_ERR_INVALID_VALUE, "gen_ray as its own launch needs the tile split (the spp split renders through the fused launch or ptb_render)");
    int rc = multi_fork();
    const int group = 8 * (int)cam.XM > 0 ? 8 * (int)cam.XM : 8;
    for (int d = 0; d < G.n_gpus && !rc; ++d) {
        DevScope scope(d);
        rc = run_gen_ray_one(rays_h, rnds_h, cam, n, d, G.n_gpus, group);
    }
    return rc ? rc : multi_join();
}

int flush_pending() {
    if (!G.pending.active) return PTB_SUCCESS;
    G.pending.active = false;
    return run_gen_ray(G.pending.rays, G.pending.rnds, G.pending.cam, G.pending.n);
}

int run_render(const RenderParams &p) {
    if (G.math == PTB_MATH_PARITY) pt_parity::launch_render(p, S());
    else pt_fast::launch_render(p, S());
    CUDA_OK(cudaGetLastError());
    G.launches++;
    return PTB_SUCCESS;
}

// wavefront pipeline: 136 bytes of scratch per pixel (ray 32, path state 96, two queue slots 8); each queue carries
// WF_QUEUE_SLACK extra bytes because its sub-wavefronts' shares are rounded up to whole blocks of 256 pixel ids
#define WF_QUEUE_SLACK 8192
int run_wavefront(RenderParams &p) {
    const size_t n = (size_t)p.W * p.H;
    if (n > D().wf_pixels) {
        cudaFree(D().wf_blob);
        D().wf_blob = nullptr; D().wf_pixels = 0;
        CUDA_OK(cudaMalloc(&D().wf_blob, n * 136 + 2 * WF_QUEUE_SLACK + 256));
        D().wf_pixels = n;
    }
    char *b = (char *)D().wf_blob;
    p.wf_ray0 = (float4 *)b; b += n * 16;
    p.wf_ray1 = (float4 *)b; b += n * 16;
    p.wf_state = (float4 *)b; b += n * 96;
    p.wf_queue[0] = (int *)b; b += n * 4 + WF_QUEUE_SLACK;
    p.wf_queue[1] = (int *)b; b += n * 4 + WF_QUEUE_SLACK;
    p.wf_ctrl = (int *)b;
    int k = G.math == PTB_MATH_PARITY ? pt_parity::launch_wavefront(p, S()) : pt_fast::launch_wavefront(p, S());
    CUDA_OK(cudaGetLastError());
    if (k < 0) return fail(PTB_ERR_DEVICE, "wavefront: a CUDA call of the iteration loop failed");
    G.launches += (unsigned long long)k;
    return PTB_SUCCESS;
}

// persistent pool kernel (pt_pool.cuh): does this launch go through it?
bool want_pool(const RenderParams &p) {
    static const bool disabled = getenv("PTB_DISABLE_POOL") && atoi(getenv("PTB_DISABLE_POOL"));     // experiments / bisecting only
    if (disabled || p.from_rays || p.accel != PT_ACCEL_LBVH) return false;
    if (G.pipe != PTB_PIPE_POOL) {
        if (G.pipe != PTB_PIPE_AUTO || G.auto_pool == 0) return false;
        // AUTO by measurement (profiles/r02_pool_shards.txt): the wavefront wins on big frames (all paths of the frame in flight,
        // coherent first bounces), the pool kernel when this GPU owns few pixels (no per-iteration tail: 1/8 of a 4K frame and below) and
        // whenever several GPUs render from one process (no host loop to keep them concurrent)
        const long long owned = (long long)p.W * p.H / (p.shard_n > 1 ? p.shard_n : 1);
        if (G.auto_pool == 2 && G.n_gpus == 1 && owned > G.auto_pool_pixels) return false;
    }
    return G.math == PTB_MATH_PARITY ? pt_parity::pool_usable(p) : pt_fast::pool_usable(p);
}
int ensure_pool_blob() {
    const size_t need = G.math == PTB_MATH_PARITY ? pt_parity::pool_scratch_bytes() : pt_fast::pool_scratch_bytes();
    if (need > D().pool_bytes) {
        cudaFree(D().pool_blob);
        D().pool_blob = nullptr; D().pool_bytes = 0;
        CUDA_OK(cudaMalloc(&D().pool_blob, need));
        D().pool_bytes = need;
    }
    return PTB_SUCCESS;
}
int run_pool(const RenderParams &p) {
    int rc = ensure_pool_blob();
    if (rc) return rc;
    const int k = G.math == PTB_MATH_PARITY ? pt_parity::launch_pool(p, D().pool_blob, S()) : pt_fast::launch_pool(p, D().pool_blob, S());
    CUDA_OK(cudaGetLastError());
    G.launches += (unsigned long long)k;
    return PTB_SUCCESS;
}

// which kernels render this launch
int run_any(RenderParams &p, bool allow_wavefront) {
    if (want_pool(p)) return run_pool(p);
    if (!p.from_rays && p.spp >= 1 && allow_wavefront && G.pipe != PTB_PIPE_MEGAKERNEL && G.pipe != PTB_PIPE_POOL && p.accel == PT_ACCEL_LBVH) return run_wavefront(p);
    return run_render(p);
}

// seeds of GPU d > 0 under the spp split: state_d = state_0 * 16807^(d * W * H) mod (2^31 - 1) — for a freshly seeded buffer that is
// the host generator's sequence continued (pixel i of GPU d gets output i + d*W*H of std::minstd_rand0, main.cpp:522-526 carried on)
__global__ void k_derive_seeds(const int *src, int *dst, size_t n, unsigned long long mult) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned long long s = (unsigned long long)(unsigned)src[i] % 2147483647ull;
    if (s == 0) s = 1;
    dst[i] = (int)((s * mult) % 2147483647ull);
}
unsigned long long powmod(unsigned long long b, unsigned long long e) {
    const unsigned long long M = 2147483647ull;
    unsigned long long r = 1;
    b %= M;
    while (e) { if (e & 1) r = r * b % M; b = b * b % M; e >>= 1; }
    return r;
}

// ONE render call on every GPU of the single-process multi-GPU mode (ptb_set_gpus).
//   tile: GPU d renders the bands of 8 scanlines d, d+n, ... with the reference's seeds; rnds / colors / tex are the PRIMARY's buffers,
//         reached by the other GPUs' kernels over NVLink (peer access): the gather rides on the kernels' own end-of-pixel stores.
//         Bitwise the 1-GPU image.
//   spp:  sample j of the call goes to GPU (sample0 + j) % n; every GPU keeps its own RNG states (derived, see above) and a
//         per-pixel SUM on its own HBM; afterwards ONE kernel on the primary adds the n sums through peer loads, scales by
//         1 / total samples and writes the mean into colors and the tone-mapped image (the fused root resolve, no NCCL).
int render_multi(size_t W, size_t H, int spp, bool from_rays) {
    const int K = PTB_KERNEL_TRACE_RAY, n = G.n_gpus;
    const bool spp_mode = G.gpu_mode == PTB_SPLIT_SPP;
    if (G.shard_n > 1) return fail(PTB_ERR_INVALID_VALUE, "ptb_set_pixel_shard and ptb_set_gpus do not combine");
    if (spp_mode && (from_rays || G.sum_mode)) return fail(PTB_ERR_INVALID_VALUE, "the spp split needs the fused launch (or ptb_render) and the running-mean accumulation");
    bool ok = true;
    const int sample0 = arg_val<int>(K, 10, &ok);
    MemObj *rnds = arg_mem(K, 8), *colors = arg_mem(K, 12), *tex = arg_mem(K, 0);
    if (!ok || !rnds || !colors) return fail(PTB_ERR_INVALID_ARGS, "trace_ray: arguments 8, 10, 12 are not set");
    const size_t npix = W * H;
    int rc = multi_fork();
    if (rc) return rc;
    for (int d = 0; d < n; ++d) {
        DevScope scope(d);
        DevCtx &c = D();
        int spp_d = spp;
        if (spp_mode) {
            const int first = (((d - sample0) % n) + n) % n;          // first j with (sample0 + j) % n == d
            spp_d = first < spp ? (spp - first + n - 1) / n : 0;
        }
        RenderParams p;
        rc = make_params(p, W, H, spp_d > 0 ? spp_d : 1, from_rays);
        if (rc) return rc;
        if (d > 0) {                                                  // materials: a local copy (an 80-byte gather per hit must not cross NVLink)
            MemObj *mats = arg_mem(K, 3);
            if (c.mats_handle != G.args[K][3].mem || c.mats_version != mats->version || c.mats_bytes < mats->bytes) {
                if (c.mats_bytes < mats->bytes) {
                    cudaFree(c.mats); c.mats = nullptr; c.mats_bytes = 0;
                    CUDA_OK(cudaMalloc(&c.mats, mats->bytes ? mats->bytes : 1));
                    c.mats_bytes = mats->bytes;
                }
                CUDA_OK(cudaMemcpyPeerAsync(c.mats, c.device, mats->dptr, G.device, mats->bytes, S()));
                c.mats_handle = G.args[K][3].mem; c.mats_version = mats->version;
            }
            p.sc.mats = (const RefMaterial *)c.mats;
        }
        if (spp_mode) {
            if (c.frame_pixels < npix) {
                cudaFree(c.rnds); cudaFree(c.sum);
                c.rnds = nullptr; c.sum = nullptr; c.frame_pixels = 0; c.rnds_version = ~0ull; c.sum_index = -1;
                if (d > 0) CUDA_OK(cudaMalloc((void **)&c.rnds, npix * sizeof(int)));
                CUDA_OK(cudaMalloc((void **)&c.sum, npix * sizeof(float4)));
                c.frame_pixels = npix;
            }
            if (d > 0 && (c.rnds_handle != G.args[K][8].mem || c.rnds_version != rnds->version)) {
                k_derive_seeds<<<(unsigned)((npix + 255) / 256), 256, 0, S()>>>((const int *)rnds->dptr, c.rnds, npix, powmod(16807ull, (unsigned long long)d * npix));
                CUDA_OK(cudaGetLastError());
                G.launches++;
                c.rnds_handle = G.args[K][8].mem; c.rnds_version = rnds->version;
            }
            // the sums are a cache of "colors x samples so far": rebuilt when the host moved the sample index or rewrote colors (checkpoint resume)
            if (sample0 == 0 || c.sum_index != sample0 || c.sum_colors_version != colors->version) {
                if (sample0 != 0 && d == 0) {
                    ResolveSrc src{};
                    src.p[0] = (const float4 *)colors->dptr; src.n = 1;
                    resolve_frame(src, (float)sample0, (int)npix, 1, 0, c.sum, nullptr, nullptr, S());
                    G.launches++;
                } else CUDA_OK(cudaMemsetAsync(c.sum, 0, npix * sizeof(float4), S()));
            }
            c.sum_index = sample0 + spp;
            if (d > 0) p.rnds = c.rnds;
            p.colors = c.sum; p.sum_mode = 1; p.tex = nullptr;
        } else {
            p.shard_rank = d; p.shard_n = n; p.shard_group = 8 * (int)W;
        }
        if (spp_d > 0) {
            rc = run_any(p, false);                                    // never the host-looped wavefront here: the GPUs must run concurrently
            if (rc) return rc;
        }
    }
    rc = multi_join();
    if (rc) return rc;
    if (spp_mode) {
        ResolveSrc src{};
        for (int d = 0; d < n; ++d) src.p[d] = G.dev[d].sum;
        src.n = n;
        const bool tex_ok = tex && tex->bytes >= npix * 16;
        resolve_frame(src, 1.0f / (float)(sample0 + spp), (int)W, (int)H, 0, (float4 *)colors->dptr, tex_ok ? (float4 *)tex->dptr : nullptr, nullptr, G.stream);
        CUDA_OK(cudaGetLastError());
        G.launches++;
        for (int d = 0; d < n; ++d) G.dev[d].sum_colors_version = colors->version;
    }
    return PTB_SUCCESS;
}

void free_dev(DevCtx &c) {
    Staged &s = c.staged;
    cudaFree(s.q0); cudaFree(s.q1); cudaFree(s.q2); cudaFree(s.f0); cudaFree(s.f1); cudaFree(s.f2); cudaFree(s.mati);
    cudaFree(s.p0); cudaFree(s.p1); cudaFree(s.p2); cudaFree(s.prim_ab); cudaFree(s.prim_shade); cudaFree(s.prim_rep); cudaFree(s.n_prims_dev);
    cudaFree(s.b0); cudaFree(s.b1); cudaFree(s.b2); cudaFree(s.big_ab); cudaFree(s.big_shade); cudaFree(s.big_rep);
    lbvh_free(&s.lbvh);
    s = Staged();
    cudaFree(c.counters); c.counters = nullptr;
    cudaFree(c.wf_blob); c.wf_blob = nullptr; c.wf_pixels = 0;
    cudaFree(c.pool_blob); c.pool_blob = nullptr; c.pool_bytes = 0;
    cudaFree(c.mats); c.mats = nullptr; c.mats_bytes = 0; c.mats_handle = 0;
    cudaFree(c.rnds); cudaFree(c.sum); c.rnds = nullptr; c.sum = nullptr; c.frame_pixels = 0; c.rnds_handle = 0; c.sum_index = -1;
    if (c.stream) cudaStreamDestroy(c.stream);
    if (c.ev_done) cudaEventDestroy(c.ev_done);
    c.stream = nullptr; c.ev_done = nullptr;
}

}  // namespace

extern "C" {

// ---------------------------------------------------------------- bring-up
int ptb_init(int device_ordinal) {
    if (G.inited) return G.err.empty() ? PTB_SUCCESS : PTB_ERR_DEVICE;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(PTB_ERR_NO_DEVICE, "no CUDA device (%s); this library has no CPU fallback", e == cudaSuccess ? "count = 0" : cudaGetErrorString(e));
    int dev = device_ordinal;
    if (dev < 0) {
        const char *lr = getenv("LOCAL_RANK");
        dev = lr ? atoi(lr) % count : 0;
    }
    if (dev >= count) return fail(PTB_ERR_NO_DEVICE, "device %d requested, %d present", dev, count);
    CUDA_OK(cudaSetDevice(dev));
    CUDA_OK(cudaGetDeviceProperties(&G.prop, dev));
    CUDA_OK(cudaStreamCreateWithFlags(&G.own_stream, cudaStreamNonBlocking));
    G.stream = G.own_stream;
    CUDA_OK(cudaMalloc((void **)&D().counters, (PT_CNT_WORDS + 2) * sizeof(unsigned long long)));
    CUDA_OK(cudaMemset(D().counters, 0, (PT_CNT_WORDS + 2) * sizeof(unsigned long long)));
    G.device = dev;
    G.dev[0].device = dev;
    G.cur = 0; G.n_gpus = 1;
    G.inited = true;
    const char *env;
    if ((env = getenv("PTB_MATH"))) G.math = atoi(env) ? PTB_MATH_FAST : PTB_MATH_PARITY;
    if ((env = getenv("PTB_ACCEL"))) G.accel = atoi(env);
    if ((env = getenv("PTB_LBVH_LEAF"))) G.lbvh_leaf = atoi(env);
    if ((env = getenv("PTB_BRUTE_MAX"))) G.brute_max = atoi(env);
    if ((env = getenv("PTB_MERGE_PRIMS"))) G.merge_prims = atoi(env) ? 1 : 0;
    if ((env = getenv("PTB_PARITY_FILTER"))) G.parity_filter = atoi(env) ? 1 : 0;
    if ((env = getenv("PTB_WF_SORT"))) G.wf_sort = atoi(env) ? 1 : 0;
    if ((env = getenv("PTB_WF_SUBS"))) G.wf_subs = atoi(env) < 1 ? 1 : (atoi(env) > 4 ? 4 : atoi(env));
    if ((env = getenv("PTB_WIDE_BVH"))) G.wide_bvh = atoi(env) ? 1 : 0;
    if ((env = getenv("PTB_TEX_NODES"))) G.tex_nodes = atoi(env) ? 1 : 0;
    if ((env = getenv("PTB_LBVH_BIG"))) G.lbvh_big = (float)atof(env);
    if ((env = getenv("PTB_AUTO_POOL"))) G.auto_pool = atoi(env) < 0 ? 0 : (atoi(env) > 2 ? 2 : atoi(env));
    if ((env = getenv("PTB_AUTO_POOL_PIXELS"))) G.auto_pool_pixels = atoll(env);
    if ((env = getenv("PTB_QUANT_NODES"))) G.quant_nodes = atoi(env) < 0 ? 0 : (atoi(env) > 2 ? 2 : atoi(env));
    if ((env = getenv("PTB_REGEN_MIN"))) G.regen_min = atoi(env) < 1 ? 1 : (atoi(env) > 32 ? 32 : atoi(env));
    return PTB_SUCCESS;
}

int ptb_shutdown(void) {
    if (!G.inited) return PTB_SUCCESS;
    cudaSetDevice(G.device);
    cudaStreamSynchronize(S());
    comm_destroy();
    for (auto &kv : G.mems) {
        if (kv.second.owned) cudaFree(kv.second.dptr);
        else if (kv.second.ipc) cudaIpcCloseMemHandle(kv.second.dptr);
    }
    G.mems.clear();
    for (int d = G.n_gpus - 1; d >= 0; --d) {
        if (G.dev[d].device < 0) continue;
        cudaSetDevice(G.dev[d].device);
        if (d > 0 && G.dev[d].stream) cudaStreamSynchronize(G.dev[d].stream);
        free_dev(G.dev[d]);
        if (d > 0) G.dev[d].device = -1;
    }
    for (int d = 1; d < PTB_MAX_GPUS; ++d) G.dev[d] = DevCtx();
    G.n_gpus = 1; G.cur = 0;
    cudaSetDevice(G.device);
    if (G.ev_ready) { cudaEventDestroy(G.ev_ready); G.ev_ready = nullptr; }
    cudaStreamDestroy(G.own_stream);
    G.own_stream = G.stream = nullptr;
    for (auto &k : G.args) for (auto &a : k) a = Arg();
    G.pending.active = false;
    G.inited = false;
    return PTB_SUCCESS;
}

int ptb_device_string(int what, char *dst, size_t cap) {
    if (!dst || !cap) return PTB_ERR_INVALID_VALUE;
    NEED_INIT();
    int drv = 0, rt = 0;
    cudaDriverGetVersion(&drv);
    cudaRuntimeGetVersion(&rt);
    switch (what) {
        case 0: snprintf(dst, cap, "%s", G.prop.name); break;
        case 1: snprintf(dst, cap, "CUDA sm_%d%d, %d SMs, %.1f GB", G.prop.major, G.prop.minor, G.prop.multiProcessorCount, G.prop.totalGlobalMem / 1e9); break;
        case 2: snprintf(dst, cap, "driver %d.%d / runtime %d.%d", drv / 1000, (drv % 100) / 10, rt / 1000, (rt % 100) / 10); break;
        case 3: snprintf(dst, cap, "prebuilt CUDA kernels for sm_100a (prog.cl is not JIT-compiled)"); break;
        default: snprintf(dst, cap, "ptb_render ptb_lbvh ptb_pixel_shard ptb_sum_accumulate"); break;
    }
    return PTB_SUCCESS;
}
int ptb_device_compute_units(void) {
    if (!G.inited && ptb_init(-1)) return 0;
    return G.prop.multiProcessorCount;
}
const char *ptb_backend_name(void) { return "cuda-sm100a"; }

// ---------------------------------------------------------------- memory
ptb_mem ptb_alloc(size_t bytes) {
    if (!G.inited && ptb_init(-1)) return 0;
    if (!G.err.empty()) return 0;
    MemObj m;
    m.bytes = bytes;
    cudaError_t e = cudaMalloc(&m.dptr, bytes ? bytes : 1);
    if (e != cudaSuccess) { fail(PTB_ERR_OUT_OF_MEMORY, "cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e)); return 0; }
    cudaMemsetAsync(m.dptr, 0, bytes, S());
    ptb_mem h = G.next++;
    G.mems[h] = m;
    return h;
}
ptb_mem ptb_alloc_image(int w, int h) {
    if (w <= 0 || h <= 0) { fail(PTB_ERR_INVALID_VALUE, "ptb_alloc_image: bad extent %dx%d", w, h); return 0; }
    ptb_mem hd = ptb_alloc((size_t)w * (size_t)h * sizeof(float4));
    MemObj *m = find(hd);
    if (m) { m->image = true; m->w = w; m->h = h; }
    return hd;
}
int ptb_retain(ptb_mem h) {
    MemObj *m = find(h);
    if (!m) return PTB_ERR_INVALID_MEM;
    m->refs++;
    return PTB_SUCCESS;
}
int ptb_release(ptb_mem h) {
    MemObj *m = find(h);
    if (!m) return PTB_ERR_INVALID_MEM;
    if (--m->refs == 0) {
        if (G.inited) {
            if (G.pending.active && (G.pending.rays == h || G.pending.rnds == h)) G.pending.active = false;
            if (m->owned) { cudaStreamSynchronize(S()); cudaFree(m->dptr); }
            else if (m->ipc) { cudaStreamSynchronize(S()); cudaIpcCloseMemHandle(m->dptr); }
        }
        for (auto &c : G.dev) {
            if (c.staged.tris_handle == h) c.staged.tris_handle = 0;
            if (c.mats_handle == h) c.mats_handle = 0;
            if (c.rnds_handle == h) c.rnds_handle = 0;
        }
        G.mems.erase(h);
    }
    return PTB_SUCCESS;
}
int ptb_write(ptb_mem h, size_t off, size_t bytes, const void *src) {
    NEED_INIT();
    MemObj *m = find(h);
    if (!m) return fail(PTB_ERR_INVALID_MEM, "ptb_write: bad handle");
    if (bytes == 0) return PTB_SUCCESS;
    if (!src || off > m->bytes || bytes > m->bytes - off) return fail(PTB_ERR_INVALID_VALUE, "ptb_write: range [%zu,%zu) outside a %zu-byte buffer", off, off + bytes, m->bytes);
    int rc = flush_pending();
    if (rc) return rc;
    CUDA_OK(cudaMemcpyAsync((char *)m->dptr + off, src, bytes, cudaMemcpyHostToDevice, S()));
    CUDA_OK(cudaStreamSynchronize(S()));
    m->version++;
    return PTB_SUCCESS;
}
int ptb_read(ptb_mem h, size_t off, size_t bytes, void *dst) {
    NEED_INIT();
    MemObj *m = find(h);
    if (!m) return fail(PTB_ERR_INVALID_MEM, "ptb_read: bad handle");
    if (bytes == 0) return PTB_SUCCESS;
    if (!dst || off > m->bytes || bytes > m->bytes - off) return fail(PTB_ERR_INVALID_VALUE, "ptb_read: range [%zu,%zu) outside a %zu-byte buffer", off, off + bytes, m->bytes);
    int rc = flush_pending();
    if (rc) return rc;
    CUDA_OK(cudaMemcpyAsync(dst, (const char *)m->dptr + off, bytes, cudaMemcpyDeviceToHost, S()));
    CUDA_OK(cudaStreamSynchronize(S()));
    return PTB_SUCCESS;
}
int ptb_ingest_mesh(ptb_mem tris_h, size_t first_tri, const float *vertices, size_t n_verts, const int *indices,
                    const int *mat_ids, size_t n_faces, const float *pos, const float *scale, float pitch_deg,
                    float yaw_deg, int mat_offset) {
    NEED_INIT();
    MemObj *m = find(tris_h);
    if (!m) return fail(PTB_ERR_INVALID_MEM, "ptb_ingest_mesh: bad handle");
    if (n_faces == 0) return PTB_SUCCESS;
    if (!vertices || !indices || !mat_ids || !pos || !scale) return fail(PTB_ERR_INVALID_VALUE, "ptb_ingest_mesh: null argument");
    const size_t cap_tris = m->bytes / sizeof(RefTriangle);
    if (first_tri > cap_tris || n_faces > cap_tris - first_tri)
        return fail(PTB_ERR_INVALID_VALUE, "ptb_ingest_mesh: triangles [%zu,%zu) outside a %zu-byte buffer", first_tri, first_tri + n_faces, m->bytes);
    for (size_t i = 0; i < 3 * n_faces; ++i)
        if (indices[i] < 0 || (size_t)indices[i] >= n_verts)
            return fail(PTB_ERR_INVALID_VALUE, "ptb_ingest_mesh: vertex index %d of face %zu outside [0,%zu)", indices[i], i / 3, n_verts);
    int rc = flush_pending();
    if (rc) return rc;
    IngestXform x;
    // main.cpp:56,64: the angle is narrowed to float before cos/sin (double functions of <math.h>)
    const float pitch = pitch_deg / 180.0f * 3.141593f, yaw = yaw_deg / 180.0f * 3.141593f;
    x.cos_pitch = cos((double)pitch); x.sin_pitch = sin((double)pitch);
    x.cos_yaw = cos((double)yaw); x.sin_yaw = sin((double)yaw);
    for (int k = 0; k < 3; ++k) { x.scale[k] = scale[k]; x.pos[k] = pos[k]; }
    x.mat_offset = mat_offset;
    float *dv = nullptr;
    int *di = nullptr, *dm = nullptr;
    const size_t vb = n_verts * 3 * sizeof(float), ib = n_faces * 3 * sizeof(int), mb = n_faces * sizeof(int);
    cudaError_t e = cudaMalloc((void **)&dv, vb);
    if (e == cudaSuccess) e = cudaMalloc((void **)&di, ib);
    if (e == cudaSuccess) e = cudaMalloc((void **)&dm, mb);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dv, vertices, vb, cudaMemcpyHostToDevice, S());
    if (e == cudaSuccess) e = cudaMemcpyAsync(di, indices, ib, cudaMemcpyHostToDevice, S());
    if (e == cudaSuccess) e = cudaMemcpyAsync(dm, mat_ids, mb, cudaMemcpyHostToDevice, S());
    if (e == cudaSuccess) {
        ingest_launch(dv, di, dm, n_faces, x, (RefTriangle *)m->dptr + first_tri, S());
        e = cudaGetLastError();
        G.launches++;
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(S());
    cudaFree(dv); cudaFree(di); cudaFree(dm);
    if (e != cudaSuccess) return fail(PTB_ERR_DEVICE, "ptb_ingest_mesh: %s", cudaGetErrorString(e));
    m->version++;
    return PTB_SUCCESS;
}
size_t ptb_mem_size(ptb_mem h) {
    MemObj *m = find(h);
    return m ? m->bytes : 0;
}
ptb_mem ptb_wrap_device(void *device_ptr, size_t bytes) {
    if (!G.inited && ptb_init(-1)) return 0;
    if (!device_ptr) return 0;
    MemObj m;
    m.dptr = device_ptr;
    m.bytes = bytes;
    m.owned = false;
    ptb_mem h = G.next++;
    G.mems[h] = m;
    return h;
}
int ptb_ipc_export(ptb_mem h, void *handle_out) {
    NEED_INIT();
    MemObj *m = find(h);
    if (!m) return fail(PTB_ERR_INVALID_MEM, "ptb_ipc_export: bad handle");
    if (!handle_out) return fail(PTB_ERR_INVALID_VALUE, "ptb_ipc_export: null output");
    if (!m->owned || m->image) return fail(PTB_ERR_INVALID_VALUE, "ptb_ipc_export: only ptb_alloc'ed buffers (a whole device allocation) can be exported");
    static_assert(sizeof(cudaIpcMemHandle_t) == PTB_IPC_HANDLE_BYTES, "IPC handle size");
    cudaIpcMemHandle_t ipc;
    CUDA_OK(cudaIpcGetMemHandle(&ipc, m->dptr));
    memcpy(handle_out, &ipc, sizeof ipc);
    return PTB_SUCCESS;
}
ptb_mem ptb_ipc_import(const void *handle_in, size_t bytes) {
    if (!G.inited && ptb_init(-1)) return 0;
    if (!handle_in || bytes == 0) { fail(PTB_ERR_INVALID_VALUE, "ptb_ipc_import: null handle or zero size"); return 0; }
    cudaIpcMemHandle_t ipc;
    memcpy(&ipc, handle_in, sizeof ipc);
    void *ptr = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&ptr, ipc, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess || !ptr) { cudaGetLastError(); fail(PTB_ERR_DEVICE, "ptb_ipc_import: cudaIpcOpenMemHandle: %s", cudaGetErrorString(e)); return 0; }
    MemObj m;
    m.dptr = ptr;
    m.bytes = bytes;
    m.owned = false;
    m.ipc = true;
    ptb_mem h = G.next++;
    G.mems[h] = m;
    return h;
}
int ptb_invalidate(ptb_mem h) {
    MemObj *m = find(h);
    if (!m) return PTB_ERR_INVALID_MEM;
    int rc = G.inited ? flush_pending() : PTB_SUCCESS;
    m->version++;
    return rc;
}
void *ptb_device_ptr(ptb_mem h) {
    MemObj *m = find(h);
    if (!m) return nullptr;
    flush_pending();
    m->version++;   // the caller may write through the pointer
    return m->dptr;
}

// ---------------------------------------------------------------- kernels
int ptb_kernel_id(const char *name) {
    if (!name) return PTB_ERR_INVALID_KERNEL;
    if (!strcmp(name, "gen_ray")) return PTB_KERNEL_GEN_RAY;
    if (!strcmp(name, "trace_ray")) return PTB_KERNEL_TRACE_RAY;
    if (!strcmp(name, "filt_im")) return PTB_KERNEL_FILT_IM;
    return PTB_ERR_INVALID_KERNEL;
}
int ptb_set_arg(int k, int i, size_t bytes, const void *value) {
    if (k < 0 || k >= PTB_KERNEL_COUNT || i < 0 || i >= 16 || !value) return fail(PTB_ERR_INVALID_VALUE, "ptb_set_arg(%d,%d): bad kernel or index", k, i);
    Arg &a = G.args[k][i];
    a.value.assign((const unsigned char *)value, (const unsigned char *)value + bytes);
    a.mem = 0;
    a.set = true;
    return PTB_SUCCESS;
}
int ptb_set_arg_mem(int k, int i, ptb_mem m) {
    if (k < 0 || k >= PTB_KERNEL_COUNT || i < 0 || i >= 16) return fail(PTB_ERR_INVALID_VALUE, "ptb_set_arg_mem(%d,%d): bad kernel or index", k, i);
    Arg &a = G.args[k][i];
    a.value.clear();
    a.mem = m;
    a.set = true;
    return PTB_SUCCESS;
}

int ptb_launch(int k, size_t gx, size_t gy) {
    NEED_INIT();
    if (gy == 0) gy = 1;
    if (k == PTB_KERNEL_GEN_RAY) {
        // gen_ray(rays, Camera, rnds) — prog.cl:384-386 ↔ main.cpp:638-640, 1-D over W*H (main.cpp:642)
        int rc = flush_pending();
        if (rc) return rc;
        bool ok = true;
        RefCamera cam = arg_val<RefCamera>(k, 1, &ok);
        if (!ok || !G.args[k][0].set || !G.args[k][2].set) return fail(PTB_ERR_INVALID_ARGS, "gen_ray: arguments 0..2 must be set");
        size_t n = gx * gy;
        if (G.fuse_gen) {
            G.pending.active = true;
            G.pending.rays = G.args[k][0].mem;
            G.pending.rnds = G.args[k][2].mem;
            G.pending.cam = cam;
            G.pending.n = n;
            return PTB_SUCCESS;
        }
        return run_gen_ray(G.args[k][0].mem, G.args[k][2].mem, cam, n);
    }
    if (k == PTB_KERNEL_TRACE_RAY) {
        const int K = PTB_KERNEL_TRACE_RAY;
        bool fused = false;
        if (G.pending.active) {
            bool ok = true;
            RefCamera cam = arg_val<RefCamera>(K, 11, &ok);
            fused = ok && G.pending.rays == G.args[K][7].mem && G.pending.rnds == G.args[K][8].mem && G.pending.n == gx * gy &&
                    memcmp(&cam, &G.pending.cam, 72) == 0 && (int)cam.XM == (int)gx;
            if (fused) G.pending.active = false;
            else { int rc = flush_pending(); if (rc) return rc; }
        }
        if (G.n_gpus > 1) return render_multi(gx, gy, 1, !fused);
        RenderParams p;
        int rc = make_params(p, gx, gy, 1, !fused);
        if (rc) return rc;
        return run_any(p, false);            // single-sample launches: the pool kernel or the megakernel
    }
    if (k == PTB_KERNEL_FILT_IM) {
        // filt_im(tex, colors) — prog.cl:391-427; launched over (W,H) (main.cpp:665-668, commented out there)
        int rc = flush_pending();
        if (rc) return rc;
        return ptb_filter_median(G.args[k][1].mem, G.args[k][0].mem, (int)gx, (int)gy);
    }
    return fail(PTB_ERR_INVALID_KERNEL, "ptb_launch: unknown kernel id %d", k);
}

int ptb_finish(void) {
    NEED_INIT();
    int rc = flush_pending();
    if (rc) return rc;
    CUDA_OK(cudaStreamSynchronize(S()));
    return PTB_SUCCESS;
}
const char *ptb_last_error(void) { return G.err.c_str(); }
void ptb_clear_error(void) { G.err.clear(); }

// ---------------------------------------------------------------- extensions
int ptb_set_accel(int accel) {
    if (accel < PTB_ACCEL_AUTO || accel > PTB_ACCEL_LBVH) return PTB_ERR_INVALID_VALUE;
    G.accel = accel;
    return PTB_SUCCESS;
}
int ptb_set_math(int math) {
    if (math != PTB_MATH_PARITY && math != PTB_MATH_FAST) return PTB_ERR_INVALID_VALUE;
    G.math = math;
    return PTB_SUCCESS;
}
int ptb_set_pipeline(int pipe) {
    if (pipe != PTB_PIPE_MEGAKERNEL && pipe != PTB_PIPE_WAVEFRONT && pipe != PTB_PIPE_AUTO && pipe != PTB_PIPE_POOL) return PTB_ERR_INVALID_VALUE;
    G.pipe = pipe;
    return PTB_SUCCESS;
}
int ptb_set_quant_nodes(int on) {
    if (on < 0 || on > 2) return PTB_ERR_INVALID_VALUE;
    G.quant_nodes = on;
    return PTB_SUCCESS;
}
int ptb_set_wf_subs(int n) {
    if (n < 1 || n > 4) return PTB_ERR_INVALID_VALUE;
    G.wf_subs = n;
    return PTB_SUCCESS;
}
int ptb_set_fuse_gen(int on) {
    int rc = G.inited ? flush_pending() : PTB_SUCCESS;
    G.fuse_gen = on ? 1 : 0;
    return rc;
}
int ptb_set_merge_prims(int on) {
    G.merge_prims = on ? 1 : 0;
    for (auto &c : G.dev) c.staged.prims_valid = false;
    return PTB_SUCCESS;
}
int ptb_set_parity_filter(int on) {
    G.parity_filter = on ? 1 : 0;
    return PTB_SUCCESS;
}
int ptb_scan_prims(void) { return D().staged.prims_valid && D().staged.n_prims > 0 ? D().staged.n_prims : (D().staged.n_tris > 0 ? D().staged.n_tris : 0); }
int ptb_set_lbvh_leaf(int max_tris) {
    if (max_tris < 1 || max_tris > 8) return PTB_ERR_INVALID_VALUE;
    G.lbvh_leaf = max_tris;
    return PTB_SUCCESS;
}
int ptb_get_accel(void) {
    const int K = PTB_KERNEL_TRACE_RAY;
    bool ok = true;
    int n = arg_val<int>(K, 2, &ok);
    int ns = arg_val<int>(K, 6, &ok);
    int a = resolve_accel(ok ? n : 0, arg_mem(K, 4) && arg_mem(K, 5) && ns > 0);
    return a == PT_ACCEL_BRUTE ? PTB_ACCEL_BRUTE : a == PT_ACCEL_REFBVH ? PTB_ACCEL_REFBVH : PTB_ACCEL_LBVH;
}
int ptb_build_accel(void) {
    NEED_INIT();
    const int K = PTB_KERNEL_TRACE_RAY;
    MemObj *tris = arg_mem(K, 1);
    bool ok = true;
    int n = arg_val<int>(K, 2, &ok);
    if (!tris || !ok) return PTB_SUCCESS;   // nothing uploaded yet: the first launch will build
    int ns = arg_val<int>(K, 6, &ok);
    const bool lbvh = resolve_accel(n, arg_mem(K, 4) && arg_mem(K, 5) && ns > 0) == PT_ACCEL_LBVH;
    for (int d = 0; d < G.n_gpus; ++d) {
        DevScope scope(d);
        int rc = stage_scene(tris, G.args[K][1].mem, n < 0 ? 0 : n);
        if (rc) return rc;
        if (lbvh) {
            rc = ensure_lbvh(0.0f);
            if (rc) return rc;
        }
        CUDA_OK(cudaStreamSynchronize(S()));
    }
    return PTB_SUCCESS;
}

// Single-process multi-GPU (SURVEY 8b's ptb_set_gpus): the next launches run on the first n GPUs of the box, counted from the
// one ptb_init chose.  See render_multi above for what each mode does.
int ptb_set_gpus(int n, int mode) {
    NEED_INIT();
    if (mode != PTB_SPLIT_TILE && mode != PTB_SPLIT_SPP) return fail(PTB_ERR_INVALID_VALUE, "ptb_set_gpus: mode must be PTB_SPLIT_TILE or PTB_SPLIT_SPP");
    int count = 0;
    CUDA_OK(cudaGetDeviceCount(&count));
    if (n < 1 || n > PTB_MAX_GPUS || n > count) return fail(PTB_ERR_INVALID_VALUE, "ptb_set_gpus: %d GPUs requested, %d present (at most %d)", n, count, PTB_MAX_GPUS);
    int rc = flush_pending();
    if (rc) return rc;
    CUDA_OK(cudaStreamSynchronize(G.stream));
    for (int d = 1; d < n; ++d) {
        DevCtx &c = G.dev[d];
        if (c.device >= 0) continue;
        const int dev = (G.device + d) % count;
        int a = 0, b = 0;
        CUDA_OK(cudaDeviceCanAccessPeer(&a, dev, G.device));
        CUDA_OK(cudaDeviceCanAccessPeer(&b, G.device, dev));
        if (!a || !b) return fail(PTB_ERR_DEVICE, "ptb_set_gpus: no peer access between GPU %d and GPU %d", G.device, dev);
        cudaError_t e = cudaDeviceEnablePeerAccess(dev, 0);                     // primary -> dev (the spp split's resolve reads the peers' sums)
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail(PTB_ERR_DEVICE, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
        cudaGetLastError();
        CUDA_OK(cudaSetDevice(dev));
        e = cudaDeviceEnablePeerAccess(G.device, 0);                            // dev -> primary (scene staging reads, frame stores)
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { cudaSetDevice(G.device); return fail(PTB_ERR_DEVICE, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e)); }
        cudaGetLastError();
        e = cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c.ev_done, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaMalloc((void **)&c.counters, (PT_CNT_WORDS + 2) * sizeof(unsigned long long));
        if (e == cudaSuccess) e = cudaMemset(c.counters, 0, (PT_CNT_WORDS + 2) * sizeof(unsigned long long));
        cudaSetDevice(G.device);
        if (e != cudaSuccess) return fail(PTB_ERR_DEVICE, "ptb_set_gpus: GPU %d: %s", dev, cudaGetErrorString(e));
        c.device = dev;
    }
    G.n_gpus = n;
    G.gpu_mode = mode;
    for (auto &c : G.dev) c.sum_index = -1;
    return PTB_SUCCESS;
}
int ptb_get_gpus(void) { return G.n_gpus; }

int ptb_render(int spp) {
    NEED_INIT();
    if (spp <= 0) return PTB_SUCCESS;
    int rc = flush_pending();
    if (rc) return rc;
    bool ok = true;
    RefCamera cam = arg_val<RefCamera>(PTB_KERNEL_TRACE_RAY, 11, &ok);
    if (!ok) return fail(PTB_ERR_INVALID_ARGS, "ptb_render: trace_ray argument 11 (Camera) is not set");
    if (G.n_gpus > 1) return render_multi((size_t)(int)cam.XM, (size_t)(int)cam.YM, spp, false);
    RenderParams p;
    rc = make_params(p, (size_t)(int)cam.XM, (size_t)(int)cam.YM, spp, false);
    if (rc) return rc;
    return run_any(p, true);                 // LBVH scenes: pool kernel or wavefront; tiny scenes: the megakernel's shared-memory scan
}

int ptb_closest_hit(ptb_mem rays_h, size_t n_rays, ptb_mem out_h) {
    NEED_INIT();
    MemObj *rays = find(rays_h), *out = find(out_h);
    if (!rays || !out) return fail(PTB_ERR_INVALID_MEM, "ptb_closest_hit: bad handle");
    if (rays->bytes < n_rays * sizeof(RefRay) || out->bytes < n_rays * 8)
        return fail(PTB_ERR_INVALID_VALUE, "ptb_closest_hit: %zu rays need %zu + %zu bytes", n_rays, n_rays * sizeof(RefRay), n_rays * 8);
    if (n_rays == 0) return PTB_SUCCESS;
    int rc = flush_pending();
    if (rc) return rc;
    bool ok = true;
    RefCamera cam = arg_val<RefCamera>(PTB_KERNEL_TRACE_RAY, 11, &ok);
    if (!ok) return fail(PTB_ERR_INVALID_ARGS, "ptb_closest_hit: bind the scene through trace_ray's arguments first (argument 11 is not set)");
    RenderParams p;
    rc = make_params(p, (size_t)(int)cam.XM, (size_t)(int)cam.YM, 1, false);
    if (rc) return rc;
    if (want_pool(p)) {
        rc = ensure_pool_blob();
        if (rc) return rc;
        const int kp = G.math == PTB_MATH_PARITY ? pt_parity::launch_pool_closest_hit(p, (const RefRay *)rays->dptr, n_rays, (float2 *)out->dptr, D().pool_blob, S())
                                                 : pt_fast::launch_pool_closest_hit(p, (const RefRay *)rays->dptr, n_rays, (float2 *)out->dptr, D().pool_blob, S());
        cudaError_t ep = cudaGetLastError();
        if (ep == cudaSuccess) ep = cudaStreamSynchronize(S());
        if (ep != cudaSuccess) return fail(PTB_ERR_DEVICE, "ptb_closest_hit: %s", cudaGetErrorString(ep));
        G.launches += (unsigned long long)kp;
        out->version++;
        return PTB_SUCCESS;
    }
    const bool wf = G.pipe != PTB_PIPE_MEGAKERNEL && p.accel == PT_ACCEL_LBVH;
    void *scratch = nullptr;
    if (wf) CUDA_OK(cudaMalloc(&scratch, n_rays * 36 + 64));
    const int k = G.math == PTB_MATH_PARITY
                      ? pt_parity::launch_closest_hit(p, (const RefRay *)rays->dptr, n_rays, (float2 *)out->dptr, wf, scratch, S())
                      : pt_fast::launch_closest_hit(p, (const RefRay *)rays->dptr, n_rays, (float2 *)out->dptr, wf, scratch, S());
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(S());
    cudaFree(scratch);
    if (e != cudaSuccess) return fail(PTB_ERR_DEVICE, "ptb_closest_hit: %s", cudaGetErrorString(e));
    G.launches += (unsigned long long)k;
    out->version++;
    return PTB_SUCCESS;
}

int ptb_set_pixel_shard(int rank, int n, int group) {
    if (n < 1 || rank < 0 || rank >= n || group < 1) return PTB_ERR_INVALID_VALUE;
    int rc = G.inited ? flush_pending() : PTB_SUCCESS;
    G.shard_rank = rank; G.shard_n = n; G.shard_group = group;
    return rc;
}
int ptb_set_accumulate_sum(int on) {
    G.sum_mode = on ? 1 : 0;
    return PTB_SUCCESS;
}
int ptb_resolve_sum(ptb_mem colors_sum, int total_spp) {
    NEED_INIT();
    MemObj *m = find(colors_sum);
    if (!m || total_spp <= 0) return fail(PTB_ERR_INVALID_VALUE, "ptb_resolve_sum: bad buffer or sample count");
    int rc = flush_pending();
    if (rc) return rc;
    const size_t n = m->bytes / sizeof(float4);
    if (n > 0x7fffffffu) return fail(PTB_ERR_INVALID_VALUE, "ptb_resolve_sum: buffer too large");
    ResolveSrc src{};
    src.p[0] = (const float4 *)m->dptr; src.n = 1;
    resolve_frame(src, 1.0f / (float)total_spp, (int)n, 1, 0, (float4 *)m->dptr, nullptr, nullptr, S());
    CUDA_OK(cudaGetLastError());
    G.launches++;
    m->version++;
    return PTB_SUCCESS;
}

int ptb_set_count_work(int on) {
    G.count_work = on ? 1 : 0;
    return PTB_SUCCESS;
}
int ptb_get_counters(ptb_counters *out) {
    NEED_INIT();
    if (!out) return PTB_ERR_INVALID_VALUE;
    int rc = flush_pending();
    if (rc) return rc;
    unsigned long long tot[PT_CNT_WORDS] = {0};
    for (int d = 0; d < G.n_gpus; ++d) {
        DevScope scope(d);
        unsigned long long h[PT_CNT_WORDS];
        CUDA_OK(cudaMemcpyAsync(h, D().counters, sizeof h, cudaMemcpyDeviceToHost, S()));
        CUDA_OK(cudaStreamSynchronize(S()));
        for (int k = 0; k < PT_CNT_WORDS; ++k) tot[k] += h[k];
    }
    out->rays = tot[PT_CNT_RAYS]; out->box_tests = tot[PT_CNT_BOX]; out->tri_tests = tot[PT_CNT_TRI];
    out->tri_stage1 = tot[PT_CNT_TRI1]; out->tri_stage2 = tot[PT_CNT_TRI2]; out->tri_stage3 = tot[PT_CNT_TRI3];
    out->pixel_samples = tot[PT_CNT_PIXEL_SAMPLES];
    return PTB_SUCCESS;
}
int ptb_reset_counters(void) {
    NEED_INIT();
    for (int d = 0; d < G.n_gpus; ++d) {
        DevScope scope(d);
        CUDA_OK(cudaMemsetAsync(D().counters, 0, PT_CNT_WORDS * sizeof(unsigned long long), S()));
    }
    return PTB_SUCCESS;
}

int ptb_resolve_rgba8(ptb_mem colors, int width, int height, int mode, unsigned char *dst_host) {
    if (!dst_host) return fail(PTB_ERR_INVALID_VALUE, "ptb_resolve_rgba8: null destination");
    return ptb_resolve_frame(colors, width, height, 1, mode, 0, dst_host);
}
// The root's epilogue in one pass: sum -> mean -> tone map -> float image + 8-bit RGBA.
int ptb_resolve_frame(ptb_mem colors, int width, int height, int divide_by, int mode, ptb_mem image, unsigned char *dst_host) {
    NEED_INIT();
    MemObj *m = find(colors), *im = image ? find(image) : nullptr;
    const size_t n = (size_t)(width > 0 ? width : 0) * (size_t)(height > 0 ? height : 0);
    if (!m || n == 0 || n > 0x7fffffffu || m->bytes < n * 16 || divide_by < 1 || (image && (!im || im->bytes < n * 16)) || (mode != 0 && mode != 1))
        return fail(PTB_ERR_INVALID_VALUE, "ptb_resolve_frame: bad arguments");
    int rc = flush_pending();
    if (rc) return rc;
    unsigned char *d = nullptr;
    if (dst_host) CUDA_OK(cudaMalloc((void **)&d, n * 4));
    ResolveSrc src{};
    src.p[0] = (const float4 *)m->dptr; src.n = 1;
    resolve_frame(src, divide_by > 1 ? 1.0f / (float)divide_by : 1.0f, width, height, mode, divide_by > 1 ? (float4 *)m->dptr : nullptr,
                  im ? (float4 *)im->dptr : nullptr, d, S());
    cudaError_t e = cudaGetLastError();
    G.launches++;
    if (divide_by > 1) m->version++;
    if (im) im->version++;
    if (e == cudaSuccess && dst_host) e = cudaMemcpyAsync(dst_host, d, n * 4, cudaMemcpyDeviceToHost, S());
    if (e == cudaSuccess && dst_host) e = cudaStreamSynchronize(S());
    cudaFree(d);
    if (e != cudaSuccess) return fail(PTB_ERR_DEVICE, "ptb_resolve_frame: %s", cudaGetErrorString(e));
    return PTB_SUCCESS;
}
// ptb_resolve_rgba8 for callers that must not allocate: same arguments, same picture, the staging buffer is a static device array
// (resolve.cu).  No cudaMalloc / cudaFree / allocating library call on this path.
int ptb_resolve_rgba8_noalloc(ptb_mem colors, int width, int height, int mode, unsigned char *dst_host) {
    NEED_INIT();
    MemObj *m = find(colors);
    const size_t n = (size_t)(width > 0 ? width : 0) * (size_t)(height > 0 ? height : 0);
    if (!m || !dst_host || n == 0 || m->bytes / 16 < n || (mode != 0 && mode != 1)) return fail(PTB_ERR_INVALID_VALUE, "ptb_resolve_rgba8_noalloc: bad arguments");
    if (n > (size_t)PTB_NOALLOC_MAX_PIXELS)
        return fail(PTB_ERR_INVALID_VALUE, "ptb_resolve_rgba8_noalloc: %d x %d pixels exceed the static scratch of PTB_NOALLOC_MAX_PIXELS = %d pixels; use ptb_resolve_rgba8",
                    width, height, PTB_NOALLOC_MAX_PIXELS);
    int rc = flush_pending();
    if (rc) return rc;
    char err[256];
    if (resolve_rgba8_noalloc((const float4 *)m->dptr, width, height, mode, dst_host, S(), err, sizeof err)) return fail(PTB_ERR_DEVICE, "%s", err);
    G.launches++;
    return PTB_SUCCESS;
}
int ptb_filter_median(ptb_mem colors, ptb_mem image, int width, int height) {
    NEED_INIT();
    MemObj *c = find(colors), *im = find(image);
    size_t n = (size_t)width * (size_t)height;
    if (!c || !im || width <= 0 || height <= 0 || c->bytes < n * 16 || im->bytes < n * 16)
        return fail(PTB_ERR_INVALID_VALUE, "ptb_filter_median: bad arguments");
    int rc = flush_pending();
    if (rc) return rc;
    filter_median((const float4 *)c->dptr, (float4 *)im->dptr, width, height, S());
    CUDA_OK(cudaGetLastError());
    G.launches++;
    im->version++;
    return PTB_SUCCESS;
}

// ---------------------------------------------------------------- one process per GPU: NCCL behind the boundary (comm.cu)
int ptb_comm_unique_id(void *id_out) {
    if (!id_out) return fail(PTB_ERR_INVALID_VALUE, "ptb_comm_unique_id: null output");
    char err[256];
    if (comm_unique_id(id_out, err, sizeof err)) return fail(PTB_ERR_DEVICE, "%s", err);
    return PTB_SUCCESS;
}
int ptb_comm_init(int rank, int nranks, const void *unique_id) {
    NEED_INIT();
    if (!unique_id || nranks < 1 || rank < 0 || rank >= nranks) return fail(PTB_ERR_INVALID_VALUE, "ptb_comm_init: bad rank %d of %d or null id", rank, nranks);
    char err[256];
    CUDA_OK(cudaSetDevice(G.device));
    if (comm_init(rank, nranks, unique_id, err, sizeof err)) return fail(PTB_ERR_DEVICE, "%s", err);
    return PTB_SUCCESS;
}
int ptb_comm_ranks(void) { return comm_ranks(); }
int ptb_comm_rank(void) { return comm_rank(); }
int ptb_comm_reduce(ptb_mem colors, int root) {
    NEED_INIT();
    MemObj *m = find(colors);
    if (!m || root < 0 || root >= comm_ranks()) return fail(PTB_ERR_INVALID_VALUE, "ptb_comm_reduce: bad buffer or root");
    int rc = flush_pending();
    if (rc) return rc;
    if (comm_ranks() == 1) return PTB_SUCCESS;
    char err[256];
    if (comm_reduce_sum((float *)m->dptr, m->bytes / sizeof(float), root, S(), err, sizeof err)) return fail(PTB_ERR_DEVICE, "%s", err);
    m->version++;
    return PTB_SUCCESS;
}
int ptb_comm_broadcast(ptb_mem buf, int root) {
    NEED_INIT();
    MemObj *m = find(buf);
    if (!m || root < 0 || root >= comm_ranks()) return fail(PTB_ERR_INVALID_VALUE, "ptb_comm_broadcast: bad buffer or root");
    int rc = flush_pending();
    if (rc) return rc;
    if (comm_ranks() == 1) return PTB_SUCCESS;
    char err[256];
    if (comm_broadcast(m->dptr, m->bytes, root, S(), err, sizeof err)) return fail(PTB_ERR_DEVICE, "%s", err);
    m->version++;
    return PTB_SUCCESS;
}
int ptb_comm_barrier(void) {
    NEED_INIT();
    int rc = flush_pending();
    if (rc) return rc;
    char err[256];
    if (comm_barrier(S(), err, sizeof err)) return fail(PTB_ERR_DEVICE, "%s", err);
    return PTB_SUCCESS;
}
int ptb_comm_destroy(void) { return comm_destroy(); }

// ---------------------------------------------------------------- checkpoint / resume (the progressive state is exactly colors + rnds + sample index)
namespace {
struct CkptHeader { char magic[4]; int32_t version, width, height, sample, iterations; uint32_t pad[2]; };
}
int ptb_save_state(const char *path) {
    NEED_INIT();
    if (!path) return fail(PTB_ERR_INVALID_VALUE, "ptb_save_state: null path");
    const int K = PTB_KERNEL_TRACE_RAY;
    bool ok = true;
    RefCamera cam = arg_val<RefCamera>(K, 11, &ok);
    const int sample = arg_val<int>(K, 10, &ok), iterations = arg_val<int>(K, 9, &ok);
    MemObj *rnds = arg_mem(K, 8), *colors = arg_mem(K, 12);
    if (!ok || !rnds || !colors) return fail(PTB_ERR_INVALID_ARGS, "ptb_save_state: bind trace_ray's arguments 8-12 first");
    const size_t n = (size_t)(int)cam.XM * (size_t)(int)cam.YM;
    if (n == 0 || rnds->bytes < n * 4 || colors->bytes < n * 16) return fail(PTB_ERR_INVALID_ARGS, "ptb_save_state: rnds/colors smaller than the camera's frame");
    std::vector<unsigned char> host(n * 20);
    int rc = ptb_read(G.args[K][8].mem, 0, n * 4, host.data());
    if (rc) return rc;
    rc = ptb_read(G.args[K][12].mem, 0, n * 16, host.data() + n * 4);
    if (rc) return rc;
    CkptHeader h{{'P', 'T', 'B', 'C'}, 1, (int)cam.XM, (int)cam.YM, sample, iterations, {0, 0}};
    FILE *f = fopen(path, "wb");
    if (!f) return fail(PTB_ERR_INVALID_VALUE, "ptb_save_state: cannot write %s", path);
    const bool wrote = fwrite(&h, sizeof h, 1, f) == 1 && fwrite(host.data(), 1, host.size(), f) == host.size();
    if (fclose(f) != 0 || !wrote) return fail(PTB_ERR_INVALID_VALUE, "ptb_save_state: short write to %s", path);
    return PTB_SUCCESS;
}
int ptb_load_state(const char *path, int *sample_out) {
    NEED_INIT();
    if (!path) return fail(PTB_ERR_INVALID_VALUE, "ptb_load_state: null path");
    const int K = PTB_KERNEL_TRACE_RAY;
    bool ok = true;
    RefCamera cam = arg_val<RefCamera>(K, 11, &ok);
    MemObj *rnds = arg_mem(K, 8), *colors = arg_mem(K, 12);
    if (!ok || !rnds || !colors) return fail(PTB_ERR_INVALID_ARGS, "ptb_load_state: bind trace_ray's arguments 8, 11, 12 first");
    FILE *f = fopen(path, "rb");
    if (!f) return fail(PTB_ERR_INVALID_VALUE, "ptb_load_state: cannot read %s", path);
    CkptHeader h;
    if (fread(&h, sizeof h, 1, f) != 1 || memcmp(h.magic, "PTBC", 4) || h.version != 1) { fclose(f); return fail(PTB_ERR_INVALID_VALUE, "ptb_load_state: %s is not a PTBC v1 file", path); }
    if (h.width != (int)cam.XM || h.height != (int)cam.YM) { fclose(f); return fail(PTB_ERR_INVALID_VALUE, "ptb_load_state: checkpoint is %dx%d, the bound camera %dx%d", h.width, h.height, (int)cam.XM, (int)cam.YM); }
    const size_t n = (size_t)h.width * (size_t)h.height;
    if (rnds->bytes < n * 4 || colors->bytes < n * 16) { fclose(f); return fail(PTB_ERR_INVALID_ARGS, "ptb_load_state: rnds/colors smaller than the checkpoint's frame"); }
    std::vector<unsigned char> host(n * 20);
    const bool got = fread(host.data(), 1, host.size(), f) == host.size();
    fclose(f);
    if (!got) return fail(PTB_ERR_INVALID_VALUE, "ptb_load_state: %s is truncated", path);
    int rc = ptb_write(G.args[K][8].mem, 0, n * 4, host.data());
    if (rc) return rc;
    rc = ptb_write(G.args[K][12].mem, 0, n * 16, host.data() + n * 4);
    if (rc) return rc;
    ptb_set_arg(K, 10, sizeof(int), &h.sample);          // the next launch continues at the checkpoint's sample index
    if (sample_out) *sample_out = h.sample;
    return PTB_SUCCESS;
}

int ptb_set_stream(void *cuda_stream) {
    NEED_INIT();
    int rc = flush_pending();
    if (rc) return rc;
    CUDA_OK(cudaStreamSynchronize(S()));
    G.stream = cuda_stream ? (cudaStream_t)cuda_stream : G.own_stream;
    return PTB_SUCCESS;
}
void *ptb_get_stream(void) { return (void *)G.stream; }
int ptb_accel_stats(int *n_nodes, int *n_tris, int *max_depth, float *build_ms) {
    const Lbvh &b = D().staged.lbvh;
    bool v = D().staged.lbvh_valid;
    if (n_nodes) *n_nodes = v ? b.n_nodes : 0;
    if (n_tris) *n_tris = v ? b.n_tris : 0;
    if (max_depth) *max_depth = v ? b.max_depth : 0;
    if (build_ms) *build_ms = v ? b.build_ms : 0.0f;
    return PTB_SUCCESS;
}
uint64_t ptb_launch_count(void) { return G.launches; }

}  // extern "C"
