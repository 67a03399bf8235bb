// pt_params.h — launch descriptors shared between the C-ABI (ptb_abi.cu) and the two kernel
// translation units (pt_kernels_parity.cu / pt_kernels_fast.cu).
#ifndef PT_PARAMS_H_
#define PT_PARAMS_H_

#include <cuda_runtime.h>
#include <stdint.h>
#include "ptb_types.h"

// closest-hit structure actually used by a launch (PTB_ACCEL_AUTO already resolved)
enum { PT_ACCEL_BRUTE = 1, PT_ACCEL_REFBVH = 2, PT_ACCEL_LBVH = 3 };

// device counters block (uint64 each)
enum {
    PT_CNT_RAYS = 0, PT_CNT_BOX, PT_CNT_TRI, PT_CNT_TRI1, PT_CNT_TRI2, PT_CNT_TRI3, PT_CNT_PIXEL_SAMPLES,
    PT_CNT_WORDS = 8
};

// Everything a path kernel reads about the scene.  Passed by value (lives in the constant bank).
struct SceneDev {
    // staged triangles in upload order, 3 float4 planes in the layout of the arithmetic flavour that
    // reads them (parity: vertices + normal; fast: normal, anchor, barycentric gradients — pt_kernels.cuh)
    const float4 *q0, *q1, *q2;
    const float4 *fq0, *fq1, *fq2;// always the fast layout (the parity flavour's conservative pre-filter reads it)
    const uint16_t *mati;
    int n_tris;
    const RefMaterial *mats;      // as uploaded (prog.cl:1-5)
    int n_mats;
    // the host-built per-object trees exactly as uploaded (prog.cl:27-30, main.cpp:263-303)
    const RefNode *kd_tree;
    const int *kd_shift;
    int n_obj;
    // GPU-built LBVH (lbvh.cu)
    const float4 *bvh;            // 4 float4 per node: both children's boxes + links (lbvh.cu)
    cudaTextureObject_t bvh_tex, bvh4_tex;   // bvh / bvh4 as linear float4 textures, 0 = fetch through the LSU
    cudaTextureObject_t bvh4q_tex;           // bvh4q as a linear uint4 texture, 0 = fetch through the LSU
    const float4 *bvh4q;          // the 4-ary nodes quantised to 64 bytes (lbvh.cu k_emit4q); null: traverse the float nodes
    int quant_wf;                 // 1: the wavefront's trace kernel walks them too (forced by ptb_set_quant_nodes(1); by measurement it prefers the float nodes)
    const float4 *bvh4;           // the same tree collapsed to 4-ary nodes, 8 float4 each (lbvh.cu); null: not usable
    const float4 *sq0, *sq1, *sq2;// triangles in Morton order
    const int *sorted_id;         // Morton position → index in upload order
    const float4 *srec;           // Morton-ordered 64-byte triangle records of this flavour's layout + upload index (lbvh.h)
    int n_nodes4;                 // wide nodes in bvh4 (128 bytes each)
    int bvh_root;                 // link of the root (may be a leaf link when n_tris == 1)
    // the same outliers as merged scan primitives (prims.cu) for the fast flavour; n_big_prims == 0: scan the triangles
    const float4 *bq0, *bq1, *bq2;
    const float2 *big_ab;
    const int *big_rep;           // primitive -> upload index
    int n_big_prims;
    int big_first, big_count;     // outlier triangles outside the tree: Morton positions [big_first, big_first + big_count)
    // merged scan primitives for the fast brute-force scan (prims.cu); n_prims == 0: not built / nothing merged
    const float4 *pq0, *pq1, *pq2;// fast layout
    const float2 *prim_ab;        // edge slopes (b, a) of the quad's two far edges; (1, 0) for a lone triangle
    const float4 *prim_shade;     // 3 per primitive: (N, bits(mati)) (X, 0) (Z, 0) — normal, material, tangent frame
    const int *prim_rep;          // primitive → upload index of the triangle that stands for it
    int n_prims;
};

struct RenderParams {
    SceneDev sc;
    RefCamera cam;                // trace_ray arg 11 / gen_ray arg 1
    // the same camera as one affine map of the jittered pixel position (fast flavour only, set by pt_cam_affine):
    // p - eye = cam_o + cam_dx * (px + rnd1) + cam_dy * (py + rnd2)   (prog.cl:82-92 with the constants folded)
    float4 cam_o, cam_dx, cam_dy;
    int *rnds;                    // per-pixel RNG state (main.cpp:509,522-526)
    float4 *colors;               // float3[W*H], 16 B each (main.cpp:520)
    float4 *tex;                  // RGBA float image or nullptr (main.cpp:512-519)
    RefRay *rays;                 // gen_ray output / trace_ray input; unused by the fused render
    int W, H;
    int iterations;               // trace_ray arg 9
    int sample0;                  // trace_ray arg 10: index of the first sample of this launch
    int spp;                      // samples rendered by this launch (1 for the reference's trace_ray)
    int from_rays;                // 1: first ray of the (single) sample comes from rays[] (drop-in trace_ray)
    int shard_rank, shard_n, shard_group;   // pixel id selected iff (id / group) % n == rank
    int sum_mode;                 // 1: colors += sample (spp-split), 0: running mean (prog.cl:379)
    int accel;                    // PT_ACCEL_*
    int count_work;               // 1: also count box / triangle tests
    int wf_subs;                  // wavefront: interleaved sub-wavefronts on separate streams (1..4; 0 = default 2)
    int wf_sort;                  // wavefront: group each shade block's surviving rays by direction octant in the next queue
    int parity_filter;            // parity flavour, brute force: conservative FFMA2 pre-filter in front of the exact test (pt_kernels.cuh)
    int regen_min;                // megakernel: lanes of a warp that must be waiting for their next sample before the camera code runs (1 = at once)
    int rng_nonneg;               // 1: every RNG state of this launch was verified to lie in [0, 2^31) (pt_rand_nonneg)
    unsigned long long *counters; // PT_CNT_WORDS words
    // wavefront pipeline scratch (pt_wavefront.cuh), one path per pixel, allocated by ptb_abi.cu; null = megakernel only
    float4 *wf_ray0, *wf_ray1;    // (P.xyz, D.x) (D.y, D.z, t_hit, bits(tri))
    float4 *wf_state;             // WF_PLANES planes of W*H float4 (pt_wavefront.cuh wf_store_hot): 0-2 hot, 3 tally, 4 colour, 5 parity fR.yz
    int *wf_queue[2];             // active pixel ids, double buffered
    int *wf_ctrl;                 // [0],[1] queue lengths, [2] fetch cursor of the trace kernel, [3] spare
};

static inline void pt_cam_affine(RenderParams &p) {
    const RefCamera &c = p.cam;
    const double X = (double)(int)c.XM, Y = (double)(int)c.YM;
    const double e[3] = {c.eye.x, c.eye.y, c.eye.z}, l[3] = {c.lookat.x, c.lookat.y, c.lookat.z};
    const double u[3] = {c.up.x, c.up.y, c.up.z}, r[3] = {c.right.x, c.right.y, c.right.z};
    float o[3], dx[3], dy[3];
    for (int k = 0; k < 3; ++k) {
        o[k] = (float)(l[k] - e[k] - r[k] - u[k]);
        dx[k] = (float)(X > 0 ? 2.0 * r[k] / X : 0.0);
        dy[k] = (float)(Y > 0 ? 2.0 * u[k] / Y : 0.0);
    }
    p.cam_o = make_float4(o[0], o[1], o[2], 0.0f);
    p.cam_dx = make_float4(dx[0], dx[1], dx[2], 0.0f);
    p.cam_dy = make_float4(dy[0], dy[1], dy[2], 0.0f);
}

// one entry per translation unit; implemented in pt_kernels.cuh
namespace pt_parity {
void launch_gen_ray(const RenderParams &p, size_t n, cudaStream_t s);
void launch_render(const RenderParams &p, cudaStream_t s);
int launch_wavefront(const RenderParams &p, cudaStream_t s);   // returns the number of kernels launched
int launch_closest_hit(const RenderParams &p, const RefRay *rays, size_t n, float2 *out, bool wavefront, void *scratch, cudaStream_t s);
// pt_pool.cuh: the persistent state-regrouping kernel for the LBVH (one launch per render)
size_t pool_scratch_bytes();
bool pool_usable(const RenderParams &p);
int launch_pool(const RenderParams &p, void *scratch, cudaStream_t s);
int launch_pool_closest_hit(const RenderParams &p, const RefRay *rays, size_t n, float2 *out, void *scratch, cudaStream_t s);
}
namespace pt_fast {
void launch_gen_ray(const RenderParams &p, size_t n, cudaStream_t s);
void launch_render(const RenderParams &p, cudaStream_t s);
int launch_wavefront(const RenderParams &p, cudaStream_t s);
int launch_closest_hit(const RenderParams &p, const RefRay *rays, size_t n, float2 *out, bool wavefront, void *scratch, cudaStream_t s);
// pt_pool.cuh: the persistent state-regrouping kernel for the LBVH (one launch per render)
size_t pool_scratch_bytes();
bool pool_usable(const RenderParams &p);
int launch_pool(const RenderParams &p, void *scratch, cudaStream_t s);
int launch_pool_closest_hit(const RenderParams &p, const RefRay *rays, size_t n, float2 *out, void *scratch, cudaStream_t s);
}

#endif  // PT_PARAMS_H_
