// lbvh.h — host interface of the GPU LBVH builder (lbvh.cu).
#ifndef PTB_LBVH_H_
#define PTB_LBVH_H_
#include <cuda_runtime.h>
#include <stddef.h>
#include <string.h>

#define LBVH_BIG_MAX 64          // outlier triangles kept out of the tree (lbvh.cu, step 2)

struct Lbvh {
    // traversal data
    float4 *bvh;                  // 4 float4 per node
    float4 *bvh4;                 // 8 float4 per 4-ary node (root = node 0), n_nodes4 of them; null when the root is a leaf
    float4 *bvh4q;                // the same wide nodes quantised to 64 bytes (8-bit child boxes against the node's own origin / step), lbvh.cu
    int n_nodes4;
    cudaTextureObject_t bvh_tex, bvh4_tex;   // the node arrays as linear float4 textures (fetches go through the TEX data pipe)
    cudaTextureObject_t bvh4q_tex;           // bvh4q as a linear uint4 texture
    int *parent, *flag4, *id4;    // binary parent links; which binary nodes are wide nodes; their compact ids
    float4 *sq0, *sq1, *sq2;      // triangles in Morton order, parity layout (vertices + normal)
    float4 *sf0, *sf1, *sf2;      // triangles in Morton order, fast layout (normal, anchor, barycentric gradients)
    int *sorted_id;               // Morton position → upload index
    float4 *rq, *rf;              // the same triangles as 64-byte records (3 float4 of the layout + (bits(upload index), 0, 0, 0)):
                                  // a leaf step fetches a triangle with two 256-bit loads, index included
    int root;                     // root link
    // kept for re-emitting with a larger inflation (far-away camera)
    void *range, *child;          // int2 per internal node
    float4 *leaf_lo, *leaf_hi, *node_lo, *node_hi;
    int n_tris, leaf_size, n_nodes, max_depth;
    int ploc_rounds;              // clustering rounds of the PLOC build (0: the Karras radix tree was used — PTB_LBVH_PLOC=0)
    int n_tree, n_big;            // Morton positions [0, n_tree) are under the tree, [n_tree, n_tris) is the outlier list
    float max_abs, reach, eps, build_ms;
    unsigned long long launches;
    // one device allocation holds every array above plus the build's scratch; it is kept across rebuilds (a re-upload of
    // the same scene rebuilds the tree without a single cudaMalloc / cudaFree) and regrown only for more triangles
    void *arena;
    size_t arena_bytes;
    int cap;                      // triangles the arena is laid out for
};

// q*: staged triangles in upload order (parity layout, holds the vertices); f*: the same triangles in the fast layout
int lbvh_build(Lbvh *b, const float4 *q0, const float4 *q1, const float4 *q2, const float4 *f0, const float4 *f1, const float4 *f2,
               int n, int leaf_size, float big_frac, cudaStream_t st, char *err, size_t errcap);
// rewrite the traversal nodes with boxes inflated for ray origins up to `reach` from the coordinate origin
int lbvh_emit(Lbvh *b, float reach, cudaStream_t st, char *err, size_t errcap);
void lbvh_free(Lbvh *b);
#endif
