// prims.h — merged scan primitives (convex planar quad = two coplanar triangles) for the fast brute-force scan; see prims.cu
#ifndef PRIMS_H_
#define PRIMS_H_
#include <cuda_runtime.h>
#include <stdint.h>

#define PRIMS_MAX_TRIS 1024   // builder is a single block, O(n^2) worst case: only for scenes the brute-force scan takes anyway

// q0..q2: staged triangles in the parity layout (vertices + normal), mati: material per triangle.
// p0..p2: primitives in the fast layout (normal, anchor, barycentric gradients), ab: the two edge slopes (b, a) of prims.cu,
// rep: upload index of the triangle that stands for the primitive (normal + material), n_prims_dev: count.
// shade: 3 float4 per primitive — (N, bits(mati)) and the tangent frame (X, Z) of prog.cl:186-204 for the stored normal.
// idx (optional): triangle i of q* is upload index idx[i] (a Morton-ordered sub-list): mati is looked up and rep reported through it.
void prims_build(const float4 *q0, const float4 *q1, const float4 *q2, const uint16_t *mati, const int *idx, int n, float4 *p0,
                 float4 *p1, float4 *p2, float2 *ab, int *rep, float4 *shade, int *n_prims_dev, cudaStream_t st);
#endif
