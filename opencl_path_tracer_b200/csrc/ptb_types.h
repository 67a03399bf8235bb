// ptb_types.h — byte layouts of the reference's kernel ABI (prog.cl:1-35 ↔ main.cpp:92-348) as the
// CUDA side sees them, plus the device-internal scene layouts this library stages them into.
//
// The host writes these records with ptb_write exactly as main.cpp:526,620,624,628,633 writes its
// cl::Buffers; nothing here is shared with the reference at source level — only the byte layout is.
#ifndef PTB_TYPES_H_
#define PTB_TYPES_H_

#include <stdint.h>

struct alignas(16) RefFloat3 {   // OpenCL float3 / cl_float3: 16 bytes, lane 3 unused
    float x, y, z, w;
};
struct RefMaterial {             // prog.cl:1-5, main.cpp:92-112 — 80 B
    RefFloat3 kd, ks, emission, F0;
    float n, shininess;
    int type;                    // 0 diffuse+lobe, 1 mirror, 2 glass, 3 emitter
    int pad_;
};
struct RefRay {                  // prog.cl:7-9, main.cpp:114-117 — 32 B
    RefFloat3 P, D;
};
struct RefTriangle {             // prog.cl:18-21, main.cpp:139-182 — 80 B
    RefFloat3 r1, r2, r3, N;
    uint16_t mati;
    uint16_t pad_[7];
};
struct RefBBox {                 // prog.cl:23-25 — 32 B
    RefFloat3 bl, tr;
};
struct RefNode {                 // prog.cl:27-30, main.cpp:184-193 — 48 B
    int from, to;                // int2 trii: leaf = [from,to), interior = from < 0
    int pad_[2];
    RefBBox bbox;
};
struct RefCamera {               // prog.cl:32-35, main.cpp:306-348 — 80 B, passed by value
    RefFloat3 eye, lookat, up, right;
    float XM, YM;
    float pad_[2];
};

#if defined(__cplusplus)
static_assert(sizeof(RefMaterial) == 80, "Material ABI");
static_assert(sizeof(RefRay) == 32, "Ray ABI");
static_assert(sizeof(RefTriangle) == 80, "Triangle ABI");
static_assert(sizeof(RefNode) == 48, "Node ABI");
static_assert(sizeof(RefCamera) == 80, "Camera ABI");
static_assert(sizeof(RefBBox) == 32, "BBox ABI");
#endif

// ---------------------------------------------------------------- device-internal layouts
//
// Staged triangles: 3 × float4 per triangle, structure-of-arrays by PLANE (three arrays of
// float4, each n_tris long) so that a warp testing consecutive triangles, and a lane fetching
// one triangle, both issue full 16-byte vector loads:
//   q0[i] = (r1.x, r1.y, r1.z, N.x)   q1[i] = (r2.x, r2.y, r2.z, N.y)   q2[i] = (r3.x, r3.y, r3.z, N.z)
// mati[i] is a separate uint16 array touched once per ray (for the winning triangle only).
//
// LBVH traversal nodes are defined next to their builder: lbvh.cu (binary: 4 float4 holding both children's boxes and
// links; 4-ary collapse: 8 float4 per node).

#endif  // PTB_TYPES_H_
