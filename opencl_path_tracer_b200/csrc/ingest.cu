// ingest.cu — see ingest.h.  Every operation is the reference's, in the reference's order and precision, written
// with explicit round-to-nearest intrinsics so that no FMA contraction can change a bit:
//   vx = -x (main.cpp:598); rotate_x then rotate_y with float*double products summed in double and narrowed
//   (main.cpp:55-70, 602-603); v*scale+pos in float (main.cpp:604-606); N = normalize((r2-r1)x(r3-r1)) in float with
//   the length through sqrt (main.cpp:146-163).
#include "ingest.h"

namespace {

struct F3 { float x, y, z; };

__device__ F3 xform(const float *verts, int vi, const IngestXform &t) {
    const float x0 = -verts[3 * (size_t)vi], y0 = verts[3 * (size_t)vi + 1], z0 = verts[3 * (size_t)vi + 2];
    // rotate_x(v, pitch)
    const float y1 = __double2float_rn(__dsub_rn(__dmul_rn((double)y0, t.cos_pitch), __dmul_rn((double)z0, t.sin_pitch)));
    const float z1 = __double2float_rn(__dadd_rn(__dmul_rn((double)y0, t.sin_pitch), __dmul_rn((double)z0, t.cos_pitch)));
    // rotate_y(v, yaw)
    const float x2 = __double2float_rn(__dadd_rn(__dmul_rn((double)x0, t.cos_yaw), __dmul_rn((double)z1, t.sin_yaw)));
    const float z2 = __double2float_rn(__dadd_rn(__dmul_rn((double)(-x0), t.sin_yaw), __dmul_rn((double)z1, t.cos_yaw)));
    F3 r;
    r.x = __fadd_rn(__fmul_rn(x2, t.scale[0]), t.pos[0]);
    r.y = __fadd_rn(__fmul_rn(y1, t.scale[1]), t.pos[1]);
    r.z = __fadd_rn(__fmul_rn(z2, t.scale[2]), t.pos[2]);
    return r;
}

__global__ void k_ingest(const float *verts, const int *idx, const int *mat_ids, size_t n_faces, IngestXform t, RefTriangle *out) {
    const size_t f = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_faces) return;
    const F3 r1 = xform(verts, idx[3 * f], t), r2 = xform(verts, idx[3 * f + 1], t), r3 = xform(verts, idx[3 * f + 2], t);
    const float ax = __fsub_rn(r2.x, r1.x), ay = __fsub_rn(r2.y, r1.y), az = __fsub_rn(r2.z, r1.z);
    const float bx = __fsub_rn(r3.x, r1.x), by = __fsub_rn(r3.y, r1.y), bz = __fsub_rn(r3.z, r1.z);
    const float nx = __fsub_rn(__fmul_rn(ay, bz), __fmul_rn(az, by));
    const float ny = __fsub_rn(__fmul_rn(az, bx), __fmul_rn(ax, bz));
    const float nz = __fsub_rn(__fmul_rn(ax, by), __fmul_rn(ay, bx));
    const float len = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(nx, nx), __fmul_rn(ny, ny)), __fmul_rn(nz, nz)));
    float4 *o = reinterpret_cast<float4 *>(out + f);            // 5 x 16 B per record
    o[0] = make_float4(r1.x, r1.y, r1.z, 0.0f);
    o[1] = make_float4(r2.x, r2.y, r2.z, 0.0f);
    o[2] = make_float4(r3.x, r3.y, r3.z, 0.0f);
    o[3] = make_float4(__fdiv_rn(nx, len), __fdiv_rn(ny, len), __fdiv_rn(nz, len), 0.0f);
    const unsigned mati = (unsigned)(unsigned short)(t.mat_offset + mat_ids[f]);   // cl_ushort conversion wraps
    o[4] = make_float4(__uint_as_float(mati), 0.0f, 0.0f, 0.0f);
}

}  // namespace

void ingest_launch(const float *verts, const int *idx, const int *mat_ids, size_t n_faces, IngestXform x, RefTriangle *out,
                   cudaStream_t st) {
    if (n_faces == 0) return;
    k_ingest<<<(unsigned)((n_faces + 255) / 256), 256, 0, st>>>(verts, idx, mat_ids, n_faces, x, out);
}
