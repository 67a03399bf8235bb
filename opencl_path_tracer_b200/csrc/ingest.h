// ingest.h — scene ingest on the device: the vertex loop of Scene::add_Obj (main.cpp:589-612) and the Triangle
// constructor (main.cpp:144-166) as one kernel writing the reference's 80-byte Triangle records.
#ifndef INGEST_H_
#define INGEST_H_
#include <cuda_runtime.h>
#include <stddef.h>
#include "ptb_types.h"

struct IngestXform {
    double cos_pitch, sin_pitch, cos_yaw, sin_yaw;   // of the float angle deg/180*3.141593, evaluated by the host's libm
    float scale[3], pos[3];
    int mat_offset;
};

// verts: 3 floats per vertex as parsed; idx: 3 vertex indices per face; mat_ids: one per face (all device pointers)
void ingest_launch(const float *verts, const int *idx, const int *mat_ids, size_t n_faces, IngestXform x, RefTriangle *out,
                   cudaStream_t st);
#endif
