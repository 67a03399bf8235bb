/*
 * pt_oracle.h — TEST INFRASTRUCTURE.  CPU restatement of the reference's path-trace hot path
 * (/root/reference/prog.cl) in plain C.  Only tests/, __graft_entry__.smoke() and bench.py's CPU
 * baseline leg may load this; the product (libptb200.so) never does.
 *
 * Parity status: PINNED.  The restatement is checked (tests/test_oracle.py) against outputs of the
 * reference's own prog.cl + main.cpp compiled where they lie (oracle/_ref/ref_cpu, built by
 * oracle/Makefile) — live when /root/reference is present, and against fixtures of those outputs
 * committed under tests/golden/ (made by tests/golden/make_golden.py).  The reference ships no tests
 * or golden vectors of its own.
 *
 * Canonical arithmetic (same choices as oracle/oclshim.h; OpenCL only bounds these in ulps):
 * IEEE single +,-,*,/ and sqrtf with no FMA contraction (-ffp-contract=off), dot = (x*x'+y*y')+z*z',
 * normalize(v) = v / sqrtf(dot(v,v)), pow/cos/sin = the correctly rounded float of the double function,
 * fmin/fmax return the non-NaN operand, double-rand call sites evaluated left to right.
 *
 * Buffers are the kernel-argument surface byte for byte (prog.cl:1-35).
 */
#ifndef PT_ORACLE_H_
#define PT_ORACLE_H_
#include <stdint.h>

typedef struct { float x, y, z, w; } pto_f3;                              /* OpenCL float3 = 16 bytes   */
typedef struct { pto_f3 kd, ks, emission, F0; float n, shininess; int32_t type, pad; } pto_material;   /* 80 B */
typedef struct { pto_f3 P, D; } pto_ray;                                  /* 32 B */
typedef struct { pto_f3 r1, r2, r3, N; uint16_t mati, pad[7]; } pto_triangle;                          /* 80 B */
typedef struct { int32_t from, to, pad[2]; pto_f3 bl, tr; } pto_node;     /* 48 B */
typedef struct { pto_f3 eye, lookat, up, right; float XM, YM, pad[2]; } pto_camera;                    /* 80 B */

typedef struct {
    uint64_t rays, box_tests, tri_tests, tri_stage1, tri_stage2, tri_stage3;
} pto_counters;

typedef struct {
    const pto_triangle *tris; int32_t tris_size;
    const pto_material *materials;
    const pto_node *kd_tree; const int32_t *kd_tree_shift; int32_t kd_tree_shift_size;
    pto_ray *rays; int32_t *rnds; pto_f3 *colors;
    float *tex;                      /* W*H*4 floats or NULL: the write_imagef target (prog.cl:380) */
    pto_camera cam;
    int32_t width, height, iterations;
    int32_t brute;                   /* 1: whole-scene first_intersect (prog.cl:318) instead of kd_intersect (prog.cl:319) */
} pto_scene;

float pto_rand(int32_t *seed);                                            /* prog.cl:72-77 */
void pto_gen_ray_pixel(const pto_scene *s, int32_t id);                   /* prog.cl:384-389 for one work-item */
void pto_trace_ray_pixel(const pto_scene *s, int32_t gx, int32_t gy, int32_t current_sample, pto_counters *c);  /* prog.cl:292-381 */
/* what main.cpp:683-687 enqueues, `spp` times starting at sample index `first_sample`, rows split over `threads` */
void pto_render(const pto_scene *s, int32_t first_sample, int32_t spp, int32_t threads, pto_counters *total);
/* closest hit only (for the LBVH == brute-force ray tests): returns t (or -1) and the triangle index */
float pto_closest_hit(const pto_scene *s, const float *origin3, const float *dir3, int32_t *tri_index);
void pto_minstd_rand0(int32_t *out, int64_t count, int64_t skip);         /* main.cpp:45,522-526 */
float pto_reinhard(const float *rgb3, float *rgba4);                      /* prog.cl:247-258,264-269 */
void pto_filmic(const float *rgb3, float *rgba4);                         /* prog.cl:259-263 */
void pto_filt_im(const pto_f3 *colors, float *tex, int32_t width, int32_t height);   /* prog.cl:391-427 over (width, height) */
int pto_struct_sizes(int32_t *out6);                                      /* Material Ray Triangle Node Camera f3 */
#endif
