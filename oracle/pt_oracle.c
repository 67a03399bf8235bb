/*
 * pt_oracle.c — TEST INFRASTRUCTURE (see pt_oracle.h).  Plain-C restatement of prog.cl's hot path.
 * Every function cites the reference lines it follows.  Written to be read next to prog.cl, but it is
 * not a copy: scalar C with explicit temporaries, one accumulating closest-hit record instead of Hit
 * structs passed by value, and the per-pixel state handled through the pto_scene view.
 */
#define _GNU_SOURCE
#include "pt_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ---------------------------------------------------------------- float3 helpers (OpenCL built-ins) */
typedef struct { float x, y, z; } v3;
static inline v3 V(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 from_f3(pto_f3 a) { return V(a.x, a.y, a.z); }
static inline pto_f3 to_f3(v3 a) { pto_f3 r = {a.x, a.y, a.z, 0.0f}; return r; }
static inline v3 add(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 sub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 mul(v3 a, v3 b) { return V(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline v3 scale(v3 a, float s) { return V(a.x * s, a.y * s, a.z * s); }
static inline v3 divs(v3 a, float s) { return V(a.x / s, a.y / s, a.z / s); }
static inline v3 neg(v3 a) { return V(-a.x, -a.y, -a.z); }
static inline v3 ssub(float s, v3 a) { return V(s - a.x, s - a.y, s - a.z); }
static inline float dot(v3 a, v3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
static inline v3 cross(v3 a, v3 b) { return V(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
static inline v3 normalize(v3 a) { return divs(a, sqrtf(dot(a, a))); }
static inline float cr_pow(float a, float b) { return (float)pow((double)a, (double)b); }
static inline float cr_cos(float a) { return (float)cos((double)a); }
static inline float cr_sin(float a) { return (float)sin((double)a); }

/* ---------------------------------------------------------------- prog.cl:72-77 */
float pto_rand(int32_t *seed) {
    uint64_t n = (uint64_t)(int64_t)(*seed);          /* ulong n = (*seed): sign-extends a negative int */
    n = (n * 48271u) % 2147483647u;
    *seed = (int32_t)n;
    return (float)n / 2147483647.0f;
}

/* main.cpp:45,522-526: default-constructed std::minstd_rand0 = Lehmer 16807 mod 2^31-1, seed 1 */
void pto_minstd_rand0(int32_t *out, int64_t count, int64_t skip) {
    uint64_t x = 1;
    for (int64_t i = 0; i < skip + count; ++i) {
        x = (x * 16807u) % 2147483647u;
        if (i >= skip) out[i - skip] = (int32_t)x;
    }
}

/* ---------------------------------------------------------------- prog.cl:82-92 + 384-389 */
void pto_gen_ray_pixel(const pto_scene *s, int32_t id) {
    float rnd1 = pto_rand(&s->rnds[id]);               /* left to right: x jitter first */
    float rnd2 = pto_rand(&s->rnds[id]);
    int X = (int)s->cam.XM, Y = (int)s->cam.YM;
    float x = (float)(id % X) + rnd1;
    float y = (float)(id / X) + rnd2;
    v3 right = scale(from_f3(s->cam.right), 2.0f * x / (float)X - 1.0f);
    v3 up = scale(from_f3(s->cam.up), 2.0f * y / (float)Y - 1.0f);
    v3 p = add(add(from_f3(s->cam.lookat), right), up);
    v3 eye = from_f3(s->cam.eye);
    s->rays[id].P = to_f3(eye);
    s->rays[id].D = to_f3(normalize(sub(p, eye)));
}

/* ---------------------------------------------------------------- prog.cl:94-122 */
typedef struct { float t; v3 P, N; int tri; } hit_t;   /* t = -1: no hit (init_Hit) */

static inline int triangle_intersect(const pto_triangle *tri, v3 P, v3 D, float *t_out, v3 *p_out, pto_counters *c) {
    v3 r1 = from_f3(tri->r1), r2 = from_f3(tri->r2), r3 = from_f3(tri->r3), N = from_f3(tri->N);
    if (c) c->tri_tests++;
    float t = dot(sub(r1, P), N) / dot(D, N);
    if (t < 0) return 0;
    if (c) c->tri_stage1++;
    v3 p = add(P, scale(D, t));
    if (!(dot(cross(sub(r2, r1), sub(p, r1)), N) >= 0)) return 0;
    if (c) c->tri_stage2++;
    if (!(dot(cross(sub(r3, r2), sub(p, r2)), N) >= 0)) return 0;
    if (c) c->tri_stage3++;
    if (!(dot(cross(sub(r1, r3), sub(p, r3)), N) >= 0)) return 0;
    *t_out = t;
    *p_out = p;
    return 1;
}

/* first_intersect over [from,to): t > 0 and strictly closer than the best of THIS call */
static hit_t first_intersect(const pto_triangle *tris, int from, int to, v3 P, v3 D, pto_counters *c) {
    hit_t best;
    best.t = -1.0f; best.P = V(0, 0, 0); best.N = V(0, 0, 0); best.tri = -1;
    for (int i = from; i < to; ++i) {
        float t;
        v3 p;
        if (triangle_intersect(&tris[i], P, D, &t, &p, c) && t > 0 && (best.t < 0 || t < best.t)) {
            best.t = t; best.P = p; best.N = from_f3(tris[i].N); best.tri = i;
        }
    }
    return best;
}

/* ---------------------------------------------------------------- prog.cl:123-143 */
static inline int bbox_intersection(const pto_node *node, v3 P, v3 D, float *tmin, float *tmax, pto_counters *c) {
    if (c) c->box_tests++;
    float tx1 = (node->bl.x - P.x) / D.x, tx2 = (node->tr.x - P.x) / D.x;
    float lo = fminf(tx1, tx2), hi = fmaxf(tx1, tx2);
    float ty1 = (node->bl.y - P.y) / D.y, ty2 = (node->tr.y - P.y) / D.y;
    lo = fmaxf(lo, fminf(ty1, ty2));
    hi = fminf(hi, fmaxf(ty1, ty2));
    float tz1 = (node->bl.z - P.z) / D.z, tz2 = (node->tr.z - P.z) / D.z;
    lo = fmaxf(lo, fminf(tz1, tz2));
    hi = fminf(hi, fmaxf(tz1, tz2));
    *tmin = lo;
    *tmax = hi;
    return hi >= lo;
}

/* ---------------------------------------------------------------- prog.cl:144-184 + 271-290 */
static hit_t kd_intersect(const pto_scene *s, v3 P, v3 D, pto_counters *c) {
    hit_t best;
    best.t = -1.0f; best.P = V(0, 0, 0); best.N = V(0, 0, 0); best.tri = -1;
    int stack[300];
    for (int o = 0; o < s->kd_tree_shift_size; ++o) {
        const int shift = s->kd_tree_shift[o];
        int ptr = 1 + shift, sp = 0;
        float tmin = 999999.0f, dist = 0.0f, tmax = -999999.0f;
        for (;;) {
            const pto_node *node = &s->kd_tree[ptr];
            int pop = 1;
            if (bbox_intersection(node, P, D, &dist, &tmax, c) && tmax >= 0 && !(dist > tmin)) {
                if (node->from < 0) {                       /* interior: push right, go left */
                    if (sp < 300) stack[sp++] = 2 * (ptr - shift) + 1 + shift;
                    ptr = 2 * (ptr - shift) + shift;
                    pop = 0;
                } else {                                    /* leaf */
                    hit_t h = first_intersect(s->tris, node->from, node->to, P, D, c);
                    if (h.t > 0 && (best.t < 0 || h.t < best.t)) { tmin = h.t; best = h; }
                }
            }
            if (pop) {
                if (sp == 0) break;
                ptr = stack[--sp];                          /* stack_pop */
            }
        }
    }
    return best;
}

float pto_closest_hit(const pto_scene *s, const float *o, const float *d, int32_t *tri_index) {
    v3 P = V(o[0], o[1], o[2]), D = V(d[0], d[1], d[2]);
    hit_t h = s->brute ? first_intersect(s->tris, 0, s->tris_size, P, D, NULL) : kd_intersect(s, P, D, NULL);
    if (tri_index) *tri_index = h.tri;
    return h.t;
}

/* ---------------------------------------------------------------- prog.cl:186-245 */
static void orthonormal_base(v3 v1, v3 *v2_out, v3 *v3_out) {
    const float E = 0.001f;
    v3 v2;
    if (fabsf(v1.x) <= E && fabsf(v1.z) <= E) {
        float rl = 1 / sqrtf(v1.y * v1.y + v1.z * v1.z);
        v2 = V(0, -v1.z * rl, v1.y * rl);
    } else {
        float rl = 1 / sqrtf(v1.x * v1.x + v1.z * v1.z);
        v2 = V(-v1.z * rl, 0, v1.x * rl);
    }
    *v2_out = v2;
    *v3_out = cross(v1, v2);
}
static void new_ray_diffuse(v3 hitP, v3 N, float rnd1, float rnd2, v3 *P, v3 *D) {
    v3 X, Y = N, Z;
    orthonormal_base(Y, &Z, &X);
    float r = sqrtf(rnd1);
    float theta = (float)(2 * M_PI * rnd2);              /* double product, narrowed */
    float x = r * cr_cos(theta), y = r * cr_sin(theta), z = sqrtf(1 - rnd1);
    *D = normalize(add(add(scale(X, x), scale(Y, z)), scale(Z, y)));
    *P = add(hitP, scale(Y, 0.001f));
}
static v3 fresnel(const pto_material *m, v3 N, v3 oldD) {
    float cosa = fabsf(dot(N, oldD));
    v3 F0 = from_f3(m->F0);
    return add(F0, scale(ssub(1, F0), cr_pow(1 - cosa, 5)));
}
static void new_ray_specular(v3 hitP, v3 N, v3 oldD, v3 *P, v3 *D) {
    float cosa = dot(N, oldD);
    *D = normalize(sub(oldD, scale(scale(N, cosa), 2.0f)));
    *P = add(hitP, scale(N, 0.001f));
}
static void new_ray_refractive(const pto_material *m, v3 hitP, v3 N, v3 oldD, int *in, float rnd, v3 *P, v3 *D) {
    float n = *in ? 1.0f / m->n : m->n;
    float cosa = dot(neg(oldD), N);
    float disc = 1.0f - (1.0f - cosa * cosa) / n / n;
    v3 F = fresnel(m, N, oldD);
    float prob = (F.x + F.y + F.z) / 3.0f;
    if (disc > 0 && rnd > prob) {
        *in = !*in;
        *P = sub(hitP, scale(N, 0.001f));
        *D = normalize(add(divs(oldD, n), scale(N, cosa / n - sqrtf(disc))));
    } else {
        new_ray_specular(hitP, N, oldD, P, D);
    }
}

/* ---------------------------------------------------------------- prog.cl:247-258, 264-269 */
static float srgb1(float a) { return a <= 0.00304f ? 12.92f * a : 1.055f * cr_pow(a, 0.4167f) - 0.055f; }
float pto_reinhard(const float *c, float *out) {
    float L = 0.2126f * c[0] + 0.7152f * c[1] + 0.0722f * c[2];
    float L2 = L / (1 + L);
    out[0] = srgb1(c[0] * L2 / L);
    out[1] = srgb1(c[1] * L2 / L);
    out[2] = srgb1(c[2] * L2 / L);
    out[3] = 1.0f;
    return L;
}

/* ---------------------------------------------------------------- prog.cl:292-381 */
void pto_trace_ray_pixel(const pto_scene *s, int32_t gx, int32_t gy, int32_t current_sample, pto_counters *c) {
    const int id = gy * s->width + gx;
    v3 fL = V(1, 1, 1), fB = V(1, 1, 1), fS = V(1, 1, 1), fR = V(1, 1, 1), color = V(0, 0, 0);
    if (current_sample == 0) s->colors[id] = to_f3(color);
    int in = 0;
    v3 eye = from_f3(s->cam.eye);
    for (int current = 0; current < s->iterations; ++current) {
        if (c) c->rays++;
        v3 P = from_f3(s->rays[id].P), D = from_f3(s->rays[id].D);
        hit_t hit = s->brute ? first_intersect(s->tris, 0, s->tris_size, P, D, c) : kd_intersect(s, P, D, c);
        if (!(hit.t > 0)) break;                               /* no environment light, prog.cl:367-376 */
        const pto_material *m = &s->materials[s->tris[hit.tri].mati];
        if (s->iterations == 1) color = add(from_f3(m->kd), from_f3(m->emission));
        v3 N = hit.N;
        if (dot(D, N) > 0) N = neg(N);
        v3 nP = P, nD = D;
        if (m->type == 0) {                                     /* prog.cl:329-340 */
            float r1 = pto_rand(&s->rnds[id]);
            float r2 = pto_rand(&s->rnds[id]);
            new_ray_diffuse(hit.P, N, r1, r2, &nP, &nD);
            float cos_theta = dot(nD, N);
            fL = mul(fL, scale(from_f3(m->kd), fmaxf(0.0f, cos_theta)));
            v3 halfway = normalize(add(normalize(sub(eye, hit.P)), nD));
            float cos_delta = dot(N, halfway);
            fB = mul(fB, scale(from_f3(m->ks), cr_pow(fmaxf(0.0f, cos_delta), m->shininess)));
        } else if (m->type == 1) {                              /* prog.cl:341-345 */
            new_ray_specular(hit.P, N, D, &nP, &nD);
            fS = mul(fS, fresnel(m, N, D));
        } else if (m->type == 2) {                              /* prog.cl:346-357 */
            int before = in;
            float rnd = pto_rand(&s->rnds[id]);
            new_ray_refractive(m, hit.P, N, D, &in, rnd, &nP, &nD);
            v3 F = fresnel(m, N, D);
            float prob = (F.x + F.y + F.z) / 3.0f;
            if (before != in) fR = scale(mul(fR, ssub(1, F)), 1 / (1 - prob));
            else fR = scale(mul(fR, F), 1 / prob);
        } else if (m->type == 3) {                              /* prog.cl:358-366 */
            float intensity = fmaxf(0.0f, dot(neg(D), N));
            float r1 = pto_rand(&s->rnds[id]);
            float r2 = pto_rand(&s->rnds[id]);
            new_ray_diffuse(hit.P, N, r1, r2, &nP, &nD);
            color = add(color, scale(mul(mul(mul(from_f3(m->emission), add(fL, fB)), fS), fR), intensity));
        }
        s->rays[id].P = to_f3(nP);
        s->rays[id].D = to_f3(nD);
    }
    v3 old = from_f3(s->colors[id]);
    v3 mean = divs(add(scale(old, (float)current_sample), color), (float)(current_sample + 1));   /* prog.cl:379 */
    s->colors[id] = to_f3(mean);
    if (s->tex) {
        float rgb[3] = {mean.x, mean.y, mean.z};
        pto_reinhard(rgb, s->tex + 4 * (size_t)id);
    }
}

/* ---------------------------------------------------------------- main.cpp:683-687, threaded over rows */
typedef struct { const pto_scene *s; int first, spp, tid, nthreads; pto_counters c; } job_t;
static void *worker(void *arg) {
    job_t *j = (job_t *)arg;
    const pto_scene *s = j->s;
    memset(&j->c, 0, sizeof j->c);
    /* pixels are independent (prog.cl touches only index id), so a thread may run all samples of its rows */
    for (int y = j->tid; y < s->height; y += j->nthreads)
        for (int k = 0; k < j->spp; ++k)
            for (int x = 0; x < s->width; ++x) {
                pto_gen_ray_pixel(s, y * s->width + x);
                pto_trace_ray_pixel(s, x, y, j->first + k, &j->c);
            }
    return NULL;
}
void pto_render(const pto_scene *s, int32_t first_sample, int32_t spp, int32_t threads, pto_counters *total) {
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    job_t jobs[256];
    pthread_t th[256];
    for (int t = 0; t < threads; ++t) {
        jobs[t].s = s; jobs[t].first = first_sample; jobs[t].spp = spp; jobs[t].tid = t; jobs[t].nthreads = threads;
        if (t > 0) pthread_create(&th[t], NULL, worker, &jobs[t]);
    }
    worker(&jobs[0]);
    for (int t = 1; t < threads; ++t) pthread_join(th[t], NULL);
    if (total)
        for (int t = 0; t < threads; ++t) {
            total->rays += jobs[t].c.rays; total->box_tests += jobs[t].c.box_tests; total->tri_tests += jobs[t].c.tri_tests;
            total->tri_stage1 += jobs[t].c.tri_stage1; total->tri_stage2 += jobs[t].c.tri_stage2; total->tri_stage3 += jobs[t].c.tri_stage3;
        }
}

/* ---------------------------------------------------------------- prog.cl:259-263 */
void pto_filmic(const float *rgb3, float *rgba4) {
    for (int k = 0; k < 3; ++k) {
        float c = fmaxf(0.0f, rgb3[k] - 0.004f);
        rgba4[k] = (c * (c * 6.2f + 0.5f)) / (c * (c * 6.2f + 1.7f) + 0.06f);
    }
    rgba4[3] = 1.0f;
}

/* ---------------------------------------------------------------- prog.cl:391-427 (launch: main.cpp:665-668, commented there)
 * One work-item per pixel over (width, height): 3x3 neighbourhood through LINEAR ids (y-1+i)*width + (x-1+j), selection
 * sort by grey value (strict '>' : the first maximum wins), element 4 of the sorted order, filmic tone map.  The guard is
 * the reference's x>0 && y>0 (x<width, y<height hold for every work-item); work-items whose neighbourhood would run past
 * the end of colors[] (the last scanline, and the last pixel of the one before) read out of range in the reference and
 * are skipped here.  tex: W*H*4 floats, untouched where nothing is written. */
void pto_filt_im(const pto_f3 *colors, float *tex, int32_t width, int32_t height) {
    const int64_t n = (int64_t)width * height;
    for (int y = 0; y < height; ++y)
        for (int x = 0; x < width; ++x) {
            if (!(x > 0 && y > 0)) continue;
            if ((int64_t)(y + 1) * width + (x + 1) >= n) continue;
            pto_f3 arr[9];
            float grey[9];
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) arr[i * 3 + j] = colors[(int64_t)(y - 1 + i) * width + (x - 1 + j)];
            for (int i = 0; i < 9; ++i) grey[i] = (arr[i].x + arr[i].y + arr[i].z) / 3.0f;
            for (int j = 9; j > 1; j--) {
                int maxi = 0;
                for (int i = 1; i < j; ++i)
                    if (grey[i] > grey[maxi]) maxi = i;
                float tg = grey[j - 1]; grey[j - 1] = grey[maxi]; grey[maxi] = tg;
                pto_f3 ta = arr[j - 1]; arr[j - 1] = arr[maxi]; arr[maxi] = ta;
            }
            const float med[3] = {arr[4].x, arr[4].y, arr[4].z};
            pto_filmic(med, tex + 4 * ((int64_t)y * width + x));
        }
}

int pto_struct_sizes(int32_t *o) {
    o[0] = sizeof(pto_material); o[1] = sizeof(pto_ray); o[2] = sizeof(pto_triangle);
    o[3] = sizeof(pto_node); o[4] = sizeof(pto_camera); o[5] = sizeof(pto_f3);
    return 6;
}
