/*
 * ptb200.h — C-ABI of the B200 path-trace device shim (libptb200.so).
 *
 * This is the drop-in boundary of the hot path.  The reference host
 * (/root/reference/main.cpp) talks to its device only through a small subset
 * of the OpenCL 1.x C++ wrapper; every entry point below replaces one of those
 * call sites (cited as main.cpp:line) or one kernel of prog.cl (prog.cl:line).
 * A header-only `CL/cl.hpp` look-alike (opencl_path_tracer_b200/shim/CL/cl.hpp)
 * forwards the cl:: classes to these functions, so main.cpp compiles and runs
 * unchanged on top of them.  INTEGRATION.md shows the binding.
 *
 * Conventions (mirroring how the reference uses OpenCL):
 *   - plain C types only; 0 == PTB_SUCCESS == CL_SUCCESS, negative = error.
 *   - the reference ignores almost every return code (main.cpp:506-682), so an
 *     error is also recorded in a sticky per-process string (ptb_last_error)
 *     and printed to stderr once; later calls become no-ops returning the error.
 *   - one in-order stream per process (the reference has one in-order
 *     cl::CommandQueue, main.cpp:506); one process drives one GPU.  Like that
 *     queue's host side the library keeps one state per process and is NOT
 *     thread-safe: call it from one thread (the reference does, from GLUT's).
 *   - buffer layouts are the reference's kernel ABI structs, byte for byte
 *     (prog.cl:1-35 / main.cpp:92-348): Material 80 B, Ray 32 B, Triangle 80 B,
 *     Node 48 B, Camera 80 B, colors float3 = 16 B per pixel, rnds int per pixel.
 *   - there is NO CPU fallback: without a CUDA device ptb_init fails.
 */
#ifndef PTB200_H_
#define PTB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PTB_SUCCESS 0
#define PTB_ERR_NO_DEVICE (-1)     /* CL_DEVICE_NOT_FOUND                       */
#define PTB_ERR_INVALID_VALUE (-30)/* CL_INVALID_VALUE                          */
#define PTB_ERR_INVALID_MEM (-38)  /* CL_INVALID_MEM_OBJECT                     */
#define PTB_ERR_INVALID_KERNEL (-46)/* CL_INVALID_KERNEL_NAME                   */
#define PTB_ERR_INVALID_ARGS (-52) /* CL_INVALID_KERNEL_ARGS                    */
#define PTB_ERR_OUT_OF_MEMORY (-4) /* CL_MEM_OBJECT_ALLOCATION_FAILURE          */
#define PTB_ERR_DEVICE (-9999)     /* a CUDA runtime call failed (see last_error)*/

/* Opaque buffer handle; 0 is the null handle.  Replaces cl::Buffer / cl::ImageGL. */
typedef uint64_t ptb_mem;

/* Kernel ids.  Replace cl::Kernel(program,"gen_ray"|"trace_ray") (main.cpp:637,647)
 * and the never-launched "filt_im" (main.cpp:665-668, prog.cl:391-427). */
enum { PTB_KERNEL_GEN_RAY = 0, PTB_KERNEL_TRACE_RAY = 1, PTB_KERNEL_FILT_IM = 2, PTB_KERNEL_COUNT = 3 };

/* ---- bring-up: replaces cl::Platform::get / getDevices / cl::Context /
 *      cl::Program::build / cl::CommandQueue (main.cpp:460-506).  Kernels are
 *      AOT-compiled for sm_100a, so "build" is a no-op that reports success. ---- */
int ptb_init(int device_ordinal);          /* idempotent; -1 = take LOCAL_RANK / device 0 */
int ptb_shutdown(void);                    /* frees every buffer and the accel structure */
/* replaces the clGetDeviceInfo dump of Scene::list_info (main.cpp:389-455).
 * what: 0 name, 1 hardware version, 2 driver version, 3 "language" version, 4 extensions */
int ptb_device_string(int what, char *dst, size_t cap);
int ptb_device_compute_units(void);        /* CL_DEVICE_MAX_COMPUTE_UNITS → SM count */
const char *ptb_backend_name(void);        /* "cuda-sm100a" for the product library */

/* ---- memory: replaces cl::Buffer(context, flags, bytes) (main.cpp:508,509,520,619,623,627,632),
 *      cl::ImageGL (main.cpp:519) and blocking enqueueWriteBuffer (main.cpp:526,620,624,628,633).
 *      Handles are reference counted like cl::Memory. ---- */
ptb_mem ptb_alloc(size_t bytes);
ptb_mem ptb_alloc_image(int width, int height); /* W×H RGBA float image, headless stand-in for the GL texture */
int ptb_retain(ptb_mem m);
int ptb_release(ptb_mem m);
int ptb_write(ptb_mem m, size_t offset, size_t bytes, const void *src);   /* blocking */
int ptb_read(ptb_mem m, size_t offset, size_t bytes, void *dst);          /* blocking; the reference's read-back is dead code (main.cpp:690,728) */
size_t ptb_mem_size(ptb_mem m);
/* Scene ingest on the device — replaces the vertex loop of Scene::add_Obj (main.cpp:589-612: negate x, rotate_x(pitch),
 * rotate_y(yaw), *scale+pos) and the Triangle constructor (main.cpp:144-166: unit normal) for one triangulated mesh:
 * vertices = attrib.vertices (3 floats each), indices = 3 vertex indices per face, mat_ids = material per face, all HOST
 * pointers as tiny_obj_loader leaves them.  Writes n_faces 80-byte Triangle records into `tris` starting at record
 * first_tri, bit-identical to what the reference host computes.  The records stay in face order (no host tree is
 * built for them: pass kd_tree_shift_size = 0 or keep the host trees for other objects only). Blocking. */
int ptb_ingest_mesh(ptb_mem tris, size_t first_tri, const float *vertices, size_t n_verts, const int *indices,
                    const int *mat_ids, size_t n_faces, const float *pos, const float *scale, float pitch_deg,
                    float yaw_deg, int mat_offset);
/* extensions for hosts that own device memory already (torch tensors, NCCL buffers).  The library cannot see writes
 * that bypass it: what it derives from a buffer's CONTENTS (the staged SoA copy + LBVH of a triangle buffer, the largest
 * material index, the "all RNG states are non-negative" scan of rnds) is computed at the first launch after a content
 * change it knows of — ptb_write, ptb_ingest_mesh, ptb_device_ptr (the caller may write through the pointer) — and
 * reused until the next one.  A host that rewrites wrapped memory itself (a torch op, a peer GPU, NCCL) calls
 * ptb_invalidate(mem) afterwards; without it the next launch renders the scene as it was when first staged. */
ptb_mem ptb_wrap_device(void *device_ptr, size_t bytes);  /* non-owning handle */
void *ptb_device_ptr(ptb_mem m);
int ptb_invalidate(ptb_mem m);                             /* "the contents changed behind the library's back" */
/* Peer memory over NVLink (one process per GPU): the owner exports a ptb_alloc'ed buffer as a 64-byte CUDA IPC handle,
 * a process on another GPU of the box imports it and binds it as a kernel argument like any other buffer.  Used by the
 * tile split: every rank's path kernel stores its pixels' colors[] straight into rank 0's frame, so the
 * gather rides on the kernel's own epilogue stores instead of a collective after it.  No equivalent in the reference
 * (single device, main.cpp:496-506). */
#define PTB_IPC_HANDLE_BYTES 64
int ptb_ipc_export(ptb_mem m, void *handle_out /* PTB_IPC_HANDLE_BYTES */);
ptb_mem ptb_ipc_import(const void *handle_in /* PTB_IPC_HANDLE_BYTES */, size_t bytes);   /* release with ptb_release */

/* ---- kernels: replaces cl::Kernel::setArg (main.cpp:638-640, 648-660) and
 *      enqueueNDRangeKernel / finish (main.cpp:642,674,675,678,680).
 *      Argument lists are prog.cl:384-386 (gen_ray) and prog.cl:292-304 (trace_ray). ---- */
int ptb_kernel_id(const char *name);       /* "gen_ray" → 0, "trace_ray" → 1, "filt_im" → 2, else <0 */
int ptb_set_arg(int kernel, int index, size_t bytes, const void *value);  /* by-value args: int, Camera */
int ptb_set_arg_mem(int kernel, int index, ptb_mem m);                    /* buffer / image args      */
int ptb_launch(int kernel, size_t global_x, size_t global_y);             /* asynchronous, in order   */
int ptb_finish(void);
const char *ptb_last_error(void);          /* "" when no error so far */
void ptb_clear_error(void);

/* ---- extensions beyond the reference surface (all optional for a drop-in host) ---- */

/* Closest-hit structure used by trace_ray / ptb_render. */
enum {
    PTB_ACCEL_AUTO = 0,        /* brute force for tiny scenes, LBVH otherwise                         */
    PTB_ACCEL_BRUTE = 1,       /* first_intersect over all triangles (prog.cl:113-122, :318)            */
    PTB_ACCEL_REFBVH = 2,      /* the uploaded Node[]/kd_tree_shift as is (prog.cl:144-184)             */
    PTB_ACCEL_LBVH = 3         /* GPU-built Morton LBVH, == brute force incl. index tie-break           */
};
/* Arithmetic variant of the path kernels. */
enum {
    PTB_MATH_PARITY = 0,       /* no FMA contraction, IEEE div/sqrt, op order of prog.cl                */
    PTB_MATH_FAST = 1          /* FMA, hoisted reciprocals, precomputed edge planes                     */
};
/* Kernel organisation of ptb_render. */
enum {
    PTB_PIPE_MEGAKERNEL = 0,   /* one persistent kernel, path regeneration inside the loop              */
    PTB_PIPE_WAVEFRONT = 1,    /* generate / intersect / shade kernels over compacted ray queues        */
    PTB_PIPE_AUTO = 2,         /* default — the faster one by measurement               */
    PTB_PIPE_POOL = 3          /* LBVH scenes: ONE persistent kernel whose warps regroup a pool of ray slots (traversal state in
                                  shared memory) by phase — node / leaf / shade — every step; no queues, no host loop:
                                  ptb_render is one launch.  Other scenes: as PTB_PIPE_MEGAKERNEL            */
};
int ptb_set_accel(int accel);
int ptb_set_math(int math);
int ptb_set_pipeline(int pipe);
/* wavefront pipeline: number of interleaved sub-wavefronts (independent ray queues on separate streams that fill each
 * other's trace tails), 1..4, default 2 (env PTB_WF_SUBS).  The image does not depend on it. */
int ptb_set_wf_subs(int n);
/* LBVH traversal (wavefront and pool kernel): read the 4-ary nodes in their 64-byte quantised form (8-bit child boxes against the node's
 * own origin and power-of-two step, rounded outward: conservative, so the closest hit is unchanged) — four 16-byte fetches per node
 * instead of five fetches of the 128-byte float node.  0 = never, 1 = always (both pipelines), 2 (default) = the pool kernel when the
 * float nodes exceed 8 MB, the wavefront never (both by measurement; env PTB_QUANT_NODES).  The image does not depend on it. */
int ptb_set_quant_nodes(int on);
/* The reference enqueues gen_ray then trace_ray for every sample (main.cpp:683-687).  With fusion on
 * (the default) a gen_ray launch is deferred and folded into the trace_ray launch that follows it on
 * the same rays/rnds buffers, so the primary rays never travel through rays[] (it is scratch in the
 * reference: rewritten by gen_ray every sample, read by nobody else).  Any read, write, finish or
 * differing launch in between runs the deferred gen_ray as its own kernel first.  0 = run the two
 * kernels literally.  Results are identical either way. */
int ptb_set_fuse_gen(int on);
int ptb_set_lbvh_leaf(int max_tris); /* 1..8 triangles per LBVH leaf (default 3); takes effect at the next build */
/* Fast flavour, brute-force scan only: merge coplanar same-material triangle pairs that form a convex quad
 * (a fan-triangulated quad, main.cpp:558) into one scan primitive (default 1).  Never used by PTB_MATH_PARITY. */
int ptb_set_merge_prims(int on);
/* Parity flavour, shared-memory brute-force scan only: put a conservative pre-filter (the fast flavour's packed test with
 * a rigorous error margin) in front of prog.cl:94-112's exact test, which then runs only on the triangles the filter
 * cannot rule out (default 1).  The result is bit-identical either way; 0 evaluates the exact test on every triangle. */
int ptb_set_parity_filter(int on);
int ptb_scan_prims(void);      /* primitives the last fast brute-force launch scanned (== triangle count when nothing merged) */
int ptb_get_accel(void);       /* the structure the next launch will really use (AUTO resolved)         */
int ptb_build_accel(void);     /* force the (lazy) staging + LBVH build now; returns when done          */

/* In-kernel sample loop: renders `spp` further samples of every pixel with the
 * arguments currently set on trace_ray (args 1-6, 8-12; `rays` is not needed),
 * starting at sample index trace_ray.arg10 and advancing the per-pixel RNG
 * states exactly as `spp` pairs of gen_ray+trace_ray launches would
 * (main.cpp:683-687).  colors[] receives the same running mean (prog.cl:379).
 * The image argument (arg 0) is written once at the end when present. */
int ptb_render(int spp);

/* Closest hit only, for arbitrary rays: `rays` holds n_rays Ray records (32 B, prog.cl:7-9), `out` receives n_rays pairs
 * {float t; int triangle} with the scene currently bound to trace_ray's arguments, the selected structure
 * (ptb_set_accel), flavour and pipeline — first_intersect / kd_intersect semantics (prog.cl:113-184): smallest t > 0, ties
 * to the lowest upload index; (-1, -1) on a miss.  Blocking.  The test surface for "every structure == brute force". */
int ptb_closest_hit(ptb_mem rays, size_t n_rays, ptb_mem out);
/* Restrict the next launches to a pixel subset — the multi-GPU tile split
 * (rank r of n renders interleaved groups of `group` consecutive pixel ids:
 * id/group % n == r).  n==1 restores the full image. */
int ptb_set_pixel_shard(int rank, int nranks, int group);
/* spp-split support: accumulate plain per-pixel sums (not the running mean) so
 * ranks can be added: colors[id] += sample; combine with ptb_resolve_sum. */
int ptb_set_accumulate_sum(int on);
int ptb_resolve_sum(ptb_mem colors_sum, int total_spp); /* colors = sum / total_spp in place */

/* Work counters of the launches since the last reset.  rays = bounce-loop
 * iterations executed (prog.cl:317-319); box/tri tests only with ptb_set_count_work(1). */
typedef struct {
    uint64_t rays;
    uint64_t box_tests;
    uint64_t tri_tests;
    uint64_t tri_stage1, tri_stage2, tri_stage3; /* tests that got past t<0 / edge 1 / edge 2 (prog.cl:100-106) */
    uint64_t pixel_samples;
} ptb_counters;
int ptb_set_count_work(int on);
int ptb_get_counters(ptb_counters *out);   /* synchronises the stream */
int ptb_reset_counters(void);

/* Output resolve: tone map prog.cl:247-269 on the linear colors buffer.
 * mode 0 = reinhard+sRGB (what trace_ray writes to the texture, prog.cl:380),
 * mode 1 = filmic (prog.cl:259-263).  Writes W*H RGBA8 (top row first, i.e. flipped:
 * the reference image is bottom-up, prog.cl:86-88) into host memory. L==0 maps to black
 * (the reference produces NaN there, prog.cl:265-268). */
int ptb_resolve_rgba8(ptb_mem colors, int width, int height, int mode, unsigned char *dst_host);
/* The root's epilogue of a render as ONE pass over the frame ("sum -> mean -> tone-map" fused): colors *= 1/divide_by
 * in place (divide_by = total samples of a sum-mode accumulation; 1 = colors already holds the mean, left untouched), the
 * tone-mapped float image into `image` (0 = none; bottom-up and with the reference's NaN for L == 0, exactly what trace_ray
 * writes to its texture, prog.cl:380) and the 8-bit RGBA picture into dst_host (NULL = none; as ptb_resolve_rgba8).  The tone
 * map keeps prog.cl's operations and order (no FMA contraction, pow as the correctly rounded float of the double function). */
int ptb_resolve_frame(ptb_mem colors, int width, int height, int divide_by, int mode, ptb_mem image, unsigned char *dst_host);
/* ptb_resolve_rgba8 WITHOUT ANY ALLOCATION.  No counterpart in the reference; added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments and the
 * same picture as ptb_resolve_rgba8; the device-side staging buffer is a static __device__ array of PTB_NOALLOC_MAX_PIXELS * 4 bytes
 * (8 MiB; every GPU that loads the library holds it until exit) instead of a per-call cudaMalloc.  width * height beyond
 * PTB_NOALLOC_MAX_PIXELS is rejected with PTB_ERR_INVALID_VALUE (use ptb_resolve_rgba8 then).  The call makes no cudaMalloc, cudaFree
 * or allocating library call; like ptb_resolve_rgba8 it returns when dst_host is filled.  Every call shares the one array, which keeps
 * the previous call's contents (each call overwrites all of the width * height pixels it uses, nothing needs clearing): calls that
 * overlap on the GPU — other host threads, other processes' streams are fine, other streams of THIS process are not — race. */
#define PTB_NOALLOC_MAX_PIXELS (2 * 1024 * 1024)
int ptb_resolve_rgba8_noalloc(ptb_mem colors, int width, int height, int mode, unsigned char *dst_host);
/* 3x3 median denoise of prog.cl:391-427 as a standalone pass: colors → image (RGBA float). */
int ptb_filter_median(ptb_mem colors, ptb_mem image, int width, int height);

/* ---- multi-GPU inside ONE process (SURVEY 8b's ptb_set_gpus; 8e).  After ptb_set_gpus(n, mode) every launch — the reference
 * host's own gen_ray / trace_ray pairs (main.cpp:683-687) as well as ptb_render — runs on the first n GPUs of the box, counted
 * from the one ptb_init chose; the host keeps talking to ONE device: all ptb_mem buffers stay on the primary GPU, the others
 * stage their own copy of the scene (+ LBVH) from the uploaded records through NVLink peer access at their first launch.
 *   PTB_SPLIT_TILE  GPU d renders the bands of 8 scanlines d, d+n, ... with the reference's per-pixel seeds and stores its pixels'
 *                   rnds / colors / image straight into the primary's buffers over NVLink (the gather rides on the path kernels'
 *                   own stores: no collective, no copy).  The frame is bitwise the 1-GPU frame.
 *   PTB_SPLIT_SPP   sample j of a call goes to GPU (sample index + j) % n (ptb_render(spp) renders spp samples IN TOTAL); every
 *                   GPU keeps per-pixel sums and its own RNG states (GPU d: state * 16807^(d*W*H) mod 2^31-1, i.e. the host
 *                   generator's sequence continued), and one kernel on the primary adds the n sums through peer loads, scales
 *                   and tone-maps (sum -> mean -> image in one pass).  Same expectation as 1 GPU, different random numbers.
 * n == 1 restores the single-GPU path.  Not combinable with ptb_set_pixel_shard / ptb_set_accumulate_sum / ptb_comm_*. */
enum { PTB_SPLIT_TILE = 0, PTB_SPLIT_SPP = 1 };
int ptb_set_gpus(int n, int mode);
int ptb_get_gpus(void);

/* ---- multi-GPU, one PROCESS per GPU.  The path shards (pixels and samples are independent, the scene is
 * read-only and uploaded by every rank); the one exchange step — the SUM of the ranks' float accumulation buffers at a
 * root — runs over NCCL on the library's stream.  No equivalent in the reference (single device, main.cpp:466,476,506).
 * Rank 0 makes the 128-byte id (ptb_comm_unique_id) and hands it to the other ranks by whatever control plane the host has
 * (a file, MPI, torch.distributed); every rank then calls ptb_comm_init after ptb_init.  NCCL is loaded at run time.
 *   spp-split:  ptb_set_accumulate_sum(1); ptb_render(spp_r) on every rank from its own seed set; ptb_comm_reduce(colors, 0);
 *               root: ptb_resolve_frame(colors, W, H, total_spp, ...)        — sum -> mean -> tone map in one pass
 *   tile-split: ptb_set_pixel_shard(rank, n, group); ptb_render(spp); ptb_comm_reduce(colors, 0) (disjoint pixel sets: x + 0 == x,
 *               an exact gather) or the peer-memory gather of ptb_ipc_export / ptb_ipc_import + ptb_comm_barrier. */
#define PTB_COMM_ID_BYTES 128
int ptb_comm_unique_id(void *id_out /* PTB_COMM_ID_BYTES */);
int ptb_comm_init(int rank, int nranks, const void *unique_id /* PTB_COMM_ID_BYTES */);
int ptb_comm_ranks(void);                      /* 1 before ptb_comm_init */
int ptb_comm_rank(void);
int ptb_comm_reduce(ptb_mem colors, int root); /* in place float SUM over all ranks into root's buffer, asynchronous on the stream */
int ptb_comm_broadcast(ptb_mem buf, int root); /* root's bytes to every rank (scene replication without a host round trip) */
int ptb_comm_barrier(void);                    /* every rank's stream has drained up to here; returns when all have */
int ptb_comm_destroy(void);

/* Checkpoint / resume of a progressive render (the reference keeps this state only in device memory and its read-back
 * is dead code, main.cpp:727-741): the per-pixel RNG states (trace_ray arg 8), the accumulation buffer (arg 12) and the
 * sample index (arg 10) of the frame the bound camera (arg 11) describes, as a "PTBC" file.  ptb_load_state writes them
 * back into the buffers bound to trace_ray, sets arg 10 and returns the sample index: the next launch continues the
 * running mean and every pixel's random stream as if nothing had happened (8 + 8 samples across a dump == 16 samples). */
int ptb_save_state(const char *path);
int ptb_load_state(const char *path, int *sample_out);

/* Stream sharing with a host framework (torch): the shim launches on this CUDA stream. */
int ptb_set_stream(void *cuda_stream);     /* NULL = the shim's own stream */
void *ptb_get_stream(void);
/* LBVH statistics of the last build: nodes, leaves, max depth, build milliseconds. */
int ptb_accel_stats(int *n_nodes, int *n_tris, int *max_depth, float *build_ms);
/* number of kernels launched by this library since init (bench.py gpu_launches) */
uint64_t ptb_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* PTB200_H_ */
