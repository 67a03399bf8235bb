// ref_backend.cpp — TEST INFRASTRUCTURE (oracle).  Not part of the product path.
//
// A CPU implementation of the C-ABI in include/ptb200.h whose "device" executes the
// REFERENCE's own kernels: /root/reference/prog.cl compiled as C++ (see Makefile for the
// mechanical transformations, oclshim.h for the built-ins).  Linked with the reference's
// main.cpp behind the cl:: stand-in it reproduces exactly what main.cpp:635-687 enqueues:
// gen_ray over W*H work-items, then trace_ray over (W,H), per sample.
//
// Three compilations of the same source live in three namespaces:
//   refcl        the shipping kernel (BVH traversal, prog.cl:319)
//   refcl_brute  the reference's commented alternative (whole-scene first_intersect, prog.cl:318)
//   refcl_count  the shipping kernel with box/triangle work counters
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include "ptb200.h"
#include "oclshim.h"

#define ORACLE_COUNT(f) (oclshim::counters().f++)

namespace refcl {
using namespace oclshim;
#define ORACLE_COUNTX(f)
#include "prog_cl.inc"
#undef ORACLE_COUNTX
}  // namespace refcl

namespace refcl_brute {
using namespace oclshim;
#define ORACLE_BRUTE 1
#define ORACLE_COUNTX(f)
#include "prog_cl.inc"
#undef ORACLE_COUNTX
#undef ORACLE_BRUTE
}  // namespace refcl_brute

namespace refcl_count {
using namespace oclshim;
#define ORACLE_COUNTX(f) (oclshim::counters().f++)
#include "prog_cl.inc"
#undef ORACLE_COUNTX
}  // namespace refcl_count

// the OpenCL address-space / kernel qualifiers were macros for the benefit of prog.cl only
#undef global
#undef kernel
#undef write_only

static_assert(sizeof(refcl::Material) == 80, "Material ABI");
static_assert(sizeof(refcl::Ray) == 32, "Ray ABI");
static_assert(sizeof(refcl::Triangle) == 80, "Triangle ABI");
static_assert(sizeof(refcl::Node) == 48, "Node ABI");
static_assert(sizeof(refcl::Camera) == 80, "Camera ABI");
static_assert(sizeof(refcl::BBox) == 32, "BBox ABI");

namespace {

struct Mem {
    std::vector<unsigned char> bytes;
    int refs = 1;
    bool is_image = false;
    oclshim::image2d img{nullptr, 0, 0};
};
// immortal: the reference's global `Scene scene` releases its buffers during static destruction
std::map<ptb_mem, Mem *> &g_mems = *new std::map<ptb_mem, Mem *>();
ptb_mem g_next = 1;
std::string &g_err = *new std::string();

struct Arg {
    std::vector<unsigned char> value;
    ptb_mem mem = 0;
    bool set = false;
};
Arg g_args[PTB_KERNEL_COUNT][16];

int g_threads = 0;
int g_mode = 0;  // 0 shipping BVH, 1 brute, 2 counting
int g_shard_rank = 0, g_shard_n = 1, g_shard_group = 1;
ptb_counters g_counters = {0, 0, 0, 0, 0, 0, 0};
uint64_t g_launches = 0;

int fail(int code, const char *msg) {
    if (g_err.empty()) {
        g_err = msg;
        fprintf(stderr, "ptb(ref-cpu): %s\n", msg);
    }
    return code;
}
Mem *find(ptb_mem m) {
    auto it = g_mems.find(m);
    return it == g_mems.end() ? nullptr : it->second;
}
template <typename T> T *buf(int kernel, int index) {
    Mem *m = find(g_args[kernel][index].mem);
    return m ? reinterpret_cast<T *>(m->bytes.data()) : nullptr;
}
template <typename T> T val(int kernel, int index) {
    T v;
    std::memset(&v, 0, sizeof(T));
    const Arg &a = g_args[kernel][index];
    if (a.value.size() >= sizeof(T)) std::memcpy(&v, a.value.data(), sizeof(T));
    return v;
}
int n_threads() {
    if (g_threads > 0) return g_threads;
    const char *e = getenv("ORACLE_THREADS");
    if (e && atoi(e) > 0) return atoi(e);
    unsigned hc = std::thread::hardware_concurrency();
    return hc ? (int)hc : 1;
}
bool pixel_selected(size_t id) {
    if (g_shard_n <= 1) return true;
    return (int)((id / (size_t)g_shard_group) % (size_t)g_shard_n) == g_shard_rank;
}

// run fn(row) for rows [0,H) over a static row-interleaved schedule, merge thread counters
template <typename F> void parallel_rows(size_t rows, F fn) {
    int nt = n_threads();
    if ((size_t)nt > rows) nt = (int)rows ? (int)rows : 1;
    std::vector<oclshim::Counters> partial((size_t)nt);
    auto worker = [&](int t) {
        oclshim::counters() = oclshim::Counters{0, 0, 0, 0, 0, 0};
        for (size_t r = (size_t)t; r < rows; r += (size_t)nt) fn(r);
        partial[(size_t)t] = oclshim::counters();
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < nt; ++t) pool.emplace_back(worker, t);
    worker(0);
    for (auto &th : pool) th.join();
    for (auto &c : partial) {
        g_counters.rays += c.rays;
        g_counters.box_tests += c.box;
        g_counters.tri_tests += c.tri;
        g_counters.tri_stage1 += c.tri1;
        g_counters.tri_stage2 += c.tri2;
        g_counters.tri_stage3 += c.tri3;
    }
}

}  // namespace

// ---------------------------------------------------------------- kernel launches
namespace {

// gen_ray(rays, camera, rnds)  — prog.cl:384-389, enqueued 1-D over W*H (main.cpp:642)
void launch_gen_ray(size_t n) {
    auto *rays = buf<refcl::Ray>(PTB_KERNEL_GEN_RAY, 0);
    refcl::Camera cam = val<refcl::Camera>(PTB_KERNEL_GEN_RAY, 1);
    int *rnds = buf<int>(PTB_KERNEL_GEN_RAY, 2);
    if (!rays || !rnds) { fail(PTB_ERR_INVALID_ARGS, "gen_ray: buffer argument missing"); return; }
    const size_t chunk = 4096;
    size_t chunks = (n + chunk - 1) / chunk;
    parallel_rows(chunks, [&](size_t c) {
        oclshim::WorkItem &wi = oclshim::work_item();
        wi.gsz[0] = n; wi.gsz[1] = 1; wi.gsz[2] = 1; wi.gid[1] = 0; wi.gid[2] = 0;
        size_t end = (c + 1) * chunk < n ? (c + 1) * chunk : n;
        for (size_t id = c * chunk; id < end; ++id) {
            if (!pixel_selected(id)) continue;
            wi.gid[0] = id;
            refcl::gen_ray(rays, cam, rnds);
        }
    });
}

// trace_ray(tex, tris, tris_size, materials, kd_tree, kd_tree_shift, kd_tree_shift_size, rays, rnds,
//           iterations, current_sample, cam, colors) — prog.cl:292-304, enqueued 2-D (W,H) (main.cpp:674,678)
template <typename NS_Triangle, typename NS_Material, typename NS_Node, typename NS_Ray, typename NS_Camera, typename Fn>
void launch_trace_ray_ns(size_t W, size_t H, Fn kernel_fn) {
    const int K = PTB_KERNEL_TRACE_RAY;
    Mem *img = find(g_args[K][0].mem);
    oclshim::image2d_t tex = (img && img->is_image) ? &img->img : nullptr;
    auto *tris = buf<NS_Triangle>(K, 1);
    int tris_size = val<int>(K, 2);
    auto *mats = buf<NS_Material>(K, 3);
    auto *nodes = buf<NS_Node>(K, 4);
    int *shift = buf<int>(K, 5);
    int shift_size = val<int>(K, 6);
    auto *rays = buf<NS_Ray>(K, 7);
    int *rnds = buf<int>(K, 8);
    int iterations = val<int>(K, 9);
    int current_sample = val<int>(K, 10);
    NS_Camera cam = val<NS_Camera>(K, 11);
    auto *colors = buf<oclshim::float3>(K, 12);
    if (!tris || !mats || !rays || !rnds || !colors) { fail(PTB_ERR_INVALID_ARGS, "trace_ray: buffer argument missing"); return; }
    parallel_rows(H, [&](size_t y) {
        oclshim::WorkItem &wi = oclshim::work_item();
        wi.gsz[0] = W; wi.gsz[1] = H; wi.gsz[2] = 1; wi.gid[1] = y; wi.gid[2] = 0;
        for (size_t x = 0; x < W; ++x) {
            if (!pixel_selected(y * W + x)) continue;
            wi.gid[0] = x;
            kernel_fn(tex, tris, tris_size, mats, nodes, shift, shift_size, rays, rnds, iterations, current_sample, cam, colors);
        }
    });
    g_counters.pixel_samples += (uint64_t)W * H / (uint64_t)(g_shard_n > 1 ? g_shard_n : 1);
}

void launch_trace_ray(size_t W, size_t H) {
    if (g_mode == 1)
        launch_trace_ray_ns<refcl_brute::Triangle, refcl_brute::Material, refcl_brute::Node, refcl_brute::Ray, refcl_brute::Camera>(
            W, H, refcl_brute::trace_ray);
    else if (g_mode == 2)
        launch_trace_ray_ns<refcl_count::Triangle, refcl_count::Material, refcl_count::Node, refcl_count::Ray, refcl_count::Camera>(
            W, H, refcl_count::trace_ray);
    else
        launch_trace_ray_ns<refcl::Triangle, refcl::Material, refcl::Node, refcl::Ray, refcl::Camera>(W, H, refcl::trace_ray);
}

// filt_im(tex, colors) — prog.cl:391-427, enqueued 2-D (W,H) in the reference's commented launch (main.cpp:665-668).
// The kernel guards x>0 && y>0 && x<width && y<height and then reads the 3x3 neighbourhood through LINEAR ids
// (y-1+i)*width + (x-1+j): on the last column that wraps into the next scanline (defined in a linear buffer), on the last
// scanline — and for the very last pixel of the scanline before it — it reads past the end of colors[] (undefined in the
// reference).  The executor runs every work-item whose nine reads stay inside the buffer and skips the others.
void launch_filt_im(size_t W, size_t H) {
    const int K = PTB_KERNEL_FILT_IM;
    Mem *img = find(g_args[K][0].mem);
    oclshim::image2d_t tex = (img && img->is_image) ? &img->img : nullptr;
    Mem *cm = find(g_args[K][1].mem);
    if (!tex || !cm || cm->bytes.size() < W * H * sizeof(oclshim::float3)) { fail(PTB_ERR_INVALID_ARGS, "filt_im: image / colors argument missing"); return; }
    auto *colors = reinterpret_cast<oclshim::float3 *>(cm->bytes.data());
    parallel_rows(H, [&](size_t y) {
        oclshim::WorkItem &wi = oclshim::work_item();
        wi.gsz[0] = W; wi.gsz[1] = H; wi.gsz[2] = 1; wi.gid[1] = y; wi.gid[2] = 0;
        for (size_t x = 0; x < W; ++x) {
            if ((y + 1) * W + (x + 1) >= W * H) continue;      // would read colors[] out of range
            wi.gid[0] = x;
            refcl::filt_im(tex, colors);
        }
    });
}

}  // namespace

// unit-level entry points used by the oracle self-checks (tests/, not-gpu): the reference's own rand()
extern "C" float ref_rand(int *seed) { return refcl::rand(seed); }
extern "C" void ref_set_threads(int n) { g_threads = n; }

// ---------------------------------------------------------------- the C-ABI (CPU flavour)
extern "C" {

int ptb_init(int) { return PTB_SUCCESS; }
int ptb_shutdown(void) {
    for (auto &kv : g_mems) delete kv.second;
    g_mems.clear();
    return PTB_SUCCESS;
}
int ptb_device_string(int what, char *dst, size_t cap) {
    const char *s[] = {"reference prog.cl compiled as C++ (host CPU)", "OpenCL-C-as-C++17", "g++ -O2 -ffp-contract=off", "OpenCL C 1.x subset", ""};
    if (!dst || !cap) return PTB_ERR_INVALID_VALUE;
    snprintf(dst, cap, "%s", s[what < 0 || what > 4 ? 4 : what]);
    return PTB_SUCCESS;
}
int ptb_device_compute_units(void) { return n_threads(); }
const char *ptb_backend_name(void) { return "ref-cpu"; }

ptb_mem ptb_alloc(size_t bytes) {
    Mem *m = new Mem();
    m->bytes.assign(bytes, 0);
    ptb_mem h = g_next++;
    g_mems[h] = m;
    return h;
}
ptb_mem ptb_alloc_image(int w, int h) {
    ptb_mem hd = ptb_alloc((size_t)w * (size_t)h * sizeof(oclshim::float4));
    Mem *m = find(hd);
    m->is_image = true;
    m->img.texels = reinterpret_cast<oclshim::float4 *>(m->bytes.data());
    m->img.width = w;
    m->img.height = h;
    return hd;
}
int ptb_retain(ptb_mem h) {
    Mem *m = find(h);
    if (!m) return PTB_ERR_INVALID_MEM;
    m->refs++;
    return PTB_SUCCESS;
}
int ptb_release(ptb_mem h) {
    Mem *m = find(h);
    if (!m) return PTB_ERR_INVALID_MEM;
    if (--m->refs == 0) {
        delete m;
        g_mems.erase(h);
    }
    return PTB_SUCCESS;
}
int ptb_write(ptb_mem h, size_t off, size_t bytes, const void *src) {
    Mem *m = find(h);
    if (!m) return fail(PTB_ERR_INVALID_MEM, "ptb_write: bad handle");
    if (bytes == 0) return PTB_SUCCESS;
    if (off + bytes > m->bytes.size()) return fail(PTB_ERR_INVALID_VALUE, "ptb_write: out of range");
    std::memcpy(m->bytes.data() + off, src, bytes);
    return PTB_SUCCESS;
}
int ptb_read(ptb_mem h, size_t off, size_t bytes, void *dst) {
    Mem *m = find(h);
    if (!m) return fail(PTB_ERR_INVALID_MEM, "ptb_read: bad handle");
    if (off + bytes > m->bytes.size()) return fail(PTB_ERR_INVALID_VALUE, "ptb_read: out of range");
    std::memcpy(dst, m->bytes.data() + off, bytes);
    return PTB_SUCCESS;
}
size_t ptb_mem_size(ptb_mem h) {
    Mem *m = find(h);
    return m ? m->bytes.size() : 0;
}
ptb_mem ptb_wrap_device(void *, size_t) { return 0; }
void *ptb_device_ptr(ptb_mem h) {
    Mem *m = find(h);
    return m ? m->bytes.data() : nullptr;
}

int ptb_kernel_id(const char *name) {
    if (!name) return PTB_ERR_INVALID_KERNEL;
    if (!strcmp(name, "gen_ray")) return PTB_KERNEL_GEN_RAY;
    if (!strcmp(name, "trace_ray")) return PTB_KERNEL_TRACE_RAY;
    if (!strcmp(name, "filt_im")) return PTB_KERNEL_FILT_IM;
    return PTB_ERR_INVALID_KERNEL;
}
int ptb_set_arg(int k, int i, size_t bytes, const void *value) {
    if (k < 0 || k >= PTB_KERNEL_COUNT || i < 0 || i >= 16) return fail(PTB_ERR_INVALID_VALUE, "ptb_set_arg: bad index");
    Arg &a = g_args[k][i];
    a.value.assign((const unsigned char *)value, (const unsigned char *)value + bytes);
    a.mem = 0;
    a.set = true;
    return PTB_SUCCESS;
}
int ptb_set_arg_mem(int k, int i, ptb_mem m) {
    if (k < 0 || k >= PTB_KERNEL_COUNT || i < 0 || i >= 16) return fail(PTB_ERR_INVALID_VALUE, "ptb_set_arg_mem: bad index");
    Arg &a = g_args[k][i];
    a.value.clear();
    a.mem = m;
    a.set = true;
    return PTB_SUCCESS;
}
int ptb_launch(int k, size_t gx, size_t gy) {
    g_launches++;
    if (k == PTB_KERNEL_GEN_RAY) launch_gen_ray(gx * (gy ? gy : 1));
    else if (k == PTB_KERNEL_TRACE_RAY) launch_trace_ray(gx, gy);
    else if (k == PTB_KERNEL_FILT_IM) launch_filt_im(gx, gy);
    else return fail(PTB_ERR_INVALID_KERNEL, "ptb_launch: kernel not available in the CPU oracle");
    return g_err.empty() ? PTB_SUCCESS : PTB_ERR_INVALID_ARGS;
}
int ptb_finish(void) { return PTB_SUCCESS; }
const char *ptb_last_error(void) { return g_err.c_str(); }
void ptb_clear_error(void) { g_err.clear(); }

// extensions: only what the oracle driver needs
int ptb_set_accel(int accel) {
    g_mode = (accel == PTB_ACCEL_BRUTE) ? 1 : (g_mode == 2 ? 2 : 0);
    return PTB_SUCCESS;
}
int ptb_get_accel(void) { return g_mode == 1 ? PTB_ACCEL_BRUTE : PTB_ACCEL_REFBVH; }
int ptb_set_math(int) { return PTB_SUCCESS; }
int ptb_set_pipeline(int) { return PTB_SUCCESS; }
int ptb_build_accel(void) { return PTB_SUCCESS; }
int ptb_render(int) { return fail(PTB_ERR_INVALID_VALUE, "ptb_render: the CPU oracle only executes the reference's two kernels"); }
int ptb_set_pixel_shard(int rank, int n, int group) {
    g_shard_rank = rank;
    g_shard_n = n < 1 ? 1 : n;
    g_shard_group = group < 1 ? 1 : group;
    return PTB_SUCCESS;
}
int ptb_set_accumulate_sum(int) { return PTB_ERR_INVALID_VALUE; }
int ptb_resolve_sum(ptb_mem, int) { return PTB_ERR_INVALID_VALUE; }
int ptb_set_count_work(int on) {
    if (g_mode != 1) g_mode = on ? 2 : 0;
    return PTB_SUCCESS;
}
int ptb_get_counters(ptb_counters *out) {
    if (out) *out = g_counters;
    return PTB_SUCCESS;
}
int ptb_reset_counters(void) {
    std::memset(&g_counters, 0, sizeof(g_counters));
    return PTB_SUCCESS;
}
int ptb_resolve_rgba8(ptb_mem, int, int, int, unsigned char *) { return PTB_ERR_INVALID_VALUE; }
int ptb_filter_median(ptb_mem colors, ptb_mem image, int w, int h) {
    ptb_set_arg_mem(PTB_KERNEL_FILT_IM, 0, image);
    ptb_set_arg_mem(PTB_KERNEL_FILT_IM, 1, colors);
    return ptb_launch(PTB_KERNEL_FILT_IM, (size_t)w, (size_t)h);
}
int ptb_set_gpus(int n, int) { return n == 1 ? PTB_SUCCESS : fail(PTB_ERR_INVALID_VALUE, "the CPU oracle is one device"); }
int ptb_set_stream(void *) { return PTB_SUCCESS; }
void *ptb_get_stream(void) { return nullptr; }
int ptb_accel_stats(int *a, int *b, int *c, float *d) {
    if (a) *a = 0;
    if (b) *b = 0;
    if (c) *c = 0;
    if (d) *d = 0;
    return PTB_SUCCESS;
}
uint64_t ptb_launch_count(void) { return g_launches; }

}  // extern "C"
