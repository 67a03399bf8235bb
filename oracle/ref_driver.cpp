// ref_driver.cpp — TEST INFRASTRUCTURE (oracle).  Not part of the product path.
//
// Headless main() for the REFERENCE host: includes /root/reference/main.cpp (two mechanically
// patched `const int` lines for the resolution, see Makefile) and replaces only the GLUT main
// loop (main.cpp:1243-1268) by a scripted one that calls the same public Scene methods the
// reference's onInitialization/onIdle call (main.cpp:745-1017, 1226): init_Scene, add_Material,
// add_Triangle, end_Obj, add_Obj, upload_Triangles, upload_Materials, render.
//
// Linked with ref_backend.cpp it is the reference-on-CPU oracle (`_ref/ref_cpu`); linked with
// libptb200.so (-DREF_DRIVER_CUDA) the very same host code drives the CUDA path (`_ref/ref_host_b200`).
#include <algorithm>
#include <cassert>
#include <cctype>
#include <chrono>
#include <cmath>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <fstream>
#include <iostream>
#include <map>
#include <queue>
#include <random>
#include <sstream>
#include <string>
#include <utility>
#include <vector>
#include <unistd.h>

static int oracle_env_int(const char *name, int dflt) {
    const char *e = getenv(name);
    return (e && atoi(e) > 0) ? atoi(e) : dflt;
}

// open up Scene's members (buffer_colors, queue, tris, kd_tree ...) for read-back and dumps
#define private public
#define main reference_main
#include "main_patched.cpp"
#undef main
#undef private

#include "ptb200.h"

namespace {

struct ObjSpec {
    std::string file;
    float pos[3] = {0, 0, 0}, scale[3] = {1, 1, 1}, pitch = 0, yaw = 0;
};

void write_file(const std::string &path, const void *data, size_t bytes) {
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) { fprintf(stderr, "cannot write %s\n", path.c_str()); exit(2); }
    fwrite(data, 1, bytes, f);
    fclose(f);
}

// PTBS triangle-soup file (written by opencl_path_tracer_b200/scenes.py):
//   "PTBS", int32 version(1), n_mats, n_tris, n_objs,
//   float mats[n_mats][16] = kd,ks,emission,N(refr. index),K(extinction) rgb each + shininess,
//   int32 type[n_mats], float verts[n_tris][9], int32 tri_mat[n_tris], int32 obj_end[n_objs]
void load_tribin(const std::string &path, bool build_bvh) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) { fprintf(stderr, "cannot read %s\n", path.c_str()); exit(2); }
    char magic[4];
    int32_t hdr[4];
    if (fread(magic, 1, 4, f) != 4 || memcmp(magic, "PTBS", 4) || fread(hdr, 4, 4, f) != 4 || hdr[0] != 1) {
        fprintf(stderr, "%s: not a PTBS v1 file\n", path.c_str());
        exit(2);
    }
    int n_mats = hdr[1], n_tris = hdr[2], n_objs = hdr[3];
    std::vector<float> mats((size_t)n_mats * 16), verts((size_t)n_tris * 9);
    std::vector<int32_t> type(n_mats), tri_mat(n_tris), obj_end(n_objs);
    bool ok = fread(mats.data(), 4, mats.size(), f) == mats.size() && fread(type.data(), 4, type.size(), f) == type.size() &&
              fread(verts.data(), 4, verts.size(), f) == verts.size() && fread(tri_mat.data(), 4, tri_mat.size(), f) == tri_mat.size() &&
              fread(obj_end.data(), 4, obj_end.size(), f) == obj_end.size();
    fclose(f);
    if (!ok) { fprintf(stderr, "%s: truncated\n", path.c_str()); exit(2); }
    int mat_offset = (int)scene.mats.size();
    for (int i = 0; i < n_mats; ++i) {
        const float *m = &mats[(size_t)i * 16];
        scene.add_Material(Material((cl_float3){m[0], m[1], m[2]}, (cl_float3){m[3], m[4], m[5]}, (cl_float3){m[6], m[7], m[8]},
                                    (cl_float3){m[9], m[10], m[11]}, (cl_float3){m[12], m[13], m[14]}, m[15], type[i]));
    }
    int next_obj = 0;
    for (int i = 0; i < n_tris; ++i) {
        const float *v = &verts[(size_t)i * 9];
        scene.add_Triangle(Triangle((cl_float3){v[0], v[1], v[2]}, (cl_float3){v[3], v[4], v[5]}, (cl_float3){v[6], v[7], v[8]},
                                    (cl_ushort)(mat_offset + tri_mat[i])));
        while (next_obj < n_objs && obj_end[next_obj] == i + 1) {
            if (build_bvh) scene.end_Obj();
            ++next_obj;
        }
    }
}

}  // namespace

int main(int argc, char **argv) {
    int W = 0, H = 0, spp = 1, iters = 4, threads = 0, accel = -1, math = -1;
    bool brute = false, count = false, no_bvh = false, use_render_ext = false;
    int shard[3] = {0, 1, 1};
    int gpus = 1, gpu_split = PTB_SPLIT_TILE, pipeline = -1;
    float fov = 39.3f, cyaw = 0.0f, cpitch = 0.0f, eye[3] = {277.5f, 277.5f, -800.0f};
    std::vector<ObjSpec> objs;
    std::vector<std::string> tribins;
    std::string out, dump, rnds_out, tex_out, filt_out, colors_in;
    auto need = [&](int &i, int n) { if (i + n >= argc) { fprintf(stderr, "missing value after %s\n", argv[i]); exit(2); } };
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "--w") { need(i, 1); W = atoi(argv[++i]); }
        else if (a == "--h") { need(i, 1); H = atoi(argv[++i]); }
        else if (a == "--spp") { need(i, 1); spp = atoi(argv[++i]); }
        else if (a == "--iterations") { need(i, 1); iters = atoi(argv[++i]); }
        else if (a == "--threads") { need(i, 1); threads = atoi(argv[++i]); }
        else if (a == "--accel") { need(i, 1); accel = atoi(argv[++i]); }
        else if (a == "--math") { need(i, 1); math = atoi(argv[++i]); }
        else if (a == "--brute") brute = true;
        else if (a == "--no-bvh") no_bvh = true;
        else if (a == "--count") count = true;
        else if (a == "--render-ext") use_render_ext = true;
        else if (a == "--fov") { need(i, 1); fov = (float)atof(argv[++i]); }
        else if (a == "--cam-yaw") { need(i, 1); cyaw = (float)atof(argv[++i]); }
        else if (a == "--cam-pitch") { need(i, 1); cpitch = (float)atof(argv[++i]); }
        else if (a == "--eye") { need(i, 3); for (int k = 0; k < 3; ++k) eye[k] = (float)atof(argv[++i]); }
        else if (a == "--shard") { need(i, 3); for (int k = 0; k < 3; ++k) shard[k] = atoi(argv[++i]); }
        else if (a == "--gpus") { need(i, 1); gpus = atoi(argv[++i]); }
        else if (a == "--split") { need(i, 1); gpu_split = std::string(argv[++i]) == "spp" ? PTB_SPLIT_SPP : PTB_SPLIT_TILE; }
        else if (a == "--pipeline") { need(i, 1); pipeline = atoi(argv[++i]); }
        else if (a == "--obj") { need(i, 1); ObjSpec o; o.file = argv[++i]; objs.push_back(o); }
        else if (a == "--pos") { need(i, 3); for (int k = 0; k < 3; ++k) objs.back().pos[k] = (float)atof(argv[++i]); }
        else if (a == "--scale") { need(i, 3); for (int k = 0; k < 3; ++k) objs.back().scale[k] = (float)atof(argv[++i]); }
        else if (a == "--pitch") { need(i, 1); objs.back().pitch = (float)atof(argv[++i]); }
        else if (a == "--yaw") { need(i, 1); objs.back().yaw = (float)atof(argv[++i]); }
        else if (a == "--tribin") { need(i, 1); tribins.push_back(argv[++i]); }
        else if (a == "--out") { need(i, 1); out = argv[++i]; }
        else if (a == "--dump") { need(i, 1); dump = argv[++i]; }
        else if (a == "--rnds-out") { need(i, 1); rnds_out = argv[++i]; }
        else if (a == "--tex-out") { need(i, 1); tex_out = argv[++i]; }
        else if (a == "--filt-out") { need(i, 1); filt_out = argv[++i]; }
        else if (a == "--colors-in") { need(i, 1); colors_in = argv[++i]; }
        else { fprintf(stderr, "unknown option %s\n", a.c_str()); return 2; }
    }
    // screen_width/height are `const int` globals initialised before main (main.cpp:20-21, patched to
    // read ORACLE_W/H): re-exec once with the environment set when the request differs.
    if ((W > 0 && W != screen_width) || (H > 0 && H != screen_height)) {
        if (getenv("ORACLE_REEXEC")) { fprintf(stderr, "resolution re-exec failed\n"); return 2; }
        char b[32];
        snprintf(b, sizeof b, "%d", W > 0 ? W : screen_width); setenv("ORACLE_W", b, 1);
        snprintf(b, sizeof b, "%d", H > 0 ? H : screen_height); setenv("ORACLE_H", b, 1);
        setenv("ORACLE_REEXEC", "1", 1);
        execv("/proc/self/exe", argv);
        perror("execv");
        return 2;
    }
    W = screen_width; H = screen_height;
    if (threads > 0) { char b[32]; snprintf(b, sizeof b, "%d", threads); setenv("ORACLE_THREADS", b, 1); }

    // headless equivalents of the interactive globals (main.cpp:27-43); forward/right/up stay 0 (Camera() side effect, main.cpp:334-336)
    iterations = iters;
    real_time = false;
    global_fov = fov; global_yaw = cyaw; global_pitch = cpitch;
    global_shift = (cl_float3){eye[0] - 500.0f, eye[1] - 500.0f, eye[2] + 1299.037842f};

    FILE *real_stdout = stdout;
    scene.init_Scene();
    for (auto &t : tribins) load_tribin(t, !no_bvh);
    for (auto &o : objs)
        scene.add_Obj(o.file, (cl_float3){o.pos[0], o.pos[1], o.pos[2]}, (cl_float3){o.scale[0], o.scale[1], o.scale[2]}, o.pitch, o.yaw);
    if (scene.kd_tree.empty()) {  // --no-bvh: keep the kernel-argument surface valid with one never-hit node
        scene.kd_tree.push_back(Node((cl_int2){-1, -1}, BBox()));
        scene.kd_tree.push_back(Node((cl_int2){0, 0}, BBox((cl_float3){1e30f, 1e30f, 1e30f}, (cl_float3){1e30f, 1e30f, 1e30f})));
        scene.kd_tree_shift.push_back(0);
    }
    scene.upload_Triangles();
    scene.upload_Materials();
    if (brute) ptb_set_accel(PTB_ACCEL_BRUTE); else if (accel >= 0) ptb_set_accel(accel);
    if (math >= 0) ptb_set_math(math);
    if (count) ptb_set_count_work(1);
    if (shard[1] > 1) ptb_set_pixel_shard(shard[0], shard[1], shard[2]);
    if (pipeline >= 0) ptb_set_pipeline(pipeline);
    // the ONE call a multi-GPU host adds: the launches of Scene::render (main.cpp:683-687) then run on `gpus` GPUs
    if (gpus > 1) ptb_set_gpus(gpus, gpu_split);
    ptb_build_accel();
    ptb_reset_counters();

    auto t0 = std::chrono::steady_clock::now();
    if (use_render_ext) {
        // same argument set-up as Scene::trace_rays (main.cpp:647-660), then the in-kernel sample loop
        scene.camera = Camera();
        cl::Kernel k(scene.program, "trace_ray");
        k.setArg(0, scene.cl_screen); k.setArg(1, scene.buffer_tris); k.setArg(2, scene.tris_size);
        k.setArg(3, scene.buffer_mats); k.setArg(4, scene.buffer_kd_tree); k.setArg(5, scene.buffer_kd_tree_shift);
        k.setArg(6, scene.kd_tree_shift_size); k.setArg(7, scene.buffer_rays); k.setArg(8, scene.buffer_rnds);
        k.setArg(9, iterations); k.setArg(10, current_sample); k.setArg(11, scene.camera); k.setArg(12, scene.buffer_colors);
        ptb_render(spp);
        current_sample += spp;
    } else {
        for (int s = 0; s < spp; ++s) scene.render();   // main.cpp:683-687
    }
    scene.queue.finish();
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    size_t n = (size_t)W * H;
    if (!out.empty()) {
        std::vector<cl_float3> colors(n);
        scene.queue.enqueueReadBuffer(scene.buffer_colors, CL_TRUE, 0, n * sizeof(cl_float3), colors.data());
        write_file(out, colors.data(), n * sizeof(cl_float3));
    }
    if (!rnds_out.empty()) {
        std::vector<cl_int> r(n);
        scene.queue.enqueueReadBuffer(scene.buffer_rnds, CL_TRUE, 0, n * sizeof(cl_int), r.data());
        write_file(rnds_out, r.data(), n * sizeof(cl_int));
    }
    if (!tex_out.empty()) {
        std::vector<cl_float4> t(n);
        scene.queue.enqueueReadBuffer(scene.cl_screen, CL_TRUE, 0, n * sizeof(cl_float4), t.data());
        write_file(tex_out, t.data(), n * sizeof(cl_float4));
    }
    if (!filt_out.empty()) {
        // the launch the reference keeps commented out at main.cpp:665-668, verbatim: filt_im(cl_screen, buffer_colors) over (W, H);
        // --colors-in replaces the rendered accumulation buffer by a raw float4 file first (arbitrary test images)
        if (!colors_in.empty()) {
            std::vector<cl_float3> c(n);
            FILE *f = fopen(colors_in.c_str(), "rb");
            if (!f || fread(c.data(), sizeof(cl_float3), n, f) != n) { fprintf(stderr, "cannot read %zu pixels from %s\n", n, colors_in.c_str()); return 2; }
            fclose(f);
            scene.queue.enqueueWriteBuffer(scene.buffer_colors, CL_TRUE, 0, n * sizeof(cl_float3), c.data());
        }
        std::vector<cl_float4> zero(n);
        memset(zero.data(), 0, n * sizeof(cl_float4));
        ptb_write(scene.cl_screen(), 0, n * sizeof(cl_float4), zero.data());      // pixels filt_im leaves alone read back as 0
        cl::Kernel kernel_filt_im = cl::Kernel(scene.program, "filt_im");
        kernel_filt_im.setArg(0, scene.cl_screen);
        kernel_filt_im.setArg(1, scene.buffer_colors);
        scene.queue.enqueueNDRangeKernel(kernel_filt_im, cl::NullRange, cl::NDRange(screen_width, screen_height), cl::NullRange);
        scene.queue.finish();
        std::vector<cl_float4> t(n);
        scene.queue.enqueueReadBuffer(scene.cl_screen, CL_TRUE, 0, n * sizeof(cl_float4), t.data());
        write_file(filt_out, t.data(), n * sizeof(cl_float4));
    }
    if (!dump.empty()) {
        // "PTBD": the kernel-argument surface exactly as the reference host uploaded it
        FILE *f = fopen(dump.c_str(), "wb");
        int32_t hdr[8] = {1, W, H, iterations, (int32_t)scene.tris.size(), (int32_t)scene.mats.size(), (int32_t)scene.kd_tree.size(),
                          (int32_t)scene.kd_tree_shift.size()};
        Camera cam = Camera();
        fwrite("PTBD", 1, 4, f);
        fwrite(hdr, 4, 8, f);
        fwrite(&cam, sizeof(Camera), 1, f);
        fwrite(scene.tris.data(), sizeof(Triangle), scene.tris.size(), f);
        fwrite(scene.mats.data(), sizeof(Material), scene.mats.size(), f);
        fwrite(scene.kd_tree.data(), sizeof(Node), scene.kd_tree.size(), f);
        fwrite(scene.kd_tree_shift.data(), sizeof(cl_int), scene.kd_tree_shift.size(), f);
        fclose(f);
    }
    ptb_counters c;
    memset(&c, 0, sizeof c);
    ptb_get_counters(&c);
    fprintf(real_stdout,
            "\n{\"backend\": \"%s\", \"w\": %d, \"h\": %d, \"spp\": %d, \"iterations\": %d, \"tris\": %zu, \"nodes\": %zu, \"objects\": %zu, "
            "\"threads\": %d, \"seconds\": %.6f, \"rays\": %llu, \"box_tests\": %llu, \"tri_tests\": %llu, \"tri_stage1\": %llu, "
            "\"tri_stage2\": %llu, \"tri_stage3\": %llu, \"error\": \"%s\"}\n",
            ptb_backend_name(), W, H, spp, iterations, scene.tris.size(), scene.kd_tree.size(), scene.kd_tree_shift.size(),
            ptb_device_compute_units(), sec, (unsigned long long)c.rays, (unsigned long long)c.box_tests, (unsigned long long)c.tri_tests,
            (unsigned long long)c.tri_stage1, (unsigned long long)c.tri_stage2, (unsigned long long)c.tri_stage3, ptb_last_error());
    return ptb_last_error()[0] ? 1 : 0;
}
