// CL/cl.hpp — the subset of the Khronos OpenCL 1.x C++ wrapper that the reference host
// uses (/root/reference/main.cpp:389-528, 618-682), implemented over the C-ABI of
// include/ptb200.h.  With this directory first on the include path the reference's
// main.cpp compiles unmodified and drives the CUDA path tracer (or, when linked against
// the CPU backend under oracle/, the reference's own prog.cl compiled as C++).
//
// Written from the call sites, not from the Khronos header: only what main.cpp touches.
#ifndef PTB200_SHIM_CL_HPP_
#define PTB200_SHIM_CL_HPP_

#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <utility>
#include <vector>
#include <type_traits>

#include "ptb200.h"
#include "../ptb_shim_state.h"

// ---------------------------------------------------------------- scalar / vector types
typedef float cl_float;
typedef int32_t cl_int;
typedef uint32_t cl_uint;
typedef uint16_t cl_ushort;
typedef uint64_t cl_ulong;
typedef uint32_t cl_bool;
typedef uint64_t cl_bitfield;
typedef cl_bitfield cl_mem_flags;
typedef cl_bitfield cl_device_type;
typedef cl_uint cl_platform_info;
typedef cl_uint cl_device_info;
typedef intptr_t cl_context_properties;
typedef struct ptb_platform_t_ *cl_platform_id;
typedef struct ptb_device_t_ *cl_device_id;

// 16-byte, 4-lane union exactly like cl_platform.h's cl_float4 (cl_float3 is a typedef of it):
// the host structs of main.cpp:92-348 get the kernel ABI layout of prog.cl:1-35 from this.
typedef union alignas(16) {
    cl_float s[4];
    struct { cl_float x, y, z, w; };
} cl_float4;
typedef cl_float4 cl_float3;
typedef union alignas(8) {
    cl_float s[2];
    struct { cl_float x, y; };
} cl_float2;
typedef union alignas(8) {
    cl_int s[2];
    struct { cl_int x, y; };
} cl_int2;

#define CL_SUCCESS 0
#define CL_TRUE 1
#define CL_FALSE 0
#define CL_DEVICE_TYPE_ALL 0xFFFFFFFFu
#define CL_DEVICE_TYPE_GPU (1u << 2)
#define CL_PLATFORM_NAME 0x0902
#define CL_DEVICE_MAX_COMPUTE_UNITS 0x1002
#define CL_DEVICE_NAME 0x102B
#define CL_DRIVER_VERSION 0x102D
#define CL_DEVICE_VERSION 0x102F
#define CL_DEVICE_EXTENSIONS 0x1030
#define CL_DEVICE_OPENCL_C_VERSION 0x103D
#define CL_CONTEXT_PLATFORM 0x1084
#define CL_GL_CONTEXT_KHR 0x2008
#define CL_WGL_HDC_KHR 0x200B
#define CL_MEM_READ_WRITE (1u << 0)
#define CL_MEM_WRITE_ONLY (1u << 1)
#define CL_MEM_READ_ONLY (1u << 2)
#define CL_PROGRAM_BUILD_LOG 0x1183

// ---------------------------------------------------------------- C API used by list_info (main.cpp:400-448)
static inline cl_int clGetPlatformIDs(cl_uint n, cl_platform_id *out, cl_uint *count) {
    if (count) *count = 1;
    if (out && n >= 1) out[0] = reinterpret_cast<cl_platform_id>(uintptr_t(1));
    return CL_SUCCESS;
}
static inline cl_int clGetDeviceIDs(cl_platform_id, cl_device_type, cl_uint n, cl_device_id *out, cl_uint *count) {
    if (count) *count = 1;
    if (out && n >= 1) out[0] = reinterpret_cast<cl_device_id>(uintptr_t(1));
    return CL_SUCCESS;
}
static inline cl_int clGetDeviceInfo(cl_device_id, cl_device_info what, size_t cap, void *dst, size_t *need) {
    ptb_init(-1);
    if (what == CL_DEVICE_MAX_COMPUTE_UNITS) {
        cl_uint cu = (cl_uint)ptb_device_compute_units();
        if (need) *need = sizeof(cu);
        if (dst && cap >= sizeof(cu)) std::memcpy(dst, &cu, sizeof(cu));
        return CL_SUCCESS;
    }
    int idx = what == CL_DEVICE_NAME ? 0 : what == CL_DEVICE_VERSION ? 1 : what == CL_DRIVER_VERSION ? 2
            : what == CL_DEVICE_OPENCL_C_VERSION ? 3 : 4;
    char tmp[512];
    tmp[0] = 0;
    ptb_device_string(idx, tmp, sizeof(tmp));
    size_t len = std::strlen(tmp) + 1;
    if (need) *need = len;
    if (dst && cap) {
        std::strncpy(static_cast<char *>(dst), tmp, cap);
        static_cast<char *>(dst)[cap - 1] = 0;
    }
    return CL_SUCCESS;
}

namespace cl {

class Device {
public:
    Device() {}
    explicit Device(cl_device_id d) : id_(d) {}
    cl_device_id operator()() const { return id_; }
    template <cl_int what> std::string getInfo() const {
        char tmp[512];
        tmp[0] = 0;
        clGetDeviceInfo(id_, what, sizeof(tmp), tmp, nullptr);
        return tmp;
    }
private:
    cl_device_id id_ = nullptr;
};

class Platform {
public:
    Platform() {}
    explicit Platform(cl_platform_id p) : id_(p) {}
    // main.cpp:461
    static cl_int get(std::vector<Platform> *out) {
        cl_int rc = ptb_init(-1);
        out->clear();
        out->push_back(Platform(reinterpret_cast<cl_platform_id>(uintptr_t(1))));
        return rc;
    }
    // main.cpp:471
    cl_int getDevices(cl_device_type, std::vector<Device> *out) const {
        out->clear();
        out->push_back(Device(reinterpret_cast<cl_device_id>(uintptr_t(1))));
        return CL_SUCCESS;
    }
    cl_platform_id operator()() const { return id_; }   // main.cpp:490
    template <cl_int what> std::string getInfo() const { return ptb_backend_name(); }
private:
    cl_platform_id id_ = nullptr;
};

class Context {
public:
    Context() {}
    // main.cpp:496 — cl::Context({device}, properties); the GL-sharing properties are ignored
    Context(const std::vector<Device> &, const cl_context_properties * = nullptr) { ptb_init(-1); }
    explicit Context(const Device &, const cl_context_properties * = nullptr) { ptb_init(-1); }
};

class Program {
public:
    typedef std::vector<std::pair<const char *, size_t> > Sources;
    Program() {}
    // main.cpp:497 — the OpenCL C source text is accepted and ignored: kernels are prebuilt
    Program(const Context &, const Sources &) {}
    // main.cpp:498
    cl_int build(const std::vector<Device> &) { return ptb_init(-1); }
    // main.cpp:499
    template <cl_int what> std::string getBuildInfo(const Device &) const { return ptb_last_error(); }
};

// ref-counted handle, like cl::Memory
class Memory {
public:
    Memory() {}
    explicit Memory(ptb_mem h) : h_(h) {}
    Memory(const Memory &o) : h_(o.h_) { if (h_) ptb_retain(h_); }
    Memory(Memory &&o) noexcept : h_(o.h_) { o.h_ = 0; }
    Memory &operator=(const Memory &o) {
        if (o.h_) ptb_retain(o.h_);
        if (h_) ptb_release(h_);
        h_ = o.h_;
        return *this;
    }
    Memory &operator=(Memory &&o) noexcept {
        if (this != &o) {
            if (h_) ptb_release(h_);
            h_ = o.h_;
            o.h_ = 0;
        }
        return *this;
    }
    ~Memory() { if (h_) ptb_release(h_); }
    ptb_mem operator()() const { return h_; }
protected:
    ptb_mem h_ = 0;
};

class Buffer : public Memory {
public:
    Buffer() {}
    // main.cpp:508,509,520,619,623,627,632
    Buffer(const Context &, cl_mem_flags, size_t bytes, void * = nullptr, cl_int *err = nullptr)
        : Memory(ptb_alloc(bytes ? bytes : 1)) {
        if (err) *err = h_ ? CL_SUCCESS : PTB_ERR_OUT_OF_MEMORY;
    }
};

// main.cpp:519 — the GL texture is replaced by a headless W×H RGBA-float device image whose
// extent is taken from the preceding glTexImage2D call (see GL/glut.h in this directory).
class ImageGL : public Memory {
public:
    ImageGL() {}
    ImageGL(const Context &, cl_mem_flags, unsigned /*target*/, int /*miplevel*/, unsigned /*texture*/, cl_int *err = nullptr)
        : Memory(ptb_alloc_image(ptb_shim_last_tex_extent()[0], ptb_shim_last_tex_extent()[1])) {
        if (err) *err = h_ ? CL_SUCCESS : PTB_ERR_OUT_OF_MEMORY;
    }
};

class NDRange {
public:
    NDRange() : dims_(0) { s_[0] = s_[1] = s_[2] = 0; }
    NDRange(size_t a) : dims_(1) { s_[0] = a; s_[1] = 1; s_[2] = 1; }
    NDRange(size_t a, size_t b) : dims_(2) { s_[0] = a; s_[1] = b; s_[2] = 1; }
    NDRange(size_t a, size_t b, size_t c) : dims_(3) { s_[0] = a; s_[1] = b; s_[2] = c; }
    size_t dimensions() const { return dims_; }
    size_t operator[](int i) const { return s_[i]; }
private:
    size_t s_[3];
    int dims_;
};
static const NDRange NullRange;

class Event {};

class Kernel {
public:
    Kernel() {}
    // main.cpp:637,647 — created fresh every sample in the reference
    Kernel(const Program &, const char *name, cl_int *err = nullptr) : id_(ptb_kernel_id(name)) {
        if (err) *err = id_ >= 0 ? CL_SUCCESS : PTB_ERR_INVALID_KERNEL;
    }
    // main.cpp:638-640, 648-660
    cl_int setArg(cl_uint index, const Buffer &b) { return ptb_set_arg_mem(id_, (int)index, b()); }
    cl_int setArg(cl_uint index, const ImageGL &b) { return ptb_set_arg_mem(id_, (int)index, b()); }
    cl_int setArg(cl_uint index, const Memory &b) { return ptb_set_arg_mem(id_, (int)index, b()); }
    template <typename T>
    typename std::enable_if<!std::is_base_of<Memory, T>::value, cl_int>::type setArg(cl_uint index, const T &value) {
        return ptb_set_arg(id_, (int)index, sizeof(T), &value);
    }
    int id() const { return id_; }
private:
    int id_ = -1;
};

class CommandQueue {
public:
    CommandQueue() {}
    CommandQueue(const Context &, const Device &, cl_bitfield = 0, cl_int *err = nullptr) {
        cl_int rc = ptb_init(-1);
        if (err) *err = rc;
    }
    // main.cpp:526,620,624,628,633 — always blocking, offset 0 in the reference
    cl_int enqueueWriteBuffer(const Buffer &b, cl_bool, size_t offset, size_t bytes, const void *src,
                              const std::vector<Event> * = nullptr, Event * = nullptr) const {
        return ptb_write(b(), offset, bytes, src);
    }
    // dead code in the reference (main.cpp:690,728); needed for a headless read-back
    cl_int enqueueReadBuffer(const Memory &b, cl_bool, size_t offset, size_t bytes, void *dst,
                             const std::vector<Event> * = nullptr, Event * = nullptr) const {
        return ptb_read(b(), offset, bytes, dst);
    }
    // main.cpp:642,674,678
    cl_int enqueueNDRangeKernel(const Kernel &k, const NDRange &, const NDRange &global, const NDRange &,
                                const std::vector<Event> * = nullptr, Event * = nullptr) const {
        size_t gx = global.dimensions() >= 1 ? global[0] : 1;
        size_t gy = global.dimensions() >= 2 ? global[1] : 1;
        return ptb_launch(k.id(), gx, gy);
    }
    // main.cpp:673,676 — nothing to acquire: the image is a plain device buffer
    cl_int enqueueAcquireGLObjects(const std::vector<Memory> *, const std::vector<Event> * = nullptr, Event * = nullptr) const {
        return CL_SUCCESS;
    }
    cl_int enqueueReleaseGLObjects(const std::vector<Memory> *, const std::vector<Event> * = nullptr, Event * = nullptr) const {
        return CL_SUCCESS;
    }
    // main.cpp:675,680
    cl_int finish() const { return ptb_finish(); }
};

}  // namespace cl
#endif  // PTB200_SHIM_CL_HPP_
