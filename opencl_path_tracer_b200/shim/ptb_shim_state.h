// Shared by the GL/glut.h and CL/cl.hpp stand-ins: remembers the extent of the last
// glTexImage2D call (main.cpp:517) so that cl::ImageGL (main.cpp:519) can size the
// headless device image that replaces the GL texture.
#ifndef PTB200_SHIM_STATE_H_
#define PTB200_SHIM_STATE_H_
inline int *ptb_shim_last_tex_extent() {
    static int wh[2] = {0, 0};
    return wh;
}
#endif
