// windows.h — inert stand-in.  The reference host includes <windows.h> (main.cpp:3) only for
// wglGetCurrentContext / wglGetCurrentDC, which feed the CL/GL-sharing context properties
// (main.cpp:491-492).  There is no graphics stack on the target box; both return null and the
// cl::Context stand-in ignores the properties.
#ifndef PTB200_SHIM_WINDOWS_H_
#define PTB200_SHIM_WINDOWS_H_
typedef void *HGLRC;
typedef void *HDC;
static inline HGLRC wglGetCurrentContext() { return nullptr; }
static inline HDC wglGetCurrentDC() { return nullptr; }
#endif
