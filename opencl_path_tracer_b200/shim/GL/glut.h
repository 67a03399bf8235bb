// GL/glut.h — inert stand-in for freeglut + OpenGL 1.x as used by the reference host
// (main.cpp:512-519 texture creation, 745-747 viewport, 1019-1039 quad draw, 1171-1268 main loop).
// No graphics stack exists on the target box ("unavailable on this box"): every call is a
// no-op, except that glTexImage2D records the texture extent for the headless device image and
// glutGet(GLUT_ELAPSED_TIME) returns real milliseconds so the status line (main.cpp:1236) still works.
#ifndef PTB200_SHIM_GLUT_H_
#define PTB200_SHIM_GLUT_H_
#include <chrono>
#include "../ptb_shim_state.h"

typedef unsigned int GLenum;
typedef unsigned int GLuint;
typedef int GLint;
typedef int GLsizei;
typedef float GLfloat;
typedef unsigned int GLbitfield;
typedef void GLvoid;

#define GL_TEXTURE_2D 0x0DE1
#define GL_TEXTURE_MIN_FILTER 0x2801
#define GL_LINEAR 0x2601
#define GL_RGBA 0x1908
#define GL_FLOAT 0x1406
#define GL_COLOR_BUFFER_BIT 0x4000
#define GL_DEPTH_BUFFER_BIT 0x0100
#define GL_QUADS 0x0007
#define GL_MODELVIEW 0x1700
#define GL_PROJECTION 0x1701
#define GLUT_RGBA 0
#define GLUT_DOUBLE 2
#define GLUT_DEPTH 16
#define GLUT_LEFT_BUTTON 0
#define GLUT_DOWN 0
#define GLUT_UP 1
#define GLUT_ELAPSED_TIME 700

static inline void glEnable(GLenum) {}
static inline void glGenTextures(GLsizei n, GLuint *t) { for (GLsizei i = 0; i < n; ++i) t[i] = 1u + (GLuint)i; }
static inline void glBindTexture(GLenum, GLuint) {}
static inline void glTexParameteri(GLenum, GLenum, GLint) {}
static inline void glTexImage2D(GLenum, GLint, GLint, GLsizei w, GLsizei h, GLint, GLenum, GLenum, const GLvoid *) {
    ptb_shim_last_tex_extent()[0] = w;
    ptb_shim_last_tex_extent()[1] = h;
}
static inline void glFinish() {}
static inline void glViewport(GLint, GLint, GLsizei, GLsizei) {}
static inline void glClearColor(GLfloat, GLfloat, GLfloat, GLfloat) {}
static inline void glClear(GLbitfield) {}
static inline void glBegin(GLenum) {}
static inline void glEnd() {}
static inline void glTexCoord2i(GLint, GLint) {}
static inline void glVertex3f(GLfloat, GLfloat, GLfloat) {}
static inline void glMatrixMode(GLenum) {}
static inline void glLoadIdentity() {}

static inline void glutInit(int *, char **) {}
static inline void glutInitWindowSize(int, int) {}
static inline void glutInitWindowPosition(int, int) {}
static inline void glutInitDisplayMode(unsigned int) {}
static inline int glutCreateWindow(const char *) { return 1; }
static inline void glutDestroyWindow(int) {}
static inline void glutDisplayFunc(void (*)(void)) {}
static inline void glutMouseFunc(void (*)(int, int, int, int)) {}
static inline void glutIdleFunc(void (*)(void)) {}
static inline void glutKeyboardFunc(void (*)(unsigned char, int, int)) {}
static inline void glutKeyboardUpFunc(void (*)(unsigned char, int, int)) {}
static inline void glutMotionFunc(void (*)(int, int)) {}
static inline void glutMainLoop() {}
static inline void glutSwapBuffers() {}
static inline void glutPostRedisplay() {}
static inline void glutFullScreen() {}
static inline void glutReshapeWindow(int, int) {}
static inline int glutGet(GLenum what) {
    static const std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    if (what != GLUT_ELAPSED_TIME) return 0;
    return (int)std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count();
}
#endif
