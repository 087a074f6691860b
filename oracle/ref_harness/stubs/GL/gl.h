/* Typedef-only stand-in so the reference's PhysicsMesh.cu (which includes
 * cuda_gl_interop.h -> GL/gl.h) compiles on a box without OpenGL headers.
 * No GL function is declared or called: RegisterNewBuffer / Sync are never
 * used by the harness.  Test infrastructure only. */
#ifndef XRT_REF_STUB_GL_H
#define XRT_REF_STUB_GL_H
typedef unsigned int GLenum;
typedef unsigned int GLuint;
typedef int GLint;
typedef int GLsizei;
typedef unsigned char GLboolean;
typedef float GLfloat;
typedef void GLvoid;
#endif
