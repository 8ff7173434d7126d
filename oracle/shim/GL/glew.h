// oracle/shim/GL/glew.h -- TEST INFRASTRUCTURE ONLY.
// No-op OpenGL surface: just enough for src/globals.hpp:32-43 (checkGlErrorFunc)
// and src/voxel_data/voxel_allocator.hpp:38-91 (updateSsboData /
// initVoxelDataAllocator) of the reference to compile unmodified.  There is no
// graphics stack on this box; nothing here talks to a GPU.
//
// The buffer calls are routed to optional hooks so a test can observe exactly
// what the reference's updateSsboData() would have sent through
// glBufferData/glBufferSubData (the boundary rc_upload_world replaces).
#pragma once
#include <stddef.h>
#include <stdint.h>

typedef unsigned int GLuint;
typedef int GLint;
typedef unsigned int GLenum;
typedef int GLsizei;
typedef char GLchar;
typedef ptrdiff_t GLsizeiptr;
typedef ptrdiff_t GLintptr;

#define GL_NO_ERROR 0
#define GL_INFO_LOG_LENGTH 0x8B84
#define GL_SHADER_STORAGE_BUFFER 0x90D2
#define GL_DYNAMIC_DRAW 0x88E8

struct ShimGlHooks {
    // (buffer id bound to GL_SHADER_STORAGE_BUFFER, byte size)
    void (*buffer_data)(GLuint buffer, GLsizeiptr size);
    // (buffer id, byte offset, byte count, source)
    void (*buffer_sub_data)(GLuint buffer, GLintptr offset, GLsizeiptr size, const void* data);
};
extern ShimGlHooks g_shimGlHooks;
extern GLuint g_shimBoundSsbo;
extern GLuint g_shimNextBufferName;

inline GLenum glGetError() { return GL_NO_ERROR; }
inline void glGetProgramiv(GLuint, GLenum, GLint* v) { if (v) *v = 1; }
inline void glGetProgramInfoLog(GLuint, GLsizei, GLsizei* len, GLchar* log) {
    if (len) *len = 0;
    if (log) log[0] = 0;
}
inline void glGenBuffers(GLsizei n, GLuint* out) {
    for (GLsizei i = 0; i < n; i++) out[i] = ++g_shimNextBufferName;
}
inline void glBindBuffer(GLenum, GLuint b) { g_shimBoundSsbo = b; }
inline void glBindBufferBase(GLenum, GLuint, GLuint) {}
inline void glBufferData(GLenum, GLsizeiptr size, const void*, GLenum) {
    if (g_shimGlHooks.buffer_data) g_shimGlHooks.buffer_data(g_shimBoundSsbo, size);
}
inline void glBufferSubData(GLenum, GLintptr off, GLsizeiptr size, const void* data) {
    if (g_shimGlHooks.buffer_sub_data) g_shimGlHooks.buffer_sub_data(g_shimBoundSsbo, off, size, data);
}
