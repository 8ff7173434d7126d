// oracle/ref_shader.cpp -- TEST INFRASTRUCTURE ONLY.
//
// The reference's fragment shader src/shaders/low_res.frag -- the ONLY implementation of the render hot path
// the reference has -- compiled by g++ from where it lies under /root/reference and run on the host cores, one
// `Shader` instance per pixel.  oracle/Makefile builds this into oracle/_ref/libref_shader.so; nothing of the
// shader is copied into this repository (REF_FRAG names a pipe: sed -f shim/glsl/frag_to_cpp.sed <the file>).
//
// Used by tests/test_ref_shader_cpu.py to hold the restated oracle (oracle_render.cpp) to the reference's own
// source pixel for pixel, and by bench.py's cpu_baseline / --impl reference legs as the "reference" CPU arm of
// the configurations the live shader covers (primary + shadow rays: BASELINE configs 1 and 2).
//
// What the host side of the reference does around the shader is restated here with citations:
//   * uniforms: src/main.cpp:83-95 (render());  originMortonPos = getMortonPos(cameraPos) src/main.cpp:45-66
//   * gl_FragCoord = pixel centre (x + 0.5, y + 0.5), y = 0 at the bottom row (OpenGL window coordinates)
//   * the two SSBOs hold the host pools at identical byte offsets: src/voxel_data/voxel_allocator.hpp:38-91
#include <cstring>
#include <omp.h>

#include "glsl_as_cpp.h"

namespace glslcpp {
static float libm_sin(float x) { return std::sin(x); }
static float libm_cos(float x) { return std::cos(x); }
static float libm_exp(float x) { return std::exp(x); }
float (*sin_hook)(float) = libm_sin;
float (*cos_hook)(float) = libm_cos;
float (*exp_hook)(float) = libm_exp;
}  // namespace glslcpp

namespace refshader {
using namespace glslcpp;

struct Shader {
#include "glsl_begin.h"
#include REF_FRAG
#include "glsl_end.h"
};

}  // namespace refshader

using refshader::Shader;

namespace {

// src/main.cpp:45-66, getMortonPos: 10 bits per axis, two bits per axis per 6-bit group
uint32_t spread10(int n) {
    n = (n | (n << 16)) & 0x030000FF;
    n = (n | (n << 8)) & 0x0300F00F;
    n = (n | (n << 4)) & 0x030C30C3;
    return (uint32_t)n;
}
uint32_t morton_of(const float* p) {
    const int x = (int)std::floor(p[0]), y = (int)std::floor(p[1]), z = (int)std::floor(p[2]);
    return (spread10(z) << 4) | (spread10(y) << 2) | spread10(x);
}

const glslcpp::u64vec2* g_nodes = nullptr;
const uint32_t* g_arrays = nullptr;

}  // namespace

static int g_lastThreads = 1;

extern "C" {

// Byte layout shared with oracle/bindings.py (RefShaderFrame).
struct RefShaderFrame {
    float origin[3];         // trueOrigin
    float camera_dir[3];     // cameraDir
    float sun_dir[3];        // sunDir
    float mouse_pos[2];      // mousePos
    float proj_plane[2];     // projPlaneSize
    float delta_time;        // deltaTime
    int32_t looking_at[3];   // lookingAtBlock
    int32_t render_pos_data; // renderPosData
    uint32_t width, height;  // resolution
    uint32_t row_begin, row_end;          // rows [row_begin, row_end)
    uint32_t col_stride, row_stride;      // every col_stride-th pixel of every row_stride-th row (bounded samples)
    int32_t threads;                      // OpenMP threads, <= 0: all
};

// What is left in the shader's globals when main() returns (for pixels that never start the shadow ray this is
// the state at the end of the primary loops).
struct RefShaderState {
    int32_t round[3];      // pos.round
    float exact[3];        // pos.exact
    uint32_t last_hit;     // lastHit
    uint32_t reflected;    // rayReflected
    uint32_t block_empty;  // block.x == -1
    float refractive_index;// currentRefractiveIndex
    float color_mod[3];    // finalColorMod
};

void refshader_set_world(const void* nodes_linear, const void* arrays_linear) {
    g_nodes = (const glslcpp::u64vec2*)nodes_linear;
    g_arrays = (const uint32_t*)arrays_linear;
}

// sin / cos / exp of the shader: null = libm.  The oracle's deterministic versions are plugged in by the tests
// that compare colours bit for bit.
void refshader_set_math(float (*s)(float), float (*c)(float), float (*e)(float)) {
    glslcpp::sin_hook = s ? s : glslcpp::libm_sin;
    glslcpp::cos_hook = c ? c : glslcpp::libm_cos;
    glslcpp::exp_hook = e ? e : glslcpp::libm_exp;
}

// rgba: height x width x 4 floats (FragColor), row 0 = bottom row; state: height x width records or null.
// Pixels skipped by the strides are left untouched.  Returns the number of fragments shaded, < 0 on error.
long long refshader_render(const RefShaderFrame* f, float* rgba, RefShaderState* state) {
    using namespace glslcpp;
    if (!g_nodes || !g_arrays || !f || !rgba || f->width == 0 || f->height == 0) return -1;
    Shader::trueOrigin = vec3(f->origin[0], f->origin[1], f->origin[2]);
    Shader::cameraDir = vec3(f->camera_dir[0], f->camera_dir[1], f->camera_dir[2]);
    Shader::mousePos = vec2(f->mouse_pos[0], f->mouse_pos[1]);
    Shader::renderPosData = f->render_pos_data;
    Shader::rootNodeIndex = 0;
    Shader::originMortonPos = morton_of(f->origin);
    Shader::projPlaneSize = vec2(f->proj_plane[0], f->proj_plane[1]);
    Shader::resolution = uvec2(f->width, f->height);
    Shader::sunDir = vec3(f->sun_dir[0], f->sun_dir[1], f->sun_dir[2]);
    Shader::lookingAtBlock = ivec3(f->looking_at[0], f->looking_at[1], f->looking_at[2]);
    Shader::deltaTime = f->delta_time;

    const uint32_t r0 = f->row_begin, r1 = f->row_end > f->height ? f->height : f->row_end;
    const uint32_t cs = f->col_stride ? f->col_stride : 1, rs = f->row_stride ? f->row_stride : 1;
    const int threads = f->threads > 0 ? f->threads : omp_get_max_threads();
    long long shaded = 0;
#pragma omp parallel num_threads(threads)
    if (omp_get_thread_num() == 0) g_lastThreads = omp_get_num_threads();  // what the run really gets (bench.py reports it)
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads) reduction(+ : shaded)
    for (uint32_t y = r0; y < r1; y += rs) {
        for (uint32_t x = 0; x < f->width; x += cs) {
            Shader s{};
            s.rc_nodes = (decltype(s.rc_nodes))g_nodes;
            s.rc_arrays = (decltype(s.rc_arrays))g_arrays;
            s.gl_FragCoord = vec4((float)x + 0.5f, (float)y + 0.5f, 0.5f, 1.0f);
            s.shader_main();
            const size_t i = (size_t)y * f->width + x;
            rgba[4 * i + 0] = s.FragColor.xy.x.v;
            rgba[4 * i + 1] = s.FragColor.xy.y.v;
            rgba[4 * i + 2] = s.FragColor.z.v;
            rgba[4 * i + 3] = s.FragColor.w.v;
            if (state) {
                RefShaderState& o = state[i];
                for (int k = 0; k < 3; ++k) {
                    o.round[k] = s.pos.round[k];
                    o.exact[k] = s.pos.exact[k].v;
                    o.color_mod[k] = s.finalColorMod[k].v;
                }
                o.last_hit = s.lastHit;
                o.reflected = s.rayReflected ? 1u : 0u;
                o.block_empty = s.block.x == ~0ull ? 1u : 0u;
                o.refractive_index = s.currentRefractiveIndex.v;
            }
            ++shaded;
        }
    }
    return shaded;
}

/* OpenMP threads the last refshader_render really ran on. */
int refshader_last_threads(void) { return g_lastThreads; }

uint32_t refshader_sizeof_frame(void) { return (uint32_t)sizeof(RefShaderFrame); }
uint32_t refshader_sizeof_state(void) { return (uint32_t)sizeof(RefShaderState); }

}  // extern "C"
