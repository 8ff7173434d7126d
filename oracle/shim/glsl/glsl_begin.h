// oracle/shim/glsl/glsl_begin.h -- TEST INFRASTRUCTURE ONLY.
// Keyword re-typing that is in force ONLY while the reference's fragment shader is being included inside
// `struct Shader { ... }` (oracle/ref_shader.cpp); glsl_end.h takes every one of them back.
//   storage / interface qualifiers: uniforms are shared by all fragments (static members), in / out and plain
//   globals are per fragment (members, re-initialised by constructing one Shader per pixel);
//   the two SSBO blocks become anonymous structs holding a pointer to an array of unknown bound, so that
//   `nodes[i]` and `children_array[a][i]` in the shader index the harness's linear copies of the pools;
//   `float` is the binary32 class F (see glsl_as_cpp.h for why); `const` globals are all literal constants in
//   this shader (low_res.frag:110,131-134,309), so static constexpr keeps them usable as array bounds.
#define uniform static inline
#define in
#define out
#define layout(...)
#define readonly
#define buffer struct
#define layoutNodes
#define layoutArrays
#define nodes (*rc_nodes)
#define children_array (*rc_arrays)
#define float F
#define const static constexpr
#define main shader_main
