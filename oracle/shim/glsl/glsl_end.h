// oracle/shim/glsl/glsl_end.h -- TEST INFRASTRUCTURE ONLY.  Undoes glsl_begin.h.
#undef uniform
#undef in
#undef out
#undef layout
#undef readonly
#undef buffer
#undef layoutNodes
#undef layoutArrays
#undef nodes
#undef children_array
#undef float
#undef const
#undef main
#undef u64
#undef i64
#undef u32
#undef u8
