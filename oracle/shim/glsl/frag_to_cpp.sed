# oracle/shim/glsl/frag_to_cpp.sed -- TEST INFRASTRUCTURE ONLY.
# The ONLY textual changes made to the reference's src/shaders/low_res.frag on its way into g++ (oracle/Makefile
# pipes the file through this script; nothing is written back anywhere).  No expression, statement or constant
# of the shader is touched:
#  1. "#version" / "#extension" are not C preprocessor directives: blanked.
s/^#(version|extension).*$//
#  2. GLSL's array-type-first declaration `uint[64] children_array[];` is spelled the C way (same object:
#     an unbounded array of 64-word child tables, low_res.frag:37-39).
s/^([[:space:]]*)uint\[64\][[:space:]]+children_array\[\];/\1uint children_array[][64];/
#  3. The shader body is compiled inside a struct (one instance per fragment, so that its globals start fresh for
#     every pixel as GLSL requires); member functions may be used before they are defined and may not be declared
#     twice, so the three forward declarations (low_res.frag:122-124) are blanked.
s/^void (getBlock|nextIntersect|nextIntersectDDA)\([^)]*\);[[:space:]]*$//
