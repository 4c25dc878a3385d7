// reduce_local_i32.cu — kernel instantiations of the single-GPU family for i32 (see reduce_launch.inl).
#define B2R_LAUNCH_DTYPE B2R_INT32
#include "reduce_launch.inl"
