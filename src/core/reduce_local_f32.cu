// reduce_local_f32.cu — kernel instantiations of the single-GPU family for f32 (see reduce_launch.inl).
#define B2R_LAUNCH_DTYPE B2R_FLOAT32
#include "reduce_launch.inl"
