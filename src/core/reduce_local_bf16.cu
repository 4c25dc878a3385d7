// reduce_local_bf16.cu — kernel instantiations of the single-GPU family for bf16 (see reduce_launch.inl).
#define B2R_LAUNCH_DTYPE B2R_BFLOAT16
#include "reduce_launch.inl"
