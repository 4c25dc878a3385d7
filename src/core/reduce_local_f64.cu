// reduce_local_f64.cu — kernel instantiations of the single-GPU family for f64 (see reduce_launch.inl).
#define B2R_LAUNCH_DTYPE B2R_FLOAT64
#include "reduce_launch.inl"
