// trace_exact.cu -- the all-FP64, reference-order build of the tracer.
// Compiled with -fmad=false (see Makefile): no multiply-add is contracted, so every
// product and sum rounds exactly as the reference's NumPy ufuncs do.
#define RT_TU_MODE 0
#include "rt_launch_impl.cuh"

namespace rt {
const Launchers& exact_launchers() { return kLaunchers; }
}  // namespace rt
