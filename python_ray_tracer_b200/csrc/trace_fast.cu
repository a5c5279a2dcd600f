// trace_fast.cu -- the restructured-arithmetic build of the tracer (FMA contraction on).
#define RT_TU_MODE 1
#include "rt_launch_impl.cuh"

namespace rt {
const Launchers& fast_launchers() { return kLaunchers; }
int max_staged_spheres() { return (int)(kMaxDynamicSmem / staged_bytes(1)); }
}  // namespace rt
