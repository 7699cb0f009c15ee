// Planar argmax + statistics kernels, f32 scores, class counts 27..32 (default CTA shape).  Split over translation
// units only to compile in parallel.
#include "argmax_planar.cuh"

namespace ssb {
int launch_planar_f32_c_d(int C, const PlanarArgs &a, cudaStream_t st) {
    switch (C) {
        case 27: return launch_planar<PlaneF32, 27, 8, 2>(to_params(a), st);
        case 28: return launch_planar<PlaneF32, 28, 8, 2>(to_params(a), st);
        case 29: return launch_planar<PlaneF32, 29, 8, 2>(to_params(a), st);
        case 30: return launch_planar<PlaneF32, 30, 8, 2>(to_params(a), st);
        case 31: return launch_planar<PlaneF32, 31, 8, 2>(to_params(a), st);
        case 32: return launch_planar<PlaneF32, 32, 8, 2>(to_params(a), st);
        default: return fail(SSB_ERR_LIMIT, "no planar kernel for C=%d", C);
    }
}
}  // namespace ssb
