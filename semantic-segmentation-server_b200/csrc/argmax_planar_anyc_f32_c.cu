// Planar argmax + statistics kernels, f32 scores, class counts 18..26 (default CTA shape).  Split over translation
// units only to compile in parallel.
#include "argmax_planar.cuh"

namespace ssb {
int launch_planar_f32_c_d(int C, const PlanarArgs &a, cudaStream_t st);
int launch_planar_f32_c_c(int C, const PlanarArgs &a, cudaStream_t st) {
    switch (C) {
        case 18: return launch_planar<PlaneF32, 18, 8, 2>(to_params(a), st);
        case 20: return launch_planar<PlaneF32, 20, 8, 2>(to_params(a), st);
        case 22: return launch_planar<PlaneF32, 22, 8, 2>(to_params(a), st);
        case 23: return launch_planar<PlaneF32, 23, 8, 2>(to_params(a), st);
        case 24: return launch_planar<PlaneF32, 24, 8, 2>(to_params(a), st);
        case 25: return launch_planar<PlaneF32, 25, 8, 2>(to_params(a), st);
        case 26: return launch_planar<PlaneF32, 26, 8, 2>(to_params(a), st);
        default: return launch_planar_f32_c_d(C, a, st);
    }
}
}  // namespace ssb
