// Planar argmax + statistics kernels, f32 scores, class counts 10..17 (default CTA shape).  Split over translation
// units only to compile in parallel.
#include "argmax_planar.cuh"

namespace ssb {
int launch_planar_f32_c_c(int C, const PlanarArgs &a, cudaStream_t st);
int launch_planar_f32_c_b(int C, const PlanarArgs &a, cudaStream_t st) {
    switch (C) {
        case 10: return launch_planar<PlaneF32, 10, 8, 2>(to_params(a), st);
        case 11: return launch_planar<PlaneF32, 11, 8, 2>(to_params(a), st);
        case 12: return launch_planar<PlaneF32, 12, 8, 2>(to_params(a), st);
        case 13: return launch_planar<PlaneF32, 13, 8, 2>(to_params(a), st);
        case 14: return launch_planar<PlaneF32, 14, 8, 2>(to_params(a), st);
        case 15: return launch_planar<PlaneF32, 15, 8, 2>(to_params(a), st);
        case 16: return launch_planar<PlaneF32, 16, 8, 2>(to_params(a), st);
        case 17: return launch_planar<PlaneF32, 17, 8, 2>(to_params(a), st);
        default: return launch_planar_f32_c_c(C, a, st);
    }
}
}  // namespace ssb
