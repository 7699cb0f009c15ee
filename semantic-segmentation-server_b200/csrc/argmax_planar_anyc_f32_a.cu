// Planar argmax + statistics kernels, f32 scores, class counts 2..9 (default CTA shape).  Split over translation
// units only to compile in parallel.
#include "argmax_planar.cuh"

namespace ssb {
int launch_planar_f32_c_b(int C, const PlanarArgs &a, cudaStream_t st);
int launch_planar_f32_other_c(int C, const PlanarArgs &a, cudaStream_t st) {
    switch (C) {
        case 2: return launch_planar<PlaneF32, 2, 8, 2>(to_params(a), st);
        case 3: return launch_planar<PlaneF32, 3, 8, 2>(to_params(a), st);
        case 4: return launch_planar<PlaneF32, 4, 8, 2>(to_params(a), st);
        case 5: return launch_planar<PlaneF32, 5, 8, 2>(to_params(a), st);
        case 6: return launch_planar<PlaneF32, 6, 8, 2>(to_params(a), st);
        case 7: return launch_planar<PlaneF32, 7, 8, 2>(to_params(a), st);
        case 8: return launch_planar<PlaneF32, 8, 8, 2>(to_params(a), st);
        case 9: return launch_planar<PlaneF32, 9, 8, 2>(to_params(a), st);
        default: return launch_planar_f32_c_b(C, a, st);
    }
}
}  // namespace ssb
