// Planar argmax + statistics kernels, f16 scores, class counts 10..17 (default CTA shape).  Split over translation
// units only to compile in parallel.
#include "argmax_planar.cuh"

namespace ssb {
int launch_planar_f16_c_c(int C, const PlanarArgs &a, cudaStream_t st);
int launch_planar_f16_c_b(int C, const PlanarArgs &a, cudaStream_t st) {
    switch (C) {
        case 10: return launch_planar<PlaneF16, 10, 4, 4>(to_params(a), st);
        case 11: return launch_planar<PlaneF16, 11, 4, 4>(to_params(a), st);
        case 12: return launch_planar<PlaneF16, 12, 4, 4>(to_params(a), st);
        case 13: return launch_planar<PlaneF16, 13, 4, 4>(to_params(a), st);
        case 14: return launch_planar<PlaneF16, 14, 4, 4>(to_params(a), st);
        case 15: return launch_planar<PlaneF16, 15, 4, 4>(to_params(a), st);
        case 16: return launch_planar<PlaneF16, 16, 4, 4>(to_params(a), st);
        case 17: return launch_planar<PlaneF16, 17, 4, 4>(to_params(a), st);
        default: return launch_planar_f16_c_c(C, a, st);
    }
}
}  // namespace ssb
