// Planar argmax + statistics kernels, f16 scores, class counts 18..26 (default CTA shape).  Split over translation
// units only to compile in parallel.
#include "argmax_planar.cuh"

namespace ssb {
int launch_planar_f16_c_d(int C, const PlanarArgs &a, cudaStream_t st);
int launch_planar_f16_c_c(int C, const PlanarArgs &a, cudaStream_t st) {
    switch (C) {
        case 18: return launch_planar<PlaneF16, 18, 4, 4>(to_params(a), st);
        case 20: return launch_planar<PlaneF16, 20, 4, 4>(to_params(a), st);
        case 22: return launch_planar<PlaneF16, 22, 4, 4>(to_params(a), st);
        case 23: return launch_planar<PlaneF16, 23, 4, 4>(to_params(a), st);
        case 24: return launch_planar<PlaneF16, 24, 4, 4>(to_params(a), st);
        case 25: return launch_planar<PlaneF16, 25, 4, 4>(to_params(a), st);
        case 26: return launch_planar<PlaneF16, 26, 4, 4>(to_params(a), st);
        default: return launch_planar_f16_c_d(C, a, st);
    }
}
}  // namespace ssb
