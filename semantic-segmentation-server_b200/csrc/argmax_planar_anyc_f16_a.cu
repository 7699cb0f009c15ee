// Planar argmax + statistics kernels, f16 scores, class counts 2..9 (default CTA shape).  Split over translation
// units only to compile in parallel.
#include "argmax_planar.cuh"

namespace ssb {
int launch_planar_f16_c_b(int C, const PlanarArgs &a, cudaStream_t st);
int launch_planar_f16_other_c(int C, const PlanarArgs &a, cudaStream_t st) {
    switch (C) {
        case 2: return launch_planar<PlaneF16, 2, 4, 4>(to_params(a), st);
        case 3: return launch_planar<PlaneF16, 3, 4, 4>(to_params(a), st);
        case 4: return launch_planar<PlaneF16, 4, 4, 4>(to_params(a), st);
        case 5: return launch_planar<PlaneF16, 5, 4, 4>(to_params(a), st);
        case 6: return launch_planar<PlaneF16, 6, 4, 4>(to_params(a), st);
        case 7: return launch_planar<PlaneF16, 7, 4, 4>(to_params(a), st);
        case 8: return launch_planar<PlaneF16, 8, 4, 4>(to_params(a), st);
        case 9: return launch_planar<PlaneF16, 9, 4, 4>(to_params(a), st);
        default: return launch_planar_f16_c_b(C, a, st);
    }
}
}  // namespace ssb
