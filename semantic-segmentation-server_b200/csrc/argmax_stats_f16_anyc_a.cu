// Streaming argmax + statistics kernel for fp16 class scores, class counts 2..9 (4 consumer warps, 2 stages).
// Split over translation units only to compile in parallel.
#include "codec_f16.cuh"

namespace ssb {
int launch_stream_f16_c_b(int C, const ArgmaxParams &p, cudaStream_t st);
int launch_stream_f16_other_c(int C, const ArgmaxParams &p, cudaStream_t st) {
    switch (C) {
        case 2: return launch_stream<CodecF16<2>, 4, 2>(p, st);
        case 3: return launch_stream<CodecF16<3>, 4, 2>(p, st);
        case 4: return launch_stream<CodecF16<4>, 4, 2>(p, st);
        case 5: return launch_stream<CodecF16<5>, 4, 2>(p, st);
        case 6: return launch_stream<CodecF16<6>, 4, 2>(p, st);
        case 7: return launch_stream<CodecF16<7>, 4, 2>(p, st);
        case 8: return launch_stream<CodecF16<8>, 4, 2>(p, st);
        case 9: return launch_stream<CodecF16<9>, 4, 2>(p, st);
        default: return launch_stream_f16_c_b(C, p, st);
    }
}
}  // namespace ssb
