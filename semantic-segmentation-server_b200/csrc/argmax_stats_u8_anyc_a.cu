// Streaming argmax + statistics kernel for u8 class scores, class counts 2..11 (the ring shape that is the
// default for 19 / 21: 4 consumer warps, 2 stages).  Split over translation units only to compile in parallel.
#include "codec_u8.cuh"

namespace ssb {
int launch_stream_u8_c_b(int C, const ArgmaxParams &p, cudaStream_t st);
int launch_stream_u8_other_c(int C, const ArgmaxParams &p, cudaStream_t st) {
    switch (C) {
        case 2: return launch_stream<CodecU8<2>, 4, 2>(p, st);
        case 3: return launch_stream<CodecU8<3>, 4, 2>(p, st);
        case 4: return launch_stream<CodecU8<4>, 4, 2>(p, st);
        case 5: return launch_stream<CodecU8<5>, 4, 2>(p, st);
        case 6: return launch_stream<CodecU8<6>, 4, 2>(p, st);
        case 7: return launch_stream<CodecU8<7>, 4, 2>(p, st);
        case 8: return launch_stream<CodecU8<8>, 4, 2>(p, st);
        case 9: return launch_stream<CodecU8<9>, 4, 2>(p, st);
        case 10: return launch_stream<CodecU8<10>, 4, 2>(p, st);
        case 11: return launch_stream<CodecU8<11>, 4, 2>(p, st);
        default: return launch_stream_u8_c_b(C, p, st);
    }
}
}  // namespace ssb
