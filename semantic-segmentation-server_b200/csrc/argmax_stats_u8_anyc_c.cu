// Streaming argmax + statistics kernel for u8 class scores, class counts 22..27 (the ring shape that is the
// default for 19 / 21: 4 consumer warps, 2 stages).  Split over translation units only to compile in parallel.
#include "codec_u8.cuh"

namespace ssb {
int launch_stream_u8_c_d(int C, const ArgmaxParams &p, cudaStream_t st);
int launch_stream_u8_c_c(int C, const ArgmaxParams &p, cudaStream_t st) {
    switch (C) {
        case 22: return launch_stream<CodecU8<22>, 4, 2>(p, st);
        case 23: return launch_stream<CodecU8<23>, 4, 2>(p, st);
        case 24: return launch_stream<CodecU8<24>, 4, 2>(p, st);
        case 25: return launch_stream<CodecU8<25>, 4, 2>(p, st);
        case 26: return launch_stream<CodecU8<26>, 4, 2>(p, st);
        case 27: return launch_stream<CodecU8<27>, 4, 2>(p, st);
        default: return launch_stream_u8_c_d(C, p, st);
    }
}
}  // namespace ssb
