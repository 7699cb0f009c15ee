// Streaming argmax + statistics kernel for u8 class scores, class counts 12..20 (the ring shape that is the
// default for 19 / 21: 4 consumer warps, 2 stages).  Split over translation units only to compile in parallel.
#include "codec_u8.cuh"

namespace ssb {
int launch_stream_u8_c_c(int C, const ArgmaxParams &p, cudaStream_t st);
int launch_stream_u8_c_b(int C, const ArgmaxParams &p, cudaStream_t st) {
    switch (C) {
        case 12: return launch_stream<CodecU8<12>, 4, 2>(p, st);
        case 13: return launch_stream<CodecU8<13>, 4, 2>(p, st);
        case 14: return launch_stream<CodecU8<14>, 4, 2>(p, st);
        case 15: return launch_stream<CodecU8<15>, 4, 2>(p, st);
        case 16: return launch_stream<CodecU8<16>, 4, 2>(p, st);
        case 17: return launch_stream<CodecU8<17>, 4, 2>(p, st);
        case 18: return launch_stream<CodecU8<18>, 4, 2>(p, st);
        case 20: return launch_stream<CodecU8<20>, 4, 2>(p, st);
        default: return launch_stream_u8_c_c(C, p, st);
    }
}
}  // namespace ssb
