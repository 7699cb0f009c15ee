// Streaming argmax + statistics kernel for u8 class scores, class counts 28..32 (the ring shape that is the
// default for 19 / 21: 4 consumer warps, 2 stages).  Split over translation units only to compile in parallel.
#include "codec_u8.cuh"

namespace ssb {
int launch_stream_u8_c_d(int C, const ArgmaxParams &p, cudaStream_t st) {
    switch (C) {
        case 28: return launch_stream<CodecU8<28>, 4, 2>(p, st);
        case 29: return launch_stream<CodecU8<29>, 4, 2>(p, st);
        case 30: return launch_stream<CodecU8<30>, 4, 2>(p, st);
        case 31: return launch_stream<CodecU8<31>, 4, 2>(p, st);
        case 32: return launch_stream<CodecU8<32>, 4, 2>(p, st);
        default: return fail(SSB_ERR_LIMIT, "no streaming kernel for C=%d", C);
    }
}
}  // namespace ssb
