// Streaming argmax + statistics kernel for fp16 class scores, class counts 18..26 (4 consumer warps, 2 stages).
// Split over translation units only to compile in parallel.
#include "codec_f16.cuh"

namespace ssb {
int launch_stream_f16_c_d(int C, const ArgmaxParams &p, cudaStream_t st);
int launch_stream_f16_c_c(int C, const ArgmaxParams &p, cudaStream_t st) {
    switch (C) {
        case 18: return launch_stream<CodecF16<18>, 4, 2>(p, st);
        case 20: return launch_stream<CodecF16<20>, 4, 2>(p, st);
        case 22: return launch_stream<CodecF16<22>, 4, 2>(p, st);
        case 23: return launch_stream<CodecF16<23>, 4, 2>(p, st);
        case 24: return launch_stream<CodecF16<24>, 4, 2>(p, st);
        case 25: return launch_stream<CodecF16<25>, 4, 2>(p, st);
        case 26: return launch_stream<CodecF16<26>, 4, 2>(p, st);
        default: return launch_stream_f16_c_d(C, p, st);
    }
}
}  // namespace ssb
