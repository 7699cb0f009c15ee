// Streaming argmax + statistics kernel for fp16 class scores, class counts 10..17 (4 consumer warps, 2 stages).
// Split over translation units only to compile in parallel.
#include "codec_f16.cuh"

namespace ssb {
int launch_stream_f16_c_c(int C, const ArgmaxParams &p, cudaStream_t st);
int launch_stream_f16_c_b(int C, const ArgmaxParams &p, cudaStream_t st) {
    switch (C) {
        case 10: return launch_stream<CodecF16<10>, 4, 2>(p, st);
        case 11: return launch_stream<CodecF16<11>, 4, 2>(p, st);
        case 12: return launch_stream<CodecF16<12>, 4, 2>(p, st);
        case 13: return launch_stream<CodecF16<13>, 4, 2>(p, st);
        case 14: return launch_stream<CodecF16<14>, 4, 2>(p, st);
        case 15: return launch_stream<CodecF16<15>, 4, 2>(p, st);
        case 16: return launch_stream<CodecF16<16>, 4, 2>(p, st);
        case 17: return launch_stream<CodecF16<17>, 4, 2>(p, st);
        default: return launch_stream_f16_c_c(C, p, st);
    }
}
}  // namespace ssb
