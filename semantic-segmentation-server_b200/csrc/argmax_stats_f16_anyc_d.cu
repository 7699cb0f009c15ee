// Streaming argmax + statistics kernel for fp16 class scores, class counts 27..32 (4 consumer warps, 2 stages).
// Split over translation units only to compile in parallel.
#include "codec_f16.cuh"

namespace ssb {
int launch_stream_f16_c_d(int C, const ArgmaxParams &p, cudaStream_t st) {
    switch (C) {
        case 27: return launch_stream<CodecF16<27>, 4, 2>(p, st);
        case 28: return launch_stream<CodecF16<28>, 4, 2>(p, st);
        case 29: return launch_stream<CodecF16<29>, 4, 2>(p, st);
        case 30: return launch_stream<CodecF16<30>, 4, 2>(p, st);
        case 31: return launch_stream<CodecF16<31>, 4, 2>(p, st);
        case 32: return launch_stream<CodecF16<32>, 4, 2>(p, st);
        default: return fail(SSB_ERR_LIMIT, "no streaming fp16 kernel for C=%d", C);
    }
}
}  // namespace ssb
