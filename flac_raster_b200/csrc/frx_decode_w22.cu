// frx_decode_w22.cu -- second copy of the frame decoder: 22 one-warp CTAs per SM (see frx_decode_frame.cuh).
#define FRX_DEC_CTAS 22
#define FRX_DEC_RINGW 24
#define FRX_DEC_NS dec_w22
#define FRX_DEC_EARLY_ISSUE 0   // 80 registers leave no room for 16 more in flight across the write-out (C2: 5.17 -> 5.09 ms without)
#include "frx_decode_frame.cuh"

cudaError_t frx_launch_decode_w22(const DecArgs &a, int out_dtype, cudaStream_t st) { return dec_w22::launch(a, out_dtype, st); }
