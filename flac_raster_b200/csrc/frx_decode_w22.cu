// frx_decode_w22.cu -- second copy of the frame decoder: 22 one-warp CTAs per SM (see frx_decode_frame.cuh).
#define FRX_DEC_CTAS 22
#define FRX_DEC_RINGW 24
#define FRX_DEC_NS dec_w22
#include "frx_decode_frame.cuh"

cudaError_t frx_launch_decode_w22(const DecArgs &a, int out_dtype, cudaStream_t st) { return dec_w22::launch(a, out_dtype, st); }
