// frx_decode_al.cu -- third copy of the frame decoder: ONE CTA of 21 warps per SM whose warps meet at every chunk (see
// dec_bar in frx_decode_frame.cuh), for calls that need between 20 and 21 warps per SM.
#define FRX_DEC_CTAS 1
#define FRX_DEC_WARPS 21
#define FRX_DEC_RINGW 32
#define FRX_DEC_NS dec_al
#include "frx_decode_frame.cuh"

cudaError_t frx_launch_decode_al(const DecArgs &a, int out_dtype, cudaStream_t st) { return dec_al::launch(a, out_dtype, st); }
