
// ==========================================================================================
// Frame encoder, round 2: ONE WARP PER FRAME, twelve independent warps per SM.
//
// A warp owns a 4096-sample frame from ticket to copy-out; nothing in the frame loop is a block-wide barrier.  A lane
// owns a contiguous block of 128 samples, so every search stage is a register sliding window over shared memory that is
// read with 16-byte loads, the level-5 Rice partitions ARE the lanes, and all reductions are shuffles.  Per warp the
// shared memory is one 16.9 KB buffer that holds, in turn and in place: the raw samples as TMA delivered them, the
// PCM, the zig-zag residuals of the winning predictor, and finally the packed frame (the packer reads a chunk of 128
// residuals into registers before it writes the bits of that chunk, and the bits never overtake the reader: checked per
// subframe from the exact bit counts, VERBATIM otherwise).  Frame sizes go through the decoupled look-back as soon as the
// exact bit count is known, so the byte offset arrives while the frame is being packed and checksummed.
// ==========================================================================================
#define FULLMASK 0xFFFFFFFFu
#ifdef FRX_DBG
// development counters (cycles, events) summed over all frame warps; read and cleared by frx_debug_read
__device__ unsigned long long g_frx_dbg[16];
#define DBG_ADD(i, v) do { if (lane == 0) atomicAdd(&g_frx_dbg[i], (unsigned long long)(v)); } while (0)
#define DBG_CLK() clock64()
#else
#define DBG_ADD(i, v) do { } while (0)
#define DBG_CLK() 0ll
#endif
#ifndef ENC_WARPS
#define ENC_WARPS 12
#endif
#define ENC_CTA (ENC_WARPS * 32)
#define SLACK_W 64                                   // words in front of the PCM area: the packed stream runs ahead by its headers
#define BLK 128                                      // samples per lane block of a full frame = samples per pack chunk
#define BLK_W (BLK + 4)                              // words per block: 4 words of padding keep 16-byte lane-strided loads conflict-free
#define PIDX(g) ((g) + 4 * ((g) >> 7))               // word of sample g inside the PCM area
#define PCM_W (FRX_MAX_BLOCK + 4 * (FRX_MAX_BLOCK / BLK))
#define WBUF_W (SLACK_W + PCM_W + 12)
// Phase alignment (kernels with ALIGNED = true, see frx_encode_kernel): the frame warps of a CTA are split into groups of
// ALIGN_GROUP warps that take their frames together and meet again at the phase boundaries of the frame (one named barrier
// per group).  Every warp of a group passes exactly EncArgs.bars_per_frame barriers per frame, whatever path its frame takes.
#ifndef ALIGN_GROUP
#define ALIGN_GROUP 6
#endif
static_assert(ENC_WARPS % ALIGN_GROUP == 0, "groups must tile the CTA");
__device__ __forceinline__ void group_bar() {
    asm volatile("bar.sync %0, %1;" :: "r"(1 + (int)(threadIdx.x >> 5) / ALIGN_GROUP), "r"(ALIGN_GROUP * 32) : "memory");
}
__device__ __forceinline__ bool group_or(bool pred) {
    int r;
    asm volatile("{ .reg .pred p, q; setp.ne.s32 q, %3, 0; bar.red.or.pred p, %1, %2, q; selp.s32 %0, 1, 0, p; }"
                 : "=r"(r) : "r"(1 + (int)(threadIdx.x >> 5) / ALIGN_GROUP), "r"(ALIGN_GROUP * 32), "r"((int)pred) : "memory");
    return r != 0;
}
#define ALIGN_BAR(n) do { if (ALIGNED) { group_bar(); (n)++; } } while (0)

__device__ __forceinline__ uint64_t warp_sum_u64(uint64_t v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULLMASK, v, o);
    return v;
}
__device__ __forceinline__ float warp_sum_f32(float v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULLMASK, v, o);
    return v;
}
__device__ __forceinline__ uint32_t zigzag(int32_t r) { return ((uint32_t)r << 1) ^ (uint32_t)(r >> 31); }
__device__ __forceinline__ uint32_t unzig_abs(uint32_t u) { return (u >> 1) + (u & 1u); }   // |r| of a zig-zag folded residual

// libFLAC-style Rice cost estimate from the partition sum (oracle/flac_oracle.c rice_bits_estimate)
__device__ __forceinline__ uint64_t rice_estimate(uint32_t k, uint32_t n, uint64_t s) {
    return 4ull + (uint64_t)(1 + k) * n + (k ? (s >> (k - 1)) : (s << 1)) - (n >> 1);
}
// smallest k with (n << k) >= s, capped at kmax  (libFLAC: rice parameter from the partition mean)
__device__ __forceinline__ uint32_t rice_param(uint32_t n, uint64_t s, uint32_t kmax) {
    if (n == 0 || s <= (uint64_t)n) return 0;
    int k = (64 - __clzll((long long)s)) - (32 - __clz((int)n)) - 1;  // floor(log2 s) - floor(log2 n) - 1
    if (k < 0) k = 0;
    while (((uint64_t)n << k) < s) k++;  // at most 2 iterations
    return (uint32_t)k > kmax ? kmax : (uint32_t)k;
}

// ------------------------------------------------------------------------------------------
// mbarrier / TMA (bulk async copy) primitives
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *b, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t parity) {
    while (!mbar_try_wait(b, parity)) {}
}
// TMA 1-D bulk copy global -> shared, completion counted in bytes on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ------------------------------------------------------------------------------------------
// per-warp shared state
// ------------------------------------------------------------------------------------------
struct FrameDesc {
    uint32_t f, fj;
    long long s0;                                // first sample of the frame inside its tile (row-major), 64-bit: tiles may hold >= 2^32 samples
    int32_t nsamp;
    frx_tile t;
    uint32_t hdr[4];
    uint32_t hdr_bits;
    int32_t tma;                                 // 1: rows are 16-byte multiples, raw samples staged by TMA bulk copies
    int32_t tile_idx;
    int16_t po_max, max_lpc;                     // largest Rice partition order / LPC order this frame can use
    FrxQuant q;
};
struct SubDec {                                  // decision record of one subframe
    uint8_t type, order, wasted, po, rice2, precision, sbps, vch;   // type: 0 CONSTANT 1 VERBATIM 2 FIXED 3 LPC; vch: virtual channel
    uint8_t fast, narrow, runs, pad1;            // fast: full frame of lean data (blocked 32-bit passes); narrow: 32-bit LPC accumulate;
                                                 // runs: every lane may pack its own block as one run (pack_runs)
    int32_t shift;
    uint32_t bits;                               // exact size of the subframe
    int16_t qlp[FRX_MAX_LPC];
    uint8_t k[64];
};
struct WarpSh {
    uint64_t mbar;
    FrameDesc fd;
    int32_t qlp[FRX_MAX_LPC];
    int32_t warm[FRX_MAX_LPC];
    uint8_t kparam[2][64];                       // candidate 0 = fixed, 1 = lpc; parameters of the candidate's best partition order
    uint64_t g_tot[5];                           // generic path: results of pass1_generic
    float g_ac[FRX_MAX_LPC + 1];
    uint32_t g_bad;
    SubDec dec[FRX_MAX_CH];
};
#define WARP_SH_BYTES ((sizeof(WarpSh) + 127) & ~(size_t)127)
#define WARP_BYTES (WARP_SH_BYTES + ((WBUF_W * 4 + 127) & ~(size_t)127))
