#ifdef FRX_DBG
__device__ unsigned long long g_frx_dbg[16];
#define DBG_ADD(i, v) atomicAdd(&g_frx_dbg[i], (unsigned long long)(v))
#define DBG_CLK() clock64()
#define DBG_PH(i) do { if (tid == 32 * DBG_PH_WARP) { const long long n_ = clock64(); atomicAdd(&g_frx_dbg[i], (unsigned long long)(n_ - dbg_ph)); dbg_ph = n_; } } while (0)
#define DBG_PH_DECL long long dbg_ph = clock64()
#ifndef DBG_PH_WARP
#define DBG_PH_WARP 2
#endif
__device__ unsigned long long *g_frx_trace;   // 8 timestamps per frame
__device__ __forceinline__ unsigned long long dbg_now() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define DBG_TRACE(f, k) do { if (g_frx_trace) g_frx_trace[(size_t)(f) * 8 + (k)] = dbg_now(); } while (0)
#define DBG_TRACEV(f, k, v) do { if (g_frx_trace) g_frx_trace[(size_t)(f) * 8 + (k)] = (v); } while (0)
#else
#define DBG_ADD(i, v) do {} while (0)
#define DBG_CLK() 0ll
#define DBG_PH(i) do {} while (0)
#define DBG_PH_DECL
#define DBG_TRACE(f, k) do {} while (0)
#define DBG_TRACEV(f, k, v) do {} while (0)
#endif
// ------------------------------------------------------------------------------------------
// shared state of one encoder CTA
// ------------------------------------------------------------------------------------------
#define FD_SLOTS 4
#define STOP_FRAME 0xFFFFFFFFu
#define TABS_BYTES (6 * FRX_ENC_THREADS * 8)     // |e| sums per thread: fixed orders 0..4 + the LPC candidate
// PCM of the current subframe in shared memory, padded so that a thread's 16-byte chunks at a 64-byte stride fall into
// distinct bank groups: sample i lives at word PCM_IDX(i); 32 zero samples sit in front (history of the first thread).
#define PCM_IDX(i) (((i) + 32) + 4 * (((i) + 32) >> 5))
#define PCM_WORDS 4672
#define PCM_BYTES (PCM_WORDS * 4)

struct FrameDesc {                               // written by the control warp, read by the compute warps
    uint32_t f, fj, s0;
    int32_t nsamp;
    frx_tile t;
    uint32_t hdr[4];
    uint32_t hdr_bits;
    int32_t tma;                                 // 1: raw samples staged by TMA bulk copies; 0: compute warps gather them
    int32_t tile_idx;
    int16_t po_max, max_lpc;                     // largest Rice partition order / LPC order this frame can use
    FrxQuant q;
};

struct EncSmall {
    uint64_t full[1], empty[1];                  // raw stage hand-over (mbarriers): one stage, refilled as soon as K1 has read it
    uint64_t sized[FD_SLOTS], placed[FD_SLOTS];  // frame size known -> look-back -> byte offset known
    FrameDesc fd[FD_SLOTS];
    uint32_t frame_bytes[FD_SLOTS];
    uint64_t out_offset[FD_SLOTS];
    // per-subframe decisions
    int wasted, magbits, fixed_order, is_const, lean;
    int lpc_ok, lpc_order, shift, precision, lpc_narrow;
    int32_t qlp[FRX_MAX_LPC];
    int32_t warm[FRX_MAX_LPC];
    uint32_t kparam[2][128];                     // candidate 0 = fixed, 1 = lpc; level L's parameters start at (1 << L) - 1
    uint64_t lvl_bits[2][8];                     // Rice search: estimated bits of partition order L (~0: not evaluated)
    uint32_t lvl_big[2][8];                      // ... and whether it needs 5-bit parameters (RICE2)
    // reductions
    uint64_t wfix[NWARPS][5];
    uint32_t wor[NWARPS][2];
    uint32_t wbad[NWARPS][2];
    float wac[NWARPS][FRX_MAX_LPC + 1];
    float autoc[FRX_MAX_LPC + 1];
    float lpc_rows[FRX_MAX_LPC][FRX_MAX_LPC];    // predictor coefficients of every order (Levinson-Durbin by-product)
    uint32_t scan[NWARPS], scan2[NWARPS];
    uint32_t crc_part[2][NWARPS];                // per frame buffer
    uint16_t crc_tab[4][256];
};
#define RES_BYTES (FRX_MAX_BLOCK * 4)             // the candidate's residuals, sample-major: res[i * 256 + tid]
#define RES_IDX(i, tid) ((i) * FRX_ENC_THREADS + (tid))
#define ENC_SMALL_BYTES ((sizeof(EncSmall) + 127) & ~(size_t)127)

// ------------------------------------------------------------------------------------------
// mbarrier / TMA (bulk async copy) / named barrier primitives
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
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
// non-blocking probe: try_wait may park the thread for a system-defined time, which a loop polling two barriers cannot afford
__device__ __forceinline__ bool mbar_test(uint64_t *b, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
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
__device__ __forceinline__ void compute_sync() { asm volatile("bar.sync 1, %0;" ::"n"(FRX_ENC_THREADS) : "memory"); }

