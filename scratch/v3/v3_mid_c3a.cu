// ------------------------------------------------------------------------------------------
// The frame encoder.  WIDE: PCM wider than 16 bits (64-bit LPC accumulate, RICE2).  HISTN: history samples a thread
// reads = the largest LPC order of the compression level (4 when the level has no LPC stage).
// 8 compute warps (one frame at a time, 16 samples per thread) + 1 control warp.  The PCM of the current subframe
// lives in padded shared memory and every stage re-reads it, so almost no register state crosses a barrier.
// ------------------------------------------------------------------------------------------
template <bool WIDE, int HISTN>
__global__ void __launch_bounds__(FRX_ENC_CTA, 2) frx_encode_kernel(const EncArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    EncSmall &sh = *reinterpret_cast<EncSmall *>(smem_raw);
    const uint32_t stage_stride = 128u + (uint32_t)FRX_MAX_BLOCK * (uint32_t)frx_dtype_size(a.dtype);
    unsigned char *raw = smem_raw + ENC_SMALL_BYTES;
    int32_t *pcm = reinterpret_cast<int32_t *>(raw + stage_stride);
    int32_t *res = reinterpret_cast<int32_t *>(reinterpret_cast<unsigned char *>(pcm) + PCM_BYTES);
    uint64_t *tabs = reinterpret_cast<uint64_t *>(reinterpret_cast<unsigned char *>(res) + RES_BYTES);
    uint32_t *bits0 = reinterpret_cast<uint32_t *>(reinterpret_cast<unsigned char *>(tabs) + TABS_BYTES);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < 2 * a.frame_words; i += FRX_ENC_CTA) bits0[i] = 0u;
    for (int i = tid; i < PCM_WORDS; i += FRX_ENC_CTA) pcm[i] = 0;
    for (int i = tid; i < 1024; i += FRX_ENC_CTA) (&sh.crc_tab[0][0])[i] = a.crc_tab[i];
    if (tid == 0) {
        mbar_init(&sh.full[0], 1); mbar_init(&sh.empty[0], NWARPS);
        for (int i = 0; i < FD_SLOTS; i++) { mbar_init(&sh.sized[i], 1); mbar_init(&sh.placed[i], 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    if (warp == NWARPS) { control_warp(a, sh, raw, lane); return; }

    const uint32_t kmax = a.bps > 16 ? 30u : 14u;  // RICE2 only for bps > 16 (libFLAC)
    const int g0 = tid * FRX_SPT;                  // first sample index of this thread
    uint32_t unit = 0, fseq = 0;
    // copy-out runs one frame behind (double-buffered frame buffer), so the look-back never stalls the compute warps
    bool pend = false;
    uint32_t pend_body = 0;

    const long long ck0 = DBG_CLK();
    DBG_PH_DECL;
    for (;;) {
        // ---- next frame: its descriptor is valid once the first raw stage of the frame is full ----
        const int slot = (int)(fseq & (FD_SLOTS - 1));
        const int buf = (int)(fseq & 1u);
        uint32_t *bits = bits0 + (size_t)buf * a.frame_words;
        mbar_wait(&sh.full[0], unit & 1u);
        const FrameDesc &fd = sh.fd[slot];
        const uint32_t f = fd.f;
        if (f == STOP_FRAME) { if (tid == 0) DBG_ADD(7, DBG_CLK() - ck0); break; }
        if (tid == 0) DBG_TRACE(f, 4);
        DBG_PH(15);
        const int n_channels = fd.t.n_channels;
        const uint32_t s0 = fd.s0;
        const int nsamp = fd.nsamp;
        const bool use_tma = fd.tma != 0;
        const int nv = min(FRX_SPT, max(0, nsamp - g0));  // valid samples of this thread
        const bool full = (nsamp == FRX_MAX_BLOCK);
        const bool plain = full && tid != 0;              // no per-sample masks needed for this thread
        uint32_t bitpos = fd.hdr_bits;
        if (tid < 4) bits[tid] = fd.hdr[tid];
        const int po_max = fd.po_max;
        const int max_lpc = min((int)fd.max_lpc, HISTN);

        // ================= per channel =================
        for (int ch = 0; ch < n_channels; ch++) {
            unsigned char *stg = raw;
            if (ch > 0) mbar_wait(&sh.full[0], unit & 1u);
            if (!use_tma) {
                gather_raw(a, fd.t, ch, s0, nsamp, stg, tid);
                compute_sync();
            }
            // ---- K1: the thread's 16 samples -> PCM -> shared memory ----
            uint32_t orv = 0, ora = 0;
            {
                uint2 oo;
                if (a.dtype == FRX_F32) oo = k1_quantize<FRX_F32, WIDE>(stg, fd.q, a.pcm_mode, tid, nv, pcm);
                else if (a.dtype == FRX_U16) oo = k1_quantize<FRX_U16, WIDE>(stg, fd.q, a.pcm_mode, tid, nv, pcm);
                else oo = k1_quantize_any<WIDE>(a.dtype, stg, fd.q, a.pcm_mode, tid, nv, pcm);
                orv = oo.x; ora = oo.y;
                // the raw stage may be refilled as soon as every warp has read its samples
                if (!use_tma) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(&sh.empty[0]);
                unit++;
            }
            orv = __reduce_or_sync(FULLMASK, orv);
            if (WIDE) ora = __reduce_or_sync(FULLMASK, ora);
            compute_sync();  // S0: PCM of the subframe is in shared memory
            DBG_PH(9);
            // ---- K2 (lean, 32-bit) fixed-order trial + K3 windowed autocorrelation partials ----
            {
                int32_t s[HISTN + FRX_SPT];
                read_pcm<HISTN>(pcm, tid, 0, s);
                uint32_t fa[5];
                if (plain) fixed_trial32<HISTN>(s, fa);
                else fixed_trial32_masked<HISTN>(s, g0, nv, fa);
#pragma unroll
                for (int o = 0; o < 5; o++) tabs[o * FRX_ENC_THREADS + tid] = fa[o];
                {
                    uint64_t w5[5];
#pragma unroll
                    for (int o = 0; o < 5; o++) w5[o] = WIDE ? warp_sum_u32_wide(fa[o]) : (uint64_t)__reduce_add_sync(FULLMASK, fa[o]);
                    if (lane == 0) {
#pragma unroll
                        for (int o = 0; o < 5; o++) sh.wfix[warp][o] = w5[o];
                        sh.wor[warp][0] = orv; sh.wor[warp][1] = ora;
                    }
                }
                if (max_lpc > 0) {
                    float xw[HISTN + FRX_SPT];
                    if (full) {
                        // window table: zero history of thread 0 comes from the zero samples in front of the PCM buffer
                        const float4 *w4 = reinterpret_cast<const float4 *>(a.window + g0 - HISTN);
#pragma unroll
                        for (int v = 0; v < (HISTN + FRX_SPT) / 4; v++) {
                            float4 wv = make_float4(0.f, 0.f, 0.f, 0.f);
                            if (tid > 0 || v >= HISTN / 4) wv = __ldg(w4 + v);
                            xw[4 * v] = (float)s[4 * v] * wv.x; xw[4 * v + 1] = (float)s[4 * v + 1] * wv.y;
                            xw[4 * v + 2] = (float)s[4 * v + 2] * wv.z; xw[4 * v + 3] = (float)s[4 * v + 3] * wv.w;
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < HISTN + FRX_SPT; i++) xw[i] = (float)s[i] * tukey_w(g0 - HISTN + i, nsamp);
                    }
                    float ac[HISTN + 1];
#pragma unroll
                    for (int lag = 0; lag <= HISTN; lag++) ac[lag] = 0.f;
#pragma unroll
                    for (int i = 0; i < FRX_SPT; i++) {
#pragma unroll
                        for (int lag = 0; lag <= HISTN; lag++) ac[lag] = fmaf(xw[HISTN + i], xw[HISTN + i - lag], ac[lag]);
                    }
#pragma unroll
                    for (int lag = 0; lag <= HISTN; lag++) {
                        float v = ac[lag];
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULLMASK, v, o);
                        if (lane == 0) sh.wac[warp][lag] = v;
                    }
                }
            }
            compute_sync();  // SA: per-warp partials and tabs[0..4] are in shared memory
            DBG_PH(10);
            // ---- every warp: fixed-order decision (cheap, redundant).  Warps 0..6: Rice search of the fixed candidate, one
            //      partition order each; warp 7: LPC design ----
            {
                uint64_t tot = 0;
                uint32_t o32 = 0;
                if (lane < 5) { for (int w2 = 0; w2 < NWARPS; w2++) tot += sh.wfix[w2][lane]; }
                else if (lane < 7) { for (int w2 = 0; w2 < NWARPS; w2++) o32 |= sh.wor[w2][lane - 5]; }
                const uint32_t all_or = __shfl_sync(FULLMASK, o32, 5), all_abs = __shfl_sync(FULLMASK, o32, 6);
                const int wasted = all_or ? (__ffs((int)all_or) - 1) : 0;
                const int magbits = WIDE ? 32 - __clz((int)all_abs) : 15;  // bits of the largest magnitude
                const bool lean = !WIDE || magbits <= 23;
                if (warp == NWARPS - 1) {
                    float acs = 0.f;
                    if (lane <= max_lpc) { for (int w2 = 0; w2 < NWARPS; w2++) acs += sh.wac[w2][lane]; }
                    if (lane <= FRX_MAX_LPC) sh.autoc[lane] = acs;
                    if (lane == 0) { sh.magbits = magbits; sh.lean = lean ? 1 : 0; sh.lpc_ok = 0; }
                    __syncwarp();
                    const long long cl0 = DBG_CLK();
                    if (max_lpc > 0 && nsamp > 4)
                        lpc_design<WIDE, HISTN>(sh, max_lpc, nsamp, wasted, a.bps - wasted, a.qlp_precision, magbits, lean, lane);
                    if (lane == 0) DBG_ADD(1, DBG_CLK() - cl0);
                } else {
                    const uint64_t f1 = __shfl_sync(FULLMASK, tot, 1);
                    uint64_t bestv = (lane < 5) ? tot : ~0ull;   // best fixed order among lanes 0..4 (lowest order wins ties)
                    int besto = lane;
                    for (int o = 1; o < 8; o <<= 1) {
                        const uint64_t ov = __shfl_xor_sync(FULLMASK, bestv, o);
                        const int oo = __shfl_xor_sync(FULLMASK, besto, o);
                        if (ov < bestv || (ov == bestv && oo < besto)) { bestv = ov; besto = oo; }
                    }
                    besto = __shfl_sync(FULLMASK, besto, 0);
                    const int fixed_order = (nsamp > 4) ? besto : -1;
                    if (warp == 0 && lane == 0) {
                        sh.wasted = wasted;
                        sh.is_const = (f1 == 0) ? 1 : 0;   // sum |s[g]-s[g-1]| == 0  <=>  all samples equal
                        sh.fixed_order = fixed_order;
                    }
                    if (fixed_order >= 0 && lean) {          // fat data: fat_search redoes the fixed candidate in 64 bits
                        const long long ca0 = DBG_CLK();
                        if (full) {
                            if (warp <= po_max) analyse_level(sh, tabs + fixed_order * FRX_ENC_THREADS, 0, warp, fixed_order, kmax, wasted);
                            if (lane == 0 && warp == 6) DBG_ADD(0, DBG_CLK() - ca0);
                            if (lane == 0 && warp == 0) DBG_ADD(2, DBG_CLK() - ca0);
                        } else if (warp == 0) analyse_partitions(sh, tabs + fixed_order * FRX_ENC_THREADS, 0, nsamp, fixed_order, po_max, kmax, wasted);
                    }
                }
            }
            compute_sync();  // SB
            DBG_PH(11);
