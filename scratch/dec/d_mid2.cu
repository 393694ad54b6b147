// ------------------------------------------------------------------------------------------
// K10-K12.  One lane owns one frame (Rice decoding and the predictor recurrence are serial inside a subframe; the
// parallelism is across frames), but a warp walks its 32 frames in lockstep, 32 samples at a time: the restored samples
// go through a shared-memory transpose so that every global store is one full line of one frame's raster row, instead of
// 32 lanes scattering single elements 16 KB apart.
// ------------------------------------------------------------------------------------------
#define DEC_WARPS 1     // one warp per CTA: a frame takes milliseconds on its lane, so the grid must fit one wave; small CTAs pack the SMs evenly
#define DEC_MAXORD 16
struct DecWarpSmem {
    uint32_t ring[DEC_RING][32];     // input rings, one column per lane
    uint32_t stage[32][33];          // [frame lane][sample in chunk], padded
    unsigned long long ptr[32];      // byte address of the first sample of the lane's chunk
    long long rstride[32];           // row stride of the lane's tile, elements
    uint32_t room[32];               // elements left in the raster row at the chunk start
    uint32_t cnt[32];                // valid samples of the lane in this chunk
    uint32_t wid[32];                // tile width
    // stereo decorrelation only (frames with channel assignment 8..10)
    uint32_t mode[32];               // 0 plain, 1 first channel -> raw PCM into the scratch row, 2/3/4 second channel of L/S, S/R, M/S
    unsigned long long ptr2[32];     // byte address of the chunk in the frame's scratch row
    long long bstride[32];           // bytes between the two band planes of the output
    float dq_range[32], dq_min[32];
    int32_t bpsv[32];
};

struct SubF {
    int type;          // 0 constant, 1 verbatim, 2 predictive, 3 nothing to do
    int order, shift, wasted, sbps;
    int32_t cval;
    RiceState rs;
};

// generic order (17..32): history ring in local memory, scattered stores (libFLAC never emits these below -l 17)
template <int DT>
__device__ __noinline__ void restore_subframe_generic(BitR &br, RiceState &rs, typename FrxElem<DT>::T *base, long long row_stride, int w,
                                                      long long px, const int32_t *coef, int order, int shift, const int32_t *warm, int nsamp,
                                                      int wasted, int bps, const FrxDequant &dq, int mode) {
    typedef typename FrxElem<DT>::T T;
    int row = (int)(px / w), col = (int)(px - (long long)row * w);
    int32_t h[32];
    for (int i = 0; i < 32; i++) h[i] = i < order ? warm[i] : 0;
    for (int n = 0; n < nsamp; n++) {
        int32_t v;
        if (n < order) v = warm[n];
        else {
            const int32_t r = rice_next(br, rs);
            long long acc = 0;
            for (int j = 0; j < order; j++) acc += (long long)coef[j] * (long long)h[(n - 1 - j) & 31];
            v = (int32_t)((long long)r + (acc >> shift));
            h[n & 31] = v;
        }
        base[(long long)row * row_stride + col] = frx_dequant<DT>((int32_t)((uint32_t)v << wasted), bps, dq, mode);
        if (++col == w) { col = 0; row++; }
    }
}

template <int DT>
__device__ __forceinline__ uint32_t dec_pack(typename FrxElem<DT>::T v) {
    typedef typename FrxElem<DT>::T T;
    if (sizeof(T) == 4) { uint32_t u; memcpy(&u, &v, 4); return u; }
    if (sizeof(T) == 2) { uint16_t u; memcpy(&u, &v, 2); return u; }
    uint8_t u; memcpy(&u, &v, 1); return u;
}
template <int DT>
__device__ __forceinline__ void dec_store(unsigned long long addr, uint32_t bits) {
    typedef typename FrxElem<DT>::T T;
    if (sizeof(T) == 4) *reinterpret_cast<uint32_t *>(addr) = bits;
    else if (sizeof(T) == 2) *reinterpret_cast<uint16_t *>(addr) = (uint16_t)bits;
    else *reinterpret_cast<uint8_t *>(addr) = (uint8_t)bits;
}

// the warp writes the 32x32 staged samples: iteration j stores frame j's chunk, lane = sample
template <int DT>
__device__ __forceinline__ void flush_chunk(DecWarpSmem &ws, int lane) {
    typedef typename FrxElem<DT>::T T;
    __syncwarp();
#pragma unroll 4
    for (int j = 0; j < 32; j++) {
        const uint32_t cnt = ws.cnt[j];
        if ((uint32_t)lane < cnt) {
            const uint32_t room = ws.room[j];
            unsigned long long addr = ws.ptr[j];
            if ((uint32_t)lane < room) addr += (unsigned long long)lane * sizeof(T);
            else {
                // the chunk runs over the end of the raster row (tile widths that are not multiples of 32)
                const uint32_t w = ws.wid[j];
                const uint32_t k = (uint32_t)lane - room;
                const uint32_t down = 1u + k / w, colx = k - (k / w) * w;
                addr = addr - (unsigned long long)(w - room) * sizeof(T) + ((unsigned long long)((long long)down * ws.rstride[j]) + colx) * sizeof(T);
            }
            dec_store<DT>(addr, ws.stage[j][lane]);
        }
    }
    __syncwarp();
}

// write-out for warps that hold frames with stereo decorrelation (libFLAC tries left/side, side/right and mid/side for
// exactly two channels; RFC 9639 section 4.2): the first channel's PCM is parked in a scratch row, the second channel's
// chunk is combined with it here, lane = sample, so both band planes are still written as full lines.
template <int DT>
__device__ __forceinline__ void flush_chunk_stereo(DecWarpSmem &ws, int lane, int pcm_mode) {
    typedef typename FrxElem<DT>::T T;
    __syncwarp();
    for (int j = 0; j < 32; j++) {
        const uint32_t cnt = ws.cnt[j];
        if ((uint32_t)lane >= cnt) continue;
        const uint32_t mode = ws.mode[j];
        const uint32_t bits = ws.stage[j][lane];
        if (mode == 1u) { *reinterpret_cast<uint32_t *>(ws.ptr2[j] + 4ull * (unsigned)lane) = bits; continue; }
        const uint32_t room = ws.room[j];
        unsigned long long addr = ws.ptr[j];
        if ((uint32_t)lane < room) addr += (unsigned long long)lane * sizeof(T);
        else {
            const uint32_t w = ws.wid[j];
            const uint32_t k = (uint32_t)lane - room;
            const uint32_t down = 1u + k / w, colx = k - (k / w) * w;
            addr = addr - (unsigned long long)(w - room) * sizeof(T) + ((unsigned long long)((long long)down * ws.rstride[j]) + colx) * sizeof(T);
        }
        if (mode == 0u) { dec_store<DT>(addr, bits); continue; }
        const int32_t s1 = (int32_t)bits;
        const int32_t s0 = *reinterpret_cast<const int32_t *>(ws.ptr2[j] + 4ull * (unsigned)lane);
        int32_t L, R;
        if (mode == 2u) { L = s0; R = s0 - s1; }                 // left, side
        else if (mode == 3u) { R = s1; L = s0 + s1; }            // side, right
        else { const int32_t mid = (int32_t)(((uint32_t)s0 << 1) | ((uint32_t)s1 & 1u)); L = (mid + s1) >> 1; R = (mid - s1) >> 1; }
        FrxDequant dq;
        dq.range_f = ws.dq_range[j]; dq.dmin_f = ws.dq_min[j];
        const int bps = ws.bpsv[j];
        dec_store<DT>(addr - (unsigned long long)ws.bstride[j], dec_pack<DT>(frx_dequant<DT>(L, bps, dq, pcm_mode)));
        dec_store<DT>(addr, dec_pack<DT>(frx_dequant<DT>(R, bps, dq, pcm_mode)));
    }
    __syncwarp();
}

// lockstep restore of one channel for the 32 frames of a warp; ORD = history length (multiple of 4 dividing 32),
// WIDE = 64-bit accumulate.  Uniform control flow: lanes whose subframe is not predictive ride along predicated.
template <int DT, int ORD, bool WIDE, bool ST>
__device__ __forceinline__ void run_chunks(BitR &br, SubF &sf, int32_t *coef, int32_t *h, DecWarpSmem &ws, int lane, int my_n, int nmax,
                                           unsigned long long base_addr, long long row_stride, int w, int row, int col, int bps,
                                           const FrxDequant &dq, int mode, uint32_t st_mode = 0u, unsigned long long scratch_addr = 0ull,
                                           long long band_stride_bytes = 0) {
    typedef typename FrxElem<DT>::T T;
    for (int n0 = 0; n0 < nmax; n0 += 32) {
        br.topup(n0 < my_n);
        // where this lane's chunk lands
        ws.ptr[lane] = base_addr + (unsigned long long)((long long)row * row_stride + col) * sizeof(T);
        ws.rstride[lane] = row_stride;
        ws.room[lane] = (uint32_t)(w - col);
        ws.wid[lane] = (uint32_t)w;
        ws.cnt[lane] = (uint32_t)min(32, max(0, my_n - n0));
        if (ST) {
            ws.mode[lane] = st_mode; ws.ptr2[lane] = scratch_addr + 4ull * (unsigned)n0; ws.bstride[lane] = band_stride_bytes;
            ws.dq_range[lane] = dq.range_f; ws.dq_min[lane] = dq.dmin_f; ws.bpsv[lane] = bps;
        }
        // a chunk past the warm-up in which every lane restores 32 predicted samples needs no per-sample predicates
        const bool steady = !ST && __all_sync(0xFFFFFFFFu, sf.type == 2 && n0 >= 32 && n0 + 32 <= my_n);
        if (steady) {
            for (int mb = 0; mb < 32; mb += ORD) {
#pragma unroll
                for (int mi = 0; mi < ORD; mi++) {
                    const int32_t r = rice_next(br, sf.rs);
                    int32_t v;
                    if (WIDE) {
                        long long acc = 0;
#pragma unroll
                        for (int j = 0; j < ORD; j++) acc += (long long)coef[j] * (long long)h[(mi - 1 - j + 2 * ORD) % ORD];
                        v = (int32_t)((long long)r + (acc >> sf.shift));
                    } else {
                        int32_t acc = 0;
#pragma unroll
                        for (int j = 0; j < ORD; j++) acc += coef[j] * h[(mi - 1 - j + 2 * ORD) % ORD];
                        v = r + (acc >> sf.shift);
                    }
                    h[mi] = v;
                    ws.stage[lane][mb + mi] = dec_pack<DT>(frx_dequant<DT>((int32_t)((uint32_t)v << sf.wasted), bps, dq, mode));
                }
            }
        } else {
        for (int mb = 0; mb < 32; mb += ORD) {
#pragma unroll
                for (int mi = 0; mi < ORD; mi++) {
                    const int n = n0 + mb + mi;
                    int32_t v = sf.cval;
                    if (n < my_n) {
                        if (sf.type == 2) {
                            if (n >= sf.order) {
                                const int32_t r = rice_next(br, sf.rs);
                                if (WIDE) {
                                    long long acc = 0;
#pragma unroll
                                    for (int j = 0; j < ORD; j++) acc += (long long)coef[j] * (long long)h[(mi - 1 - j + 2 * ORD) % ORD];
                                    v = (int32_t)((long long)r + (acc >> sf.shift));
                                } else {
                                    int32_t acc = 0;
#pragma unroll
                                    for (int j = 0; j < ORD; j++) acc += coef[j] * h[(mi - 1 - j + 2 * ORD) % ORD];
                                    v = r + (acc >> sf.shift);
                                }
                                h[mi] = v;
                            } else v = h[mi];   // warm-up sample (n < order <= ORD, so n == mi)
                        } else if (sf.type == 1) v = br.sget(sf.sbps);
                    }
                    {
                        const uint32_t rawv = (uint32_t)v << sf.wasted;
                        const uint32_t outv = dec_pack<DT>(frx_dequant<DT>((int32_t)rawv, bps, dq, mode));
                        ws.stage[lane][mb + mi] = (ST && st_mode) ? rawv : outv;
                    }
                }
            }
        }
        if (ST) flush_chunk_stereo<DT>(ws, lane, mode); else flush_chunk<DT>(ws, lane);
        col += 32;
        while (col >= w) { col -= w; row++; }
    }
}

