
// ------------------------------------------------------------------------------------------
// K1: raw samples -> PCM in the warp buffer.  Virtual channel vch: 0..7 = band first_band + vch; for two-channel stereo
// trials 8 = mid ((L+R)>>1), 9 = side (L-R).
// ------------------------------------------------------------------------------------------
template <int DT>
__device__ __forceinline__ int32_t quant_one(typename FrxElem<DT>::T v, const FrxQuant &qz, int mode) {
    if (mode == FRX_PCM_REFERENCE) return frx_quant_ref<DT>(v, qz);
    if (mode == FRX_PCM_LOSSLESS) return frx_quant_lossless<DT>(v);
    return frx_quant_spatial<DT>(v);
}

// in-place expansion of one block: raw elements sit end-aligned in the block's first 512 bytes + 16, PCM words grow from
// its start (a 16-byte read is always ahead of the words written so far).  FULL: all 128 samples are valid.
// F32SAFE: float32 reference mode with the stream-wide guarantee that x * 8388607 fits int32 (no per-sample range check).
// INT16: 8/16-bit reference mode of a tile under FRX_QF_INT16 (frx_common.cuh).
template <int DT, bool FULL, bool F32SAFE, bool INT16 = false>
__device__ __forceinline__ void quantize_block_t(uint32_t *blk, const FrxQuant &qz, int mode, int nvalid, uint32_t &orv, uint32_t &ora) {
    typedef typename FrxElem<DT>::T T;
    constexpr int ESZ = (int)sizeof(T);
    constexpr int EPV = 16 / ESZ;                           // elements per 16-byte load
    const unsigned char *raw = reinterpret_cast<const unsigned char *>(blk) + (BLK_W * 4 - 16 - BLK * ESZ);
#pragma unroll 2
    for (int t = 0; t < BLK / EPV; t++) {
        const uint4 v = *reinterpret_cast<const uint4 *>(raw + 16 * t);
        T e[EPV];
        memcpy(e, &v, 16);
        int32_t x[EPV];
        if (INT16) {
            // integer form of the 16-bit quantiser for all the elements of the load, one (rare) branch for those whose
            // quotient is exact and needs the float64 chain
            bool exact = false;
#pragma unroll
            for (int i = 0; i < EPV; i++) x[i] = frx_quant16_int<DT>(e[i], qz, exact);
            if (exact) {
#pragma unroll
                for (int i = 0; i < EPV; i++) x[i] = frx_quant_ref_chain<DT>(e[i], qz);
            }
        }
#pragma unroll
        for (int i = 0; i < EPV; i++) {
            if (INT16) { }
            else if (F32SAFE) x[i] = __float2int_rz(__fmul_rn((float)e[i], 8388607.0f));
            else x[i] = quant_one<DT>(e[i], qz, mode);
            if (!FULL && t * EPV + i >= nvalid) x[i] = 0;
            orv |= (uint32_t)x[i];
            ora |= (uint32_t)(x[i] ^ (x[i] >> 31));
        }
#pragma unroll
        for (int i = 0; i < EPV / 4; i++)
            *reinterpret_cast<uint4 *>(blk + t * EPV + 4 * i) = make_uint4((uint32_t)x[4 * i], (uint32_t)x[4 * i + 1], (uint32_t)x[4 * i + 2], (uint32_t)x[4 * i + 3]);   // same type as the read: ordered
    }
}
template <int DT>
__device__ __forceinline__ void quantize_block(uint32_t *blk, const FrxQuant &qz, int mode, int nvalid, uint32_t &orv, uint32_t &ora) {
    if (nvalid == BLK) {
        if (DT == FRX_F32 && mode == FRX_PCM_REFERENCE && (qz.flags & FRX_QF_F32_SAFE)) quantize_block_t<DT, true, DT == FRX_F32>(blk, qz, mode, nvalid, orv, ora);
        else if ((DT == FRX_U8 || DT == FRX_I16 || DT == FRX_U16) && mode == FRX_PCM_REFERENCE && (qz.flags & FRX_QF_INT16))
            quantize_block_t<DT, true, false, DT == FRX_U8 || DT == FRX_I16 || DT == FRX_U16>(blk, qz, mode, nvalid, orv, ora);
        else quantize_block_t<DT, true, false>(blk, qz, mode, nvalid, orv, ora);
    } else quantize_block_t<DT, false, false>(blk, qz, mode, nvalid, orv, ora);
}
__device__ __noinline__ void quantize_block_any(int dtype, uint32_t *blk, const FrxQuant &qz, int mode, int nvalid, uint32_t &orv, uint32_t &ora) {
    switch (dtype) {   // the dtypes the benchmarks do not use share one out-of-line copy
        case FRX_U8: quantize_block<FRX_U8>(blk, qz, mode, nvalid, orv, ora); break;
        case FRX_I16: quantize_block<FRX_I16>(blk, qz, mode, nvalid, orv, ora); break;
        case FRX_I32: quantize_block<FRX_I32>(blk, qz, mode, nvalid, orv, ora); break;
        default: quantize_block<FRX_U32>(blk, qz, mode, nvalid, orv, ora); break;
    }
}

template <int DT>
__device__ __forceinline__ int32_t gather_one(const EncArgs &a, const FrameDesc &d, int band, long long t) {
    typedef typename FrxElem<DT>::T T;
    const long long row = t / d.t.w, col = t - row * d.t.w;
    const T *p = (const T *)a.raster + (long long)band * a.band_stride + ((long long)d.t.row_off + row) * a.row_stride + d.t.col_off + col;
    return quant_one<DT>(__ldg(p), d.q, a.pcm_mode);
}
__device__ __noinline__ int32_t gather_any(const EncArgs &a, const FrameDesc &d, int band, long long t) {
    switch (a.dtype) {
        case FRX_U8: return gather_one<FRX_U8>(a, d, band, t);
        case FRX_I16: return gather_one<FRX_I16>(a, d, band, t);
        case FRX_U16: return gather_one<FRX_U16>(a, d, band, t);
        case FRX_I32: return gather_one<FRX_I32>(a, d, band, t);
        case FRX_U32: return gather_one<FRX_U32>(a, d, band, t);
        default: return gather_one<FRX_F32>(a, d, band, t);
    }
}

// Mid / side of a two-channel frame, aligned rows: every lane reads the 128 samples of its block from both band planes with
// 16-byte loads, quantises both and stores mid or side as PCM (no raw staging: the buffer holds one channel at a time)
template <int DT>
__device__ __forceinline__ void load_ms_block(const EncArgs &a, const FrameDesc &d, uint32_t *blk, int lane, int vch, int nvalid, uint32_t &orv, uint32_t &ora) {
    typedef typename FrxElem<DT>::T T;
    constexpr int EPV = 16 / (int)sizeof(T);
    const long long w = d.t.w;
    const long long t0 = d.s0 + (long long)lane * BLK;
    long long row = t0 / w, col = t0 - row * w;
    const T *L = (const T *)a.raster + (long long)d.t.first_band * a.band_stride + (long long)d.t.row_off * a.row_stride + d.t.col_off;
    const T *R = L + a.band_stride;
#pragma unroll 1
    for (int i = 0; i < BLK; i += EPV) {
        int32_t x[EPV];
        if (i < nvalid) {
            const long long e = row * a.row_stride + col;
            const uint4 vl = __ldg(reinterpret_cast<const uint4 *>(L + e)), vr = __ldg(reinterpret_cast<const uint4 *>(R + e));
            T el[EPV], er[EPV];
            memcpy(el, &vl, 16); memcpy(er, &vr, 16);
            int32_t ql[EPV], qr[EPV];
            if ((DT == FRX_U8 || DT == FRX_I16 || DT == FRX_U16) && a.pcm_mode == FRX_PCM_REFERENCE && (d.q.flags & FRX_QF_INT16)) {
                bool exact = false;
#pragma unroll
                for (int j = 0; j < EPV; j++) { ql[j] = frx_quant16_int<DT>(el[j], d.q, exact); qr[j] = frx_quant16_int<DT>(er[j], d.q, exact); }
                if (exact) {
#pragma unroll
                    for (int j = 0; j < EPV; j++) { ql[j] = frx_quant_ref_chain<DT>(el[j], d.q); qr[j] = frx_quant_ref_chain<DT>(er[j], d.q); }
                }
            } else {
#pragma unroll
                for (int j = 0; j < EPV; j++) { ql[j] = quant_one<DT>(el[j], d.q, a.pcm_mode); qr[j] = quant_one<DT>(er[j], d.q, a.pcm_mode); }
            }
#pragma unroll
            for (int j = 0; j < EPV; j++) {
                const int32_t l = ql[j], r = qr[j];
                x[j] = vch == 8 ? ((l + r) >> 1) : (l - r);
                if (i + j >= nvalid) x[j] = 0;
                orv |= (uint32_t)x[j];
                ora |= (uint32_t)(x[j] ^ (x[j] >> 31));
            }
            col += EPV;
            if (col >= w) { col = 0; row++; }
        } else {
#pragma unroll
            for (int j = 0; j < EPV; j++) x[j] = 0;
        }
#pragma unroll
        for (int j = 0; j < EPV / 4; j++)
            *reinterpret_cast<uint4 *>(blk + i + 4 * j) = make_uint4((uint32_t)x[4 * j], (uint32_t)x[4 * j + 1], (uint32_t)x[4 * j + 2], (uint32_t)x[4 * j + 3]);
    }
}
__device__ __noinline__ void load_ms_block_any(const EncArgs &a, const FrameDesc &d, uint32_t *blk, int lane, int vch, int nvalid, uint32_t &orv, uint32_t &ora) {
    switch (a.dtype) {
        case FRX_U8: load_ms_block<FRX_U8>(a, d, blk, lane, vch, nvalid, orv, ora); break;
        case FRX_I16: load_ms_block<FRX_I16>(a, d, blk, lane, vch, nvalid, orv, ora); break;
        case FRX_U16: load_ms_block<FRX_U16>(a, d, blk, lane, vch, nvalid, orv, ora); break;
        case FRX_I32: load_ms_block<FRX_I32>(a, d, blk, lane, vch, nvalid, orv, ora); break;
        case FRX_U32: load_ms_block<FRX_U32>(a, d, blk, lane, vch, nvalid, orv, ora); break;
        default: load_ms_block<FRX_F32>(a, d, blk, lane, vch, nvalid, orv, ora); break;
    }
}

// Stage virtual channel `vch` of the frame as PCM in buf[SLACK_W + PIDX(g)].  Returns (OR of the samples, OR of their
// magnitudes), warp-uniform.  The samples behind a short frame are zero up to the next block boundary + one block.
__device__ __noinline__ uint2 load_channel(const EncArgs &a, WarpSh &ws, uint32_t *buf, int vch, uint32_t &phase, int lane) {
    const FrameDesc &d = ws.fd;
    const int nsamp = d.nsamp;
    uint32_t *pcm = buf + SLACK_W;
    uint32_t orv = 0, ora = 0;
    const int nblk = (nsamp + BLK - 1) / BLK;
    if (d.tma && vch < 8) {
        const int esz = frx_dtype_size(a.dtype);
        const long long w = d.t.w;
        const unsigned char *band = (const unsigned char *)a.raster +
            ((long long)(d.t.first_band + vch) * a.band_stride + (long long)d.t.row_off * a.row_stride + d.t.col_off) * esz;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // earlier generic accesses to the buffer before the async writes
        __syncwarp();
        if (lane == 0) mbar_arrive_expect_tx(&ws.mbar, (uint32_t)nsamp * (uint32_t)esz);
        __syncwarp();
        if (lane < nblk) {
            const long long g0 = (long long)lane * BLK, g1 = min((long long)nsamp, g0 + BLK);
            long long t0 = d.s0 + g0;
            const long long t1 = d.s0 + g1;
            long long row = t0 / w;
            unsigned char *dst0 = reinterpret_cast<unsigned char *>(pcm + lane * BLK_W) + (BLK_W * 4 - 16 - BLK * esz);
            while (t0 < t1) {
                const long long rend = min(t1, (row + 1) * w);
                bulk_g2s(dst0 + (t0 - d.s0 - g0) * esz, band + (row * a.row_stride + (t0 - row * w)) * esz, (uint32_t)((rend - t0) * esz), &ws.mbar);
                t0 = rend; row++;
            }
        }
        mbar_wait(&ws.mbar, phase);
        phase ^= 1u;
        if (lane < nblk) {
            const int nvalid = min(BLK, nsamp - lane * BLK);
            uint32_t *blk = pcm + lane * BLK_W;
            if (a.dtype == FRX_F32) quantize_block<FRX_F32>(blk, d.q, a.pcm_mode, nvalid, orv, ora);
            else if (a.dtype == FRX_U16) quantize_block<FRX_U16>(blk, d.q, a.pcm_mode, nvalid, orv, ora);
            else quantize_block_any(a.dtype, blk, d.q, a.pcm_mode, nvalid, orv, ora);
        } else if (lane == nblk) {
            uint32_t *blk = pcm + lane * BLK_W;
            for (int i = 0; i < BLK; i += 4) *reinterpret_cast<int4 *>(blk + i) = make_int4(0, 0, 0, 0);
        }
    } else if (d.tma) {
        // mid / side (vch >= 8) of aligned rows
        uint32_t o1 = 0, o2 = 0;
        if (lane <= nblk && lane < 32) load_ms_block_any(a, d, pcm + lane * BLK_W, lane, vch, max(0, min(BLK, nsamp - lane * BLK)), o1, o2);
        orv = o1; ora = o2;
    } else {
        const int b0 = d.t.first_band;
        for (int g = lane; g < nsamp; g += 32) {
            int32_t x;
            if (vch < 8) x = gather_any(a, d, b0 + vch, d.s0 + g);
            else {
                const int32_t l = gather_any(a, d, b0, d.s0 + g), r = gather_any(a, d, b0 + 1, d.s0 + g);
                x = vch == 8 ? ((l + r) >> 1) : (l - r);
            }
            pcm[PIDX(g)] = (uint32_t)x;
            orv |= (uint32_t)x;
            ora |= (uint32_t)(x ^ (x >> 31));
        }
        const int gend = min(FRX_MAX_BLOCK, nblk * BLK + BLK);
        for (int g = nsamp + lane; g < gend; g += 32) pcm[PIDX(g)] = 0u;
    }
    orv = __reduce_or_sync(FULLMASK, orv);
    ora = __reduce_or_sync(FULLMASK, ora);
    __syncwarp();
    return make_uint2(orv, ora);
}

// ------------------------------------------------------------------------------------------
// Rice search of one candidate from its partition sums.  F = the level whose partitions the lanes hold one each (5 for
// full frames): lane p holds the sum of level-F partition p in sumA + sumB; for po_max == 6 sumA / sumB are the two
// level-6 partitions 2p, 2p+1.  Coarser levels are built with shuffles; levels above po_max are not evaluated.  Writes the
// best order's parameters to kparam[] and returns (estimated bits incl. the 6-bit method/order field, best order, RICE2).
// ------------------------------------------------------------------------------------------
struct RiceChoice { uint64_t bits; int po; bool rice2; };
// sum of per-lane bit estimates: one redux when they are small (always, unless a partition is hopeless), exact otherwise
__device__ __forceinline__ uint64_t warp_sum_bits(uint64_t b) {
    if (__all_sync(FULLMASK, b < (1ull << 26))) return (uint64_t)__reduce_add_sync(FULLMASK, (uint32_t)b);
    return warp_sum_u64(b);
}
__device__ __noinline__ RiceChoice analyse_partitions(uint8_t *kparam, uint64_t sumA, uint64_t sumB, int nsamp, int pred_order, int po_max,
                                                      int F, uint32_t kmax) {
    const int lane = threadIdx.x & 31;
    uint64_t sum = sumA + sumB;
    uint64_t best = ~0ull;
    int best_po = 0;
    uint32_t kbest0 = 0, kbest1 = 0;
    if (po_max == 6) {
        const uint32_t n0 = (uint32_t)(nsamp >> 6) - (lane == 0 ? (uint32_t)pred_order : 0u), n1 = (uint32_t)(nsamp >> 6);
        const uint32_t k0 = rice_param(n0, sumA, kmax), k1 = rice_param(n1, sumB, kmax);
        uint64_t b = rice_estimate(k0, n0, sumA) + rice_estimate(k1, n1, sumB);
        b = warp_sum_bits(b) + 6;
        best = b; best_po = 6; kbest0 = k0; kbest1 = k1;
    }
    for (int L = F; L >= 0; L--) {
        const int step = 1 << (F - L);                   // lanes that are multiples of step hold level-L partitions
        const bool mine = (lane & (step - 1)) == 0 && lane < (1 << F);
        const int p = lane >> (F - L);
        const uint32_t n = (uint32_t)(nsamp >> L) - (p == 0 ? (uint32_t)pred_order : 0u);
        uint32_t k = 0;
        uint64_t b = 0;
        if (L <= po_max) {
            if (mine) { k = rice_param(n, sum, kmax); b = rice_estimate(k, n, sum); }
            b = warp_sum_bits(b) + 6;
            if (b < best) { best = b; best_po = L; kbest0 = k; }
        }
        if (L > 0) {
            const uint64_t o = __shfl_down_sync(FULLMASK, sum, step);
            sum += o;                                    // valid in lanes that are multiples of 2*step
        }
    }
    if (best_po == 6) { kparam[2 * lane] = (uint8_t)kbest0; kparam[2 * lane + 1] = (uint8_t)kbest1; }
    else {
        const int step = 1 << (F - best_po);
        if ((lane & (step - 1)) == 0 && lane < (1 << F)) kparam[lane >> (F - best_po)] = (uint8_t)kbest0;
    }
    const uint32_t big = __ballot_sync(FULLMASK, kbest0 >= 15u || (best_po == 6 && kbest1 >= 15u));
    __syncwarp();
    RiceChoice rc;
    rc.bits = best; rc.po = best_po; rc.rice2 = big != 0u;
    return rc;
}
