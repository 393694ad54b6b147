// One warp analyses one candidate: partition sums from the per-thread |r| sums, Rice parameter per partition,
// estimated bits per partition order, best order.  Lane p owns finest-level partition p (2p, 2p+1 when po_max == 6);
// coarser levels are built with shuffles, so partition q of level L lives in lane q << (F - L).
__device__ __noinline__ void analyse_partitions(EncSmall &sh, const uint64_t *tabs, int cand, int nsamp, int pred_order, int po_max, uint32_t kmax, int rsh) {
    const int lane = threadIdx.x & 31;
    const int F = po_max > 5 ? 5 : po_max;               // finest level held one-per-lane
    const int nthr = (nsamp + FRX_SPT - 1) / FRX_SPT;
    uint64_t sum = 0, sumB = 0;                          // sumB: second half when po_max == 6
    uint64_t best = ~0ull;
    int best_po = 0;
    uint32_t kbest0 = 0, kbest1 = 0;
    if (po_max == 6) {
        const int tpp = (nsamp >> 6) / FRX_SPT;          // threads per level-6 partition
        for (int t = 0; t < tpp; t++) { sum += tabs[(2 * lane) * tpp + t] >> rsh; sumB += tabs[(2 * lane + 1) * tpp + t] >> rsh; }
        uint32_t n0 = (uint32_t)(nsamp >> 6) - (lane == 0 ? (uint32_t)pred_order : 0u), n1 = (uint32_t)(nsamp >> 6);
        uint32_t k0 = rice_param(n0, sum, kmax), k1 = rice_param(n1, sumB, kmax);
        uint64_t b = rice_estimate(k0, n0, sum) + rice_estimate(k1, n1, sumB);
        b = warp_sum_u64(b) + 6;
        best = b; best_po = 6; kbest0 = k0; kbest1 = k1;
        sum += sumB;
    } else if (F == 0) {
        for (int t = lane; t < nthr; t += 32) sum += tabs[t] >> rsh;
        sum = warp_sum_u64(sum);
    } else {
        const int tpp = (nsamp >> F) / FRX_SPT;          // threads per finest partition (>= 1)
        if (lane < (1 << F))
            for (int t = 0; t < tpp; t++) sum += tabs[lane * tpp + t] >> rsh;
    }
    // levels F .. 0
    for (int L = F; L >= 0; L--) {
        const int step = 1 << (F - L);                   // lanes that are multiples of step hold level-L partitions
        const bool mine = (lane & (step - 1)) == 0 && lane < (1 << F);
        const int p = lane >> (F - L);
        uint32_t n = (uint32_t)(nsamp >> L) - (p == 0 ? (uint32_t)pred_order : 0u);
        uint32_t k = 0;
        uint64_t b = 0;
        if (mine) { k = rice_param(n, sum, kmax); b = rice_estimate(k, n, sum); }
        b = warp_sum_u64(b) + 6;
        if (b < best) { best = b; best_po = L; kbest0 = k; }
        if (L > 0) {
            uint64_t o = __shfl_down_sync(FULLMASK, sum, step);
            sum += o;                                    // valid in lanes that are multiples of 2*step
        }
    }
    // publish in the per-level format pick_level() reads: only the winning level carries a finite cost
    uint32_t *kp = &sh.kparam[cand][(1 << best_po) - 1];
    if (best_po == 6) { kp[2 * lane] = kbest0; kp[2 * lane + 1] = kbest1; }
    else {
        const int step = 1 << (F - best_po);
        if ((lane & (step - 1)) == 0 && lane < (1 << F)) kp[lane >> (F - best_po)] = kbest0;
    }
    uint32_t big = __ballot_sync(FULLMASK, kbest0 >= 15u || (best_po == 6 && kbest1 >= 15u));
    if (lane < 8) { sh.lvl_bits[cand][lane] = (lane == best_po) ? best : ~0ull; sh.lvl_big[cand][lane] = big ? 1u : 0u; }
    __syncwarp();
}

// The same search for ONE partition order of a full 4096-sample frame (2^L partitions of 256 >> L threads each): the compute
// warps run it side by side, warp L on level L.  The per-thread |r| sums must fit 32 bits (the lean paths guarantee it).
__device__ __noinline__ void analyse_level(EncSmall &sh, const uint64_t *tabs, int cand, int L, int pred_order, uint32_t kmax, int rsh) {
    const int lane = threadIdx.x & 31;
    const uint32_t nfull = (uint32_t)FRX_MAX_BLOCK >> L;
    uint32_t *kp = &sh.kparam[cand][(1 << L) - 1];
    uint64_t bits = 0;
    uint32_t big = 0;
    if (L <= 3) {
        // a row of 32 threads never straddles a partition: reduce the eight rows (lane t keeps row t), then groups of rows
        uint64_t mine = 0;
#pragma unroll
        for (int t = 0; t < 8; t++) {
            const uint64_t s = warp_sum_u32_wide((uint32_t)(tabs[32 * t + lane] >> rsh));
            if (lane == t) mine = s;
        }
        const int g = 8 >> L;
        for (int o = 1; o < g; o <<= 1) mine += __shfl_xor_sync(FULLMASK, mine, o);
        if (lane < 8 && (lane & (g - 1)) == 0) {
            const int p = lane >> (3 - L);
            const uint32_t n = nfull - (p == 0 ? (uint32_t)pred_order : 0u);
            const uint32_t k = rice_param(n, mine, kmax);
            bits = rice_estimate(k, n, mine);
            kp[p] = k;
            big = k >= 15u ? 1u : 0u;
        }
    } else {
        const int T = FRX_ENC_THREADS >> L;            // threads per partition: 16, 8, 4
        for (int p = lane; p < (1 << L); p += 32) {
            uint64_t s = 0;
            for (int i = 0; i < T; i++) s += tabs[p * T + i] >> rsh;
            const uint32_t n = nfull - (p == 0 ? (uint32_t)pred_order : 0u);
            const uint32_t k = rice_param(n, s, kmax);
            bits += rice_estimate(k, n, s);
            kp[p] = k;
            big |= k >= 15u ? 1u : 0u;
        }
    }
    bits = warp_sum_u64(bits) + 6;
    const uint32_t anybig = __ballot_sync(FULLMASK, big != 0u);
    if (lane == 0) { sh.lvl_bits[cand][L] = bits; sh.lvl_big[cand][L] = anybig ? 1u : 0u; }
}

// best partition order of a candidate: highest order first, a lower one only when strictly cheaper (libFLAC's loop).
// Called by whole warps: lane L looks at order L, one redux picks the minimum, the highest lane holding it wins.
__device__ __forceinline__ void pick_level(const EncSmall &sh, int cand, int po_max, int &po, uint64_t &bits, bool &rice2) {
    const int lane = threadIdx.x & 31;
    uint32_t key = 0xFFFFFFFFu, big = 0;
    if (lane <= po_max) {
        const uint64_t b = sh.lvl_bits[cand][lane];
        key = b < 0xFFFFFFFEull ? (uint32_t)b : 0xFFFFFFFEu;
        big = sh.lvl_big[cand][lane];
    }
    const uint32_t m = __reduce_min_sync(FULLMASK, key);
    const int bp = 31 - __clz((int)(__ballot_sync(FULLMASK, key == m) & 0xFFu));
    po = bp; bits = m; rice2 = __shfl_sync(FULLMASK, big, bp) != 0u;
}
