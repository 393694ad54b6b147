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
    // publish parameters of the best level
    if (best_po == 6) { sh.kparam[cand][2 * lane] = kbest0; sh.kparam[cand][2 * lane + 1] = kbest1; }
    else {
        const int step = 1 << (F - best_po);
        if ((lane & (step - 1)) == 0 && lane < (1 << F)) sh.kparam[cand][lane >> (F - best_po)] = kbest0;
    }
    uint32_t big = __ballot_sync(FULLMASK, kbest0 >= 15u || (best_po == 6 && kbest1 >= 15u));
    if (lane == 0) { sh.cand_po[cand] = best_po; sh.cand_bits[cand] = best; sh.cand_rice2[cand] = big ? 1 : 0; }
    __syncwarp();
}

