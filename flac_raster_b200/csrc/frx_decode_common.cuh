// frx_decode_common.cuh -- helpers shared by the decode translation units.
#pragma once
#include "frx_internal.cuh"

#define FRX_NOPOS 0xFFFFFFFFFFFFFFFFull

__device__ __forceinline__ int frx_size_code_bps(uint32_t c) {
    switch (c) { case 1: return 8; case 2: return 12; case 4: return 16; case 5: return 20; case 6: return 24; case 7: return 32; case 0: return 0; }
    return -1;
}

__device__ __forceinline__ int dec_find_tile(const uint32_t *prefix, int n, uint32_t f) {
    int lo = 0, hi = n;
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (prefix[mid] <= f) lo = mid; else hi = mid;
    }
    return lo;
}

// ---- K9a: one warp per tile, chunked shuffle scan over the tile's frame sizes ----
