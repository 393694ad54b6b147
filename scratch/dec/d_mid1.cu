// ------------------------------------------------------------------------------------------
// bit reader.  A lane walks its own frame, but a warp runs its 32 lanes in lockstep: if every lane fetched from global
// memory whenever its own window ran dry, the warp would stall on 32 unrelated misses per few samples.  So each lane
// owns a ring of DEC_RING words in shared memory (column `lane` of ring[word][lane], conflict-free), topped up for all
// lanes at once at every chunk boundary with aligned 16-byte loads; the 64-bit window is fed from the ring.
// ------------------------------------------------------------------------------------------
#define DEC_RING 32
__device__ __noinline__ uint4 dec_load16_clamped(const uint8_t *np, const uint8_t *lo, const uint8_t *hi) {
    uint32_t w[4];
    for (int k = 0; k < 4; k++) {
        w[k] = 0;
        for (int i = 0; i < 4; i++) {
            const uint8_t *p = np + 4 * k + i;
            const uint32_t b = (p >= lo && p < hi) ? (uint32_t)__ldg(p) : 0u;
            w[k] |= b << (8 * i);
        }
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}
struct BitR {
    const uint8_t *lo, *hi;     // valid byte range of the buffer
    const uint8_t *np;          // next 16-byte aligned block to fetch
    uint32_t *ring;             // &ring[0][lane]; word i of the stream sits at ring[(i % DEC_RING) * 32]
    uint32_t rd, wr;            // words consumed / fetched
    uint64_t acc;               // left aligned
    int cnt;                    // valid bits in acc
    uint32_t start_bits;        // stream position of the frame's first bit, in bits from ring word 0
    __device__ __forceinline__ uint4 load16() const {
        if (np >= lo && np + 16 <= hi) return __ldg(reinterpret_cast<const uint4 *>(np));
        return dec_load16_clamped(np, lo, hi);
    }
    __device__ __forceinline__ void push16(const uint4 v) {
        ring[((wr + 0u) % DEC_RING) * 32] = frx_bswap32(v.x);
        ring[((wr + 1u) % DEC_RING) * 32] = frx_bswap32(v.y);
        ring[((wr + 2u) % DEC_RING) * 32] = frx_bswap32(v.z);
        ring[((wr + 3u) % DEC_RING) * 32] = frx_bswap32(v.w);
        wr += 4; np += 16;
    }
    // chunk boundary: every lane fills its ring (uniform trip count so the loads of a round issue together)
    __device__ __forceinline__ void topup(bool want) {
        for (;;) {
            const uint32_t free_blocks = want ? (DEC_RING - (wr - rd)) >> 2 : 0u;
            if (!__any_sync(0xFFFFFFFFu, free_blocks != 0u)) break;
            uint4 v[4];
#pragma unroll
            for (int k = 0; k < 4; k++) if ((uint32_t)k < free_blocks) { const uint8_t *p = np; np += 16 * k; v[k] = load16(); np = p; }
#pragma unroll
            for (int k = 0; k < 4; k++) if ((uint32_t)k < free_blocks) push16(v[k]);
            if (!__any_sync(0xFFFFFFFFu, free_blocks > 4u)) break;
        }
    }
    __device__ __forceinline__ uint32_t pop() {
        if (rd == wr) push16(load16());      // ring ran dry inside a chunk (more than ~4 bytes per sample): fetch on demand
        const uint32_t w = ring[(rd % DEC_RING) * 32];
        rd++;
        return w;
    }
    __device__ __forceinline__ void init(const uint8_t *base, uint64_t n_bytes, uint64_t pos, uint32_t *ring_col, bool live) {
        lo = base; hi = base + n_bytes;
        const uintptr_t addr = (uintptr_t)(base + pos);
        const int lead = (int)(addr & 15u);
        np = (const uint8_t *)(addr - lead);
        ring = ring_col; rd = 0; wr = 0; start_bits = (uint32_t)lead * 8u; acc = 0; cnt = 0;
        topup(live);
        if (live) {
            rd = (uint32_t)(lead >> 2);
            const int lb = 8 * (lead & 3);
            acc = (uint64_t)pop() << (32 + lb);
            cnt = 32 - lb;
            refill();
        }
    }
    // bits consumed since init: everything popped, minus what still sits in the window
    __device__ __forceinline__ uint32_t consumed() const { return rd * 32u - (uint32_t)cnt - start_bits; }
    __device__ __forceinline__ void refill() {
        if (cnt <= 32) {
            acc |= (uint64_t)pop() << (32 - cnt);
            cnt += 32;
        }
    }
    // n in 0..32
    __device__ __forceinline__ uint32_t get(int n) {
        refill();
        const uint32_t v = n ? (uint32_t)(acc >> (64 - n)) : 0u;
        acc <<= n;
        cnt -= n;
        return v;
    }
    __device__ __forceinline__ int32_t sget(int n) {  // signed, n in 1..32
        const uint32_t v = get(n);
        return n >= 32 ? (int32_t)v : ((int32_t)(v << (32 - n)) >> (32 - n));
    }
    __device__ __forceinline__ uint32_t unary() {
        uint32_t z = 0;
        for (;;) {
            refill();
            if (acc == 0) {
                z += (uint32_t)cnt; cnt = 0;
                if (np > hi + 256) return z;  // ran off the buffer: malformed
                continue;
            }
            const int lz = __clzll((long long)acc);
            z += (uint32_t)lz;
            acc = lz == 63 ? 0ull : (acc << (lz + 1));  // a set bit exists within cnt bits, so lz < cnt
            cnt -= lz + 1;
            return z;
        }
    }
};

struct RiceState {
    uint32_t k;        // current parameter
    uint32_t left;     // samples left in the current partition
    int pbits;         // 4 or 5
    uint32_t esc;      // 15 or 31
    int rawbits;       // >= 0: escape partition with this many raw bits
    uint32_t psize;    // samples per partition
};

__device__ __forceinline__ int32_t rice_next(BitR &br, RiceState &rs) {
    if (rs.left == 0) {
        rs.k = br.get(rs.pbits);
        rs.left = rs.psize;
        rs.rawbits = -1;
        if (rs.k == rs.esc) rs.rawbits = (int)br.get(5);
    }
    rs.left--;
    if (rs.rawbits >= 0) return rs.rawbits ? br.sget(rs.rawbits) : 0;
    const uint32_t z = br.unary();
    const uint32_t u = (z << rs.k) | br.get((int)rs.k);
    return (int32_t)(u >> 1) ^ -(int32_t)(u & 1u);
}

__device__ __forceinline__ int ilog2_u32(uint32_t v) { return 31 - __clz((int)v); }

