"""Assemble csrc/frx_encode.cu from the round-1 file's K0 head and launcher tail (git show) and the v2 parts."""
import subprocess, sys
from pathlib import Path
root = Path(__file__).resolve().parents[2]
r1 = subprocess.run(["git", "show", "ed1290f:flac_raster_b200/csrc/frx_encode.cu"], capture_output=True, text=True, cwd=root, check=True).stdout.split("\n")
def find(s):
    for i, l in enumerate(r1):
        if s in l:
            return i
    raise KeyError(s)
head = "\n".join(r1[:find("// bit writer into the zero-initialised shared frame buffer") - 1])
tail = "\n".join(r1[find("// K0+K1 alone (frx_quantize): interleaved int32 PCM of one tile") - 1:])
head = (Path(__file__).parent / "header.txt").read_text() + head[head.index("#include <math_constants.h>"):]
head = head.replace("    q[i] = o;\n", "    o.rcp = o.range != 0.0 ? 1.0 / o.range : 0.0;   // correctly rounded (IEEE division)\n    q[i] = o;\n")
assert "o.rcp" in head
head = head.replace("    q[i] = o;\n", """    o.irange = 0; o.rcpf = 0.f;
    if ((dtype == FRX_U8 || dtype == FRX_U16 || dtype == FRX_I16) && !o.degenerate && o.imax - o.imin <= (dtype == FRX_I16 ? 32767 : 65535)) {
        // neither max - min nor any x - min wraps in the raster dtype: the integer form of the quantiser applies
        o.irange = (int)(o.imax - o.imin); o.rcpf = __fdiv_rz(1.0f, (float)o.irange); o.flags |= FRX_QF_INT16;
    }
    q[i] = o;\n""")
assert "o.irange" in head
head = head.replace("#define HIST FRX_MAX_LPC\n#define NWARPS (FRX_ENC_THREADS / 32)   // compute warps\n", "")
old = tail[tail.index("size_t frx_enc_smem_bytes"):tail.index("size_t frx_minmax_raw_size()")]
tail = tail.replace(old, """size_t frx_enc_smem_bytes(int frame_words, int dtype) {
    (void)frame_words; (void)dtype;
    return 2048 + (size_t)ENC_WARPS * WARP_BYTES;
}
int frx_enc_warps_per_cta() { return ENC_WARPS; }

""")
old = tail[tail.index("template <bool WIDE, int HISTN>\nstatic cudaError_t launch_encode_t"):tail.index("cudaError_t frx_launch_encode(")]
tail = tail.replace(old, """template <bool WIDE, int HISTN>
static cudaError_t launch_encode_t(const EncArgs &a, int sm_count, cudaStream_t st) {
    const size_t smem = frx_enc_smem_bytes(0, 0);
    cudaError_t e = cudaFuncSetAttribute(frx_encode_kernel<WIDE, HISTN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(frx_encode_kernel<WIDE, HISTN>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return e;
    // persistent: one CTA of twelve frame warps per SM, tickets hand out the frames in order
    long long grid = sm_count;
    const long long need = ((long long)a.total_frames + ENC_WARPS - 1) / ENC_WARPS;
    if (grid > need) grid = need;
    if (grid < 1) grid = 1;
    frx_encode_kernel<WIDE, HISTN><<<(unsigned)grid, ENC_CTA, smem, st>>>(a);
    return cudaGetLastError();
}

""")
tail = tail[:tail.index('#ifdef FRX_DBG\nextern "C" void frx_debug_trace')]
mid = "".join((Path(__file__).parent / f"mid_{i}.cu").read_text() for i in range(1, 9))
mid9 = (Path(__file__).parent / "mid_9.cu").read_text()
(root / "flac_raster_b200" / "csrc" / "frx_encode.cu").write_text(head + mid + "\n" + tail + mid9)
print("assembled")
