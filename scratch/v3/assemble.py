import os
d = os.path.dirname(os.path.abspath(__file__))
R = lambda n: open(os.path.join(d, n)).read()
head = R('v1_head.cu')
head = head.replace("#define NWARPS (FRX_ENC_THREADS / 32)\n#define PCM_PAD(i) ((i) + ((i) >> 4))\n", "#define NWARPS (FRX_ENC_THREADS / 32)   // compute warps\n")
head = head.replace('''//   K1..K8 frx_encode_kernel     one CTA per 4096-sample frame, persistent grid, ticket-ordered:''', '''//   K1..K8 frx_encode_kernel     one CTA per 4096-sample frame, persistent grid, ticket-ordered; 8 compute warps plus one
//        control warp that takes tickets, stages the next raw samples with TMA bulk copies (mbarrier pipeline), builds the
//        frame header and runs the look-back while the compute warps search, pack and checksum:''')
head = head.replace("    o.fmin = 0.f; o.fmax = 0.f; o.range = 1.0; o.degenerate = 0; o.pad = 0;", "    o.fmin = 0.f; o.fmax = 0.f; o.range = 1.0; o.degenerate = 0; o.flags = 0;")
head = head.replace('''        else { o.fmin = frx_float_unordered(mm[i].fmin_o); o.fmax = frx_float_unordered(mm[i].fmax_o); }''', '''        else {
            o.fmin = frx_float_unordered(mm[i].fmin_o); o.fmax = frx_float_unordered(mm[i].fmax_o);
            const double m = fmax(fabs((double)o.fmin), fabs((double)o.fmax));
            if (m * 8388607.0 < 2147483000.0) o.flags |= FRX_QF_F32_SAFE;
        }''')
tail = R('v1_tail.cu')
tail = tail.replace('''size_t frx_enc_smem_bytes(int frame_words) {
    return ENC_SMALL_BYTES + ENC_UNION_BYTES + sizeof(uint32_t) * (size_t)frame_words;
}''', '''size_t frx_enc_smem_bytes(int frame_words, int dtype) {
    return ENC_SMALL_BYTES + (128 + (size_t)FRX_MAX_BLOCK * frx_dtype_size(dtype)) + PCM_BYTES + RES_BYTES + TABS_BYTES + 2 * sizeof(uint32_t) * (size_t)frame_words;
}''')
tail = tail.replace("size_t smem = frx_enc_smem_bytes(a.frame_words);", "size_t smem = frx_enc_smem_bytes(a.frame_words, a.dtype);")
tail = tail.replace("cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, frx_encode_kernel<WIDE, HISTN>, FRX_ENC_THREADS, smem);", "cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, frx_encode_kernel<WIDE, HISTN>, FRX_ENC_CTA, smem);")
tail = tail.replace("frx_encode_kernel<WIDE, HISTN><<<(unsigned)grid, FRX_ENC_THREADS, smem, st>>>(a);", "frx_encode_kernel<WIDE, HISTN><<<(unsigned)grid, FRX_ENC_CTA, smem, st>>>(a);")
parts = [head, R('v3_mid_a.cu'), R('v1_analyse.cu'), R('v3_mid_b1.cu'), R('v3_mid_b2.cu'), R('v1_resid.cu'), R('v3_mid_c1.cu'), R('v3_mid_c2.cu'),
         R('v3_mid_c3a.cu'), R('v3_mid_c3b.cu'), tail]
open(os.path.join(d, '..', '..', 'flac_raster_b200', 'csrc', 'frx_encode.cu'), 'w').write("".join(parts))
