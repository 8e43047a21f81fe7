// Bit-parallel string kernels of the PROCLU / flankalign path: one pair per thread.
//
//  vsk_lcs_kernel  consensus LCS filter: LCS_multiple_word / LCS_single_word
//                  (reference src/psearch1.91/bitwise LCS multiple word.c:29-310,
//                  bitwise LCS single word.c:20-199)
//  vsk_ed_kernel   flank edit distance with a free end on the text side:
//                  Edit_Distance_multiple_word_NoEndPenaltySeq1, both copies
//                  (src/psearch1.91/bitwise edit distance ... .c:29-392 and
//                   src/newflankalign/bitwise edit distance ... .c:29-425)
//
// Layout is the reference's: bit g of the vectors is DP column g (bit 0 = the
// empty prefix), so a text of N characters needs N+1 bits, W = (N+64)/64 words.
// The reference loops rows outside and words inside; here the loops are swapped
// (word outside, rows inside) so one word's state lives in registers for the
// whole column block, and the carries that cross from word j-1 to word j are
// kept as one bit per row in a small per-thread array.  Word-level semantics
// are unchanged, so the psearch copy's stale sum-carry (its .c:278-283) and its
// unmasked final scan (.c:352-380) reproduce bit for bit.
#include "vsk_bits.cuh"

__global__ void __launch_bounds__(128)
vsk_lcs_kernel(const uint8_t *__restrict__ seqA, const int64_t *__restrict__ offA,
               const uint8_t *__restrict__ seqB, const int64_t *__restrict__ offB,
               const int32_t *__restrict__ ia, const int32_t *__restrict__ ib, int64_t n,
               int32_t *__restrict__ out) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int a = ia[p], b = ib[p];
    out[p] = vsk_lcs_device(seqA + offA[a], (int)(offA[a + 1] - offA[a]), seqB + offB[b],
                            (int)(offB[b + 1] - offB[b]));
}

template <int VARIANT>
__global__ void __launch_bounds__(128)
vsk_ed_kernel(const uint8_t *__restrict__ tbuf, const int64_t *__restrict__ toff,
              const int32_t *__restrict__ tlen, const uint8_t *__restrict__ pbuf,
              const int64_t *__restrict__ poff, const int32_t *__restrict__ plen, int64_t n,
              int32_t *__restrict__ out) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    out[p] = vsk_ed_device<VARIANT>(tbuf + toff[p], tlen[p], pbuf + poff[p], plen[p]);
}

int vsk_lcs_launch(vsk_ctx *ctx, const uint8_t *d_seqA, const int64_t *d_offA, const uint8_t *d_seqB,
                   const int64_t *d_offB, const int32_t *d_ia, const int32_t *d_ib, int64_t n,
                   int32_t *d_out) {
    if (n <= 0) return VSK_OK;
    const int64_t grid = (n + 127) / 128;
    vsk_lcs_kernel<<<(unsigned)grid, 128, 0, ctx->stream>>>(d_seqA, d_offA, d_seqB, d_offB, d_ia, d_ib,
                                                            n, d_out);
    ctx->launches++;
    VSK_CUDA(ctx, cudaGetLastError());
    return VSK_OK;
}

int vsk_ed_launch(vsk_ctx *ctx, const uint8_t *d_text, const int64_t *d_toff, const int32_t *d_tlen,
                  const uint8_t *d_pat, const int64_t *d_poff, const int32_t *d_plen, int64_t n,
                  int variant, int32_t *d_out) {
    if (n <= 0) return VSK_OK;
    const int64_t grid = (n + 127) / 128;
    if (variant == VSK_ED_PSEARCH)
        vsk_ed_kernel<VSK_ED_PSEARCH><<<(unsigned)grid, 128, 0, ctx->stream>>>(
            d_text, d_toff, d_tlen, d_pat, d_poff, d_plen, n, d_out);
    else
        vsk_ed_kernel<VSK_ED_FLANKALIGN><<<(unsigned)grid, 128, 0, ctx->stream>>>(
            d_text, d_toff, d_tlen, d_pat, d_poff, d_plen, n, d_out);
    ctx->launches++;
    VSK_CUDA(ctx, cudaGetLastError());
    return VSK_OK;
}
