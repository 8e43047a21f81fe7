// Device functions of the bit-parallel string kernels (see vsk_bits.cu).
#pragma once
#include "vsk_internal.cuh"

#define VSK_MAXSTR 6080               // longest string (bits incl. bit 0 must fit)
#define VSK_CW (VSK_MAXSTR / 64 + 1)  // words of a per-row carry bit array

__device__ __forceinline__ int vsk_code(unsigned char c) {
    switch (c) {
        case 'A': return 0;
        case 'C': return 1;
        case 'G': return 2;
        case 'T': return 3;
        case 'N': return 4;
    }
    return -1;
}

// match vectors of word j: bit b set in peq[c] iff text[64*j + b - 1] has code c
__device__ __forceinline__ void vsk_build_peq(const uint8_t *__restrict__ text, int N, int j,
                                              uint64_t (&peq)[5]) {
#pragma unroll
    for (int c = 0; c < 5; ++c) peq[c] = 0;
    const int g0 = j * 64;
    for (int b = 0; b < 64; ++b) {
        const int g = g0 + b;
        if (g >= 1 && g <= N) {
            const int c = vsk_code(text[g - 1]);
            const uint64_t bit = 1ull << b;
            peq[0] |= (c == 0) ? bit : 0ull;
            peq[1] |= (c == 1) ? bit : 0ull;
            peq[2] |= (c == 2) ? bit : 0ull;
            peq[3] |= (c == 3) ? bit : 0ull;
            peq[4] |= (c == 4) ? bit : 0ull;
        }
    }
}

__device__ __forceinline__ uint64_t vsk_pick(const uint64_t (&peq)[5], int c) {
    uint64_t m = peq[4];
    m = (c == 3) ? peq[3] : m;
    m = (c == 2) ? peq[2] : m;
    m = (c == 1) ? peq[1] : m;
    m = (c == 0) ? peq[0] : m;
    return m;
}

__device__ __forceinline__ bool vsk_all_valid(const uint8_t *s, int n) {
    bool ok = true;
    for (int i = 0; i < n; ++i) ok &= (vsk_code(s[i]) >= 0);
    return ok;
}


// LCS length as LCS_multiple_word(s1, s2, n1, m1) returns it (s1 = string1).
__device__ __forceinline__ int vsk_lcs_device(const uint8_t *__restrict__ s1, int n1,
                                              const uint8_t *__restrict__ s2, int m1) {
    // the longer string is bit-indexed; on a tie string1 (multiple word.c:76-86)
    const uint8_t *text, *pat;
    int N, M;
    if (n1 >= m1) { text = s1; N = n1; pat = s2; M = m1; }
    else { text = s2; N = m1; pat = s1; M = n1; }
    if (!vsk_all_valid(text, N) || !vsk_all_valid(pat, M)) return -1;
    // single-word copy starts its bitmask at 0x10 (single word.c:59): for N in
    // 61..63 the characters from position 60 on are shifted out of the word.
    if (N <= 63 && N > 60) N = 60;
    const int W = (N + 64) >> 6;
    uint64_t carry[VSK_CW];
    int count = 0;
    for (int j = 0; j < W; ++j) {
        uint64_t peq[5];
        vsk_build_peq(text, N, j, peq);
        uint64_t V = ~0ull;
        uint64_t cw_in = 0, cw_out = 0;
        for (int i = 0; i < M; ++i) {
            if ((i & 63) == 0) {
                if (j > 0) cw_in = carry[i >> 6];
                cw_out = 0;
            }
            const uint64_t m = vsk_pick(peq, vsk_code(pat[i]));
            const uint64_t x = V & m;
            uint64_t s = V + x;
            uint64_t co = (s < x) ? 1ull : 0ull;
            const uint64_t cin = (cw_in >> (i & 63)) & 1ull;
            const uint64_t s2v = s + cin;
            co |= (s2v < s) ? 1ull : 0ull;
            V = s2v | (V & ~m);
            cw_out |= co << (i & 63);
            if ((i & 63) == 63 || i == M - 1) carry[i >> 6] = cw_out;
        }
        uint64_t z = ~V;
        const int top = N - j * 64;  // highest valid bit index in this word
        if (top < 63) z &= (2ull << top) - 1ull;
        count += __popcll(z);
    }
    return count;
}

// Edit_Distance_multiple_word_NoEndPenaltySeq1(text, pat, N, M), either copy.
// complement of a flank character as flankalign builds leftcp/rightcp (flankalign.c:523-526)
__device__ __forceinline__ uint8_t vsk_comp_char(uint8_t c) {
    switch (c) {
        case 'A': return 'T';
        case 'C': return 'G';
        case 'G': return 'C';
        case 'T': return 'A';
    }
    return c;
}

// COMP: the pattern is read through vsk_comp_char (complemented on the fly)
template <int VARIANT, bool COMP = false>
__device__ __forceinline__ int vsk_ed_device(const uint8_t *__restrict__ text, int N,
                                             const uint8_t *__restrict__ pat, int M) {
    if (M == 0) return 0;
    if (!vsk_all_valid(text, N) || !vsk_all_valid(pat, M)) return -1;
    const int W = (N + 64) >> 6;
    uint64_t c_sum[VSK_CW], c_v1[VSK_CW], c_v0[VSK_CW];
    int score = M, best = M;
    for (int j = 0; j < W; ++j) {
        uint64_t peq[5];
        vsk_build_peq(text, N, j, peq);
        uint64_t PD = (j == 0) ? 1ull : 0ull, PS = PD;
        uint64_t in_sum = 0, in_v1 = ~0ull, in_v0 = 0;  // word 0: sum carry 0, +1 shift carry 1
        uint64_t o_sum = 0, o_v1 = 0, o_v0 = 0;
        for (int i = 0; i < M; ++i) {
            const int sh = i & 63;
            if (sh == 0) {
                if (j > 0) { in_sum = c_sum[i >> 6]; in_v1 = c_v1[i >> 6]; in_v0 = c_v0[i >> 6]; }
                o_sum = o_v1 = o_v0 = 0;
            }
            const uint64_t m = vsk_pick(peq, vsk_code(COMP ? vsk_comp_char(pat[i]) : pat[i]));
            const uint64_t cs = (in_sum >> sh) & 1ull;
            const uint64_t mm = m | PD;
            const uint64_t ris = ~mm;
            const uint64_t nmi = ris | PS;
            const uint64_t x = nmi ^ PD;
            uint64_t sum, cso;
            if (VARIANT == VSK_ED_PSEARCH) {
                sum = nmi + PS + cs;
                cso = cs | ((sum < nmi) ? 1ull : 0ull);
            } else {
                sum = nmi + cs;
                cso = (sum < nmi) ? 1ull : 0ull;
                sum += PS;
                cso |= (sum < PS) ? 1ull : 0ull;
            }
            const uint64_t sr = sum & ris;
            const uint64_t v0 = sr ^ x;
            const uint64_t v1 = PD | (sr & PS);
            const uint64_t v0s = (v0 << 1) + ((in_v0 >> sh) & 1ull);
            const uint64_t v1s = (v1 << 1) + ((in_v1 >> sh) & 1ull);
            o_sum |= cso << sh;
            o_v0 |= (v0 >> 63) << sh;
            o_v1 |= (v1 >> 63) << sh;
            PD = mm & v1s;
            PS = v1s | (mm & v0s);
            if (sh == 63 || i == M - 1) { c_sum[i >> 6] = o_sum; c_v1[i >> 6] = o_v1; c_v0[i >> 6] = o_v0; }
        }
        // final-row scan of this word (free end on the text side)
        uint64_t d = PD, ins = ~PS;
        if (VARIANT == VSK_ED_FLANKALIGN && j == W - 1) {
            const int junk = 63 + (W - 1) * 64 - N;
            const uint64_t mask = ~0ull >> junk;
            d &= mask;
            ins &= mask;
        }
        for (int b = (j == 0) ? 1 : 0; b < 64; ++b) {
            score += (int)((ins >> b) & 1ull) - (int)((d >> b) & 1ull);
            best = min(best, score);
        }
    }
    return best;
}
