// Profile-to-profile global alignment kernels (the PROCLU DP).
//
// Replaces GetFixedProfileAP (reference src/psearch1.91/profile.h:2040-2227) and
// GetFixedProfileAPNarrowband (profile.h:1733-2038).  One pair per warp.  The
// (L1+1)x(L2+1) matrix is swept in strips of 32*CPL columns; inside a strip
// lane k owns CPL consecutive columns and trails lane k-1 by one row, so the 32
// lanes sit on an anti-diagonal wavefront and the cell to the left of a lane's
// strip arrives by __shfl_up_sync.  The previous row of a lane's columns lives
// in registers.  The reference stores a direction per cell and walks back from
// the corner summing dashnum*cost and dashnum (profile.h:2180-2210); here each
// cell carries those two sums (and the path length) of its chosen predecessor
// forward, so nothing but the running wavefront is ever stored.
//
// Substitution costs: the reference gathers from the 501,501-byte triangular
// table per cell.  Per pair the warp builds the two local alphabets (distinct
// composition ids of each profile) and a compact na x nb byte table in shared
// memory; the per-cell gather is then a conflict-light LDS.  Pairs whose compact
// table would not fit fall back to gathering from the full symmetric table in
// global memory (L1/L2 resident).
//
// Narrowband: the band of the reference (per-row lstart/rend from IEEE double
// row*svector, profile.h:1835-1839) is applied as a mask: out-of-band cells are
// VSK_BIG, exactly the value of the sentinels the reference plants beside the
// band (profile.h:1850-1877), so they never win or tie against in-band cells.
#include <algorithm>

#include <cstdlib>

#include <mutex>

#include "vsk_internal.cuh"

#define DP_WARPS 4
#define DP_TBYTES 8192
#define FULLMASK 0xffffffffu

struct DpWarpSmem {
    uint8_t T[DP_TBYTES];   // compact substitution table [na][nb]
    uint8_t mapL[1024];     // global id -> local row index
    uint32_t bmL[32];       // presence bitmaps over ids 0..1023
    uint32_t bmT[32];
    uint16_t preT[32];      // exclusive popcount prefix of bmT
    uint16_t invL[256];     // local -> global id
    uint16_t invT[256];
};

#define DP_SMEM_BYTES (2048 + DP_WARPS * sizeof(DpWarpSmem))

template <int CPL, bool MASKED, bool SMEMT>
__device__ __forceinline__ void dp_row_cells(int (&Hd)[CPL], int (&Hp1)[CPL], int (&Hp2)[CPL],
                                             const int (&tb)[CPL], const int (&gt)[CPL],
                                             const int (&ci1)[CPL], const int (&ci2)[CPL], int dgd,
                                             int dgp1, int dgp2, int lfd, int lfp1, int lfp2,
                                             const uint8_t *__restrict__ trow, int gl, int ndl,
                                             int ri1, int ri2, int lo, int hi) {
#pragma unroll
    for (int j = 0; j < CPL; ++j) {
        const int upd = Hd[j], upp1 = Hp1[j], upp2 = Hp2[j];
        const int dlt = SMEMT ? (int)trow[tb[j]] : (int)__ldg(trow + tb[j]);
        const int sa = dgd + dlt;  // diagonal
        const int f = upd + gl;    // down  (gap in the read / top profile)
        const int e = lfd + gt[j]; // right (gap in the reference / left profile)
        // min3switch (profile.h:1098-1099): diag if sa<=f && sa<=e; else down if f<=e; else right
        const bool pF = f <= e;
        const int m1 = pF ? f : e;
        const bool pD = sa <= m1;
        int d = pD ? sa : m1;
        int dn = ndl + ci2[j] - VSK_LEN1;  // NONDASHES[l] + NONDASHES[t]
        dn = dn < 10 ? dn : 10;
        const int p1d = dgp1 + dn * dlt, p2d = dgp2 + dn + VSK_LEN1;
        const int p1f = upp1 + ri1, p2f = upp2 + ri2;
        const int p1e = lfp1 + ci1[j], p2e = lfp2 + ci2[j];
        const int p1 = pD ? p1d : (pF ? p1f : p1e);
        const int p2 = pD ? p2d : (pF ? p2f : p2e);
        if (MASKED) {
            if (j < lo || j > hi) d = VSK_BIG;
        }
        dgd = upd; dgp1 = upp1; dgp2 = upp2;
        Hd[j] = d; Hp1[j] = p1; Hp2[j] = p2;
        lfd = d; lfp1 = p1; lfp2 = p2;
    }
}

// Row record of the reference-side profile, built once per pair: bits 0..15 = offset of the
// row inside the substitution table (local row index * nb, or the composition id when the
// table is the global one), 16..23 = gap cost d(0,l), 24..31 = NONDASHES[l].
__device__ __forceinline__ uint32_t dp_pack_row(int off, int gn) {
    return (uint32_t)off | ((uint32_t)(gn & 255) << 16) | ((uint32_t)(gn >> 8) << 24);
}

template <int CPL, bool SMEMT>
__device__ __forceinline__ void dp_pair(const vsk_dp_job &job, const uint16_t *__restrict__ Tp,
                                        const uint8_t *__restrict__ dist, const uint16_t *s_gapnd,
                                        DpWarpSmem &ws, int nb, int4 *bscr, const uint32_t *rowrec,
                                        int4 *out) {
    const int lane = threadIdx.x & 31;
    const int l1 = job.l1, l2 = job.l2, br = job.br;
    const bool banded = br >= 0;
    const double sv = job.svector;
    const int npass = (l2 + 32 * CPL - 1) / (32 * CPL);
    int totd = 0, totp1 = 0, totp2 = 0;  // row-0 prefix over the strips already done
    int Hd[CPL], Hp1[CPL], Hp2[CPL];
    int prev_r_hi = l1;  // rows [.., prev_r_hi) of the previous strip left a boundary cell in bscr

    for (int pass = 0; pass < npass; ++pass) {
        const int cbase = pass * 32 * CPL + lane * CPL;  // matrix column left of this lane's strip
        int tb[CPL], gt[CPL], ci1[CPL], ci2[CPL];
        int sd = 0, s1 = 0, s2 = 0;
#pragma unroll
        for (int j = 0; j < CPL; ++j) {
            const int c = cbase + j;
            const int id = (c < l2) ? (int)Tp[c] : 0;  // id 0: gap cost 0, no non-dash symbols
            const int gn = s_gapnd[id];
            const int g = gn & 255, nd = gn >> 8;
            gt[j] = g;
            ci1[j] = nd * g;
            ci2[j] = nd + VSK_LEN1;
            if (SMEMT) {
                const int w = id >> 5;
                tb[j] = ws.preT[w] + __popc(ws.bmT[w] & ((1u << (id & 31)) - 1u));
            } else {
                tb[j] = id;
            }
            sd += g; s1 += nd * g; s2 += nd + VSK_LEN1;
            Hd[j] = sd; Hp1[j] = s1; Hp2[j] = s2;
        }
        // row 0 (profile.h:2074-2079): cumulative gap costs; exclusive warp scan
        int xd = sd, x1 = s1, x2 = s2;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int a = __shfl_up_sync(FULLMASK, xd, o);
            const int b = __shfl_up_sync(FULLMASK, x1, o);
            const int c = __shfl_up_sync(FULLMASK, x2, o);
            if (lane >= o) { xd += a; x1 += b; x2 += c; }
        }
        const int ed = xd - sd + totd, e1 = x1 - s1 + totp1, e2 = x2 - s2 + totp2;
#pragma unroll
        for (int j = 0; j < CPL; ++j) {
            Hd[j] += ed; Hp1[j] += e1; Hp2[j] += e2;
            if (banded && cbase + j + 1 > br) Hd[j] = VSK_BIG;  // profile.h:1801-1802
        }
        int dgd = ed, dgp1 = e1, dgp2 = e2;  // (row 0, column cbase)
        if (banded && cbase > br) dgd = VSK_BIG;
        totd += __shfl_sync(FULLMASK, xd, 31);
        totp1 += __shfl_sync(FULLMASK, x1, 31);
        totp2 += __shfl_sync(FULLMASK, x2, 31);

        // rows [r_lo, r_hi) (0-based) are the only ones whose band can touch this strip's columns
        // [c0+1, c0+32*CPL]; outside them every cell of the strip is out of band.  The bounds are
        // conservative (the per-cell mask below stays exact): center = (int)(row*sv) lies in
        // (row*sv - 1, row*sv].
        int r_lo = 0, r_hi = l1;
        if (banded) {
            const int c0 = pass * 32 * CPL;
            const double a = (double)(c0 + 1 - br), b = (double)(min(l2, c0 + 32 * CPL) + br + 1);
            r_lo = max(0, (int)floor(a / sv) - 2);
            r_hi = min(l1, (int)ceil(b / sv) + 2);
            if (r_hi < r_lo) r_hi = r_lo;
        }
        if (r_lo > 0) {  // the row above the first processed one is entirely out of band here
#pragma unroll
            for (int j = 0; j < CPL; ++j) Hd[j] = VSK_BIG;
            dgd = VSK_BIG;
        }
        const bool rd_scratch = (lane == 0) && (pass > 0);
        const bool wr_scratch = (lane == 31) && (pass < npass - 1);
        int4 nxt = make_int4(VSK_BIG, 0, 0, 0);
        if (rd_scratch && r_lo < prev_r_hi) nxt = __ldcg(bscr + r_lo);
        // column 0 (profile.h:2084-2090) for lane 0 of the first strip is the running sum in dg*
        uint32_t rec_next = __ldcg(rowrec + min(max(r_lo - lane, 0), l1 - 1));

        auto step = [&](const int s, const bool check) {
            const int r = s - lane;  // 0-based row of the reference (left) profile
            int lfd = __shfl_up_sync(FULLMASK, Hd[CPL - 1], 1);
            int lfp1 = __shfl_up_sync(FULLMASK, Hp1[CPL - 1], 1);
            int lfp2 = __shfl_up_sync(FULLMASK, Hp2[CPL - 1], 1);
            const bool active = !check || ((r >= r_lo) && (r < r_hi));
            const uint32_t rec = rec_next;
            rec_next = __ldcg(rowrec + min(max(r + 1, 0), l1 - 1));
            const int gl = (rec >> 16) & 255, ndl = rec >> 24;
            const int ri1 = ndl * gl, ri2 = ndl + VSK_LEN1;
            const uint8_t *trow = SMEMT ? (ws.T + (rec & 0xffffu)) : (dist + (rec & 0xffffu) * VSK_DROW);
            int lo = 0, hi = CPL;
            bool need_mask = false;
            if (active) {
                if (lane == 0) {
                    if (pass == 0) {
                        lfd = dgd + gl; lfp1 = dgp1 + ri1; lfp2 = dgp2 + ri2;
                    } else {
                        lfd = nxt.x; lfp1 = nxt.y; lfp2 = nxt.z;
                        nxt = (r + 1 < prev_r_hi) ? __ldcg(bscr + r + 1) : make_int4(VSK_BIG, 0, 0, 0);
                    }
                }
                if (banded) {  // profile.h:1835-1839
                    const int center = __double2int_rz(__dmul_rn((double)(r + 1), sv));
                    const int lstart = max(1, center - br);
                    const int rend = min(l2, center + br);
                    lo = lstart - (cbase + 1);
                    hi = rend - (cbase + 1);
                    need_mask = (lo > 0) || (hi < CPL - 1);
                }
            }
            const bool anymask = banded && __any_sync(FULLMASK, need_mask);
            if (active) {
                if (anymask)
                    dp_row_cells<CPL, true, SMEMT>(Hd, Hp1, Hp2, tb, gt, ci1, ci2, dgd, dgp1, dgp2,
                                                   lfd, lfp1, lfp2, trow, gl, ndl, ri1, ri2, lo, hi);
                else
                    dp_row_cells<CPL, false, SMEMT>(Hd, Hp1, Hp2, tb, gt, ci1, ci2, dgd, dgp1, dgp2,
                                                    lfd, lfp1, lfp2, trow, gl, ndl, ri1, ri2, lo, hi);
                dgd = lfd; dgp1 = lfp1; dgp2 = lfp2;
                if (wr_scratch) bscr[r] = make_int4(Hd[CPL - 1], Hp1[CPL - 1], Hp2[CPL - 1], 0);
            }
        };
        // pipeline fill, steady state (every lane active, no range checks), drain
        const int s_fill_end = min(r_lo + 31, r_hi + 31);
        int s = r_lo;
        for (; s < s_fill_end; ++s) step(s, true);
        for (; s < r_hi; ++s) step(s, false);
        for (; s < r_hi + 31; ++s) step(s, true);
        prev_r_hi = r_hi;
        __syncwarp();  // scratch written by lane 31 is read by lane 0 in the next strip
    }
    // the corner (l1, l2) is in band, so the last strip ran through row l1-1
    const int cl = (l2 - 1) - (npass - 1) * 32 * CPL;
    if (lane == cl / CPL) {
        const int jr = cl % CPL;
#pragma unroll
        for (int j = 0; j < CPL; ++j)
            if (j == jr)
                out[job.out_idx] = make_int4(Hd[j], Hp1[j], Hp2[j] & (VSK_LEN1 - 1), Hp2[j] >> 17);
    }
}

__global__ void __launch_bounds__(DP_WARPS * 32)
vsk_dp_kernel(const uint16_t *__restrict__ left_ids, const uint16_t *__restrict__ top_ids,
              const vsk_dp_job *__restrict__ jobs, int n_jobs, const uint8_t *__restrict__ dist,
              const uint16_t *__restrict__ gapnd, int4 *scratch, int scratch_stride, int *queue,
              int4 *out, const int *__restrict__ n_jobs_dev, int skip_dir) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint16_t *s_gapnd = reinterpret_cast<uint16_t *>(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (n_jobs_dev) n_jobs = *n_jobs_dev;  // pairs handed back by the direction-bit path
    DpWarpSmem &ws = reinterpret_cast<DpWarpSmem *>(smem + 2048)[warp];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) s_gapnd[i] = gapnd[i];
    __syncthreads();
    // per-warp scratch slice: scratch_stride boundary cells (int4) followed by scratch_stride row records
    int4 *my_bscr = scratch + (size_t)(blockIdx.x * DP_WARPS + warp) * (scratch_stride + scratch_stride / 4);
    uint32_t *my_rowrec = reinterpret_cast<uint32_t *>(my_bscr + scratch_stride);

    for (;;) {
        int ji = 0;
        if (lane == 0) ji = atomicAdd(queue, 1);
        ji = __shfl_sync(FULLMASK, ji, 0);
        if (ji >= n_jobs) break;
        const vsk_dp_job job = jobs[ji];
        if (skip_dir && VSK_DP_DIRJOB(job.l2)) continue;  // done by vsk_dp_dir_kernel + vsk_dp_trace_kernel
        const uint16_t *L = left_ids + job.left_off;
        const uint16_t *Tp = top_ids + job.top_off;
        const int l1 = job.l1, l2 = job.l2;

        // local alphabets
        ws.bmL[lane] = 0;
        ws.bmT[lane] = 0;
        __syncwarp();
        for (int i = lane; i < l1; i += 32) {
            const int id = L[i];
            atomicOr(&ws.bmL[id >> 5], 1u << (id & 31));
        }
        for (int i = lane; i < l2; i += 32) {
            const int id = Tp[i];
            atomicOr(&ws.bmT[id >> 5], 1u << (id & 31));
        }
        __syncwarp();
        const uint32_t bl = ws.bmL[lane], bt = ws.bmT[lane];
        int xl = __popc(bl), xt = __popc(bt);
        const int cl = xl, ct = xt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int a = __shfl_up_sync(FULLMASK, xl, o);
            const int b = __shfl_up_sync(FULLMASK, xt, o);
            if (lane >= o) { xl += a; xt += b; }
        }
        const int na = __shfl_sync(FULLMASK, xl, 31), nb = __shfl_sync(FULLMASK, xt, 31);
        const bool smemt = (na <= 256) && (nb <= 256) && (na * nb <= DP_TBYTES);
        if (smemt) {
            ws.preT[lane] = (uint16_t)(xt - ct);
            uint32_t b = bl;
            int rk = xl - cl;
            while (b) {
                const int id = lane * 32 + __ffs(b) - 1;
                b &= b - 1;
                ws.mapL[id] = (uint8_t)rk;
                ws.invL[rk] = (uint16_t)id;
                ++rk;
            }
            b = bt;
            rk = xt - ct;
            while (b) {
                const int id = lane * 32 + __ffs(b) - 1;
                b &= b - 1;
                ws.invT[rk] = (uint16_t)id;
                ++rk;
            }
            __syncwarp();
#pragma unroll 4
            for (int a = 0; a < na; ++a) {
                const uint8_t *drow = dist + (int)ws.invL[a] * VSK_DROW;
                for (int c = lane; c < nb; c += 32) ws.T[a * nb + c] = __ldg(drow + ws.invT[c]);
            }
            __syncwarp();
        }
        // row records of the reference-side profile
        for (int i = lane; i < l1; i += 32) {
            const int id = L[i];
            my_rowrec[i] = dp_pack_row(smemt ? (int)ws.mapL[id] * nb : id, s_gapnd[id]);
        }
        __syncwarp();
#define DP_DISPATCH(CPLV)                                                                         \
    do {                                                                                          \
        if (smemt) dp_pair<CPLV, true>(job, Tp, dist, s_gapnd, ws, nb, my_bscr, my_rowrec, out);  \
        else dp_pair<CPLV, false>(job, Tp, dist, s_gapnd, ws, nb, my_bscr, my_rowrec, out);       \
    } while (0)
        if (l2 <= 32) DP_DISPATCH(1);
        else if (l2 <= 64) DP_DISPATCH(2);
        else if (l2 <= 128) DP_DISPATCH(4);
        else DP_DISPATCH(8);
#undef DP_DISPATCH
        __syncwarp();
    }
}

// ---- quarter-warp kernel for medium pairs ------------------------------------------------
// Pairs with at most 128 columns waste most of a warp's pipeline (31 fill steps, a per-pair
// table build).  Here 8 lanes own a pair (four pairs per warp, each group with its own job
// loop and its own sync mask), one strip of 8*CPL columns, and the substitution costs come from
// the dense K x K table of the data set in shared memory, so there is no per-pair set-up.
#define Q8_WARPS 4
#define Q8_G 8

template <int CPL>
__device__ __forceinline__ void dp_q8_pair(const vsk_dp_job &job, const uint16_t *__restrict__ L,
                                           const uint16_t *__restrict__ Tp, const uint16_t *s_gapnd,
                                           const uint8_t *s_dense, const uint8_t *s_ktab, int K,
                                           unsigned gmask, int glane, int4 *out) {
    const int l1 = job.l1, l2 = job.l2, br = job.br;
    const bool banded = br >= 0;
    const double sv = job.svector;
    const int cbase = glane * CPL;
    int Hd[CPL], Hp1[CPL], Hp2[CPL], tb[CPL], gt[CPL], ci1[CPL], ci2[CPL];
    int sd = 0, s1 = 0, s2 = 0;
#pragma unroll
    for (int j = 0; j < CPL; ++j) {
        const int c = cbase + j;
        const int id = (c < l2) ? (int)Tp[c] : 0;
        const int gn = s_gapnd[id];
        const int g = gn & 255, nd = gn >> 8;
        gt[j] = g;
        ci1[j] = nd * g;
        ci2[j] = nd + VSK_LEN1;
        tb[j] = s_dense[id];
        sd += g; s1 += nd * g; s2 += nd + VSK_LEN1;
        Hd[j] = sd; Hp1[j] = s1; Hp2[j] = s2;
    }
    int xd = sd, x1 = s1, x2 = s2;
#pragma unroll
    for (int o = 1; o < Q8_G; o <<= 1) {
        const int a = __shfl_up_sync(gmask, xd, o, Q8_G);
        const int b = __shfl_up_sync(gmask, x1, o, Q8_G);
        const int c = __shfl_up_sync(gmask, x2, o, Q8_G);
        if (glane >= o) { xd += a; x1 += b; x2 += c; }
    }
    const int ed = xd - sd, e1 = x1 - s1, e2 = x2 - s2;
#pragma unroll
    for (int j = 0; j < CPL; ++j) {
        Hd[j] += ed; Hp1[j] += e1; Hp2[j] += e2;
        if (banded && cbase + j + 1 > br) Hd[j] = VSK_BIG;
    }
    int dgd = ed, dgp1 = e1, dgp2 = e2;
    if (banded && cbase > br) dgd = VSK_BIG;
    int id_next = L[min(max(-glane, 0), l1 - 1)];

    auto step = [&](const int s, const bool check) {
        const int r = s - glane;
        int lfd = __shfl_up_sync(gmask, Hd[CPL - 1], 1, Q8_G);
        int lfp1 = __shfl_up_sync(gmask, Hp1[CPL - 1], 1, Q8_G);
        int lfp2 = __shfl_up_sync(gmask, Hp2[CPL - 1], 1, Q8_G);
        const bool active = !check || ((r >= 0) && (r < l1));
        const int id = id_next;
        id_next = L[min(max(r + 1, 0), l1 - 1)];
        const int gn = s_gapnd[id];
        const int gl = gn & 255, ndl = gn >> 8;
        const int ri1 = ndl * gl, ri2 = ndl + VSK_LEN1;
        const uint8_t *trow = s_ktab + (int)s_dense[id] * K;
        int lo = 0, hi = CPL;
        if (active) {
            if (glane == 0) { lfd = dgd + gl; lfp1 = dgp1 + ri1; lfp2 = dgp2 + ri2; }
            if (banded) {  // profile.h:1835-1839
                const int center = __double2int_rz(__dmul_rn((double)(r + 1), sv));
                lo = max(1, center - br) - (cbase + 1);
                hi = min(l2, center + br) - (cbase + 1);
            }
            // the four groups of a warp run different pairs: choosing the masked code per step would
            // make them diverge, so a banded pair always takes it
            if (banded)
                dp_row_cells<CPL, true, true>(Hd, Hp1, Hp2, tb, gt, ci1, ci2, dgd, dgp1, dgp2, lfd, lfp1, lfp2, trow,
                                              gl, ndl, ri1, ri2, lo, hi);
            else
                dp_row_cells<CPL, false, true>(Hd, Hp1, Hp2, tb, gt, ci1, ci2, dgd, dgp1, dgp2, lfd, lfp1, lfp2, trow,
                                               gl, ndl, ri1, ri2, lo, hi);
            dgd = lfd; dgp1 = lfp1; dgp2 = lfp2;
        }
    };
    const int s_fill_end = min(Q8_G - 1, l1 + Q8_G - 1);
    int s = 0;
    for (; s < s_fill_end; ++s) step(s, true);
    for (; s < l1; ++s) step(s, false);
    for (; s < l1 + Q8_G - 1; ++s) step(s, true);
    const int cl = l2 - 1;
    if (glane == cl / CPL) {
        const int jr = cl % CPL;
#pragma unroll
        for (int j = 0; j < CPL; ++j)
            if (j == jr)
                out[job.out_idx] = make_int4(Hd[j], Hp1[j], Hp2[j] & (VSK_LEN1 - 1), Hp2[j] >> 17);
    }
}

__global__ void __launch_bounds__(Q8_WARPS * 32)
vsk_dp_q8_kernel(const uint16_t *__restrict__ left_ids, const uint16_t *__restrict__ top_ids,
                 const vsk_dp_job *__restrict__ jobs, int n_jobs, const uint8_t *__restrict__ dense,
                 const uint8_t *__restrict__ ktab, int K, const uint16_t *__restrict__ gapnd, int *queue,
                 int4 *__restrict__ out) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint16_t *s_gapnd = reinterpret_cast<uint16_t *>(smem);
    uint8_t *s_dense = smem + 2048;
    uint8_t *s_ktab = s_dense + 1024;
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) { s_gapnd[i] = gapnd[i]; s_dense[i] = dense[i]; }
    for (int i = threadIdx.x; i < K * K; i += blockDim.x) s_ktab[i] = ktab[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, glane = lane & (Q8_G - 1);
    const unsigned gmask = 0xffu << (lane & ~(Q8_G - 1));
    for (;;) {
        int ji = 0;
        if (glane == 0) ji = atomicAdd(queue, 1);
        ji = __shfl_sync(gmask, ji, 0, Q8_G);
        if (ji >= n_jobs) break;
        const vsk_dp_job job = jobs[ji];
        const uint16_t *L = left_ids + job.left_off;
        const uint16_t *Tp = top_ids + job.top_off;
        if (job.l2 <= 32) dp_q8_pair<4>(job, L, Tp, s_gapnd, s_dense, s_ktab, K, gmask, glane, out);
        else if (job.l2 <= 64) dp_q8_pair<8>(job, L, Tp, s_gapnd, s_dense, s_ktab, K, gmask, glane, out);
        else dp_q8_pair<16>(job, L, Tp, s_gapnd, s_dense, s_ktab, K, gmask, glane, out);
        __syncwarp(gmask);
    }
}

// ---- thread-per-pair kernel for tiny pairs ---------------------------------------------
// Most pairs PROCLU aligns are tiny (patterns of 7-30 columns): on the warp-per-pair kernel they
// pay 31 pipeline-fill steps and a per-pair table build for a few hundred cells.  Here one thread
// owns one pair (full matrix only, <= 32 columns, <= 64 rows): its previous row lives in shared
// memory laid out [column][thread] (conflict free), the cell state is packed in two words
// (w0 = distance:15 | length:7 | sumw:10 -- the bounds follow from the size limits -- and
// w1 = dsum), and substitution costs come from a dense K x K table over the composition ids that
// actually occur in the two profile sets, resident in shared memory.
#define TINY_THREADS 128
#define TINY_MAXCOLS 32
#define TINY_MAXROWS 64
#define TINY_DMASK 0x7fffu
#define TINY_LEN1 (1u << 15)
#define TINY_W1 (1u << 22)

__global__ void __launch_bounds__(TINY_THREADS)
vsk_dp_tiny_kernel(const uint16_t *__restrict__ left_ids, const uint16_t *__restrict__ top_ids,
                   const vsk_dp_job *__restrict__ jobs, int n_jobs, const uint8_t *__restrict__ dense,
                   const uint8_t *__restrict__ ktab, int K, const uint16_t *__restrict__ gapnd,
                   int4 *__restrict__ out) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint32_t *s_w0 = reinterpret_cast<uint32_t *>(smem);                 // [33][128]
    uint32_t *s_w1 = s_w0 + (TINY_MAXCOLS + 1) * TINY_THREADS;           // [33][128]
    uint32_t *s_col = s_w1 + (TINY_MAXCOLS + 1) * TINY_THREADS;          // [32][128]
    uint16_t *s_gapnd = reinterpret_cast<uint16_t *>(s_col + TINY_MAXCOLS * TINY_THREADS);  // [1024]
    uint8_t *s_dense = reinterpret_cast<uint8_t *>(s_gapnd + 1024);      // [1024]
    uint8_t *s_ktab = s_dense + 1024;                                    // [K*K]
    const int tid = threadIdx.x;
    for (int i = tid; i < 1024; i += TINY_THREADS) { s_gapnd[i] = gapnd[i]; s_dense[i] = dense[i]; }
    for (int i = tid; i < K * K; i += TINY_THREADS) s_ktab[i] = ktab[i];
    __syncthreads();
    const int ji = blockIdx.x * TINY_THREADS + tid;
    if (ji >= n_jobs) return;
    const vsk_dp_job job = jobs[ji];
    const uint16_t *L = left_ids + job.left_off;
    const uint16_t *T = top_ids + job.top_off;
    const int l1 = job.l1, l2 = job.l2;
    // column constants and row 0 (profile.h:2074-2079)
    uint32_t w0 = 0, w1 = 0;
    s_w0[tid] = 0;
    s_w1[tid] = 0;
    for (int c = 0; c < l2; ++c) {
        const int id = T[c];
        const uint32_t gn = s_gapnd[id];
        const uint32_t g = gn & 255u, nd = gn >> 8;
        s_col[c * TINY_THREADS + tid] = (uint32_t)s_dense[id] | (g << 8) | (nd << 16);
        w0 += g + TINY_LEN1 + nd * TINY_W1;
        w1 += nd * g;
        s_w0[(c + 1) * TINY_THREADS + tid] = w0;
        s_w1[(c + 1) * TINY_THREADS + tid] = w1;
    }
    uint32_t lf0 = 0, lf1 = 0;
    for (int r = 0; r < l1; ++r) {
        const int id = L[r];
        const uint32_t gn = s_gapnd[id];
        const uint32_t gl = gn & 255u, ndl = gn >> 8;
        const uint8_t *krow = s_ktab + (int)s_dense[id] * K;
        const uint32_t rinc0 = gl + TINY_LEN1 + ndl * TINY_W1, rinc1 = ndl * gl;
        // column 0 (profile.h:2084-2090)
        uint32_t dg0 = s_w0[tid], dg1 = s_w1[tid];
        lf0 = dg0 + rinc0;
        lf1 = dg1 + rinc1;
        s_w0[tid] = lf0;
        s_w1[tid] = lf1;
        for (int c = 1; c <= l2; ++c) {
            const uint32_t cc = s_col[(c - 1) * TINY_THREADS + tid];
            const uint32_t up0 = s_w0[c * TINY_THREADS + tid], up1 = s_w1[c * TINY_THREADS + tid];
            const uint32_t gt = (cc >> 8) & 255u, ndt = cc >> 16;
            const uint32_t dlt = krow[cc & 255u];
            const uint32_t sa = (dg0 & TINY_DMASK) + dlt;
            const uint32_t f = (up0 & TINY_DMASK) + gl;
            const uint32_t e = (lf0 & TINY_DMASK) + gt;
            // min3switch (profile.h:1098-1099)
            const bool pF = f <= e;
            const uint32_t m1 = pF ? f : e;
            const bool pD = sa <= m1;
            uint32_t dn = ndl + ndt;
            dn = dn < 10u ? dn : 10u;
            uint32_t n0, n1;
            if (pD) { n0 = dg0 + dlt + TINY_LEN1 + dn * TINY_W1; n1 = dg1 + dn * dlt; }
            else if (pF) { n0 = up0 + rinc0; n1 = up1 + rinc1; }
            else { n0 = lf0 + gt + TINY_LEN1 + ndt * TINY_W1; n1 = lf1 + ndt * gt; }
            dg0 = up0; dg1 = up1;
            lf0 = n0; lf1 = n1;
            s_w0[c * TINY_THREADS + tid] = n0;
            s_w1[c * TINY_THREADS + tid] = n1;
        }
    }
    out[job.out_idx] = make_int4((int)(lf0 & TINY_DMASK), (int)lf1, (int)(lf0 >> 22), (int)((lf0 >> 15) & 127u));
}

// ---- tiny pairs, direction-bit variant ------------------------------------------------------------
// Same pairs as vsk_dp_tiny_kernel (<= 32 columns, <= 64 rows, full matrix), thread per pair, but like the
// long-pair path it keeps only the distance per cell and two direction bits (bit 1: not diagonal, bit 0:
// right rather than down; min3switch, profile.h:1098-1099), then walks the path back from the corner and
// sums what the reference's traceback sums (profile.h:2135-2210).  About 13 instead of 30 instructions per
// cell.  Previous row (u16: distances stay below 255 * 96) and column constants live in shared memory as
// [column][thread]; the direction words of the <= 64 rows (two u32 per row: columns 0-15 and 16-31) in
// the thread's local memory.
#define TD_THREADS 128

// DENSE: costs from the K x K table of the data set's compositions in shared memory; otherwise (more than 200
// distinct compositions) the column constant is the id itself and costs come from the global table.
template <bool DENSE>
__global__ void __launch_bounds__(TD_THREADS)
vsk_dp_tinydir_kernel(const uint16_t *__restrict__ left_ids, const uint16_t *__restrict__ top_ids,
                      const vsk_dp_job *__restrict__ jobs, int n_jobs, const uint8_t *__restrict__ dense,
                      const uint8_t *__restrict__ ktab, int K, const uint8_t *__restrict__ dist,
                      const uint16_t *__restrict__ gapnd, int4 *__restrict__ out) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint16_t *s_h = reinterpret_cast<uint16_t *>(smem);                    // [33][TD_THREADS]
    uint16_t *s_col = s_h + (TINY_MAXCOLS + 1) * TD_THREADS;               // [32][TD_THREADS]: dense index | gap << 8 (DENSE), else the id
    uint16_t *s_gapnd = s_col + TINY_MAXCOLS * TD_THREADS;                 // [1024]
    uint8_t *s_dense = reinterpret_cast<uint8_t *>(s_gapnd + 1024);        // [1024]
    uint8_t *s_ktab = s_dense + 1024;                                      // [K*K]
    const int tid = threadIdx.x;
    for (int i = tid; i < 1024; i += TD_THREADS) s_gapnd[i] = gapnd[i];
    if (DENSE) {
        for (int i = tid; i < 1024; i += TD_THREADS) s_dense[i] = dense[i];
        for (int i = tid; i < K * K; i += TD_THREADS) s_ktab[i] = ktab[i];
    }
    __syncthreads();
    const int ji = blockIdx.x * TD_THREADS + tid;
    if (ji >= n_jobs) return;
    const vsk_dp_job job = jobs[ji];
    const uint16_t *L = left_ids + job.left_off;
    const uint16_t *T = top_ids + job.top_off;
    const int l1 = job.l1, l2 = job.l2;
    const int n0 = min(l2, 16), n1 = l2 - n0;
    uint32_t dirs[TINY_MAXROWS][2];
    // row 0: cumulative gap costs (profile.h:2074-2079)
    int acc = 0;
    s_h[tid] = 0;
    for (int c = 0; c < l2; ++c) {
        const int id = T[c];
        const int g = s_gapnd[id] & 255;
        s_col[c * TD_THREADS + tid] = DENSE ? (uint16_t)(s_dense[id] | (g << 8)) : (uint16_t)id;
        acc += g;
        s_h[(c + 1) * TD_THREADS + tid] = (uint16_t)acc;
    }
    int lf = 0;
    for (int r = 0; r < l1; ++r) {
        const int id = L[r];
        const int gl = s_gapnd[id] & 255;
        const uint8_t *krow = DENSE ? (s_ktab + (int)s_dense[id] * K) : (dist + id * VSK_DROW);
        int dg = s_h[tid];
        lf = dg + gl;  // column 0 (profile.h:2084-2090)
        s_h[tid] = (uint16_t)lf;
        uint32_t w0 = 0, w1 = 0;
        for (int c = 1; c <= n0; ++c) {
            const int cc = s_col[(c - 1) * TD_THREADS + tid];
            const int up = s_h[c * TD_THREADS + tid];
            const int sa = dg + (DENSE ? (int)krow[cc & 255] : (int)__ldg(krow + cc)), f = up + gl;
            const int e = lf + (DENSE ? (cc >> 8) : (int)(s_gapnd[cc] & 255));
            const int d = min(min(sa, f), e);
            w0 = __funnelshift_l((uint32_t)(d - sa), w0, 1);
            w0 = __funnelshift_l((uint32_t)(e - f), w0, 1);
            dg = up;
            lf = d;
            s_h[c * TD_THREADS + tid] = (uint16_t)d;
        }
        for (int c = n0 + 1; c <= l2; ++c) {
            const int cc = s_col[(c - 1) * TD_THREADS + tid];
            const int up = s_h[c * TD_THREADS + tid];
            const int sa = dg + (DENSE ? (int)krow[cc & 255] : (int)__ldg(krow + cc)), f = up + gl;
            const int e = lf + (DENSE ? (cc >> 8) : (int)(s_gapnd[cc] & 255));
            const int d = min(min(sa, f), e);
            w1 = __funnelshift_l((uint32_t)(d - sa), w1, 1);
            w1 = __funnelshift_l((uint32_t)(e - f), w1, 1);
            dg = up;
            lf = d;
            s_h[c * TD_THREADS + tid] = (uint16_t)d;
        }
        dirs[r][0] = w0;
        dirs[r][1] = w1;
    }
    // traceback from (l1, l2)
    int r = l1, c = l2, p1 = 0, sumw = 0, len = 0;
    while (r > 0 && c > 0) {
        const int j = c - 1;  // column index; cell j of word 0 sits at bits 2*(n0-1-j), of word 1 at 2*(n1-1-(j-16))
        const uint32_t bits = (j < 16) ? (dirs[r - 1][0] >> (2 * (n0 - 1 - j))) & 3u : (dirs[r - 1][1] >> (2 * (n1 - 1 - (j - 16)))) & 3u;
        const int idl = L[r - 1], idt = T[c - 1];
        const int gnl = s_gapnd[idl], gnt = s_gapnd[idt];
        if (!(bits & 2u)) {
            int dn = (gnl >> 8) + (gnt >> 8);
            dn = dn < 10 ? dn : 10;
            p1 += dn * (DENSE ? (int)s_ktab[(int)s_dense[idl] * K + s_dense[idt]] : (int)__ldg(dist + idl * VSK_DROW + idt));
            sumw += dn;
            --r;
            --c;
        } else if (!(bits & 1u)) {
            p1 += (gnl >> 8) * (gnl & 255);
            sumw += gnl >> 8;
            --r;
        } else {
            p1 += (gnt >> 8) * (gnt & 255);
            sumw += gnt >> 8;
            --c;
        }
        ++len;
    }
    for (; r > 0; --r, ++len) {  // column 0: gaps in the top profile
        const int gnl = s_gapnd[L[r - 1]];
        p1 += (gnl >> 8) * (gnl & 255);
        sumw += gnl >> 8;
    }
    for (; c > 0; --c, ++len) {  // row 0: gaps in the left profile
        const int gnt = s_gapnd[T[c - 1]];
        p1 += (gnt >> 8) * (gnt & 255);
        sumw += gnt >> 8;
    }
    out[job.out_idx] = make_int4(lf, p1, sumw, len);
}

__global__ void vsk_id_presence_kernel(const uint16_t *__restrict__ ids, int64_t n, uint32_t *__restrict__ out) {
    __shared__ uint32_t bm[32];
    if (threadIdx.x < 32) bm[threadIdx.x] = 0;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int id = ids[i];
        const uint32_t bit = 1u << (id & 31);
        if (!(bm[id >> 5] & bit)) atomicOr(&bm[id >> 5], bit);
    }
    __syncthreads();
    if (threadIdx.x < 32 && bm[threadIdx.x]) atomicOr(&out[threadIdx.x], bm[threadIdx.x]);
}

// the same on a stream of the caller's, without waiting: the bits land in ctx->h_present_pin (page-locked, so
// that the call does not block) once the stream has got past the copy; the caller moves them to their place
int vsk_id_presence_async(vsk_ctx *ctx, const uint16_t *d_ids, int64_t n, vsk_idset *out, cudaStream_t st) {
    memset(out, 0, sizeof *out);
    if (!ctx->h_present_pin) VSK_CUDA(ctx, cudaHostAlloc((void **)&ctx->h_present_pin, sizeof(out->bits), cudaHostAllocDefault));
    memset(ctx->h_present_pin, 0, sizeof(out->bits));
    if (n <= 0) return VSK_OK;
    uint32_t *d = reinterpret_cast<uint32_t *>(ctx->d_queue) + 16;
    VSK_CUDA(ctx, cudaMemsetAsync(d, 0, sizeof(out->bits), st));
    const int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 8);
    vsk_id_presence_kernel<<<grid, 256, 0, st>>>(d_ids, n, d);
    ctx->launches++;
    VSK_CUDA(ctx, cudaGetLastError());
    VSK_CUDA(ctx, cudaMemcpyAsync(ctx->h_present_pin, d, sizeof(out->bits), cudaMemcpyDeviceToHost, st));
    return VSK_OK;
}

int vsk_id_presence(vsk_ctx *ctx, const uint16_t *d_ids, int64_t n, vsk_idset *out) {
    memset(out, 0, sizeof *out);
    if (n <= 0) return VSK_OK;
    uint32_t *d = reinterpret_cast<uint32_t *>(ctx->d_queue) + 16;  // 32 spare words of the counter block
    VSK_CUDA(ctx, cudaMemsetAsync(d, 0, sizeof(out->bits), ctx->stream));
    const int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 8);
    vsk_id_presence_kernel<<<grid, 256, 0, ctx->stream>>>(d_ids, n, d);
    ctx->launches++;
    VSK_CUDA(ctx, cudaGetLastError());
    VSK_CUDA(ctx, cudaMemcpyAsync(out->bits, d, sizeof(out->bits), cudaMemcpyDeviceToHost, ctx->stream));
    VSK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return VSK_OK;
}

// dense tables for the id set `ids` (cached in the context); returns K, or 0 when too large
static int tiny_tables(vsk_ctx *ctx, const vsk_idset *ids, int *K_out) {
    *K_out = 0;
    if (!ids) return VSK_OK;
    if (ctx->tiny_valid && memcmp(&ctx->tiny_key, ids, sizeof *ids) == 0) {
        *K_out = ctx->tiny_K;
        return VSK_OK;
    }
    std::vector<int> present;
    for (int id = 0; id < VSK_NIDS; ++id)
        if (ids->bits[id >> 5] & (1u << (id & 31))) present.push_back(id);
    const int K = (int)present.size();
    ctx->tiny_key = *ids;
    ctx->tiny_valid = true;
    ctx->tiny_K = 0;
    if (K == 0 || K > 200) return VSK_OK;  // K*K bytes must fit beside the row buffers in shared memory
    std::vector<uint8_t> dense(1024, 0), ktab((size_t)K * K);
    for (int a = 0; a < K; ++a) {
        dense[(size_t)present[(size_t)a]] = (uint8_t)a;
        for (int b = 0; b < K; ++b)
            ktab[(size_t)a * K + b] = ctx->h_dist[(size_t)present[(size_t)a] * VSK_DROW + present[(size_t)b]];
    }
    if (!ctx->d_tiny_dense) {
        VSK_CUDA(ctx, cudaMalloc(&ctx->d_tiny_dense, 1024));
        VSK_CUDA(ctx, cudaMalloc(&ctx->d_tiny_ktab, 200 * 200));
    }
    VSK_CUDA(ctx, cudaMemcpyAsync(ctx->d_tiny_dense, dense.data(), 1024, cudaMemcpyHostToDevice, ctx->stream));
    VSK_CUDA(ctx, cudaMemcpyAsync(ctx->d_tiny_ktab, ktab.data(), ktab.size(), cudaMemcpyHostToDevice, ctx->stream));
    VSK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->tiny_K = K;
    *K_out = K;
    return VSK_OK;
}

// kernel attributes are per device: one bit per device ordinal
static uint64_t g_dp_attr_set = 0;
static uint64_t g_tiny_attr_set = 0;
static uint64_t g_q8_attr_set = 0;

int vsk_dp_launch(vsk_ctx *ctx, const uint16_t *d_left_ids, const uint16_t *d_top_ids,
                  const vsk_dp_job *d_jobs, int64_t n_jobs, int64_t n_q8, int64_t n_tiny, int32_t max_l1,
                  const vsk_idset *ids, int4 *d_out) {
    if (n_jobs <= 0) return VSK_OK;
    if (n_jobs > 0x7fffffff) return vsk_fail(ctx, VSK_ERR_ARG, "too many DP pairs in one launch");
    int K = 0;
    const double h0 = vsk_now_ms();
    if (n_tiny > 0 || n_q8 > 0) {
        int rc = tiny_tables(ctx, ids, &K);
        if (rc) return rc;
    }
    const double h1 = vsk_now_ms();
    vsk_host_acc(ctx, "dp_tables", h1 - h0);
    static const bool use_dir = [] {
        const char *e = getenv("VSK_DP_DIR");
        return !(e && e[0] == '0');
    }();
    static const bool use_dir_q8 = [] {
        const char *e = getenv("VSK_DP_DIRQ8");
        return !(e && e[0] == '0');
    }();
    static const bool use_tinydir = [] {
        const char *e = getenv("VSK_DP_TINYDIR");
        return !(e && e[0] == '0');
    }();
    // K == 0: the data set uses more than 200 distinct compositions (no dense table).  The direction-bit small-pair
    // kernels then read the global table; the older sum-carrying kernels (run-time switches) need the dense table,
    // so with one of them selected everything goes through the warp-per-pair kernels.
    if (K == 0 && (!use_tinydir || !(use_dir && use_dir_q8))) n_tiny = n_q8 = 0;
    const int64_t n_warp = n_jobs - n_tiny - n_q8;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ctx->profile) {  // owned by the context from here on (a failure below does not leak them)
        VSK_CUDA(ctx, cudaEventCreate(&e0));
        ctx->prof_events.push_back(e0);
        if (cudaEventCreate(&e1) != cudaSuccess) {
            ctx->prof_events.pop_back();
            cudaEventDestroy(e0);
            return vsk_fail(ctx, VSK_ERR_CUDA, "cudaEventCreate failed");
        }
        ctx->prof_events.push_back(e1);
        VSK_CUDA(ctx, cudaEventRecord(e0, ctx->stream));
    }
    const bool dir_q8 = use_dir && use_dir_q8 && n_q8 > 0;  // the 8-lane jobs take the direction-bit path too
    const bool dir_any = use_dir && (n_warp > 0 || dir_q8);
    int64_t grid_w = 0;
    int stride_w = 0;
    if (n_warp > 0 || dir_any) {  // vsk_dp_kernel also takes the pairs the direction-bit path hands back
        if (!((g_dp_attr_set >> (ctx->device & 63)) & 1)) {
            VSK_CUDA(ctx, cudaFuncSetAttribute(vsk_dp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               (int)DP_SMEM_BYTES));
            g_dp_attr_set |= 1ull << (ctx->device & 63);
        }
        if (!ctx->occ_dp) {
            VSK_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctx->occ_dp, vsk_dp_kernel, DP_WARPS * 32,
                                                                         DP_SMEM_BYTES));
            if (ctx->occ_dp < 1) ctx->occ_dp = 1;
        }
        const int per_sm = ctx->occ_dp;
        int64_t grid = (int64_t)ctx->sm_count * per_sm;  // persistent: one wave, work pulled from a queue
        const int64_t need = dir_any ? grid : (n_warp + DP_WARPS - 1) / DP_WARPS;
        if (grid > need) grid = need;
        const int stride = ((max_l1 + 31) / 32) * 32;
        const size_t bytes = (size_t)grid * DP_WARPS * (stride + stride / 4) * sizeof(int4);  // boundary + row records
        if (bytes > ctx->dp_scratch_bytes) {
            if (ctx->d_dp_scratch) VSK_CUDA(ctx, cudaFree(ctx->d_dp_scratch));
            ctx->d_dp_scratch = nullptr;
            ctx->dp_scratch_bytes = 0;
            VSK_CUDA(ctx, cudaMalloc(&ctx->d_dp_scratch, bytes));
            ctx->dp_scratch_bytes = bytes;
        }
        grid_w = grid;
        stride_w = stride;
        if (dir_any) {  // the direction-bit path forks here and overlaps the kernels below
            const double hb = vsk_now_ms();
            int rc = vsk_dp_dir_begin(ctx, d_jobs, n_warp, dir_q8 ? n_q8 : 0);
            if (rc) return rc;
            vsk_host_acc(ctx, "dp_plan", vsk_now_ms() - hb);
        }
        if (n_warp > 0) {
            VSK_CUDA(ctx, cudaMemsetAsync(ctx->d_queue, 0, sizeof(int), ctx->stream));
            vsk_dp_kernel<<<(unsigned)grid, DP_WARPS * 32, DP_SMEM_BYTES, ctx->stream>>>(
                d_left_ids, d_top_ids, d_jobs, (int)n_warp, ctx->d_dist, ctx->d_gapnd,
                (int4 *)ctx->d_dp_scratch, stride, ctx->d_queue, d_out, nullptr, use_dir ? 1 : 0);
            ctx->launches++;
            VSK_CUDA(ctx, cudaGetLastError());
        }
    }
    auto small_pairs = [&]() -> int {
        if (n_q8 > 0 && !dir_q8) {
            const size_t smem = 2048 + 1024 + (size_t)K * K;
            if (!((g_q8_attr_set >> (ctx->device & 63)) & 1)) {
                VSK_CUDA(ctx, cudaFuncSetAttribute(vsk_dp_q8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                   2048 + 1024 + 200 * 200));
                g_q8_attr_set |= 1ull << (ctx->device & 63);
            }
            int per_sm = 0;
            VSK_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, vsk_dp_q8_kernel, Q8_WARPS * 32, smem));
            if (per_sm < 1) per_sm = 1;
            int64_t grid = (int64_t)ctx->sm_count * per_sm;
            const int64_t need = (n_q8 + Q8_WARPS * 4 - 1) / (Q8_WARPS * 4);
            if (grid > need) grid = need;
            VSK_CUDA(ctx, cudaMemsetAsync(ctx->d_queue + 1, 0, sizeof(int), ctx->stream));
            vsk_dp_q8_kernel<<<(unsigned)grid, Q8_WARPS * 32, smem, ctx->stream>>>(
                d_left_ids, d_top_ids, d_jobs + n_warp, (int)n_q8, ctx->d_tiny_dense, ctx->d_tiny_ktab, K, ctx->d_gapnd,
                ctx->d_queue + 1, d_out);
            ctx->launches++;
            VSK_CUDA(ctx, cudaGetLastError());
        }
        if (n_tiny > 0 && use_tinydir) {
            const size_t smem = (size_t)(2 * TINY_MAXCOLS + 1) * TD_THREADS * 2 + 2048 + 1024 + (size_t)K * K;
            static uint64_t attr_td = 0;
            if (!((attr_td >> (ctx->device & 63)) & 1)) {
                VSK_CUDA(ctx, cudaFuncSetAttribute(vsk_dp_tinydir_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                   (int)((2 * TINY_MAXCOLS + 1) * TD_THREADS * 2 + 2048 + 1024 + 200 * 200)));
                attr_td |= 1ull << (ctx->device & 63);
            }
            const unsigned gridt = (unsigned)((n_tiny + TD_THREADS - 1) / TD_THREADS);
            if (K > 0)
                vsk_dp_tinydir_kernel<true><<<gridt, TD_THREADS, smem, ctx->stream>>>(
                    d_left_ids, d_top_ids, d_jobs + n_warp + n_q8, (int)n_tiny, ctx->d_tiny_dense, ctx->d_tiny_ktab, K,
                    ctx->d_dist, ctx->d_gapnd, d_out);
            else
                vsk_dp_tinydir_kernel<false><<<gridt, TD_THREADS, smem, ctx->stream>>>(
                    d_left_ids, d_top_ids, d_jobs + n_warp + n_q8, (int)n_tiny, nullptr, nullptr, 0, ctx->d_dist,
                    ctx->d_gapnd, d_out);
            ctx->launches++;
            VSK_CUDA(ctx, cudaGetLastError());
        } else if (n_tiny > 0) {
            const size_t smem = (size_t)(2 * (TINY_MAXCOLS + 1) + TINY_MAXCOLS) * TINY_THREADS * 4 + 2048 + 1024 +
                                (size_t)K * K;
            if (!((g_tiny_attr_set >> (ctx->device & 63)) & 1)) {
                VSK_CUDA(ctx, cudaFuncSetAttribute(vsk_dp_tiny_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                   (int)((2 * (TINY_MAXCOLS + 1) + TINY_MAXCOLS) * TINY_THREADS * 4 +
                                                         2048 + 1024 + 200 * 200)));
                g_tiny_attr_set |= 1ull << (ctx->device & 63);
            }
            vsk_dp_tiny_kernel<<<(unsigned)((n_tiny + TINY_THREADS - 1) / TINY_THREADS), TINY_THREADS, smem, ctx->stream>>>(
                d_left_ids, d_top_ids, d_jobs + n_warp + n_q8, (int)n_tiny, ctx->d_tiny_dense, ctx->d_tiny_ktab, K,
                ctx->d_gapnd, d_out);
            ctx->launches++;
            VSK_CUDA(ctx, cudaGetLastError());
        }
        return VSK_OK;
    };
    // Where the small-pair kernels go relative to the sweeps of the direction-bit path (VSK_DP_SMALL_ORDER):
    // 0 = enqueued first on ctx->stream, running beside the first sweeps; 1 = the sweeps wait for them;
    // 2 = after the last sweep, beside the 8-lane kernels and traces that end the stage; 3 = after the join
    // (default).  Measured on the 1 M-read bench, DP stage: 224 / 227 / 222 / 217 ms -- the thread-per-pair
    // kernel and the sweep slow each other down when they share SMs, so they run one after the other.
    static const int small_order = [] {
        const char *e = getenv("VSK_DP_SMALL_ORDER");
        const int v = e ? atoi(e) : 3;
        return (v >= 0 && v <= 3) ? v : 3;
    }();
    if (!dir_any) {
        if (int rc = small_pairs()) return rc;
    } else {
        if (small_order <= 1) {
            if (int rc = small_pairs()) return rc;
            if (small_order == 1) VSK_CUDA(ctx, cudaEventRecord(ctx->dd_ev[0], ctx->stream));  // fork point moves behind them
        }
        vsk_dp_job *d_redo = nullptr;
        int *d_redo_n = nullptr;
        int rc = vsk_dp_dir_launch(ctx, d_left_ids, d_top_ids, d_jobs, n_warp, dir_q8 ? n_q8 : 0, K, max_l1, d_out,
                                   &d_redo, &d_redo_n);
        if (rc) return rc;
        if (small_order == 2) {
            if ((rc = vsk_dp_dir_wait_sweeps(ctx))) return rc;
            if ((rc = small_pairs())) return rc;
        }
        if ((rc = vsk_dp_dir_join(ctx))) return rc;
        if (small_order == 3 && (rc = small_pairs())) return rc;
        if (d_redo) {  // pairs whose corner the band does not reach
            VSK_CUDA(ctx, cudaMemsetAsync(ctx->d_queue, 0, sizeof(int), ctx->stream));
            vsk_dp_kernel<<<(unsigned)grid_w, DP_WARPS * 32, DP_SMEM_BYTES, ctx->stream>>>(
                d_left_ids, d_top_ids, d_redo, 0, ctx->d_dist, ctx->d_gapnd, (int4 *)ctx->d_dp_scratch, stride_w,
                ctx->d_queue, d_out, d_redo_n, 0);
            ctx->launches++;
            VSK_CUDA(ctx, cudaGetLastError());
        }
    }
    if (ctx->profile) {
        VSK_CUDA(ctx, cudaEventRecord(e1, ctx->stream));
    }
    vsk_host_acc(ctx, "dp_launch_total", vsk_now_ms() - h0);
    return VSK_OK;
}

// ---- pair DP for callers that must not allocate -------------------------------------
// Some hosts of this library cannot call cudaMalloc at the point where they need an
// alignment: a service that has handed the whole GPU to its own pool allocator, a
// latency-critical section between two of the caller's launches where the implicit
// device synchronisation of cudaMalloc/cudaFree is not acceptable, or a process whose
// memory budget was fixed at start-up.  vsk_pair_dp_static gives the results of
// vsk_pair_dp through vsk_dp_kernel alone (the one-kernel variant, which has no buffer of
// direction words to hold), and everything it needs (the boundary-column scratch of
// the column strips, the device copy of the pair descriptors, the work counter) lives
// in one file-scope device array of fixed size that every GPU loading libvsk.so holds
// until the process exits.
#define VSK_STATIC_CTAS 128

struct VskStaticScratch {
    // per warp: VSK_STATIC_MAX_ROWS boundary cells (int4) + as many row records (uint32)
    int4 boundary[(size_t)VSK_STATIC_CTAS * DP_WARPS * (VSK_STATIC_MAX_ROWS + VSK_STATIC_MAX_ROWS / 4)];  // 20 MiB
    vsk_dp_job jobs[VSK_STATIC_MAX_PAIRS];  // 1.25 MiB
    int queue[64];
};
static_assert(sizeof(vsk_dp_desc) == sizeof(vsk_dp_job), "public descriptor mirrors the kernel's job record");

static __device__ __align__(256) unsigned char vsk_dp_static_scratch[sizeof(VskStaticScratch)];

// In host code the name above is only a stand-in for the device array; the address a
// kernel can use has to be asked for, once per device, with that device current.
static VskStaticScratch *g_static_dev_addr[64];
static std::mutex g_static_mutex;

// Per-device set-up of vsk_pair_dp_static, done by vsk_init for the context's device (with that
// device current): looks up the device address of the fixed array and sets the kernel attribute.
// Under CUDA's lazy module loading this is also what makes the array and the kernel image
// resident, so after vsk_init the entry point itself neither loads nor allocates anything.
// Guarded: two host threads may create contexts (or make their first call) at the same time.
int vsk_dp_static_setup(vsk_ctx *ctx) {
    const int dev = ctx->device;
    if (dev < 0 || dev >= 64) return VSK_OK;  // the entry point refuses such a device itself
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (g_static_dev_addr[dev]) return VSK_OK;
    void *addr = nullptr;
    VSK_CUDA(ctx, cudaGetSymbolAddress(&addr, vsk_dp_static_scratch));
    VSK_CUDA(ctx, cudaFuncSetAttribute(vsk_dp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)DP_SMEM_BYTES));
    g_static_dev_addr[dev] = static_cast<VskStaticScratch *>(addr);
    return VSK_OK;
}

extern "C" int vsk_pair_dp_static(vsk_ctx *ctx, const uint16_t *d_left_ids, int64_t n_left_ids,
                                  const uint16_t *d_top_ids, int64_t n_top_ids, const vsk_dp_desc *pairs,
                                  int64_t n, vsk_dp_result *d_out) {
    if (!ctx || !d_left_ids || !d_top_ids || (n > 0 && (!pairs || !d_out))) return VSK_ERR_ARG;
    if (n < 0 || n > VSK_STATIC_MAX_PAIRS)
        return vsk_fail(ctx, VSK_ERR_ARG, "vsk_pair_dp_static: %lld pairs in one call, the fixed scratch holds %d",
                        (long long)n, VSK_STATIC_MAX_PAIRS);
    for (int64_t i = 0; i < n; ++i) {
        const vsk_dp_desc &p = pairs[i];
        if (p.l1 < 1 || p.l1 > VSK_STATIC_MAX_ROWS)
            return vsk_fail(ctx, VSK_ERR_ARG,
                            "vsk_pair_dp_static: pair %lld has %d rows, the fixed scratch holds %d per pair",
                            (long long)i, p.l1, VSK_STATIC_MAX_ROWS);
        if (p.l2 < 1 || p.l2 > VSK_MAXPROF)
            return vsk_fail(ctx, VSK_ERR_ARG, "vsk_pair_dp_static: pair %lld has %d columns (allowed 1..%d)",
                            (long long)i, p.l2, VSK_MAXPROF);
        if (p.left_off < 0 || p.left_off + p.l1 > n_left_ids || p.top_off < 0 || p.top_off + p.l2 > n_top_ids ||
            p.out_idx < 0 || p.out_idx >= n)
            return vsk_fail(ctx, VSK_ERR_ARG, "vsk_pair_dp_static: pair %lld is out of range", (long long)i);
        // a band must be the one vsk_band_params gives for these lengths (the kernel's row ranges rely on it)
        if (p.br >= 0 && (p.svector != (double)p.l2 / (double)p.l1 || p.br < 7 || p.br > 2 * VSK_MAXPROF))
            return vsk_fail(ctx, VSK_ERR_ARG,
                            "vsk_pair_dp_static: pair %lld: br/svector are not what vsk_band_params returns for %d x %d",
                            (long long)i, p.l1, p.l2);
    }
    if (n == 0) return VSK_OK;
    const int dev = ctx->device;
    if (dev < 0 || dev >= 64) return vsk_fail(ctx, VSK_ERR_ARG, "vsk_pair_dp_static: device %d not supported", dev);
    VSK_CUDA(ctx, cudaSetDevice(dev));
    if (!g_static_dev_addr[dev]) {  // vsk_init normally did this already
        const int rc = vsk_dp_static_setup(ctx);
        if (rc) return rc;
    }
    VskStaticScratch *S = g_static_dev_addr[dev];
    // the array keeps what the previous call left in it: the work counter must start at zero;
    // descriptors are overwritten for [0,n); boundary cells are always written before they are read
    VSK_CUDA(ctx, cudaMemsetAsync(S->queue, 0, sizeof(int), ctx->stream));
    VSK_CUDA(ctx, cudaMemcpyAsync(S->jobs, pairs, sizeof(vsk_dp_job) * (size_t)n, cudaMemcpyHostToDevice,
                                  ctx->stream));
    int64_t grid = (n + DP_WARPS - 1) / DP_WARPS;
    if (grid > VSK_STATIC_CTAS) grid = VSK_STATIC_CTAS;
    vsk_dp_kernel<<<(unsigned)grid, DP_WARPS * 32, DP_SMEM_BYTES, ctx->stream>>>(
        d_left_ids, d_top_ids, S->jobs, (int)n, ctx->d_dist, ctx->d_gapnd, S->boundary, VSK_STATIC_MAX_ROWS,
        S->queue, reinterpret_cast<int4 *>(d_out), nullptr, 0);
    ctx->launches++;
    VSK_CUDA(ctx, cudaGetLastError());
    VSK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return VSK_OK;
}

extern "C" int vsk_band_params(int32_t l1, int32_t l2, int homology, double *svector, int32_t *br) {
    if (l1 < 1 || l2 < 1 || !svector || !br) return VSK_ERR_ARG;
    int b = 0;
    vsk_band_host(l1, l2, homology, svector, &b);
    *br = b;
    return VSK_OK;
}
