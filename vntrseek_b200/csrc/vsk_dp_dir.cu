// Profile DP for the long pairs, in two kernels: a forward sweep that keeps only the distance
// per cell and writes the chosen direction (2 bits per cell) to HBM, and a trace that walks the
// directions back from the corner and sums dashnum*cost / dashnum along the path -- literally what
// GetFixedProfileAP / GetFixedProfileAPNarrowband do (reference src/psearch1.91/profile.h:2092-2210,
// 1828-2009), except that the matrix holds 2-bit directions instead of 8-byte cells.
//
// (For narrowband pairs the sweep covers a superset of the band and the trace verifies that the path
// stays inside it; see vsk_dp_trace_kernel.)
//
// Why: the one-kernel variant in vsk_dp.cu carries the two path sums with every cell, which is half
// of its ~20 instructions per cell and three quarters of its per-lane state.  Dropping them lets a
// lane own 16 columns (strips of 512) at fewer registers, i.e. fewer instructions per cell and less
// per-step overhead.  The trace costs L1+L2 dependent steps per pair, run one pair per thread.
//
// Direction words: one uint32 per (row, lane) of a strip = the CPL (8, 12 or 16 per strip, arranged per pair by
// dd_plan so that the strips fit the pair) cells of that lane; cell j holds "right" in bit 2(CPL-1-j) and,
// looked at only when that bit is clear, "down rather than diagonal" in bit 2(CPL-1-j)+1 (dd_row_cells).
// They are stored lane-major: the words
// of one lane (strip*32 + lane) for rows 0..l1-1 are contiguous (padded to 8 rows), because that is
// how the trace reads them -- a path stays in one lane's columns for about CPL rows, so eight of its
// steps share a 32-byte sector.  A lane parks its words in a 16-row ring in shared memory and every
// eighth step all lanes store their last complete group of 8 rows as one sector.
#include <algorithm>
#include <cstdlib>

#include <cub/cub.cuh>

#include "vsk_internal.cuh"

#define DD_WARPS 4
#define DD_TBYTES 8192
#define DD_MAXCPL 16
#define DD_FULL 0xffffffffu

struct DdWarpSmem {
    uint8_t T[DD_TBYTES];   // compact substitution table [na][nb]
    uint32_t bmL[32], bmT[32];
    uint16_t preT[32];
    uint32_t win[64];       // sliding window of row records, slot = row & 63 (steady phase)
    union {
        struct {                 // while the pair is set up
            uint8_t mapL[1024];  // composition id -> local row index
            uint16_t invL[256], invT[256];
        } su;
        uint32_t ring[32][16];   // during the sweep: the last 16 direction words of each lane, [lane][row & 15]
    } u;
};
#define DD_SMEM_BYTES (2048 + DD_WARPS * sizeof(DdWarpSmem))

// Column strips of a pair: nm strips of 32*cm columns (cm columns per lane), then -- if cl != 0 -- one last strip
// of 32*cl columns.  The cheapest arrangement in issued instructions per wavefront step (9 per cell + about 41
// that do not depend on the strip width): either equal strips of 8, 12 or 16 columns per lane, or strips of 16
// (or 12) with a narrower or wider last one, which cuts the padding of the last strip (about 5 % of the sweep
// on the bench's job mix).  `force` (tuning aid, VSK_DD_FORCE): 8|12|16 = equal strips of that width, 1 = equal
// strips only (the width still chosen per pair), 0 = free choice.
struct DdPlan {
    int cm, nm, cl;
};
#define DD_CELL_COST 9
#define DD_STEP_COST 41
__host__ __device__ inline DdPlan dd_plan(int l2, int force) {
    DdPlan p = {16, 1, 0};
    int best = 0x7fffffff;
#pragma unroll
    for (int c = 16; c >= 8; c -= 4) {
        const int n = (l2 + 32 * c - 1) / (32 * c);
        const int cost = n * (c * DD_CELL_COST + DD_STEP_COST);  // x (l1 + 31) steps either way
        if (force > 1 ? (c == force) : (cost < best)) {
            best = cost;
            p.cm = c;
            p.nm = n;
            p.cl = 0;
        }
    }
    if (force == 0) {
        for (int cm = 16; cm >= 12; cm -= 4) {
            const int w = 32 * cm;
            for (int k = 1; k * w < l2; ++k) {
                const int rem = l2 - k * w;
                for (int cl = 8; cl <= 16; cl += 4) {
                    if (32 * cl < rem) continue;
                    const int cost = k * (cm * DD_CELL_COST + DD_STEP_COST) + cl * DD_CELL_COST + DD_STEP_COST;
                    if (cl != cm && cost < best) {
                        best = cost;
                        p.cm = cm;
                        p.nm = k;
                        p.cl = cl;
                    }
                    break;  // the narrowest last strip that fits
                }
            }
        }
    }
    return p;
}
__host__ __device__ inline int dd_strips(const DdPlan &p) { return p.nm + (p.cl ? 1 : 0); }

// row record: bits 0..15 table row offset (local index * nb, or the id for the global table),
// 16..23 gap cost d(0,l)
__device__ __forceinline__ uint32_t dd_pack_row(int off, int gn) {
    return (uint32_t)off | ((uint32_t)(gn & 255) << 16);
}

// local alphabets, compact table, row records (same scheme as vsk_dp_kernel); returns nb, smemt
__device__ __forceinline__ bool dd_setup(const uint16_t *__restrict__ L, const uint16_t *__restrict__ Tp, int l1, int l2,
                                         const uint8_t *__restrict__ dist, const uint16_t *s_gapnd, DdWarpSmem &ws,
                                         uint32_t *rowrec, int &nb_out) {
    const int lane = threadIdx.x & 31;
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
        const int a = __shfl_up_sync(DD_FULL, xl, o);
        const int b = __shfl_up_sync(DD_FULL, xt, o);
        if (lane >= o) { xl += a; xt += b; }
    }
    const int na = __shfl_sync(DD_FULL, xl, 31), nb = __shfl_sync(DD_FULL, xt, 31);
    const bool smemt = (na <= 256) && (nb <= 256) && (na * nb <= DD_TBYTES);
    if (smemt) {
        ws.preT[lane] = (uint16_t)(xt - ct);
        uint32_t b = bl;
        int rk = xl - cl;
        while (b) {
            const int id = lane * 32 + __ffs(b) - 1;
            b &= b - 1;
            ws.u.su.mapL[id] = (uint8_t)rk;
            ws.u.su.invL[rk] = (uint16_t)id;
            ++rk;
        }
        b = bt;
        rk = xt - ct;
        while (b) {
            const int id = lane * 32 + __ffs(b) - 1;
            b &= b - 1;
            ws.u.su.invT[rk] = (uint16_t)id;
            ++rk;
        }
        __syncwarp();
#pragma unroll 4
        for (int a = 0; a < na; ++a) {
            const uint8_t *drow = dist + (int)ws.u.su.invL[a] * VSK_DROW;
            for (int c = lane; c < nb; c += 32) ws.T[a * nb + c] = __ldg(drow + ws.u.su.invT[c]);
        }
        __syncwarp();
    }
    for (int i = lane; i < l1; i += 32) {
        const int id = L[i];
        rowrec[i] = dd_pack_row(smemt ? (int)ws.u.su.mapL[id] * nb : id, s_gapnd[id]);
    }
    __syncwarp();
    nb_out = nb;
    return smemt;
}

// One row of a lane's CPL cells.  Hin = the previous row of these columns, Hout = this row (the two may be
// the same array).
template <int CPL, bool SMEMT>
__device__ __forceinline__ uint32_t dd_row_cells(const int (&Hin)[CPL], int (&Hout)[CPL], const int (&tb)[CPL],
                                                 const int (&gt)[CPL], int dgd, int lfd,
                                                 const uint8_t *__restrict__ trow, int gl) {
    uint32_t w = 0;
    const uint32_t sbase = SMEMT ? (uint32_t)__cvta_generic_to_shared(trow) : 0u;
#pragma unroll
    for (int j = 0; j < CPL; ++j) {
        const int upd = Hin[j];
        int dlt;
        if (SMEMT) {
            // the table address as a multiply-add: it issues on the FMA pipe, the ALU pipe is the busier one
            uint32_t a;
            asm("mad.lo.u32 %0, %1, 1, %2;" : "=r"(a) : "r"(tb[j]), "r"(sbase));
            asm("ld.shared.u8 %0, [%1];" : "=r"(dlt) : "r"(a));
        } else {
            dlt = (int)__ldg(trow + tb[j]);
        }
        const int sa = dgd + dlt;                     // diagonal
        const int t = __viaddmin_s32(upd, gl, sa);    // min(down, diagonal)
        const int d = __viaddmin_s32(lfd, gt[j], t);  // min(right, t)
        // min3switch (profile.h:1098-1099): diag if sa<=f && sa<=e; else down if f<=e; else right.
        // With t = min(f, sa): right iff e < t (then e < sa and e < f); otherwise down iff f < sa (then
        // f <= e), otherwise diagonal.  Bit 1 = sign of t - sa (f < sa), bit 0 = sign of d - t (e < t).
        w = __funnelshift_l((uint32_t)(t - sa), w, 1);
        w = __funnelshift_l((uint32_t)(d - t), w, 1);
        dgd = upd;
        Hout[j] = d;
        lfd = d;
    }
    return w;
}

// stores rows g0..g0+7 (g0 a multiple of 8) of this lane's ring as one 32-byte sector
__device__ __forceinline__ void dd_flush8(const DdWarpSmem &ws, int lane, uint32_t *__restrict__ dlane, int g0) {
    const uint4 a = *reinterpret_cast<const uint4 *>(&ws.u.ring[lane][g0 & 15]);
    const uint4 b = *reinterpret_cast<const uint4 *>(&ws.u.ring[lane][(g0 & 15) + 4]);
    uint4 *dst = reinterpret_cast<uint4 *>(dlane + g0);
    __stcs(dst, a);
    __stcs(dst + 1, b);
}

// One strip of a pair: the 32*CPL columns from column col0 on (strip number `strip` of the pair; first / last:
// the pair's first / last strip).  totd (gap costs of the columns so far) and prev_r_hi (rows the previous strip
// left in the boundary scratch) pass from strip to strip.
template <int CPL, bool SMEMT>
__device__ __forceinline__ void dd_pair(const vsk_dp_job &job, const uint16_t *__restrict__ Tp,
                                        const uint8_t *__restrict__ dist, const uint16_t *s_gapnd, DdWarpSmem &ws,
                                        int nb, int *bscr, const uint32_t *rowrec, uint32_t *__restrict__ dir,
                                        int4 *out, const int col0, const int strip, const bool first,
                                        const bool last, int &totd, int &prev_r_hi) {
    const int lane = threadIdx.x & 31;
    const int l1 = job.l1, l2 = job.l2, br = job.br;
    const bool banded = br >= 0;
    const double sv = job.svector;
    int Hd[CPL];

    {
        const int cbase = col0 + lane * CPL;
        int tb[CPL], gt[CPL];
        int sd = 0;
#pragma unroll
        for (int j = 0; j < CPL; ++j) {
            const int c = cbase + j;
            const int id = (c < l2) ? (int)Tp[c] : 0;
            const int g = s_gapnd[id] & 255;
            gt[j] = g;
            if (SMEMT) {
                const int w = id >> 5;
                tb[j] = ws.preT[w] + __popc(ws.bmT[w] & ((1u << (id & 31)) - 1u));
            } else {
                tb[j] = id;
            }
            sd += g;
            Hd[j] = sd;
        }
        int xd = sd;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int a = __shfl_up_sync(DD_FULL, xd, o);
            if (lane >= o) xd += a;
        }
        const int ed = xd - sd + totd;
#pragma unroll
        for (int j = 0; j < CPL; ++j) {
            Hd[j] += ed;
            if (banded && cbase + j + 1 > br) Hd[j] = VSK_BIG;  // profile.h:1801-1802
        }
        int dgd = ed;
        if (banded && cbase > br) dgd = VSK_BIG;
        totd += __shfl_sync(DD_FULL, xd, 31);

        // rows whose band can touch this strip (conservative; the per-cell mask stays exact)
        int r_lo = 0, r_hi = l1;
        if (banded) {
            const int c0 = col0;
            const double a = (double)(c0 + 1 - br), b = (double)(min(l2, c0 + (32 * CPL)) + br + 1);
            r_lo = max(0, (int)floor(a / sv) - 2);
            r_hi = min(l1, (int)ceil(b / sv) + 2);
            if (r_hi < r_lo) r_hi = r_lo;
        }
        if (r_lo > 0) {
#pragma unroll
            for (int j = 0; j < CPL; ++j) Hd[j] = VSK_BIG;
            dgd = VSK_BIG;
        }
        const bool wr_scratch = (lane == 31) && !last;
        int nxt = VSK_BIG;
        if (lane == 0 && !first && r_lo < prev_r_hi) nxt = __ldcg(bscr + r_lo);
        uint32_t rec_next = __ldcg(rowrec + min(max(r_lo - lane, 0), l1 - 1));
        const int l1p = (l1 + 7) & ~7;
        uint32_t *dlane = dir + (size_t)(strip * 32 + lane) * l1p;  // this lane's words, row-indexed
        const int g_lo = r_lo & ~7;

        // a step with range checks: pipeline fill and drain (lanes whose row is outside [r_lo, r_hi) idle)
        auto step = [&](const int s) {
            const int r = s - lane;
            int lfd = __shfl_up_sync(DD_FULL, Hd[CPL - 1], 1);
            const bool active = (r >= r_lo) && (r < r_hi);
            const uint32_t rec = rec_next;
            rec_next = __ldcg(rowrec + min(max(r + 1, 0), l1 - 1));
            const int gl = (int)__byte_perm(rec, 0u, 0x4432);  // rec >> 16, kept out of the per-cell adds
            const uint8_t *trow = SMEMT ? (ws.T + (rec & 0xffffu)) : (dist + (rec & 0xffffu) * VSK_DROW);
            if (active) {
                if (lane == 0) {
                    if (first) {
                        lfd = dgd + gl;  // column 0, profile.h:2084-2090
                    } else {
                        lfd = nxt;
                        nxt = (r + 1 < prev_r_hi) ? __ldcg(bscr + r + 1) : VSK_BIG;
                    }
                }
                const uint32_t w = dd_row_cells<CPL, SMEMT>(Hd, Hd, tb, gt, dgd, lfd, trow, gl);
                dgd = lfd;
                ws.u.ring[lane][r & 15] = w;
                if (wr_scratch) bscr[r] = Hd[CPL - 1];
            }
            if ((s & 7) == 7) {  // the lane's last complete group of 8 rows goes out
                const int g0 = ((min(r, r_hi - 1) + 1) & ~7) - 8;
                if (g0 >= g_lo) dd_flush8(ws, lane, dlane, g0);
            }
        };
        // Steps r_lo .. r_hi+30.  The middle ones, where every lane has a row, run in the steady loop below in
        // blocks of 8 steps (two rows per trip, four trips, then the flush of the block's direction words); the
        // steady phase starts on a multiple of 8 steps, so that the flush and the window refill need no test
        // per step.  What is left over on either side goes through the checked step.
        const int s_fill_end = min(r_lo + 31, r_hi + 31);
        const int s_steady_begin = min((s_fill_end + 7) & ~7, r_hi + 31);
        int s = r_lo;
#pragma unroll 1
        for (int phase = 0; phase < 2; ++phase) {  // one copy of the checked step serves fill and drain
            const int lim = phase ? r_hi + 31 : s_steady_begin;
#pragma unroll 1
            for (; s < lim; ++s) step(s);
            if (phase == 0 && s + 8 <= r_hi) {
                // steady phase: every lane has a row in [r_lo, r_hi) and every step sends a row out; the
                // addresses advance by one row per step (the scratch has a row of slack past l1)
                int *bp = bscr + (s - lane);                   // lane 31 leaves its row's last cell here
                // Row records come through a 64-row window in shared memory and the left neighbours of
                // lane 0 through a register per lane, both fetched a block of 32 steps ahead, so that no
                // step waits for global memory.
                __syncwarp();
                ws.win[(s - 31 + lane) & 63] = __ldcg(rowrec + s - 31 + lane);
                ws.win[(s + 1 + lane) & 63] = __ldcg(rowrec + s + 1 + lane);
                int bcur = first ? 0 : __ldcg(bscr + s + 1 + lane), bpre = 0;
                uint32_t wpre = 0;
                __syncwarp();
                rec_next = ws.win[(s - lane) & 63];
                int k = 0;
                // one steady step: previous row of the lane's columns in Hin, this row into Hout.  The loop
                // runs it twice per trip with the two arrays swapping roles, so no row is copied.
                auto steady = [&](const int (&Hin)[CPL], int (&Hout)[CPL]) {
                    const int r = s - lane;
                    int lfd = __shfl_up_sync(DD_FULL, Hin[CPL - 1], 1);
                    const uint32_t rec = rec_next;
                    rec_next = ws.win[(r + 1) & 63];
                    const int gl = (int)__byte_perm(rec, 0u, 0x4432);  // rec >> 16, kept out of the per-cell adds
                    const uint8_t *trow = SMEMT ? (ws.T + (rec & 0xffffu)) : (dist + (rec & 0xffffu) * VSK_DROW);
                    const int v = __shfl_sync(DD_FULL, bcur, k & 31);  // previous strip's last cell of row s+1
                    if (lane == 0) lfd = first ? dgd + gl : nxt;       // column 0 (profile.h:2084-2090) : boundary
                    nxt = (s + 1 < prev_r_hi) ? v : VSK_BIG;
                    const uint32_t w = dd_row_cells<CPL, SMEMT>(Hin, Hout, tb, gt, dgd, lfd, trow, gl);
                    dgd = lfd;
                    ws.u.ring[lane][r & 15] = w;
                    if (wr_scratch) *bp = Hout[CPL - 1];
                    ++bp;
                    ++s;
                    ++k;
                };
                int Hx[CPL];
#pragma unroll 1
                while (s + 8 <= r_hi) {
                    if ((k & 31) == 0) {  // rows s+33 .. s+64: needed from step s+32 on
                        wpre = __ldcg(rowrec + s + 33 + lane);
                        if (!first) bpre = __ldcg(bscr + s + 33 + lane);
                    }
#pragma unroll 1
                    for (int q = 0; q < 4; ++q) {
                        steady(Hd, Hx);
                        steady(Hx, Hd);
                    }
                    {  // s is a multiple of 8 again: the lane's last complete group of 8 rows goes out
                        const int g0 = ((s - lane) & ~7) - 8;
                        if (g0 >= g_lo) dd_flush8(ws, lane, dlane, g0);
                    }
                    if ((k & 31) == 0) {  // rows <= s-32 are done with: their slots take rows s+1 .. s+32
                        __syncwarp();
                        ws.win[(s + 1 + lane) & 63] = wpre;
                        bcur = bpre;
                        __syncwarp();
                    }
                }
            }
        }
        if (r_hi > r_lo) {  // the rows after the last complete group (and that group once more)
            const int ga = (r_hi - 1) & ~7;
            if (ga - 8 >= g_lo) dd_flush8(ws, lane, dlane, ga - 8);
            dd_flush8(ws, lane, dlane, ga);
        }
        prev_r_hi = r_hi;
        __syncwarp();
    }
    if (last) {
        const int cl = (l2 - 1) - col0;
        if (lane == cl / CPL) {
            const int jr = cl % CPL;
#pragma unroll
            for (int j = 0; j < CPL; ++j)
                if (j == jr) out[job.out_idx] = make_int4(Hd[j], 0, 0, 0);
        }
    }
}

// forward sweep: jobs[0..n) with their direction-word offsets dir_off[] into dirbuf
__global__ void __maxnreg__(120)
vsk_dp_dir_kernel(const uint16_t *__restrict__ left_ids, const uint16_t *__restrict__ top_ids,
                  const vsk_dp_job *__restrict__ jobs, const int64_t *__restrict__ dir_off,
                  const int *__restrict__ bounds, int batch, int64_t batch_words, int job_end, int force,
                  const uint8_t *__restrict__ dist, const uint16_t *__restrict__ gapnd, int *scratch, int scratch_stride,
                  uint32_t *__restrict__ dirbuf, int *queue, int4 *out) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint16_t *s_gapnd = reinterpret_cast<uint16_t *>(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    DdWarpSmem &ws = reinterpret_cast<DdWarpSmem *>(smem + 2048)[warp];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) s_gapnd[i] = gapnd[i];
    __syncthreads();
    // per-warp scratch slice: scratch_stride boundary distances + scratch_stride row records
    int *my_bscr = scratch + (size_t)(blockIdx.x * DD_WARPS + warp) * 2 * scratch_stride;
    uint32_t *my_rowrec = reinterpret_cast<uint32_t *>(my_bscr + scratch_stride);
    const int lo_job = bounds[batch], hi_job = min(bounds[batch + 1], job_end);
    for (;;) {
        int ji = 0;
        if (lane == 0) ji = lo_job + atomicAdd(queue + batch, 1);
        ji = __shfl_sync(DD_FULL, ji, 0);
        if (ji >= hi_job) break;
        const vsk_dp_job job = jobs[ji];
        if (!VSK_DP_DIRJOB(job.l2)) continue;
        const uint16_t *L = left_ids + job.left_off;
        const uint16_t *Tp = top_ids + job.top_off;
        int nb = 0;
        const bool smemt = dd_setup(L, Tp, job.l1, job.l2, dist, s_gapnd, ws, my_rowrec, nb);
        uint32_t *dir = dirbuf + (dir_off[ji] - (int64_t)batch * batch_words);
        int totd = 0, prev_r_hi = job.l1;
        // strip by strip; ONE dispatch inside the loop, so that every instantiation of dd_pair exists once in
        // the kernel (code size).  The plan is recomputed per strip instead of being kept in
        // registers across the sweep (the kernel runs at its register limit).
        int ns = 1;
#pragma unroll 1
        for (int strip = 0; strip < ns; ++strip) {
            const DdPlan plan = dd_plan(job.l2, force);
            ns = dd_strips(plan);
            const bool tail = strip >= plan.nm;  // the last strip of a mixed plan
            const int cpl = tail ? plan.cl : plan.cm;
            const int col0 = tail ? plan.nm * 32 * plan.cm : strip * 32 * plan.cm;
            const bool sf = (strip == 0), sl = (strip == ns - 1);
#define DD_DISPATCH(C)                                                                                              \
    do {                                                                                                            \
        if (smemt)                                                                                                  \
            dd_pair<C, true>(job, Tp, dist, s_gapnd, ws, nb, my_bscr, my_rowrec, dir, out, col0, strip, sf, sl, totd, \
                             prev_r_hi);                                                                            \
        else                                                                                                        \
            dd_pair<C, false>(job, Tp, dist, s_gapnd, ws, nb, my_bscr, my_rowrec, dir, out, col0, strip, sf, sl,     \
                              totd, prev_r_hi);                                                                     \
    } while (0)
            if (cpl == 16) DD_DISPATCH(16);
            else if (cpl == 12) DD_DISPATCH(12);
            else DD_DISPATCH(8);
#undef DD_DISPATCH
            __syncwarp();
        }
    }
}

// ---- the same sweep for pairs of at most 128 columns: 8 lanes per pair -------------------------
// Four pairs per warp, each group of 8 lanes with its own job loop and sync mask (as
// vsk_dp_q8_kernel); one strip of 8*CPL columns (CPL 4/8/16 by the number of columns); costs from
// the dense K x K table of the data set in shared memory, so no per-pair set-up.  Direction words
// as above with 8 lanes instead of strips*32: word (row, lane) at lane*l1p + row.
#define DQ_WARPS 4
#define DQ_G 8

__host__ __device__ inline int dq_cpl(int l2) { return l2 <= 32 ? 4 : l2 <= 64 ? 8 : 16; }

// DENSE: costs from the K x K table of the data set's compositions in shared memory; otherwise (more than 200
// distinct compositions in the data set) from the symmetric table in global memory.
template <int CPL, bool DENSE>
__device__ __forceinline__ void dq_pair(const vsk_dp_job &job, const uint16_t *__restrict__ L,
                                        const uint16_t *__restrict__ Tp, const uint16_t *s_gapnd,
                                        const uint8_t *s_dense, const uint8_t *s_ktab, int K,
                                        const uint8_t *__restrict__ dist, uint32_t *ring,
                                        unsigned gmask, int glane, uint32_t *__restrict__ dir, int4 *out) {
    const int l1 = job.l1, l2 = job.l2, br = job.br;
    const bool banded = br >= 0;
    const int cbase = glane * CPL;
    int Hd[CPL], tb[CPL], gt[CPL];
    int sd = 0;
#pragma unroll
    for (int j = 0; j < CPL; ++j) {
        const int c = cbase + j;
        const int id = (c < l2) ? (int)Tp[c] : 0;
        gt[j] = s_gapnd[id] & 255;
        tb[j] = DENSE ? (int)s_dense[id] : id;
        sd += gt[j];
        Hd[j] = sd;
    }
    int xd = sd;
#pragma unroll
    for (int o = 1; o < DQ_G; o <<= 1) {
        const int a = __shfl_up_sync(gmask, xd, o, DQ_G);
        if (glane >= o) xd += a;
    }
    const int ed = xd - sd;
#pragma unroll
    for (int j = 0; j < CPL; ++j) {
        Hd[j] += ed;
        if (banded && cbase + j + 1 > br) Hd[j] = VSK_BIG;  // profile.h:1801-1802
    }
    int dgd = ed;
    if (banded && cbase > br) dgd = VSK_BIG;
    int id_next = L[min(max(-glane, 0), l1 - 1)];
    const int l1p = (l1 + 7) & ~7;
    uint32_t *dlane = dir + (size_t)glane * l1p;

    auto step = [&](const int s, const bool check) {
        const int r = s - glane;
        int lfd = __shfl_up_sync(gmask, Hd[CPL - 1], 1, DQ_G);
        const bool active = !check || ((r >= 0) && (r < l1));
        const int id = id_next;
        id_next = L[min(max(r + 1, 0), l1 - 1)];
        const int gl = s_gapnd[id] & 255;
        const uint8_t *trow = DENSE ? (s_ktab + (int)s_dense[id] * K) : (dist + id * VSK_DROW);
        if (active) {
            if (glane == 0) lfd = dgd + gl;  // column 0, profile.h:2084-2090
            const uint32_t w = dd_row_cells<CPL, DENSE>(Hd, Hd, tb, gt, dgd, lfd, trow, gl);
            dgd = lfd;
            ring[r & 15] = w;
        }
        if ((s & 7) == 7) {  // the lane's last complete group of 8 rows goes out as one sector
            const int g0 = ((min(r, l1 - 1) + 1) & ~7) - 8;
            if (g0 >= 0) {
                const uint4 a = *reinterpret_cast<const uint4 *>(ring + (g0 & 15));
                const uint4 b = *reinterpret_cast<const uint4 *>(ring + (g0 & 15) + 4);
                uint4 *dst = reinterpret_cast<uint4 *>(dlane + g0);
                __stcs(dst, a);
                __stcs(dst + 1, b);
            }
        }
    };
    const int s_fill_end = min(DQ_G - 1, l1 + DQ_G - 1);
    int s = 0;
    for (; s < s_fill_end; ++s) step(s, true);
    {  // steady phase (every lane of the group has a row): blocks of 8 steps from a multiple of 8 on, two rows per
       // trip with the two register rows swapping roles, the block's direction words flushed after it
        const int s_al = min((s + 7) & ~7, l1);
        for (; s < s_al; ++s) step(s, false);
        auto lean = [&](const int (&Hin)[CPL], int (&Hout)[CPL]) {
            const int r = s - glane;
            int lfd = __shfl_up_sync(gmask, Hin[CPL - 1], 1, DQ_G);
            const int id = id_next;
            id_next = L[min(r + 1, l1 - 1)];
            const int gl = s_gapnd[id] & 255;
            const uint8_t *trow = DENSE ? (s_ktab + (int)s_dense[id] * K) : (dist + id * VSK_DROW);
            if (glane == 0) lfd = dgd + gl;  // column 0, profile.h:2084-2090
            const uint32_t w = dd_row_cells<CPL, DENSE>(Hin, Hout, tb, gt, dgd, lfd, trow, gl);
            dgd = lfd;
            ring[r & 15] = w;
            ++s;
        };
        int Hx[CPL];
#pragma unroll 1
        while (s + 8 <= l1) {
#pragma unroll 1
            for (int q = 0; q < 4; ++q) {
                lean(Hd, Hx);
                lean(Hx, Hd);
            }
            const int g0 = ((s - glane) & ~7) - 8;
            if (g0 >= 0) {
                const uint4 a = *reinterpret_cast<const uint4 *>(ring + (g0 & 15));
                const uint4 b = *reinterpret_cast<const uint4 *>(ring + (g0 & 15) + 4);
                uint4 *dst = reinterpret_cast<uint4 *>(dlane + g0);
                __stcs(dst, a);
                __stcs(dst + 1, b);
            }
        }
    }
    for (; s < l1; ++s) step(s, false);
    for (; s < l1 + DQ_G - 1; ++s) step(s, true);
    {  // the rows after the last complete group (and that group once more)
        const int ga = (l1 - 1) & ~7;
#pragma unroll
        for (int g0 = ga - 8; g0 <= ga; g0 += 8) {
            if (g0 < 0) continue;
            const uint4 a = *reinterpret_cast<const uint4 *>(ring + (g0 & 15));
            const uint4 b = *reinterpret_cast<const uint4 *>(ring + (g0 & 15) + 4);
            uint4 *dst = reinterpret_cast<uint4 *>(dlane + g0);
            __stcs(dst, a);
            __stcs(dst + 1, b);
        }
    }
    const int cl = l2 - 1;
    if (glane == cl / CPL) {
        const int jr = cl % CPL;
#pragma unroll
        for (int j = 0; j < CPL; ++j)
            if (j == jr) out[job.out_idx] = make_int4(Hd[j], 0, 0, 0);
    }
}

// jobs[q_begin..q_end) are the 8-lane jobs; a batch takes the part of its range that lies in there
template <bool DENSE>
__global__ void __launch_bounds__(DQ_WARPS * 32)
vsk_dp_q8dir_kernel(const uint16_t *__restrict__ left_ids, const uint16_t *__restrict__ top_ids,
                    const vsk_dp_job *__restrict__ jobs, const int64_t *__restrict__ dir_off,
                    const int *__restrict__ bounds, int batch, int64_t batch_words, int q_begin, int q_end,
                    const uint8_t *__restrict__ dense, const uint8_t *__restrict__ ktab, int K,
                    const uint8_t *__restrict__ dist, const uint16_t *__restrict__ gapnd,
                    uint32_t *__restrict__ dirbuf, int *queue, int4 *out) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint16_t *s_gapnd = reinterpret_cast<uint16_t *>(smem);
    uint32_t *s_ring = reinterpret_cast<uint32_t *>(smem + 2048);  // [thread][16]
    uint8_t *s_dense = smem + 2048 + DQ_WARPS * 32 * 64;
    uint8_t *s_ktab = s_dense + 1024;
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) s_gapnd[i] = gapnd[i];
    if (DENSE) {
        for (int i = threadIdx.x; i < 1024; i += blockDim.x) s_dense[i] = dense[i];
        for (int i = threadIdx.x; i < K * K; i += blockDim.x) s_ktab[i] = ktab[i];
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, glane = lane & (DQ_G - 1);
    const unsigned gmask = 0xffu << (lane & ~(DQ_G - 1));
    uint32_t *ring = s_ring + threadIdx.x * 16;
    const int lo_job = max(bounds[batch], q_begin), hi_job = min(bounds[batch + 1], q_end);
    for (;;) {
        int ji = 0;
        if (glane == 0) ji = lo_job + atomicAdd(queue, 1);
        ji = __shfl_sync(gmask, ji, 0, DQ_G);
        if (ji >= hi_job) break;
        const vsk_dp_job job = jobs[ji];
        const uint16_t *L = left_ids + job.left_off;
        const uint16_t *Tp = top_ids + job.top_off;
        uint32_t *dir = dirbuf + (dir_off[ji] - (int64_t)batch * batch_words);
        if (job.l2 <= 32) dq_pair<4, DENSE>(job, L, Tp, s_gapnd, s_dense, s_ktab, K, dist, ring, gmask, glane, dir, out);
        else if (job.l2 <= 64) dq_pair<8, DENSE>(job, L, Tp, s_gapnd, s_dense, s_ktab, K, dist, ring, gmask, glane, dir, out);
        else dq_pair<16, DENSE>(job, L, Tp, s_gapnd, s_dense, s_ktab, K, dist, ring, gmask, glane, dir, out);
        __syncwarp(gmask);
    }
}

// Trace: one thread per pair walks the stored directions from (l1,l2) to (0,0) and sums what the
// reference's traceback sums (profile.h:2135-2210): dashnum*cost, dashnum, path length.
//
// Narrowband pairs: the forward sweep does not mask the cells outside the band (it only skips the
// rows of a strip the band cannot touch and closes row 0 past column br), i.e. it solves the DP on a
// region that contains the band.  If the path it finds stays inside the band, the banded DP finds
// the same path with the same values: restricting the region can only raise the value of a
// candidate that is off the path, and min3switch's preference order (diagonal, down, right) then
// picks the same candidate.  The trace checks every cell of the path against the band; a pair whose
// path leaves it -- or whose corner is unreachable (distance >= BIG) -- is handed back on `redo`
// and goes through vsk_dp_kernel, which masks cell by cell.
#define DT_THREADS 64
#define DT_AHEAD 24

__global__ void __launch_bounds__(DT_THREADS)
vsk_dp_trace_kernel(const uint16_t *__restrict__ left_ids, const uint16_t *__restrict__ top_ids,
                    const vsk_dp_job *__restrict__ jobs, const int64_t *__restrict__ dir_off,
                    const int *__restrict__ bounds, int batch, int64_t batch_words, int force,
                    const uint8_t *__restrict__ dist, const uint16_t *__restrict__ gapnd,
                    const uint32_t *__restrict__ dirbuf, int4 *out, vsk_dp_job *redo, int *redo_count, int q_begin) {
    __shared__ uint16_t s_gapnd[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) s_gapnd[i] = gapnd[i];
    __syncthreads();
    const int lo_job = bounds[batch], hi_job = bounds[batch + 1];
    for (int ji = lo_job + blockIdx.x * blockDim.x + threadIdx.x; ji < hi_job; ji += gridDim.x * blockDim.x) {
        const int64_t off0 = dir_off[ji];
        if (dir_off[ji + 1] == off0) continue;  // not a job of this path
        const vsk_dp_job job = jobs[ji];
        const int d = out[job.out_idx].x;
        if (d >= VSK_BIG) {
            redo[atomicAdd(redo_count, 1)] = job;
            continue;
        }
        const uint16_t *L = left_ids + job.left_off;
        const uint16_t *Tp = top_ids + job.top_off;
        const uint32_t *dir = dirbuf + (off0 - (int64_t)batch * batch_words);
        const int l1p = (job.l1 + 7) & ~7;
        // columns [0, main_cols) sit in lanes of cm columns, the rest (the last strip of a mixed plan) in lanes of
        // cl columns that start at lane index main_lanes
        int cm, cl = 0, main_cols = 0x7fffffff, main_lanes = 0x7fffffff;
        if (ji >= q_begin) {
            cm = dq_cpl(job.l2);  // 8-lane job
        } else {
            const DdPlan plan = dd_plan(job.l2, force);
            cm = plan.cm;
            if (plan.cl) {
                cl = plan.cl;
                main_lanes = plan.nm * 32;
                main_cols = main_lanes * cm;
            }
        }
        int r = job.l1, c = job.l2;
        int p1 = 0, sumw = 0, len = 0;
        // position of column c-1 inside the strips: strip*32 + lane, cell
        int cpl, sl, j;
        if (c - 1 < main_cols) {
            cpl = cm;
            sl = (c - 1) / cm;
            j = (c - 1) - sl * cm;
        } else {
            cpl = cl;
            sl = (c - 1 - main_cols) / cl;
            j = (c - 1 - main_cols) - sl * cl;
            sl += main_lanes;
        }
        uint32_t w = 0;
        int wr = -1, wl = -1;  // (row, strip*32+lane) of the cached word
        const bool banded = job.br >= 0;
        int blo = 1, bhi = job.l2;  // in-band columns of row r (profile.h:1835-1839)
        bool off_band = false;
        while (r > 0 && c > 0) {
            if (wr != r) {
                if (banded) {
                    const int center = __double2int_rz(__dmul_rn((double)r, job.svector));
                    blo = max(1, center - job.br);
                    bhi = min(job.l2, center + job.br);
                }
            }
            if (c < blo || c > bhi) {
                off_band = true;
                break;
            }
            if (wr != r || wl != sl) {
                const uint32_t *wp = dir + (size_t)sl * l1p + (r - 1);
                if (((r - 1) & 7) == 7 || wl != sl) {
                    // entering a new 32-byte sector (8 rows of one lane): the path runs roughly
                    // diagonally, so pull the sector it will most likely need DT_AHEAD rows from now
                    // into L2 while this one is in flight
                    const int pr = r - 1 - DT_AHEAD;
                    if (pr >= 0) {
                        const int pcol = max(c - 1 - DT_AHEAD, 0);
                        const int pc = pcol < main_cols ? pcol / cm : main_lanes + (pcol - main_cols) / cl;
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(dir + (size_t)pc * l1p + pr));
                    }
                }
                w = __ldg(wp);
                wr = r;
                wl = sl;
            }
            const uint32_t bits = (w >> (2 * (cpl - 1 - j))) & 3u;  // bit 0: right; else bit 1: down; else diagonal
            const int idl = L[r - 1], idt = Tp[c - 1];
            const int gnl = s_gapnd[idl], gnt = s_gapnd[idt];
            bool left = false;
            if (bits == 0u) {
                int dn = (gnl >> 8) + (gnt >> 8);
                dn = dn < 10 ? dn : 10;
                p1 += dn * (int)__ldg(dist + idl * VSK_DROW + idt);
                sumw += dn;
                --r;
                left = true;
            } else if (bits == 2u) {
                p1 += (gnl >> 8) * (gnl & 255);
                sumw += gnl >> 8;
                --r;
            } else {
                p1 += (gnt >> 8) * (gnt & 255);
                sumw += gnt >> 8;
                left = true;
            }
            if (left) {
                --c;
                if (--j < 0) {
                    --sl;
                    if (sl < main_lanes) cpl = cm;  // out of the last strip of a mixed plan
                    j = cpl - 1;
                }
            }
            ++len;
        }
        if (off_band || (banded && r == 0 && c > job.br)) {  // row 0 is open up to column br (profile.h:1801-1802)
            redo[atomicAdd(redo_count, 1)] = job;
            continue;
        }
        for (; r > 0; --r, ++len) {  // column 0: gaps in the top profile
            const int gnl = s_gapnd[L[r - 1]];
            p1 += (gnl >> 8) * (gnl & 255);
            sumw += gnl >> 8;
        }
        for (; c > 0; --c, ++len) {  // row 0: gaps in the left profile
            const int gnt = s_gapnd[Tp[c - 1]];
            p1 += (gnt >> 8) * (gnt & 255);
            sumw += gnt >> 8;
        }
        out[job.out_idx] = make_int4(d, p1, sumw, len);
    }
}

// direction words a job needs (0 = not a job for this path); jobs[q_begin..n) are the 8-lane jobs
__global__ void vsk_dd_words_kernel(const vsk_dp_job *__restrict__ jobs, int n, int q_begin, int force,
                                    int64_t *__restrict__ words) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int l1 = jobs[i].l1, l2 = jobs[i].l2;
    const int64_t l1p = (l1 + 7) & ~7;
    if (i >= q_begin) {
        words[i] = DQ_G * l1p;
        return;
    }
    words[i] = VSK_DP_DIRJOB(l2) ? (int64_t)dd_strips(dd_plan(l2, force)) * 32 * l1p : 0;
}

// bounds[b] = first job whose direction words start at or after b*batch_words (jobs of a batch are
// [bounds[b], bounds[b+1])); bounds[nb] = n
__global__ void vsk_dd_bounds_kernel(const int64_t *__restrict__ off, int n, int64_t batch_words, int nb,
                                     int *__restrict__ bounds) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b > nb) return;
    if (b == nb) {
        bounds[b] = n;
        return;
    }
    const int64_t want = (int64_t)b * batch_words;
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (off[mid] >= want) hi = mid;
        else lo = mid + 1;
    }
    bounds[b] = lo;
}

static uint64_t g_dd_attr_set = 0;  // one bit per device ordinal

// grow-only buffer k of the context
static int dd_reserve(vsk_ctx *ctx, int k, size_t bytes, void **out) {
    if (bytes > ctx->dd_cap[k]) {
        if (ctx->dd_mem[k]) VSK_CUDA(ctx, cudaFree(ctx->dd_mem[k]));  // waits for the device
        ctx->dd_mem[k] = nullptr;
        ctx->dd_cap[k] = 0;
        const size_t want = bytes + bytes / 8 + 256;
        VSK_CUDA(ctx, cudaMalloc(&ctx->dd_mem[k], want));
        ctx->dd_cap[k] = want;
    }
    *out = ctx->dd_mem[k];
    return VSK_OK;
}

extern "C" int vsk_dd_plan(int32_t l2, int force, int32_t plan[3]) {
    if (!plan || l2 < 1 || !(force == 0 || force == 1 || force == 8 || force == 12 || force == 16)) return VSK_ERR_ARG;
    const DdPlan p = dd_plan(l2, force);
    plan[0] = p.cm;
    plan[1] = p.nm;
    plan[2] = p.cl;
    return VSK_OK;
}

static int dd_force() {  // tuning aid: VSK_DD_FORCE=8|12|16 fixes the columns per lane of the sweep, 1 = equal strips only
    static const int force = [] {
        const char *e = getenv("VSK_DD_FORCE");
        const int c = e ? atoi(e) : 0;
        return (c == 8 || c == 12 || c == 16 || c == 1) ? c : 0;
    }();
    return force;
}

// Sizes the direction words of the jobs (the one host round trip of the path) and marks the fork point.  It runs
// BEFORE the other DP kernels are enqueued: the device is idle for the few microseconds of the two small
// kernels, and from then on everything of the DP stage is queued without a host synchronisation, so a host
// thread that is late (other processes on the cores) cannot leave the device without work in mid-stage.
int vsk_dp_dir_begin(vsk_ctx *ctx, const vsk_dp_job *d_jobs, int64_t n_warp, int64_t n_q8) {
    for (int i = 0; i < 2; ++i)
        if (!ctx->dd_stream[i]) VSK_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->dd_stream[i], cudaStreamNonBlocking));
    for (int i = 0; i < 7; ++i)
        if (!ctx->dd_ev[i]) VSK_CUDA(ctx, cudaEventCreateWithFlags(&ctx->dd_ev[i], cudaEventDisableTiming));
    const int64_t n = n_warp + n_q8;
    ctx->dd_plan_n = n;
    ctx->dd_plan_total = ctx->dd_plan_off_q8 = 0;
    if (n > 0) {
        int rc;
        int64_t *d_words = nullptr, *d_off = nullptr;
        void *d_tmp = nullptr;
        size_t tmp_bytes = 0;
        cudaStream_t st = ctx->stream;
        VSK_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, d_words, d_off, (int)(n + 1), st));
        if ((rc = dd_reserve(ctx, 0, (size_t)(n + 1) * sizeof(int64_t), (void **)&d_words))) return rc;
        if ((rc = dd_reserve(ctx, 1, (size_t)(n + 1) * sizeof(int64_t), (void **)&d_off))) return rc;
        if ((rc = dd_reserve(ctx, 2, tmp_bytes + 16, &d_tmp))) return rc;
        VSK_CUDA(ctx, cudaMemsetAsync(d_words + n, 0, sizeof(int64_t), st));
        vsk_dd_words_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d_jobs, (int)n, (int)n_warp, dd_force(), d_words);
        ctx->launches++;
        VSK_CUDA(ctx, cub::DeviceScan::ExclusiveSum(d_tmp, tmp_bytes, d_words, d_off, (int)(n + 1), st));
        ctx->launches += 2;
        VSK_CUDA(ctx, cudaMemcpyAsync(&ctx->dd_plan_total, d_off + n, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
        VSK_CUDA(ctx, cudaMemcpyAsync(&ctx->dd_plan_off_q8, d_off + n_warp, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
        VSK_CUDA(ctx, cudaStreamSynchronize(st));
    }
    VSK_CUDA(ctx, cudaEventRecord(ctx->dd_ev[0], ctx->stream));
    return VSK_OK;
}

// Runs the jobs of d_jobs[0..n) that satisfy VSK_DP_DIRJOB through the two kernels above, in batches
// whose direction words fit the buffer; the batches alternate between two side streams (each with
// its own half of the buffer) that fork from the vsk_dp_dir_begin mark and join ctx->stream at the
// end.  *redo / *redo_count (device, owned by the context) receive the pairs that must go through
// vsk_dp_kernel; both stay null when no job qualified.
int vsk_dp_dir_launch(vsk_ctx *ctx, const uint16_t *d_left_ids, const uint16_t *d_top_ids, const vsk_dp_job *d_jobs,
                      int64_t n_warp, int64_t n_q8, int K, int32_t max_l1, int4 *d_out, vsk_dp_job **redo,
                      int **redo_count) {
    *redo = nullptr;
    *redo_count = nullptr;
    const int64_t n = n_warp + n_q8;  // the 8-lane jobs follow the warp-per-pair jobs
    if (n <= 0) return VSK_OK;
    cudaStream_t s0 = ctx->dd_stream[0], s1 = ctx->dd_stream[1];
    int rc;
    if (ctx->dd_plan_n != n) return vsk_fail(ctx, VSK_ERR_ARG, "vsk_dp_dir_launch without its vsk_dp_dir_begin");
    ctx->dd_plan_n = -1;
    const int force = dd_force();
    int64_t *d_off = (int64_t *)ctx->dd_mem[1];           // prefix sums of the jobs' direction words (vsk_dp_dir_begin)
    const int64_t total = ctx->dd_plan_total, off_q8 = ctx->dd_plan_off_q8;  // all words; words before the first 8-lane job
    VSK_CUDA(ctx, cudaStreamWaitEvent(s0, ctx->dd_ev[0], 0));
    ctx->dd_sweeps_marked = 0;
    ctx->dd_join_pending = false;
    if (total == 0) {
        VSK_CUDA(ctx, cudaEventRecord(ctx->dd_ev[2], s0));
        VSK_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->dd_ev[2], 0));
        return VSK_OK;
    }
    // batch size: VSK_DIR_BATCH_MB of direction words per stream (default 12 GiB), at most a sixth of the
    // memory that is free or already ours
    const double hs0 = vsk_now_ms();
    const int64_t max_job_words = (int64_t)((VSK_MAXPROF + 255) / 256) * ((max_l1 + 7) & ~7) * 32;  // strips are >= 256 columns
    // The free-memory query goes through the driver's global lock (slow when something else polls the driver), so
    // its answer is kept; it is asked again only when the buffer would have to grow.
    int64_t batch_words = 0, half_words = 0;
    int nb = 0, nhalf = 0;
    for (int attempt = 0; attempt < 2; ++attempt) {
        if (!ctx->dd_batch_words || attempt == 1) {
            int64_t batch_mb = 12288;
            if (const char *e = getenv("VSK_DIR_BATCH_MB")) batch_mb = std::max<int64_t>(16, atoll(e));
            size_t free_b = 0, total_b = 0;
            VSK_CUDA(ctx, cudaMemGetInfo(&free_b, &total_b));
            ctx->dd_batch_words = std::min<int64_t>(batch_mb << 18, (int64_t)((free_b + ctx->dd_cap[3]) / 6 / 4)) & ~(int64_t)7;
            attempt = 1;
        }
        batch_words = std::max<int64_t>(ctx->dd_batch_words, max_job_words);
        nb = (int)((total + batch_words - 1) / batch_words);
        half_words = (nb == 1) ? total : batch_words + max_job_words;
        nhalf = nb == 1 ? 1 : 2;
        if ((size_t)half_words * nhalf * 4 <= ctx->dd_cap[3]) break;  // fits the buffer we have
    }

    if (!((g_dd_attr_set >> (ctx->device & 63)) & 1)) {
        VSK_CUDA(ctx, cudaFuncSetAttribute(vsk_dp_dir_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)DD_SMEM_BYTES));
        g_dd_attr_set |= 1ull << (ctx->device & 63);
    }
    const size_t dd_smem = DD_SMEM_BYTES;
    if (!ctx->occ_dd) {
        VSK_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctx->occ_dd, vsk_dp_dir_kernel, DD_WARPS * 32,
                                                                     dd_smem));
        if (ctx->occ_dd < 1) ctx->occ_dd = 1;
    }
    const int grid = ctx->sm_count * ctx->occ_dd;
    // the 8-lane kernel
    const bool dense = K > 0;  // K == 0: no dense table for this data set, the 8-lane kernel reads the global table
    const size_t dq_smem = 2048 + DQ_WARPS * 32 * 64 + 1024 + (size_t)K * K;
    int grid8 = 0;
    if (n_q8 > 0) {
        static uint64_t attr8 = 0;
        if (!((attr8 >> (ctx->device & 63)) & 1)) {
            VSK_CUDA(ctx, cudaFuncSetAttribute(vsk_dp_q8dir_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               2048 + DQ_WARPS * 32 * 64 + 1024 + 200 * 200));
            attr8 |= 1ull << (ctx->device & 63);
        }
        // occupancy at the largest table (200 x 200): a smaller K could only raise it, the grid is persistent anyway
        int &per_sm8 = ctx->occ_dq[dense ? 1 : 0];
        if (!per_sm8) {
            const size_t worst = 2048 + DQ_WARPS * 32 * 64 + 1024 + (dense ? 200 * 200 : 0);
            if (dense)
                VSK_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm8, vsk_dp_q8dir_kernel<true>, DQ_WARPS * 32, worst));
            else
                VSK_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm8, vsk_dp_q8dir_kernel<false>, DQ_WARPS * 32, worst));
            if (per_sm8 < 1) per_sm8 = 1;
        }
        grid8 = ctx->sm_count * per_sm8;
    }
    const int stride = ((max_l1 + 31) / 32) * 32 + 128;  // slack: the steady phase fetches up to 96 rows past the last
    const size_t scr_ints = (size_t)grid * DD_WARPS * 2 * stride;

    uint32_t *d_dir = nullptr;
    int *d_small = nullptr, *d_scr = nullptr;
    if ((rc = dd_reserve(ctx, 3, (size_t)half_words * nhalf * 4, (void **)&d_dir))) return rc;
    if ((rc = dd_reserve(ctx, 4, scr_ints * 2 * sizeof(int), (void **)&d_scr))) return rc;
    if ((rc = dd_reserve(ctx, 5, (size_t)(3 * (nb + 1) + 1) * sizeof(int), (void **)&d_small))) return rc;
    if ((rc = dd_reserve(ctx, 6, (size_t)n * sizeof(vsk_dp_job), (void **)redo))) return rc;
    int *d_bounds = d_small, *d_queue = d_small + (nb + 1), *d_queue8 = d_small + 2 * (nb + 1);
    *redo_count = d_small + 3 * (nb + 1);
    VSK_CUDA(ctx, cudaMemsetAsync(d_queue, 0, (size_t)(2 * (nb + 1) + 1) * sizeof(int), s0));  // queue heads + redo count
    vsk_dd_bounds_kernel<<<(nb + 1 + 127) / 128, 128, 0, s0>>>(d_off, (int)n, batch_words, nb, d_bounds);
    ctx->launches++;
    VSK_CUDA(ctx, cudaEventRecord(ctx->dd_ev[1], s0));
    VSK_CUDA(ctx, cudaStreamWaitEvent(s1, ctx->dd_ev[1], 0));  // s1 starts after the batch bounds exist

    const int h_q8_first = (int)(off_q8 / batch_words);  // first batch with 8-lane jobs
    vsk_host_acc(ctx, "dp_setup", vsk_now_ms() - hs0);
    for (int b = 0; b < nb; ++b) {
        cudaStream_t st = (b & 1) ? s1 : s0;
        uint32_t *dir = d_dir + (size_t)(b & 1) * half_words;
        int *scr = d_scr + (size_t)(b & 1) * scr_ints;
        if (n_warp > 0 && b <= h_q8_first) {
            cudaEvent_t pe0 = nullptr, pe1 = nullptr;
            if (ctx->profile) {
                VSK_CUDA(ctx, cudaEventCreate(&pe0));
                VSK_CUDA(ctx, cudaEventCreate(&pe1));
                VSK_CUDA(ctx, cudaEventRecord(pe0, st));
            }
            vsk_dp_dir_kernel<<<grid, DD_WARPS * 32, dd_smem, st>>>(d_left_ids, d_top_ids, d_jobs, d_off, d_bounds, b,
                                                                    batch_words, (int)n_warp, force, ctx->d_dist,
                                                                    ctx->d_gapnd, scr, stride, dir, d_queue, d_out);
            ctx->launches++;
            if (ctx->profile) {
                VSK_CUDA(ctx, cudaEventRecord(pe1, st));
                ctx->prof_sweep_events.push_back(pe0);
                ctx->prof_sweep_events.push_back(pe1);
            }
            VSK_CUDA(ctx, cudaEventRecord(ctx->dd_ev[4 + (b & 1)], st));  // the stream's last sweep so far
            ctx->dd_sweeps_marked |= 1 << (b & 1);
        }
        if (n_q8 > 0 && b >= h_q8_first) {
            if (dense)
                vsk_dp_q8dir_kernel<true><<<grid8, DQ_WARPS * 32, dq_smem, st>>>(
                    d_left_ids, d_top_ids, d_jobs, d_off, d_bounds, b, batch_words, (int)n_warp, (int)n,
                    ctx->d_tiny_dense, ctx->d_tiny_ktab, K, ctx->d_dist, ctx->d_gapnd, dir, d_queue8 + b, d_out);
            else
                vsk_dp_q8dir_kernel<false><<<grid8, DQ_WARPS * 32, dq_smem, st>>>(
                    d_left_ids, d_top_ids, d_jobs, d_off, d_bounds, b, batch_words, (int)n_warp, (int)n, nullptr,
                    nullptr, 0, ctx->d_dist, ctx->d_gapnd, dir, d_queue8 + b, d_out);
            ctx->launches++;
        }
        vsk_dp_trace_kernel<<<ctx->sm_count * 16, DT_THREADS, 0, st>>>(d_left_ids, d_top_ids, d_jobs, d_off, d_bounds, b,
                                                                       batch_words, force, ctx->d_dist, ctx->d_gapnd,
                                                                       dir, d_out, *redo, *redo_count, (int)n_warp);
        ctx->launches++;
    }
    VSK_CUDA(ctx, cudaGetLastError());
    VSK_CUDA(ctx, cudaEventRecord(ctx->dd_ev[2], s0));
    VSK_CUDA(ctx, cudaEventRecord(ctx->dd_ev[3], s1));
    ctx->dd_join_pending = true;
    return VSK_OK;
}

// ctx->stream waits for the sweeps (not for the 8-lane kernels and the traces behind them)
int vsk_dp_dir_wait_sweeps(vsk_ctx *ctx) {
    for (int i = 0; i < 2; ++i)
        if (ctx->dd_sweeps_marked & (1 << i)) VSK_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->dd_ev[4 + i], 0));
    return VSK_OK;
}

// join: ctx->stream continues after both side streams
int vsk_dp_dir_join(vsk_ctx *ctx) {
    if (!ctx->dd_join_pending) return VSK_OK;
    ctx->dd_join_pending = false;
    VSK_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->dd_ev[2], 0));
    VSK_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->dd_ev[3], 0));
    return VSK_OK;
}
