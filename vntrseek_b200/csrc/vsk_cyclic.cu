// Cyclic profile alignment on the GPU: GetBestProfileAP (reference src/psearch1.91/profile.h:1586-1675)
// with its bounded aligner GetBoundedProfileAP (:1281-1585) and the path -> limits step PAPToLimit
// (:1231-1279).  No program of the reference calls it (SURVEY.md banner, point 1); it is the aligner
// VNTRseek's description means by "rotation-aware", built here as SURVEY 8(f) row 4 and reachable
// through vsk_pair_dp_cyclic only (PROCLU does not use it, as the reference's psearch does not).
//
// What is computed, per pair (left profile: l1 rows, top profile: l2 columns): the global alignment of
// the left profile against every rotation m of the top profile, i.e. against columns m+1..m+l2 of the
// doubled top profile, each restricted to the region between the paths already found for the two
// rotations that enclose m; the rotation with the smallest distanceweighted/(255*weights) wins, ties
// going to the rotation visited first (m = 0, then midpoints depth-first, left half first).
//
// Mapping: one warp per pair, rotations one after the other in the reference's order (each needs its
// predecessors' paths).  Inside a rotation the rows go top to bottom; a row's cells are taken 32 at a
// time, one per lane: with A[c] = min(diagonal, down) the row recurrence H[c] = min(A[c], H[c-1] + g[c])
// has the closed form H[c] = G[c] + min_{k<=c}(A[k] - G[k]) over the prefix sums G of the column gap
// costs, which is a prefix-minimum across the lanes (5 shuffles); the direction of a cell then follows
// from the three candidates exactly as min3switch orders them (profile.h:1098-1099).  Directions go to a
// byte matrix in global memory; lane 0 walks them back for the weighted sums and the path's limits.
#include <algorithm>
#include <vector>

#include "vsk_internal.cuh"

#define CY_WARPS 4
#define CY_SLOTS 16       // limit arrays alive at once: the recursion's depth + 3 (reference: paLOGDEPTH 16)
#define CY_VLN 600000000  // PA_VLN, profile.h:1091
#define CY_FULL 0xffffffffu

struct CyJob {
    int64_t left_off, top_off;
    int32_t l1, l2, out_idx, pad_;
};

struct CyWarpSmem {
    int G2[2 * VSK_CYCLIC_MAX + 2];        // prefix sums of the doubled top profile's gap costs
    int row[2][VSK_CYCLIC_MAX + 2];        // distances of the previous / current row, by column - start
    uint16_t top2[2 * VSK_CYCLIC_MAX];     // doubled top profile
    int8_t slot_of[VSK_CYCLIC_MAX + 1];    // rotation -> limits slot (-1 = none)
    int8_t free_slots[CY_SLOTS];
    int16_t stack_lo[32], stack_hi[32];
};

struct CyAp {
    int distance, dsum, sumw, length;
};

// One bounded alignment of rows 1..l1 against columns start+1..end of the doubled top profile.
// mins/maxs/own: limit arrays (int2: x = left-most, y = right-most column of a path in that row).
__device__ CyAp cy_bounded(const uint16_t *__restrict__ L, int l1, int start, int end, const int2 *mins,
                           const int2 *maxs, int2 *own, CyWarpSmem &ws, const uint8_t *__restrict__ dist,
                           const uint16_t *s_gapnd, uint8_t *__restrict__ dirm) {
    const int lane = threadIdx.x & 31;
    const int width = end - start + 1;
    int *prev = ws.row[0], *cur = ws.row[1];
    // row 0: cumulative gap costs (profile.h:1296-1317)
    int maxr = min(maxs[0].y, end);
    for (int c = start + lane; c <= maxr; c += 32) {
        prev[c - start] = ws.G2[c] - ws.G2[start];
        dirm[c - start] = (c == start) ? 0 : 3;
    }
    {
        const int nxt = min(maxs[1].y, end);
        for (int i = maxr + 1 + lane; i <= nxt; i += 32) prev[i - start] = CY_VLN;
    }
    for (int r = lane; r <= l1; r += 32) own[r] = make_int2(0x7fffffff, -1);
    __syncwarp();
    int colsum = 0;  // the start column while no row has been cut off on the left (profile.h:1322-1336)
    for (int r = 1; r <= l1; ++r) {
        const int l = L[r - 1];
        const int gl = s_gapnd[l] & 255;
        const uint8_t *drow = dist + l * VSK_DROW;
        const int minr = max(mins[r].x, start);
        maxr = min(maxs[r].y, end);
        uint8_t *drowm = dirm + (size_t)r * width;
        int c0, carry;
        if (minr > start) {  // left cut-off row: the cell before its first one is PA_VLN (profile.h:1348-1352)
            c0 = minr;
            carry = CY_VLN;
            if (lane == 0) cur[minr - 1 - start] = CY_VLN;
        } else {
            colsum += gl;
            c0 = start + 1;
            carry = colsum;
            if (lane == 0) {
                cur[0] = colsum;
                drowm[0] = 2;
            }
        }
        for (int base = c0; base <= maxr; base += 32) {
            const int c = base + lane;
            const bool valid = c <= maxr;
            int sa = 0, f = 0, g = 0, G = 0, A = 0x7fffffff;
            if (valid) {
                const int t = ws.top2[c - 1];
                G = ws.G2[c];
                g = G - ws.G2[c - 1];
                sa = prev[c - 1 - start] + (int)__ldg(drow + t);
                f = prev[c - start] + gl;
                A = min(sa, f);
            }
            int v = valid ? A - G : 0x7fffffff;  // H[c] - G[c] = min over k <= c of (A[k] - G[k])
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(CY_FULL, v, o);
                if (lane >= o) v = min(v, u);
            }
            v = min(v, carry - ws.G2[base - 1]);
            const int H = valid ? v + G : 0;
            int Hl = __shfl_up_sync(CY_FULL, H, 1);
            if (lane == 0) Hl = carry;
            if (valid) {
                const int e = Hl + g;
                cur[c - start] = H;
                drowm[c - start] = (sa <= f && sa <= e) ? 1 : (f <= e ? 2 : 3);
            }
            carry = __shfl_sync(CY_FULL, H, min(31, maxr - base));
        }
        if (r < l1) {  // cells the next row may look at beyond this row's end (profile.h:1410-1417)
            const int nxt = min(maxs[r + 1].y, end);
            for (int i = maxr + 1 + lane; i <= nxt; i += 32) cur[i - start] = CY_VLN;
        }
        __syncwarp();
        int *tmp = prev;
        prev = cur;
        cur = tmp;
    }
    CyAp ap;
    ap.distance = prev[end - start];
    ap.dsum = ap.sumw = ap.length = 0;
    __threadfence_block();
    __syncwarp();
    if (lane == 0) {  // traceback (profile.h:1429-1497) and the path's limits (PAPToLimit, :1231-1279)
        int r = l1, c = end;
        for (;;) {
            const int d = dirm[(size_t)r * width + (c - start)];
            int2 o = own[r];
            o.x = min(o.x, c);
            o.y = max(o.y, c);
            own[r] = o;
            if (d == 1) {
                const int lid = L[r - 1], t = ws.top2[c - 1];
                const int dn = min(10, (int)(s_gapnd[lid] >> 8) + (int)(s_gapnd[t] >> 8));
                ap.dsum += dn * (int)__ldg(dist + lid * VSK_DROW + t);
                ap.sumw += dn;
                --r;
                --c;
            } else if (d == 3) {
                const int gt = s_gapnd[ws.top2[c - 1]];
                ap.dsum += (gt >> 8) * (gt & 255);
                ap.sumw += gt >> 8;
                --c;
            } else if (d == 2) {
                const int gn = s_gapnd[L[r - 1]];
                ap.dsum += (gn >> 8) * (gn & 255);
                ap.sumw += gn >> 8;
                --r;
            } else {
                break;
            }
            ++ap.length;
        }
    }
    ap.dsum = __shfl_sync(CY_FULL, ap.dsum, 0);
    ap.sumw = __shfl_sync(CY_FULL, ap.sumw, 0);
    ap.length = __shfl_sync(CY_FULL, ap.length, 0);
    __threadfence_block();
    __syncwarp();
    return ap;
}

// ratio the reference compares (profile.h:1637-1640), in IEEE double without contraction
__device__ __forceinline__ double cy_ratio(const CyAp &a) {
    return __ddiv_rn((double)a.dsum, __dmul_rn(255.0, (double)a.sumw));
}

__global__ void __launch_bounds__(CY_WARPS * 32)
vsk_cyclic_kernel(const uint16_t *__restrict__ left_ids, const uint16_t *__restrict__ top_ids,
                  const CyJob *__restrict__ jobs, int n, const uint8_t *__restrict__ dist,
                  const uint16_t *__restrict__ gapnd, uint8_t *scratch, size_t dir_bytes, size_t lim_stride, int *queue,
                  vsk_cyclic_result *out) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint16_t *s_gapnd = reinterpret_cast<uint16_t *>(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    CyWarpSmem &ws = reinterpret_cast<CyWarpSmem *>(smem + 2048)[warp];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) s_gapnd[i] = gapnd[i];
    __syncthreads();
    const size_t per_warp = dir_bytes + (size_t)CY_SLOTS * lim_stride * sizeof(int2);
    uint8_t *my = scratch + (size_t)(blockIdx.x * CY_WARPS + warp) * per_warp;
    uint8_t *dirm = my;
    int2 *lims = reinterpret_cast<int2 *>(my + dir_bytes);
    for (;;) {
        int ji = 0;
        if (lane == 0) ji = atomicAdd(queue, 1);
        ji = __shfl_sync(CY_FULL, ji, 0);
        if (ji >= n) break;
        const CyJob job = jobs[ji];
        const uint16_t *L = left_ids + job.left_off;
        const uint16_t *Tp = top_ids + job.top_off;
        const int l1 = job.l1, l2 = job.l2;
        // doubled top profile (DoubleProfile, profile.h:1566-1584) and the prefix sums of its gap costs
        for (int i = lane; i < 2 * l2; i += 32) ws.top2[i] = Tp[i % l2];
        for (int i = lane; i <= l2; i += 32) ws.slot_of[i] = -1;
        __syncwarp();
        if (lane == 0) {
            int acc = 0;
            ws.G2[0] = 0;
            for (int i = 1; i <= 2 * l2; ++i) {
                acc += s_gapnd[ws.top2[i - 1]] & 255;
                ws.G2[i] = acc;
            }
            for (int k = 0; k < CY_SLOTS; ++k) ws.free_slots[k] = (int8_t)(CY_SLOTS - 1 - k);  // pop from the end
        }
        __syncwarp();
        int nfree = CY_SLOTS;
        auto slot_ptr = [&](int s) { return lims + (size_t)s * lim_stride; };
        auto take = [&]() { return (int)ws.free_slots[--nfree]; };
        auto give = [&](int s) {
            __syncwarp();
            if (lane == 0) ws.free_slots[nfree] = (int8_t)s;
            ++nfree;
            __syncwarp();
        };
        // rectangular limits for the first alignment (init_rectangular_limits, profile.h:1219-1229)
        const int s0 = take(), sN = take();
        for (int r = lane; r <= l1 + 1; r += 32) {
            slot_ptr(s0)[r] = make_int2(0, 0);
            slot_ptr(sN)[r] = make_int2(l2, l2);
        }
        __syncwarp();
        const int sown = take();
        CyAp best = cy_bounded(L, l1, 0, l2, slot_ptr(s0), slot_ptr(sN), slot_ptr(sown), ws, dist, s_gapnd, dirm);
        int best_start = 0;
        // the first path becomes the limits of rotation 0 and, one period to the right, of rotation l2
        for (int r = lane; r <= l1; r += 32) {
            const int2 o = slot_ptr(sown)[r];
            slot_ptr(sN)[r] = make_int2(o.x + l2, o.y + l2);
        }
        give(s0);
        if (lane == 0) {
            ws.slot_of[0] = (int8_t)sown;
            ws.slot_of[l2] = (int8_t)sN;
            ws.stack_lo[0] = 0;
            ws.stack_hi[0] = (int16_t)l2;
        }
        __syncwarp();
        int sp = 0;
        while (sp >= 0) {
            const int lo = ws.stack_lo[sp], hi = ws.stack_hi[sp];
            --sp;
            __syncwarp();
            if (hi - lo > 1) {
                const int m = (lo + hi) / 2;
                const int sm = take();
                const CyAp ap = cy_bounded(L, l1, m, m + l2, slot_ptr(ws.slot_of[lo]), slot_ptr(ws.slot_of[hi]),
                                           slot_ptr(sm), ws, dist, s_gapnd, dirm);
                if (cy_ratio(ap) < cy_ratio(best)) {
                    best = ap;
                    best_start = m;
                }
                if (lane == 0) {
                    ws.slot_of[m] = (int8_t)sm;
                    ws.stack_lo[sp + 1] = (int16_t)m;  // the left half is popped first
                    ws.stack_hi[sp + 1] = (int16_t)hi;
                    ws.stack_lo[sp + 2] = (int16_t)lo;
                    ws.stack_hi[sp + 2] = (int16_t)m;
                }
                sp += 2;
                __syncwarp();
            } else {  // removelimits(left), profile.h:1649
                const int s = ws.slot_of[lo];
                __syncwarp();
                if (lane == 0) ws.slot_of[lo] = -1;
                give(s);
            }
        }
        if (lane == 0) {
            vsk_cyclic_result r;
            r.distance = best.distance;
            r.dsum = best.dsum;
            r.sumw = best.sumw;
            r.length = best.length;
            r.start = best_start;
            r.pad_[0] = r.pad_[1] = r.pad_[2] = 0;
            out[job.out_idx] = r;
        }
        __syncwarp();
    }
}

extern "C" int vsk_pair_dp_cyclic(vsk_ctx *ctx, const vsk_profset *left_set, const vsk_profset *top_set, int64_t n,
                                  const int32_t *left_idx, const int32_t *top_idx, vsk_cyclic_result *out) {
    if (!ctx || !left_set || !top_set || n < 0 || (n > 0 && (!left_idx || !top_idx || !out))) return VSK_ERR_ARG;
    if (n == 0) return VSK_OK;
    if (n > 0x7fffffff) return vsk_fail(ctx, VSK_ERR_ARG, "too many pairs in one call");
    VSK_CUDA(ctx, cudaSetDevice(ctx->device));
    std::vector<CyJob> jobs((size_t)n);
    int max1 = 1, max2 = 1;
    for (int64_t i = 0; i < n; ++i) {
        const int32_t a = left_idx[i], b = top_idx[i];
        if (a < 0 || a >= left_set->n || b < 0 || b >= top_set->n)
            return vsk_fail(ctx, VSK_ERR_ARG, "pair %lld: profile index out of range", (long long)i);
        CyJob &j = jobs[(size_t)i];
        j.left_off = left_set->h_off[(size_t)a];
        j.top_off = top_set->h_off[(size_t)b];
        j.l1 = (int32_t)(left_set->h_off[(size_t)a + 1] - j.left_off);
        j.l2 = (int32_t)(top_set->h_off[(size_t)b + 1] - j.top_off);
        j.out_idx = (int32_t)i;
        j.pad_ = 0;
        if (j.l1 < 1 || j.l2 < 1 || j.l1 > VSK_CYCLIC_MAX || j.l2 > VSK_CYCLIC_MAX)
            return vsk_fail(ctx, VSK_ERR_ARG, "pair %lld: %d x %d columns (the cyclic aligner takes 1..%d)", (long long)i,
                            j.l1, j.l2, VSK_CYCLIC_MAX);
        max1 = std::max(max1, j.l1);
        max2 = std::max(max2, j.l2);
    }
    // heaviest first: l1 * l2 * log2(l2) cells
    std::stable_sort(jobs.begin(), jobs.end(), [](const CyJob &x, const CyJob &y) {
        return (int64_t)x.l1 * x.l2 * (32 - __builtin_clz((unsigned)x.l2)) > (int64_t)y.l1 * y.l2 * (32 - __builtin_clz((unsigned)y.l2));
    });
    const size_t smem = 2048 + CY_WARPS * sizeof(CyWarpSmem);
    static uint64_t attr_set = 0;
    if (!((attr_set >> (ctx->device & 63)) & 1)) {
        VSK_CUDA(ctx, cudaFuncSetAttribute(vsk_cyclic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set |= 1ull << (ctx->device & 63);
    }
    int per_sm = 0;
    VSK_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, vsk_cyclic_kernel, CY_WARPS * 32, smem));
    if (per_sm < 1) per_sm = 1;
    int64_t grid = (int64_t)ctx->sm_count * per_sm;
    grid = std::min<int64_t>(grid, (n + CY_WARPS - 1) / CY_WARPS);
    const size_t dir_bytes = (((size_t)(max1 + 1) * (size_t)(max2 + 1)) + 255) & ~(size_t)255;
    const size_t lim_stride = (size_t)((max1 + 2 + 31) & ~31);
    const size_t per_warp = dir_bytes + (size_t)CY_SLOTS * lim_stride * sizeof(int2);
    cudaStream_t st = ctx->stream;
    uint8_t *d_scr = nullptr;
    CyJob *d_jobs = nullptr;
    vsk_cyclic_result *d_out = nullptr;
    int *d_queue = nullptr;
    int rc = VSK_OK;
    auto cleanup = [&]() {
        for (void *p : {(void *)d_scr, (void *)d_jobs, (void *)d_out, (void *)d_queue})
            if (p) cudaFreeAsync(p, st);
    };
#define CY_TRY(call)                                                                              \
    do {                                                                                          \
        cudaError_t e__ = (call);                                                                 \
        if (e__ != cudaSuccess) {                                                                 \
            rc = vsk_fail(ctx, VSK_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
            cleanup();                                                                            \
            return rc;                                                                            \
        }                                                                                         \
    } while (0)
    CY_TRY(cudaMallocFromPoolAsync(&d_scr, per_warp * (size_t)grid * CY_WARPS, ctx->pool, st));
    CY_TRY(cudaMallocFromPoolAsync(&d_jobs, sizeof(CyJob) * (size_t)n, ctx->pool, st));
    CY_TRY(cudaMallocFromPoolAsync(&d_out, sizeof(vsk_cyclic_result) * (size_t)n, ctx->pool, st));
    CY_TRY(cudaMallocFromPoolAsync(&d_queue, sizeof(int), ctx->pool, st));
    CY_TRY(cudaMemcpyAsync(d_jobs, jobs.data(), sizeof(CyJob) * (size_t)n, cudaMemcpyHostToDevice, st));
    CY_TRY(cudaMemsetAsync(d_queue, 0, sizeof(int), st));
    vsk_cyclic_kernel<<<(unsigned)grid, CY_WARPS * 32, smem, st>>>(left_set->d_ids, top_set->d_ids, d_jobs, (int)n,
                                                                  ctx->d_dist, ctx->d_gapnd, d_scr, dir_bytes, lim_stride,
                                                                  d_queue, d_out);
    ctx->launches++;
    CY_TRY(cudaGetLastError());
    CY_TRY(cudaMemcpyAsync(out, d_out, sizeof(vsk_cyclic_result) * (size_t)n, cudaMemcpyDeviceToHost, st));
    CY_TRY(cudaStreamSynchronize(st));
#undef CY_TRY
    cleanup();
    return VSK_OK;
}
