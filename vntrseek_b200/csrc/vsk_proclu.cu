// PROCLU on the device: the per-read pipeline of psearch
// (doSearchSimilarities, reference src/psearch1.91/psearch.c:777-845;
//  _HCLUST_find_candidates hclust.h:273-494; _HCLUST_align_candidates hclust.h:497-857)
// as a sequence of batch stages over all reads of a chunk:
//
//   seeds    one warp per read profile: seed codes of the consensus -> bucket of the
//            reference index -> window of entries by pattern size (+ the one hit
//            past `hi`) -> (read profile, ref profile) keys
//   dedup    radix sort + unique of the keys (candsort + the "skip same ref" test,
//            psearch.c:101-114,820-821, hclust.h:522-526)
//   LCS      bit-parallel LCS of the two consensus strings against
//            min((int)(.85*shortest+.5), shortest-1)            (hclust.h:533-545)
//   DP       full or narrowband profile alignment, heaviest pairs first (vsk_dp.cu),
//            similarity = 1 - dsum/(255*sumw) >= cutoff/100      (hclust.h:550-593)
//   flanks   every (ref rotlist member x read rotlist member): orientation, two
//            edit distances, thresholds                          (hclust.h:604-697)
//   links    accepted (refkey, readkey, dir) records, in the order the reference
//            would have called ClusterBaseAddLink, back to the host.
#include <cub/cub.cuh>
#include <thrust/iterator/transform_output_iterator.h>

#include <algorithm>
#include <cstdlib>

#include "vsk_bits.cuh"
#include "vsk_host.hpp"

// ---------------------------------------------------------------- device images
struct DevTR {
    int32_t n = 0;
    int64_t *prof_off = nullptr, *cons_off = nullptr, *fl_off = nullptr;
    uint16_t *ids = nullptr;
    uint8_t *cons = nullptr, *fl = nullptr;
    int32_t *leftlen = nullptr, *rightlen = nullptr, *patlen = nullptr, *key = nullptr;
    int32_t *rot_off = nullptr, *rot_mem = nullptr;
    uint8_t *dir = nullptr, *master = nullptr;
    // consensus in bit-parallel form, 64 characters per block (built on the device after upload):
    // cons_planes[4*b + c] = bit i set iff character i of block b is "ACGT"[c];
    // cons_codes[2*b + k]  = 2-bit codes of characters 32k..32k+31 of block b
    int64_t *cons_blk = nullptr;  // [2n+1] first block of every profile's consensus
    uint64_t *cons_planes = nullptr, *cons_codes = nullptr;
    uint64_t *cons_cnt = nullptr;  // [2n] number of A, C, G, T in the consensus, 16 bits each
};

struct DevIndex {
    const uint64_t *codes[15];
    const int64_t *boff[15];
    int32_t ncodes[15];
    // lut[t][code >> lut_shift[t]] = first code with that prefix (VSK_LUT_BITS bits, one more element
    // at the end): the search for a code starts in its prefix's range instead of the whole array
    const int32_t *lut[15];
    int32_t lut_shift[15];
    const int16_t *e_pmin;
    const int32_t *e_prof, *e_first;
    const int64_t *e_gcum;  // exclusive prefix of the multiplicities over all entries (+1 element)
};

#define VSK_LUT_BITS 16

struct DevLuts {
    const int16_t *trange_of, *err_of;  // [5001]
    const int32_t *lcs_min;             // [VSK_MAXPROF+1]
    const int32_t *flank_maxerr;        // [VSK_MAXPROF+1]
};

struct vsk_reads {
    vsk_ctx *ctx = nullptr;
    const vsk_trset *host = nullptr;
    DevTR d;
    std::vector<void *> allocs;
    int64_t h2d_bytes = 0;
    vsk_idset present;  // composition ids that occur in this set
    mutable bool checked = false;  // the pattern-size / tuple-size checks passed (done once per upload)
    // the composition ids and the flanks (what the scan needs last: profile DP, flank test) may still be on
    // their way on the context's copy stream while the seed / LCS stages run; ctx->copy_done marks their arrival
    mutable bool deferred = false;
};

struct vsk_refs {
    vsk_ctx *ctx = nullptr;
    const vsk_trset *host = nullptr;
    vsk_reads *tr = nullptr;  // device image of the TR set
    DevIndex idx;
    DevLuts luts;
    std::vector<void *> allocs;
    int64_t total_hits = 0, entries = 0;
    double build_seconds = 0;
};

__constant__ int c_seed_off[15][32];
__constant__ int c_tsize[15];
__constant__ int c_nones[15];

template <typename T>
static int upload_vec(vsk_ctx *ctx, std::vector<void *> &allocs, const std::vector<T> &h, T **d,
                      int64_t *bytes = nullptr, size_t pad = 8) {
    VSK_CUDA(ctx, cudaMallocFromPoolAsync((void **)d, sizeof(T) * (h.size() + pad), ctx->pool, ctx->stream));
    allocs.push_back(*d);
    if (!h.empty())
        VSK_CUDA(ctx, cudaMemcpyAsync(*d, h.data(), sizeof(T) * h.size(), cudaMemcpyHostToDevice, ctx->stream));
    if (bytes) *bytes += (int64_t)(sizeof(T) * h.size());
    return VSK_OK;
}

__device__ __forceinline__ int base_code(uint8_t c) {
    return (c == 'C') ? 1 : (c == 'G') ? 2 : (c == 'T') ? 3 : 0;
}

// one warp per profile: consensus ASCII -> match planes + 2-bit codes, 64 characters per block
__global__ void __launch_bounds__(256) vsk_pack_consensus_kernel(DevTR t, int64_t nprof) {
    const int64_t p = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (p >= nprof) return;
    const uint8_t *seq = t.cons + t.cons_off[p];
    const int len = (int)(t.cons_off[p + 1] - t.cons_off[p]);
    const int64_t b0 = t.cons_blk[p];
    const int nblk = (len + 63) >> 6;
    uint64_t cnt = 0;
    for (int b = 0; b < nblk; ++b) {
        uint32_t pl[2][4];
        uint64_t code[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int i = b * 64 + h * 32 + lane;
            const int c = (i < len) ? base_code(seq[i]) : -1;
#pragma unroll
            for (int k = 0; k < 4; ++k) pl[h][k] = __ballot_sync(0xffffffffu, c == k);
            // 2-bit codes: lane l contributes bits 2l, 2l+1
            const uint32_t lo = __ballot_sync(0xffffffffu, c >= 0 && (c & 1)), hi = __ballot_sync(0xffffffffu, c >= 0 && (c & 2));
            uint64_t w = 0;
            for (int l = 0; l < 32; ++l) w |= ((uint64_t)((lo >> l) & 1u) | ((uint64_t)((hi >> l) & 1u) << 1)) << (2 * l);
            code[h] = w;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) cnt += (uint64_t)(__popc(pl[0][k]) + __popc(pl[1][k])) << (16 * k);
        if (lane < 4) t.cons_planes[4 * (b0 + b) + lane] = (uint64_t)pl[0][lane] | ((uint64_t)pl[1][lane] << 32);
        if (lane < 2) t.cons_codes[2 * (b0 + b) + lane] = code[lane];
    }
    if (lane == 0) t.cons_cnt[p] = cnt;
}

// ---- decoding a raw TR set on the device (vsk_trset_load_leb36_raw) --------------------------------
// What ReadProfileWithRC does per record after tokenising the line (profile.h:546-729): two LEB36
// characters per column -> composition id (LEB36ToInt, :786-807), the consensus = majority symbol of every
// column that is not dash-majority (getConcensusFromProfile, :280-322), and the flanks upper-cased with
// non-ACGT -> N, the left one reversed, the reverse-complement profile's flanks complemented and swapped
// (:667-695).  c_cons_sym[id] = 0..3 for A,C,G,T, 4 for a dash-majority column (first maximum wins).
__constant__ int8_t c_cons_sym[1024];

__device__ __forceinline__ int leb36_digit_dev(int c) { return (c < 65) ? (c & 0xF) : ((c & 0x1F) + 9); }

// warp per profile.  pass 0: conslen[p] = consensus length (and the id range check); pass 1: ids and
// consensus characters are written (cons_off known by then).
template <int PASS>
__global__ void __launch_bounds__(256)
vsk_leb36_decode_kernel(const uint8_t *__restrict__ text, const int64_t *__restrict__ raw_prof,
                        const int64_t *__restrict__ prof_off, const int64_t *__restrict__ cons_off, int64_t nprof,
                        uint16_t *__restrict__ ids, uint8_t *__restrict__ cons, int32_t *__restrict__ conslen,
                        int *__restrict__ err) {
    const int64_t p = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (p >= nprof) return;
    const uint8_t *src = text + raw_prof[p];
    const int64_t o = prof_off[p];
    const int len = (int)(prof_off[p + 1] - o);
    int64_t cpos = PASS ? cons_off[p] : 0;
    int total = 0;
    for (int base = 0; base < len; base += 32) {
        const int c = base + lane;
        int v = 0, sym = 4;
        if (c < len) {
            v = leb36_digit_dev(src[2 * c]) + 36 * leb36_digit_dev(src[2 * c + 1]);
            if (v < 0 || v > 1000) {
                *err = 1;
                v = 0;
            }
            sym = c_cons_sym[v];
        }
        const unsigned keep = __ballot_sync(0xffffffffu, sym != 4);
        if (PASS && c < len) {
            ids[o + c] = (uint16_t)v;
            if (sym != 4) cons[cpos + __popc(keep & ((1u << lane) - 1u))] = (uint8_t)"ACGT"[sym];
        }
        cpos += __popc(keep);
        total += __popc(keep);
    }
    if (!PASS && lane == 0) conslen[p] = total;
}

__device__ __forceinline__ uint8_t flank_clean(uint8_t c) {  // profile.h:667-685
    if (c >= 'a' && c <= 'z') c -= 32;
    return (c == 'A' || c == 'C' || c == 'G' || c == 'T') ? c : (uint8_t)'N';
}
__device__ __forceinline__ uint8_t flank_comp(uint8_t c) {
    return c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : c == 'T' ? 'A' : c;
}

// warp per record: the four flank strings of its two profiles
__global__ void __launch_bounds__(256)
vsk_flank_decode_kernel(const uint8_t *__restrict__ text, const int64_t *__restrict__ raw_flank,
                        const int32_t *__restrict__ raw_flen, const int64_t *__restrict__ fl_off, int64_t n,
                        uint8_t *__restrict__ fl) {
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (i >= n) return;
    const uint8_t *L = text + raw_flank[2 * i], *R = text + raw_flank[2 * i + 1];
    const int ll = raw_flen[2 * i], rl = raw_flen[2 * i + 1];
    uint8_t *fwd_l = fl + fl_off[4 * i], *fwd_r = fl + fl_off[4 * i + 1], *rc_l = fl + fl_off[4 * i + 2],
            *rc_r = fl + fl_off[4 * i + 3];
    for (int k = lane; k < ll; k += 32) {
        const uint8_t c = flank_clean(L[ll - 1 - k]);  // forward left = the file's left flank reversed
        fwd_l[k] = c;
        rc_r[k] = flank_comp(c);
    }
    for (int k = lane; k < rl; k += 32) {
        const uint8_t c = flank_clean(R[k]);
        fwd_r[k] = c;
        rc_l[k] = flank_comp(c);
    }
}

static int upload_raw(vsk_ctx *ctx, std::vector<void *> &allocs, const uint8_t *h, size_t n, uint8_t **d,
                      int64_t *bytes) {
    VSK_CUDA(ctx, cudaMallocFromPoolAsync((void **)d, n + 64, ctx->pool, ctx->stream));
    allocs.push_back(*d);
    if (n) VSK_CUDA(ctx, cudaMemcpyAsync(*d, h, n, cudaMemcpyHostToDevice, ctx->stream));
    if (bytes) *bytes += (int64_t)n;
    return VSK_OK;
}


// The kernels hold per-string state sized for VSK_MAXPROF columns / characters (look-up tables,
// carry arrays, the packed path-weight field): longer records are refused here, loudly, instead of
// being computed wrongly.  (The reference's own limit is MAXPROFILESIZE 5000 path steps, profile.h:1090.)
int vsk_trset_check_limits(vsk_ctx *ctx, const vsk_trset *ts) {
    for (int32_t i = 0; i < ts->n; ++i)
        for (int o = 0; o < 2; ++o) {
            const size_t p = (size_t)(2 * i + o);
            const int64_t plen = ts->prof_off[p + 1] - ts->prof_off[p];
            const int64_t clen = ts->cons_off[p + 1] - ts->cons_off[p];
            if (plen > VSK_MAXPROF || clen > VSK_MAXPROF)
                return vsk_fail(ctx, VSK_ERR_ARG, "profile %d has %lld columns (limit %d)", ts->key[(size_t)i],
                                (long long)std::max(plen, clen), VSK_MAXPROF);
            const int64_t ll = ts->fl_off[2 * p + 1] - ts->fl_off[2 * p], rl = ts->fl_off[2 * p + 2] - ts->fl_off[2 * p + 1];
            if (ll > VSK_MAXPROF || rl > VSK_MAXPROF || ts->leftlen[p] > VSK_MAXPROF || ts->rightlen[p] > VSK_MAXPROF)
                return vsk_fail(ctx, VSK_ERR_ARG, "profile %d has a flank of %lld characters (limit %d)",
                                ts->key[(size_t)i], (long long)std::max(ll, rl), VSK_MAXPROF);
        }
    return VSK_OK;
}

int vsk_trset_check_limits(vsk_ctx *ctx, const vsk_trset *ts);
// Uploads a raw TR set: the file's text goes to the device and is decoded there; the consensus lengths come
// back so that the host knows cons_off (first upload only).
static int decode_raw_trset(vsk_ctx *ctx, const vsk_trset *ts, vsk_reads *r) {
    static uint64_t sym_set = 0;  // per device
    if (!((sym_set >> (ctx->device & 63)) & 1)) {
        int8_t sym[1024];
        memset(sym, 4, sizeof sym);
        int id = 0;
        for (int a = 0; a <= 10; ++a)
            for (int c = 0; c <= 10; ++c)
                for (int g = 0; g <= 10; ++g)
                    for (int t = 0; t <= 10; ++t) {
                        const int d = 10 - a - c - g - t;
                        if (d < 0) continue;
                        const int cnt[5] = {a, c, g, t, d};
                        int sel = 0, mx = 0;  // getConcensusFromProfile, profile.h:291-298
                        for (int j = 0; j < 5; ++j)
                            if (cnt[j] > mx) { mx = cnt[j]; sel = j; }
                        sym[id++] = (int8_t)sel;
                    }
        VSK_CUDA(ctx, cudaMemcpyToSymbol(c_cons_sym, sym, sizeof sym));
        sym_set |= 1ull << (ctx->device & 63);
    }
    cudaStream_t st = ctx->stream;
    const int64_t np = 2 * (int64_t)ts->n;
    int rc;
    uint8_t *d_text = nullptr;
    int64_t *d_raw_prof = nullptr, *d_raw_flank = nullptr;
    int32_t *d_raw_flen = nullptr, *d_conslen = nullptr;
    int *d_err = nullptr;
    std::vector<void *> tmp;  // freed when the decode is done
    if ((rc = upload_raw(ctx, tmp, reinterpret_cast<const uint8_t *>(ts->raw_text.data()), ts->raw_text.size(), &d_text,
                         &r->h2d_bytes)) ||
        (rc = upload_vec(ctx, tmp, ts->raw_prof, &d_raw_prof, &r->h2d_bytes)) ||
        (rc = upload_vec(ctx, tmp, ts->raw_flank, &d_raw_flank, &r->h2d_bytes)) ||
        (rc = upload_vec(ctx, tmp, ts->raw_flen, &d_raw_flen, &r->h2d_bytes)) ||
        (rc = upload_vec(ctx, r->allocs, ts->prof_off, &r->d.prof_off, &r->h2d_bytes)) ||
        (rc = upload_vec(ctx, r->allocs, ts->fl_off, &r->d.fl_off, &r->h2d_bytes))) {
        for (void *p : tmp) cudaFreeAsync(p, st);
        return rc;
    }
    auto fail = [&](int code) {
        for (void *p : tmp) cudaFreeAsync(p, st);
        return code;
    };
#define RAW_CUDA(call)                                                                                          \
    do {                                                                                                        \
        cudaError_t e__ = (call);                                                                               \
        if (e__ != cudaSuccess)                                                                                 \
            return fail(vsk_fail(ctx, VSK_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__))); \
    } while (0)
    RAW_CUDA(cudaMallocFromPoolAsync((void **)&d_conslen, sizeof(int32_t) * (size_t)(np + 1), ctx->pool, st));
    tmp.push_back(d_conslen);
    RAW_CUDA(cudaMallocFromPoolAsync((void **)&d_err, sizeof(int), ctx->pool, st));
    tmp.push_back(d_err);
    RAW_CUDA(cudaMemsetAsync(d_err, 0, sizeof(int), st));
    RAW_CUDA(cudaMallocFromPoolAsync((void **)&r->d.ids, sizeof(uint16_t) * (size_t)(ts->prof_off.back() + 8), ctx->pool, st));
    r->allocs.push_back(r->d.ids);
    RAW_CUDA(cudaMallocFromPoolAsync((void **)&r->d.fl, (size_t)ts->fl_off.back() + 64, ctx->pool, st));
    r->allocs.push_back(r->d.fl);
    const unsigned gp = (unsigned)((np * 32 + 255) / 256), gr = (unsigned)(((int64_t)ts->n * 32 + 255) / 256);
    if (np) {
        vsk_leb36_decode_kernel<0><<<gp, 256, 0, st>>>(d_text, d_raw_prof, r->d.prof_off, nullptr, np, nullptr, nullptr,
                                                      d_conslen, d_err);
        vsk_flank_decode_kernel<<<gr, 256, 0, st>>>(d_text, d_raw_flank, d_raw_flen, r->d.fl_off, ts->n, r->d.fl);
        ctx->launches += 2;
    }
    std::vector<int32_t> conslen((size_t)np);
    int herr = 0;
    if (np) RAW_CUDA(cudaMemcpyAsync(conslen.data(), d_conslen, sizeof(int32_t) * (size_t)np, cudaMemcpyDeviceToHost, st));
    RAW_CUDA(cudaMemcpyAsync(&herr, d_err, sizeof(int), cudaMemcpyDeviceToHost, st));
    RAW_CUDA(cudaStreamSynchronize(st));
    if (herr) return fail(vsk_fail(ctx, VSK_ERR_ARG, "composition id outside 0..1000 in the profile file"));
    vsk_trset *mt = const_cast<vsk_trset *>(ts);
    if (!ts->raw_decoded) {
        for (int64_t p = 0; p < np; ++p) mt->cons_off[(size_t)p + 1] = mt->cons_off[(size_t)p] + conslen[(size_t)p];
        mt->raw_decoded = true;
    }
    if ((rc = upload_vec(ctx, r->allocs, ts->cons_off, &r->d.cons_off, &r->h2d_bytes))) return fail(rc);
    RAW_CUDA(cudaMallocFromPoolAsync((void **)&r->d.cons, (size_t)ts->cons_off.back() + 64, ctx->pool, st));
    r->allocs.push_back(r->d.cons);
    if (np) {
        vsk_leb36_decode_kernel<1><<<gp, 256, 0, st>>>(d_text, d_raw_prof, r->d.prof_off, r->d.cons_off, np, r->d.ids,
                                                      r->d.cons, nullptr, d_err);
        ctx->launches++;
    }
    RAW_CUDA(cudaGetLastError());
#undef RAW_CUDA
    for (void *p : tmp) cudaFreeAsync(p, st);
    return vsk_trset_check_limits(ctx, ts);  // the consensus lengths are known now
}

static int upload_trset(vsk_ctx *ctx, const vsk_trset *ts, vsk_reads **out, bool defer_big = false) {
    *out = nullptr;
    if (int lim = vsk_trset_check_limits(ctx, ts)) return lim;
    vsk_reads *r = new vsk_reads();
    r->ctx = ctx;
    r->host = ts;
    r->d.n = ts->n;
    int rc;
    if (!ts->rot_flat || (!ts->raw && !ts->blk_ready)) vsk_trset_prepare(const_cast<vsk_trset *>(ts));
#define UP(field, vec) \
    if ((rc = upload_vec(ctx, r->allocs, vec, &r->d.field, &r->h2d_bytes))) { *out = r; return rc; }
#define UPB(field, vec)                                                                                   \
    if ((rc = upload_raw(ctx, r->allocs, reinterpret_cast<const uint8_t *>((vec).data()), (vec).size(),    \
                         &r->d.field, &r->h2d_bytes))) { *out = r; return rc; }
    if (ts->raw) {
        if ((rc = decode_raw_trset(ctx, ts, r))) { *out = r; return rc; }
    } else {
        UP(prof_off, ts->prof_off) UP(cons_off, ts->cons_off) UP(fl_off, ts->fl_off)
        UPB(cons, ts->cons)
        if (!defer_big) {
            UP(ids, ts->ids) UPB(fl, ts->fl)
        } else {
            // ids and flanks: allocated in stream order here, copied on the copy stream, so that the copy (the
            // larger half of the set) overlaps the seed and LCS stages, which read neither
            if (!ctx->copy_stream) VSK_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
            if (!ctx->copy_done) VSK_CUDA(ctx, cudaEventCreateWithFlags(&ctx->copy_done, cudaEventDisableTiming));
            if (!ctx->copy_go) VSK_CUDA(ctx, cudaEventCreateWithFlags(&ctx->copy_go, cudaEventDisableTiming));
            VSK_CUDA(ctx, cudaMallocFromPoolAsync((void **)&r->d.ids, sizeof(uint16_t) * (ts->ids.size() + 8), ctx->pool, ctx->stream));
            r->allocs.push_back(r->d.ids);
            VSK_CUDA(ctx, cudaMallocFromPoolAsync((void **)&r->d.fl, ts->fl.size() + 64, ctx->pool, ctx->stream));
            r->allocs.push_back(r->d.fl);
            VSK_CUDA(ctx, cudaEventRecord(ctx->copy_go, ctx->stream));
            VSK_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->copy_go, 0));
            r->deferred = true;  // from here on the copy stream owns the two buffers until copy_done
            if (!ts->ids.empty())
                VSK_CUDA(ctx, cudaMemcpyAsync(r->d.ids, ts->ids.data(), sizeof(uint16_t) * ts->ids.size(), cudaMemcpyHostToDevice, ctx->copy_stream));
            if (!ts->fl.empty())
                VSK_CUDA(ctx, cudaMemcpyAsync(r->d.fl, ts->fl.data(), ts->fl.size(), cudaMemcpyHostToDevice, ctx->copy_stream));
            r->h2d_bytes += (int64_t)(sizeof(uint16_t) * ts->ids.size() + ts->fl.size());
            if ((rc = vsk_id_presence_async(ctx, r->d.ids, ts->prof_off.back(), &r->present, ctx->copy_stream))) { *out = r; return rc; }
            VSK_CUDA(ctx, cudaEventRecord(ctx->copy_done, ctx->copy_stream));
        }
    }
    UP(leftlen, ts->leftlen) UP(rightlen, ts->rightlen) UP(patlen, ts->patlen)
    UP(key, ts->key) UP(rot_off, ts->rot_off) UP(rot_mem, ts->rot_mem) UP(dir, ts->dir) UP(master, ts->master)
#undef UP
#undef UPB
    {
        const size_t np = (size_t)2 * ts->n;
        if (!ts->blk_ready) vsk_trset_prepare(const_cast<vsk_trset *>(ts));  // raw set: cons_off is known only now
        const int64_t nblk = ts->cons_blk[np];
        if ((rc = upload_vec(ctx, r->allocs, ts->cons_blk, &r->d.cons_blk, &r->h2d_bytes))) { *out = r; return rc; }
        VSK_CUDA(ctx, cudaMallocFromPoolAsync((void **)&r->d.cons_planes, sizeof(uint64_t) * 4 * (size_t)(nblk + 1), ctx->pool, ctx->stream));
        r->allocs.push_back(r->d.cons_planes);
        VSK_CUDA(ctx, cudaMallocFromPoolAsync((void **)&r->d.cons_codes, sizeof(uint64_t) * 2 * (size_t)(nblk + 1), ctx->pool, ctx->stream));
        r->allocs.push_back(r->d.cons_codes);
        VSK_CUDA(ctx, cudaMallocFromPoolAsync((void **)&r->d.cons_cnt, sizeof(uint64_t) * (np + 1), ctx->pool, ctx->stream));
        r->allocs.push_back(r->d.cons_cnt);
        if (np) {
            vsk_pack_consensus_kernel<<<(unsigned)((np * 32 + 255) / 256), 256, 0, ctx->stream>>>(r->d, (int64_t)np);
            ctx->launches++;
            VSK_CUDA(ctx, cudaGetLastError());
        }
    }
    VSK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *out = r;
    if (r->deferred) return VSK_OK;  // the presence bits arrive with copy_done
    return vsk_id_presence(ctx, r->d.ids, ts->prof_off.back(), &r->present);
}

// the scan is about to touch the composition ids / flanks (or the set is going away): wait for their copy
static int reads_arrived(vsk_ctx *ctx, const vsk_reads *r) {
    if (r && r->deferred) {
        VSK_CUDA(ctx, cudaEventSynchronize(ctx->copy_done));
        memcpy(const_cast<vsk_reads *>(r)->present.bits, ctx->h_present_pin, sizeof(r->present.bits));
        r->deferred = false;
    }
    return VSK_OK;
}

static void free_reads(vsk_reads *r) {
    if (!r) return;
    if (r->deferred) cudaEventSynchronize(r->ctx->copy_done);
    for (void *p : r->allocs) cudaFreeAsync(p, r->ctx->stream);
    delete r;
}

extern "C" int vsk_reads_upload(vsk_ctx *ctx, const vsk_trset *reads, vsk_reads **out) {
    if (!ctx || !reads || !out) return VSK_ERR_ARG;
    VSK_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = upload_trset(ctx, reads, out);
    if (rc) { free_reads(*out); *out = nullptr; }
    return rc;
}

extern "C" void vsk_reads_free(vsk_ctx *ctx, vsk_reads *r) {
    if (ctx) cudaSetDevice(ctx->device);
    free_reads(r);
}

extern "C" int vsk_refs_create(vsk_ctx *ctx, const vsk_trset *refs, vsk_refs **out) {
    return vsk_refs_create_cached(ctx, refs, nullptr, out);
}

extern "C" int vsk_refs_create_cached(vsk_ctx *ctx, const vsk_trset *refs, const char *cache_path, vsk_refs **out) {
    if (!ctx || !refs || !out) return VSK_ERR_ARG;
    *out = nullptr;
    if (refs->raw) return vsk_fail(ctx, VSK_ERR_ARG, "a raw (GPU-decoded) TR set cannot be the reference set: the seed index needs its consensus strings on the host");
    VSK_CUDA(ctx, cudaSetDevice(ctx->device));
    vsk_refindex_host hi;
    std::string err;
    int rc = vsk_refindex_get(*refs, cache_path, hi, err, nullptr);
    if (rc) return vsk_fail(ctx, rc, "%s", err.c_str());
    vsk_refs *R = new vsk_refs();
    R->ctx = ctx;
    R->host = refs;
    R->total_hits = hi.total_hits;
    R->entries = (int64_t)hi.e_prof.size();
#define FAIL(x) do { rc = (x); if (rc) { vsk_refs_free(ctx, R); return rc; } } while (0)
    FAIL(upload_trset(ctx, refs, &R->tr));
    memset(&R->idx, 0, sizeof R->idx);
    for (int t = 1; t <= 14; ++t) {
        uint64_t *dc;
        int64_t *db;
        FAIL(upload_vec(ctx, R->allocs, hi.codes[t], &dc));
        if (hi.bucket_off[t].empty()) hi.bucket_off[t].push_back(0);
        FAIL(upload_vec(ctx, R->allocs, hi.bucket_off[t], &db));
        R->idx.codes[t] = dc;
        R->idx.boff[t] = db;
        R->idx.ncodes[t] = (int32_t)hi.codes[t].size();
        {  // prefix table over the sorted codes
            const std::vector<uint64_t> &cv = hi.codes[t];
            int shift = 0;
            if (!cv.empty())
                while ((cv.back() >> shift) >> VSK_LUT_BITS) ++shift;
            std::vector<int32_t> lut(((size_t)1 << VSK_LUT_BITS) + 1, (int32_t)cv.size());
            size_t k = 0;
            for (size_t pfx = 0; pfx < ((size_t)1 << VSK_LUT_BITS); ++pfx) {
                while (k < cv.size() && (cv[k] >> shift) < pfx) ++k;
                lut[pfx] = (int32_t)k;
            }
            int32_t *dl;
            FAIL(upload_vec(ctx, R->allocs, lut, &dl));
            R->idx.lut[t] = dl;
            R->idx.lut_shift[t] = shift;
        }
    }
    std::vector<int64_t> gcum(hi.e_mult.size() + 1, 0);
    for (size_t i = 0; i < hi.e_mult.size(); ++i) gcum[i + 1] = gcum[i] + hi.e_mult[i];
    int16_t *dp;
    int32_t *dpr, *dfi;
    int64_t *dg;
    FAIL(upload_vec(ctx, R->allocs, hi.e_pmin, &dp));
    FAIL(upload_vec(ctx, R->allocs, hi.e_prof, &dpr));
    FAIL(upload_vec(ctx, R->allocs, hi.e_first, &dfi));
    FAIL(upload_vec(ctx, R->allocs, gcum, &dg));
    R->idx.e_pmin = dp;
    R->idx.e_prof = dpr;
    R->idx.e_first = dfi;
    R->idx.e_gcum = dg;
    // lookup tables whose entries come from double arithmetic the reference does on the host
    const vsk_seedinfo &si = vsk_seeds();
    std::vector<int16_t> tr(si.trange_of, si.trange_of + 5001), er(si.err_of, si.err_of + 5001);
    std::vector<int32_t> lcsmin(VSK_MAXPROF + 1), fmax(VSK_MAXPROF + 1);
    for (int s = 0; s <= VSK_MAXPROF; ++s) {
        const int m = (int)(s * .85 + .5);          // LCS_CUTOFF, hclust.h:538-541
        lcsmin[(size_t)s] = std::min(m, s - 1);
        fmax[(size_t)s] = std::min(8, (int)(0.4 * s + .01));  // hclust.h:664-667
    }
    int16_t *dtr, *der;
    int32_t *dl, *df;
    FAIL(upload_vec(ctx, R->allocs, tr, &dtr));
    FAIL(upload_vec(ctx, R->allocs, er, &der));
    FAIL(upload_vec(ctx, R->allocs, lcsmin, &dl));
    FAIL(upload_vec(ctx, R->allocs, fmax, &df));
    R->luts.trange_of = dtr;
    R->luts.err_of = der;
    R->luts.lcs_min = dl;
    R->luts.flank_maxerr = df;
    cudaError_t e = cudaMemcpyToSymbol(c_seed_off, si.off, sizeof(si.off));
    if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_tsize, si.tsize, sizeof(si.tsize));
    if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_nones, si.nones, sizeof(si.nones));
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) {
        vsk_refs_free(ctx, R);
        return vsk_fail(ctx, VSK_ERR_CUDA, "refs upload: %s", cudaGetErrorString(e));
    }
#undef FAIL
    *out = R;
    return VSK_OK;
}

extern "C" void vsk_refs_free(vsk_ctx *ctx, vsk_refs *R) {
    if (!R) return;
    if (ctx) cudaSetDevice(ctx->device);
    free_reads(R->tr);
    for (void *p : R->allocs) cudaFreeAsync(p, R->ctx->stream);
    cudaStreamSynchronize(R->ctx->stream);
    delete R;
}

extern "C" int64_t vsk_refs_seed_hits(const vsk_refs *R) { return R ? R->total_hits : 0; }

// ---------------------------------------------------------------- kernels

// window of one seed code inside the index: distinct entries [s,e2) plus, when an
// entry past `hi` exists, the head hit of that next pattern-size group (hclust.h:380-421)
struct SeedWindow {
    int64_t s, e2;
    int extra;
};

__device__ __forceinline__ bool seed_lookup(const DevIndex &idx, int t, uint64_t code, int lo, int hi,
                                            SeedWindow &w) {
    const uint64_t *codes = idx.codes[t];
    const uint64_t pfx = code >> idx.lut_shift[t];
    if (pfx >> VSK_LUT_BITS) return false;  // above every code of the index
    int a = idx.lut[t][pfx], b = idx.lut[t][pfx + 1];
    while (a < b) {
        const int m = (a + b) >> 1;
        if (codes[m] < code) a = m + 1;
        else b = m;
    }
    if (a >= idx.ncodes[t] || codes[a] != code) return false;
    const int64_t bs = idx.boff[t][a], be = idx.boff[t][a + 1];
    int64_t x = bs, y = be;
    while (x < y) {  // first entry with pmin >= lo
        const int64_t m = (x + y) >> 1;
        if (idx.e_pmin[m] < lo) x = m + 1;
        else y = m;
    }
    w.s = x;
    y = be;
    while (x < y) {  // first entry with pmin > hi
        const int64_t m = (x + y) >> 1;
        if (idx.e_pmin[m] <= hi) x = m + 1;
        else y = m;
    }
    w.e2 = x;
    w.extra = (x < be) ? 1 : 0;
    if (w.s >= be) return false;  // no hit with pmin >= lo: start_hit stays NULL
    return true;
}

template <bool FILL>
__global__ void __launch_bounds__(256)
vsk_seed_kernel(DevTR reads, DevIndex idx, DevLuts luts, int32_t p_begin, int32_t p_end,
                int64_t *__restrict__ cnt_emit, unsigned long long *__restrict__ raw_total,
                const int64_t *__restrict__ off_emit, int64_t base, int key_shift,
                uint64_t *__restrict__ keys, uint64_t *__restrict__ wcache) {
    // wcache[consensus position] = what the count pass found for the window that ends there
    // (first entry << 26 | entries << 1 | extra; 0 = nothing to emit), so that the fill pass does not
    // search again
    const int lane = threadIdx.x & 31;
    const int32_t p = p_begin + (int32_t)((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5);
    if (p >= p_end) return;
    const int32_t q = p >> 1;
    int64_t emit = 0, raw = 0;
    int t = 0;
    const int pmin = reads.patlen[q];
    if (reads.master[q]) t = luts.trange_of[pmin];
    if (t > 0) {
        const int e = luts.err_of[pmin];
        const int lo = pmin - e, hi = pmin + e;
        const uint8_t *seq = reads.cons + reads.cons_off[p];
        const int len = (int)(reads.cons_off[p + 1] - reads.cons_off[p]);
        const int ts = c_tsize[t], nones = c_nones[t];
        int64_t wbase = FILL ? (off_emit[p] - base) : 0;
        for (int i0 = ts - 1; i0 < len; i0 += 32) {
            const int i = i0 + lane;
            SeedWindow w;
            w.s = w.e2 = 0;
            w.extra = 0;
            bool found = false;
            int mult = 1;
            if (FILL) {
                const uint64_t pk = (i < len) ? wcache[reads.cons_off[p] + i] : 0ull;
                found = pk != 0ull;
                w.s = (int64_t)(pk >> 26);
                w.e2 = w.s + (int64_t)((pk >> 1) & 0x1ffffffull);
                w.extra = (int)(pk & 1ull);
            } else {
                uint64_t code = ~0ull;  // no window position (cannot equal a real code: those use < 60 bits)
                if (i < len) {
                    uint32_t c1 = 0, c2 = 0;
                    for (int k = 0; k < nones; ++k) {
                        const uint32_t b = (uint32_t)base_code(seq[i + c_seed_off[t][k]]);
                        if (k <= 14) c1 = c1 * 4 + b;
                        else c2 = c2 * 4 + b;
                    }
                    code = ((uint64_t)c1 << 32) | c2;
                }
                // periodic (low-complexity) consensus strings repeat the same seed code every few positions:
                // only the first lane of each group of equal codes looks the bucket up and emits keys; the
                // CandidatesConsidered counter still counts the window once per position (`mult` lanes)
                const unsigned same = __match_any_sync(0xffffffffu, code);
                const bool leader = (i < len) && (lane == __ffs(same) - 1);
                mult = __popc(same);
                if (leader) found = seed_lookup(idx, t, code, lo, hi, w);
                if (i < len)
                    wcache[reads.cons_off[p] + i] =
                        found ? (((uint64_t)w.s << 26) | ((uint64_t)(w.e2 - w.s) << 1) | (uint64_t)w.extra) : 0ull;
            }
            const int64_t nd = found ? (w.e2 - w.s + w.extra) : 0;
            if (!FILL) {
                emit += nd;
                if (found) raw += (int64_t)mult * (idx.e_gcum[w.e2] - idx.e_gcum[w.s] + w.extra);
            } else {
                // exclusive prefix of nd over the warp
                int64_t incl = nd;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int64_t v = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o) incl += v;
                }
                int64_t dst = wbase + incl - nd;
                if (found) {
                    // sort key: read profile (relative to the chunk) above the reference profile
                    const uint64_t hi32 = (uint64_t)(uint32_t)(p - p_begin) << key_shift;
                    for (int64_t k = w.s; k < w.e2; ++k) keys[dst++] = hi32 | (uint32_t)idx.e_prof[k];
                    if (w.extra) keys[dst] = hi32 | (uint32_t)idx.e_first[w.e2];
                }
                wbase += __shfl_sync(0xffffffffu, incl, 31);
            }
        }
    }
    if (!FILL) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            emit += __shfl_down_sync(0xffffffffu, emit, o);
            raw += __shfl_down_sync(0xffffffffu, raw, o);
        }
        if (lane == 0) {
            cnt_emit[p - p_begin] = emit;
            if (raw) atomicAdd(raw_total, (unsigned long long)raw);
        }
    }
}

// LCS filter on the packed consensus strings.  Same result as LCS_multiple_word /
// LCS_single_word on the ASCII strings (the consensus holds only A,C,G,T): any exact bit-vector
// LCS gives the same length, so the vectors here index character i at bit i of block i/64; the
// single-word copy's 0x10 bitmask quirk (single word.c:59) is the truncation of the longer string
// to 60 characters when its length is 61..63.
__device__ __forceinline__ uint64_t lcs_plane(const uint64_t (&pl)[4], uint32_t code) {
    const uint64_t lo = (code & 1u) ? pl[1] : pl[0], hi = (code & 1u) ? pl[3] : pl[2];
    return (code & 2u) ? hi : lo;
}

__global__ void __launch_bounds__(128)
vsk_lcs_filter_kernel(DevTR refs, DevTR reads, DevLuts luts, const uint64_t *__restrict__ pairs, int64_t n,
                      uint8_t *__restrict__ flags, uint32_t *__restrict__ long_list, unsigned *__restrict__ long_count) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool in_range = i < n;
    const uint64_t key = in_range ? pairs[i] : 0ull;
    const uint32_t p = (uint32_t)(key >> 32), h = (uint32_t)key;
    const int nr = in_range ? (int)(refs.cons_off[h + 1] - refs.cons_off[h]) : 0;
    const int mr = in_range ? (int)(reads.cons_off[p + 1] - reads.cons_off[p]) : 0;
    // the longer string is bit-indexed; on a tie string1 = the reference (multiple word.c:76-86)
    const bool ref_is_text = nr >= mr;
    int N = ref_is_text ? nr : mr;
    const int M = ref_is_text ? mr : nr;
    const int shortest = M;
    if (N <= 63 && N > 60) N = 60;
    // a common subsequence cannot use more copies of a base than either string has: pairs whose
    // composition bound is already below the threshold fail without running the recurrence
    bool decided = false;
    if (in_range) {
        const uint64_t ca = refs.cons_cnt[h], cb = reads.cons_cnt[p];
        int bound = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) bound += min((int)((ca >> (16 * k)) & 0xffffu), (int)((cb >> (16 * k)) & 0xffffu));
        if (bound < luts.lcs_min[shortest]) {
            flags[i] = 0;
            decided = true;
        }
    }
    // more than one block: left to vsk_lcs_sys_kernel, which keeps this kernel's warps uniform.  Three
    // lists by the number of 64-column blocks W (2 / 3..8 / more), one counter update per warp each;
    // list 1 grows from the back of the buffer list 0 grows from the front of
    const bool is_long = in_range && !decided && N > 64;
    const int W = (N + 63) >> 6;
    const int cls = (W <= 2) ? 0 : (W <= 8) ? 1 : 2;
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const unsigned lmask = __ballot_sync(0xffffffffu, is_long && cls == c);
        if (!lmask) continue;
        const int leader = __ffs(lmask) - 1;
        unsigned base = 0;
        if (lane == leader) base = atomicAdd(long_count + c, (unsigned)__popc(lmask));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (is_long && cls == c) {
            const int64_t k = (int64_t)base + __popc(lmask & ((1u << lane) - 1u));
            const int64_t at = (c == 0) ? k : (c == 1) ? n - 1 - k : n + k;
            long_list[at] = (uint32_t)i;
            flags[i] = 0;
        }
    }
    if (!in_range || is_long || decided) return;
    const uint64_t *planes = ref_is_text ? refs.cons_planes + 4 * refs.cons_blk[h] : reads.cons_planes + 4 * reads.cons_blk[p];
    const uint64_t *codes = ref_is_text ? reads.cons_codes + 2 * reads.cons_blk[p] : refs.cons_codes + 2 * refs.cons_blk[h];
    uint64_t pl[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) pl[k] = planes[k];
    const uint64_t valid = (N == 64) ? ~0ull : ((1ull << N) - 1ull);
    uint64_t V = ~0ull, cw = 0;
    for (int r = 0; r < M; ++r) {
        if ((r & 31) == 0) cw = codes[r >> 5];
        const uint64_t m = lcs_plane(pl, (uint32_t)(cw >> (2 * (r & 31))) & 3u) & valid;
        const uint64_t x = V & m;
        V = (V + x) | (V & ~m);
    }
    const int count = __popcll(~V & valid);
    flags[i] = (count >= luts.lcs_min[shortest]) ? 1 : 0;  // hclust.h:537-545
}

// The multi-block pairs deferred by vsk_lcs_filter_kernel.  A group of G lanes works on one pair:
// lane g owns block (64 columns) g of the bit vector and runs row s-g at step s, so the carry of
// the row's addition moves one lane per step (shuffle) -- a systolic pass of M+G-1 steps; strings
// of more than 64*G columns take several passes, the carries out of the last lane parked one bit
// per row in shared memory (G = 32 only).  Persistent lanes with flattened state: the loop body is
// one step, so groups that are at different places of different pairs execute the same
// instructions; a group pulls its next pair from the list's counter when it is done.
template <int G>
__global__ void __launch_bounds__(128)
vsk_lcs_sys_kernel(DevTR refs, DevTR reads, DevLuts luts, const uint64_t *__restrict__ pairs,
                   const uint32_t *__restrict__ list, int list_dir, const unsigned *__restrict__ list_count,
                   unsigned *__restrict__ head, uint8_t *__restrict__ flags) {
    constexpr int NB = (G == 32) ? (VSK_MAXSTR / 64 + 2) : 1;
    __shared__ uint64_t s_bnd[(G == 32) ? 4 : 1][2][NB];
    const unsigned n = *list_count;
    const int lane = threadIdx.x & 31, gl = lane & (G - 1);
    const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << (lane & ~(G - 1)));
    uint64_t *bnd_in = s_bnd[(G == 32) ? (threadIdx.x >> 5) : 0][0], *bnd_out = s_bnd[(G == 32) ? (threadIdx.x >> 5) : 0][1];
    const uint64_t *planes = nullptr, *codes = nullptr;
    uint64_t pl[4] = {0, 0, 0, 0}, V = 0, valid = 0, cw = 0, cout_w = 0, bin_w = 0;
    int N = 0, M = 0, W = 0, pass = 0, s = 0, count = 0, need = 0;
    unsigned co_prev = 0;
    uint32_t item = 0;
    bool have = false;
    for (;;) {
        if (!have) {
            unsigned k = 0;
            if (gl == 0) k = atomicAdd(head, 1u);
            k = __shfl_sync(gmask, k, 0, G);
            if (k >= n) break;
            item = list[(int64_t)list_dir * (int64_t)k];
            const uint32_t p = (uint32_t)(pairs[item] >> 32), h = (uint32_t)pairs[item];
            const int nr = (int)(refs.cons_off[h + 1] - refs.cons_off[h]);
            const int mr = (int)(reads.cons_off[p + 1] - reads.cons_off[p]);
            const bool ref_is_text = nr >= mr;  // on a tie string1 = the reference (multiple word.c:76-86)
            planes = ref_is_text ? refs.cons_planes + 4 * refs.cons_blk[h] : reads.cons_planes + 4 * reads.cons_blk[p];
            codes = ref_is_text ? reads.cons_codes + 2 * reads.cons_blk[p] : refs.cons_codes + 2 * refs.cons_blk[h];
            N = ref_is_text ? nr : mr;
            M = ref_is_text ? mr : nr;
            need = luts.lcs_min[M];  // hclust.h:537-545
            W = (N + 63) >> 6;
            pass = 0;
            s = 0;
            count = 0;
            have = true;
        }
        if (s == 0) {  // start of a pass
            const int j = pass * G + gl;
            const int top = N - j * 64;
#pragma unroll
            for (int k = 0; k < 4; ++k) pl[k] = (top > 0) ? planes[4 * j + k] : 0ull;
            valid = (top >= 64) ? ~0ull : (top > 0) ? ((1ull << top) - 1ull) : 0ull;
            V = ~0ull;
            co_prev = 0;
        }
        const int r = s - gl;
        const unsigned cin_lane = __shfl_up_sync(gmask, co_prev, 1, G);
        if (r >= 0 && r < M) {
            if ((r & 31) == 0) cw = codes[r >> 5];
            unsigned cin = cin_lane;
            if (gl == 0) {
                cin = 0;
                if (G == 32 && pass > 0) {
                    if ((r & 63) == 0) bin_w = bnd_in[r >> 6];
                    cin = (unsigned)(bin_w >> (r & 63)) & 1u;
                }
            }
            const uint64_t m = lcs_plane(pl, (uint32_t)(cw >> (2 * (r & 31))) & 3u) & valid;
            const uint64_t x = V & m;
            const uint64_t s1 = V + x;
            unsigned co = (s1 < x) ? 1u : 0u;
            const uint64_t s2 = s1 + cin;
            co |= (s2 < s1) ? 1u : 0u;
            V = s2 | (V & ~m);
            co_prev = co;
            if (G == 32 && gl == G - 1) {  // carries into the next pass, one bit per row
                if ((r & 63) == 0) cout_w = 0;
                cout_w |= (uint64_t)co << (r & 63);
                if ((r & 63) == 63 || r == M - 1) bnd_out[r >> 6] = cout_w;
            }
        }
        if (++s == M + G - 1) {  // end of the pass
            int c = __popcll(~V & valid);
#pragma unroll
            for (int o = G >> 1; o > 0; o >>= 1) c += __shfl_xor_sync(gmask, c, o, G);
            count += c;
            ++pass;
            s = 0;
            // the LCS length is the sum over blocks of the matched columns: stop as soon as the
            // threshold is reached, or cannot be reached even if every remaining column matched
            const int remaining = max(N - 64 * G * pass, 0);
            if (pass * G >= W || count >= need || count + remaining < need) {
                if (gl == 0) flags[item] = (count >= need) ? 1 : 0;
                have = false;
            }
            if (G == 32) {
                __syncwarp();
                uint64_t *t = bnd_in;
                bnd_in = bnd_out;
                bnd_out = t;
            }
        }
    }
}

// DP job per surviving pair (hclust.h:550-577) + its scheduling key
__global__ void __launch_bounds__(128)
vsk_dp_jobs_kernel(DevTR refs, DevTR reads, const uint64_t *__restrict__ pairs, int64_t n, double tan_angle,
                   int count_cells,
                   vsk_dp_job *__restrict__ jobs, uint32_t *__restrict__ cost, uint8_t *__restrict__ uncertain,
                   unsigned long long *__restrict__ stats /* [0]=cells [1]=n_uncertain [2]=max_l1 [6]=id+result bytes */) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long cells = 0, bytes = 0, wcells = 0, wbytes = 0;
    unsigned nunc = 0, ntiny = 0;
    int maxl1 = 0;
    if (i < n) {
        const uint32_t p = (uint32_t)(pairs[i] >> 32), h = (uint32_t)pairs[i];
        vsk_dp_job j;
        j.left_off = refs.prof_off[h];
        j.top_off = reads.prof_off[p];
        j.l1 = (int32_t)(refs.prof_off[h + 1] - j.left_off);
        j.l2 = (int32_t)(reads.prof_off[p + 1] - j.top_off);
        j.out_idx = (int32_t)i;
        j.br = -1;
        j.svector = 0.0;
        uint8_t unc = 0;
        cells = (unsigned long long)j.l1 * (unsigned long long)j.l2;
        if (min(refs.patlen[h >> 1], reads.patlen[p >> 1]) >= 30) {
            const double big = (double)max(j.l1, j.l2), small = (double)min(j.l1, j.l2);
            const double sizeerror = __dmul_rn(__dsub_rn(__ddiv_rn(big, small), 1.0), 100.0);
            if (!(sizeerror > 100.0)) {
                // profile.h:1752-1770.  sqrt, /, * are IEEE-exact on both sides; atan/sin may differ
                // from the host libm in the last place, so a band radius that lands within 1e-7 of
                // an integer is handed to the host to settle (`uncertain`).
                const double minlen = small;
                const double virtdiagonal = sqrt(__dmul_rn(__dmul_rn(minlen, minlen), 2.0));
                const double topangle = atan(__ddiv_rn((double)j.l1, (double)j.l2));
                const double midradius = __dmul_rn(tan_angle, __ddiv_rn(virtdiagonal, 2.0));
                const double bandradius = __ddiv_rn(midradius, sin(topangle));
                const double v = (bandradius >= 6.0) ? bandradius : 6.0;
                j.br = (int)v + 1;
                const double fr = v - floor(v);
                if (v > 6.0 && (fr < 1e-7 || fr > 1.0 - 1e-7)) unc = 1;
                j.svector = __ddiv_rn((double)j.l2, (double)j.l1);
                if (count_cells) {  // band cells: sum over rows of (rend - lstart + 1), profile.h:1835-1839
                    cells = 0;
                    for (int row = 1; row <= j.l1; ++row) {
                        const int center = __double2int_rz(__dmul_rn((double)row, j.svector));
                        const int a = max(1, center - j.br), b = min(j.l2, center + j.br);
                        if (b >= a) cells += (unsigned long long)(b - a + 1);
                    }
                }
            }
        }
        jobs[i] = j;
        // warp-per-pair jobs sort before the tiny (thread-per-pair) ones, each group heaviest first
        cost[i] = vsk_dp_sort_key(j.l1, j.l2, j.br);
        // low half: tiny (thread-per-pair) jobs, high half: quarter-warp jobs
        ntiny = VSK_DP_TINY(j.l1, j.l2, j.br) ? 1u : (VSK_DP_Q8(j.l1, j.l2, j.br) ? 0x10000u : 0u);
        uncertain[i] = unc;
        nunc = unc;
        maxl1 = j.l1;
        bytes = (unsigned long long)(2 * (j.l1 + j.l2) + 16);  // u16 ids in, one int4 out
        if (VSK_DP_DIRJOB(j.l2)) {  // the share of the warp-per-pair forward sweep (the dominant kernel)
            wcells = cells;
            wbytes = bytes;
        }
    }
    // one atomic per warp instead of one per thread
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        cells += __shfl_down_sync(0xffffffffu, cells, o);
        bytes += __shfl_down_sync(0xffffffffu, bytes, o);
        wcells += __shfl_down_sync(0xffffffffu, wcells, o);
        wbytes += __shfl_down_sync(0xffffffffu, wbytes, o);
        nunc += __shfl_down_sync(0xffffffffu, nunc, o);
        ntiny += __shfl_down_sync(0xffffffffu, ntiny, o);
        maxl1 = max(maxl1, __shfl_down_sync(0xffffffffu, maxl1, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&stats[0], cells);
        if (nunc) atomicAdd(&stats[1], (unsigned long long)nunc);
        atomicMax(&stats[2], (unsigned long long)maxl1);
        atomicAdd(&stats[6], bytes);
        if (wcells) {
            atomicAdd(&stats[9], wcells);
            atomicAdd(&stats[10], wbytes);
        }
        if (ntiny & 0xffffu) atomicAdd(&stats[7], (unsigned long long)(ntiny & 0xffffu));
        if (ntiny >> 16) atomicAdd(&stats[8], (unsigned long long)(ntiny >> 16));
    }
}

__global__ void vsk_gather_jobs_kernel(const vsk_dp_job *__restrict__ jobs, const uint32_t *__restrict__ order,
                                       int64_t n, vsk_dp_job *__restrict__ sorted) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) sorted[i] = jobs[order[i]];
}

__global__ void vsk_iota_kernel(uint32_t *v, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] = (uint32_t)i;
}

// similarity test (hclust.h:593) and number of flank checks the pair expands to (hclust.h:604-607)
__global__ void __launch_bounds__(128)
vsk_dp_pass_kernel(DevTR refs, DevTR reads, const uint64_t *__restrict__ pairs, const vsk_dp_job *__restrict__ jobs,
                   const int4 *__restrict__ res, int64_t n, double target, int32_t *__restrict__ nexp,
                   unsigned long long *__restrict__ stats /* [3]=passed [4]=abort */) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned passed = 0, aborts = 0;
    if (i < n) {
        const int4 r = res[i];
        const double sim = __dsub_rn(1.0, __ddiv_rn((double)r.y, __dmul_rn(255.0, (double)r.z)));
        if (jobs[i].br >= 0 && r.w > 5000) aborts = 1;  // profile.h:1960-1965 exit(1)
        int ne = 0;
        if (sim >= target) {
            const uint32_t p = (uint32_t)(pairs[i] >> 32), h = (uint32_t)pairs[i];
            const int rq = p >> 1, rr = h >> 1;
            ne = (refs.rot_off[rr + 1] - refs.rot_off[rr]) * (reads.rot_off[rq + 1] - reads.rot_off[rq]);
            passed = 1;
        }
        nexp[i] = ne;
    }
    passed = __reduce_add_sync(0xffffffffu, passed);
    aborts = __reduce_add_sync(0xffffffffu, aborts);
    if ((threadIdx.x & 31) == 0) {
        if (passed) atomicAdd(&stats[3], (unsigned long long)passed);
        if (aborts) atomicAdd(&stats[4], (unsigned long long)aborts);
    }
}

template <bool RMODE>
__global__ void __launch_bounds__(128)
vsk_flank_kernel(DevTR refs, DevTR reads, DevLuts luts, const uint64_t *__restrict__ pairs,
                 const int4 *__restrict__ res, const int64_t *__restrict__ exp_off, int64_t npairs, int64_t nexp,
                 int maxerrors, vsk_link_rec *__restrict__ recs, uint8_t *__restrict__ flags,
                 unsigned long long *__restrict__ stats /* [5]=ED error */) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nexp) return;
    // pair that owns expansion e: last i with exp_off[i] <= e
    int64_t a = 0, b = npairs;
    while (a < b) {
        const int64_t m = (a + b) >> 1;
        if (exp_off[m + 1] <= e) a = m + 1;
        else b = m;
    }
    const int64_t i = a;
    const int k = (int)(e - exp_off[i]);
    const uint32_t p = (uint32_t)(pairs[i] >> 32), h = (uint32_t)pairs[i];
    const int rq = p >> 1, rr = h >> 1;
    const int nread = reads.rot_off[rq + 1] - reads.rot_off[rq];
    const int ma = refs.rot_mem[refs.rot_off[rr] + k / nread];    // ref rotlist member (outer loop)
    const int mb = reads.rot_mem[reads.rot_off[rq] + k % nread];  // read rotlist member (inner loop)
    // hclust.h:631-651
    const int newdir = (int)(p & 1) ^ (int)(h & 1) ^ (int)refs.dir[ma] ^ (int)reads.dir[mb];
    const int rp = 2 * ma + newdir;  // ref_ptr: forward or rc profile of the member
    const int dp = 2 * mb;           // read_ptr: always the member's forward profile
    const int rl = refs.leftlen[rp], rrl = refs.rightlen[rp];
    const int ql = reads.leftlen[dp], qr = reads.rightlen[dp];
    const int maxl = maxerrors ? maxerrors : luts.flank_maxerr[ql];
    const int maxr = maxerrors ? maxerrors : luts.flank_maxerr[qr];
    const uint8_t *rfl = refs.fl + refs.fl_off[2 * rp], *rfr = refs.fl + refs.fl_off[2 * rp + 1];
    const uint8_t *qfl = reads.fl + reads.fl_off[2 * dp], *qfr = reads.fl + reads.fl_off[2 * dp + 1];
    int lerr, rerr;
    if (!RMODE) {  // refs vs reads (hclust.h:656-697): the read flank is cut to the reference flank's length
        lerr = vsk_ed_device<VSK_ED_PSEARCH>(rfl, rl, qfl, min(ql, rl));
        rerr = vsk_ed_device<VSK_ED_PSEARCH>(rfr, rrl, qfr, min(qr, rrl));
    } else {       // refs vs refs (hclust.h:700-751): full lengths, then once more with the roles swapped
        lerr = vsk_ed_device<VSK_ED_PSEARCH>(rfl, rl, qfl, ql);
        rerr = vsk_ed_device<VSK_ED_PSEARCH>(rfr, rrl, qfr, qr);
        if (lerr != -1 && rerr != -1 && !(lerr <= maxl && rerr <= maxr)) {
            lerr = vsk_ed_device<VSK_ED_PSEARCH>(qfl, ql, rfl, rl);
            rerr = vsk_ed_device<VSK_ED_PSEARCH>(qfr, qr, rfr, rrl);
        }
    }
    if (lerr == -1 || rerr == -1) atomicAdd(&stats[5], 1ull);  // hclust.h:690-694 quits
    vsk_link_rec r;
    r.refkey = refs.key[ma];
    r.readkey = reads.key[mb];
    r.dir = newdir;
    r.lerr = lerr;
    r.rerr = rerr;
    r.dsum = res[i].y;
    r.sumw = res[i].z;
    recs[e] = r;
    flags[e] = (lerr <= maxl && rerr <= maxr) ? 1 : 0;
}

// ---------------------------------------------------------------- host driver
namespace {
struct Scratch {  // stream-ordered temporaries of one run
    vsk_ctx *ctx;
    std::vector<void *> ptrs;
    explicit Scratch(vsk_ctx *c) : ctx(c) {}
    ~Scratch() {
        for (void *p : ptrs) cudaFreeAsync(p, ctx->stream);
    }
    template <typename T>
    cudaError_t get(T **p, size_t n) {
        cudaError_t e = cudaMallocFromPoolAsync((void **)p, sizeof(T) * (n ? n : 1), ctx->pool, ctx->stream);
        if (e == cudaSuccess) ptrs.push_back(*p);
        return e;
    }
};
inline unsigned blocks(int64_t n, int per) { return (unsigned)((n + per - 1) / per); }
}  // namespace

struct vsk_proclu_result {
    // accepted links in a page-locked buffer of the context (vsk_pin_get), returned when the result is freed
    vsk_ctx *ctx = nullptr;
    vsk_link_rec *links = nullptr;
    size_t n_links = 0, cap = 0;  // records, bytes
    vsk_proclu_stats stats;
    ~vsk_proclu_result() {
        if (links) vsk_pin_put(ctx, links, cap);
    }
};

// ---- candsort + dedup without a sort ----------------------------------------------------
// The keys of one read profile are contiguous (the seed kernel writes them at the profile's
// offset), so "sort by (read, ref) and drop duplicates" is, per read profile, "which reference
// profiles occur": a block sets one bit per key in a shared-memory bitmap over the reference
// profiles and reads the set bits back in ascending order (clearing them for the next profile).
// The unique keys go back to the start of the profile's own segment, widened to
// (read profile << 32 | reference profile); vsk_compact_kernel then closes the gaps.
#define DEDUP_THREADS 256

// bitmap of `nwords` 32-bit words + a summary with one bit per bitmap word (only the words that
// were touched are visited when the bits are read back); thread t owns SW consecutive summary words
__global__ void __launch_bounds__(DEDUP_THREADS)
vsk_dedup_kernel(uint64_t *__restrict__ keys, const int64_t *__restrict__ off_emit, int64_t base, int32_t p_begin,
                 int32_t nprof, int nwords, uint64_t refmask, int *__restrict__ counter, int64_t *__restrict__ ucount) {
    extern __shared__ uint32_t s_bm[];
    __shared__ int s_q;
    __shared__ int s_wsum[DEDUP_THREADS / 32];
    const int nsum = (nwords + 31) >> 5;
    const int SW = (nsum + DEDUP_THREADS - 1) / DEDUP_THREADS;
    uint32_t *bm = s_bm, *sum = s_bm + nwords;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < nwords + nsum; i += DEDUP_THREADS) s_bm[i] = 0;
    const int sw0 = min(tid * SW, nsum), sw1 = min(sw0 + SW, nsum);
    for (;;) {
        __syncthreads();  // everybody is done with s_q (and the bitmap is clean)
        if (tid == 0) {   // next profile that has keys
            int q;
            for (;;) {
                q = atomicAdd(counter, 1);
                if (q >= nprof || off_emit[p_begin + q + 1] != off_emit[p_begin + q]) break;
                ucount[q] = 0;
            }
            s_q = q;
        }
        __syncthreads();
        const int q = s_q;
        if (q >= nprof) break;
        const int64_t s0 = off_emit[p_begin + q] - base;
        const int64_t n = off_emit[p_begin + q + 1] - base - s0;
        for (int64_t i = tid; i < n; i += DEDUP_THREADS) {
            const uint32_t ref = (uint32_t)(keys[s0 + i] & refmask);
            const uint32_t w = ref >> 5;
            atomicOr(&bm[w], 1u << (ref & 31));
            atomicOr(&sum[w >> 5], 1u << (w & 31));
        }
        __syncthreads();
        int cnt = 0;
        for (int sw = sw0; sw < sw1; ++sw) {
            uint32_t m = sum[sw];
            while (m) {
                cnt += __popc(bm[sw * 32 + __ffs(m) - 1]);
                m &= m - 1;
            }
        }
        int incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) s_wsum[warp] = incl;
        __syncthreads();
        int woff = 0, total = 0;
#pragma unroll
        for (int w = 0; w < DEDUP_THREADS / 32; ++w) {
            const int v = s_wsum[w];
            if (w < warp) woff += v;
            total += v;
        }
        if (cnt) {
            const uint64_t hi = (uint64_t)(uint32_t)(p_begin + q) << 32;
            int64_t dst = s0 + woff + incl - cnt;
            for (int sw = sw0; sw < sw1; ++sw) {
                uint32_t m = sum[sw];
                if (!m) continue;
                sum[sw] = 0;
                while (m) {
                    const int w = sw * 32 + __ffs(m) - 1;
                    m &= m - 1;
                    uint32_t b = bm[w];
                    bm[w] = 0;
                    while (b) {
                        keys[dst++] = hi | (uint64_t)((uint32_t)w * 32u + (uint32_t)(__ffs(b) - 1));
                        b &= b - 1;
                    }
                }
            }
        }
        if (tid == 0) ucount[q] = total;
    }
}

// warp per read profile: the unique keys of the profile move to their place in the dense pair list
__global__ void __launch_bounds__(256)
vsk_compact_kernel(const uint64_t *__restrict__ keys, const int64_t *__restrict__ off_emit, int64_t base,
                   int32_t p_begin, int32_t nprof, const int64_t *__restrict__ ucount,
                   const int64_t *__restrict__ uoff, uint64_t *__restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t q = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (q >= nprof) return;
    const int64_t n = ucount[q];
    if (n == 0) return;
    const uint64_t *src = keys + (off_emit[p_begin + q] - base);
    uint64_t *dst = out + uoff[q];
    for (int64_t i = lane; i < n; i += 32) dst[i] = src[i];
}

struct WidenKey {  // chunk-relative sort key -> (read profile << 32) | reference profile
    int shift;
    uint32_t p_begin;
    uint64_t mask;
    __host__ __device__ uint64_t operator()(uint64_t k) const {
        return ((uint64_t)((uint32_t)(k >> shift) + p_begin) << 32) | (k & mask);
    }
};

static int run_chunk(vsk_ctx *ctx, const vsk_refs *R, const vsk_reads *Q, const vsk_proclu_params *prm,
                     int32_t p_begin, int32_t p_end, const int64_t *d_off_emit, int64_t base, int64_t nkeys,
                     uint64_t *d_wcache, vsk_proclu_result *out) {
    if (nkeys == 0) return VSK_OK;
    cudaStream_t st = ctx->stream;
    Scratch S(ctx);
    const DevTR &refs = R->tr->d, &reads = Q->d;
    uint64_t *keys, *keys2, *pairs, *pairs2;
    int64_t *d_n;
    unsigned long long *d_stats;
    VSK_CUDA(ctx, S.get(&keys, (size_t)nkeys));
    VSK_CUDA(ctx, S.get(&keys2, (size_t)nkeys));
    VSK_CUDA(ctx, S.get(&d_n, 4));
    VSK_CUDA(ctx, S.get(&d_stats, 12));
    VSK_CUDA(ctx, cudaMemsetAsync(d_stats, 0, 12 * sizeof(unsigned long long), st));
    const int64_t nprof = p_end - p_begin;
    vsk_stage_mark(ctx, "chunk_setup");
    // Sort keys are as narrow as the chunk allows (fewer radix passes): the read profile relative
    // to the chunk above the reference profile; the unique pass widens them to (read << 32 | ref).
    int key_shift = 1;
    while (key_shift < 32 && ((uint64_t)(2 * (int64_t)refs.n - 1) >> key_shift)) ++key_shift;
    int end_bit = key_shift + 1;
    while (end_bit < 64 && ((uint64_t)(nprof - 1) >> (end_bit - key_shift))) ++end_bit;
    vsk_seed_kernel<true><<<blocks(nprof * 32, 256), 256, 0, st>>>(reads, R->idx, R->luts, p_begin, p_end,
                                                                   nullptr, nullptr, d_off_emit, base, key_shift, keys,
                                                                   d_wcache);
    ctx->launches++;
    vsk_stage_mark(ctx, "seed_fill");
    // candsort + dedup: per-profile bitmap when the reference profiles fit one in shared memory,
    // else radix sort + unique
    static const bool use_bitmap = [] {
        const char *e = getenv("VSK_DEDUP_BITMAP");
        return !(e && e[0] == '0');
    }();
    const int64_t bm_words = (2 * (int64_t)refs.n + 31) / 32;
    size_t tb = 0;
    void *tmp = nullptr;
    int64_t npairs = 0;
    if (use_bitmap && bm_words <= 48000 && nprof < 0x7fffffff) {
        const size_t smem = (size_t)(bm_words + (bm_words + 31) / 32) * sizeof(uint32_t);
        static uint64_t attr_set = 0;  // one bit per device ordinal
        if (!((attr_set >> (ctx->device & 63)) & 1)) {
            VSK_CUDA(ctx, cudaFuncSetAttribute(vsk_dedup_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               (48000 + 1500) * (int)sizeof(uint32_t)));
            attr_set |= 1ull << (ctx->device & 63);
        }
        int per_sm = 0;
        VSK_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, vsk_dedup_kernel, DEDUP_THREADS, smem));
        if (per_sm < 1) per_sm = 1;
        int64_t *ucount, *uoff;
        int *counter;
        VSK_CUDA(ctx, S.get(&ucount, (size_t)nprof + 1));
        VSK_CUDA(ctx, S.get(&uoff, (size_t)nprof + 1));
        VSK_CUDA(ctx, S.get(&counter, 4));
        VSK_CUDA(ctx, cudaMemsetAsync(counter, 0, sizeof(int), st));
        VSK_CUDA(ctx, cudaMemsetAsync(ucount + nprof, 0, sizeof(int64_t), st));
        vsk_dedup_kernel<<<(unsigned)(ctx->sm_count * per_sm), DEDUP_THREADS, smem, st>>>(
            keys, d_off_emit, base, p_begin, (int32_t)nprof, (int)bm_words, ((uint64_t)1 << key_shift) - 1, counter, ucount);
        VSK_CUDA(ctx, cudaGetLastError());
        cub::DeviceScan::ExclusiveSum(nullptr, tb, ucount, uoff, (int)nprof + 1, st);
        VSK_CUDA(ctx, S.get((uint8_t **)&tmp, tb));
        VSK_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, tb, ucount, uoff, (int)nprof + 1, st));
        vsk_compact_kernel<<<blocks(nprof * 32, 256), 256, 0, st>>>(keys, d_off_emit, base, p_begin, (int32_t)nprof,
                                                                   ucount, uoff, keys2);
        ctx->launches += 2;
        pairs = keys2;
        pairs2 = keys;
        vsk_stage_mark(ctx, "sort_unique");
        VSK_CUDA(ctx, cudaMemcpyAsync(&npairs, uoff + nprof, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    } else {
        auto wide = thrust::make_transform_output_iterator(
            keys, WidenKey{key_shift, (uint32_t)p_begin, ((uint64_t)1 << key_shift) - 1});
        size_t tb2 = 0;
        cub::DeviceRadixSort::SortKeys(nullptr, tb, keys, keys2, nkeys, 0, end_bit, st);
        cub::DeviceSelect::Unique(nullptr, tb2, keys2, wide, d_n, nkeys, st);
        tb = std::max(tb, tb2);
        VSK_CUDA(ctx, S.get((uint8_t **)&tmp, tb));
        VSK_CUDA(ctx, cub::DeviceRadixSort::SortKeys(tmp, tb, keys, keys2, nkeys, 0, end_bit, st));
        VSK_CUDA(ctx, cub::DeviceSelect::Unique(tmp, tb, keys2, wide, d_n, nkeys, st));
        pairs = keys;
        pairs2 = keys2;
        vsk_stage_mark(ctx, "sort_unique");
        VSK_CUDA(ctx, cudaMemcpyAsync(&npairs, d_n, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    }
    VSK_CUDA(ctx, cudaStreamSynchronize(st));
    out->stats.lcs_alignments += (uint64_t)npairs;
    if (npairs == 0) return VSK_OK;
    // LCS filter
    uint8_t *flags;
    VSK_CUDA(ctx, S.get(&flags, (size_t)npairs));
    uint32_t *long_list;
    unsigned *long_ctr;
    VSK_CUDA(ctx, S.get(&long_list, 2 * (size_t)npairs));
    VSK_CUDA(ctx, S.get(&long_ctr, 6));
    VSK_CUDA(ctx, cudaMemsetAsync(long_ctr, 0, 6 * sizeof(unsigned), st));
    vsk_lcs_filter_kernel<<<blocks(npairs, 128), 128, 0, st>>>(refs, reads, R->luts, pairs, npairs, flags, long_list,
                                                              long_ctr);
    // heaviest class first
    vsk_lcs_sys_kernel<32><<<(unsigned)ctx->sm_count * 8, 128, 0, st>>>(refs, reads, R->luts, pairs, long_list + npairs, 1,
                                                                       long_ctr + 2, long_ctr + 5, flags);
    vsk_lcs_sys_kernel<8><<<(unsigned)ctx->sm_count * 8, 128, 0, st>>>(refs, reads, R->luts, pairs,
                                                                      long_list + npairs - 1, -1, long_ctr + 1,
                                                                      long_ctr + 4, flags);
    vsk_lcs_sys_kernel<2><<<(unsigned)ctx->sm_count * 8, 128, 0, st>>>(refs, reads, R->luts, pairs, long_list, 1,
                                                                      long_ctr, long_ctr + 3, flags);
    ctx->launches += 4;
    size_t tb3 = 0;
    cub::DeviceSelect::Flagged(nullptr, tb3, pairs, flags, pairs2, d_n, npairs, st);
    if (tb3 > tb) { VSK_CUDA(ctx, S.get((uint8_t **)&tmp, tb3)); tb = tb3; }
    VSK_CUDA(ctx, cub::DeviceSelect::Flagged(tmp, tb, pairs, flags, pairs2, d_n, npairs, st));
    vsk_stage_mark(ctx, "lcs_filter");
    int64_t ndp = 0;
    VSK_CUDA(ctx, cudaMemcpyAsync(&ndp, d_n, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    VSK_CUDA(ctx, cudaStreamSynchronize(st));
    out->stats.profile_alignments += (uint64_t)ndp;
    if (ndp == 0) return VSK_OK;
    // DP jobs, heaviest first
    vsk_dp_job *jobs, *sorted;
    uint32_t *cost, *cost2, *order, *order2;
    uint8_t *unc;
    int4 *res;
    VSK_CUDA(ctx, S.get(&jobs, (size_t)ndp));
    VSK_CUDA(ctx, S.get(&sorted, (size_t)ndp));
    VSK_CUDA(ctx, S.get(&cost, (size_t)ndp));
    VSK_CUDA(ctx, S.get(&cost2, (size_t)ndp));
    VSK_CUDA(ctx, S.get(&order, (size_t)ndp));
    VSK_CUDA(ctx, S.get(&order2, (size_t)ndp));
    VSK_CUDA(ctx, S.get(&unc, (size_t)ndp));
    VSK_CUDA(ctx, S.get(&res, (size_t)ndp));
    const int homology_int = (int)((prm->cutoff / 100.0) * 100);  // TARGET_HOMOLOGY_INT, psearch.c:1003
    const double tan_angle = tan(acos(homology_int / 100.0));
    vsk_dp_jobs_kernel<<<blocks(ndp, 128), 128, 0, st>>>(refs, reads, pairs2, ndp, tan_angle, prm->count_cells, jobs,
                                                        cost, unc, d_stats);
    vsk_iota_kernel<<<blocks(ndp, 256), 256, 0, st>>>(order, ndp);
    ctx->launches += 2;
    unsigned long long hs[12];
    VSK_CUDA(ctx, cudaMemcpyAsync(hs, d_stats, sizeof hs, cudaMemcpyDeviceToHost, st));
    VSK_CUDA(ctx, cudaStreamSynchronize(st));
    if (hs[1]) {  // settle band radii the device could not decide
        std::vector<uint8_t> hu((size_t)ndp);
        std::vector<vsk_dp_job> hj((size_t)ndp);
        VSK_CUDA(ctx, cudaMemcpyAsync(hu.data(), unc, (size_t)ndp, cudaMemcpyDeviceToHost, st));
        VSK_CUDA(ctx, cudaMemcpyAsync(hj.data(), jobs, sizeof(vsk_dp_job) * (size_t)ndp, cudaMemcpyDeviceToHost, st));
        VSK_CUDA(ctx, cudaStreamSynchronize(st));
        for (int64_t i = 0; i < ndp; ++i)
            if (hu[(size_t)i]) {
                vsk_band_host(hj[(size_t)i].l1, hj[(size_t)i].l2, homology_int, &hj[(size_t)i].svector,
                              &hj[(size_t)i].br);
                VSK_CUDA(ctx, cudaMemcpyAsync(jobs + i, &hj[(size_t)i], sizeof(vsk_dp_job), cudaMemcpyHostToDevice, st));
            }
        VSK_CUDA(ctx, cudaStreamSynchronize(st));
    }
    out->stats.dp_cells += (int64_t)hs[0];
    if (const char *dump = getenv("VSK_DUMP_JOBS")) {  // debugging aid: append (l1, l2, br) of every DP job
        std::vector<vsk_dp_job> hj((size_t)ndp);
        VSK_CUDA(ctx, cudaMemcpyAsync(hj.data(), jobs, sizeof(vsk_dp_job) * (size_t)ndp, cudaMemcpyDeviceToHost, st));
        VSK_CUDA(ctx, cudaStreamSynchronize(st));
        if (FILE *f = fopen(dump, "ab")) {
            for (const vsk_dp_job &j : hj) {
                const int32_t rec[3] = {j.l1, j.l2, j.br};
                fwrite(rec, sizeof rec, 1, f);
            }
            fclose(f);
        }
    }
    out->stats.dp_alg_bytes += (int64_t)hs[6];
    out->stats.dp_cells_sweep += (int64_t)hs[9];
    out->stats.dp_alg_bytes_sweep += (int64_t)hs[10];
    size_t tb4 = 0;
    cub::DeviceRadixSort::SortPairsDescending(nullptr, tb4, cost, cost2, order, order2, ndp, 0, 22, st);
    if (tb4 > tb) { VSK_CUDA(ctx, S.get((uint8_t **)&tmp, tb4)); tb = tb4; }
    VSK_CUDA(ctx, cub::DeviceRadixSort::SortPairsDescending(tmp, tb, cost, cost2, order, order2, ndp, 0, 22, st));
    vsk_gather_jobs_kernel<<<blocks(ndp, 256), 256, 0, st>>>(jobs, order2, ndp, sorted);
    ctx->launches++;
    vsk_stage_mark(ctx, "dp_schedule");
    int rc = reads_arrived(ctx, Q);  // ids, flanks and the presence bits of a set still in flight
    if (rc) return rc;
    vsk_idset ids;
    for (int k = 0; k < 32; ++k) ids.bits[k] = R->tr->present.bits[k] | Q->present.bits[k];
    rc = vsk_dp_launch(ctx, refs.ids, reads.ids, sorted, ndp, (int64_t)hs[8], (int64_t)hs[7], (int32_t)hs[2], &ids, res);
    if (rc) return rc;
    vsk_stage_mark(ctx, "dp_kernel");
    // similarity test + flank expansion
    int32_t *nexp;
    int64_t *exp_off;
    VSK_CUDA(ctx, S.get(&nexp, (size_t)ndp + 1));
    VSK_CUDA(ctx, S.get(&exp_off, (size_t)ndp + 1));
    VSK_CUDA(ctx, cudaMemsetAsync(nexp + ndp, 0, sizeof(int32_t), st));
    const double target = prm->cutoff / 100.0;
    vsk_dp_pass_kernel<<<blocks(ndp, 128), 128, 0, st>>>(refs, reads, pairs2, jobs, res, ndp, target, nexp, d_stats);
    ctx->launches++;
    size_t tb5 = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tb5, nexp, exp_off, ndp + 1, st);
    if (tb5 > tb) { VSK_CUDA(ctx, S.get((uint8_t **)&tmp, tb5)); tb = tb5; }
    VSK_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, tb, nexp, exp_off, ndp + 1, st));
    int64_t ne = 0;
    VSK_CUDA(ctx, cudaMemcpyAsync(&ne, exp_off + ndp, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    VSK_CUDA(ctx, cudaMemcpyAsync(hs, d_stats, sizeof hs, cudaMemcpyDeviceToHost, st));
    VSK_CUDA(ctx, cudaStreamSynchronize(st));
    out->stats.profile_alignments_success += (uint64_t)hs[3];
    if (hs[4]) return vsk_fail(ctx, VSK_ERR_ABORT, "ERROR! Cannot find start!");
    if (ne == 0) return VSK_OK;
    vsk_link_rec *recs, *recs2;
    uint8_t *lflags;
    VSK_CUDA(ctx, S.get(&recs, (size_t)ne));
    VSK_CUDA(ctx, S.get(&recs2, (size_t)ne));
    VSK_CUDA(ctx, S.get(&lflags, (size_t)ne));
    if (prm->option_r)
        vsk_flank_kernel<true><<<blocks(ne, 128), 128, 0, st>>>(refs, reads, R->luts, pairs2, res, exp_off, ndp, ne,
                                                                0, recs, lflags, d_stats);
    else
        vsk_flank_kernel<false><<<blocks(ne, 128), 128, 0, st>>>(refs, reads, R->luts, pairs2, res, exp_off, ndp, ne,
                                                                 prm->maxerrors, recs, lflags, d_stats);
    ctx->launches++;
    size_t tb6 = 0;
    cub::DeviceSelect::Flagged(nullptr, tb6, recs, lflags, recs2, d_n, ne, st);
    if (tb6 > tb) { VSK_CUDA(ctx, S.get((uint8_t **)&tmp, tb6)); tb = tb6; }
    VSK_CUDA(ctx, cub::DeviceSelect::Flagged(tmp, tb, recs, lflags, recs2, d_n, ne, st));
    int64_t nl = 0;
    VSK_CUDA(ctx, cudaMemcpyAsync(&nl, d_n, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    VSK_CUDA(ctx, cudaMemcpyAsync(hs, d_stats, sizeof hs, cudaMemcpyDeviceToHost, st));
    VSK_CUDA(ctx, cudaStreamSynchronize(st));
    if (hs[5])
        return vsk_fail(ctx, VSK_ERR_ABORT, "Edit_Distance_multiple_word_NoEndPenaltySeq1 returned -1. Aborting!");
    vsk_stage_mark(ctx, "flank_links");
    const size_t old = out->n_links;
    // device copy of the link table (the chunks append to it): grow-only buffer of the context
    if ((old + (size_t)nl) * sizeof(vsk_link_rec) > ctx->d_links_cap) {
        const size_t want = std::max<size_t>(2 * (old + (size_t)nl) * sizeof(vsk_link_rec), (size_t)1 << 20);
        void *nb = nullptr;
        VSK_CUDA(ctx, cudaMalloc(&nb, want));
        if (old) VSK_CUDA(ctx, cudaMemcpyAsync(nb, ctx->d_links, old * sizeof(vsk_link_rec), cudaMemcpyDeviceToDevice, st));
        VSK_CUDA(ctx, cudaStreamSynchronize(st));
        if (ctx->d_links) VSK_CUDA(ctx, cudaFree(ctx->d_links));
        ctx->d_links = nb;
        ctx->d_links_cap = want;
    }
    if (nl)
        VSK_CUDA(ctx, cudaMemcpyAsync((vsk_link_rec *)ctx->d_links + old, recs2, sizeof(vsk_link_rec) * (size_t)nl,
                                      cudaMemcpyDeviceToDevice, st));
    out->n_links = old + (size_t)nl;
    if (prm->links_on_device) {  // the caller reads them through vsk_proclu_result_links_dev
        VSK_CUDA(ctx, cudaStreamSynchronize(st));
        vsk_stage_mark(ctx, "links_d2h");
        return VSK_OK;
    }
    if ((old + (size_t)nl) * sizeof(vsk_link_rec) > out->cap) {
        size_t ncap = 0;
        vsk_link_rec *nb = (vsk_link_rec *)vsk_pin_get(ctx, 2 * (old + (size_t)nl) * sizeof(vsk_link_rec), &ncap);
        if (!nb) return vsk_fail(ctx, VSK_ERR_CUDA, "page-locked buffer for %zu links", old + (size_t)nl);
        if (old) memcpy(nb, out->links, old * sizeof(vsk_link_rec));
        if (out->links) vsk_pin_put(ctx, out->links, out->cap);
        out->links = nb;
        out->cap = ncap;
    }
    if (nl)
        VSK_CUDA(ctx, cudaMemcpyAsync(out->links + old, recs2, sizeof(vsk_link_rec) * (size_t)nl,
                                      cudaMemcpyDeviceToHost, st));
    VSK_CUDA(ctx, cudaStreamSynchronize(st));
    out->stats.d2h_bytes += (int64_t)(sizeof(vsk_link_rec) * (size_t)nl);
    vsk_stage_mark(ctx, "links_d2h");
    return VSK_OK;
}

extern "C" int vsk_proclu_run_resident(vsk_ctx *ctx, const vsk_refs *R, const vsk_reads *Q,
                                       const vsk_proclu_params *prm, vsk_proclu_result **out) {
    if (!ctx || !R || !Q || !prm || !out) return VSK_ERR_ARG;
    *out = nullptr;
    if (prm->option_r && prm->maxerrors != 0)  // psearch.c:995-999
        return vsk_fail(ctx, VSK_ERR_ARG, "maxerror must be 0 because it will be dynamically generated based on flank length.");
    VSK_CUDA(ctx, cudaSetDevice(ctx->device));
    const vsk_trset *hq = Q->host;
    const vsk_seedinfo &si = vsk_seeds();
    for (int32_t q = 0; q < hq->n && !Q->checked; ++q) {  // the reference quits on these (psearch.c:784-787, hclust.h:866-869)
        const int pmin = hq->patlen[(size_t)q];
        if (pmin < 7 || pmin > 5000) return vsk_fail(ctx, VSK_ERR_ABORT, "pmin(%d) must be in 7-5000 range!", pmin);
        if (!hq->master[(size_t)q]) continue;
        const int t = si.trange_of[pmin];
        for (int o = 0; o < 2 && t > 0; ++o) {
            const int64_t len = hq->cons_off[(size_t)(2 * q + o + 1)] - hq->cons_off[(size_t)(2 * q + o)];
            if (si.tsize[t] > len)
                return vsk_fail(ctx, VSK_ERR_ABORT, "Tuple size (%d) is larger than sequence(%d). Aborting!",
                                si.tsize[t], (int)len);
        }
    }
    Q->checked = true;
    if (2 * (int64_t)R->tr->d.n >= (1 << 25))  // a seed bucket holds at most one entry per reference profile
        return vsk_fail(ctx, VSK_ERR_ARG, "more than 2^24 reference TRs: the seed-window records do not hold them");
    vsk_proclu_result *res = new vsk_proclu_result();
    res->ctx = ctx;
    memset(&res->stats, 0, sizeof res->stats);
    const int32_t nprof = 2 * Q->d.n;
    if (nprof == 0) { *out = res; return VSK_OK; }
    cudaStream_t st = ctx->stream;
    int rc = VSK_OK;
    {
        Scratch S(ctx);
        int64_t *cnt, *off;
        unsigned long long *raw;
        cudaError_t e;
        uint64_t *wcache;  // one word per consensus character of the reads: the seed windows the count pass found
        if ((e = S.get(&cnt, (size_t)nprof + 1)) != cudaSuccess || (e = S.get(&off, (size_t)nprof + 1)) != cudaSuccess ||
            (e = S.get(&raw, 1)) != cudaSuccess || (e = S.get(&wcache, (size_t)hq->cons_off.back() + 1)) != cudaSuccess) {
            delete res;
            return vsk_fail(ctx, VSK_ERR_CUDA, "alloc: %s", cudaGetErrorString(e));
        }
        vsk_stage_mark(ctx, nullptr);
        cudaMemsetAsync(raw, 0, sizeof(unsigned long long), st);
        cudaMemsetAsync(cnt + nprof, 0, sizeof(int64_t), st);
        vsk_seed_kernel<false><<<blocks((int64_t)nprof * 32, 256), 256, 0, st>>>(Q->d, R->idx, R->luts, 0, nprof, cnt,
                                                                                raw, nullptr, 0, 0, nullptr, wcache);
        ctx->launches++;
        size_t tb = 0;
        cub::DeviceScan::ExclusiveSum(nullptr, tb, cnt, off, nprof + 1, st);
        void *tmp;
        S.get((uint8_t **)&tmp, tb);
        cub::DeviceScan::ExclusiveSum(tmp, tb, cnt, off, nprof + 1, st);
        vsk_stage_mark(ctx, "seed_count");
        size_t hoff_cap = 0;
        int64_t *hoff = (int64_t *)vsk_pin_get(ctx, sizeof(int64_t) * ((size_t)nprof + 1), &hoff_cap);
        if (!hoff) {
            delete res;
            return vsk_fail(ctx, VSK_ERR_CUDA, "page-locked buffer for the key offsets");
        }
        struct PinGuard {
            vsk_ctx *c; void *p; size_t cap;
            ~PinGuard() { vsk_pin_put(c, p, cap); }
        } hoff_guard{ctx, hoff, hoff_cap};
        unsigned long long hraw = 0;
        cudaMemcpyAsync(hoff, off, sizeof(int64_t) * ((size_t)nprof + 1), cudaMemcpyDeviceToHost, st);
        cudaMemcpyAsync(&hraw, raw, sizeof hraw, cudaMemcpyDeviceToHost, st);
        e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) {
            delete res;
            return vsk_fail(ctx, VSK_ERR_CUDA, "seed count: %s", cudaGetErrorString(e));
        }
        res->stats.candidates_considered = hraw;
        res->stats.candidate_keys = hoff[(size_t)nprof];
        // chunks of read profiles whose keys fit the budget
        int64_t budget = prm->key_budget > 0 ? prm->key_budget : (int64_t)256 << 20;
        if (prm->key_budget <= 0)
            if (const char *e = getenv("VSK_KEY_BUDGET")) budget = std::max<int64_t>(1024, atoll(e));
        int32_t pb = 0;
        int64_t masters_total = 0, masters_done = 0;
        if (prm->progress)
            for (int32_t q = 0; q < hq->n; ++q) masters_total += hq->master[(size_t)q] ? 1 : 0;
        while (pb < nprof && rc == VSK_OK) {
            int32_t pe = pb + 2;
            while (pe < nprof && hoff[(size_t)pe + 2] - hoff[(size_t)pb] <= budget) pe += 2;
            rc = run_chunk(ctx, R, Q, prm, pb, pe, off, hoff[(size_t)pb], hoff[(size_t)pe] - hoff[(size_t)pb], wcache, res);
            res->stats.chunks++;
            if (prm->progress && rc == VSK_OK) {
                for (int32_t q = pb / 2; q < pe / 2; ++q) masters_done += hq->master[(size_t)q] ? 1 : 0;
                prm->progress(prm->progress_user, masters_done, masters_total, res->stats.profile_alignments);
            }
            pb = pe;
        }
    }
    if (rc) { delete res; return rc; }
    res->stats.confirmed = (uint64_t)res->n_links;
    *out = res;
    return VSK_OK;
}

extern "C" int vsk_proclu_run(vsk_ctx *ctx, const vsk_refs *R, const vsk_trset *reads, const vsk_proclu_params *prm,
                              vsk_proclu_result **out) {
    if (!ctx || !R || !reads || !out) return VSK_ERR_ARG;
    vsk_reads *Q = nullptr;
    VSK_CUDA(ctx, cudaSetDevice(ctx->device));
    vsk_stage_mark(ctx, nullptr);
    int rc = upload_trset(ctx, reads, &Q, /*defer_big=*/!reads->raw);
    vsk_stage_mark(ctx, "upload");
    if (rc) { free_reads(Q); return rc; }
    const int64_t h2d = Q->h2d_bytes;
    rc = vsk_proclu_run_resident(ctx, R, Q, prm, out);
    if (rc == VSK_OK) (*out)->stats.h2d_bytes = h2d;
    vsk_reads_free(ctx, Q);
    return rc;
}

extern "C" int64_t vsk_proclu_result_links(const vsk_proclu_result *r, const vsk_link **links) {
    if (!r) return 0;
    if (links) *links = reinterpret_cast<const vsk_link *>(r->links);
    return (int64_t)r->n_links;
}
extern "C" int64_t vsk_proclu_result_links_dev(const vsk_proclu_result *r, const vsk_link **d_links) {
    if (!r) return 0;
    if (d_links) *d_links = r->n_links ? reinterpret_cast<const vsk_link *>(r->ctx->d_links) : nullptr;
    return (int64_t)r->n_links;
}
extern "C" void vsk_proclu_result_stats(const vsk_proclu_result *r, vsk_proclu_stats *s) {
    if (r && s) *s = r->stats;
}
extern "C" void vsk_proclu_result_free(vsk_proclu_result *r) { delete r; }
