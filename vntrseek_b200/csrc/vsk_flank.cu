// flankalign on the GPU (reference src/newflankalign/flankalign.c).
//
// The reference parses one `.flanks` cluster block at a time, forks a child per
// block, sorts reads and references by pattern size and, for every read x
// reference whose pattern sizes differ by at most PATLEN_SIZE_ERR %, computes
// four no-end-penalty edit distances (forward flanks, and the complement of the
// opposite flanks), applies the thresholds and prints `ref:lerr:rerr`.
//
// Here the host parses and orders (including the reference's GFSR-driven
// unstable quicksort, so the order of equal pattern sizes -- and therefore the
// output bytes -- match), enumerates the (read, reference) pairs of many blocks
// into one batch, and one kernel thread per pair does the four edit distances
// and the decision (flankalign.c:573-749).  Results come back as one int per
// pair and the host writes the `<a>_<b>.map` files in block order.
#include <ctype.h>
#include <errno.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <string>

#include "vsk_bits.cuh"
#include "vsk_host.hpp"

struct vsk_flankrec {
    int64_t left_off, right_off;  // into the batch string pool
    int32_t leftlen, rightlen;
};

// one pair: 4 x Edit_Distance_multiple_word_NoEndPenaltySeq1 (newflankalign copy) + decision.
// out: bit 0..7 lerr, 8..15 rerr, 16 accepted, 17 edit-distance error (-1) in the forward pair,
// 18 error in the swapped pair
__global__ void __launch_bounds__(128)
vsk_flankalign_kernel(const uint8_t *__restrict__ pool, const vsk_flankrec *__restrict__ refs,
                      const vsk_flankrec *__restrict__ reads, const int32_t *__restrict__ pair_read,
                      const int32_t *__restrict__ pair_ref, int64_t n, int maxerrors,
                      const int32_t *__restrict__ maxerr_lut, uint32_t *__restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const vsk_flankrec rd = reads[pair_read[i]], rf = refs[pair_ref[i]];
    int mel1, mer1, mel2, mer2, large;
    if (maxerrors != 0) {
        mel1 = mer1 = mel2 = mer2 = large = maxerrors;  // flankalign.c:622-628
    } else {
        mel1 = maxerr_lut[rd.leftlen];   // min(8,(int)(0.4*leftlen+.01)), flankalign.c:633-636
        mer1 = maxerr_lut[rd.rightlen];
        mel2 = mer1;
        mer2 = mel1;
        large = max(mel1, mer1);
    }
    // forward: ref left vs read left, ref right vs read right (flankalign.c:639-654)
    int refleft = min(rf.leftlen, rd.leftlen + large + 2);
    int refright = min(rf.rightlen, rd.rightlen + large + 2);
    const int lerr1 = vsk_ed_device<VSK_ED_FLANKALIGN>(pool + rf.left_off, refleft, pool + rd.left_off,
                                                       min(rd.leftlen, refleft));
    const int rerr1 = vsk_ed_device<VSK_ED_FLANKALIGN>(pool + rf.right_off, refright, pool + rd.right_off,
                                                       min(rd.rightlen, refright));
    uint32_t res = 0;
    if (lerr1 < 0 || rerr1 < 0) {
        out[i] = 1u << 17;
        return;
    }
    // swapped: ref left vs complement(read right), ref right vs complement(read left) (:674-689)
    refleft = min(rf.leftlen, rd.rightlen + large + 2);
    refright = min(rf.rightlen, rd.leftlen + large + 2);
    const int lerr2 = vsk_ed_device<VSK_ED_FLANKALIGN, true>(pool + rf.left_off, refleft, pool + rd.right_off,
                                                             min(rd.rightlen, refleft));
    const int rerr2 = vsk_ed_device<VSK_ED_FLANKALIGN, true>(pool + rf.right_off, refright, pool + rd.left_off,
                                                             min(rd.leftlen, refright));
    if (lerr2 < 0 || rerr2 < 0) {
        out[i] = 1u << 18;
        return;
    }
    const bool p1 = (lerr1 <= mel1 && rerr1 <= mer1), p2 = (lerr2 <= mel2 && rerr2 <= mer2);
    if (p1 || p2) {  // flankalign.c:708-738
        int lerr, rerr;
        if (p1 && p2) {
            if (lerr1 + rerr1 <= lerr2 + rerr2) { lerr = lerr1; rerr = rerr1; }
            else { lerr = lerr2; rerr = rerr2; }
        } else if (p1) { lerr = lerr1; rerr = rerr1; }
        else { lerr = lerr2; rerr = rerr2; }
        res = (uint32_t)(lerr & 255) | ((uint32_t)(rerr & 255) << 8) | (1u << 16);
    }
    out[i] = res;
}

// ------------------------------------------------------------------ host side
namespace {
// easylife's generator + quicksort over (patsize, index) -- see vsk_index.cu for the
// derivation; flankalign children start from the pristine generator state
// (EasyListQuickSort reads then refs, flankalign.c:559-560)
struct Gfsr2 {
    uint32_t s[99];
    Gfsr2() {
        static const uint32_t init[99] = {
            40, 1387256442, 539505633, 7126687, 2115653676, 480642437, 1403109719, 898019591, 1609472695,
            742049136, 964528840, 1774590149, 531014893, 1478060509, 224730595, 1413365137, 1415397063,
            370513614, 1981855272, 1672294721, 1559669404, 1992066581, 440083042, 1552169384, 949029171,
            1848294689, 1014369863, 1226252978, 199445637, 552539314, 101995811, 1795618857, 1468200845,
            403608434, 466262418, 1783034892, 2125486341, 1437171068, 839437811, 685760609, 311739045,
            1876584692, 223544964, 667792106, 1829604735, 887026472, 688815796, 1153871680, 1135467106,
            1975710098, 1393037901, 330755675, 804762632, 393596594, 1695657725, 50479950, 1039358666,
            1885424316, 400881551, 142829986, 187416368, 821029919, 1292641081, 415120294, 1104581275,
            1258423968, 304285054, 400491932, 2014625087, 1619263031, 750624285, 1996732699, 97476312,
            1250544934, 2145510054, 1510875684, 262891578, 616032534, 1316668730, 1500747974, 2138561534,
            809719156, 1605036043, 510086967, 317411066, 54278455, 2052774305, 439191668, 1881943474,
            1397167115, 2046084812, 644321591, 328615697, 1004646018, 1110120728, 2007784487, 992677826,
            1756605308, 796797739};
        memcpy(s, init, sizeof s);
    }
    int range(int lo, int hi) {
        uint32_t p = s[0];
        if (++p > 98) p = 1;
        uint32_t k = p + 27;
        if (k > 98) k -= 98;
        s[p] ^= s[k];
        s[0] = p;
        const double d = ((double)s[p] / 2147483647.0) * (hi - lo + 1) + lo - .5;
        int i = (int)d;
        const double g = d - i;
        if (g >= .5) i++;
        else if (g <= -.5) i--;
        return i < lo ? lo : (i > hi ? hi : i);
    }
};
struct PItem {
    int patsize;
    int32_t idx;
};
struct PSort {
    Gfsr2 rng;
    static void insertion(PItem *a, int n) {
        for (int j = 1; j < n; ++j) {
            const PItem key = a[j];
            int i = j - 1;
            while (i >= 0 && a[i].patsize > key.patsize) { a[i + 1] = a[i]; --i; }
            a[i + 1] = key;
        }
    }
    int partition(PItem *a, int i, int k) {
        int r0 = rng.range(i, k), r1 = rng.range(i, k), r2 = rng.range(i, k);
        if (r0 > r1) std::swap(r0, r1);
        if (r1 > r2) std::swap(r1, r2);
        const int pval = a[r1].patsize;
        --i; ++k;
        for (;;) {
            do { --k; } while (a[k].patsize > pval);
            do { ++i; } while (a[i].patsize < pval);
            if (i >= k) break;
            std::swap(a[i], a[k]);
        }
        return k;
    }
    void helper(PItem *a, int i, int k) {
        while (i < k) {
            const int j = partition(a, i, k);
            if (j - i >= 9) helper(a, i, j);
            else insertion(a + i, j - i + 1);
            i = j + 1;
        }
    }
    void sort(std::vector<PItem> &v) { if (v.size() > 1) helper(v.data(), 0, (int)v.size() - 1); }
};

struct Block {
    int a1, a2;
    std::vector<int32_t> reads;           // record indices (batch-wide), in sorted order
    std::vector<int32_t> per_read_pairs;  // number of pairs of each sorted read
    int64_t pair_begin = 0;               // first pair of this block in the batch
};

struct Batch {
    std::string pool;
    std::vector<vsk_flankrec> ref_recs, read_recs;
    std::vector<long long> ref_ids, read_ids;
    std::vector<int32_t> pair_read, pair_ref;
    std::vector<Block> blocks;
    void clear() {
        pool.clear(); ref_recs.clear(); read_recs.clear(); ref_ids.clear(); read_ids.clear();
        pair_read.clear(); pair_ref.clear(); blocks.clear();
    }
};

inline void normalise(char *b, char *e) {  // toupper + non-ACGT -> N (flankalign.c:340-346 etc.)
    for (; b < e; ++b) {
        *b = (char)toupper((unsigned char)*b);
        if (*b != 'A' && *b != 'C' && *b != 'G' && *b != 'T') *b = 'N';
    }
}

// walk to the next ',' (or the end), optionally normalising on the way; terminates the
// field and returns the start of the next one -- the reference's src1..src4 walks
inline char *next_field(char *p, bool norm) {
    char *b = p;
    while (*p != ',' && *p != '\0') ++p;
    if (norm) normalise(b, p);
    if (*p == ',') { *p = '\0'; ++p; }
    return p;
}
}  // namespace

static int flush_batch(vsk_ctx *ctx, Batch &B, int maxerrors, vsk_flank_stats *st) {
    const int64_t n = (int64_t)B.pair_read.size();
    std::vector<uint32_t> out((size_t)n);
    if (n > 0) {
        cudaStream_t s = ctx->stream;
        uint8_t *d_pool = nullptr;
        vsk_flankrec *d_refs = nullptr, *d_reads = nullptr;
        int32_t *d_pr = nullptr, *d_pf = nullptr, *d_lut = nullptr;
        uint32_t *d_out = nullptr;
        int maxlen = 1;
        for (const vsk_flankrec &r : B.read_recs) maxlen = std::max(maxlen, std::max(r.leftlen, r.rightlen));
        for (const vsk_flankrec &r : B.ref_recs) maxlen = std::max(maxlen, std::max(r.leftlen, r.rightlen));
        if (maxlen > VSK_MAXPROF) return vsk_fail(ctx, VSK_ERR_ARG, "flank longer than %d characters", VSK_MAXPROF);
        std::vector<int32_t> lut((size_t)maxlen + 1);
        for (int k = 0; k <= maxlen; ++k) lut[(size_t)k] = std::min(8, (int)(0.4 * k + .01));
        VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_pool, B.pool.size() + 16, ctx->pool, s));
        VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_refs, sizeof(vsk_flankrec) * (B.ref_recs.size() + 1), ctx->pool, s));
        VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_reads, sizeof(vsk_flankrec) * (B.read_recs.size() + 1), ctx->pool, s));
        VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_pr, sizeof(int32_t) * (size_t)n, ctx->pool, s));
        VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_pf, sizeof(int32_t) * (size_t)n, ctx->pool, s));
        VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_lut, sizeof(int32_t) * lut.size(), ctx->pool, s));
        VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_out, sizeof(uint32_t) * (size_t)n, ctx->pool, s));
        VSK_CUDA(ctx, cudaMemcpyAsync(d_pool, B.pool.data(), B.pool.size(), cudaMemcpyHostToDevice, s));
        VSK_CUDA(ctx, cudaMemcpyAsync(d_refs, B.ref_recs.data(), sizeof(vsk_flankrec) * B.ref_recs.size(), cudaMemcpyHostToDevice, s));
        VSK_CUDA(ctx, cudaMemcpyAsync(d_reads, B.read_recs.data(), sizeof(vsk_flankrec) * B.read_recs.size(), cudaMemcpyHostToDevice, s));
        VSK_CUDA(ctx, cudaMemcpyAsync(d_pr, B.pair_read.data(), sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, s));
        VSK_CUDA(ctx, cudaMemcpyAsync(d_pf, B.pair_ref.data(), sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, s));
        VSK_CUDA(ctx, cudaMemcpyAsync(d_lut, lut.data(), sizeof(int32_t) * lut.size(), cudaMemcpyHostToDevice, s));
        cudaEvent_t e0, e1;
        VSK_CUDA(ctx, cudaEventCreate(&e0));
        VSK_CUDA(ctx, cudaEventCreate(&e1));
        VSK_CUDA(ctx, cudaEventRecord(e0, s));
        vsk_flankalign_kernel<<<(unsigned)((n + 127) / 128), 128, 0, s>>>(d_pool, d_refs, d_reads, d_pr, d_pf, n,
                                                                       maxerrors, d_lut, d_out);
        ctx->launches++;
        VSK_CUDA(ctx, cudaGetLastError());
        VSK_CUDA(ctx, cudaEventRecord(e1, s));
        VSK_CUDA(ctx, cudaMemcpyAsync(out.data(), d_out, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToHost, s));
        VSK_CUDA(ctx, cudaStreamSynchronize(s));
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        if (st) {
            st->kernel_ms += ms;
            st->h2d_bytes += (int64_t)(B.pool.size() + sizeof(vsk_flankrec) * (B.ref_recs.size() + B.read_recs.size()) + 8 * (size_t)n);
            st->d2h_bytes += (int64_t)(4 * (size_t)n);
        }
        for (void *p : {(void *)d_pool, (void *)d_refs, (void *)d_reads, (void *)d_pr, (void *)d_pf, (void *)d_lut, (void *)d_out})
            cudaFreeAsync(p, s);
    }
    // write the .map files in block order (flankalign.c:562-579,741-746); cwd is the output dir
    std::string text;
    char tmp[160];
    for (const Block &b : B.blocks) {
        text.clear();
        snprintf(tmp, sizeof tmp, "@%d_%d\t1\t1\t1\t1\t1", b.a1, b.a2);
        text += tmp;
        int64_t p = b.pair_begin;
        bool aborted = false;
        for (size_t r = 0; r < b.reads.size() && !aborted; ++r) {
            const vsk_flankrec &rr = B.read_recs[(size_t)b.reads[r]];
            snprintf(tmp, sizeof tmp, "\n%lld\t'\t%d\t%d\t1\t1\t1\t", B.read_ids[(size_t)b.reads[r]], rr.leftlen, rr.rightlen);
            text += tmp;
            bool comma = false;
            for (int k = 0; k < b.per_read_pairs[r]; ++k, ++p) {
                const uint32_t o = out[(size_t)p];
                if (o & (3u << 17)) {  // the reference child quits at this point (flankalign.c:655-658,690-693)
                    printf("\n\nERROR!!!!:\n\n\nFlankAlign - wrong sizes for narrowband alignment, seq1 must be larger. Aborting!\n\n");
                    aborted = true;
                    break;
                }
                if (o & (1u << 16)) {
                    snprintf(tmp, sizeof tmp, "%s%lld:%d:%d", comma ? "," : "", B.ref_ids[(size_t)B.pair_ref[(size_t)p]],
                             (int)(o & 255), (int)((o >> 8) & 255));
                    text += tmp;
                    comma = true;
                    if (st) st->accepted++;
                }
            }
        }
        snprintf(tmp, sizeof tmp, "%d_%d.map", b.a1, b.a2);
        FILE *fp = fopen(tmp, "w");
        if (!fp) return vsk_fail(ctx, VSK_ERR_ABORT, "FlankAlign - can't open output file. Aborting!");
        fwrite(text.data(), 1, text.size(), fp);
        fclose(fp);
        printf("processing: %d_%d (gpu) \n", b.a1, b.a2);
    }
    B.clear();
    return VSK_OK;
}

// flankalign.exe outdir xmldir inputfile MAXERRORS MAX_FLANK_CONSIDERED MAXTHREADS PATLEN_SIZE_ERR
// (flankalign.c:214-823).  The caller has already chdir'ed into outdir.
extern "C" int vsk_flankalign_run(vsk_ctx *ctx, const char *inputfile, int maxerrors, int maxflank, int paterr,
                                  int64_t pairs_per_batch, vsk_flank_stats *stats) {
    if (!ctx || !inputfile || maxflank < 1 || paterr < 1) return VSK_ERR_ARG;
    VSK_CUDA(ctx, cudaSetDevice(ctx->device));
    if (stats) memset(stats, 0, sizeof *stats);
    if (pairs_per_batch <= 0) pairs_per_batch = 4 << 20;
    FILE *fp = (inputfile[0] == '-') ? stdin : fopen(inputfile, "r");
    if (!fp) return vsk_fail(ctx, VSK_ERR_ABORT, "FlankAlign - Could not open input file. Aborting!");
    char *line = nullptr;
    size_t cap = 0;
    ssize_t got;
    int a1 = 0, a2 = 0;
    bool hasdata = false;
    while ((got = getline(&line, &cap, fp)) >= 0) {  // flankalign.c:299-312
        if (line[0] == '@') {
            sscanf(line, "@(%d_%d)", &a1, &a2);
            if (a1 <= 0 || a2 <= 0) {
                printf("Can't read cluster number.");
                free(line);
                return VSK_ERR_ABORT;
            }
            hasdata = true;
            break;
        }
    }
    Batch B;
    int rc = VSK_OK;
    struct Tmp { long long id; int patsize; int32_t rec; };
    std::vector<Tmp> refs, reads;
    while (hasdata && rc == VSK_OK) {
        refs.clear();
        reads.clear();
        // the header line itself parses as rid 0 and is skipped; so does the '*' line
        for (;;) {
            const long long rid = strtoll(line, nullptr, 10);
            if (rid < 0) {  // -id,FL,SEQ,FR,PAT (flankalign.c:329-390)
                char *src1 = next_field(line, false);
                char *src2 = next_field(src1, true);   // FL normalised
                char *src3 = next_field(src2, false);  // SEQ untouched
                char *src4 = next_field(src3, true);   // FR normalised
                for (char *e = src4 + strlen(src4); e > src4 && (e[-1] == '\n' || e[-1] == '\r'); --e) e[-1] = '\0';
                vsk_flankrec r;
                r.leftlen = (int32_t)strlen(src1);
                r.rightlen = (int32_t)strlen(src3);
                r.left_off = (int64_t)B.pool.size();
                {
                    const std::string fl(src1);
                    B.pool.append(fl.rbegin(), fl.rend());  // left = GetReverse(FL)
                }
                r.right_off = (int64_t)B.pool.size();
                B.pool.append(src3);
                refs.push_back(Tmp{rid, (int)strlen(src4), (int32_t)B.ref_recs.size()});
                B.ref_recs.push_back(r);
                B.ref_ids.push_back(rid);
            } else if (rid > 0) {  // id,first,last,DNA,PAT (flankalign.c:392-535)
                char *src1 = next_field(line, false);
                char *src2 = next_field(src1, false);
                char *src3 = next_field(src2, false);
                char *src4 = next_field(src3, true);  // DNA normalised
                for (char *e = src4 + strlen(src4); e > src4 && (e[-1] == '\n' || e[-1] == '\r'); --e) e[-1] = '\0';
                const long start = strtol(src1, nullptr, 10), end = strtol(src2, nullptr, 10);
                const long length = (long)strlen(src3);
                vsk_flankrec r;
                r.left_off = (int64_t)B.pool.size();
                for (long i = start - 2; i >= 0; --i)
                    if (i < length) B.pool.push_back(src3[i]);
                    else { B.pool.resize((size_t)r.left_off); break; }  // past the terminator: strlen would stop at once
                r.leftlen = (int32_t)(B.pool.size() - (size_t)r.left_off);
                if (r.leftlen > maxflank) { B.pool.resize((size_t)r.left_off + (size_t)maxflank); r.leftlen = maxflank; }
                r.right_off = (int64_t)B.pool.size();
                for (long i = end; i < length; ++i) B.pool.push_back(src3[i]);
                r.rightlen = (int32_t)(B.pool.size() - (size_t)r.right_off);
                if (r.rightlen > maxflank) { B.pool.resize((size_t)r.right_off + (size_t)maxflank); r.rightlen = maxflank; }
                reads.push_back(Tmp{rid, (int)strlen(src4), (int32_t)B.read_recs.size()});
                B.read_recs.push_back(r);
                B.read_ids.push_back(rid);
            }
            got = getline(&line, &cap, fp);
            if (got < 0) { hasdata = false; break; }
            if (line[0] == '@') break;
        }
        // child: EasyListQuickSort(read_list) then (ref_list) from the pristine generator (flankalign.c:559-560)
        PSort sorter;
        std::vector<PItem> rd(reads.size()), rf(refs.size());
        for (size_t i = 0; i < reads.size(); ++i) rd[i] = PItem{reads[i].patsize, (int32_t)i};
        for (size_t i = 0; i < refs.size(); ++i) rf[i] = PItem{refs[i].patsize, (int32_t)i};
        sorter.sort(rd);
        sorter.sort(rf);
        Block blk;
        blk.a1 = a1;
        blk.a2 = a2;
        blk.pair_begin = (int64_t)B.pair_read.size();
        size_t windowstart = 0;
        for (size_t i = 0; i < rd.size(); ++i) {  // flankalign.c:573-616
            const Tmp &q = reads[(size_t)rd[i].idx];
            blk.reads.push_back(q.rec);
            int np = 0;
            for (size_t j = windowstart; j < rf.size(); ++j) {
                const Tmp &r = refs[(size_t)rf[j].idx];
                double sizeerror;
                if (r.patsize > q.patsize) sizeerror = (r.patsize / (double)q.patsize - 1.0) * 100;
                else sizeerror = (q.patsize / (double)r.patsize - 1.0) * 100;
                if (sizeerror > paterr) {
                    if (r.patsize < q.patsize) { windowstart = j; continue; }
                    else break;
                }
                B.pair_read.push_back(q.rec);
                B.pair_ref.push_back(r.rec);
                ++np;
            }
            blk.per_read_pairs.push_back(np);
        }
        if (stats) {
            stats->blocks++;
            stats->pairs += (int64_t)B.pair_read.size() - blk.pair_begin;
        }
        B.blocks.push_back(std::move(blk));
        if ((int64_t)B.pair_read.size() >= pairs_per_batch) rc = flush_batch(ctx, B, maxerrors, stats);
        if (!hasdata) break;
        a1 = a2 = 0;  // flankalign.c:795-802
        sscanf(line, "@(%d_%d)", &a1, &a2);
        if (a1 <= 0 || a2 <= 0) {
            printf("Can't read cluster number. \n%s\n", line);
            rc = VSK_ERR_ABORT;
        }
    }
    if (rc == VSK_OK) rc = flush_batch(ctx, B, maxerrors, stats);
    free(line);
    if (fp != stdin) fclose(fp);
    if (rc == VSK_OK) printf("\ndone!!! processed: %lld, ThreadCounter: 0\n", stats ? (long long)stats->blocks : 0LL);
    return rc;
}

// kernel-level entry for tests and the bench: pairs over explicit flank records
extern "C" int vsk_flank_pairs(vsk_ctx *ctx, const char *pool, int64_t pool_bytes, const vsk_flank *refs, int32_t nref,
                               const vsk_flank *reads, int32_t nread, const int32_t *pair_read, const int32_t *pair_ref,
                               int64_t npairs, int maxerrors, uint32_t *out, float *kernel_ms) {
    if (!ctx || npairs < 0) return VSK_ERR_ARG;
    Batch B;
    B.pool.assign(pool, pool + pool_bytes);
    static_assert(sizeof(vsk_flank) == sizeof(vsk_flankrec), "public flank record mirrors the kernel's");
    B.ref_recs.assign(reinterpret_cast<const vsk_flankrec *>(refs), reinterpret_cast<const vsk_flankrec *>(refs) + nref);
    B.read_recs.assign(reinterpret_cast<const vsk_flankrec *>(reads), reinterpret_cast<const vsk_flankrec *>(reads) + nread);
    for (int64_t i = 0; i < npairs; ++i)
        if (pair_read[i] < 0 || pair_read[i] >= nread || pair_ref[i] < 0 || pair_ref[i] >= nref)
            return vsk_fail(ctx, VSK_ERR_ARG, "flank pair %lld out of range", (long long)i);
    for (const auto *v : {&B.ref_recs, &B.read_recs})
        for (const vsk_flankrec &r : *v)
            if (r.leftlen < 0 || r.rightlen < 0 || r.left_off < 0 || r.right_off < 0 || r.left_off + r.leftlen > pool_bytes ||
                r.right_off + r.rightlen > pool_bytes)
                return vsk_fail(ctx, VSK_ERR_ARG, "flank record outside the string pool");
    // reuse the batch path without blocks: run the kernel and copy the raw results
    VSK_CUDA(ctx, cudaSetDevice(ctx->device));
    B.pair_read.assign(pair_read, pair_read + npairs);
    B.pair_ref.assign(pair_ref, pair_ref + npairs);
    B.ref_ids.assign((size_t)nref, 0);
    B.read_ids.assign((size_t)nread, 0);
    // one pseudo block that owns all pairs so flush_batch has nothing to write
    vsk_flank_stats st;
    memset(&st, 0, sizeof st);
    std::vector<uint32_t> res((size_t)npairs);
    {
        // inline copy of the device part of flush_batch, keeping the results
        const int64_t n = npairs;
        if (n > 0) {
            cudaStream_t s = ctx->stream;
            uint8_t *d_pool = nullptr;
            vsk_flankrec *d_refs = nullptr, *d_reads = nullptr;
            int32_t *d_pr = nullptr, *d_pf = nullptr, *d_lut = nullptr;
            uint32_t *d_out = nullptr;
            int maxlen = 1;
            for (const vsk_flankrec &r : B.read_recs) maxlen = std::max(maxlen, std::max(r.leftlen, r.rightlen));
            for (const vsk_flankrec &r : B.ref_recs) maxlen = std::max(maxlen, std::max(r.leftlen, r.rightlen));
            if (maxlen > VSK_MAXPROF) return vsk_fail(ctx, VSK_ERR_ARG, "flank longer than %d characters", VSK_MAXPROF);
            std::vector<int32_t> lut((size_t)maxlen + 1);
            for (int k = 0; k <= maxlen; ++k) lut[(size_t)k] = std::min(8, (int)(0.4 * k + .01));
            VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_pool, B.pool.size() + 16, ctx->pool, s));
            VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_refs, sizeof(vsk_flankrec) * (B.ref_recs.size() + 1), ctx->pool, s));
            VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_reads, sizeof(vsk_flankrec) * (B.read_recs.size() + 1), ctx->pool, s));
            VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_pr, sizeof(int32_t) * (size_t)n, ctx->pool, s));
            VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_pf, sizeof(int32_t) * (size_t)n, ctx->pool, s));
            VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_lut, sizeof(int32_t) * lut.size(), ctx->pool, s));
            VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_out, sizeof(uint32_t) * (size_t)n, ctx->pool, s));
            VSK_CUDA(ctx, cudaMemcpyAsync(d_pool, B.pool.data(), B.pool.size(), cudaMemcpyHostToDevice, s));
            VSK_CUDA(ctx, cudaMemcpyAsync(d_refs, B.ref_recs.data(), sizeof(vsk_flankrec) * B.ref_recs.size(), cudaMemcpyHostToDevice, s));
            VSK_CUDA(ctx, cudaMemcpyAsync(d_reads, B.read_recs.data(), sizeof(vsk_flankrec) * B.read_recs.size(), cudaMemcpyHostToDevice, s));
            VSK_CUDA(ctx, cudaMemcpyAsync(d_pr, B.pair_read.data(), sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, s));
            VSK_CUDA(ctx, cudaMemcpyAsync(d_pf, B.pair_ref.data(), sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, s));
            VSK_CUDA(ctx, cudaMemcpyAsync(d_lut, lut.data(), sizeof(int32_t) * lut.size(), cudaMemcpyHostToDevice, s));
            int rc2 = vsk_timer_start(ctx);
            if (rc2) return rc2;
            vsk_flankalign_kernel<<<(unsigned)((n + 127) / 128), 128, 0, s>>>(d_pool, d_refs, d_reads, d_pr, d_pf, n,
                                                                           maxerrors, d_lut, d_out);
            ctx->launches++;
            VSK_CUDA(ctx, cudaGetLastError());
            float ms = 0;
            rc2 = vsk_timer_stop(ctx, &ms);
            if (rc2) return rc2;
            if (kernel_ms) *kernel_ms = ms;
            VSK_CUDA(ctx, cudaMemcpyAsync(out, d_out, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToHost, s));
            VSK_CUDA(ctx, cudaStreamSynchronize(s));
            for (void *p : {(void *)d_pool, (void *)d_refs, (void *)d_reads, (void *)d_pr, (void *)d_pf, (void *)d_lut, (void *)d_out})
                cudaFreeAsync(p, s);
        }
    }
    return VSK_OK;
}
