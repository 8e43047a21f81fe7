// refflankalign on the GPU (reference src/newrefflankalign/refflankalign.c).
//
// Reference-vs-reference flank comparison inside a cluster block: references sorted by
// pattern size, every pair i<j until the size error exceeds PATLEN_SIZE_ERR %, four banded
// unit-cost alignments of equal-length flanks per pair
// (narrowbandUnitCostDistanceSameLengthBestScoreLastRowColumnLowMemUsesPointers,
// narrowbandDistanceAlignment.c:533-776), left/right "connected" marks, the XML report and
// representatives.txt.  Host: parsing, the reference's sort order, MinimumRepresentation, the
// report.  Device: one thread per pair, the band held relative to the diagonal.
#include <ctype.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>

#include "vsk_host.hpp"
#include "vsk_internal.cuh"

#define RF_MAXBAND 31  // largest MAXERRORS the kernel holds in its per-thread band (2k+3 cells)

struct RfRec {
    int64_t left_off, right_off;  // into the string pool; left is stored reversed
    int32_t leftlen, rightlen;    // after the REFLEN cut
};

__device__ __forceinline__ uint8_t rf_comp(uint8_t c) {  // GetComplement, refflankalign.c (A<->T, C<->G)
    switch (c) {
        case 'A': return 'T';
        case 'C': return 'G';
        case 'G': return 'C';
        case 'T': return 'A';
    }
    return c;
}

// banded unit-cost DP, -1 unless len1 == len2; cells with |row-col| <= k, 1000 outside;
// score = min over the last row (band cells + the sentinels beside them) and over the last
// column cells of the rows whose band reaches it.  COMP1: s1 is read through rf_comp.
template <bool COMP1>
__device__ int rf_nbdist(const uint8_t *__restrict__ s1, const uint8_t *__restrict__ s2, int len1, int len2, int k) {
    if (len1 != len2) return -1;
    int band[2 * RF_MAXBAND + 3];  // band[d] = cell (row, row - k + d) of the previous row
    const int width = 2 * k + 1;
    // row 0
    for (int d = 0; d < width; ++d) {
        const int col = d - k;
        band[d] = (col >= 0 && col <= len2) ? col : 1000;
    }
    int score = 1000000;
    if (k >= len2) score = min(score, len2);  // row 0 reaches the last column
    for (int row = 1; row <= len1; ++row) {
        const uint8_t a = COMP1 ? rf_comp(s1[row - 1]) : s1[row - 1];
        int left = 1000;  // cell to the left of the band's first cell
        for (int d = 0; d < width; ++d) {
            const int col = row - k + d;
            const int diag = band[d];                                 // (row-1, col-1)
            const int up = (d + 1 < width) ? band[d + 1] : 1000;      // (row-1, col)
            int v;
            if (col < 0 || col > len2) v = 1000;
            else if (col == 0) v = row;
            else {
                const int m = (a == s2[col - 1]) ? 0 : 1;
                v = min(min(diag + m, left + 1), up + 1);
            }
            band[d] = v;
            left = v;
        }
        if (row + k >= len2) score = min(score, band[len2 - row + k]);
    }
    // last row: every band cell, and the 1000 sentinel left of the band when the band does not
    // start at column 0
    for (int d = 0; d < width; ++d) {
        const int col = len1 - k + d;
        if (col >= 0 && col <= len2) score = min(score, band[d]);
    }
    if (len1 - k > 0) score = min(score, 1000);
    return score;
}

// out[4*i..] = lerr1, rerr1, lerr2, rerr2 of pair i (outer a = earlier in sorted order)
__global__ void __launch_bounds__(128)
vsk_refflank_kernel(const uint8_t *__restrict__ pool, const RfRec *__restrict__ recs, const int32_t *__restrict__ pa,
                    const int32_t *__restrict__ pb, int64_t n, int maxerr, int32_t *__restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const RfRec A = recs[pa[i]], B = recs[pb[i]];  // A = read_ptr (outer), B = ref_ptr (inner)
    // refflankalign.c:700-739
    out[4 * i + 0] = rf_nbdist<false>(pool + A.left_off, pool + B.left_off, A.leftlen, max(B.leftlen, A.leftlen), maxerr);
    out[4 * i + 1] = rf_nbdist<false>(pool + A.right_off, pool + B.right_off, A.rightlen, max(B.rightlen, A.rightlen), maxerr);
    out[4 * i + 2] = rf_nbdist<true>(pool + A.right_off, pool + B.left_off, A.rightlen, max(B.leftlen, A.rightlen), maxerr);
    out[4 * i + 3] = rf_nbdist<true>(pool + A.left_off, pool + B.right_off, A.leftlen, max(B.rightlen, A.leftlen), maxerr);
}

// ------------------------------------------------------------------ host side
namespace {
struct RGfsr {  // easylife's generator (easylife.c:107-141), pristine per child process
    uint32_t s[99];
    RGfsr() {
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
struct RItem {
    int patsize;
    int32_t idx;
};
struct RSort {  // EasyListQuickSort (easylife.c:884-1018) on pattern size
    RGfsr rng;
    static void insertion(RItem *a, int n) {
        for (int j = 1; j < n; ++j) {
            const RItem key = a[j];
            int i = j - 1;
            while (i >= 0 && a[i].patsize > key.patsize) { a[i + 1] = a[i]; --i; }
            a[i + 1] = key;
        }
    }
    int partition(RItem *a, int i, int k) {
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
    void helper(RItem *a, int i, int k) {
        while (i < k) {
            const int j = partition(a, i, k);
            if (j - i >= 9) helper(a, i, j);
            else insertion(a + i, j - i + 1);
            i = j + 1;
        }
    }
    void sort(std::vector<RItem> &v) { if (v.size() > 1) helper(v.data(), 0, (int)v.size() - 1); }
};

inline void rf_normalise(char *b, char *e) {
    for (; b < e; ++b) {
        *b = (char)toupper((unsigned char)*b);
        if (*b != 'A' && *b != 'C' && *b != 'G' && *b != 'T') *b = 'N';
    }
}
inline char *rf_next_field(char *p, bool norm) {
    char *b = p;
    while (*p != ',' && *p != '\0') ++p;
    if (norm) rf_normalise(b, p);
    if (*p == ',') { *p = '\0'; ++p; }
    return p;
}

// MinimumRepresentation (refflankalign.c:196-264): smallest rotation of the pattern, or of its
// reverse complement when one of ITS rotations (the unrotated reverse complement itself is never
// tried: the loops start at one rotation) is smaller still
std::string min_representation(const std::string &pat) {
    const size_t len = pat.size();
    if (len == 0) return pat;
    auto compl_rev = [](const std::string &s) {
        std::string r(s.rbegin(), s.rend());
        for (char &c : r) {
            switch (c) {
                case 'A': c = 'T'; break;
                case 'C': c = 'G'; break;
                case 'G': c = 'C'; break;
                case 'T': c = 'A'; break;
                case 'a': c = 't'; break;
                case 'c': c = 'g'; break;
                case 'g': c = 'c'; break;
                case 't': c = 'a'; break;
            }
        }
        return r;
    };
    std::string best = pat, cur = pat, rc = compl_rev(pat);
    for (size_t i = 1; i < len; ++i) {
        std::rotate(cur.begin(), cur.begin() + 1, cur.end());
        if (strcmp(cur.c_str(), best.c_str()) < 0) best = cur;
    }
    for (size_t i = 1; i < len; ++i) {
        std::rotate(rc.begin(), rc.begin() + 1, rc.end());
        if (strcmp(rc.c_str(), best.c_str()) < 0) best = rc;
    }
    return best;
}

struct RfBlock {
    int a1, a2;
    long long read_count;
    std::vector<long long> id;        // sorted order
    std::vector<std::string> pattern, sequence;
    std::vector<int32_t> rec;
    std::vector<int32_t> npairs;      // pairs owned by each outer reference
    int64_t pair_begin;
};
}  // namespace

static int rf_flush(vsk_ctx *ctx, std::string &pool, std::vector<RfRec> &recs, std::vector<int32_t> &pa,
                    std::vector<int32_t> &pb, std::vector<RfBlock> &blocks, int maxerr, int reflen, const char *outdir,
                    const char *xmldir, std::vector<std::string> &rep_lines) {
    const int64_t n = (int64_t)pa.size();
    std::vector<int32_t> out((size_t)n * 4 + 4);
    if (n > 0) {
        cudaStream_t s = ctx->stream;
        uint8_t *d_pool = nullptr;
        RfRec *d_recs = nullptr;
        int32_t *d_pa = nullptr, *d_pb = nullptr, *d_out = nullptr;
        VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_pool, pool.size() + 16, ctx->pool, s));
        VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_recs, sizeof(RfRec) * (recs.size() + 1), ctx->pool, s));
        VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_pa, sizeof(int32_t) * (size_t)n, ctx->pool, s));
        VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_pb, sizeof(int32_t) * (size_t)n, ctx->pool, s));
        VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_out, sizeof(int32_t) * 4 * (size_t)n, ctx->pool, s));
        VSK_CUDA(ctx, cudaMemcpyAsync(d_pool, pool.data(), pool.size(), cudaMemcpyHostToDevice, s));
        VSK_CUDA(ctx, cudaMemcpyAsync(d_recs, recs.data(), sizeof(RfRec) * recs.size(), cudaMemcpyHostToDevice, s));
        VSK_CUDA(ctx, cudaMemcpyAsync(d_pa, pa.data(), sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, s));
        VSK_CUDA(ctx, cudaMemcpyAsync(d_pb, pb.data(), sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, s));
        vsk_refflank_kernel<<<(unsigned)((n + 127) / 128), 128, 0, s>>>(d_pool, d_recs, d_pa, d_pb, n, maxerr, d_out);
        ctx->launches++;
        VSK_CUDA(ctx, cudaGetLastError());
        VSK_CUDA(ctx, cudaMemcpyAsync(out.data(), d_out, sizeof(int32_t) * 4 * (size_t)n, cudaMemcpyDeviceToHost, s));
        VSK_CUDA(ctx, cudaStreamSynchronize(s));
        for (void *p : {(void *)d_pool, (void *)d_recs, (void *)d_pa, (void *)d_pb, (void *)d_out}) cudaFreeAsync(p, s);
    }
    char tmp[4200];
    for (const RfBlock &b : blocks) {
        const size_t nr = b.id.size();
        std::vector<int> count(nr, 0), scoresum(nr, 0), comparisons(nr, 0);
        std::vector<char> leftcon(nr, 0), rightcon(nr, 0);
        snprintf(tmp, sizeof tmp, "%s/%d_%d.xml", xmldir, b.a1, b.a2);
        FILE *fp = fopen(tmp, "w");
        if (!fp) return vsk_fail(ctx, VSK_ERR_ABORT, "RefFlankAlign - can't open output XML file (%s). Aborting!", tmp);
        fprintf(fp, "<?xml-stylesheet type=\"text/xsl\" href=\"../family.xsl\"?>\n");
        fprintf(fp, "<cluster trim=\"%d\" read_count=\"%lld\" ref_count=\"%zu\" report=\"all\" maxedits=\"%d\" id=\"%d_%d\">\n",
                reflen, b.read_count, nr, maxerr, b.a1, b.a2);
        if (nr == 0) {  // refflankalign.c:646-652: the child writes this and quits
            fprintf(fp, "\n\nRefFlankAlign - empty reference list, this should not happen(%d_%d). Aborting!\n\n", b.a1, b.a2);
            fclose(fp);
            continue;
        }
        int64_t p = b.pair_begin;
        bool aborted = false;
        for (size_t i = 0; i < nr && !aborted; ++i) {
            for (int k = 0; k < b.npairs[i]; ++k, ++p) {
                const size_t j = i + 1 + (size_t)k;
                const int lerr1 = out[4 * p], rerr1 = out[4 * p + 1], lerr2 = out[4 * p + 2], rerr2 = out[4 * p + 3];
                if (lerr1 < 0 || rerr1 < 0 || lerr2 < 0 || rerr2 < 0) {  // refflankalign.c:714-720,741-747
                    fprintf(fp, "\n\nRefFlankAlign - wrong sizes for narrowband alignment, seq2 must be larger. Aborting!\n\n");
                    aborted = true;
                    break;
                }
                const int errors = std::min(std::max(lerr1, rerr1), std::max(lerr2, rerr2));
                scoresum[i] += std::min(lerr1 + rerr1, lerr2 + rerr2);
                comparisons[i]++;
                if (lerr1 <= maxerr || lerr2 <= maxerr) { leftcon[j] = 1; leftcon[i] = 1; }
                if (rerr1 <= maxerr || rerr2 <= maxerr) { rightcon[j] = 1; rightcon[i] = 1; }
                if (errors <= maxerr) count[i]++;
            }
        }
        if (aborted) {
            fclose(fp);
            continue;
        }
        int dist = 0, undist = 0, leftc = 0, rightc = 0, anyc = 0;
        for (size_t i = 0; i < nr; ++i) {
            if (leftcon[i]) leftc++;
            if (rightcon[i]) rightc++;
            if (!leftcon[i] && !rightcon[i]) dist++;
            if (leftcon[i] && rightcon[i]) undist++;
            if (leftcon[i] || rightcon[i]) anyc++;
        }
        auto nodes = [&](const char *indent, int cnt, auto pred) {
            fprintf(fp, "%s<nodes count=\"%d\">", indent, cnt);
            int w = 0;
            for (size_t i = 0; i < nr; ++i)
                if (pred(i)) {
                    if (w >= 1) fprintf(fp, " ");
                    fprintf(fp, "%lld", b.id[i]);
                    w++;
                }
            fprintf(fp, "</nodes>\n");
        };
        fprintf(fp, "  <classes>\n");
        fprintf(fp, "    <class component_count=\"%d\" name=\"distinguishable\" ref_count=\"%d\" density=\"None\">\n", dist ? 1 : 0, dist);
        if (dist) nodes("      ", dist, [&](size_t i) { return !leftcon[i] && !rightcon[i]; });
        fprintf(fp, "    </class>\n");
        auto klass = [&](const char *name, int cnt, auto pred) {
            fprintf(fp, "    <class name=\"%s\" ref_count=\"%d\" density=\"None\" component_count=\"%d\">\n", name, cnt, cnt ? 1 : 0);
            if (cnt) {
                fprintf(fp, "      <component density=\"None\">\n");
                nodes("        ", cnt, pred);
                fprintf(fp, "      </component>\n");
            }
            fprintf(fp, "    </class>\n");
        };
        klass("left-connected", leftc, [&](size_t i) { return (bool)leftcon[i]; });
        klass("right-connected", rightc, [&](size_t i) { return (bool)rightcon[i]; });
        klass("left-or-right connected", anyc, [&](size_t i) { return leftcon[i] || rightcon[i]; });
        klass("undistinguishable", undist, [&](size_t i) { return leftcon[i] && rightcon[i]; });
        fprintf(fp, "  </classes>\n");
        fprintf(fp, "  <references>\n");
        int largest = 0, ssum = 0;
        long long totlen = 0;
        size_t largest_i = 0;
        for (size_t i = 0; i < nr; ++i) {
            if (count[i] > largest) { largest = count[i]; largest_i = i; }
            ssum += scoresum[i];
            const RfRec &r = recs[(size_t)b.rec[i]];
            totlen += (long long)comparisons[i] * (r.leftlen + r.rightlen);
            fprintf(fp, "    <reference>\n      <pattern>%s</pattern>\n      <array>%s</array>\n    </reference>\n",
                    b.pattern[i].c_str(), b.sequence[i].c_str());
        }
        char scorebf[40];
        if (totlen == 0) strcpy(scorebf, "None");
        else snprintf(scorebf, sizeof scorebf, "%.2lf", 1.0 - ssum / totlen);  // integer division, refflankalign.c:965
        {   // the parent copies each .rep line back with fgets(tempbuf, 1000, ...) (refflankalign.c:1053-1055)
            std::string l = std::to_string(b.a1) + "_" + std::to_string(b.a2) + "\t" + scorebf + "\t" + b.pattern[largest_i];
            if (l.size() > 999) l.resize(999);
            rep_lines.push_back(l);
        }
        fprintf(fp, "  </references>\n</cluster>\n");
        fclose(fp);
        printf("processing: %d_%d (gpu)\n", b.a1, b.a2);
    }
    (void)outdir;
    pool.clear(); recs.clear(); pa.clear(); pb.clear(); blocks.clear();
    return VSK_OK;
}

// refflankalign.exe outdir xmldir inputfile MAXERRORS MAXTHREADS PATLEN_SIZE_ERR REFLEN
extern "C" int vsk_refflankalign_run(vsk_ctx *ctx, const char *outdir, const char *xmldir, const char *inputfile,
                                     int maxerrors, int paterr, int reflen, int64_t *n_blocks, int64_t *n_pairs) {
    if (!ctx || !outdir || !xmldir || !inputfile || paterr < 1 || reflen < 1) return VSK_ERR_ARG;
    if (maxerrors < 0 || maxerrors > RF_MAXBAND)
        return vsk_fail(ctx, VSK_ERR_ARG, "MAXERRORS %d outside 0..%d supported by the GPU build", maxerrors, RF_MAXBAND);
    VSK_CUDA(ctx, cudaSetDevice(ctx->device));
    std::string rep_path = std::string(outdir) + "/representatives.txt";
    FILE *RFP = fopen(rep_path.c_str(), "w");
    if (!RFP) return vsk_fail(ctx, VSK_ERR_ABORT, "RefFlankAlign - can't open output representative file (%s). Aborting!", rep_path.c_str());
    FILE *fp = (inputfile[0] == '-') ? stdin : fopen(inputfile, "r");
    if (!fp) { fclose(RFP); return vsk_fail(ctx, VSK_ERR_ABORT, "RefFlankAlign - Could not open input file. Aborting!"); }
    char *line = nullptr;
    size_t cap = 0;
    int a1 = 0, a2 = 0;
    bool hasdata = false;
    while (getline(&line, &cap, fp) >= 0)
        if (line[0] == '@') {
            sscanf(line, "@(%d_%d)", &a1, &a2);
            if (a1 <= 0 || a2 <= 0) { printf("Can't read cluster number."); free(line); fclose(RFP); return VSK_ERR_ABORT; }
            hasdata = true;
            break;
        }
    std::string pool;
    std::vector<RfRec> recs;
    std::vector<int32_t> pa, pb;
    std::vector<RfBlock> blocks;
    std::vector<std::string> rep_lines;
    int rc = VSK_OK;
    int64_t nblocks = 0, npairs = 0;
    struct Tmp { long long id; int patsize; int32_t rec; std::string pattern, sequence; };
    while (hasdata && rc == VSK_OK) {
        std::vector<Tmp> refs;
        long long read_count = 0;
        for (;;) {
            const long long rid = strtoll(line, nullptr, 10);
            if (rid < 0) {  // refflankalign.c:432-505
                char *src1 = rf_next_field(line, false);
                char *src2 = rf_next_field(src1, true);   // left flank, normalised
                char *src3 = rf_next_field(src2, false);  // array sequence
                char *src4 = rf_next_field(src3, true);   // right flank, normalised
                for (char *e = src4 + strlen(src4); e > src4 && (e[-1] == '\n' || e[-1] == '\r'); --e) e[-1] = '\0';
                RfRec r;
                r.leftlen = std::min((int32_t)strlen(src1), (int32_t)reflen);
                r.rightlen = std::min((int32_t)strlen(src3), (int32_t)reflen);
                r.left_off = (int64_t)pool.size();
                {
                    const std::string fl(src1);
                    pool.append(fl.rbegin(), fl.rend());
                }
                r.right_off = (int64_t)pool.size();
                pool.append(src3);
                refs.push_back(Tmp{rid, (int)strlen(src4), (int32_t)recs.size(), std::string(src4), std::string(src2)});
                recs.push_back(r);
            } else if (rid > 0) {
                read_count++;
            }
            if (getline(&line, &cap, fp) < 0) { hasdata = false; break; }
            if (line[0] == '@') break;
        }
        RSort sorter;
        std::vector<RItem> order(refs.size());
        for (size_t i = 0; i < refs.size(); ++i) order[i] = RItem{refs[i].patsize, (int32_t)i};
        sorter.sort(order);
        RfBlock blk;
        blk.a1 = a1;
        blk.a2 = a2;
        blk.read_count = read_count;
        blk.pair_begin = (int64_t)pa.size();
        for (size_t i = 0; i < order.size(); ++i) {
            const Tmp &t = refs[(size_t)order[i].idx];
            blk.id.push_back(t.id);
            blk.pattern.push_back(min_representation(t.pattern));
            blk.sequence.push_back(t.sequence);
            blk.rec.push_back(t.rec);
        }
        for (size_t i = 0; i < order.size(); ++i) {  // refflankalign.c:654-698
            const Tmp &A = refs[(size_t)order[i].idx];
            int np = 0;
            for (size_t j = i + 1; j < order.size(); ++j) {
                const Tmp &Bj = refs[(size_t)order[j].idx];
                double sizeerror;
                if (Bj.patsize > A.patsize) sizeerror = (Bj.patsize / (double)A.patsize - 1.0) * 100;
                else sizeerror = (A.patsize / (double)Bj.patsize - 1.0) * 100;
                if (sizeerror > paterr) break;
                pa.push_back(A.rec);
                pb.push_back(Bj.rec);
                ++np;
            }
            blk.npairs.push_back(np);
        }
        npairs += (int64_t)pa.size() - blk.pair_begin;
        nblocks++;
        blocks.push_back(std::move(blk));
        if (pa.size() >= (size_t)(2 << 20)) rc = rf_flush(ctx, pool, recs, pa, pb, blocks, maxerrors, reflen, outdir, xmldir, rep_lines);
        if (!hasdata) break;
        a1 = a2 = 0;
        sscanf(line, "@(%d_%d)", &a1, &a2);
        if (a1 <= 0 || a2 <= 0) {
            printf("Can't read cluster number. \n%s\n", line);
            rc = VSK_ERR_ABORT;
        }
    }
    if (rc == VSK_OK) rc = rf_flush(ctx, pool, recs, pa, pb, blocks, maxerrors, reflen, outdir, xmldir, rep_lines);
    free(line);
    if (fp != stdin) fclose(fp);
    for (const std::string &l : rep_lines) fprintf(RFP, "%s\n", l.c_str());
    fclose(RFP);
    if (n_blocks) *n_blocks = nblocks;
    if (n_pairs) *n_pairs = npairs;
    if (rc == VSK_OK) printf("\ndone!!! processed: %lld, ThreadCounter: 0\n", (long long)nblocks);
    return rc;
}

// kernel-level entry: the banded distance of explicit string pairs (tests)
extern "C" int vsk_nbdist_pairs(vsk_ctx *ctx, const char *a, const int64_t *a_off, const char *b, const int64_t *b_off,
                                int64_t n, int maxerr, int32_t *out) {
    if (!ctx || n < 0 || maxerr < 0 || maxerr > RF_MAXBAND) return VSK_ERR_ARG;
    if (n == 0) return VSK_OK;
    std::string pool(a, a + a_off[n]);
    const int64_t boff = (int64_t)pool.size();
    pool.append(b, b + b_off[n]);
    // pair i: record 2i = (a_i as both flanks), 2i+1 = (b_i as both flanks); only lerr1 is reported
    std::vector<RfRec> recs;
    std::vector<int32_t> pa, pb;
    for (int64_t i = 0; i < n; ++i) {
        RfRec ra, rb;
        ra.left_off = ra.right_off = a_off[i];
        ra.leftlen = ra.rightlen = (int32_t)(a_off[i + 1] - a_off[i]);
        rb.left_off = rb.right_off = boff + b_off[i];
        rb.leftlen = rb.rightlen = (int32_t)(b_off[i + 1] - b_off[i]);
        recs.push_back(ra);
        recs.push_back(rb);
        pa.push_back((int32_t)(2 * i));
        pb.push_back((int32_t)(2 * i + 1));
    }
    VSK_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    uint8_t *d_pool = nullptr;
    RfRec *d_recs = nullptr;
    int32_t *d_pa = nullptr, *d_pb = nullptr, *d_out = nullptr;
    std::vector<int32_t> res((size_t)n * 4);
    VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_pool, pool.size() + 16, ctx->pool, s));
    VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_recs, sizeof(RfRec) * recs.size(), ctx->pool, s));
    VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_pa, sizeof(int32_t) * (size_t)n, ctx->pool, s));
    VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_pb, sizeof(int32_t) * (size_t)n, ctx->pool, s));
    VSK_CUDA(ctx, cudaMallocFromPoolAsync(&d_out, sizeof(int32_t) * 4 * (size_t)n, ctx->pool, s));
    VSK_CUDA(ctx, cudaMemcpyAsync(d_pool, pool.data(), pool.size(), cudaMemcpyHostToDevice, s));
    VSK_CUDA(ctx, cudaMemcpyAsync(d_recs, recs.data(), sizeof(RfRec) * recs.size(), cudaMemcpyHostToDevice, s));
    VSK_CUDA(ctx, cudaMemcpyAsync(d_pa, pa.data(), sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, s));
    VSK_CUDA(ctx, cudaMemcpyAsync(d_pb, pb.data(), sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, s));
    vsk_refflank_kernel<<<(unsigned)((n + 127) / 128), 128, 0, s>>>(d_pool, d_recs, d_pa, d_pb, n, maxerr, d_out);
    ctx->launches++;
    VSK_CUDA(ctx, cudaGetLastError());
    VSK_CUDA(ctx, cudaMemcpyAsync(res.data(), d_out, sizeof(int32_t) * 4 * (size_t)n, cudaMemcpyDeviceToHost, s));
    VSK_CUDA(ctx, cudaStreamSynchronize(s));
    for (void *p : {(void *)d_pool, (void *)d_recs, (void *)d_pa, (void *)d_pb, (void *)d_out}) cudaFreeAsync(p, s);
    for (int64_t i = 0; i < n; ++i)
        out[i] = (a_off[i + 1] - a_off[i] != b_off[i + 1] - b_off[i]) ? -1 : res[(size_t)4 * i];  // len1 != len2 (.c:549-551)
    return VSK_OK;
}
