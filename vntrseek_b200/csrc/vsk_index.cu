// Host build of the reference seed index (psearch.c:161-766 + hclust.h:98-270),
// reproducing the reference's bucket order exactly.
//
// Why the order matters: a read's candidate window inside a seed bucket runs from
// the first hit with pmin >= lo up to AND INCLUDING the first hit with pmin > hi
// (hclust.h:380-421).  Which hit is "first" inside an equal-pmin group is decided
// by the reference's unstable quicksort (EasyArrayQuickSort, easylife.c:884-981),
// whose median-of-three pivots come from a fixed-state GFSR generator
// (easylife.c:107-141) shared by all sort calls of the process.  So the sort,
// the generator and the order in which buckets are sorted (ascending code for
// TRANGE <= 4; the 2-key hash's walker list, doublehash.h:334-392, for the rest)
// are restated here.  Only the identity of the first hit of every pmin group
// survives into the device index; the rest of a bucket is a set.
#include <time.h>
#include <math.h>
#include <string.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <thread>

#include "vsk_host.hpp"
#include "vsk_internal.cuh"

static const char *kSeeds[15] = {
    nullptr,
    "111*1",
    "1111*11",
    "11111*111",
    "11111*1111",
    "111111*11111*11",
    "111111*111111*11",
    "111*11111*111*111*11",
    "11*111*11111**11**111*11*11",
    "11*111*11111**11**111*11*11",
    "11*111*11111**11**111*11*11",
    "11*111*11111**11**111*11*11",
    "1**11***1111**11*11**1*11*11*111**11",
    "1**11***1111**11*11**1*11*11*111**11**111",
    "1**11***1111**11*11**1*11*11*111**11**111*1*11",
};

const vsk_seedinfo &vsk_seeds() {
    static vsk_seedinfo si;
    static bool init = false;
    if (!init) {
        memset(&si, 0, sizeof si);
        for (int t = 1; t <= 14; ++t) {
            const int len = (int)strlen(kSeeds[t]);
            si.tsize[t] = len;
            int k = 0;
            for (int i = 0; i < len; ++i)
                if (kSeeds[t][i] == '1') si.off[t][k++] = i - (len - 1);  // retrieve_seed_info, hclust.h:1135-1159
            si.nones[t] = k;
        }
        static const int lim[14] = {6, 10, 17, 25, 30, 39, 59, 89, 134, 201, 303, 474, 715, 1111};
        for (int p = 0; p <= 5000; ++p) {  // psearch.c:284-318
            int t = 0;
            for (int k = 0; k < 14; ++k)
                if (p >= lim[k]) t = k + 1;
            si.trange_of[p] = (int16_t)t;
            si.err_of[p] = (int16_t)(int)(p * 0.10 + 1.0);  // PATLEN_SIZE_ERR_FRACTION, TRUNC_ROUND_CEIL
        }
        init = true;
    }
    return si;
}

// ---- the reference's generator and sort, restated -----------------------------
namespace {
struct Gfsr {
    uint32_t s[99];
    Gfsr() {
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
    double next() {
        uint32_t p = s[0];
        if (++p > 98) p = 1;
        uint32_t k = p + 27;
        if (k > 98) k -= 98;
        s[p] ^= s[k];
        s[0] = p;
        return (double)s[p] / 2147483647.0;
    }
    int range(int lo, int hi) {  // EASYLIFE_INT_RANDRANGE + _el_roundG
        const double d = next() * (hi - lo + 1) + lo - .5;
        int i = (int)d;
        const double g = d - i;
        if (g >= .5) i++;
        else if (g <= -.5) i--;
        if (i < lo) i = lo;
        if (i > hi) i = hi;
        return i;
    }
};

struct SortItem {
    int16_t pmin;
    int32_t prof;
};

struct EasySort {
    Gfsr rng;
    static void insertion(SortItem *a, int n) {
        for (int j = 1; j < n; ++j) {
            const SortItem key = a[j];
            int i = j - 1;
            while (i >= 0 && a[i].pmin > key.pmin) { a[i + 1] = a[i]; --i; }
            a[i + 1] = key;
        }
    }
    int partition(SortItem *a, int i, int k) {
        int r0 = rng.range(i, k), r1 = rng.range(i, k), r2 = rng.range(i, k);
        if (r0 > r1) std::swap(r0, r1);
        if (r1 > r2) std::swap(r1, r2);
        const int pval = a[r1].pmin;
        --i;
        ++k;
        for (;;) {
            do { --k; } while (a[k].pmin > pval);
            do { ++i; } while (a[i].pmin < pval);
            if (i >= k) break;
            std::swap(a[i], a[k]);
        }
        return k;
    }
    void helper(SortItem *a, int i, int k) {
        while (i < k) {
            const int j = partition(a, i, k);
            if (j - i >= 9) helper(a, i, j);
            else insertion(a + i, j - i + 1);
            i = j + 1;
        }
    }
    void sort(SortItem *a, int n) { helper(a, 0, n - 1); }
};

inline uint32_t rjhash(uint32_t a) {  // doublehash.h:82-90
    a = (a + 0x7ed55d16) + (a << 12);
    a = (a ^ 0xc761c23c) ^ (a >> 19);
    a = (a + 0x165667b1) + (a << 5);
    a = (a + 0xd3a2646c) ^ (a << 9);
    a = (a + 0xfd7046c5) + (a << 3);
    a = (a ^ 0xb55a4f09) ^ (a >> 16);
    return a;
}

// open-addressing map (trange, code) -> dense bucket id in creation order
struct BucketMap {
    std::vector<uint64_t> keys;
    std::vector<int32_t> vals;
    std::vector<uint8_t> tr;
    size_t mask = 0, used = 0;
    void init(size_t cap) {
        size_t c = 1;
        while (c < cap) c <<= 1;
        keys.assign(c, 0);
        vals.assign(c, -1);
        tr.assign(c, 0);
        mask = c - 1;
        used = 0;
    }
    void grow() {
        std::vector<uint64_t> ok;
        std::vector<int32_t> ov;
        std::vector<uint8_t> ot;
        ok.swap(keys); ov.swap(vals); ot.swap(tr);
        init(ok.size() * 2);
        for (size_t i = 0; i < ok.size(); ++i)
            if (ov[i] >= 0) { bool nw; *find(ot[i], ok[i], nw) = ov[i]; }
    }
    int32_t *find(int t, uint64_t code, bool &is_new) {
        if ((used + 1) * 10 > (mask + 1) * 7) grow();
        uint64_t h = (code ^ ((uint64_t)t << 58)) * 0x9E3779B97F4A7C15ull;
        size_t i = (size_t)(h >> 20) & mask;
        for (;;) {
            if (vals[i] < 0) { keys[i] = code; tr[i] = (uint8_t)t; is_new = true; ++used; return &vals[i]; }
            if (keys[i] == code && tr[i] == t) { is_new = false; return &vals[i]; }
            i = (i + 1) & mask;
        }
    }
};

inline int base_index(char c) {
    switch (c) {
        case 'A': return 0;
        case 'C': return 1;
        case 'G': return 2;
        case 'T': return 3;
    }
    return -1;
}
}  // namespace

// seed code of the window that ends at position i (hclust.h:157-205): the first
// 15 ones go to seedcode (high word), the rest to seedcode2 (low word)
static inline uint64_t seed_code(const char *seq, int i, const vsk_seedinfo &si, int t) {
    uint32_t c1 = 0, c2 = 0;
    for (int k = 0; k < si.nones[t]; ++k) {
        const int b = base_index(seq[i + si.off[t][k]]);
        if (k <= 14) c1 = c1 * 4 + (uint32_t)b;
        else c2 = c2 * 4 + (uint32_t)b;
    }
    return ((uint64_t)c1 << 32) | c2;
}

static double now_s() {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + ts.tv_nsec * 1e-9;
}

// The straightforward one-thread build: hits appended bucket by bucket through a hash map, the way
// the reference does it.  Kept as the cross-check of the threaded build below (vsk_index_cache_verify
// rebuilds with this one; VSK_INDEX_SERIAL=1 selects it everywhere).
static int build_refindex_serial(const vsk_trset &refs, vsk_refindex_host &out, std::string &err) {
    const vsk_seedinfo &si = vsk_seeds();
    const bool timing = getenv("VSK_INDEX_TIMING") != nullptr;
    double t_phase = now_s();
    auto phase = [&](const char *name) {
        if (!timing) return;
        const double t = now_s();
        fprintf(stderr, "[vsk index] %-10s %.2f s\n", name, t - t_phase);
        t_phase = t;
    };
    struct Bucket {
        int t;
        uint64_t code;
        int64_t count = 0, start = 0, fill = 0;
    };
    std::vector<Bucket> buckets;
    std::vector<int32_t> hit_bucket;  // per hit, in insertion order
    std::vector<SortItem> hit_item;
    BucketMap map;
    map.init(1 << 20);
    uint64_t tuples = 0;

    auto seed_sequence = [&](int32_t prof, int t, int pmin) {
        const char *seq = refs.cons.data() + refs.cons_off[prof];
        const int len = (int)(refs.cons_off[prof + 1] - refs.cons_off[prof]);
        const int ts = si.tsize[t];
        for (int i = ts - 1; i < len; ++i) {
            const uint64_t code = seed_code(seq, i, si, t);
            bool is_new;
            int32_t *slot = map.find(t, code, is_new);
            if (is_new) {
                *slot = (int32_t)buckets.size();
                Bucket b;
                b.t = t;
                b.code = code;
                buckets.push_back(b);
            }
            buckets[(size_t)*slot].count++;
            hit_bucket.push_back(*slot);
            hit_item.push_back(SortItem{(int16_t)pmin, prof});
            ++tuples;
        }
    };

    for (int32_t r = 0; r < refs.n; ++r) {
        const int pmin = refs.patlen[r];
        if (pmin < 7 || pmin > 5000) {  // psearch.c:518-521
            err = "pmin(" + std::to_string(pmin) + ") must be in 7-5000 range!";
            return VSK_ERR_ABORT;
        }
        if (!refs.master[r]) continue;  // psearch.c:452-471: non-masters leave the scan list
        const int t = si.trange_of[pmin];
        if (t <= 0) continue;
        for (int o = 0; o < 2; ++o) {
            const int len = (int)(refs.cons_off[2 * r + o + 1] - refs.cons_off[2 * r + o]);
            if (si.tsize[t] > len) {  // seqdupPlusTuple, hclust.h:866-869
                err = "Tuple size (" + std::to_string(si.tsize[t]) + ") is larger than sequence(" +
                      std::to_string(len) + "). Aborting!";
                return VSK_ERR_ABORT;
            }
        }
        seed_sequence(2 * r, t, pmin);
        seed_sequence(2 * r + 1, t, pmin);
        const int e = si.err_of[pmin];
        const int lower = pmin - e;
        if (lower >= 7 && si.trange_of[lower] <= t - 1) {  // psearch.c:552-577
            for (int o = 0; o < 2; ++o)
                if (si.tsize[t - 1] > (int)(refs.cons_off[2 * r + o + 1] - refs.cons_off[2 * r + o])) {
                    err = "Tuple size is larger than sequence. Aborting!";
                    return VSK_ERR_ABORT;
                }
            seed_sequence(2 * r, t - 1, pmin);
            seed_sequence(2 * r + 1, t - 1, pmin);
        }
        const int higher = pmin + e;
        if (higher <= 5000 && si.trange_of[higher] >= t + 1) {  // psearch.c:583-607
            for (int o = 0; o < 2; ++o)
                if (si.tsize[t + 1] > (int)(refs.cons_off[2 * r + o + 1] - refs.cons_off[2 * r + o])) {
                    err = "Tuple size is larger than sequence. Aborting!";
                    return VSK_ERR_ABORT;
                }
            seed_sequence(2 * r, t + 1, pmin);
            seed_sequence(2 * r + 1, t + 1, pmin);
        }
    }

    phase("seed");
    // scatter hits into bucket-contiguous storage, insertion order kept
    const size_t nb = buckets.size(), nh = hit_item.size();
    int64_t acc = 0;
    for (size_t b = 0; b < nb; ++b) { buckets[b].start = acc; acc += buckets[b].count; }
    std::vector<SortItem> items(nh);
    for (size_t h = 0; h < nh; ++h) {
        Bucket &b = buckets[(size_t)hit_bucket[h]];
        items[(size_t)(b.start + b.fill++)] = hit_item[h];
    }
    std::vector<int32_t>().swap(hit_bucket);
    std::vector<SortItem>().swap(hit_item);

    phase("scatter");
    if (timing) {
        size_t multi = 0, multi_hits = 0;
        for (size_t b = 0; b < nb; ++b) if (buckets[b].count > 1) { ++multi; multi_hits += (size_t)buckets[b].count; }
        fprintf(stderr, "[vsk index] hits %zu buckets %zu multi-hit buckets %zu (hits in them %zu)\n", nh, nb, multi, multi_hits);
    }
    // the order in which the reference sorts its buckets (psearch.c:623-648)
    std::vector<int32_t> order;
    order.reserve(nb);
    {
        std::vector<int32_t> byt[15];
        for (size_t b = 0; b < nb; ++b) byt[buckets[b].t].push_back((int32_t)b);  // creation order
        const size_t hsize = 1162687;  // EasyLifeGetPrime(1000000), easylife.c:1620-1645
        for (int t = 1; t <= 14; ++t) {
            std::vector<int32_t> &v = byt[t];
            if (t <= 4) {
                std::sort(v.begin(), v.end(),
                          [&](int32_t a, int32_t b) { return buckets[(size_t)a].code < buckets[(size_t)b].code; });
                order.insert(order.end(), v.begin(), v.end());
            } else {
                // walker = rack slots in reverse order of first use; chain = reverse creation order
                std::unordered_map<uint32_t, std::vector<int32_t>> racks;
                std::vector<uint32_t> first_use;
                for (int32_t b : v) {
                    const uint64_t code = buckets[(size_t)b].code;
                    const uint32_t hk = (uint32_t)((uint32_t)(rjhash((uint32_t)(code >> 32)) + rjhash((uint32_t)code)) % hsize);
                    auto it = racks.find(hk);
                    if (it == racks.end()) { first_use.push_back(hk); racks[hk].push_back(b); }
                    else it->second.push_back(b);
                }
                for (size_t k = first_use.size(); k-- > 0;) {
                    const std::vector<int32_t> &chain = racks[first_use[k]];
                    for (size_t c = chain.size(); c-- > 0;) order.push_back(chain[c]);
                }
            }
        }
    }
    phase("order");
    EasySort sorter;
    for (int32_t b : order) {
        Bucket &bk = buckets[(size_t)b];
        if (bk.count > 1) sorter.sort(items.data() + bk.start, (int)bk.count);
    }

    phase("sort");
    // flatten: per TRANGE codes ascending; entries = distinct (pmin, prof) + multiplicity + group head
    out = vsk_refindex_host();
    out.tuples = tuples;
    out.total_hits = (int64_t)nh;
    std::vector<int32_t> byt[15];
    for (size_t b = 0; b < nb; ++b) byt[buckets[b].t].push_back((int32_t)b);
    std::vector<std::pair<int32_t, int32_t>> tmp;  // (prof, count) scratch
    for (int t = 1; t <= 14; ++t) {
        std::vector<int32_t> &v = byt[t];
        std::sort(v.begin(), v.end(),
                  [&](int32_t a, int32_t b) { return buckets[(size_t)a].code < buckets[(size_t)b].code; });
        out.codes[t].reserve(v.size());
        out.bucket_off[t].reserve(v.size() + 1);
        for (int32_t b : v) {
            const Bucket &bk = buckets[(size_t)b];
            out.codes[t].push_back(bk.code);
            out.bucket_off[t].push_back((int64_t)out.e_prof.size());
            const SortItem *a = items.data() + bk.start;
            int64_t g = 0, cum = 0;
            while (g < bk.count) {
                int64_t h = g;
                while (h < bk.count && a[h].pmin == a[g].pmin) ++h;
                const int32_t first = a[g].prof;
                std::vector<int32_t> profs;
                profs.reserve((size_t)(h - g));
                for (int64_t k = g; k < h; ++k) profs.push_back(a[k].prof);
                std::sort(profs.begin(), profs.end());
                for (size_t k = 0; k < profs.size();) {
                    size_t m = k;
                    while (m < profs.size() && profs[m] == profs[k]) ++m;
                    out.e_pmin.push_back(a[g].pmin);
                    out.e_prof.push_back(profs[k]);
                    out.e_mult.push_back((int32_t)(m - k));
                    out.e_first.push_back(first);
                    out.e_cum.push_back(cum);
                    cum += (int64_t)(m - k);
                    k = m;
                }
                g = h;
            }
        }
        out.bucket_off[t].push_back((int64_t)out.e_prof.size());
    }
    phase("flatten");
    return VSK_OK;
}

// ---- the same index, built on several threads -------------------------------------------------
// Nothing in the result depends on how the hits get into their buckets, only on (a) the hits of a
// bucket in insertion order, (b) the order in which the reference sorts its buckets (it decides
// where the shared GFSR generator stands when a bucket is partitioned) and (c) the final layout.
// So: every hit becomes a record (key = TRANGE | seed code, stamp = its insertion number), the
// records are split into key ranges by sampled splitters (stable scatter), each range is sorted
// stably by key on its own thread -- a bucket is then a run of equal keys, in insertion order,
// whose first stamp says when the reference created it -- the (few) buckets with more than one
// hit are put in the reference's sorting order and partitioned one after the other with the
// restated generator (the only serial step: 0.6 s for 250 K references), and the flattened arrays
// are written range by range.
namespace {
struct HitRec {
    uint64_t key;    // trange << 60 | seedcode << 30 | seedcode2 : sorts like (trange, code)
    uint32_t stamp;  // insertion number of the hit
    int32_t prof;
    int16_t pmin;
};

inline uint64_t rec_code(uint64_t key) {  // back to the (seedcode << 32 | seedcode2) the index stores
    return (((key >> 30) & 0x3fffffffull) << 32) | (key & 0x3fffffffull);
}

struct MultiBucket {
    int64_t start;   // first record
    int32_t count;
    uint32_t stamp;  // creation = stamp of its first hit
    uint64_t key;
};

// runs fn(i) for i in [0, n) on `nt` threads, work handed out in chunks
template <typename F>
void parallel_for(int64_t n, int nt, int64_t chunk, F fn) {
    if (nt <= 1 || n <= chunk) {
        for (int64_t i = 0; i < n; ++i) fn(i);
        return;
    }
    std::atomic<int64_t> next(0);
    std::vector<std::thread> th;
    auto body = [&] {
        for (;;) {
            const int64_t b = next.fetch_add(chunk);
            if (b >= n) break;
            const int64_t e = std::min(n, b + chunk);
            for (int64_t i = b; i < e; ++i) fn(i);
        }
    };
    for (int k = 1; k < nt; ++k) th.emplace_back(body);
    body();
    for (auto &t : th) t.join();
}
}  // namespace

int vsk_build_refindex_host(const vsk_trset &refs, vsk_refindex_host &out, std::string &err) {
    if (const char *e = getenv("VSK_INDEX_SERIAL"))
        if (e[0] == '1') return build_refindex_serial(refs, out, err);
    const vsk_seedinfo &si = vsk_seeds();
    const bool timing = getenv("VSK_INDEX_TIMING") != nullptr;
    int nt = (int)std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 64u);
    if (const char *e = getenv("VSK_INDEX_THREADS")) nt = std::max(1, std::min(256, atoi(e)));
    double t_phase = now_s();
    auto phase = [&](const char *name) {
        if (!timing) return;
        const double t = now_s();
        fprintf(stderr, "[vsk index] %-10s %.2f s\n", name, t - t_phase);
        t_phase = t;
    };

    // ---- which (profile, TRANGE) seedings happen, in the reference's order, and where their hits start
    struct Seeding { int32_t prof; int16_t t, pmin; int64_t base; };
    std::vector<Seeding> seedings;
    int64_t nh = 0;
    auto add_seeding = [&](int32_t prof, int t, int pmin) {
        const int len = (int)(refs.cons_off[prof + 1] - refs.cons_off[prof]);
        seedings.push_back(Seeding{prof, (int16_t)t, (int16_t)pmin, nh});
        nh += std::max(0, len - si.tsize[t] + 1);
    };
    for (int32_t r = 0; r < refs.n; ++r) {
        const int pmin = refs.patlen[r];
        if (pmin < 7 || pmin > 5000) {  // psearch.c:518-521
            err = "pmin(" + std::to_string(pmin) + ") must be in 7-5000 range!";
            return VSK_ERR_ABORT;
        }
        if (!refs.master[r]) continue;  // psearch.c:452-471: non-masters leave the scan list
        const int t = si.trange_of[pmin];
        if (t <= 0) continue;
        for (int o = 0; o < 2; ++o) {
            const int len = (int)(refs.cons_off[2 * r + o + 1] - refs.cons_off[2 * r + o]);
            if (si.tsize[t] > len) {  // seqdupPlusTuple, hclust.h:866-869
                err = "Tuple size (" + std::to_string(si.tsize[t]) + ") is larger than sequence(" +
                      std::to_string(len) + "). Aborting!";
                return VSK_ERR_ABORT;
            }
        }
        add_seeding(2 * r, t, pmin);
        add_seeding(2 * r + 1, t, pmin);
        const int e = si.err_of[pmin];
        const int lower = pmin - e;
        if (lower >= 7 && si.trange_of[lower] <= t - 1) {  // psearch.c:552-577
            for (int o = 0; o < 2; ++o)
                if (si.tsize[t - 1] > (int)(refs.cons_off[2 * r + o + 1] - refs.cons_off[2 * r + o])) {
                    err = "Tuple size is larger than sequence. Aborting!";
                    return VSK_ERR_ABORT;
                }
            add_seeding(2 * r, t - 1, pmin);
            add_seeding(2 * r + 1, t - 1, pmin);
        }
        const int higher = pmin + e;
        if (higher <= 5000 && si.trange_of[higher] >= t + 1) {  // psearch.c:583-607
            for (int o = 0; o < 2; ++o)
                if (si.tsize[t + 1] > (int)(refs.cons_off[2 * r + o + 1] - refs.cons_off[2 * r + o])) {
                    err = "Tuple size is larger than sequence. Aborting!";
                    return VSK_ERR_ABORT;
                }
            add_seeding(2 * r, t + 1, pmin);
            add_seeding(2 * r + 1, t + 1, pmin);
        }
    }
    if (nh >= ((int64_t)1 << 32)) {
        err = "more than 2^32 seed hits in the reference set";
        return VSK_ERR_ARG;
    }
    out = vsk_refindex_host();
    out.tuples = (uint64_t)nh;
    out.total_hits = nh;
    if (nh == 0) return VSK_OK;

    // ---- hit records, in insertion order
    std::vector<HitRec> rec((size_t)nh);
    parallel_for((int64_t)seedings.size(), nt, 256, [&](int64_t s) {
        const Seeding &sd = seedings[(size_t)s];
        const char *seq = refs.cons.data() + refs.cons_off[sd.prof];
        const int len = (int)(refs.cons_off[sd.prof + 1] - refs.cons_off[sd.prof]);
        const int ts = si.tsize[sd.t];
        int64_t h = sd.base;
        for (int i = ts - 1; i < len; ++i, ++h) {
            const uint64_t code = seed_code(seq, i, si, sd.t);
            HitRec &x = rec[(size_t)h];
            x.key = ((uint64_t)sd.t << 60) | ((code >> 32) << 30) | (code & 0x3fffffffull);
            x.stamp = (uint32_t)h;
            x.prof = sd.prof;
            x.pmin = sd.pmin;
        }
    });
    std::vector<Seeding>().swap(seedings);
    phase("seed");

    // ---- key ranges from sampled splitters; stable scatter into them
    const int want_parts = (int)std::max<int64_t>(1, std::min<int64_t>(4096, nh / 8192));
    std::vector<uint64_t> split;  // range p holds keys in (split[p-1], split[p]]; the last is open
    {
        const int64_t ns = std::min<int64_t>(nh, (int64_t)want_parts * 64);
        std::vector<uint64_t> sample((size_t)ns);
        for (int64_t k = 0; k < ns; ++k) sample[(size_t)k] = rec[(size_t)(k * nh / ns)].key;
        std::sort(sample.begin(), sample.end());
        for (int p = 1; p < want_parts; ++p) {
            const uint64_t v = sample[(size_t)((int64_t)p * ns / want_parts)];
            if (split.empty() || v > split.back()) split.push_back(v);
        }
    }
    const int P = (int)split.size() + 1;
    auto part_of = [&](uint64_t key) { return (int)(std::lower_bound(split.begin(), split.end(), key) - split.begin()); };
    const int ns_slices = nt;  // one slice of the insertion order per thread
    std::vector<int64_t> hist((size_t)ns_slices * P, 0);
    std::vector<uint16_t> pid((size_t)nh);
    parallel_for(ns_slices, nt, 1, [&](int64_t sl) {
        const int64_t b = nh * sl / ns_slices, e = nh * (sl + 1) / ns_slices;
        int64_t *hrow = hist.data() + (size_t)sl * P;
        for (int64_t h = b; h < e; ++h) {
            const int p = part_of(rec[(size_t)h].key);
            pid[(size_t)h] = (uint16_t)p;
            ++hrow[p];
        }
    });
    std::vector<int64_t> part_off((size_t)P + 1, 0);
    {
        int64_t acc = 0;
        for (int p = 0; p < P; ++p) {
            part_off[(size_t)p] = acc;
            for (int sl = 0; sl < ns_slices; ++sl) {
                const int64_t c = hist[(size_t)sl * P + p];
                hist[(size_t)sl * P + p] = acc;  // where slice sl writes its first record of range p
                acc += c;
            }
        }
        part_off[(size_t)P] = acc;
    }
    std::vector<HitRec> srt((size_t)nh);
    parallel_for(ns_slices, nt, 1, [&](int64_t sl) {
        const int64_t b = nh * sl / ns_slices, e = nh * (sl + 1) / ns_slices;
        int64_t *hrow = hist.data() + (size_t)sl * P;
        for (int64_t h = b; h < e; ++h) srt[(size_t)hrow[pid[(size_t)h]]++] = rec[(size_t)h];
    });
    std::vector<HitRec>().swap(rec);
    std::vector<uint16_t>().swap(pid);
    phase("scatter");

    // ---- every range sorted stably by key; first use of the 2-key hash's racks; buckets with several hits
    const uint32_t hsize = 1162687;  // EasyLifeGetPrime(1000000), easylife.c:1620-1645
    std::vector<std::atomic<uint32_t>> rack_first((size_t)15 * hsize);
    parallel_for((int64_t)rack_first.size(), nt, 1 << 16, [&](int64_t i) { rack_first[(size_t)i].store(0xffffffffu, std::memory_order_relaxed); });
    auto rack_of = [&](uint64_t key) {
        const uint64_t code = rec_code(key);
        return (uint32_t)((uint32_t)(rjhash((uint32_t)(code >> 32)) + rjhash((uint32_t)code)) % hsize);
    };
    std::vector<std::vector<MultiBucket>> multi((size_t)P);
    std::vector<int64_t> part_entries((size_t)P, 0);     // entries the range will emit (after the replay)
    std::vector<int64_t> part_codes((size_t)P * 15, 0);  // buckets of the range per TRANGE
    parallel_for(P, nt, 1, [&](int64_t p) {
        HitRec *a = srt.data() + part_off[(size_t)p], *z = srt.data() + part_off[(size_t)p + 1];
        std::stable_sort(a, z, [](const HitRec &x, const HitRec &y) { return x.key < y.key; });
        for (HitRec *g = a; g < z;) {
            HitRec *h = g + 1;
            while (h < z && h->key == g->key) ++h;
            const int t = (int)(g->key >> 60);
            ++part_codes[(size_t)p * 15 + t];
            if (t > 4) {  // doublehash.h:334-392: racks are walked in reverse order of first use
                std::atomic<uint32_t> &rf = rack_first[(size_t)t * hsize + rack_of(g->key)];
                uint32_t cur = rf.load(std::memory_order_relaxed);
                while (g->stamp < cur && !rf.compare_exchange_weak(cur, g->stamp, std::memory_order_relaxed)) {}
            }
            if (h - g > 1)
                multi[(size_t)p].push_back(MultiBucket{(int64_t)(g - srt.data()), (int32_t)(h - g), g->stamp, g->key});
            g = h;
        }
    });
    phase("sort");

    // ---- the reference's order over the buckets it actually has to partition (psearch.c:623-648)
    std::vector<MultiBucket> order;
    {
        size_t tot = 0;
        for (auto &v : multi) tot += v.size();
        order.reserve(tot);
        for (auto &v : multi) {  // ranges are in key order: ascending (TRANGE, code)
            order.insert(order.end(), v.begin(), v.end());
            std::vector<MultiBucket>().swap(v);
        }
        // TRANGE <= 4: direct arrays, ascending code -- as they stand.  TRANGE > 4: racks in reverse order
        // of first use, inside a rack the chain in reverse creation order.
        auto first_big = std::partition_point(order.begin(), order.end(), [](const MultiBucket &b) { return (b.key >> 60) <= 4; });
        const int64_t nbig = (int64_t)(order.end() - first_big);
        MultiBucket *big = order.data() + (first_big - order.begin());
        std::vector<uint32_t> rfu((size_t)nbig);  // first use of the bucket's rack
        parallel_for(nbig, nt, 4096, [&](int64_t i) {
            const int t = (int)(big[i].key >> 60);
            rfu[(size_t)i] = rack_first[(size_t)t * hsize + rack_of(big[i].key)].load(std::memory_order_relaxed);
        });
        std::vector<int64_t> perm((size_t)nbig);
        for (int64_t i = 0; i < nbig; ++i) perm[(size_t)i] = i;
        // ascending TRANGE, then descending rack first use, then descending creation
        std::sort(perm.begin(), perm.end(), [&](int64_t x, int64_t y) {
            const int tx = (int)(big[x].key >> 60), ty = (int)(big[y].key >> 60);
            if (tx != ty) return tx < ty;
            if (rfu[(size_t)x] != rfu[(size_t)y]) return rfu[(size_t)x] > rfu[(size_t)y];
            return big[x].stamp > big[y].stamp;
        });
        std::vector<MultiBucket> sorted((size_t)nbig);
        for (int64_t i = 0; i < nbig; ++i) sorted[(size_t)i] = big[perm[(size_t)i]];
        std::copy(sorted.begin(), sorted.end(), big);
    }
    std::vector<std::atomic<uint32_t>>().swap(rack_first);
    phase("order");

    // ---- the reference's sort, bucket after bucket, with the one generator
    {
        EasySort sorter;
        std::vector<SortItem> tmp;
        for (const MultiBucket &b : order) {
            tmp.resize((size_t)b.count);
            HitRec *a = srt.data() + b.start;
            for (int32_t k = 0; k < b.count; ++k) tmp[(size_t)k] = SortItem{a[k].pmin, a[k].prof};
            sorter.sort(tmp.data(), b.count);
            for (int32_t k = 0; k < b.count; ++k) { a[k].pmin = tmp[(size_t)k].pmin; a[k].prof = tmp[(size_t)k].prof; }
        }
    }
    std::vector<MultiBucket>().swap(order);
    phase("replay");

    // ---- flatten: per TRANGE codes ascending; entries = distinct (pmin, prof) + multiplicity + group head
    // pass A: how many entries every range emits
    parallel_for(P, nt, 1, [&](int64_t p) {
        const HitRec *a = srt.data() + part_off[(size_t)p], *z = srt.data() + part_off[(size_t)p + 1];
        int64_t ne = 0;
        std::vector<int32_t> profs;
        for (const HitRec *g = a; g < z;) {
            const HitRec *h = g + 1;
            while (h < z && h->key == g->key) ++h;
            if (h - g == 1) ++ne;
            else
                for (const HitRec *u = g; u < h;) {  // groups of equal pmin; distinct profiles inside
                    const HitRec *w = u;
                    profs.clear();
                    while (w < h && w->pmin == u->pmin) profs.push_back((w++)->prof);
                    std::sort(profs.begin(), profs.end());
                    ne += (int64_t)(std::unique(profs.begin(), profs.end()) - profs.begin());
                    u = w;
                }
            g = h;
        }
        part_entries[(size_t)p] = ne;
    });
    std::vector<int64_t> ent_off((size_t)P + 1, 0), code_off((size_t)(P + 1) * 15, 0);
    for (int p = 0; p < P; ++p) {
        ent_off[(size_t)p + 1] = ent_off[(size_t)p] + part_entries[(size_t)p];
        for (int t = 0; t < 15; ++t)
            code_off[(size_t)(p + 1) * 15 + t] = code_off[(size_t)p * 15 + t] + part_codes[(size_t)p * 15 + t];
    }
    const int64_t ne_total = ent_off[(size_t)P];
    out.e_pmin.resize((size_t)ne_total);
    out.e_prof.resize((size_t)ne_total);
    out.e_mult.resize((size_t)ne_total);
    out.e_first.resize((size_t)ne_total);
    out.e_cum.resize((size_t)ne_total);
    for (int t = 1; t <= 14; ++t) {
        const int64_t nc = code_off[(size_t)P * 15 + t];
        out.codes[t].resize((size_t)nc);
        out.bucket_off[t].resize((size_t)nc + 1);
    }
    // pass B: fill.  A TRANGE's buckets are consecutive in key order, so its bucket_off ends where the next
    // TRANGE's entries begin; that last element is written after the loop.
    parallel_for(P, nt, 1, [&](int64_t p) {
        const HitRec *a = srt.data() + part_off[(size_t)p], *z = srt.data() + part_off[(size_t)p + 1];
        int64_t e = ent_off[(size_t)p];
        int64_t ci[15];
        for (int t = 0; t < 15; ++t) ci[t] = code_off[(size_t)p * 15 + t];
        std::vector<int32_t> profs;
        for (const HitRec *g = a; g < z;) {
            const HitRec *h = g + 1;
            while (h < z && h->key == g->key) ++h;
            const int t = (int)(g->key >> 60);
            out.codes[t][(size_t)ci[t]] = rec_code(g->key);
            out.bucket_off[t][(size_t)ci[t]] = e;
            ++ci[t];
            int64_t cum = 0;
            for (const HitRec *u = g; u < h;) {
                const HitRec *w = u;
                profs.clear();
                while (w < h && w->pmin == u->pmin) profs.push_back((w++)->prof);
                std::sort(profs.begin(), profs.end());
                for (size_t k = 0; k < profs.size();) {
                    size_t m = k;
                    while (m < profs.size() && profs[m] == profs[k]) ++m;
                    out.e_pmin[(size_t)e] = u->pmin;
                    out.e_prof[(size_t)e] = profs[k];
                    out.e_mult[(size_t)e] = (int32_t)(m - k);
                    out.e_first[(size_t)e] = u->prof;  // the group's first hit in the reference's bucket order
                    out.e_cum[(size_t)e] = cum;
                    cum += (int64_t)(m - k);
                    ++e;
                    k = m;
                }
                u = w;
            }
            g = h;
        }
    });
    {
        int64_t acc = 0;  // entries before TRANGE t = entries of all smaller TRANGEs
        for (int t = 1; t <= 14; ++t) {
            // number of entries of TRANGE t: from the first bucket of the next used TRANGE, or the total
            int64_t end = ne_total;
            for (int u = t + 1; u <= 14; ++u)
                if (!out.codes[u].empty()) { end = out.bucket_off[u][0]; break; }
            if (!out.codes[t].empty()) out.bucket_off[t][out.codes[t].size()] = end;
            else out.bucket_off[t][0] = acc;
            if (!out.codes[t].empty()) acc = end;
        }
    }
    phase("flatten");
    return VSK_OK;
}

// ---- index cache ------------------------------------------------------------------------
// The index is a function of the reference set alone, and building it is serial host work (the
// reference's own sort has to be replayed hit by hit, see the top of this file): ~48 s for 250 K
// references, paid by every psearch process although all of them use the same reference file.
// vsk_refs_create_cached keeps the flattened index in a file next to wherever the caller wants it:
// magic, version, a fingerprint of everything the build reads (number of records, pattern
// lengths, master flags, consensus strings), the arrays, an end mark.  A file that does not
// match in any of these is ignored and rewritten (temporary file + rename, so concurrent
// processes never see a partial one).
static const char kIdxMagic[8] = {'V', 'S', 'K', 'I', 'D', 'X', '2', 0};
static const uint64_t kIdxEnd = 0x58444e4549444e45ull;

static uint64_t fp_mix(uint64_t h, const void *data, size_t bytes) {
    const uint8_t *p = (const uint8_t *)data;
    size_t i = 0;
    for (; i + 8 <= bytes; i += 8) {
        uint64_t w;
        memcpy(&w, p + i, 8);
        h = (h ^ w) * 0x9e3779b97f4a7c15ull;
        h ^= h >> 29;
    }
    for (; i < bytes; ++i) h = (h ^ p[i]) * 0x100000001b3ull;
    return h;
}

uint64_t vsk_refs_fingerprint(const vsk_trset &refs) {
    uint64_t h = 0xcbf29ce484222325ull;
    const int64_t n = refs.n;
    h = fp_mix(h, &n, sizeof n);
    h = fp_mix(h, refs.patlen.data(), refs.patlen.size() * sizeof(int32_t));
    h = fp_mix(h, refs.master.data(), refs.master.size());
    h = fp_mix(h, refs.cons_off.data(), refs.cons_off.size() * sizeof(int64_t));
    h = fp_mix(h, refs.cons.data(), refs.cons.size());
    return h;
}

template <typename T>
static bool put_vec(FILE *f, const std::vector<T> &v) {
    const uint64_t n = v.size();
    if (fwrite(&n, sizeof n, 1, f) != 1) return false;
    return n == 0 || fwrite(v.data(), sizeof(T), n, f) == n;
}

template <typename T>
static bool get_vec(FILE *f, std::vector<T> &v, uint64_t limit) {
    uint64_t n = 0;
    if (fread(&n, sizeof n, 1, f) != 1 || n > limit) return false;
    v.resize(n);
    return n == 0 || fread(v.data(), sizeof(T), n, f) == n;
}

int vsk_refindex_save(const vsk_refindex_host &hi, uint64_t fingerprint, const char *path) {
    const std::string tmp = std::string(path) + ".tmp" + std::to_string((long long)getpid());
    FILE *f = fopen(tmp.c_str(), "wb");
    if (!f) return VSK_ERR_ARG;
    bool ok = fwrite(kIdxMagic, 8, 1, f) == 1 && fwrite(&fingerprint, 8, 1, f) == 1 &&
              fwrite(&hi.total_hits, 8, 1, f) == 1 && fwrite(&hi.tuples, 8, 1, f) == 1;
    for (int t = 1; t <= 14 && ok; ++t) ok = put_vec(f, hi.codes[t]) && put_vec(f, hi.bucket_off[t]);
    ok = ok && put_vec(f, hi.e_pmin) && put_vec(f, hi.e_prof) && put_vec(f, hi.e_mult) && put_vec(f, hi.e_first) &&
         put_vec(f, hi.e_cum) && fwrite(&kIdxEnd, 8, 1, f) == 1;
    ok = (fclose(f) == 0) && ok;
    if (!ok || rename(tmp.c_str(), path) != 0) {
        remove(tmp.c_str());
        return VSK_ERR_ARG;
    }
    return VSK_OK;
}

// VSK_OK = loaded; anything else = no usable file (absent, other reference set, truncated, older format)
int vsk_refindex_load(vsk_refindex_host &hi, uint64_t fingerprint, const char *path) {
    FILE *f = fopen(path, "rb");
    if (!f) return VSK_ERR_ARG;
    fseek(f, 0, SEEK_END);
    const uint64_t limit = (uint64_t)ftell(f);  // no array can have more elements than the file has bytes
    fseek(f, 0, SEEK_SET);
    char magic[8];
    uint64_t fp = 0, end = 0;
    hi = vsk_refindex_host();
    bool ok = fread(magic, 8, 1, f) == 1 && memcmp(magic, kIdxMagic, 8) == 0 && fread(&fp, 8, 1, f) == 1 &&
              fp == fingerprint && fread(&hi.total_hits, 8, 1, f) == 1 && fread(&hi.tuples, 8, 1, f) == 1;
    for (int t = 1; t <= 14 && ok; ++t) ok = get_vec(f, hi.codes[t], limit) && get_vec(f, hi.bucket_off[t], limit);
    ok = ok && get_vec(f, hi.e_pmin, limit) && get_vec(f, hi.e_prof, limit) && get_vec(f, hi.e_mult, limit) &&
         get_vec(f, hi.e_first, limit) && get_vec(f, hi.e_cum, limit) && fread(&end, 8, 1, f) == 1 && end == kIdxEnd;
    fclose(f);
    if (ok) {  // shape checks: the device code indexes these arrays without looking
        const size_t ne = hi.e_prof.size();
        ok = hi.e_pmin.size() == ne && hi.e_mult.size() == ne && hi.e_first.size() == ne && hi.e_cum.size() == ne;
        for (int t = 1; t <= 14 && ok; ++t)
            ok = hi.bucket_off[t].size() == hi.codes[t].size() + 1 || (hi.codes[t].empty() && hi.bucket_off[t].size() <= 1);
    }
    if (!ok) hi = vsk_refindex_host();
    return ok ? VSK_OK : VSK_ERR_ARG;
}

// the index of `refs`: from `cache_path` when it holds this reference set's, else built (and the
// file written; a failure to write is not an error).  *from_cache tells which.
int vsk_refindex_get(const vsk_trset &refs, const char *cache_path, vsk_refindex_host &hi, std::string &err,
                     int *from_cache) {
    if (from_cache) *from_cache = 0;
    uint64_t fp = 0;
    if (cache_path && *cache_path) {
        fp = vsk_refs_fingerprint(refs);
        if (vsk_refindex_load(hi, fp, cache_path) == VSK_OK) {
            if (from_cache) *from_cache = 1;
            return VSK_OK;
        }
    }
    const int rc = vsk_build_refindex_host(refs, hi, err);
    if (rc) return rc;
    if (cache_path && *cache_path) vsk_refindex_save(hi, fp, cache_path);
    return VSK_OK;
}

extern "C" int vsk_index_cache_build(const vsk_trset *refs, const char *path, int *from_cache) {
    if (!refs || !path || !*path) return VSK_ERR_ARG;
    vsk_refindex_host hi;
    std::string err;
    return vsk_refindex_get(*refs, path, hi, err, from_cache);
}

// VSK_OK when `path` holds the index of `refs` and it equals, array by array, a fresh build
extern "C" int vsk_index_cache_verify(const vsk_trset *refs, const char *path) {
    if (!refs || !path) return VSK_ERR_ARG;
    vsk_refindex_host a, b;
    std::string err;
    if (vsk_refindex_load(a, vsk_refs_fingerprint(*refs), path) != VSK_OK) return VSK_ERR_ARG;
    const int rc = build_refindex_serial(*refs, b, err);  // the independent one-thread build
    if (rc) return rc;
    bool same = a.total_hits == b.total_hits && a.tuples == b.tuples && a.e_pmin == b.e_pmin && a.e_prof == b.e_prof &&
                a.e_mult == b.e_mult && a.e_first == b.e_first && a.e_cum == b.e_cum;
    for (int t = 1; t <= 14 && same; ++t) {
        std::vector<int64_t> x = a.bucket_off[t], y = b.bucket_off[t];
        if (x.empty()) x.push_back(0);  // an unused TRANGE is stored either way
        if (y.empty()) y.push_back(0);
        same = a.codes[t] == b.codes[t] && (a.codes[t].empty() || x == y);
    }
    return same ? VSK_OK : VSK_ERR_ABORT;
}
