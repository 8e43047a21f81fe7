// Host I/O of libvsk: .leb36 / .rotindex readers producing vsk_trset
// (ReadProfileWithRC, reference src/psearch1.91/profile.h:448-732; LoadRotated,
// psearch.c:1086-1201).  Pure host code; lives in a .cu file only so that one
// compiler builds the whole library.
#include <ctype.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <thread>

#include "vsk_host.hpp"
#include "vsk_internal.cuh"

// composition tables (InitializeCompositions, profile.h:837-886)
struct CompTables {
    int8_t sym[1001];  // 0..3 = A,C,G,T majority, 4 = dash (first-max-wins over A,C,G,T,dash)
    CompTables() {
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
    }
};
static const CompTables &comp_tables() {
    static CompTables t;
    return t;
}

static inline int leb36_digit(unsigned char c) {  // LEB36ToInt, profile.h:786-807
    return (c < 65) ? (c & 0xF) : ((c & 0x1F) + 9);
}

static inline char complement_char(char c) {  // init_complement_ascii, hclust.h:935-965
    switch (c) {
        case 'A': return 'T';
        case 'C': return 'G';
        case 'G': return 'C';
        case 'T': return 'A';
        case 'a': return 't';
        case 'c': return 'g';
        case 'g': return 'c';
        case 't': return 'a';
    }
    return c;
}

static void finish_record(vsk_trset &ts, int32_t key, int32_t patlen, const std::vector<uint16_t> &fwd,
                          const std::vector<uint16_t> &rc, std::string left, std::string right,
                          int maxflank, bool option_r, int reflen) {
    const CompTables &ct = comp_tables();
    ts.key.push_back(key);
    ts.patlen.push_back(patlen);
    const std::vector<uint16_t> *pp[2] = {&fwd, &rc};
    for (int o = 0; o < 2; ++o) {
        ts.ids.insert(ts.ids.end(), pp[o]->begin(), pp[o]->end());
        ts.prof_off.push_back((int64_t)ts.ids.size());
        for (uint16_t id : *pp[o]) {
            const int s = ct.sym[id];
            if (s != 4) ts.cons.push_back("ACGT"[s]);
        }
        ts.cons_off.push_back((int64_t)ts.cons.size());
    }
    // flanks: upper-case, non-ACGT -> N (profile.h:667-685)
    for (std::string *s : {&left, &right})
        for (char &c : *s) {
            c = (char)toupper((unsigned char)c);
            if (!strchr("ACGT", c) || c == 0) c = 'N';
        }
    std::string fl(left.rbegin(), left.rend());  // forward left = GetReverse(leftflank)
    const std::string &fr = right;
    std::string rl(fr), rr(fl);                  // rc left = complement(fwd right), rc right = complement(fwd left)
    for (char &c : rl) c = complement_char(c);
    for (char &c : rr) c = complement_char(c);
    const std::string *parts[4] = {&fl, &fr, &rl, &rr};
    for (int k = 0; k < 4; ++k) {
        ts.fl.insert(ts.fl.end(), parts[k]->begin(), parts[k]->end());
        ts.fl_off.push_back((int64_t)ts.fl.size());
    }
    for (int o = 0; o < 2; ++o) {
        int ll = (int)parts[2 * o]->size(), rl2 = (int)parts[2 * o + 1]->size();
        if (option_r) {  // profile.h:711-715
            ll = std::min(ll, reflen);
            rl2 = std::min(rl2, reflen);
        } else if (ts.is_read) {  // profile.h:723-728
            ll = std::min(ll, maxflank);
            rl2 = std::min(rl2, maxflank);
        }
        ts.leftlen.push_back(ll);
        ts.rightlen.push_back(rl2);
    }
}

// raw mode: the record's offsets and lengths only (what finish_record derives from lengths alone)
static void finish_record_raw(vsk_trset &ts, int32_t key, int32_t patlen, int64_t off_fwd, int64_t len_fwd,
                              int64_t off_rc, int64_t len_rc, int64_t off_left, int32_t len_left, int64_t off_right,
                              int32_t len_right, int maxflank, bool option_r, int reflen) {
    ts.key.push_back(key);
    ts.patlen.push_back(patlen);
    ts.raw_prof.push_back(off_fwd);
    ts.raw_prof.push_back(off_rc);
    ts.prof_off.push_back(ts.prof_off.back() + len_fwd);
    ts.prof_off.push_back(ts.prof_off.back() + len_rc);
    ts.cons_off.push_back(0);  // known after the device has decoded the profiles
    ts.cons_off.push_back(0);
    ts.raw_flank.push_back(off_left);
    ts.raw_flank.push_back(off_right);
    ts.raw_flen.push_back(len_left);
    ts.raw_flen.push_back(len_right);
    // forward: left (reversed), right; rc: left = complement(right), right = complement(left)
    const int32_t parts[4] = {len_left, len_right, len_right, len_left};
    for (int k = 0; k < 4; ++k) ts.fl_off.push_back(ts.fl_off.back() + parts[k]);
    for (int o = 0; o < 2; ++o) {
        int ll = parts[2 * o], rl2 = parts[2 * o + 1];
        if (option_r) {  // profile.h:711-715
            ll = std::min(ll, reflen);
            rl2 = std::min(rl2, reflen);
        } else if (ts.is_read) {  // profile.h:723-728
            ll = std::min(ll, maxflank);
            rl2 = std::min(rl2, maxflank);
        }
        ts.leftlen.push_back(ll);
        ts.rightlen.push_back(rl2);
    }
}

static void init_rot(vsk_trset &ts) {
    ts.n = (int32_t)ts.key.size();
    ts.master.assign(ts.n, 1);
    ts.dir.assign(ts.n, 0);
    ts.rotlist.assign(ts.n, std::vector<int32_t>());
    ts.by_key.clear();
    ts.by_key.reserve((size_t)ts.n * 2);
    for (int32_t i = 0; i < ts.n; ++i) {
        ts.rotlist[i].push_back(i);
        ts.by_key.emplace(ts.key[i], i);
    }
}

static vsk_trset *new_trset(bool is_read) {
    vsk_trset *ts = new vsk_trset();
    ts->is_read = is_read;
    ts->prof_off.push_back(0);
    ts->cons_off.push_back(0);
    ts->fl_off.push_back(0);
    return ts;
}

// Parses the records in [p, end) (*end == 0) into ts.  Returns VSK_OK or VSK_ERR_ARG (bad composition
// id, message in ts.err).  `stopped` = the range did not parse to its end (the reference stops at the
// first record it cannot parse, psearch.c:332-345); `ragged` = a record was not on one line or the
// range ended inside a record, i.e. line starts are not record starts in this file.
static int parse_leb36_range(char *p, const char *path, int maxflank, bool option_r, int reflen, vsk_trset &ts,
                             bool &stopped, bool &ragged, const char *raw_base = nullptr) {
    stopped = ragged = false;
    auto next_token = [&](char *&tok, size_t &len) -> bool {
        while (*p && isspace((unsigned char)*p)) ++p;
        if (!*p) return false;
        tok = p;
        while (*p && !isspace((unsigned char)*p)) ++p;
        len = (size_t)(p - tok);
        return true;
    };
    std::vector<uint16_t> fwd, rc;
    for (;;) {
        char *tok[12];
        size_t len[12];
        int got = 0;
        while (got < 12 && next_token(tok[got], len[got])) ++got;
        if (got < 12) {
            if (got > 0) stopped = ragged = true;  // the text ends inside a record
            break;
        }
        if (memchr(tok[0], '\n', (size_t)(tok[11] + len[11] - tok[0]))) ragged = true;
        stopped = true;  // until this record is accepted
        char *e;
        const long key = strtol(tok[0], &e, 10);
        if (e == tok[0]) break;
        const long patlen = strtol(tok[1], &e, 10);
        if (e == tok[1]) break;
        const long proflen = strtol(tok[3], &e, 10);
        if (e == tok[3]) break;
        const long proflenrc = strtol(tok[4], &e, 10);
        if (e == tok[4]) break;
        if (proflen <= 0 || proflenrc <= 0) break;
        if (len[5] != (size_t)(2 * proflen) || len[6] != (size_t)(2 * proflenrc)) break;
        if (raw_base) {  // raw mode: keep where the fields are, decode on the device
            const char *bar = (const char *)memchr(tok[11], '|', len[11]);
            if (!bar) break;  // "No flanks specified!" -> NULL -> stop
            finish_record_raw(ts, (int32_t)key, (int32_t)patlen, tok[5] - raw_base, proflen, tok[6] - raw_base, proflenrc,
                              tok[11] - raw_base, (int32_t)(bar - tok[11]), bar + 1 - raw_base,
                              (int32_t)(tok[11] + len[11] - (bar + 1)), maxflank, option_r, reflen);
            stopped = false;
            continue;
        }
        fwd.resize((size_t)proflen);
        rc.resize((size_t)proflenrc);
        bool bad = false;
        for (long i = 0; i < proflen; ++i) {
            const int v = leb36_digit((unsigned char)tok[5][2 * i]) + 36 * leb36_digit((unsigned char)tok[5][2 * i + 1]);
            if (v < 0 || v > 1000) bad = true;
            fwd[(size_t)i] = (uint16_t)v;
        }
        for (long i = 0; i < proflenrc; ++i) {
            const int v = leb36_digit((unsigned char)tok[6][2 * i]) + 36 * leb36_digit((unsigned char)tok[6][2 * i + 1]);
            if (v < 0 || v > 1000) bad = true;
            rc[(size_t)i] = (uint16_t)v;
        }
        if (bad) {
            ts.err = "composition id outside 0..1000 in " + std::string(path);
            return VSK_ERR_ARG;
        }
        // flanks token LEFT|RIGHT
        std::string fl(tok[11], len[11]);
        const size_t bar = fl.find('|');
        if (bar == std::string::npos) break;  // "No flanks specified!" -> NULL -> stop
        finish_record(ts, (int32_t)key, (int32_t)patlen, fwd, rc, fl.substr(0, bar), fl.substr(bar + 1), maxflank,
                      option_r, reflen);
        stopped = false;
    }
    return VSK_OK;
}

// appends part b to a (offsets rebased)
static void append_trset(vsk_trset &a, const vsk_trset &b) {
    const int64_t id0 = (int64_t)a.ids.size(), c0 = (int64_t)a.cons.size(), f0 = (int64_t)a.fl.size();
    a.key.insert(a.key.end(), b.key.begin(), b.key.end());
    a.patlen.insert(a.patlen.end(), b.patlen.begin(), b.patlen.end());
    a.ids.insert(a.ids.end(), b.ids.begin(), b.ids.end());
    a.cons.insert(a.cons.end(), b.cons.begin(), b.cons.end());
    a.fl.insert(a.fl.end(), b.fl.begin(), b.fl.end());
    for (size_t i = 1; i < b.prof_off.size(); ++i) a.prof_off.push_back(b.prof_off[i] + id0);
    for (size_t i = 1; i < b.cons_off.size(); ++i) a.cons_off.push_back(b.cons_off[i] + c0);
    for (size_t i = 1; i < b.fl_off.size(); ++i) a.fl_off.push_back(b.fl_off[i] + f0);
    a.leftlen.insert(a.leftlen.end(), b.leftlen.begin(), b.leftlen.end());
    a.rightlen.insert(a.rightlen.end(), b.rightlen.begin(), b.rightlen.end());
    a.raw_prof.insert(a.raw_prof.end(), b.raw_prof.begin(), b.raw_prof.end());  // offsets into the one shared text
    a.raw_flank.insert(a.raw_flank.end(), b.raw_flank.begin(), b.raw_flank.end());
    a.raw_flen.insert(a.raw_flen.end(), b.raw_flen.begin(), b.raw_flen.end());
}

static int load_leb36(const char *path, int is_read, int maxflank, int option_r, int reflen, bool raw, vsk_trset **out);

extern "C" int vsk_trset_load_leb36(const char *path, int is_read, int maxflank, int option_r, int reflen,
                                    vsk_trset **out) {
    return load_leb36(path, is_read, maxflank, option_r, reflen, false, out);
}

// The same file, tokenised only: the profiles and flanks stay text and are decoded by the GPU when the set is
// uploaded (vsk_reads_upload / vsk_proclu_run).  For read sets; a reference set needs its consensus strings
// on the host (seed index) and is refused by vsk_refs_create in this form.
extern "C" int vsk_trset_load_leb36_raw(const char *path, int is_read, int maxflank, int option_r, int reflen,
                                        vsk_trset **out) {
    return load_leb36(path, is_read, maxflank, option_r, reflen, true, out);
}

static int load_leb36(const char *path, int is_read, int maxflank, int option_r, int reflen, bool raw, vsk_trset **out) {
    if (!path || !out) return VSK_ERR_ARG;
    *out = nullptr;
    FILE *fp = fopen(path, "rb");
    if (!fp) return VSK_ERR_ARG;
    fseek(fp, 0, SEEK_END);
    const long sz = ftell(fp);
    fseek(fp, 0, SEEK_SET);
    std::vector<char> buf((size_t)sz + 1);
    if (sz > 0 && fread(buf.data(), 1, (size_t)sz, fp) != (size_t)sz) { fclose(fp); return VSK_ERR_ARG; }
    fclose(fp);
    buf[(size_t)sz] = 0;
    vsk_trset *ts = new_trset(is_read != 0);
    // Text parsing is the slowest host step of a run once the scan is on the GPU (6 us per record on
    // one core), so large files are cut at line starts and parsed by several threads, the parts
    // joined in order.  That is the serial result when every record sits on its own line (what
    // trf2proclu-ngs writes); a file where that does not hold is detected and parsed serially.
    int nthreads = 1;
    if (const char *e = getenv("VSK_LOAD_THREADS")) nthreads = atoi(e);
    else if (sz > (4 << 20)) nthreads = (int)std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 16u);
    nthreads = (int)std::max<long>(1, std::min<long>(nthreads, sz / (1 << 20)));
    bool done = false;
    if (nthreads > 1) {
        std::vector<char *> cut((size_t)nthreads + 1);
        cut[0] = buf.data();
        cut[(size_t)nthreads] = buf.data() + sz;
        for (int k = 1; k < nthreads; ++k) {
            char *q = buf.data() + (size_t)((double)sz * k / nthreads);
            char *nl = (char *)memchr(q, '\n', (size_t)(buf.data() + sz - q));
            cut[(size_t)k] = nl ? nl : buf.data() + sz;
        }
        std::vector<vsk_trset *> part((size_t)nthreads);
        std::vector<int> rcs((size_t)nthreads, VSK_OK);
        std::vector<char> stopped((size_t)nthreads, 0), ragged((size_t)nthreads, 0), saved((size_t)nthreads + 1, 0);
        for (int k = 1; k <= nthreads; ++k) {  // a range ends in a NUL where its last newline was
            saved[(size_t)k] = *cut[(size_t)k];
            *cut[(size_t)k] = 0;
        }
        std::vector<std::thread> th;
        for (int k = 0; k < nthreads; ++k)
            th.emplace_back([&, k] {
                part[(size_t)k] = new_trset(is_read != 0);
                bool st = false, rg = false;
                char *b = cut[(size_t)k] + (k > 0 && cut[(size_t)k] < buf.data() + sz ? 1 : 0);
                if (b > cut[(size_t)k + 1]) b = cut[(size_t)k + 1];
                rcs[(size_t)k] = parse_leb36_range(b, path, maxflank, option_r != 0, reflen, *part[(size_t)k], st, rg,
                                                   raw ? buf.data() : nullptr);
                stopped[(size_t)k] = st;
                ragged[(size_t)k] = rg;
            });
        for (auto &t : th) t.join();
        bool any_ragged = false;
        for (int k = 0; k < nthreads; ++k) any_ragged = any_ragged || ragged[(size_t)k];
        if (!any_ragged) {
            done = true;
            for (int k = 0; k < nthreads; ++k) {
                append_trset(*ts, *part[(size_t)k]);
                if (rcs[(size_t)k] != VSK_OK) {
                    ts->err = part[(size_t)k]->err;
                    for (auto *q : part) delete q;
                    *out = ts;
                    return rcs[(size_t)k];
                }
                if (stopped[(size_t)k]) break;  // nothing after the first record that does not parse
            }
        }
        for (auto *q : part) delete q;
        if (!done)
            // back to the text as read; descending, so that two cuts on the same newline (a line longer
            // than a range) put back the character the first of them saved, not the NUL the second saw
            for (int k = nthreads - 1; k >= 1; --k) *cut[(size_t)k] = saved[(size_t)k];
    }
    if (!done) {
        bool st, rg;
        const int rc = parse_leb36_range(buf.data(), path, maxflank, option_r != 0, reflen, *ts, st, rg,
                                         raw ? buf.data() : nullptr);
        if (rc != VSK_OK) {
            *out = ts;
            return rc;
        }
    }
    if (raw) {
        ts->raw = true;
        ts->raw_text.swap(buf);
    }
    init_rot(*ts);
    *out = ts;
    return VSK_OK;
}

extern "C" int vsk_trset_from_arrays(int32_t n, const int32_t *key, const int32_t *patlen,
                                     const int64_t *prof_off, const uint16_t *ids,
                                     const int64_t *flank_off, const char *flanks, int is_read,
                                     int maxflank, int option_r, int reflen, vsk_trset **out) {
    if (!out || n < 0) return VSK_ERR_ARG;
    vsk_trset *ts = new_trset(is_read != 0);
    std::vector<uint16_t> fwd, rc;
    for (int32_t i = 0; i < n; ++i) {
        fwd.assign(ids + prof_off[2 * i], ids + prof_off[2 * i + 1]);
        rc.assign(ids + prof_off[2 * i + 1], ids + prof_off[2 * i + 2]);
        for (uint16_t v : fwd) if (v > 1000) { delete ts; return VSK_ERR_ARG; }
        for (uint16_t v : rc) if (v > 1000) { delete ts; return VSK_ERR_ARG; }
        std::string l(flanks + flank_off[2 * i], flanks + flank_off[2 * i + 1]);
        std::string r(flanks + flank_off[2 * i + 1], flanks + flank_off[2 * i + 2]);
        finish_record(*ts, key[i], patlen[i], fwd, rc, l, r, maxflank, option_r != 0, reflen);
    }
    init_rot(*ts);
    *out = ts;
    return VSK_OK;
}

// LoadRotated (psearch.c:1086-1201): each line "key{'|"} key{'|"} ...", the first is
// the master; the others leave the scan list, join the master's rotlist and get
// dir = firstdir ? !d : d.
extern "C" int vsk_trset_apply_rotindex(vsk_trset *ts, const char *path) {
    if (!ts || !path) return VSK_ERR_ARG;
    FILE *fp = fopen(path, "r");
    if (!fp) return VSK_OK;  // "no rotated file is fine"
    // the lists are about to change: page-locked ranges (vsk_trset_pin) are released, pin again afterwards
    for (void *p : ts->pinned) cudaHostUnregister(p);
    ts->pinned.clear();
    std::vector<char> line(1 << 20);
    std::string all;
    size_t got;
    while ((got = fread(line.data(), 1, line.size(), fp)) > 0) all.append(line.data(), got);
    fclose(fp);
    size_t pos = 0;
    const size_t N = all.size();
    while (pos < N) {
        size_t eol = all.find_first_of("\r\n", pos);
        if (eol == std::string::npos) eol = N;
        const char *s = all.c_str() + pos, *end = all.c_str() + eol;
        int count = 0, firstdir = 0;
        int32_t master = -1;
        while (s < end) {
            while (s < end && (*s == ' ' || *s == '\t')) ++s;
            if (s >= end) break;
            char *e;
            const long k = strtol(s, &e, 10);
            if (e == s) { ts->err = "Invalid format detected in the redundancy file!"; return VSK_ERR_ABORT; }
            s = e;
            while (s < end && (*s == ' ' || *s == '\t')) ++s;
            if (s >= end || (*s != '\'' && *s != '"')) {
                ts->err = "Unexpected value qualifier detected in the redundancy file!";
                return VSK_ERR_ABORT;
            }
            const int d = (*s == '\'') ? 0 : 1;
            ++s;
            auto it = ts->by_key.find((int32_t)k);
            if (it == ts->by_key.end()) {
                ts->err = "Error looking up key " + std::to_string(k) + " in profileHash!";
                return VSK_ERR_ABORT;
            }
            ++count;
            if (count == 1) {
                master = it->second;
                firstdir = d;
                ts->dir[master] = 0;
            } else {
                const int32_t m = it->second;
                ts->rotlist[master].push_back(m);
                ts->rot_flat = false;
                ts->master[m] = 0;
                ts->dir[m] = (uint8_t)(firstdir ? !d : d);
            }
        }
        pos = eol + 1;
    }
    return VSK_OK;
}

void vsk_trset_prepare(vsk_trset *ts) {
    if (!ts->rot_flat) {  // rotlists flattened once (cache inside the TR set)
        ts->rot_off.assign((size_t)ts->n + 1, 0);
        ts->rot_mem.clear();
        for (int32_t i = 0; i < ts->n; ++i) {
            ts->rot_off[(size_t)i] = (int32_t)ts->rot_mem.size();
            ts->rot_mem.insert(ts->rot_mem.end(), ts->rotlist[(size_t)i].begin(), ts->rotlist[(size_t)i].end());
        }
        ts->rot_off[(size_t)ts->n] = (int32_t)ts->rot_mem.size();
        ts->rot_flat = true;
    }
    if (!ts->blk_ready && (!ts->raw || ts->raw_decoded)) {
        const size_t np = (size_t)2 * ts->n;
        ts->cons_blk.assign(np + 1, 0);
        for (size_t p = 0; p < np; ++p)
            ts->cons_blk[p + 1] = ts->cons_blk[p] + (ts->cons_off[p + 1] - ts->cons_off[p] + 63) / 64;
        ts->blk_ready = true;
    }
}

// Page-lock the large arrays of a TR set so that vsk_proclu_run's host->device copies run at
// full PCIe rate (one-time cost; undone by vsk_trset_free).
extern "C" int vsk_trset_pin(vsk_trset *ts) {
    if (!ts) return VSK_ERR_ARG;
    if (!ts->pinned.empty()) return VSK_OK;
    vsk_trset_prepare(ts);
    auto pin = [&](void *p, size_t bytes) {
        if (bytes < (1u << 16)) return;
        if (cudaHostRegister(p, bytes, cudaHostRegisterPortable) == cudaSuccess) ts->pinned.push_back(p);
        else cudaGetLastError();
    };
    pin(ts->ids.data(), ts->ids.size() * sizeof(uint16_t));
    pin(ts->cons.data(), ts->cons.size());
    pin(ts->fl.data(), ts->fl.size());
    pin(ts->prof_off.data(), ts->prof_off.size() * sizeof(int64_t));
    pin(ts->cons_off.data(), ts->cons_off.size() * sizeof(int64_t));
    pin(ts->fl_off.data(), ts->fl_off.size() * sizeof(int64_t));
    pin(ts->leftlen.data(), ts->leftlen.size() * sizeof(int32_t));
    pin(ts->rightlen.data(), ts->rightlen.size() * sizeof(int32_t));
    pin(ts->patlen.data(), ts->patlen.size() * sizeof(int32_t));
    pin(ts->key.data(), ts->key.size() * sizeof(int32_t));
    if (ts->rot_flat) {
        pin(ts->rot_off.data(), ts->rot_off.size() * sizeof(int32_t));
        pin(ts->rot_mem.data(), ts->rot_mem.size() * sizeof(int32_t));
    }
    if (ts->blk_ready) pin(ts->cons_blk.data(), ts->cons_blk.size() * sizeof(int64_t));
    pin(ts->master.data(), ts->master.size());
    pin(ts->dir.data(), ts->dir.size());
    return VSK_OK;
}

extern "C" void vsk_trset_free(vsk_trset *ts) {
    if (!ts) return;
    for (void *p : ts->pinned) cudaHostUnregister(p);
    delete ts;
}
extern "C" int32_t vsk_trset_size(const vsk_trset *ts) { return ts ? ts->n : 0; }
extern "C" const char *vsk_trset_error(const vsk_trset *ts) { return ts ? ts->err.c_str() : ""; }

// psearch.c:388-440: one key space for references and reads
extern "C" int vsk_trsets_check_unique(const vsk_trset *a, const vsk_trset *b, int32_t *dup) {
    std::unordered_map<int32_t, int> seen;
    seen.reserve((size_t)(a ? a->n : 0) + (size_t)(b ? b->n : 0));
    for (const vsk_trset *ts : {a, b}) {
        if (!ts) continue;
        for (int32_t k : ts->key)
            if (!seen.emplace(k, 1).second) {
                if (dup) *dup = k;
                return VSK_ERR_ABORT;
            }
    }
    return VSK_OK;
}

// order-sensitive hash of everything a TR set holds (records, profiles, consensus, flanks, flank
// lengths, redundancy state): two sets with the same value went through the loader the same way
extern "C" uint64_t vsk_trset_checksum(const vsk_trset *ts) {
    if (!ts) return 0;
    uint64_t h = 0xcbf29ce484222325ull;
    auto mix = [&](const void *d, size_t bytes) {
        const uint8_t *p = (const uint8_t *)d;
        for (size_t i = 0; i < bytes; ++i) h = (h ^ p[i]) * 0x100000001b3ull;
        h ^= bytes;
        h *= 0x9e3779b97f4a7c15ull;
    };
    const int64_t n = ts->n;
    mix(&n, sizeof n);
    mix(ts->key.data(), ts->key.size() * 4);
    mix(ts->patlen.data(), ts->patlen.size() * 4);
    mix(ts->prof_off.data(), ts->prof_off.size() * 8);
    mix(ts->ids.data(), ts->ids.size() * 2);
    mix(ts->cons_off.data(), ts->cons_off.size() * 8);
    mix(ts->cons.data(), ts->cons.size());
    mix(ts->fl_off.data(), ts->fl_off.size() * 8);
    mix(ts->fl.data(), ts->fl.size());
    mix(ts->leftlen.data(), ts->leftlen.size() * 4);
    mix(ts->rightlen.data(), ts->rightlen.size() * 4);
    mix(ts->master.data(), ts->master.size());
    mix(ts->dir.data(), ts->dir.size());
    for (const auto &r : ts->rotlist) mix(r.data(), r.size() * 4);
    return h;
}

// ---- packed TR sets ----------------------------------------------------------------------
// With the scan on the GPU and the seed index cached, reading the .leb36 TEXT is the slowest step
// of a psearch process (ReadProfileWithRC, profile.h:448-732: about 2 us per record per core --
// LEB36 digits, consensus, flank clean-up -- against 0.35 us per read for the scan).  The arrays
// that parse produces are kept in a binary file next to the text: magic, the parse parameters, a
// fingerprint of the text file (size, modification time, hash of its first and last 64 KB), the
// arrays, an end mark.  A file that does not match in any of these is ignored and rewritten.  The
// redundancy state (.rotindex) is not part of it: vsk_trset_apply_rotindex runs after either load.
static const char kTrsMagic[8] = {'V', 'S', 'K', 'T', 'R', 'S', '1', 0};
static const uint64_t kTrsEnd = 0x444e455352544b53ull;

struct TrsHeader {
    char magic[8];
    int32_t is_read, maxflank, option_r, reflen;
    uint64_t src_size, src_mtime_ns, src_hash;
    int64_t n;
};

static bool source_fingerprint(const char *path, TrsHeader &h) {
    struct stat st;
    if (stat(path, &st) != 0) return false;
    h.src_size = (uint64_t)st.st_size;
    h.src_mtime_ns = (uint64_t)st.st_mtim.tv_sec * 1000000000ull + (uint64_t)st.st_mtim.tv_nsec;
    FILE *f = fopen(path, "rb");
    if (!f) return false;
    std::vector<uint8_t> buf(1 << 16);
    uint64_t hash = 0xcbf29ce484222325ull;
    auto mix = [&](size_t got) {
        for (size_t i = 0; i < got; ++i) hash = (hash ^ buf[i]) * 0x100000001b3ull;
    };
    mix(fread(buf.data(), 1, buf.size(), f));
    if (st.st_size > (off_t)buf.size()) {
        fseek(f, -(long)buf.size(), SEEK_END);
        mix(fread(buf.data(), 1, buf.size(), f));
    }
    fclose(f);
    h.src_hash = hash;
    return true;
}

template <typename T>
static bool trs_put(FILE *f, const std::vector<T> &v) {
    const uint64_t n = v.size();
    return fwrite(&n, sizeof n, 1, f) == 1 && (n == 0 || fwrite(v.data(), sizeof(T), n, f) == n);
}
template <typename T>
static bool trs_get(FILE *f, std::vector<T> &v, uint64_t limit) {
    uint64_t n = 0;
    if (fread(&n, sizeof n, 1, f) != 1 || n > limit) return false;
    v.resize(n);
    return n == 0 || fread(v.data(), sizeof(T), n, f) == n;
}

static int trset_save_packed(const vsk_trset &ts, const TrsHeader &h, const char *path) {
    const std::string tmp = std::string(path) + ".tmp" + std::to_string((long long)getpid());
    FILE *f = fopen(tmp.c_str(), "wb");
    if (!f) return VSK_ERR_ARG;
    bool ok = fwrite(&h, sizeof h, 1, f) == 1 && trs_put(f, ts.key) && trs_put(f, ts.patlen) && trs_put(f, ts.prof_off) &&
              trs_put(f, ts.ids) && trs_put(f, ts.cons_off) && trs_put(f, ts.cons) && trs_put(f, ts.fl_off) &&
              trs_put(f, ts.fl) && trs_put(f, ts.leftlen) && trs_put(f, ts.rightlen) && fwrite(&kTrsEnd, 8, 1, f) == 1;
    ok = (fclose(f) == 0) && ok;
    if (!ok || rename(tmp.c_str(), path) != 0) {
        remove(tmp.c_str());
        return VSK_ERR_ARG;
    }
    return VSK_OK;
}

static vsk_trset *trset_load_packed(const TrsHeader &want, const char *path) {
    FILE *f = fopen(path, "rb");
    if (!f) return nullptr;
    fseek(f, 0, SEEK_END);
    const uint64_t limit = (uint64_t)ftell(f);
    fseek(f, 0, SEEK_SET);
    TrsHeader h;
    vsk_trset *ts = new vsk_trset();
    uint64_t end = 0;
    bool ok = fread(&h, sizeof h, 1, f) == 1 && memcmp(&h, &want, sizeof h) == 0 && trs_get(f, ts->key, limit) &&
              trs_get(f, ts->patlen, limit) && trs_get(f, ts->prof_off, limit) && trs_get(f, ts->ids, limit) &&
              trs_get(f, ts->cons_off, limit) && trs_get(f, ts->cons, limit) && trs_get(f, ts->fl_off, limit) &&
              trs_get(f, ts->fl, limit) && trs_get(f, ts->leftlen, limit) && trs_get(f, ts->rightlen, limit) &&
              fread(&end, 8, 1, f) == 1 && end == kTrsEnd;
    fclose(f);
    if (ok) {  // shapes the rest of the library indexes without looking
        const size_t n = ts->key.size();
        ok = (int64_t)n == h.n && ts->patlen.size() == n && ts->prof_off.size() == 2 * n + 1 &&
             ts->cons_off.size() == 2 * n + 1 && ts->fl_off.size() == 4 * n + 1 && ts->leftlen.size() == 2 * n &&
             ts->rightlen.size() == 2 * n && ts->prof_off.front() == 0 && ts->prof_off.back() == (int64_t)ts->ids.size() &&
             ts->cons_off.back() == (int64_t)ts->cons.size() && ts->fl_off.back() == (int64_t)ts->fl.size();
        for (size_t i = 0; ok && i + 1 < ts->prof_off.size(); ++i) ok = ts->prof_off[i] <= ts->prof_off[i + 1];
        for (size_t i = 0; ok && i + 1 < ts->cons_off.size(); ++i) ok = ts->cons_off[i] <= ts->cons_off[i + 1];
        for (size_t i = 0; ok && i + 1 < ts->fl_off.size(); ++i) ok = ts->fl_off[i] <= ts->fl_off[i + 1];
        for (size_t i = 0; ok && i < ts->ids.size(); ++i) ok = ts->ids[i] <= 1000;
    }
    if (!ok) {
        delete ts;
        return nullptr;
    }
    ts->is_read = want.is_read != 0;
    init_rot(*ts);
    return ts;
}

extern "C" int vsk_trset_load_leb36_cached(const char *path, int is_read, int maxflank, int option_r, int reflen,
                                           const char *cache_path, vsk_trset **out, int *from_cache) {
    if (from_cache) *from_cache = 0;
    if (!path || !out) return VSK_ERR_ARG;
    if (!cache_path || !*cache_path) return vsk_trset_load_leb36(path, is_read, maxflank, option_r, reflen, out);
    TrsHeader h;
    memset(&h, 0, sizeof h);
    memcpy(h.magic, kTrsMagic, 8);
    h.is_read = is_read != 0;
    h.maxflank = maxflank;
    h.option_r = option_r != 0;
    h.reflen = reflen;
    if (!source_fingerprint(path, h)) return VSK_ERR_ARG;
    // n is part of the header that must match: read it from the candidate file itself
    {
        TrsHeader got;
        FILE *f = fopen(cache_path, "rb");
        if (f) {
            if (fread(&got, sizeof got, 1, f) == 1) h.n = got.n;
            fclose(f);
        }
    }
    if (vsk_trset *ts = trset_load_packed(h, cache_path)) {
        *out = ts;
        if (from_cache) *from_cache = 1;
        return VSK_OK;
    }
    const int rc = vsk_trset_load_leb36(path, is_read, maxflank, option_r, reflen, out);
    if (rc != VSK_OK) return rc;
    h.n = (*out)->n;
    trset_save_packed(**out, h, cache_path);  // a failure to write is not an error
    return VSK_OK;
}
