// Host I/O of libvsk: .leb36 / .rotindex readers producing vsk_trset
// (ReadProfileWithRC, reference src/psearch1.91/profile.h:448-732; LoadRotated,
// psearch.c:1086-1201).  Pure host code; lives in a .cu file only so that one
// compiler builds the whole library.
#include <ctype.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

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

extern "C" int vsk_trset_load_leb36(const char *path, int is_read, int maxflank, int option_r, int reflen,
                                    vsk_trset **out) {
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
    char *p = buf.data();
    auto next_token = [&](char *&tok, size_t &len) -> bool {
        while (*p && isspace((unsigned char)*p)) ++p;
        if (!*p) return false;
        tok = p;
        while (*p && !isspace((unsigned char)*p)) ++p;
        len = (size_t)(p - tok);
        return true;
    };
    std::vector<uint16_t> fwd, rc;
    for (;;) {  // the reference stops at the first record it cannot parse (psearch.c:332-345)
        char *tok[12];
        size_t len[12];
        bool ok = true;
        for (int k = 0; k < 12 && ok; ++k) ok = next_token(tok[k], len[k]);
        if (!ok) break;
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
            ts->err = "composition id outside 0..1000 in " + std::string(path);
            *out = ts;
            return VSK_ERR_ARG;
        }
        // flanks token LEFT|RIGHT
        std::string fl(tok[11], len[11]);
        const size_t bar = fl.find('|');
        if (bar == std::string::npos) break;  // "No flanks specified!" -> NULL -> stop
        finish_record(*ts, (int32_t)key, (int32_t)patlen, fwd, rc, fl.substr(0, bar), fl.substr(bar + 1),
                      maxflank, option_r != 0, reflen);
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

// Page-lock the large arrays of a TR set so that vsk_proclu_run's host->device copies run at
// full PCIe rate (one-time cost; undone by vsk_trset_free).
extern "C" int vsk_trset_pin(vsk_trset *ts) {
    if (!ts) return VSK_ERR_ARG;
    if (!ts->pinned.empty()) return VSK_OK;
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
