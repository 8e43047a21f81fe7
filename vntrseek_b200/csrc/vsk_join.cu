// join_clusters: the merge of per-file PROCLU cluster files into one file (SURVEY 8(f) row 1;
// reference src/join_clusters/joinc.c:128-391 + the loader :393-571).  Host code only: the work
// is parsing, one hash over the reference ids and list splicing -- there is nothing to put on the
// GPU -- but it is the last serial step behind the multi-GPU run, so it is written to run at memory
// speed (files read whole, flat arrays, an open-addressing table) and to give the reference's
// output byte for byte, including its order rules:
//   * files in strcmp order of their paths (joinc.c:228-238);
//   * a cluster's members in REVERSE file order (the loader pushes on a list head, :466, :520-535),
//     split into references (negative ids) and reads;
//   * cluster i absorbs, for each of its references in turn, the not-yet-absorbed cluster that
//     owns that reference, whole chain at a time, appended at the end of i's chain (:246-311);
//     the absorbed chain is flipped when the two clusters disagree on that reference's direction;
//   * output per surviving cluster: the references of the chain, duplicates inside the cluster
//     dropped, then the reads, each preceded by a comma (:320-383).
#include <dirent.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <strings.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/vsk.h"

namespace {

struct JoinCluster {
    int64_t ref_begin, ref_end;    // into ref_id / ref_dir
    int64_t read_begin, read_end;  // into read_id / read_dir
    int32_t next = -1;             // chain of absorbed clusters
    bool alive = true;             // cid != 0 in the reference
};

struct JoinState {
    std::vector<JoinCluster> cl;
    std::vector<int32_t> ref_id, read_id;
    std::vector<char> ref_dir, read_dir;
    std::string err;
};

// reference id -> cluster, open addressing (ids are arbitrary ints, 0x80000000 = empty)
struct IdTable {
    std::vector<int32_t> key, val;
    uint32_t mask = 0;
    void init(size_t n) {
        size_t cap = 16;
        while (cap < 2 * n + 16) cap <<= 1;
        key.assign(cap, INT32_MIN);
        val.assign(cap, -1);
        mask = (uint32_t)(cap - 1);
    }
    static uint32_t mix(uint32_t a) {
        a ^= a >> 16;
        a *= 0x7feb352du;
        a ^= a >> 15;
        a *= 0x846ca68bu;
        a ^= a >> 16;
        return a;
    }
    int32_t get(int32_t k) const {
        for (uint32_t h = mix((uint32_t)k) & mask;; h = (h + 1) & mask) {
            if (key[h] == k) return val[h];
            if (key[h] == INT32_MIN) return -1;
        }
    }
    void set(int32_t k, int32_t v) {
        for (uint32_t h = mix((uint32_t)k) & mask;; h = (h + 1) & mask) {
            if (key[h] == k || key[h] == INT32_MIN) {
                key[h] = k;
                val[h] = v;
                return;
            }
        }
    }
};

int fail(JoinState &S, int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    S.err = buf;
    return code;
}

inline bool is_ws(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\v' || c == '\f'; }

// LoadClustersFromFile (joinc.c:393-571) on the file's bytes: `@` then members `<int><'|">`, an
// optional `#comment` token ends the cluster; anything that is not a digit or '-' after a member
// ends it too (and must be the next `@`).
int parse_file(JoinState &S, const std::string &path) {
    FILE *fp = fopen(path.c_str(), "rb");
    if (!fp) return fail(S, VSK_ERR_ABORT, "Unable to open intermediate cluster file!");
    std::vector<char> buf;
    fseek(fp, 0, SEEK_END);
    const long sz = ftell(fp);
    fseek(fp, 0, SEEK_SET);
    buf.resize(sz > 0 ? (size_t)sz : 0);
    if (sz > 0 && fread(buf.data(), 1, (size_t)sz, fp) != (size_t)sz) {
        fclose(fp);
        return fail(S, VSK_ERR_ABORT, "Unable to open intermediate cluster file!");
    }
    fclose(fp);
    const char *p = buf.data(), *end = p + buf.size();
    std::vector<std::pair<int32_t, char>> mem;  // members in file order
    for (;;) {
        while (p < end && is_ws(*p)) ++p;
        if (p >= end) break;
        if (*p != '@') return fail(S, VSK_ERR_ABORT, "Invalid format detected in cluster file 1!");
        ++p;
        mem.clear();
        for (;;) {
            while (p < end && is_ws(*p)) ++p;  // "%d"
            const char *q = p;
            bool neg = false;
            if (q < end && (*q == '-' || *q == '+')) { neg = (*q == '-'); ++q; }
            if (q >= end || *q < '0' || *q > '9')
                return fail(S, VSK_ERR_ABORT, "Invalid format detected in cluster file 2!");
            long long v = 0;
            while (q < end && *q >= '0' && *q <= '9') { v = v * 10 + (*q - '0'); ++q; }
            p = q;
            while (p < end && is_ws(*p)) ++p;  // " %c"
            if (p >= end || (*p != '\'' && *p != '"'))
                return fail(S, VSK_ERR_ABORT, "Unexpected value qualifier detected in cluster file!");
            mem.emplace_back((int32_t)(neg ? -v : v), *p);
            ++p;
            while (p < end && is_ws(*p)) ++p;  // " %c"
            if (p >= end) break;
            if (*p == '#') {  // comment: one whitespace-delimited token ("%s")
                ++p;
                while (p < end && is_ws(*p)) ++p;
                while (p < end && !is_ws(*p)) ++p;
                break;
            }
            if ((*p >= '0' && *p <= '9') || *p == '-') continue;
            break;
        }
        JoinCluster c;
        c.ref_begin = (int64_t)S.ref_id.size();
        c.read_begin = (int64_t)S.read_id.size();
        for (size_t i = mem.size(); i-- > 0;) {  // the reference's list is built head first
            if (mem[i].first >= 0) {
                S.read_id.push_back(mem[i].first);
                S.read_dir.push_back(mem[i].second);
            } else {
                S.ref_id.push_back(mem[i].first);
                S.ref_dir.push_back(mem[i].second);
            }
        }
        c.ref_end = (int64_t)S.ref_id.size();
        c.read_end = (int64_t)S.read_id.size();
        S.cl.push_back(c);
    }
    return VSK_OK;
}

}  // namespace

static int join_files(const std::vector<std::string> &paths, const char *out_path, vsk_join_stats *stats,
                      FILE *log, JoinState &S) {
    // joinc.c:149-156: the output file is opened (and so truncated) before anything is read
    FILE *fpo = fopen(out_path, "w");
    if (!fpo) return fail(S, VSK_ERR_ARG, "Error, cannot open file '%s' for writing!", out_path);
    for (const std::string &p : paths) {
        const size_t before = S.cl.size();
        if (log) fprintf(log, "\nprocessing %s...", p.c_str());
        const int rc = parse_file(S, p);
        if (rc) {
            fclose(fpo);
            return rc;
        }
        if (log) fprintf(log, "%d clusters\n", (int)(S.cl.size() - before));
    }
    const int32_t n = (int32_t)S.cl.size();
    vsk_join_stats st;
    memset(&st, 0, sizeof st);
    st.clusters_read = n;
    st.clusters_after_join = n;
    st.refs_read = (int64_t)S.ref_id.size();
    st.reads_read = (int64_t)S.read_id.size();
    if (S.ref_id.size() >= 0x7fffffffu || S.cl.size() >= 0x7fffffffu) {
        fclose(fpo);
        return fail(S, VSK_ERR_ARG, "more than 2^31 clusters or reference occurrences");
    }
    if (n == 0) {  // joinc.c:184-185: nothing to do, the (empty) output file stays
        fclose(fpo);
        if (stats) *stats = st;
        return VSK_OK;
    }
    // ---- join (joinc.c:246-311) ----
    // The reference re-points the hash entry of every reference of an absorbed chain to the new head
    // and flips the chain's direction marks in place, i.e. it walks the absorbed chain at every merge
    // (quadratic when clusters keep absorbing one growing chain).  The same result in near-linear
    // time: a union-find over the clusters with the flip kept as a parity on the parent link
    // (current mark = original mark XOR parity to the root), a tail pointer per head for the splice,
    // and, per reference id, its FIRST occurrence in its component's chain -- the one the
    // reference's flip test finds (:271-285); chains only grow at the end, so it stays the first.
    std::vector<int32_t> parent((size_t)n), tail((size_t)n);
    std::vector<uint8_t> par((size_t)n, 0);
    for (int32_t i = 0; i < n; ++i) parent[(size_t)i] = tail[(size_t)i] = i;
    std::vector<int32_t> stack;
    auto find = [&](int32_t t, uint8_t &parity) -> int32_t {  // root of t, parity of t relative to it
        stack.clear();
        int32_t r = t;
        while (parent[(size_t)r] != r) {
            stack.push_back(r);
            r = parent[(size_t)r];
        }
        uint8_t acc = 0;  // from the node nearest the root outwards
        for (size_t k = stack.size(); k-- > 0;) {
            const int32_t x = stack[k];
            acc ^= par[(size_t)x];
            par[(size_t)x] = acc;
            parent[(size_t)x] = r;
        }
        parity = acc;
        return r;
    };
    IdTable occ;  // reference id -> index into ref_id of its first occurrence in its component's chain
    occ.init(S.ref_id.size());
    std::vector<int32_t> owner(S.ref_id.size());  // cluster of every reference occurrence
    for (int32_t i = 0; i < n; ++i)
        for (int64_t k = S.cl[(size_t)i].ref_begin; k < S.cl[(size_t)i].ref_end; ++k) owner[(size_t)k] = i;
    for (int32_t i = 0; i < n; ++i) {
        const JoinCluster &ci = S.cl[(size_t)i];
        for (int64_t j = ci.ref_begin; j < ci.ref_end; ++j) {
            const int32_t id = S.ref_id[(size_t)j];
            const int32_t at = occ.get(id);
            bool mine = false;
            if (at >= 0) {
                uint8_t p = 0;
                const int32_t o = find(owner[(size_t)at], p);
                if (o != i) {  // another, not yet absorbed cluster owns this reference: absorb its chain
                    const char cur = ((S.ref_dir[(size_t)at] == '\'') != (p != 0)) ? '\'' : '"';
                    const bool flip = cur != S.ref_dir[(size_t)j];
                    if (flip && log) fprintf(log, "\nflipped c:%d (%d)!", i, id);
                    S.cl[(size_t)tail[(size_t)i]].next = o;
                    tail[(size_t)i] = tail[(size_t)o];
                    parent[(size_t)o] = i;
                    par[(size_t)o] = flip ? 1 : 0;
                    S.cl[(size_t)o].alive = false;
                    st.clusters_after_join--;
                } else {
                    mine = owner[(size_t)at] == i;  // an earlier occurrence in this very cluster stays the first
                }
            }
            if (!mine) occ.set(id, (int32_t)j);
        }
    }
    // the marks as the reference leaves them
    for (int32_t t = 0; t < n; ++t) {
        uint8_t p = 0;
        find(t, p);
        if (!p) continue;
        const JoinCluster &ct = S.cl[(size_t)t];
        for (int64_t k = ct.ref_begin; k < ct.ref_end; ++k) S.ref_dir[(size_t)k] = (S.ref_dir[(size_t)k] == '\'') ? '"' : '\'';
        for (int64_t k = ct.read_begin; k < ct.read_end; ++k) S.read_dir[(size_t)k] = (S.read_dir[(size_t)k] == '\'') ? '"' : '\'';
    }
    IdTable seen;
    // ---- write (joinc.c:320-383) ----
    if (log) fprintf(log, "\n\nWriting final cluster file...");
    seen.init(S.ref_id.size());
    std::string line;
    char num[16];
    auto put_int = [&](int32_t v) {
        const int len = snprintf(num, sizeof num, "%d", v);
        line.append(num, (size_t)len);
    };
    for (int32_t i = 0; i < n; ++i) {
        if (!S.cl[(size_t)i].alive) continue;
        line.clear();
        bool comma = false;
        for (int32_t t = i; t >= 0; t = S.cl[(size_t)t].next) {
            const JoinCluster &ct = S.cl[(size_t)t];
            for (int64_t k = ct.ref_begin; k < ct.ref_end; ++k) {
                const int32_t id = S.ref_id[(size_t)k];
                const int32_t o = seen.get(id);
                if (o >= 0 && o != i && log)
                    fprintf(log, "\n\nError: duplicate ids found in clusters! (%d)\n", id);
                if (o != i) {  // not yet written in this cluster
                    if (comma) line.push_back(',');
                    put_int(id);
                    line.push_back(S.ref_dir[(size_t)k]);
                    st.refs_written++;
                    comma = true;
                }
                seen.set(id, i);
            }
        }
        int depth = 0;
        for (int32_t t = i; t >= 0; t = S.cl[(size_t)t].next) {
            ++depth;
            const JoinCluster &ct = S.cl[(size_t)t];
            for (int64_t k = ct.read_begin; k < ct.read_end; ++k) {
                line.push_back(',');
                put_int(S.read_id[(size_t)k]);
                line.push_back(S.read_dir[(size_t)k]);
                st.reads_written++;
            }
        }
        if (depth > st.maxdepth) st.maxdepth = depth;
        line.push_back('\n');
        fwrite(line.data(), 1, line.size(), fpo);
    }
    if (fclose(fpo) != 0) return fail(S, VSK_ERR_ARG, "write error on '%s'", out_path);
    if (stats) *stats = st;
    return VSK_OK;
}

static void copy_err(const JoinState &S, char *err, size_t errlen) {
    if (err && errlen) {
        strncpy(err, S.err.c_str(), errlen - 1);
        err[errlen - 1] = 0;
    }
}

extern "C" int vsk_join_clusters_files(const char *const *paths, int32_t n_paths, const char *out_path,
                                       vsk_join_stats *stats, char *err, size_t errlen) {
    if (err && errlen) err[0] = 0;
    if (!out_path || n_paths < 0 || (n_paths && !paths)) return VSK_ERR_ARG;
    std::vector<std::string> v;
    for (int32_t i = 0; i < n_paths; ++i) v.emplace_back(paths[i]);
    std::sort(v.begin(), v.end(), [](const std::string &a, const std::string &b) { return strcmp(a.c_str(), b.c_str()) < 0; });
    JoinState S;
    const int rc = join_files(v, out_path, stats, nullptr, S);
    if (rc) copy_err(S, err, errlen);
    return rc;
}

// the directory scan of joinc.c:160-181 / 208-238: names longer than 6 characters that end in
// ".clu" (any case); path = folder + "/" + name
extern "C" int vsk_join_clusters_dir(const char *dir, const char *out_path, vsk_join_stats *stats, int verbose,
                                     char *err, size_t errlen) {
    if (err && errlen) err[0] = 0;
    if (!dir || !out_path) return VSK_ERR_ARG;
    std::vector<std::string> v;
    if (DIR *d = opendir(dir)) {
        while (struct dirent *de = readdir(d)) {
            const size_t len = strlen(de->d_name);
            if (len > 6 && 0 == strcasecmp(".clu", de->d_name + len - 4)) {
                std::string p(dir);
                if (p.empty() || p.back() != '/') p.push_back('/');
                p += de->d_name;
                v.push_back(p);
            }
        }
        closedir(d);
    }
    std::sort(v.begin(), v.end(), [](const std::string &a, const std::string &b) { return strcmp(a.c_str(), b.c_str()) < 0; });
    JoinState S;
    const int rc = join_files(v, out_path, stats, verbose ? stdout : nullptr, S);
    if (rc) copy_err(S, err, errlen);
    return rc;
}
