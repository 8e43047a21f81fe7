// Best-reference assignment from flankalign's .map files (reference
// perl/run_rankflankmap.pl:96-231, the part before the database): per read line, every
// `ref:lerr:rerr` hit scores 1 - (lerr+rerr)/(leftlen+rightlen) (0 when the read has no flanks),
// hits under 0.90 are dropped, the best score wins and ties are all kept.  Host code (a parse and a
// running maximum per line); the rows the reference inserts into its `rankflank` and `map` tables
// are written as tab-separated files it can bulk-load instead.
//
// Perl semantics that matter here: `split` drops trailing empty fields (a read line without hits
// has 7 fields and is skipped, :110-111); the hit must match /^-?(\d+):(\d+):(\d+)/ (the sign of
// the reference id is dropped, anything after the third number is ignored, :126); the arithmetic
// is IEEE double; `index($file,"_map")` is false only for names that START with "_map" (:99).
#include <dirent.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/vsk.h"

namespace {

struct Field {
    const char *b, *e;
};

// Perl's split(sep, line): fields between separators, trailing empty fields removed
void perl_split(const char *b, const char *e, char sep, std::vector<Field> &out) {
    out.clear();
    const char *s = b;
    for (const char *p = b;; ++p) {
        if (p == e || *p == sep) {
            out.push_back({s, p});
            if (p == e) break;
            s = p + 1;
        }
    }
    while (!out.empty() && out.back().b == out.back().e) out.pop_back();
}

// numeric value of a string the way Perl numifies an integer-looking field (leading blanks, sign, digits)
double perl_num(Field f) {
    std::string s(f.b, f.e);
    return strtod(s.c_str(), nullptr);
}

// /^-?(\d+):(\d+):(\d+)/
bool match_hit(Field f, Field &id, double &e1, double &e2) {
    const char *p = f.b;
    if (p < f.e && *p == '-') ++p;
    const char *part[3][2];
    for (int k = 0; k < 3; ++k) {
        part[k][0] = p;
        while (p < f.e && *p >= '0' && *p <= '9') ++p;
        part[k][1] = p;
        if (part[k][0] == p) return false;
        if (k < 2) {
            if (p >= f.e || *p != ':') return false;
            ++p;
        }
    }
    id = {part[0][0], part[0][1]};
    e1 = strtod(std::string(part[1][0], part[1][1]).c_str(), nullptr);
    e2 = strtod(std::string(part[2][0], part[2][1]).c_str(), nullptr);
    return true;
}

}  // namespace

static int rank_one_file(const std::string &path, FILE *frank, FILE *fmap, vsk_rank_stats &st) {
    FILE *fp = fopen(path.c_str(), "rb");
    if (!fp) return VSK_ERR_ARG;
    std::vector<char> buf;
    fseek(fp, 0, SEEK_END);
    const long sz = ftell(fp);
    fseek(fp, 0, SEEK_SET);
    buf.resize(sz > 0 ? (size_t)sz : 0);
    const size_t got = sz > 0 ? fread(buf.data(), 1, (size_t)sz, fp) : 0;
    fclose(fp);
    if (got != buf.size()) return VSK_ERR_ARG;
    std::vector<Field> f, hits, best;
    const char *p = buf.data(), *end = p + buf.size();
    while (p < end) {
        const char *eol = (const char *)memchr(p, '\n', (size_t)(end - p));
        const char *le = eol ? eol : end;  // chomp
        perl_split(p, le, '\t', f);
        p = eol ? eol + 1 : end;
        if (f.size() < 8) continue;
        st.read_lines++;
        const double flanks = perl_num(f[2]) + perl_num(f[3]);
        perl_split(f[7].b, f[7].e, ',', hits);
        double bestscore = 0;
        best.clear();
        for (const Field &h : hits) {
            Field id;
            double e1, e2;
            if (!match_hit(h, id, e1, e2)) continue;
            const double score = (flanks == 0) ? 0.0 : 1 - ((e1 + e2) / flanks);
            if (score >= 0.90) {
                if (score > bestscore) {
                    best.clear();
                    best.push_back(id);
                    bestscore = score;
                } else if (score == bestscore) {
                    best.push_back(id);
                }
            }
            st.map_rows++;
            if (fmap) {
                fwrite(id.b, 1, (size_t)(id.e - id.b), fmap);
                fputc('\t', fmap);
                fwrite(f[0].b, 1, (size_t)(f[0].e - f[0].b), fmap);
                fputc('\n', fmap);
            }
        }
        for (const Field &id : best) {
            st.rank_rows++;
            fwrite(id.b, 1, (size_t)(id.e - id.b), frank);
            fputc('\t', frank);
            fwrite(f[0].b, 1, (size_t)(f[0].e - f[0].b), frank);
            fprintf(frank, "\t%.15g\t%d\n", bestscore, (int)best.size() - 1);
        }
        if (!best.empty()) st.reads_ranked++;
    }
    return VSK_OK;
}

// rows of the reference's `rankflank` table (refid, readid, score, ties) -> rank_tsv and, when
// map_tsv is given, of its `map` table (refid, readid); files of the folder in name order
extern "C" int vsk_rankflank_dir(const char *mapdir, const char *rank_tsv, const char *map_tsv, vsk_rank_stats *stats) {
    if (!mapdir || !rank_tsv) return VSK_ERR_ARG;
    std::vector<std::string> names;
    DIR *d = opendir(mapdir);
    if (!d) return VSK_ERR_ARG;
    while (struct dirent *de = readdir(d)) {
        if (0 == strncmp(de->d_name, "_map", 4)) continue;  // run_rankflankmap.pl:99
        std::string p = std::string(mapdir) + "/" + de->d_name;
        struct stat sb;
        if (stat(p.c_str(), &sb) != 0 || !S_ISREG(sb.st_mode)) continue;  // ".", "..": nothing to read
        names.push_back(de->d_name);
    }
    closedir(d);
    std::sort(names.begin(), names.end());
    FILE *frank = fopen(rank_tsv, "w");
    if (!frank) return VSK_ERR_ARG;
    FILE *fmap = map_tsv ? fopen(map_tsv, "w") : nullptr;
    if (map_tsv && !fmap) {
        fclose(frank);
        return VSK_ERR_ARG;
    }
    vsk_rank_stats st;
    memset(&st, 0, sizeof st);
    int rc = VSK_OK;
    for (const std::string &n : names) {
        st.files++;
        rc = rank_one_file(std::string(mapdir) + "/" + n, frank, fmap, st);
        if (rc) break;
    }
    if (fmap) fclose(fmap);
    fclose(frank);
    if (stats) *stats = st;
    return rc;
}
