// Host clustering and output: ClusterBaseAddLink / cl_mergeclusters /
// ClusterBasePrint (reference src/psearch1.91/cluster.h:268-332,452-541) over the
// link records the device produced, in call order.  Host-only code.
#include <stdio.h>
#include <stdlib.h>

#include <unordered_map>

#include "vsk_host.hpp"

namespace {
struct KeyRec {
    int32_t key;
    int32_t dir;
    int32_t cluster;
    int8_t leftcon, rightcon;  // -r mode, hclust.h:825-847
};
struct Cluster {
    std::vector<int32_t> items;  // indices into keys, in append order
    bool alive = true;
};
}  // namespace

static int clusters_write(const vsk_link *links, int64_t n, const char *clu_path, const char *map_path,
                          uint64_t *trcount, int option_r);

extern "C" int vsk_clusters_write(const vsk_link *links, int64_t n, const char *clu_path, const char *map_path,
                                  uint64_t *trcount) {
    return clusters_write(links, n, clu_path, map_path, trcount, 0);
}

// -r (ref-to-ref) flavour: members print as `key+` (a flank matched another TR exactly) or `key-`
// (cluster.h:515-522), after the leftcon/rightcon marks of hclust.h:825-847 (MAXERRORS is 0 there)
extern "C" int vsk_clusters_write_r(const vsk_link *links, int64_t n, const char *clu_path, const char *map_path,
                                    uint64_t *trcount) {
    return clusters_write(links, n, clu_path, map_path, trcount, 1);
}

static int clusters_write(const vsk_link *links, int64_t n, const char *clu_path, const char *map_path,
                          uint64_t *trcount, int option_r) {
    std::vector<KeyRec> keys;
    std::vector<Cluster> clusters;  // creation order == print order
    std::unordered_map<int32_t, int32_t> by_key;
    auto find = [&](int32_t k) -> int32_t {
        auto it = by_key.find(k);
        return it == by_key.end() ? -1 : it->second;
    };
    auto insert = [&](int32_t cl, int32_t key, int dir) {
        keys.push_back(KeyRec{key, dir, cl, 0, 0});
        by_key.emplace(key, (int32_t)keys.size() - 1);
        clusters[(size_t)cl].items.push_back((int32_t)keys.size() - 1);
    };
    FILE *mf = nullptr;
    if (map_path) {
        mf = fopen(map_path, "w");
        if (!mf) return VSK_ERR_ARG;
    }
    for (int64_t i = 0; i < n; ++i) {
        const vsk_link &l = links[i];
        const int dir = l.dir;
        if (mf) {
            const double sim = (1.0 - (l.dsum) / (255.0 * l.sumw));
            fprintf(mf, "%d%c=>%d:%.2lf:%d:%d\n", l.refkey, (0 == dir) ? '\'' : '\"', l.readkey, sim, l.lerr, l.rerr);
        }
        if (l.refkey == l.readkey) continue;  // ClusterBaseAddLink returns at once; so does the -r marking
        const int32_t k1 = find(l.refkey), k2 = find(l.readkey);
        if (k1 < 0 && k2 < 0) {
            clusters.emplace_back();
            const int32_t c = (int32_t)clusters.size() - 1;
            insert(c, l.refkey, 0);
            insert(c, l.readkey, dir);
        } else if (k1 >= 0 && k2 >= 0) {
            int32_t c1 = keys[(size_t)k1].cluster, c2 = keys[(size_t)k2].cluster;
            if (c1 != c2) {
                const int flip = keys[(size_t)k1].dir ^ keys[(size_t)k2].dir ^ dir;
                if (clusters[(size_t)c1].items.size() < clusters[(size_t)c2].items.size()) std::swap(c1, c2);
                for (int32_t it : clusters[(size_t)c2].items) {
                    keys[(size_t)it].cluster = c1;
                    keys[(size_t)it].dir ^= flip;
                }
                auto &a = clusters[(size_t)c1].items;
                auto &b = clusters[(size_t)c2].items;
                a.insert(a.end(), b.begin(), b.end());
                std::vector<int32_t>().swap(b);
                clusters[(size_t)c2].alive = false;
            }
        } else if (k1 >= 0) {
            insert(keys[(size_t)k1].cluster, l.readkey, keys[(size_t)k1].dir == 0 ? dir : !dir);
        } else {
            insert(keys[(size_t)k2].cluster, l.refkey, keys[(size_t)k2].dir == 0 ? dir : !dir);
        }
        if (option_r && abs(l.refkey) != abs(l.readkey)) {
            KeyRec &a = keys[(size_t)find(l.refkey)], &b = keys[(size_t)find(l.readkey)];
            if (l.lerr <= 0) a.leftcon = b.leftcon = 1;
            if (l.rerr <= 0) a.rightcon = b.rightcon = 1;
        }
    }
    if (mf) fclose(mf);
    if (trcount) *trcount = (uint64_t)keys.size();
    if (clu_path) {
        FILE *fp = fopen(clu_path, "w");
        if (!fp) return VSK_ERR_ARG;
        for (const Cluster &c : clusters) {
            if (!c.alive) continue;
            fputc('@', fp);
            for (int32_t it : c.items) {
                const KeyRec &k = keys[(size_t)it];
                if (option_r) fprintf(fp, (k.leftcon || k.rightcon) ? " %d+" : " %d-", k.key);
                else fprintf(fp, k.dir ? " %d\"" : " %d'", k.key);
            }
            fputc('\n', fp);
        }
        fclose(fp);
    }
    return VSK_OK;
}
