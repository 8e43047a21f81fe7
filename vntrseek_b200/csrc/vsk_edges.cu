// Edges mode of psearch (doEdgesBruteForce, reference src/psearch1.91/psearch.c:1223-1405):
// for every "id1,id2" line four full-matrix profile alignments (fwd/rc x fwd/rc),
// the orientation from the first two, the best similarity of the four.
#include <stdio.h>

#include "vsk_host.hpp"
#include "vsk_internal.cuh"

extern "C" int vsk_edges_run(vsk_ctx *ctx, const vsk_trset *A, const vsk_trset *B, const char *edgein_path,
                             const char *out_path, int64_t *n_edges) {
    if (!ctx || !A || !B || !edgein_path || !out_path) return VSK_ERR_ARG;
    FILE *fi = fopen(edgein_path, "r");
    if (!fi) return vsk_fail(ctx, VSK_ERR_ARG, "Unable to open input file '%s'!", edgein_path);
    FILE *fo = fopen(out_path, "w");
    if (!fo) { fclose(fi); return vsk_fail(ctx, VSK_ERR_ARG, "Unable to create output file '%s'", out_path); }
    // one profile set: A's profiles then B's
    std::vector<int64_t> off;
    std::vector<uint16_t> ids(A->ids);
    ids.insert(ids.end(), B->ids.begin(), B->ids.end());
    off.assign(A->prof_off.begin(), A->prof_off.end());
    for (size_t i = 1; i < B->prof_off.size(); ++i) off.push_back(A->prof_off.back() + B->prof_off[i]);
    std::unordered_map<int32_t, int32_t> by_key;  // key -> first profile index in the merged set
    for (int32_t i = 0; i < A->n; ++i)
        if (!by_key.emplace(A->key[(size_t)i], 2 * i).second) {
            fclose(fi); fclose(fo);
            return vsk_fail(ctx, VSK_ERR_ABORT, "ERROR: duplicate index!");
        }
    for (int32_t i = 0; i < B->n; ++i)
        if (!by_key.emplace(B->key[(size_t)i], 2 * (A->n + i)).second) {
            fclose(fi); fclose(fo);
            return vsk_fail(ctx, VSK_ERR_ABORT, "ERROR: duplicate index!");
        }
    std::vector<int32_t> li, ti, k1, k2;
    char buffer[1010];
    int rc = VSK_OK;
    while (fgets(buffer, 1000, fi)) {
        int id1, id2;
        if (2 != sscanf(buffer, "%d,%d", &id1, &id2)) continue;
        auto a = by_key.find(id1), b = by_key.find(id2);
        if (a == by_key.end() || b == by_key.end()) {
            rc = vsk_fail(ctx, VSK_ERR_ABORT, "ERROR: could not lookup profile id: %d!", a == by_key.end() ? id1 : id2);
            break;
        }
        const int32_t pa = a->second, pb = b->second;
        const int32_t l4[4] = {pa, pa + 1, pa, pa + 1}, t4[4] = {pb, pb, pb + 1, pb + 1};
        li.insert(li.end(), l4, l4 + 4);
        ti.insert(ti.end(), t4, t4 + 4);
        k1.push_back(id1);
        k2.push_back(id2);
    }
    fclose(fi);
    if (rc == VSK_OK) {
        vsk_profset *ps = nullptr;
        rc = vsk_profset_upload(ctx, (int32_t)off.size() - 1, off.data(), ids.data(), &ps);
        std::vector<vsk_dp_result> res(li.size() + 1);
        if (rc == VSK_OK)
            rc = vsk_pair_dp(ctx, ps, ps, (int64_t)li.size(), li.data(), ti.data(), nullptr, 0, res.data());
        vsk_profset_free(ctx, ps);
        if (rc == VSK_OK) {
            double totsim = 0;
            long long count = 0;
            for (size_t e = 0; e < k1.size(); ++e) {
                double s[4];
                for (int k = 0; k < 4; ++k)
                    s[k] = (1.0 - (res[4 * e + k].dsum) / (255.0 * res[4 * e + k].sumw));
                const int dir = (s[0] >= s[1]) ? 0 : 1;             // psearch.c:1362-1368
                double best = (s[0] >= s[1]) ? s[0] : s[1];
                const double best2 = (s[2] >= s[3]) ? s[2] : s[3];  // psearch.c:1381-1387
                best = (best >= best2) ? best : best2;              // max(), profile.h:91
                ++count;
                totsim += best;
                fprintf(fo, "%d%c,%d,%.4lf\n", k1[e], (dir == 0) ? '\'' : '\"', k2[e], best);
            }
            fprintf(fo, "#ave: %.4lf\n", totsim / count);
            if (n_edges) *n_edges = count;
        }
    }
    fclose(fo);
    return rc;
}
