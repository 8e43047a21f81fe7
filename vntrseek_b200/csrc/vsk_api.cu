// extern "C" entry points of libvsk.so (declared in include/vsk.h): context,
// uploads, and the kernel-level batch calls.  The PROCLU / flankalign pipelines
// live in vsk_proclu.cu / vsk_flank.cu.
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>

#include <algorithm>
#include <numeric>

#include "vsk_internal.cuh"

int vsk_fail(vsk_ctx *ctx, int code, const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf;
    else fprintf(stderr, "libvsk: %s\n", buf);
    return code;
}

void vsk_band_host(int leftlen, int toplen, int homology, double *svector, int *br) {
    // same libm calls, same order as profile.h:1754-1769
    const double minlen = (toplen <= leftlen) ? toplen : leftlen;
    const double virtdiagonal = sqrt(pow(minlen, 2) * 2);
    const double angle = acos(homology / 100.0);
    const double topangle = atan((double)leftlen / (double)toplen);
    const double midradius = tan(angle) * (virtdiagonal / 2.0);
    const double bandradius = midradius / sin(topangle);
    const double startradius = 6;
    *svector = (double)toplen / (double)leftlen;
    *br = (int)((bandradius >= startradius) ? bandradius : startradius) + 1;
}

extern "C" int vsk_load_distance_file(const char *path, uint8_t *out) {
    FILE *fp = fopen(path, "rb");
    if (!fp) return VSK_ERR_ARG;
    int32_t header[3];
    int rc = VSK_OK;
    if (fread(header, sizeof(int32_t), 3, fp) != 3 || header[0] != 777 || header[1] != 10 ||
        header[2] != 5 || fread(out, 1, 501501, fp) != 501501)
        rc = VSK_ERR_ARG;
    fclose(fp);
    return rc;
}

static int init_device(vsk_ctx *ctx, const uint8_t *sm) {
    const int device = ctx->device;
    VSK_CUDA(ctx, cudaSetDevice(device));
    cudaDeviceProp prop;
    VSK_CUDA(ctx, cudaGetDeviceProperties(&prop, device));
    ctx->sm_count = prop.multiProcessorCount;
    VSK_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    {  // freed stream-ordered memory stays in the context's pool: the work buffers are re-used launch after launch
        cudaMemPoolProps pp;
        memset(&pp, 0, sizeof pp);
        pp.allocType = cudaMemAllocationTypePinned;
        pp.handleTypes = cudaMemHandleTypeNone;
        pp.location.type = cudaMemLocationTypeDevice;
        pp.location.id = device;
        VSK_CUDA(ctx, cudaMemPoolCreate(&ctx->pool, &pp));
        uint64_t thr = ~0ull;
        VSK_CUDA(ctx, cudaMemPoolSetAttribute(ctx->pool, cudaMemPoolAttrReleaseThreshold, &thr));
    }
    // symmetric expansion of the triangular table (EVALDISTANCE, profile.h:1100-1103)
    std::vector<uint8_t> full((size_t)VSK_NIDS * VSK_DROW, 0);
    for (int r = 0; r < VSK_NIDS; ++r)
        for (int c = r; c < VSK_NIDS; ++c) {
            const uint8_t v = sm[1000 * r - (r * r - r) / 2 + c];
            full[(size_t)r * VSK_DROW + c] = v;
            full[(size_t)c * VSK_DROW + r] = v;
        }
    // NONDASHES in the reference's enumeration order (profile.h:840-881)
    int id = 0;
    for (int a = 0; a <= 10; ++a)
        for (int c = 0; c <= 10; ++c)
            for (int g = 0; g <= 10; ++g)
                for (int t = 0; t <= 10; ++t)
                    if (a + c + g + t <= 10) ctx->h_nd[id++] = (uint8_t)(a + c + g + t);
    std::vector<uint16_t> gapnd(1024, 0);
    for (int i = 0; i < VSK_NIDS; ++i) {
        ctx->h_gap[i] = full[i];  // d(0, i)
        gapnd[i] = (uint16_t)(ctx->h_gap[i] | (ctx->h_nd[i] << 8));
    }
    VSK_CUDA(ctx, cudaMalloc(&ctx->d_dist, full.size()));
    VSK_CUDA(ctx, cudaMalloc(&ctx->d_gapnd, gapnd.size() * sizeof(uint16_t)));
    VSK_CUDA(ctx, cudaMalloc(&ctx->d_queue, 256));
    ctx->h_dist = full;
    VSK_CUDA(ctx, cudaMemcpy(ctx->d_dist, full.data(), full.size(), cudaMemcpyHostToDevice));
    VSK_CUDA(ctx, cudaMemcpy(ctx->d_gapnd, gapnd.data(), gapnd.size() * sizeof(uint16_t),
                             cudaMemcpyHostToDevice));
    // vsk_pair_dp_static must not load a module or allocate when it is called: its fixed device
    // array and its kernel become resident here
    return vsk_dp_static_setup(ctx);
}

extern "C" int vsk_init(vsk_ctx **out, int device, const uint8_t *sm) {
    if (!out || !sm) return VSK_ERR_ARG;
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return vsk_fail(nullptr, VSK_ERR_CUDA, "no CUDA device (%s); libvsk has no CPU path",
                        cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return vsk_fail(nullptr, VSK_ERR_ARG, "bad device %d", device);
    vsk_ctx *ctx = new vsk_ctx();
    ctx->device = device;
    const int rc = init_device(ctx, sm);
    if (rc != VSK_OK) {  // nobody can ask this context for its error text: print it, release what was made
        fprintf(stderr, "libvsk: vsk_init: %s\n", ctx->err.c_str());
        vsk_destroy(ctx);
        return rc;
    }
    *out = ctx;
    return VSK_OK;
}

extern "C" void vsk_destroy(vsk_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaFree(ctx->d_dist);
    cudaFree(ctx->d_gapnd);
    cudaFree(ctx->d_queue);
    cudaFree(ctx->d_dp_scratch);
    cudaFree(ctx->d_tiny_dense);
    cudaFree(ctx->d_tiny_ktab);
    cudaFree(ctx->d_links);
    for (int i = 0; i < 2; ++i)
        if (ctx->dd_stream[i]) cudaStreamDestroy(ctx->dd_stream[i]);
    for (cudaEvent_t e : ctx->dd_ev)
        if (e) cudaEventDestroy(e);
    for (auto &b : ctx->pin_free) cudaFreeHost(b.first);
    for (void *m : ctx->dd_mem) cudaFree(m);
    if (ctx->h_present_pin) cudaFreeHost(ctx->h_present_pin);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->copy_go) cudaEventDestroy(ctx->copy_go);
    if (ctx->copy_done) cudaEventDestroy(ctx->copy_done);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    if (ctx->pool) cudaMemPoolDestroy(ctx->pool);
    delete ctx;
}

void *vsk_pin_get(vsk_ctx *ctx, size_t bytes, size_t *cap) {
    if (bytes == 0) bytes = 1;
    size_t best = ctx->pin_free.size();
    for (size_t i = 0; i < ctx->pin_free.size(); ++i)  // smallest buffer that fits
        if (ctx->pin_free[i].second >= bytes && (best == ctx->pin_free.size() || ctx->pin_free[i].second < ctx->pin_free[best].second))
            best = i;
    if (best < ctx->pin_free.size()) {
        void *p = ctx->pin_free[best].first;
        *cap = ctx->pin_free[best].second;
        ctx->pin_free.erase(ctx->pin_free.begin() + (long)best);
        return p;
    }
    const size_t want = ((bytes + bytes / 4) + 65535) & ~(size_t)65535;
    void *p = nullptr;
    if (cudaHostAlloc(&p, want, cudaHostAllocDefault) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    *cap = want;
    return p;
}

void vsk_pin_put(vsk_ctx *ctx, void *p, size_t cap) {
    if (!p) return;
    if (ctx->pin_free.size() >= 8) {  // keep the list short: drop the smallest
        size_t small = 0;
        for (size_t i = 1; i < ctx->pin_free.size(); ++i)
            if (ctx->pin_free[i].second < ctx->pin_free[small].second) small = i;
        if (ctx->pin_free[small].second < cap) {
            cudaFreeHost(ctx->pin_free[small].first);
            ctx->pin_free[small] = {p, cap};
        } else {
            cudaFreeHost(p);
        }
        return;
    }
    ctx->pin_free.emplace_back(p, cap);
}

extern "C" const char *vsk_last_error(const vsk_ctx *ctx) { return ctx ? ctx->err.c_str() : "no context"; }
extern "C" uint64_t vsk_launch_count(const vsk_ctx *ctx) { return ctx ? ctx->launches : 0; }

extern "C" int vsk_sync(vsk_ctx *ctx) {
    if (!ctx) return VSK_ERR_ARG;
    VSK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return VSK_OK;
}

extern "C" int vsk_profset_upload(vsk_ctx *ctx, int32_t n, const int64_t *off, const uint16_t *ids,
                                  vsk_profset **out) {
    if (!ctx || !out || n < 0 || !off) return VSK_ERR_ARG;
    *out = nullptr;
    for (int32_t i = 0; i < n; ++i) {
        const int64_t len = off[i + 1] - off[i];
        if (len < 1 || len > VSK_MAXPROF)
            return vsk_fail(ctx, VSK_ERR_ARG, "profile %d has length %lld (allowed 1..%d)", i,
                            (long long)len, VSK_MAXPROF);
    }
    const int64_t total = n ? off[n] - off[0] : 0;
    for (int64_t i = 0; i < total; ++i)
        if (ids[off[0] + i] > 1000)
            return vsk_fail(ctx, VSK_ERR_ARG, "composition id %d out of range 0..1000", (int)ids[off[0] + i]);
    vsk_profset *ps = new vsk_profset();
    ps->n = n;
    ps->total = total;
    ps->h_off.resize((size_t)n + 1);
    for (int32_t i = 0; i <= n; ++i) ps->h_off[i] = off[i] - off[0];
    VSK_CUDA(ctx, cudaSetDevice(ctx->device));
    VSK_CUDA(ctx, cudaMalloc(&ps->d_off, sizeof(int64_t) * ((size_t)n + 1)));
    VSK_CUDA(ctx, cudaMalloc(&ps->d_ids, sizeof(uint16_t) * (size_t)(total + 8)));
    VSK_CUDA(ctx, cudaMemcpyAsync(ps->d_off, ps->h_off.data(), sizeof(int64_t) * ((size_t)n + 1),
                                  cudaMemcpyHostToDevice, ctx->stream));
    if (total)
        VSK_CUDA(ctx, cudaMemcpyAsync(ps->d_ids, ids + off[0], sizeof(uint16_t) * (size_t)total,
                                      cudaMemcpyHostToDevice, ctx->stream));
    VSK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    memset(&ps->present, 0, sizeof ps->present);
    for (int64_t i = 0; i < total; ++i) ps->present.bits[ids[off[0] + i] >> 5] |= 1u << (ids[off[0] + i] & 31);
    *out = ps;
    return VSK_OK;
}

extern "C" void vsk_profset_free(vsk_ctx *ctx, vsk_profset *ps) {
    if (!ps) return;
    if (ctx) cudaSetDevice(ctx->device);
    cudaFree(ps->d_off);
    cudaFree(ps->d_ids);
    delete ps;
}

// ---- pair DP -----------------------------------------------------------------
struct vsk_dp_batch {
    int64_t n = 0, n_tiny = 0, n_q8 = 0;
    int64_t cells = 0;
    int32_t max_l1 = 0;
    vsk_idset ids;
    const vsk_profset *left = nullptr, *top = nullptr;
    vsk_dp_job *d_jobs = nullptr;
    int4 *d_out = nullptr;
};

static int64_t band_cells(int l1, int l2, int br, double sv) {
    int64_t n = 0;
    for (int row = 1; row <= l1; ++row) {
        const int center = (int)(row * sv);
        const int a = std::max(1, center - br), b = std::min(l2, center + br);
        if (b >= a) n += b - a + 1;
    }
    return n;
}

extern "C" void vsk_dp_batch_free(vsk_ctx *ctx, vsk_dp_batch *b);

extern "C" int vsk_dp_batch_prepare(vsk_ctx *ctx, const vsk_profset *L, const vsk_profset *T, int64_t n,
                                    const int32_t *li, const int32_t *ti, const uint8_t *mode,
                                    int homology, vsk_dp_batch **out) {
    if (!ctx || !L || !T || !out || n < 0) return VSK_ERR_ARG;
    *out = nullptr;
    std::vector<vsk_dp_job> jobs((size_t)n);
    std::vector<int64_t> cost((size_t)n);
    int64_t cells = 0;
    int32_t max_l1 = 1;
    for (int64_t i = 0; i < n; ++i) {
        if (li[i] < 0 || li[i] >= L->n || ti[i] < 0 || ti[i] >= T->n)
            return vsk_fail(ctx, VSK_ERR_ARG, "pair %lld: profile index out of range", (long long)i);
        vsk_dp_job &j = jobs[(size_t)i];
        j.left_off = L->h_off[li[i]];
        j.top_off = T->h_off[ti[i]];
        j.l1 = (int32_t)(L->h_off[li[i] + 1] - j.left_off);
        j.l2 = (int32_t)(T->h_off[ti[i] + 1] - j.top_off);
        j.out_idx = (int32_t)i;
        j.br = -1;
        j.svector = 0.0;
        if (mode && mode[i]) {
            vsk_band_host(j.l1, j.l2, homology, &j.svector, &j.br);
            cells += band_cells(j.l1, j.l2, j.br, j.svector);
        } else {
            cells += (int64_t)j.l1 * j.l2;
        }
        // warp-per-pair jobs first (cost ~ strips * (rows + pipeline fill), heaviest first), tiny jobs last
        cost[(size_t)i] = (int64_t)vsk_dp_sort_key(j.l1, j.l2, j.br);
        max_l1 = std::max(max_l1, j.l1);
    }
    std::vector<int64_t> order((size_t)n);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(),
                     [&](int64_t a, int64_t b) { return cost[(size_t)a] > cost[(size_t)b]; });
    std::vector<vsk_dp_job> sorted((size_t)n);
    for (int64_t i = 0; i < n; ++i) sorted[(size_t)i] = jobs[(size_t)order[(size_t)i]];
    vsk_dp_batch *b = new vsk_dp_batch();
    b->n = n;
    for (int64_t i = 0; i < n; ++i)
    {
        if (VSK_DP_TINY(jobs[(size_t)i].l1, jobs[(size_t)i].l2, jobs[(size_t)i].br)) b->n_tiny++;
        else if (VSK_DP_Q8(jobs[(size_t)i].l1, jobs[(size_t)i].l2, jobs[(size_t)i].br)) b->n_q8++;
    }
    for (int k = 0; k < 32; ++k) b->ids.bits[k] = L->present.bits[k] | T->present.bits[k];
    b->cells = cells;
    b->max_l1 = max_l1;
    b->left = L;
    b->top = T;
    const int rc = [&]() -> int {  // the batch is released below if any of these fails
        VSK_CUDA(ctx, cudaSetDevice(ctx->device));
        VSK_CUDA(ctx, cudaMalloc(&b->d_jobs, sizeof(vsk_dp_job) * (size_t)std::max<int64_t>(n, 1)));
        VSK_CUDA(ctx, cudaMalloc(&b->d_out, sizeof(int4) * (size_t)std::max<int64_t>(n, 1)));
        if (n)
            VSK_CUDA(ctx, cudaMemcpyAsync(b->d_jobs, sorted.data(), sizeof(vsk_dp_job) * (size_t)n,
                                          cudaMemcpyHostToDevice, ctx->stream));
        VSK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        return VSK_OK;
    }();
    if (rc != VSK_OK) {
        vsk_dp_batch_free(ctx, b);
        return rc;
    }
    *out = b;
    return VSK_OK;
}

extern "C" int vsk_dp_batch_run(vsk_ctx *ctx, vsk_dp_batch *b) {
    if (!ctx || !b) return VSK_ERR_ARG;
    return vsk_dp_launch(ctx, b->left->d_ids, b->top->d_ids, b->d_jobs, b->n, b->n_q8, b->n_tiny, b->max_l1, &b->ids, b->d_out);
}

extern "C" int vsk_dp_batch_fetch(vsk_ctx *ctx, vsk_dp_batch *b, vsk_dp_result *out) {
    if (!ctx || !b || !out) return VSK_ERR_ARG;
    if (b->n)
        VSK_CUDA(ctx, cudaMemcpyAsync(out, b->d_out, sizeof(int4) * (size_t)b->n, cudaMemcpyDeviceToHost,
                                      ctx->stream));
    VSK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return VSK_OK;
}

extern "C" int64_t vsk_dp_batch_cells(const vsk_dp_batch *b) { return b ? b->cells : 0; }

extern "C" void vsk_dp_batch_free(vsk_ctx *ctx, vsk_dp_batch *b) {
    if (!b) return;
    if (ctx) cudaSetDevice(ctx->device);
    cudaFree(b->d_jobs);
    cudaFree(b->d_out);
    delete b;
}

extern "C" int vsk_pair_dp(vsk_ctx *ctx, const vsk_profset *L, const vsk_profset *T, int64_t n,
                           const int32_t *li, const int32_t *ti, const uint8_t *mode, int homology,
                           vsk_dp_result *out) {
    vsk_dp_batch *b = nullptr;
    int rc = vsk_dp_batch_prepare(ctx, L, T, n, li, ti, mode, homology, &b);
    if (rc == VSK_OK) rc = vsk_dp_batch_run(ctx, b);
    if (rc == VSK_OK) rc = vsk_dp_batch_fetch(ctx, b, out);
    vsk_dp_batch_free(ctx, b);
    return rc;
}

// ---- string kernels -----------------------------------------------------------
// device buffers of one call: freed when the call returns, on every path
struct DevBufs {
    std::vector<void *> v;
    ~DevBufs() {
        for (void *p : v) cudaFree(p);
    }
};

template <typename T>
static int dev_copy(vsk_ctx *ctx, DevBufs &bufs, T **d, const T *h, size_t n) {
    VSK_CUDA(ctx, cudaMalloc((void **)d, sizeof(T) * std::max<size_t>(n, 1)));
    bufs.v.push_back(*d);
    if (n) VSK_CUDA(ctx, cudaMemcpyAsync(*d, h, sizeof(T) * n, cudaMemcpyHostToDevice, ctx->stream));
    return VSK_OK;
}

extern "C" int vsk_lcs_pairs(vsk_ctx *ctx, const char *a, const int64_t *a_off, const char *b,
                             const int64_t *b_off, int64_t n, const int32_t *ia, const int32_t *ib,
                             int32_t *out) {
    if (!ctx || n < 0) return VSK_ERR_ARG;
    if (n == 0) return VSK_OK;
    int32_t ma = 0, mb = 0;
    for (int64_t i = 0; i < n; ++i) { ma = std::max(ma, ia[i]); mb = std::max(mb, ib[i]); }
    for (int64_t i = 0; i < n; ++i) {
        if (a_off[ia[i] + 1] - a_off[ia[i]] > VSK_MAXPROF || b_off[ib[i] + 1] - b_off[ib[i]] > VSK_MAXPROF)
            return vsk_fail(ctx, VSK_ERR_ARG, "LCS string longer than %d", VSK_MAXPROF);
    }
    VSK_CUDA(ctx, cudaSetDevice(ctx->device));
    uint8_t *da = nullptr, *db = nullptr;
    int64_t *doa = nullptr, *dob = nullptr;
    int32_t *dia = nullptr, *dib = nullptr, *dout = nullptr;
    int rc;
    DevBufs bufs;
    if ((rc = dev_copy(ctx, bufs, &da, (const uint8_t *)a, (size_t)a_off[ma + 1]))) return rc;
    if ((rc = dev_copy(ctx, bufs, &db, (const uint8_t *)b, (size_t)b_off[mb + 1]))) return rc;
    if ((rc = dev_copy(ctx, bufs, &doa, a_off, (size_t)ma + 2))) return rc;
    if ((rc = dev_copy(ctx, bufs, &dob, b_off, (size_t)mb + 2))) return rc;
    if ((rc = dev_copy(ctx, bufs, &dia, ia, (size_t)n))) return rc;
    if ((rc = dev_copy(ctx, bufs, &dib, ib, (size_t)n))) return rc;
    VSK_CUDA(ctx, cudaMalloc(&dout, sizeof(int32_t) * (size_t)n));
    bufs.v.push_back(dout);
    rc = vsk_lcs_launch(ctx, da, doa, db, dob, dia, dib, n, dout);
    if (rc == VSK_OK) {
        cudaError_t e = cudaMemcpyAsync(out, dout, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) rc = vsk_fail(ctx, VSK_ERR_CUDA, "lcs: %s", cudaGetErrorString(e));
    }
    return rc;
}

extern "C" int vsk_ed_pairs(vsk_ctx *ctx, const char *text, const int64_t *t_off, const int32_t *t_len,
                            const char *pat, const int64_t *p_off, const int32_t *p_len, int64_t n,
                            int variant, int32_t *out) {
    if (!ctx || n < 0 || (variant != VSK_ED_PSEARCH && variant != VSK_ED_FLANKALIGN)) return VSK_ERR_ARG;
    if (n == 0) return VSK_OK;
    int64_t tmax = 0, pmax = 0;
    for (int64_t i = 0; i < n; ++i) {
        if (t_len[i] < 0 || p_len[i] < 0 || t_len[i] > VSK_MAXPROF || p_len[i] > VSK_MAXPROF)
            return vsk_fail(ctx, VSK_ERR_ARG, "flank length out of range 0..%d", VSK_MAXPROF);
        tmax = std::max(tmax, t_off[i] + t_len[i]);
        pmax = std::max(pmax, p_off[i] + p_len[i]);
    }
    VSK_CUDA(ctx, cudaSetDevice(ctx->device));
    uint8_t *dt = nullptr, *dp = nullptr;
    int64_t *dto = nullptr, *dpo = nullptr;
    int32_t *dtl = nullptr, *dpl = nullptr, *dout = nullptr;
    int rc;
    DevBufs bufs;
    if ((rc = dev_copy(ctx, bufs, &dt, (const uint8_t *)text, (size_t)tmax))) return rc;
    if ((rc = dev_copy(ctx, bufs, &dp, (const uint8_t *)pat, (size_t)pmax))) return rc;
    if ((rc = dev_copy(ctx, bufs, &dto, t_off, (size_t)n))) return rc;
    if ((rc = dev_copy(ctx, bufs, &dpo, p_off, (size_t)n))) return rc;
    if ((rc = dev_copy(ctx, bufs, &dtl, t_len, (size_t)n))) return rc;
    if ((rc = dev_copy(ctx, bufs, &dpl, p_len, (size_t)n))) return rc;
    VSK_CUDA(ctx, cudaMalloc(&dout, sizeof(int32_t) * (size_t)n));
    bufs.v.push_back(dout);
    rc = vsk_ed_launch(ctx, dt, dto, dtl, dp, dpo, dpl, n, variant, dout);
    if (rc == VSK_OK) {
        cudaError_t e = cudaMemcpyAsync(out, dout, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) rc = vsk_fail(ctx, VSK_ERR_CUDA, "ed: %s", cudaGetErrorString(e));
    }
    return rc;
}

// ---- timing helpers (CUDA events on the library's own stream) --------------------
extern "C" int vsk_timer_start(vsk_ctx *ctx) {
    if (!ctx) return VSK_ERR_ARG;
    if (!ctx->t0) {
        VSK_CUDA(ctx, cudaEventCreate(&ctx->t0));
        VSK_CUDA(ctx, cudaEventCreate(&ctx->t1));
    }
    VSK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    VSK_CUDA(ctx, cudaEventRecord(ctx->t0, ctx->stream));
    return VSK_OK;
}

extern "C" int vsk_timer_stop(vsk_ctx *ctx, float *ms) {
    if (!ctx || !ms || !ctx->t0) return VSK_ERR_ARG;
    VSK_CUDA(ctx, cudaEventRecord(ctx->t1, ctx->stream));
    VSK_CUDA(ctx, cudaEventSynchronize(ctx->t1));
    VSK_CUDA(ctx, cudaEventElapsedTime(ms, ctx->t0, ctx->t1));
    return VSK_OK;
}

void vsk_stage_mark(vsk_ctx *ctx, const char *name) {
    if (!ctx->profile) return;
    cudaEvent_t e;
    if (cudaEventCreate(&e) != cudaSuccess) return;
    cudaEventRecord(e, ctx->stream);
    ctx->stage_marks.emplace_back(name, e);
}

// "stage=ms;stage=ms;..." accumulated since the last call (profile mode)
extern "C" int vsk_profile_stages(vsk_ctx *ctx, char *buf, size_t n) {
    if (!ctx || !buf || !n) return VSK_ERR_ARG;
    VSK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    std::vector<std::pair<std::string, double>> acc;
    for (size_t i = 1; i < ctx->stage_marks.size(); ++i) {
        if (!ctx->stage_marks[i].first) continue;
        float ms = 0;
        cudaEventElapsedTime(&ms, ctx->stage_marks[i - 1].second, ctx->stage_marks[i].second);
        bool found = false;
        for (auto &a : acc)
            if (a.first == ctx->stage_marks[i].first) { a.second += ms; found = true; }
        if (!found) acc.emplace_back(ctx->stage_marks[i].first, ms);
    }
    for (auto &m : ctx->stage_marks) cudaEventDestroy(m.second);
    ctx->stage_marks.clear();
    std::string out;
    char tmp[96];
    for (auto &a : acc) {
        snprintf(tmp, sizeof tmp, "%s=%.3f;", a.first.c_str(), a.second);
        out += tmp;
    }
    for (auto &a : ctx->host_acc) {  // host wall time inside named sections (not part of the device timeline)
        snprintf(tmp, sizeof tmp, "host_%s=%.3f;", a.first, a.second);
        out += tmp;
    }
    ctx->host_acc.clear();
    snprintf(buf, n, "%s", out.c_str());
    return VSK_OK;
}

extern "C" int vsk_profile_enable(vsk_ctx *ctx, int on) {
    if (!ctx) return VSK_ERR_ARG;
    ctx->profile = on != 0;
    return VSK_OK;
}

extern "C" int vsk_profile_read(vsk_ctx *ctx, double *dp_ms, int64_t *dp_launches) {
    if (!ctx) return VSK_ERR_ARG;
    VSK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    double tot = 0;
    for (size_t i = 0; i + 1 < ctx->prof_events.size(); i += 2) {
        float ms = 0;
        // a launch that failed half-way leaves its second event unrecorded: skipped
        if (cudaEventElapsedTime(&ms, ctx->prof_events[i], ctx->prof_events[i + 1]) == cudaSuccess) tot += ms;
        else cudaGetLastError();
        cudaEventDestroy(ctx->prof_events[i]);
        cudaEventDestroy(ctx->prof_events[i + 1]);
    }
    if (dp_ms) *dp_ms = tot;
    if (dp_launches) *dp_launches = (int64_t)(ctx->prof_events.size() / 2);
    ctx->prof_events.clear();
    return VSK_OK;
}

extern "C" int vsk_profile_read_sweep(vsk_ctx *ctx, double *sweep_ms, int64_t *sweep_launches, double *busy_ms) {
    if (!ctx) return VSK_ERR_ARG;
    VSK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int i = 0; i < 2; ++i)
        if (ctx->dd_stream[i]) VSK_CUDA(ctx, cudaStreamSynchronize(ctx->dd_stream[i]));
    double tot = 0;
    const size_t n = ctx->prof_sweep_events.size() / 2;
    std::vector<std::pair<double, double>> iv;  // [start, end) of every launch, ms since the first start event
    for (size_t i = 0; i < n; ++i) {
        float ms = 0, t0 = 0;
        VSK_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->prof_sweep_events[2 * i], ctx->prof_sweep_events[2 * i + 1]));
        VSK_CUDA(ctx, cudaEventElapsedTime(&t0, ctx->prof_sweep_events[0], ctx->prof_sweep_events[2 * i]));
        tot += ms;
        iv.emplace_back((double)t0, (double)t0 + ms);
    }
    for (cudaEvent_t e : ctx->prof_sweep_events) cudaEventDestroy(e);
    // launches of consecutive batches run on two streams and overlap: the time during which at least one
    // of them was running is the length of the union of their intervals
    std::sort(iv.begin(), iv.end());
    double busy = 0, cur_a = 0, cur_b = -1;
    for (auto &x : iv) {
        if (cur_b < cur_a || x.first > cur_b) {
            if (cur_b >= cur_a) busy += cur_b - cur_a;
            cur_a = x.first;
            cur_b = x.second;
        } else if (x.second > cur_b) {
            cur_b = x.second;
        }
    }
    if (cur_b >= cur_a) busy += cur_b - cur_a;
    if (sweep_ms) *sweep_ms = tot;
    if (sweep_launches) *sweep_launches = (int64_t)n;
    if (busy_ms) *busy_ms = busy;
    ctx->prof_sweep_events.clear();
    return VSK_OK;
}

// Integer-pipe peak on this device: independent chains of add / min / multiply-add.  ptxas turns
// one chain step (a+b; min(a, c+i); a*3+k) into TWO instructions, VIADDMNMX (DPX: add and min in
// one, ALU pipe) and IMAD (FMA pipe; the +k folds into the next step's immediate), so the figure
// counts 3 integer operations per 2 issued instructions: above the plain issue limit of one
// instruction per clock per scheduler (profiles/r02_int_peak_sass.txt).
__global__ void __launch_bounds__(256) vsk_int_peak_kernel(int *out, int iters, int seed) {
    int a[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = seed + threadIdx.x * (k + 1);
    const int b = seed | 1, c = seed ^ 0x5555;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            a[k] = a[k] + b;          // IADD3
            a[k] = min(a[k], c + i);  // VIMNMX (+ IADD on the uniform path)
            a[k] = a[k] * 3 + k;      // IMAD
        }
    }
    int s = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) s ^= a[k];
    if (s == 0x7fffffff) out[0] = s;
}

extern "C" int vsk_int_peak(vsk_ctx *ctx, double *gops) {
    if (!ctx || !gops) return VSK_ERR_ARG;
    const int iters = 20000, grid = ctx->sm_count * 8;
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        float ms;
        int rc = vsk_timer_start(ctx);
        if (rc) return rc;
        vsk_int_peak_kernel<<<grid, 256, 0, ctx->stream>>>(ctx->d_queue + 4, iters, rep + 1);
        VSK_CUDA(ctx, cudaGetLastError());
        rc = vsk_timer_stop(ctx, &ms);
        if (rc) return rc;
        if (rep > 0 && ms < best) best = ms;
    }
    // 3 integer instructions per chain step, 8 chains, 256 threads
    *gops = 3.0 * 8 * 256 * (double)grid * iters / (best * 1e-3) / 1e9;
    return VSK_OK;
}
