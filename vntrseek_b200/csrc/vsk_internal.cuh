// Internal declarations shared by the libvsk translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <time.h>
#include <vector>

#include "../../include/vsk.h"

#define VSK_NIDS 1001
#define VSK_DROW 1024        // padded row stride of the symmetric distance table
#define VSK_BIG 100000000    // out-of-band sentinel (reference: profile.h:1802)
#define VSK_LEN1 (1 << 17)   // path-length unit packed above sumw (sumw <= 10*(L1+L2) < 2^17)
#define VSK_MAXPROF 6000     // longest profile accepted by the DP kernels

// presence bitmap over composition ids 0..1023
struct vsk_idset {
    uint32_t bits[32];
};

struct vsk_ctx {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    // stream-ordered allocations come from a pool of this context (freed memory stays in it between
    // runs); the device's default pool, which a host application may use itself, is left alone
    cudaMemPool_t pool = nullptr;
    // copy stream of vsk_proclu_run: the composition ids and flanks of a read set travel on it while the first
    // stages of the scan already run (copy_go: buffers allocated; copy_done: data and presence bits arrived)
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t copy_go = nullptr, copy_done = nullptr;
    uint32_t *h_present_pin = nullptr;  // page-locked landing place of the presence bits computed on the copy stream
    uint8_t *d_dist = nullptr;    // [VSK_NIDS][VSK_DROW] symmetric distances
    uint16_t *d_gapnd = nullptr;  // [1024]: d(0,id) | NONDASHES[id] << 8
    uint8_t h_nd[VSK_NIDS];
    uint8_t h_gap[VSK_NIDS];
    // DP scratch (boundary columns between column strips), grown on demand
    void *d_dp_scratch = nullptr;
    size_t dp_scratch_bytes = 0;
    int *d_queue = nullptr;       // dynamic work counter
    // side streams of the direction-bit DP path: batches alternate between them so that the tail of
    // one forward sweep is filled by the next (created on first use)
    cudaStream_t dd_stream[2] = {nullptr, nullptr};
    cudaEvent_t dd_ev[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    int dd_sweeps_marked = 0;
    bool dd_join_pending = false;
    int64_t dd_plan_total = 0, dd_plan_off_q8 = 0, dd_plan_n = -1;  // vsk_dp_dir_begin -> vsk_dp_dir_launch
    // ... and its work buffers (direction words, boundary scratch, offsets, redo list): grow-only,
    // kept for the life of the context so that no launch waits for an allocation
    void *dd_mem[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    size_t dd_cap[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    // accepted links of the last PROCLU run, kept on the device as well (grow-only; what
    // vsk_proclu_result_links_dev hands out, e.g. as the send buffer of the multi-GPU all-gather)
    void *d_links = nullptr;
    size_t d_links_cap = 0;       // bytes
    std::vector<uint8_t> h_dist;  // host copy of the symmetric table
    // dense-alphabet tables of the thread-per-pair DP kernel, cached per id set
    vsk_idset tiny_key;
    bool tiny_valid = false;
    int tiny_K = 0;
    uint8_t *d_tiny_dense = nullptr;  // [1024] composition id -> dense index
    uint8_t *d_tiny_ktab = nullptr;   // [K*K] distances between present ids
    // page-locked host buffers handed out by vsk_pin_get and returned by vsk_pin_put (results of a
    // run live in one until they are freed, so the buffer of one run serves the next)
    std::vector<std::pair<void *, size_t>> pin_free;
    std::string err;
    uint64_t launches = 0;        // kernels launched through this context
    // optional per-launch timing of the DP kernel (bench.py's roofline leg)
    bool profile = false;
    std::vector<cudaEvent_t> prof_events;  // start/stop pairs
    std::vector<cudaEvent_t> prof_sweep_events;  // start/stop pairs around every vsk_dp_dir_kernel launch
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    // stage marks of the PROCLU pipeline (profile mode): name of the stage that ENDS at the event
    std::vector<std::pair<const char *, cudaEvent_t>> stage_marks;
    // host wall time spent in named sections (profile mode), reported by vsk_profile_stages as host_<name>
    std::vector<std::pair<const char *, double>> host_acc;
    // answers of driver queries that do not change while the context lives
    int occ_dp = 0, occ_dd = 0, occ_dq[2] = {0, 0};
    int64_t dd_batch_words = 0;  // batch size of the direction-word buffer (chosen at the first sizing)
};

inline double vsk_now_ms() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}
inline void vsk_host_acc(vsk_ctx *ctx, const char *name, double ms) {
    if (!ctx->profile) return;
    for (auto &a : ctx->host_acc)
        if (a.first == name) { a.second += ms; return; }
    ctx->host_acc.emplace_back(name, ms);
}

// profile mode: drop an event that closes stage `name` (nullptr = start of a run)
void vsk_stage_mark(vsk_ctx *ctx, const char *name);

struct vsk_profset {
    vsk_idset present;
    int32_t n = 0;
    int64_t total = 0;
    int64_t *d_off = nullptr;   // [n+1]
    uint16_t *d_ids = nullptr;  // [total]
    std::vector<int64_t> h_off; // host copy of the offsets (lengths are needed for scheduling)
};

int vsk_fail(vsk_ctx *ctx, int code, const char *fmt, ...);

// page-locked host memory from the context's free list (nullptr when the allocation fails)
void *vsk_pin_get(vsk_ctx *ctx, size_t bytes, size_t *cap);
void vsk_pin_put(vsk_ctx *ctx, void *p, size_t cap);

#define VSK_CUDA(ctx, call)                                                              \
    do {                                                                                 \
        cudaError_t e__ = (call);                                                        \
        if (e__ != cudaSuccess)                                                          \
            return vsk_fail((ctx), VSK_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, \
                            cudaGetErrorString(e__));                                    \
    } while (0)

// Band geometry of the narrowband aligner computed with the host libm in the
// same operation order as the reference (profile.h:1752-1770).
void vsk_band_host(int leftlen, int toplen, int homology, double *svector, int *br);

// kernels / launchers -----------------------------------------------------
struct vsk_dp_job {   // one pair as the kernel sees it (sorted by cost, heaviest first); vsk_dp_desc mirrors it
    int64_t left_off, top_off;
    int32_t l1, l2;
    int32_t br;       // band radius, <0 = full matrix
    int32_t out_idx;  // where the result goes
    double svector;
};

// Jobs are ordered: the warp-per-pair jobs (heaviest first), then `n_tiny` jobs that satisfy
// VSK_DP_TINY, which go to the thread-per-pair kernel when `ids` (the composition ids that occur
// in both profile sets) is given and small enough for the dense table to fit in shared memory.
#define VSK_DP_TINY(l1, l2, br) ((br) < 0 && (l2) <= 32 && (l1) <= 64)
// ... preceded by `n_q8` jobs for the quarter-warp kernel: at most 128 columns and not tiny
#define VSK_DP_Q8(l1, l2, br) ((l2) <= 128 && !VSK_DP_TINY(l1, l2, br))
// scheduling key (sorted descending): class, then size
__host__ __device__ inline uint32_t vsk_dp_sort_key(int l1, int l2, int br) {
    if (VSK_DP_TINY(l1, l2, br)) return (uint32_t)(l1 * 64 + l2);
    if (VSK_DP_Q8(l1, l2, br)) return (1u << 20) | ((uint32_t)((l2 > 64) ? 2 : (l2 > 32) ? 1 : 0) << 14) | (uint32_t)min(l1, 16383);
    return (2u << 20) + (uint32_t)(((l2 + 255) >> 8) * (l1 + 31));
}
int vsk_dp_launch(vsk_ctx *ctx, const uint16_t *d_left_ids, const uint16_t *d_top_ids,
                  const vsk_dp_job *d_jobs, int64_t n_jobs, int64_t n_q8, int64_t n_tiny, int32_t max_l1,
                  const vsk_idset *ids, int4 *d_out);
// Pairs with more than 128 columns go through the direction-bit kernels of vsk_dp_dir.cu (forward
// sweep + trace); vsk_dp_kernel skips them when told so and takes the pairs handed back on `redo`.
#define VSK_DP_DIRJOB(l2) ((l2) > 128)
// vsk_dp_dir_begin marks the point of ctx->stream the side streams start from (call it before the
// other DP kernels are enqueued so that they overlap); vsk_dp_dir_launch joins back into ctx->stream.
// *redo / *redo_count point into the context's buffers (valid until the next call).
int vsk_dp_dir_begin(vsk_ctx *ctx, const vsk_dp_job *d_jobs, int64_t n_warp, int64_t n_q8);
// d_jobs[0..n_warp) are the warp-per-pair jobs, d_jobs[n_warp..n_warp+n_q8) the 8-lane jobs (dense
// K x K table of the context).
int vsk_dp_dir_wait_sweeps(vsk_ctx *ctx);
int vsk_dp_dir_join(vsk_ctx *ctx);
int vsk_dp_dir_launch(vsk_ctx *ctx, const uint16_t *d_left_ids, const uint16_t *d_top_ids, const vsk_dp_job *d_jobs,
                      int64_t n_warp, int64_t n_q8, int K, int32_t max_l1, int4 *d_out, vsk_dp_job **redo,
                      int **redo_count);
// per-device set-up of vsk_pair_dp_static (address of its fixed device array, kernel attribute)
int vsk_dp_static_setup(vsk_ctx *ctx);
// which composition ids occur in d_ids[0..n)
int vsk_id_presence(vsk_ctx *ctx, const uint16_t *d_ids, int64_t n, vsk_idset *out);
int vsk_id_presence_async(vsk_ctx *ctx, const uint16_t *d_ids, int64_t n, vsk_idset *out, cudaStream_t st);

int vsk_lcs_launch(vsk_ctx *ctx, const uint8_t *d_seqA, const int64_t *d_offA, const uint8_t *d_seqB,
                   const int64_t *d_offB, const int32_t *d_ia, const int32_t *d_ib, int64_t n,
                   int32_t *d_out);

int vsk_ed_launch(vsk_ctx *ctx, const uint8_t *d_text, const int64_t *d_toff, const int32_t *d_tlen,
                  const uint8_t *d_pat, const int64_t *d_poff, const int32_t *d_plen, int64_t n,
                  int variant, int32_t *d_out);
