/*
 * vsk.h -- C ABI of libvsk.so, the B200 (sm_100a) implementation of VNTRseek's
 * alignment core: PROCLU (psearch) and flank alignment (flankalign).
 *
 * The reference has no FFI for this path: its boundary is argv + text files
 * (SURVEY.md section 8b).  The drop-in binaries in vntrseek_b200/csrc/ (same argv,
 * same files) are thin C front-ends over the entry points below; each entry
 * point names the reference function(s) it replaces.  All pointers are plain
 * host pointers unless the name says `_dev`; sizes are explicit; nothing in a
 * signature is a torch or C++ type.  Every function returns VSK_OK (0) or a
 * negative VSK_ERR_* code; vsk_last_error() gives the text.  There is no CPU
 * fallback: without a CUDA device vsk_init fails.
 *
 * Citations are relative to the reference tree (yzhernand/VNTRseek).
 */
#ifndef VSK_H
#define VSK_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VSK_OK 0
#define VSK_ERR_ARG (-1)    /* bad argument / malformed input            */
#define VSK_ERR_CUDA (-2)   /* CUDA runtime error (text in last_error)   */
#define VSK_ERR_NOMEM (-3)
#define VSK_ERR_ABORT (-4)  /* condition on which the reference exit(1)s */

typedef struct vsk_ctx vsk_ctx;
typedef struct vsk_profset vsk_profset;

/* ---- context -------------------------------------------------------------
 * dist501501 = the payload of eucledian.dst after its 12-byte header
 * (LoadDistanceTable, src/psearch1.91/profile.h:1035-1079).  The library
 * expands it to a symmetric matrix in HBM once (EVALDISTANCE, profile.h:1100)
 * and derives NONDASHES (InitializeCompositions, profile.h:837-886).        */
int vsk_init(vsk_ctx **ctx, int device, const uint8_t *dist501501);
void vsk_destroy(vsk_ctx *ctx);
const char *vsk_last_error(const vsk_ctx *ctx);
/* number of kernels this context has launched so far */
uint64_t vsk_launch_count(const vsk_ctx *ctx);
int vsk_sync(vsk_ctx *ctx);

/* measurement helpers: CUDA events on the library's stream; per-launch timing of
 * the DP kernel; integer-pipe peak (IADD3/VIMNMX/IMAD chains) in Gop/s          */
int vsk_timer_start(vsk_ctx *ctx);
int vsk_timer_stop(vsk_ctx *ctx, float *ms);
int vsk_profile_enable(vsk_ctx *ctx, int on);
int vsk_profile_read(vsk_ctx *ctx, double *dp_kernel_ms, int64_t *dp_kernel_launches);
int vsk_int_peak(vsk_ctx *ctx, double *gops);
/* profile mode: the forward-sweep kernel alone (vsk_dp_dir_kernel), events on the side stream each
 * launch runs on: summed duration and number of its launches, and `busy_ms` = the time during
 * which at least one launch was running (consecutive batches run on two streams and overlap, so
 * the sum of the durations exceeds it).                                                        */
int vsk_profile_read_sweep(vsk_ctx *ctx, double *sweep_ms, int64_t *sweep_launches, double *busy_ms);
/* profile mode: "stage=ms;..." of the PROCLU pipeline stages since the last call */
int vsk_profile_stages(vsk_ctx *ctx, char *buf, size_t n);

/* Reads eucledian.dst (header 777,10,5 checked) into out[501501]. */
int vsk_load_distance_file(const char *path, uint8_t *out501501);

/* ---- profile sets ---------------------------------------------------------
 * A profile = PROFILE.indices (profile.h:111-139): composition ids 0..1000.
 * off[n_prof+1] are element offsets into ids.  Uploaded once, kept in HBM.   */
int vsk_profset_upload(vsk_ctx *ctx, int32_t n_prof, const int64_t *off, const uint16_t *ids,
                       vsk_profset **out);
void vsk_profset_free(vsk_ctx *ctx, vsk_profset *ps);

/* ---- pair DP ---------------------------------------------------------------
 * Replaces GetFixedProfileAP (profile.h:2040-2227; mode 0) and
 * GetFixedProfileAPNarrowband(p1,p2,sm,homology) (profile.h:1733-2038; mode 1)
 * for n independent pairs: rows = left_set[left_idx[i]] (p1, the reference TR),
 * columns = top_set[top_idx[i]] (p2, the read TR).
 * out[i] = {PAP.distance, PAP.distanceweighted, PAP.weights, PAP.length}
 * (profile.h:141-149); similarity = 1.0 - dsum/(255.0*sumw) (profile.h:2220).
 * This is also the whole of edges mode (doEdgesBruteForce, psearch.c:1223-1405). */
typedef struct {
    int32_t distance, dsum, sumw, length;
} vsk_dp_result;

int vsk_pair_dp(vsk_ctx *ctx, const vsk_profset *left_set, const vsk_profset *top_set, int64_t n,
                const int32_t *left_idx, const int32_t *top_idx, const uint8_t *mode, int homology,
                vsk_dp_result *out);

/* Same, split so that the pair list stays resident in HBM between runs
 * (bench.py's device-resident leg): prepare uploads and schedules the pairs,
 * run launches the kernel and leaves the results in HBM, fetch copies them out. */
typedef struct vsk_dp_batch vsk_dp_batch;
int vsk_dp_batch_prepare(vsk_ctx *ctx, const vsk_profset *left_set, const vsk_profset *top_set,
                         int64_t n, const int32_t *left_idx, const int32_t *top_idx,
                         const uint8_t *mode, int homology, vsk_dp_batch **out);
int vsk_dp_batch_run(vsk_ctx *ctx, vsk_dp_batch *b);
int vsk_dp_batch_fetch(vsk_ctx *ctx, vsk_dp_batch *b, vsk_dp_result *out);
/* DP cells the reference evaluates for this batch (GCUPS numerator) */
int64_t vsk_dp_batch_cells(const vsk_dp_batch *b);
void vsk_dp_batch_free(vsk_ctx *ctx, vsk_dp_batch *b);

/* ---- cyclic pair DP -----------------------------------------------------------------
 * GetBestProfileAP (profile.h:1586-1675, with GetBoundedProfileAP :1281-1585 and PAPToLimit
 * :1231-1279): the left profile against EVERY rotation of the top profile, each rotation aligned
 * between the paths of the rotations that enclose it, best weighted distance wins (ties: the
 * rotation the reference visits first).  No program of the reference calls this aligner; it is
 * offered for rotation-invariant matching and is not part of vsk_proclu_run.  Profiles of
 * 1..VSK_CYCLIC_MAX columns.  out[i].start = the winning rotation (PAP.start).               */
#define VSK_CYCLIC_MAX 1024
typedef struct {
    int32_t distance, dsum, sumw, length, start;
    int32_t pad_[3];
} vsk_cyclic_result;
int vsk_pair_dp_cyclic(vsk_ctx *ctx, const vsk_profset *left_set, const vsk_profset *top_set, int64_t n,
                       const int32_t *left_idx, const int32_t *top_idx, vsk_cyclic_result *out);

/* ---- pair DP without any allocation ------------------------------------------------
 * (No counterpart in the reference: it exists for hosts of this library that cannot
 * allocate device memory when they need an alignment -- the GPU is owned by the
 * caller's own pool allocator, the call sits in a latency-critical section where the
 * device-wide synchronisation of cudaMalloc/cudaFree is unacceptable, or the memory
 * budget was fixed at start-up.)
 * Same results as vsk_pair_dp (it runs the one-kernel DP, which holds no direction words).  The caller passes the inputs and the output
 * as DEVICE pointers (composition ids of both profile sets, n results) and the pair
 * descriptors as a HOST array; the boundary-column scratch, the device copy of the
 * descriptors and the work counter live in one fixed device array inside libvsk.so
 * (about 21.3 MiB, held by every GPU that loads the library until the process exits).
 * Limits: at most VSK_STATIC_MAX_PAIRS pairs per call and VSK_STATIC_MAX_ROWS rows
 * (l1, the reference-side profile) per pair; larger requests return VSK_ERR_ARG with a
 * message and do nothing.  The call makes no cudaMalloc / cudaFree / allocating library
 * call; it returns after the kernel has finished.
 * vsk_init does the per-device set-up for its device (it asks CUDA for the device address of that
 * array and sets one kernel attribute, which under lazy module loading is also what makes the
 * array and the kernel resident), so the entry point itself neither loads nor allocates; should a
 * context reach this call without it (it cannot through vsk_init), the first call does the same
 * set-up -- make that call before starting any stream capture.
 * The inputs must be COMPLETE when the call is made: the library launches on its own
 * non-blocking stream, which is not ordered after the caller's streams -- synchronise the stream
 * (or event) that produced d_left_ids / d_top_ids first.  The ids must be 0..1000 (what
 * vsk_profset_upload checks for host data; here they are device data the call cannot look at), and
 * br / svector of a banded pair must be exactly what vsk_band_params returns for its lengths
 * (checked: svector == (double)l2/l1, 7 <= br).
 * All calls on a device share the array and it keeps the previous call's contents (the
 * work counter is cleared by every call): two calls that overlap on the same GPU --
 * from two host threads or two contexts -- race; serialise them.                     */
#define VSK_STATIC_MAX_PAIRS 32768
#define VSK_STATIC_MAX_ROWS 2048
typedef struct {
    int64_t left_off, top_off; /* element offsets of the two profiles in the id arrays     */
    int32_t l1, l2;            /* rows (reference TR profile) / columns (read TR profile)  */
    int32_t br;                /* band radius from vsk_band_params, or -1 = full matrix    */
    int32_t out_idx;           /* slot of this pair in d_out, 0..n-1                        */
    double svector;            /* from vsk_band_params (ignored when br < 0)              */
} vsk_dp_desc;
int vsk_pair_dp_static(vsk_ctx *ctx, const uint16_t *d_left_ids, int64_t n_left_ids,
                       const uint16_t *d_top_ids, int64_t n_top_ids, const vsk_dp_desc *pairs,
                       int64_t n, vsk_dp_result *d_out);
/* band geometry of GetFixedProfileAPNarrowband (profile.h:1752-1770), host libm */
int vsk_band_params(int32_t l1, int32_t l2, int homology, double *svector, int32_t *br);
/* Diagnostic (no GPU needed): how the forward sweep of the profile DP (the library's replacement of
 * GetFixedProfileAP's row loop, profile.h:2092-2128) cuts the l2 columns of a pair into strips:
 * plan[0] columns per lane in each of plan[1] strips of 32 lanes, then -- if plan[2] != 0 -- one last
 * strip of plan[2] columns per lane.  force: 0 = the library's choice, 1 = equal strips only,
 * 8|12|16 = equal strips of that width (the values of VSK_DD_FORCE).                               */
int vsk_dd_plan(int32_t l2, int force, int32_t plan[3]);

/* ---- string sets (consensus sequences, flanks) -----------------------------
 * ASCII, concatenated; off[n+1] byte offsets.                                */

/* LCS length of n pairs as psearch computes it: LCS_multiple_word /
 * LCS_single_word (src/psearch1.91/bitwise LCS multiple word.c:29,
 * bitwise LCS single word.c:20, incl. the 0x10 bitmask quirk).  a = string1
 * (reference consensus), b = string2 (read consensus).  out[i] = -1 on a
 * non-ACGTN character.                                                        */
int vsk_lcs_pairs(vsk_ctx *ctx, const char *a, const int64_t *a_off, const char *b,
                  const int64_t *b_off, int64_t n, const int32_t *ia, const int32_t *ib,
                  int32_t *out);

/* Edit_Distance_multiple_word_NoEndPenaltySeq1(text, pattern, N, M):
 * variant 0 = psearch copy (src/psearch1.91/bitwise edit distance alignment
 * multiple word no end penalty.c:29), variant 1 = newflankalign copy
 * (src/newflankalign/ same file name :29).  text i = text[t_off[i] .. +t_len[i]),
 * pattern i likewise.                                                         */
#define VSK_ED_PSEARCH 0
#define VSK_ED_FLANKALIGN 1
int vsk_ed_pairs(vsk_ctx *ctx, const char *text, const int64_t *t_off, const int32_t *t_len,
                 const char *pat, const int64_t *p_off, const int32_t *p_len, int64_t n,
                 int variant, int32_t *out);


/* ---- TR sets (.leb36 files in memory) ----------------------------------------
 * ReadProfileWithRC (profile.h:448-732): record i owns profile 2i (as written)
 * and 2i+1 (its reverse complement, as written), the consensus of each without
 * dash columns, and the flanks (left reversed; rc = complemented and swapped;
 * upper-cased, non-ACGT -> N; read flanks cut to maxflank, -r mode to reflen). */
typedef struct vsk_trset vsk_trset;
int vsk_trset_load_leb36(const char *path, int is_read, int maxflank, int option_r, int reflen,
                         vsk_trset **out);
/* The same with the parsed arrays kept in a binary file: parsing the text
 * (LEB36 digits, consensus, flank clean-up: profile.h:448-732,786-807) costs about 2 us per record
 * per core, more than the GPU scan costs per read.  cache_path (NULL = no cache) is read when it
 * was written for exactly this text file (size, modification time, hash of its first and last
 * 64 KB) and these parse parameters, else the text is parsed and the file (re)written through a
 * temporary + rename.  The redundancy state is not cached: apply the .rotindex afterwards as usual. */
int vsk_trset_load_leb36_cached(const char *path, int is_read, int maxflank, int option_r, int reflen,
                                const char *cache_path, vsk_trset **out, int *from_cache);
/* The same file tokenised only (second half): the LEB36 profile fields and the flank
 * text stay as they are in the file and are decoded by the GPU when the set is uploaded
 * (vsk_reads_upload / vsk_proclu_run): LEB36ToInt (profile.h:786-807), the consensus
 * (getConcensusFromProfile, :280-322) and the flank clean-up / reversal / complement (:667-695).  The
 * host keeps keys, pattern sizes and lengths; the consensus lengths arrive with the first upload.  For
 * read sets: vsk_refs_create needs the consensus strings on the host and refuses a raw set.        */
int vsk_trset_load_leb36_raw(const char *path, int is_read, int maxflank, int option_r, int reflen,
                             vsk_trset **out);
/* same from arrays: prof_off[2n+1] into ids; flank_off[2n+1] into flanks (left,
 * right per record, as they stand in the file: left NOT reversed)             */
int vsk_trset_from_arrays(int32_t n, const int32_t *key, const int32_t *patlen,
                          const int64_t *prof_off, const uint16_t *ids, const int64_t *flank_off,
                          const char *flanks, int is_read, int maxflank, int option_r, int reflen,
                          vsk_trset **out);
/* LoadRotated (psearch.c:1086-1201); a missing file is fine (returns VSK_OK). */
int vsk_trset_apply_rotindex(vsk_trset *ts, const char *path);
/* psearch.c:397-401,424-428: a key may appear once across both files */
int vsk_trsets_check_unique(const vsk_trset *a, const vsk_trset *b, int32_t *dup_key);
/* page-lock the set's large arrays (one-time) so vsk_proclu_run's H2D copies run at PCIe rate */
int vsk_trset_pin(vsk_trset *ts);
int32_t vsk_trset_size(const vsk_trset *ts);
/* Diagnostic: order-sensitive hash of everything the set holds (records, profiles, consensus,
 * flanks, redundancy state).  Large .leb36 files are parsed by several host threads
 * (VSK_LOAD_THREADS, default = cores up to 16; 1 = the serial parser); equal checksums mean the
 * two ways loaded the same set.                                                              */
uint64_t vsk_trset_checksum(const vsk_trset *ts);
const char *vsk_trset_error(const vsk_trset *ts);
void vsk_trset_free(vsk_trset *ts);

/* ---- PROCLU ---------------------------------------------------------------------
 * vsk_refs_create = everything doSearchSimilarities does to the reference set
 * before the scan (psearch.c:507-766): seed every master profile into its TRANGE
 * (and neighbours), order the buckets as the reference's sort leaves them, and
 * place TR set + index in HBM.  The TR set must outlive the handle.            */
typedef struct vsk_refs vsk_refs;
typedef struct vsk_reads vsk_reads;
typedef struct vsk_proclu_result vsk_proclu_result;
int vsk_refs_create(vsk_ctx *ctx, const vsk_trset *refs, vsk_refs **out);
/* The same with the flattened index kept in a file.  The reference rebuilds its index in every
 * psearch process (psearch.c:507-766); here the build is ~48 s of serial host work for 250 K
 * references (its sort order has to be replayed for the one-hit-beyond-`hi` rule), and every
 * process of a run uses the same reference file.  cache_path (NULL = no cache): read when it
 * holds the index of exactly this reference set (fingerprint of record count, pattern lengths,
 * master flags after the rotindex, consensus strings), else the index is built and the file
 * written through a temporary + rename.  vsk_index_cache_build does the host part alone (e.g.
 * once at install time, no GPU needed); vsk_index_cache_verify rebuilds and compares.        */
int vsk_refs_create_cached(vsk_ctx *ctx, const vsk_trset *refs, const char *cache_path, vsk_refs **out);
int vsk_index_cache_build(const vsk_trset *refs, const char *path, int *from_cache);
int vsk_index_cache_verify(const vsk_trset *refs, const char *path);
void vsk_refs_free(vsk_ctx *ctx, vsk_refs *refs);
int64_t vsk_refs_seed_hits(const vsk_refs *refs);

typedef struct {
    int32_t cutoff;      /* argv CUTOFF 70..100 (psearch.c:909-912)                  */
    int32_t maxerrors;   /* argv MAXERRORS, 0 = from flank length (hclust.h:660-668)  */
    int32_t option_r;    /* -r: ref-to-ref flank rules (hclust.h:700-751); needs maxerrors 0 */
    int32_t count_cells; /* 1: stats.dp_cells counts the reference's band cells exactly (slower) */
    int64_t key_budget;  /* candidate keys per chunk, 0 = default (or $VSK_KEY_BUDGET)  */
    int32_t links_on_device; /* 1: the links stay in HBM only (vsk_proclu_result_links_dev), no D2H */
    int32_t reserved_;
    /* optional: called after every chunk of reads with the scan-list position reached (read TRs that
     * are their redundancy group's master, the reads the reference's loop visits, psearch.c:777-845),
     * the size of that list, and stProfileAlignments so far -- what psearch's per-percent progress
     * lines report (psearch.c:836-844)                                                             */
    void (*progress)(void *user, int64_t reads_done, int64_t reads_total, uint64_t profile_alignments);
    void *progress_user;
} vsk_proclu_params;

typedef struct {                          /* BLASTSTATS, hclust.h:43-51 */
    uint64_t candidates_considered;       /* stCandidatesConsidered                  */
    uint64_t lcs_alignments;              /* stLCSalignments                         */
    uint64_t profile_alignments;          /* stProfileAlignments                     */
    uint64_t profile_alignments_success;  /* stProfileAlignmentsSuccess              */
    uint64_t confirmed;                   /* stPairsConfirmedAfterFlankAlignments    */
    int64_t dp_cells;                     /* DP cells the reference would evaluate   */
    int64_t candidate_keys;               /* keys materialised before dedup          */
    int64_t chunks;
    int64_t h2d_bytes, d2h_bytes;
    int64_t dp_alg_bytes;                 /* sum over DP pairs of 2*(L1+L2)+16 bytes  */
    int64_t dp_cells_sweep;               /* dp_cells of the pairs over 128 columns (forward-sweep kernel) */
    int64_t dp_alg_bytes_sweep;           /* dp_alg_bytes of those pairs                                   */
} vsk_proclu_stats;

typedef struct {       /* one ClusterBaseAddLink call (hclust.h:813-822), in call order */
    int32_t refkey, readkey;
    int32_t dir, lerr, rerr;
    int32_t dsum, sumw;  /* bestsim = 1.0 - dsum/(255.0*sumw) */
} vsk_link;

/* The per-read scan (psearch.c:777-845) for a whole read file: host buffers in,
 * links out (H2D of the reads and D2H of the links inside the call).           */
int vsk_proclu_run(vsk_ctx *ctx, const vsk_refs *refs, const vsk_trset *reads,
                   const vsk_proclu_params *prm, vsk_proclu_result **out);
/* same with the read set already resident in HBM */
int vsk_reads_upload(vsk_ctx *ctx, const vsk_trset *reads, vsk_reads **out);
void vsk_reads_free(vsk_ctx *ctx, vsk_reads *reads);
int vsk_proclu_run_resident(vsk_ctx *ctx, const vsk_refs *refs, const vsk_reads *reads,
                            const vsk_proclu_params *prm, vsk_proclu_result **out);
int64_t vsk_proclu_result_links(const vsk_proclu_result *r, const vsk_link **links);
/* The same table in HBM (a DEVICE pointer; it belongs to the context and is overwritten by the
 * context's next run): the send buffer of the multi-GPU merge, so that the per-rank tables go
 * from the kernel that compacts them straight into the all-gather (the analogue of join_clusters
 * reading the per-file .clu, src/join_clusters/joinc.c:246-317) without a trip through the host. */
int64_t vsk_proclu_result_links_dev(const vsk_proclu_result *r, const vsk_link **d_links);
void vsk_proclu_result_stats(const vsk_proclu_result *r, vsk_proclu_stats *stats);
/* The links live in page-locked memory owned by the context: free every result before vsk_destroy. */
void vsk_proclu_result_free(vsk_proclu_result *r);

/* ClusterBaseAddLink over the links in order + ClusterBasePrint (cluster.h:268-332,
 * 452-541) -> `<reads>.clu`; map_path != NULL also writes the -m map lines
 * (hclust.h:818-822).  *trcount = CLUSTERBASE.trcount ("total clustered").      */
int vsk_clusters_write(const vsk_link *links, int64_t n, const char *clu_path, const char *map_path,
                       uint64_t *trcount);
/* -r flavour: leftcon/rightcon marks (hclust.h:825-847) and `key+` / `key-` members (cluster.h:515-522) */
int vsk_clusters_write_r(const vsk_link *links, int64_t n, const char *clu_path, const char *map_path,
                         uint64_t *trcount);

/* ---- join_clusters -------------------------------------------------------------------
 * The merge of the per-file PROCLU cluster files into one file (reference
 * src/join_clusters/joinc.c:128-391, loader :393-571): clusters that share a reference TR
 * (negative id) are chained together, the absorbed chain flipped when the two disagree on
 * that reference's direction; output = the reference's bytes (member order included).
 * Host code, no vsk_ctx and no GPU needed.  `paths` are sorted by strcmp as the reference
 * sorts its file list; vsk_join_clusters_dir takes the folder and picks the files the
 * reference picks (names longer than 6 characters ending in ".clu", any case).
 * Errors on which the reference exit(1)s give VSK_ERR_ABORT with its message in `err`.   */
typedef struct {
    int32_t clusters_read, clusters_after_join, maxdepth, pad_;
    int64_t refs_read, refs_written, reads_read, reads_written;
} vsk_join_stats;
int vsk_join_clusters_files(const char *const *paths, int32_t n_paths, const char *out_path,
                            vsk_join_stats *stats, char *err, size_t errlen);
int vsk_join_clusters_dir(const char *dir, const char *out_path, vsk_join_stats *stats, int verbose,
                          char *err, size_t errlen);

/* ---- best-reference table ------------------------------------------------------------
 * What perl/run_rankflankmap.pl:96-231 computes from flankalign's .map files before it
 * talks to the database: per read line the hits `ref:lerr:rerr` score
 * 1-(lerr+rerr)/(leftlen+rightlen) (0 for a read without flanks), hits under 0.90 are
 * dropped, the best score wins, ties are kept.  rank_tsv gets the rows of its `rankflank`
 * table (refid, readid, score printed %.15g as Perl prints it, ties), map_tsv (may be
 * NULL) those of its `map` table (refid, readid); files of the folder whose names do not
 * start with "_map", in name order.  Host code, no vsk_ctx.                             */
typedef struct {
    int64_t files, read_lines, map_rows, rank_rows, reads_ranked;
} vsk_rank_stats;
int vsk_rankflank_dir(const char *mapdir, const char *rank_tsv, const char *map_tsv, vsk_rank_stats *stats);

/* ---- flankalign ---------------------------------------------------------------------
 * vsk_flankalign_run = the whole of flankalign.exe after argv (reference
 * src/newflankalign/flankalign.c:288-823): parse the `.flanks` blocks, order reads and
 * references of a block by pattern size with the reference's own sort, enumerate the
 * read x reference pairs inside PATLEN_SIZE_ERR %, four edit distances per pair on the
 * GPU (Edit_Distance_multiple_word_NoEndPenaltySeq1, newflankalign copy), thresholds,
 * and write `<a>_<b>.map` into the current directory.                              */
typedef struct {
    int64_t blocks, pairs, accepted;
    int64_t h2d_bytes, d2h_bytes;
    double kernel_ms;
} vsk_flank_stats;
int vsk_flankalign_run(vsk_ctx *ctx, const char *inputfile, int maxerrors, int max_flank_considered,
                       int patlen_size_err, int64_t pairs_per_batch, vsk_flank_stats *stats);

/* The per-pair logic of flankalign.c:573-749 on explicit records: left/right are offsets
 * into `pool` (left flanks already reversed, as FLANK.left holds them).
 * out[i]: bits 0-7 lerr, 8-15 rerr, bit 16 accepted, bits 17/18 edit distance returned -1. */
typedef struct {
    int64_t left_off, right_off;
    int32_t leftlen, rightlen;
} vsk_flank;
int vsk_flank_pairs(vsk_ctx *ctx, const char *pool, int64_t pool_bytes, const vsk_flank *refs, int32_t nref,
                    const vsk_flank *reads, int32_t nread, const int32_t *pair_read, const int32_t *pair_ref,
                    int64_t npairs, int maxerrors, uint32_t *out, float *kernel_ms);

/* ---- refflankalign ----------------------------------------------------------------------
 * Everything refflankalign.exe does after argv (reference
 * src/newrefflankalign/refflankalign.c:379-1074): `<xmldir>/<a>_<b>.xml` per block and
 * `<outdir>/representatives.txt`.  MAXERRORS <= 31 on the GPU build.                    */
int vsk_refflankalign_run(vsk_ctx *ctx, const char *outdir, const char *xmldir, const char *inputfile,
                          int maxerrors, int patlen_size_err, int reflen, int64_t *n_blocks, int64_t *n_pairs);
/* narrowbandUnitCostDistanceSameLengthBestScoreLastRowColumnLowMemUsesPointers(a_i, b_i, maxerr)
 * (src/newrefflankalign/narrowbandDistanceAlignment.c:533-776) for n string pairs           */
int vsk_nbdist_pairs(vsk_ctx *ctx, const char *a, const int64_t *a_off, const char *b, const int64_t *b_off,
                     int64_t n, int maxerr, int32_t *out);

/* Edges mode (doEdgesBruteForce, psearch.c:1223-1405): `id1,id2` lines in, one
 * `id1{'|"},id2,%.4lf` line per edge + `#ave:` out.                              */
int vsk_edges_run(vsk_ctx *ctx, const vsk_trset *a, const vsk_trset *b, const char *edgein_path,
                  const char *out_path, int64_t *n_edges);

#ifdef __cplusplus
}
#endif
#endif /* VSK_H */
