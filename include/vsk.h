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

#ifdef __cplusplus
}
#endif
#endif /* VSK_H */
