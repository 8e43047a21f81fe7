/*
 * psearch.exe -- drop-in replacement for the reference's PROCLU binary
 * (reference src/psearch1.91/psearch.c:869-1083): same argv, same input files,
 * same .clu / .map / .edges outputs, same exit codes, the counters block on
 * stderr in the same format (it lands in <reads>.proclu_log, run_proclu.pl:169).
 * All computation happens in libvsk.so on the GPU; this file is argv + printing.
 *
 * GPU selection: environment variable VSK_DEVICE (default 0).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../../include/vsk.h"



static void die(const char *msg) {
    fprintf(stderr, "\n\nERROR!!!!:\n%s", msg);
    fflush(stderr);
    exit(1);
}

/* VSK_TRSET_CACHE=1: keep the parsed arrays of every .leb36 file in <file>.vsktr next to it (read
 * back when the text has not changed since); unset = parse the text in every process as the reference does */
static int load_profiles(const char *path, int is_read, int maxflank, int option_r, int reflen, vsk_trset **out) {
    const char *e = getenv("VSK_TRSET_CACHE");
    const char *g = getenv("VSK_GPU_DECODE"); /* =1: read files are only tokenised here, decoded on the GPU */
    if (is_read && g && *g && strcmp(g, "0") != 0)
        return vsk_trset_load_leb36_raw(path, is_read, maxflank, option_r, reflen, out);
    if (e && *e && strcmp(e, "0") != 0) {
        char cache[1200];
        snprintf(cache, sizeof cache, "%s.vsktr", path);
        return vsk_trset_load_leb36_cached(path, is_read, maxflank, option_r, reflen, cache, out, NULL);
    }
    return vsk_trset_load_leb36(path, is_read, maxflank, option_r, reflen, out);
}

/* The reference's per-percent progress lines (psearch.c:836-844): "\nP% (ProfileAlignments: N, total
 * time: T secs)", one for every percent of the scan list, P counting up from 0.  The scan runs chunk by
 * chunk on the GPU, so the lines of all the marks a chunk crosses appear together when it ends, each with
 * the ProfileAlignments count reached at that point. */
static time_t progress_start;
static int progress_percent = 0;
static void progress_lines(void *user, int64_t done, int64_t total, uint64_t profile_alignments) {
    (void) user;
    if (total <= 0) return;
    /* the reference prints at most one line per read: marks are P = 0 .. min(100, reads) */
    while (progress_percent <= 100 && progress_percent < 101 &&
           (int) (100 * (2 * done) / (double) (2 * total)) >= progress_percent && progress_percent <= done) {
        fprintf(stderr, "\n%d%% (ProfileAlignments: %llu, total time: %.1lf secs)", progress_percent,
                (unsigned long long) profile_alignments, (double) (time(NULL) - progress_start));
        progress_percent++;
    }
    fflush(stderr);
}

int main(int argc, char **argv) {
    char outputfile[1100], mapfile[1100], edgesfile[1100], buffer[1100];
    const char *edgesIn = NULL;
    int cutoff2 = 0, maxerrors, maxflank, reflen = 50, option_r = 0, option_m = 0, i, rc;
    static uint8_t table[501501];
    vsk_ctx *ctx = NULL;
    vsk_trset *refs = NULL, *reads = NULL;
    time_t startTime;

    if (argc < 7) {
        printf("\nPSEARCH v1.91 (B200) - Finds clusters from list of profiles.");
        printf("\n\nUsage:  %s REFPROFILES READPROFILES DISTANCEFILE CUTOFF(70-100) "
               "MAXERRORS MAXFLANKCONSIDERED [EDGESINFILE] [OPTION]\n\n", argv[0]);
        printf("\tOPTIONS: \n");
        printf("\t\t-r ref-to-ref alignments, only one flank has to match, 50 is default length "
               "but may be followed by integer flank length, ex -r 50 \n");
        printf("\t\t-m produce map file  \n\n");
        exit(1);
    }
    snprintf(outputfile, sizeof outputfile, "%s.clu", argv[2]);
    snprintf(mapfile, sizeof mapfile, "%s.map", argv[2]);
    snprintf(edgesfile, sizeof edgesfile, "%s.edges", argv[2]);
    sscanf(argv[4], "%d", &cutoff2);
    if (0 == cutoff2 || cutoff2 < 70 || cutoff2 > 100) {
        printf("\nERROR: Please enter similarity cutoff (70-100)!");
        exit(1);
    }
    maxerrors = atoi(argv[5]);
    maxflank = atoi(argv[6]);
    if (maxflank > 1000000) maxflank = 1000000;
    /* option parsing exactly as psearch.c:919-981 */
    if (argc >= 8 && (0 == strcmp(argv[7], "-r") || 0 == strcmp(argv[7], "-R"))) {
        option_r = 1;
        if (argc >= 9 && 0 != atoi(argv[8])) reflen = atoi(argv[8]);
    } else {
        edgesIn = (argc >= 8) ? argv[7] : NULL;
        if (edgesIn && (strlen(edgesIn) < 7 || 0 != strcmp(edgesIn + strlen(edgesIn) - 7, ".edgein"))) {
            printf("\nERROR: Please make sure edgesin file has extension .edgein!");
            exit(1);
        }
    }
    for (i = 8; i <= 9; i++)
        if (argc >= i + 1 && (0 == strcmp(argv[i], "-r") || 0 == strcmp(argv[i], "-R"))) {
            option_r = 1;
            if (argc >= i + 2 && 0 != atoi(argv[i + 1])) reflen = atoi(argv[i + 1]);
        }
    for (i = 7; i <= 9; i++)
        if (argc >= i + 1 && (0 == strcmp(argv[i], "-m") || 0 == strcmp(argv[i], "-M"))) option_m = 1;
    if (option_m) {
        FILE *mf = fopen(mapfile, "w");
        if (mf == NULL) {
            fprintf(stderr, "\nERROR: Unable to open mapfile file '%s'", mapfile);
            exit(1);
        }
        fclose(mf);
    }
    if (maxerrors != 0 && option_r) {
        fprintf(stderr, "\nERROR: maxerror must be 0 because it will be dynamically generated based on flank length.");
        exit(1);
    }
    startTime = time(NULL);
    progress_start = startTime;
    {
        FILE *f = fopen(argv[1], "r");
        if (!f) { fprintf(stderr, "\nERROR: Unable to open input file '%s'", argv[1]); exit(1); }
        fclose(f);
        f = fopen(argv[2], "r");
        if (!f) { fprintf(stderr, "\nERROR: Unable to open input file '%s'", argv[2]); exit(1); }
        fclose(f);
    }
    if (vsk_load_distance_file(argv[3], table) != VSK_OK) {
        fprintf(stderr, "\nERROR: Unable to load distance table '%s'", argv[3]);
        exit(1);
    }
    if (vsk_init(&ctx, getenv("VSK_DEVICE") ? atoi(getenv("VSK_DEVICE")) : 0, table) != VSK_OK) {
        fprintf(stderr, "\nERROR: no usable CUDA device (this psearch has no CPU path)\n");
        exit(1);
    }

    if (edgesIn) { /* psearch.c:1032-1045: both files are loaded as reference-type sets */
        int64_t n = 0;
        if (vsk_trset_load_leb36(argv[1], 0, maxflank, option_r, reflen, &refs) != VSK_OK ||
            vsk_trset_load_leb36(argv[2], 0, maxflank, option_r, reflen, &reads) != VSK_OK)
            die("unable to read profiles");
        fprintf(stdout, "\nLoaded %d profiles from 1st file", (int) vsk_trset_size(refs));
        fprintf(stdout, "\nLoaded %d profiles from 2nd file", (int) vsk_trset_size(reads));
        rc = vsk_edges_run(ctx, refs, reads, edgesIn, edgesfile, &n);
        if (rc != VSK_OK) die(vsk_last_error(ctx));
        fprintf(stdout, "\nAlignment complete\n");
        exit(0);
    }

    fprintf(stderr, "\nReading reference profiles...");
    if (load_profiles(argv[1], 0, maxflank, option_r, reflen, &refs) != VSK_OK)
        die(refs ? vsk_trset_error(refs) : "unable to read reference profiles");
    fprintf(stderr, "(total time: %.1lf secs)", (double) (time(NULL) - startTime));
    fprintf(stderr, "\nLoaded %d profiles from reference file!", (int) vsk_trset_size(refs));
    fprintf(stderr, "\nReading read profiles...");
    if (load_profiles(argv[2], 1, maxflank, option_r, reflen, &reads) != VSK_OK)
        die(reads ? vsk_trset_error(reads) : "unable to read read profiles");
    fprintf(stderr, "(total time: %.1lf secs)", (double) (time(NULL) - startTime));
    fprintf(stderr, "\nLoaded %d profiles from reads file!", (int) vsk_trset_size(reads));
    {
        int dup = 0;
        if (vsk_trsets_check_unique(refs, reads, &dup) != VSK_OK) {
            printf("\nERROR: duplicate index %d!", dup);
            exit(1);
        }
    }
    snprintf(buffer, sizeof buffer, "%s.rotindex", argv[1]);
    if (vsk_trset_apply_rotindex(refs, buffer) != VSK_OK) die(vsk_trset_error(refs));
    snprintf(buffer, sizeof buffer, "%s.rotindex", argv[2]);
    if (vsk_trset_apply_rotindex(reads, buffer) != VSK_OK) die(vsk_trset_error(reads));

    {
        vsk_refs *R = NULL;
        vsk_proclu_result *res = NULL;
        vsk_proclu_params prm;
        vsk_proclu_stats st;
        const vsk_link *links = NULL;
        int64_t nl;
        uint64_t trcount = 0;
        fprintf(stderr, "\nSeeding reference profiles' concensus sequences...");
        {   /* VSK_INDEX_CACHE: unset = build the index in this process; "1" = keep it in
               <REFPROFILES>.vskidx; anything else = that file */
            const char *ic = getenv("VSK_INDEX_CACHE");
            char icpath[1200];
            const char *cache = NULL;
            if (ic && *ic && strcmp(ic, "0") != 0) {
                if (0 == strcmp(ic, "1")) {
                    snprintf(icpath, sizeof icpath, "%s.vskidx", argv[1]);
                    cache = icpath;
                } else {
                    cache = ic;
                }
            }
            if (vsk_refs_create_cached(ctx, refs, cache, &R) != VSK_OK) die(vsk_last_error(ctx));
        }
        fprintf(stderr, "(total time: %.1lf secs)", (double) (time(NULL) - startTime));
        fprintf(stderr, "\nScanning reads...");
        memset(&prm, 0, sizeof prm);
        prm.progress = progress_lines;
        prm.cutoff = cutoff2;
        prm.maxerrors = maxerrors;
        prm.option_r = option_r;
        if (vsk_proclu_run(ctx, R, reads, &prm, &res) != VSK_OK) die(vsk_last_error(ctx));
        vsk_proclu_result_stats(res, &st);
        nl = vsk_proclu_result_links(res, &links);
        /* the files are written before the counters block is printed (one pass over the links gives
           both "total clustered" and the files); the block itself reads as the reference's */
        rc = option_r ? vsk_clusters_write_r(links, nl, outputfile, option_m ? mapfile : NULL, &trcount)
                      : vsk_clusters_write(links, nl, outputfile, option_m ? mapfile : NULL, &trcount);
        fprintf(stderr,
                "\nscan finished!\n\t"
                "CandidatesConsidered: %llu\n\t"
                "LCSalignments: %llu\n\t"
                "ProfileAlignments: %llu\n\t"
                "ProfileAlignmentsSuccess: %llu\n\t"
                "pairs confirmed after flank LCS alignments: %llu\n\t"
                "total clustered: %llu\n\t"
                "total time: %.1lf secs\n\t"
                "writing results ...",
                (unsigned long long) st.candidates_considered, (unsigned long long) st.lcs_alignments,
                (unsigned long long) st.profile_alignments, (unsigned long long) st.profile_alignments_success,
                (unsigned long long) st.confirmed, (unsigned long long) trcount,
                (double) (time(NULL) - startTime));
        fflush(stderr);
        if (rc != VSK_OK) {
            fprintf(stderr, "\nERROR: Unable to create output file '%s'", outputfile);
            exit(1);
        }
        vsk_proclu_result_free(res);
        vsk_refs_free(ctx, R);
    }
    fprintf(stderr, "done (total time: %.1lf secs)\n\n", (double) (time(NULL) - startTime));
    fflush(stderr);
    vsk_trset_free(refs);
    vsk_trset_free(reads);
    vsk_destroy(ctx);
    return 0;
}
