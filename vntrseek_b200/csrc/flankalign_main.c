/*
 * flankalign.exe -- drop-in replacement for the reference's flank aligner
 * (reference src/newflankalign/flankalign.c:214-823): same argv
 *   outdir xmldir inputfile MAXERRORS MAX_FLANK_CONSIDERED MAXTHREADS PATLEN_SIZE_ERR
 * same `.flanks` input (or `-` for stdin), same `<outdir>/<a>_<b>.map` files.
 * MAXTHREADS is validated and otherwise unused: the blocks are batched onto the GPU
 * instead of being forked.  GPU selection: environment variable VSK_DEVICE.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include "../../include/vsk.h"

static void quit(const char *msg) {
    printf("\n\nERROR!!!!:\n%s", msg);
    fflush(stdout);
    exit(1);
}

int main(int argc, char **argv) {
    int maxerrors = 0, maxflank = 1000, maxthreads = 1, paterr = 0;
    static uint8_t table[501501]; /* flankalign needs no distance table; the context wants one */
    const char *dst;
    char inpath[4096];
    vsk_ctx *ctx = NULL;
    vsk_flank_stats st;
    if (argc < 8)
        quit("\n\nFlankAlign - Please use ./flankalign  outdir xmldir inputfile MAXERRORS "
             "MAX_FLANK_CONSIDERED MAXTHREADS PATLEN_SIZE_ERR\n\n");
    maxerrors = atoi(argv[4]);
    maxflank = atoi(argv[5]);
    if (maxflank < 1) quit("\n\nFlankAlign - ? should be more than or equal to one!\n\n");
    maxthreads = atoi(argv[6]);
    if (maxthreads < 1) quit("\n\nFlankAlign - maxthreads should be more than or equal to one!\n\n");
    paterr = atoi(argv[7]);
    if (paterr < 1) quit("\n\nFlankAlign - PATLEN_SIZE_ERR must be at least 1%!\n\n");
    /* the input path is relative to the caller's cwd in the reference too, but it is opened
     * AFTER the chdir into outdir (flankalign.c:278-294); keep that order */
    if (chdir(argv[1])) {
        mkdir(argv[1], 0755);
        if (chdir(argv[1])) quit("\n\nFlankAlign - could not enter or create output dir!\n\n");
    }
    snprintf(inpath, sizeof inpath, "%s", argv[3]);
    dst = getenv("VSK_DISTANCE_FILE");
    if (dst) vsk_load_distance_file(dst, table);
    if (vsk_init(&ctx, getenv("VSK_DEVICE") ? atoi(getenv("VSK_DEVICE")) : 0, table) != VSK_OK)
        quit("\n\nFlankAlign - no usable CUDA device (this build has no CPU path)\n\n");
    if (inpath[0] != '-') {
        FILE *f = fopen(inpath, "r");
        if (!f) quit("\n\nFlankAlign - Could not open input file. Aborting!\n\n");
        fclose(f);
    }
    if (vsk_flankalign_run(ctx, inpath, maxerrors, maxflank, paterr, 0, &st) != VSK_OK) {
        printf("\n\nERROR!!!!:\n%s\n", vsk_last_error(ctx));
        exit(1);
    }
    vsk_destroy(ctx);
    return 0;
}
