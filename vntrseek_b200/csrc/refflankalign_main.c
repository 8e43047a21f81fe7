/*
 * refflankalign.exe -- drop-in replacement for the reference's reference-vs-reference flank
 * aligner (reference src/newrefflankalign/refflankalign.c:327-1074): same argv
 *   outdir xmldir inputfile MAXERRORS MAXTHREADS PATLEN_SIZE_ERR REFLEN
 * same `.flanks` input, same `<xmldir>/<a>_<b>.xml` and `<outdir>/representatives.txt`.
 * MAXTHREADS is validated and otherwise unused.  GPU selection: VSK_DEVICE.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/vsk.h"

static void quit(const char *msg) {
    printf("\n\nERROR!!!!:\n%s", msg);
    fflush(stdout);
    exit(1);
}

int main(int argc, char **argv) {
    static uint8_t table[501501];
    vsk_ctx *ctx = NULL;
    int maxerrors, maxthreads, paterr, reflen;
    if (argc < 8)
        quit("\n\nRefFlankAlign - Please use ./Refflankalign  outdir xmldir inputfile MAXERRORS MAXTHREADS "
             "PATLEN_SIZE_ERR REFLEN\n\n");
    maxerrors = atoi(argv[4]);
    maxthreads = atoi(argv[5]);
    if (maxthreads < 1) quit("\n\nRefFlankAlign - maxthreads should be more than or equal to one!\n\n");
    paterr = atoi(argv[6]);
    if (paterr < 1) quit("\n\nRefFlankAlign - PATLEN_SIZE_ERR must be at least 1%!\n\n");
    reflen = atoi(argv[7]);
    if (reflen < 1) quit("\n\nRefFlankAlign - REFLEN must be at least 1%!\n\n");
    if (vsk_init(&ctx, getenv("VSK_DEVICE") ? atoi(getenv("VSK_DEVICE")) : 0, table) != VSK_OK)
        quit("\n\nRefFlankAlign - no usable CUDA device (this build has no CPU path)\n\n");
    if (vsk_refflankalign_run(ctx, argv[1], argv[2], argv[3], maxerrors, paterr, reflen, NULL, NULL) != VSK_OK) {
        printf("\n\nERROR!!!!:\n%s\n", vsk_last_error(ctx));
        exit(1);
    }
    vsk_destroy(ctx);
    return 0;
}
