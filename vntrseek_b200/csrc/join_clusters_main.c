/*
 * join_clusters.exe -- drop-in replacement for the reference's cluster-file merge
 * (reference src/join_clusters/joinc.c:128-391): same argv (INPUTFOLDER outputfile), same
 * choice of input files (names longer than 6 characters ending in .clu, in strcmp order of
 * their paths), byte-identical output file, same summary line on stdout.  The reference also
 * prints one progress line per cluster ("<i> <percent>%"); this program does not.
 * All of the work is host code in libvsk.so (vsk_join_clusters_dir); no GPU is touched.
 */
#include <stdio.h>
#include <stdlib.h>

#include "../../include/vsk.h"

int main(int argc, char **argv) {
    vsk_join_stats st;
    char err[600];
    int rc;
    if (argc < 3) {
        printf("\njoin_clusters v1.00 (B200 build) - joins all .CLU files in a directory based "
               "on negative indices found in a cluster, assigns unique ids.");
        printf("\n\nUsage:\n\n%s INPUTFOLDER outputfile\n", argv[0]);
        exit(1);
    }
    rc = vsk_join_clusters_dir(argv[1], argv[2], &st, 1, err, sizeof err);
    if (rc != VSK_OK) {
        printf("\n\n%s\n", err);
        exit(1);
    }
    if (st.clusters_read == 0) return 0; /* joinc.c:184-185 */
    printf("\nTotal clusters read: %d, after join: %d, total refs(written) %lld "
           "(%lld),  total reads(written) %lld (%lld) maxdepth=%d\n",
           st.clusters_read, st.clusters_after_join, (long long)st.refs_read, (long long)st.refs_written,
           (long long)st.reads_read, (long long)st.reads_written, st.maxdepth);
    return 0;
}
