/*
 * ref_harness.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Thin C shim that exposes the *unmodified* reference functions of the
 * PROCLU / flank-alignment hot path as plain C symbols so that the oracle
 * restatement (oracle/vsk_oracle.c) and the CUDA kernels can be checked
 * against the real thing.  The reference sources are compiled where they lie
 * under /root/reference (see oracle/Makefile); nothing is copied.
 *
 *   GetFixedProfileAP              src/psearch1.91/profile.h:2040
 *   GetFixedProfileAPNarrowband    src/psearch1.91/profile.h:1733
 *   LCS_multiple_word              src/psearch1.91/bitwise LCS multiple word.c:29
 *   Edit_Distance_..._NoEndPenaltySeq1 (psearch copy)
 *        src/psearch1.91/bitwise edit distance alignment multiple word no end penalty.c:29
 *   Edit_Distance_..._NoEndPenaltySeq1 (newflankalign copy)
 *        src/newflankalign/bitwise edit distance alignment multiple word no end penalty.c:29
 *   narrowbandUnitCostDistanceSameLengthBestScoreLastRowColumnLowMemUsesPointers
 *        src/newrefflankalign/narrowbandDistanceAlignment.c:533
 */
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* the five externs profile.h expects (normally psearch.c:44-46,1204 and
 * hclust.h:1008-1074 supply them) */
int  REFLEN             = 50;
int  MAXFLANKCONSIDERED = 1000000;
char OPTION             = 0;

void doCriticalErrorAndQuit( const char *format, ... ) {
    va_list argp;
    va_start( argp, format );
    vfprintf( stderr, format, argp );
    va_end( argp );
    exit( 1 );
}

char *GetReverse( char *s ) {
    int   n = (int) strlen( s ), i;
    char *b = (char *) malloc( n + 1 );
    for ( i = 0; i < n; i++ )
        b[i] = s[n - 1 - i];
    b[n] = 0;
    return b;
}

char *GetComplement( char *s ) {
    int   n = (int) strlen( s ), i;
    char *b = (char *) malloc( n + 1 );
    for ( i = 0; i < n; i++ ) {
        char c = s[i];
        switch ( c ) {
        case 'A': c = 'T'; break;
        case 'C': c = 'G'; break;
        case 'G': c = 'C'; break;
        case 'T': c = 'A'; break;
        case 'a': c = 't'; break;
        case 'c': c = 'g'; break;
        case 'g': c = 'c'; break;
        case 't': c = 'a'; break;
        }
        b[i] = c;
    }
    b[n] = 0;
    return b;
}

#include "easylife.h"
#include "profile.h"

int LCS_multiple_word( char *string1, char *string2, int n, int m );
int ED_psearch( char *stringN, char *stringM, int N, int M );
int ED_flankalign( char *stringN, char *stringM, int N, int M );
int narrowbandUnitCostDistanceSameLengthBestScoreLastRowColumnLowMemUsesPointers(
  char *seq1, char *seq2, int len1, int len2, int maxerr );

static unsigned char *g_sm = NULL;

int ref_init( const char *distfile ) {
    InitializeCompositions();
    g_sm = LoadDistanceTable( (char *) distfile );
    return g_sm ? 0 : -1;
}

/* mode 0 = GetFixedProfileAP, 1 = GetFixedProfileAPNarrowband(homology).
 * out = {distance, distanceweighted, weights, length}; returns similarity */
double ref_dp( const int *left, int l1, const int *top, int l2, int mode,
  int homology, int *out ) {
    PROFILE p1, p2;
    PAP *   ap;
    double  sim;
    memset( &p1, 0, sizeof p1 );
    memset( &p2, 0, sizeof p2 );
    p1.proflen = l1;
    p1.indices = (int *) left;
    p2.proflen = l2;
    p2.indices = (int *) top;
    ap = mode ? GetFixedProfileAPNarrowband( &p1, &p2, g_sm, homology )
              : GetFixedProfileAP( &p1, &p2, g_sm );
    out[0] = ap->distance;
    out[1] = ap->distanceweighted;
    out[2] = ap->weights;
    out[3] = ap->length;
    sim    = ap->similarity;
    free( ap );
    return sim;
}

int ref_lcs( const char *s1, const char *s2, int n, int m ) {
    return LCS_multiple_word( (char *) s1, (char *) s2, n, m );
}
int ref_ed_psearch( const char *sN, const char *sM, int N, int M ) {
    return ED_psearch( (char *) sN, (char *) sM, N, M );
}
int ref_ed_flankalign( const char *sN, const char *sM, int N, int M ) {
    return ED_flankalign( (char *) sN, (char *) sM, N, M );
}
int ref_nbdist( const char *s1, const char *s2, int l1, int l2, int maxerr ) {
    return narrowbandUnitCostDistanceSameLengthBestScoreLastRowColumnLowMemUsesPointers(
      (char *) s1, (char *) s2, l1, l2, maxerr );
}
int ref_nondashes( int id ) { return NONDASHES[id]; }
int ref_evaldistance( int r, int c ) { return EVALDISTANCE( r, c, g_sm ); }

/* n pairs in one call (the loop in C, one thread): ids as uint16 with offsets, pair k =
 * (left profile li[k], top profile ti[k]); out[4k..4k+3] as ref_dp */
void ref_dp_batch( const unsigned short *lids, const long long *loff, const unsigned short *tids,
  const long long *toff, const int *li, const int *ti, const unsigned char *mode, long long n,
  int homology, int *out ) {
    long long k;
    for ( k = 0; k < n; k++ ) {
        int  l1 = (int) ( loff[li[k] + 1] - loff[li[k]] ), l2 = (int) ( toff[ti[k] + 1] - toff[ti[k]] ), i;
        int *a = (int *) malloc( sizeof( int ) * ( l1 + l2 + 2 ) ), *b = a + l1 + 1;
        for ( i = 0; i < l1; i++ )
            a[i] = lids[loff[li[k]] + i];
        for ( i = 0; i < l2; i++ )
            b[i] = tids[toff[ti[k]] + i];
        ref_dp( a, l1, b, l2, mode[k], homology, out + 4 * k );
        free( a );
    }
}

/* GetBestProfileAP (profile.h:1586-1675): the cyclic aligner (no caller in the reference's programs).
 * left = p1 (slength rows), top = p2 (plength columns, every rotation of it is tried);
 * out = {distance, distanceweighted, weights, length, start}; returns similarity */
double ref_cyclic( const int *left, int l1, const int *top, int l2, int *out ) {
    PROFILE p1, p2;
    PAP *   ap;
    double  sim;
    memset( &p1, 0, sizeof p1 );
    memset( &p2, 0, sizeof p2 );
    p1.proflen = l1;
    p1.indices = (int *) left;
    p2.proflen = l2;
    p2.indices = (int *) top;
    ap         = GetBestProfileAP( &p1, &p2, g_sm );
    out[0]     = ap->distance;
    out[1]     = ap->distanceweighted;
    out[2]     = ap->weights;
    out[3]     = ap->length;
    out[4]     = ap->start;
    sim        = ap->similarity;
    FreePAP( ap );
    return sim;
}
