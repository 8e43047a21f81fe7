/*
 * vsk_oracle.c -- CPU restatement of the VNTRseek PROCLU / flank-alignment
 * hot-path arithmetic.  TEST INFRASTRUCTURE ONLY: imported by tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg as the *checker*;
 * never linked into, called from, or shipped with the product library.
 *
 * Parity status: PINNED.  Every function here is checked (tests/test_oracle_*.py)
 * against the unmodified reference functions compiled from /root/reference by
 * oracle/Makefile (`make ref` -> oracle/_ref/libref_harness.so) and against the
 * golden vectors under tests/golden/ that were generated from those same
 * reference functions (tests/golden/make_golden.py).  The reference ships no
 * golden vectors of its own for this path.
 *
 * Citations are relative to /root/reference/.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define VO_NIDS 1001
#define VO_BIG 100000000 /* out-of-band sentinel, src/psearch1.91/profile.h:1802 */

static uint8_t vo_dist[VO_NIDS][VO_NIDS]; /* symmetric expansion of eucledian.dst */
static uint8_t vo_nd[VO_NIDS];            /* NONDASHES, profile.h:837-886 */
static int     vo_ready = 0;

/* Expand the 501,501-byte upper-triangular table (profile.h:1035-1079,
 * EVALDISTANCE profile.h:1100-1103) into a full symmetric matrix and rebuild
 * NONDASHES by enumerating compositions in the reference's nested a,c,g,t
 * order (profile.h:840-881). */
void vo_init( const uint8_t *sm ) {
    int r, c, id = 0, a, g, t;
    for ( r = 0; r < VO_NIDS; r++ )
        for ( c = r; c < VO_NIDS; c++ ) {
            uint8_t v     = sm[1000 * r - ( r * r - r ) / 2 + c];
            vo_dist[r][c] = v;
            vo_dist[c][r] = v;
        }
    for ( a = 0; a <= 10; a++ )
        for ( c = 0; c <= 10; c++ )
            for ( g = 0; g <= 10; g++ )
                for ( t = 0; t <= 10; t++ )
                    if ( a + c + g + t <= 10 )
                        vo_nd[id++] = (uint8_t) ( a + c + g + t );
    vo_ready = 1;
}

int vo_nondashes( int id ) { return vo_nd[id]; }
int vo_distance( int r, int c ) { return vo_dist[r][c]; }

/* Band geometry of GetFixedProfileAPNarrowband (profile.h:1752-1770). Same
 * libm calls in the same order so the doubles agree bit for bit. */
void vo_band( int leftlen, int toplen, int homology, double *svector,
  int *bandradius_int ) {
    double minlen       = ( toplen <= leftlen ) ? toplen : leftlen;
    double virtdiagonal = sqrt( pow( minlen, 2 ) * 2 );
    double angle        = acos( homology / 100.0 );
    double topangle     = atan( (double) leftlen / (double) toplen );
    double midradius    = tan( angle ) * ( virtdiagonal / 2.0 );
    double bandradius   = midradius / sin( topangle );
    double startradius  = 6;
    *svector            = (double) toplen / (double) leftlen;
    *bandradius_int =
      (int) ( ( bandradius >= startradius ) ? bandradius : startradius ) + 1;
}

typedef struct {
    int32_t d;   /* min-cost distance                      */
    int32_t ds;  /* sum dashnum*cost along the chosen path */
    int32_t w;   /* sum dashnum along the chosen path      */
    int32_t len; /* path length                            */
} vo_cell;

/*
 * Global profile-to-profile alignment with the path-weighted re-score.
 *   mode 0: GetFixedProfileAP            (profile.h:2040-2227)
 *   mode 1: GetFixedProfileAPNarrowband  (profile.h:1733-2038), band as a mask
 * rows = `left` (p1, the reference TR), columns = `top` (p2, the read TR).
 * The reference fills an (L1+1)x(L2+1) matrix of (distance, direction) and
 * walks the directions back from the corner accumulating dashnum*cost and
 * dashnum (profile.h:2180-2210).  Because every cell has exactly one chosen
 * predecessor, those sums can be carried forward with the cell instead, which
 * needs two rows of state and no traceback.  Tie-break = min3switch
 * (profile.h:1098-1099): diagonal if sa<=f && sa<=e, else down if f<=e, else
 * right.  Out-of-band cells are VO_BIG (the reference writes 100000000
 * sentinels next to the band, profile.h:1850-1877, and never reads any other
 * out-of-band cell), so they can never win or tie against an in-band value.
 * out = {distance, distanceweighted, weights, length}.
 */
int vo_dp( const int32_t *left, int l1, const int32_t *top, int l2, int mode,
  int homology, int32_t *out ) {
    vo_cell *prev, *cur, *tmp;
    double   svector = 0;
    int      br = 0, row, col;
    if ( !vo_ready || l1 < 1 || l2 < 1 )
        return -1;
    if ( mode )
        vo_band( l1, l2, homology, &svector, &br );
    prev = (vo_cell *) malloc( sizeof( vo_cell ) * ( l2 + 1 ) );
    cur  = (vo_cell *) malloc( sizeof( vo_cell ) * ( l2 + 1 ) );
    memset( &prev[0], 0, sizeof( vo_cell ) );
    for ( col = 1; col <= l2; col++ ) { /* profile.h:2074-2079 / 1794-1809 */
        int t = top[col - 1], g = vo_dist[0][t], n = vo_nd[t];
        prev[col].d   = prev[col - 1].d + g;
        prev[col].ds  = prev[col - 1].ds + n * g;
        prev[col].w   = prev[col - 1].w + n;
        prev[col].len = prev[col - 1].len + 1;
        if ( mode && col > br )
            prev[col].d = VO_BIG;
    }
    for ( row = 1; row <= l1; row++ ) {
        int l = left[row - 1], gl = vo_dist[0][l], nl = vo_nd[l];
        int lstart = 1, rend = l2;
        if ( mode ) { /* profile.h:1835-1839 */
            int center = (int) ( row * svector );
            lstart     = center - br;
            if ( lstart < 1 )
                lstart = 1;
            rend = center + br;
            if ( rend > l2 )
                rend = l2;
        }
        /* column 0: profile.h:2084-2090 */
        cur[0].d   = prev[0].d + gl;
        cur[0].ds  = prev[0].ds + nl * gl;
        cur[0].w   = prev[0].w + nl;
        cur[0].len = prev[0].len + 1;
        for ( col = 1; col <= l2; col++ ) {
            int t = top[col - 1];
            if ( col < lstart || col > rend ) {
                cur[col].d  = VO_BIG;
                cur[col].ds = cur[col].w = cur[col].len = 0;
                continue;
            }
            {
                int gt = vo_dist[0][t], nt = vo_nd[t], dlt = vo_dist[l][t];
                int e  = cur[col - 1].d + gt;
                int f  = prev[col].d + gl;
                int sa = prev[col - 1].d + dlt;
                if ( sa <= f && sa <= e ) {
                    int dn       = nl + nt > 10 ? 10 : nl + nt;
                    cur[col].d   = sa;
                    cur[col].ds  = prev[col - 1].ds + dn * dlt;
                    cur[col].w   = prev[col - 1].w + dn;
                    cur[col].len = prev[col - 1].len + 1;
                } else if ( f <= e ) {
                    cur[col].d   = f;
                    cur[col].ds  = prev[col].ds + nl * gl;
                    cur[col].w   = prev[col].w + nl;
                    cur[col].len = prev[col].len + 1;
                } else {
                    cur[col].d   = e;
                    cur[col].ds  = cur[col - 1].ds + nt * gt;
                    cur[col].w   = cur[col - 1].w + nt;
                    cur[col].len = cur[col - 1].len + 1;
                }
            }
        }
        tmp  = prev;
        prev = cur;
        cur  = tmp;
    }
    out[0] = prev[l2].d;
    out[1] = prev[l2].ds;
    out[2] = prev[l2].w;
    out[3] = prev[l2].len;
    free( prev );
    free( cur );
    return 0;
}

/* similarity exactly as profile.h:2016-2017 / 2220-2221 */
double vo_similarity( int dsum, int sumw ) {
    return ( 1.0 - ( dsum ) / ( 255.0 * sumw ) );
}

/* number of DP cells the reference evaluates for a pair (work unit of GCUPS):
 * full = l1*l2; narrowband = sum over rows of (rend-lstart+1). */
long long vo_cells( int l1, int l2, int mode, int homology ) {
    double    sv;
    int       br, row;
    long long n = 0;
    if ( !mode )
        return (long long) l1 * l2;
    vo_band( l1, l2, homology, &sv, &br );
    for ( row = 1; row <= l1; row++ ) {
        int center = (int) ( row * sv ), a = center - br, b = center + br;
        if ( a < 1 )
            a = 1;
        if ( b > l2 )
            b = l2;
        if ( b >= a )
            n += b - a + 1;
    }
    return n;
}

static int vo_code( char c ) {
    switch ( c ) {
    case 'A': return 0;
    case 'C': return 1;
    case 'G': return 2;
    case 'T': return 3;
    case 'N': return 4;
    }
    return -1;
}

/*
 * LCS length of the two consensus strings as psearch computes it:
 * LCS_multiple_word (src/psearch1.91/bitwise LCS multiple word.c:29-310)
 * dispatching to LCS_single_word (bitwise LCS single word.c:20-199) when the
 * longer string has <= 63 characters.  The multi-word routine is an exact
 * bit-parallel LCS.  The single-word routine starts its column bitmask at
 * 0x10 instead of 0x2 (single word.c:59), so characters 60.. of the longer
 * string are shifted out of the 64-bit word: for longer-length 61..63 the
 * result is LCS(longer[0:60], shorter).  On n == m the "longer" string is
 * string1 (multiple word.c:76-86).  Returns -1 on a non-ACGTN character.
 */
int vo_lcs( const char *s1, const char *s2, int n, int m ) {
    const char *a, *b; /* a = bit-indexed (longer), b = iterated */
    int         N, M, W, i, j, cnt = 0;
    uint64_t *  peq, *v;
    if ( n >= m ) {
        a = s1; N = n; b = s2; M = m;
    } else {
        a = s2; N = m; b = s1; M = n;
    }
    for ( i = 0; i < N; i++ )
        if ( vo_code( a[i] ) < 0 )
            return -1;
    for ( i = 0; i < M; i++ )
        if ( vo_code( b[i] ) < 0 )
            return -1;
    if ( N <= 63 && N > 60 )
        N = 60; /* the 0x10 quirk */
    W   = ( N + 63 ) / 64;
    if ( W <= 0 || M <= 0 )
        return 0;
    peq = (uint64_t *) calloc( (size_t) 5 * W, sizeof( uint64_t ) );
    v   = (uint64_t *) malloc( sizeof( uint64_t ) * W );
    for ( i = 0; i < N; i++ )
        peq[vo_code( a[i] ) * W + i / 64] |= 1ULL << ( i % 64 );
    for ( j = 0; j < W; j++ )
        v[j] = ~0ULL;
    for ( i = 0; i < M; i++ ) {
        const uint64_t *mv    = peq + vo_code( b[i] ) * W;
        unsigned        carry = 0;
        for ( j = 0; j < W; j++ ) {
            uint64_t x = v[j] & mv[j];
            uint64_t s = v[j] + x;
            unsigned c = s < x;
            s += carry;
            c |= ( carry && s == 0 );
            v[j]  = s | ( v[j] & ~mv[j] );
            carry = c;
        }
    }
    for ( j = 0; j < W; j++ ) {
        uint64_t z = ~v[j];
        if ( j == W - 1 && ( N % 64 ) )
            z &= ( 1ULL << ( N % 64 ) ) - 1;
        cnt += __builtin_popcountll( z );
    }
    free( peq );
    free( v );
    return cnt;
}

/* threshold of hclust.h:537-545 */
int vo_lcs_min( int n, int m ) {
    int shortest = n < m ? n : m;
    int minlen   = (int) ( shortest * .85 + .5 );
    return minlen < shortest - 1 ? minlen : shortest - 1;
}

/*
 * Edit_Distance_multiple_word_NoEndPenaltySeq1: unit-cost edit distance of
 * pattern P (read flank, M chars) against text T (reference flank, N chars)
 * with a free end on the text side (minimum over the final row).
 *   variant 0 = psearch copy
 *     (src/psearch1.91/bitwise edit distance alignment multiple word no end penalty.c:29-392)
 *     stale carry (never cleared inside a row, :278-283) and no junk-bit mask
 *     on the final scan (:352-380)
 *   variant 1 = newflankalign copy
 *     (src/newflankalign/bitwise edit ... .c:29-425): carry fixed (:274-280),
 *     junk bits of the last word masked (:109-110,385-389)
 * Bit i of the vectors is column i of the DP (bit 0 = the empty prefix), so
 * N+1 bit positions are needed.  Word-level semantics are kept so that the
 * psearch copy's deviations for N >= 128 reproduce.  Returns -1 on a
 * non-ACGTN character (first in T, then in P as it is reached).
 */
int vo_ed( const char *T, const char *P, int N, int M, int variant ) {
    int       W, i, j, score, best;
    uint64_t *peq, *pd, *ps;
    if ( M == 0 )
        return 0;
    W   = ( N + 64 ) / 64;
    peq = (uint64_t *) calloc( (size_t) 5 * W, sizeof( uint64_t ) );
    pd  = (uint64_t *) calloc( W, sizeof( uint64_t ) );
    ps  = (uint64_t *) calloc( W, sizeof( uint64_t ) );
    for ( i = 1; i <= N; i++ ) {
        int c = vo_code( T[i - 1] );
        if ( c < 0 ) {
            free( peq ); free( pd ); free( ps );
            return -1;
        }
        peq[c * W + i / 64] |= 1ULL << ( i % 64 );
    }
    pd[0] = 1;
    ps[0] = 1;
    for ( i = 0; i < M; i++ ) {
        int             c = vo_code( P[i] );
        const uint64_t *mv;
        uint64_t        csum = 0, cv1 = 1, cv0 = 0;
        if ( c < 0 ) {
            free( peq ); free( pd ); free( ps );
            return -1;
        }
        mv = peq + c * W;
        for ( j = 0; j < W; j++ ) {
            uint64_t ov1 = cv1, ov0 = cv0;
            uint64_t mm  = mv[j] | pd[j];
            uint64_t ris = ~mm;
            uint64_t nmi = ris | ps[j];
            uint64_t x   = nmi ^ pd[j];
            uint64_t sum, sr, v0, v1, v0s, v1s;
            if ( variant == 0 ) {
                sum = nmi + ps[j] + csum;
                if ( sum < nmi )
                    csum = 1;
            } else {
                sum  = nmi + csum;
                csum = ( sum < nmi ) ? 1 : 0;
                sum += ps[j];
                csum |= ( sum < ps[j] ) ? 1 : 0;
            }
            sr    = sum & ris;
            v0    = sr ^ x;
            v1    = pd[j] | ( sr & ps[j] );
            cv0   = v0 >> 63;
            cv1   = v1 >> 63;
            v0s   = ( v0 << 1 ) + ov0;
            v1s   = ( v1 << 1 ) + ov1;
            pd[j] = mm & v1s;
            ps[j] = v1s | ( mm & v0s );
        }
    }
    score = best = M;
    for ( j = 0; j < W; j++ ) {
        uint64_t d = pd[j], ins = ~ps[j];
        if ( variant == 1 && j == W - 1 ) {
            int      junk = 63 + ( W - 1 ) * 64 - N;
            uint64_t mask = ~0ULL >> junk;
            d &= mask;
            ins &= mask;
        }
        for ( i = 0; i < 64; i++ ) {
            if ( i | j ) {
                score -= (int) ( d & 1 );
                score += (int) ( ins & 1 );
                if ( score < best )
                    best = score;
            }
            d >>= 1;
            ins >>= 1;
        }
    }
    free( peq ); free( pd ); free( ps );
    return best;
}

/* flank error thresholds, hclust.h:664-667 / flankalign.c:633-636 */
int vo_flank_maxerr( int readflanklen ) {
    int v = (int) ( 0.4 * readflanklen + .01 );
    return v < 8 ? v : 8;
}

/*
 * refflankalign's banded unit-cost DP
 * (src/newrefflankalign/narrowbandDistanceAlignment.c:533-776): -1 unless the
 * lengths are equal; cells with |row-col| <= maxerr, everything else 1000;
 * score = min over the last row (band cells plus the 1000 sentinels beside
 * them) and over the last-column cells of the rows whose band reaches it.
 */
int vo_nbdist( const char *s1, const char *s2, int len1, int len2, int maxerr ) {
    int *prev, *cur, *tmp, row, col, score;
    if ( len1 != len2 )
        return -1;
    prev = (int *) malloc( sizeof( int ) * ( len2 + 2 ) );
    cur  = (int *) malloc( sizeof( int ) * ( len2 + 2 ) );
    for ( col = 0; col <= len2; col++ )
        prev[col] = ( col <= maxerr ) ? col : 1000;
    score = 1000000;
    if ( maxerr >= len2 && prev[len2] < score ) /* row 0 reaches last column */
        score = prev[len2];
    for ( row = 1; row <= len1; row++ ) {
        for ( col = 0; col <= len2; col++ ) {
            int dcol = row > col ? row - col : col - row;
            if ( dcol > maxerr )
                cur[col] = 1000;
            else if ( col == 0 )
                cur[col] = row;
            else {
                int dg = prev[col - 1] + ( s1[row - 1] == s2[col - 1] ? 0 : 1 );
                int lf = cur[col - 1] + 1, up = prev[col] + 1;
                int mn = dg <= lf ? dg : lf;
                cur[col] = mn <= up ? mn : up;
            }
        }
        if ( row + maxerr >= len2 && cur[len2] < score )
            score = cur[len2];
        tmp = prev; prev = cur; cur = tmp;
    }
    /* last row: band cells and, when they exist, the 1000 sentinels */
    for ( col = 0; col <= len2; col++ ) {
        int dcol = len1 > col ? len1 - col : col - len1;
        int v    = dcol <= maxerr ? prev[col] : 1000;
        if ( dcol <= maxerr + 1 && v < score )
            score = v;
    }
    free( prev );
    free( cur );
    return score;
}

/* ---- batched forms of the functions above (arrays in, arrays out; pthreads over the pairs) ----
 * Same arithmetic, one call per pair; they exist so that the kernel-level parity tests can
 * check >= 10^6 pairs per kernel in seconds.                      */
#include <pthread.h>
#include <unistd.h>

typedef struct {
    void ( *fn )( int64_t k, void *arg );
    void *   arg;
    int64_t  n, chunk;
    int64_t *next;
} vo_pf;

static void *vo_pf_worker( void *p ) {
    vo_pf *w = (vo_pf *) p;
    for ( ;; ) {
        int64_t k0 = __atomic_fetch_add( w->next, w->chunk, __ATOMIC_RELAXED ), k;
        if ( k0 >= w->n )
            break;
        for ( k = k0; k < w->n && k < k0 + w->chunk; k++ )
            w->fn( k, w->arg );
    }
    return NULL;
}

static void vo_parallel_for( int64_t n, int64_t chunk, void ( *fn )( int64_t, void * ), void *arg ) {
    long      nc = sysconf( _SC_NPROCESSORS_ONLN );
    int       nt = (int) ( nc < 1 ? 1 : nc > 64 ? 64 : nc ), i;
    int64_t   next = 0;
    vo_pf     w    = { fn, arg, n, chunk, &next };
    pthread_t th[64];
    if ( n < 4 * chunk )
        nt = 1;
    for ( i = 1; i < nt; i++ )
        if ( pthread_create( &th[i], NULL, vo_pf_worker, &w ) != 0 ) {
            nt = i;
            break;
        }
    vo_pf_worker( &w );
    for ( i = 1; i < nt; i++ )
        pthread_join( th[i], NULL );
}

typedef struct {
    const uint16_t *lids, *tids;
    const int64_t * loff, *toff;
    const int32_t * li, *ti;
    const uint8_t * mode;
    int             homology;
    int32_t *       out;
} vo_dpb;

static void vo_dp_one( int64_t k, void *p ) {
    vo_dpb * q  = (vo_dpb *) p;
    int      l1 = (int) ( q->loff[q->li[k] + 1] - q->loff[q->li[k]] );
    int      l2 = (int) ( q->toff[q->ti[k] + 1] - q->toff[q->ti[k]] ), i;
    int32_t *a = (int32_t *) malloc( sizeof( int32_t ) * ( l1 + l2 + 2 ) ), *b = a + l1 + 1;
    for ( i = 0; i < l1; i++ )
        a[i] = q->lids[q->loff[q->li[k]] + i];
    for ( i = 0; i < l2; i++ )
        b[i] = q->tids[q->toff[q->ti[k]] + i];
    if ( vo_dp( a, l1, b, l2, q->mode[k], q->homology, q->out + 4 * k ) != 0 )
        q->out[4 * k] = q->out[4 * k + 1] = q->out[4 * k + 2] = q->out[4 * k + 3] = -1;
    free( a );
}

void vo_dp_batch( const uint16_t *lids, const int64_t *loff, const uint16_t *tids,
  const int64_t *toff, const int32_t *li, const int32_t *ti, const uint8_t *mode, int64_t n,
  int homology, int32_t *out ) {
    vo_dpb q = { lids, tids, loff, toff, li, ti, mode, homology, out };
    vo_parallel_for( n, 64, vo_dp_one, &q );
}

typedef struct {
    const char *   a, *b;
    const int64_t *aoff, *boff;
    const int32_t *ia, *ib, *alen, *blen;
    int            par;
    int32_t *      out;
} vo_strb;

static void vo_lcs_one( int64_t k, void *p ) {
    vo_strb *q = (vo_strb *) p;
    q->out[k]  = vo_lcs( q->a + q->aoff[q->ia[k]], q->b + q->boff[q->ib[k]],
      (int) ( q->aoff[q->ia[k] + 1] - q->aoff[q->ia[k]] ), (int) ( q->boff[q->ib[k] + 1] - q->boff[q->ib[k]] ) );
}

void vo_lcs_batch( const char *a, const int64_t *aoff, const char *b, const int64_t *boff,
  const int32_t *ia, const int32_t *ib, int64_t n, int32_t *out ) {
    vo_strb q = { a, b, aoff, boff, ia, ib, NULL, NULL, 0, out };
    vo_parallel_for( n, 256, vo_lcs_one, &q );
}

static void vo_ed_one( int64_t k, void *p ) {
    vo_strb *q = (vo_strb *) p;
    q->out[k]  = vo_ed( q->a + q->aoff[k], q->b + q->boff[k], q->alen[k], q->blen[k], q->par );
}

void vo_ed_batch( const char *text, const int64_t *toff, const int32_t *tlen, const char *pat,
  const int64_t *poff, const int32_t *plen, int64_t n, int variant, int32_t *out ) {
    vo_strb q = { text, pat, toff, poff, NULL, NULL, tlen, plen, variant, out };
    vo_parallel_for( n, 256, vo_ed_one, &q );
}

static void vo_nb_one( int64_t k, void *p ) {
    vo_strb *q = (vo_strb *) p;
    q->out[k]  = vo_nbdist( q->a + q->aoff[k], q->b + q->boff[k], (int) ( q->aoff[k + 1] - q->aoff[k] ),
      (int) ( q->boff[k + 1] - q->boff[k] ), q->par );
}

void vo_nbdist_batch( const char *a, const int64_t *aoff, const char *b, const int64_t *boff, int64_t n,
  int maxerr, int32_t *out ) {
    vo_strb q = { a, b, aoff, boff, NULL, NULL, NULL, NULL, maxerr, out };
    vo_parallel_for( n, 256, vo_nb_one, &q );
}

/* ---- cyclic alignment: GetBestProfileAP (profile.h:1586-1675) -------------------------------
 * Dead code in the reference's programs, but it is the aligner VNTRseek's description means by
 * "rotation-aware": the left profile (l1 rows) is aligned globally against every rotation of the
 * top profile (l2 columns), i.e. against the windows [m, m+l2] of the doubled top profile, and the
 * rotation with the smallest weighted distance per weight wins (strict <, in the order the
 * reference visits them: m = 0 first, then midpoints depth-first, left half before right half).
 * Rotation m is aligned only between the paths found for the rotations `lo` < m < `hi` that
 * enclose it (GetBoundedProfileAP, profile.h:1281-1585): row r may use columns
 * max(left-most column of lo's path in row r, m) .. min(right-most column of hi's path, m+l2);
 * the cell left of a cut-off row and the cells right of a row's end up to the next row's end are
 * PA_VLN.  Recurrence and tie-break as vo_dp; the weighted sums come from the traceback.
 * out = {distance, distanceweighted, weights, length, start}.  Returns 0, or -1 on bad input.   */
#define VO_VLN 600000000
typedef struct {
    int left, right;
} vo_lim;

typedef struct {
    int distance, dsum, sumw, length;
} vo_pap;

/* one bounded alignment; writes the limits of its own path to `own` (PAPToLimit, profile.h:1231-1279) */
static vo_pap vo_bounded( const int32_t *left, const int32_t *top2, int l1, int start, int end,
  const vo_lim *mins, const vo_lim *maxs, int *D, uint8_t *dir, vo_lim *own ) {
    const int width = end - start + 1;
    int       r, c, i, max, min;
    vo_pap    ap;
#define DD( r, c ) D[(size_t) ( r ) * width + ( ( c ) - start )]
#define RR( r, c ) dir[(size_t) ( r ) * width + ( ( c ) - start )]
    DD( 0, start ) = 0;
    RR( 0, start ) = 0;
    max            = maxs[0].right < end ? maxs[0].right : end;
    for ( c = start + 1; c <= max; c++ ) {
        DD( 0, c ) = vo_dist[0][top2[c - 1]] + DD( 0, c - 1 );
        RR( 0, c ) = 3;
    }
    {
        int nxt = maxs[1].right < end ? maxs[1].right : end;
        for ( i = nxt; i > max; i-- )
            DD( 0, i ) = VO_VLN;
    }
    for ( r = 1; r <= l1; r++ ) { /* start column, while the left bound does not cut it off */
        DD( r, start ) = vo_dist[0][left[r - 1]] + DD( r - 1, start );
        RR( r, start ) = 2;
        if ( r < l1 ) {
            min = mins[r + 1].left > start ? mins[r + 1].left : start;
            if ( min > start )
                break;
        }
    }
    for ( r = 1; r <= l1; r++ ) {
        const int l = left[r - 1], gl = vo_dist[0][l];
        min = mins[r].left > start ? mins[r].left : start;
        max = maxs[r].right < end ? maxs[r].right : end;
        if ( min > start ) {
            DD( r, min - 1 ) = VO_VLN;
            c                = min;
        } else {
            c = min + 1;
        }
        for ( ; c <= max; c++ ) {
            const int t = top2[c - 1];
            const int e = DD( r, c - 1 ) + vo_dist[0][t], f = DD( r - 1, c ) + gl, sa = DD( r - 1, c - 1 ) + vo_dist[l][t];
            if ( sa <= f && sa <= e ) {
                DD( r, c ) = sa;
                RR( r, c ) = 1;
            } else if ( f <= e ) {
                DD( r, c ) = f;
                RR( r, c ) = 2;
            } else {
                DD( r, c ) = e;
                RR( r, c ) = 3;
            }
        }
        if ( r < l1 ) {
            int nxt = maxs[r + 1].right < end ? maxs[r + 1].right : end;
            for ( i = nxt; i > max; i-- )
                DD( r, i ) = VO_VLN;
        }
    }
    /* traceback: length, weighted sums (profile.h:1429-1497), and the path's limits */
    ap.distance = DD( l1, end );
    ap.length = ap.dsum = ap.sumw = 0;
    r = l1;
    c = end;
    for ( i = 0; i <= l1; i++ ) {
        own[i].left  = 0x7fffffff;
        own[i].right = -1;
    }
    for ( ;; ) {
        const int d = RR( r, c );
        if ( c < own[r].left )
            own[r].left = c;
        if ( c > own[r].right )
            own[r].right = c;
        if ( d == 1 ) {
            const int l = left[r - 1], t = top2[c - 1];
            int       dn = vo_nd[l] + vo_nd[t];
            if ( dn > 10 )
                dn = 10;
            ap.dsum += dn * vo_dist[l][t];
            ap.sumw += dn;
            r--;
            c--;
        } else if ( d == 3 ) {
            const int t = top2[c - 1];
            ap.dsum += vo_nd[t] * vo_dist[0][t];
            ap.sumw += vo_nd[t];
            c--;
        } else if ( d == 2 ) {
            const int l = left[r - 1];
            ap.dsum += vo_nd[l] * vo_dist[l][0];
            ap.sumw += vo_nd[l];
            r--;
        } else
            break;
        ap.length++;
    }
#undef DD
#undef RR
    return ap;
}

int vo_cyclic( const int32_t *left, int l1, const int32_t *top, int l2, int32_t *out ) {
    int32_t *top2;
    int *    D;
    uint8_t *dir;
    vo_lim **lim; /* limits of the rotations on the current path of the recursion, by rotation */
    int      stackl[64], stackr[64], sp = -1, m, i, best_start = 0;
    vo_pap   best, ap;
    if ( !vo_ready || l1 < 1 || l2 < 1 )
        return -1;
    top2 = (int32_t *) malloc( sizeof( int32_t ) * 2 * l2 );
    for ( i = 0; i < 2 * l2; i++ )
        top2[i] = top[i % l2];
    D   = (int *) malloc( sizeof( int ) * (size_t) ( l1 + 1 ) * ( l2 + 1 ) );
    dir = (uint8_t *) malloc( (size_t) ( l1 + 1 ) * ( l2 + 1 ) );
    lim = (vo_lim **) calloc( (size_t) l2 + 1, sizeof( vo_lim * ) );
    lim[0]  = (vo_lim *) malloc( sizeof( vo_lim ) * ( l1 + 2 ) );
    lim[l2] = (vo_lim *) malloc( sizeof( vo_lim ) * ( l1 + 2 ) );
    for ( i = 0; i <= l1 + 1; i++ ) { /* init_rectangular_limits, profile.h:1219-1229 */
        lim[0][i].left = lim[0][i].right = 0;
        lim[l2][i].left = lim[l2][i].right = l2;
    }
    {
        vo_lim *own = (vo_lim *) malloc( sizeof( vo_lim ) * ( l1 + 2 ) );
        best        = vo_bounded( left, top2, l1, 0, l2, lim[0], lim[l2], D, dir, own );
        own[l1 + 1] = own[l1];
        free( lim[0] );
        lim[0] = own;
        for ( i = 0; i <= l1 + 1; i++ ) { /* the same path, one period to the right */
            lim[l2][i].left  = lim[0][i].left + l2;
            lim[l2][i].right = lim[0][i].right + l2;
        }
    }
    stackl[++sp] = 0;
    stackr[sp]   = l2;
    while ( sp >= 0 ) {
        const int lo = stackl[sp], hi = stackr[sp];
        sp--;
        if ( hi - lo > 1 ) {
            vo_lim *own = (vo_lim *) malloc( sizeof( vo_lim ) * ( l1 + 2 ) );
            m           = ( lo + hi ) / 2;
            ap          = vo_bounded( left, top2, l1, m, m + l2, lim[lo], lim[hi], D, dir, own );
            own[l1 + 1] = own[l1];
            lim[m]      = own;
            if ( ( ap.dsum / ( 255.0 * ap.sumw ) ) < ( best.dsum / ( 255.0 * best.sumw ) ) ) {
                best       = ap;
                best_start = m;
            }
            stackl[++sp] = m; /* the left half is popped first */
            stackr[sp]   = hi;
            stackl[++sp] = lo;
            stackr[sp]   = m;
        } else { /* removelimits(left), profile.h:1649 */
            free( lim[lo] );
            lim[lo] = NULL;
        }
    }
    free( lim[l2] );
    free( lim );
    free( D );
    free( dir );
    free( top2 );
    out[0] = best.distance;
    out[1] = best.dsum;
    out[2] = best.sumw;
    out[3] = best.length;
    out[4] = best_start;
    return 0;
}
