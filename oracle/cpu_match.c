/*
 * cpu_match.c -- TEST INFRASTRUCTURE (checker only; the product never calls or links this).
 *
 * CPU restatement of the reference's brute-force matcher, FeaturesDev::match ->
 * compute_distance (reference src/popsift/features.cu:163-225):
 *
 *   l2_in_t0 (features.cu:163-182): lane k of a 32-lane warp takes elements 4k..4k+3 of both descriptors,
 *       m = l - r;  res = m.x*m.x + m.y*m.y + m.z*m.z + m.w*m.w
 *   which the reference build contracts to  fma(m.w,m.w, fma(m.z,m.z, fma(m.x,m.x, m.y*m.y)))  (SASS of
 *   oracle/_ref/obj/features.o: FMUL on the .y difference, then three FFMA with .x, .z, .w), followed by
 *       res += shuffle_down(res, 16); ... 8; 4; 2; 1      -> lane 0 holds the sum
 *   compute_distance (features.cu:184-225): for i in 0..r_len-1 in order, strict `<` updates of (best, second);
 *       accept = best / second < 0.8f.
 *
 * Parity: pinned against the reference's own compute_distance kernel run on a B200 (tests/golden/ref/match_*.bin,
 * produced by oracle/ref_driver --match-desc, tools/run_ref_golden.py).
 *
 * usage: cpu_match L.f32 R.f32 OUT.bin     (raw float32 files of n*128 values; OUT: l_len x {int idx1, idx2, accept})
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

static float* read_desc( const char* fn, int* n )
{
    FILE* f = fopen( fn, "rb" );
    if( !f ) { fprintf( stderr, "cpu_match: cannot open %s\n", fn ); exit( 2 ); }
    fseek( f, 0, SEEK_END ); long sz = ftell( f ); fseek( f, 0, SEEK_SET );
    *n = (int)( sz / ( 128 * sizeof(float) ) );
    float* p = (float*)malloc( sz ? sz : 1 );
    if( fread( p, 1, sz, f ) != (size_t)sz ) { fprintf( stderr, "cpu_match: short read\n" ); exit( 2 ); }
    fclose( f );
    return p;
}

/* features.cu:163-182, lane 0's value */
static float l2_in_t0( const float* l, const float* r )
{
    volatile float lane[32];                     /* volatile: no re-association, no contraction beyond the explicit fmaf */
    for( int k = 0; k < 32; k++ ) {
        const float mx = l[4 * k + 0] - r[4 * k + 0], my = l[4 * k + 1] - r[4 * k + 1];
        const float mz = l[4 * k + 2] - r[4 * k + 2], mw = l[4 * k + 3] - r[4 * k + 3];
        float res = my * my;
        res = fmaf( mx, mx, res );
        res = fmaf( mz, mz, res );
        res = fmaf( mw, mw, res );
        lane[k] = res;
    }
    for( int s = 16; s >= 1; s >>= 1 )           /* shuffle_down by s: only lanes < s still feed lane 0 */
        for( int k = 0; k < s; k++ ) lane[k] = lane[k] + lane[k + s];
    return lane[0];
}

int main( int argc, char** argv )
{
    if( argc != 4 ) { fprintf( stderr, "usage: cpu_match L.f32 R.f32 OUT.bin\n" ); return 2; }
    int nl, nr;
    float* l = read_desc( argv[1], &nl );
    float* r = read_desc( argv[2], &nr );
    FILE* out = fopen( argv[3], "wb" );
    if( !out ) { fprintf( stderr, "cpu_match: cannot write %s\n", argv[3] ); return 2; }
    for( int idx = 0; idx < nl; idx++ ) {
        float v1 = INFINITY, v2 = INFINITY;      /* features.cu:190-193 */
        int   i1 = 0, i2 = 0;
        for( int i = 0; i < nr; i++ ) {
            const float res = l2_in_t0( l + 128 * (size_t)idx, r + 128 * (size_t)i );
            if( res < v1 ) { v2 = v1; i2 = i1; v1 = res; i1 = i; }          /* features.cu:203-216 */
            else if( res < v2 ) { v2 = res; i2 = i; }
        }
        int rec[3] = { i1, i2, ( v1 / v2 < 0.8f ) ? 1 : 0 };                /* features.cu:222-223 */
        fwrite( rec, sizeof(int), 3, out );
    }
    fclose( out );
    free( l ); free( r );
    return 0;
}
