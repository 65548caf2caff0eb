/* tables.cpp -- host-side constant tables (Gauss filters, descriptor weights, sizes).
 *
 * These feed every kernel and must be bit-identical to what the reference uploads, so the
 * float/double mix of each expression is kept (reference gauss_filter.cu:127-257 init_filter,
 * :302-328 spans, :342-372 computeBlurTable; sift_constants.cu:22-53; sift_conf.cu:275-278;
 * popsift.cpp:89-120).  Compile with -ffp-contract=off.
 */
#include "common.h"
#include <cmath>
#include <algorithm>

namespace psift {

static int span_for( int gauss_mode, float sigma )
{
    /* one-sided width INCLUDING the centre; capped at 31 so that taps[span] is a zero slot */
    int s;
    switch( gauss_mode ) {
    case PSIFT_GAUSS_OPENCV_COMPUTE:                           /* gauss_filter.cu:322-328 */
        s = ( ( int( roundf( 2.0f * 4.0f * sigma + 1.0f ) ) | 1 ) >> 1 ) + 1;
        break;
    case PSIFT_GAUSS_VLFEAT_RELATIVE:                          /* gauss_filter.cu:311-319 */
        s = std::min<int>( ceilf( 4.0f * sigma ) + 1, TAP_ALIGN - 1 );
        if( ( s & 1 ) == 0 ) s += 1;
        break;
    case PSIFT_GAUSS_FIXED9:  s = 5; break;
    case PSIFT_GAUSS_FIXED15: s = 8; break;
    default:                                                   /* VLFeat: gauss_filter.cu:302-308 */
        s = ceilf( 4.0f * sigma ) + 1;
        break;
    }
    return std::min<int>( s, TAP_ALIGN - 1 );
}

/* taps = exp(-x^2/(2 sigma^2)) evaluated in double, stored as float, summed in double with
 * the float taps, then each float tap divided by the double sum (gauss_filter.cu:352-368). */
static void fill_taps( int gauss_mode, float sigma, Taps& t )
{
    const int spn = span_for( gauss_mode, sigma );
    t.span = spn;
    double sum = 1.0;
    t.g[0] = 1.0f;
    for( int x = 1; x < spn; x++ ) {
        const double q   = double( x ) / sigma;
        const float  val = (float)exp( -0.5 * pow( q, 2.0 ) );
        t.g[x] = val;
        sum += 2.0f * val;
    }
    for( int x = 0; x < spn; x++ )         t.g[x] = (float)( t.g[x] / sum );
    for( int x = spn; x < TAP_ALIGN; x++ ) t.g[x] = 0.0f;
}

bool build_gauss_tables( const psift_config& cfg, GaussTables& out, const char** err )
{
    const float sigma0 = cfg.sigma;
    const int   levels = cfg.levels;
    if( sigma0 > 2.0 )        { *err = "sigma > 2.0 is not supported (reference gauss_filter.cu:131)"; return false; }
    if( levels + 3 > MAX_LVL ) { *err = "too many levels (reference gauss_filter.cu:138: at most 12 filter stages)"; return false; }
    if( levels < 1 )          { *err = "levels must be positive"; return false; }

    out.stages = levels + 3;
    const bool  has_blur     = cfg.assume_initial_blur != 0;
    const float initial_blur = has_blur ? cfg.initial_blur * powf( 2.0f, cfg.upscale_factor ) : 0.0f;

    for( int l = 0; l < MAX_LVL; l++ ) { out.inc_sigma[l] = out.abs_o0_sigma[l] = out.abs_oN_sigma[l] = 0.0f; }
    for( int o = 0; o < MAX_OCT; o++ ) { out.dd_sigma[o] = 0.0f; }

    /* inc: gauss_filter.cu:177-186 */
    out.inc_sigma[0] = has_blur ? sqrtf( fabsf( sigma0 * sigma0 - initial_blur * initial_blur ) ) : sigma0;
    for( int l = 1; l < out.stages; l++ ) {
        const float sigmaP = sigma0 * powf( 2.0f, (float)( l - 1 ) / (float)levels );
        const float sigmaS = sigma0 * powf( 2.0f, (float)( l     ) / (float)levels );
        out.inc_sigma[l] = sqrtf( sigmaS * sigmaS - sigmaP * sigmaP );
    }
    /* abs_o0: gauss_filter.cu:194-199 */
    for( int l = 0; l < out.stages; l++ ) {
        const float sigmaS = sigma0 * powf( 2.0f, (float)l / (float)levels );
        out.abs_o0_sigma[l] = sqrtf( fabsf( sigmaS * sigmaS - initial_blur * initial_blur ) );
    }
    /* abs_oN: gauss_filter.cu:209-216 */
    for( int l = 1; l < out.stages; l++ ) {
        const float sigmaS = sigma0 * powf( 2.0f, (float)l / (float)levels );
        out.abs_oN_sigma[l] = sqrtf( sigmaS * sigmaS - sigma0 * sigma0 );
    }
    /* dd: gauss_filter.cu:227-237 */
    for( int o = 0; o < MAX_OCT; o++ ) {
        const float oct_sigma = scalbnf( sigma0, o );
        const float b = sqrtf( fabsf( oct_sigma * oct_sigma - initial_blur * initial_blur ) );
        out.dd_sigma[o] = scalbnf( b, -o );
    }
    /* the reference computes a table for every slot, including the unused sigma==0 ones */
    for( int l = 0; l < MAX_LVL; l++ ) {
        fill_taps( cfg.gauss_mode, out.inc_sigma[l],    out.inc[l] );
        fill_taps( cfg.gauss_mode, out.abs_o0_sigma[l], out.abs_o0[l] );
        fill_taps( cfg.gauss_mode, out.abs_oN_sigma[l], out.abs_oN[l] );
    }
    for( int o = 0; o < MAX_OCT; o++ ) fill_taps( cfg.gauss_mode, out.dd_sigma[o], out.dd[o] );
    return true;
}

void build_consts( const psift_config& cfg, Consts& c, DescTables& t )
{
    c.sigma0           = cfg.sigma;
    c.sigma_k          = powf( 2.0f, 1.0f / cfg.levels );                 /* sift_constants.cu:27 */
    c.edge_limit       = cfg.edge_limit;
    c.threshold        = ( cfg.threshold * 0.5f * 255.0f / cfg.levels );   /* sift_conf.cu:275-278 */
    c.max_extrema      = cfg.max_extrema;
    c.max_orientations = cfg.max_extrema + cfg.max_extrema / 4;           /* sift_constants.cu:31 */
    c.norm_multi       = cfg.norm_multi;
    c.sift_mode        = cfg.sift_mode;
    c.desc_mode        = cfg.desc_mode;
    c.norm_mode        = cfg.norm_mode;
    c.up_fac           = (int)cfg.upscale_factor;                          /* sift_pyramid.cu:283,297 float -> int */
    c.grid_size        = cfg.filter_grid_size;
    c.deterministic    = cfg.deterministic_order != 0;

    const float dn_step = 1.0f / 8.0f;                                     /* sift_constants.cu:33-41 */
    const float dn_base = 0.5f * dn_step - 20.0f * dn_step;
    for( int y = 0; y < 40; y++ )
        for( int x = 0; x < 40; x++ ) {
            const float dnx = dn_base + x * dn_step;
            const float dny = dn_base + y * dn_step;
            t.desc_gauss[y][x] = expf( -scalbnf( dnx * dnx + dny * dny, -3 ) );
        }
    for( int i = 0; i < 16; i++ ) {                                        /* sift_constants.cu:43-46 */
        const float nx = -1.0f + 1.0f / 16.0f + i * 1.0f / 8.0f;
        t.desc_tile[i] = 1.0f - fabsf( nx );
    }
}

void octave0_size( const psift_config& cfg, int w, int h, int& w0, int& h0 )
{
    const float scaleFactor = 1.0f / powf( 2.0f, -cfg.upscale_factor );    /* popsift.cpp:97-98 */
    w0 = (int)ceilf( w * scaleFactor );
    h0 = (int)ceilf( h * scaleFactor );
}

int octave_count( const psift_config& cfg, int w, int h )
{
    if( cfg.octaves >= 0 ) return std::max( cfg.octaves, 1 );
    const float scaleFactor = 1.0f / powf( 2.0f, -cfg.upscale_factor );
    /* NB: adds the scale FACTOR, not the exponent (popsift.cpp:107-110) */
    int oct = std::max( int( floor( logf( (float)std::min( w, h ) ) / logf( 2.0f ) ) - 3.0f + scaleFactor ), 1 );
    return std::min( oct, MAX_OCT );
}

} // namespace psift
