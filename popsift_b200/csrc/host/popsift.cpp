/* PopSift / SiftJob -- host-side job queue and multi-GPU dispatcher.
 *
 * Reference: src/popsift/popsift.cpp (constructor :16-38, configure :63-87, private_init :89-120,
 * uninit :122-137, enqueue :139-169, uploadImages :171-180, extractDownloadLoop :182-213,
 * SiftJob :238-313).  The reference runs two boost threads (upload, extract) around ONE pyramid
 * on device 0.  Here there is one worker thread per GPU; each owns a psift context and keeps
 * several frames in flight on it.  The image stream is sharded across GPUs by whichever worker
 * dequeues a job first: no collective, no peer traffic, results come back through the job's
 * own future.  All CUDA work goes through the C ABI (psift.h).
 */
#include <popsift/popsift.h>
#if defined( __x86_64__ ) && defined( __SSE2__ )
#include <emmintrin.h>
#endif
#include <cstdint>
#include <psift.h>

#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <iostream>
#include <mutex>
#include <thread>
#include <sys/mman.h>
#include <atomic>

using popsift::FeaturesHost;
using popsift::FeaturesBase;

namespace popsift {
/* host/debug_dump.cpp: Config::LogMode::All */
void debug_dump_all( void* ctx, const Config& conf, int n_octaves, int n_levels, FeaturesHost* features, const char* basename );
}

/* ------------------------------------------------------------------ SiftJob */

/* The job's private copy of the image.  A fresh 2 MB malloc costs ~500 page faults when it is first written - that, not
 * the copy, bounds enqueue() at about a thousand 1080p frames per second - so large copies are 2 MB-aligned and
 * advised as huge pages (one fault per 2 MB where transparent huge pages are available; harmless where they are not):
 * 1 140 -> 1 435 frames/s through enqueue()/get() on one B200 (tests/cpp/api_demo, 1 500 frames enqueued up front). */
static unsigned char* job_alloc( size_t n )
{
    const size_t huge = (size_t)2 << 20;
    if( n >= huge / 2 ) {
        void* p = 0;
        const size_t bytes = ( n + huge - 1 ) / huge * huge;
        if( posix_memalign( &p, huge, bytes ) == 0 ) {
            madvise( p, bytes, MADV_HUGEPAGE );
            return (unsigned char*)p;
        }
    }
    return (unsigned char*)malloc( n ? n : 1 );
}

/* Page-locked blocks for the jobs' private copies, RECYCLED: enqueue() copies the caller's image straight into one (the
 * one copy its contract needs, SiftJob::SiftJob popsift.cpp:238-253), the worker uploads from it without a second,
 * staging copy (psift_submit_pinned_u8), and the block returns to the pool as soon as the frame has been retired.  A block
 * is page-locked once, when the pool grows; the pool holds at most POOL_BYTES in all, jobs beyond that (a caller that
 * enqueues thousands of frames up front) get ordinary memory and the staging path.  Size classes of 2 MB. */
namespace {
struct PinnedPool {
    static const size_t GRAIN = (size_t)2 << 20, POOL_BYTES = (size_t)768 << 20;
    std::mutex mu;
    struct E { size_t k; void* p; };
    std::vector<E> idle;
    size_t         total = 0;          /* bytes page-locked so far (idle + in use) */
    static size_t klass( size_t n ) { return ( std::max<size_t>( n, 1 ) + GRAIN - 1 ) / GRAIN * GRAIN; }
    void* get( size_t n )
    {
        const size_t k = klass( n );
        {
            std::lock_guard<std::mutex> lk( mu );
            for( size_t i = 0; i < idle.size(); i++ ) if( idle[i].k == k ) { void* p = idle[i].p; idle[i] = idle.back(); idle.pop_back(); return p; }
            if( total + k > POOL_BYTES ) return 0;
            total += k;
        }
        void* p = 0;
        if( psift_host_alloc_pinned( &p, k ) != PSIFT_OK ) { std::lock_guard<std::mutex> lk( mu ); total -= k; return 0; }
        return p;
    }
    void put( void* p, size_t n ) { std::lock_guard<std::mutex> lk( mu ); idle.push_back( E{ klass( n ), p } ); }
};
PinnedPool& pinned_pool() { static PinnedPool* p = new PinnedPool; return *p; }     /* never torn down: jobs may outlive every PopSift */
}

/* The private copy of a frame: read once by the DMA engine and never by this CPU again, so it is written with
 * streaming stores where the machine has them - no read-for-ownership of the destination, no eviction of the caller's
 * working set (with eight GPUs fed from one host the copies of all workers add up to the memory system's limit). */
static void copy_image( unsigned char* dst, const void* src_v, size_t n )
{
#if defined( __x86_64__ ) && defined( __SSE2__ )
    const unsigned char* src = (const unsigned char*)src_v;
    if( n >= ( 256u << 10 ) && ( ( (uintptr_t)dst & 15u ) == 0 ) ) {
        size_t i = 0;
        for( ; i + 64 <= n; i += 64 ) {
            const __m128i a = _mm_loadu_si128( (const __m128i*)( src + i ) ),      b = _mm_loadu_si128( (const __m128i*)( src + i + 16 ) );
            const __m128i c = _mm_loadu_si128( (const __m128i*)( src + i + 32 ) ), d = _mm_loadu_si128( (const __m128i*)( src + i + 48 ) );
            _mm_stream_si128( (__m128i*)( dst + i ), a );      _mm_stream_si128( (__m128i*)( dst + i + 16 ), b );
            _mm_stream_si128( (__m128i*)( dst + i + 32 ), c ); _mm_stream_si128( (__m128i*)( dst + i + 48 ), d );
        }
        _mm_sfence();
        if( i < n ) memcpy( dst + i, src + i, n - i );
        return;
    }
#endif
    memcpy( dst, src_v, n );
}

void SiftJob::allocImage( size_t n, const void* src )
{
    _imageData = (unsigned char*)pinned_pool().get( n );
    _pinned = _imageData != 0;
    if( !_imageData ) _imageData = job_alloc( n );
    if( _imageData ) copy_image( _imageData, src, n );
    else std::cerr << "popsift: SiftJob: out of memory for a " << _w << "x" << _h << " image" << std::endl;
}

void SiftJob::releaseImage()
{
    if( !_imageData ) return;
    const size_t n = (size_t)_w * _h * ( _is_float ? sizeof(float) : 1 );
    if( _pinned ) pinned_pool().put( _imageData, n ); else free( _imageData );
    _imageData = 0; _pinned = false;
}

SiftJob::SiftJob( int w, int h, const unsigned char* imageData )
    : _w( w ), _h( h ), _imageData( 0 ), _is_float( false ), _pinned( false ), _img( 0 )
{
    _f = _p.get_future().share();
    allocImage( (size_t)w * h, imageData );
}

SiftJob::SiftJob( int w, int h, const float* imageData )
    : _w( w ), _h( h ), _imageData( 0 ), _is_float( true ), _pinned( false ), _img( 0 )
{
    _f = _p.get_future().share();
    allocImage( (size_t)w * h * sizeof(float), imageData );
}

SiftJob::~SiftJob( ) { releaseImage(); }

void SiftJob::setImg( popsift::ImageBase* img ) { _img = img; }
popsift::ImageBase* SiftJob::getImg() { return _img; }
void SiftJob::setFeatures( FeaturesBase* f ) { _p.set_value( f ); }
FeaturesHost* SiftJob::get()     { return getHost(); }
FeaturesBase* SiftJob::getBase() { return _f.get(); }
FeaturesHost* SiftJob::getHost() { return dynamic_cast<FeaturesHost*>( _f.get() ); }
popsift::FeaturesDev* SiftJob::getDev() { return dynamic_cast<popsift::FeaturesDev*>( _f.get() ); }

/* ------------------------------------------------------------------ PopSift */

static std::atomic<int> g_device_limit( 0 );
static std::atomic<int> g_device_first( 0 );

struct PopSiftImpl
{
    std::mutex              mu;
    std::condition_variable cv;
    std::deque<SiftJob*>    queue;
    bool                    stop = false;
    bool                    started = false;
    std::atomic<long>       processed{ 0 };
    int                     first_device = 0;              /* GPUs first_device .. first_device + ctx.size() - 1 */
    int                     active = 0;                    /* jobs popped from the queue and not yet retired (under mu) */
    std::vector<psift_ctx*> ctx;
    std::vector<psift_ctx*> tiled_ctx;                     /* extractTiled()'s own contexts, band r on tiled_ctx[r] */
    psift_config            tiled_cfg;                     /* what they were last configured with */
    std::mutex              tiled_mu;
    std::vector<std::thread> workers;
    psift_config            cfg;
    popsift::Config         conf;                          /* as configured (log mode, upscale factor for the debug dumps) */
    bool                    matching = false;              /* Config::MatchingMode: results stay on the device */

    void worker( int idx );
    static FeaturesHost* convert( const psift_result& r, psift_result_block* block = 0 );
};

/* psift results -> FeaturesHost.  With a result block (psift_wait_take) the descriptors stay where the GPU put them and
 * the FeaturesHost owns the block; only the 72-byte Feature records (host pointers into that block) are written here. */
FeaturesHost* PopSiftImpl::convert( const psift_result& r, psift_result_block* block )
{
    FeaturesHost* fh = block ? new FeaturesHost( r.n_feat, r.n_desc, (popsift::Descriptor*)r.desc, block )
                             : new FeaturesHost( r.n_feat, r.n_desc );
    if( fh->getFeatureCount() != r.n_feat ) return fh;
    popsift::Descriptor* dbase = fh->getDescriptors();
    if( !block && r.n_desc > 0 ) memcpy( dbase, r.desc, sizeof(popsift::Descriptor) * (size_t)r.n_desc );
    popsift::Feature* f = fh->getFeatures();
    for( int i = 0; i < r.n_feat; i++ ) {
        const psift_feature& s = r.feat[i];
        f[i].debug_octave = s.debug_octave;
        f[i].xpos = s.xpos; f[i].ypos = s.ypos; f[i].sigma = s.sigma; f[i].num_ori = s.num_ori;
        for( int k = 0; k < ORIENTATION_MAX_COUNT; k++ ) {
            f[i].orientation[k] = s.orientation[k];
            f[i].desc[k] = ( s.desc_idx[k] >= 0 ) ? dbase + s.desc_idx[k] : 0;
        }
    }
    return fh;
}

void PopSiftImpl::worker( int idx )
{
    psift_ctx* c = ctx[idx];
    std::deque<std::pair<SiftJob*, psift_job*> > inflight;
    const int in_flight_max = std::max( 1, psift_slot_count( c ) );   /* the context's frame slots (PSIFT_SLOTS) */
    SiftJob* held = 0;                                                /* popped, but the context had no free slot yet */
    for( ;; ) {
        /* fill the pipeline: block only when nothing is in flight */
        while( (int)inflight.size() < in_flight_max ) {
            SiftJob* job = held;
            held = 0;
            if( !job ) {
                std::unique_lock<std::mutex> lk( mu );
                if( inflight.empty() ) cv.wait( lk, [&] { return stop || !queue.empty(); } );
                if( !queue.empty() ) { job = queue.front(); queue.pop_front(); active++; }
            }
            if( !job ) break;
            psift_job* pj = 0;
            int rc = job->data() == 0 ? PSIFT_E_INVALID
                   : job->isFloat() ? ( job->isPinned() ? psift_submit_pinned_f32( c, job->width(), job->height(), (const float*)job->data(), 0, &pj )
                                                        : psift_submit_f32( c, job->width(), job->height(), (const float*)job->data(), 0, &pj ) )
                                    : ( job->isPinned() ? psift_submit_pinned_u8( c, job->width(), job->height(), job->data(), 0, &pj )
                                                        : psift_submit_u8 ( c, job->width(), job->height(), job->data(), 0, &pj ) );
            if( rc == PSIFT_E_STATE && !inflight.empty() ) {
                /* every slot is in flight (another user of the context, or fewer slots than expected): not a failure -
                 * retire the oldest frame below and submit this job again */
                held = job;
                break;
            }
            if( rc != PSIFT_OK ) {
                std::cerr << "popsift: extraction failed: " << psift_last_error() << std::endl;
                { std::lock_guard<std::mutex> lk( mu ); active--; }
                job->setFeatures( matching ? (FeaturesBase*)new popsift::FeaturesDev( ) : (FeaturesBase*)new FeaturesHost( 0, 0 ) );
                continue;
            }
            inflight.push_back( std::make_pair( job, pj ) );
        }
        if( inflight.empty() ) {
            std::unique_lock<std::mutex> lk( mu );
            if( stop && queue.empty() ) return;
            continue;
        }
        /* retire the oldest frame */
        SiftJob* job = inflight.front().first; psift_job* pj = inflight.front().second;
        inflight.pop_front();
        psift_result r;
        psift_result_block* rb = 0;
        FeaturesBase* fh;
        if( matching ) {
            /* matchPrepareLoop, popsift.cpp:215-236: clone on the device, nothing goes to the host */
            psift_dev_features* df = 0;
            if( psift_wait_dev( pj, &df ) == PSIFT_OK ) fh = new popsift::FeaturesDev( df );
            else {
                std::cerr << "popsift: extraction failed: " << psift_last_error() << std::endl;
                fh = new popsift::FeaturesDev( );
            }
        } else if( psift_wait_take( pj, &r, &rb ) == PSIFT_OK ) {
            fh = convert( r, rb );
            if( r.n_desc == 0 && r.n_feat != 0 ) std::cerr << "Warning: no descriptors extracted" << std::endl;
            bool log_all;
            { std::lock_guard<std::mutex> lk( mu ); log_all = conf.getLogMode() == popsift::Config::All; }
            if( log_all )          /* popsift.cpp:201-209; the planes are still in the slot */
                popsift::debug_dump_all( c, conf, r.n_octaves, r.n_levels, static_cast<FeaturesHost*>( fh ), "pyramid" );
        } else {
            std::cerr << "popsift: extraction failed: " << psift_last_error() << std::endl;
            fh = new FeaturesHost( 0, 0 );
        }
        psift_job_free( pj );
        job->releaseImage();                 /* uploaded long ago: the page-locked copy can serve the next enqueue() */
        processed++;
        { std::lock_guard<std::mutex> lk( mu ); active--; }
        job->setFeatures( fh );
    }
}

void PopSift::setDeviceLimit( int n ) { g_device_limit = n; }
void PopSift::setDevices( int first, int count ) { g_device_first = std::max( 0, first ); g_device_limit = count; }

PopSift::PopSift( ImageMode imode )
    : _impl( new PopSiftImpl ), _last_init_w( 0 ), _last_init_h( 0 ), _image_mode( imode )
{
    configure( _config, true );
}

PopSift::PopSift( const popsift::Config& config, popsift::Config::ProcessingMode mode, ImageMode imode )
    : _impl( new PopSiftImpl ), _last_init_w( 0 ), _last_init_h( 0 ), _image_mode( imode )
{
    _impl->matching = ( mode == popsift::Config::MatchingMode );   /* popsift.cpp:26-36 */
    configure( config, true );
}

PopSift::~PopSift() { uninit(); delete _impl; }

int PopSift::deviceCount() const { return (int)_impl->ctx.size(); }

bool PopSift::configure( const popsift::Config& config, bool force )
{
    PopSiftImpl* I = _impl;
    if( I->processed.load() > 0 && !force ) return false;          /* popsift.cpp:65-67 */
    if( !I->ctx.empty() ) {
        /* re-configuration: drain first (queued jobs AND jobs a worker has popped but not yet retired), so that no
         * frame runs under a configuration other than the one it was enqueued with */
        for( ;; ) {
            { std::lock_guard<std::mutex> lk( I->mu ); if( I->queue.empty() && I->active == 0 ) break; }
            std::this_thread::yield();
        }
    }
    {
        std::lock_guard<std::mutex> lk( I->mu );                     /* workers read conf (log mode) under the same lock */
        _config = config;
        _config.levels = std::max( 2, config.levels );               /* popsift.cpp:71 */
        _config.toC( &I->cfg );
        I->conf = _config;
    }
    if( I->ctx.empty() ) {
        int n = psift_device_count();
        const char* env = getenv( "POPSIFT_NUM_GPUS" );
        int lim = g_device_limit.load();
        if( env && atoi( env ) > 0 ) lim = atoi( env );
        const int first = std::min( g_device_first.load(), std::max( 0, n - 1 ) );
        n -= first;
        I->first_device = first;
        if( lim > 0 && lim < n ) n = lim;
        if( n <= 0 ) {
            std::cerr << "popsift: no CUDA device found (this library has no CPU path)" << std::endl;
            return false;
        }
        for( int d = 0; d < n; d++ ) {
            psift_ctx* c = 0;
            if( psift_create( first + d, &I->cfg, &c ) != PSIFT_OK ) {
                std::cerr << "popsift: cannot create a context on device " << d << ": " << psift_last_error() << std::endl;
                continue;
            }
            I->ctx.push_back( c );
        }
        if( I->ctx.empty() ) return false;
        I->stop = false;
        for( size_t k = 0; k < I->ctx.size(); k++ ) I->workers.emplace_back( &PopSiftImpl::worker, I, (int)k );
        I->started = true;
        return true;
    }
    bool ok = true;
    for( auto c : I->ctx ) {
        int rc;
        while( ( rc = psift_configure( c, &I->cfg ) ) == PSIFT_E_STATE ) std::this_thread::yield();
        if( rc != PSIFT_OK ) { std::cerr << "popsift: configure failed: " << psift_last_error() << std::endl; ok = false; }
    }
    return ok;
}

void PopSift::uninit( )
{
    PopSiftImpl* I = _impl;
    if( !I->started ) return;
    { std::lock_guard<std::mutex> lk( I->mu ); I->stop = true; }
    I->cv.notify_all();
    for( auto& t : I->workers ) if( t.joinable() ) t.join();
    I->workers.clear();
    for( auto c : I->ctx ) psift_destroy( c );
    I->ctx.clear();
    { std::lock_guard<std::mutex> lk( I->tiled_mu ); for( auto c : I->tiled_ctx ) psift_destroy( c ); I->tiled_ctx.clear(); }
    I->started = false;
}

FeaturesHost* PopSift::extractTiled( int w, int h, const unsigned char* imageData, int bands )
{
    PopSiftImpl* I = _impl;
    if( !I->started || !imageData ) { std::cerr << "popsift: extractTiled on an uninitialised PopSift object" << std::endl; return new FeaturesHost( 0, 0 ); }
    std::lock_guard<std::mutex> lk( I->tiled_mu );
    const int n_dev = (int)I->ctx.size();
    const int g_tiled_first = I->first_device;
    const int n = bands > 0 ? bands : n_dev;
    if( !I->tiled_ctx.empty() && memcmp( &I->tiled_cfg, &I->cfg, sizeof I->cfg ) != 0 )
        for( auto c : I->tiled_ctx )                                  /* follow configure() */
            if( psift_configure( c, &I->cfg ) != PSIFT_OK ) {
                std::cerr << "popsift: tiled extraction: configure failed: " << psift_last_error() << std::endl;
                return new FeaturesHost( 0, 0 );
            }
    I->tiled_cfg = I->cfg;
    while( (int)I->tiled_ctx.size() < n ) {
        psift_ctx* c = 0;
        if( psift_create( g_tiled_first + (int)I->tiled_ctx.size() % n_dev, &I->cfg, &c ) != PSIFT_OK ) {
            std::cerr << "popsift: cannot create a context for a band: " << psift_last_error() << std::endl;
            return new FeaturesHost( 0, 0 );
        }
        I->tiled_ctx.push_back( c );
    }
    psift_banded_result r;
    if( psift_extract_banded( I->tiled_ctx.data(), n, w, h, imageData, 0, 0, &r ) != PSIFT_OK ) {
        std::cerr << "popsift: tiled extraction failed: " << psift_last_error() << std::endl;
        return new FeaturesHost( 0, 0 );
    }
    psift_result pr;
    memset( &pr, 0, sizeof pr );
    pr.n_feat = r.n_feat; pr.n_desc = r.n_desc; pr.feat = r.feat; pr.desc = r.desc; pr.ext = r.ext;
    FeaturesHost* fh = PopSiftImpl::convert( pr );
    psift_banded_free( &r );
    return fh;
}

SiftJob* PopSift::enqueue( int w, int h, const unsigned char* imageData )
{
    if( _image_mode != ByteImages ) {
        std::cerr << "popsift: image mode error: byte image sent to a PopSift object configured for float images" << std::endl;
        return 0;                                                    /* the reference exits here, popsift.cpp:143-148 */
    }
    if( !_impl->started ) { std::cerr << "popsift: enqueue on an uninitialised PopSift object" << std::endl; return 0; }
    SiftJob* job = new SiftJob( w, h, imageData );
    { std::lock_guard<std::mutex> lk( _impl->mu ); _impl->queue.push_back( job ); }
    _impl->cv.notify_one();
    return job;
}

SiftJob* PopSift::enqueue( int w, int h, const float* imageData )
{
    if( _image_mode != FloatImages ) {
        std::cerr << "popsift: image mode error: float image sent to a PopSift object configured for byte images" << std::endl;
        return 0;
    }
    if( !_impl->started ) { std::cerr << "popsift: enqueue on an uninitialised PopSift object" << std::endl; return 0; }
    SiftJob* job = new SiftJob( w, h, imageData );
    { std::lock_guard<std::mutex> lk( _impl->mu ); _impl->queue.push_back( job ); }
    _impl->cv.notify_one();
    return job;
}
