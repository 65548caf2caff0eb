/* <popsift/popsift.h> -- PopSift / SiftJob, source compatible with the reference
 * (reference src/popsift/popsift.h:40-167), re-designed for B200:
 *   - no Boost: std::thread + a mutex/condvar queue (the reference needs boost::sync_queue);
 *   - one worker thread and one device context PER GPU; enqueue() is thread safe and shards
 *     the image stream across all visible GPUs (the reference is single-GPU, device_prop.cu:24-26);
 *   - several frames in flight per GPU; enqueue() copies the image (popsift.cpp:245-247);
 *   - the destructor joins the workers (the reference's is empty, popsift.cpp:59-61).
 * All CUDA work goes through the C ABI in <psift.h>. */
#pragma once
#include <future>
#include <memory>
#include <vector>

#include "sift_conf.h"
#include "sift_constants.h"
#include "features.h"

namespace popsift
{
    class ImageBase;       /* opaque: kept so that signatures naming it compile */
    class Pyramid;
}
struct PopSiftImpl;

class SiftJob
{
    std::promise<popsift::FeaturesBase*> _p;
    std::shared_future<popsift::FeaturesBase*> _f;   /* shared: get()/getDev()/getBase() may be called more than once */
    int                 _w;
    int                 _h;
    unsigned char*      _imageData;     /* private copy, byte or float pixels */
    bool                _is_float;
    bool                _pinned;        /* _imageData is a page-locked block of the library's recycled pool */
    popsift::ImageBase* _img;
    void allocImage( size_t bytes, const void* src );
public:
    /** byte image, value range 0..255 */
    SiftJob( int w, int h, const unsigned char* imageData );
    /** float image, value range [0..1[ */
    SiftJob( int w, int h, const float* imageData );
    ~SiftJob( );

    popsift::FeaturesHost* get();      /* same as getHost() */
    popsift::FeaturesBase* getBase();
    popsift::FeaturesHost* getHost();
    popsift::FeaturesDev*  getDev();

    void setImg( popsift::ImageBase* img );
    popsift::ImageBase* getImg();
    void setFeatures( popsift::FeaturesBase* f );

    /* used by the workers */
    int  width() const  { return _w; }
    int  height() const { return _h; }
    bool isFloat() const { return _is_float; }
    const unsigned char* data() const { return _imageData; }
    bool isPinned() const { return _pinned; }
    void releaseImage();               /* the upload is complete: the private copy goes back to the pool */
};

class PopSift
{
public:
    enum ImageMode { ByteImages, FloatImages };

    PopSift( ImageMode imode = ByteImages );
    PopSift( const popsift::Config&          config,
             popsift::Config::ProcessingMode mode  = popsift::Config::ExtractingMode,
             ImageMode                       imode = ByteImages );
    ~PopSift();

    /** Provide the configuration if you used the default constructor.  Returns false once
     *  images have been processed, unless force is set (reference popsift.cpp:63-87). */
    bool configure( const popsift::Config& config, bool force = false );

    void uninit( );

    SiftJob* enqueue( int w, int h, const unsigned char* imageData );
    SiftJob* enqueue( int w, int h, const float* imageData );

    /** deprecated */
    inline void uninit( int /*pipe*/ ) { uninit(); }
    /** deprecated */
    inline bool init( int /*pipe*/, int w, int h ) { _last_init_w = w; _last_init_h = h; return true; }
    /** deprecated */
    inline popsift::FeaturesBase* execute( int /*pipe*/, const unsigned char* imageData )
    {
        SiftJob* j = enqueue( _last_init_w, _last_init_h, imageData );
        if( !j ) return 0;
        popsift::FeaturesBase* f = j->getBase();
        delete j;
        return f;
    }

    /* extensions */
    int  deviceCount() const;                 /* GPUs this object dispatches to */
    static void setDeviceLimit( int n );      /* use at most n GPUs (0 = all); call before constructing */
    static void setDevices( int first, int count );   /* use GPUs first .. first+count-1 (count 0 = all from first); call before constructing */
    /* One giant image cut into row bands over all GPUs of this object (the reference runs such an
     * image un-tiled on one GPU): synchronous, returns the merged features, identical to the un-tiled result.  Uses its
     * own contexts (created on first use), so it may be called while enqueue()d jobs are in flight.  The caller owns
     * the result.  bands = 0: one band per GPU. */
    popsift::FeaturesHost* extractTiled( int w, int h, const unsigned char* imageData, int bands = 0 );

private:
    PopSiftImpl*    _impl;
    popsift::Config _config;
    int             _last_init_w;
    int             _last_init_h;
    ImageMode       _image_mode;
};
