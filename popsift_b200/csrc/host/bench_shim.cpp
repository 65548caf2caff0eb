/* psift_cxx_stream -- drives the drop-in C++ classes exactly as a caller would (PopSift::enqueue() / SiftJob::get(),
 * reference popsift.cpp:139-169, 300-313; the calling pattern of src/application/main.cpp:240-262) from one or more
 * producer threads, so that bench.py and the tests can measure and check that path through ctypes.
 *
 * Every producer enqueues its share of the frames and keeps at most `window` jobs outstanding: when the window is
 * full it get()s the oldest job, walks its features (a checksum over positions and descriptor bytes, so the result is
 * really touched) and deletes it -- frames are enqueued at the rate they retire, not thousands up front.  The caller's
 * frame buffers are ordinary (pageable) memory: enqueue() makes its private copy, as its contract says.
 */
#include <popsift/popsift.h>
#include <popsift/features.h>
#include <psift.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>

namespace {

struct Gate {                 /* all producers pass together: used between the warm-up and the timed frames */
    std::mutex mu; std::condition_variable cv; int waiting = 0, generation = 0, n = 1;
    void pass()
    {
        std::unique_lock<std::mutex> lk( mu );
        const int g = generation;
        if( ++waiting == n ) { waiting = 0; generation++; cv.notify_all(); }
        else cv.wait( lk, [&] { return generation != g; } );
    }
};

struct Tally { long long feat = 0, desc = 0; double check = 0.0; bool failed = false; };

void retire( SiftJob* job, Tally& t )
{
    popsift::Features* f = job->get();
    if( !f ) { t.failed = true; delete job; return; }
    t.feat += f->getFeatureCount(); t.desc += f->getDescriptorCount();
    double c = 0.0;
    for( popsift::Features::F_const_iterator it = f->begin(); it != f->end(); ++it ) {
        c += it->xpos + it->ypos;
        for( int k = 0; k < it->num_ori; k++ ) c += it->desc[k]->features[0] + it->desc[k]->features[127];
    }
    t.check += c;
    delete f;
    delete job;
}

void produce( PopSift* sift, int me, int n_prod, int w, int h, const unsigned char* const* frames, int n_distinct,
              int warm, int timed, int window, Gate* gate, Tally* out,
              std::chrono::steady_clock::time_point* t_start, std::chrono::steady_clock::time_point* t_stop )
{
    std::deque<SiftJob*> q;
    Tally warm_t, t;
    auto run = [&]( int n_frames, Tally& tally ) {
        for( int i = me; i < n_frames; i += n_prod ) {
            SiftJob* j = sift->enqueue( w, h, frames[i % n_distinct] );
            if( !j ) { tally.failed = true; break; }
            q.push_back( j );
            if( (int)q.size() >= window ) { retire( q.front(), tally ); q.pop_front(); }
        }
        while( !q.empty() ) { retire( q.front(), tally ); q.pop_front(); }
    };
    run( warm, warm_t );
    gate->pass();
    if( me == 0 ) *t_start = std::chrono::steady_clock::now();
    gate->pass();
    run( timed, t );
    gate->pass();
    if( me == 0 ) *t_stop = std::chrono::steady_clock::now();
    t.failed = t.failed || warm_t.failed;
    *out = t;
}

} // namespace

extern "C" int psift_cxx_stream( int first_device, int n_devices, int n_producers, int w, int h,
                                 const unsigned char* const* frames, int n_distinct, int warm_frames, int timed_frames,
                                 int window, double* seconds, long long* n_feat, long long* n_desc, double* checksum,
                                 int* devices_used )
{
    if( !frames || n_distinct < 1 || w < 1 || h < 1 || timed_frames < 1 || n_producers < 1 || window < 1 ) return PSIFT_E_INVALID;
    PopSift::setDevices( first_device, n_devices );
    popsift::Config config;                                   /* the reference's defaults */
    PopSift sift( config, popsift::Config::ExtractingMode, PopSift::ByteImages );
    PopSift::setDevices( 0, 0 );
    if( sift.deviceCount() < 1 ) return PSIFT_E_NODEVICE;
    if( devices_used ) *devices_used = sift.deviceCount();

    Gate gate; gate.n = n_producers;
    std::vector<Tally> tallies( n_producers );
    std::vector<std::thread> th;
    std::chrono::steady_clock::time_point t0, t1;
    for( int p = 0; p < n_producers; p++ )
        th.emplace_back( produce, &sift, p, n_producers, w, h, frames, n_distinct, warm_frames, timed_frames, window, &gate, &tallies[p], &t0, &t1 );
    for( auto& t : th ) t.join();
    sift.uninit();

    long long nf = 0, nd = 0; double cs = 0.0; bool failed = false;
    for( auto& t : tallies ) { nf += t.feat; nd += t.desc; cs += t.check; failed = failed || t.failed; }
    if( seconds ) *seconds = std::chrono::duration<double>( t1 - t0 ).count();
    if( n_feat ) *n_feat = nf;
    if( n_desc ) *n_desc = nd;
    if( checksum ) *checksum = cs;
    return failed ? PSIFT_E_STATE : PSIFT_OK;
}
