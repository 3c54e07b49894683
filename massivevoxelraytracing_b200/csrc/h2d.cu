// h2d.cu -- host->device upload of the caller's (pageable) vertex arrays for IntersectorOctreeGPU::build
// (IntersectorOctreeGPU.hpp:68-70: three oroMemcpyHtoD of std::vector data per build).
//
// cudaMemcpyAsync from pageable memory is staged by the driver through one small pinned buffer and ran at ~11 GB/s on the
// B200 box (565 MB of triangles = 50 ms, 3x the whole device-side build).  Here W worker threads copy 8 MiB chunks into
// their own pinned double buffers and put them on the wire from their own streams, so host memcpy and PCIe transfers of
// different chunks overlap.  Sources that are already pinned / registered skip the staging.
#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include <functional>
#include "internal.cuh"

namespace hako
{
struct Stager
{
	static const int MAX_WORKERS = 8, NBUF = 2;
	static const size_t CHUNK = 8u << 20;
	int workers = 0;
	void* buf[MAX_WORKERS][NBUF] = {};
	cudaStream_t stream[MAX_WORKERS] = {};
	cudaEvent_t ev[MAX_WORKERS][NBUF] = {};
	cudaEvent_t start = nullptr;
	cudaStream_t joinStream = nullptr; // uploadMark: gathers the worker streams behind one event
	cudaEvent_t markEv[MAX_WORKERS] = {};
};

static int makeStager( hako_ctx* ctx )
{
	if( ctx->stager ) return 0;
	Stager* st = new Stager();
	// measured on the B200 box (16 vCPU, 565 MB): 1 worker 47.7 ms, 2: 29.2, 4: 14.2, 8: 12.8 (plain pageable copy: 50.2)
	const unsigned hc = std::thread::hardware_concurrency();
	int w = hc >= 16 ? 8 : 4;
	// one process per GPU on the same host (torchrun exports LOCAL_WORLD_SIZE): the ranks stage their uploads at the same time
	// and share the cores
	if( const char* e = getenv( "LOCAL_WORLD_SIZE" ) )
	{
		const int ranks = atoi( e );
		if( ranks > 1 && hc ) w = std::max( 2, std::min( w, (int)hc / ranks ) );
	}
	if( const char* e = getenv( "HAKO_H2D_WORKERS" ) ) w = atoi( e );
	if( hc && (unsigned)w > hc ) w = (int)hc;
	st->workers = std::max( 1, std::min( w, (int)Stager::MAX_WORKERS ) );
	ctx->stager = st;
	HAKO_CUDA( cudaEventCreateWithFlags( &st->start, cudaEventDisableTiming ) );
	HAKO_CUDA( cudaStreamCreateWithFlags( &st->joinStream, cudaStreamNonBlocking ) );
	for( int t = 0; t < st->workers; t++ )
	{
		HAKO_CUDA( cudaStreamCreateWithFlags( &st->stream[t], cudaStreamNonBlocking ) );
		HAKO_CUDA( cudaEventCreateWithFlags( &st->markEv[t], cudaEventDisableTiming ) );
		for( int b = 0; b < Stager::NBUF; b++ )
		{
			HAKO_CUDA( cudaMallocHost( &st->buf[t][b], Stager::CHUNK ) );
			HAKO_CUDA( cudaEventCreateWithFlags( &st->ev[t][b], cudaEventDisableTiming ) );
		}
	}
	return 0;
}

void destroyStager( hako_ctx* ctx )
{
	Stager* st = ctx->stager;
	if( !st ) return;
	for( int t = 0; t < Stager::MAX_WORKERS; t++ )
	{
		for( int b = 0; b < Stager::NBUF; b++ )
		{
			if( st->buf[t][b] ) cudaFreeHost( st->buf[t][b] );
			if( st->ev[t][b] ) cudaEventDestroy( st->ev[t][b] );
		}
		if( st->stream[t] ) cudaStreamDestroy( st->stream[t] );
		if( st->markEv[t] ) cudaEventDestroy( st->markEv[t] );
	}
	if( st->start ) cudaEventDestroy( st->start );
	if( st->joinStream ) cudaStreamDestroy( st->joinStream );
	delete st;
	ctx->stager = nullptr;
}

static bool isPinned( const void* p )
{
	cudaPointerAttributes a;
	if( cudaPointerGetAttributes( &a, p ) != cudaSuccess )
	{
		cudaGetLastError();
		return false;
	}
	return a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged;
}

// Puts the copies in flight on the stager's own streams (ordered behind everything enqueued on `s` so far) and returns when the
// last chunk has been handed to the DMA engine.  `s` does NOT wait for them until uploadFinish( ctx, s ) is called, so a caller
// may run this on a helper thread and keep launching kernels that do not need the data on `s`.
int uploadBegin( hako_ctx* ctx, cudaStream_t s, const H2DSegment* segs, int n, bool orderBehindStream, int groupSegs, const std::function<void( int )>& onGroup )
{
	// groups: every `groupSegs` consecutive segments form a group; onGroup( g ) is called (from whichever worker thread gets
	// there) as soon as the last chunk of group g has been handed to the DMA engine -- groups may complete out of order
	struct Job
	{
		char* dst;
		const char* src;
		size_t bytes;
		int group;
	};
	HAKO_TRY( makeStager( ctx ) );
	Stager* st = ctx->stager;
	if( orderBehindStream )
	{
		// the destination was allocated stream-ordered on s: the worker streams start behind that point.  (A helper thread that
		// continues an upload whose first part already did this passes false: it must not touch `s` while the caller enqueues on it.)
		HAKO_CUDA( cudaEventRecord( st->start, s ) );
		for( int t = 0; t < st->workers; t++ ) HAKO_CUDA( cudaStreamWaitEvent( st->stream[t], st->start, 0 ) );
	}
	std::vector<Job> jobs;
	const int nGroups = groupSegs > 0 ? ( n + groupSegs - 1 ) / groupSegs : 0;
	std::vector<std::atomic<int>> groupLeft( nGroups );
	for( auto& g : groupLeft ) g.store( 0 );
	for( int i = 0; i < n; i++ )
	{
		const int group = groupSegs > 0 ? i / groupSegs : 0;
		if( segs[i].bytes == 0 ) continue;
		if( segs[i].bytes < ( 1u << 20 ) || isPinned( segs[i].src ) )
		{
			// small or already pinned: straight onto the wire (worker stream 0; its buffer-0 event is re-recorded below)
			HAKO_CUDA( cudaMemcpyAsync( segs[i].dst, segs[i].src, segs[i].bytes, cudaMemcpyHostToDevice, st->stream[0] ) );
			continue;
		}
		for( size_t off = 0; off < segs[i].bytes; off += Stager::CHUNK )
		{
			jobs.push_back( { (char*)segs[i].dst + off, (const char*)segs[i].src + off, std::min( Stager::CHUNK, segs[i].bytes - off ), group } );
			if( nGroups ) groupLeft[group].fetch_add( 1 );
		}
	}
	// groups without staged chunks (empty, small or pinned segments) are complete once their direct copies are enqueued: now
	if( onGroup )
		for( int g = 0; g < nGroups; g++ )
			if( groupLeft[g].load() == 0 ) onGroup( g );
	const int W = std::max( 1, std::min<int>( st->workers, (int)jobs.size() ) );
	std::atomic<int> err{ 0 };
	const int device = ctx->device;
	auto work = [&]( int t ) {
		if( cudaSetDevice( device ) != cudaSuccess )
		{
			err.store( (int)cudaGetLastError() );
			return;
		}
		int k = 0;
		for( size_t j = (size_t)t; j < jobs.size(); j += (size_t)W, k++ )
		{
			const int b = k % Stager::NBUF;
			cudaError_t e = cudaEventSynchronize( st->ev[t][b] ); // the chunk that used this buffer two rounds ago has left
			if( e == cudaSuccess )
			{
				memcpy( st->buf[t][b], jobs[j].src, jobs[j].bytes );
				e = cudaMemcpyAsync( jobs[j].dst, st->buf[t][b], jobs[j].bytes, cudaMemcpyHostToDevice, st->stream[t] );
			}
			if( e == cudaSuccess ) e = cudaEventRecord( st->ev[t][b], st->stream[t] );
			if( e != cudaSuccess )
			{
				err.store( (int)e );
				return;
			}
			if( nGroups && groupLeft[jobs[j].group].fetch_sub( 1 ) == 1 && onGroup ) onGroup( jobs[j].group );
		}
	};
	std::vector<std::thread> threads;
	for( int t = 1; t < W; t++ ) threads.emplace_back( work, t );
	work( 0 );
	for( auto& th : threads ) th.join();
	if( err.load() ) return cudaFail( (cudaError_t)err.load(), "staged H2D", __FILE__, __LINE__ );
	// direct copies (and an empty job list) are covered by re-recording stream 0's events
	if( jobs.empty() || W == 1 )
		for( int b = 0; b < Stager::NBUF; b++ ) HAKO_CUDA( cudaEventRecord( st->ev[0][b], st->stream[0] ) );
	return 0;
}
// `s` continues behind the last chunk of every worker (both buffers' events cover the tail)
int uploadFinish( hako_ctx* ctx, cudaStream_t s )
{
	Stager* st = ctx->stager;
	if( !st ) return 0;
	for( int t = 0; t < st->workers; t++ )
		for( int b = 0; b < Stager::NBUF; b++ ) HAKO_CUDA( cudaStreamWaitEvent( s, st->ev[t][b], 0 ) );
	return 0;
}
// records `ev` behind every copy begun so far (a partial uploadFinish that another thread can wait on later)
int uploadMark( hako_ctx* ctx, cudaEvent_t ev )
{
	Stager* st = ctx->stager;
	if( !st ) return 0;
	// a point in every worker stream behind whatever has been enqueued on it so far (callers order themselves after the
	// enqueues they care about); safe while the workers keep enqueueing later chunks
	for( int t = 0; t < st->workers; t++ )
	{
		HAKO_CUDA( cudaEventRecord( st->markEv[t], st->stream[t] ) );
		HAKO_CUDA( cudaStreamWaitEvent( st->joinStream, st->markEv[t], 0 ) );
	}
	HAKO_CUDA( cudaEventRecord( ev, st->joinStream ) );
	return 0;
}
int uploadSegments( hako_ctx* ctx, cudaStream_t s, const H2DSegment* segs, int n )
{
	HAKO_TRY( uploadBegin( ctx, s, segs, n, true, 0, std::function<void( int )>() ) );
	return uploadFinish( ctx, s );
}
} // namespace hako
