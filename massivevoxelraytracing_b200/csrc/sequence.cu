// sequence.cu -- the input stage and the frame loop around the hot path.
//
//   hako_bounds_device / hako_scene_params   getBoundingBox (voxUtil.hpp:66-77) as a device reduction over resident vertex data,
//                                            and the dps rule of voxRT.cpp:188-190: origin = bbox lower, dps = max extent / gridRes
//   hako_seq_*                               the RTCamp frame loop (RTCamp.cpp:122-196): per frame updateScene -> clear -> N steps ->
//                                            resolve -> copy into one of a few pooled frame buffers -> a host thread downloads the
//                                            frame and writes NNN.png while the next frame is already building / rendering.
//                                            Headless: there is no window to present to.
#include <condition_variable>
#include <cstdlib>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "internal.cuh"

using namespace hako;

extern "C" int hako_image_save_png( const char* path, const uint8_t* rgba, int w, int h );

namespace
{
// order-preserving float <-> int map, so min / max can be integer atomics
__device__ __forceinline__ int orderedInt( float f )
{
	const int i = __float_as_int( f );
	return i >= 0 ? i : i ^ 0x7FFFFFFF;
}
__host__ __device__ __forceinline__ float orderedFloat( int i )
{
	const int j = i >= 0 ? i : i ^ 0x7FFFFFFF;
#ifdef __CUDA_ARCH__
	return __int_as_float( j );
#else
	float f;
	memcpy( &f, &j, 4 );
	return f;
#endif
}
__global__ void __launch_bounds__( 256 ) boundsKernel( const float* __restrict__ pos, uint64_t nVertices, int* __restrict__ out /* lo xyz, hi xyz (ordered ints) */ )
{
	float lo[3] = { HAKO_MAX_FLOAT, HAKO_MAX_FLOAT, HAKO_MAX_FLOAT }, hi[3] = { -HAKO_MAX_FLOAT, -HAKO_MAX_FLOAT, -HAKO_MAX_FLOAT };
	for( uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nVertices; i += (uint64_t)gridDim.x * blockDim.x )
	{
#pragma unroll
		for( int a = 0; a < 3; a++ )
		{
			const float v = pos[i * 3 + a];
			lo[a] = smin( lo[a], v ); // glm::min / glm::max are the same compare-selects (voxUtil.hpp:72-73)
			hi[a] = smax( hi[a], v );
		}
	}
#pragma unroll
	for( int a = 0; a < 3; a++ )
	{
		for( int o = 16; o; o >>= 1 )
		{
			lo[a] = smin( lo[a], __shfl_xor_sync( 0xFFFFFFFFu, lo[a], o ) );
			hi[a] = smax( hi[a], __shfl_xor_sync( 0xFFFFFFFFu, hi[a], o ) );
		}
		if( ( threadIdx.x & 31u ) == 0 )
		{
			atomicMin( out + a, orderedInt( lo[a] ) );
			atomicMax( out + 3 + a, orderedInt( hi[a] ) );
		}
	}
}
} // namespace

struct hako_seq
{
	hako_pt* pt = nullptr;
	hako_ctx* ctx = nullptr;
	int w = 0, h = 0;
	// Two streams (RTCamp.cpp:138-196 runs everything on one): frame f's clear / steps / resolve / download are only ENQUEUED on
	// `stream`, so the host returns at once and updateScene( f + 1 ) -- voxelize, sort, unique, octree on `buildStream`, with its
	// host waits -- runs on the GPU beside the rendering of frame f.  HAKO_SEQ_OVERLAP=0: one stream, the A/B reference.
	cudaStream_t stream = nullptr;		// render
	cudaStream_t buildStream = nullptr; // build (== stream when overlap is off)
	bool overlap = true;
	hako_octree* octree = nullptr; // the scene of the frame in flight
	struct Slot
	{
		uint8_t* dev = nullptr;
		uint8_t* host = nullptr; // pinned
		cudaEvent_t ready = nullptr;
	};
	std::vector<Slot> slots;
	// free slots / queued write jobs
	std::mutex mu;
	std::condition_variable cv;
	std::deque<int> freeSlots;
	struct Job
	{
		int slot;
		std::string path;
	};
	std::deque<Job> jobs;
	int inFlight = 0;
	bool quit = false;
	int writeErrors = 0;
	std::vector<std::thread> writers;
	cudaEvent_t ev[4] = {};
};

static void writerLoop( hako_seq* q )
{
	cudaSetDevice( q->ctx->device );
	for( ;; )
	{
		hako_seq::Job job;
		{
			std::unique_lock<std::mutex> lk( q->mu );
			q->cv.wait( lk, [&] { return q->quit || !q->jobs.empty(); } );
			if( q->jobs.empty() ) return;
			job = q->jobs.front();
			q->jobs.pop_front();
		}
		hako_seq::Slot& s = q->slots[job.slot];
		bool ok = cudaEventSynchronize( s.ready ) == cudaSuccess; // the frame has landed in the pinned buffer
		if( ok && !job.path.empty() ) ok = hako_image_save_png( job.path.c_str(), s.host, q->w, q->h ) == HAKO_OK;
		{
			std::lock_guard<std::mutex> lk( q->mu );
			if( !ok ) q->writeErrors++;
			q->freeSlots.push_back( job.slot );
			q->inFlight--;
		}
		q->cv.notify_all();
	}
}

extern "C" {
int hako_pt_rgba8_device( hako_pt* pt, void** rgba8_d, int* w, int* h );
hako_ctx* hako_pt_context( hako_pt* pt );

int hako_bounds_device( hako_ctx* ctx, hako_stream stream, const float* pos_d, uint64_t nVertices, float lower[3], float upper[3] )
{
	if( !ctx || !lower || !upper || ( nVertices && !pos_d ) )
	{
		setError( "hako_bounds_device: null argument" );
		return HAKO_E_INVALID;
	}
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	int* out = nullptr;
	HAKO_TRY( dalloc( ctx, &out, 6, s ) );
	// getBoundingBox starts from +-FLT_MAX (voxUtil.hpp:68-69)
	int init[6];
	for( int a = 0; a < 3; a++ )
	{
		const float big = HAKO_MAX_FLOAT, small = -HAKO_MAX_FLOAT;
		int bi, si;
		memcpy( &bi, &big, 4 ), memcpy( &si, &small, 4 );
		init[a] = bi;
		init[3 + a] = si ^ 0x7FFFFFFF;
	}
	int* pinned = (int*)ctx->pinned;
	memcpy( pinned, init, sizeof( init ) );
	cudaError_t e = cudaMemcpyAsync( out, pinned, sizeof( init ), cudaMemcpyHostToDevice, s );
	if( e == cudaSuccess && nVertices )
	{
		const uint32_t blocks = (uint32_t)std::min<uint64_t>( (uint64_t)ctx->smCount * 8, ( nVertices + 255 ) / 256 );
		boundsKernel<<<blocks, 256, 0, s>>>( pos_d, nVertices, out );
		ctx->launches.fetch_add( 1 );
		e = cudaGetLastError();
	}
	if( e == cudaSuccess ) e = cudaMemcpyAsync( pinned + 8, out, sizeof( init ), cudaMemcpyDeviceToHost, s );
	if( e == cudaSuccess ) e = cudaStreamSynchronize( s );
	dfree( out, s );
	if( e != cudaSuccess ) return cudaFail( e, "bounds", __FILE__, __LINE__ );
	for( int a = 0; a < 3; a++ )
	{
		lower[a] = orderedFloat( pinned[8 + a] );
		upper[a] = orderedFloat( pinned[8 + 3 + a] );
	}
	return HAKO_OK;
}

// voxRT.cpp:188-190 (the same in voxRTGPU.cpp / voxPTGPU.cpp): origin = bbox_lower, dps = max( size ) / (float)gridRes
int hako_scene_params( const float lower[3], const float upper[3], int gridRes, float origin[3], float* dps )
{
	if( !lower || !upper || !origin || !dps || gridRes <= 0 )
	{
		setError( "hako_scene_params: bad argument" );
		return HAKO_E_INVALID;
	}
	const float sx = upper[0] - lower[0], sy = upper[1] - lower[1], sz = upper[2] - lower[2];
	origin[0] = lower[0], origin[1] = lower[1], origin[2] = lower[2];
	*dps = smax( smax( sx, sy ), sz ) / (float)gridRes;
	return HAKO_OK;
}

int hako_seq_create( hako_pt* pt, int nBuffers, int nWriters, hako_seq** out )
{
	if( !pt || !out )
	{
		setError( "hako_seq_create: null argument" );
		return HAKO_E_INVALID;
	}
	*out = nullptr;
	void* rgba = nullptr;
	int w = 0, h = 0;
	hako_pt_rgba8_device( pt, &rgba, &w, &h );
	if( !rgba )
	{
		setError( "hako_seq_create: the path tracer has no frame buffer (call hako_pt_resize first)" );
		return HAKO_E_STATE;
	}
	hako_seq* q = new hako_seq();
	q->pt = pt;
	q->ctx = hako_pt_context( pt );
	q->w = w, q->h = h;
	cudaSetDevice( q->ctx->device );
	nBuffers = std::max( 1, std::min( nBuffers > 0 ? nBuffers : 4, 16 ) ); // RTCamp.cpp:127-131 pools 4
	nWriters = std::max( 1, std::min( nWriters > 0 ? nWriters : 2, 16 ) );
	const char* ov = getenv( "HAKO_SEQ_OVERLAP" );
	q->overlap = !( ov && atoi( ov ) == 0 );
	cudaError_t e = cudaStreamCreateWithFlags( &q->stream, cudaStreamNonBlocking );
	if( e == cudaSuccess && q->overlap ) e = cudaStreamCreateWithFlags( &q->buildStream, cudaStreamNonBlocking );
	if( !q->overlap ) q->buildStream = q->stream;
	for( int i = 0; i < 4 && e == cudaSuccess; i++ ) e = cudaEventCreate( &q->ev[i] );
	q->slots.resize( nBuffers );
	for( int i = 0; i < nBuffers && e == cudaSuccess; i++ )
	{
		e = cudaMalloc( (void**)&q->slots[i].dev, (size_t)w * h * 4 );
		if( e == cudaSuccess ) e = cudaMallocHost( (void**)&q->slots[i].host, (size_t)w * h * 4 );
		if( e == cudaSuccess ) e = cudaEventCreateWithFlags( &q->slots[i].ready, cudaEventDisableTiming );
		q->freeSlots.push_back( i );
	}
	if( e != cudaSuccess )
	{
		cudaFail( e, "hako_seq_create", __FILE__, __LINE__ );
		for( auto& s : q->slots )
		{
			if( s.dev ) cudaFree( s.dev );
			if( s.host ) cudaFreeHost( s.host );
			if( s.ready ) cudaEventDestroy( s.ready );
		}
		delete q;
		return HAKO_E_CUDA;
	}
	for( int i = 0; i < nWriters; i++ ) q->writers.emplace_back( writerLoop, q );
	*out = q;
	return HAKO_OK;
}

// one RTCamp frame (RTCamp.cpp:140-194).  Returns once the frame's work is enqueued and a pooled buffer was available;
// the PNG (pngPath may be NULL: download only) is written by a pool thread.  updateMs / renderMs (may be NULL) are the device
// times of the scene update and of clear + steps + resolve, read back synchronously only when asked for.
int hako_seq_render_frame( hako_seq* q, const float* pos, const float* col, const float* emi, uint32_t nTriangles, const float origin[3], float dps, int gridRes,
						   const hako_camera* cam, int iterations, int maxDepth, const char* pngPath, float* updateMs, float* renderMs )
{
	if( !q || !cam || !origin )
	{
		setError( "hako_seq_render_frame: null argument" );
		return HAKO_E_INVALID;
	}
	HAKO_CUDA( cudaSetDevice( q->ctx->device ) );
	cudaStream_t s = q->stream, bs = q->buildStream;
	HAKO_CUDA( cudaEventRecord( q->ev[0], bs ) );
	hako_octree* next = nullptr;
	HAKO_TRY( hako_octree_build( q->ctx, bs, pos, col, emi, nTriangles, origin, dps, gridRes, &next ) ); // PathTracer::updateScene; the previous frame may still be rendering
	HAKO_CUDA( cudaEventRecord( q->ev[1], bs ) );
	HAKO_CUDA( cudaStreamWaitEvent( s, q->ev[1], 0 ) );
	if( q->octree ) hako_octree_destroy_async( q->octree, s ); // stream-ordered behind the previous frame's kernels: no device-wide wait
	q->octree = next;
	HAKO_CUDA( cudaEventRecord( q->ev[3], s ) );
	HAKO_TRY( hako_pt_clear( q->pt, s ) );
	for( int it = 0; it < iterations; it++ ) HAKO_TRY( hako_pt_step( q->pt, q->octree, s, cam, it, 0, q->h, maxDepth ) );
	HAKO_TRY( hako_pt_resolve( q->pt, s ) );
	int slot = -1;
	{
		std::unique_lock<std::mutex> lk( q->mu );
		q->cv.wait( lk, [&] { return !q->freeSlots.empty(); } ); // RTCamp.cpp:168-170 spins on try_pop
		slot = q->freeSlots.front();
		q->freeSlots.pop_front();
	}
	void* rgba = nullptr;
	int w, h;
	hako_pt_rgba8_device( q->pt, &rgba, &w, &h );
	hako_seq::Slot& sl = q->slots[slot];
	cudaError_t e = cudaMemcpyAsync( sl.dev, rgba, (size_t)w * h * 4, cudaMemcpyDeviceToDevice, s ); // :172
	if( e == cudaSuccess ) e = cudaEventRecord( q->ev[2], s );
	if( e == cudaSuccess ) e = cudaMemcpyAsync( sl.host, sl.dev, (size_t)w * h * 4, cudaMemcpyDeviceToHost, s );
	if( e == cudaSuccess ) e = cudaEventRecord( sl.ready, s );
	if( e != cudaSuccess )
	{
		std::lock_guard<std::mutex> lk( q->mu );
		q->freeSlots.push_back( slot );
		return cudaFail( e, "hako_seq_render_frame", __FILE__, __LINE__ );
	}
	{
		std::lock_guard<std::mutex> lk( q->mu );
		q->jobs.push_back( { slot, pngPath ? pngPath : "" } );
		q->inFlight++;
	}
	q->cv.notify_all();
	if( updateMs || renderMs )
	{
		HAKO_CUDA( cudaEventSynchronize( q->ev[2] ) );
		if( updateMs ) cudaEventElapsedTime( updateMs, q->ev[0], q->ev[1] );
		if( renderMs ) cudaEventElapsedTime( renderMs, q->ev[3], q->ev[2] );
	}
	return HAKO_OK;
}

// taskGroup.waitForAllElementsToFinish (RTCamp.cpp:196); returns HAKO_E_INVALID if any PNG could not be written
int hako_seq_wait( hako_seq* q )
{
	if( !q ) return HAKO_E_INVALID;
	std::unique_lock<std::mutex> lk( q->mu );
	q->cv.wait( lk, [&] { return q->inFlight == 0; } );
	if( q->writeErrors )
	{
		setError( "hako_seq_wait: %d frame(s) could not be written", q->writeErrors );
		q->writeErrors = 0;
		return HAKO_E_INVALID;
	}
	return HAKO_OK;
}

// the most recent frame's host copy of a slot is only valid until the slot is reused; tests read the last written file instead
void hako_seq_destroy( hako_seq* q )
{
	if( !q ) return;
	hako_seq_wait( q );
	{
		std::lock_guard<std::mutex> lk( q->mu );
		q->quit = true;
	}
	q->cv.notify_all();
	for( auto& t : q->writers ) t.join();
	cudaSetDevice( q->ctx->device );
	cudaStreamSynchronize( q->stream );
	if( q->octree ) hako_octree_destroy( q->octree );
	for( auto& s : q->slots )
	{
		cudaFree( s.dev );
		cudaFreeHost( s.host );
		cudaEventDestroy( s.ready );
	}
	for( int i = 0; i < 4; i++ ) cudaEventDestroy( q->ev[i] );
	cudaStreamSynchronize( q->stream );
	if( q->overlap && q->buildStream ) cudaStreamDestroy( q->buildStream );
	cudaStreamDestroy( q->stream );
	delete q;
}
}
