// context.cu -- hako_ctx: device selection, the stream-ordered memory pool, error strings, launch accounting.
// Replaces oroInitialize / oroCtxCreate / the nvrtc `Shader` compile (voxRTGPU.cpp:25-68, hipUtil.hpp:112-157):
// the kernels are compiled ahead of time for sm_100a, so creating a context is only "pick the device".
#include <cstdarg>
#include <cstdio>
#include <cstring>

#include "internal.cuh"

namespace hako
{
static thread_local char g_error[512] = "";

void setError( const char* fmt, ... )
{
	va_list ap;
	va_start( ap, fmt );
	vsnprintf( g_error, sizeof( g_error ), fmt, ap );
	va_end( ap );
}
int cudaFail( cudaError_t e, const char* what, const char* file, int line )
{
	const char* base = strrchr( file, '/' );
	setError( "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString( e ), base ? base + 1 : file, line, what );
	return HAKO_E_CUDA;
}
} // namespace hako

extern "C" {

const char* hako_last_error( void ) { return hako::g_error; }
int hako_version( void ) { return 100; }

int hako_ctx_create( int device, hako_ctx** out )
{
	if( !out ) return HAKO_E_INVALID;
	*out = nullptr;
	int count = 0;
	cudaError_t e = cudaGetDeviceCount( &count );
	if( e != cudaSuccess || count == 0 )
	{
		hako::setError( "no CUDA device usable (%s); libhako_b200 has no CPU fallback", e == cudaSuccess ? "device count is 0" : cudaGetErrorString( e ) );
		return HAKO_E_CUDA;
	}
	if( device < 0 || device >= count )
	{
		hako::setError( "device %d out of range [0,%d)", device, count );
		return HAKO_E_INVALID;
	}
	cudaDeviceProp prop;
	HAKO_CUDA( cudaGetDeviceProperties( &prop, device ) );
	if( prop.major != 10 )
	{
		hako::setError( "device %d is sm_%d%d; this library ships sm_100a code only", device, prop.major, prop.minor );
		return HAKO_E_CUDA;
	}
	HAKO_CUDA( cudaSetDevice( device ) );
	hako_ctx* c = new hako_ctx();
	c->device = device;
	c->smCount = prop.multiProcessorCount;
	c->l2Bytes = (size_t)prop.l2CacheSize;
	HAKO_CUDA( cudaDeviceGetDefaultMemPool( &c->pool, device ) );
	uint64_t keep = ~0ull; // never trim: steady-state rebuilds reuse the same blocks
	HAKO_CUDA( cudaMemPoolSetAttribute( c->pool, cudaMemPoolAttrReleaseThreshold, &keep ) );
	for( int i = 0; i < 8; i++ ) HAKO_CUDA( cudaEventCreate( &c->ev[i] ) );
	HAKO_CUDA( cudaMallocHost( (void**)&c->pinned, 64 * sizeof( uint64_t ) ) );
	HAKO_CUDA( cudaStreamCreateWithFlags( &c->copyStream, cudaStreamNonBlocking ) );
	for( int i = 0; i < 2; i++ ) HAKO_CUDA( cudaStreamCreateWithFlags( &c->bandStream[i], cudaStreamNonBlocking ) );
	for( int i = 0; i < 17; i++ ) HAKO_CUDA( cudaEventCreateWithFlags( &c->bandEv[i], cudaEventDisableTiming ) );
	*out = c;
	return HAKO_OK;
}

void hako_ctx_destroy( hako_ctx* c )
{
	if( !c ) return;
	cudaSetDevice( c->device );
	cudaDeviceSynchronize();
	hako::destroyStager( c );
	for( int i = 0; i < 8; i++ )
		if( c->ev[i] ) cudaEventDestroy( c->ev[i] );
	if( c->pinned ) cudaFreeHost( c->pinned );
	for( int i = 0; i < 17; i++ )
		if( c->bandEv[i] ) cudaEventDestroy( c->bandEv[i] );
	if( c->copyStream ) cudaStreamDestroy( c->copyStream );
	for( int i = 0; i < 2; i++ )
		if( c->bandStream[i] ) cudaStreamDestroy( c->bandStream[i] );
	delete c;
}

int hako_ctx_synchronize( hako_ctx* c, hako_stream stream )
{
	if( !c ) return HAKO_E_INVALID;
	HAKO_CUDA( cudaSetDevice( c->device ) );
	HAKO_CUDA( cudaStreamSynchronize( (cudaStream_t)stream ) );
	return HAKO_OK;
}

int hako_ctx_build_stats( const hako_ctx* c, hako_build_stats* out )
{
	if( !c || !out ) return HAKO_E_INVALID;
	*out = c->stats;
	return HAKO_OK;
}

uint64_t hako_ctx_launch_count( const hako_ctx* c ) { return c ? c->launches.load() : 0; }
}
