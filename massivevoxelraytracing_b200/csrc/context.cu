// context.cu -- hako_ctx: device selection, the stream-ordered memory pool, error strings, launch accounting.
// Replaces oroInitialize / oroCtxCreate / the nvrtc `Shader` compile (voxRTGPU.cpp:25-68, hipUtil.hpp:112-157):
// the kernels are compiled ahead of time for sm_100a, so creating a context is only "pick the device".
#include <algorithm>
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

namespace hako
{
// resident-set read sweep: every thread streams 16-byte words of a buffer that fits L2, bypassing L1 (ld.global.cg)
__global__ void __launch_bounds__( 256 ) l2ReadKernel( const uint4* __restrict__ p, uint64_t n16, int reps, uint32_t* __restrict__ sink )
{
	uint32_t acc = 0;
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for( int r = 0; r < reps; r++ )
	{
		// a different starting phase per repetition so that consecutive sweeps do not hit the same slice at the same time; four
		// independent 16-byte loads in flight per thread and iteration, no integer division in the loop
		const uint64_t off = ( (uint64_t)r * 4099u * 64u ) % n16;
		uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
		for( ; i + 3 * stride < n16; i += 4 * stride )
		{
			uint64_t a0 = i + off, a1 = a0 + stride, a2 = a1 + stride, a3 = a2 + stride;
			a0 -= a0 >= n16 ? n16 : 0, a1 -= a1 >= n16 ? n16 : 0, a2 -= a2 >= n16 ? n16 : 0, a3 -= a3 >= n16 ? n16 : 0;
			const uint4 v0 = __ldcg( p + a0 ), v1 = __ldcg( p + a1 ), v2 = __ldcg( p + a2 ), v3 = __ldcg( p + a3 );
			acc += ( v0.x ^ v0.y ^ v0.z ^ v0.w ) + ( v1.x ^ v1.y ^ v1.z ^ v1.w ) + ( v2.x ^ v2.y ^ v2.z ^ v2.w ) + ( v3.x ^ v3.y ^ v3.z ^ v3.w );
		}
		for( ; i < n16; i += stride )
		{
			uint64_t a0 = i + off;
			a0 -= a0 >= n16 ? n16 : 0;
			const uint4 v = __ldcg( p + a0 );
			acc += v.x ^ v.y ^ v.z ^ v.w;
		}
	}
	if( acc == 0x12345678u ) *sink = acc; // never true for the zero-filled buffer's checksum pattern; keeps the loads alive
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
	// A PRIVATE stream-ordered pool: steady-state rebuilds reuse its blocks (it is never trimmed while the context lives), the
	// memory goes back to the driver with the context, and the process's default pool -- which torch's or anybody else's
	// cudaMallocAsync uses -- is left alone.  One host thread per context: stats, the pinned words and the internal streams are
	// not synchronised between threads (use one hako_ctx per thread, as tests/cpp/shim_demo.cpp does per GPU).
	cudaError_t e2 = cudaSuccess;
	{
		cudaMemPoolProps props = {};
		props.allocType = cudaMemAllocationTypePinned;
		props.handleTypes = cudaMemHandleTypeNone;
		props.location.type = cudaMemLocationTypeDevice;
		props.location.id = device;
		e2 = cudaMemPoolCreate( &c->pool, &props );
		uint64_t keep = ~0ull;
		if( e2 == cudaSuccess ) e2 = cudaMemPoolSetAttribute( c->pool, cudaMemPoolAttrReleaseThreshold, &keep );
	}
	for( int i = 0; i < 8 && e2 == cudaSuccess; i++ ) e2 = cudaEventCreate( &c->ev[i] );
	if( e2 == cudaSuccess ) e2 = cudaMallocHost( (void**)&c->pinned, 64 * sizeof( uint64_t ) );
	if( e2 == cudaSuccess ) e2 = cudaStreamCreateWithFlags( &c->copyStream, cudaStreamNonBlocking );
	for( int i = 0; i < 2 && e2 == cudaSuccess; i++ ) e2 = cudaStreamCreateWithFlags( &c->bandStream[i], cudaStreamNonBlocking );
	for( int i = 0; i < 17 && e2 == cudaSuccess; i++ ) e2 = cudaEventCreateWithFlags( &c->bandEv[i], cudaEventDisableTiming );
	if( e2 != cudaSuccess )
	{
		const int r = hako::cudaFail( e2, "hako_ctx_create", __FILE__, __LINE__ );
		hako_ctx_destroy( c ); // releases whatever was created
		return r;
	}
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
	if( c->pool ) cudaMemPoolDestroy( c->pool );
	delete c;
}

int hako_ctx_synchronize( hako_ctx* c, hako_stream stream )
{
	if( !c ) return HAKO_E_INVALID;
	HAKO_CUDA( cudaSetDevice( c->device ) );
	HAKO_CUDA( cudaStreamSynchronize( (cudaStream_t)stream ) );
	return HAKO_OK;
}

int hako_ctx_build_stats( const hako_ctx* cc, hako_build_stats* out )
{
	if( !cc || !out ) return HAKO_E_INVALID;
	hako_ctx* c = const_cast<hako_ctx*>( cc );
	if( c->statsPending )
	{
		// the build returned without a final host wait: its last event may still be in flight
		HAKO_CUDA( cudaSetDevice( c->device ) );
		HAKO_CUDA( cudaEventSynchronize( c->ev[5] ) );
		hako_build_stats& st = c->stats;
		cudaEventElapsedTime( &st.ms_h2d, c->ev[0], c->ev[1] );
		cudaEventElapsedTime( &st.ms_voxelize, c->ev[1], c->ev[2] );
		cudaEventElapsedTime( &st.ms_sort, c->ev[2], c->ev[3] );
		cudaEventElapsedTime( &st.ms_unique, c->ev[3], c->ev[4] );
		cudaEventElapsedTime( &st.ms_octree, c->ev[4], c->ev[5] );
		cudaEventElapsedTime( &st.ms_total, c->ev[0], c->ev[5] );
		c->statsPending = false;
	}
	*out = c->stats;
	return HAKO_OK;
}

uint64_t hako_ctx_launch_count( const hako_ctx* c ) { return c ? c->launches.load() : 0; }

// Grow the context's memory pool to `bytes` now (one allocation, released back INTO the pool: its release threshold keeps the memory
// mapped), so that the first build of an application does not pay for the mapping on its critical path.
int hako_ctx_reserve( hako_ctx* c, hako_stream stream, uint64_t bytes )
{
	if( !c ) return HAKO_E_INVALID;
	if( bytes == 0 ) return HAKO_OK;
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( c->device ) );
	void* p = nullptr;
	cudaError_t e = cudaMallocFromPoolAsync( &p, (size_t)bytes, c->pool, s );
	if( e != cudaSuccess )
	{
		cudaGetLastError();
		hako::setError( "hako_ctx_reserve: %llu bytes: %s", (unsigned long long)bytes, cudaGetErrorString( e ) );
		return HAKO_E_CUDA;
	}
	HAKO_CUDA( cudaFreeAsync( p, s ) );
	HAKO_CUDA( cudaStreamSynchronize( s ) );
	return HAKO_OK;
}

// The L2 roof of this device, measured ("B200 L2 peak is not in MEASURED_PEAKS -- measure once with a resident-set
// microbenchmark and record"): read bandwidth over a buffer of setBytes (0: a quarter of the L2) that is swept `reps` times.
int hako_ctx_l2_probe( hako_ctx* c, hako_stream stream, uint64_t setBytes, int reps, uint64_t* l2Bytes, float* readGBs )
{
	if( !c || !readGBs ) return HAKO_E_INVALID;
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( c->device ) );
	if( l2Bytes ) *l2Bytes = c->l2Bytes;
	if( setBytes == 0 ) setBytes = c->l2Bytes / 4;
	setBytes &= ~(uint64_t)4095;
	if( setBytes < 4096 || reps < 1 ) return HAKO_E_INVALID;
	uint4* buf = nullptr;
	uint32_t* sink = nullptr;
	HAKO_TRY( hako::dalloc( c, &buf, setBytes / 16, s ) );
	int r = hako::dalloc( c, &sink, 1, s );
	float best = 0.0f;
	do
	{
		if( r ) break;
		cudaError_t e = cudaMemsetAsync( buf, 0, setBytes, s );
		if( e != cudaSuccess )
		{
			r = hako::cudaFail( e, "memset", __FILE__, __LINE__ );
			break;
		}
		const uint32_t grid = (uint32_t)c->smCount * 8u;
		for( int k = 0; k < 4 && r == 0; k++ ) // first launch warms the set into L2; best of the next three
		{
			e = cudaEventRecord( c->ev[6], s );
			hako::l2ReadKernel<<<grid, 256, 0, s>>>( buf, setBytes / 16, k == 0 ? 1 : reps, sink );
			c->launches.fetch_add( 1 );
			if( e == cudaSuccess ) e = cudaGetLastError();
			if( e == cudaSuccess ) e = cudaEventRecord( c->ev[7], s );
			if( e == cudaSuccess ) e = cudaEventSynchronize( c->ev[7] );
			float ms = 0.0f;
			if( e == cudaSuccess ) e = cudaEventElapsedTime( &ms, c->ev[6], c->ev[7] );
			if( e != cudaSuccess )
			{
				r = hako::cudaFail( e, "l2 probe", __FILE__, __LINE__ );
				break;
			}
			if( k > 0 && ms > 0.0f ) best = std::max( best, (float)( (double)setBytes * reps / ( ms * 1e6 ) ) );
		}
	} while( 0 );
	hako::dfree( buf, s ), hako::dfree( sink, s );
	*readGBs = best;
	return r;
}
}
