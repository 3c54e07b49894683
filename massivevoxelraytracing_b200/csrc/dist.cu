// dist.cu -- the multi-GPU layer behind the C ABI (hako_dist_*): one hako_ctx per GPU, NCCL over NVLink 5 / NVSwitch.
//
// The reference has no multi-GPU path (its only distribution is usecase2_submission/run.py:13-27, two hosts splitting a frame
// range over SSH).  The partition is SURVEY.md 8(e)'s: the octree is read-only after the build and rays are independent, so
//   * the octree is built on one rank and REPLICATED (hako_dist_broadcast_octree: one small header broadcast, then nodes / masks /
//     attributes in one NCCL group = one fused broadcast kernel, straight between the library-owned arrays);
//   * primary rays shard by rows / views with no collective; path tracing shards spp (rank r renders iterations r, r + G, ...) and
//     the per-rank float4 accumulators are summed with ONE ncclReduce (hako_dist_reduce_accum; w carries the spp count);
//   * the build itself shards by triangles with one all-to-all of (key, attribute) pairs by Morton range (hako_dist_build_sharded).
// A C++ host written like voxPTGPU.cpp:88-193 or RTCamp.cpp:138-196 reaches all of it without Python: one process per GPU under any
// launcher (the 128-byte id of hako_dist_unique_id travels by whatever channel the launcher has), or one thread per GPU in one
// process (tests/cpp/shim_demo.cpp).  massivevoxelraytracing_b200/dist.py is a caller of these entry points.
//
// libnccl is dlopen'ed at hako_dist_init: a process that has already loaded one (torch ships its own libnccl.so.2) keeps using that
// copy, and libhako_b200.so still loads on a machine without NCCL.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "internal.cuh"

using namespace hako;

namespace
{
struct Nccl
{
	void* lib = nullptr;
	decltype( &ncclGetUniqueId ) GetUniqueId = nullptr;
	decltype( &ncclCommInitRank ) CommInitRank = nullptr;
	decltype( &ncclCommDestroy ) CommDestroy = nullptr;
	decltype( &ncclGetErrorString ) GetErrorString = nullptr;
	decltype( &ncclBroadcast ) Broadcast = nullptr;
	decltype( &ncclReduce ) Reduce = nullptr;
	decltype( &ncclAllReduce ) AllReduce = nullptr;
	decltype( &ncclAllGather ) AllGather = nullptr;
	decltype( &ncclSend ) Send = nullptr;
	decltype( &ncclRecv ) Recv = nullptr;
	decltype( &ncclGroupStart ) GroupStart = nullptr;
	decltype( &ncclGroupEnd ) GroupEnd = nullptr;
	decltype( &ncclGetVersion ) GetVersion = nullptr;
};
Nccl g_nccl;
std::mutex g_ncclMutex;

int loadNccl()
{
	std::lock_guard<std::mutex> lk( g_ncclMutex );
	if( g_nccl.lib ) return 0;
	void* h = dlopen( "libnccl.so.2", RTLD_NOW | RTLD_NOLOAD ); // the copy the process already uses (torch's), if any
	if( !h ) h = dlopen( "libnccl.so.2", RTLD_NOW );
	if( !h ) h = dlopen( "libnccl.so", RTLD_NOW );
	if( !h )
	{
		setError( "hako_dist: libnccl.so.2 not found (%s)", dlerror() );
		return HAKO_E_STATE;
	}
	bool ok = true;
#define L( field, name ) ok = ok && ( g_nccl.field = ( decltype( g_nccl.field ) )dlsym( h, name ) ) != nullptr
	L( GetUniqueId, "ncclGetUniqueId" );
	L( CommInitRank, "ncclCommInitRank" );
	L( CommDestroy, "ncclCommDestroy" );
	L( GetErrorString, "ncclGetErrorString" );
	L( Broadcast, "ncclBroadcast" );
	L( Reduce, "ncclReduce" );
	L( AllReduce, "ncclAllReduce" );
	L( AllGather, "ncclAllGather" );
	L( Send, "ncclSend" );
	L( Recv, "ncclRecv" );
	L( GroupStart, "ncclGroupStart" );
	L( GroupEnd, "ncclGroupEnd" );
	L( GetVersion, "ncclGetVersion" );
#undef L
	if( !ok )
	{
		setError( "hako_dist: libnccl.so.2 lacks a required symbol" );
		return HAKO_E_STATE;
	}
	g_nccl.lib = h;
	return 0;
}
int ncclFail( ncclResult_t r, const char* what, int line )
{
	setError( "NCCL error %d (%s) at dist.cu:%d: %s", (int)r, g_nccl.GetErrorString ? g_nccl.GetErrorString( r ) : "?", line, what );
	return HAKO_E_CUDA;
}
#define HAKO_NCCL( call )                                              \
	do                                                                 \
	{                                                                  \
		ncclResult_t r__ = ( call );                                   \
		if( r__ != ncclSuccess ) return ncclFail( r__, #call, __LINE__ ); \
	} while( 0 )
#define CHECK_ARG( cond, msg )           \
	do                                   \
	{                                    \
		if( !( cond ) )                  \
		{                                \
			hako::setError( "%s", msg ); \
			return HAKO_E_INVALID;       \
		}                                \
	} while( 0 )

constexpr int SPLITTER_SAMPLES = 2048;

// evenly spaced samples of a rank's sorted keys, followed by the key count
__global__ void sampleKeysKernel( const uint64_t* __restrict__ keys, uint64_t n, int S, uint64_t* __restrict__ out )
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if( i < S ) out[i] = n ? keys[( (uint64_t)i * ( n - 1 ) ) / (uint64_t)( S - 1 )] : 0ull;
	if( i == S ) out[S] = n;
}
// counts[j] = number of keys in [ splitter[j-1], splitter[j] ) of a sorted array (lower bounds of the splitters)
__global__ void splitCountsKernel( const uint64_t* __restrict__ keys, uint64_t n, const uint64_t* __restrict__ splitters, int world, uint64_t* __restrict__ counts )
{
	__shared__ uint64_t s_cut[65];
	const int j = threadIdx.x;
	if( j <= world )
	{
		uint64_t lo = 0, hi = n;
		if( j == 0 ) hi = 0;
		else if( j == world ) lo = n;
		else
		{
			const uint64_t key = splitters[j - 1];
			while( lo < hi )
			{
				const uint64_t mid = ( lo + hi ) >> 1;
				if( keys[mid] < key ) lo = mid + 1;
				else hi = mid;
			}
		}
		s_cut[j] = lo;
	}
	__syncthreads();
	if( j < world ) counts[j] = s_cut[j + 1] - s_cut[j];
}

// weighted quantiles of the gathered samples == dist.py: choose_splitters (a sample of rank r stands for count[r] / S pairs)
std::vector<uint64_t> chooseSplitters( const uint64_t* gathered /* [world][S + 1] */, int world, int S )
{
	struct E
	{
		uint64_t key;
		double w;
	};
	std::vector<E> all;
	for( int r = 0; r < world; r++ )
	{
		const uint64_t cnt = gathered[(size_t)r * ( S + 1 ) + S];
		if( cnt == 0 ) continue;
		for( int i = 0; i < S; i++ ) all.push_back( { gathered[(size_t)r * ( S + 1 ) + i], (double)cnt / S } );
	}
	std::vector<uint64_t> cuts( std::max( world - 1, 0 ), 0ull );
	if( all.empty() ) return cuts;
	std::stable_sort( all.begin(), all.end(), []( const E& a, const E& b ) { return a.key < b.key; } );
	std::vector<double> cw( all.size() );
	double acc = 0.0;
	for( size_t i = 0; i < all.size(); i++ ) cw[i] = ( acc += all[i].w );
	for( int j = 1; j < world; j++ )
	{
		const double target = acc * j / world;
		size_t idx = std::lower_bound( cw.begin(), cw.end(), target ) - cw.begin();
		idx = std::min( idx, all.size() - 1 );
		cuts[j - 1] = all[idx].key;
		if( j > 1 ) cuts[j - 1] = std::max( cuts[j - 1], cuts[j - 2] );
	}
	return cuts;
}

double nowMs()
{
	return std::chrono::duration<double, std::milli>( std::chrono::steady_clock::now().time_since_epoch() ).count();
}
} // namespace

struct hako_dist
{
	hako_ctx* ctx = nullptr;
	ncclComm_t comm = nullptr;
	int rank = 0, world = 1;
	uint64_t* dev = nullptr;	// small device scratch: samples, counts, headers
	uint64_t* host = nullptr; // pinned mirror
	size_t words = 0;
};

extern "C" {

int hako_dist_unique_id( uint8_t id[HAKO_DIST_ID_BYTES] )
{
	CHECK_ARG( id, "hako_dist_unique_id: null argument" );
	static_assert( HAKO_DIST_ID_BYTES == NCCL_UNIQUE_ID_BYTES, "id size" );
	HAKO_TRY( loadNccl() );
	ncclUniqueId u;
	HAKO_NCCL( g_nccl.GetUniqueId( &u ) );
	memcpy( id, u.internal, NCCL_UNIQUE_ID_BYTES );
	return HAKO_OK;
}

void hako_dist_destroy( hako_dist* d )
{
	if( !d ) return;
	cudaSetDevice( d->ctx->device );
	if( d->comm ) g_nccl.CommDestroy( d->comm );
	if( d->dev ) cudaFree( d->dev );
	if( d->host ) cudaFreeHost( d->host );
	delete d;
}

int hako_dist_init( hako_ctx* ctx, const uint8_t id[HAKO_DIST_ID_BYTES], int rank, int world, hako_dist** out )
{
	CHECK_ARG( ctx && id && out, "hako_dist_init: null argument" );
	CHECK_ARG( world >= 1 && world <= 64 && rank >= 0 && rank < world, "hako_dist_init: bad rank / world size (1..64 ranks)" );
	*out = nullptr;
	HAKO_TRY( loadNccl() );
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	hako_dist* d = new hako_dist();
	d->ctx = ctx;
	d->rank = rank;
	d->world = world;
	d->words = (size_t)world * ( SPLITTER_SAMPLES + 1 ) + (size_t)world * world + 4096;
	ncclUniqueId u;
	memcpy( u.internal, id, NCCL_UNIQUE_ID_BYTES );
	ncclResult_t nr = g_nccl.CommInitRank( &d->comm, world, u, rank );
	cudaError_t e = cudaSuccess;
	if( nr == ncclSuccess ) e = cudaMalloc( (void**)&d->dev, d->words * 8 );
	if( nr == ncclSuccess && e == cudaSuccess ) e = cudaMallocHost( (void**)&d->host, d->words * 8 );
	// connections are set up lazily by the first collective of each kind: do that here, not inside a caller's timed region
	if( nr == ncclSuccess && e == cudaSuccess ) e = cudaMemset( d->dev, 0, d->words * 8 );
	if( nr == ncclSuccess && e == cudaSuccess ) nr = g_nccl.AllReduce( d->dev, d->dev, 256, ncclUint64, ncclSum, d->comm, 0 );
	if( nr == ncclSuccess && e == cudaSuccess ) nr = g_nccl.Broadcast( d->dev, d->dev, 256, ncclUint64, 0, d->comm, 0 );
	if( nr == ncclSuccess && e == cudaSuccess ) nr = g_nccl.Reduce( d->dev, d->dev, 256, ncclFloat, ncclSum, 0, d->comm, 0 );
	if( nr == ncclSuccess && e == cudaSuccess ) nr = g_nccl.AllGather( d->dev, d->dev + 1024, 4, ncclUint64, d->comm, 0 ); // out of place: world * 4 <= 256 words
	if( nr == ncclSuccess && e == cudaSuccess && world > 1 )
	{
		nr = g_nccl.GroupStart();
		for( int p = 0; p < world && nr == ncclSuccess; p++ )
		{
			if( p == rank ) continue;
			nr = g_nccl.Send( d->dev + 2048, 8, ncclUint64, p, d->comm, 0 );
			if( nr == ncclSuccess ) nr = g_nccl.Recv( d->dev + 2048 + 8 * ( p + 1 ), 8, ncclUint64, p, d->comm, 0 );
		}
		ncclResult_t ge = g_nccl.GroupEnd();
		if( nr == ncclSuccess ) nr = ge;
	}
	if( nr == ncclSuccess && e == cudaSuccess ) e = cudaStreamSynchronize( 0 );
	if( nr != ncclSuccess || e != cudaSuccess )
	{
		int r = nr != ncclSuccess ? ncclFail( nr, "hako_dist_init", __LINE__ ) : cudaFail( e, "hako_dist_init", __FILE__, __LINE__ );
		hako_dist_destroy( d );
		return r;
	}
	*out = d;
	return HAKO_OK;
}

int hako_dist_info( const hako_dist* d, int* rank, int* world, int* ncclVersion )
{
	CHECK_ARG( d, "hako_dist_info: null argument" );
	if( rank ) *rank = d->rank;
	if( world ) *world = d->world;
	if( ncclVersion ) g_nccl.GetVersion( ncclVersion );
	return HAKO_OK;
}

int hako_dist_barrier( hako_dist* d, hako_stream stream )
{
	CHECK_ARG( d, "hako_dist_barrier: null argument" );
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( d->ctx->device ) );
	HAKO_NCCL( g_nccl.AllReduce( d->dev, d->dev, 1, ncclUint64, ncclSum, d->comm, s ) );
	HAKO_CUDA( cudaStreamSynchronize( s ) );
	return HAKO_OK;
}

// max over the ranks of a host float (device-timed milliseconds): "time every multi-GPU number as the max over ranks"
int hako_dist_max_f32( hako_dist* d, hako_stream stream, float* value )
{
	CHECK_ARG( d && value, "hako_dist_max_f32: null argument" );
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( d->ctx->device ) );
	*(float*)d->host = *value;
	HAKO_CUDA( cudaMemcpyAsync( d->dev, d->host, 4, cudaMemcpyHostToDevice, s ) );
	HAKO_NCCL( g_nccl.AllReduce( d->dev, d->dev, 1, ncclFloat, ncclMax, d->comm, s ) );
	HAKO_CUDA( cudaMemcpyAsync( d->host, d->dev, 4, cudaMemcpyDeviceToHost, s ) );
	HAKO_CUDA( cudaStreamSynchronize( s ) );
	*value = *(float*)d->host;
	return HAKO_OK;
}

// Replicate the octree of rank `root`.  src: the built octree on root (ignored elsewhere).  *out: on root src itself, elsewhere a
// new replica owned by the caller (hako_octree_destroy).  One header broadcast + host wait (the receivers must size their arrays),
// then nodes + masks + attributes as ONE NCCL group on `stream`; the call returns with that broadcast enqueued (stream-ordered).
int hako_dist_broadcast_octree( hako_dist* d, hako_stream stream, int root, hako_octree* src, hako_octree** out )
{
	CHECK_ARG( d && out && root >= 0 && root < d->world, "hako_dist_broadcast_octree: bad argument" );
	CHECK_ARG( d->rank != root || src, "hako_dist_broadcast_octree: the root rank passes the octree to replicate" );
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( d->ctx->device ) );
	*out = nullptr;
	hako_octree_info info;
	memset( &info, 0, sizeof( info ) );
	static_assert( sizeof( hako_octree_info ) % 8 == 0 && sizeof( hako_octree_info ) <= 4096, "header size" );
	if( d->rank == root )
	{
		HAKO_TRY( hako_octree_info_get( src, &info ) );
		if( !src->attrs ) info.reserved = 1; // no attribute array on the root
		memcpy( d->host, &info, sizeof( info ) );
		HAKO_CUDA( cudaMemcpyAsync( d->dev, d->host, sizeof( info ), cudaMemcpyHostToDevice, s ) );
	}
	HAKO_NCCL( g_nccl.Broadcast( d->dev, d->dev, sizeof( info ) / 8, ncclUint64, root, d->comm, s ) );
	hako_octree* o = src;
	if( d->rank != root )
	{
		HAKO_CUDA( cudaMemcpyAsync( d->host, d->dev, sizeof( info ), cudaMemcpyDeviceToHost, s ) );
		HAKO_CUDA( cudaStreamSynchronize( s ) );
		memcpy( &info, d->host, sizeof( info ) );
		const bool noAttrs = info.reserved == 1;
		info.reserved = 0;
		HAKO_TRY( hako_octree_create_replica( d->ctx, &info, &o ) );
		if( noAttrs && o->attrs )
		{
			cudaFree( o->attrs );
			o->attrs = nullptr;
		}
	}
	const uint64_t nodeBytes = (uint64_t)o->numberOfNodes * sizeof( TraceNode ), attrBytes = o->attrs ? (uint64_t)o->numberOfVoxels * 8 : 0;
	ncclResult_t nr = g_nccl.GroupStart();
	if( nr == ncclSuccess && nodeBytes ) nr = g_nccl.Broadcast( o->nodes, o->nodes, nodeBytes / 8, ncclUint64, root, d->comm, s );
	if( nr == ncclSuccess && o->numberOfNodes ) nr = g_nccl.Broadcast( o->masks, o->masks, o->numberOfNodes, ncclUint8, root, d->comm, s );
	if( nr == ncclSuccess && attrBytes ) nr = g_nccl.Broadcast( o->attrs, o->attrs, attrBytes / 8, ncclUint64, root, d->comm, s );
	ncclResult_t ge = g_nccl.GroupEnd();
	if( nr == ncclSuccess ) nr = ge;
	if( nr != ncclSuccess )
	{
		if( d->rank != root ) hako_octree_destroy( o );
		return ncclFail( nr, "octree broadcast", __LINE__ );
	}
	*out = o;
	return HAKO_OK;
}

// Sum the per-rank float4 accumulators of `pt` onto rank `root`, in place: the ONE collective of spp-sharded path tracing.
int hako_dist_reduce_accum( hako_dist* d, hako_stream stream, hako_pt* pt, int root )
{
	CHECK_ARG( d && pt && root >= 0 && root < d->world, "hako_dist_reduce_accum: bad argument" );
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( d->ctx->device ) );
	void* acc = nullptr;
	uint64_t bytes = 0;
	HAKO_TRY( hako_pt_accum_device( pt, &acc, &bytes ) );
	if( !acc || !bytes )
	{
		setError( "hako_dist_reduce_accum: the path tracer has no frame buffer (hako_pt_resize first)" );
		return HAKO_E_STATE;
	}
	HAKO_NCCL( g_nccl.Reduce( acc, acc, bytes / 4, ncclFloat, ncclSum, root, d->comm, s ) );
	return HAKO_OK;
}
// any device buffer of floats (row-sharded tiles, test plumbing)
int hako_dist_reduce_f32( hako_dist* d, hako_stream stream, float* buf_d, uint64_t count, int root )
{
	CHECK_ARG( d && ( buf_d || !count ) && root >= 0 && root < d->world, "hako_dist_reduce_f32: bad argument" );
	HAKO_CUDA( cudaSetDevice( d->ctx->device ) );
	if( count ) HAKO_NCCL( g_nccl.Reduce( buf_d, buf_d, count, ncclFloat, ncclSum, root, d->comm, (cudaStream_t)stream ) );
	return HAKO_OK;
}
int hako_dist_broadcast_bytes( hako_dist* d, hako_stream stream, void* buf_d, uint64_t bytes, int root )
{
	CHECK_ARG( d && ( buf_d || !bytes ) && root >= 0 && root < d->world, "hako_dist_broadcast_bytes: bad argument" );
	HAKO_CUDA( cudaSetDevice( d->ctx->device ) );
	if( bytes ) HAKO_NCCL( g_nccl.Broadcast( buf_d, buf_d, bytes, ncclUint8, root, d->comm, (cudaStream_t)stream ) );
	return HAKO_OK;
}

// The build sharded over the ranks (SURVEY.md 8(e) "sharding the build by Morton range"): this rank's triangles (HOST arrays, in the
// frame of the whole scene) -> the octree of the WHOLE scene, resident on every rank and byte-identical to hako_octree_build's.
//   voxelize + sort the shard | splitters = weighted quantiles of 2 048 key samples per rank (one tiny all-gather) | ONE all-to-all
//   of the (key, packed attribute) pairs by Morton range | sort + unique / integer-mean merge of the owned range | all-gather of
//   the unique voxels straight into every rank's resident arrays | DAG nodes (global numbering, hence replicated).
// timesMs (optional, 6 floats): voxelize+sort, splitters, exchange, sort+unique, gather, nodes -- host wall clock, synchronised.
int hako_dist_build_sharded( hako_dist* d, hako_stream stream, const float* pos, const float* col, const float* emi, uint32_t nTriangles,
							 const float origin[3], float dps, int gridRes, hako_octree** out, float* timesMs )
{
	CHECK_ARG( d && out && origin, "hako_dist_build_sharded: null argument" );
	*out = nullptr;
	hako_ctx* ctx = d->ctx;
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	const int world = d->world, rank = d->rank, S = SPLITTER_SAMPLES;
	double t0 = nowMs();
	auto mark = [&]( int i ) {
		if( !timesMs ) return;
		cudaStreamSynchronize( s );
		const double t1 = nowMs();
		timesMs[i] = (float)( t1 - t0 );
		t0 = t1;
	};
	hako_pairs* pairs = nullptr;
	HAKO_TRY( hako_voxelize_sorted( ctx, stream, pos, col, emi, nTriangles, origin, dps, gridRes, &pairs ) );
	mark( 0 );
	uint64_t *rk = nullptr, *rv = nullptr;
	hako_octree* o = nullptr;
	int r = 0;
	do
	{
		cudaError_t e;
		ncclResult_t nr;
#define CK( call )                                    \
	if( ( e = ( call ) ) != cudaSuccess )             \
	{                                                 \
		r = cudaFail( e, #call, __FILE__, __LINE__ ); \
		break;                                        \
	}
#define NK( call )                              \
	if( ( nr = ( call ) ) != ncclSuccess )      \
	{                                           \
		r = ncclFail( nr, #call, __LINE__ );    \
		break;                                  \
	}
		const uint64_t n = pairs->n;
		// ---- splitters ----
		// scratch layout: [0, (S + 1) * world) the gathered samples; a rank's own samples live in its slot (in-place all-gather)
		uint64_t* gathered = d->dev;
		uint64_t* mine = d->dev + (size_t)rank * ( S + 1 );
		sampleKeysKernel<<<divUp( S + 1, 256 ), 256, 0, s>>>( pairs->keys, n, S, mine );
		ctx->launches.fetch_add( 1 );
		NK( g_nccl.AllGather( mine, gathered, S + 1, ncclUint64, d->comm, s ) );
		CK( cudaMemcpyAsync( d->host, gathered, (size_t)world * ( S + 1 ) * 8, cudaMemcpyDeviceToHost, s ) );
		CK( cudaStreamSynchronize( s ) );
		const std::vector<uint64_t> splitters = chooseSplitters( d->host, world, S );
		uint64_t* spl_d = d->dev + (size_t)world * ( S + 1 );		   // [world - 1] (<= 64)
		uint64_t* cnt_d = spl_d + 64;								   // [world][world]
		uint64_t* spl_h = d->host + (size_t)world * ( S + 1 );
		uint64_t* cnt_h = spl_h + 64;
		for( int j = 0; j + 1 < world; j++ ) spl_h[j] = splitters[j];
		if( world > 1 ) CK( cudaMemcpyAsync( spl_d, spl_h, (size_t)( world - 1 ) * 8, cudaMemcpyHostToDevice, s ) );
		splitCountsKernel<<<1, 128, 0, s>>>( pairs->keys, n, spl_d, world, cnt_d + (size_t)rank * world );
		ctx->launches.fetch_add( 1 );
		NK( g_nccl.AllGather( cnt_d + (size_t)rank * world, cnt_d, world, ncclUint64, d->comm, s ) );
		CK( cudaMemcpyAsync( cnt_h, cnt_d, (size_t)world * world * 8, cudaMemcpyDeviceToHost, s ) );
		CK( cudaStreamSynchronize( s ) );
		std::vector<uint64_t> sendCnt( world ), recvCnt( world ), sendOff( world + 1, 0 ), recvOff( world + 1, 0 );
		for( int p = 0; p < world; p++ )
		{
			sendCnt[p] = cnt_h[(size_t)rank * world + p];
			recvCnt[p] = cnt_h[(size_t)p * world + rank];
			sendOff[p + 1] = sendOff[p] + sendCnt[p];
			recvOff[p + 1] = recvOff[p] + recvCnt[p];
		}
		if( sendOff[world] != n )
		{
			setError( "hako_dist_build_sharded: split counts do not add up (%llu of %llu pairs)", (unsigned long long)sendOff[world], (unsigned long long)n );
			r = HAKO_E_STATE;
			break;
		}
		const uint64_t nRecv = recvOff[world];
		if( nRecv >= 0xFFFFFFFFull )
		{
			setError( "hako_dist_build_sharded: %llu pairs in this rank's Morton range exceed the 32-bit counters", (unsigned long long)nRecv );
			r = HAKO_E_LIMIT;
			break;
		}
		mark( 1 );
		// ---- the exchange: ONE all-to-all (keys and values in the same NCCL group) ----
		if( ( r = dalloc( ctx, &rk, nRecv, s ) ) ) break;
		if( ( r = dalloc( ctx, &rv, nRecv, s ) ) ) break;
		NK( g_nccl.GroupStart() );
		nr = ncclSuccess;
		for( int p = 0; p < world && nr == ncclSuccess; p++ )
		{
			if( p == rank ) continue;
			if( sendCnt[p] ) nr = g_nccl.Send( pairs->keys + sendOff[p], sendCnt[p], ncclUint64, p, d->comm, s );
			if( nr == ncclSuccess && sendCnt[p] ) nr = g_nccl.Send( pairs->values + sendOff[p], sendCnt[p], ncclUint64, p, d->comm, s );
			if( nr == ncclSuccess && recvCnt[p] ) nr = g_nccl.Recv( rk + recvOff[p], recvCnt[p], ncclUint64, p, d->comm, s );
			if( nr == ncclSuccess && recvCnt[p] ) nr = g_nccl.Recv( rv + recvOff[p], recvCnt[p], ncclUint64, p, d->comm, s );
		}
		{
			ncclResult_t ge = g_nccl.GroupEnd();
			if( nr == ncclSuccess ) nr = ge;
		}
		if( nr != ncclSuccess )
		{
			r = ncclFail( nr, "all-to-all of voxel pairs", __LINE__ );
			break;
		}
		if( sendCnt[rank] )
		{
			CK( cudaMemcpyAsync( rk + recvOff[rank], pairs->keys + sendOff[rank], sendCnt[rank] * 8, cudaMemcpyDeviceToDevice, s ) );
			CK( cudaMemcpyAsync( rv + recvOff[rank], pairs->values + sendOff[rank], sendCnt[rank] * 8, cudaMemcpyDeviceToDevice, s ) );
		}
		// the shard's dumped voxels go back to the pool behind the exchange
		dfree( pairs->keys, s ), dfree( pairs->values, s );
		pairs->keys = pairs->values = nullptr;
		mark( 2 );
		// ---- merge the owned range ----
		uint64_t U = 0;
		uint32_t he = 0;
		if( ( r = hako_sort_unique_device( ctx, stream, rk, rv, nRecv, gridRes, &U, &he ) ) ) break;
		mark( 3 );
		// ---- gather the unique voxels of every range into the resident arrays of the octree ----
		uint64_t* meta_d = d->dev;	// [world][2]
		uint64_t* meta_h = d->host;
		meta_h[2 * rank] = U, meta_h[2 * rank + 1] = he;
		CK( cudaMemcpyAsync( meta_d + 2 * rank, meta_h + 2 * rank, 16, cudaMemcpyHostToDevice, s ) );
		NK( g_nccl.AllGather( meta_d + 2 * rank, meta_d, 2, ncclUint64, d->comm, s ) );
		CK( cudaMemcpyAsync( meta_h, meta_d, (size_t)world * 16, cudaMemcpyDeviceToHost, s ) );
		CK( cudaStreamSynchronize( s ) );
		std::vector<uint64_t> offs( world + 1, 0 );
		uint32_t anyEmission = 0;
		for( int p = 0; p < world; p++ )
		{
			offs[p + 1] = offs[p] + meta_h[2 * p];
			anyEmission |= meta_h[2 * p + 1] ? 1u : 0u;
		}
		const uint64_t total = offs[world];
		if( total >= 0xFFFFFFFFull )
		{
			setError( "hako_dist_build_sharded: %llu voxels exceed the 32-bit voxel index", (unsigned long long)total );
			r = HAKO_E_LIMIT;
			break;
		}
		void *m_d = nullptr, *a_d = nullptr;
		if( ( r = hako_octree_create_for_voxels( ctx, stream, total, origin, dps, gridRes, &o, &m_d, &a_d ) ) ) break;
		uint64_t* M = (uint64_t*)m_d;
		uint64_t* A = (uint64_t*)a_d;
		if( U )
		{
			CK( cudaMemcpyAsync( M + offs[rank], rk, U * 8, cudaMemcpyDeviceToDevice, s ) );
			CK( cudaMemcpyAsync( A + offs[rank], rv, U * 8, cudaMemcpyDeviceToDevice, s ) );
		}
		dfree( rk, s ), dfree( rv, s );
		rk = rv = nullptr;
		if( world > 1 )
		{
			// every range goes from its owner straight to its place in every other rank's arrays.  "p2p": a send / receive mesh (all
			// NVLink paths of the switch busy at once); "bcast": one in-place broadcast per range, grouped
			static const char* modeEnv = getenv( "HAKO_SHARD_GATHER" );
			const bool p2p = !( modeEnv && strcmp( modeEnv, "bcast" ) == 0 );
			NK( g_nccl.GroupStart() );
			nr = ncclSuccess;
			for( int p = 0; p < world && nr == ncclSuccess; p++ )
			{
				const uint64_t cp = meta_h[2 * p];
				if( p2p )
				{
					if( p == rank ) continue;
					if( U ) nr = g_nccl.Send( M + offs[rank], U, ncclUint64, p, d->comm, s );
					if( nr == ncclSuccess && U ) nr = g_nccl.Send( A + offs[rank], U, ncclUint64, p, d->comm, s );
					if( nr == ncclSuccess && cp ) nr = g_nccl.Recv( M + offs[p], cp, ncclUint64, p, d->comm, s );
					if( nr == ncclSuccess && cp ) nr = g_nccl.Recv( A + offs[p], cp, ncclUint64, p, d->comm, s );
				}
				else if( cp )
				{
					nr = g_nccl.Broadcast( M + offs[p], M + offs[p], cp, ncclUint64, p, d->comm, s );
					if( nr == ncclSuccess ) nr = g_nccl.Broadcast( A + offs[p], A + offs[p], cp, ncclUint64, p, d->comm, s );
				}
			}
			ncclResult_t ge = g_nccl.GroupEnd();
			if( nr == ncclSuccess ) nr = ge;
			if( nr != ncclSuccess )
			{
				r = ncclFail( nr, "gather of the unique voxels", __LINE__ );
				break;
			}
		}
		mark( 4 );
		if( ( r = hako_octree_build_nodes( o, stream, 1, (int)anyEmission ) ) ) break;
		mark( 5 );
		ctx->stats.dumpedVoxels = n;
		ctx->stats.uniqueVoxels = total;
#undef CK
#undef NK
	} while( 0 );
	if( pairs )
	{
		dfree( pairs->keys, s ), dfree( pairs->values, s );
		pairs->keys = pairs->values = nullptr;
		hako_pairs_destroy( pairs );
	}
	dfree( rk, s ), dfree( rv, s );
	if( r )
	{
		cudaStreamSynchronize( s );
		if( o ) hako_octree_destroy( o );
		return r;
	}
	*out = o;
	return HAKO_OK;
}

} // extern "C"
