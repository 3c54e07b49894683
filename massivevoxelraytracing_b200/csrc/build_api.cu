// build_api.cu -- C ABI of the build path: IntersectorOctreeGPU::build (IntersectorOctreeGPU.hpp:40-241),
// IntersectorOctree::build / buildDAGReference (IntersectorOctree.hpp:221-235) and the individual stages.
//
// Launch sequence of the reference: voxCount -> [sync] -> voxelize -> sortPairs (3 kernels x P) -> [sync] -> unique -> [sync]
// -> octreeTaskInit -> [sync] -> bottomUpOctreeBuild x log2(N) (+ memset of the hash table per level) -> [sync] -> D2D compaction
// -> embedMasks, with ~12 oroMalloc/oroFree.  Here: columns -> [read] -> count -> [read D] -> emit -> histogram + P onesweep
// passes -> unique -> [read U] -> level histogram -> [read] -> 4 kernels per level -> [read node count] -> D2D copy.
// Allocations come from the context's stream-ordered pool.
#include <condition_variable>
#include <mutex>
#include <thread>

#include "internal.cuh"

using namespace hako;

namespace hako
{
int uniqueMergeLaunch( hako_ctx* ctx, cudaStream_t s, const uint64_t* keys, const uint64_t* values, uint64_t n,
					   uint64_t* keysOut, hako_voxel_attrib* attrsOut, unsigned long long* counters_d );
}

#define CHECK_ARG( cond, msg )           \
	do                                   \
	{                                    \
		if( !( cond ) )                  \
		{                                \
			hako::setError( "%s", msg ); \
			return HAKO_E_INVALID;       \
		}                                \
	} while( 0 )

static bool isPow2( int v ) { return v > 0 && ( v & ( v - 1 ) ) == 0; }
static int log2i( int v )
{
	int l = 0;
	while( ( 1 << l ) < v ) l++;
	return l;
}
// IntersectorOctreeGPU.hpp:118 sorts on the next multiple of 8 of 3 * log2( gridRes ) bits; the keys have no bits above
// 3 * log2( gridRes ), so sorting on exactly that window gives the same order and lets the sort pick 9-bit digits
static int sortEndBit( int gridRes ) { return 3 * log2i( gridRes ); }

static void initOctree( hako_octree* o, hako_ctx* ctx, const float origin[3], float dps, int gridRes )
{
	o->ctx = ctx;
	o->lower = { origin[0], origin[1], origin[2] };
	// IntersectorOctreeGPU.hpp:63-65 / IntersectorOctree.hpp:224: upper = origin + dps * gridRes
	const float ext = dps * (float)gridRes;
	o->upper = { origin[0] + ext, origin[1] + ext, origin[2] + ext };
	o->dps = dps;
	o->gridRes = gridRes;
}

static int buildFromDeviceTriangles( hako_ctx* ctx, cudaStream_t s, const float* pos_d, const float* col_d, const float* emi_d, uint32_t nTriangles,
									 const float origin[3], float dps, int gridRes, hako_octree** out,
									 const std::function<int( int )>& beforeAttributesAreRead = std::function<int( int )>(), int emitSlices = 1 )
{
	hako_build_stats& st = ctx->stats;
	const uint64_t launches0 = ctx->launches.load();
	hako_octree* o = new hako_octree();
	initOctree( o, ctx, origin, dps, gridRes );
	uint64_t *keys = nullptr, *values = nullptr, *keysTmp = nullptr, *valuesTmp = nullptr;
	bool mergeCountersPending = false;
	uint64_t levelBoundaries[22] = {};
	bool haveLevels = false;
	int r = 0;
	do
	{
		cudaError_t e;
#define CK( call )                                    \
	if( ( e = ( call ) ) != cudaSuccess )             \
	{                                                 \
		r = cudaFail( e, #call, __FILE__, __LINE__ ); \
		break;                                        \
	}
		CK( cudaEventRecord( ctx->ev[1], s ) );
		uint64_t D = 0;
		if( ( r = voxelizeDevice( ctx, s, pos_d, col_d, emi_d, nTriangles, o->lower, dps, gridRes, 1 /* sixSeparating: voxKernel.cu:68,109 */, &keys, &values, &D,
								   beforeAttributesAreRead, emitSlices ) ) )
			break;
		st.dumpedVoxels = D;
		CK( cudaEventRecord( ctx->ev[2], s ) );
		uint64_t U = 0;
		if( D )
		{
			if( ( r = dalloc( ctx, &keysTmp, D, s ) ) ) break;
			if( ( r = dalloc( ctx, &valuesTmp, D, s ) ) ) break;
			int inTmp = 0;
			if( ( r = sortPairsDevice( ctx, s, keys, values, keysTmp, valuesTmp, D, 0, sortEndBit( gridRes ), &inTmp ) ) ) break;
			CK( cudaEventRecord( ctx->ev[3], s ) );
			uint64_t* sk = inTmp ? keysTmp : keys;
			uint64_t* sv = inTmp ? valuesTmp : values;
			// the idle pair of work buffers goes back first; then the number of runs is counted (one read of the keys) so that
			// the merge can write straight into the exact-size resident arrays -- no D-sized output pair, no copies
			dfree( inTmp ? keys : keysTmp, s ), dfree( inTmp ? values : valuesTmp, s );
			if( inTmp )
				keys = values = nullptr;
			else
				keysTmp = valuesTmp = nullptr;
			uint64_t runs = 0;
			if( ( r = countRunsDevice( ctx, s, sk, D, &runs, levelBoundaries ) ) ) break;
			haveLevels = true;
			if( ( r = dalloc( ctx, &o->mortons, runs, s ) ) ) break;
			if( ( r = dalloc( ctx, &o->attrs, runs, s ) ) ) break;
			// the merge's own counters (U again, hasEmission) are read with the end-of-build synchronisation: no host wait here  (pinned slots 32 / 33: the octree stage uses 0..21)
			unsigned long long* uc = nullptr;
			if( ( r = dalloc( ctx, &uc, 2, s ) ) ) break;
			r = uniqueMergeLaunch( ctx, s, sk, sv, D, o->mortons, o->attrs, uc );
			if( r == 0 )
			{
				e = cudaMemcpyAsync( ctx->pinned + 32, uc, 16, cudaMemcpyDeviceToHost, s );
				if( e != cudaSuccess ) r = cudaFail( e, "D2H unique counters", __FILE__, __LINE__ );
			}
			dfree( uc, s );
			if( r ) break;
			U = runs;
			mergeCountersPending = true;
			dfree( sk, s ), dfree( sv, s );
			keys = values = keysTmp = valuesTmp = nullptr;
		}
		else
		{
			CK( cudaEventRecord( ctx->ev[3], s ) );
		}
		st.uniqueVoxels = U;
		if( U >= 0xFFFFFFFFull )
		{
			setError( "build: %llu voxels exceed the 32-bit voxel index", (unsigned long long)U );
			r = HAKO_E_LIMIT;
			break;
		}
		o->numberOfVoxels = (uint32_t)U;
		CK( cudaEventRecord( ctx->ev[4], s ) );
		if( ( r = buildOctreeDevice( ctx, s, o->mortons, U, gridRes, 1, &o->nodes, &o->masks, &o->numberOfNodes, nullptr, &o->wide, haveLevels ? levelBoundaries : nullptr ) ) ) break;
		st.numberOfNodes = o->numberOfNodes;
		CK( cudaEventRecord( ctx->ev[5], s ) );
		// No host wait here: the octree stage's own read of the final node count (its last host wait) has drained everything the
		// host looks at -- the merge's counters below included; what may still be in flight on `s` is the copy of the nodes into
		// their exact-size arrays, stream-ordered like the reference's final embedMasks launch (IntersectorOctreeGPU.hpp:230-238).
		// The per-stage times are read from the events when hako_ctx_build_stats asks for them.
		if( U && o->numberOfNodes == 0 )
		{
			CK( cudaStreamSynchronize( s ) ); // a tree without nodes (gridRes 1): nothing waited yet
			st.hostSyncs++;
		}
		if( mergeCountersPending )
		{
			if( ctx->pinned[32] != U )
			{
				setError( "build: unique produced %llu voxels, %llu runs were counted", (unsigned long long)ctx->pinned[32], (unsigned long long)U );
				r = HAKO_E_STATE;
				break;
			}
			o->hasEmission = ctx->pinned[33] ? 1u : 0u;
		}
		ctx->statsPending = true;
#undef CK
	} while( 0 );
	dfree( keys, s ), dfree( values, s ), dfree( keysTmp, s ), dfree( valuesTmp, s );
	st.kernelLaunches = (uint32_t)( ctx->launches.load() - launches0 );
	if( r )
	{
		hako_octree_destroy( o );
		return r;
	}
	*out = o;
	return HAKO_OK;
}

// The caller's three HOST vertex arrays go up with the call (IntersectorOctreeGPU.hpp:68-70) and `body` runs on the device copies:
// body( pos_d, col_d, emi_d, beforeAttributesAreRead, slices ) -- see voxelizeDevice for the hook.
static int withStagedTriangles( hako_ctx* ctx, cudaStream_t s, const float* pos, const float* col, const float* emi, uint32_t nTriangles,
								const std::function<int( const float*, const float*, const float*, const std::function<int( int )>&, int )>& body )
{
	float* d = nullptr;
	const uint64_t nf = (uint64_t)nTriangles * 9;
	HAKO_TRY( dalloc( ctx, &d, nf * 3, s ) );
	int r = 0;
	// The positions go up first; the colour / emission arrays follow on a helper thread, in SLICES triangle ranges (colour and
	// emission of a range together), while the counting pass of the voxelizer (which reads positions only) and then the emit
	// pass of the earlier ranges already run: slice k of the emit pass is ordered behind the copies of ranges 0..k only.
	const int slicesEnv = getenv( "HAKO_BUILD_SLICES" ) ? atoi( getenv( "HAKO_BUILD_SLICES" ) ) : 6;
	const int SLICES = std::max( 1, std::min( { slicesEnv, 16, (int)std::max<uint32_t>( 1u, nTriangles / 65536u ) } ) ); // = voxelizeDevice's clamp
	std::thread helper;
	int helperResult = 0;
	std::mutex mtx;
	std::condition_variable cv;
	int slicesUp = 0; // ranges whose copies are enqueued and whose event is recorded
	bool helperDone = false;
	cudaEvent_t sliceEv[16] = {};
	if( nf && nf * 4 < ( 4u << 20 ) )
	{
		// small meshes: three plain copies, no staging threads
		cudaError_t e = cudaMemcpyAsync( d, pos, nf * 4, cudaMemcpyHostToDevice, s );
		if( e == cudaSuccess ) e = cudaMemcpyAsync( d + nf, col, nf * 4, cudaMemcpyHostToDevice, s );
		if( e == cudaSuccess ) e = cudaMemcpyAsync( d + 2 * nf, emi, nf * 4, cudaMemcpyHostToDevice, s );
		if( e != cudaSuccess ) r = cudaFail( e, "H2D vertices", __FILE__, __LINE__ );
	}
	else if( nf )
	{
		const H2DSegment segPos = { d, pos, nf * 4 };
		r = uploadSegments( ctx, s, &segPos, 1 );
		for( int k = 0; k < SLICES && r == 0; k++ )
		{
			cudaError_t e = cudaEventCreateWithFlags( &sliceEv[k], cudaEventDisableTiming );
			if( e != cudaSuccess ) r = cudaFail( e, "cudaEventCreate", __FILE__, __LINE__ );
		}
		if( r == 0 )
			helper = std::thread( [&]() {
				// ONE staged upload of all ranges (range-major: colour k, emission k, colour k+1, ...) keeps every worker busy; a
				// range is published when the contiguous prefix of completed ranges reaches it
				std::vector<H2DSegment> segs;
				for( int k = 0; k < SLICES; k++ )
				{
					const uint64_t f0 = (uint64_t)nTriangles * k / SLICES * 9, f1 = (uint64_t)nTriangles * ( k + 1 ) / SLICES * 9;
					segs.push_back( { d + nf + f0, col + f0, ( f1 - f0 ) * 4 } );
					segs.push_back( { d + 2 * nf + f0, emi + f0, ( f1 - f0 ) * 4 } );
				}
				bool done[16] = {};
				int markResult = 0;
				auto onRange = [&]( int g ) {
					std::lock_guard<std::mutex> lk( mtx );
					done[g] = true;
					while( slicesUp < SLICES && done[slicesUp] && markResult == 0 )
					{
						cudaSetDevice( ctx->device ); // worker threads of the stager have set it; the direct-copy path runs here
						markResult = uploadMark( ctx, sliceEv[slicesUp] );
						if( markResult == 0 ) slicesUp++;
					}
					cv.notify_all();
				};
				int hr = cudaSetDevice( ctx->device ) == cudaSuccess ? 0 : HAKO_E_CUDA;
				if( hr == 0 ) hr = uploadBegin( ctx, s, segs.data(), (int)segs.size(), false, 2, onRange );
				{
					std::lock_guard<std::mutex> lk( mtx );
					helperResult = hr ? hr : markResult;
					helperDone = true;
				}
				cv.notify_all();
			} );
	}
	// beforeEmit( k ): ranges 0..k are needed; -1: everything
	auto join = [&]( int k ) -> int {
		if( !helper.joinable() && !helperDone ) return 0; // small mesh: plain copies on `s`
		const int need = k < 0 ? SLICES : std::min( SLICES, k + 1 );
		{
			std::unique_lock<std::mutex> lk( mtx );
			cv.wait( lk, [&]() { return slicesUp >= need || helperDone; } );
			if( helperResult )
			{
				setError( "staged upload of the colour / emission arrays failed" );
				return helperResult;
			}
		}
		HAKO_CUDA( cudaStreamWaitEvent( s, sliceEv[need - 1], 0 ) );
		return 0;
	};
	if( r == 0 ) r = body( d, d + nf, d + 2 * nf, join, SLICES );
	if( helper.joinable() ) helper.join();
	for( int k = 0; k < 16; k++ )
		if( sliceEv[k] ) cudaEventDestroy( sliceEv[k] );
	dfree( d, s );
	return r;
}


extern "C" {

int hako_octree_build_device( hako_ctx* ctx, hako_stream stream, const float* pos_d, const float* col_d, const float* emi_d,
							  uint32_t nTriangles, const float origin[3], float dps, int gridRes, hako_octree** out )
{
	CHECK_ARG( ctx && out && origin, "hako_octree_build_device: null argument" );
	CHECK_ARG( nTriangles == 0 || ( pos_d && col_d && emi_d ), "hako_octree_build_device: null vertex data" );
	CHECK_ARG( isPow2( gridRes ) && gridRes <= ( 1 << 21 ), "hako_octree_build: gridRes must be a power of two <= 2^21 (IntersectorOctreeGPU.hpp:48-51 aborts)" );
	CHECK_ARG( dps > 0.0f, "hako_octree_build: dps must be positive" );
	*out = nullptr;
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	ctx->stats = hako_build_stats{};
	ctx->stats.nTriangles = nTriangles;
	HAKO_CUDA( cudaEventRecord( ctx->ev[0], s ) );
	return buildFromDeviceTriangles( ctx, s, pos_d, col_d, emi_d, nTriangles, origin, dps, gridRes, out );
}

int hako_octree_build( hako_ctx* ctx, hako_stream stream, const float* pos, const float* col, const float* emi,
					   uint32_t nTriangles, const float origin[3], float dps, int gridRes, hako_octree** out )
{
	CHECK_ARG( ctx && out && origin, "hako_octree_build: null argument" );
	CHECK_ARG( nTriangles == 0 || ( pos && col && emi ), "hako_octree_build: null vertex data" );
	CHECK_ARG( isPow2( gridRes ) && gridRes <= ( 1 << 21 ), "hako_octree_build: gridRes must be a power of two <= 2^21 (IntersectorOctreeGPU.hpp:48-51 aborts)" );
	CHECK_ARG( dps > 0.0f, "hako_octree_build: dps must be positive" );
	*out = nullptr;
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	ctx->stats = hako_build_stats{};
	ctx->stats.nTriangles = nTriangles;
	HAKO_CUDA( cudaEventRecord( ctx->ev[0], s ) );
	return withStagedTriangles( ctx, s, pos, col, emi, nTriangles, [&]( const float* p_d, const float* c_d, const float* e_d, const std::function<int( int )>& join, int slices ) {
		return buildFromDeviceTriangles( ctx, s, p_d, c_d, e_d, nTriangles, origin, dps, gridRes, out, join, slices );
	} );
}

int hako_octree_build_from_morton( hako_ctx* ctx, hako_stream stream, const uint64_t* mortons, const hako_voxel_attrib* attrs,
								   uint64_t n, const float origin[3], float dps, int gridRes, int dag, hako_octree** out )
{
	CHECK_ARG( ctx && out && origin && ( n == 0 || mortons ), "hako_octree_build_from_morton: null argument" );
	CHECK_ARG( isPow2( gridRes ) && gridRes <= ( 1 << 21 ), "hako_octree_build_from_morton: gridRes must be a power of two <= 2^21" );
	CHECK_ARG( n < 0xFFFFFFFFull, "hako_octree_build_from_morton: more than 2^32 - 1 voxels" );
	*out = nullptr;
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	ctx->stats = hako_build_stats{};
	hako_octree* o = new hako_octree();
	initOctree( o, ctx, origin, dps, gridRes );
	o->numberOfVoxels = (uint32_t)n;
	int r = 0;
	do
	{
		cudaError_t e;
		if( ( r = dalloc( ctx, &o->mortons, n, s ) ) ) break;
		if( n && ( e = cudaMemcpyAsync( o->mortons, mortons, n * 8, cudaMemcpyHostToDevice, s ) ) != cudaSuccess )
		{
			r = cudaFail( e, "H2D mortons", __FILE__, __LINE__ );
			break;
		}
		if( attrs && n )
		{
			if( ( r = dalloc( ctx, &o->attrs, n, s ) ) ) break;
			if( ( e = cudaMemcpyAsync( o->attrs, attrs, n * 8, cudaMemcpyHostToDevice, s ) ) != cudaSuccess )
			{
				r = cudaFail( e, "H2D attrs", __FILE__, __LINE__ );
				break;
			}
			for( uint64_t i = 0; i < n && !o->hasEmission; i++ ) o->hasEmission = ( attrs[i].emission[0] | attrs[i].emission[1] | attrs[i].emission[2] ) != 0;
		}
		uint32_t launches = 0;
		if( ( r = buildOctreeDevice( ctx, s, o->mortons, n, gridRes, dag, &o->nodes, &o->masks, &o->numberOfNodes, &launches, &o->wide ) ) ) break;
		ctx->stats.uniqueVoxels = n;
		ctx->stats.numberOfNodes = o->numberOfNodes;
		ctx->stats.kernelLaunches = launches;
		if( ( e = cudaStreamSynchronize( s ) ) != cudaSuccess ) r = cudaFail( e, "sync", __FILE__, __LINE__ );
	} while( 0 );
	if( r )
	{
		hako_octree_destroy( o );
		return r;
	}
	*out = o;
	return HAKO_OK;
}

// ---- stages of a build sharded over GPUs (massivevoxelraytracing_b200/dist.py: build_sharded) ----
// rank r voxelizes and sorts ITS triangles; the (key, packed attribute) pairs are exchanged by Morton range; every rank merges the
// range it owns; the unique voxels are gathered and the DAG octree is built on every rank from the same sorted list.

// stage 1: a shard of the triangle soup (HOST arrays, in the frame of the whole scene) -> sorted dumped voxels on the device
int hako_voxelize_sorted( hako_ctx* ctx, hako_stream stream, const float* pos, const float* col, const float* emi, uint32_t nTriangles,
						  const float origin[3], float dps, int gridRes, hako_pairs** out )
{
	CHECK_ARG( ctx && out && origin, "hako_voxelize_sorted: null argument" );
	CHECK_ARG( nTriangles == 0 || ( pos && col && emi ), "hako_voxelize_sorted: null vertex data" );
	CHECK_ARG( isPow2( gridRes ) && gridRes <= ( 1 << 21 ) && dps > 0.0f, "hako_voxelize_sorted: bad grid" );
	*out = nullptr;
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	hako_pairs* p = new hako_pairs();
	p->ctx = ctx;
	uint64_t *keys = nullptr, *values = nullptr, *keysTmp = nullptr, *valuesTmp = nullptr;
	uint64_t D = 0;
	// same staged, range-overlapped upload as hako_octree_build
	int r = withStagedTriangles( ctx, s, pos, col, emi, nTriangles, [&]( const float* p_d, const float* c_d, const float* e_d, const std::function<int( int )>& join, int slices ) {
		return voxelizeDevice( ctx, s, p_d, c_d, e_d, nTriangles, V3{ origin[0], origin[1], origin[2] }, dps, gridRes, 1, &keys, &values, &D, join, slices );
	} );
	do
	{
		if( r ) break;
		if( D )
		{
			if( ( r = dalloc( ctx, &keysTmp, D, s ) ) ) break;
			if( ( r = dalloc( ctx, &valuesTmp, D, s ) ) ) break;
			int inTmp = 0;
			if( ( r = sortPairsDevice( ctx, s, keys, values, keysTmp, valuesTmp, D, 0, sortEndBit( gridRes ), &inTmp ) ) ) break;
			if( inTmp ) std::swap( keys, keysTmp ), std::swap( values, valuesTmp );
		}
		cudaError_t e = cudaStreamSynchronize( s );
		if( e != cudaSuccess )
		{
			r = cudaFail( e, "hako_voxelize_sorted", __FILE__, __LINE__ );
			break;
		}
		p->keys = keys, p->values = values, p->n = D;
		keys = values = nullptr;
	} while( 0 );
	dfree( keys, s ), dfree( values, s ), dfree( keysTmp, s ), dfree( valuesTmp, s );
	if( r )
	{
		delete p;
		return r;
	}
	*out = p;
	return HAKO_OK;
}
int hako_pairs_info( const hako_pairs* p, uint64_t* n, void** keys_d, void** values_d )
{
	CHECK_ARG( p, "hako_pairs_info: null argument" );
	if( n ) *n = p->n;
	if( keys_d ) *keys_d = p->keys;
	if( values_d ) *values_d = p->values;
	return HAKO_OK;
}
void hako_pairs_destroy( hako_pairs* p )
{
	if( !p ) return;
	cudaSetDevice( p->ctx->device );
	dfree( p->keys, (cudaStream_t) nullptr ), dfree( p->values, (cudaStream_t) nullptr );
	delete p;
}

// stage 2: pairs received from every rank for the Morton range this rank owns (DEVICE arrays, a concatenation of sorted runs)
// -> sorted unique keys + integer-mean attributes, in place at the front of the same arrays
int hako_sort_unique_device( hako_ctx* ctx, hako_stream stream, uint64_t* keys_d, uint64_t* values_d, uint64_t n, int gridRes, uint64_t* U, uint32_t* hasEmission )
{
	CHECK_ARG( ctx && U && ( n == 0 || ( keys_d && values_d ) ), "hako_sort_unique_device: null argument" );
	CHECK_ARG( isPow2( gridRes ) && gridRes <= ( 1 << 21 ), "hako_sort_unique_device: bad grid" );
	CHECK_ARG( n < 0xFFFFFFFFull, "hako_sort_unique_device: more than 2^32 - 1 pairs" );
	CHECK_ARG( ( ( (uintptr_t)keys_d | (uintptr_t)values_d ) & 15u ) == 0, "hako_sort_unique_device: keys_d / values_d must be 16-byte aligned (the merge kernel reads pairs of keys)" );
	*U = 0;
	if( hasEmission ) *hasEmission = 0;
	if( n == 0 ) return HAKO_OK;
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	uint64_t *keysTmp = nullptr, *valuesTmp = nullptr;
	int r = 0;
	do
	{
		if( ( r = dalloc( ctx, &keysTmp, n, s ) ) ) break;
		if( ( r = dalloc( ctx, &valuesTmp, n, s ) ) ) break;
		int inTmp = 0;
		if( ( r = sortPairsDevice( ctx, s, keys_d, values_d, keysTmp, valuesTmp, n, 0, sortEndBit( gridRes ), &inTmp ) ) ) break;
		uint64_t* sk = inTmp ? keysTmp : keys_d;
		uint64_t* sv = inTmp ? valuesTmp : values_d;
		uint64_t* uk = inTmp ? keys_d : keysTmp;
		uint64_t* ua = inTmp ? values_d : valuesTmp;
		uint32_t he = 0;
		if( ( r = uniqueMergeDevice( ctx, s, sk, sv, n, uk, (hako_voxel_attrib*)ua, U, &he ) ) ) break; // synchronises to read U
		if( hasEmission ) *hasEmission = he;
		if( uk != keys_d )
		{
			cudaError_t e = cudaMemcpyAsync( keys_d, uk, *U * 8, cudaMemcpyDeviceToDevice, s );
			if( e == cudaSuccess ) e = cudaMemcpyAsync( values_d, ua, *U * 8, cudaMemcpyDeviceToDevice, s );
			if( e != cudaSuccess )
			{
				r = cudaFail( e, "hako_sort_unique_device", __FILE__, __LINE__ );
				break;
			}
		}
	} while( 0 );
	dfree( keysTmp, s ), dfree( valuesTmp, s );
	return r;
}

// stage 3: the gathered sorted unique voxels (DEVICE arrays) -> octree; == hako_octree_build_from_morton without the upload
int hako_octree_build_from_morton_device( hako_ctx* ctx, hako_stream stream, const uint64_t* mortons_d, const hako_voxel_attrib* attrs_d,
										  uint64_t n, const float origin[3], float dps, int gridRes, int dag, int hasEmission, hako_octree** out )
{
	CHECK_ARG( ctx && out && origin && ( n == 0 || mortons_d ), "hako_octree_build_from_morton_device: null argument" );
	CHECK_ARG( isPow2( gridRes ) && gridRes <= ( 1 << 21 ), "hako_octree_build_from_morton_device: gridRes must be a power of two <= 2^21" );
	CHECK_ARG( n < 0xFFFFFFFFull, "hako_octree_build_from_morton_device: more than 2^32 - 1 voxels" );
	*out = nullptr;
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	ctx->stats = hako_build_stats{};
	hako_octree* o = new hako_octree();
	initOctree( o, ctx, origin, dps, gridRes );
	o->numberOfVoxels = (uint32_t)n;
	o->hasEmission = hasEmission ? 1u : 0u;
	int r = 0;
	do
	{
		cudaError_t e = cudaSuccess;
		if( ( r = dalloc( ctx, &o->mortons, n, s ) ) ) break;
		if( n ) e = cudaMemcpyAsync( o->mortons, mortons_d, n * 8, cudaMemcpyDeviceToDevice, s );
		if( e == cudaSuccess && attrs_d && n )
		{
			if( ( r = dalloc( ctx, &o->attrs, n, s ) ) ) break;
			e = cudaMemcpyAsync( o->attrs, attrs_d, n * 8, cudaMemcpyDeviceToDevice, s );
		}
		if( e != cudaSuccess )
		{
			r = cudaFail( e, "D2D voxels", __FILE__, __LINE__ );
			break;
		}
		uint32_t launches = 0;
		if( ( r = buildOctreeDevice( ctx, s, o->mortons, n, gridRes, dag, &o->nodes, &o->masks, &o->numberOfNodes, &launches, &o->wide ) ) ) break;
		ctx->stats.uniqueVoxels = n;
		ctx->stats.numberOfNodes = o->numberOfNodes;
		ctx->stats.kernelLaunches = launches;
		if( ( e = cudaStreamSynchronize( s ) ) != cudaSuccess ) r = cudaFail( e, "sync", __FILE__, __LINE__ );
	} while( 0 );
	if( r )
	{
		hako_octree_destroy( o );
		return r;
	}
	*out = o;
	return HAKO_OK;
}

// stage 3, without a staging copy: the octree object first (its resident key / attribute arrays sized for the gathered voxels, to
// be filled in place by the collective), then the nodes
int hako_octree_create_for_voxels( hako_ctx* ctx, hako_stream stream, uint64_t n, const float origin[3], float dps, int gridRes, hako_octree** out,
								   void** mortons_d, void** attrs_d )
{
	CHECK_ARG( ctx && out && origin, "hako_octree_create_for_voxels: null argument" );
	CHECK_ARG( isPow2( gridRes ) && gridRes <= ( 1 << 21 ), "hako_octree_create_for_voxels: gridRes must be a power of two <= 2^21" );
	CHECK_ARG( n < 0xFFFFFFFFull, "hako_octree_create_for_voxels: more than 2^32 - 1 voxels" );
	*out = nullptr;
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	hako_octree* o = new hako_octree();
	initOctree( o, ctx, origin, dps, gridRes );
	o->numberOfVoxels = (uint32_t)n;
	int r = dalloc( ctx, &o->mortons, n, s );
	if( r == 0 ) r = dalloc( ctx, &o->attrs, n, s );
	if( r == 0 )
	{
		cudaError_t e = cudaStreamSynchronize( s ); // the arrays are handed to another library's streams
		if( e != cudaSuccess ) r = cudaFail( e, "hako_octree_create_for_voxels", __FILE__, __LINE__ );
	}
	if( r )
	{
		hako_octree_destroy( o );
		return r;
	}
	if( mortons_d ) *mortons_d = o->mortons;
	if( attrs_d ) *attrs_d = o->attrs;
	*out = o;
	return HAKO_OK;
}
int hako_octree_build_nodes( hako_octree* o, hako_stream stream, int dag, int hasEmission )
{
	CHECK_ARG( o && !o->nodes, "hako_octree_build_nodes: null octree or nodes already built" );
	hako_ctx* ctx = o->ctx;
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	ctx->stats = hako_build_stats{};
	o->hasEmission = hasEmission ? 1u : 0u;
	uint32_t launches = 0;
	HAKO_TRY( buildOctreeDevice( ctx, s, o->mortons, o->numberOfVoxels, o->gridRes, dag, &o->nodes, &o->masks, &o->numberOfNodes, &launches, &o->wide ) );
	ctx->stats.uniqueVoxels = o->numberOfVoxels;
	ctx->stats.numberOfNodes = o->numberOfNodes;
	ctx->stats.kernelLaunches = launches;
	HAKO_CUDA( cudaStreamSynchronize( s ) );
	return HAKO_OK;
}

int hako_voxelize( hako_ctx* ctx, hako_stream stream, const float* pos, const float* col, const float* emi, uint32_t nTriangles,
				   const float origin[3], float dps, int gridRes, int sixSeparating,
				   uint64_t* keys, hako_voxel_attrib* attrs, uint64_t capacity, uint64_t* D )
{
	CHECK_ARG( ctx && origin && D, "hako_voxelize: null argument" );
	CHECK_ARG( nTriangles == 0 || pos, "hako_voxelize: null vertex data" );
	CHECK_ARG( gridRes > 0 && dps > 0.0f, "hako_voxelize: bad grid" );
	CHECK_ARG( !keys || ( col && emi && attrs ), "hako_voxelize: emitting needs colours, emissions and an attribute buffer" );
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	*D = 0;
	const uint64_t nf = (uint64_t)nTriangles * 9;
	float* d = nullptr;
	HAKO_TRY( dalloc( ctx, &d, nf * 3, s ) );
	uint64_t *k_d = nullptr, *v_d = nullptr;
	int r = 0;
	do
	{
		cudaError_t e = cudaSuccess;
		if( nf )
		{
			e = cudaMemcpyAsync( d, pos, nf * 4, cudaMemcpyHostToDevice, s );
			if( e == cudaSuccess && keys ) e = cudaMemcpyAsync( d + nf, col, nf * 4, cudaMemcpyHostToDevice, s );
			if( e == cudaSuccess && keys ) e = cudaMemcpyAsync( d + 2 * nf, emi, nf * 4, cudaMemcpyHostToDevice, s );
		}
		if( e != cudaSuccess )
		{
			r = cudaFail( e, "H2D vertices", __FILE__, __LINE__ );
			break;
		}
		const V3 org = { origin[0], origin[1], origin[2] };
		if( ( r = voxelizeDevice( ctx, s, d, d + nf, d + 2 * nf, nTriangles, org, dps, gridRes, sixSeparating, keys ? &k_d : nullptr, keys ? &v_d : nullptr, D ) ) ) break;
		if( keys && *D )
		{
			if( *D > capacity )
			{
				setError( "hako_voxelize: capacity %llu < %llu dumped voxels", (unsigned long long)capacity, (unsigned long long)*D );
				r = HAKO_E_INVALID;
				break;
			}
			e = cudaMemcpyAsync( keys, k_d, *D * 8, cudaMemcpyDeviceToHost, s );
			if( e == cudaSuccess ) e = cudaMemcpyAsync( attrs, v_d, *D * 8, cudaMemcpyDeviceToHost, s );
			if( e != cudaSuccess )
			{
				r = cudaFail( e, "D2H voxels", __FILE__, __LINE__ );
				break;
			}
		}
		if( ( e = cudaStreamSynchronize( s ) ) != cudaSuccess ) r = cudaFail( e, "sync", __FILE__, __LINE__ );
	} while( 0 );
	dfree( d, s ), dfree( k_d, s ), dfree( v_d, s );
	return r;
}

int hako_sort_pairs_device( hako_ctx* ctx, hako_stream stream, uint64_t* keys_d, uint64_t* values_d, uint64_t n, int beginBit, int endBit )
{
	CHECK_ARG( ctx && ( n == 0 || keys_d ), "hako_sort_pairs_device: null argument" );
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	if( n == 0 ) return HAKO_OK;
	uint64_t *kt = nullptr, *vt = nullptr;
	int r = dalloc( ctx, &kt, n, s );
	if( !r && values_d ) r = dalloc( ctx, &vt, n, s );
	if( !r ) r = sortPairsDevice( ctx, s, keys_d, values_d, kt, vt, n, beginBit, endBit, nullptr );
	dfree( kt, s ), dfree( vt, s );
	return r;
}

int hako_sort_pairs( hako_ctx* ctx, hako_stream stream, uint64_t* keys, uint64_t* values, uint64_t n, int beginBit, int endBit )
{
	CHECK_ARG( ctx && ( n == 0 || keys ), "hako_sort_pairs: null argument" );
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	if( n == 0 ) return HAKO_OK;
	uint64_t *k_d = nullptr, *v_d = nullptr;
	int r = dalloc( ctx, &k_d, n, s );
	if( !r && values ) r = dalloc( ctx, &v_d, n, s );
	do
	{
		if( r ) break;
		cudaError_t e = cudaMemcpyAsync( k_d, keys, n * 8, cudaMemcpyHostToDevice, s );
		if( e == cudaSuccess && values ) e = cudaMemcpyAsync( v_d, values, n * 8, cudaMemcpyHostToDevice, s );
		if( e != cudaSuccess )
		{
			r = cudaFail( e, "H2D", __FILE__, __LINE__ );
			break;
		}
		if( ( r = hako_sort_pairs_device( ctx, stream, k_d, v_d, n, beginBit, endBit ) ) ) break;
		e = cudaMemcpyAsync( keys, k_d, n * 8, cudaMemcpyDeviceToHost, s );
		if( e == cudaSuccess && values ) e = cudaMemcpyAsync( values, v_d, n * 8, cudaMemcpyDeviceToHost, s );
		if( e == cudaSuccess ) e = cudaStreamSynchronize( s );
		if( e != cudaSuccess ) r = cudaFail( e, "D2H", __FILE__, __LINE__ );
	} while( 0 );
	dfree( k_d, s ), dfree( v_d, s );
	return r;
}

int hako_unique_merge( hako_ctx* ctx, hako_stream stream, uint64_t* keys, hako_voxel_attrib* attrs, uint64_t n, uint64_t* U, uint32_t* hasEmission )
{
	CHECK_ARG( ctx && U && ( n == 0 || ( keys && attrs ) ), "hako_unique_merge: null argument" );
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	*U = 0;
	if( hasEmission ) *hasEmission = 0;
	if( n == 0 ) return HAKO_OK;
	uint64_t *k_d = nullptr, *v_d = nullptr, *ko_d = nullptr, *vo_d = nullptr;
	int r = dalloc( ctx, &k_d, n, s );
	if( !r ) r = dalloc( ctx, &v_d, n, s );
	if( !r ) r = dalloc( ctx, &ko_d, n, s );
	if( !r ) r = dalloc( ctx, &vo_d, n, s );
	do
	{
		if( r ) break;
		cudaError_t e = cudaMemcpyAsync( k_d, keys, n * 8, cudaMemcpyHostToDevice, s );
		if( e == cudaSuccess ) e = cudaMemcpyAsync( v_d, attrs, n * 8, cudaMemcpyHostToDevice, s );
		if( e != cudaSuccess )
		{
			r = cudaFail( e, "H2D", __FILE__, __LINE__ );
			break;
		}
		uint32_t he = 0;
		if( ( r = uniqueMergeDevice( ctx, s, k_d, v_d, n, ko_d, (hako_voxel_attrib*)vo_d, U, &he ) ) ) break;
		if( hasEmission ) *hasEmission = he;
		e = cudaMemcpyAsync( keys, ko_d, *U * 8, cudaMemcpyDeviceToHost, s );
		if( e == cudaSuccess ) e = cudaMemcpyAsync( attrs, vo_d, *U * 8, cudaMemcpyDeviceToHost, s );
		if( e == cudaSuccess ) e = cudaStreamSynchronize( s );
		if( e != cudaSuccess ) r = cudaFail( e, "D2H", __FILE__, __LINE__ );
	} while( 0 );
	dfree( k_d, s ), dfree( v_d, s ), dfree( ko_d, s ), dfree( vo_d, s );
	return r;
}
}
