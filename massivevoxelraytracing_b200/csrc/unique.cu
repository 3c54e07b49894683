// unique.cu -- sorted (key, attribute) pairs -> first key of each run + integer-mean attributes.
//
// == mergeVoxels (voxRT.cpp:14-51): sums of the u8 colour / emission channels over a run divided by the run length in
// integer arithmetic (order-independent, so sort stability is irrelevant); alpha = count / count = 1 as on the CPU path
// (the reference's GPU `unique` kernel, voxKernel.cu:170-243, writes 255 there; it is never read).
// The reference compacts in place through StreamCompaction::filter, whose blocks wait for each other in index order
// (StreamCompaction.hpp:121-138,186-196).  Here: one fused kernel, out of place -- head flags, block scan, decoupled
// look-back for the tile prefix, and the head thread of each run walks its run and writes the merged voxel.
#include <algorithm>
#include <mutex>

#include "internal.cuh"
#include "scan.cuh"

namespace hako
{
namespace
{
constexpr int UQ_BLOCK = 256;
constexpr int UQ_ITEMS = 8;
constexpr int UQ_TILE = UQ_BLOCK * UQ_ITEMS;
} // namespace

__global__ void __launch_bounds__( UQ_BLOCK, 4 ) uniqueMergeKernel( const uint64_t* __restrict__ keys, const uint64_t* __restrict__ values, uint64_t n,
																 uint64_t* __restrict__ keysOut, uint64_t* __restrict__ attrsOut, ScanState st,
																 unsigned long long* __restrict__ counters /* [0] = U, [1] = hasEmission */ )
{
	const uint32_t tile = scanTakeTicket( st );
	const uint64_t base = (uint64_t)tile * UQ_TILE + (uint64_t)threadIdx.x * UQ_ITEMS;
	uint64_t k[UQ_ITEMS], v[UQ_ITEMS];
	uint32_t heads = 0, nHeads = 0;
	uint64_t prev = ( base > 0 && base <= n ) ? keys[base - 1] : 0ull;
	// the thread's own values are loaded up front (independent loads); only a run that continues past the thread's last item
	// is walked through memory
	if( base + UQ_ITEMS <= n )
	{
		// 64 contiguous, 64-byte aligned bytes per thread and array: 16-byte loads halve the L1 requests of the strided pattern
#pragma unroll
		for( int i = 0; i < UQ_ITEMS; i += 2 )
		{
			const ulonglong2 kk = *reinterpret_cast<const ulonglong2*>( keys + base + i );
			const ulonglong2 vv = *reinterpret_cast<const ulonglong2*>( values + base + i );
			k[i] = kk.x, k[i + 1] = kk.y, v[i] = vv.x, v[i + 1] = vv.y;
		}
	}
	else
	{
#pragma unroll
		for( int i = 0; i < UQ_ITEMS; i++ )
		{
			k[i] = ( base + i < n ) ? keys[base + i] : 0ull;
			v[i] = ( base + i < n ) ? values[base + i] : 0ull;
		}
	}
#pragma unroll
	for( int i = 0; i < UQ_ITEMS; i++ )
	{
		const uint64_t idx = base + i;
		const bool head = ( idx < n ) && ( idx == 0 || k[i] != prev );
		prev = k[i];
		if( head )
		{
			heads |= 1u << i;
			nHeads++;
		}
	}
	uint32_t total;
	const uint32_t excl = blockExclusiveScan<UQ_BLOCK>( nHeads, &total );
	// warp 0 resolves the tile's global offset while the other warps already merge their runs (which needs only the tile-local slots);
	// everyone meets at the barrier in front of the write-out
	__shared__ unsigned long long s_prefix;
	if( threadIdx.x < 32 )
	{
		const uint64_t pre = scanTilePrefixWarp( st, tile, total );
		if( threadIdx.x == 0 ) s_prefix = pre;
	}
	// The merged voxels of the tile are staged in shared memory in output order and written out as one contiguous, coalesced
	// block (a thread's own outputs are a strided dribble of up to 8 entries).
	__shared__ uint64_t s_outK[UQ_TILE];
	__shared__ uint64_t s_outA[UQ_TILE];
	uint32_t slot = excl;
	uint32_t anyEmission = 0;
	// One pass over the thread's items with running sums: a head closes the run before it (if this thread owns it) and opens
	// the next; items in front of the thread's first head belong to a run of an earlier thread, which walks into them below.
	uint32_t c0 = 0, c1 = 0, c2 = 0, e0 = 0, e1 = 0, e2 = 0, cnt = 0;
	uint64_t kcur = 0;
	bool open = false;
	auto add = [&]( const uint64_t x ) {
		c0 += (uint32_t)( x ) & 0xFFu;
		c1 += (uint32_t)( x >> 8 ) & 0xFFu;
		c2 += (uint32_t)( x >> 16 ) & 0xFFu;
		e0 += (uint32_t)( x >> 32 ) & 0xFFu;
		e1 += (uint32_t)( x >> 40 ) & 0xFFu;
		e2 += (uint32_t)( x >> 48 ) & 0xFFu;
		cnt++;
	};
	auto emit = [&]() {
		// integer means (voxRT.cpp:38-47); a run of one voxel -- the common case -- needs no division
		uint32_t q0 = c0, q1 = c1, q2 = c2, r0 = e0, r1 = e1, r2 = e2;
		if( cnt > 1 )
		{
			if( cnt <= 4096u )
			{
				// six quotients by one divisor: m = ceil( 2^32 / cnt ), q = floor( c * m / 2^32 ).  With e = m * cnt - 2^32 in [0, cnt)
				// the product overshoots c / cnt by c * e / ( cnt * 2^32 ) < 1 / cnt whenever c * cnt < 2^32, and c <= 255 * cnt, so
				// the floor is exact for cnt <= 4 103 (one u32 division per run instead of six: they were ~40 % of this kernel's
				// instructions)
				const uint32_t m = 0xFFFFFFFFu / cnt + 1u;
				q0 = __umulhi( q0, m ), q1 = __umulhi( q1, m ), q2 = __umulhi( q2, m );
				r0 = __umulhi( r0, m ), r1 = __umulhi( r1, m ), r2 = __umulhi( r2, m );
			}
			else
				q0 /= cnt, q1 /= cnt, q2 /= cnt, r0 /= cnt, r1 /= cnt, r2 /= cnt;
		}
		anyEmission |= r0 | r1 | r2;
		s_outK[slot] = kcur;
		s_outA[slot] = (uint64_t)q0 | ( (uint64_t)q1 << 8 ) | ( (uint64_t)q2 << 16 ) | ( 1ull << 24 ) | ( (uint64_t)r0 << 32 ) | ( (uint64_t)r1 << 40 ) | ( (uint64_t)r2 << 48 ) | ( 1ull << 56 );
		slot++;
	};
#pragma unroll
	for( int i = 0; i < UQ_ITEMS; i++ )
	{
		if( ( heads >> i ) & 1u )
		{
			if( open ) emit();
			kcur = k[i];
			c0 = c1 = c2 = e0 = e1 = e2 = cnt = 0;
			open = true;
			add( v[i] );
		}
		else if( open && base + i < n )
			add( v[i] );
	}
	if( open )
	{
		// the thread's last run may continue in the items of the following threads (voxKernel.cu:203-226 walks the whole run this way)
		for( uint64_t j = base + UQ_ITEMS; j < n; j++ )
		{
			if( keys[j] != kcur ) break;
			add( values[j] );
		}
		emit();
	}
	__syncthreads();
	const uint64_t prefix = s_prefix;
	for( uint32_t i = threadIdx.x; i < total; i += UQ_BLOCK )
	{
		keysOut[prefix + i] = s_outK[i];
		attrsOut[prefix + i] = s_outA[i];
	}
	// hasEmission (voxKernel.cu:228-231 atomicExch): one atomic per warp at most
	if( __any_sync( 0xFFFFFFFFu, anyEmission != 0 ) && ( threadIdx.x & 31u ) == 0 ) atomicOr( &counters[1], 1ull );
	if( tile == gridDim.x - 1 && threadIdx.x == 0 ) counters[0] = prefix + total;
}

// One read of the SORTED dumped keys (duplicates still in) gives everything the host must know before the rest of the build can be
// sized: the number of distinct keys U = 1 + #{ i : keys[i] != keys[i-1] } (exact-size outputs of the merge), and for every octree
// level the number of boundaries between neighbouring distinct keys whose highest differing bit lies in that level's digit -- the
// per-level group counts that octreeTaskInit (voxKernel.cu:245-267) gets with one atomicInc per boundary.  Duplicates contribute
// nothing to either, so the histogram is that of the unique list.  counters: [0] = U - 1, [1 + l] = boundaries with top bit in digit l.
__global__ void __launch_bounds__( 256 ) countRunsKernel( const uint64_t* __restrict__ keys, uint64_t n, unsigned long long* __restrict__ counters )
{
	__shared__ uint32_t s_hist[22];
	if( threadIdx.x < 22 ) s_hist[threadIdx.x] = 0;
	__syncthreads();
	uint32_t mine = 0;
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for( uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x + 1; i < n; i += stride )
	{
		const uint64_t x = keys[i] ^ keys[i - 1];
		if( x )
		{
			mine++;
			atomicAdd( &s_hist[( 63 - __clzll( (long long)x ) ) / 3], 1u );
		}
	}
	for( int o = 16; o; o >>= 1 ) mine += __shfl_down_sync( 0xFFFFFFFFu, mine, o );
	if( ( threadIdx.x & 31u ) == 0 && mine ) atomicAdd( &counters[0], (unsigned long long)mine );
	__syncthreads();
	if( threadIdx.x < 22 && s_hist[threadIdx.x] ) atomicAdd( &counters[1 + threadIdx.x], (unsigned long long)s_hist[threadIdx.x] );
}
// *U = distinct keys; levelBoundaries (may be null): 22 counts for buildOctreeDevice.  Synchronises (the one host read of this stage).
int countRunsDevice( hako_ctx* ctx, cudaStream_t s, const uint64_t* keys, uint64_t n, uint64_t* U, uint64_t* levelBoundaries )
{
	*U = 0;
	if( levelBoundaries )
		for( int i = 0; i < 22; i++ ) levelBoundaries[i] = 0;
	if( n == 0 ) return 0;
	unsigned long long* counter = nullptr;
	HAKO_TRY( dalloc( ctx, &counter, 23, s ) );
	cudaError_t e = cudaMemsetAsync( counter, 0, 23 * 8, s );
	if( e == cudaSuccess )
	{
		const uint32_t blocks = (uint32_t)std::min<uint64_t>( (uint64_t)ctx->smCount * 8, ( n + 255 ) / 256 );
		countRunsKernel<<<blocks, 256, 0, s>>>( keys, n, counter );
		ctx->launches.fetch_add( 1, std::memory_order_relaxed );
		e = cudaGetLastError();
	}
	if( e == cudaSuccess ) e = cudaMemcpyAsync( ctx->pinned, counter, 23 * 8, cudaMemcpyDeviceToHost, s );
	if( e == cudaSuccess ) e = cudaStreamSynchronize( s );
	ctx->stats.hostSyncs++;
	dfree( counter, s );
	if( e != cudaSuccess ) return cudaFail( e, "countRunsDevice", __FILE__, __LINE__ );
	*U = ctx->pinned[0] + 1;
	if( levelBoundaries )
		for( int i = 0; i < 22; i++ ) levelBoundaries[i] = ctx->pinned[1 + i];
	return 0;
}

// keysOut / attrsOut must hold n entries (or exactly the number of runs).  counters_d: 2 x u64 (U, hasEmission), written stream-ordered.
int uniqueMergeLaunch( hako_ctx* ctx, cudaStream_t s, const uint64_t* keys, const uint64_t* values, uint64_t n,
					   uint64_t* keysOut, hako_voxel_attrib* attrsOut, unsigned long long* counters_d )
{
	HAKO_CUDA( cudaMemsetAsync( counters_d, 0, 16, s ) );
	if( n == 0 ) return 0;
	const uint64_t tiles = ( n + UQ_TILE - 1 ) / UQ_TILE;
	void* mem = nullptr;
	HAKO_CUDA( cudaMallocFromPoolAsync( &mem, scanStateBytes( tiles ), ctx->pool, s ) );
	int r = 0;
	do
	{
		const cudaError_t e = cudaMemsetAsync( mem, 0, scanStateBytes( tiles ), s );
		if( e != cudaSuccess )
		{
			r = cudaFail( e, "memset scan state", __FILE__, __LINE__ );
			break;
		}
		HAKO_LAUNCH_R( ctx, r, uniqueMergeKernel, (uint32_t)tiles, UQ_BLOCK, 0, s, keys, values, n, keysOut, (uint64_t*)attrsOut, makeScanState( mem, tiles ), counters_d );
	} while( 0 );
	cudaFreeAsync( mem, s ); // also when the launch failed
	return r;
}

int uniqueMergeDevice( hako_ctx* ctx, cudaStream_t s, const uint64_t* keys, const uint64_t* values, uint64_t n,
					   uint64_t* keysOut, hako_voxel_attrib* attrsOut, uint64_t* U, uint32_t* hasEmission )
{
	unsigned long long* counters = nullptr;
	HAKO_TRY( dalloc( ctx, &counters, 2, s ) );
	int r = uniqueMergeLaunch( ctx, s, keys, values, n, keysOut, attrsOut, counters );
	if( r == 0 )
	{
		cudaError_t e = cudaMemcpyAsync( ctx->pinned, counters, 16, cudaMemcpyDeviceToHost, s );
		if( e == cudaSuccess ) e = cudaStreamSynchronize( s );
		ctx->stats.hostSyncs++;
		if( e != cudaSuccess )
			r = cudaFail( e, "read unique count", __FILE__, __LINE__ );
		else
		{
			*U = ctx->pinned[0];
			*hasEmission = ctx->pinned[1] != 0;
		}
	}
	dfree( counters, s );
	return r;
}
} // namespace hako

// ---- hako_unique_merge_noalloc: the merge for callers that may not allocate (CUDA graph capture, serving loops) --------------------
// No counterpart in the reference.  Same kernel (uniqueMergeKernel) as every other merge
// of the library; the only scratch it needs -- the descriptors and the ticket of its decoupled look-back scan -- lives in a file-scope
// device array instead of the context's memory pool.  Every GPU that loads the library holds these 64.3 KB until the process exits.
// ONE array per device: all calls share it, it keeps the last call's contents (cleared here, stream-ordered, before every launch), and
// two calls that overlap on the GPU (different streams, no dependency between them) race on it.
namespace
{
constexpr uint64_t UQ_NOALLOC_MAX_TILES = ( HAKO_UNIQUE_NOALLOC_MAX_ITEMS + hako::UQ_TILE - 1 ) / hako::UQ_TILE;
constexpr uint64_t UQ_NOALLOC_BYTES = ( ( UQ_NOALLOC_MAX_TILES + 2 ) * 8 + 255 ) & ~255ull; // scanStateBytes( max tiles ), rounded up
static __device__ __align__( 256 ) unsigned char g_uniqueNoAllocScratch[UQ_NOALLOC_BYTES];

// In host code the array's name is a host-side stand-in, not a device address: the address comes from cudaGetSymbolAddress, once per
// device, with that device current.
int uniqueNoAllocScratch( int device, void** out )
{
	static void* addr[64] = {};
	static std::mutex lock;
	if( device < 0 || device >= 64 )
	{
		hako::setError( "hako_unique_merge_noalloc: device index %d out of range", device );
		return HAKO_E_INVALID;
	}
	std::lock_guard<std::mutex> g( lock );
	if( !addr[device] )
	{
		void* p = nullptr;
		const cudaError_t e = cudaGetSymbolAddress( &p, g_uniqueNoAllocScratch );
		if( e != cudaSuccess ) return hako::cudaFail( e, "cudaGetSymbolAddress( g_uniqueNoAllocScratch )", __FILE__, __LINE__ );
		addr[device] = p;
	}
	*out = addr[device];
	return HAKO_OK;
}
} // namespace

extern "C" int hako_unique_merge_noalloc( hako_ctx* ctx, hako_stream stream, const uint64_t* keys_d, const hako_voxel_attrib* attrs_d, uint64_t n,
										  uint64_t* keysOut_d, hako_voxel_attrib* attrsOut_d, uint64_t* counts_d )
{
	if( !ctx || !counts_d || ( n && ( !keys_d || !attrs_d || !keysOut_d || !attrsOut_d ) ) )
	{
		hako::setError( "hako_unique_merge_noalloc: null argument" );
		return HAKO_E_INVALID;
	}
	if( n > HAKO_UNIQUE_NOALLOC_MAX_ITEMS )
	{
		hako::setError( "hako_unique_merge_noalloc: %llu items exceed the static scratch (HAKO_UNIQUE_NOALLOC_MAX_ITEMS = %llu); use hako_sort_unique_device",
						(unsigned long long)n, (unsigned long long)HAKO_UNIQUE_NOALLOC_MAX_ITEMS );
		return HAKO_E_LIMIT;
	}
	if( ( ( (uintptr_t)keys_d | (uintptr_t)attrs_d ) & 15u ) || ( ( (uintptr_t)keysOut_d | (uintptr_t)attrsOut_d | (uintptr_t)counts_d ) & 7u ) )
	{
		hako::setError( "hako_unique_merge_noalloc: keys_d / attrs_d must be 16-byte aligned, the outputs 8-byte aligned" );
		return HAKO_E_INVALID;
	}
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	void* scratch = nullptr;
	HAKO_TRY( uniqueNoAllocScratch( ctx->device, &scratch ) );
	HAKO_CUDA( cudaMemsetAsync( counts_d, 0, 16, s ) ); // [0] = U, [1] = hasEmission
	if( n == 0 ) return HAKO_OK;
	const uint64_t tiles = ( n + hako::UQ_TILE - 1 ) / hako::UQ_TILE;
	HAKO_CUDA( cudaMemsetAsync( scratch, 0, hako::scanStateBytes( tiles ), s ) ); // the array keeps the last call's descriptors and ticket
	HAKO_LAUNCH( ctx, hako::uniqueMergeKernel, (uint32_t)tiles, hako::UQ_BLOCK, 0, s, keys_d, (const uint64_t*)attrs_d, n, keysOut_d, (uint64_t*)attrsOut_d,
				 hako::makeScanState( scratch, tiles ), (unsigned long long*)counts_d );
	return HAKO_OK;
}
