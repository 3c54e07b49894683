// unique.cu -- sorted (key, attribute) pairs -> first key of each run + integer-mean attributes.
//
// == mergeVoxels (voxRT.cpp:14-51): sums of the u8 colour / emission channels over a run divided by the run length in
// integer arithmetic (order-independent, so sort stability is irrelevant); alpha = count / count = 1 as on the CPU path
// (the reference's GPU `unique` kernel, voxKernel.cu:170-243, writes 255 there; it is never read).
// The reference compacts in place through StreamCompaction::filter, whose blocks wait for each other in index order
// (StreamCompaction.hpp:121-138,186-196).  Here: one fused kernel, out of place -- head flags, block scan, decoupled
// look-back for the tile prefix, and the head thread of each run walks its run and writes the merged voxel.
#include <algorithm>

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

__global__ void __launch_bounds__( UQ_BLOCK ) uniqueMergeKernel( const uint64_t* __restrict__ keys, const uint64_t* __restrict__ values, uint64_t n,
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
	const uint64_t prefix = scanTilePrefix( st, tile, total );
	uint64_t out = prefix + excl;
	uint32_t anyEmission = 0;
#pragma unroll
	for( int i = 0; i < UQ_ITEMS; i++ )
	{
		if( !( heads & ( 1u << i ) ) ) continue;
		// walk the run (voxKernel.cu:203-226 does the same serial walk from the first element)
		uint32_t c[3] = { 0, 0, 0 }, e[3] = { 0, 0, 0 }, cnt = 0;
		auto add = [&]( const uint64_t x ) {
			c[0] += (uint32_t)( x ) & 0xFFu;
			c[1] += (uint32_t)( x >> 8 ) & 0xFFu;
			c[2] += (uint32_t)( x >> 16 ) & 0xFFu;
			e[0] += (uint32_t)( x >> 32 ) & 0xFFu;
			e[1] += (uint32_t)( x >> 40 ) & 0xFFu;
			e[2] += (uint32_t)( x >> 48 ) & 0xFFu;
			cnt++;
		};
		// the part of the run inside this thread's items: registers
		bool open = true;
#pragma unroll
		for( int j = 0; j < UQ_ITEMS; j++ )
		{
			if( j < i || !open ) continue;
			if( j > i && ( ( heads >> j ) & 1u || base + j >= n ) )
				open = false;
			else
				add( v[j] );
		}
		// ... and whatever continues in the items of the following threads (voxKernel.cu:203-226 walks the whole run this way)
		if( open )
			for( uint64_t j = base + UQ_ITEMS; j < n; j++ )
			{
				if( keys[j] != k[i] ) break;
				add( values[j] );
			}
		const uint32_t er = e[0] / cnt, eg = e[1] / cnt, eb = e[2] / cnt;
		const uint64_t a = (uint64_t)( c[0] / cnt ) | ( (uint64_t)( c[1] / cnt ) << 8 ) | ( (uint64_t)( c[2] / cnt ) << 16 ) | ( 1ull << 24 ) |
						   ( (uint64_t)er << 32 ) | ( (uint64_t)eg << 40 ) | ( (uint64_t)eb << 48 ) | ( 1ull << 56 );
		anyEmission |= er | eg | eb;
		keysOut[out] = k[i];
		attrsOut[out] = a;
		out++;
	}
	// hasEmission (voxKernel.cu:228-231 atomicExch): one atomic per warp at most
	if( __any_sync( 0xFFFFFFFFu, anyEmission != 0 ) && ( threadIdx.x & 31u ) == 0 ) atomicOr( &counters[1], 1ull );
	if( tile == gridDim.x - 1 && threadIdx.x == 0 ) counters[0] = prefix + total;
}

// number of distinct keys of a sorted array (one read of the keys): lets the caller allocate the exact-size outputs of the merge
__global__ void __launch_bounds__( 256 ) countRunsKernel( const uint64_t* __restrict__ keys, uint64_t n, unsigned long long* __restrict__ count )
{
	uint32_t mine = 0;
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for( uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride ) mine += ( i == 0 || keys[i] != keys[i - 1] ) ? 1u : 0u;
	for( int o = 16; o; o >>= 1 ) mine += __shfl_down_sync( 0xFFFFFFFFu, mine, o );
	if( ( threadIdx.x & 31u ) == 0 && mine ) atomicAdd( count, (unsigned long long)mine );
}
int countRunsDevice( hako_ctx* ctx, cudaStream_t s, const uint64_t* keys, uint64_t n, uint64_t* U )
{
	*U = 0;
	if( n == 0 ) return 0;
	unsigned long long* counter = nullptr;
	HAKO_TRY( dalloc( ctx, &counter, 1, s ) );
	cudaError_t e = cudaMemsetAsync( counter, 0, 8, s );
	if( e == cudaSuccess )
	{
		const uint32_t blocks = (uint32_t)std::min<uint64_t>( (uint64_t)ctx->smCount * 8, ( n + 255 ) / 256 );
		countRunsKernel<<<blocks, 256, 0, s>>>( keys, n, counter );
		ctx->launches.fetch_add( 1, std::memory_order_relaxed );
		e = cudaGetLastError();
	}
	if( e == cudaSuccess ) e = cudaMemcpyAsync( ctx->pinned, counter, 8, cudaMemcpyDeviceToHost, s );
	if( e == cudaSuccess ) e = cudaStreamSynchronize( s );
	ctx->stats.hostSyncs++;
	dfree( counter, s );
	if( e != cudaSuccess ) return cudaFail( e, "countRunsDevice", __FILE__, __LINE__ );
	*U = ctx->pinned[0];
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
	HAKO_CUDA( cudaMemsetAsync( mem, 0, scanStateBytes( tiles ), s ) );
	HAKO_LAUNCH( ctx, uniqueMergeKernel, (uint32_t)tiles, UQ_BLOCK, 0, s, keys, values, n, keysOut, (uint64_t*)attrsOut, makeScanState( mem, tiles ), counters_d );
	cudaFreeAsync( mem, s );
	return 0;
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
