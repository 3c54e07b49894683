// scan.cuh -- single-pass device-wide exclusive scan with decoupled look-back, as a device function that fused
// kernels call once per tile.  It replaces two serialising constructs of the reference:
//   - StreamCompaction::filter's CAS chain on a {prefix, nextBlock} word + granteeBlockExecutionOrder spin
//     (StreamCompaction.hpp:121-138,186-196), which orders blocks strictly one after another;
//   - thrs prefixSumExclusiveInplace's g_iterator chain (tinyhipradixsort.hpp:182-250).
// Here a tile publishes its aggregate immediately and only *reads* its predecessors, so tiles overlap.
//
// Descriptor word (u64): bits 63..62 = status (0 not ready, 1 aggregate, 2 inclusive prefix), bits 61..0 = value.
// Tile ids come from an atomic ticket so that a tile's predecessors are always resident or finished (forward progress).
#pragma once

#include "common.cuh"

namespace hako
{
#define SCAN_FLAG_AGG ( 1ull << 62 )
#define SCAN_FLAG_INC ( 2ull << 62 )
#define SCAN_VALUE_MASK ( ( 1ull << 62 ) - 1 )

struct ScanState
{
	unsigned long long* desc; // [numTiles], zero-initialised
	uint32_t* ticket;		  // zero-initialised
};
// bytes to allocate (and zero) for a scan over `tiles` tiles: descriptors, then the ticket
inline uint64_t scanStateBytes( uint64_t tiles ) { return ( tiles + 2 ) * 8; }
inline ScanState makeScanState( void* mem, uint64_t tiles )
{
	ScanState s;
	s.desc = (unsigned long long*)mem;
	s.ticket = (uint32_t*)( s.desc + tiles + 1 );
	return s;
}

__device__ __forceinline__ uint32_t scanTakeTicket( const ScanState& st )
{
	__shared__ uint32_t s_ticket;
	if( threadIdx.x == 0 ) s_ticket = atomicAdd( st.ticket, 1u );
	__syncthreads();
	return s_ticket;
}

__device__ __forceinline__ unsigned long long ldVolatile( const unsigned long long* p ) { return *(const volatile unsigned long long*)p; }
__device__ __forceinline__ void stVolatile( unsigned long long* p, unsigned long long v ) { *(volatile unsigned long long*)p = v; }

// The look-back itself, for ONE warp (all 32 lanes call it with the tile's total): publishes the tile's aggregate, walks back over the
// predecessors 32 at a time, publishes the inclusive prefix and returns the exclusive one (to every lane).  A kernel whose other warps
// have work that does not need the prefix calls this from warp 0 and synchronises later (uniqueMergeKernel).
__device__ __forceinline__ uint64_t scanTilePrefixWarp( const ScanState& st, uint32_t tile, uint64_t tileTotal )
{
	const uint32_t lane = threadIdx.x & 31u;
	if( tile == 0 )
	{
		if( lane == 0 ) stVolatile( &st.desc[0], SCAN_FLAG_INC | tileTotal );
		return 0;
	}
	if( lane == 0 ) stVolatile( &st.desc[tile], SCAN_FLAG_AGG | tileTotal );
	uint64_t prefix = 0;
	int64_t base = (int64_t)tile - 1; // lane l inspects tile base - l
	for( ;; )
	{
		const int64_t idx = base - lane;
		unsigned long long d = SCAN_FLAG_INC; // virtual tiles before 0: inclusive prefix 0
		uint32_t incMask, firstInc;
		// poll the window until every predecessor UP TO the nearest inclusive one has published something; tiles behind that one are
		// not needed, so a straggler among them must not hold this tile (and through it its successors) up
		for( ;; )
		{
			if( idx >= 0 ) d = ldVolatile( &st.desc[idx] );
			const uint32_t status = (uint32_t)( d >> 62 );
			incMask = __ballot_sync( 0xFFFFFFFFu, status == 2u );
			const uint32_t pendMask = __ballot_sync( 0xFFFFFFFFu, status == 0u );
			firstInc = incMask ? ( __ffs( incMask ) - 1 ) : 32u;
			const uint32_t need = firstInc >= 31u ? 0xFFFFFFFFu : ( ( 2u << firstInc ) - 1u ); // lanes 0 .. firstInc
			if( ( pendMask & need ) == 0u ) break;
		}
		uint64_t v = ( lane <= firstInc ) ? ( d & SCAN_VALUE_MASK ) : 0ull;
#pragma unroll
		for( int o = 16; o > 0; o >>= 1 ) v += __shfl_xor_sync( 0xFFFFFFFFu, v, o );
		prefix += v;
		if( incMask ) break;
		base -= 32;
	}
	if( lane == 0 ) stVolatile( &st.desc[tile], SCAN_FLAG_INC | ( prefix + tileTotal ) );
	return prefix;
}

// Called by ALL threads of the block with the tile's total (valid in thread 0 at least); returns the exclusive prefix of the
// tile (sum of the totals of tiles 0..tile-1) to all threads.  Warp 0 does the look-back.
__device__ __forceinline__ uint64_t scanTilePrefix( const ScanState& st, uint32_t tile, uint64_t tileTotal )
{
	__shared__ unsigned long long s_prefix;
	if( threadIdx.x < 32 )
	{
		tileTotal = __shfl_sync( 0xFFFFFFFFu, tileTotal, 0 );
		const uint64_t prefix = scanTilePrefixWarp( st, tile, tileTotal );
		if( threadIdx.x == 0 ) s_prefix = prefix;
	}
	__syncthreads();
	return s_prefix;
}

// Block-wide exclusive scan of one uint32 per thread (BLOCK threads, BLOCK a multiple of 32, <= 1024).
// Returns the exclusive prefix; *total receives the block sum (all threads).
template <int BLOCK>
__device__ __forceinline__ uint32_t blockExclusiveScan( uint32_t v, uint32_t* total )
{
	__shared__ uint32_t s_warp[BLOCK / 32];
	__shared__ uint32_t s_total;
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	uint32_t inc = v;
#pragma unroll
	for( int o = 1; o < 32; o <<= 1 )
	{
		uint32_t n = __shfl_up_sync( 0xFFFFFFFFu, inc, o );
		if( lane >= o ) inc += n;
	}
	if( lane == 31 ) s_warp[warp] = inc;
	__syncthreads();
	if( warp == 0 )
	{
		uint32_t w = ( lane < BLOCK / 32 ) ? s_warp[lane] : 0u;
		uint32_t winc = w;
#pragma unroll
		for( int o = 1; o < 32; o <<= 1 )
		{
			uint32_t n = __shfl_up_sync( 0xFFFFFFFFu, winc, o );
			if( lane >= o ) winc += n;
		}
		if( lane < BLOCK / 32 ) s_warp[lane] = winc - w;
		if( lane == 31 ) s_total = winc;
	}
	__syncthreads();
	const uint32_t r = s_warp[warp] + inc - v;
	*total = s_total;
	__syncthreads(); // s_warp / s_total may be reused by the next call
	return r;
}
} // namespace hako
