// sort.cu -- stable LSD radix sort of (u64 key, u64 value) pairs on a bit window, "onesweep" style.
//
// Stands in for thrs::RadixSort::sortPairs / sortKeys (tinyhipradixsort.hpp:849-944), which runs three kernels per
// 8-bit digit (blockCount :119-149, a chained prefix scan over 256 x nBlocks counters :182-250, reorderKeyPair :252-475)
// and therefore reads the keys twice per pass.  Here:
//   - digits are 8 or 9 bits wide, whichever needs fewer passes for the bit window (36 bits at 4096^3: 4 x 9 instead of 5 x 8;
//     33 bits at 2048^3: 4 instead of 5; 42 / 45 bits at 2^14 / 2^15: 5 instead of 6);
//   - ONE histogram kernel reads the keys once and produces the digit histograms of all passes;
//   - each pass is ONE kernel: a tile ranks its keys (warp match + popc, stable), publishes its per-digit counts,
//     resolves its global offsets by decoupled look-back over the preceding tiles, and scatters through shared
//     memory so that global stores are runs of consecutive addresses.
// Algorithmic traffic: n * (8 + P * 32) bytes for pairs, vs n * P * (8 + 32) for the reference.
#include <algorithm>
#include <cstdlib>
#include <type_traits>

#include "internal.cuh"

namespace hako
{
namespace
{
constexpr int SORT_BLOCK = 256;
constexpr int SORT_WARPS = SORT_BLOCK / 32;
constexpr int SORT_ITEMS = 16;
constexpr int SORT_TILE = SORT_BLOCK * SORT_ITEMS; // 4096 pairs
constexpr int MAX_PASSES = 8;

#ifndef HAKO_SORT_PREFETCH
#define HAKO_SORT_PREFETCH 1
#endif
#define LB_FLAG_AGG ( 1ull << 62 )
#define LB_FLAG_INC ( 2ull << 62 )
#define LB_VALUE_MASK ( ( 1ull << 62 ) - 1 )
#ifndef LB_BACKOFF_NS
#define LB_BACKOFF_NS 0 // sweep 0..1000 ns: no effect (profiles/r2n_sort_backoff.log) -- the look-back waits on round trips, not on unpublished tiles
#endif
#ifndef HAKO_SORT_LB
#define HAKO_SORT_LB 2 // predecessors fetched per look-back round trip, both digits of a thread side by side: 1 / 2 / 4 / 8 = 4.14 / 4.01 / 4.31 / 5.69 ms (profiles/r2p_sort_lookback.log): deeper batches over-fetch 4 KB descriptors
#endif

__device__ __forceinline__ uint32_t digitOf( uint64_t key, int shift, uint32_t mask ) { return (uint32_t)( key >> shift ) & mask; }

// ---- histograms of every pass in one read of the keys ----
template <int RBITS>
__global__ void __launch_bounds__( SORT_BLOCK ) sortHistogramKernel( const uint64_t* __restrict__ keys, uint64_t n, int beginBit, int endBit, int passes,
																	 unsigned long long* __restrict__ hist /* [passes][RADIX] */ )
{
	constexpr int RADIX = 1 << RBITS;
	__shared__ uint32_t s_hist[MAX_PASSES * RADIX];
	for( int i = threadIdx.x; i < passes * RADIX; i += SORT_BLOCK ) s_hist[i] = 0;
	__syncthreads();
	const uint64_t stride = (uint64_t)gridDim.x * SORT_BLOCK;
	for( uint64_t i = (uint64_t)blockIdx.x * SORT_BLOCK + threadIdx.x; i < n; i += stride )
	{
		const uint64_t k = keys[i];
		for( int p = 0; p < passes; p++ )
		{
			const int shift = beginBit + p * RBITS;
			const int bits = min( RBITS, endBit - shift );
			atomicAdd( &s_hist[p * RADIX + digitOf( k, shift, ( 1u << bits ) - 1u )], 1u );
		}
	}
	__syncthreads();
	for( int i = threadIdx.x; i < passes * RADIX; i += SORT_BLOCK )
	{
		const uint32_t c = s_hist[i];
		if( c ) atomicAdd( &hist[i], (unsigned long long)c );
	}
}
// exclusive scan of each pass's bins (one block of RADIX threads per pass)
template <int RBITS>
__global__ void __launch_bounds__( 1 << RBITS ) sortScanHistogramKernel( unsigned long long* __restrict__ hist )
{
	constexpr int RADIX = 1 << RBITS;
	__shared__ unsigned long long s[RADIX];
	unsigned long long* h = hist + (size_t)blockIdx.x * RADIX;
	const unsigned long long v = h[threadIdx.x];
	s[threadIdx.x] = v;
	__syncthreads();
	for( int o = 1; o < RADIX; o <<= 1 )
	{
		unsigned long long t = ( threadIdx.x >= o ) ? s[threadIdx.x - o] : 0ull;
		__syncthreads();
		s[threadIdx.x] += t;
		__syncthreads();
	}
	h[threadIdx.x] = s[threadIdx.x] - v;
}

// ---- one pass ----
template <int RBITS, bool PAIRS, int MIN_CTAS>
__global__ void __launch_bounds__( SORT_BLOCK, MIN_CTAS ) onesweepKernel( const uint64_t* __restrict__ keysIn, const uint64_t* __restrict__ valuesIn,
															   uint64_t* __restrict__ keysOut, uint64_t* __restrict__ valuesOut, uint64_t n,
															   int shift, uint32_t mask, const unsigned long long* __restrict__ digitBase /* [RADIX] exclusive */,
															   unsigned long long* __restrict__ lookback /* [tiles][RADIX], zeroed */, uint32_t* __restrict__ ticket )
{
	constexpr int RADIX = 1 << RBITS;
	constexpr int DPT = RADIX / SORT_BLOCK; // digits per thread in the per-digit phase: thread t owns digits [t * DPT, (t + 1) * DPT)
	__shared__ uint64_t s_data[SORT_TILE];				   // 32 KB: keys, then values, in sorted-by-digit order
	__shared__ uint16_t s_warpHist[SORT_WARPS][RADIX];	   // 4 / 8 KB: per-warp digit counts (<= 512) -> per-warp exclusive offsets (<= 4096)
	__shared__ uint32_t s_tileBase[RADIX];				   // exclusive offset of each digit inside the sorted tile
	__shared__ unsigned long long s_globalBase[RADIX];	   // global output index of the tile's first key of each digit
	__shared__ uint32_t s_tile;

	if( threadIdx.x == 0 ) s_tile = atomicAdd( ticket, 1u );
	for( int i = threadIdx.x; i < SORT_WARPS * RADIX; i += SORT_BLOCK ) ( &s_warpHist[0][0] )[i] = 0;
	__syncthreads();
	const uint32_t tile = s_tile;
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	const uint64_t tileStart = (uint64_t)tile * SORT_TILE;
	const uint32_t tileCount = (uint32_t)min( (uint64_t)SORT_TILE, n - tileStart );

	// warp-striped arrangement: warp w owns items [w * 512, (w + 1) * 512) of the tile, item i of lane l is w * 512 + i * 32 + l.
	// Within a warp the index order (i, l) is the input order, across warps the warp order is -- ranking in that order is stable.
	uint64_t key[SORT_ITEMS];
	uint16_t rank[SORT_ITEMS]; // position inside the sorted tile
	const uint32_t warpBase = warp * ( SORT_ITEMS * 32 );
#pragma unroll
	for( int i = 0; i < SORT_ITEMS; i++ )
	{
		const uint32_t idx = warpBase + i * 32 + lane;
		key[i] = ( idx < tileCount ) ? keysIn[tileStart + idx] : ~0ull;
	}
	// The values are needed only after the keys have left (one shared buffer, 64 registers): ask L2 for their lines now, so that
	// the loads further down find them there instead of queueing behind the scatter traffic in DRAM (profiles/r1q: 30 % of the
	// warp-stall samples waited for those loads): 4.46 -> 4.18 ms at the bench scene.  (The same for the KEYS of a tile 592
	// ahead changed nothing.)  One prefetch per 128-byte line = lanes 0 and 16 of every item row.
	if( PAIRS && HAKO_SORT_PREFETCH && ( lane & 15u ) == 0 )
	{
#pragma unroll
		for( int i = 0; i < SORT_ITEMS; i++ )
		{
			const uint32_t idx = warpBase + i * 32 + lane;
			if( idx < tileCount ) asm volatile( "prefetch.global.L2 [%0];" ::"l"( valuesIn + tileStart + idx ) );
		}
	}
	// rank inside the warp: keys with the same digit get consecutive ranks in (i, lane) order
#pragma unroll
	for( int i = 0; i < SORT_ITEMS; i++ )
	{
		const uint32_t idx = warpBase + i * 32 + lane;
		const bool valid = idx < tileCount;
		// invalid (padding) items sit behind the tile's last key; they get their own match class and are never written
		const uint32_t d = valid ? digitOf( key[i], shift, mask ) : ( RADIX - 1 );
		const uint32_t peers = __match_any_sync( 0xFFFFFFFFu, d | ( valid ? 0u : 0x10000u ) );
		const uint32_t before = __popc( peers & ( ( 1u << lane ) - 1u ) );
		uint32_t base = 0;
		if( valid ) base = s_warpHist[warp][d];
		__syncwarp();
		if( valid && before == 0 ) s_warpHist[warp][d] = (uint16_t)( base + __popc( peers ) );
		__syncwarp();
		rank[i] = (uint16_t)( base + before );
	}
	__syncthreads();

	// per digit (thread t owns DPT consecutive digits): exclusive scan over the warps, tile total, look-back
	{
		uint32_t total[DPT];
		unsigned long long excl[DPT];
		uint32_t mine = 0;
#pragma unroll
		for( int j = 0; j < DPT; j++ )
		{
			const uint32_t d = threadIdx.x * DPT + j;
			uint32_t tot = 0;
#pragma unroll
			for( int w = 0; w < SORT_WARPS; w++ )
			{
				const uint32_t c = s_warpHist[w][d];
				s_warpHist[w][d] = (uint16_t)tot;
				tot += c;
			}
			total[j] = tot;
			mine += tot;
			// publish this tile's count for digit d (the predecessors are summed below, after every digit is published)
			unsigned long long* my = lookback + (size_t)tile * RADIX + d;
			*(volatile unsigned long long*)my = ( tile == 0 ? LB_FLAG_INC : LB_FLAG_AGG ) | (unsigned long long)tot;
		}
#pragma unroll
		for( int j = 0; j < DPT; j++ )
		{
			const uint32_t d = threadIdx.x * DPT + j;
			excl[j] = 0;
			if( tile != 0 )
			{
				int64_t t = (int64_t)tile - 1;
				for( ;; )
				{
					unsigned long long v;
					do
					{
						v = *(const volatile unsigned long long*)( lookback + (size_t)t * RADIX + d );
					} while( ( v >> 62 ) == 0 );
					excl[j] += v & LB_VALUE_MASK;
					if( ( v >> 62 ) == 2 ) break;
					t--;
				}
				*(volatile unsigned long long*)( lookback + (size_t)tile * RADIX + d ) = LB_FLAG_INC | ( excl[j] + total[j] );
			}
		}
		// block-wide exclusive scan of the tile totals over the digits (thread-local run of DPT digits, then over threads)
		uint32_t inc = mine;
#pragma unroll
		for( int o = 1; o < 32; o <<= 1 )
		{
			const uint32_t nb = __shfl_up_sync( 0xFFFFFFFFu, inc, o );
			if( lane >= o ) inc += nb;
		}
		__shared__ uint32_t s_warpTotals[SORT_WARPS];
		if( lane == 31 ) s_warpTotals[warp] = inc;
		__syncthreads();
		uint32_t wbase = 0;
#pragma unroll
		for( int w = 0; w < SORT_WARPS; w++ )
			if( w < (int)warp ) wbase += s_warpTotals[w];
		uint32_t tb = wbase + inc - mine;
#pragma unroll
		for( int j = 0; j < DPT; j++ )
		{
			const uint32_t d = threadIdx.x * DPT + j;
			s_tileBase[d] = tb;
			s_globalBase[d] = digitBase[d] + excl[j] - tb; // so that output index = s_globalBase[digit] + position in sorted tile
			tb += total[j];
		}
	}
	__syncthreads();

	// scatter keys into the sorted tile
#pragma unroll
	for( int i = 0; i < SORT_ITEMS; i++ )
	{
		const uint32_t idx = warpBase + i * 32 + lane;
		if( idx < tileCount )
		{
			const uint32_t d = digitOf( key[i], shift, mask );
			const uint32_t p = s_tileBase[d] + s_warpHist[warp][d] + rank[i];
			rank[i] = (uint16_t)p;
			s_data[p] = key[i];
		}
	}
	__syncthreads();
	// coalesced write-out; remember each position's destination for the values
	uint64_t dst[SORT_ITEMS];
#pragma unroll
	for( int i = 0; i < SORT_ITEMS; i++ )
	{
		const uint32_t p = i * SORT_BLOCK + threadIdx.x;
		if( p < tileCount )
		{
			const uint64_t k = s_data[p];
			dst[i] = s_globalBase[digitOf( k, shift, mask )] + p;
			keysOut[dst[i]] = k;
		}
	}
	if( PAIRS )
	{
		__syncthreads();
#pragma unroll
		for( int i = 0; i < SORT_ITEMS; i++ )
		{
			const uint32_t idx = warpBase + i * 32 + lane;
			if( idx < tileCount ) s_data[rank[i]] = valuesIn[tileStart + idx];
		}
		__syncthreads();
#pragma unroll
		for( int i = 0; i < SORT_ITEMS; i++ )
		{
			const uint32_t p = i * SORT_BLOCK + threadIdx.x;
			if( p < tileCount ) valuesOut[dst[i]] = s_data[p];
		}
	}
}
// ---- one pass, second generation: the tile's keys AND values are fetched by two bulk-async copies (cp.async.bulk, the 1-D TMA
// path) issued by one thread at kernel entry, each into its own shared buffer and each completing on its own mbarrier.  The value
// tile therefore arrives while the keys are ranked and the look-back runs -- in the first-generation kernel the value loads could
// only be issued after the key tile had left the single shared buffer, and 30 % of the warp-stall samples waited for them
// (profiles/r1q_onesweep_stalls.txt).  Ranking, look-back and the two-step scatter are unchanged, so the result is the same
// stable order.  Needs 16-byte aligned arrays (the bulk copy's granularity); sortPairsDevice falls back otherwise.
__device__ __forceinline__ uint32_t smemAddr( const void* p ) { return (uint32_t)__cvta_generic_to_shared( p ); }
__device__ __forceinline__ void mbarInit( uint64_t* bar, uint32_t count )
{
	asm volatile( "mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"( smemAddr( bar ) ), "r"( count ) );
}
__device__ __forceinline__ void mbarExpectTx( uint64_t* bar, uint32_t bytes )
{
	asm volatile( "mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"( smemAddr( bar ) ), "r"( bytes ) : "memory" );
}
__device__ __forceinline__ void bulkLoad( void* dstSmem, const void* srcGlobal, uint32_t bytes, uint64_t* bar )
{
	asm volatile( "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"( smemAddr( dstSmem ) ), "l"( srcGlobal ), "r"( bytes ),
				  "r"( smemAddr( bar ) )
				  : "memory" );
}
__device__ __forceinline__ void mbarWait( uint64_t* bar, uint32_t parity )
{
	asm volatile(
		"{\n"
		".reg .pred p;\n"
		"WAIT_%=:\n"
		"mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
		"@p bra DONE_%=;\n"
		"bra WAIT_%=;\n"
		"DONE_%=:\n"
		"}\n" ::"r"( smemAddr( bar ) ),
		"r"( parity )
		: "memory" );
}

#ifdef HAKO_SORT_PROFILE
// phase clocks of the bulk pass, summed over tiles by thread 0 (tools/sort_phase_probe.py; never compiled into the shipped library)
__device__ unsigned long long g_sortProf[16];
#define SORT_PROF_MARK( slot )                                                  \
	do                                                                          \
	{                                                                           \
		if( threadIdx.x == 0 )                                                  \
		{                                                                       \
			const long long now_ = clock64();                                   \
			atomicAdd( &g_sortProf[slot], (unsigned long long)( now_ - prof_t ) ); \
			prof_t = now_;                                                      \
		}                                                                       \
	} while( 0 )
#else
#define SORT_PROF_MARK( slot ) \
	do                         \
	{                          \
	} while( 0 )
#endif

template <int RBITS, int BLOCK, int ITEMS>
struct SortV2Smem
{
	static constexpr int RADIX = 1 << RBITS;
	static constexpr int WARPS = BLOCK / 32;
	static constexpr int TILE = BLOCK * ITEMS;
	static constexpr size_t OFF_KEYS = 0;								 // u64[TILE]: key tile, later the sorted staging buffer
	static constexpr size_t OFF_VALUES = OFF_KEYS + (size_t)TILE * 8;	 // u64[TILE]: value tile in input order
	static constexpr size_t OFF_GBASE = OFF_VALUES + (size_t)TILE * 8;	 // u64[RADIX]
	static constexpr size_t OFF_TBASE = OFF_GBASE + (size_t)RADIX * 8;	 // u32[RADIX]
	static constexpr size_t OFF_WHIST = OFF_TBASE + (size_t)RADIX * 4;	 // u16[WARPS][RADIX]
	static constexpr size_t OFF_WTOT = OFF_WHIST + (size_t)WARPS * RADIX * 2; // u32[WARPS]
	static constexpr size_t OFF_BAR = ( OFF_WTOT + (size_t)WARPS * 4 + 15 ) & ~(size_t)15; // 2 mbarriers + the tile index
	static constexpr size_t BYTES = OFF_BAR + 32;
};

// D32: 32-bit look-back descriptors (2 flag bits + 30 value bits; the launcher picks them when n < 2^30): half the descriptor traffic.
// RANK bit 1: the lanes that share a digit come from one ballot per digit bit instead of MATCH.ANY; bit 0: the class leader bumps its
// 16-bit counter with a plain load / add / store instead of a packed shared-memory atomic.  72.7 M pairs x 36 bits on the B200
// (profiles/r2s..r2u_sort_knobs.log, tools/sort_phase_probe.py): 64-bit + match + atomic 4.00 ms, D32 3.82, D32 + ballots + plain 3.43;
// LB = 4 with D32 3.53-3.78; MATCH.ANY for all 15 rows up front (so that their latencies overlap), then the counter chain: 4.00 -- the
// instruction itself is the cost (profiles/r3e_rank1.log).  A tile's life is then ~35 000 cycles: ranking 9 600, look-back 11 400 (8.6 round trips, 2 polls of
// descriptors not yet published), the data movement itself (tile in, two shared-memory scatters, two write-outs) the remaining 40 %.
template <int RBITS, bool PAIRS, int BLOCK, int ITEMS, int MIN_CTAS, bool D32 = false, int LB = HAKO_SORT_LB, int RANK = 0>
__global__ void __launch_bounds__( BLOCK, MIN_CTAS ) onesweepBulkKernel( const uint64_t* __restrict__ keysIn, const uint64_t* __restrict__ valuesIn,
																 uint64_t* __restrict__ keysOut, uint64_t* __restrict__ valuesOut, uint64_t n,
																 int shift, uint32_t mask, const unsigned long long* __restrict__ digitBase /* [RADIX] exclusive */,
																 void* __restrict__ lookbackRaw /* [tiles][RADIX] descriptors, zeroed */, uint32_t* __restrict__ ticket, uint32_t backoffNs )
{
	using DescT = typename std::conditional<D32, uint32_t, unsigned long long>::type;
	constexpr int FSH = D32 ? 30 : 62;
	constexpr DescT F_AGG = (DescT)1 << FSH, F_INC = (DescT)2 << FSH, V_MASK = ( (DescT)1 << FSH ) - 1;
	DescT* __restrict__ lookback = (DescT*)lookbackRaw;
	using L = SortV2Smem<RBITS, BLOCK, ITEMS>;
	constexpr int RADIX = L::RADIX, WARPS = L::WARPS, TILE = L::TILE;
	constexpr int DPT = ( RADIX + BLOCK - 1 ) / BLOCK; // digits per thread in the per-digit phase (threads beyond RADIX / DPT idle there)
	extern __shared__ __align__( 128 ) unsigned char s_raw[];
	uint64_t* s_keys = (uint64_t*)( s_raw + L::OFF_KEYS );
	uint64_t* s_values = (uint64_t*)( s_raw + L::OFF_VALUES );
	unsigned long long* s_globalBase = (unsigned long long*)( s_raw + L::OFF_GBASE );
	uint32_t* s_tileBase = (uint32_t*)( s_raw + L::OFF_TBASE );
	uint16_t* s_warpHist = (uint16_t*)( s_raw + L::OFF_WHIST ); // [WARPS][RADIX]
	uint32_t* s_warpTotals = (uint32_t*)( s_raw + L::OFF_WTOT );
	uint64_t* s_bar = (uint64_t*)( s_raw + L::OFF_BAR );
	uint32_t* s_tile = (uint32_t*)( s_bar + 2 );

#ifdef HAKO_SORT_PROFILE
	long long prof_t = clock64();
	uint32_t prof_rounds = 0, prof_spins = 0;
#endif
	if( threadIdx.x == 0 )
	{
		mbarInit( &s_bar[0], 1 );
		mbarInit( &s_bar[1], 1 );
		asm volatile( "fence.mbarrier_init.release.cluster;" ::: "memory" );
		const uint32_t t = atomicAdd( ticket, 1u );
		*s_tile = t;
		const uint64_t start = (uint64_t)t * TILE;
		const uint32_t cnt = (uint32_t)min( (uint64_t)TILE, n - start );
		const uint32_t bytes = ( cnt * 8u ) & ~15u; // whole 16-byte units; an odd last key is fetched below
		if( bytes )
		{
			mbarExpectTx( &s_bar[0], bytes );
			bulkLoad( s_keys, keysIn + start, bytes, &s_bar[0] );
			if( PAIRS )
			{
				mbarExpectTx( &s_bar[1], bytes );
				bulkLoad( s_values, valuesIn + start, bytes, &s_bar[1] );
			}
		}
	}
	for( int i = threadIdx.x; i < WARPS * RADIX / 2; i += BLOCK ) ( (uint32_t*)s_warpHist )[i] = 0;
	__syncthreads();
	const uint32_t tile = *s_tile;
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	const uint64_t tileStart = (uint64_t)tile * TILE;
	const uint32_t tileCount = (uint32_t)min( (uint64_t)TILE, n - tileStart );
	const bool bulk = ( ( tileCount * 8u ) & ~15u ) != 0;
	const bool oddTail = ( tileCount & 1u ) != 0; // the last pair of an odd tile does not fill a 16-byte unit

	// warp-striped arrangement: warp w owns items [w * ITEMS * 32, (w + 1) * ITEMS * 32), item i of lane l is w * ITEMS * 32 + i * 32 + l
	uint64_t key[ITEMS];
	uint16_t rank[ITEMS];
	const uint32_t warpBase = warp * ( ITEMS * 32 );
	SORT_PROF_MARK( 0 ); // ticket + barrier set-up + histogram clear
	if( bulk ) mbarWait( &s_bar[0], 0 );
	SORT_PROF_MARK( 1 ); // key tile arrived
#pragma unroll
	for( int i = 0; i < ITEMS; i++ )
	{
		const uint32_t idx = warpBase + i * 32 + lane;
		uint64_t k = ~0ull;
		if( idx < tileCount ) k = ( oddTail && idx == tileCount - 1 ) ? keysIn[tileStart + idx] : s_keys[idx];
		key[i] = k;
	}
	uint16_t* myHist = s_warpHist + warp * RADIX;
	uint32_t* myHist32 = (uint32_t*)myHist; // two 16-bit counters per word (a warp holds at most ITEMS * 32 <= 512 keys of a digit)
	SORT_PROF_MARK( 11 ); // keys shared -> registers
#pragma unroll
	for( int i = 0; i < ITEMS; i++ )
	{
		const uint32_t idx = warpBase + i * 32 + lane;
		const bool valid = idx < tileCount;
		// invalid (padding) items sit behind the tile's last key; they get their own match class and are never written
		const uint32_t d = valid ? digitOf( key[i], shift, mask ) : ( RADIX - 1 );
		uint32_t peers;
		if( RANK & 2 )
		{
			// the lanes that hold the same digit, from one ballot per digit bit instead of match.any
			peers = __ballot_sync( 0xFFFFFFFFu, valid );
			peers = valid ? peers : ~peers;
#pragma unroll
			for( int b = 0; b < RBITS; b++ )
			{
				const bool bit = ( d >> b ) & 1u;
				const uint32_t m = __ballot_sync( 0xFFFFFFFFu, bit );
				peers &= bit ? m : ~m;
			}
		}
		else
			peers = __match_any_sync( 0xFFFFFFFFu, d | ( valid ? 0u : 0x10000u ) );
		const uint32_t before = __popc( peers & ( ( 1u << lane ) - 1u ) );
		const int leader = __ffs( peers ) - 1;
		// the class leader reserves the class's ranks with ONE shared-memory atomic that returns the running count: no
		// load -> add -> store chain through shared memory per item, so the 15 / 16 items of a thread pipeline
		const uint32_t sh = ( d & 1u ) * 16u;
		uint32_t old = 0;
		if( !( RANK & 1 ) )
		{
			if( valid && (int)lane == leader ) old = atomicAdd( &myHist32[d >> 1], (uint32_t)__popc( peers ) << sh );
			__syncwarp();
			old = __shfl_sync( 0xFFFFFFFFu, old, leader );
			rank[i] = (uint16_t)( ( ( old >> sh ) & 0xFFFFu ) + before );
		}
		else
		{
			if( valid && (int)lane == leader )
			{
				old = myHist[d];
				myHist[d] = (uint16_t)( old + __popc( peers ) );
			}
			__syncwarp();
			old = __shfl_sync( 0xFFFFFFFFu, old, leader );
			rank[i] = (uint16_t)( old + before );
		}
	}
	SORT_PROF_MARK( 15 ); // thread 0's warp ranked
	__syncthreads(); // every key is in registers: s_keys becomes the sorted staging buffer
	SORT_PROF_MARK( 2 ); // every warp ranked

	{
		uint32_t total[DPT];
		unsigned long long excl[DPT];
		uint32_t mine = 0;
		const bool owner = threadIdx.x * DPT < RADIX;
#pragma unroll
		for( int j = 0; j < DPT; j++ )
		{
			const uint32_t d = threadIdx.x * DPT + j;
			total[j] = 0;
			excl[j] = 0;
			if( !owner ) continue;
			uint32_t tot = 0;
#pragma unroll
			for( int w = 0; w < WARPS; w++ )
			{
				const uint32_t c = s_warpHist[w * RADIX + d];
				s_warpHist[w * RADIX + d] = (uint16_t)tot;
				tot += c;
			}
			total[j] = tot;
			mine += tot;
			*(volatile DescT*)( lookback + (size_t)tile * RADIX + d ) = ( tile == 0 ? F_INC : F_AGG ) | (DescT)tot;
		}
		SORT_PROF_MARK( 3 ); // column sums + aggregate published
		// Decoupled look-back, LB predecessors per round trip and the thread's DPT digits side by side: the loads of a round are
		// independent, so a chain of AGGREGATE descriptors costs one L2 latency per LB tiles instead of one per tile and digit
		// (descriptors behind the first INCLUSIVE one are fetched for nothing).  HAKO_SORT_LOOKBACK sweep: profiles/r2o, r2p.
#ifdef HAKO_SORT_PROFILE
		// timing experiment only (WRONG output): no look-back at all, a made-up prefix that keeps the write addresses spread like the real ones
		if( backoffNs == 0xFFFFFFFFu )
		{
#pragma unroll
			for( int j = 0; j < DPT; j++ ) excl[j] = (unsigned long long)tile * ( TILE / RADIX );
		}
		else
#endif
		if( owner && tile != 0 )
		{
			bool done[DPT];
#pragma unroll
			for( int j = 0; j < DPT; j++ ) done[j] = false;
			int64_t t = (int64_t)tile - 1;
			bool all = false;
			while( !all )
			{
				DescT v[DPT][LB];
#pragma unroll
				for( int j = 0; j < DPT; j++ )
#pragma unroll
					for( int k = 0; k < LB; k++ )
					{
						v[j][k] = F_INC; // in front of tile 0 (or a digit that is done): inclusive prefix 0
						if( !done[j] && t - k >= 0 ) v[j][k] = *(const volatile DescT*)( lookback + (size_t)( t - k ) * RADIX + threadIdx.x * DPT + j );
					}
				all = true;
#pragma unroll
				for( int j = 0; j < DPT; j++ )
				{
#pragma unroll
					for( int k = 0; k < LB; k++ )
					{
						if( done[j] ) continue;
						DescT vv = v[j][k];
						while( ( vv >> FSH ) == 0 )
						{
							if( backoffNs ) __nanosleep( backoffNs );
#ifdef HAKO_SORT_PROFILE
							if( threadIdx.x == 0 ) prof_spins++;
#endif
							vv = *(const volatile DescT*)( lookback + (size_t)( t - k ) * RADIX + threadIdx.x * DPT + j );
						}
						excl[j] += vv & V_MASK;
						if( ( vv >> FSH ) == 2 ) done[j] = true;
					}
					all = all && done[j];
				}
				t -= LB;
#ifdef HAKO_SORT_PROFILE
				prof_rounds++;
#endif
			}
#pragma unroll
			for( int j = 0; j < DPT; j++ )
				*(volatile DescT*)( lookback + (size_t)tile * RADIX + threadIdx.x * DPT + j ) = F_INC | (DescT)( excl[j] + total[j] );
		}
		SORT_PROF_MARK( 4 ); // thread 0's own look-back
#ifdef HAKO_SORT_PROFILE
		if( threadIdx.x == 0 ) atomicAdd( &g_sortProf[12], (unsigned long long)prof_rounds ), atomicAdd( &g_sortProf[14], (unsigned long long)prof_spins );
#endif
		uint32_t inc = mine;
#pragma unroll
		for( int o = 1; o < 32; o <<= 1 )
		{
			const uint32_t nb = __shfl_up_sync( 0xFFFFFFFFu, inc, o );
			if( lane >= o ) inc += nb;
		}
		if( lane == 31 ) s_warpTotals[warp] = inc;
		__syncthreads();
		uint32_t wbase = 0;
#pragma unroll
		for( int w = 0; w < WARPS; w++ )
			if( w < (int)warp ) wbase += s_warpTotals[w];
		uint32_t tb = wbase + inc - mine;
		if( owner )
		{
#pragma unroll
			for( int j = 0; j < DPT; j++ )
			{
				const uint32_t d = threadIdx.x * DPT + j;
				s_tileBase[d] = tb;
				s_globalBase[d] = digitBase[d] + excl[j] - tb;
				tb += total[j];
			}
		}
	}
	__syncthreads();
	SORT_PROF_MARK( 5 ); // everyone's look-back + tile bases
#pragma unroll
	for( int i = 0; i < ITEMS; i++ )
	{
		const uint32_t idx = warpBase + i * 32 + lane;
		if( idx < tileCount )
		{
			const uint32_t d = digitOf( key[i], shift, mask );
			const uint32_t p = s_tileBase[d] + myHist[d] + rank[i];
			rank[i] = (uint16_t)p;
			s_keys[p] = key[i];
		}
	}
	__syncthreads();
	SORT_PROF_MARK( 6 ); // keys in sorted order in shared memory
	uint64_t dst[ITEMS];
#pragma unroll
	for( int i = 0; i < ITEMS; i++ )
	{
		const uint32_t p = i * BLOCK + threadIdx.x;
		if( p < tileCount )
		{
			const uint64_t k = s_keys[p];
			dst[i] = s_globalBase[digitOf( k, shift, mask )] + p;
			keysOut[dst[i]] = k;
		}
	}
	SORT_PROF_MARK( 7 ); // key stores issued
	if( PAIRS )
	{
		// the value tile has been in flight since kernel entry
		if( bulk ) mbarWait( &s_bar[1], 0 );
		__syncthreads(); // the sorted keys have left s_keys
		SORT_PROF_MARK( 8 );
#pragma unroll
		for( int i = 0; i < ITEMS; i++ )
		{
			const uint32_t idx = warpBase + i * 32 + lane;
			if( idx < tileCount ) s_keys[rank[i]] = ( oddTail && idx == tileCount - 1 ) ? valuesIn[tileStart + idx] : s_values[idx];
		}
		__syncthreads();
		SORT_PROF_MARK( 9 ); // values in sorted order in shared memory
#pragma unroll
		for( int i = 0; i < ITEMS; i++ )
		{
			const uint32_t p = i * BLOCK + threadIdx.x;
			if( p < tileCount ) valuesOut[dst[i]] = s_keys[p];
		}
		SORT_PROF_MARK( 10 ); // value stores issued
	}
#ifdef HAKO_SORT_PROFILE
	if( threadIdx.x == 0 ) atomicAdd( &g_sortProf[13], 1ull );
#endif
}

// ---- cluster pass (HAKO_SORT_VARIANT=4): CL consecutive tiles per thread-block cluster ---------------------------------------
// The bulk pass above spends a third of a tile's life in the look-back (tools/sort_phase_probe.py): ~9 round trips over predecessors
// that have published their counts but are themselves still looking back, plus polls of tiles that are still ranking.  Here the CL CTAs
// of a cluster take CL consecutive tiles, exchange their per-digit counts through distributed shared memory (one cluster barrier), and
// the decoupled look-back runs over CLUSTERS: CL times fewer descriptors, CL times more keys per round trip.  Every CTA of the cluster
// walks the (same) cluster chain itself -- no second barrier; CTA 0 publishes.
// MEASURED AND NOT ADOPTED: correct (stable) but slower than the bulk pass -- 2 / 4 / 8 CTAs per cluster 4.52 / 4.49 / 4.78 ms against 3.44 ms
// for 72.7 M pairs x 36 bits (profiles/r2v_sort_cluster.log): the tile loads cannot start before the whole cluster runs (barrier 0), and
// barrier 1 makes every tile wait for the slowest ranking of its cluster, which costs more than the look-back round trips it saves.
__device__ __forceinline__ uint32_t clusterCtaRank()
{
	uint32_t r;
	asm volatile( "mov.u32 %0, %%cluster_ctarank;" : "=r"( r ) );
	return r;
}
__device__ __forceinline__ void clusterArrive() { asm volatile( "barrier.cluster.arrive.release.aligned;" ::: "memory" ); }
__device__ __forceinline__ void clusterWait() { asm volatile( "barrier.cluster.wait.acquire.aligned;" ::: "memory" ); }
__device__ __forceinline__ uint32_t mapToCta( const void* smemPtr, uint32_t rank )
{
	uint32_t out;
	asm volatile( "mapa.shared::cluster.u32 %0, %1, %2;" : "=r"( out ) : "r"( smemAddr( smemPtr ) ), "r"( rank ) );
	return out;
}
__device__ __forceinline__ uint32_t ldCluster32( uint32_t addr )
{
	uint32_t v;
	asm volatile( "ld.shared::cluster.u32 %0, [%1];" : "=r"( v ) : "r"( addr ) : "memory" );
	return v;
}
__device__ __forceinline__ uint2 ldCluster64( uint32_t addr )
{
	uint2 v;
	asm volatile( "ld.shared::cluster.v2.u32 {%0, %1}, [%2];" : "=r"( v.x ), "=r"( v.y ) : "r"( addr ) : "memory" );
	return v;
}

template <int RBITS, bool PAIRS, int BLOCK, int ITEMS, int MIN_CTAS, int CL, bool D32>
__global__ void __launch_bounds__( BLOCK, MIN_CTAS ) onesweepClusterKernel( const uint64_t* __restrict__ keysIn, const uint64_t* __restrict__ valuesIn,
																	uint64_t* __restrict__ keysOut, uint64_t* __restrict__ valuesOut, uint64_t n,
																	int shift, uint32_t mask, const unsigned long long* __restrict__ digitBase /* [RADIX] exclusive */,
																	void* __restrict__ lookbackRaw /* [clusters][RADIX] descriptors, zeroed */, uint32_t* __restrict__ ticket )
{
	using L = SortV2Smem<RBITS, BLOCK, ITEMS>;
	using DescT = typename std::conditional<D32, uint32_t, unsigned long long>::type;
	constexpr int FSH = D32 ? 30 : 62;
	constexpr DescT F_AGG = (DescT)1 << FSH, F_INC = (DescT)2 << FSH, V_MASK = ( (DescT)1 << FSH ) - 1;
	constexpr int RADIX = L::RADIX, WARPS = L::WARPS, TILE = L::TILE;
	constexpr int DPT = RADIX / BLOCK; // digits per thread in the per-digit phase
	static_assert( RADIX % BLOCK == 0 && DPT == 2, "the cluster pass keeps two digits per thread (9-bit digits, 256 threads)" );
	constexpr int LB = HAKO_SORT_LB;
	DescT* __restrict__ lookback = (DescT*)lookbackRaw;
	extern __shared__ __align__( 128 ) unsigned char s_raw[];
	uint64_t* s_keys = (uint64_t*)( s_raw + L::OFF_KEYS );
	uint64_t* s_values = (uint64_t*)( s_raw + L::OFF_VALUES );
	unsigned long long* s_globalBase = (unsigned long long*)( s_raw + L::OFF_GBASE );
	uint32_t* s_tileBase = (uint32_t*)( s_raw + L::OFF_TBASE );
	uint16_t* s_warpHist = (uint16_t*)( s_raw + L::OFF_WHIST ); // [WARPS][RADIX]
	uint32_t* s_warpTotals = (uint32_t*)( s_raw + L::OFF_WTOT );
	uint64_t* s_bar = (uint64_t*)( s_raw + L::OFF_BAR );
	uint32_t* s_ticket = (uint32_t*)( s_bar + 2 );
	// the tile's per-digit counts, read by the other CTAs of the cluster (their own 2 KB behind the bulk pass's layout: a CTA that is
	// ahead must not overwrite them while a neighbour still reads)
	uint32_t* s_counts = (uint32_t*)( s_raw + L::BYTES );

	const uint32_t crank = clusterCtaRank();
	if( threadIdx.x == 0 )
	{
		mbarInit( &s_bar[0], 1 );
		mbarInit( &s_bar[1], 1 );
		asm volatile( "fence.mbarrier_init.release.cluster;" ::: "memory" );
		if( crank == 0 ) *s_ticket = atomicAdd( ticket, 1u );
	}
	for( int i = threadIdx.x; i < WARPS * RADIX / 2; i += BLOCK ) ( (uint32_t*)s_warpHist )[i] = 0;
	// barrier 0: every CTA of the cluster is running (its shared memory may be addressed) and CTA 0's ticket is visible
	clusterArrive();
	clusterWait();
	const uint32_t cluster = ldCluster32( mapToCta( s_ticket, 0 ) );
	const uint32_t tile = cluster * CL + crank;
	const uint64_t tileStart = (uint64_t)tile * TILE;
	const uint32_t tileCount = tileStart < n ? (uint32_t)min( (uint64_t)TILE, n - tileStart ) : 0u;
	const uint32_t bulkBytes = ( tileCount * 8u ) & ~15u; // whole 16-byte units; an odd last key is fetched below
	if( threadIdx.x == 0 && bulkBytes )
	{
		mbarExpectTx( &s_bar[0], bulkBytes );
		bulkLoad( s_keys, keysIn + tileStart, bulkBytes, &s_bar[0] );
		if( PAIRS )
		{
			mbarExpectTx( &s_bar[1], bulkBytes );
			bulkLoad( s_values, valuesIn + tileStart, bulkBytes, &s_bar[1] );
		}
	}
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	const bool bulk = bulkBytes != 0;
	const bool oddTail = ( tileCount & 1u ) != 0;

	uint64_t key[ITEMS];
	uint16_t rank[ITEMS];
	const uint32_t warpBase = warp * ( ITEMS * 32 );
	if( bulk ) mbarWait( &s_bar[0], 0 );
#pragma unroll
	for( int i = 0; i < ITEMS; i++ )
	{
		const uint32_t idx = warpBase + i * 32 + lane;
		uint64_t k = ~0ull;
		if( idx < tileCount ) k = ( oddTail && idx == tileCount - 1 ) ? keysIn[tileStart + idx] : s_keys[idx];
		key[i] = k;
	}
	uint16_t* myHist = s_warpHist + warp * RADIX;
#pragma unroll
	for( int i = 0; i < ITEMS; i++ )
	{
		const uint32_t idx = warpBase + i * 32 + lane;
		const bool valid = idx < tileCount;
		const uint32_t d = valid ? digitOf( key[i], shift, mask ) : ( RADIX - 1 );
		uint32_t peers = __ballot_sync( 0xFFFFFFFFu, valid );
		peers = valid ? peers : ~peers;
#pragma unroll
		for( int b = 0; b < RBITS; b++ )
		{
			const bool bit = ( d >> b ) & 1u;
			const uint32_t m = __ballot_sync( 0xFFFFFFFFu, bit );
			peers &= bit ? m : ~m;
		}
		const uint32_t before = __popc( peers & ( ( 1u << lane ) - 1u ) );
		const int leader = __ffs( peers ) - 1;
		uint32_t old = 0;
		if( valid && (int)lane == leader )
		{
			old = myHist[d];
			myHist[d] = (uint16_t)( old + __popc( peers ) );
		}
		__syncwarp();
		old = __shfl_sync( 0xFFFFFFFFu, old, leader );
		rank[i] = (uint16_t)( old + before );
	}
	__syncthreads(); // every key is in registers: s_keys becomes the sorted staging buffer

	uint32_t total[DPT];
	uint32_t mine = 0;
#pragma unroll
	for( int j = 0; j < DPT; j++ )
	{
		const uint32_t d = threadIdx.x * DPT + j;
		uint32_t tot = 0;
#pragma unroll
		for( int w = 0; w < WARPS; w++ )
		{
			const uint32_t c = s_warpHist[w * RADIX + d];
			s_warpHist[w * RADIX + d] = (uint16_t)tot;
			tot += c;
		}
		total[j] = tot;
		mine += tot;
	}
	*(uint2*)( s_counts + threadIdx.x * DPT ) = make_uint2( total[0], total[1] );
	// barrier 1: the counts of all CL tiles are in their CTAs' shared memory
	clusterArrive();
	// (between arrive and wait: the tile-local bases, which need nothing from the other CTAs)
	uint32_t inc = mine;
#pragma unroll
	for( int o = 1; o < 32; o <<= 1 )
	{
		const uint32_t nb = __shfl_up_sync( 0xFFFFFFFFu, inc, o );
		if( lane >= o ) inc += nb;
	}
	if( lane == 31 ) s_warpTotals[warp] = inc;
	__syncthreads();
	uint32_t wbase = 0;
#pragma unroll
	for( int w = 0; w < WARPS; w++ )
		if( w < (int)warp ) wbase += s_warpTotals[w];
	const uint32_t tb0 = wbase + inc - mine, tb1 = tb0 + total[0];
	s_tileBase[threadIdx.x * DPT] = tb0;
	s_tileBase[threadIdx.x * DPT + 1] = tb1;
	clusterWait();

	// counts of the cluster's tiles for this thread's digits: in front of this tile, and all of them
	uint32_t ahead[DPT] = { 0u, 0u }, whole[DPT] = { 0u, 0u };
#pragma unroll
	for( int r = 0; r < CL; r++ )
	{
		const uint2 c = ldCluster64( mapToCta( s_counts + threadIdx.x * DPT, (uint32_t)r ) );
		if( r < (int)crank ) ahead[0] += c.x, ahead[1] += c.y;
		whole[0] += c.x, whole[1] += c.y;
	}
	// barrier 2 (arrive now, wait before exit): nobody leaves while its counts may still be read
	clusterArrive();
	if( crank == 0 )
	{
#pragma unroll
		for( int j = 0; j < DPT; j++ )
			*(volatile DescT*)( lookback + (size_t)cluster * RADIX + threadIdx.x * DPT + j ) = ( cluster == 0 ? F_INC : F_AGG ) | (DescT)whole[j];
	}
	unsigned long long excl[DPT] = { 0ull, 0ull };
	if( cluster != 0 )
	{
		bool done[DPT] = { false, false };
		int64_t t = (int64_t)cluster - 1;
		bool all = false;
		while( !all )
		{
			DescT v[DPT][LB];
#pragma unroll
			for( int j = 0; j < DPT; j++ )
#pragma unroll
				for( int k = 0; k < LB; k++ )
				{
					v[j][k] = F_INC; // in front of cluster 0 (or a digit that is done): inclusive prefix 0
					if( !done[j] && t - k >= 0 ) v[j][k] = *(const volatile DescT*)( lookback + (size_t)( t - k ) * RADIX + threadIdx.x * DPT + j );
				}
			all = true;
#pragma unroll
			for( int j = 0; j < DPT; j++ )
			{
#pragma unroll
				for( int k = 0; k < LB; k++ )
				{
					if( done[j] ) continue;
					DescT vv = v[j][k];
					while( ( vv >> FSH ) == 0 ) vv = *(const volatile DescT*)( lookback + (size_t)( t - k ) * RADIX + threadIdx.x * DPT + j );
					excl[j] += vv & V_MASK;
					if( ( vv >> FSH ) == 2 ) done[j] = true;
				}
				all = all && done[j];
			}
			t -= LB;
		}
		if( crank == 0 )
		{
#pragma unroll
			for( int j = 0; j < DPT; j++ )
				*(volatile DescT*)( lookback + (size_t)cluster * RADIX + threadIdx.x * DPT + j ) = F_INC | (DescT)( excl[j] + whole[j] );
		}
	}
	s_globalBase[threadIdx.x * DPT] = digitBase[threadIdx.x * DPT] + excl[0] + ahead[0] - tb0;
	s_globalBase[threadIdx.x * DPT + 1] = digitBase[threadIdx.x * DPT + 1] + excl[1] + ahead[1] - tb1;
	__syncthreads();
#pragma unroll
	for( int i = 0; i < ITEMS; i++ )
	{
		const uint32_t idx = warpBase + i * 32 + lane;
		if( idx < tileCount )
		{
			const uint32_t d = digitOf( key[i], shift, mask );
			const uint32_t p = s_tileBase[d] + myHist[d] + rank[i];
			rank[i] = (uint16_t)p;
			s_keys[p] = key[i];
		}
	}
	__syncthreads();
	uint64_t dst[ITEMS];
#pragma unroll
	for( int i = 0; i < ITEMS; i++ )
	{
		const uint32_t p = i * BLOCK + threadIdx.x;
		if( p < tileCount )
		{
			const uint64_t k = s_keys[p];
			dst[i] = s_globalBase[digitOf( k, shift, mask )] + p;
			keysOut[dst[i]] = k;
		}
	}
	if( PAIRS )
	{
		if( bulk ) mbarWait( &s_bar[1], 0 );
		__syncthreads(); // the sorted keys have left s_keys
#pragma unroll
		for( int i = 0; i < ITEMS; i++ )
		{
			const uint32_t idx = warpBase + i * 32 + lane;
			if( idx < tileCount ) s_keys[rank[i]] = ( oddTail && idx == tileCount - 1 ) ? valuesIn[tileStart + idx] : s_values[idx];
		}
		__syncthreads();
#pragma unroll
		for( int i = 0; i < ITEMS; i++ )
		{
			const uint32_t p = i * BLOCK + threadIdx.x;
			if( p < tileCount ) valuesOut[dst[i]] = s_keys[p];
		}
	}
	clusterWait(); // barrier 2
}

template <int RBITS, bool PAIRS, int CL, bool D32>
int launchClusterOne( hako_ctx* ctx, cudaStream_t s, uint32_t tiles, const uint64_t* kin, const uint64_t* vin, uint64_t* kout, uint64_t* vout, uint64_t n, int shift, uint32_t mask,
					  const unsigned long long* digitBase, unsigned long long* lookback, uint32_t* ticket )
{
	auto k = onesweepClusterKernel<RBITS, PAIRS, 256, 15, 3, CL, D32>;
	constexpr size_t bytes = SortV2Smem<RBITS, 256, 15>::BYTES + ( (size_t)4 << RBITS );
	static bool configured[16] = {}; // per device and instantiation
	const int dev = ctx->device & 15;
	if( !configured[dev] )
	{
		HAKO_CUDA( cudaFuncSetAttribute( k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes ) );
		configured[dev] = true;
	}
	cudaLaunchConfig_t cfg = {};
	cfg.gridDim = dim3( ( ( tiles + CL - 1 ) / CL ) * CL );
	cfg.blockDim = dim3( 256 );
	cfg.dynamicSmemBytes = bytes;
	cfg.stream = s;
	cudaLaunchAttribute attr[1];
	attr[0].id = cudaLaunchAttributeClusterDimension;
	attr[0].val.clusterDim.x = CL, attr[0].val.clusterDim.y = 1, attr[0].val.clusterDim.z = 1;
	cfg.attrs = attr;
	cfg.numAttrs = 1;
	void* lb = (void*)lookback;
	HAKO_CUDA( cudaLaunchKernelEx( &cfg, k, kin, vin, kout, vout, n, shift, mask, digitBase, lb, ticket ) );
	ctx->launches.fetch_add( 1, std::memory_order_relaxed );
	return 0;
}
template <int RBITS, bool PAIRS>
int launchClusterRP( hako_ctx* ctx, cudaStream_t s, int cl, bool d32, uint32_t tiles, const uint64_t* kin, const uint64_t* vin, uint64_t* kout, uint64_t* vout, uint64_t n, int shift,
					 uint32_t mask, const unsigned long long* digitBase, unsigned long long* lookback, uint32_t* ticket )
{
#define HAKO_CL1( C, D ) return launchClusterOne<RBITS, PAIRS, C, D>( ctx, s, tiles, kin, vin, kout, vout, n, shift, mask, digitBase, lookback, ticket )
	if( cl == 2 )
	{
		if( d32 ) HAKO_CL1( 2, true );
		HAKO_CL1( 2, false );
	}
	if( cl == 4 )
	{
		if( d32 ) HAKO_CL1( 4, true );
		HAKO_CL1( 4, false );
	}
	if( d32 ) HAKO_CL1( 8, true );
	HAKO_CL1( 8, false );
#undef HAKO_CL1
}

// variants of the bulk-copy pass: block size x items per thread x resident CTAs per SM (HAKO_SORT_VARIANT)
//   1: 256 x 15 x 3 (3 840-pair tiles, 71 KB)   2: 512 x 8 x 2 (4 096-pair tiles, 83 KB)   3: 256 x 16 x 2 (4 096, 79 KB)
//   4: the cluster pass above.  Smaller tiles at higher occupancy were measured and dropped: 256 x 10 x 4 3.95 ms, 256 x 7 x 5 4.82 ms
//   against 3.45 ms for variant 1 (profiles/r2w_sort_tiles.log); the cluster pass 4.49-4.78 ms (profiles/r2v_sort_cluster.log).
//   Not staging the value tile (L2 prefetch at entry, late loads into registers, a u16 digit per sorted position instead of the
//   destinations in registers) buys a 4th / 5th CTA per SM and nothing else: 256 x 15 x 4 3.55 ms, 256 x 12 x 5 3.88 ms
//   (profiles/r3b_sort_latev.log; the code is in the history, not in the tree).  More tiles in flight only lengthen every look-back
//   walk: the pass is bound by how fast inclusive prefixes propagate from tile to tile, not by occupancy.
int sortVariantTile( int variant ) { return ( variant == 1 || variant == 4 ) ? 256 * 15 : ( variant == 0 ? SORT_TILE : 4096 ); }
template <int RBITS, bool PAIRS, int BLOCK, int ITEMS, int MIN_CTAS, bool D32 = false, int LB = HAKO_SORT_LB, int RANK = 0>
int launchBulkOne( hako_ctx* ctx, cudaStream_t s, uint32_t tiles, const uint64_t* kin, const uint64_t* vin, uint64_t* kout, uint64_t* vout, uint64_t n, int shift, uint32_t mask,
				   const unsigned long long* digitBase, unsigned long long* lookback, uint32_t* ticket )
{
	auto k = onesweepBulkKernel<RBITS, PAIRS, BLOCK, ITEMS, MIN_CTAS, D32, LB, RANK>;
	constexpr size_t bytes = SortV2Smem<RBITS, BLOCK, ITEMS>::BYTES;
	static bool configured[16] = {}; // per device and instantiation
	int dev = ctx->device & 15;
	if( !configured[dev] )
	{
		HAKO_CUDA( cudaFuncSetAttribute( k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes ) );
		configured[dev] = true;
	}
	static const int backoffEnv = getenv( "HAKO_SORT_BACKOFF" ) ? atoi( getenv( "HAKO_SORT_BACKOFF" ) ) : LB_BACKOFF_NS; // ns a look-back poll sleeps when its predecessor is not ready
	HAKO_LAUNCH( ctx, k, tiles, BLOCK, bytes, s, kin, vin, kout, vout, n, shift, mask, digitBase, (void*)lookback, ticket, (uint32_t)( backoffEnv > 0 ? backoffEnv : ( backoffEnv == -1 ? 0xFFFFFFFFu : 0u ) ) ); // -1: profiling builds only, see the kernel
	return 0;
}
template <int RBITS, bool PAIRS>
int launchBulkRP( hako_ctx* ctx, cudaStream_t s, int variant, uint32_t tiles, const uint64_t* kin, const uint64_t* vin, uint64_t* kout, uint64_t* vout, uint64_t n, int shift,
				  uint32_t mask, const unsigned long long* digitBase, unsigned long long* lookback, uint32_t* ticket )
{
	if( variant == 1 )
	{
		// A/B knobs of the default variant: HAKO_SORT_D32=0 forces 64-bit descriptors, HAKO_SORT_RANK=0 the MATCH.ANY + atomic ranking
		static const bool d32Env = getenv( "HAKO_SORT_D32" ) ? atoi( getenv( "HAKO_SORT_D32" ) ) != 0 : true;
		static const int rankEnv = getenv( "HAKO_SORT_RANK" ) ? atoi( getenv( "HAKO_SORT_RANK" ) ) & 3 : 3;
		const bool d32 = d32Env && n < ( 1ull << 30 );
#define HAKO_BULK1( D, A ) return launchBulkOne<RBITS, PAIRS, 256, 15, 3, D, HAKO_SORT_LB, A>( ctx, s, tiles, kin, vin, kout, vout, n, shift, mask, digitBase, lookback, ticket )
		if( d32 )
		{
			if( rankEnv == 3 ) HAKO_BULK1( true, 3 );
			HAKO_BULK1( true, 0 );
		}
		if( rankEnv == 3 ) HAKO_BULK1( false, 3 );
		HAKO_BULK1( false, 0 );
#undef HAKO_BULK1
	}
	if( variant == 3 ) return launchBulkOne<RBITS, PAIRS, 256, 16, 2>( ctx, s, tiles, kin, vin, kout, vout, n, shift, mask, digitBase, lookback, ticket );
	return launchBulkOne<RBITS, PAIRS, 512, 8, 2>( ctx, s, tiles, kin, vin, kout, vout, n, shift, mask, digitBase, lookback, ticket );
}
int launchBulkPass( hako_ctx* ctx, cudaStream_t s, int variant, int rbits, bool pairs, uint32_t tiles, const uint64_t* kin, const uint64_t* vin, uint64_t* kout, uint64_t* vout,
					uint64_t n, int shift, uint32_t mask, const unsigned long long* digitBase, unsigned long long* lookback, uint32_t* ticket )
{
	if( variant == 4 )
	{
		// cluster pass: 9-bit digits only (two digits per thread); HAKO_SORT_CLUSTER = CTAs per cluster (2 / 4 / 8)
		static const int clEnv = getenv( "HAKO_SORT_CLUSTER" ) ? atoi( getenv( "HAKO_SORT_CLUSTER" ) ) : 8;
		static const bool d32Env = getenv( "HAKO_SORT_D32" ) ? atoi( getenv( "HAKO_SORT_D32" ) ) != 0 : true;
		const bool d32 = d32Env && n < ( 1ull << 30 );
		if( rbits == 9 )
			return pairs ? launchClusterRP<9, true>( ctx, s, clEnv, d32, tiles, kin, vin, kout, vout, n, shift, mask, digitBase, lookback, ticket )
						 : launchClusterRP<9, false>( ctx, s, clEnv, d32, tiles, kin, vin, kout, vout, n, shift, mask, digitBase, lookback, ticket );
		variant = 1;
	}
	if( rbits == 9 )
		return pairs ? launchBulkRP<9, true>( ctx, s, variant, tiles, kin, vin, kout, vout, n, shift, mask, digitBase, lookback, ticket )
					 : launchBulkRP<9, false>( ctx, s, variant, tiles, kin, vin, kout, vout, n, shift, mask, digitBase, lookback, ticket );
	return pairs ? launchBulkRP<8, true>( ctx, s, variant, tiles, kin, vin, kout, vout, n, shift, mask, digitBase, lookback, ticket )
				 : launchBulkRP<8, false>( ctx, s, variant, tiles, kin, vin, kout, vout, n, shift, mask, digitBase, lookback, ticket );
}
} // namespace

// Sorts in place from the caller's point of view unless resultInTmp is given: with an odd number of passes the result
// lands in (keysTmp, valuesTmp); *resultInTmp tells the caller, otherwise it is copied back.
int sortPairsDevice( hako_ctx* ctx, cudaStream_t s, uint64_t* keys, uint64_t* values, uint64_t* keysTmp, uint64_t* valuesTmp,
					 uint64_t n, int beginBit, int endBit, int* resultInTmp )
{
	if( resultInTmp ) *resultInTmp = 0;
	if( beginBit < 0 || endBit > 64 || endBit < beginBit )
	{
		setError( "sort: bad bit window [%d, %d)", beginBit, endBit );
		return HAKO_E_INVALID;
	}
	// digit width: 9 bits when that saves a pass, else 8
	const int window = endBit - beginBit;
	const char* rbEnv = getenv( "HAKO_SORT_RBITS" ); // tuning knob: force the digit width
	const int rbits = rbEnv ? ( atoi( rbEnv ) == 9 ? 9 : 8 ) : ( ( window + 8 ) / 9 < ( window + 7 ) / 8 ? 9 : 8 );
	const int RADIX = 1 << rbits;
	const int passes = ( window + rbits - 1 ) / rbits;
	ctx->stats.sortPasses = (uint32_t)passes;
	if( n == 0 || passes == 0 ) return 0;
	// bulk-copy passes need 16-byte aligned arrays; a caller's own device pointers (hako_sort_pairs_device) may not be
	const bool aligned16 = ( ( (uintptr_t)keys | (uintptr_t)keysTmp | (uintptr_t)values | (uintptr_t)valuesTmp ) & 15u ) == 0;
	const char* varEnv = getenv( "HAKO_SORT_VARIANT" ); // 0: first-generation kernel (plain loads, 4 096-pair tiles)
	int variant = varEnv ? atoi( varEnv ) : 1; // B200, 72.7 M pairs x 36 bits: variant 0 4.46 ms, 1 4.18, 2 4.62, 3 5.23 (profiles/r2e_sort_variants.log)
	if( variant < 0 || variant > 4 || !aligned16 ) variant = 0;
	const uint64_t tileSize = (uint64_t)sortVariantTile( variant );
	const uint64_t tiles = ( n + tileSize - 1 ) / tileSize;
	if( tiles >= 0x7FFFFFFFull )
	{
		setError( "sort: too many tiles" );
		return HAKO_E_LIMIT;
	}
	unsigned long long* hist = nullptr;
	unsigned long long* lookback = nullptr;
	uint32_t* tickets = nullptr;
	int r = 0;
	do
	{
		if( ( r = dalloc( ctx, &hist, (uint64_t)passes * RADIX, s ) ) ) break;
		if( ( r = dalloc( ctx, &lookback, (uint64_t)passes * tiles * RADIX, s ) ) ) break;
		if( ( r = dalloc( ctx, &tickets, (uint64_t)passes, s ) ) ) break;
		cudaError_t e = cudaMemsetAsync( hist, 0, (size_t)passes * RADIX * 8, s );
		if( e == cudaSuccess ) e = cudaMemsetAsync( lookback, 0, (size_t)passes * tiles * RADIX * 8, s );
		if( e == cudaSuccess ) e = cudaMemsetAsync( tickets, 0, (size_t)passes * 4, s );
		if( e != cudaSuccess )
		{
			r = cudaFail( e, "memset sort state", __FILE__, __LINE__ );
			break;
		}
		const uint32_t histBlocks = (uint32_t)min( (uint64_t)ctx->smCount * 8, ( n + SORT_BLOCK * 8 - 1 ) / ( SORT_BLOCK * 8 ) );
		if( rbits == 9 )
		{
			HAKO_LAUNCH_R( ctx, r, sortHistogramKernel<9>, histBlocks, SORT_BLOCK, 0, s, keys, n, beginBit, endBit, passes, hist );
			HAKO_LAUNCH_R( ctx, r, sortScanHistogramKernel<9>, passes, RADIX, 0, s, hist );
		}
		else
		{
			HAKO_LAUNCH_R( ctx, r, sortHistogramKernel<8>, histBlocks, SORT_BLOCK, 0, s, keys, n, beginBit, endBit, passes, hist );
			HAKO_LAUNCH_R( ctx, r, sortScanHistogramKernel<8>, passes, RADIX, 0, s, hist );
		}
		uint64_t *kin = keys, *vin = values, *kout = keysTmp, *vout = valuesTmp;
		for( int p = 0; p < passes; p++ )
		{
			const int shift = beginBit + p * rbits;
			const int bits = ( endBit - shift < rbits ) ? ( endBit - shift ) : rbits;
			const uint32_t mask = ( 1u << bits ) - 1u;
			auto k8p = onesweepKernel<8, true, 2>;
			auto k8k = onesweepKernel<8, false, 2>;
			auto k9p = onesweepKernel<9, true, 2>;
			auto k9k = onesweepKernel<9, false, 2>;
			auto k = rbits == 9 ? ( values ? k9p : k9k ) : ( values ? k8p : k8k );
			const char* occEnv = getenv( "HAKO_SORT_MIN_CTAS" ); // tuning knob
			const int occ = occEnv ? atoi( occEnv ) : 4; // B200: 2 CTAs/SM 5.43 ms, 3: 5.08, 4: 4.50 (72.7 M pairs, 36 bits; profiles/r1g)
			if( occ == 3 ) k = rbits == 9 ? ( values ? onesweepKernel<9, true, 3> : onesweepKernel<9, false, 3> ) : ( values ? onesweepKernel<8, true, 3> : onesweepKernel<8, false, 3> );
			if( occ == 4 ) k = rbits == 9 ? ( values ? onesweepKernel<9, true, 4> : onesweepKernel<9, false, 4> ) : ( values ? onesweepKernel<8, true, 4> : onesweepKernel<8, false, 4> );
			if( variant == 0 )
				HAKO_LAUNCH_R( ctx, r, k, (uint32_t)tiles, SORT_BLOCK, 0, s, kin, vin, kout, vout, n, shift, mask, hist + (size_t)p * RADIX,
							 lookback + (size_t)p * tiles * RADIX, tickets + p );
			else if( ( r = launchBulkPass( ctx, s, variant, rbits, values != nullptr, (uint32_t)tiles, kin, vin, kout, vout, n, shift, mask, hist + (size_t)p * RADIX,
										  lookback + (size_t)p * tiles * RADIX, tickets + p ) ) )
				break;
			uint64_t* t = kin;
			kin = kout, kout = t;
			t = vin;
			vin = vout, vout = t;
		}
		if( r ) break;
		if( kin != keys )
		{
			if( resultInTmp )
				*resultInTmp = 1;
			else
			{
				e = cudaMemcpyAsync( keys, kin, n * 8, cudaMemcpyDeviceToDevice, s );
				if( e == cudaSuccess && values ) e = cudaMemcpyAsync( values, vin, n * 8, cudaMemcpyDeviceToDevice, s );
				if( e != cudaSuccess ) r = cudaFail( e, "copy back", __FILE__, __LINE__ );
			}
		}
	} while( 0 );
	dfree( hist, s ), dfree( lookback, s ), dfree( tickets, s );
	return r;
}
} // namespace hako

#ifdef HAKO_SORT_PROFILE
extern "C" int hako_debug_sort_profile( unsigned long long* out16, int reset )
{
	if( out16 && cudaMemcpyFromSymbol( out16, hako::g_sortProf, sizeof( hako::g_sortProf ) ) != cudaSuccess ) return -1;
	if( reset )
	{
		unsigned long long z[16] = {};
		if( cudaMemcpyToSymbol( hako::g_sortProf, z, sizeof( z ) ) != cudaSuccess ) return -1;
	}
	return 0;
}
#endif
