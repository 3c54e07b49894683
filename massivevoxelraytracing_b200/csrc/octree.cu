// octree.cu -- bottom-up sparse-voxel-octree build from a sorted unique Morton list, with DAG compression that
// reproduces the CPU reference's node numbering exactly.
//
// Reference semantics: buildOctreeDAGReference (IntersectorOctree.hpp:11-123) -- per level, group consecutive tasks by
// morton >> 3, node = { mask, children[8], nVoxelsPSum[8] }, dedupe through std::map<OctreeNode,int> whose order compares
// mask + children only (voxCommon.hpp:163-180), new nodes numbered in first-occurrence order, root last; then embedMasks
// (voxCommon.hpp:183-195).  buildOctreeNaive (:125-212) is the same without dedupe and with nVoxelsPSum left 0.
// The reference's GPU build (bottomUpOctreeBuild, voxKernel.cu:269-424) numbers nodes by an atomicInc race behind a CAS-locked
// linear-probing table, so its trees are isomorphic to the CPU's but not identical.  Here numbering is deterministic:
//
//   per level:  head flags + scan -> group index g            (groupBuildKernel, decoupled look-back)
//               key(g) = children[8] (the embedded child masks make `mask` a function of the children)
//               table[hash(key)] <- atomicMin(g)               (dagInsertKernel: the smallest g of every key wins)
//               first(g) = (table[key(g)] == g); rank = exclusive scan of first  (dagLookupKernel, fused scan, writes nodes)
//               node index of g = levelBase + rank[ rep(g) ]   (dagAssignKernel)
//
// which is exactly "first occurrence in group order gets the next index".  Equal keys cannot occur across levels (leaf-level
// children are all 0xFFFFFFFF, upper levels hold real indices), so per-level dedupe == the global std::map.
// Level 0 is special-cased: its key is the 8-bit mask alone, so first occurrences are found with a 256-entry atomicMin
// table and the <= 255 leaf nodes are rebuilt from their masks (nVoxelsPSum = popcount prefix) -- no per-group node is stored.
// Child pointers carry the child's mask from the start (index | mask << 24), so there is no separate embedMasks pass.
#include <cstdlib>

#include "internal.cuh"
#include "scan.cuh"

namespace hako
{
namespace
{
constexpr int OB_BLOCK = 256;
constexpr int OB_ITEMS = 4;
constexpr int OB_TILE = OB_BLOCK * OB_ITEMS;
constexpr uint32_t SENTINEL = 0xFFFFFFFFu;
constexpr int MAX_LEVELS = 21;

struct LevelState
{
	uint32_t levelBase[MAX_LEVELS + 2]; // levelBase[l] = number of nodes created below level l (device-resident, DAG mode)
};

// ---- per-level group counts in one pass (octreeTaskInit, voxKernel.cu:245-267, does this with one atomicInc per boundary) ----
__global__ void __launch_bounds__( 256 ) levelHistKernel( const uint64_t* __restrict__ mortons, uint64_t n, unsigned long long* __restrict__ hist /* [22] */ )
{
	__shared__ uint32_t s_hist[22];
	if( threadIdx.x < 22 ) s_hist[threadIdx.x] = 0;
	__syncthreads();
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for( uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x + 1; i < n; i += stride )
	{
		const uint64_t x = mortons[i] ^ mortons[i - 1];
		const int top = 63 - __clzll( (long long)x ); // x != 0: keys are unique
		atomicAdd( &s_hist[top / 3], 1u );
	}
	__syncthreads();
	if( threadIdx.x < 22 && s_hist[threadIdx.x] ) atomicAdd( &hist[threadIdx.x], (unsigned long long)s_hist[threadIdx.x] );
}

__device__ __forceinline__ void storeNode( TraceNode* dst, const uint32_t children[8], const uint32_t psum[8] )
{
	uint4* d = (uint4*)dst;
#pragma unroll
	for( int q = 0; q < 4; q++ ) d[q] = make_uint4( children[2 * q], psum[2 * q], children[2 * q + 1], psum[2 * q + 1] );
}

// ---- group construction ----
// LEAF: tasks are the voxels themselves (morton = mortons[i], child = sentinel, 1 voxel each).
// DAG : write the group's node to groupNode[g] (temp) for dedupe; otherwise (naive) write the final node at naiveBase + g.
template <bool LEAF, bool DAG>
__global__ void __launch_bounds__( OB_BLOCK ) groupBuildKernel( const uint64_t* __restrict__ taskMorton, const uint32_t* __restrict__ taskChild, const uint32_t* __restrict__ taskVox,
															   uint64_t n, ScanState st,
															   uint64_t* __restrict__ nextMorton, uint32_t* __restrict__ nextChild, uint32_t* __restrict__ nextVox,
															   TraceNode* __restrict__ groupNode, uint8_t* __restrict__ groupMask,
															   uint32_t* __restrict__ leafFirst /* [256] DAG leaf level */,
															   TraceNode* __restrict__ nodes, uint8_t* __restrict__ masks, uint32_t naiveBase, int wide )
{
	__shared__ uint32_t s_first[256];
	if( LEAF && DAG )
	{
		s_first[threadIdx.x] = SENTINEL;
	}
	const uint32_t tile = scanTakeTicket( st ); // has a __syncthreads
	const uint64_t base = (uint64_t)tile * OB_TILE + (uint64_t)threadIdx.x * OB_ITEMS;
	uint64_t m[OB_ITEMS];
	uint32_t heads = 0, nHeads = 0;
	uint64_t prev = ( base > 0 && base <= n ) ? taskMorton[base - 1] : 0ull;
#pragma unroll
	for( int i = 0; i < OB_ITEMS; i++ )
	{
		const uint64_t idx = base + i;
		m[i] = ( idx < n ) ? taskMorton[idx] : 0ull;
		const bool head = ( idx < n ) && ( idx == 0 || ( m[i] >> 3 ) != ( prev >> 3 ) );
		prev = m[i];
		if( head )
		{
			heads |= 1u << i;
			nHeads++;
		}
	}
	uint32_t total;
	const uint32_t excl = blockExclusiveScan<OB_BLOCK>( nHeads, &total );
	const uint64_t prefix = scanTilePrefix( st, tile, total );
	uint32_t g = (uint32_t)prefix + excl;
#pragma unroll
	for( int i = 0; i < OB_ITEMS; i++ )
	{
		if( !( heads & ( 1u << i ) ) ) continue;
		const uint64_t parent = m[i] >> 3;
		uint32_t mask = 0;
		uint32_t children[8], cnt[8];
#pragma unroll
		for( int q = 0; q < 8; q++ )
		{
			children[q] = SENTINEL;
			cnt[q] = 0;
		}
		// IntersectorOctree.hpp:76-87: walk the (at most 8) tasks of the group
		for( uint64_t j = base + i; j < n; j++ )
		{
			const uint64_t mj = ( j == base + i ) ? m[i] : taskMorton[j];
			if( ( mj >> 3 ) != parent ) break;
			const uint32_t s = (uint32_t)( mj & 7u );
			mask |= 1u << s;
			const uint32_t ch = LEAF ? SENTINEL : taskChild[j];
			const uint32_t nv = LEAF ? 1u : taskVox[j];
#pragma unroll
			for( int q = 0; q < 8; q++ )
			{
				if( q == (int)s )
				{
					children[q] = ch;
					cnt[q] = nv;
				}
			}
		}
		// :88-95 exclusive prefix over the 8 slots
		uint32_t psum[8];
		uint32_t sum = 0;
#pragma unroll
		for( int q = 0; q < 8; q++ )
		{
			psum[q] = DAG ? sum : 0u; // buildOctreeNaive leaves nVoxelsPSum at 0 (IntersectorOctree.hpp:195)
			sum += cnt[q];
		}
		nextMorton[g] = parent;
		nextVox[g] = sum;
		if( DAG )
		{
			groupMask[g] = (uint8_t)mask;
			if( LEAF )
			{
				if( s_first[mask] > g ) atomicMin( &s_first[mask], g );
			}
			else
			{
				storeNode( &groupNode[g], children, psum );
			}
		}
		else
		{
			const uint32_t idx = naiveBase + g;
			storeNode( &nodes[idx], children, psum );
			masks[idx] = (uint8_t)mask;
			nextChild[g] = wide ? idx : ( idx | ( mask << 24 ) );
		}
		g++;
	}
	if( LEAF && DAG )
	{
		__syncthreads();
		const uint32_t f = s_first[threadIdx.x];
		if( f != SENTINEL ) atomicMin( &leafFirst[threadIdx.x], f );
	}
}

// ---- DAG, leaf level: the groups of the voxels themselves ----
// Same result as groupBuildKernel<true, true> (parent key, voxel count and child mask of every group of <= 8 sibling voxels, plus the
// first group of each of the 255 masks), restructured like uniqueMergeKernel: 8 keys per thread in registers and ONE pass over them with a
// running mask (the generic kernel re-reads the <= 8 members of every group from global memory), warp 0 resolves the tile's offset by
// look-back while the other warps already build their masks, and the groups leave through shared memory as contiguous, coalesced rows.
constexpr int LG_ITEMS = 8;
constexpr int LG_TILE = OB_BLOCK * LG_ITEMS;
__global__ void __launch_bounds__( OB_BLOCK, 4 ) leafGroupDagKernel( const uint64_t* __restrict__ mortons, uint64_t n, ScanState st, uint64_t* __restrict__ nextMorton,
																	 uint32_t* __restrict__ nextVox, uint8_t* __restrict__ groupMask, uint32_t* __restrict__ leafFirst /* [256] */ )
{
	__shared__ uint32_t s_first[256];
	__shared__ uint64_t s_parent[LG_TILE];
	__shared__ uint8_t s_mask[LG_TILE];
	__shared__ unsigned long long s_prefix;
	s_first[threadIdx.x] = SENTINEL;
	const uint32_t tile = scanTakeTicket( st ); // has a __syncthreads
	const uint64_t base = (uint64_t)tile * LG_TILE + (uint64_t)threadIdx.x * LG_ITEMS;
	uint64_t m[LG_ITEMS];
	if( base + LG_ITEMS <= n )
	{
#pragma unroll
		for( int i = 0; i < LG_ITEMS; i += 2 )
		{
			const ulonglong2 mm = *reinterpret_cast<const ulonglong2*>( mortons + base + i );
			m[i] = mm.x, m[i + 1] = mm.y;
		}
	}
	else
	{
#pragma unroll
		for( int i = 0; i < LG_ITEMS; i++ ) m[i] = ( base + i < n ) ? mortons[base + i] : 0ull;
	}
	uint32_t heads = 0, nHeads = 0;
	uint64_t prev = ( base > 0 && base <= n ) ? mortons[base - 1] : 0ull;
#pragma unroll
	for( int i = 0; i < LG_ITEMS; i++ )
	{
		const uint64_t idx = base + i;
		const bool head = ( idx < n ) && ( idx == 0 || ( m[i] >> 3 ) != ( prev >> 3 ) );
		prev = m[i];
		if( head )
		{
			heads |= 1u << i;
			nHeads++;
		}
	}
	uint32_t total;
	const uint32_t excl = blockExclusiveScan<OB_BLOCK>( nHeads, &total );
	if( threadIdx.x < 32 )
	{
		const uint64_t pre = scanTilePrefixWarp( st, tile, total );
		if( threadIdx.x == 0 ) s_prefix = pre;
	}
	uint32_t slot = excl, mask = 0;
	uint64_t parent = 0;
	bool open = false;
#pragma unroll
	for( int i = 0; i < LG_ITEMS; i++ )
	{
		if( ( heads >> i ) & 1u )
		{
			if( open )
			{
				s_parent[slot] = parent;
				s_mask[slot] = (uint8_t)mask;
				slot++;
			}
			parent = m[i] >> 3;
			mask = 0;
			open = true;
		}
		if( open && base + i < n ) mask |= 1u << (uint32_t)( m[i] & 7u );
	}
	if( open )
	{
		// the thread's last group may continue in the keys of the following threads (at most 7 more)
		for( uint64_t j = base + LG_ITEMS; j < n; j++ )
		{
			const uint64_t mj = mortons[j];
			if( ( mj >> 3 ) != parent ) break;
			mask |= 1u << (uint32_t)( mj & 7u );
		}
		s_parent[slot] = parent;
		s_mask[slot] = (uint8_t)mask;
	}
	__syncthreads();
	const uint64_t prefix = s_prefix;
	for( uint32_t i = threadIdx.x; i < total; i += OB_BLOCK )
	{
		const uint32_t g = (uint32_t)prefix + i;
		const uint32_t mk = s_mask[i];
		nextMorton[g] = s_parent[i];
		nextVox[g] = (uint32_t)__popc( mk ); // one voxel per set child bit
		groupMask[g] = (uint8_t)mk;
		if( s_first[mk] > g ) atomicMin( &s_first[mk], g );
	}
	__syncthreads();
	const uint32_t f = s_first[threadIdx.x];
	if( f != SENTINEL ) atomicMin( &leafFirst[threadIdx.x], f );
}

// ---- DAG, levels above the leaves: the groups of the previous level's groups ----
// Same result as groupBuildKernel<false, true> (the temporary 64-byte node, parent key, voxel count and child mask of every group of
// <= 8 sibling tasks).  There a thread loops over its four items and, for every item that heads a group, walks that group's tasks from
// global memory: 7 of 32 lanes are active in that walk (profiles/r4e_r4e_octree.json), and the kernel is issue-bound.  Here every thread
// walks its four tasks ONCE, in registers, with all lanes active, filling the slots of the node under construction; a head closes the
// node before it (the only divergent part: prefix over the eight counts + the stores), and the last node of a thread continues into the
// <= 7 tasks of the following threads.
__global__ void __launch_bounds__( OB_BLOCK ) upperGroupDagKernel( const uint64_t* __restrict__ taskMorton, const uint32_t* __restrict__ taskChild, const uint32_t* __restrict__ taskVox,
																   uint64_t n, ScanState st, uint64_t* __restrict__ nextMorton, uint32_t* __restrict__ nextVox,
																   TraceNode* __restrict__ groupNode, uint8_t* __restrict__ groupMask )
{
	const uint32_t tile = scanTakeTicket( st ); // has a __syncthreads
	const uint64_t base = (uint64_t)tile * OB_TILE + (uint64_t)threadIdx.x * OB_ITEMS;
	static_assert( OB_ITEMS == 4, "vector loads below assume four tasks per thread" );
	uint64_t m[OB_ITEMS];
	uint32_t c[OB_ITEMS], v[OB_ITEMS];
	if( base + OB_ITEMS <= n )
	{
		const ulonglong2 m01 = *reinterpret_cast<const ulonglong2*>( taskMorton + base ), m23 = *reinterpret_cast<const ulonglong2*>( taskMorton + base + 2 );
		const uint4 cc = *reinterpret_cast<const uint4*>( taskChild + base ), vv = *reinterpret_cast<const uint4*>( taskVox + base );
		m[0] = m01.x, m[1] = m01.y, m[2] = m23.x, m[3] = m23.y;
		c[0] = cc.x, c[1] = cc.y, c[2] = cc.z, c[3] = cc.w;
		v[0] = vv.x, v[1] = vv.y, v[2] = vv.z, v[3] = vv.w;
	}
	else
	{
#pragma unroll
		for( int i = 0; i < OB_ITEMS; i++ )
		{
			const bool in = base + i < n;
			m[i] = in ? taskMorton[base + i] : 0ull;
			c[i] = in ? taskChild[base + i] : SENTINEL;
			v[i] = in ? taskVox[base + i] : 0u;
		}
	}
	uint32_t heads = 0, nHeads = 0;
	uint64_t prev = ( base > 0 && base <= n ) ? taskMorton[base - 1] : 0ull;
#pragma unroll
	for( int i = 0; i < OB_ITEMS; i++ )
	{
		const uint64_t idx = base + i;
		const bool head = ( idx < n ) && ( idx == 0 || ( m[i] >> 3 ) != ( prev >> 3 ) );
		prev = m[i];
		if( head )
		{
			heads |= 1u << i;
			nHeads++;
		}
	}
	uint32_t total;
	const uint32_t excl = blockExclusiveScan<OB_BLOCK>( nHeads, &total );
	const uint64_t prefix = scanTilePrefix( st, tile, total );
	uint32_t g = (uint32_t)prefix + excl;

	uint32_t children[8], cnt[8];
	uint32_t mask = 0;
	uint64_t parent = 0;
	bool open = false;
	auto place = [&]( const uint64_t mj, const uint32_t ch, const uint32_t nv ) {
		const uint32_t s = (uint32_t)( mj & 7u );
		mask |= 1u << s;
#pragma unroll
		for( int q = 0; q < 8; q++ )
		{
			if( q == (int)s )
			{
				children[q] = ch;
				cnt[q] = nv;
			}
		}
	};
	auto flush = [&]() {
		// IntersectorOctree.hpp:88-95: exclusive prefix of the voxel counts over the 8 slots
		uint32_t psum[8];
		uint32_t sum = 0;
#pragma unroll
		for( int q = 0; q < 8; q++ )
		{
			psum[q] = sum;
			sum += cnt[q];
		}
		nextMorton[g] = parent;
		nextVox[g] = sum;
		groupMask[g] = (uint8_t)mask;
		storeNode( &groupNode[g], children, psum );
		g++;
	};
#pragma unroll
	for( int i = 0; i < OB_ITEMS; i++ )
	{
		if( ( heads >> i ) & 1u )
		{
			if( open ) flush();
			parent = m[i] >> 3;
			mask = 0;
#pragma unroll
			for( int q = 0; q < 8; q++ )
			{
				children[q] = SENTINEL;
				cnt[q] = 0;
			}
			open = true;
		}
		if( open && base + i < n ) place( m[i], c[i], v[i] );
	}
	if( open )
	{
		for( uint64_t j = base + OB_ITEMS; j < n; j++ )
		{
			const uint64_t mj = taskMorton[j];
			if( ( mj >> 3 ) != parent ) break;
			place( mj, taskChild[j], taskVox[j] );
		}
		flush();
	}
}

// ---- DAG, leaf level: <= 255 distinct masks ----
__global__ void __launch_bounds__( 256 ) leafRankKernel( const uint32_t* __restrict__ leafFirst, uint32_t* __restrict__ leafIndexOfMask,
														 TraceNode* __restrict__ nodes, uint8_t* __restrict__ masks, LevelState* __restrict__ ls )
{
	__shared__ uint32_t s_first[256];
	const uint32_t mMask = threadIdx.x;
	const uint32_t f = leafFirst[mMask];
	s_first[mMask] = f;
	__syncthreads();
	uint32_t rank = 0, present = 0;
	for( int q = 0; q < 256; q++ )
	{
		const uint32_t o = s_first[q];
		if( o != SENTINEL )
		{
			present++;
			if( o < f ) rank++;
		}
	}
	if( f != SENTINEL )
	{
		leafIndexOfMask[mMask] = rank;
		uint32_t children[8], psum[8];
#pragma unroll
		for( int q = 0; q < 8; q++ )
		{
			children[q] = SENTINEL;
			psum[q] = __popc( mMask & ( ( 1u << q ) - 1u ) ); // one voxel per present child
		}
		storeNode( &nodes[rank], children, psum );
		masks[rank] = (uint8_t)mMask;
	}
	if( threadIdx.x == 0 )
	{
		ls->levelBase[0] = 0;
		ls->levelBase[1] = present;
	}
}
__global__ void __launch_bounds__( 256 ) leafAssignKernel( const uint8_t* __restrict__ groupMask, uint32_t G, const uint32_t* __restrict__ leafIndexOfMask, uint32_t* __restrict__ nextChild, int wide )
{
	__shared__ uint32_t s_idx[256];
	s_idx[threadIdx.x] = leafIndexOfMask[threadIdx.x];
	__syncthreads();
	const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	if( g >= G ) return;
	const uint32_t mask = groupMask[g];
	nextChild[g] = wide ? s_idx[mask] : ( s_idx[mask] | ( mask << 24 ) );
}

// ---- DAG, upper levels ----
__device__ __forceinline__ void loadKey( const TraceNode* node, uint32_t key[8] )
{
	const uint4* p = (const uint4*)node;
#pragma unroll
	for( int q = 0; q < 4; q++ )
	{
		const uint4 v = p[q];
		key[2 * q] = v.x;
		key[2 * q + 1] = v.z;
	}
}
__device__ __forceinline__ uint32_t hashKey( const uint32_t key[8] )
{
	Murmur32 h( 0x2545F491u );
#pragma unroll
	for( int q = 0; q < 8; q++ ) h.combine( key[q] );
	return h.getHash();
}
__device__ __forceinline__ bool sameKey( const uint32_t a[8], const uint32_t b[8] )
{
	bool eq = true;
#pragma unroll
	for( int q = 0; q < 8; q++ ) eq = eq && ( a[q] == b[q] );
	return eq;
}
__global__ void __launch_bounds__( 256 ) dagInsertKernel( const TraceNode* __restrict__ groupNode, uint32_t G, uint32_t* __restrict__ table, uint32_t tableMask )
{
	const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	if( g >= G ) return;
	uint32_t key[8];
	loadKey( &groupNode[g], key );
	uint32_t h = hashKey( key ) & tableMask;
	for( ;; )
	{
		uint32_t v = *(volatile uint32_t*)&table[h];
		if( v == SENTINEL )
		{
			v = atomicCAS( &table[h], SENTINEL, g );
			if( v == SENTINEL ) return; // claimed an empty slot
		}
		// the slot belongs to group v's key (it only ever changes to a smaller group with the SAME key)
		uint32_t other[8];
		loadKey( &groupNode[v], other );
		if( sameKey( key, other ) )
		{
			if( g < v ) atomicMin( &table[h], g );
			return;
		}
		h = ( h + 1 ) & tableMask;
	}
}
__global__ void __launch_bounds__( OB_BLOCK ) dagLookupKernel( const TraceNode* __restrict__ groupNode, const uint8_t* __restrict__ groupMask, uint32_t G,
															  const uint32_t* __restrict__ table, uint32_t tableMask, ScanState st,
															  uint32_t* __restrict__ rep, uint32_t* __restrict__ rankOf,
															  TraceNode* __restrict__ nodes, uint8_t* __restrict__ masks, LevelState* __restrict__ ls, int level )
{
	const uint32_t tile = scanTakeTicket( st );
	const uint32_t g = tile * OB_BLOCK + threadIdx.x;
	uint32_t isFirst = 0;
	if( g < G )
	{
		uint32_t key[8];
		loadKey( &groupNode[g], key );
		uint32_t h = hashKey( key ) & tableMask;
		uint32_t r;
		for( ;; )
		{
			r = table[h];
			if( r == g ) break;
			uint32_t other[8];
			loadKey( &groupNode[r], other );
			if( sameKey( key, other ) ) break;
			h = ( h + 1 ) & tableMask;
		}
		rep[g] = r;
		isFirst = ( r == g ) ? 1u : 0u;
	}
	uint32_t total;
	const uint32_t excl = blockExclusiveScan<OB_BLOCK>( isFirst, &total );
	const uint64_t prefix = scanTilePrefix( st, tile, total );
	const uint32_t base = ls->levelBase[level];
	if( isFirst )
	{
		const uint32_t rank = (uint32_t)prefix + excl;
		rankOf[g] = rank;
		const uint4* src = (const uint4*)&groupNode[g];
		uint4* dst = (uint4*)&nodes[base + rank];
#pragma unroll
		for( int q = 0; q < 4; q++ ) dst[q] = src[q];
		masks[base + rank] = groupMask[g];
	}
	if( tile == gridDim.x - 1 && threadIdx.x == 0 ) ls->levelBase[level + 1] = base + (uint32_t)( prefix + total );
}
__global__ void __launch_bounds__( 256 ) dagAssignKernel( const uint32_t* __restrict__ rep, const uint32_t* __restrict__ rankOf, const uint8_t* __restrict__ groupMask, uint32_t G,
														  const LevelState* __restrict__ ls, int level, uint32_t* __restrict__ nextChild, int wide )
{
	const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	if( g >= G ) return;
	const uint32_t idx = ls->levelBase[level] + rankOf[rep[g]];
	nextChild[g] = wide ? idx : ( idx | ( (uint32_t)groupMask[g] << 24 ) );
}
// embedMasks (voxCommon.hpp:183-195, voxKernel.cu:426-435) for a tree that was built with plain 32-bit child indices because its
// node count before deduplication did not fit 24 bits, but whose final count does
__global__ void __launch_bounds__( 256 ) embedMasksKernel( TraceNode* __restrict__ nodes, const uint8_t* __restrict__ masks, uint32_t nNodes )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if( ( i >> 3 ) >= nNodes ) return;
	uint2* slot = &nodes[i >> 3].slot[i & 7u];
	const uint32_t c = slot->x;
	if( c != SENTINEL ) slot->x = c | ( (uint32_t)masks[c] << 24 );
}

int zeroedScanState( hako_ctx* ctx, cudaStream_t s, uint64_t tiles, void** mem, ScanState* st )
{
	HAKO_CUDA( cudaMallocFromPoolAsync( mem, scanStateBytes( tiles ), ctx->pool, s ) );
	HAKO_CUDA( cudaMemsetAsync( *mem, 0, scanStateBytes( tiles ), s ) );
	*st = makeScanState( *mem, tiles );
	return 0;
}
} // namespace

int buildOctreeDevice( hako_ctx* ctx, cudaStream_t s, const uint64_t* mortons, uint64_t n, int gridRes, int dag,
					   TraceNode** nodesOut, uint8_t** masksOut, uint32_t* nNodesOut, uint32_t* launchesOut, int* wideOut, const uint64_t* levelBoundaries )
{
	if( wideOut ) *wideOut = 0;
	*nodesOut = nullptr;
	*masksOut = nullptr;
	*nNodesOut = 0;
	int levels = 0;
	while( ( 1 << levels ) < gridRes ) levels++;
	if( n == 0 || levels == 0 ) return 0; // `while( 1 < wide )` never runs: no nodes
	if( levels > MAX_LEVELS )
	{
		setError( "octree: gridRes %d exceeds 21 bits per axis (morton.hpp:73)", gridRes );
		return HAKO_E_LIMIT;
	}
	const uint64_t launches0 = ctx->launches.load();

	// 1. group counts per level: from the caller (the merge stage counted them in its own read of the keys), else one pass + host read
	uint64_t bounds[22];
	if( levelBoundaries )
		for( int j = 0; j < 22; j++ ) bounds[j] = levelBoundaries[j];
	else
	{
		unsigned long long* hist = nullptr;
		HAKO_TRY( dalloc( ctx, &hist, 22, s ) );
		HAKO_CUDA( cudaMemsetAsync( hist, 0, 22 * 8, s ) );
		const uint32_t blocks = (uint32_t)min( (uint64_t)ctx->smCount * 8, ( n + 255 ) / 256 );
		levelHistKernel<<<blocks, 256, 0, s>>>( mortons, n, hist );
		ctx->launches.fetch_add( 1, std::memory_order_relaxed );
		const cudaError_t le = cudaGetLastError();
		if( le != cudaSuccess )
		{
			dfree( hist, s );
			return cudaFail( le, "levelHistKernel", __FILE__, __LINE__ );
		}
		HAKO_CUDA( cudaMemcpyAsync( ctx->pinned, hist, 22 * 8, cudaMemcpyDeviceToHost, s ) );
		HAKO_CUDA( cudaStreamSynchronize( s ) );
		ctx->stats.hostSyncs++;
		dfree( hist, s );
		for( int j = 0; j < 22; j++ ) bounds[j] = ctx->pinned[j];
	}
	uint64_t G[MAX_LEVELS + 1]; // G[l] = groups (= nodes before dedupe) at level l
	for( int l = levels - 1; l >= 0; l-- )
	{
		// a boundary between i-1 and i exists at level l iff top/3 >= l + 1
		uint64_t b = 0;
		for( int j = l + 1; j < 22; j++ ) b += bounds[j];
		G[l] = 1 + b;
	}
	uint64_t pool = 0;
	for( int l = 0; l < levels; l++ ) pool += ( dag && l == 0 ) ? ( G[0] < 255 ? G[0] : 255 ) : G[l];
	uint64_t maxUpper = 0; // largest group count of a level >= 1 (temp node storage / hash table sizing)
	for( int l = 1; l < levels; l++ ) maxUpper = G[l] > maxUpper ? G[l] : maxUpper;
	if( pool >= 0xFFFFFFFFull )
	{
		setError( "octree: %llu nodes before deduplication exceed the 32-bit node index", (unsigned long long)pool );
		return HAKO_E_LIMIT;
	}
	// The reference keeps a child's mask in the top byte of its 32-bit pointer (voxCommon.hpp:183-195), so 2^24 - 1 nodes is the
	// ceiling of its format.  When the node count BEFORE deduplication does not fit, the tree is built with plain 32-bit child
	// indices; if the final count fits 24 bits after all the masks are embedded at the end (one more pass), otherwise the tree
	// stays in the wide layout (the traversal then reads the child's mask from the side array).  HAKO_FORCE_WIDE=1: tests.
	const char* forceEnv = getenv( "HAKO_FORCE_WIDE" );
	const bool forceWide = forceEnv && atoi( forceEnv ) != 0;
	const int wide = ( pool >= 0xFFFFFFull || forceWide ) ? 1 : 0;

	TraceNode* nodes = nullptr;
	uint8_t* masks = nullptr;
	uint64_t* mortonA = nullptr, *mortonB = nullptr;
	uint32_t *childA = nullptr, *childB = nullptr, *voxA = nullptr, *voxB = nullptr;
	TraceNode* groupNode = nullptr;
	uint8_t* groupMask = nullptr;
	uint32_t *table = nullptr, *rep = nullptr, *rankOf = nullptr, *leafTables = nullptr;
	LevelState* ls = nullptr;
	void* scanMem = nullptr;
	uint32_t tableSize = 1;
	while( tableSize < 2 * maxUpper ) tableSize <<= 1;
	int r = 0;
	do
	{
		if( ( r = dalloc( ctx, &nodes, pool, s ) ) ) break;
		if( ( r = dalloc( ctx, &masks, pool, s ) ) ) break;
		if( ( r = dalloc( ctx, &mortonA, G[0], s ) ) ) break;
		if( ( r = dalloc( ctx, &childA, G[0], s ) ) ) break;
		if( ( r = dalloc( ctx, &voxA, G[0], s ) ) ) break;
		const uint64_t g1 = levels > 1 ? G[1] : 1;
		if( ( r = dalloc( ctx, &mortonB, g1, s ) ) ) break;
		if( ( r = dalloc( ctx, &childB, g1, s ) ) ) break;
		if( ( r = dalloc( ctx, &voxB, g1, s ) ) ) break;
		if( dag )
		{
			if( ( r = dalloc( ctx, &groupMask, G[0], s ) ) ) break;
			if( ( r = dalloc( ctx, &groupNode, maxUpper, s ) ) ) break;
			if( ( r = dalloc( ctx, &table, (uint64_t)tableSize, s ) ) ) break;
			if( ( r = dalloc( ctx, &rep, maxUpper, s ) ) ) break;
			if( ( r = dalloc( ctx, &rankOf, maxUpper, s ) ) ) break;
			if( ( r = dalloc( ctx, &leafTables, 512, s ) ) ) break;
			if( ( r = dalloc( ctx, &ls, 1, s ) ) ) break;
		}
		cudaError_t e = cudaSuccess;
#define CK( call )                                                \
	if( ( e = ( call ) ) != cudaSuccess )                         \
	{                                                             \
		r = cudaFail( e, #call, __FILE__, __LINE__ );             \
		break;                                                    \
	}
		auto kLeafDag = groupBuildKernel<true, true>;
		auto kUpperDag = groupBuildKernel<false, true>;
		auto kLeafNaive = groupBuildKernel<true, false>;
		auto kUpperNaive = groupBuildKernel<false, false>;
		const uint64_t* tM = mortons;
		const uint32_t *tC = nullptr, *tV = nullptr;
		uint64_t* nM = mortonA;
		uint32_t *nC = childA, *nV = voxA;
		uint64_t nTasks = n;
		uint32_t naiveBase = 0;
		for( int l = 0; l < levels && r == 0; l++ )
		{
			static const bool genericLeaf = getenv( "HAKO_OCTREE_GENERIC_LEAF" ) && atoi( getenv( "HAKO_OCTREE_GENERIC_LEAF" ) ) != 0; // A/B knob
			const bool fastLeaf = dag && l == 0 && !genericLeaf && ( ( (uintptr_t)mortons & 15u ) == 0 );
			const uint64_t tileSize = fastLeaf ? LG_TILE : OB_TILE;
			const uint64_t tiles = ( nTasks + tileSize - 1 ) / tileSize;
			ScanState st;
			if( ( r = zeroedScanState( ctx, s, tiles, &scanMem, &st ) ) ) break;
			const uint32_t Gl = (uint32_t)G[l];
			if( dag )
			{
				if( l == 0 )
				{
					CK( cudaMemsetAsync( leafTables, 0xFF, 512 * 4, s ) );
					if( fastLeaf )
						HAKO_LAUNCH_R( ctx, r, leafGroupDagKernel, (uint32_t)tiles, OB_BLOCK, 0, s, tM, nTasks, st, nM, nV, groupMask, leafTables );
					else
						HAKO_LAUNCH_R( ctx, r, kLeafDag, (uint32_t)tiles, OB_BLOCK, 0, s, tM, tC, tV, nTasks, st, nM, nC, nV, groupNode, groupMask,
									 leafTables, nodes, masks, 0u, wide );
					HAKO_LAUNCH_R( ctx, r, leafRankKernel, 1, 256, 0, s, leafTables, leafTables + 256, nodes, masks, ls );
					HAKO_LAUNCH_R( ctx, r, leafAssignKernel, divUp( Gl, 256 ), 256, 0, s, groupMask, Gl, leafTables + 256, nC, wide );
				}
				else
				{
					static const bool genericUpper = getenv( "HAKO_OCTREE_GENERIC_UPPER" ) && atoi( getenv( "HAKO_OCTREE_GENERIC_UPPER" ) ) != 0; // A/B knob
					if( !genericUpper )
						HAKO_LAUNCH_R( ctx, r, upperGroupDagKernel, (uint32_t)tiles, OB_BLOCK, 0, s, tM, tC, tV, nTasks, st, nM, nV, groupNode, groupMask );
					else
						HAKO_LAUNCH_R( ctx, r, kUpperDag, (uint32_t)tiles, OB_BLOCK, 0, s, tM, tC, tV, nTasks, st, nM, nC, nV, groupNode, groupMask,
									 (uint32_t*)nullptr, nodes, masks, 0u, wide );
					uint32_t tsz = 1;
					while( tsz < 2 * (uint64_t)Gl ) tsz <<= 1;
					CK( cudaMemsetAsync( table, 0xFF, (size_t)tsz * 4, s ) );
					HAKO_LAUNCH_R( ctx, r, dagInsertKernel, divUp( Gl, 256 ), 256, 0, s, groupNode, Gl, table, tsz - 1 );
					cudaFreeAsync( scanMem, s );
					scanMem = nullptr;
					const uint64_t tiles2 = ( (uint64_t)Gl + OB_BLOCK - 1 ) / OB_BLOCK;
					ScanState st2;
					if( ( r = zeroedScanState( ctx, s, tiles2, &scanMem, &st2 ) ) ) break;
					HAKO_LAUNCH_R( ctx, r, dagLookupKernel, (uint32_t)tiles2, OB_BLOCK, 0, s, groupNode, groupMask, Gl, table, tsz - 1, st2, rep, rankOf, nodes, masks, ls, l );
					HAKO_LAUNCH_R( ctx, r, dagAssignKernel, divUp( Gl, 256 ), 256, 0, s, rep, rankOf, groupMask, Gl, ls, l, nC, wide );
				}
			}
			else
			{
				if( l == 0 )
					HAKO_LAUNCH_R( ctx, r, kLeafNaive, (uint32_t)tiles, OB_BLOCK, 0, s, tM, tC, tV, nTasks, st, nM, nC, nV, groupNode, groupMask,
								 (uint32_t*)nullptr, nodes, masks, naiveBase, wide );
				else
					HAKO_LAUNCH_R( ctx, r, kUpperNaive, (uint32_t)tiles, OB_BLOCK, 0, s, tM, tC, tV, nTasks, st, nM, nC, nV, groupNode, groupMask,
								 (uint32_t*)nullptr, nodes, masks, naiveBase, wide );
				naiveBase += Gl;
			}
			cudaFreeAsync( scanMem, s );
			scanMem = nullptr;
			// the groups of this level are the tasks of the next
			tM = nM, tC = nC, tV = nV;
			nTasks = Gl;
			if( nM == mortonA )
				nM = mortonB, nC = childB, nV = voxB;
			else
				nM = mortonA, nC = childA, nV = voxA;
		}
		if( r ) break;
		// 2. final node count
		uint32_t nNodes = naiveBase;
		if( dag )
		{
			CK( cudaMemcpyAsync( ctx->pinned, &ls->levelBase[levels], 4, cudaMemcpyDeviceToHost, s ) );
			CK( cudaStreamSynchronize( s ) );
			ctx->stats.hostSyncs++;
			nNodes = *(uint32_t*)ctx->pinned;
		}
		int stayWide = 0;
		if( wide )
		{
			if( nNodes < 0xFFFFFFu && !forceWide )
				HAKO_LAUNCH_R( ctx, r, embedMasksKernel, divUp( (uint64_t)nNodes * 8, 256 ), 256, 0, s, nodes, masks, nNodes );
			else
				stayWide = 1;
		}
		if( wideOut ) *wideOut = stayWide;
		if( stayWide && !wideOut )
		{
			setError( "octree: %u nodes exceed the 24-bit node index (voxCommon.hpp:183-195)", nNodes );
			r = HAKO_E_LIMIT;
			break;
		}
		// 3. exact-size resident arrays (the pool was sized before dedupe)
		if( (uint64_t)nNodes == pool )
		{
			*nodesOut = nodes, *masksOut = masks;
			nodes = nullptr, masks = nullptr;
		}
		else
		{
			if( ( r = dalloc( ctx, nodesOut, nNodes, s ) ) ) break;
			if( ( r = dalloc( ctx, masksOut, nNodes, s ) ) ) break;
			CK( cudaMemcpyAsync( *nodesOut, nodes, (size_t)nNodes * sizeof( TraceNode ), cudaMemcpyDeviceToDevice, s ) );
			CK( cudaMemcpyAsync( *masksOut, masks, nNodes, cudaMemcpyDeviceToDevice, s ) );
		}
		*nNodesOut = nNodes;
#undef CK
	} while( 0 );
	dfree( nodes, s ), dfree( masks, s ), dfree( mortonA, s ), dfree( mortonB, s ), dfree( childA, s ), dfree( childB, s ), dfree( voxA, s ), dfree( voxB, s );
	dfree( groupNode, s ), dfree( groupMask, s ), dfree( table, s ), dfree( rep, s ), dfree( rankOf, s ), dfree( leafTables, s ), dfree( ls, s );
	if( scanMem ) cudaFreeAsync( scanMem, s );
	if( r )
	{
		dfree( *nodesOut, s ), dfree( *masksOut, s );
		*nodesOut = nullptr, *masksOut = nullptr;
	}
	if( launchesOut ) *launchesOut = (uint32_t)( ctx->launches.load() - launches0 );
	return r;
}
} // namespace hako
