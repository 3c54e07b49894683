// traverse.cuh -- Revelles parametric ray/octree traversal on the 64-byte trace layout.
//
// Follows octreeTraverse_EfficientParametric (voxCommon.hpp:231-423, SMALL_STACK + ENABLE_EMBEDED_MASK branch)
// operation for operation in fp32 (the library is built with -fmad=false, IEEE division), so t / nMajor / vIndex
// are bit-identical to the CPU IntersectorOctree::intersect (IntersectorOctree.hpp:248-257).
//
// What is different from the reference's GPU half:
//   - no DynamicAllocatorGPU (IntersectorOctreeGPU.hpp:277-361, a 1 GiB global stack pool with CAS locks): the stack is
//     per-thread, 6 words per entry (one 16-B + one 8-B local store), at most one entry per tree level;
//   - `scale` is not stored: it is 2^-level, rebuilt exactly from the level counter;
//   - one descent = ONE 8-byte read-only load (child pointer + voxel prefix sum share a slot) instead of two 4-byte
//     loads 32 bytes apart in a 68-byte node.
#pragma once

#include "common.cuh"

namespace hako
{
#define HAKO_STACK_DEPTH 22 // <= one push per level, 21-bit Morton axes (morton.hpp:73)

// 2^-level, exact (what `scale *= 0.5f` per descent produces, voxCommon.hpp:385)
__device__ __forceinline__ float levelScale( uint32_t level ) { return __uint_as_float( ( 127u - level ) << 23 ); }

// The traversal as a resumable object: begin() sets a ray up (mirroring, clamped 1/rd, root slab test), step() advances it
// by ONE uniform step.  Persistent kernels keep one Traversal per lane and refill a lane the moment its ray ends, so a warp
// never waits for its slowest ray; octreeTraverseImpl() below is the plain "run to completion" wrapper.
//
// One uniform step per call (instead of the reference's visit loop + inner sibling loop + pop): every lane of a warp
// executes the same straight-line body whatever its depth in the tree, so the warp does not split into the "computing t0 /
// walking siblings / popping" groups the nested form produces.  Each value is the same fp32 expression of ( t1, dt, level,
// childMask ) as in voxCommon.hpp:317-411, so recomputing it per step is bit-neutral.
enum : int
{
	TRAV_CONTINUE = 0,
	TRAV_HIT = 1,
	TRAV_MISS = 2
};
// the per-thread stack lives in its own object (local memory); keeping it out of Traversal lets the scalar state stay in registers
struct TravStack
{
	uint4 a[HAKO_STACK_DEPTH]; // node, tx1, ty1, tz1
	uint2 b[HAKO_STACK_DEPTH]; // childMask | level << 8, nVoxelSkipped
	__device__ __forceinline__ void push( int& sp, const uint4 na, const uint2 nb )
	{
		a[sp] = na;
		b[sp] = nb;
		sp++;
	}
	__device__ __forceinline__ void pop( int& sp, uint4& oa, uint2& ob )
	{
		sp--;
		oa = a[sp];
		ob = b[sp];
	}
};
// The same stack in SHARED memory, entry-major and thread-minor: entry k of thread t is a[k * BLOCK + t] / b[k * BLOCK + t], so a thread's
// banks depend on t only and a warp's pushes and pops are conflict-free whatever the depths of its lanes.  Local memory interleaves the
// lanes of a warp word by word instead: the 24-byte entries of lanes at d different depths cost 6 d L1 tag look-ups per push and per pop,
// and in the trace waves of the path tracer that traffic is 85 % of all L1 sectors (profiles/r3s).  `depth` entries per thread (the tree's
// number of levels: at most one push per level) bound the threads per SM: depth x 24 B x threads <= 227 KB, 768 threads up to 4096^3.
// PACKED squeezes an entry into 22 bytes (the child mask in progress is 3 bits, the level 5) at the price of a few instructions per push
// and pop, which keeps 768 threads at 8192^3.  Tried and much slower than either: the 8-byte half of an entry left in local memory, and
// only the first K entries in shared memory with the rest in local memory (profiles/r4l, r4m).
template <bool PACKED>
struct TravStackShared
{
	uint4* a;	 // node, tx1, ty1, tz1
	uint2* b;	 // !PACKED: childMask | level << 8 [| mask << 16], nVoxelSkipped
	uint32_t* s; // PACKED: nVoxelSkipped
	uint16_t* c; // PACKED: childMask (3 bits) | level (5 bits) | mask of the node (wide trees, 8 bits)
	uint32_t stride; // threads per CTA
	static constexpr int ENTRY_BYTES = PACKED ? 22 : 24;
	// pointers into a CTA's dynamic shared memory for `depth` entries per thread
	__device__ __forceinline__ void bind( unsigned char* smem, int depth, uint32_t threads, uint32_t tid )
	{
		stride = threads;
		a = (uint4*)smem + tid;
		b = (uint2*)( smem + (size_t)depth * threads * 16 ) + tid;
		s = (uint32_t*)( smem + (size_t)depth * threads * 16 ) + tid;
		c = (uint16_t*)( smem + (size_t)depth * threads * 20 ) + tid;
	}
	__device__ __forceinline__ void push( int& sp, const uint4 na, const uint2 nb )
	{
		const uint32_t i = (uint32_t)sp * stride;
		a[i] = na;
		if( PACKED )
		{
			s[i] = nb.y;
			c[i] = (uint16_t)( ( nb.x & 7u ) | ( ( ( nb.x >> 8 ) & 31u ) << 3 ) | ( ( nb.x >> 16 ) << 8 ) );
		}
		else
			b[i] = nb;
		sp++;
	}
	__device__ __forceinline__ void pop( int& sp, uint4& oa, uint2& ob )
	{
		sp--;
		const uint32_t i = (uint32_t)sp * stride;
		oa = a[i];
		if( PACKED )
		{
			const uint32_t p = c[i];
			ob = make_uint2( ( p & 7u ) | ( ( ( p >> 3 ) & 31u ) << 8 ) | ( ( p >> 8 ) << 16 ), s[i] );
		}
		else
			ob = b[i];
	}
};
// Optional shared-memory copy of the LAST `count` nodes of the array = the top of the tree (numbering is bottom-up, the root is
// last): north_star's "octree top staged in shared memory".  Built per CTA by stageTop(); step<.., TOP = true> reads a node's slot
// from it when node >= base.  Measured A/B in DESIGN.md section 7a (HAKO_SMEM_TOP); off by default.
struct TopCache
{
	const uint2* slots; // [count][8]
	uint32_t base;		// numberOfNodes - count
};
template <int TOPK>
__device__ __forceinline__ TopCache stageTop( const OctreeView& v, uint2* s_top )
{
	const uint32_t count = v.numberOfNodes < (uint32_t)TOPK ? v.numberOfNodes : (uint32_t)TOPK;
	const uint32_t base = v.numberOfNodes - count;
	const uint2* src = &v.nodes[base].slot[0];
	for( uint32_t i = threadIdx.x; i < count * 8u; i += blockDim.x ) s_top[i] = __ldg( src + i );
	__syncthreads();
	return TopCache{ s_top, base };
}

struct Traversal
{
	int sp;
	float dtx, dty, dtz;
	float tx1, ty1, tz1;
	uint32_t node, level, childMask, skipped, vMask;
	bool ordinary; // every root slab distance is below 1e37 in magnitude: step<true> is bit-identical to step<false>
	uint32_t nmask; // WIDE trees only: the mask of `node` (narrow trees carry it in the top byte of `node`)
	// results of a TRAV_HIT step
	float t;
	int nMajor;

	// false: the ray misses the root box (or the tree is empty).
	// WIDE: the tree has plain 32-bit child indices (>= 2^24 - 1 nodes); a node's mask comes from v.masks.
	template <bool WIDE = false>
	__device__ __forceinline__ bool begin( const OctreeView& v, V3 ro, V3 rd )
	{
		if( v.numberOfNodes == 0 ) return false;
		// voxCommon.hpp:240-260: mirror negative directions
		V3 inv = { 1.0f / rd.x, 1.0f / rd.y, 1.0f / rd.z };
		vMask = 0;
		if( inv.x < 0.0f )
		{
			vMask |= 1u;
			inv.x = -inv.x;
			ro.x = v.lower.x + v.upper.x - ro.x;
		}
		if( inv.y < 0.0f )
		{
			vMask |= 2u;
			inv.y = -inv.y;
			ro.y = v.lower.y + v.upper.y - ro.y;
		}
		if( inv.z < 0.0f )
		{
			vMask |= 4u;
			inv.z = -inv.z;
			ro.z = v.lower.z + v.upper.z - ro.z;
		}
		// :265-269: clamp 1/rd so that (bound - ro) * inv never overflows to inf (no inf * 0)
		const V3 lo = v.lower - ro, up = v.upper - ro;
		inv.x = smin( inv.x, HAKO_MAX_FLOAT / smax( smax( sabs( lo.x ), sabs( up.x ) ), 1.0f ) );
		inv.y = smin( inv.y, HAKO_MAX_FLOAT / smax( smax( sabs( lo.y ), sabs( up.y ) ), 1.0f ) );
		inv.z = smin( inv.z, HAKO_MAX_FLOAT / smax( smax( sabs( lo.z ), sabs( up.z ) ), 1.0f ) );
		// :271-278 root slab test
		const float t0x = lo.x * inv.x, t0y = lo.y * inv.y, t0z = lo.z * inv.z;
		tx1 = up.x * inv.x, ty1 = up.y * inv.y, tz1 = up.z * inv.z;
		if( min3( tx1, ty1, tz1 ) < max3( t0x, t0y, t0z ) ) return false;
		dtx = tx1 - t0x, dty = ty1 - t0y, dtz = tz1 - t0z;
		const float big = 1.0e37f;
		ordinary = sabs( t0x ) < big && sabs( t0y ) < big && sabs( t0z ) < big && sabs( tx1 ) < big && sabs( ty1 ) < big && sabs( tz1 ) < big;
		sp = 0;
		const uint32_t root = v.numberOfNodes - 1; // the root is the last node (IntersectorOctree.hpp:252)
		nmask = (uint32_t)__ldg( v.masks + root );
		node = WIDE ? root : ( root | ( nmask << 24 ) );
		level = 0;
		childMask = 0xFFu; // "not chosen yet" (reference: 0xFFFFFFFF)
		skipped = 0;
		return true;
	}

	// hit: t / nMajor are set, `skipped` is vIndex.
	// FINITE = true may be used when begin() found every root slab distance below 1e37 in magnitude (`ordinary`): all later t
	// values lie inside the root interval, so none is inf / NaN and fmaxf / fminf select the same operand as the reference's
	// compare-select max / min except for the sign of a zero, which no consumer below can see (==, <, and t only when
	// 0 < lmax): one FMNMX3 instead of two FSETP + FSEL pairs, ~6 of ~95 instructions per step.  Rays with huge slab distances
	// (a zero or denormal direction component against the clamped 1/rd) can produce inf - inf = NaN, where the two forms
	// differ; they take the exact path (tests/test_gpu_traverse.py::test_random_rays_and_shadow has such rays).
	template <bool FINITE = false, bool WIDE = false, bool TOP = false, class STACK = TravStack>
	__device__ __forceinline__ int step( const OctreeView& v, const bool shadow, STACK& stk, const TopCache tc = TopCache{ nullptr, 0xFFFFFFFFu } )
	{
		const float scale = levelScale( level );
		// :317-319 t0 is recomputed from t1 each visit
		const float tx0 = tx1 - dtx * scale;
		const float ty0 = ty1 - dty * scale;
		const float tz0 = tz1 - dtz * scale;
		const float lmax = FINITE ? fmaxf( fmaxf( tx0, ty0 ), tz0 ) : max3( tx0, ty0, tz0 );

		bool pop = false;
		if( node == 0xFFFFFFFFu )
		{
			// :322-336 leaf voxel
			if( 0.0f < lmax )
			{
				t = lmax;
				nMajor = ( lmax == tx0 ) ? 1 : ( ( lmax == ty0 ) ? 2 : 0 );
				return TRAV_HIT;
			}
			pop = true;
		}
		else
		{
			// :338-348
			const float txM = 0.5f * ( tx0 + tx1 );
			const float tyM = 0.5f * ( ty0 + ty1 );
			const float tzM = 0.5f * ( tz0 + tz1 );
			if( childMask == 0xFFu ) childMask = ( txM < lmax ? 1u : 0u ) | ( tyM < lmax ? 2u : 0u ) | ( tzM < lmax ? 4u : 0u );
			const float x1 = ( childMask & 1u ) ? tx1 : txM;
			const float y1 = ( childMask & 2u ) ? ty1 : tyM;
			const float z1 = ( childMask & 4u ) ? tz1 : tzM;
			// :365-411 the child the ray is in now; the exit plane decides the next sibling
			const float umin = FINITE ? fminf( fminf( x1, y1 ), z1 ) : min3( x1, y1, z1 );
			const uint32_t mv = ( umin == x1 ) ? 1u : ( ( umin == y1 ) ? 2u : 4u );
			const bool hasNext = ( childMask & mv ) == 0;
			const uint32_t ci = childMask ^ vMask;
			childMask |= mv;
			const uint32_t present = WIDE ? ( ( nmask >> ci ) & 1u ) : ( ( node >> ( 24 + ci ) ) & 1u );
			if( present && !( umin < 0.0f ) )
			{
				if( hasNext )
				{
					stk.push( sp, make_uint4( node, __float_as_uint( tx1 ), __float_as_uint( ty1 ), __float_as_uint( tz1 ) ),
							  make_uint2( childMask | ( level << 8 ) | ( WIDE ? nmask << 16 : 0u ), skipped ) );
				}
				const uint32_t nidx = WIDE ? node : ( node & 0xFFFFFFu );
				const uint2 slot = ( TOP && nidx >= tc.base ) ? tc.slots[( nidx - tc.base ) * 8u + ci] : ldg_u2( &v.nodes[nidx].slot[ci] );
				node = slot.x;
				if( WIDE ) nmask = ( node != 0xFFFFFFFFu ) ? (uint32_t)__ldg( v.masks + node ) : 0u; // the second dependent load of the wide layout
				tx1 = x1;
				ty1 = y1;
				tz1 = z1;
				level++;
				childMask = 0xFFu;
				if( !shadow ) skipped += slot.y; // :388-391
			}
			else
				pop = !hasNext; // otherwise stay in this node: the next step looks at the sibling behind plane `mv`
		}
		if( pop )
		{
			// :414-422
			if( sp == 0 ) return TRAV_MISS;
			uint4 a;
			uint2 b;
			stk.pop( sp, a, b );
			node = a.x;
			tx1 = __uint_as_float( a.y );
			ty1 = __uint_as_float( a.z );
			tz1 = __uint_as_float( a.w );
			childMask = b.x & 0xFFu;
			level = WIDE ? ( ( b.x >> 8 ) & 0xFFu ) : ( b.x >> 8 );
			if( WIDE ) nmask = b.x >> 16;
			skipped = b.y;
		}
		return TRAV_CONTINUE;
	}
};

// returns true on a hit (and then writes t / nMajor / vIndex); a miss leaves the outputs untouched (voxRT.cpp:313-325).
// `shadow` may be a compile-time constant (the template wrapper below) or a per-lane runtime value.
template <bool TOP = false>
__device__ __forceinline__ bool octreeTraverseImpl( const OctreeView& v, V3 ro, V3 rd, const bool shadow, float* tOut, int* nMajorOut, uint32_t* vIndexOut,
													const TopCache tc = TopCache{ nullptr, 0xFFFFFFFFu } )
{
	Traversal tr;
	TravStack stk;
	int r;
	if( v.wide )
	{
		// trees beyond the reference format's 2^24 - 1 nodes
		if( !tr.template begin<true>( v, ro, rd ) ) return false;
		if( tr.ordinary )
		{
			do
			{
				r = tr.template step<true, true, TOP>( v, shadow, stk, tc );
			} while( r == TRAV_CONTINUE );
		}
		else
		{
			do
			{
				r = tr.template step<false, true, TOP>( v, shadow, stk, tc );
			} while( r == TRAV_CONTINUE );
		}
	}
	else
	{
		if( !tr.template begin<false>( v, ro, rd ) ) return false;
		if( tr.ordinary )
		{
			do
			{
				r = tr.template step<true, false, TOP>( v, shadow, stk, tc );
			} while( r == TRAV_CONTINUE );
		}
		else
		{
			do
			{
				r = tr.template step<false, false, TOP>( v, shadow, stk, tc );
			} while( r == TRAV_CONTINUE );
		}
	}
	if( r == TRAV_MISS ) return false;
	*tOut = tr.t;
	*nMajorOut = tr.nMajor;
	*vIndexOut = tr.skipped;
	return true;
}

template <bool SHADOW>
__device__ __forceinline__ bool octreeTraverse( const OctreeView& v, V3 ro, V3 rd, float* tOut, int* nMajorOut, uint32_t* vIndexOut )
{
	return octreeTraverseImpl( v, ro, rd, SHADOW, tOut, nMajorOut, vIndexOut );
}

// getHitN, voxCommon.hpp:564-577
__device__ __forceinline__ V3 getHitN( int major, V3 rd )
{
	switch( major )
	{
	case 0:
		return { 0.0f, 0.0f, 0.0f < rd.z ? -1.0f : 1.0f };
	case 1:
		return { 0.0f < rd.x ? -1.0f : 1.0f, 0.0f, 0.0f };
	case 2:
		return { 0.0f, 0.0f < rd.y ? -1.0f : 1.0f, 0.0f };
	}
	return { 0.0f, 0.0f, 0.0f };
}

// CameraPinhole::shoot, renderCommon.hpp:37-49: ((right * mix) * W) / H + up * mix + front
__device__ __forceinline__ void cameraShoot( const hako_camera& cam, int x, int y, float ox, float oy, int w, int h, V3* ro, V3* rd )
{
	const float xf = ( x + ox ) / w;
	const float yf = ( y + oy ) / h;
	const float a = mixf( -cam.tanHthetaY, cam.tanHthetaY, xf );
	const float b = mixf( cam.tanHthetaY, -cam.tanHthetaY, yf );
	rd->x = cam.right[0] * a * w / h + cam.up[0] * b + cam.front[0];
	rd->y = cam.right[1] * a * w / h + cam.up[1] * b + cam.front[1];
	rd->z = cam.right[2] * a * w / h + cam.up[2] * b + cam.front[2];
	*ro = { cam.o[0], cam.o[1], cam.o[2] };
}
// CameraPinhole::shootThinLens, renderCommon.hpp:50-75
__device__ __forceinline__ void cameraShootThinLens( const hako_camera& cam, int x, int y, float ox, float oy, int w, int h, float u0, float u1, V3* ro, V3* rd )
{
	const float xf = ( x + ox ) / w;
	const float yf = ( y + oy ) / h;
	const float fx = cam.focus * mixf( -cam.tanHthetaY, cam.tanHthetaY, xf ) * w / h;
	const float fy = cam.focus * mixf( cam.tanHthetaY, -cam.tanHthetaY, yf );
	const float fz = cam.focus;
	const float lx = mixf( -cam.lensR, cam.lensR, u0 );
	const float ly = mixf( -cam.lensR, cam.lensR, u1 );
	const float lz = 0.0f;
	const float dx = fx - lx, dy = fy - ly, dz = fz - lz;
	rd->x = cam.right[0] * dx + cam.up[0] * dy + cam.front[0] * dz;
	rd->y = cam.right[1] * dx + cam.up[1] * dy + cam.front[1] * dz;
	rd->z = cam.right[2] * dx + cam.up[2] * dy + cam.front[2] * dz;
	ro->x = cam.o[0] + cam.right[0] * lx + cam.up[0] * ly + cam.front[0] * lz;
	ro->y = cam.o[1] + cam.right[1] * lx + cam.up[1] * ly + cam.front[1] * lz;
	ro->z = cam.o[2] + cam.right[2] * lx + cam.up[2] * ly + cam.front[2] * lz;
}
} // namespace hako
