// voxelize.cu -- triangle -> voxel 6-separating / conservative overlap + Morton keys + barycentric colour.
//
// Arithmetic follows VTContext (voxelization.hpp:124-337), the voxel loop of voxRT.cpp:198-240 (== voxCount / voxelize,
// voxKernel.cu:58-166) and closestBarycentricCoordinateOnTriangle (vectorMath.hpp:249-284) operation for operation,
// including the x86 float->int cast semantics the CPU path relies on (f2i_x86).
//
// Parallel decomposition (different from the reference, which runs one thread per triangle twice and reserves output
// ranges with a global atomicAdd, so that its emission order is a race):
//   1. triColumns     one thread per triangle: number of x-columns of its clamped bounding box
//   2. scan           column offsets
//   3. columnVoxels<false>  one thread per (triangle, x) column: count the voxels of the column
//   4. scan           voxel offsets (+ total D, the single host read of this stage)
//   5. columnVoxels<true>   same enumeration, emit key + packed attribute at offset
// A column is a pure function of (triangle, x) (voxelization.hpp:230-336), so slicing by x is parity-safe, it bounds
// the work per thread (a 4096^3 bunny triangle covers ~7k voxels) and the emission order is deterministic: triangle-major,
// then x, y, z -- exactly the order of the CPU loop.
#include <algorithm>
#include <cstdlib>

#include "internal.cuh"
#include "scan.cuh"

namespace hako
{
namespace
{
// voxelization.hpp:15-48
__device__ __forceinline__ V2 proj2( V3 p, int axis )
{
	if( axis == 0 ) return { p.x, p.y };
	if( axis == 1 ) return { p.y, p.z };
	return { p.z, p.x };
}
__device__ __forceinline__ float projR( V3 p, int axis ) { return axis == 0 ? p.z : ( axis == 1 ? p.x : p.y ); }
__device__ __forceinline__ I2 proj2i( I3 p, int axis )
{
	if( axis == 0 ) return { p.x, p.y };
	if( axis == 1 ) return { p.y, p.z };
	return { p.z, p.x };
}
__device__ __forceinline__ int projRi( I3 p, int axis ) { return axis == 0 ? p.z : ( axis == 1 ? p.x : p.y ); }
// voxelization.hpp:96-122
__device__ __forceinline__ V3 unproj( V2 p, float r, int axis )
{
	if( axis == 0 ) return { p.x, p.y, r };
	if( axis == 1 ) return { r, p.x, p.y };
	return { p.y, r, p.x };
}
__device__ __forceinline__ I3 unproji( int px, int py, int r, int axis )
{
	if( axis == 0 ) return { px, py, r };
	if( axis == 1 ) return { r, px, py };
	return { py, r, px };
}
// voxelization.hpp:84-94
__device__ __forceinline__ int majorAxis( V3 d )
{
	float x = sabs( d.x ), y = sabs( d.y ), z = sabs( d.z );
	if( x < y ) return y < z ? 0 : 2;
	return x < z ? 0 : 1;
}

struct Edges
{
	float nx[3], ny[3], dc[3];
};
// voxelization.hpp:162-199: edge normals and d constants of one 2-D projection
__device__ __forceinline__ Edges edgeSetup( V3 v0, V3 v1, V3 v2, V3 n, int axis, bool sixSep, float dps )
{
	Edges E;
	const V2 vs[3] = { proj2( v0, axis ), proj2( v1, axis ), proj2( v2, axis ) };
	const float sgn = 0.0f < projR( n, axis ) ? 1.0f : -1.0f;
#pragma unroll
	for( int e = 0; e < 3; e++ )
	{
		const V2 a = vs[e];
		const V2 b = vs[( e + 1 ) % 3];
		const V2 d = { b.x - a.x, b.y - a.y };
		const V2 ne = { -d.y * sgn, d.x * sgn };
		E.nx[e] = ne.x;
		E.ny[e] = ne.y;
		float c;
		if( !sixSep )
		{
			c = smax( ne.x * dps, 0.0f ) + smax( ne.y * dps, 0.0f ) - dot( ne, a );
		}
		else
		{
			const V2 h = { dps * 0.5f - a.x, dps * 0.5f - a.y };
			c = dot( ne, h ) + 0.5f * dps * smax( sabs( ne.x ), sabs( ne.y ) );
		}
		E.dc[e] = c;
	}
	return E;
}

struct TriBox
{
	int major;
	I2 lo_xy, hi_xy;
	int lo_z, hi_z;
	V3 n;
};
// voxelization.hpp:146-160: normal, major axis, clamped integer bounding box in projected coordinates
__device__ __forceinline__ TriBox triBox( V3 v0, V3 v1, V3 v2, V3 origin, float dps, int gridRes )
{
	TriBox b;
	const V3 e01 = v1 - v0;
	const V3 e12 = v2 - v1;
	b.n = cross( e01, e12 );
	b.major = majorAxis( b.n );
	const V3 bl = { smin( smin( v0.x, v1.x ), v2.x ), smin( smin( v0.y, v1.y ), v2.y ), smin( smin( v0.z, v1.z ), v2.z ) };
	const V3 bu = { smax( smax( v0.x, v1.x ), v2.x ), smax( smax( v0.y, v1.y ), v2.y ), smax( smax( v0.z, v1.z ), v2.z ) };
	I3 lo = { f2i_x86( floorf( ( bl.x - origin.x ) / dps ) ), f2i_x86( floorf( ( bl.y - origin.y ) / dps ) ), f2i_x86( floorf( ( bl.z - origin.z ) / dps ) ) };
	I3 hi = { f2i_x86( floorf( ( bu.x - origin.x ) / dps ) ), f2i_x86( floorf( ( bu.y - origin.y ) / dps ) ), f2i_x86( floorf( ( bu.z - origin.z ) / dps ) ) };
	lo = { smax( lo.x, 0 ), smax( lo.y, 0 ), smax( lo.z, 0 ) };
	hi = { smin( hi.x, gridRes - 1 ), smin( hi.y, gridRes - 1 ), smin( hi.z, gridRes - 1 ) };
	b.lo_xy = proj2i( lo, b.major );
	b.hi_xy = proj2i( hi, b.major );
	b.lo_z = projRi( lo, b.major );
	b.hi_z = projRi( hi, b.major );
	return b;
}

__device__ __forceinline__ V3 loadV3( const float* p, uint64_t i ) { return { p[i * 3], p[i * 3 + 1], p[i * 3 + 2] }; }

// vectorMath.hpp:249-284
__device__ __forceinline__ V3 closestBary( V3 v0, V3 v1, V3 v2, V3 P )
{
	const V3 d0 = v0 - P, d1 = v1 - P, d2 = v2 - P;
	const V3 e0 = v2 - v0, e1 = v0 - v1, e2 = v1 - v2;
	const V3 Ng = cross( e2, e0 );
	float U = dot( cross( d2, d0 ), Ng );
	float V = dot( cross( d0, d1 ), Ng );
	float W = dot( cross( d1, d2 ), Ng );
	if( U < 0.0f )
	{
		V = dot( -d0, e0 );
		W = dot( d2, e0 );
	}
	else if( V < 0.0f )
	{
		W = dot( -d1, e1 );
		U = dot( d0, e1 );
	}
	else if( W < 0.0f )
	{
		U = dot( -d2, e2 );
		V = dot( d1, e2 );
	}
	const V3 bc = { smax( 0.0f, U ), smax( 0.0f, V ), smax( 0.0f, W ) };
	const float s = bc.x + bc.y + bc.z;
	return { bc.x / s, bc.y / s, bc.z / s };
}

// the same with the per-triangle part ( e0, e1, e2, Ng ) computed once per triangle: identical expressions, so identical bits
struct BaryTri
{
	V3 e0, e1, e2, Ng;
};
__device__ __forceinline__ BaryTri baryTri( V3 v0, V3 v1, V3 v2 )
{
	BaryTri t;
	t.e0 = v2 - v0, t.e1 = v0 - v1, t.e2 = v1 - v2;
	t.Ng = cross( t.e2, t.e0 );
	return t;
}
__device__ __forceinline__ V3 closestBaryPre( V3 v0, V3 v1, V3 v2, const BaryTri& t, V3 P )
{
	const V3 d0 = v0 - P, d1 = v1 - P, d2 = v2 - P;
	float U = dot( cross( d2, d0 ), t.Ng );
	float V = dot( cross( d0, d1 ), t.Ng );
	float W = dot( cross( d1, d2 ), t.Ng );
	if( U < 0.0f )
	{
		V = dot( -d0, t.e0 );
		W = dot( d2, t.e0 );
	}
	else if( V < 0.0f )
	{
		W = dot( -d1, t.e1 );
		U = dot( d0, t.e1 );
	}
	else if( W < 0.0f )
	{
		U = dot( -d2, t.e2 );
		V = dot( d1, t.e2 );
	}
	const V3 bc = { smax( 0.0f, U ), smax( 0.0f, V ), smax( 0.0f, W ) };
	const float s = bc.x + bc.y + bc.z;
	return { bc.x / s, bc.y / s, bc.z / s };
}

// (uint8_t)( v ) with x86 semantics, cheaper than f2u8_x86: cvt.rzi saturates, so its low byte differs from cvttss2si's only for
// v >= 2^31 (0x7FFFFFFF -> 0xFF against the indefinite value 0x80000000 -> 0x00); NaN gives 0 either way, v < -2^31 gives 0x80000000
__device__ __forceinline__ uint32_t f2u8_fast( float v ) { return v < 2147483648.0f ? ( (uint32_t)__float2int_rz( v ) & 0xFFu ) : 0u; }
__device__ __forceinline__ uint32_t packRGB255fast( V3 c )
{
	return f2u8_fast( c.x * 255.0f + 0.5f ) | ( f2u8_fast( c.y * 255.0f + 0.5f ) << 8 ) | ( f2u8_fast( c.z * 255.0f + 0.5f ) << 16 ) | 0xFF000000u;
}
__device__ __forceinline__ uint32_t packRGB255( V3 c )
{
	// (uint8_t)( v * 255.0f + 0.5f ) with x86 cast semantics; alpha = 255 (voxRT.cpp:230-234, voxKernel.cu:151-158)
	return (uint32_t)f2u8_x86( c.x * 255.0f + 0.5f ) | ( (uint32_t)f2u8_x86( c.y * 255.0f + 0.5f ) << 8 ) | ( (uint32_t)f2u8_x86( c.z * 255.0f + 0.5f ) << 16 ) | 0xFF000000u;
}
} // namespace

// A triangle is "small" when the (x, y) footprint of its clamped box holds at most SMALL_AREA cells: the warp-cooperative kernel
// (32 triangles per warp, warpVoxelsKernel below) takes those -- the multi-million-triangle regime of the scanned models -- and the
// per-column kernels take the rest, where one triangle alone is thousands of columns (bunny at 4096^3: ~7 k voxels per triangle).
constexpr int64_t SMALL_AREA = 4096;
__device__ __forceinline__ bool isSmallTri( const TriBox& b )
{
	const int64_t nx = (int64_t)b.hi_xy.x - b.lo_xy.x + 1, ny = (int64_t)b.hi_xy.y - b.lo_xy.y + 1;
	return nx <= 0 || ny <= 0 || nx * ny <= SMALL_AREA;
}

// ---- 1. columns per triangle (largeOnly: small triangles get no column, they belong to warpVoxelsKernel) ----
__global__ void __launch_bounds__( 256 ) triColumnsKernel( const float* __restrict__ pos, uint32_t nTriangles, V3 origin, float dps, int gridRes, uint32_t* __restrict__ nCols, int largeOnly )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if( i >= nTriangles ) return;
	const TriBox b = triBox( loadV3( pos, (uint64_t)i * 3 ), loadV3( pos, (uint64_t)i * 3 + 1 ), loadV3( pos, (uint64_t)i * 3 + 2 ), origin, dps, gridRes );
	uint32_t n = ( b.hi_xy.x >= b.lo_xy.x ) ? (uint32_t)( b.hi_xy.x - b.lo_xy.x + 1 ) : 0u;
	if( largeOnly && isSmallTri( b ) ) n = 0;
	nCols[i] = n;
}

// ---- generic u32 exclusive scan kernel (in place allowed), 256 threads x 8 items ----
#define SCAN_ITEMS 8
__global__ void __launch_bounds__( 256 ) scanU32Kernel( const uint32_t* __restrict__ in, uint32_t* __restrict__ out, uint64_t n, ScanState st, uint64_t* __restrict__ totalOut )
{
	const uint32_t tile = scanTakeTicket( st );
	const uint64_t base = (uint64_t)tile * ( 256 * SCAN_ITEMS ) + (uint64_t)threadIdx.x * SCAN_ITEMS;
	uint32_t v[SCAN_ITEMS];
	uint32_t sum = 0;
#pragma unroll
	for( int k = 0; k < SCAN_ITEMS; k++ )
	{
		v[k] = ( base + k < n ) ? in[base + k] : 0u;
		sum += v[k];
	}
	uint32_t total;
	const uint32_t excl = blockExclusiveScan<256>( sum, &total );
	const uint64_t prefix = scanTilePrefix( st, tile, total );
	uint32_t run = (uint32_t)prefix + excl;
#pragma unroll
	for( int k = 0; k < SCAN_ITEMS; k++ )
	{
		if( base + k < n ) out[base + k] = run;
		run += v[k];
	}
	if( totalOut && tile == gridDim.x - 1 && threadIdx.x == 0 ) *totalOut = prefix + total;
}
int scanU32( hako_ctx* ctx, cudaStream_t s, const uint32_t* in, uint32_t* out, uint64_t n, uint64_t* totalOut_d )
{
	if( n == 0 )
	{
		if( totalOut_d ) HAKO_CUDA( cudaMemsetAsync( totalOut_d, 0, 8, s ) );
		return 0;
	}
	const uint64_t tiles = ( n + 256 * SCAN_ITEMS - 1 ) / ( 256 * SCAN_ITEMS );
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
		HAKO_LAUNCH_R( ctx, r, scanU32Kernel, (uint32_t)tiles, 256, 0, s, in, out, n, makeScanState( mem, tiles ), totalOut_d );
	} while( 0 );
	cudaFreeAsync( mem, s ); // also when the launch failed
	return r;
}

// ---- 3./5. per-column enumeration ----
#ifndef HAKO_EMIT_CTAS
#define HAKO_EMIT_CTAS 6 // B200, bench scene: 4 / 5 / 6 / 8 CTAs per SM = 7.58 / 6.86 / 6.57 / 6.67 ms for the voxelize stage (80 B of spills at 6)
#endif
#ifndef HAKO_COUNT_CTAS
#define HAKO_COUNT_CTAS 12 // 8 -> 12: another 0.15 ms (192 B of spills)
#endif
template <bool EMIT>
__global__ void __launch_bounds__( 128, EMIT ? HAKO_EMIT_CTAS : HAKO_COUNT_CTAS ) columnVoxelsKernel( const float* __restrict__ pos, const float* __restrict__ col, const float* __restrict__ emi,
															uint32_t nTriangles, const uint32_t* __restrict__ colOffset, uint32_t nColumns,
															V3 origin, float dps, int gridRes, int sixSeparating,
															uint32_t* __restrict__ colCount, const uint32_t* __restrict__ voxOffset, uint32_t totalVoxels,
															uint64_t* __restrict__ keys, uint64_t* __restrict__ values, uint32_t blockBase )
{
	// column -> triangle: colOffset is the exclusive scan of columns per triangle; the owner of column c is the last
	// triangle whose offset is <= c.  A 128-column block spans at most 128 owners, so their offsets are staged once
	// in shared memory (zero-column triangles in between are handled by the global fallback search).
	__shared__ uint32_t s_off[129];
	__shared__ uint32_t s_triFirst;
	const uint32_t c0 = ( blockIdx.x + blockBase ) * 128u; // blockBase: the emit pass may be launched in slices (voxelizeDevice)
	if( threadIdx.x == 0 )
	{
		uint32_t lo = 0, hi = nTriangles; // last tri with colOffset[tri] <= c0
		while( hi - lo > 1 )
		{
			const uint32_t mid = ( lo + hi ) >> 1;
			if( colOffset[mid] <= c0 )
				lo = mid;
			else
				hi = mid;
		}
		s_triFirst = lo;
	}
	__syncthreads();
	const uint32_t triFirst = s_triFirst;
	for( uint32_t j = threadIdx.x; j < 129; j += 128 )
	{
		const uint32_t tIdx = triFirst + 1 + j;
		s_off[j] = ( tIdx < nTriangles ) ? colOffset[tIdx] : 0xFFFFFFFFu;
	}
	__syncthreads();
	// EMIT: a warp whose 32 columns hold at most EMIT_QUEUE voxels enumerates them into a shared-memory queue (cheap, divergent:
	// lanes own 0..dozens of voxels) and then shades them with all 32 lanes (closest-point barycentrics, colour mix, Morton
	// key: ~250 instructions per voxel) and stores keys / values as consecutive words.  Inactive lanes of the last block
	// run on the last column with empty loops so that the whole warp reaches the queue drain.
	constexpr uint32_t EMIT_QUEUE = 384;
	__shared__ uint64_t s_qVoxel[EMIT ? 4 : 1][EMIT ? EMIT_QUEUE : 1];
	__shared__ uint8_t s_qOwner[EMIT ? 4 : 1][EMIT ? EMIT_QUEUE : 1];	 // the lane whose triangle the voxel belongs to
	__shared__ float s_triData[EMIT ? 4 : 1][EMIT ? 32 : 1][27];		 // per lane: positions, colours, emissions of its triangle (odd pitch: no bank conflicts)
	const bool active = c0 + threadIdx.x < nColumns;
	if( !EMIT && !active ) return;
	const uint32_t c = active ? c0 + threadIdx.x : nColumns - 1;
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
	uint32_t warpOut0 = 0, warpTotal = 0;
	bool useQueue = false;
	if( EMIT )
	{
		const uint32_t cw0 = c0 + warp * 32u, cw1 = cw0 + 32u;
		warpOut0 = cw0 < nColumns ? voxOffset[cw0] : totalVoxels;
		warpTotal = ( cw1 < nColumns ? voxOffset[cw1] : totalVoxels ) - warpOut0;
		useQueue = warpTotal <= EMIT_QUEUE;
	}

	uint32_t tri;
	if( c < s_off[128] )
	{
		// first j with s_off[j] > c  => owner = triFirst + j
		uint32_t lo = 0, hi = 128;
		while( lo < hi )
		{
			const uint32_t mid = ( lo + hi ) >> 1;
			if( s_off[mid] > c )
				hi = mid;
			else
				lo = mid + 1;
		}
		tri = triFirst + lo;
	}
	else
	{
		uint32_t lo = triFirst, hi = nTriangles;
		while( hi - lo > 1 )
		{
			const uint32_t mid = ( lo + hi ) >> 1;
			if( colOffset[mid] <= c )
				lo = mid;
			else
				hi = mid;
		}
		tri = lo;
	}

	const V3 v0 = loadV3( pos, (uint64_t)tri * 3 ), v1 = loadV3( pos, (uint64_t)tri * 3 + 1 ), v2 = loadV3( pos, (uint64_t)tri * 3 + 2 );
	const bool six = sixSeparating != 0;
	const TriBox b = triBox( v0, v1, v2, origin, dps, gridRes );
	const int major = b.major;
	const int x = b.lo_xy.x + (int)( c - colOffset[tri] );

	// edge functions: the major projection bounds y; the two others are the 2-D overlap tests (voxelization.hpp:311-336)
	const int ax1 = ( major == 0 ) ? 1 : 0;
	const int ax2 = ( major == 2 ) ? 1 : 2;
	const Edges EM = edgeSetup( v0, v1, v2, b.n, major, six, dps );
	const Edges EA = edgeSetup( v0, v1, v2, b.n, ax1, six, dps );
	const Edges EB = edgeSetup( v0, v1, v2, b.n, ax2, six, dps );

	// voxelization.hpp:200-223 plane coefficients
	const V2 org_xy = proj2( origin, major );
	const float org_z = projR( origin, major );
	const V2 v0xy = proj2( v0, major );
	const float v0z = projR( v0, major );
	const V2 nxy = proj2( b.n, major );
	const float nz = projR( b.n, major );
	const float kx = -nxy.x / nz;
	const float ky = -nxy.y / nz;
	const float K = -kx * v0xy.x - ky * v0xy.y + v0z;
	const float cmax = K + dps * ( smax( kx, 0.0f ) + smax( ky, 0.0f ) );
	const float cmin = K + dps * ( smin( kx, 0.0f ) + smin( ky, 0.0f ) );
	const float csix = K + 0.5f * dps * ( kx + ky );

	// yRangeInclusive( x ), voxelization.hpp:230-269
	int ly = 1, uy = -1;
	{
		const float xc = org_xy.x + x * dps;
		float miny = -HAKO_MAX_FLOAT, maxy = HAKO_MAX_FLOAT;
		bool empty = false;
#pragma unroll
		for( int e = 0; e < 3; e++ )
		{
			const float ex = EM.nx[e], ey = EM.ny[e], cc = EM.dc[e];
			if( ey == 0.0f )
			{
				if( -ex * xc <= cc ) continue;
				empty = true;
				break;
			}
			const float k = -( xc * ex + cc ) / ey;
			if( 0.0f < ey )
				miny = smax( miny, k );
			else
				maxy = smin( maxy, k );
		}
		if( !empty )
		{
			const float minIf = smax( ( miny - org_xy.y ) / dps, -2147483648.0f );
			const float maxIf = smin( ( maxy - org_xy.y ) / dps, 2147483520.0f );
			ly = smax( f2i_x86( ceilf( minIf ) ), b.lo_xy.y );
			uy = smin( f2i_x86( floorf( maxIf ) ), b.hi_xy.y );
		}
	}

	V3 c0v = {}, c1v = {}, c2v = {}, e0v = {}, e1v = {}, e2v = {};
	uint32_t out = 0;
	if( EMIT )
	{
		if( !useQueue )
		{
			c0v = loadV3( col, (uint64_t)tri * 3 ), c1v = loadV3( col, (uint64_t)tri * 3 + 1 ), c2v = loadV3( col, (uint64_t)tri * 3 + 2 );
			e0v = loadV3( emi, (uint64_t)tri * 3 ), e1v = loadV3( emi, (uint64_t)tri * 3 + 1 ), e2v = loadV3( emi, (uint64_t)tri * 3 + 2 );
		}
		else
		{
			// the drain reads the triangle through shared memory: 27 words per voxel from L1 would make it LSU-bound
			float* d = s_triData[warp][lane];
			d[0] = v0.x, d[1] = v0.y, d[2] = v0.z, d[3] = v1.x, d[4] = v1.y, d[5] = v1.z, d[6] = v2.x, d[7] = v2.y, d[8] = v2.z;
#pragma unroll
			for( int k = 0; k < 9; k++ )
			{
				d[9 + k] = col[(uint64_t)tri * 9 + k];
				d[18 + k] = emi[(uint64_t)tri * 9 + k];
			}
		}
		out = voxOffset[c];
		if( !active ) uy = ly - 1; // padding lane: nothing to enumerate
	}
	uint32_t count = 0;
	const bool flat = b.lo_z == b.hi_z; // voxelization.hpp:313-316: a one-slab box overlaps by construction
	for( int y = ly; y <= uy; y++ )
	{
		// zRangeInclusive( x, y ), voxelization.hpp:270-299 (no pre-clamp before the casts: x86 semantics)
		const V2 o = { org_xy.x + dps * x, org_xy.y + dps * y };
		const float var = kx * o.x + ky * o.y;
		int zmin, zmax;
		if( six )
		{
			const float idx = ( ( var + csix ) - org_z ) / dps;
			const float zf = floorf( idx );
			const int z = f2i_x86( zf );
			zmin = ( idx == zf ) ? z - 1 : z;
			zmax = z;
		}
		else
		{
			zmin = f2i_x86( floorf( ( ( var + cmin ) - org_z ) / dps ) );
			zmax = f2i_x86( floorf( ( ( var + cmax ) - org_z ) / dps ) );
		}
		zmin = smax( zmin, b.lo_z );
		zmax = smin( zmax, b.hi_z );
		for( int z = zmin; z <= zmax; z++ )
		{
			const float r = org_z + (float)z * dps;
			const V3 p = unproj( o, r, major ); // the voxel's lower corner (voxelization.hpp:300-305)
			if( !flat )
			{
				const V2 pa = proj2( p, ax1 ), pb = proj2( p, ax2 );
				bool ok = true;
#pragma unroll
				for( int e = 0; e < 3; e++ )
				{
					if( EA.nx[e] * pa.x + EA.ny[e] * pa.y + EA.dc[e] < 0.0f ) ok = false;
					if( EB.nx[e] * pb.x + EB.ny[e] * pb.y + EB.dc[e] < 0.0f ) ok = false;
				}
				if( !ok ) continue;
			}
			if( EMIT && useQueue )
			{
				const I3 ci = unproji( x, y, z, major );
				s_qVoxel[warp][out - warpOut0] = (uint64_t)(uint32_t)ci.x | ( (uint64_t)(uint32_t)ci.y << 21 ) | ( (uint64_t)(uint32_t)ci.z << 42 );
				s_qOwner[warp][out - warpOut0] = (uint8_t)lane;
				out++;
			}
			else if( EMIT )
			{
				const I3 ci = unproji( x, y, z, major );
				const V3 bc = closestBary( v0, v1, v2, p );
				// note the rotation: bc.x * c1 + bc.y * c2 + bc.z * c0 (voxKernel.cu:148, voxRT.cpp:229)
				const V3 bcol = bc.x * c1v + bc.y * c2v + bc.z * c0v;
				const V3 bemi = bc.x * e1v + bc.y * e2v + bc.z * e0v;
				keys[out] = mortonEncode( (uint32_t)ci.x, (uint32_t)ci.y, (uint32_t)ci.z );
				values[out] = (uint64_t)packRGB255( bcol ) | ( (uint64_t)packRGB255( bemi ) << 32 );
				out++;
			}
			count++;
		}
	}
	if( !EMIT ) colCount[c] = count;
	if( EMIT && useQueue )
	{
		__syncwarp();
		for( uint32_t i = lane; i < warpTotal; i += 32u )
		{
			const uint64_t q = s_qVoxel[warp][i];
			const float* d = s_triData[warp][s_qOwner[warp][i]];
			const uint32_t cx = (uint32_t)q & 0x1FFFFFu, cy = (uint32_t)( q >> 21 ) & 0x1FFFFFu, cz = (uint32_t)( q >> 42 ) & 0x1FFFFFu;
			// the voxel's lower corner: origin + dps * index per axis, the same products and sums as in the enumeration
			// ( o = org_xy + dps * {x, y}, r = org_z + (float)z * dps, voxelization.hpp:300-305 )
			const V3 p = { origin.x + dps * (int)cx, origin.y + dps * (int)cy, origin.z + dps * (int)cz };
			const V3 bc = closestBary( V3{ d[0], d[1], d[2] }, V3{ d[3], d[4], d[5] }, V3{ d[6], d[7], d[8] }, p );
			const V3 k0 = { d[9], d[10], d[11] }, k1 = { d[12], d[13], d[14] }, k2 = { d[15], d[16], d[17] };
			const V3 m0 = { d[18], d[19], d[20] }, m1 = { d[21], d[22], d[23] }, m2 = { d[24], d[25], d[26] };
			const V3 bcol = bc.x * k1 + bc.y * k2 + bc.z * k0; // voxKernel.cu:148, voxRT.cpp:229
			const V3 bemi = bc.x * m1 + bc.y * m2 + bc.z * m0;
			keys[warpOut0 + i] = mortonEncode( cx, cy, cz );
			values[warpOut0 + i] = (uint64_t)packRGB255( bcol ) | ( (uint64_t)packRGB255( bemi ) << 32 );
		}
	}
}

// ---- warp-cooperative voxelization of small triangles: ONE set-up per triangle, all 32 lanes busy in every phase ----
// A warp owns 32 consecutive triangles.
//   set-up   lane = triangle: box, the 9 edge functions of the three projections, plane coefficients (VTContext, voxelization.hpp:
//            146-224) -- computed ONCE and parked in shared memory (39 words), instead of once per (triangle, x) column and pass;
//   columns  lane = (triangle, x) of the warp's column list (warp scan + 5-step search): yRangeInclusive (:230-269);
//   rows     lane = (triangle, x, y) of the chunk's row list: zRangeInclusive (:270-299) + the two 2-D overlap tests (:311-336);
//   voxels   COUNT: summed.  EMIT: queued in shared memory in (triangle, x, y, z) order -- the CPU loop's order, voxRT.cpp:198-240 --
//            and shaded by all 32 lanes when the queue is half full (closest-point barycentrics, colour mix, Morton key).
// Every value is the same fp32 expression as in columnVoxelsKernel, so the voxel set and the attributes are bit-identical.
constexpr int VW_WARPS = 4;	   // warps per block
constexpr int VW_CTX = 40;	   // words per triangle context (16-byte aligned rows)
constexpr int VW_QUEUE = 256;  // queue entries per warp; a row chunk adds at most 32 x 4 voxels
enum
{
	// 40 words per triangle, grouped so that each phase fetches what it needs with 16-byte shared loads (the lanes of a chunk mostly
	// read the same few triangles, i.e. broadcasts): [0..3] header, [4..7] z box + plane slopes, [8..11] plane constants, [12..38] edges
	VC_FLAGS = 0, // major | flat << 2
	VC_LOX = 1,
	VC_LOY = 2,
	VC_HIY = 3,
	VC_LOZ = 4,
	VC_HIZ = 5,
	VC_KX = 6,
	VC_KY = 7,
	VC_CSIX = 8,
	VC_CMIN = 9,
	VC_CMAX = 10,
	VC_EM = 12, // nx[3], ny[3], dc[3] of the major projection
	VC_EA = 21, // first minor projection
	VC_EB = 30	// second minor projection
};
__device__ __forceinline__ uint32_t warpInclusiveScan( uint32_t v, uint32_t lane )
{
#pragma unroll
	for( int o = 1; o < 32; o <<= 1 )
	{
		const uint32_t n = __shfl_up_sync( 0xFFFFFFFFu, v, o );
		if( lane >= o ) v += n;
	}
	return v;
}
// largest i in [0, 31] with pref[i] <= j (pref non-decreasing, pref[0] <= j)
__device__ __forceinline__ uint32_t lastNotAbove( const uint32_t* pref, uint32_t j )
{
	uint32_t lo = 0;
#pragma unroll
	for( int step = 16; step >= 1; step >>= 1 )
		if( pref[lo + step] <= j ) lo += step;
	return lo;
}

// MODE 0: count.  MODE 1: emit (enumerate again, queue, shade).  MODE 2: count AND record: every voxel found is written as one
// 8-byte record ( owner lane | x | y | z ) into the warp's slab of `slabCap` records, so that the emit pass is pure shading
// (shadeRecordsKernel) instead of a second enumeration; a warp that finds more voxels than its slab holds keeps counting and is
// emitted by MODE 1 (`onlyAbove` = slabCap makes MODE 1 skip the warps whose records fit).
constexpr int VR_COORD_BITS = 19; // record = lane (5 bits) | x, y, z (19 bits each): grids up to 524288^3
template <int MODE>
__global__ void __launch_bounds__( VW_WARPS * 32, MODE == 1 ? 4 : 8 ) warpVoxelsKernel( const float* __restrict__ pos, const float* __restrict__ col, const float* __restrict__ emi,
																	   uint32_t nTriangles, V3 origin, float dps, int gridRes, int sixSeparating,
																	   uint32_t* __restrict__ warpCount, const uint32_t* __restrict__ warpOffset,
																	   uint64_t* __restrict__ keys, uint64_t* __restrict__ values, uint32_t blockBase,
																	   uint64_t* __restrict__ slab, uint32_t slabCap, uint32_t onlyAbove )
{
	constexpr bool EMIT = MODE == 1;
	constexpr bool RECORD = MODE == 2;
	__shared__ __align__( 16 ) float s_ctx[VW_WARPS][32][VW_CTX];
	__shared__ uint32_t s_colPref[VW_WARPS][32];
	__shared__ uint32_t s_rowPref[VW_WARPS][32];
	// positions, colours, emissions of the warp's triangles + Ng of closestBary: 32-word rows read with eight 16-byte loads per
	// voxel (the drain is bound by shared-memory load issue: 27 scalar loads per voxel cost more than the arithmetic they feed)
	__shared__ __align__( 16 ) float s_tri[EMIT ? VW_WARPS : 1][EMIT ? 32 : 1][32];
	__shared__ uint64_t s_q[EMIT ? VW_WARPS : 1][EMIT ? VW_QUEUE : 1];
	__shared__ uint8_t s_qOwner[EMIT ? VW_WARPS : 1][EMIT ? VW_QUEUE : 1];
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
	const uint32_t gw = ( blockIdx.x + blockBase ) * VW_WARPS + warp; // global warp = group of 32 triangles
	const uint32_t tri = gw * 32u + lane;
	if( EMIT && onlyAbove )
	{
		// hybrid: this launch only serves the warps whose voxels did not fit their record slab (warp-uniform test)
		if( gw * 32u >= nTriangles || warpOffset[gw + 1] - warpOffset[gw] <= onlyAbove ) return;
	}
	const bool six = sixSeparating != 0;
	float* ctx = s_ctx[warp][lane];
	uint32_t* ictx = (uint32_t*)ctx;

	// ---- set-up: lane = triangle ----
	uint32_t nCols = 0;
	if( tri < nTriangles )
	{
		const V3 v0 = loadV3( pos, (uint64_t)tri * 3 ), v1 = loadV3( pos, (uint64_t)tri * 3 + 1 ), v2 = loadV3( pos, (uint64_t)tri * 3 + 2 );
		const TriBox b = triBox( v0, v1, v2, origin, dps, gridRes );
		if( isSmallTri( b ) && b.hi_xy.x >= b.lo_xy.x )
		{
			nCols = (uint32_t)( b.hi_xy.x - b.lo_xy.x + 1 );
			const int major = b.major;
			const int ax1 = ( major == 0 ) ? 1 : 0;
			const int ax2 = ( major == 2 ) ? 1 : 2;
			const Edges EM = edgeSetup( v0, v1, v2, b.n, major, six, dps );
			const Edges EA = edgeSetup( v0, v1, v2, b.n, ax1, six, dps );
			const Edges EB = edgeSetup( v0, v1, v2, b.n, ax2, six, dps );
			// voxelization.hpp:200-223 plane coefficients
			const V2 v0xy = proj2( v0, major );
			const float v0z = projR( v0, major );
			const V2 nxy = proj2( b.n, major );
			const float nz = projR( b.n, major );
			const float kx = -nxy.x / nz;
			const float ky = -nxy.y / nz;
			const float K = -kx * v0xy.x - ky * v0xy.y + v0z;
			ictx[VC_FLAGS] = (uint32_t)major | ( b.lo_z == b.hi_z ? 4u : 0u );
			ictx[VC_LOX] = (uint32_t)b.lo_xy.x;
			ictx[VC_LOY] = (uint32_t)b.lo_xy.y;
			ictx[VC_HIY] = (uint32_t)b.hi_xy.y;
			ictx[VC_LOZ] = (uint32_t)b.lo_z;
			ictx[VC_HIZ] = (uint32_t)b.hi_z;
#pragma unroll
			for( int e = 0; e < 3; e++ )
			{
				ctx[VC_EM + e] = EM.nx[e], ctx[VC_EM + 3 + e] = EM.ny[e], ctx[VC_EM + 6 + e] = EM.dc[e];
				ctx[VC_EA + e] = EA.nx[e], ctx[VC_EA + 3 + e] = EA.ny[e], ctx[VC_EA + 6 + e] = EA.dc[e];
				ctx[VC_EB + e] = EB.nx[e], ctx[VC_EB + 3 + e] = EB.ny[e], ctx[VC_EB + 6 + e] = EB.dc[e];
			}
			ctx[VC_KX] = kx;
			ctx[VC_KY] = ky;
			ctx[VC_CSIX] = K + 0.5f * dps * ( kx + ky );
			ctx[VC_CMIN] = K + dps * ( smin( kx, 0.0f ) + smin( ky, 0.0f ) );
			ctx[VC_CMAX] = K + dps * ( smax( kx, 0.0f ) + smax( ky, 0.0f ) );
			if( EMIT )
			{
				float* d = s_tri[warp][lane];
				d[0] = v0.x, d[1] = v0.y, d[2] = v0.z, d[3] = v1.x, d[4] = v1.y, d[5] = v1.z, d[6] = v2.x, d[7] = v2.y, d[8] = v2.z;
#pragma unroll
				for( int k = 0; k < 9; k++ )
				{
					d[9 + k] = col[(uint64_t)tri * 9 + k];
					d[18 + k] = emi[(uint64_t)tri * 9 + k];
				}
				const BaryTri bt = baryTri( v0, v1, v2 );
				d[27] = bt.Ng.x, d[28] = bt.Ng.y, d[29] = bt.Ng.z; // the edges are only needed when the corner projects outside the triangle
			}
		}
	}
	const uint32_t colIncl = warpInclusiveScan( nCols, lane );
	const uint32_t colTotal = __shfl_sync( 0xFFFFFFFFu, colIncl, 31 );
	s_colPref[warp][lane] = colIncl - nCols;
	__syncwarp();
	uint32_t total = 0;	  // COUNT: voxels of this lane; EMIT: voxels already written by the warp
	uint32_t qCount = 0;  // EMIT: voxels waiting in the queue (warp-uniform)
	const uint32_t outBase = EMIT ? warpOffset[gw] : 0u;
	// shades queue entries [0, qCount) with all lanes and stores them at outBase + total
	auto drain = [&]() {
		__syncwarp();
		for( uint32_t i = lane; i < qCount; i += 32u )
		{
			const uint64_t q = s_q[EMIT ? warp : 0][EMIT ? i : 0];
			const float4* dv = (const float4*)s_tri[EMIT ? warp : 0][EMIT ? s_qOwner[warp][i] : 0];
			float d[32];
#pragma unroll
			for( int k = 0; k < 8; k++ )
			{
				const float4 f = dv[k];
				d[4 * k] = f.x, d[4 * k + 1] = f.y, d[4 * k + 2] = f.z, d[4 * k + 3] = f.w;
			}
			const uint32_t cx = (uint32_t)q & 0x1FFFFFu, cy = (uint32_t)( q >> 21 ) & 0x1FFFFFu, cz = (uint32_t)( q >> 42 ) & 0x1FFFFFu;
			// the voxel's lower corner: origin + dps * index per axis -- the products and sums of the enumeration
			// ( o = org_xy + dps * {x, y}, r = org_z + (float)z * dps, voxelization.hpp:300-305 )
			const V3 p = { origin.x + dps * (int)cx, origin.y + dps * (int)cy, origin.z + dps * (int)cz };
			const V3 q0 = { d[0], d[1], d[2] }, q1 = { d[3], d[4], d[5] }, q2 = { d[6], d[7], d[8] };
			const BaryTri bt = { q2 - q0, q0 - q1, q1 - q2, V3{ d[27], d[28], d[29] } };
			const V3 bc = closestBaryPre( q0, q1, q2, bt, p );
			const V3 k0 = { d[9], d[10], d[11] }, k1 = { d[12], d[13], d[14] }, k2 = { d[15], d[16], d[17] };
			const V3 m0 = { d[18], d[19], d[20] }, m1 = { d[21], d[22], d[23] }, m2 = { d[24], d[25], d[26] };
			const V3 bcol = bc.x * k1 + bc.y * k2 + bc.z * k0; // voxKernel.cu:148, voxRT.cpp:229
			const V3 bemi = bc.x * m1 + bc.y * m2 + bc.z * m0;
			const uint64_t o = (uint64_t)outBase + total + i;
			keys[o] = mortonEncode( cx, cy, cz );
			values[o] = (uint64_t)packRGB255fast( bcol ) | ( (uint64_t)packRGB255fast( bemi ) << 32 );
		}
		total += qCount;
		qCount = 0;
		__syncwarp();
	};

	for( uint32_t cbase = 0; cbase < colTotal; cbase += 32u )
	{
		// ---- columns: lane = (triangle, x) ----
		const uint32_t j = cbase + lane;
		const bool hasCol = j < colTotal;
		const uint32_t owner = hasCol ? lastNotAbove( s_colPref[warp], j ) : 0u;
		const float4* ocv = (const float4*)s_ctx[warp][owner];
		const float4 oh = ocv[0]; // flags, lo_x, lo_y, hi_y
		float em[12];			  // edges of the major projection: words 12..20
#pragma unroll
		for( int k = 0; k < 3; k++ )
		{
			const float4 f = ocv[3 + k];
			em[4 * k] = f.x, em[4 * k + 1] = f.y, em[4 * k + 2] = f.z, em[4 * k + 3] = f.w;
		}
		const int major = (int)( __float_as_uint( oh.x ) & 3u );
		const V2 org_xy = proj2( origin, major );
		const int x = (int)__float_as_uint( oh.y ) + (int)( j - s_colPref[warp][owner] );
		int ly = 1, uy = -1;
		if( hasCol )
		{
			// yRangeInclusive( x ), voxelization.hpp:230-269
			const float xc = org_xy.x + x * dps;
			float miny = -HAKO_MAX_FLOAT, maxy = HAKO_MAX_FLOAT;
			bool empty = false;
#pragma unroll
			for( int e = 0; e < 3; e++ )
			{
				const float ex = em[e], ey = em[3 + e], cc = em[6 + e];
				if( ey == 0.0f )
				{
					if( -ex * xc <= cc ) continue;
					empty = true;
					break;
				}
				const float k = -( xc * ex + cc ) / ey;
				if( 0.0f < ey )
					miny = smax( miny, k );
				else
					maxy = smin( maxy, k );
			}
			if( !empty )
			{
				const float minIf = smax( ( miny - org_xy.y ) / dps, -2147483648.0f );
				const float maxIf = smin( ( maxy - org_xy.y ) / dps, 2147483520.0f );
				ly = smax( f2i_x86( ceilf( minIf ) ), (int)__float_as_uint( oh.z ) );
				uy = smin( f2i_x86( floorf( maxIf ) ), (int)__float_as_uint( oh.w ) );
			}
		}
		const uint32_t rows = ( hasCol && uy >= ly ) ? (uint32_t)( uy - ly + 1 ) : 0u;
		const uint32_t rowIncl = warpInclusiveScan( rows, lane );
		const uint32_t rowTotal = __shfl_sync( 0xFFFFFFFFu, rowIncl, 31 );
		const uint32_t rowExcl = rowIncl - rows;
		__syncwarp();
		s_rowPref[warp][lane] = rowExcl;
		__syncwarp();

		for( uint32_t rbase = 0; rbase < rowTotal; rbase += 32u )
		{
			// ---- rows: lane = (triangle, x, y) ----
			const uint32_t r = rbase + lane;
			const bool hasRow = r < rowTotal;
			const uint32_t cl = hasRow ? lastNotAbove( s_rowPref[warp], r ) : 0u;
			const uint32_t rOwner = __shfl_sync( 0xFFFFFFFFu, owner, cl );
			const int rx = __shfl_sync( 0xFFFFFFFFu, x, cl );
			const int rly = __shfl_sync( 0xFFFFFFFFu, ly, cl );
			const uint32_t rEx = __shfl_sync( 0xFFFFFFFFu, rowExcl, cl );
			const int y = rly + (int)( r - rEx );
			const float4* rcv = (const float4*)s_ctx[warp][rOwner];
			const float4 rh = rcv[0], rz = rcv[1], rk = rcv[2]; // header | lo_z, hi_z, kx, ky | csix, cmin, cmax
			float ed[20];										 // words 20..39: EA at ed[1..9], EB at ed[10..18]
#pragma unroll
			for( int k = 0; k < 5; k++ )
			{
				const float4 f = rcv[5 + k];
				ed[4 * k] = f.x, ed[4 * k + 1] = f.y, ed[4 * k + 2] = f.z, ed[4 * k + 3] = f.w;
			}
			const uint32_t flags = __float_as_uint( rh.x );
			const int rMajor = (int)( flags & 3u );
			const bool flat = ( flags & 4u ) != 0; // voxelization.hpp:313-316: a one-slab box overlaps by construction
			const V2 rorg = proj2( origin, rMajor );
			const float org_z = projR( origin, rMajor );
			const int ax1 = ( rMajor == 0 ) ? 1 : 0;
			const int ax2 = ( rMajor == 2 ) ? 1 : 2;
			int zmin = 0, zmax = -1;
			uint32_t okMask = 0; // bit k: voxel z = zmin + k passes the overlap tests
			const V2 o = { rorg.x + dps * rx, rorg.y + dps * y };
			if( hasRow )
			{
				// zRangeInclusive( x, y ), voxelization.hpp:270-299 (no pre-clamp before the casts: x86 semantics)
				const float var = rz.z * o.x + rz.w * o.y;
				if( six )
				{
					const float idx = ( ( var + rk.x ) - org_z ) / dps;
					const float zf = floorf( idx );
					const int z = f2i_x86( zf );
					zmin = ( idx == zf ) ? z - 1 : z;
					zmax = z;
				}
				else
				{
					zmin = f2i_x86( floorf( ( ( var + rk.y ) - org_z ) / dps ) );
					zmax = f2i_x86( floorf( ( ( var + rk.z ) - org_z ) / dps ) );
				}
				zmin = smax( zmin, (int)__float_as_uint( rz.x ) );
				zmax = smin( zmax, (int)__float_as_uint( rz.y ) );
				// |kx|, |ky| <= 1 on the major axis, so a row holds at most 4 voxels; 8 keeps a chunk within the queue whatever happens
				for( int z = zmin, k = 0; z <= zmax && k < 8; z++, k++ )
				{
					bool ok = true;
					if( !flat )
					{
						const float rr = org_z + (float)z * dps;
						const V3 p = unproj( o, rr, rMajor ); // the voxel's lower corner (voxelization.hpp:300-305)
						const V2 pa = proj2( p, ax1 ), pb = proj2( p, ax2 );
#pragma unroll
						for( int e = 0; e < 3; e++ )
						{
							if( ed[1 + e] * pa.x + ed[4 + e] * pa.y + ed[7 + e] < 0.0f ) ok = false;
							if( ed[10 + e] * pb.x + ed[13 + e] * pb.y + ed[16 + e] < 0.0f ) ok = false;
						}
					}
					if( ok ) okMask |= 1u << k;
				}
			}
			const uint32_t nv = __popc( okMask );
			if( RECORD )
			{
				// `total` is warp-uniform here: the voxels found so far; records beyond the slab are dropped (the count goes on)
				const uint32_t vIncl = warpInclusiveScan( nv, lane );
				const uint32_t vTotal = __shfl_sync( 0xFFFFFFFFu, vIncl, 31 );
				uint32_t at = total + vIncl - nv;
				for( uint32_t m = okMask; m; m &= m - 1u )
				{
					const int z = zmin + ( __ffs( m ) - 1 );
					const I3 ci = unproji( rx, y, z, rMajor );
					if( at < slabCap )
						slab[(uint64_t)gw * slabCap + at] = (uint64_t)rOwner | ( (uint64_t)(uint32_t)ci.x << 5 ) | ( (uint64_t)(uint32_t)ci.y << ( 5 + VR_COORD_BITS ) ) |
															( (uint64_t)(uint32_t)ci.z << ( 5 + 2 * VR_COORD_BITS ) );
					at++;
				}
				total += vTotal;
			}
			else if( !EMIT )
				total += nv;
			else
			{
				const uint32_t vIncl = warpInclusiveScan( nv, lane );
				const uint32_t vTotal = __shfl_sync( 0xFFFFFFFFu, vIncl, 31 );
				if( qCount + vTotal > VW_QUEUE ) drain();
				uint32_t at = qCount + vIncl - nv;
				for( uint32_t m = okMask; m; m &= m - 1u )
				{
					const int z = zmin + ( __ffs( m ) - 1 );
					const I3 ci = unproji( rx, y, z, rMajor );
					s_q[EMIT ? warp : 0][EMIT ? at : 0] = (uint64_t)(uint32_t)ci.x | ( (uint64_t)(uint32_t)ci.y << 21 ) | ( (uint64_t)(uint32_t)ci.z << 42 );
					s_qOwner[EMIT ? warp : 0][EMIT ? at : 0] = (uint8_t)rOwner;
					at++;
				}
				qCount += vTotal;
				if( qCount >= VW_QUEUE / 2 ) drain();
			}
		}
	}
	if( EMIT )
	{
		if( qCount ) drain();
	}
	else
	{
		if( !RECORD )
		{
#pragma unroll
			for( int o = 16; o > 0; o >>= 1 ) total += __shfl_xor_sync( 0xFFFFFFFFu, total, o );
		}
		if( lane == 0 ) warpCount[gw] = total;
	}
}

// ---- the emit pass of the record path: pure shading.  A warp loads its 32 triangles' positions, colours, emissions (+ Ng) into
// shared memory, then walks its slab: one record = one voxel -> closest-point barycentrics, colour mix, Morton key, two stores. ----
__global__ void __launch_bounds__( VW_WARPS * 32, 8 ) shadeRecordsKernel( const float* __restrict__ pos, const float* __restrict__ col, const float* __restrict__ emi, uint32_t nTriangles,
														 V3 origin, float dps, const uint32_t* __restrict__ warpOffset, const uint64_t* __restrict__ slab, uint32_t slabCap,
														 uint64_t* __restrict__ keys, uint64_t* __restrict__ values, uint32_t blockBase )
{
	__shared__ __align__( 16 ) float s_tri[VW_WARPS][32][32];
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
	const uint32_t gw = ( blockIdx.x + blockBase ) * VW_WARPS + warp;
	if( gw * 32u >= nTriangles ) return;
	const uint32_t outBase = warpOffset[gw];
	const uint32_t count = warpOffset[gw + 1] - outBase;
	if( count == 0 || count > slabCap ) return; // nothing to do, or the warp overflowed its slab and is emitted by warpVoxelsKernel<1>
	const uint32_t tri = gw * 32u + lane;
	if( tri < nTriangles )
	{
		const V3 v0 = loadV3( pos, (uint64_t)tri * 3 ), v1 = loadV3( pos, (uint64_t)tri * 3 + 1 ), v2 = loadV3( pos, (uint64_t)tri * 3 + 2 );
		float* d = s_tri[warp][lane];
		d[0] = v0.x, d[1] = v0.y, d[2] = v0.z, d[3] = v1.x, d[4] = v1.y, d[5] = v1.z, d[6] = v2.x, d[7] = v2.y, d[8] = v2.z;
#pragma unroll
		for( int k = 0; k < 9; k++ )
		{
			d[9 + k] = col[(uint64_t)tri * 9 + k];
			d[18 + k] = emi[(uint64_t)tri * 9 + k];
		}
		const BaryTri bt = baryTri( v0, v1, v2 );
		d[27] = bt.Ng.x, d[28] = bt.Ng.y, d[29] = bt.Ng.z;
	}
	__syncwarp();
	const uint64_t* rec = slab + (uint64_t)gw * slabCap;
	constexpr uint32_t CM = ( 1u << VR_COORD_BITS ) - 1u;
	for( uint32_t i = lane; i < count; i += 32u )
	{
		const uint64_t q = __ldcs( rec + i );
		const float4* dv = (const float4*)s_tri[warp][(uint32_t)q & 31u];
		float d[32];
#pragma unroll
		for( int k = 0; k < 8; k++ )
		{
			const float4 f = dv[k];
			d[4 * k] = f.x, d[4 * k + 1] = f.y, d[4 * k + 2] = f.z, d[4 * k + 3] = f.w;
		}
		const uint32_t cx = (uint32_t)( q >> 5 ) & CM, cy = (uint32_t)( q >> ( 5 + VR_COORD_BITS ) ) & CM, cz = (uint32_t)( q >> ( 5 + 2 * VR_COORD_BITS ) ) & CM;
		// the voxel's lower corner: origin + dps * index per axis -- the products and sums of the enumeration (voxelization.hpp:300-305)
		const V3 p = { origin.x + dps * (int)cx, origin.y + dps * (int)cy, origin.z + dps * (int)cz };
		const V3 q0 = { d[0], d[1], d[2] }, q1 = { d[3], d[4], d[5] }, q2 = { d[6], d[7], d[8] };
		const BaryTri bt = { q2 - q0, q0 - q1, q1 - q2, V3{ d[27], d[28], d[29] } };
		const V3 bc = closestBaryPre( q0, q1, q2, bt, p );
		const V3 k0 = { d[9], d[10], d[11] }, k1 = { d[12], d[13], d[14] }, k2 = { d[15], d[16], d[17] };
		const V3 m0 = { d[18], d[19], d[20] }, m1 = { d[21], d[22], d[23] }, m2 = { d[24], d[25], d[26] };
		const V3 bcol = bc.x * k1 + bc.y * k2 + bc.z * k0; // voxKernel.cu:148, voxRT.cpp:229
		const V3 bemi = bc.x * m1 + bc.y * m2 + bc.z * m0;
		const uint64_t o = (uint64_t)outBase + i;
		keys[o] = mortonEncode( cx, cy, cz );
		values[o] = (uint64_t)packRGB255fast( bcol ) | ( (uint64_t)packRGB255fast( bemi ) << 32 );
	}
}

// warps whose voxel count (difference of neighbouring offsets) exceeds the slab capacity -> out[0]; the largest count -> out[1]
__global__ void __launch_bounds__( 256 ) countAboveKernel( const uint32_t* __restrict__ offsets, uint32_t n, uint32_t cap, unsigned long long* __restrict__ out )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t cnt = i < n ? offsets[i + 1] - offsets[i] : 0u;
	const uint32_t m = __ballot_sync( 0xFFFFFFFFu, cnt > cap );
	if( ( threadIdx.x & 31u ) == 0 && m ) atomicAdd( out, (unsigned long long)__popc( m ) );
	const uint32_t mx = __reduce_max_sync( 0xFFFFFFFFu, cnt );
	if( ( threadIdx.x & 31u ) == 0 && mx ) atomicMax( out + 1, (unsigned long long)mx ); // the largest warp: sizes the next build's slabs
}

int voxelizeDevice( hako_ctx* ctx, cudaStream_t s, const float* pos, const float* col, const float* emi, uint32_t nTriangles,
					V3 origin, float dps, int gridRes, int sixSeparating, uint64_t** keysOut, uint64_t** valuesOut, uint64_t* D,
					const std::function<int( int )>& beforeEmit, int emitSlices )
{
	// Output layout: the voxels of the small triangles first (warpVoxelsKernel; triangle-major, then x, y, z: the CPU loop's order,
	// voxRT.cpp:198-240), then those of the large triangles (columnVoxelsKernel, same order among themselves).  The order is
	// deterministic but unspecified for callers (the reference's is an atomicAdd race, voxKernel.cu:121); sort + unique do not care.
	bool emitHookRun = false;
	*D = 0;
	if( keysOut ) *keysOut = nullptr;
	if( valuesOut ) *valuesOut = nullptr;
	if( nTriangles == 0 ) return beforeEmit ? beforeEmit( -1 ) : 0;
	static const bool columnsOnly = getenv( "HAKO_VOX_COLUMNS_ONLY" ) && atoi( getenv( "HAKO_VOX_COLUMNS_ONLY" ) ) != 0; // A/B knob: the round-1 path
	uint32_t* colOffset = nullptr;
	uint32_t* colCount = nullptr;
	uint32_t* warpCount = nullptr;
	uint64_t* slab = nullptr;	// voxel records of the small triangles (record path)
	uint64_t* totals = nullptr; // [0] columns of the large triangles, [1] voxels of the small triangles, [2] voxels of the large ones, [3] warps beyond their slab, [4] largest warp
	int r = 0;
	do
	{
		const uint32_t nWarpBlocks = columnsOnly ? 0u : divUp( nTriangles, 32u * VW_WARPS );
		const uint32_t nWarps = nWarpBlocks * VW_WARPS;
		if( ( r = dalloc( ctx, &colOffset, (uint64_t)nTriangles + 1, s ) ) ) break;
		if( ( r = dalloc( ctx, &warpCount, (uint64_t)nWarps + 1, s ) ) ) break;
		if( ( r = dalloc( ctx, &totals, 5, s ) ) ) break;
		// Record path: the counting pass keeps every voxel it finds (8 bytes) in a per-warp slab, and the emit pass only shades.
		// Slab capacity: what a 4 GiB temporary (at most an eighth of the free memory) gives every warp, at most 4 096 records; when
		// this context has built the same mesh before (RTCamp rebuilds it every frame), 1.25 x the largest warp of that build,
		// scaled by the change of resolution.  Scenes too dense for 256 records per warp (32768^3: ~18 k voxels per warp)
		// enumerate twice as before, and so does any single warp that overflows its slab.
		// HAKO_VOX_RECORD=0: always enumerate twice (A/B reference).
		static const bool recordOff = getenv( "HAKO_VOX_RECORD" ) && atoi( getenv( "HAKO_VOX_RECORD" ) ) == 0;
		uint32_t slabCap = 0;
		if( nWarps && keysOut && !recordOff && gridRes <= ( 1 << VR_COORD_BITS ) )
		{
			// cudaMemGetInfo takes 0.1 ms when the device is idle and 1 - 20 ms when frees of the previous build are still in flight
			// (tools/build_device_probe.py, profiles/r4b_vox_timing.log): ask when a mesh is seen for the first time, reuse the answer
			// while the same mesh is rebuilt (a slab that then fails to allocate falls back to two enumerations below)
			const bool sameMesh = ctx->lastVoxTriangles == nTriangles && ctx->lastVoxGrid > 0 && ctx->lastVoxPerWarp > 0.0;
			if( !sameMesh || ctx->voxFreeBytes == 0 )
			{
				size_t freeB = 0, totalB = 0;
				cudaMemGetInfo( &freeB, &totalB );
				ctx->voxFreeBytes = (uint64_t)freeB;
			}
			const uint64_t hardBudget = std::min<uint64_t>( 4ull << 30, ctx->voxFreeBytes / 8 );
			uint64_t cap = hardBudget / ( 8ull * nWarps );
			if( sameMesh )
			{
				const double scale = (double)gridRes / (double)ctx->lastVoxGrid;
				cap = std::min<uint64_t>( (uint64_t)( 1.25 * ctx->lastVoxPerWarp * scale * scale ) + 64, cap );
			}
			cap = std::min<uint64_t>( 4096, cap ) & ~31ull;
			if( cap >= 256 ) slabCap = (uint32_t)cap;
			static const int capEnv = getenv( "HAKO_VOX_SLAB_CAP" ) ? atoi( getenv( "HAKO_VOX_SLAB_CAP" ) ) : 0; // tests: force overflowing warps
			if( capEnv > 0 ) slabCap = (uint32_t)capEnv;
		}
		if( slabCap && dalloc( ctx, &slab, (uint64_t)nWarps * slabCap, s ) )
		{
			// no room for the records (the free-memory figure may be stale): enumerate twice instead, and ask again next time
			cudaGetLastError();
			slab = nullptr;
			slabCap = 0;
			ctx->voxFreeBytes = 0;
		}
		cudaError_t e = cudaMemsetAsync( totals, 0, 40, s );
		if( e != cudaSuccess )
		{
			r = cudaFail( e, "memset totals", __FILE__, __LINE__ );
			break;
		}
		auto wCount = warpVoxelsKernel<0>;
		auto wEmit = warpVoxelsKernel<1>;
		auto wRecord = warpVoxelsKernel<2>;
		if( nWarps )
		{
			if( slabCap )
				HAKO_LAUNCH_R( ctx, r, wRecord, nWarpBlocks, VW_WARPS * 32, 0, s, pos, col, emi, nTriangles, origin, dps, gridRes, sixSeparating, warpCount,
							   (const uint32_t*)nullptr, (uint64_t*)nullptr, (uint64_t*)nullptr, 0u, slab, slabCap, 0u );
			else
				HAKO_LAUNCH_R( ctx, r, wCount, nWarpBlocks, VW_WARPS * 32, 0, s, pos, col, emi, nTriangles, origin, dps, gridRes, sixSeparating, warpCount,
							   (const uint32_t*)nullptr, (uint64_t*)nullptr, (uint64_t*)nullptr, 0u, (uint64_t*)nullptr, 0u, 0u );
			if( ( r = scanU32( ctx, s, warpCount, warpCount, nWarps, totals + 1 ) ) ) break;
			// warpCount[nWarps] = total, so that a warp's count is warpCount[w + 1] - warpCount[w]; then the warps beyond their slab
			e = cudaMemcpyAsync( warpCount + nWarps, totals + 1, 4, cudaMemcpyDeviceToDevice, s );
			if( e != cudaSuccess )
			{
				r = cudaFail( e, "copy total", __FILE__, __LINE__ );
				break;
			}
			if( slabCap ) HAKO_LAUNCH_R( ctx, r, countAboveKernel, divUp( nWarps, 256 ), 256, 0, s, warpCount, nWarps, slabCap, (unsigned long long*)( totals + 3 ) );
		}
		HAKO_LAUNCH_R( ctx, r, triColumnsKernel, divUp( nTriangles, 256 ), 256, 0, s, pos, nTriangles, origin, dps, gridRes, colOffset, columnsOnly ? 0 : 1 );
		if( ( r = scanU32( ctx, s, colOffset, colOffset, nTriangles, totals ) ) ) break;
		e = cudaMemcpyAsync( ctx->pinned, totals, 40, cudaMemcpyDeviceToHost, s );
		if( e == cudaSuccess ) e = cudaStreamSynchronize( s );
		ctx->stats.hostSyncs++;
		if( e != cudaSuccess )
		{
			r = cudaFail( e, "read column / voxel totals", __FILE__, __LINE__ );
			break;
		}
		const uint64_t nColumns = ctx->pinned[0];
		const uint64_t smallVoxels = ctx->pinned[1];
		const uint64_t overflowWarps = ctx->pinned[3];
		if( nWarps )
		{
			ctx->lastVoxTriangles = nTriangles;
			ctx->lastVoxGrid = gridRes;
			ctx->lastVoxPerWarp = slabCap ? (double)ctx->pinned[4] : 0.0; // the LARGEST warp of this build (0: not measured)
		}
		if( nColumns >= 0xFFFFFFFFull )
		{
			setError( "voxelize: %llu (triangle, x) columns exceed the 32-bit work index", (unsigned long long)nColumns );
			r = HAKO_E_LIMIT;
			break;
		}
		auto kCount = columnVoxelsKernel<false>;
		auto kEmit = columnVoxelsKernel<true>;
		uint64_t largeVoxels = 0;
		if( nColumns )
		{
			// the large triangles (none in the scanned-model regime): one more pass and one more host read
			if( ( r = dalloc( ctx, &colCount, nColumns, s ) ) ) break;
			HAKO_LAUNCH_R( ctx, r, kCount, divUp( nColumns, 128 ), 128, 0, s, pos, col, emi, nTriangles, colOffset, (uint32_t)nColumns,
						 origin, dps, gridRes, sixSeparating, colCount, (const uint32_t*)nullptr, 0u, (uint64_t*)nullptr, (uint64_t*)nullptr, 0u );
			if( ( r = scanU32( ctx, s, colCount, colCount, nColumns, totals + 2 ) ) ) break;
			e = cudaMemcpyAsync( ctx->pinned, totals + 2, 8, cudaMemcpyDeviceToHost, s );
			if( e == cudaSuccess ) e = cudaStreamSynchronize( s );
			ctx->stats.hostSyncs++;
			if( e != cudaSuccess )
			{
				r = cudaFail( e, "read voxel total", __FILE__, __LINE__ );
				break;
			}
			largeVoxels = ctx->pinned[0];
		}
		const uint64_t total = smallVoxels + largeVoxels;
		*D = total;
		if( total >= 0xFFFFFFFFull )
		{
			// the reference counts dumped voxels in a uint32 (voxKernel.cu:86, IntersectorOctreeGPU.hpp:92)
			setError( "voxelize: %llu dumped voxels exceed the reference's 32-bit voxel counters", (unsigned long long)total );
			r = HAKO_E_LIMIT;
			break;
		}
		if( total == 0 || !keysOut ) break;
		if( ( r = dalloc( ctx, keysOut, total, s ) ) ) break;
		if( ( r = dalloc( ctx, valuesOut, total, s ) ) ) break;
		emitHookRun = true;
		// Emit slices: slice k covers the triangles [ T_k, T_k+1 ), T_k = nTriangles * k / emitSlices -- the ranges in which the
		// caller uploads the attribute arrays; beforeEmit( k ) returns once the attributes of slices 0..k are on their way (ordered
		// into `s`), so the upload overlaps the emission of earlier slices.  A block of 128 triangles belongs to the last slice any
		// of its triangles belongs to.
		emitSlices = std::max( 1, std::min( { emitSlices, 16, (int)std::max<uint32_t>( 1u, nTriangles / 65536u ) } ) );
		uint32_t blockBase = 0;
		for( int k = 0; k < emitSlices && r == 0; k++ )
		{
			const bool last = k + 1 == emitSlices;
			const uint32_t blockEnd = last ? nWarpBlocks : std::min<uint32_t>( nWarpBlocks, (uint32_t)( (uint64_t)nTriangles * ( k + 1 ) / emitSlices / ( 32u * VW_WARPS ) ) );
			if( beforeEmit && ( r = beforeEmit( last ? -1 : k ) ) ) break; // -1: everything
			if( blockEnd > blockBase && smallVoxels )
			{
				if( slabCap )
				{
					HAKO_LAUNCH_R( ctx, r, shadeRecordsKernel, blockEnd - blockBase, VW_WARPS * 32, 0, s, pos, col, emi, nTriangles, origin, dps, warpCount, slab, slabCap,
								   *keysOut, *valuesOut, blockBase );
					if( overflowWarps ) // the few warps whose voxels did not fit their slab: enumerate again (the others return at once)
						HAKO_LAUNCH_R( ctx, r, wEmit, blockEnd - blockBase, VW_WARPS * 32, 0, s, pos, col, emi, nTriangles, origin, dps, gridRes, sixSeparating,
									   (uint32_t*)nullptr, warpCount, *keysOut, *valuesOut, blockBase, (uint64_t*)nullptr, 0u, slabCap );
				}
				else
					HAKO_LAUNCH_R( ctx, r, wEmit, blockEnd - blockBase, VW_WARPS * 32, 0, s, pos, col, emi, nTriangles, origin, dps, gridRes, sixSeparating, (uint32_t*)nullptr,
								   warpCount, *keysOut, *valuesOut, blockBase, (uint64_t*)nullptr, 0u, 0u );
			}
			blockBase = std::max( blockBase, blockEnd );
		}
		if( r ) break;
		if( largeVoxels )
			HAKO_LAUNCH_R( ctx, r, kEmit, divUp( nColumns, 128 ), 128, 0, s, pos, col, emi, nTriangles, colOffset, (uint32_t)nColumns, origin, dps, gridRes, sixSeparating,
						 (uint32_t*)nullptr, colCount, (uint32_t)largeVoxels, *keysOut + smallVoxels, *valuesOut + smallVoxels, 0u );
	} while( 0 );
	if( beforeEmit && !emitHookRun )
	{
		const int rh = beforeEmit( -1 ); // the caller's upload must be joined on every path
		if( r == 0 ) r = rh;
	}
	dfree( colOffset, s ), dfree( colCount, s ), dfree( warpCount, s ), dfree( totals, s ), dfree( slab, s );
	if( r && keysOut )
	{
		dfree( *keysOut, s ), dfree( *valuesOut, s );
		*keysOut = *valuesOut = nullptr;
	}
	return r;
}
} // namespace hako
