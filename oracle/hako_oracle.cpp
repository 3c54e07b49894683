/*
 * hako_oracle.cpp -- CPU ORACLE ("port").  TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A restatement of the reference's CPU path for the massive-voxel hot path, written so that every
 * fp32 operation happens in the reference's order (compile with -ffp-contract=off, no -mfma):
 *
 *   voxelization.hpp:84-94,146-336  triangle -> voxel 6-separating / conservative overlap
 *   morton.hpp:5-16,70-77,101-115   Morton encode (naive / PDEP / magic bits)
 *   voxRT.cpp:14-51                 mergeVoxels   (sort + unique + integer mean)
 *   voxRT.cpp:198-240               voxel loop + barycentric colour
 *   vectorMath.hpp:100-120,249-284  ss_min/ss_max/ss_abs, closestBarycentricCoordinateOnTriangle
 *   IntersectorOctree.hpp:11-123    buildOctreeDAGReference
 *   IntersectorOctree.hpp:125-212   buildOctreeNaive
 *   IntersectorOctree.hpp:237-257   embedMasks, intersect
 *   voxCommon.hpp:231-423           octreeTraverse_EfficientParametric (SMALL_STACK branch)
 *   renderCommon.hpp:21-75          CameraPinhole
 *
 * Parity status: PINNED -- checked bit-for-bit against the reference's own headers compiled in the
 * build container (oracle/ref_driver.cpp -> oracle/_ref/libhako_ref.so, tests/test_oracle_vs_ref.py)
 * and against the golden values of SURVEY.md section 8(c) (tests/golden/).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.
 */
#include "hako_oracle.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <map>
#include <thread>
#include <vector>

#include <immintrin.h>

namespace
{
struct V2
{
	float x, y;
};
struct V3
{
	float x, y, z;
};
struct I2
{
	int x, y;
};
struct I3
{
	int x, y, z;
};

// vectorMath.hpp:100-114 -- compare-select, NOT fmaxf/fminf (NaN behaviour differs)
template <class T>
inline T smax( T x, T y ) { return ( x < y ) ? y : x; }
template <class T>
inline T smin( T x, T y ) { return ( y < x ) ? y : x; }
inline float sabs( float x ) { return x >= 0.0f ? x : -x; }

// vectorMath.hpp:66-77 (_mm_floor_ss / _mm_ceil_ss) == floorf / ceilf
inline float sfloor( float v ) { return floorf( v ); }
inline float sceil( float v ) { return ceilf( v ); }

// (int)float as compiled for x86-64: cvttss2si => INT_MIN for NaN / out of range.
inline int f2i( float v ) { return _mm_cvttss_si32( _mm_set_ss( v ) ); }
// (uint8_t)float as compiled for x86-64: cvttss2si then truncate to 8 bits.
inline uint8_t f2u8( float v ) { return (uint8_t)( (uint32_t)f2i( v ) & 0xFFu ); }

inline V3 operator-( V3 a, V3 b ) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
inline V3 operator+( V3 a, V3 b ) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
inline V3 operator*( V3 a, float b ) { return { a.x * b, a.y * b, a.z * b }; }
inline V3 operator*( float a, V3 b ) { return { a * b.x, a * b.y, a * b.z }; }
inline V3 operator-( V3 a ) { return { -a.x, -a.y, -a.z }; }
inline float dot( V3 a, V3 b ) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float dot( V2 a, V2 b ) { return a.x * b.x + a.y * b.y; }
inline V3 cross( V3 a, V3 b ) { return { a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; }
inline float max3( float a, float b, float c ) { return smax( smax( a, b ), c ); }
inline float min3( float a, float b, float c ) { return smin( smin( a, b ), c ); }

// voxelization.hpp:15-48
inline V2 proj2( V3 p, int axis )
{
	if( axis == 0 ) return { p.x, p.y };
	if( axis == 1 ) return { p.y, p.z };
	return { p.z, p.x };
}
inline float projR( V3 p, int axis ) { return axis == 0 ? p.z : ( axis == 1 ? p.x : p.y ); }
inline I2 proj2i( I3 p, int axis )
{
	if( axis == 0 ) return { p.x, p.y };
	if( axis == 1 ) return { p.y, p.z };
	return { p.z, p.x };
}
inline int projRi( I3 p, int axis ) { return axis == 0 ? p.z : ( axis == 1 ? p.x : p.y ); }
// voxelization.hpp:96-122
inline V3 unproj( V2 p, float r, int axis )
{
	if( axis == 0 ) return { p.x, p.y, r };
	if( axis == 1 ) return { r, p.x, p.y };
	return { p.y, r, p.x };
}
inline I3 unproji( int px, int py, int r, int axis )
{
	if( axis == 0 ) return { px, py, r };
	if( axis == 1 ) return { r, px, py };
	return { py, r, px };
}
// voxelization.hpp:84-94
inline int majorAxis( V3 d )
{
	float x = sabs( d.x ), y = sabs( d.y ), z = sabs( d.z );
	if( x < y ) return y < z ? 0 : 2;
	return x < z ? 0 : 1;
}

// voxelization.hpp:124-337
struct TriSetup
{
	int major;
	I2 lo_xy, hi_xy;
	int lo_z, hi_z;
	float dc[3][3], nx[3][3], ny[3][3];
	V2 org_xy;
	float org_z;
	float kx, ky, cmax, cmin, csix;

	TriSetup( V3 v0, V3 v1, V3 v2, bool sixSep, V3 origin, float dps, int gridRes )
	{
		V3 e01 = v1 - v0;
		V3 e12 = v2 - v1;
		V3 n = cross( e01, e12 );
		major = majorAxis( n );

		V3 bl = { smin( smin( v0.x, v1.x ), v2.x ), smin( smin( v0.y, v1.y ), v2.y ), smin( smin( v0.z, v1.z ), v2.z ) };
		V3 bu = { smax( smax( v0.x, v1.x ), v2.x ), smax( smax( v0.y, v1.y ), v2.y ), smax( smax( v0.z, v1.z ), v2.z ) };
		V3 lf = { sfloor( ( bl.x - origin.x ) / dps ), sfloor( ( bl.y - origin.y ) / dps ), sfloor( ( bl.z - origin.z ) / dps ) };
		V3 uf = { sfloor( ( bu.x - origin.x ) / dps ), sfloor( ( bu.y - origin.y ) / dps ), sfloor( ( bu.z - origin.z ) / dps ) };
		I3 lo = { f2i( lf.x ), f2i( lf.y ), f2i( lf.z ) };
		I3 hi = { f2i( uf.x ), f2i( uf.y ), f2i( uf.z ) };
		lo = { smax( lo.x, 0 ), smax( lo.y, 0 ), smax( lo.z, 0 ) };
		hi = { smin( hi.x, gridRes - 1 ), smin( hi.y, gridRes - 1 ), smin( hi.z, gridRes - 1 ) };

		lo_xy = proj2i( lo, major );
		hi_xy = proj2i( hi, major );
		lo_z = projRi( lo, major );
		hi_z = projRi( hi, major );

		for( int axis = 0; axis < 3; axis++ )
		{
			V2 vs[3] = { proj2( v0, axis ), proj2( v1, axis ), proj2( v2, axis ) };
			float rem = projR( n, axis );
			float sgn = 0.0f < rem ? 1.0f : -1.0f;
			for( int e = 0; e < 3; e++ )
			{
				V2 a = vs[e];
				V2 b = vs[( e + 1 ) % 3];
				V2 d = { b.x - a.x, b.y - a.y };
				V2 ne = { -d.y * sgn, d.x * sgn };
				nx[axis][e] = ne.x;
				ny[axis][e] = ne.y;
				float c;
				if( !sixSep )
				{
					c = smax( ne.x * dps, 0.0f ) + smax( ne.y * dps, 0.0f ) - dot( ne, a );
				}
				else
				{
					V2 h = { dps * 0.5f - a.x, dps * 0.5f - a.y };
					c = dot( ne, h ) + 0.5f * dps * smax( sabs( ne.x ), sabs( ne.y ) );
				}
				dc[axis][e] = c;
			}
		}
		org_xy = proj2( origin, major );
		org_z = projR( origin, major );
		V2 v0xy = proj2( v0, major );
		float v0z = projR( v0, major );
		V2 nxy = proj2( n, major );
		float nz = projR( n, major );
		kx = -nxy.x / nz;
		ky = -nxy.y / nz;
		float K = -kx * v0xy.x - ky * v0xy.y + v0z;
		cmax = K + dps * ( smax( kx, 0.0f ) + smax( ky, 0.0f ) );
		cmin = K + dps * ( smin( kx, 0.0f ) + smin( ky, 0.0f ) );
		csix = K + 0.5f * dps * ( kx + ky );
	}
	I2 yRange( int x, float dps ) const
	{
		float xc = org_xy.x + x * dps;
		float miny = -3.402823466e+38F, maxy = 3.402823466e+38F;
		for( int e = 0; e < 3; e++ )
		{
			float ex = nx[major][e], ey = ny[major][e], c = dc[major][e];
			if( ey == 0.0f )
			{
				if( -ex * xc <= c ) continue;
				return { 1, -1 };
			}
			float k = -( xc * ex + c ) / ey;
			if( 0.0f < ey )
				miny = smax( miny, k );
			else
				maxy = smin( maxy, k );
		}
		float minIf = smax( ( miny - org_xy.y ) / dps, -2147483648.0f );
		float maxIf = smin( ( maxy - org_xy.y ) / dps, 2147483520.0f );
		int ly = f2i( sceil( minIf ) );
		int uy = f2i( sfloor( maxIf ) );
		ly = smax( ly, lo_xy.y );
		uy = smin( uy, hi_xy.y );
		return { ly, uy };
	}
	I2 zRange( int x, int y, float dps, bool sixSep ) const
	{
		V2 o = { org_xy.x + dps * x, org_xy.y + dps * y };
		float var = kx * o.x + ky * o.y;
		int zmin, zmax;
		if( sixSep )
		{
			float tsix = var + csix;
			float idx = ( tsix - org_z ) / dps;
			float zf = sfloor( idx );
			int z = f2i( zf );
			zmin = idx == zf ? z - 1 : z;
			zmax = z;
		}
		else
		{
			float tmax = var + cmax;
			float tmin = var + cmin;
			zmin = f2i( sfloor( ( tmin - org_z ) / dps ) );
			zmax = f2i( sfloor( ( tmax - org_z ) / dps ) );
		}
		zmin = smax( zmin, lo_z );
		zmax = smin( zmax, hi_z );
		return { zmin, zmax };
	}
	V3 corner( int x, int y, int z, float dps ) const
	{
		V2 pp = { org_xy.x + dps * x, org_xy.y + dps * y };
		float r = org_z + (float)z * dps;
		return unproj( pp, r, major );
	}
	bool overlaps( V3 p ) const
	{
		if( lo_z == hi_z ) return true;
		for( int axis = 0; axis < 3; axis++ )
		{
			if( axis == major ) continue;
			V2 pp = proj2( p, axis );
			for( int e = 0; e < 3; e++ )
			{
				float d = nx[axis][e] * pp.x + ny[axis][e] * pp.y + dc[axis][e];
				if( d < 0.0f ) return false;
			}
		}
		return true;
	}
};

// vectorMath.hpp:249-284
inline V3 closestBary( V3 v0, V3 v1, V3 v2, V3 P )
{
	V3 d0 = v0 - P, d1 = v1 - P, d2 = v2 - P;
	V3 e0 = v2 - v0, e1 = v0 - v1, e2 = v1 - v2;
	V3 Ng = cross( e2, e0 );
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
	V3 bc = { smax( 0.0f, U ), smax( 0.0f, V ), smax( 0.0f, W ) };
	float s = bc.x + bc.y + bc.z;
	return { bc.x / s, bc.y / s, bc.z / s };
}

struct Node
{
	uint8_t mask;
	uint32_t children[8];
	uint32_t psum[8];
};
// voxCommon.hpp:163-180 -- ordering used by std::map<OctreeNode,int>: (mask, children), psum is not part of the key
struct NodeLess
{
	bool operator()( const Node& a, const Node& b ) const
	{
		if( a.mask != b.mask ) return a.mask < b.mask;
		for( int i = 0; i < 8; i++ )
		{
			if( a.children[i] == b.children[i] ) continue;
			return a.children[i] < b.children[i];
		}
		return false;
	}
};
struct Task
{
	uint64_t morton;
	uint32_t child;
	uint32_t nVox;
};

// voxCommon.hpp:202-212
struct StackEl
{
	uint32_t node;
	float tx1, ty1, tz1;
	float scale;
	uint32_t childMask;
	uint32_t skipped;
};

struct TravStats
{
	uint64_t visits = 0, descents = 0, pushes = 0, maxsp = 0;
};

// voxCommon.hpp:231-423 (SMALL_STACK + ENABLE_EMBEDED_MASK)
template <bool STATS>
inline void traverse( const Node* nodes, uint32_t nodeIndex, V3 ro, V3 rd, V3 lower, V3 upper,
					  float* t, int* nMajor, uint32_t* vIndex, bool shadow, TravStats* st )
{
	StackEl stack[32];
	V3 inv = { 1.0f / rd.x, 1.0f / rd.y, 1.0f / rd.z };
	uint32_t vMask = 0;
	if( inv.x < 0.0f )
	{
		vMask |= 1u;
		inv.x = -inv.x;
		ro.x = lower.x + upper.x - ro.x;
	}
	if( inv.y < 0.0f )
	{
		vMask |= 2u;
		inv.y = -inv.y;
		ro.y = lower.y + upper.y - ro.y;
	}
	if( inv.z < 0.0f )
	{
		vMask |= 4u;
		inv.z = -inv.z;
		ro.z = lower.z + upper.z - ro.z;
	}
	V3 lo = lower - ro, up = upper - ro;
	V3 bound = {
		FLT_MAX / smax( smax( sabs( lo.x ), sabs( up.x ) ), 1.0f ),
		FLT_MAX / smax( smax( sabs( lo.y ), sabs( up.y ) ), 1.0f ),
		FLT_MAX / smax( smax( sabs( lo.z ), sabs( up.z ) ), 1.0f ) };
	inv = { smin( inv.x, bound.x ), smin( inv.y, bound.y ), smin( inv.z, bound.z ) };

	V3 t0 = { lo.x * inv.x, lo.y * inv.y, lo.z * inv.z };
	V3 t1 = { up.x * inv.x, up.y * inv.y, up.z * inv.z };
	float lmaxTop = max3( t0.x, t0.y, t0.z );
	if( min3( t1.x, t1.y, t1.z ) < lmaxTop ) return;

	nodeIndex |= (uint32_t)nodes[nodeIndex].mask << 24;

	int sp = 0;
	StackEl cur = { nodeIndex, t1.x, t1.y, t1.z, 1.0f, 0xFFFFFFFFu, 0 };
	V3 dt = { t1.x - t0.x, t1.y - t0.y, t1.z - t0.z };

	for( ;; )
	{
		if( STATS ) st->visits++;
		float tx0 = cur.tx1 - dt.x * cur.scale;
		float ty0 = cur.ty1 - dt.y * cur.scale;
		float tz0 = cur.tz1 - dt.z * cur.scale;
		float lmax = max3( tx0, ty0, tz0 );
		bool popNow = false;

		if( cur.node == 0xFFFFFFFFu )
		{
			if( 0.0f < lmax )
			{
				*t = lmax;
				*nMajor = lmax == tx0 ? 1 : ( lmax == ty0 ? 2 : 0 );
				*vIndex = cur.skipped;
				break;
			}
			popNow = true;
		}
		if( !popNow )
		{
			float txM = 0.5f * ( tx0 + cur.tx1 );
			float tyM = 0.5f * ( ty0 + cur.ty1 );
			float tzM = 0.5f * ( tz0 + cur.tz1 );
			if( cur.childMask == 0xFFFFFFFFu )
			{
				cur.childMask = ( txM < lmax ? 1u : 0u ) | ( tyM < lmax ? 2u : 0u ) | ( tzM < lmax ? 4u : 0u );
			}
			uint32_t mask = cur.node >> 24;
			const Node& node = nodes[cur.node & 0xFFFFFFu];
			float x1 = ( cur.childMask & 1u ) ? cur.tx1 : txM;
			float y1 = ( cur.childMask & 2u ) ? cur.ty1 : tyM;
			float z1 = ( cur.childMask & 4u ) ? cur.tz1 : tzM;
			bool descended = false;
			for( ;; )
			{
				float umin = min3( x1, y1, z1 );
				uint32_t mv = umin == x1 ? 1u : ( umin == y1 ? 2u : 4u );
				bool hasNext = ( cur.childMask & mv ) == 0;
				uint32_t ci = cur.childMask ^ vMask;
				cur.childMask |= mv;
				bool behind = umin < 0.0f;
				if( ( mask & ( 1u << ci ) ) && !behind )
				{
					if( hasNext )
					{
						stack[sp++] = cur;
						if( STATS )
						{
							st->pushes++;
							if( (uint64_t)sp > st->maxsp ) st->maxsp = sp;
						}
					}
					cur.node = node.children[ci];
					cur.tx1 = x1;
					cur.ty1 = y1;
					cur.tz1 = z1;
					cur.scale *= 0.5f;
					cur.childMask = 0xFFFFFFFFu;
					if( !shadow ) cur.skipped += node.psum[ci];
					if( STATS ) st->descents++;
					descended = true;
					break;
				}
				if( !hasNext ) break;
				if( mv == 1u )
					x1 = cur.tx1;
				else if( mv == 2u )
					y1 = cur.ty1;
				else
					z1 = cur.tz1;
			}
			if( descended ) continue;
		}
		if( sp )
			cur = stack[--sp];
		else
			break;
	}
}
} // namespace

struct OrcOctree
{
	std::vector<Node> nodes;
	V3 lower, upper;
};

extern "C" {

const char* orc_impl_name( void ) { return "port"; }

// morton.hpp:70-77
uint64_t orc_morton_encode( uint32_t x, uint32_t y, uint32_t z )
{
	return _pdep_u64( x & 0x1FFFFF, 0x1249249249249249ULL ) |
		   _pdep_u64( y & 0x1FFFFF, 0x1249249249249249ULL << 1 ) |
		   _pdep_u64( z & 0x1FFFFF, 0x1249249249249249ULL << 2 );
}
// morton.hpp:5-16
uint64_t orc_morton_encode_naive( uint32_t x, uint32_t y, uint32_t z )
{
	uint64_t code = 0;
	for( uint64_t i = 0; i < 21; ++i )
	{
		code |= ( (uint64_t)( ( x >> i ) & 1u ) << ( 3 * i ) ) | ( (uint64_t)( ( y >> i ) & 1u ) << ( 3 * i + 1 ) ) |
				( (uint64_t)( ( z >> i ) & 1u ) << ( 3 * i + 2 ) );
	}
	return code;
}
// morton.hpp:93-98
void orc_morton_decode( uint64_t m, uint32_t* x, uint32_t* y, uint32_t* z )
{
	*x = (uint32_t)_pext_u64( m, 0x1249249249249249ULL );
	*y = (uint32_t)_pext_u64( m, 0x1249249249249249ULL << 1 );
	*z = (uint32_t)_pext_u64( m, 0x1249249249249249ULL << 2 );
}
int orc_sort_bits_morton( uint32_t gridRes ) { return __builtin_ctz( gridRes ) * 3; }

uint64_t orc_voxelize( const float* pos, const float* col, const float* emi, uint32_t nTriangles,
					   const float origin[3], float dps, int gridRes, int sixSeparating,
					   uint64_t* keys, uint8_t* colors, uint8_t* emissions, uint64_t capacity )
{
	const V3* P = (const V3*)pos;
	const V3* C = (const V3*)col;
	const V3* E = (const V3*)emi;
	V3 org = { origin[0], origin[1], origin[2] };
	bool six = sixSeparating != 0;
	uint64_t n = 0;
	for( uint32_t i = 0; i < nTriangles; i++ )
	{
		V3 v0 = P[i * 3], v1 = P[i * 3 + 1], v2 = P[i * 3 + 2];
		TriSetup ctx( v0, v1, v2, six, org, dps, gridRes );
		for( int x = ctx.lo_xy.x; x <= ctx.hi_xy.x; x++ )
		{
			I2 yr = ctx.yRange( x, dps );
			for( int y = yr.x; y <= yr.y; y++ )
			{
				I2 zr = ctx.zRange( x, y, dps, six );
				for( int z = zr.x; z <= zr.y; z++ )
				{
					V3 p = ctx.corner( x, y, z, dps );
					if( !ctx.overlaps( p ) ) continue;
					if( n < capacity )
					{
						I3 c = unproji( x, y, z, ctx.major );
						keys[n] = orc_morton_encode( c.x, c.y, c.z );
						if( colors || emissions )
						{
							V3 bc = closestBary( v0, v1, v2, p );
							if( colors )
							{
								V3 c0 = C[i * 3], c1 = C[i * 3 + 1], c2 = C[i * 3 + 2];
								V3 b = bc.x * c1 + bc.y * c2 + bc.z * c0;
								colors[n * 4 + 0] = f2u8( b.x * 255.0f + 0.5f );
								colors[n * 4 + 1] = f2u8( b.y * 255.0f + 0.5f );
								colors[n * 4 + 2] = f2u8( b.z * 255.0f + 0.5f );
								colors[n * 4 + 3] = 255;
							}
							if( emissions )
							{
								V3 e0 = E[i * 3], e1 = E[i * 3 + 1], e2 = E[i * 3 + 2];
								V3 b = bc.x * e1 + bc.y * e2 + bc.z * e0;
								emissions[n * 4 + 0] = f2u8( b.x * 255.0f + 0.5f );
								emissions[n * 4 + 1] = f2u8( b.y * 255.0f + 0.5f );
								emissions[n * 4 + 2] = f2u8( b.z * 255.0f + 0.5f );
								emissions[n * 4 + 3] = 255;
							}
						}
					}
					n++;
				}
			}
		}
	}
	return n;
}

// voxRT.cpp:14-51.  std::map<key, {uvec4 colour sum, uvec4 emission sum}>; w accumulates the count;
// output = sum / count in integer arithmetic (alpha = count / count = 1).  A stable index sort by key
// visits the same groups in the same (ascending key) order the map iteration does.
uint64_t orc_merge_voxels( uint64_t* keys, uint8_t* colors, uint8_t* emissions, uint64_t n )
{
	std::vector<uint32_t> order( n );
	for( uint64_t i = 0; i < n; i++ ) order[i] = (uint32_t)i;
	std::stable_sort( order.begin(), order.end(), [&]( uint32_t a, uint32_t b ) { return keys[a] < keys[b]; } );
	std::vector<uint64_t> ok;
	std::vector<uint8_t> oc, oe;
	uint64_t i = 0;
	while( i < n )
	{
		uint64_t k = keys[order[i]];
		uint32_t cs[4] = { 0, 0, 0, 0 }, es[4] = { 0, 0, 0, 0 };
		uint64_t j = i;
		for( ; j < n && keys[order[j]] == k; j++ )
		{
			const uint8_t* c = colors + (uint64_t)order[j] * 4;
			const uint8_t* e = emissions + (uint64_t)order[j] * 4;
			cs[0] += c[0], cs[1] += c[1], cs[2] += c[2], cs[3] += 1;
			es[0] += e[0], es[1] += e[1], es[2] += e[2], es[3] += 1;
		}
		ok.push_back( k );
		for( int q = 0; q < 4; q++ )
		{
			oc.push_back( (uint8_t)( cs[q] / cs[3] ) );
			oe.push_back( (uint8_t)( es[q] / es[3] ) );
		}
		i = j;
	}
	memcpy( keys, ok.data(), ok.size() * 8 );
	memcpy( colors, oc.data(), oc.size() );
	memcpy( emissions, oe.data(), oe.size() );
	return ok.size();
}

static void embedMasks( std::vector<Node>& nodes )
{
	// IntersectorOctree.hpp:237-246 + voxCommon.hpp:183-195
	for( size_t i = 0; i < nodes.size(); i++ )
	{
		for( int c = 0; c < 8; c++ )
		{
			uint32_t child = nodes[i].children[c];
			if( child == 0xFFFFFFFFu ) continue;
			child |= (uint32_t)nodes[child].mask << 24;
			nodes[i].children[c] = child;
		}
	}
}

static OrcOctree* buildOctree( const uint64_t* mortons, uint64_t n, const float origin[3], float dps, int gridRes, bool dag )
{
	OrcOctree* o = new OrcOctree();
	o->lower = { origin[0], origin[1], origin[2] };
	// IntersectorOctree.hpp:224: origin + vec3(dps) * (float)gridRes
	float ext = dps * (float)gridRes;
	o->upper = { origin[0] + ext, origin[1] + ext, origin[2] + ext };

	std::vector<Task> cur( n ), next;
	for( uint64_t i = 0; i < n; i++ ) cur[i] = { mortons[i], 0xFFFFFFFFu, 1u };
	std::map<Node, int, NodeLess> seen;
	int wide = gridRes;
	while( 1 < wide )
	{
		size_t i = 0;
		while( i < cur.size() )
		{
			uint64_t parent = cur[i].morton >> 3;
			size_t j = i;
			Node node;
			node.mask = 0;
			for( int q = 0; q < 8; q++ )
			{
				node.children[q] = 0xFFFFFFFFu;
				node.psum[q] = 0;
			}
			for( ; j < cur.size() && ( cur[j].morton >> 3 ) == parent; j++ )
			{
				uint32_t s = (uint32_t)( cur[j].morton & 7 );
				node.mask |= (uint8_t)( 1u << s );
				node.children[s] = cur[j].child;
				node.psum[s] = dag ? cur[j].nVox : 0u; // naive build: "not supported yet" (IntersectorOctree.hpp:195)
			}
			uint32_t total = 0;
			if( dag )
			{
				for( int q = 0; q < 8; q++ )
				{
					uint32_t c = node.psum[q];
					node.psum[q] = total;
					total += c;
				}
			}
			uint32_t idx;
			if( dag )
			{
				auto it = seen.find( node );
				if( it == seen.end() )
				{
					idx = (uint32_t)o->nodes.size();
					o->nodes.push_back( node );
					seen[node] = (int)idx;
				}
				else
				{
					idx = (uint32_t)it->second;
				}
			}
			else
			{
				idx = (uint32_t)o->nodes.size();
				o->nodes.push_back( node );
			}
			next.push_back( { parent, idx, total } );
			i = j;
		}
		cur.clear();
		std::swap( cur, next );
		wide /= 2;
	}
	embedMasks( o->nodes );
	return o;
}

OrcOctree* orc_octree_build_dag( const uint64_t* mortons, uint64_t n, const float origin[3], float dps, int gridRes )
{
	return buildOctree( mortons, n, origin, dps, gridRes, true );
}
OrcOctree* orc_octree_build_naive( const uint64_t* mortons, uint64_t n, const float origin[3], float dps, int gridRes )
{
	return buildOctree( mortons, n, origin, dps, gridRes, false );
}
void orc_octree_free( OrcOctree* o ) { delete o; }
uint32_t orc_octree_num_nodes( const OrcOctree* o ) { return (uint32_t)o->nodes.size(); }
void orc_octree_bounds( const OrcOctree* o, float lower[3], float upper[3] )
{
	lower[0] = o->lower.x, lower[1] = o->lower.y, lower[2] = o->lower.z;
	upper[0] = o->upper.x, upper[1] = o->upper.y, upper[2] = o->upper.z;
}
void orc_octree_copy_nodes( const OrcOctree* o, OrcNode68* out )
{
	for( size_t i = 0; i < o->nodes.size(); i++ )
	{
		memset( &out[i], 0, sizeof( OrcNode68 ) );
		out[i].mask = o->nodes[i].mask;
		memcpy( out[i].children, o->nodes[i].children, 32 );
		memcpy( out[i].nVoxelsPSum, o->nodes[i].psum, 32 );
	}
}

void orc_intersect_batch( const OrcOctree* o, const float* ro, const float* rd, uint64_t n, int isShadowRay,
						  float* t, int* nMajor, uint32_t* vIndex, uint64_t* stats )
{
	if( o->nodes.empty() ) return;
	TravStats st;
	uint32_t root = (uint32_t)o->nodes.size() - 1;
	for( uint64_t i = 0; i < n; i++ )
	{
		V3 a = { ro[i * 3], ro[i * 3 + 1], ro[i * 3 + 2] };
		V3 d = { rd[i * 3], rd[i * 3 + 1], rd[i * 3 + 2] };
		if( stats )
			traverse<true>( o->nodes.data(), root, a, d, o->lower, o->upper, &t[i], &nMajor[i], &vIndex[i], isShadowRay != 0, &st );
		else
			traverse<false>( o->nodes.data(), root, a, d, o->lower, o->upper, &t[i], &nMajor[i], &vIndex[i], isShadowRay != 0, nullptr );
	}
	if( stats )
	{
		stats[0] = st.visits, stats[1] = st.descents, stats[2] = st.pushes, stats[3] = st.maxsp;
	}
}

// glm::lookAt (RH) + glm::perspectiveFov (RH, -1..1) + CameraPinhole::initFromPerspective (renderCommon.hpp:21-35)
void orc_camera_lookat( const float eye[3], const float center[3], const float upv[3], float fovy,
						int width, int height, float zNear, float zFar, float focus, float lensR, OrcCamera* out )
{
	(void)zNear;
	(void)zFar;
	V3 e = { eye[0], eye[1], eye[2] }, c = { center[0], center[1], center[2] }, u0 = { upv[0], upv[1], upv[2] };
	auto normalize = []( V3 v ) {
		float inv = 1.0f / sqrtf( v.x * v.x + v.y * v.y + v.z * v.z ); // glm::inversesqrt
		return V3{ v.x * inv, v.y * inv, v.z * inv };
	};
	// glm::cross( x, y ) = ( x.y*y.z - y.y*x.z, x.z*y.x - y.z*x.x, x.x*y.y - y.x*x.y )
	auto gcross = []( V3 x, V3 y ) { return V3{ x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y }; };
	V3 f = normalize( c - e );
	V3 s = normalize( gcross( f, u0 ) );
	V3 u = gcross( s, f );
	// view matrix columns: view[i] = ( s[i], u[i], -f[i] ), view[3] = ( -dot(s,e), -dot(u,e), dot(f,e) )
	V3 tr = { -dot( s, e ), -dot( u, e ), dot( f, e ) };
	// vT = transpose( mat3( view ) ): vT[0] = s, vT[1] = u, vT[2] = -f
	V3 nf = { -f.x, -f.y, -f.z };
	out->front[0] = -nf.x, out->front[1] = -nf.y, out->front[2] = -nf.z;
	out->up[0] = u.x, out->up[1] = u.y, out->up[2] = u.z;
	out->right[0] = s.x, out->right[1] = s.y, out->right[2] = s.z;
	// m = vT * view[3]  (glm mat3 * vec3: m[0][r]*v.x + m[1][r]*v.y + m[2][r]*v.z)
	V3 m = { s.x * tr.x + u.x * tr.y + nf.x * tr.z, s.y * tr.x + u.y * tr.y + nf.y * tr.z, s.z * tr.x + u.z * tr.y + nf.z * tr.z };
	out->o[0] = -m.x, out->o[1] = -m.y, out->o[2] = -m.z;
	// glm::perspectiveFovRH_NO: h = cos(fov/2) / sin(fov/2); proj[1][1] = h
	float h = cosf( 0.5f * fovy ) / sinf( 0.5f * fovy );
	(void)width;
	(void)height;
	out->tanHthetaY = 1.0f / h;
	out->lensR = lensR;
	out->focus = focus;
}

static inline float mixf( float a, float b, float t ) { return a + ( b - a ) * t; }

void orc_camera_shoot( const OrcCamera* cam, int x, int y, float ox, float oy, int w, int h, float ro[3], float rd[3] )
{
	float xf = ( x + ox ) / w;
	float yf = ( y + oy ) / h;
	float a = mixf( -cam->tanHthetaY, cam->tanHthetaY, xf );
	float b = mixf( cam->tanHthetaY, -cam->tanHthetaY, yf );
	for( int k = 0; k < 3; k++ )
	{
		// ((right * mix) * W) / H + up * mix + front   (renderCommon.hpp:42-45)
		rd[k] = cam->right[k] * a * w / h + cam->up[k] * b + cam->front[k];
		ro[k] = cam->o[k];
	}
}
void orc_camera_shoot_thin_lens( const OrcCamera* cam, int x, int y, float ox, float oy, int w, int h, float u0, float u1, float ro[3], float rd[3] )
{
	float xf = ( x + ox ) / w;
	float yf = ( y + oy ) / h;
	float fx = cam->focus * mixf( -cam->tanHthetaY, cam->tanHthetaY, xf ) * w / h;
	float fy = cam->focus * mixf( cam->tanHthetaY, -cam->tanHthetaY, yf );
	float fz = cam->focus;
	float lx = mixf( -cam->lensR, cam->lensR, u0 );
	float ly = mixf( -cam->lensR, cam->lensR, u1 );
	float lz = 0.0f;
	float dx = fx - lx, dy = fy - ly, dz = fz - lz;
	for( int k = 0; k < 3; k++ )
	{
		rd[k] = cam->right[k] * dx + cam->up[k] * dy + cam->front[k] * dz;
		ro[k] = cam->o[k] + cam->right[k] * lx + cam->up[k] * ly + cam->front[k] * lz;
	}
}

void orc_render_primary( const OrcOctree* o, const OrcCamera* cam, int w, int h, int y0, int y1, int nThreads,
						 float* t, int* nMajor, uint32_t* vIndex, uint64_t* stats )
{
	if( nThreads < 1 ) nThreads = 1;
	std::vector<TravStats> sts( nThreads );
	uint32_t root = o->nodes.empty() ? 0 : (uint32_t)o->nodes.size() - 1;
	auto work = [&]( int tid ) {
		for( int j = y0 + tid; j < y1; j += nThreads )
		{
			for( int i = 0; i < w; i++ )
			{
				float ro[3], rd[3];
				orc_camera_shoot( cam, i, j, 0.5f, 0.5f, w, h, ro, rd );
				size_t p = (size_t)j * w + i;
				float tt = FLT_MAX;
				int nm = -1;
				uint32_t vi = 0xFFFFFFFFu;
				if( !o->nodes.empty() )
				{
					if( stats )
						traverse<true>( o->nodes.data(), root, { ro[0], ro[1], ro[2] }, { rd[0], rd[1], rd[2] }, o->lower, o->upper, &tt, &nm, &vi, false, &sts[tid] );
					else
						traverse<false>( o->nodes.data(), root, { ro[0], ro[1], ro[2] }, { rd[0], rd[1], rd[2] }, o->lower, o->upper, &tt, &nm, &vi, false, nullptr );
				}
				t[p] = tt;
				nMajor[p] = nm;
				vIndex[p] = vi;
			}
		}
	};
	if( nThreads == 1 )
	{
		work( 0 );
	}
	else
	{
		std::vector<std::thread> th;
		for( int k = 0; k < nThreads; k++ ) th.emplace_back( work, k );
		for( auto& x : th ) x.join();
	}
	if( stats )
	{
		stats[0] = stats[1] = stats[2] = stats[3] = 0;
		for( auto& s : sts )
		{
			stats[0] += s.visits, stats[1] += s.descents, stats[2] += s.pushes;
			stats[3] = std::max<uint64_t>( stats[3], s.maxsp );
		}
	}
}

// diagnostic: loop visits of every primary ray (how long the longest rays of a frame are; tools/ray_length_probe.py)
void orc_render_primary_visits( const OrcOctree* o, const OrcCamera* cam, int w, int h, int nThreads, uint32_t* visits )
{
	if( nThreads < 1 ) nThreads = 1;
	uint32_t root = o->nodes.empty() ? 0 : (uint32_t)o->nodes.size() - 1;
	auto work = [&]( int tid ) {
		for( int j = tid; j < h; j += nThreads )
			for( int i = 0; i < w; i++ )
			{
				float ro[3], rd[3];
				orc_camera_shoot( cam, i, j, 0.5f, 0.5f, w, h, ro, rd );
				float tt = FLT_MAX;
				int nm = -1;
				uint32_t vi = 0xFFFFFFFFu;
				TravStats st;
				if( !o->nodes.empty() )
					traverse<true>( o->nodes.data(), root, { ro[0], ro[1], ro[2] }, { rd[0], rd[1], rd[2] }, o->lower, o->upper, &tt, &nm, &vi, false, &st );
				visits[(size_t)j * w + i] = (uint32_t)st.visits;
			}
	};
	std::vector<std::thread> th;
	for( int k = 0; k < nThreads; k++ ) th.emplace_back( work, k );
	for( auto& x : th ) x.join();
}

uint64_t orc_fnv1a64( const void* p, uint64_t bytes )
{
	const uint8_t* b = (const uint8_t*)p;
	uint64_t h = 0xcbf29ce484222325ULL;
	for( uint64_t i = 0; i < bytes; i++ )
	{
		h ^= b[i];
		h *= 0x100000001b3ULL;
	}
	return h;
}
}
