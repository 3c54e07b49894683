// common.cuh -- shared device/host helpers for libhako_b200 (sm_100a only).
//
// FP discipline (hard part 2): the whole library is compiled with -fmad=false and without
// fast-math; min/max are compare-selects exactly as the reference's ss_min/ss_max (vectorMath.hpp:100-109),
// never fminf/fmaxf (NaN semantics differ); float->int casts reproduce x86 cvttss2si (INT_MIN on NaN/overflow)
// because the CPU reference relies on it (voxelization.hpp:157-158,283,291-292).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/hako.h"

#define HAKO_MAX_FLOAT 3.402823466e+38F
#define HAKO_PI 3.14159265358979323846f

namespace hako
{
// ---- error plumbing (host) ----
void setError( const char* fmt, ... );
int cudaFail( cudaError_t e, const char* what, const char* file, int line );

#define HAKO_CUDA( call )                                                     \
	do                                                                        \
	{                                                                         \
		cudaError_t e__ = ( call );                                           \
		if( e__ != cudaSuccess ) return hako::cudaFail( e__, #call, __FILE__, __LINE__ ); \
	} while( 0 )
#define HAKO_TRY( call )         \
	do                           \
	{                            \
		int r__ = ( call );      \
		if( r__ != 0 ) return r__; \
	} while( 0 )

inline uint32_t divUp( uint64_t a, uint32_t b ) { return (uint32_t)( ( a + b - 1 ) / b ); }

// ---- device math ----
template <class T>
__host__ __device__ __forceinline__ T smax( T x, T y ) { return ( x < y ) ? y : x; }
template <class T>
__host__ __device__ __forceinline__ T smin( T x, T y ) { return ( y < x ) ? y : x; }
__host__ __device__ __forceinline__ float sabs( float x ) { return x >= 0.0f ? x : -x; } // vectorMath.hpp:110-114 (keeps -0.0f, maps NaN to -NaN)

// (int)float as x86-64 compiles it (cvttss2si): NaN and out-of-range give INT_MIN.  CUDA's cvt.rzi saturates
// and maps NaN to 0, so the indefinite value has to be produced by hand.
__device__ __forceinline__ int f2i_x86( float v )
{
	// -2147483648.0f is exactly representable; anything >= 2^31, < -2^31 or NaN is "indefinite"
	if( !( v >= -2147483648.0f && v < 2147483648.0f ) ) return (int)0x80000000;
	return __float2int_rz( v );
}
// (uint8_t)float on x86-64: cvttss2si to 32 bits, then truncation to the low byte.
__device__ __forceinline__ uint8_t f2u8_x86( float v ) { return (uint8_t)( (uint32_t)f2i_x86( v ) & 0xFFu ); }

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
__host__ __device__ __forceinline__ V3 operator-( V3 a, V3 b ) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
__host__ __device__ __forceinline__ V3 operator+( V3 a, V3 b ) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
__host__ __device__ __forceinline__ V3 operator*( V3 a, V3 b ) { return { a.x * b.x, a.y * b.y, a.z * b.z }; }
__host__ __device__ __forceinline__ V3 operator*( V3 a, float b ) { return { a.x * b, a.y * b, a.z * b }; }
__host__ __device__ __forceinline__ V3 operator*( float a, V3 b ) { return { a * b.x, a * b.y, a * b.z }; }
__host__ __device__ __forceinline__ V3 operator/( V3 a, float b ) { return { a.x / b, a.y / b, a.z / b }; }
__host__ __device__ __forceinline__ V3 operator-( V3 a ) { return { -a.x, -a.y, -a.z }; }
__host__ __device__ __forceinline__ float dot( V3 a, V3 b ) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__host__ __device__ __forceinline__ float dot( V2 a, V2 b ) { return a.x * b.x + a.y * b.y; }
__host__ __device__ __forceinline__ V3 cross( V3 a, V3 b ) { return { a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; }
__host__ __device__ __forceinline__ float max3( float a, float b, float c ) { return smax( smax( a, b ), c ); }
__host__ __device__ __forceinline__ float min3( float a, float b, float c ) { return smin( smin( a, b ), c ); }
__host__ __device__ __forceinline__ float mixf( float a, float b, float t ) { return a + ( b - a ) * t; } // vectorMath.hpp:237-240

// ---- Morton, magic-bits form (morton.hpp:101-115 == voxKernel.cu:11-26; equal to PDEP by unittest.cpp:183-216) ----
__host__ __device__ __forceinline__ uint64_t splitBy3( uint32_t a )
{
	uint64_t x = a & 0x1FFFFF;
	x = ( x | x << 32 ) & 0x1f00000000ffffull;
	x = ( x | x << 16 ) & 0x1f0000ff0000ffull;
	x = ( x | x << 8 ) & 0x100f00f00f00f00full;
	x = ( x | x << 4 ) & 0x10c30c30c30c30c3ull;
	x = ( x | x << 2 ) & 0x1249249249249249ull;
	return x;
}
__host__ __device__ __forceinline__ uint64_t mortonEncode( uint32_t x, uint32_t y, uint32_t z )
{
	return splitBy3( x ) | splitBy3( y ) << 1 | splitBy3( z ) << 2;
}

// ---- MurmurHash32 (voxCommon.hpp:45-109) ----
__host__ __device__ __forceinline__ uint32_t rotl32( uint32_t x, int r ) { return ( x << r ) | ( x >> ( 32 - r ) ); }
__host__ __device__ __forceinline__ uint32_t fmix32( uint32_t h )
{
	h ^= h >> 16;
	h *= 0x85ebca6bu;
	h ^= h >> 13;
	h *= 0xc2b2ae35u;
	h ^= h >> 16;
	return h;
}
struct Murmur32
{
	uint32_t h1, len;
	__host__ __device__ explicit Murmur32( uint32_t seed ) : h1( seed ), len( 0 ) {}
	__host__ __device__ __forceinline__ void combine( uint32_t k1 )
	{
		k1 *= 0xcc9e2d51u;
		k1 = rotl32( k1, 15 );
		k1 *= 0x1b873593u;
		h1 ^= k1;
		h1 = rotl32( h1, 13 );
		h1 = h1 * 5 + 0xe6546b64u;
		len++;
	}
	__host__ __device__ __forceinline__ uint32_t getHash() const { return fmix32( h1 ^ ( len * 4 ) ); }
};
__host__ __device__ __forceinline__ uint32_t hashCombine( uint32_t a, uint32_t b )
{
	Murmur32 h( a );
	h.combine( b );
	return h.getHash();
}

// ---- the resident octree as the kernels see it ----
// Trace layout (derived, 64 B, 64-B aligned): slot[c] = { children[c] (index | mask<<24, or 0xFFFFFFFF), nVoxelsPSum[c] }.
// One descent = one 8-byte load = one 32-B sector; two nodes per 128-B line.  The node's own mask byte is kept in a
// side array (needed by the canonical export and for the root).
struct alignas( 64 ) TraceNode
{
	uint2 slot[8];
};
struct OctreeView
{
	const TraceNode* nodes;
	const uint8_t* masks;
	const hako_voxel_attrib* attrs;
	uint32_t numberOfNodes;
	uint32_t numberOfVoxels;
	V3 lower, upper;
	float emissionScale;
	uint32_t hasEmission;
	uint32_t wide; // child slots hold plain 32-bit indices; masks[child] is the child's mask
};

// read-only 8-byte load through the non-coherent path (the tree is immutable while tracing)
__device__ __forceinline__ uint2 ldg_u2( const uint2* p ) { return __ldg( p ); }
} // namespace hako
