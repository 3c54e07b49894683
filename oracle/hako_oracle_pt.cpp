/*
 * hako_oracle_pt.cpp -- CPU ORACLE ("port"), path-tracing half.  TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Restates, with the host-side math the reference's headers use on the CPU (cosf / sinf / sqrtf / powf,
 * vectorMath.hpp:93-97):
 *   voxCommon.hpp:45-109        MurmurHash32 / hashCombine
 *   renderCommon.hpp:86-117     PCG32, uniformf
 *   renderCommon.hpp:119-151    GetOrthonormalBasis, sampleLambertian
 *   renderCommon.hpp:168-202    luminance, getSpherical, upper_bound_f
 *   renderCommon.hpp:354-470    HDRI::sampleNearest / importanceSample / SAT accessors
 *   pmjSampler.hpp:14-58        GetPMJ02Samples;  :62-102 Owen scrambling;  :114-134 table setup;  :158-173 sample2d
 *   voxKernel.cu:485-608        HDRIstoreImportance / buildSATh / buildSATv / buildSAT2u32 (sequential double sums)
 *   voxKernel.cu:610-777        renderPT (sample mapping, dimension order and estimator)
 *   voxKernel.cu:779-795        renderResolve
 *
 * Parity status: the pieces the reference can run on the host (PMJ table, sample2d, scrambles, Murmur, importanceSample,
 * sampleLambertian, camera) are PINNED against oracle/_ref (tests/test_oracle_pt.py).  HDRI::sampleNearest, the SAT build
 * and renderPT exist only as device code in the reference (renderCommon.hpp:353-366, voxKernel.cu), so their restatement
 * here is "parity unpinned" beyond those pieces; image parity is RMSE-level as BASELINE.json:north_star states.
 */
#include "hako_oracle.h"

#include <cfloat>
#include <cmath>
#include <cstring>
#include <immintrin.h>
#include <thread>
#include <vector>

extern "C" {
void orc_intersect_batch( const OrcOctree* o, const float* ro, const float* rd, uint64_t n, int isShadowRay,
						  float* t, int* nMajor, uint32_t* vIndex, uint64_t* stats );
}

namespace
{
const float PI = 3.14159265358979323846f;

struct F3
{
	float x, y, z;
};
inline F3 operator+( F3 a, F3 b ) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
inline F3 operator*( F3 a, F3 b ) { return { a.x * b.x, a.y * b.y, a.z * b.z }; }
inline F3 operator*( F3 a, float b ) { return { a.x * b, a.y * b, a.z * b }; }
inline F3 operator/( F3 a, float b ) { return { a.x / b, a.y / b, a.z / b }; }
inline float dot( F3 a, F3 b ) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float smaxf( float x, float y ) { return ( x < y ) ? y : x; }
inline float sminf( float x, float y ) { return ( y < x ) ? y : x; }
inline float sclamp( float x, float a, float b ) { return sminf( smaxf( x, a ), b ); } // vectorMath.hpp ss_clamp
inline float mixf( float a, float b, float t ) { return a + ( b - a ) * t; }

inline uint32_t rotl32( uint32_t x, int r ) { return ( x << r ) | ( x >> ( 32 - r ) ); }
inline uint32_t fmix32( uint32_t h )
{
	h ^= h >> 16;
	h *= 0x85ebca6bu;
	h ^= h >> 13;
	h *= 0xc2b2ae35u;
	h ^= h >> 16;
	return h;
}
struct Murmur
{
	uint32_t h1, len;
	explicit Murmur( uint32_t seed ) : h1( seed ), len( 0 ) {}
	void combine( uint32_t k1 )
	{
		k1 *= 0xcc9e2d51u;
		k1 = rotl32( k1, 15 );
		k1 *= 0x1b873593u;
		h1 ^= k1;
		h1 = rotl32( h1, 13 );
		h1 = h1 * 5 + 0xe6546b64u;
		len++;
	}
	uint32_t getHash() const { return fmix32( h1 ^ ( len * 4 ) ); }
};
inline uint32_t hashCombine( uint32_t a, uint32_t b )
{
	Murmur h( a );
	h.combine( b );
	return h.getHash();
}
inline uint32_t hashCombine( uint32_t a, uint32_t b, uint32_t c )
{
	Murmur h( a );
	h.combine( b );
	h.combine( c );
	return h.getHash();
}

struct PCG32
{
	uint64_t state, inc;
	void setup( uint64_t seed, uint64_t stream )
	{
		state = 0;
		inc = stream * 2 + 1;
		nextU32();
		state += seed;
		nextU32();
	}
	uint32_t nextU32()
	{
		uint64_t old = state;
		state = old * 6364136223846793005ULL + inc;
		uint32_t xorshifted = (uint32_t)( ( ( old >> 18u ) ^ old ) >> 27u );
		uint32_t rot = (uint32_t)( old >> 59u );
		return ( xorshifted >> rot ) | ( xorshifted << ( ( -rot ) & 31 ) );
	}
};
inline float uniformf( uint32_t x )
{
	uint32_t bits = ( x >> 9 ) | 0x3f800000u;
	float v;
	memcpy( &v, &bits, 4 );
	return v - 1.0f;
}

// pmjSampler.hpp:14-58
void pmj02( int num_samples, float* samples, PCG32& rng )
{
	static const uint32_t xors[2][32] = {
		{ 0x0, 0x0, 0x2, 0x6, 0x6, 0xe, 0x36, 0x4e, 0x16, 0x2e, 0x276, 0x6ce, 0x716, 0xc2e, 0x3076, 0x40ce, 0x116, 0x22e, 0x20676, 0x60ece, 0x61716, 0xe2c2e, 0x367076, 0x4ec0ce, 0x170116, 0x2c022e, 0x2700676, 0x6c00ece, 0x7001716, 0xc002c2e, 0x30007076, 0x4000c0ce },
		{ 0x0, 0x1, 0x3, 0x3, 0x7, 0x1b, 0x27, 0xb, 0x17, 0x13b, 0x367, 0x38b, 0x617, 0x183b, 0x2067, 0x8b, 0x117, 0x1033b, 0x30767, 0x30b8b, 0x71617, 0x1b383b, 0x276067, 0xb808b, 0x160117, 0x138033b, 0x3600767, 0x3800b8b, 0x6001617, 0x1800383b, 0x20006067, 0x808b } };
	auto u = [&]() { return uniformf( rng.nextU32() ); };
	for( int d = 0; d < 2; d++ ) samples[d] = u();
	for( int log_n = 0; ( 1 << log_n ) < num_samples; log_n++ )
	{
		int prev_len = 1 << log_n;
		int n_strata = prev_len * 2;
		float i_strata = 1.0f / n_strata;
		for( int i = 0; i < prev_len && ( prev_len + i ) < num_samples; i++ )
		{
			const int px = i ^ (int)xors[0][log_n];
			const int pxs = (int)( samples[px * 2] * n_strata );
			const int xs = pxs ^ 1;
			const int py = i ^ (int)xors[1][log_n];
			const int pys = (int)( samples[py * 2 + 1] * n_strata );
			const int ys = pys ^ 1;
			float* s = &samples[( prev_len + i ) * 2];
			// argument evaluation order of GetPMJ02Point( ..., uniformFloat(), uniformFloat(), sample ): g++ evaluates
			// right to left, so xi1 is drawn BEFORE xi0 (pinned against oracle/_ref)
			float xi1 = u();
			float xi0 = u();
			s[0] = ( xi0 + xs ) * i_strata;
			s[1] = ( xi1 + ys ) * i_strata;
		}
	}
}

inline uint32_t laine_karras( uint32_t x, uint32_t seed )
{
	x += seed;
	x ^= x * 0x6c50b47cu;
	x ^= x * 0xb82f1e52u;
	x ^= x * 0xc7afe638u;
	x ^= x * 0x8d22f6e6u;
	return x;
}
inline uint32_t reverseBits( uint32_t v )
{
	v = ( ( v >> 1 ) & 0x55555555u ) | ( ( v & 0x55555555u ) << 1 );
	v = ( ( v >> 2 ) & 0x33333333u ) | ( ( v & 0x33333333u ) << 2 );
	v = ( ( v >> 4 ) & 0x0F0F0F0Fu ) | ( ( v & 0x0F0F0F0Fu ) << 4 );
	v = ( ( v >> 8 ) & 0x00FF00FFu ) | ( ( v & 0x00FF00FFu ) << 8 );
	v = ( v >> 16 ) | ( v << 16 );
	return v;
}
inline uint32_t nus( uint32_t x, uint32_t seed ) { return reverseBits( laine_karras( reverseBits( x ), seed ) ); }
inline float scramble_f32( float x, uint32_t seed )
{
	x += 1.0f;
	uint32_t bits;
	memcpy( &bits, &x, 4 );
	uint32_t s = 0x3f800000u | ( nus( bits & 0x7FFFFFu, seed ) & 0x7FFFFFu );
	float r;
	memcpy( &r, &s, 4 );
	return r - 1.0f;
}
// pmjSampler.hpp:158-173
inline void sample2d( const float* table, uint32_t sampleIdx, uint32_t dimension, uint32_t stream, float* x, float* y )
{
	sampleIdx = nus( sampleIdx, hashCombine( stream, dimension, 31082745u ) ) & 4095u;
	dimension = nus( dimension, hashCombine( stream, 54761983u ) ) & 127u;
	uint32_t head = dimension * 2 * 4096;
	float a = table[head + sampleIdx * 2 + 0];
	float b = table[head + sampleIdx * 2 + 1];
	*x = scramble_f32( a, hashCombine( stream, dimension, 83927105u ) );
	*y = scramble_f32( b, hashCombine( stream, dimension, 12654890u ) );
}

// renderCommon.hpp:119-151
inline void orthoBasis( F3 z, F3* xa, F3* ya )
{
	const float sign = copysignf( 1.0f, z.z );
	const float a = -1.0f / ( sign + z.z );
	const float b = z.x * z.y * a;
	*xa = { 1.0f + sign * z.x * z.x * a, sign * b, -sign * z.x };
	*ya = { b, sign + z.y * z.y * a, -z.y };
}
inline F3 sampleLambertian( float a, float b, F3 Ng )
{
	float r = sqrtf( a );
	float theta = b * PI * 2.0f;
	float x = r * cosf( theta );
	float y = r * sinf( theta );
	float z = sqrtf( smaxf( 1.0f - a, 0.0f ) );
	F3 xa, ya;
	orthoBasis( Ng, &xa, &ya );
	return xa * x + ya * y + Ng * z;
}
inline float luminance( float r, float g, float b ) { return 0.2126f * r + 0.7152f * g + 0.0722f * b; }

struct Hdri
{
	const float* pixels; // RGBA32F
	const float* primary;
	const uint32_t* sats; // [7][w*h]: global, +x, -x, +y, -y, +z, -z
	int w, h, wP, hP;
	float scale;
	bool enabled() const { return 0.0f < scale && pixels; }
	uint32_t psH( const uint32_t* sat, uint32_t x ) const { return x <= 0 ? 0 : sat[(size_t)w * ( h - 1 ) + x - 1]; }
	uint32_t psV( const uint32_t* sat, uint32_t x, uint32_t y ) const
	{
		if( y <= 0 ) return 0;
		uint32_t s0 = x <= 0 ? 0 : sat[(size_t)w * ( y - 1 ) + ( x - 1 )];
		uint32_t s1 = sat[(size_t)w * ( y - 1 ) + x];
		return s1 - s0;
	}
	uint32_t count( const uint32_t* sat, uint32_t x, uint32_t y ) const
	{
		uint32_t a = ( x <= 0 || y <= 0 ) ? 0 : sat[(size_t)w * ( y - 1 ) + ( x - 1 )];
		uint32_t b = ( y <= 0 ) ? 0 : sat[(size_t)w * ( y - 1 ) + x];
		uint32_t c = ( x <= 0 ) ? 0 : sat[(size_t)w * y + ( x - 1 )];
		uint32_t d = sat[(size_t)w * y + x];
		return ( d - b ) + ( a - c );
	}
	// renderCommon.hpp:367-435
	void importanceSample( F3* dir, F3* L, float* srPDF, F3 N, bool axisAligned, float u0, float u1, float u2, float u3, uint32_t* XY ) const
	{
		const uint32_t* sat = sats;
		const size_t n = (size_t)w * h;
		if( axisAligned )
		{
			const float k = 0.8f;
			if( k < N.x )
				sat = sats + 1 * n;
			else if( N.x < -k )
				sat = sats + 2 * n;
			else if( k < N.y )
				sat = sats + 3 * n;
			else if( N.y < -k )
				sat = sats + 4 * n;
			else if( k < N.z )
				sat = sats + 5 * n;
			else if( N.z < -k )
				sat = sats + 6 * n;
		}
		int i = 0, j = w;
		while( i < j )
		{
			const int m = ( i + j ) / 2;
			const float v = (float)psH( sat, m ) / (float)0xFFFFFFFFu;
			if( v <= u0 )
				i = m + 1;
			else
				j = m;
		}
		uint32_t X = (uint32_t)i - 1;
		uint32_t vol = psH( sat, X + 1 ) - psH( sat, X );
		i = 0, j = h;
		while( i < j )
		{
			const int m = ( i + j ) / 2;
			const float v = (float)psV( sat, X, m ) / (float)vol;
			if( v <= u1 )
				i = m + 1;
			else
				j = m;
		}
		uint32_t Y = (uint32_t)i - 1;
		float pSelection = (float)count( sat, X, Y ) / (float)0xFFFFFFFF;
		float dTheta = PI / (float)h;
		float dPhi = 2.0f * PI / (float)w;
		float theta = Y * dTheta;
		float dH = 2.0f * sinf( dTheta * 0.5f ) * sinf( dTheta * 0.5f + theta );
		float sr = dH * dPhi;
		float sY = mixf( cosf( theta ), cosf( theta + dTheta ), u2 );
		float phi = dPhi * ( (float)X + u3 ) + PI;
		float sX = cosf( phi );
		float sZ = sinf( phi );
		float sinTheta = sqrtf( smaxf( 1.0f - sY * sY, 0.0f ) );
		*dir = { sX * sinTheta, sY, sZ * sinTheta };
		*srPDF = pSelection / sr;
		const float* c = pixels + ( (size_t)Y * w + X ) * 4;
		*L = F3{ c[0], c[1], c[2] } * scale;
		if( XY )
		{
			XY[0] = X;
			XY[1] = Y;
		}
	}
	// renderCommon.hpp:354-365 (device-only in the reference)
	F3 sampleNearest( F3 d, bool isPrimary ) const
	{
		if( !pixels ) return F3{ 0.0f, 0.0f, 0.0f }; // no environment loaded (the reference would dereference null)
		int ww = isPrimary ? wP : w;
		int hh = isPrimary ? hP : h;
		float phi = atan2f( d.z, d.x ) + PI;
		float theta = atan2f( sqrtf( d.x * d.x + d.z * d.z ), d.y );
		float u = phi / ( PI * 2.0f ), v = theta / PI;
		int x = (int)sclamp( u * ww, 0.0f, (float)( ww - 1.0f ) );
		int y = (int)sclamp( v * hh, 0.0f, (float)( hh - 1.0f ) );
		size_t index = (size_t)y * ww + x;
		const float* c = ( isPrimary && primary ) ? primary + index * 4 : pixels + index * 4;
		return F3{ c[0], c[1], c[2] } * scale;
	}
};

// ---- SAT build: voxKernel.cu:485-608 ----
// HDRIstoreImportance (:485-524) with host math
void storeImportance( const float* rgba, int w, int h, double* sat, int cosWeighted, F3 axis )
{
	for( int pixelY = 0; pixelY < h; pixelY++ )
		for( int pixelX = 0; pixelX < w; pixelX++ )
		{
			size_t pixelIdx = (size_t)pixelY * w + pixelX;
			float dTheta = PI / (float)h;
			float dPhi = 2.0f * PI / (float)w;
			float theta = pixelY * dTheta;
			float dH = 2.0f * sinf( dTheta * 0.5f ) * sinf( dTheta * 0.5f + theta );
			float dW = dPhi;
			float sr = dH * dW;
			const float* c = rgba + pixelIdx * 4;
			float wgt = 1.0f;
			if( cosWeighted )
			{
				float sY = mixf( cosf( theta ), cosf( theta + dTheta ), 0.5f );
				float phi = dPhi * ( (float)pixelX + 0.5f ) + PI;
				float sX = cosf( phi );
				float sZ = sinf( phi );
				float sinTheta = sqrtf( smaxf( 1.0f - sY * sY, 0.0f ) );
				F3 dirCenter = { sX * sinTheta, sY, sZ * sinTheta };
				wgt = smaxf( dot( axis, dirCenter ), 0.0f );
			}
			sat[pixelIdx] = luminance( c[0], c[1], c[2] ) * sr * wgt; // float product, widened on store
		}
}
// prefixSumInclusive<double,512> (:526-555): Hillis-Steele over one 512-element tile, in the reference's summation order
const int SAT_BLOCK = 512;
double scanTile( double prefix, double* s )
{
	double tmp[SAT_BLOCK];
	for( int offset = 1; offset < SAT_BLOCK; offset <<= 1 )
	{
		for( int i = 0; i < SAT_BLOCK; i++ ) tmp[i] = ( offset <= i ) ? s[i] + s[i - offset] : s[i];
		memcpy( s, tmp, sizeof( tmp ) );
	}
	double sum = s[SAT_BLOCK - 1];
	for( int i = 0; i < SAT_BLOCK; i++ ) s[i] += prefix;
	return sum;
}
// buildSATh / buildSATv (:557-598): `stride` walks a row (1) or a column (w)
void scanLine( double* line, int n, size_t stride )
{
	double s[SAT_BLOCK];
	double prefix = 0.0;
	for( int i = 0; i < n; i += SAT_BLOCK )
	{
		for( int k = 0; k < SAT_BLOCK; k++ ) s[k] = ( i + k < n ) ? line[(size_t)( i + k ) * stride] : 0.0;
		prefix += scanTile( prefix, s );
		for( int k = 0; k < SAT_BLOCK && i + k < n; k++ ) line[(size_t)( i + k ) * stride] = s[k];
	}
}
void buildSat( const float* rgba, int w, int h, uint32_t* out, int cosWeighted, F3 axis, int nThreads )
{
	std::vector<double> sat( (size_t)w * h );
	storeImportance( rgba, w, h, sat.data(), cosWeighted, axis );
	auto par = [&]( int n, auto fn ) {
		std::vector<std::thread> th;
		for( int k = 0; k < nThreads; k++ )
			th.emplace_back( [&, k]() { for( int i = k; i < n; i += nThreads ) fn( i ); } );
		for( auto& t : th ) t.join();
	};
	par( h, [&]( int y ) { scanLine( sat.data() + (size_t)y * w, w, 1 ); } );
	par( w, [&]( int x ) { scanLine( sat.data() + x, h, (size_t)w ); } );
	// buildSAT2u32 (:600-608)
	size_t n = (size_t)w * h;
	double sum = sat[n - 1];
	for( size_t i = 0; i < n; i++ ) out[i] = (uint32_t)( sat[i] / ( sum ) * (double)0xFFFFFFFFu );
}

// getHitN, voxCommon.hpp:564-577
inline F3 getHitN( int major, F3 rd )
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
inline F3 rawReflectance( const uint8_t* c ) { return { (float)c[0] / 255.0f, (float)c[1] / 255.0f, (float)c[2] / 255.0f }; } // renderCommon.hpp:160-166

struct PtScene
{
	const OrcOctree* octree;
	const uint8_t* attrs; // VoxelAttirb[nVoxels]: color uchar4, emission uchar4 (voxCommon.hpp:121-125)
	uint32_t hasEmission;
	float emissionScale; // IntersectorOctreeGPU.hpp:273
	Hdri hdri;
	const float* pmj;
	int maxDepth;
};
inline void isect( const PtScene& s, F3 ro, F3 rd, bool shadow, float* t, int* nMajor, uint32_t* vIndex, uint64_t* rays )
{
	orc_intersect_batch( s.octree, &ro.x, &rd.x, 1, shadow ? 1 : 0, t, nMajor, vIndex, nullptr );
	( *rays )++;
}
inline F3 voxelEmission( const PtScene& s, uint32_t vIndex, bool withScale )
{
	return rawReflectance( s.attrs + (size_t)vIndex * 8 + 4 ) * ( withScale ? s.emissionScale : 1.0f );
}

// one (pixel, spp) path of renderPT, voxKernel.cu:633-761; returns L
F3 tracePath( const PtScene& s, const OrcCamera& cam, uint32_t x, uint32_t y, uint32_t pixelIdx, uint32_t spp, int w, int h, uint64_t* rays )
{
	Murmur hash( 0 );
	hash.combine( pixelIdx );
	int dim = 0;
	const uint32_t stream = hash.getHash();
	auto SAMPLE_2D = [&]( float* a, float* b ) { sample2d( s.pmj, spp, (uint32_t)( dim++ ), stream, a, b ); };

	float c0, c1, l0, l1;
	SAMPLE_2D( &c0, &c1 );
	SAMPLE_2D( &l0, &l1 );
	F3 ro, rd;
	orc_camera_shoot_thin_lens( &cam, (int)x, (int)y, c0, c1, w, h, l0, l1, &ro.x, &rd.x );

	F3 T = { 1.0f, 1.0f, 1.0f };
	F3 L = { 0.0f, 0.0f, 0.0f };
	float t = FLT_MAX;
	int nMajor = 0;
	uint32_t vIndex = 0;
	isect( s, ro, rd, false, &t, &nMajor, &vIndex, rays );
	if( t == FLT_MAX )
		L = L + T * s.hdri.sampleNearest( rd, true );
	else
		L = L + T * voxelEmission( s, vIndex, false );

	for( int depth = 0; depth < s.maxDepth && t != FLT_MAX; depth++ )
	{
		F3 R = rawReflectance( s.attrs + (size_t)vIndex * 8 );
		F3 hitN = getHitN( nMajor, rd );
		F3 hitP = ro + rd * t;
		if( s.hdri.enabled() )
		{
			float u0, u1, u2, u3;
			SAMPLE_2D( &u0, &u1 );
			SAMPLE_2D( &u2, &u3 );
			F3 dir, emissive;
			float p;
			s.hdri.importanceSample( &dir, &emissive, &p, hitN, true, u0, u1, u2, u3, nullptr );
			float t2 = FLT_MAX;
			int nm2 = 0;
			uint32_t vi2 = 0;
			isect( s, hitP, dir, true, &t2, &nm2, &vi2, rays );
			if( t2 == FLT_MAX ) L = L + T * ( R / PI ) * smaxf( dot( hitN, dir ), 0.0f ) * emissive / p;
		}
		T = T * R;
		const int nSampleExtraDirect = s.hasEmission ? 1 : 0; // EXTRA_IMPLICIT_SAMPLING, renderCommon.hpp:15
		for( int k = 0; depth == 0 && k < nSampleExtraDirect; k++ )
		{
			float a, b;
			SAMPLE_2D( &a, &b );
			F3 dir = sampleLambertian( a, b, hitN );
			float t2 = FLT_MAX;
			int nm2 = 0;
			uint32_t vi2 = 0;
			isect( s, hitP, dir, false, &t2, &nm2, &vi2, rays );
			if( t2 != FLT_MAX ) L = L + T * voxelEmission( s, vi2, true ) / (float)( 1 + nSampleExtraDirect );
		}
		float a, b;
		SAMPLE_2D( &a, &b );
		F3 dir = sampleLambertian( a, b, hitN );
		ro = hitP;
		rd = dir;
		t = FLT_MAX;
		isect( s, ro, rd, false, &t, &nMajor, &vIndex, rays );
		if( t != FLT_MAX )
		{
			F3 Le = voxelEmission( s, vIndex, true );
			L = L + T * Le * ( depth == 0 ? 1.0f / (float)( 1 + nSampleExtraDirect ) : 1.0f );
		}
	}
	return L;
}
} // namespace

extern "C" {
uint32_t orc_murmur_combine( uint32_t seed, const uint32_t* values, int n )
{
	Murmur h( seed );
	for( int i = 0; i < n; i++ ) h.combine( values[i] );
	return h.getHash();
}
void orc_pcg32_stream( uint64_t seed, uint64_t stream, int n, uint32_t* out )
{
	PCG32 rng;
	rng.setup( seed, stream );
	for( int i = 0; i < n; i++ ) out[i] = rng.nextU32();
}
uint32_t orc_nested_uniform_scramble( uint32_t x, uint32_t seed ) { return nus( x, seed ); }
float orc_scramble_f32( float x, uint32_t seed ) { return scramble_f32( x, seed ); }
// PMJSampler::setup, pmjSampler.hpp:114-134: 128 sequences x 4096 samples x 2 floats from PCG32( 0, 2525 )
void orc_pmj_table( float* out )
{
	PCG32 rng;
	rng.setup( 0, 2525 );
	for( int i = 0; i < 128; i++ ) pmj02( 4096, out + (size_t)2 * 4096 * i, rng );
}
void orc_pmj_sample2d( const float* table, uint32_t sampleIdx, uint32_t dimension, uint32_t stream, float* xy )
{
	sample2d( table, sampleIdx, dimension, stream, &xy[0], &xy[1] );
}
// HDRI::load, renderCommon.hpp:214-314: sats = [7][w*h] (global, +x, -x, +y, -y, +z, -z)
void orc_hdri_build_sats( const float* rgba, int w, int h, uint32_t* sats )
{
	int nThreads = (int)std::thread::hardware_concurrency();
	if( nThreads < 1 ) nThreads = 1;
	const F3 axes[6] = { { 1, 0, 0 }, { -1, 0, 0 }, { 0, 1, 0 }, { 0, -1, 0 }, { 0, 0, 1 }, { 0, 0, -1 } };
	const size_t n = (size_t)w * h;
	buildSat( rgba, w, h, sats, 0, F3{ 0, 0, 0 }, nThreads );
	for( int i = 0; i < 6; i++ ) buildSat( rgba, w, h, sats + n * ( i + 1 ), 1, axes[i], nThreads );
}
void orc_hdri_importance_sample( const float* rgba, const uint32_t* sats, int w, int h, float scale, const float* N, int axisAligned,
								 float u0, float u1, float u2, float u3, float* dir, float* L, float* srPDF, uint32_t* XY )
{
	Hdri hd{ rgba, nullptr, sats, w, h, w, h, scale };
	F3 d, l;
	hd.importanceSample( &d, &l, srPDF, F3{ N[0], N[1], N[2] }, axisAligned != 0, u0, u1, u2, u3, XY );
	dir[0] = d.x, dir[1] = d.y, dir[2] = d.z;
	L[0] = l.x, L[1] = l.y, L[2] = l.z;
}
void orc_hdri_sample_nearest( const float* rgba, int w, int h, float scale, const float* d, float* out )
{
	Hdri hd{ rgba, nullptr, nullptr, w, h, w, h, scale };
	F3 c = hd.sampleNearest( F3{ d[0], d[1], d[2] }, false );
	out[0] = c.x, out[1] = c.y, out[2] = c.z;
}
void orc_sample_lambertian( float a, float b, const float* Ng, float* out )
{
	F3 d = sampleLambertian( a, b, F3{ Ng[0], Ng[1], Ng[2] } );
	out[0] = d.x, out[1] = d.y, out[2] = d.z;
}

// renderPT (voxKernel.cu:610-777), rows [y0,y1) of one step (= 16 spp, spp = iteration*16 + localSpp).
// accum: float4[w*h] (rgb sum, w = spp count), accumulated INTO (`+=`, :770-774).  The reference sums the 16 samples of a
// pixel with shared-memory atomicAdd (order unspecified); here they are summed in localSpp order, then added to the pixel.
// hdriRgba == NULL or hdriScale <= 0 disables the environment (HDRI::isEnabled).  stats (optional): [0] += rays traced.
void orc_pt_render( const OrcOctree* octree, const uint8_t* attrs, uint32_t hasEmission, const OrcCamera* cam, int w, int h, int y0, int y1,
					int iteration, const float* hdriRgba, const uint32_t* sats, int hw, int hh, float hdriScale,
					const float* primaryRgba, int wP, int hP, const float* pmjTable, int maxDepth, float* accum, uint64_t* stats, int nThreads )
{
	PtScene s;
	s.octree = octree;
	s.attrs = attrs;
	s.hasEmission = hasEmission;
	s.emissionScale = 7.5f;
	s.hdri = Hdri{ hdriRgba, primaryRgba, sats, hw, hh, primaryRgba ? wP : hw, primaryRgba ? hP : hh, hdriRgba ? hdriScale : 0.0f };
	s.pmj = pmjTable;
	s.maxDepth = maxDepth;
	if( nThreads < 1 ) nThreads = 1;
	std::vector<uint64_t> rays( nThreads, 0 );
	std::vector<std::thread> th;
	for( int k = 0; k < nThreads; k++ )
		th.emplace_back( [&, k]() {
			for( int y = y0 + k; y < y1; y += nThreads )
				for( int x = 0; x < w; x++ )
				{
					const uint32_t pixelIdx = (uint32_t)y * w + x;
					float lx = 0.0f, ly = 0.0f, lz = 0.0f;
					for( uint32_t localSpp = 0; localSpp < 16; localSpp++ )
					{
						F3 L = tracePath( s, *cam, x, y, pixelIdx, (uint32_t)iteration * 16 + localSpp, w, h, &rays[k] );
						lx += L.x, ly += L.y, lz += L.z;
					}
					float* px = accum + (size_t)pixelIdx * 4;
					px[0] += lx, px[1] += ly, px[2] += lz, px[3] += 16.0f;
				}
		} );
	for( auto& t : th ) t.join();
	if( stats )
		for( int k = 0; k < nThreads; k++ ) stats[0] += rays[k];
}
// renderResolve, voxKernel.cu:779-795 (powf on the host, vectorMath.hpp:93-97)
void orc_pt_resolve( const float* accum, uint64_t n, uint8_t* rgba )
{
	for( uint64_t i = 0; i < n; i++ )
	{
		const float* v = accum + i * 4;
		for( int c = 0; c < 3; c++ )
		{
			float f = 255 * powf( v[c] / v[3], 1.0f / 2.2f ) + 0.5f;
			int q = _mm_cvttss_si32( _mm_set_ss( f ) );
			rgba[i * 4 + c] = (uint8_t)( q < 255 ? q : 255 );
		}
		rgba[i * 4 + 3] = 255;
	}
}
}
