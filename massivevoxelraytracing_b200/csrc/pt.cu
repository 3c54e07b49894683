// pt.cu -- the path tracer behind hako_pt_* (PathTracer, PathTracer.hpp:14-170; renderPT / renderResolve / the HDRI
// SAT kernels, voxKernel.cu:485-795; PMJSampler, pmjSampler.hpp; HDRI, renderCommon.hpp:204-481).
//
// What is kept from the reference: the sample mapping (stream = Murmur(pixelIdx), spp = iteration*16 + localSpp, the
// dimension order of SURVEY.md appendix A 28-30), the estimator, and every fp32 expression in its operation order.
// What is redesigned for B200:
//   - no megakernel with four divergent traversal call sites and a 1 GiB global stack pool: a PERSISTENT kernel runs a
//     per-lane path state machine in which every loop iteration traces exactly one ray per lane -- primary, shadow,
//     extra-implicit and bounce rays of different lanes go through ONE traversal call site, so a warp stays converged
//     across ray kinds;
//   - lanes whose path ended are refilled at once: warp ballot -> rank by popc -> indices from a per-warp chunk that is
//     fetched from a global work counter (one atomic per 256 paths);
//   - radiance is written per path and reduced per pixel in localSpp order by a second kernel, so a step is
//     DETERMINISTIC (the reference's shared-memory atomicAdd order is not) and equals the oracle's summation order;
//   - theta-only trigonometry of the lat-long map is tabulated per row on the host (exactly the values the CPU code
//     computes), the SAT build is bit-exact with the oracle (same fp32 products, same Hillis-Steele double sums).
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "internal.cuh"
#include "scan.cuh"
#include "traverse.cuh"

using namespace hako;

#define PT_SPP 16		  // nBatchSpp, voxKernel.cu:631
#define PT_BLOCK 256
#define PT_CHUNK 256	  // paths fetched from the global counter per warp-level atomic
#define PMJ_LENGTH 4096	  // pmjSampler.hpp:107-111
#define PMJ_NSEQ 128
#define PT_COUNTERS 64	  // [0] work, [1] rays traced, [2 + v] live paths at vertex v, [32 + v] ray work counter of wave v, [63] traversal steps of the trace waves
#define PT_LANES 4 // concurrent row bands of a wavefront step; each has its own PT_COUNTERS block and wave buffers
#define PT_MAX_WAVES 28

struct hako_pt
{
	hako_ctx* ctx = nullptr;
	int w = 0, h = 0;
	float4* accum = nullptr; // rgb sum, w = spp count (PathTracer.hpp:81-97)
	uchar4* rgba8 = nullptr;
	float* pmj = nullptr; // [128][4096][2]
	// environment (HDRI, renderCommon.hpp:472-480)
	float4* hdriPixels = nullptr;
	float4* hdriPrimary = nullptr;
	uint32_t* sats = nullptr; // [7][hw*hh]: global, +x, -x, +y, -y, +z, -z
	float4* rowTab = nullptr; // per row: { sr, cos(theta), cos(theta+dTheta), 0 }
	int hw = 0, hh = 0, wP = 0, hP = 0;
	float hdriScale = 1.75f; // HDRI::m_scale, renderCommon.hpp:480
	// per-path radiance of the step in flight
	float4* pathL = nullptr;
	uint64_t pathCap = 0;
	unsigned long long* counters = nullptr; // [0] work counter, [1] rays traced
	uint32_t* hitQueue = nullptr;			// path indices whose primary ray hit (wavefront modes), capacity pathCap
	// full wavefront mode: ping-pong vertex lists, ray directions, ray results; capacity waveCap live paths (3 rays each).
	// One set per LANE: the row bands of a step run on up to PT_LANES internal streams at once, because every trace wave ends
	// with a tail in which a few long rays hold the kernel open (a ray is a serial chain of dependent node fetches) while most
	// SMs idle; the waves of another band fill those tails (tools/pt_overlap_probe.py: 17.6 -> 15.8 ms per 1080p step).
	void* waveV[PT_LANES][2] = {};
	float4* waveDir[PT_LANES][2] = {};
	float4* waveRes[PT_LANES] = {};
	cudaStream_t laneStream[PT_LANES] = {};
	cudaEvent_t laneEv[PT_LANES + 1] = {};
	int lanes = 3; // HAKO_PT_LANES
	uint64_t waveCap = 0;
	int waveRefillMin = 24, waveBurst = 16; // traceWaveKernel knobs (HAKO_PT_REFILL_MIN / HAKO_PT_BURST); B200 sweep in profiles/r1f_pt_sweep.txt
	int waveTraceCtas = 5;
	int waveSmemStack = 1;	   // HAKO_PT_SMEM_STACK=0: traversal stack of the trace waves in local memory (A/B reference)
	int waveRefillMinSmem = 4; // refill threshold of the trace waves when their stack is in shared memory (refills are cheap there)
	// ray binning in front of every trace wave (HAKO_PT_BIN = bits per axis of the origin grid, 0 = off): the rays of a wave are
	// counting-sorted by ( kind, direction octant, Morton code of the origin cell ) and the trace kernel draws them in that order
	uint32_t* binKeys[PT_LANES] = {};
	uint32_t* binOrder[PT_LANES] = {};
	uint32_t* binHist[PT_LANES] = {}; // scan state, then one counter per bin
	uint64_t binCap = 0;
	int binBits = 0, binCellMajor = 0;
	uint64_t waveBandPaths = 16u << 20;	  // paths per internal row band of a step (bounds the wavefront buffers)
	// 0: single persistent kernel (pathTraceKernel<false>); 1: primaryWaveKernel + pathTraceKernel<true>;
	// 2: full wavefront (primaryWaveKernel, then shadeWaveKernel / traceWaveKernel per vertex).  HAKO_PT_WAVEFRONT overrides
	// (A/B knob, read at create time).
	int wavefront = 2;
};

namespace
{
struct PtEnv
{
	const float4* pixels;
	const float4* primary;
	const uint32_t* sats;
	const float4* rowTab;
	int w, h, wP, hP;
	float scale;
};
struct PtParams
{
	OctreeView tree;
	PtEnv env;
	hako_camera cam;
	const float* pmj;
	float4* pathL;
	unsigned long long* counters; // [0] work counter, [1] rays traced, [2] hit-queue length
	uint32_t* hitQueue;
	int w, h, y0;
	uint32_t nPaths;
	uint32_t iteration;
	int maxDepth;
};

// ---- PMJ02 lookup with Owen scrambling, pmjSampler.hpp:62-102,158-173 ----
__device__ __forceinline__ uint32_t hashCombine3( uint32_t a, uint32_t b, uint32_t c )
{
	Murmur32 h( a );
	h.combine( b );
	h.combine( c );
	return h.getHash();
}
__device__ __forceinline__ uint32_t laineKarras( uint32_t x, uint32_t seed )
{
	x += seed;
	x ^= x * 0x6c50b47cu;
	x ^= x * 0xb82f1e52u;
	x ^= x * 0xc7afe638u;
	x ^= x * 0x8d22f6e6u;
	return x;
}
__device__ __forceinline__ uint32_t nestedUniformScramble( uint32_t x, uint32_t seed ) { return __brev( laineKarras( __brev( x ), seed ) ); }
__device__ __forceinline__ float scrambleF32( float x, uint32_t seed )
{
	x += 1.0f;
	const uint32_t bits = __float_as_uint( x );
	return __uint_as_float( 0x3f800000u | ( nestedUniformScramble( bits & 0x7FFFFFu, seed ) & 0x7FFFFFu ) ) - 1.0f;
}
__device__ __forceinline__ float2 pmjSample2d( const float* __restrict__ table, uint32_t sampleIdx, uint32_t dimension, uint32_t stream )
{
	sampleIdx = nestedUniformScramble( sampleIdx, hashCombine3( stream, dimension, 31082745u ) ) & ( PMJ_LENGTH - 1 );
	dimension = nestedUniformScramble( dimension, hashCombine( stream, 54761983u ) ) & ( PMJ_NSEQ - 1 );
	const float2 s = __ldg( (const float2*)table + ( dimension * PMJ_LENGTH + sampleIdx ) );
	return make_float2( scrambleF32( s.x, hashCombine3( stream, dimension, 83927105u ) ), scrambleF32( s.y, hashCombine3( stream, dimension, 12654890u ) ) );
}

// ---- sampleLambertian / GetOrthonormalBasis, renderCommon.hpp:119-151 ----
__device__ __forceinline__ V3 sampleLambertian( float a, float b, V3 Ng )
{
	const float r = sqrtf( a );
	const float theta = b * HAKO_PI * 2.0f;
	const float x = r * cosf( theta );
	const float y = r * sinf( theta );
	const float z = sqrtf( smax( 1.0f - a, 0.0f ) );
	const float sign = copysignf( 1.0f, Ng.z );
	const float aa = -1.0f / ( sign + Ng.z );
	const float bb = Ng.x * Ng.y * aa;
	const V3 xaxis = { 1.0f + sign * Ng.x * Ng.x * aa, sign * bb, -sign * Ng.x };
	const V3 yaxis = { bb, sign + Ng.y * Ng.y * aa, -Ng.y };
	return xaxis * x + yaxis * y + Ng * z;
}

// ---- HDRI, renderCommon.hpp:354-470 ----
__device__ __forceinline__ uint32_t satPsH( const PtEnv& e, const uint32_t* __restrict__ sat, uint32_t x )
{
	return x == 0 ? 0u : __ldg( sat + (size_t)e.w * ( e.h - 1 ) + x - 1 );
}
__device__ __forceinline__ uint32_t satPsV( const PtEnv& e, const uint32_t* __restrict__ sat, uint32_t x, uint32_t y )
{
	if( y == 0 ) return 0u;
	const uint32_t s0 = x == 0 ? 0u : __ldg( sat + (size_t)e.w * ( y - 1 ) + ( x - 1 ) );
	const uint32_t s1 = __ldg( sat + (size_t)e.w * ( y - 1 ) + x );
	return s1 - s0;
}
__device__ __forceinline__ uint32_t satCount( const PtEnv& e, const uint32_t* __restrict__ sat, uint32_t x, uint32_t y )
{
	const uint32_t a = ( x == 0 || y == 0 ) ? 0u : __ldg( sat + (size_t)e.w * ( y - 1 ) + ( x - 1 ) );
	const uint32_t b = ( y == 0 ) ? 0u : __ldg( sat + (size_t)e.w * ( y - 1 ) + x );
	const uint32_t c = ( x == 0 ) ? 0u : __ldg( sat + (size_t)e.w * y + ( x - 1 ) );
	const uint32_t d = __ldg( sat + (size_t)e.w * y + x );
	return ( d - b ) + ( a - c );
}
__device__ __forceinline__ void hdriImportanceSample( const PtEnv& e, V3* dir, V3* L, float* srPDF, V3 N, float u0, float u1, float u2, float u3 )
{
	const size_t n = (size_t)e.w * e.h;
	const uint32_t* sat = e.sats;
	const float k = 0.8f; // axis-aligned tables, :372-399
	if( k < N.x )
		sat += 1 * n;
	else if( N.x < -k )
		sat += 2 * n;
	else if( k < N.y )
		sat += 3 * n;
	else if( N.y < -k )
		sat += 4 * n;
	else if( k < N.z )
		sat += 5 * n;
	else if( N.z < -k )
		sat += 6 * n;

	// upper_bound_f (:187-202) over the marginal, then over the column
	int i = 0, j = e.w;
	while( i < j )
	{
		const int m = ( i + j ) / 2;
		const float v = (float)satPsH( e, sat, m ) / (float)0xFFFFFFFFu;
		if( v <= u0 )
			i = m + 1;
		else
			j = m;
	}
	const uint32_t X = (uint32_t)i - 1;
	const uint32_t vol = satPsH( e, sat, X + 1 ) - satPsH( e, sat, X );
	i = 0, j = e.h;
	while( i < j )
	{
		const int m = ( i + j ) / 2;
		const float v = (float)satPsV( e, sat, X, m ) / (float)vol;
		if( v <= u1 )
			i = m + 1;
		else
			j = m;
	}
	const uint32_t Y = (uint32_t)i - 1;
	const float pSelection = (float)satCount( e, sat, X, Y ) / (float)0xFFFFFFFFu;
	const float dPhi = 2.0f * HAKO_PI / (float)e.w;
	const float4 row = __ldg( e.rowTab + Y ); // sr, cos(theta), cos(theta+dTheta): theta-only terms, tabulated by the host
	const float sY = mixf( row.y, row.z, u2 );
	const float phi = dPhi * ( (float)X + u3 ) + HAKO_PI;
	const float sX = cosf( phi );
	const float sZ = sinf( phi );
	const float sinTheta = sqrtf( smax( 1.0f - sY * sY, 0.0f ) );
	*dir = { sX * sinTheta, sY, sZ * sinTheta };
	*srPDF = pSelection / row.x;
	const float4 c = __ldg( e.pixels + (size_t)Y * e.w + X );
	*L = V3{ c.x, c.y, c.z } * e.scale;
}
__device__ __forceinline__ V3 hdriSampleNearest( const PtEnv& e, V3 d, bool isPrimary )
{
	const int w = isPrimary ? e.wP : e.w;
	const int h = isPrimary ? e.hP : e.h;
	const float phi = atan2f( d.z, d.x ) + HAKO_PI; // getSpherical, :176-181
	const float theta = atan2f( sqrtf( d.x * d.x + d.z * d.z ), d.y );
	const float u = phi / ( HAKO_PI * 2.0f ), v = theta / HAKO_PI;
	const int x = f2i_x86( smin( smax( u * w, 0.0f ), (float)( w - 1.0f ) ) );
	const int y = f2i_x86( smin( smax( v * h, 0.0f ), (float)( h - 1.0f ) ) );
	const size_t index = (size_t)y * w + x;
	const float4 c = ( isPrimary && e.primary ) ? __ldg( e.primary + index ) : __ldg( e.pixels + index );
	return V3{ c.x, c.y, c.z } * e.scale;
}
// rawReflectance, renderCommon.hpp:160-166, on a packed uchar4
__device__ __forceinline__ V3 rawReflectance( uint32_t c ) { return { (float)( c & 0xFFu ) / 255.0f, (float)( ( c >> 8 ) & 0xFFu ) / 255.0f, (float)( ( c >> 16 ) & 0xFFu ) / 255.0f }; }
// VoxelAttirb (voxCommon.hpp:121-125) as one 8-byte read-only load: .x = color, .y = emission
__device__ __forceinline__ uint2 loadAttrib( const OctreeView& v, uint32_t vIndex ) { return __ldg( (const uint2*)v.attrs + vIndex ); }
} // namespace

namespace
{
enum : uint32_t
{
	RAY_PRIMARY = 0,
	RAY_SHADOW = 1,
	RAY_EXTRA = 2,
	RAY_BOUNCE = 3,
	LANE_IDLE = 4
};
enum : uint32_t
{
	NEXT_TRACE = 0,
	NEXT_SHADE,
	NEXT_GEN_EXTRA,
	NEXT_GEN_BOUNCE,
	NEXT_FINISH
};

// per-lane path state of the persistent kernels
struct PathState
{
	uint32_t kind, path, spp, stream, dim;
	int depth;
	V3 o, d, T, L, pending, hitN;
};

// path index -> (pixel, localSpp) keeps the reference's "16 adjacent threads share a pixel" (voxKernel.cu:635-637);
// camera + lens dimensions, thin-lens primary ray (:640-672)
__device__ __forceinline__ void startPath( const PtParams& p, PathState& s, uint32_t path )
{
	s.path = path;
	const uint32_t localPixel = path / PT_SPP, localSpp = path % PT_SPP;
	const uint32_t x = localPixel % (uint32_t)p.w, y = (uint32_t)p.y0 + localPixel / (uint32_t)p.w;
	const uint32_t pixelIdx = y * (uint32_t)p.w + x;
	s.spp = p.iteration * PT_SPP + localSpp;
	Murmur32 hs( 0 );
	hs.combine( pixelIdx );
	s.stream = hs.getHash();
	s.dim = 0;
	const float2 cam_u01 = pmjSample2d( p.pmj, s.spp, s.dim++, s.stream );
	const float2 lens_u01 = pmjSample2d( p.pmj, s.spp, s.dim++, s.stream );
	cameraShootThinLens( p.cam, (int)x, (int)y, cam_u01.x, cam_u01.y, p.w, p.h, lens_u01.x, lens_u01.y, &s.o, &s.d );
	s.T = { 1.0f, 1.0f, 1.0f };
	s.L = { 0.0f, 0.0f, 0.0f };
	s.depth = 0;
	s.kind = RAY_PRIMARY;
}

// ---- the pieces of one path vertex (voxKernel.cu:678-759), shared by the persistent kernel and the wavefront kernels ----
__device__ __forceinline__ V3 emissionOf( const PtParams& p, uint32_t vIndex ) { return rawReflectance( loadAttrib( p.tree, vIndex ).y ); }
// a camera ray that hits an emitter sees it unscaled, :686-688
__device__ __forceinline__ void addPrimaryHit( const PtParams& p, PathState& s, uint32_t vIndex ) { s.L = s.L + s.T * emissionOf( p, vIndex ); }
// :712-715
__device__ __forceinline__ void addShadowResult( PathState& s, bool hit )
{
	if( !hit ) s.L = s.L + s.pending;
}
// :733-736
__device__ __forceinline__ void addExtraResult( const PtParams& p, const uint32_t nExtra, PathState& s, bool hit, uint32_t vIndex )
{
	if( hit ) s.L = s.L + s.T * ( emissionOf( p, vIndex ) * p.tree.emissionScale ) / (float)( 1u + nExtra );
}
// :750-758; s.depth is the index of the vertex the bounce ray left from
__device__ __forceinline__ void addBounceResult( const PtParams& p, const uint32_t nExtra, PathState& s, bool hit, uint32_t vIndex )
{
	if( hit ) s.L = s.L + s.T * ( emissionOf( p, vIndex ) * p.tree.emissionScale ) * ( s.depth == 0 ? 1.0f / (float)( 1u + nExtra ) : 1.0f );
}
// the surface point the path ray (s.o, s.d) just hit, :693-716: moves s.o to hitP, sets hitN, multiplies the throughput, and --
// when the environment is on -- draws the light sample: s.pending is its contribution, `shadowDir` the shadow ray (from hitP)
__device__ __forceinline__ void shadeSurface( const PtParams& p, const bool envOn, PathState& s, float t, int nMajor, uint32_t vIndex, V3* shadowDir )
{
	const V3 R = rawReflectance( loadAttrib( p.tree, vIndex ).x );
	s.hitN = getHitN( nMajor, s.d );
	s.o = s.o + s.d * t; // hitP: every following ray of this vertex starts here, without an epsilon
	if( envOn )
	{
		const float2 u01 = pmjSample2d( p.pmj, s.spp, s.dim++, s.stream );
		const float2 u23 = pmjSample2d( p.pmj, s.spp, s.dim++, s.stream );
		V3 dir, emissive;
		float pdf;
		hdriImportanceSample( p.env, &dir, &emissive, &pdf, s.hitN, u01.x, u01.y, u23.x, u23.y );
		s.pending = s.T * ( R / HAKO_PI ) * smax( dot( s.hitN, dir ), 0.0f ) * emissive / pdf;
		*shadowDir = dir;
	}
	s.T = s.T * R;
}
// the cosine-weighted direction of the extra-implicit (:720-731) and bounce (:741-748) rays
__device__ __forceinline__ V3 lambertDir( const PtParams& p, PathState& s )
{
	const float2 u01 = pmjSample2d( p.pmj, s.spp, s.dim++, s.stream );
	return sampleLambertian( u01.x, u01.y, s.hitN );
}

// consume the result of the ray a lane just traced and set up its next ray (or finish the path: kind = LANE_IDLE)
__device__ __forceinline__ void shadeVertex( const PtParams& p, const bool envOn, const uint32_t nExtra, PathState& s, const bool hit, const float t,
											 const int nMajor, const uint32_t vIndex )
{
	uint32_t next;
	if( s.kind == RAY_PRIMARY )
	{
		if( !hit )
		{
			if( p.env.pixels ) s.L = s.L + s.T * hdriSampleNearest( p.env, s.d, true ); // :678-683
			next = NEXT_FINISH;
		}
		else
		{
			addPrimaryHit( p, s, vIndex );
			next = NEXT_SHADE;
		}
	}
	else if( s.kind == RAY_SHADOW )
	{
		addShadowResult( s, hit );
		next = ( s.depth == 0 && nExtra ) ? NEXT_GEN_EXTRA : NEXT_GEN_BOUNCE;
	}
	else if( s.kind == RAY_EXTRA )
	{
		addExtraResult( p, nExtra, s, hit, vIndex );
		next = NEXT_GEN_BOUNCE;
	}
	else
	{
		addBounceResult( p, nExtra, s, hit, vIndex );
		s.depth++;
		next = ( hit && s.depth < p.maxDepth ) ? NEXT_SHADE : NEXT_FINISH;
	}

	if( next == NEXT_SHADE )
	{
		V3 dir;
		shadeSurface( p, envOn, s, t, nMajor, vIndex, &dir );
		if( envOn )
		{
			s.d = dir;
			s.kind = RAY_SHADOW;
			next = NEXT_TRACE;
		}
		else
			next = ( s.depth == 0 && nExtra ) ? NEXT_GEN_EXTRA : NEXT_GEN_BOUNCE;
	}
	if( next == NEXT_GEN_EXTRA || next == NEXT_GEN_BOUNCE )
	{
		s.d = lambertDir( p, s );
		s.kind = ( next == NEXT_GEN_EXTRA ) ? RAY_EXTRA : RAY_BOUNCE;
	}
	else if( next == NEXT_FINISH )
	{
		p.pathL[s.path] = make_float4( s.L.x, s.L.y, s.L.z, 0.0f );
		s.kind = LANE_IDLE;
	}
}

// warp-ballot refill with persistent work fetch: idle lanes are ranked by popc and take consecutive work indices from a
// per-warp chunk fetched with one atomicAdd on the global work counter.  Warp-collective (all 32 lanes must call it).
// FROM_QUEUE = false: work index = path index, the lane starts the path's primary ray.
// FROM_QUEUE = true : work index -> hitQueue[] -> a path whose primary ray already hit (primaryWaveKernel left the hit record
//                     in pathL[path]); the lane re-derives the camera ray and goes straight to shading that vertex.
template <bool FROM_QUEUE>
__device__ __forceinline__ void refillIdleLanes( const PtParams& p, const bool envOn, const uint32_t nExtra, const uint32_t nWork, PathState& s, uint32_t lane,
												 uint32_t& wNext, uint32_t& wEnd, bool& drained )
{
	const uint32_t idleMask = __ballot_sync( 0xFFFFFFFFu, s.kind == LANE_IDLE );
	if( !idleMask || drained ) return;
	if( wNext == wEnd )
	{
		uint32_t base = 0;
		if( lane == 0 ) base = (uint32_t)atomicAdd( p.counters, (unsigned long long)PT_CHUNK );
		base = __shfl_sync( 0xFFFFFFFFu, base, 0 );
		if( base >= nWork )
		{
			drained = true;
			return;
		}
		wNext = base;
		wEnd = min( base + PT_CHUNK, nWork );
	}
	const uint32_t rank = __popc( idleMask & ( ( 1u << lane ) - 1u ) );
	const uint32_t avail = wEnd - wNext;
	if( s.kind == LANE_IDLE && rank < avail )
	{
		if( FROM_QUEUE )
		{
			const uint32_t path = __ldg( p.hitQueue + wNext + rank );
			const float4 rec = __ldcs( p.pathL + path ); // { t, vIndex, nMajor, - } of the primary hit
			startPath( p, s, path );
			shadeVertex( p, envOn, nExtra, s, true, rec.x, __float_as_int( rec.z ), __float_as_uint( rec.y ) );
		}
		else
			startPath( p, s, wNext + rank );
	}
	wNext += min( (uint32_t)__popc( idleMask ), avail );
}

// renderPT, voxKernel.cu:610-777, as a persistent per-lane state machine (see the header of this file): every loop
// iteration traces one whole ray per lane through ONE traversal call site, then shades and sets up the lane's next ray.
// FROM_QUEUE = false: the whole path, primary ray included (single-kernel mode).
// FROM_QUEUE = true : the continuation of the paths whose primary ray hit (second wave of the wavefront mode).
template <bool FROM_QUEUE>
__global__ void __launch_bounds__( PT_BLOCK, 4 ) pathTraceKernel( const PtParams p )
{
	const uint32_t lane = threadIdx.x & 31u;
	const bool envOn = 0.0f < p.env.scale && p.env.pixels != nullptr; // HDRI::isEnabled
	const uint32_t nExtra = p.tree.hasEmission ? 1u : 0u;			  // EXTRA_IMPLICIT_SAMPLING, :720
	const uint32_t nWork = FROM_QUEUE ? (uint32_t)p.counters[2] : p.nPaths;

	uint32_t wNext = 0, wEnd = 0; // warp-uniform chunk of work indices
	bool drained = false;
	PathState s = {};
	s.kind = LANE_IDLE;
	unsigned long long rays = 0;

	for( ;; )
	{
		refillIdleLanes<FROM_QUEUE>( p, envOn, nExtra, nWork, s, lane, wNext, wEnd, drained );
		if( __all_sync( 0xFFFFFFFFu, s.kind == LANE_IDLE ) )
		{
			if( drained ) break;
			continue;
		}
		if( s.kind != LANE_IDLE )
		{
			// ---- extend: ONE traversal call site for every ray kind ----
			float t = HAKO_MAX_FLOAT;
			int nMajor = 0;
			uint32_t vIndex = 0;
			const bool hit = octreeTraverseImpl( p.tree, s.o, s.d, s.kind == RAY_SHADOW, &t, &nMajor, &vIndex );
			rays++;
			shadeVertex( p, envOn, nExtra, s, hit, t, nMajor, vIndex );
		}
	}

	// one atomic per warp for the ray counter
	for( int off = 16; off; off >>= 1 ) rays += __shfl_down_sync( 0xFFFFFFFFu, rays, off );
	if( lane == 0 && rays ) atomicAdd( p.counters + 1, rays );
}

// First wave of the wavefront mode: the primary rays of ALL paths of the step, one thread per path, plain launch.
// Primary rays are coherent (the 16 samples of a pixel sit in adjacent lanes, a block is a 4x4 pixel tile) and short; in the
// single persistent kernel they shared warps with the long incoherent shadow / bounce rays of the ~20 % of paths that hit
// something, so most lanes of most warps waited (11 of 32 lanes active on average, profiles/r1b).  Here a miss finishes the
// path at once (environment lookup, :678-683); a hit leaves { t, vIndex, nMajor } in pathL[path] and the path index in
// hitQueue -- warp-ballot compaction, one atomicAdd per warp -- for continuePaths (pathTraceKernel<true>).
__global__ void __launch_bounds__( PT_BLOCK ) primaryWaveKernel( const PtParams p, const int rows )
{
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t tilesX = ( (uint32_t)p.w + 3u ) / 4u;
	const uint32_t tx = blockIdx.x % tilesX, ty = blockIdx.x / tilesX;
	const uint32_t pix = threadIdx.x / PT_SPP; // 0..15 inside the 4x4 tile
	const uint32_t x = tx * 4u + ( pix & 3u ), yl = ty * 4u + ( pix >> 2 );
	const bool inside = x < (uint32_t)p.w && yl < (uint32_t)rows;
	const bool envOn = 0.0f < p.env.scale && p.env.pixels != nullptr;
	const uint32_t nExtra = p.tree.hasEmission ? 1u : 0u;

	bool hit = false;
	uint32_t path = 0;
	if( inside )
	{
		path = ( yl * (uint32_t)p.w + x ) * PT_SPP + ( threadIdx.x % PT_SPP );
		PathState s;
		startPath( p, s, path );
		float t = HAKO_MAX_FLOAT;
		int nMajor = 0;
		uint32_t vIndex = 0;
		hit = octreeTraverseImpl( p.tree, s.o, s.d, false, &t, &nMajor, &vIndex );
		if( hit )
			__stcs( p.pathL + path, make_float4( t, __uint_as_float( vIndex ), __int_as_float( nMajor ), 0.0f ) );
		else
			shadeVertex( p, envOn, nExtra, s, false, t, nMajor, vIndex ); // finishes the path: writes pathL[path]
	}
	// ray compaction: ballot -> rank -> one atomicAdd per warp
	const uint32_t hitMask = __ballot_sync( 0xFFFFFFFFu, hit );
	const uint32_t insideMask = __ballot_sync( 0xFFFFFFFFu, inside );
	if( hitMask )
	{
		unsigned long long base = 0;
		if( lane == 0 ) base = atomicAdd( p.counters + 2, (unsigned long long)__popc( hitMask ) );
		base = __shfl_sync( 0xFFFFFFFFu, base, 0 );
		if( hit ) p.hitQueue[base + __popc( hitMask & ( ( 1u << lane ) - 1u ) )] = path;
	}
	if( lane == 0 && insideMask ) atomicAdd( p.counters + 1, (unsigned long long)__popc( insideMask ) );
}

// ================= full wavefront mode (pt->wavefront == 2) =================
// After primaryWaveKernel the live paths are a compact list.  Per path vertex v = 0, 1, ...:
//   shadeWaveKernel( v )  one thread per live path: consumes the results of the rays of vertex v-1 (shadow / extra / bounce),
//                         finishes the path or shades the new surface point and writes ALL rays of vertex v at once -- the
//                         directions of the shadow, extra-implicit and bounce rays depend only on the sampler, not on each
//                         other's results (:697-748) -- into a dense ray list (R(v) rays per live path); survivors are
//                         compacted by warp ballot into the next vertex list;
//   traceWaveKernel( v )  persistent, traversal only: a lane whose ray ended takes the next ray from the list as soon as enough
//                         lanes of its warp are idle (ballot -> rank -> per-warp chunk from a global counter), so long rays no
//                         longer hold 31 lanes hostage.  (profiles/r1e: in the single persistent kernel a warp round lasts 233
//                         steps while the mean secondary ray takes 92; 12 of 32 lanes are active on average.)
// The arithmetic per path is the same sequence of fp32 expressions as in shadeVertex, so the accumulators are bit-identical
// across the three modes (tests/test_gpu_pt.py::test_pt_modes_identical).
struct WaveVertex // 64 B
{
	float4 a; // hitP.xyz, path index
	float4 b; // T.xyz
	float4 c; // L.xyz
	float4 d; // pending.xyz
};
struct WaveParams
{
	const WaveVertex* vin; // vertex list of v-1 (v > 0)
	WaveVertex* vout;	   // vertex list of v
	const float4* dirIn;   // ray directions of vertex v-1: [k * nIn + slot] (kind-major: a warp of the trace kernel gets rays of ONE kind)
	const float4* resIn;   // their results: { t, vIndex, nMajor, hit }
	float4* dirOut;		   // ray directions of vertex v: [k * stride + slot], stride = nIn (an upper bound of the survivors)
	const unsigned long long* nIn; // live paths at vertex v; nIn[-1] = live paths at v-1 = the stride of dirIn / resIn
	unsigned long long* nOut;
	int v;	  // index of the vertex being created (0 = the primary hit)
	int rIn;  // rays per live path at vertex v-1
	int rOut; // rays per live path at vertex v
};
// rays per vertex: [shadow] [extra, first vertex only] bounce
__host__ __device__ __forceinline__ int waveRaysPerVertex( bool envOn, uint32_t nExtra, int v ) { return ( envOn ? 1 : 0 ) + ( ( v == 0 && nExtra ) ? 1 : 0 ) + 1; }
// sampler dimension at the start of vertex v: 2 camera dimensions, then per vertex 2 (light sample) + [1 extra] + 1 bounce
__device__ __forceinline__ uint32_t waveDimAt( bool envOn, uint32_t nExtra, int v ) { return 2u + (uint32_t)v * ( envOn ? 3u : 1u ) + ( ( v > 0 && nExtra ) ? 1u : 0u ); }

__global__ void __launch_bounds__( 256 ) shadeWaveKernel( const PtParams p, const WaveParams wp )
{
	const uint32_t lane = threadIdx.x & 31u;
	const bool envOn = 0.0f < p.env.scale && p.env.pixels != nullptr;
	const uint32_t nExtra = p.tree.hasEmission ? 1u : 0u;
	const uint32_t nIn = (uint32_t)*wp.nIn;
	const size_t strideIn = wp.v > 0 ? (size_t)wp.nIn[-1] : 0, strideOut = nIn;
	// whole warps iterate together (ballot compaction below)
	for( uint32_t base = ( blockIdx.x * blockDim.x + threadIdx.x ) & ~31u; base < nIn; base += gridDim.x * blockDim.x )
	{
		const uint32_t i = base + lane;
		bool alive = false;
		PathState s;
		float t = 0.0f;
		int nMajor = 0;
		uint32_t vIndex = 0;
		if( i < nIn )
		{
			if( wp.v == 0 )
			{
				const uint32_t path = __ldg( p.hitQueue + i );
				const float4 rec = __ldcs( p.pathL + path ); // { t, vIndex, nMajor, - } of the primary hit
				startPath( p, s, path );
				t = rec.x, vIndex = __float_as_uint( rec.y ), nMajor = __float_as_int( rec.z );
				addPrimaryHit( p, s, vIndex );
				alive = true;
			}
			else
			{
				const WaveVertex vx = wp.vin[i];
				s.path = __float_as_uint( vx.a.w );
				s.o = { vx.a.x, vx.a.y, vx.a.z };
				s.T = { vx.b.x, vx.b.y, vx.b.z };
				s.L = { vx.c.x, vx.c.y, vx.c.z };
				s.pending = { vx.d.x, vx.d.y, vx.d.z };
				s.depth = wp.v - 1;
				int k = 0;
				if( envOn )
				{
					const float4 r = __ldcs( wp.resIn + (size_t)( k++ ) * strideIn + i );
					addShadowResult( s, r.w != 0.0f );
				}
				if( s.depth == 0 && nExtra )
				{
					const float4 r = __ldcs( wp.resIn + (size_t)( k++ ) * strideIn + i );
					addExtraResult( p, nExtra, s, r.w != 0.0f, __float_as_uint( r.y ) );
				}
				const float4 r = __ldcs( wp.resIn + (size_t)k * strideIn + i );
				const float4 bd = __ldcs( wp.dirIn + (size_t)k * strideIn + i );
				const bool hit = r.w != 0.0f;
				addBounceResult( p, nExtra, s, hit, __float_as_uint( r.y ) );
				s.depth++;
				if( hit && s.depth < p.maxDepth )
				{
					alive = true;
					s.d = { bd.x, bd.y, bd.z };
					t = r.x, vIndex = __float_as_uint( r.y ), nMajor = __float_as_int( r.z );
					// re-derive the sampler state of the path
					const uint32_t localPixel = s.path / PT_SPP, localSpp = s.path % PT_SPP;
					const uint32_t x = localPixel % (uint32_t)p.w, y = (uint32_t)p.y0 + localPixel / (uint32_t)p.w;
					s.spp = p.iteration * PT_SPP + localSpp;
					Murmur32 hs( 0 );
					hs.combine( y * (uint32_t)p.w + x );
					s.stream = hs.getHash();
				}
				else
					p.pathL[s.path] = make_float4( s.L.x, s.L.y, s.L.z, 0.0f );
			}
		}
		// survivors: ballot -> rank -> one atomicAdd per warp
		const uint32_t aliveMask = __ballot_sync( 0xFFFFFFFFu, alive );
		if( !aliveMask ) continue;
		unsigned long long slot0 = 0;
		if( lane == 0 ) slot0 = atomicAdd( wp.nOut, (unsigned long long)__popc( aliveMask ) );
		slot0 = __shfl_sync( 0xFFFFFFFFu, slot0, 0 );
		if( alive )
		{
		const size_t slot = (size_t)slot0 + __popc( aliveMask & ( ( 1u << lane ) - 1u ) );
		s.dim = waveDimAt( envOn, nExtra, wp.v );
		s.depth = wp.v;
		V3 dir = {};
		shadeSurface( p, envOn, s, t, nMajor, vIndex, &dir );
		float4* dout = wp.dirOut + slot;
		if( envOn )
		{
			*dout = make_float4( dir.x, dir.y, dir.z, 0.0f );
			dout += strideOut;
		}
		if( wp.v == 0 && nExtra )
		{
			const V3 e = lambertDir( p, s );
			*dout = make_float4( e.x, e.y, e.z, 0.0f );
			dout += strideOut;
		}
		const V3 bnc = lambertDir( p, s );
		*dout = make_float4( bnc.x, bnc.y, bnc.z, 0.0f );
		WaveVertex o;
		o.a = make_float4( s.o.x, s.o.y, s.o.z, __uint_as_float( s.path ) );
		o.b = make_float4( s.T.x, s.T.y, s.T.z, 0.0f );
		o.c = make_float4( s.L.x, s.L.y, s.L.z, 0.0f );
		o.d = make_float4( s.pending.x, s.pending.y, s.pending.z, 0.0f );
		wp.vout[slot] = o;
		}
	}
}

struct TraceWaveParams
{
	const WaveVertex* vertices;
	const float4* dirs;
	float4* results;
	const unsigned long long* nVertices; // nVertices[-1] = live paths one vertex earlier = pitch of dirs / results
	unsigned long long* work; // ray index counter of this wave (zeroed)
	unsigned long long* rays; // total rays traced
	unsigned long long* steps; // total traversal steps (node visits) of the rays traced by this kernel
	int raysPerVertex;
	int shadowFirst; // ray 0 of every vertex is a shadow ray (any-hit semantics: no voxel-rank accumulation)
	int refillMin;	 // refill when at least this many lanes of the warp are idle
	int burst;		 // traversal steps between two refill checks
	const uint32_t* order; // binned wave: position in the work queue -> ray index (nullptr: identity)
};

// ---- ray binning: a counting sort of the wave's ray indices --------------------------------------------------------------
// The results of a ray are written to the slot its (kind, vertex) pair owns, whoever traces it, so the order in which the trace
// kernel draws rays changes which lanes share a warp and nothing else: the accumulators stay bit-identical.
#define BIN_SCAN_TILE 4096 // bins per CTA of binScanKernel (1024 threads x 4)
struct BinParams
{
	const WaveVertex* vertices;
	const float4* dirs;
	const unsigned long long* nVertices; // nVertices[-1] = pitch of dirs
	int raysPerVertex;
	int cellBits;  // per axis
	int cellMajor; // key = kind | cell | octant instead of kind | octant | cell
	V3 lower;
	V3 scale; // cells per world unit
	uint32_t* keys;	 // [nRays]
	uint32_t* hist;	 // [bins]: counts, then (after binScanKernel) the running cursors
	uint32_t* order; // [nRays]
	uint32_t bins;
	ScanState scan;
};
__device__ __forceinline__ uint32_t binSpread3( uint32_t v ) // 6 bits -> every third bit
{
	v = ( v | ( v << 8 ) ) & 0x0000F00Fu;
	v = ( v | ( v << 4 ) ) & 0x000C30C3u;
	v = ( v | ( v << 2 ) ) & 0x00249249u;
	return v;
}
__global__ void __launch_bounds__( 256 ) binCountKernel( const BinParams bp )
{
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t nVertices = (uint32_t)*bp.nVertices;
	const uint32_t stride = (uint32_t)bp.nVertices[-1];
	const uint32_t nRays = nVertices * (uint32_t)bp.raysPerVertex;
	const int cmax = ( 1 << bp.cellBits ) - 1;
	for( uint32_t base = ( blockIdx.x * blockDim.x + threadIdx.x ) & ~31u; base < nRays; base += gridDim.x * blockDim.x )
	{
		const uint32_t r = base + lane;
		const bool valid = r < nRays;
		uint32_t key = 0xFFFFFFFFu;
		if( valid )
		{
			const uint32_t k = r / nVertices, slot = r - k * nVertices;
			const float4 o = __ldg( &bp.vertices[slot].a );
			const float4 d = __ldg( bp.dirs + (size_t)k * stride + slot );
			const int cx = min( max( (int)( ( o.x - bp.lower.x ) * bp.scale.x ), 0 ), cmax );
			const int cy = min( max( (int)( ( o.y - bp.lower.y ) * bp.scale.y ), 0 ), cmax );
			const int cz = min( max( (int)( ( o.z - bp.lower.z ) * bp.scale.z ), 0 ), cmax );
			const uint32_t cell = binSpread3( (uint32_t)cx ) | ( binSpread3( (uint32_t)cy ) << 1 ) | ( binSpread3( (uint32_t)cz ) << 2 );
			const uint32_t oct = ( d.x < 0.0f ? 1u : 0u ) | ( d.y < 0.0f ? 2u : 0u ) | ( d.z < 0.0f ? 4u : 0u );
			const int cb3 = 3 * bp.cellBits;
			key = bp.cellMajor ? ( ( ( k << cb3 ) | cell ) << 3 ) | oct : ( ( ( k << 3 ) | oct ) << cb3 ) | cell;
			bp.keys[r] = key;
		}
		const uint32_t peers = __match_any_sync( 0xFFFFFFFFu, key );
		if( valid && lane == (uint32_t)( __ffs( peers ) - 1 ) ) atomicAdd( bp.hist + key, (uint32_t)__popc( peers ) );
	}
}
// exclusive scan of the bin counters in place (decoupled look-back over 4 096-bin tiles)
__global__ void __launch_bounds__( 1024 ) binScanKernel( const BinParams bp )
{
	const uint32_t tile = scanTakeTicket( bp.scan );
	const uint32_t i0 = tile * BIN_SCAN_TILE + threadIdx.x * 4u;
	uint32_t c[4];
#pragma unroll
	for( int j = 0; j < 4; j++ ) c[j] = ( i0 + j < bp.bins ) ? bp.hist[i0 + j] : 0u;
	uint32_t total = 0;
	const uint32_t excl = blockExclusiveScan<1024>( c[0] + c[1] + c[2] + c[3], &total );
	const uint32_t prefix = (uint32_t)scanTilePrefix( bp.scan, tile, (uint64_t)total );
	uint32_t run = prefix + excl;
#pragma unroll
	for( int j = 0; j < 4; j++ )
	{
		if( i0 + j < bp.bins ) bp.hist[i0 + j] = run;
		run += c[j];
	}
}
__global__ void __launch_bounds__( 256 ) binScatterKernel( const BinParams bp )
{
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t nRays = (uint32_t)*bp.nVertices * (uint32_t)bp.raysPerVertex;
	for( uint32_t base = ( blockIdx.x * blockDim.x + threadIdx.x ) & ~31u; base < nRays; base += gridDim.x * blockDim.x )
	{
		const uint32_t r = base + lane;
		const bool valid = r < nRays;
		const uint32_t key = valid ? __ldcs( bp.keys + r ) : 0xFFFFFFFFu;
		const uint32_t peers = __match_any_sync( 0xFFFFFFFFu, key );
		const int leader = __ffs( peers ) - 1;
		uint32_t pos = 0;
		if( valid && (int)lane == leader ) pos = atomicAdd( bp.hist + key, (uint32_t)__popc( peers ) );
		pos = __shfl_sync( 0xFFFFFFFFu, pos, leader );
		if( valid ) bp.order[pos + __popc( peers & ( ( 1u << lane ) - 1u ) )] = r;
	}
}
template <int MIN_CTAS, int TOPK = 0>
__global__ void __launch_bounds__( 256, MIN_CTAS ) traceWaveKernel( const OctreeView tree, const TraceWaveParams tp )
{
	// TOPK > 0 (HAKO_SMEM_TOP, A/B variant): the persistent CTA keeps the top TOPK nodes of the tree in shared memory
	__shared__ uint2 s_top[TOPK > 0 ? TOPK * 8 : 1];
	TopCache tc{ nullptr, 0xFFFFFFFFu };
	if( TOPK > 0 ) tc = stageTop<( TOPK > 0 ? TOPK : 1 )>( tree, s_top );
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t nVertices = (uint32_t)*tp.nVertices;
	const uint32_t stride = (uint32_t)tp.nVertices[-1]; // live paths one vertex earlier: the pitch of the kind-major ray arrays
	const uint32_t nRays = nVertices * (uint32_t)tp.raysPerVertex;
	uint32_t wNext = 0, wEnd = 0;
	bool drained = false;
	Traversal tr;
	TravStack stk;
	bool flying = false, shadow = false;
	uint32_t addr = 0; // k * stride + slot of the ray in flight
	uint32_t traced = 0;
	unsigned long long steps = 0;
	for( ;; )
	{
		const uint32_t idleMask = __ballot_sync( 0xFFFFFFFFu, !flying );
		if( idleMask == 0xFFFFFFFFu || ( !drained && __popc( idleMask ) >= tp.refillMin ) )
		{
			if( !drained && wNext == wEnd )
			{
				uint32_t base = 0;
				if( lane == 0 ) base = (uint32_t)atomicAdd( tp.work, (unsigned long long)PT_CHUNK );
				base = __shfl_sync( 0xFFFFFFFFu, base, 0 );
				if( base >= nRays )
					drained = true;
				else
				{
					wNext = base;
					wEnd = min( base + PT_CHUNK, nRays );
				}
			}
			if( drained )
			{
				if( idleMask == 0xFFFFFFFFu ) break;
			}
			else
			{
				const uint32_t rank = __popc( idleMask & ( ( 1u << lane ) - 1u ) );
				const uint32_t avail = wEnd - wNext;
				if( !flying && rank < avail )
				{
					const uint32_t ray = tp.order ? __ldg( tp.order + wNext + rank ) : wNext + rank;
					const uint32_t k = ray / nVertices, slot = ray - k * nVertices;
					addr = k * stride + slot;
					const float4 o = __ldg( &tp.vertices[slot].a );
					const float4 d = __ldcs( tp.dirs + addr );
					shadow = tp.shadowFirst && k == 0;
					flying = tree.wide ? tr.template begin<true>( tree, V3{ o.x, o.y, o.z }, V3{ d.x, d.y, d.z } )
									   : tr.template begin<false>( tree, V3{ o.x, o.y, o.z }, V3{ d.x, d.y, d.z } );
					if( !flying ) __stcs( tp.results + addr, make_float4( 0.0f, 0.0f, 0.0f, 0.0f ) ); // misses the root box
					traced++;
				}
				wNext += min( (uint32_t)__popc( idleMask ), avail );
			}
		}
		if( flying )
		{
			int it = 0;
#pragma unroll 1
			for( ; it < tp.burst; it++ )
			{
				const int r = tree.wide ? ( tr.ordinary ? tr.template step<true, true, ( TOPK > 0 )>( tree, shadow, stk, tc ) : tr.template step<false, true, ( TOPK > 0 )>( tree, shadow, stk, tc ) )
										: ( tr.ordinary ? tr.template step<true, false, ( TOPK > 0 )>( tree, shadow, stk, tc ) : tr.template step<false, false, ( TOPK > 0 )>( tree, shadow, stk, tc ) );
				if( r == TRAV_CONTINUE ) continue;
				flying = false;
				if( r == TRAV_HIT )
					__stcs( tp.results + addr, make_float4( tr.t, __uint_as_float( tr.skipped ), __int_as_float( tr.nMajor ), 1.0f ) );
				else
					__stcs( tp.results + addr, make_float4( 0.0f, 0.0f, 0.0f, 0.0f ) );
				it++;
				break;
			}
			steps += (uint32_t)it; // one add per burst, not per step
		}
	}
	for( int off = 16; off; off >>= 1 ) traced += __shfl_down_sync( 0xFFFFFFFFu, traced, off );
	if( lane == 0 && traced ) atomicAdd( tp.rays, (unsigned long long)traced );
	for( int off = 16; off; off >>= 1 ) steps += __shfl_down_sync( 0xFFFFFFFFu, steps, off );
	if( lane == 0 && steps ) atomicAdd( tp.steps, steps );
}

// traceWaveKernel with the traversal stack in shared memory (TravStackShared, traverse.cuh): `depth` entries per thread, dynamic shared
// memory = depth x 24 B (22 B PACKED) x 256 threads per CTA, as many CTAs per SM as that leaves room for.
template <int MIN_CTAS, bool PACKED>
__global__ void __launch_bounds__( 256, MIN_CTAS ) traceWaveSmemKernel( const OctreeView tree, const TraceWaveParams tp, const int depth )
{
	extern __shared__ __align__( 16 ) unsigned char s_stack[];
	constexpr int TOPK = 0;
	// TOPK > 0 (HAKO_SMEM_TOP, A/B variant): the persistent CTA keeps the top TOPK nodes of the tree in shared memory
	__shared__ uint2 s_top[TOPK > 0 ? TOPK * 8 : 1];
	TopCache tc{ nullptr, 0xFFFFFFFFu };
	if( TOPK > 0 ) tc = stageTop<( TOPK > 0 ? TOPK : 1 )>( tree, s_top );
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t nVertices = (uint32_t)*tp.nVertices;
	const uint32_t stride = (uint32_t)tp.nVertices[-1]; // live paths one vertex earlier: the pitch of the kind-major ray arrays
	const uint32_t nRays = nVertices * (uint32_t)tp.raysPerVertex;
	uint32_t wNext = 0, wEnd = 0;
	bool drained = false;
	Traversal tr;
	TravStackShared<PACKED> stk;
	stk.bind( s_stack, depth, 256, threadIdx.x );
	bool flying = false, shadow = false;
	uint32_t addr = 0; // k * stride + slot of the ray in flight
	uint32_t traced = 0;
	unsigned long long steps = 0;
	for( ;; )
	{
		const uint32_t idleMask = __ballot_sync( 0xFFFFFFFFu, !flying );
		if( idleMask == 0xFFFFFFFFu || ( !drained && __popc( idleMask ) >= tp.refillMin ) )
		{
			if( !drained && wNext == wEnd )
			{
				uint32_t base = 0;
				if( lane == 0 ) base = (uint32_t)atomicAdd( tp.work, (unsigned long long)PT_CHUNK );
				base = __shfl_sync( 0xFFFFFFFFu, base, 0 );
				if( base >= nRays )
					drained = true;
				else
				{
					wNext = base;
					wEnd = min( base + PT_CHUNK, nRays );
				}
			}
			if( drained )
			{
				if( idleMask == 0xFFFFFFFFu ) break;
			}
			else
			{
				const uint32_t rank = __popc( idleMask & ( ( 1u << lane ) - 1u ) );
				const uint32_t avail = wEnd - wNext;
				if( !flying && rank < avail )
				{
					const uint32_t ray = tp.order ? __ldg( tp.order + wNext + rank ) : wNext + rank;
					const uint32_t k = ray / nVertices, slot = ray - k * nVertices;
					addr = k * stride + slot;
					const float4 o = __ldg( &tp.vertices[slot].a );
					const float4 d = __ldcs( tp.dirs + addr );
					shadow = tp.shadowFirst && k == 0;
					flying = tree.wide ? tr.template begin<true>( tree, V3{ o.x, o.y, o.z }, V3{ d.x, d.y, d.z } )
									   : tr.template begin<false>( tree, V3{ o.x, o.y, o.z }, V3{ d.x, d.y, d.z } );
					if( !flying ) __stcs( tp.results + addr, make_float4( 0.0f, 0.0f, 0.0f, 0.0f ) ); // misses the root box
					traced++;
				}
				wNext += min( (uint32_t)__popc( idleMask ), avail );
			}
		}
		if( flying )
		{
			int it = 0;
#pragma unroll 1
			for( ; it < tp.burst; it++ )
			{
				const int r = tree.wide ? ( tr.ordinary ? tr.template step<true, true, ( TOPK > 0 ), TravStackShared<PACKED>>( tree, shadow, stk, tc ) : tr.template step<false, true, ( TOPK > 0 ), TravStackShared<PACKED>>( tree, shadow, stk, tc ) )
										: ( tr.ordinary ? tr.template step<true, false, ( TOPK > 0 ), TravStackShared<PACKED>>( tree, shadow, stk, tc ) : tr.template step<false, false, ( TOPK > 0 ), TravStackShared<PACKED>>( tree, shadow, stk, tc ) );
				if( r == TRAV_CONTINUE ) continue;
				flying = false;
				if( r == TRAV_HIT )
					__stcs( tp.results + addr, make_float4( tr.t, __uint_as_float( tr.skipped ), __int_as_float( tr.nMajor ), 1.0f ) );
				else
					__stcs( tp.results + addr, make_float4( 0.0f, 0.0f, 0.0f, 0.0f ) );
				it++;
				break;
			}
			steps += (uint32_t)it; // one add per burst, not per step
		}
	}
	for( int off = 16; off; off >>= 1 ) traced += __shfl_down_sync( 0xFFFFFFFFu, traced, off );
	if( lane == 0 && traced ) atomicAdd( tp.rays, (unsigned long long)traced );
	for( int off = 16; off; off >>= 1 ) steps += __shfl_down_sync( 0xFFFFFFFFu, steps, off );
	if( lane == 0 && steps ) atomicAdd( tp.steps, steps );
}

// per-pixel reduction in localSpp order, then `frameBuffer += `, voxKernel.cu:763-774
__global__ void __launch_bounds__( 256 ) accumulateKernel( const float4* __restrict__ pathL, float4* __restrict__ accum, uint32_t nPixels )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if( i >= nPixels ) return;
	float x = 0.0f, y = 0.0f, z = 0.0f;
#pragma unroll
	for( int k = 0; k < PT_SPP; k++ )
	{
		const float4 l = __ldcs( pathL + (size_t)i * PT_SPP + k );
		x += l.x, y += l.y, z += l.z;
	}
	float4 a = accum[i];
	a.x += x, a.y += y, a.z += z, a.w += (float)PT_SPP;
	accum[i] = a;
}

// renderResolve, voxKernel.cu:779-795
__global__ void __launch_bounds__( 256 ) resolveKernel( uchar4* __restrict__ out, const float4* __restrict__ accum, uint32_t n )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if( i >= n ) return;
	const float4 v = accum[i];
	const int r = f2i_x86( 255 * powf( v.x / v.w, 1.0f / 2.2f ) + 0.5f );
	const int g = f2i_x86( 255 * powf( v.y / v.w, 1.0f / 2.2f ) + 0.5f );
	const int b = f2i_x86( 255 * powf( v.z / v.w, 1.0f / 2.2f ) + 0.5f );
	out[i] = make_uchar4( (uint8_t)min( r, 255 ), (uint8_t)min( g, 255 ), (uint8_t)min( b, 255 ), 255 );
}
} // namespace

// ---- HDRI::load, renderCommon.hpp:214-314: 1 + 6 summed-area tables ----
namespace
{
#define SAT_BLOCK 512 // SAT_BLOCK_SIZE, voxKernel.cu:555

// prefixSumInclusive<double, 512> (voxKernel.cu:526-553): Hillis-Steele over one 512-element tile, double-buffered in
// shared memory.  The summation ORDER is the reference's (and the oracle's), so the tables are bit-identical.
__device__ __forceinline__ double scanTile( double prefix, double value, double ( *buf )[SAT_BLOCK], double* sumOut )
{
	int cur = 0;
	buf[0][threadIdx.x] = value;
	__syncthreads();
	for( uint32_t offset = 1; offset < SAT_BLOCK; offset <<= 1 )
	{
		double x = buf[cur][threadIdx.x];
		if( offset <= threadIdx.x ) x += buf[cur][threadIdx.x - offset];
		buf[cur ^ 1][threadIdx.x] = x;
		cur ^= 1;
		__syncthreads();
	}
	*sumOut = buf[cur][SAT_BLOCK - 1];
	const double r = buf[cur][threadIdx.x] + prefix;
	__syncthreads();
	return r;
}

// HDRIstoreImportance (voxKernel.cu:485-524) fused into buildSATh (:557-577): one block per row.
// rowTab: { sr, sY = mix(cos(theta), cos(theta+dTheta), 0.5), sinTheta(sY), - };  colTab: { cos(phi), sin(phi) } of the
// texel centre -- host-computed so the importance is the same fp32 value the CPU code produces.
__global__ void __launch_bounds__( SAT_BLOCK ) satRowKernel( const float4* __restrict__ pixels, int w, int h, const float4* __restrict__ rowImp,
															 const float2* __restrict__ colTab, int cosWeighted, V3 axis, double* __restrict__ sat )
{
	__shared__ double buf[2][SAT_BLOCK];
	const int Y = blockIdx.x;
	const float4 row = rowImp[Y];
	double prefix = 0.0;
	for( int i = 0; i < w; i += SAT_BLOCK )
	{
		const int X = i + (int)threadIdx.x;
		double v = 0.0;
		if( X < w )
		{
			const float4 c = pixels[(size_t)Y * w + X];
			float wgt = 1.0f;
			if( cosWeighted )
			{
				const float2 cs = colTab[X];
				const V3 dirCenter = { cs.x * row.z, row.y, cs.y * row.z };
				wgt = smax( dot( axis, dirCenter ), 0.0f );
			}
			v = (double)( ( 0.2126f * c.x + 0.7152f * c.y + 0.0722f * c.z ) * row.x * wgt );
		}
		double sum;
		const double r = scanTile( prefix, v, buf, &sum );
		prefix += sum;
		if( X < w ) sat[(size_t)Y * w + X] = r;
	}
}
// buildSATv (:578-598): one block per column
__global__ void __launch_bounds__( SAT_BLOCK ) satColKernel( int w, int h, double* __restrict__ sat )
{
	__shared__ double buf[2][SAT_BLOCK];
	const int X = blockIdx.x;
	double prefix = 0.0;
	for( int i = 0; i < h; i += SAT_BLOCK )
	{
		const int Y = i + (int)threadIdx.x;
		const double v = Y < h ? sat[(size_t)Y * w + X] : 0.0;
		double sum;
		const double r = scanTile( prefix, v, buf, &sum );
		prefix += sum;
		if( Y < h ) sat[(size_t)Y * w + X] = r;
	}
}
// buildSAT2u32 (:600-608)
__global__ void __launch_bounds__( 256 ) satToU32Kernel( uint32_t* __restrict__ out, const double* __restrict__ sat, uint32_t n )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	const double sum = sat[n - 1];
	if( i < n )
	{
		const double q = sat[i] / ( sum ) * (double)0xFFFFFFFFu;
		// (uint32_t)double on x86-64 is cvttsd2si to 64 bits + truncation: NaN / out of range -> low word of INT64_MIN = 0
		out[i] = ( q >= 0.0 && q < 9223372036854775808.0 ) ? (uint32_t)(unsigned long long)q : 0u;
	}
}
} // namespace

// ---- host side: PMJ02 table (PMJSampler::setup, pmjSampler.hpp:114-134; GetPMJ02Samples :14-58) ----
namespace
{
struct HostPCG32 // renderCommon.hpp:86-111
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
		const uint64_t old = state;
		state = old * 6364136223846793005ULL + inc;
		const uint32_t xorshifted = (uint32_t)( ( ( old >> 18u ) ^ old ) >> 27u );
		const uint32_t rot = (uint32_t)( old >> 59u );
		return ( xorshifted >> rot ) | ( xorshifted << ( ( 0u - rot ) & 31u ) );
	}
	float uniformf() // renderCommon.hpp:112-117
	{
		const uint32_t bits = ( nextU32() >> 9 ) | 0x3f800000u;
		float v;
		memcpy( &v, &bits, 4 );
		return v - 1.0f;
	}
};
// one (0,2)-sequence: each new sample goes to the unoccupied half of the sub-stratum of an earlier sample, chosen by
// the xor tables of Helmer's stochastic generation (the table values are the published constants of that method)
void generatePMJ02( int nSamples, float* samples, HostPCG32& rng )
{
	static const uint32_t xors[2][32] = {
		{ 0x0, 0x0, 0x2, 0x6, 0x6, 0xe, 0x36, 0x4e, 0x16, 0x2e, 0x276, 0x6ce, 0x716, 0xc2e, 0x3076, 0x40ce, 0x116, 0x22e, 0x20676, 0x60ece, 0x61716, 0xe2c2e, 0x367076, 0x4ec0ce, 0x170116, 0x2c022e, 0x2700676, 0x6c00ece, 0x7001716, 0xc002c2e, 0x30007076, 0x4000c0ce },
		{ 0x0, 0x1, 0x3, 0x3, 0x7, 0x1b, 0x27, 0xb, 0x17, 0x13b, 0x367, 0x38b, 0x617, 0x183b, 0x2067, 0x8b, 0x117, 0x1033b, 0x30767, 0x30b8b, 0x71617, 0x1b383b, 0x276067, 0xb808b, 0x160117, 0x138033b, 0x3600767, 0x3800b8b, 0x6001617, 0x1800383b, 0x20006067, 0x808b } };
	samples[0] = rng.uniformf();
	samples[1] = rng.uniformf();
	for( int logN = 0; ( 1 << logN ) < nSamples; logN++ )
	{
		const int prevLen = 1 << logN;
		const int nStrata = prevLen * 2;
		const float invStrata = 1.0f / nStrata;
		for( int i = 0; i < prevLen && ( prevLen + i ) < nSamples; i++ )
		{
			const int xs = (int)( samples[( i ^ (int)xors[0][logN] ) * 2] * nStrata ) ^ 1;
			const int ys = (int)( samples[( i ^ (int)xors[1][logN] ) * 2 + 1] * nStrata ) ^ 1;
			// the reference passes two uniformFloat() calls as function arguments; with the ABI it is built for here the
			// second one is evaluated first, which is what the oracle is pinned to
			const float xi1 = rng.uniformf();
			const float xi0 = rng.uniformf();
			samples[( prevLen + i ) * 2 + 0] = ( xi0 + xs ) * invStrata;
			samples[( prevLen + i ) * 2 + 1] = ( xi1 + ys ) * invStrata;
		}
	}
}

int ptFreeHdri( hako_pt* pt, cudaStream_t s )
{
	dfree( pt->hdriPixels, s );
	dfree( pt->hdriPrimary, s );
	dfree( pt->sats, s );
	dfree( pt->rowTab, s );
	pt->hdriPixels = pt->hdriPrimary = nullptr;
	pt->sats = nullptr;
	pt->rowTab = nullptr;
	pt->hw = pt->hh = pt->wP = pt->hP = 0;
	return 0;
}
} // namespace

#define PT_ARG( cond, msg )              \
	do                                   \
	{                                    \
		if( !( cond ) )                  \
		{                                \
			hako::setError( "%s", msg ); \
			return HAKO_E_INVALID;       \
		}                                \
	} while( 0 )

extern "C" {

int hako_pt_create( hako_ctx* ctx, hako_stream stream, hako_pt** out )
{
	PT_ARG( ctx && out, "hako_pt_create: null argument" );
	*out = nullptr;
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	hako_pt* pt = new hako_pt();
	pt->ctx = ctx;
	if( const char* e = getenv( "HAKO_PT_WAVEFRONT" ) ) pt->wavefront = atoi( e );
	if( const char* e = getenv( "HAKO_PT_REFILL_MIN" ) ) pt->waveRefillMin = pt->waveRefillMinSmem = atoi( e );
	if( const char* e = getenv( "HAKO_PT_BURST" ) ) pt->waveBurst = atoi( e );
	if( const char* e = getenv( "HAKO_PT_TRACE_CTAS" ) ) pt->waveTraceCtas = atoi( e );
	if( const char* e = getenv( "HAKO_PT_SMEM_STACK" ) ) pt->waveSmemStack = atoi( e );
	if( const char* e = getenv( "HAKO_PT_BAND_PATHS" ) ) pt->waveBandPaths = strtoull( e, nullptr, 10 );
	if( const char* e = getenv( "HAKO_PT_LANES" ) ) pt->lanes = atoi( e );
	if( const char* e = getenv( "HAKO_PT_BIN" ) ) pt->binBits = std::max( 0, std::min( 6, atoi( e ) ) );
	if( const char* e = getenv( "HAKO_PT_BIN_CELL_MAJOR" ) ) pt->binCellMajor = atoi( e ) != 0;
	pt->lanes = std::max( 1, std::min( PT_LANES, pt->lanes ) );
	std::vector<float> table( (size_t)2 * PMJ_LENGTH * PMJ_NSEQ );
	HostPCG32 rng;
	rng.setup( 0, 2525 ); // pmjSampler.hpp:116-117
	for( int i = 0; i < PMJ_NSEQ; i++ ) generatePMJ02( PMJ_LENGTH, table.data() + (size_t)2 * PMJ_LENGTH * i, rng );
	int r = dalloc( ctx, &pt->pmj, table.size(), s );
	if( r == 0 ) r = dalloc( ctx, &pt->counters, PT_COUNTERS * PT_LANES, s );
	if( r == 0 )
	{
		cudaError_t e = cudaMemcpyAsync( pt->pmj, table.data(), table.size() * sizeof( float ), cudaMemcpyHostToDevice, s );
		if( e == cudaSuccess ) e = cudaMemsetAsync( pt->counters, 0, PT_COUNTERS * PT_LANES * sizeof( unsigned long long ), s );
		for( int i = 0; i < PT_LANES && e == cudaSuccess; i++ ) e = cudaStreamCreateWithFlags( &pt->laneStream[i], cudaStreamNonBlocking );
		for( int i = 0; i <= PT_LANES && e == cudaSuccess; i++ ) e = cudaEventCreateWithFlags( &pt->laneEv[i], cudaEventDisableTiming );
		if( e == cudaSuccess ) e = cudaStreamSynchronize( s ); // `table` dies with this scope
		if( e != cudaSuccess ) r = cudaFail( e, "PMJ upload", __FILE__, __LINE__ );
	}
	if( r )
	{
		hako_pt_destroy( pt );
		return r;
	}
	*out = pt;
	return HAKO_OK;
}

void hako_pt_destroy( hako_pt* pt )
{
	if( !pt ) return;
	cudaSetDevice( pt->ctx->device );
	cudaStream_t s = nullptr;
	ptFreeHdri( pt, s );
	dfree( pt->accum, s );
	dfree( pt->rgba8, s );
	dfree( pt->pmj, s );
	dfree( pt->pathL, s );
	dfree( pt->hitQueue, s );
	for( int k = 0; k < PT_LANES; k++ )
	{
		if( pt->laneStream[k] ) cudaStreamSynchronize( pt->laneStream[k] );
		for( int i = 0; i < 2; i++ ) dfree( pt->waveV[k][i], s ), dfree( pt->waveDir[k][i], s );
		dfree( pt->waveRes[k], s );
		dfree( pt->binKeys[k], s ), dfree( pt->binOrder[k], s ), dfree( pt->binHist[k], s );
		if( pt->laneStream[k] ) cudaStreamDestroy( pt->laneStream[k] );
	}
	for( int k = 0; k <= PT_LANES; k++ )
		if( pt->laneEv[k] ) cudaEventDestroy( pt->laneEv[k] );
	dfree( pt->counters, s );
	cudaStreamSynchronize( s );
	delete pt;
}

int hako_pt_clear( hako_pt* pt, hako_stream stream )
{
	PT_ARG( pt, "hako_pt_clear: null path tracer" );
	if( !pt->accum )
	{
		hako::setError( "hako_pt_clear: no frame buffer (call hako_pt_resize first)" );
		return HAKO_E_STATE;
	}
	HAKO_CUDA( cudaSetDevice( pt->ctx->device ) );
	HAKO_CUDA( cudaMemsetAsync( pt->accum, 0, (size_t)pt->w * pt->h * sizeof( float4 ), (cudaStream_t)stream ) );
	return HAKO_OK;
}

int hako_pt_resize( hako_pt* pt, hako_stream stream, int w, int h )
{
	PT_ARG( pt && w > 0 && h > 0, "hako_pt_resize: bad argument" );
	if( (uint64_t)w * h * PT_SPP > 0xFFFFFFFFull )
	{
		hako::setError( "hako_pt_resize: %dx%d x 16 spp exceeds the 32-bit path index", w, h );
		return HAKO_E_LIMIT;
	}
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( pt->ctx->device ) );
	if( pt->accum && pt->w == w && pt->h == h ) return HAKO_OK;
	dfree( pt->accum, s );
	dfree( pt->rgba8, s );
	pt->accum = nullptr;
	pt->rgba8 = nullptr;
	HAKO_TRY( dalloc( pt->ctx, &pt->accum, (uint64_t)w * h, s ) );
	HAKO_TRY( dalloc( pt->ctx, &pt->rgba8, (uint64_t)w * h, s ) );
	pt->w = w;
	pt->h = h;
	return hako_pt_clear( pt, stream ); // PathTracer.hpp:84-88
}

int hako_pt_load_hdri( hako_pt* pt, hako_stream stream, const float* rgba, int w, int h, const float* rgbaPrimary, int wP, int hP )
{
	PT_ARG( pt && rgba && w > 0 && h > 0, "hako_pt_load_hdri: bad argument" );
	PT_ARG( !rgbaPrimary || ( wP > 0 && hP > 0 ), "hako_pt_load_hdri: bad primary image size" );
	hako_ctx* ctx = pt->ctx;
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	ptFreeHdri( pt, s );
	const size_t n = (size_t)w * h;

	// theta-only and phi-only terms, with the host's libm exactly as the CPU code evaluates them
	// (voxKernel.cu:495-518 for the tables, renderCommon.hpp:414-424 for sampling)
	std::vector<float4> rowTab( h ), rowImp( h );
	std::vector<float2> colTab( w );
	const float dTheta = HAKO_PI / (float)h;
	const float dPhi = 2.0f * HAKO_PI / (float)w;
	for( uint32_t Y = 0; Y < (uint32_t)h; Y++ )
	{
		const float theta = Y * dTheta;
		const float dH = 2.0f * sinf( dTheta * 0.5f ) * sinf( dTheta * 0.5f + theta );
		const float sr = dH * dPhi;
		const float c0 = cosf( theta ), c1 = cosf( theta + dTheta );
		const float sY = mixf( c0, c1, 0.5f );
		rowTab[Y] = make_float4( sr, c0, c1, 0.0f );
		rowImp[Y] = make_float4( sr, sY, sqrtf( smax( 1.0f - sY * sY, 0.0f ) ), 0.0f );
	}
	for( int X = 0; X < w; X++ )
	{
		const float phi = dPhi * ( (float)X + 0.5f ) + HAKO_PI;
		colTab[X] = make_float2( cosf( phi ), sinf( phi ) );
	}

	float4* rowImp_d = nullptr;
	float2* colTab_d = nullptr;
	double* satF64 = nullptr;
	int r = 0;
	do
	{
		if( ( r = dalloc( ctx, &pt->hdriPixels, n, s ) ) ) break;
		if( ( r = dalloc( ctx, &pt->sats, 7 * n, s ) ) ) break;
		if( ( r = dalloc( ctx, &pt->rowTab, (uint64_t)h, s ) ) ) break;
		if( ( r = dalloc( ctx, &rowImp_d, (uint64_t)h, s ) ) ) break;
		if( ( r = dalloc( ctx, &colTab_d, (uint64_t)w, s ) ) ) break;
		if( ( r = dalloc( ctx, &satF64, n, s ) ) ) break;
		cudaError_t e = cudaMemcpyAsync( pt->hdriPixels, rgba, n * sizeof( float4 ), cudaMemcpyHostToDevice, s );
		if( e == cudaSuccess ) e = cudaMemcpyAsync( pt->rowTab, rowTab.data(), h * sizeof( float4 ), cudaMemcpyHostToDevice, s );
		if( e == cudaSuccess ) e = cudaMemcpyAsync( rowImp_d, rowImp.data(), h * sizeof( float4 ), cudaMemcpyHostToDevice, s );
		if( e == cudaSuccess ) e = cudaMemcpyAsync( colTab_d, colTab.data(), w * sizeof( float2 ), cudaMemcpyHostToDevice, s );
		if( e == cudaSuccess && rgbaPrimary )
		{
			if( ( r = dalloc( ctx, &pt->hdriPrimary, (uint64_t)wP * hP, s ) ) ) break;
			e = cudaMemcpyAsync( pt->hdriPrimary, rgbaPrimary, (size_t)wP * hP * sizeof( float4 ), cudaMemcpyHostToDevice, s ); // loadPrimary, :315-325
		}
		if( e != cudaSuccess )
		{
			r = cudaFail( e, "H2D environment", __FILE__, __LINE__ );
			break;
		}
		const V3 axes[7] = { { 0, 0, 0 }, { 1, 0, 0 }, { -1, 0, 0 }, { 0, 1, 0 }, { 0, -1, 0 }, { 0, 0, 1 }, { 0, 0, -1 } }; // :272-279
		r = [&]() -> int {
			for( int k = 0; k < 7; k++ )
			{
				HAKO_LAUNCH( ctx, satRowKernel, h, SAT_BLOCK, 0, s, pt->hdriPixels, w, h, rowImp_d, colTab_d, k == 0 ? 0 : 1, axes[k], satF64 );
				HAKO_LAUNCH( ctx, satColKernel, w, SAT_BLOCK, 0, s, w, h, satF64 );
				HAKO_LAUNCH( ctx, satToU32Kernel, divUp( n, 256 ), 256, 0, s, pt->sats + n * k, satF64, (uint32_t)n );
			}
			return 0;
		}();
	} while( 0 );
	dfree( rowImp_d, s );
	dfree( colTab_d, s );
	dfree( satF64, s );
	if( r == 0 )
	{
		cudaError_t e = cudaStreamSynchronize( s ); // renderCommon.hpp:313; also: the host tables above die with this scope
		if( e != cudaSuccess ) r = cudaFail( e, "environment build", __FILE__, __LINE__ );
	}
	if( r )
	{
		cudaStreamSynchronize( s );
		ptFreeHdri( pt, s );
		return r;
	}
	pt->hw = w, pt->hh = h;
	pt->wP = rgbaPrimary ? wP : w;
	pt->hP = rgbaPrimary ? hP : h;
	return HAKO_OK;
}

int hako_pt_step( hako_pt* pt, const hako_octree* o, hako_stream stream, const hako_camera* cam, int iteration, int y0, int y1, int maxDepth )
{
	PT_ARG( pt && o && cam, "hako_pt_step: null argument" );
	if( !pt->accum )
	{
		hako::setError( "hako_pt_step: no frame buffer (call hako_pt_resize first)" );
		return HAKO_E_STATE;
	}
	PT_ARG( 0 <= y0 && y0 <= y1 && y1 <= pt->h, "hako_pt_step: row range outside the frame" );
	PT_ARG( iteration >= 0 && maxDepth >= 0, "hako_pt_step: negative iteration / maxDepth" );
	if( o->numberOfVoxels && !o->attrs )
	{
		hako::setError( "hako_pt_step: the octree has no voxel attributes (imported without attrs)" );
		return HAKO_E_STATE;
	}
	if( y0 == y1 ) return HAKO_OK;
	hako_ctx* ctx = pt->ctx;
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );

	const uint32_t nPixels = (uint32_t)( y1 - y0 ) * (uint32_t)pt->w;
	const uint64_t nPaths = (uint64_t)nPixels * PT_SPP;
	if( pt->pathCap < nPaths )
	{
		dfree( pt->pathL, s );
		dfree( pt->hitQueue, s );
		pt->pathL = nullptr;
		pt->hitQueue = nullptr;
		pt->pathCap = 0;
		HAKO_TRY( dalloc( ctx, &pt->pathL, nPaths, s ) );
		HAKO_TRY( dalloc( ctx, &pt->hitQueue, nPaths, s ) );
		pt->pathCap = nPaths;
	}
	HAKO_CUDA( cudaMemsetAsync( pt->counters, 0, sizeof( unsigned long long ), s ) );	  // the work counter; [1] keeps counting rays
	HAKO_CUDA( cudaMemsetAsync( pt->counters + 2, 0, sizeof( unsigned long long ), s ) ); // hit-queue length

	PtParams p;
	p.tree = o->view();
	p.env = PtEnv{ pt->hdriPixels, pt->hdriPrimary, pt->sats, pt->rowTab, pt->hw, pt->hh, pt->wP, pt->hP, pt->hdriPixels ? pt->hdriScale : 0.0f };
	p.cam = *cam;
	p.pmj = pt->pmj;
	p.pathL = pt->pathL;
	p.counters = pt->counters;
	p.hitQueue = pt->hitQueue;
	p.w = pt->w, p.h = pt->h, p.y0 = y0;
	p.nPaths = (uint32_t)nPaths;
	p.iteration = (uint32_t)iteration;
	p.maxDepth = maxDepth;

	const bool envOn = p.env.pixels != nullptr && 0.0f < p.env.scale;
	const uint32_t nExtra = o->hasEmission ? 1u : 0u;
	const int waves = std::max( maxDepth, 1 ); // vertices per path: the primary hit is always shaded (voxKernel.cu:691-693)
	if( pt->wavefront == 2 && waves <= PT_MAX_WAVES )
	{
		// full wavefront, in row bands of at most waveBandPaths paths so that the live-path buffers stay bounded even when
		// every primary ray hits; band b runs on lane b % lanes (own stream, counters and wave buffers), the lanes side by side
		const int lanes = pt->lanes;
		const uint32_t rowsTotal = (uint32_t)( y1 - y0 );
		const uint32_t rowsCap = (uint32_t)std::max<uint64_t>( 4, ( pt->waveBandPaths / ( (uint64_t)pt->w * PT_SPP ) ) & ~3ull );
		const uint32_t rowsPerBand = std::min( rowsCap, std::max( 4u, ( ( rowsTotal + lanes - 1 ) / lanes + 3u ) & ~3u ) );
		const uint64_t bandCap = std::min<uint64_t>( nPaths, (uint64_t)rowsPerBand * pt->w * PT_SPP );
		if( pt->waveCap < bandCap )
		{
			for( int k = 0; k < PT_LANES; k++ )
			{
				for( int i = 0; i < 2; i++ )
				{
					dfree( pt->waveV[k][i], s ), dfree( pt->waveDir[k][i], s );
					pt->waveV[k][i] = nullptr, pt->waveDir[k][i] = nullptr;
				}
				dfree( pt->waveRes[k], s );
				pt->waveRes[k] = nullptr;
			}
			pt->waveCap = 0;
			for( int k = 0; k < lanes; k++ )
			{
				for( int i = 0; i < 2; i++ )
				{
					WaveVertex* v = nullptr;
					HAKO_TRY( dalloc( ctx, &v, bandCap, s ) );
					pt->waveV[k][i] = v;
					HAKO_TRY( dalloc( ctx, &pt->waveDir[k][i], bandCap * 3, s ) );
				}
				HAKO_TRY( dalloc( ctx, &pt->waveRes[k], bandCap * 3, s ) );
			}
			pt->waveCap = bandCap;
		}
		// bins of a trace wave: ( up to 3 ray kinds ) x 8 octants x cells, scanned in tiles of BIN_SCAN_TILE
		const uint32_t bins = pt->binBits ? ( 3u * 8u ) << ( 3 * pt->binBits ) : 0u;
		const uint32_t binTiles = ( bins + BIN_SCAN_TILE - 1 ) / BIN_SCAN_TILE;
		const uint64_t binStateWords = ( scanStateBytes( binTiles ) + 3 ) / 4;
		if( bins && pt->binCap < bandCap )
		{
			for( int k = 0; k < PT_LANES; k++ )
			{
				dfree( pt->binKeys[k], s ), dfree( pt->binOrder[k], s ), dfree( pt->binHist[k], s );
				pt->binKeys[k] = pt->binOrder[k] = pt->binHist[k] = nullptr;
			}
			pt->binCap = 0;
			for( int k = 0; k < lanes; k++ )
			{
				HAKO_TRY( dalloc( ctx, &pt->binKeys[k], bandCap * 3, s ) );
				HAKO_TRY( dalloc( ctx, &pt->binOrder[k], bandCap * 3, s ) );
				HAKO_TRY( dalloc( ctx, &pt->binHist[k], binStateWords + bins, s ) );
			}
			pt->binCap = bandCap;
		}
		// occupancy target of the trace kernel: 5 CTAs/SM = 48 registers, 6 = 40, 8 = 32 (HAKO_PT_TRACE_CTAS)
		auto kTrace5 = traceWaveKernel<5>;
		auto kTrace6 = traceWaveKernel<6>;
		auto kTrace8 = traceWaveKernel<8>;
		auto kTraceSmem = traceWaveSmemKernel<3, false>;
		auto kTraceSmemPacked = traceWaveSmemKernel<3, true>;
		auto kTrace5Top = traceWaveKernel<5, 512>; // HAKO_SMEM_TOP=512: top 512 nodes in shared memory (A/B)
		static const int topEnv = getenv( "HAKO_SMEM_TOP" ) ? atoi( getenv( "HAKO_SMEM_TOP" ) ) : 0;
		const int occ = pt->waveTraceCtas >= 8 ? 8 : ( pt->waveTraceCtas >= 6 ? 6 : 5 );
		const uint32_t traceGrid = (uint32_t)occ * ctx->smCount;
		const uint32_t shadeGrid = 8u * ctx->smCount;
		// fork: the lanes start after everything queued on `s` (allocations, the previous step's accumulate)
		HAKO_CUDA( cudaEventRecord( pt->laneEv[PT_LANES], s ) );
		for( int k = 0; k < lanes; k++ ) HAKO_CUDA( cudaStreamWaitEvent( pt->laneStream[k], pt->laneEv[PT_LANES], 0 ) );
		// the bands are enqueued inside a callable so that a failure in the middle still reaches the join below: the lanes must never
		// be left forked from `s`
		const auto enqueueBands = [&]() -> int {
		int band = 0;
		for( int b0 = y0; b0 < y1; b0 += (int)rowsPerBand, band++ )
		{
			const int lane = band % lanes;
			cudaStream_t ls = pt->laneStream[lane];
			unsigned long long* counters = pt->counters + (size_t)lane * PT_COUNTERS;
			const int b1 = std::min( y1, b0 + (int)rowsPerBand );
			PtParams q = p;
			q.y0 = b0;
			q.pathL = pt->pathL + (size_t)( b0 - y0 ) * pt->w * PT_SPP;
			q.hitQueue = pt->hitQueue + (size_t)( b0 - y0 ) * pt->w * PT_SPP;
			q.counters = counters;
			q.nPaths = (uint32_t)( (uint64_t)( b1 - b0 ) * pt->w * PT_SPP );
			// live counts [2 .. 2 + waves] and the waves' ray counters [32 .. 32 + waves)
			HAKO_CUDA( cudaMemsetAsync( counters + 2, 0, ( PT_COUNTERS - 3 ) * sizeof( unsigned long long ), ls ) ); // [1] rays and [63] steps keep counting
			const uint32_t tiles = ( ( (uint32_t)pt->w + 3u ) / 4u ) * ( ( (uint32_t)( b1 - b0 ) + 3u ) / 4u );
			HAKO_LAUNCH( ctx, primaryWaveKernel, tiles, PT_BLOCK, 0, ls, q, b1 - b0 );
			for( int v = 0; v <= waves; v++ )
			{
				// shade( v ): consumes the rays of vertex v-1; for v == waves every path ends (depth == maxDepth)
				WaveParams wp;
				wp.vin = (const WaveVertex*)pt->waveV[lane][( v + 1 ) & 1];
				wp.vout = (WaveVertex*)pt->waveV[lane][v & 1];
				wp.dirIn = pt->waveDir[lane][( v + 1 ) & 1];
				wp.dirOut = pt->waveDir[lane][v & 1];
				wp.resIn = pt->waveRes[lane];
				wp.nIn = counters + 2 + v;
				wp.nOut = counters + 2 + v + 1;
				wp.v = v;
				wp.rIn = v > 0 ? waveRaysPerVertex( envOn, nExtra, v - 1 ) : 0;
				wp.rOut = waveRaysPerVertex( envOn, nExtra, v );
				HAKO_LAUNCH( ctx, shadeWaveKernel, shadeGrid, 256, 0, ls, q, wp );
				if( v == waves ) break;
				TraceWaveParams tp;
				tp.vertices = wp.vout;
				tp.dirs = wp.dirOut;
				tp.results = pt->waveRes[lane];
				tp.nVertices = wp.nOut;
				tp.work = counters + 32 + v;
				tp.rays = counters + 1;
				tp.steps = counters + PT_COUNTERS - 1;
				tp.raysPerVertex = wp.rOut;
				tp.shadowFirst = envOn ? 1 : 0;
				tp.refillMin = std::max( 1, std::min( 32, pt->waveRefillMin ) );
				tp.burst = std::max( 1, pt->waveBurst );
				tp.order = nullptr;
				if( bins )
				{
					BinParams bp;
					bp.vertices = tp.vertices, bp.dirs = tp.dirs, bp.nVertices = tp.nVertices, bp.raysPerVertex = tp.raysPerVertex;
					bp.cellBits = pt->binBits, bp.cellMajor = pt->binCellMajor;
					bp.lower = q.tree.lower;
					const float cells = (float)( 1 << pt->binBits );
					bp.scale = V3{ cells / ( q.tree.upper.x - q.tree.lower.x ), cells / ( q.tree.upper.y - q.tree.lower.y ), cells / ( q.tree.upper.z - q.tree.lower.z ) };
					bp.keys = pt->binKeys[lane], bp.order = pt->binOrder[lane];
					bp.hist = pt->binHist[lane] + binStateWords;
					bp.bins = bins;
					bp.scan = makeScanState( pt->binHist[lane], binTiles );
					HAKO_CUDA( cudaMemsetAsync( pt->binHist[lane], 0, ( binStateWords + bins ) * sizeof( uint32_t ), ls ) );
					HAKO_LAUNCH( ctx, binCountKernel, shadeGrid, 256, 0, ls, bp );
					HAKO_LAUNCH( ctx, binScanKernel, binTiles, 1024, 0, ls, bp );
					HAKO_LAUNCH( ctx, binScatterKernel, shadeGrid, 256, 0, ls, bp );
					tp.order = bp.order;
				}
				int smemDepth = 0;
				if( pt->waveSmemStack )
				{
					// one push per level at most
					int lv = 0;
					while( ( 1 << lv ) < o->gridRes ) lv++;
					if( o->gridRes > 0 && lv >= 1 && lv <= HAKO_STACK_DEPTH ) smemDepth = lv;
				}
				// whole 24-byte entries when three CTAs of them fit an SM (up to 4096^3), else 22-byte entries; trees too deep for three CTAs
				// of those keep the stack in local memory (two CTAs per SM measured no faster than the local stack at 5)
				auto ctasFor = []( size_t bytes ) { return (uint32_t)std::min<size_t>( 4, ( 227u * 1024u ) / ( bytes + 1024 ) ); };
				const size_t smemBytes24 = (size_t)smemDepth * 24 * 256, smemBytes22 = (size_t)smemDepth * 22 * 256;
				const bool smemPacked = smemDepth && ctasFor( smemBytes24 ) < 3;
				const size_t smemBytes = smemPacked ? smemBytes22 : smemBytes24;
				const uint32_t smemPerSm = smemDepth ? ctasFor( smemBytes ) : 0u;
				const bool smemStack = smemDepth && smemPerSm >= 3;
				if( smemStack )
				{
					tp.refillMin = std::max( 1, std::min( 32, pt->waveRefillMinSmem ) );
					static size_t configured[16][2] = {};
					if( configured[ctx->device & 15][smemPacked] < smemBytes )
					{
						if( smemPacked )
							HAKO_CUDA( cudaFuncSetAttribute( kTraceSmemPacked, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes ) );
						else
							HAKO_CUDA( cudaFuncSetAttribute( kTraceSmem, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes ) );
						configured[ctx->device & 15][smemPacked] = smemBytes;
					}
					if( smemPacked )
						HAKO_LAUNCH( ctx, kTraceSmemPacked, smemPerSm * ctx->smCount, 256, smemBytes, ls, q.tree, tp, smemDepth );
					else
						HAKO_LAUNCH( ctx, kTraceSmem, smemPerSm * ctx->smCount, 256, smemBytes, ls, q.tree, tp, smemDepth );
				}
				else if( topEnv == 512 )
					HAKO_LAUNCH( ctx, kTrace5Top, traceGrid, 256, 0, ls, q.tree, tp );
				else if( occ == 8 )
					HAKO_LAUNCH( ctx, kTrace8, traceGrid, 256, 0, ls, q.tree, tp );
				else if( occ == 6 )
					HAKO_LAUNCH( ctx, kTrace6, traceGrid, 256, 0, ls, q.tree, tp );
				else
					HAKO_LAUNCH( ctx, kTrace5, traceGrid, 256, 0, ls, q.tree, tp );
			}
		}
		return 0;
		};
		const int bandsResult = enqueueBands();
		// join: the accumulate below (and whatever the caller queues next) waits for every lane -- also when a band failed
		cudaError_t je = cudaSuccess;
		for( int k = 0; k < lanes; k++ )
		{
			cudaError_t e1 = cudaEventRecord( pt->laneEv[k], pt->laneStream[k] );
			if( e1 == cudaSuccess ) e1 = cudaStreamWaitEvent( s, pt->laneEv[k], 0 );
			if( e1 != cudaSuccess && je == cudaSuccess ) je = e1;
		}
		if( bandsResult ) return bandsResult;
		HAKO_CUDA( je );
	}
	else
	{
	// persistent grid: as many CTAs as are co-resident on the chip
	static int ctasPerSm[2] = { 0, 0 };
	auto kWhole = pathTraceKernel<false>;
	auto kContinue = pathTraceKernel<true>;
	const int variant = pt->wavefront ? 1 : 0;
	if( ctasPerSm[variant] == 0 )
	{
		if( variant )
			HAKO_CUDA( cudaOccupancyMaxActiveBlocksPerMultiprocessor( &ctasPerSm[1], kContinue, PT_BLOCK, 0 ) );
		else
			HAKO_CUDA( cudaOccupancyMaxActiveBlocksPerMultiprocessor( &ctasPerSm[0], kWhole, PT_BLOCK, 0 ) );
		if( ctasPerSm[variant] < 1 ) ctasPerSm[variant] = 1;
	}
	const uint32_t grid = (uint32_t)std::min<uint64_t>( (uint64_t)ctasPerSm[variant] * ctx->smCount, ( nPaths + PT_BLOCK - 1 ) / PT_BLOCK );
	if( variant )
	{
		// wave 1: every primary ray (dense launch, 4x4-pixel tiles); wave 2: the paths that hit, from the compacted queue
		const uint32_t tiles = ( ( (uint32_t)pt->w + 3u ) / 4u ) * ( ( (uint32_t)( y1 - y0 ) + 3u ) / 4u );
		HAKO_LAUNCH( ctx, primaryWaveKernel, tiles, PT_BLOCK, 0, s, p, y1 - y0 );
		HAKO_LAUNCH( ctx, kContinue, grid, PT_BLOCK, 0, s, p );
	}
	else
		HAKO_LAUNCH( ctx, kWhole, grid, PT_BLOCK, 0, s, p );
	}
	HAKO_LAUNCH( ctx, accumulateKernel, divUp( nPixels, 256 ), 256, 0, s, pt->pathL, pt->accum + (size_t)y0 * pt->w, nPixels );
	return HAKO_OK;
}

int hako_pt_resolve( hako_pt* pt, hako_stream stream )
{
	PT_ARG( pt, "hako_pt_resolve: null path tracer" );
	if( !pt->accum )
	{
		hako::setError( "hako_pt_resolve: no frame buffer" );
		return HAKO_E_STATE;
	}
	HAKO_CUDA( cudaSetDevice( pt->ctx->device ) );
	const uint32_t n = (uint32_t)pt->w * pt->h;
	HAKO_LAUNCH( pt->ctx, resolveKernel, divUp( n, 256 ), 256, 0, (cudaStream_t)stream, pt->rgba8, pt->accum, n );
	return HAKO_OK;
}

int hako_pt_read_rgba8( hako_pt* pt, hako_stream stream, uint8_t* rgba )
{
	PT_ARG( pt && rgba, "hako_pt_read_rgba8: null argument" );
	HAKO_TRY( hako_pt_resolve( pt, stream ) ); // toImageAsync = resolve + D2H, PathTracer.hpp:128-137
	HAKO_CUDA( cudaMemcpyAsync( rgba, pt->rgba8, (size_t)pt->w * pt->h * 4, cudaMemcpyDeviceToHost, (cudaStream_t)stream ) );
	HAKO_CUDA( cudaStreamSynchronize( (cudaStream_t)stream ) );
	return HAKO_OK;
}

// the reference's toImageAsync proper (PathTracer.hpp:118-129): resolve + D2H enqueued, NO host wait -- the caller synchronises the
// stream (voxPTGPU.cpp:188-189).  rgba should be pinned for the copy to be truly asynchronous.
int hako_pt_read_rgba8_async( hako_pt* pt, hako_stream stream, uint8_t* rgba )
{
	PT_ARG( pt && rgba, "hako_pt_read_rgba8_async: null argument" );
	HAKO_TRY( hako_pt_resolve( pt, stream ) );
	HAKO_CUDA( cudaMemcpyAsync( rgba, pt->rgba8, (size_t)pt->w * pt->h * 4, cudaMemcpyDeviceToHost, (cudaStream_t)stream ) );
	return HAKO_OK;
}

int hako_pt_read_accum( hako_pt* pt, hako_stream stream, float* rgbw )
{
	PT_ARG( pt && rgbw, "hako_pt_read_accum: null argument" );
	if( !pt->accum )
	{
		hako::setError( "hako_pt_read_accum: no frame buffer" );
		return HAKO_E_STATE;
	}
	HAKO_CUDA( cudaSetDevice( pt->ctx->device ) );
	HAKO_CUDA( cudaMemcpyAsync( rgbw, pt->accum, (size_t)pt->w * pt->h * sizeof( float4 ), cudaMemcpyDeviceToHost, (cudaStream_t)stream ) );
	HAKO_CUDA( cudaStreamSynchronize( (cudaStream_t)stream ) );
	return HAKO_OK;
}

int hako_pt_accum_device( hako_pt* pt, void** accum_d, uint64_t* bytes )
{
	PT_ARG( pt && accum_d && bytes, "hako_pt_accum_device: null argument" );
	*accum_d = pt->accum;
	*bytes = (uint64_t)pt->w * pt->h * sizeof( float4 );
	return HAKO_OK;
}

int hako_pt_rgba8_device( hako_pt* pt, void** rgba8_d, int* w, int* h )
{
	PT_ARG( pt && rgba8_d && w && h, "hako_pt_rgba8_device: null argument" );
	*rgba8_d = pt->rgba8;
	*w = pt->w, *h = pt->h;
	return HAKO_OK;
}
hako_ctx* hako_pt_context( hako_pt* pt ) { return pt ? pt->ctx : nullptr; }

int hako_pt_ray_count( hako_pt* pt, hako_stream stream, uint64_t* raysTraced, int reset )
{
	PT_ARG( pt && raysTraced, "hako_pt_ray_count: null argument" );
	HAKO_CUDA( cudaSetDevice( pt->ctx->device ) );
	// every lane counts in its own block (steps join their lanes into the caller's stream before they return)
	unsigned long long v[PT_LANES] = {};
	for( int k = 0; k < PT_LANES; k++ )
	{
		HAKO_CUDA( cudaMemcpyAsync( &v[k], pt->counters + k * PT_COUNTERS + 1, sizeof( v[k] ), cudaMemcpyDeviceToHost, (cudaStream_t)stream ) );
		if( reset ) HAKO_CUDA( cudaMemsetAsync( pt->counters + k * PT_COUNTERS + 1, 0, sizeof( v[k] ), (cudaStream_t)stream ) );
	}
	HAKO_CUDA( cudaStreamSynchronize( (cudaStream_t)stream ) );
	*raysTraced = 0;
	for( int k = 0; k < PT_LANES; k++ ) *raysTraced += v[k];
	return HAKO_OK;
}

int hako_pt_trace_steps( hako_pt* pt, hako_stream stream, uint64_t* steps, int reset )
{
	PT_ARG( pt && steps, "hako_pt_trace_steps: null argument" );
	HAKO_CUDA( cudaSetDevice( pt->ctx->device ) );
	unsigned long long v[PT_LANES] = {};
	for( int k = 0; k < PT_LANES; k++ )
	{
		HAKO_CUDA( cudaMemcpyAsync( &v[k], pt->counters + k * PT_COUNTERS + PT_COUNTERS - 1, sizeof( v[k] ), cudaMemcpyDeviceToHost, (cudaStream_t)stream ) );
		if( reset ) HAKO_CUDA( cudaMemsetAsync( pt->counters + k * PT_COUNTERS + PT_COUNTERS - 1, 0, sizeof( v[k] ), (cudaStream_t)stream ) );
	}
	HAKO_CUDA( cudaStreamSynchronize( (cudaStream_t)stream ) );
	*steps = 0;
	for( int k = 0; k < PT_LANES; k++ ) *steps += v[k];
	return HAKO_OK;
}

int hako_pt_pmj_table( hako_pt* pt, hako_stream stream, float* samples )
{
	PT_ARG( pt && samples, "hako_pt_pmj_table: null argument" );
	HAKO_CUDA( cudaSetDevice( pt->ctx->device ) );
	HAKO_CUDA( cudaMemcpyAsync( samples, pt->pmj, sizeof( float ) * 2 * PMJ_LENGTH * PMJ_NSEQ, cudaMemcpyDeviceToHost, (cudaStream_t)stream ) );
	HAKO_CUDA( cudaStreamSynchronize( (cudaStream_t)stream ) );
	return HAKO_OK;
}

int hako_pt_hdri_sats( hako_pt* pt, hako_stream stream, uint32_t* sats )
{
	PT_ARG( pt && sats, "hako_pt_hdri_sats: null argument" );
	if( !pt->sats )
	{
		hako::setError( "hako_pt_hdri_sats: no environment loaded" );
		return HAKO_E_STATE;
	}
	HAKO_CUDA( cudaSetDevice( pt->ctx->device ) );
	HAKO_CUDA( cudaMemcpyAsync( sats, pt->sats, sizeof( uint32_t ) * 7 * pt->hw * pt->hh, cudaMemcpyDeviceToHost, (cudaStream_t)stream ) );
	HAKO_CUDA( cudaStreamSynchronize( (cudaStream_t)stream ) );
	return HAKO_OK;
}
}
