/*
 * ref_driver_pt.cpp -- the path-tracing pieces of the REFERENCE that run on the host, behind the oracle ABI.
 * TEST INFRASTRUCTURE (oracle/_ref/libhako_ref.so, built only where /root/reference exists).  Glue only:
 *   MurmurHash32                          voxCommon.hpp:45-109
 *   PCG32                                 renderCommon.hpp:86-111
 *   nested_uniform_scramble/scramble_f32  pmjSampler.hpp:62-102
 *   PMJSampler::setup(false)/sample2d     pmjSampler.hpp:114-173
 *   HDRI::importanceSample                renderCommon.hpp:367-435 (host pointers planted in the public members)
 *   sampleLambertian                      renderCommon.hpp:119-151
 * HDRI::sampleNearest, the SAT build kernels, renderPT and renderResolve are device-only in the reference and are
 * therefore NOT exported here (the port restates them, "parity unpinned").
 */
#include <map>
#include <string.h>
#include <vector>

#include <glm/glm.hpp>
#include <glm/ext.hpp>

#include "voxCommon.hpp"
#include "renderCommon.hpp"
#define private public /* PMJSampler::m_samples (test glue only) */
#include "pmjSampler.hpp"
#undef private

#include "hako_oracle.h"

// PMJSampler::setup( isGPU = true ) and HDRI::load reference the Orochi GPU runtime; those branches are never taken here
// (host tables only), so the four symbols they pull in are satisfied by traps.
oroError OROAPI oroMalloc( oroDeviceptr*, size_t ) { __builtin_trap(); }
oroError OROAPI oroFree( oroDeviceptr ) { __builtin_trap(); }
oroError OROAPI oroMemcpyHtoDAsync( oroDeviceptr, const void*, size_t, oroStream ) { __builtin_trap(); }
oroError OROAPI oroStreamSynchronize( oroStream ) { __builtin_trap(); }

static PMJSampler& pmj()
{
	static PMJSampler s;
	static bool ready = false;
	if( !ready )
	{
		s.setup( false, 0 );
		ready = true;
	}
	return s;
}

extern "C" {
uint32_t orc_murmur_combine( uint32_t seed, const uint32_t* values, int n )
{
	MurmurHash32 h( seed );
	for( int i = 0; i < n; i++ ) h.combine( values[i] );
	return h.getHash();
}
void orc_pcg32_stream( uint64_t seed, uint64_t stream, int n, uint32_t* out )
{
	PCG32 rng;
	rng.setup( seed, stream );
	for( int i = 0; i < n; i++ ) out[i] = rng.nextU32();
}
uint32_t orc_nested_uniform_scramble( uint32_t x, uint32_t seed ) { return nested_uniform_scramble( x, seed ); }
float orc_scramble_f32( float x, uint32_t seed ) { return scramble_f32( x, seed ); }
void orc_pmj_table( float* out )
{
	memcpy( out, pmj().m_samples, sizeof( float ) * 2 * PMJSampler::LENGTH * PMJSampler::N_SEQUENCE );
}
void orc_pmj_sample2d( const float* /*table: the reference owns its own*/, uint32_t sampleIdx, uint32_t dimension, uint32_t stream, float* xy )
{
	float2 s = pmj().sample2d( sampleIdx, dimension, stream );
	xy[0] = s.x;
	xy[1] = s.y;
}
void orc_hdri_importance_sample( const float* rgba, const uint32_t* sats, int w, int h, float scale, const float* N, int axisAligned,
								 float u0, float u1, float u2, float u3, float* dir, float* L, float* srPDF, uint32_t* XY )
{
	HDRI hd;
	const size_t n = (size_t)w * h;
	hd.m_pixels = (float4*)rgba;
	hd.m_sat = (uint32_t*)sats;
	for( int i = 0; i < 6; i++ ) hd.m_sats[i] = (uint32_t*)sats + n * ( i + 1 );
	hd.m_width = w;
	hd.m_height = h;
	hd.m_scale = scale;
	float3 d, l;
	hd.importanceSample( &d, &l, srPDF, float3{ N[0], N[1], N[2] }, axisAligned != 0, u0, u1, u2, u3 );
	dir[0] = d.x, dir[1] = d.y, dir[2] = d.z;
	L[0] = l.x, L[1] = l.y, L[2] = l.z;
	if( XY ) XY[0] = XY[1] = 0xFFFFFFFFu; /* the reference does not report the texel */
	hd.m_pixels = 0; /* not ours to free */
	hd.m_sat = 0;
	for( int i = 0; i < 6; i++ ) hd.m_sats[i] = 0;
}
void orc_sample_lambertian( float a, float b, const float* Ng, float* out )
{
	float3 d = sampleLambertian( a, b, float3{ Ng[0], Ng[1], Ng[2] } );
	out[0] = d.x, out[1] = d.y, out[2] = d.z;
}
}
