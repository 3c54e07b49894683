/*
 * ref_driver.cpp -- exposes the REFERENCE'S OWN CPU code through the oracle ABI (hako_oracle.h).
 * TEST INFRASTRUCTURE.  Built only where /root/reference exists (the build container) by
 * oracle/Makefile into oracle/_ref/libhako_ref.so; the reference headers are compiled from where
 * they lie (a staged, scope-patched copy of voxCommon.hpp / hipUtil.hpp lives in the git-ignored
 * oracle/_ref/stage/, see the Makefile).  Nothing of the reference is copied into the repository.
 *
 * This file contains only glue: argument marshalling around the reference's functions
 *   VTContext                              voxelization.hpp:124-337
 *   encode2mortonCode_PDEP / _Naive / ...  morton.hpp
 *   buildOctreeDAGReference / Naive        IntersectorOctree.hpp:11-212
 *   IntersectorOctree::intersect           IntersectorOctree.hpp:248-257
 *   CameraPinhole                          renderCommon.hpp:17-84
 * and the driver loops of voxRT.cpp:14-51 (mergeVoxels, reproduced through std::map exactly as the
 * reference does), :198-240 (voxel loop) and :307-347 (render loop), which live in a main() that
 * cannot be linked (window / GL shell).
 */
#include <assert.h>
#include <float.h>
#include <map>
#include <memory>
#include <string.h>
#include <thread>
#include <vector>

#include <glm/glm.hpp>
#include <glm/ext.hpp>

#include "morton.hpp"
#include "voxelization.hpp"
#include "voxCommon.hpp"
#include "IntersectorOctree.hpp"
#include "renderCommon.hpp"

#include "hako_oracle.h"

struct OrcOctree
{
	IntersectorOctree isect;
};

static_assert( sizeof( OrcCamera ) == sizeof( CameraPinhole ), "camera layout" );

extern "C" {

const char* orc_impl_name( void ) { return "reference"; }

uint64_t orc_morton_encode( uint32_t x, uint32_t y, uint32_t z ) { return encode2mortonCode_PDEP( x, y, z ); }
uint64_t orc_morton_encode_naive( uint32_t x, uint32_t y, uint32_t z ) { return encode2mortonCode_Naive( x, y, z ); }
void orc_morton_decode( uint64_t m, uint32_t* x, uint32_t* y, uint32_t* z ) { decodeMortonCode_PEXT( m, x, y, z ); }
int orc_sort_bits_morton( uint32_t gridRes ) { return numberOfSortBitsMorton( gridRes ); }

uint64_t orc_voxelize( const float* pos, const float* col, const float* emi, uint32_t nTriangles,
					   const float origin[3], float dps, int gridRes, int sixSeparating,
					   uint64_t* keys, uint8_t* colors, uint8_t* emissions, uint64_t capacity )
{
	const float3* P = (const float3*)pos;
	const float3* C = (const float3*)col;
	const float3* E = (const float3*)emi;
	bool six = sixSeparating != 0;
	uint64_t n = 0;
	for( uint32_t i = 0; i < nTriangles; i++ )
	{
		float3 v0 = P[i * 3], v1 = P[i * 3 + 1], v2 = P[i * 3 + 2];
		VTContext context( v0, v1, v2, six, { origin[0], origin[1], origin[2] }, dps, gridRes );
		int2 xrange = context.xRangeInclusive();
		for( int x = xrange.x; x <= xrange.y; x++ )
		{
			int2 yrange = context.yRangeInclusive( x, dps );
			for( int y = yrange.x; y <= yrange.y; y++ )
			{
				int2 zrange = context.zRangeInclusive( x, y, dps, six );
				for( int z = zrange.x; z <= zrange.y; z++ )
				{
					float3 p = context.p( x, y, z, dps );
					if( context.intersect( p ) )
					{
						if( n < capacity )
						{
							int3 c = context.i( x, y, z );
							keys[n] = encode2mortonCode_PDEP( c.x, c.y, c.z );
							if( colors || emissions )
							{
								float3 bc = closestBarycentricCoordinateOnTriangle( v0, v1, v2, p );
								if( colors )
								{
									float3 bColor = bc.x * C[i * 3 + 1] + bc.y * C[i * 3 + 2] + bc.z * C[i * 3];
									glm::u8vec4 vc = { bColor.x * 255.0f + 0.5f, bColor.y * 255.0f + 0.5f, bColor.z * 255.0f + 0.5f, 255 };
									memcpy( colors + n * 4, &vc, 4 );
								}
								if( emissions )
								{
									float3 bEmission = bc.x * E[i * 3 + 1] + bc.y * E[i * 3 + 2] + bc.z * E[i * 3];
									glm::u8vec4 ve = { bEmission.x * 255.0f + 0.5f, bEmission.y * 255.0f + 0.5f, bEmission.z * 255.0f + 0.5f, 255 };
									memcpy( emissions + n * 4, &ve, 4 );
								}
							}
						}
						n++;
					}
				}
			}
		}
	}
	return n;
}

uint64_t orc_merge_voxels( uint64_t* keys, uint8_t* colors, uint8_t* emissions, uint64_t n )
{
	struct Attribute
	{
		glm::uvec4 color;
		glm::uvec4 emission;
	};
	std::map<uint64_t, Attribute> voxels;
	const glm::u8vec4* values = (const glm::u8vec4*)colors;
	const glm::u8vec4* emis = (const glm::u8vec4*)emissions;
	for( uint64_t i = 0; i < n; i++ )
	{
		auto key = keys[i];
		auto value = values[i];
		auto emission = emis[i];
		auto it = voxels.find( key );
		if( it == voxels.end() )
		{
			voxels[key] = { { value.x, value.y, value.z, 1 }, { emission.x, emission.y, emission.z, 1 } };
		}
		else
		{
			it->second.color += glm::vec4{ value.x, value.y, value.z, 1 };
			it->second.emission += glm::vec4{ emission.x, emission.y, emission.z, 1 };
		}
	}
	uint64_t k = 0;
	for( auto kv : voxels )
	{
		keys[k] = kv.first;
		auto v = glm::u8vec4( kv.second.color / kv.second.color.w );
		memcpy( colors + k * 4, &v, 4 );
		auto e = glm::u8vec4( kv.second.emission / kv.second.emission.w );
		memcpy( emissions + k * 4, &e, 4 );
		k++;
	}
	return k;
}

static std::vector<uint64_t> toVec( const uint64_t* m, uint64_t n ) { return std::vector<uint64_t>( m, m + n ); }

OrcOctree* orc_octree_build_dag( const uint64_t* mortons, uint64_t n, const float origin[3], float dps, int gridRes )
{
	OrcOctree* o = new OrcOctree();
	o->isect.buildDAGReference( toVec( mortons, n ), glm::vec3( origin[0], origin[1], origin[2] ), dps, gridRes );
	return o;
}
OrcOctree* orc_octree_build_naive( const uint64_t* mortons, uint64_t n, const float origin[3], float dps, int gridRes )
{
	OrcOctree* o = new OrcOctree();
	o->isect.build( toVec( mortons, n ), glm::vec3( origin[0], origin[1], origin[2] ), dps, gridRes );
	// buildOctreeNaive leaves nVoxelsPSum of absent children uninitialised; zero them for comparisons.
	for( auto& nd : o->isect.m_nodes )
		for( int i = 0; i < 8; i++ ) nd.nVoxelsPSum[i] = 0;
	return o;
}
void orc_octree_free( OrcOctree* o ) { delete o; }
uint32_t orc_octree_num_nodes( const OrcOctree* o ) { return (uint32_t)o->isect.m_nodes.size(); }
void orc_octree_bounds( const OrcOctree* o, float lower[3], float upper[3] )
{
	for( int i = 0; i < 3; i++ )
	{
		lower[i] = o->isect.m_lower[i];
		upper[i] = o->isect.m_upper[i];
	}
}
void orc_octree_copy_nodes( const OrcOctree* o, OrcNode68* out )
{
	const auto& ns = o->isect.m_nodes;
	for( size_t i = 0; i < ns.size(); i++ )
	{
		memset( &out[i], 0, sizeof( OrcNode68 ) );
		out[i].mask = ns[i].mask;
		memcpy( out[i].children, ns[i].children, 32 );
		memcpy( out[i].nVoxelsPSum, ns[i].nVoxelsPSum, 32 );
	}
}

void orc_intersect_batch( const OrcOctree* o, const float* ro, const float* rd, uint64_t n, int isShadowRay,
						  float* t, int* nMajor, uint32_t* vIndex, uint64_t* stats )
{
	OrcOctree* oo = const_cast<OrcOctree*>( o );
	if( oo->isect.m_nodes.empty() ) return;
	for( uint64_t i = 0; i < n; i++ )
	{
		float3 a = { ro[i * 3], ro[i * 3 + 1], ro[i * 3 + 2] };
		float3 d = { rd[i * 3], rd[i * 3 + 1], rd[i * 3 + 2] };
		if( isShadowRay )
		{
			StackElement stack[32];
			const auto& I = oo->isect;
			octreeTraverse_EfficientParametric( I.m_nodes.data(), (uint32_t)I.m_nodes.size() - 1, stack, a, d,
												{ I.m_lower.x, I.m_lower.y, I.m_lower.z }, { I.m_upper.x, I.m_upper.y, I.m_upper.z },
												&t[i], &nMajor[i], &vIndex[i], true );
		}
		else
		{
			oo->isect.intersect( a, d, &t[i], &nMajor[i], &vIndex[i] );
		}
	}
	if( stats ) stats[0] = stats[1] = stats[2] = stats[3] = 0; // the reference has no counters
}

void orc_camera_lookat( const float eye[3], const float center[3], const float up[3], float fovy,
						int width, int height, float zNear, float zFar, float focus, float lensR, OrcCamera* out )
{
	// libs/prlib/src/pr.cpp:1104-1112 (projection), :1154-1156 (view)
	glm::mat4 proj = glm::perspectiveFov( fovy, (float)width, (float)height, zNear, zFar );
	glm::mat4 view = glm::lookAt( glm::vec3( eye[0], eye[1], eye[2] ), glm::vec3( center[0], center[1], center[2] ), glm::vec3( up[0], up[1], up[2] ) );
	CameraPinhole cam;
	cam.initFromPerspective( view, proj, focus, lensR );
	memcpy( out, &cam, sizeof( cam ) );
}
void orc_camera_shoot( const OrcCamera* cam, int x, int y, float ox, float oy, int w, int h, float ro[3], float rd[3] )
{
	CameraPinhole c;
	memcpy( &c, cam, sizeof( c ) );
	float3 o, d;
	c.shoot( &o, &d, x, y, ox, oy, w, h );
	ro[0] = o.x, ro[1] = o.y, ro[2] = o.z;
	rd[0] = d.x, rd[1] = d.y, rd[2] = d.z;
}
void orc_camera_shoot_thin_lens( const OrcCamera* cam, int x, int y, float ox, float oy, int w, int h, float u0, float u1, float ro[3], float rd[3] )
{
	CameraPinhole c;
	memcpy( &c, cam, sizeof( c ) );
	float3 o, d;
	c.shootThinLens( &o, &d, x, y, ox, oy, w, h, u0, u1 );
	ro[0] = o.x, ro[1] = o.y, ro[2] = o.z;
	rd[0] = d.x, rd[1] = d.y, rd[2] = d.z;
}

void orc_render_primary( const OrcOctree* o, const OrcCamera* cam, int w, int h, int y0, int y1, int nThreads,
						 float* t, int* nMajor, uint32_t* vIndex, uint64_t* stats )
{
	OrcOctree* oo = const_cast<OrcOctree*>( o );
	CameraPinhole pinhole;
	memcpy( &pinhole, cam, sizeof( pinhole ) );
	if( nThreads < 1 ) nThreads = 1;
	// voxRT.cpp:307-347 renderLine
	auto renderLine = [&]( int j ) {
		for( int i = 0; i < w; ++i )
		{
			float3 ro, rd;
			pinhole.shoot( &ro, &rd, i, j, 0.5f, 0.5f, w, h );
			float tt = FLT_MAX;
			int nm = -1;
			uint32_t vi = 0xFFFFFFFFu;
			if( !oo->isect.m_nodes.empty() )
			{
				oo->isect.intersect( ro, rd, &tt, &nm, &vi );
			}
			size_t p = (size_t)j * w + i;
			t[p] = tt;
			nMajor[p] = nm;
			vIndex[p] = vi;
		}
	};
	if( nThreads == 1 )
	{
		for( int j = y0; j < y1; j++ ) renderLine( j );
	}
	else
	{
		std::vector<std::thread> th;
		for( int k = 0; k < nThreads; k++ )
			th.emplace_back( [&, k]() { for( int j = y0 + k; j < y1; j += nThreads ) renderLine( j ); } );
		for( auto& x : th ) x.join();
	}
	if( stats ) stats[0] = stats[1] = stats[2] = stats[3] = 0;
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

// The reference headers define non-inline functions (embedMask, octreeTraverse_EfficientParametric), so the
// path-tracing glue has to live in the same translation unit.
#include "ref_driver_pt.cpp"
