// IntersectorOctreeGPU.hpp -- header-only C++ shim over the C ABI (hako.h) that keeps the reference's class names,
// method names and argument order, so callers written against
//     IntersectorOctreeGPU   (reference IntersectorOctreeGPU.hpp:21-275)
//     IntersectorOctree      (reference IntersectorOctree.hpp:214-266)
// port by swapping the include (voxRTGPU.cpp:104-106,171,191-203; voxRT.cpp:263-267,288,322).
//
// Differences a caller sees
//   - `Shader* voxKernel` is accepted and ignored (kernels are compiled ahead of time for sm_100a; there is no nvrtc);
//   - `oroStream` is a cudaStream_t passed as void*;
//   - vertex containers are templates over any 12-byte {x,y,z} float struct (glm::vec3 works; glm is not required);
//   - the device-side intersect() is not a __device__ member: rays are traced by the library's own kernels
//     (render / intersectBatch / PathTracer::step).  The POD members are still readable (m_numberOfNodes, ...);
//   - errors throw hako::Error instead of __debugbreak()/abort() (hipUtil.hpp:18-22, IntersectorOctreeGPU.hpp:48-51).
#pragma once

#include <cfloat>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "hako.h"

namespace hako
{
struct Error : std::runtime_error
{
	int code;
	Error( int c, const char* msg ) : std::runtime_error( msg ? msg : "hako error" ), code( c ) {}
};
inline void check( int code )
{
	if( code != HAKO_OK ) throw Error( code, hako_last_error() );
}

// one per device; replaces oroInitialize / oroCtxCreate / Shader (voxRTGPU.cpp:25-68)
class Context
{
public:
	explicit Context( int device = 0 ) { check( hako_ctx_create( device, &m_ctx ) ); }
	~Context() { hako_ctx_destroy( m_ctx ); }
	Context( const Context& ) = delete;
	Context& operator=( const Context& ) = delete;
	hako_ctx* get() const { return m_ctx; }
	void synchronize( hako_stream stream = nullptr ) { check( hako_ctx_synchronize( m_ctx, stream ) ); }
	void reserve( uint64_t bytes, hako_stream stream = nullptr ) { check( hako_ctx_reserve( m_ctx, stream, bytes ) ); } // size the memory pool at start-up
	hako_build_stats buildStats() const
	{
		hako_build_stats s;
		check( hako_ctx_build_stats( m_ctx, &s ) );
		return s;
	}

private:
	hako_ctx* m_ctx = nullptr;
};
} // namespace hako

struct Shader; // the reference's nvrtc wrapper (hipUtil.hpp:106-186); only ever passed through as a pointer

class IntersectorOctreeBase
{
public:
	explicit IntersectorOctreeBase( hako::Context* ctx ) : m_ctx( ctx ) {}
	~IntersectorOctreeBase() { cleanUp(); }
	IntersectorOctreeBase( const IntersectorOctreeBase& ) = delete;
	IntersectorOctreeBase& operator=( const IntersectorOctreeBase& ) = delete;

	// IntersectorOctreeGPU::cleanUp, IntersectorOctreeGPU.hpp:26-38
	void cleanUp()
	{
		hako_octree_destroy( m_octree );
		m_octree = nullptr;
		m_numberOfNodes = m_numberOfVoxels = 0;
	}
	const hako_octree* handle() const { return m_octree; }
	bool hasEmission() const { return m_hasEmission != 0; }
	uint64_t getMemoryConsumption() const { return m_deviceBytes; } // IntersectorOctree.hpp:258-261

	// IntersectorOctree::intersect (IntersectorOctree.hpp:248-257) / IntersectorOctreeGPU::intersect (:243-251) over n rays.
	// Contract as in the reference: preset t[i] = FLT_MAX; a miss leaves t / nMajor / vIndex untouched (voxRT.cpp:313-325).
	void intersect( const float* ro, const float* rd, uint64_t n, float* t, int* nMajor, uint32_t* vIndex, bool isShadowRay = false, hako_stream stream = nullptr ) const
	{
		hako::check( hako_intersect_batch( m_octree, stream, ro, rd, n, isShadowRay ? 1 : 0, t, nMajor, vIndex ) );
	}
	// the `render` kernel (voxKernel.cu:437-483; voxRTGPU.cpp:191-209): pixel-centre primary rays, RGBA8 out (HOST)
	void render( hako_stream stream, const hako_camera& camera, int width, int height, bool showVertexColor, uint8_t* rgba, uint32_t* vIndex = nullptr, float* t = nullptr,
				 int* nMajor = nullptr ) const
	{
		hako::check( hako_render_primary( m_octree, stream, &camera, width, height, 0, height, showVertexColor ? 1 : 0, rgba, t, nMajor, vIndex ) );
	}
	// a run of frames, the orbit loop of voxRTGPU.cpp:150-208: frame i+1 is traced while frame i is copied out.  Outputs are
	// nViews x height x width (HOST), frame-major; any may be null.
	void renderFrames( hako_stream stream, const hako_camera* cameras, int nViews, int width, int height, bool showVertexColor, uint8_t* rgba,
					   uint32_t* vIndex = nullptr, float* t = nullptr, int* nMajor = nullptr ) const
	{
		hako::check( hako_render_primary_views( m_octree, stream, cameras, nViews, width, height, showVertexColor ? 1 : 0, rgba, t, nMajor, vIndex ) );
	}
	// canonical reference layout for parity / tooling: OctreeNode[ ] (68 B), sorted unique Morton keys, VoxelAttirb[ ]
	void download( std::vector<hako_node68>* nodes, std::vector<uint64_t>* mortons = nullptr, std::vector<hako_voxel_attrib>* attrs = nullptr, hako_stream stream = nullptr ) const
	{
		if( nodes ) nodes->resize( m_numberOfNodes );
		if( mortons ) mortons->resize( m_numberOfVoxels );
		if( attrs ) attrs->resize( m_numberOfVoxels );
		hako::check( hako_octree_export( m_octree, stream, nodes ? nodes->data() : nullptr, mortons ? mortons->data() : nullptr, attrs ? attrs->data() : nullptr ) );
	}

	// the POD members of the reference struct (IntersectorOctreeGPU.hpp:265-274 / IntersectorOctree.hpp:262-265)
	uint32_t m_numberOfNodes = 0;
	uint32_t m_numberOfVoxels = 0;
	float m_lower[3] = {};
	float m_upper[3] = {};
	float m_dps = 0.0f;
	float m_emissionScale = 7.5f;
	uint32_t m_hasEmission = 0;

protected:
	void adopt( hako_octree* o )
	{
		cleanUp();
		m_octree = o;
		hako_octree_info i;
		hako::check( hako_octree_info_get( o, &i ) );
		m_numberOfNodes = i.numberOfNodes;
		m_numberOfVoxels = i.numberOfVoxels;
		for( int k = 0; k < 3; k++ ) m_lower[k] = i.lower[k], m_upper[k] = i.upper[k];
		m_dps = i.dps;
		m_emissionScale = i.emissionScale;
		m_hasEmission = i.hasEmission;
		m_deviceBytes = i.deviceBytes;
	}
	hako::Context* m_ctx;
	hako_octree* m_octree = nullptr;
	uint64_t m_deviceBytes = 0;
};

// reference IntersectorOctreeGPU.hpp:21-275
class IntersectorOctreeGPU : public IntersectorOctreeBase
{
public:
	using IntersectorOctreeBase::IntersectorOctreeBase;

	// build( vertices, vcolors, vemissions, voxKernel, stream, origin, dps, gridRes ), IntersectorOctreeGPU.hpp:40-47.
	// Three vertices per triangle, as trianglesFlattened (voxUtil.hpp:8-65) produces.
	template <class Vec3>
	void build( const std::vector<Vec3>& vertices, const std::vector<Vec3>& vcolors, const std::vector<Vec3>& vemissions, Shader* /*voxKernel*/, hako_stream stream, const Vec3& origin,
				float dps, int gridRes )
	{
		static_assert( sizeof( Vec3 ) == 3 * sizeof( float ), "vertex type must be three packed floats (glm::vec3)" );
		if( vertices.size() % 3 != 0 || vcolors.size() != vertices.size() || vemissions.size() != vertices.size() )
			throw hako::Error( HAKO_E_INVALID, "IntersectorOctreeGPU::build: vertex / colour / emission arrays differ in length" );
		hako_octree* o = nullptr;
		hako::check( hako_octree_build( m_ctx->get(), stream, (const float*)vertices.data(), (const float*)vcolors.data(), (const float*)vemissions.data(),
										(uint32_t)( vertices.size() / 3 ), (const float*)&origin, dps, gridRes, &o ) );
		adopt( o );
	}
	// vertex data already resident in HBM (per-frame rebuilds skip the H2D of IntersectorOctreeGPU.hpp:68-70)
	void buildDevice( const float* pos_d, const float* col_d, const float* emi_d, uint32_t nTriangles, hako_stream stream, const float origin[3], float dps, int gridRes )
	{
		hako_octree* o = nullptr;
		hako::check( hako_octree_build_device( m_ctx->get(), stream, pos_d, col_d, emi_d, nTriangles, origin, dps, gridRes, &o ) );
		adopt( o );
	}
};

// reference IntersectorOctree.hpp:214-266 (the CPU class's surface; the work happens on the GPU)
class IntersectorOctree : public IntersectorOctreeBase
{
public:
	using IntersectorOctreeBase::IntersectorOctreeBase;

	// build (naive, :221-227) / buildDAGReference (:229-235): sorted unique Morton list -> octree
	template <class Vec3>
	void build( const std::vector<uint64_t>& mortonVoxels, const Vec3& origin, float dps, int gridRes, hako_stream stream = nullptr )
	{
		buildImpl( mortonVoxels, (const float*)&origin, dps, gridRes, 0, stream );
	}
	template <class Vec3>
	void buildDAGReference( const std::vector<uint64_t>& mortonVoxels, const Vec3& origin, float dps, int gridRes, hako_stream stream = nullptr )
	{
		buildImpl( mortonVoxels, (const float*)&origin, dps, gridRes, 1, stream );
	}
	// single-ray form with the reference's signature (:248-257)
	void intersect( const float ro[3], const float rd[3], float* t, int* nMajor, uint32_t* vIndex ) const
	{
		IntersectorOctreeBase::intersect( ro, rd, 1, t, nMajor, vIndex, false, nullptr );
	}
	using IntersectorOctreeBase::intersect;

private:
	void buildImpl( const std::vector<uint64_t>& m, const float* origin, float dps, int gridRes, int dag, hako_stream stream )
	{
		hako_octree* o = nullptr;
		hako::check( hako_octree_build_from_morton( m_ctx->get(), stream, m.data(), nullptr, m.size(), origin, dps, gridRes, dag, &o ) );
		adopt( o );
	}
};
