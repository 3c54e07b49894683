// PathTracer.hpp -- header-only C++ shim over the C ABI (hako.h) with the reference's PathTracer surface
// (reference PathTracer.hpp:14-170): setup / cleanUp / resizeFrameBufferIfNeeded / clearFrameBuffer / loadHDRI /
// updateScene / step / resolve / toImageAsync / getSteps / getNumberOfVoxels / getOctreeBytes, in the call order of
// voxPTGPU.cpp:88-91,169,177-193 and RTCamp.cpp:111-123,147-169.
//
// Differences a caller sees
//   - setup() takes no kernel path / include dir / isNvidia flag (no nvrtc), and allocates no 1 GiB stack pool;
//   - loadHDRI( stream, file, filePrimary ) as in the reference (Radiance .hdr, decoded by hako_image_load_hdr exactly as
//     stbi_loadf( .., 4 ) does, libs/prlib/src/prp.cpp:353-366), plus an overload taking decoded RGBA32F pixels;
//   - step() takes the pinhole camera directly (hako_camera == CameraPinhole).  A pr::Camera3D caller builds it with
//     hako_camera_lookat( origin, lookat, up, fovy, width, height, zNear, zFar, focus, lensR ), which is
//     pr::GetCameraMatrix + CameraPinhole::initFromPerspective (PathTracer.hpp:152-156);
//   - toImageAsync() fills a caller buffer (width * height * 4 bytes) instead of a pr::Image2DRGBA8 and returns after
//     the copy has landed;
//   - m_maxDepth is a member (reference: the literal 8 at voxKernel.cu:691; BASELINE config 3 uses 4).
#pragma once

#include "IntersectorOctreeGPU.hpp"

class PathTracer
{
public:
	explicit PathTracer( hako::Context* ctx ) : m_ctx( ctx ), m_intersectorOctreeGPU( ctx ) {}
	~PathTracer() { cleanUp(); }
	PathTracer( const PathTracer& ) = delete;
	PathTracer& operator=( const PathTracer& ) = delete;

	int getSteps() const { return m_steps; }
	uint64_t getNumberOfVoxels() const { return m_intersectorOctreeGPU.m_numberOfVoxels; }
	uint64_t getOctreeBytes() const { return (uint64_t)m_intersectorOctreeGPU.m_numberOfNodes * sizeof( hako_node68 ); } // PathTracer.hpp:38-41

	void setup( hako_stream stream ) // PathTracer.hpp:43-69
	{
		cleanUp();
		hako::check( hako_pt_create( m_ctx->get(), stream, &m_pt ) );
	}
	void cleanUp() // :70-79
	{
		m_intersectorOctreeGPU.cleanUp();
		hako_pt_destroy( m_pt );
		m_pt = nullptr;
		m_width = m_height = m_steps = 0;
	}
	void resizeFrameBufferIfNeeded( hako_stream stream, int width, int height ) // :81-97
	{
		if( width != m_width || height != m_height ) m_steps = 0; // a new buffer is cleared
		hako::check( hako_pt_resize( m_pt, stream, width, height ) );
		m_width = width;
		m_height = height;
	}
	void clearFrameBuffer( hako_stream stream ) // :98-103
	{
		m_steps = 0;
		hako::check( hako_pt_clear( m_pt, stream ) );
	}
	void loadHDRI( hako_stream stream, const float* rgba, int width, int height, const float* rgbaPrimary = nullptr, int widthPrimary = 0, int heightPrimary = 0 ) // :104-116
	{
		hako::check( hako_pt_load_hdri( m_pt, stream, rgba, width, height, rgbaPrimary, widthPrimary, heightPrimary ) );
	}
	void loadHDRI( hako_stream stream, const char* file, const char* filePrimary = nullptr ) // :104-116, the reference's signature
	{
		float *a = nullptr, *b = nullptr;
		int w = 0, h = 0, wp = 0, hp = 0;
		hako::check( hako_image_load_hdr( file, &a, &w, &h ) );
		if( filePrimary )
		{
			const int r = hako_image_load_hdr( filePrimary, &b, &wp, &hp );
			if( r )
			{
				hako_image_free( a );
				hako::check( r );
			}
		}
		const int r = hako_pt_load_hdri( m_pt, stream, a, w, h, b, wp, hp );
		hako_image_free( a );
		hako_image_free( b );
		hako::check( r );
	}
	template <class Vec3>
	void updateScene( const std::vector<Vec3>& vertices, const std::vector<Vec3>& vcolors, const std::vector<Vec3>& vemissions, hako_stream stream, const Vec3& origin, float dps,
					  int gridRes ) // :139-148
	{
		m_intersectorOctreeGPU.build( vertices, vcolors, vemissions, (Shader*)nullptr, stream, origin, dps, gridRes );
	}
	// one call = 16 spp over the whole frame (:150-169)
	void step( hako_stream stream, const hako_camera& pinhole )
	{
		hako::check( hako_pt_step( m_pt, m_intersectorOctreeGPU.handle(), stream, &pinhole, m_steps++, 0, m_height, m_maxDepth ) );
	}
	// tile / spp sharding across GPUs: rows [y0, y1) of iteration `iteration` (spp = iteration*16 + localSpp, voxKernel.cu:642)
	void stepTile( hako_stream stream, const hako_camera& pinhole, int iteration, int y0, int y1 )
	{
		hako::check( hako_pt_step( m_pt, m_intersectorOctreeGPU.handle(), stream, &pinhole, iteration, y0, y1, m_maxDepth ) );
	}
	void resolve( hako_stream stream ) { hako::check( hako_pt_resolve( m_pt, stream ) ); } // :118-127
	void toImageAsync( hako_stream stream, uint8_t* rgba ) { hako::check( hako_pt_read_rgba8( m_pt, stream, rgba ) ); } // :128-137
	void readAccumulators( hako_stream stream, float* rgbw ) { hako::check( hako_pt_read_accum( m_pt, stream, rgbw ) ); }

	hako::Context* m_ctx;
	IntersectorOctreeGPU m_intersectorOctreeGPU;
	hako_pt* m_pt = nullptr;
	int m_width = 0;
	int m_height = 0;
	int m_steps = 0;
	int m_maxDepth = 8;
};
