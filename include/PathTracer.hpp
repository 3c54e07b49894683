// PathTracer.hpp -- header-only C++ shim over the C ABI (hako.h) with the reference's PathTracer surface
// (reference PathTracer.hpp:14-170): setup / cleanUp / resizeFrameBufferIfNeeded / clearFrameBuffer / loadHDRI /
// updateScene / step / resolve / toImageAsync / getSteps / getNumberOfVoxels / getOctreeBytes, in the call order of
// voxPTGPU.cpp:88-91,169,177-193 and RTCamp.cpp:111-123,147-169.
//
// Every method keeps the reference's argument list, so voxPTGPU.cpp:88-91,177-193 compiles with only the include swapped:
//   - PathTracer() default-constructs (device 0, or $HAKO_DEVICE) as the reference's does; PathTracer( hako::Context* ) shares a context;
//   - setup( stream, kernel, includeDir, isNvidia ) (:43): the three nvrtc arguments are accepted and ignored (ahead-of-time sm_100a
//     code, no 1 GiB stack pool); setup( stream ) is the short form;
//   - loadHDRI( stream, file, filePrimary ) (:104): Radiance .hdr decoded by hako_image_load_hdr exactly as stbi_loadf( .., 4 )
//     (libs/prlib/src/prp.cpp:353-366); an overload takes decoded RGBA32F pixels;
//   - step( stream, camera, focus, lensR ) (:150): `camera` is any pr::Camera3D-like struct (origin, lookat, up, fovy, zNear, zFar;
//     libs/prlib/src/pr.hpp:189-198) and goes through hako_camera_lookat == pr::GetCameraMatrix + CameraPinhole::initFromPerspective
//     (:152-156); step( stream, hako_camera ) takes the pinhole directly;
//   - toImageAsync( stream, image* ) (:118): `image` is any pr::Image2DRGBA8-like object (allocate( w, h ), data()); resolve and
//     the D2H copy are ENQUEUED, the caller synchronises the stream as voxPTGPU.cpp:188-189 does; toImage( stream, uint8_t* ) waits;
//   - m_maxDepth is a member (reference: the literal 8 at voxKernel.cu:691; BASELINE config 3 uses 4).
#pragma once

#include <cstdlib>
#include <memory>

#include "IntersectorOctreeGPU.hpp"

class PathTracer
{
	static hako::Context* ownContext( std::unique_ptr<hako::Context>& slot )
	{
		const char* dev = getenv( "HAKO_DEVICE" );
		slot.reset( new hako::Context( dev ? atoi( dev ) : 0 ) );
		return slot.get();
	}
	std::unique_ptr<hako::Context> m_ownCtx; // declared first: it must outlive every member that uses it

public:
	PathTracer() : m_ctx( ownContext( m_ownCtx ) ), m_intersectorOctreeGPU( m_ctx ) {} // `PathTracer pt;`, voxPTGPU.cpp:88
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
	// the reference's argument list (:43): kernel source path, include directory and vendor flag fed nvrtc; nothing to compile here
	void setup( hako_stream stream, const char* /*kernel*/, const char* /*includeDir*/, bool /*isNvidia*/ ) { setup( stream ); }
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
	// the reference's argument list (:150): a pr::Camera3D-like camera -> view / projection matrices -> CameraPinhole (:152-156)
	template <class Camera3D>
	void step( hako_stream stream, const Camera3D& camera, float focus, float lensR )
	{
		const float eye[3] = { camera.origin.x, camera.origin.y, camera.origin.z };
		const float center[3] = { camera.lookat.x, camera.lookat.y, camera.lookat.z };
		const float up[3] = { camera.up.x, camera.up.y, camera.up.z };
		hako_camera pinhole;
		hako::check( hako_camera_lookat( eye, center, up, camera.fovy, m_width, m_height, camera.zNear, camera.zFar, focus, lensR, &pinhole ) );
		step( stream, pinhole );
	}
	// tile / spp sharding across GPUs: rows [y0, y1) of iteration `iteration` (spp = iteration*16 + localSpp, voxKernel.cu:642)
	void stepTile( hako_stream stream, const hako_camera& pinhole, int iteration, int y0, int y1 )
	{
		hako::check( hako_pt_step( m_pt, m_intersectorOctreeGPU.handle(), stream, &pinhole, iteration, y0, y1, m_maxDepth ) );
	}
	void resolve( hako_stream stream ) { hako::check( hako_pt_resolve( m_pt, stream ) ); } // :118-127
	// :118-129 with the reference's argument list: output->allocate( w, h ), resolve, D2H ENQUEUED into output->data(); the caller
	// synchronises the stream before reading the image (voxPTGPU.cpp:188-189)
	template <class Image2DRGBA8>
	void toImageAsync( hako_stream stream, Image2DRGBA8* output )
	{
		output->allocate( m_width, m_height );
		static_assert( sizeof( *output->data() ) == 4, "image pixels must be 4 bytes (glm::u8vec4)" );
		hako::check( hako_pt_read_rgba8_async( m_pt, stream, (uint8_t*)output->data() ) );
	}
	void toImageAsync( hako_stream stream, uint8_t* rgba ) { hako::check( hako_pt_read_rgba8_async( m_pt, stream, rgba ) ); }
	void toImage( hako_stream stream, uint8_t* rgba ) { hako::check( hako_pt_read_rgba8( m_pt, stream, rgba ) ); } // returns after the copy has landed
	void readAccumulators( hako_stream stream, float* rgbw ) { hako::check( hako_pt_read_accum( m_pt, stream, rgbw ) ); }

	hako::Context* m_ctx;
	IntersectorOctreeGPU m_intersectorOctreeGPU;
	hako_pt* m_pt = nullptr;
	int m_width = 0;
	int m_height = 0;
	int m_steps = 0;
	int m_maxDepth = 8;
};
