// shim_demo.cpp -- a voxRTGPU / voxPTGPU-style caller written against the drop-in C++ shims (include/*.hpp):
// the same calls, in the same order, as voxRTGPU.cpp:104-209 and voxPTGPU.cpp:88-193, headless (no window).
// Built and linked by tests/test_abi.py (CPU) and executed by tests/test_gpu_shim.py (GPU).
#include <cmath>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

#include "IntersectorOctreeGPU.hpp"
#include "PathTracer.hpp"

struct vec3 // stand-in for glm::vec3
{
	float x, y, z;
};

// a small closed shape: an octahedron with a bright top, plus a ground quad
static void makeScene( std::vector<vec3>& v, std::vector<vec3>& c, std::vector<vec3>& e )
{
	const vec3 P[6] = { { 1, 0, 0 }, { -1, 0, 0 }, { 0, 1, 0 }, { 0, -1, 0 }, { 0, 0, 1 }, { 0, 0, -1 } };
	const int F[8][3] = { { 0, 2, 4 }, { 4, 2, 1 }, { 1, 2, 5 }, { 5, 2, 0 }, { 4, 3, 0 }, { 1, 3, 4 }, { 5, 3, 1 }, { 0, 3, 5 } };
	for( int f = 0; f < 8; f++ )
		for( int k = 0; k < 3; k++ )
		{
			const vec3 p = P[F[f][k]];
			v.push_back( p );
			c.push_back( { 0.5f + 0.5f * p.x, 0.5f + 0.5f * p.y, 0.5f + 0.5f * p.z } );
			e.push_back( p.y > 0.5f ? vec3{ 0.9f, 0.8f, 0.6f } : vec3{ 0, 0, 0 } );
		}
	const vec3 Q[4] = { { -1.5f, -1.0f, -1.5f }, { 1.5f, -1.0f, -1.5f }, { 1.5f, -1.0f, 1.5f }, { -1.5f, -1.0f, 1.5f } };
	const int G[2][3] = { { 0, 1, 2 }, { 0, 2, 3 } };
	for( int f = 0; f < 2; f++ )
		for( int k = 0; k < 3; k++ )
		{
			v.push_back( Q[G[f][k]] );
			c.push_back( { 0.7f, 0.7f, 0.7f } );
			e.push_back( { 0, 0, 0 } );
		}
}

// stand-ins for pr::Camera3D (libs/prlib/src/pr.hpp:189-198) and pr::Image2DRGBA8, as voxPTGPU.cpp uses them
struct Camera3D
{
	vec3 origin = { 4, 4, 4 };
	vec3 lookat = { 0, 0, 0 };
	vec3 up = { 0, 1, 0 };
	float fovy = 45.0f * 3.14159265f / 180.0f;
	float zNear = 0.01f;
	float zFar = 1000.0f;
	bool zUp = false;
	float perspective = 1.0f;
};
struct Image2DRGBA8
{
	struct px
	{
		uint8_t x, y, z, w;
	};
	void allocate( int w, int h )
	{
		m_w = w, m_h = h;
		m_data.resize( (size_t)w * h );
	}
	px* data() { return m_data.data(); }
	int width() const { return m_w; }
	int height() const { return m_h; }
	std::vector<px> m_data;
	int m_w = 0, m_h = 0;
};

// One thread per GPU, the north-star partition through the C ABI alone (no Python, no torch): rank 0 builds, the octree is
// replicated with hako_dist_broadcast_octree, each rank path-traces its spp shard (iterations r, r + G, ...), ONE
// hako_dist_reduce_accum sums the accumulators on rank 0, which compares them with its own 1-GPU render of all iterations.
static int multiGpuSection( const std::vector<vec3>& vertices, const std::vector<vec3>& vcolors, const std::vector<vec3>& vemissions, vec3 origin, float dps, int gridRes,
							const std::vector<float>& env, int ew, int eh, const hako_camera& cam, int W, int H )
{
	const int world = 2, iterations = 4;
	hako_ctx* probe = nullptr;
	if( hako_ctx_create( 1, &probe ) != HAKO_OK )
	{
		printf( "multi-GPU section skipped: %s\n", hako_last_error() );
		return 0;
	}
	hako_ctx_destroy( probe );
	uint8_t id[HAKO_DIST_ID_BYTES];
	hako::check( hako_dist_unique_id( id ) );
	std::vector<float> reduced( (size_t)W * H * 4 ), single( (size_t)W * H * 4 );
	int status[2] = { 0, 0 };
	auto rankMain = [&]( int rank ) {
		try
		{
			hako::Context ctx( rank );
			hako_dist* d = nullptr;
			hako::check( hako_dist_init( ctx.get(), id, rank, world, &d ) );
			IntersectorOctreeGPU built( &ctx );
			if( rank == 0 ) built.build( vertices, vcolors, vemissions, (Shader*)nullptr, nullptr, origin, dps, gridRes );
			hako_octree* tree = nullptr;
			hako::check( hako_dist_broadcast_octree( d, nullptr, 0, rank == 0 ? const_cast<hako_octree*>( built.handle() ) : nullptr, &tree ) );
			hako_pt* pt = nullptr;
			hako::check( hako_pt_create( ctx.get(), nullptr, &pt ) );
			hako::check( hako_pt_resize( pt, nullptr, W, H ) );
			hako::check( hako_pt_load_hdri( pt, nullptr, env.data(), ew, eh, env.data(), ew, eh ) );
			for( int it = rank; it < iterations; it += world ) hako::check( hako_pt_step( pt, tree, nullptr, &cam, it, 0, H, 4 ) );
			hako::check( hako_dist_reduce_accum( d, nullptr, pt, 0 ) );
			if( rank == 0 )
			{
				hako::check( hako_pt_read_accum( pt, nullptr, reduced.data() ) );
				hako::check( hako_pt_clear( pt, nullptr ) );
				for( int it = 0; it < iterations; it++ ) hako::check( hako_pt_step( pt, tree, nullptr, &cam, it, 0, H, 4 ) );
				hako::check( hako_pt_read_accum( pt, nullptr, single.data() ) );
			}
			hako::check( hako_dist_barrier( d, nullptr ) );
			hako_pt_destroy( pt );
			if( rank != 0 ) hako_octree_destroy( tree );
			hako_dist_destroy( d );
		}
		catch( const hako::Error& err )
		{
			fprintf( stderr, "rank %d: hako error %d: %s\n", rank, err.code, err.what() );
			status[rank] = 1;
		}
	};
	std::thread t1( rankMain, 1 );
	rankMain( 0 );
	t1.join();
	if( status[0] || status[1] ) return 5;
	double worst = 0.0;
	for( size_t i = 0; i < reduced.size(); i++ )
	{
		if( ( i & 3 ) == 3 )
		{
			if( reduced[i] != single[i] || reduced[i] != 16.0f * iterations ) return 6; // spp counts add up exactly
			continue;
		}
		const double den = std::fabs( single[i] ) > 1e-3 ? std::fabs( single[i] ) : 1e-3;
		worst = std::fmax( worst, std::fabs( reduced[i] - single[i] ) / den );
	}
	printf( "2 GPUs, spp-sharded + ncclReduce vs 1 GPU: worst relative difference %.3g\n", worst );
	return worst <= 1e-5 ? 0 : 7;
}

int main( int argc, char** argv )
{
	const int gridRes = 128, W = 256, H = 192;
	try
	{
		hako::Context ctx( 0 );
		std::vector<vec3> vertices, vcolors, vemissions;
		makeScene( vertices, vcolors, vemissions );
		const vec3 origin = { -1.5f, -1.0f, -1.5f };
		const float dps = 3.0f / gridRes;

		// ---- voxRTGPU: build + primary rays ----
		IntersectorOctreeGPU intersector( &ctx );
		intersector.build( vertices, vcolors, vemissions, (Shader*)nullptr, nullptr, origin, dps, gridRes );
		const float eye[3] = { 4, 4, 4 }, center[3] = { 0, 0, 0 }, up[3] = { 0, 1, 0 };
		hako_camera cam;
		hako::check( hako_camera_lookat( eye, center, up, 45.0f * 3.14159265f / 180.0f, W, H, 0.01f, 1000.0f, 6.0f, 0.02f, &cam ) );
		std::vector<uint8_t> image( (size_t)W * H * 4 );
		std::vector<uint32_t> ids( (size_t)W * H );
		intersector.render( nullptr, cam, W, H, true, image.data(), ids.data() );
		size_t hits = 0;
		for( uint32_t id : ids ) hits += id != 0xFFFFFFFFu;
		printf( "voxels %u nodes %u hits %zu hasEmission %d\n", intersector.m_numberOfVoxels, intersector.m_numberOfNodes, hits, (int)intersector.hasEmission() );

		// ---- a run of frames in one call (pipelined trace / download): frame 0 must equal the single-frame call ----
		{
			const int nViews = 5;
			std::vector<hako_camera> cams( nViews );
			for( int k = 0; k < nViews; k++ )
			{
				const float e[3] = { 4.0f - 0.5f * k, 4.0f, 4.0f + 0.5f * k };
				hako::check( hako_camera_lookat( e, center, up, 45.0f * 3.14159265f / 180.0f, W, H, 0.01f, 1000.0f, 6.0f, 0.02f, &cams[k] ) );
			}
			std::vector<uint8_t> frames( (size_t)nViews * W * H * 4 );
			std::vector<uint32_t> frameIds( (size_t)nViews * W * H );
			intersector.renderFrames( nullptr, cams.data(), nViews, W, H, true, frames.data(), frameIds.data() );
			if( memcmp( frames.data(), image.data(), image.size() ) != 0 || memcmp( frameIds.data(), ids.data(), ids.size() * 4 ) != 0 )
			{
				printf( "renderFrames: frame 0 differs from render()\n" );
				return 3;
			}
			size_t h4 = 0;
			for( size_t i = (size_t)4 * W * H; i < (size_t)5 * W * H; i++ ) h4 += frameIds[i] != 0xFFFFFFFFu;
			printf( "frames %d, hits in frame 4: %zu\n", nViews, h4 );
		}

		// ---- IntersectorOctree (CPU class surface): rebuild from the sorted unique Morton list, cast one ray ----
		std::vector<uint64_t> mortons;
		std::vector<hako_node68> nodes;
		intersector.download( &nodes, &mortons );
		IntersectorOctree cpuStyle( &ctx );
		cpuStyle.buildDAGReference( mortons, origin, dps, gridRes );
		const float ro[3] = { 4, 4, 4 }, rd[3] = { -1, -1, -1 };
		float t = FLT_MAX;
		int nMajor = -1;
		uint32_t vIndex = 0xFFFFFFFFu;
		cpuStyle.intersect( ro, rd, &t, &nMajor, &vIndex );
		printf( "dag nodes %u (gpu build %u) centre ray t %.6f nMajor %d vIndex %u\n", cpuStyle.m_numberOfNodes, (unsigned)nodes.size(), t, nMajor, vIndex );
		if( cpuStyle.m_numberOfNodes != nodes.size() || t == FLT_MAX ) return 2;

		// ---- voxPTGPU: setup, environment, scene, 4 steps (= 64 spp), resolve ----
		PathTracer pt( &ctx );
		pt.setup( nullptr );
		pt.resizeFrameBufferIfNeeded( nullptr, W, H );
		const int ew = 64, eh = 32;
		std::vector<float> env( (size_t)ew * eh * 4 );
		for( int y = 0; y < eh; y++ )
			for( int x = 0; x < ew; x++ )
			{
				float* p = &env[( (size_t)y * ew + x ) * 4];
				const float sky = 0.2f + 0.8f * ( 1.0f - (float)y / eh );
				p[0] = 0.6f * sky, p[1] = 0.8f * sky, p[2] = sky, p[3] = 1.0f;
			}
		pt.loadHDRI( nullptr, env.data(), ew, eh );
		pt.updateScene( vertices, vcolors, vemissions, nullptr, origin, dps, gridRes );
		for( int i = 0; i < 4; i++ ) pt.step( nullptr, cam );
		pt.toImage( nullptr, image.data() );
		std::vector<float> acc( (size_t)W * H * 4 );
		pt.readAccumulators( nullptr, acc.data() );
		double sum = 0;
		for( size_t i = 0; i < (size_t)W * H; i++ ) sum += acc[i * 4] + acc[i * 4 + 1] + acc[i * 4 + 2];
		printf( "steps %d spp %.0f mean radiance %.6f\n", pt.getSteps(), acc[3], sum / ( 3.0 * W * H * acc[3] ) );
		if( acc[3] != 64.0f || !( sum > 0 ) ) return 3;
		if( argc > 1 )
		{
			FILE* f = fopen( argv[1], "wb" );
			if( f )
			{
				fprintf( f, "P6\n%d %d\n255\n", W, H );
				for( size_t i = 0; i < (size_t)W * H; i++ ) fwrite( &image[i * 4], 1, 3, f );
				fclose( f );
			}
		}
		// ---- the same with the reference's own argument lists, as voxPTGPU.cpp:88-91,177-193 is written ----
		{
			PathTracer pt2; // default-constructed, voxPTGPU.cpp:88
			pt2.setup( nullptr, "../voxKernel.cu", "../", true );
			pt2.resizeFrameBufferIfNeeded( nullptr, W, H );
			pt2.loadHDRI( nullptr, env.data(), ew, eh );
			pt2.updateScene( vertices, vcolors, vemissions, nullptr, origin, dps, gridRes );
			Camera3D camera;
			static Image2DRGBA8 img;
			for( int i = 0; i < 4; i++ ) pt2.step( nullptr, camera, 6.0f, 0.02f );
			pt2.toImageAsync( nullptr, &img );
			hako::check( hako_ctx_synchronize( pt2.m_ctx->get(), nullptr ) ); // oroStreamSynchronize( stream ), voxPTGPU.cpp:189
			if( img.width() != W || img.height() != H || memcmp( img.data(), image.data(), image.size() ) != 0 )
			{
				printf( "reference-signature PathTracer calls: image differs from the hako_camera path\n" );
				return 8;
			}
			printf( "reference-signature calls ok (%d steps)\n", pt2.getSteps() );
		}
		{
			const int mg = multiGpuSection( vertices, vcolors, vemissions, origin, dps, gridRes, env, ew, eh, cam, W, H );
			if( mg ) return mg;
		}
		// error behaviour: gridRes must be a power of two (IntersectorOctreeGPU.hpp:48-51 abort()) -> hako::Error here
		try
		{
			intersector.build( vertices, vcolors, vemissions, (Shader*)nullptr, nullptr, origin, dps, 100 );
			return 4;
		}
		catch( const hako::Error& err )
		{
			printf( "expected error: %s\n", err.what() );
		}
		pt.cleanUp();
	}
	catch( const hako::Error& err )
	{
		fprintf( stderr, "hako error %d: %s\n", err.code, err.what() );
		return 1;
	}
	printf( "shim demo ok\n" );
	return 0;
}
