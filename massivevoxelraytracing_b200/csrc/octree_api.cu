// octree_api.cu -- C ABI for the octree object: import / export / info / intersect / primary rays.
// (IntersectorOctreeGPU.hpp:21-275 and IntersectorOctree.hpp:214-266 are the reference surfaces.)
#include <cstdlib>
#include <cstring>
#include <vector>

#include "internal.cuh"

using namespace hako;

#define CHECK_ARG( cond, msg )      \
	do                              \
	{                               \
		if( !( cond ) )             \
		{                           \
			hako::setError( "%s", msg ); \
			return HAKO_E_INVALID;  \
		}                           \
	} while( 0 )

extern "C" {

int hako_octree_import( hako_ctx* ctx, hako_stream stream, const hako_node68* nodes, uint32_t nNodes,
						const hako_voxel_attrib* attrs, uint32_t nVoxels, const float lower[3], const float upper[3], float dps,
						hako_octree** out )
{
	CHECK_ARG( ctx && out && lower && upper, "hako_octree_import: null argument" );
	CHECK_ARG( nNodes == 0 || nodes, "hako_octree_import: nodes is null" );
	if( nNodes >= 0xFFFFFFu )
	{
		hako::setError( "hako_octree_import: %u nodes exceed the 24-bit node index (voxCommon.hpp:183-195)", nNodes );
		return HAKO_E_LIMIT;
	}
	// The traversal indexes nodes[child & 0xFFFFFF] and pushes one stack entry per level without bounds checks, so a tree from
	// outside is validated before it is uploaded: every child index is smaller than its parent's (the bottom-up numbering of
	// IntersectorOctree.hpp:11-123; it also rules out cycles), the mask embedded in a child pointer is the child's own mask byte
	// (voxCommon.hpp:183-195), a present child has its bit in the parent's mask, and the tree is at most 21 levels deep.
	int depthOfRoot = 0;
	{
		std::vector<uint8_t> depth( nNodes, 1 );
		for( uint32_t i = 0; i < nNodes; i++ )
		{
			uint32_t dmax = 0;
			for( int c = 0; c < 8; c++ )
			{
				const uint32_t ch = nodes[i].children[c];
				if( ch == 0xFFFFFFFFu ) continue;
				const uint32_t idx = ch & 0xFFFFFFu;
				if( idx >= i )
				{
					hako::setError( "hako_octree_import: node %u child %d points at node %u (children must precede their parent)", i, c, idx );
					return HAKO_E_INVALID;
				}
				if( ( ch >> 24 ) != nodes[idx].mask )
				{
					hako::setError( "hako_octree_import: node %u child %d carries mask 0x%02x but node %u has mask 0x%02x (embedMasks, voxCommon.hpp:183-195)", i, c, ch >> 24, idx,
									nodes[idx].mask );
					return HAKO_E_INVALID;
				}
				if( !( ( nodes[i].mask >> c ) & 1u ) )
				{
					hako::setError( "hako_octree_import: node %u has child %d but its mask 0x%02x does not", i, c, nodes[i].mask );
					return HAKO_E_INVALID;
				}
				dmax = depth[idx] > dmax ? depth[idx] : dmax;
			}
			if( dmax + 1 > 21 )
			{
				hako::setError( "hako_octree_import: the tree is deeper than 21 levels at node %u (morton.hpp:73: 21 bits per axis)", i );
				return HAKO_E_INVALID;
			}
			depth[i] = (uint8_t)( dmax + 1 );
		}
		if( nNodes ) depthOfRoot = depth[nNodes - 1];
	}
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	hako_octree* o = new hako_octree();
	o->ctx = ctx;
	o->numberOfNodes = nNodes;
	o->numberOfVoxels = nVoxels;
	o->gridRes = depthOfRoot ? ( 1 << depthOfRoot ) : 0; // levels below the root = log2( gridRes )
	o->lower = { lower[0], lower[1], lower[2] };
	o->upper = { upper[0], upper[1], upper[2] };
	o->dps = dps;
	int r = 0;
	hako_node68* staged = nullptr;
	do
	{
		if( ( r = dalloc( ctx, &o->nodes, nNodes, s ) ) ) break;
		if( ( r = dalloc( ctx, &o->masks, nNodes, s ) ) ) break;
		if( ( r = dalloc( ctx, &staged, nNodes, s ) ) ) break;
		if( nNodes )
		{
			cudaError_t e = cudaMemcpyAsync( staged, nodes, (size_t)nNodes * sizeof( hako_node68 ), cudaMemcpyHostToDevice, s );
			if( e != cudaSuccess )
			{
				r = cudaFail( e, "H2D nodes", __FILE__, __LINE__ );
				break;
			}
			if( ( r = traceImportNodes( ctx, s, staged, nNodes, o->nodes, o->masks ) ) ) break;
		}
		if( attrs && nVoxels )
		{
			if( ( r = dalloc( ctx, &o->attrs, nVoxels, s ) ) ) break;
			cudaError_t e = cudaMemcpyAsync( o->attrs, attrs, (size_t)nVoxels * sizeof( hako_voxel_attrib ), cudaMemcpyHostToDevice, s );
			if( e != cudaSuccess )
			{
				r = cudaFail( e, "H2D attrs", __FILE__, __LINE__ );
				break;
			}
			for( uint32_t i = 0; i < nVoxels && !o->hasEmission; i++ )
				o->hasEmission = ( attrs[i].emission[0] | attrs[i].emission[1] | attrs[i].emission[2] ) != 0;
		}
	} while( 0 );
	dfree( staged, s );
	if( r == 0 )
	{
		cudaError_t e = cudaStreamSynchronize( s ); // the host buffers may be reused by the caller on return
		if( e != cudaSuccess ) r = cudaFail( e, "sync", __FILE__, __LINE__ );
	}
	if( r )
	{
		hako_octree_destroy( o );
		return r;
	}
	*out = o;
	return HAKO_OK;
}

int hako_octree_info_get( const hako_octree* o, hako_octree_info* out )
{
	CHECK_ARG( o && out, "hako_octree_info_get: null argument" );
	out->numberOfNodes = o->numberOfNodes;
	out->numberOfVoxels = o->numberOfVoxels;
	out->lower[0] = o->lower.x, out->lower[1] = o->lower.y, out->lower[2] = o->lower.z;
	out->upper[0] = o->upper.x, out->upper[1] = o->upper.y, out->upper[2] = o->upper.z;
	out->dps = o->dps;
	out->emissionScale = o->emissionScale;
	out->hasEmission = o->hasEmission;
	out->gridRes = o->gridRes;
	out->deviceBytes = (uint64_t)o->numberOfNodes * ( sizeof( TraceNode ) + 1 ) + ( o->attrs ? (uint64_t)o->numberOfVoxels * 8 : 0 ) +
					   ( o->mortons ? (uint64_t)o->numberOfVoxels * 8 : 0 );
	out->wideIndex = (uint32_t)o->wide;
	out->reserved = 0;
	return HAKO_OK;
}

int hako_octree_export( const hako_octree* o, hako_stream stream, hako_node68* nodes, uint64_t* mortons, hako_voxel_attrib* attrs )
{
	CHECK_ARG( o, "hako_octree_export: null octree" );
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( o->ctx->device ) );
	if( nodes && o->numberOfNodes && o->wide )
	{
		hako::setError( "hako_octree_export: a wide-index tree (%u nodes) has no representation in the reference's 68-byte layout (24-bit child index)", o->numberOfNodes );
		return HAKO_E_LIMIT;
	}
	if( nodes && o->numberOfNodes )
	{
		hako_node68* staged = nullptr;
		HAKO_TRY( dalloc( o->ctx, &staged, o->numberOfNodes, s ) );
		int r = traceExportNodes( o->ctx, s, o->nodes, o->masks, o->numberOfNodes, staged );
		if( r == 0 )
		{
			cudaError_t e = cudaMemcpyAsync( nodes, staged, (size_t)o->numberOfNodes * sizeof( hako_node68 ), cudaMemcpyDeviceToHost, s );
			if( e != cudaSuccess ) r = cudaFail( e, "D2H nodes", __FILE__, __LINE__ );
		}
		dfree( staged, s );
		HAKO_TRY( r );
	}
	if( mortons && o->numberOfVoxels )
	{
		if( !o->mortons )
		{
			hako::setError( "hako_octree_export: this octree holds no Morton keys (imported tree)" );
			return HAKO_E_STATE;
		}
		HAKO_CUDA( cudaMemcpyAsync( mortons, o->mortons, (size_t)o->numberOfVoxels * 8, cudaMemcpyDeviceToHost, s ) );
	}
	if( attrs && o->numberOfVoxels )
	{
		if( !o->attrs )
		{
			hako::setError( "hako_octree_export: this octree holds no voxel attributes" );
			return HAKO_E_STATE;
		}
		HAKO_CUDA( cudaMemcpyAsync( attrs, o->attrs, (size_t)o->numberOfVoxels * 8, cudaMemcpyDeviceToHost, s ) );
	}
	HAKO_CUDA( cudaStreamSynchronize( s ) );
	return HAKO_OK;
}

void hako_octree_destroy( hako_octree* o )
{
	if( !o ) return;
	cudaSetDevice( o->ctx->device );
	// plain cudaFree is legal for pool allocations and synchronises, so no in-flight kernel loses its tree
	if( o->nodes ) cudaFree( o->nodes );
	if( o->masks ) cudaFree( o->masks );
	if( o->attrs ) cudaFree( o->attrs );
	if( o->mortons ) cudaFree( o->mortons );
	delete o;
}

// the same, stream-ordered: the arrays go back to the pool behind the work already queued on `stream` (no device-wide wait)
void hako_octree_destroy_async( hako_octree* o, hako_stream stream )
{
	if( !o ) return;
	cudaSetDevice( o->ctx->device );
	cudaStream_t s = (cudaStream_t)stream;
	dfree( o->nodes, s ), dfree( o->masks, s ), dfree( o->attrs, s ), dfree( o->mortons, s );
	delete o;
}

int hako_octree_device_arrays( const hako_octree* o, void** nodes_d, uint64_t* nodeBytes, void** masks_d, uint64_t* maskBytes,
							   void** attrs_d, uint64_t* attrBytes )
{
	CHECK_ARG( o, "hako_octree_device_arrays: null octree" );
	if( nodes_d ) *nodes_d = o->nodes;
	if( nodeBytes ) *nodeBytes = (uint64_t)o->numberOfNodes * sizeof( TraceNode );
	if( masks_d ) *masks_d = o->masks;
	if( maskBytes ) *maskBytes = o->numberOfNodes;
	if( attrs_d ) *attrs_d = o->attrs;
	if( attrBytes ) *attrBytes = o->attrs ? (uint64_t)o->numberOfVoxels * 8 : 0;
	return HAKO_OK;
}

int hako_octree_create_replica( hako_ctx* ctx, const hako_octree_info* info, hako_octree** out )
{
	CHECK_ARG( ctx && info && out, "hako_octree_create_replica: null argument" );
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	hako_octree* o = new hako_octree();
	o->ctx = ctx;
	o->numberOfNodes = info->numberOfNodes;
	o->numberOfVoxels = info->numberOfVoxels;
	o->lower = { info->lower[0], info->lower[1], info->lower[2] };
	o->upper = { info->upper[0], info->upper[1], info->upper[2] };
	o->dps = info->dps;
	o->emissionScale = info->emissionScale;
	o->hasEmission = info->hasEmission;
	o->gridRes = info->gridRes;
	o->wide = (int)info->wideIndex;
	int r = dalloc( ctx, &o->nodes, o->numberOfNodes, 0 );
	if( !r ) r = dalloc( ctx, &o->masks, o->numberOfNodes, 0 );
	if( !r ) r = dalloc( ctx, &o->attrs, o->numberOfVoxels, 0 );
	if( !r )
	{
		cudaError_t e = cudaStreamSynchronize( 0 );
		if( e != cudaSuccess ) r = cudaFail( e, "sync", __FILE__, __LINE__ );
	}
	if( r )
	{
		hako_octree_destroy( o );
		return r;
	}
	*out = o;
	return HAKO_OK;
}

int hako_intersect_batch_device( const hako_octree* o, hako_stream stream, const float* ro_d, const float* rd_d, uint64_t n, int isShadowRay,
								 float* t_d, int32_t* nMajor_d, uint32_t* vIndex_d )
{
	CHECK_ARG( o && ( n == 0 || ( ro_d && rd_d && t_d && nMajor_d && vIndex_d ) ), "hako_intersect_batch_device: null argument" );
	HAKO_CUDA( cudaSetDevice( o->ctx->device ) );
	return launchIntersectBatch( o->ctx, (cudaStream_t)stream, o->view(), ro_d, rd_d, n, isShadowRay, t_d, nMajor_d, vIndex_d );
}

int hako_intersect_batch( const hako_octree* o, hako_stream stream, const float* ro, const float* rd, uint64_t n, int isShadowRay,
						  float* t, int32_t* nMajor, uint32_t* vIndex )
{
	CHECK_ARG( o && ( n == 0 || ( ro && rd && t && nMajor && vIndex ) ), "hako_intersect_batch: null argument" );
	if( n == 0 ) return HAKO_OK;
	hako_ctx* ctx = o->ctx;
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	float *ro_d = nullptr, *rd_d = nullptr, *t_d = nullptr;
	int32_t* nm_d = nullptr;
	uint32_t* vi_d = nullptr;
	int r = 0;
	do
	{
		if( ( r = dalloc( ctx, &ro_d, n * 3, s ) ) ) break;
		if( ( r = dalloc( ctx, &rd_d, n * 3, s ) ) ) break;
		if( ( r = dalloc( ctx, &t_d, n, s ) ) ) break;
		if( ( r = dalloc( ctx, &nm_d, n, s ) ) ) break;
		if( ( r = dalloc( ctx, &vi_d, n, s ) ) ) break;
		cudaError_t e;
#define CP( dst, src, bytes, kind )                                                   \
	if( ( e = cudaMemcpyAsync( dst, src, bytes, kind, s ) ) != cudaSuccess )          \
	{                                                                                 \
		r = cudaFail( e, "cudaMemcpyAsync " #dst, __FILE__, __LINE__ );               \
		break;                                                                        \
	}
		CP( ro_d, ro, n * 12, cudaMemcpyHostToDevice );
		CP( rd_d, rd, n * 12, cudaMemcpyHostToDevice );
		// a miss leaves the caller's preset values untouched, so they travel too
		CP( t_d, t, n * 4, cudaMemcpyHostToDevice );
		CP( nm_d, nMajor, n * 4, cudaMemcpyHostToDevice );
		CP( vi_d, vIndex, n * 4, cudaMemcpyHostToDevice );
		if( ( r = launchIntersectBatch( ctx, s, o->view(), ro_d, rd_d, n, isShadowRay, t_d, nm_d, vi_d ) ) ) break;
		CP( t, t_d, n * 4, cudaMemcpyDeviceToHost );
		CP( nMajor, nm_d, n * 4, cudaMemcpyDeviceToHost );
		CP( vIndex, vi_d, n * 4, cudaMemcpyDeviceToHost );
		if( ( e = cudaStreamSynchronize( s ) ) != cudaSuccess ) r = cudaFail( e, "sync", __FILE__, __LINE__ );
	} while( 0 );
	dfree( ro_d, s ), dfree( rd_d, s ), dfree( t_d, s ), dfree( nm_d, s ), dfree( vi_d, s );
	return r;
}

int hako_render_primary_device( const hako_octree* o, hako_stream stream, const hako_camera* cam, int w, int h, int y0, int y1,
								int showVertexColor, uint8_t* rgba_d, float* t_d, int32_t* nMajor_d, uint32_t* vIndex_d )
{
	CHECK_ARG( o && cam, "hako_render_primary_device: null argument" );
	CHECK_ARG( w > 0 && h > 0 && 0 <= y0 && y0 <= y1 && y1 <= h, "hako_render_primary_device: bad image / row range" );
	CHECK_ARG( !( showVertexColor && rgba_d && !o->attrs ), "hako_render_primary_device: showVertexColor needs voxel attributes" );
	HAKO_CUDA( cudaSetDevice( o->ctx->device ) );
	return launchRenderPrimary( o->ctx, (cudaStream_t)stream, o->view(), *cam, w, h, y0, y1, showVertexColor, rgba_d, t_d, nMajor_d, vIndex_d );
}

int hako_render_primary( const hako_octree* o, hako_stream stream, const hako_camera* cam, int w, int h, int y0, int y1,
						 int showVertexColor, uint8_t* rgba, float* t, int32_t* nMajor, uint32_t* vIndex )
{
	CHECK_ARG( o && cam, "hako_render_primary: null argument" );
	CHECK_ARG( w > 0 && h > 0 && 0 <= y0 && y0 <= y1 && y1 <= h, "hako_render_primary: bad image / row range" );
	CHECK_ARG( !( showVertexColor && rgba && !o->attrs ), "hako_render_primary: showVertexColor needs voxel attributes" );
	hako_ctx* ctx = o->ctx;
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	const uint64_t n = (uint64_t)w * ( y1 - y0 );
	if( n == 0 ) return HAKO_OK;
	uint8_t* rgba_d = nullptr;
	float* t_d = nullptr;
	int32_t* nm_d = nullptr;
	uint32_t* vi_d = nullptr;
	int r = 0;
	do
	{
		if( rgba && ( r = dalloc( ctx, &rgba_d, n * 4, s ) ) ) break;
		if( t && ( r = dalloc( ctx, &t_d, n, s ) ) ) break;
		if( nMajor && ( r = dalloc( ctx, &nm_d, n, s ) ) ) break;
		if( vIndex && ( r = dalloc( ctx, &vi_d, n, s ) ) ) break;
		// Row bands: band k is copied to the host on ctx->copyStream while later bands are being traced, so the call costs
		// about max( trace, D2H ) instead of their sum.  Band kernels alternate between two internal streams: a band is a
		// single wave of CTAs whose tail would otherwise serialise with the next band.
		static const int bandsEnv = getenv( "HAKO_BANDS" ) ? atoi( getenv( "HAKO_BANDS" ) ) : 0;
		int bands = bandsEnv > 0 ? bandsEnv : ( y1 - y0 ) / 256;
		bands = bands < 1 ? 1 : ( bands > 16 ? 16 : bands );
		const int rowsPerBand = ( ( y1 - y0 + bands - 1 ) / bands + 15 ) / 16 * 16; // whole 16-row tiles
		cudaError_t e = cudaEventRecord( ctx->bandEv[16], s );						 // the device buffers were allocated on `s`
		for( int i = 0; i < 2 && e == cudaSuccess; i++ ) e = cudaStreamWaitEvent( ctx->bandStream[i], ctx->bandEv[16], 0 );
		int k = 0;
		for( int b0 = y0; b0 < y1 && r == 0 && e == cudaSuccess; b0 += rowsPerBand, k++ )
		{
			const int b1 = b0 + rowsPerBand < y1 ? b0 + rowsPerBand : y1;
			const uint64_t off = (uint64_t)( b0 - y0 ) * w, cnt = (uint64_t)( b1 - b0 ) * w;
			cudaStream_t bs = ctx->bandStream[k & 1];
			if( ( r = launchRenderPrimary( ctx, bs, o->view(), *cam, w, h, b0, b1, showVertexColor, rgba_d ? rgba_d + off * 4 : nullptr, t_d ? t_d + off : nullptr,
										   nm_d ? nm_d + off : nullptr, vi_d ? vi_d + off : nullptr ) ) )
				break;
			if( ( e = cudaEventRecord( ctx->bandEv[k], bs ) ) != cudaSuccess ) break;
			if( ( e = cudaStreamWaitEvent( ctx->copyStream, ctx->bandEv[k], 0 ) ) != cudaSuccess ) break;
			if( rgba && ( e = cudaMemcpyAsync( rgba + off * 4, rgba_d + off * 4, cnt * 4, cudaMemcpyDeviceToHost, ctx->copyStream ) ) != cudaSuccess ) break;
			if( t && ( e = cudaMemcpyAsync( t + off, t_d + off, cnt * 4, cudaMemcpyDeviceToHost, ctx->copyStream ) ) != cudaSuccess ) break;
			if( nMajor && ( e = cudaMemcpyAsync( nMajor + off, nm_d + off, cnt * 4, cudaMemcpyDeviceToHost, ctx->copyStream ) ) != cudaSuccess ) break;
			if( vIndex && ( e = cudaMemcpyAsync( vIndex + off, vi_d + off, cnt * 4, cudaMemcpyDeviceToHost, ctx->copyStream ) ) != cudaSuccess ) break;
		}
		// every band kernel is ordered before its copy, so the last copy closes the call: ONE host wait on that event; the
		// frees below are stream-ordered on `s`, which is made to wait for the same event
		if( e == cudaSuccess ) e = cudaEventRecord( ctx->bandEv[16], ctx->copyStream );
		if( e == cudaSuccess ) e = cudaStreamWaitEvent( s, ctx->bandEv[16], 0 );
		if( e == cudaSuccess ) e = cudaEventSynchronize( ctx->bandEv[16] );
		if( e != cudaSuccess && r == 0 ) r = cudaFail( e, "band pipeline", __FILE__, __LINE__ );
		if( r )
		{
			// error path: drain everything before the buffers go back to the pool
			for( int i = 0; i < 2; i++ ) cudaStreamSynchronize( ctx->bandStream[i] );
			cudaStreamSynchronize( ctx->copyStream );
			cudaStreamSynchronize( s );
		}
	} while( 0 );
#undef CP
	dfree( rgba_d, s ), dfree( t_d, s ), dfree( nm_d, s ), dfree( vi_d, s );
	return r;
}

// A run of frames (the orbit loop of voxRTGPU.cpp:150-208, headless): frame i is traced while frame i-1 is on its way to the
// host.  A ring of RING device frames; frame kernels alternate between the two band streams, copies are queued on the copy
// stream, and the host waits once, for the last copy.
int hako_render_primary_views( const hako_octree* o, hako_stream stream, const hako_camera* cams, int nViews, int w, int h,
							   int showVertexColor, uint8_t* rgba, float* t, int32_t* nMajor, uint32_t* vIndex )
{
	CHECK_ARG( o && ( cams || nViews == 0 ), "hako_render_primary_views: null argument" );
	CHECK_ARG( w > 0 && h > 0 && nViews >= 0, "hako_render_primary_views: bad image / view count" );
	CHECK_ARG( !( showVertexColor && rgba && !o->attrs ), "hako_render_primary_views: showVertexColor needs voxel attributes" );
	if( nViews == 0 ) return HAKO_OK;
	hako_ctx* ctx = o->ctx;
	cudaStream_t s = (cudaStream_t)stream;
	HAKO_CUDA( cudaSetDevice( ctx->device ) );
	const uint64_t n = (uint64_t)w * h;
	enum { RING = 4 };
	const int ring = nViews < RING ? nViews : RING;
	uint8_t* rgba_d = nullptr;
	float* t_d = nullptr;
	int32_t* nm_d = nullptr;
	uint32_t* vi_d = nullptr;
	cudaEvent_t traced[RING] = {}, copied[RING] = {};
	int r = 0;
	cudaError_t e = cudaSuccess;
	do
	{
		if( rgba && ( r = dalloc( ctx, &rgba_d, n * 4 * ring, s ) ) ) break;
		if( t && ( r = dalloc( ctx, &t_d, n * ring, s ) ) ) break;
		if( nMajor && ( r = dalloc( ctx, &nm_d, n * ring, s ) ) ) break;
		if( vIndex && ( r = dalloc( ctx, &vi_d, n * ring, s ) ) ) break;
		for( int i = 0; i < ring && e == cudaSuccess; i++ )
		{
			e = cudaEventCreateWithFlags( &traced[i], cudaEventDisableTiming );
			if( e == cudaSuccess ) e = cudaEventCreateWithFlags( &copied[i], cudaEventDisableTiming );
		}
		if( e == cudaSuccess ) e = cudaEventRecord( ctx->bandEv[16], s ); // the device frames were allocated on `s`
		for( int i = 0; i < 2 && e == cudaSuccess; i++ ) e = cudaStreamWaitEvent( ctx->bandStream[i], ctx->bandEv[16], 0 );
		for( int v = 0; v < nViews && r == 0 && e == cudaSuccess; v++ )
		{
			const int slot = v % ring;
			const uint64_t dOff = (uint64_t)slot * n, hOff = (uint64_t)v * n;
			cudaStream_t bs = ctx->bandStream[v & 1];
			if( v >= ring && ( e = cudaStreamWaitEvent( bs, copied[slot], 0 ) ) != cudaSuccess ) break; // the slot's previous frame has left
			if( ( r = launchRenderPrimary( ctx, bs, o->view(), cams[v], w, h, 0, h, showVertexColor, rgba_d ? rgba_d + dOff * 4 : nullptr, t_d ? t_d + dOff : nullptr,
										   nm_d ? nm_d + dOff : nullptr, vi_d ? vi_d + dOff : nullptr ) ) )
				break;
			if( ( e = cudaEventRecord( traced[slot], bs ) ) != cudaSuccess ) break;
			if( ( e = cudaStreamWaitEvent( ctx->copyStream, traced[slot], 0 ) ) != cudaSuccess ) break;
			if( rgba && ( e = cudaMemcpyAsync( rgba + hOff * 4, rgba_d + dOff * 4, n * 4, cudaMemcpyDeviceToHost, ctx->copyStream ) ) != cudaSuccess ) break;
			if( t && ( e = cudaMemcpyAsync( t + hOff, t_d + dOff, n * 4, cudaMemcpyDeviceToHost, ctx->copyStream ) ) != cudaSuccess ) break;
			if( nMajor && ( e = cudaMemcpyAsync( nMajor + hOff, nm_d + dOff, n * 4, cudaMemcpyDeviceToHost, ctx->copyStream ) ) != cudaSuccess ) break;
			if( vIndex && ( e = cudaMemcpyAsync( vIndex + hOff, vi_d + dOff, n * 4, cudaMemcpyDeviceToHost, ctx->copyStream ) ) != cudaSuccess ) break;
			if( ( e = cudaEventRecord( copied[slot], ctx->copyStream ) ) != cudaSuccess ) break;
		}
		if( e == cudaSuccess ) e = cudaEventRecord( ctx->bandEv[16], ctx->copyStream );
		if( e == cudaSuccess ) e = cudaStreamWaitEvent( s, ctx->bandEv[16], 0 );
		if( e == cudaSuccess ) e = cudaEventSynchronize( ctx->bandEv[16] );
		if( e != cudaSuccess && r == 0 ) r = cudaFail( e, "frame pipeline", __FILE__, __LINE__ );
		if( r )
		{
			for( int i = 0; i < 2; i++ ) cudaStreamSynchronize( ctx->bandStream[i] );
			cudaStreamSynchronize( ctx->copyStream );
			cudaStreamSynchronize( s );
		}
	} while( 0 );
	for( int i = 0; i < RING; i++ )
	{
		if( traced[i] ) cudaEventDestroy( traced[i] );
		if( copied[i] ) cudaEventDestroy( copied[i] );
	}
	dfree( rgba_d, s ), dfree( t_d, s ), dfree( nm_d, s ), dfree( vi_d, s );
	return r;
}
}
