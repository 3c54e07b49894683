// render.cu -- primary-ray casting and batched intersection (the `render` kernel, voxKernel.cu:437-483, and the
// CPU renderLine loop, voxRT.cpp:307-347), plus conversion between the canonical 68-byte OctreeNode array
// (voxCommon.hpp:133-138) and the resident 64-byte trace layout.
#include <cstdlib>

#include "internal.cuh"
#include "traverse.cuh"

namespace hako
{
// ---- layout conversion ----
__global__ void __launch_bounds__( 256 ) importNodesKernel( const uint32_t* __restrict__ nodes68, uint32_t n, TraceNode* __restrict__ out, uint8_t* __restrict__ masks )
{
	// one thread per (node, slot): 17 words in, 16 words out per node
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	uint32_t node = i >> 3, c = i & 7u;
	if( node >= n ) return;
	const uint32_t* src = nodes68 + (uint64_t)node * 17u;
	out[node].slot[c] = make_uint2( src[1 + c], src[9 + c] );
	if( c == 0 ) masks[node] = (uint8_t)( src[0] & 0xFFu );
}
__global__ void __launch_bounds__( 256 ) exportNodesKernel( const TraceNode* __restrict__ nodes, const uint8_t* __restrict__ masks, uint32_t n, uint32_t* __restrict__ out68 )
{
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	uint32_t node = i >> 3, c = i & 7u;
	if( node >= n ) return;
	uint2 s = nodes[node].slot[c];
	uint32_t* dst = out68 + (uint64_t)node * 17u;
	dst[1 + c] = s.x;
	dst[9 + c] = s.y;
	if( c == 0 ) dst[0] = masks[node]; // padding bytes are written as zero
}
int traceImportNodes( hako_ctx* ctx, cudaStream_t s, const hako_node68* nodes68_d, uint32_t n, TraceNode* out, uint8_t* masks )
{
	if( n == 0 ) return 0;
	HAKO_LAUNCH( ctx, importNodesKernel, divUp( (uint64_t)n * 8, 256 ), 256, 0, s, (const uint32_t*)nodes68_d, n, out, masks );
	return 0;
}
int traceExportNodes( hako_ctx* ctx, cudaStream_t s, const TraceNode* nodes, const uint8_t* masks, uint32_t n, hako_node68* out68_d )
{
	if( n == 0 ) return 0;
	HAKO_LAUNCH( ctx, exportNodesKernel, divUp( (uint64_t)n * 8, 256 ), 256, 0, s, nodes, masks, n, (uint32_t*)out68_d );
	return 0;
}

// ---- batched intersection: IntersectorOctreeGPU::intersect over an array of rays ----
template <bool SHADOW>
__global__ void __launch_bounds__( 256 ) intersectBatchKernel( OctreeView v, const float* __restrict__ ro, const float* __restrict__ rd, uint64_t n,
															   float* __restrict__ t, int32_t* __restrict__ nMajor, uint32_t* __restrict__ vIndex )
{
	uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if( i >= n ) return;
	V3 o = { ro[i * 3], ro[i * 3 + 1], ro[i * 3 + 2] };
	V3 d = { rd[i * 3], rd[i * 3 + 1], rd[i * 3 + 2] };
	float tt;
	int nm;
	uint32_t vi;
	if( octreeTraverse<SHADOW>( v, o, d, &tt, &nm, &vi ) )
	{
		t[i] = tt;
		nMajor[i] = nm;
		vIndex[i] = vi;
	}
}
int launchIntersectBatch( hako_ctx* ctx, cudaStream_t s, const OctreeView& v, const float* ro, const float* rd, uint64_t n, int shadow,
						  float* t, int32_t* nMajor, uint32_t* vIndex )
{
	if( n == 0 ) return 0;
	auto kShadow = intersectBatchKernel<true>;
	auto kClosest = intersectBatchKernel<false>;
	if( shadow )
		HAKO_LAUNCH( ctx, kShadow, divUp( n, 256 ), 256, 0, s, v, ro, rd, n, t, nMajor, vIndex );
	else
		HAKO_LAUNCH( ctx, kClosest, divUp( n, 256 ), 256, 0, s, v, ro, rd, n, t, nMajor, vIndex );
	return 0;
}

// ---- primary rays ----
// One thread per pixel; a warp covers an 8x4 pixel tile (coherent rays walk the same nodes, and each output row
// segment is one full 32-byte sector); a block is WARPS_X x WARPS_Y such tiles.
// Measured alternatives (B200, bench workload, 4096^3 stand-in): persistent warps that refill finished lanes from a tile
// counter reach 4.1-6.3 Grays/s against 9.3 for this plain launch -- refilled lanes lose coherence with their warp, the
// resumable state costs 16 more registers, and the hardware CTA scheduler already balances 65 k warps well.
template <int WARPS_X, int WARPS_Y, int TOPK>
__global__ void __launch_bounds__( WARPS_X * WARPS_Y * 32 ) renderPrimaryKernel( OctreeView v, hako_camera cam, int w, int h, int y0, int y1, int showVertexColor,
															  uchar4* __restrict__ rgba, float* __restrict__ tOut, int32_t* __restrict__ nMajorOut, uint32_t* __restrict__ vIndexOut )
{
	// TOPK > 0: the top TOPK nodes of the tree are copied to shared memory by every CTA (A/B variant, HAKO_SMEM_TOP)
	__shared__ uint2 s_top[TOPK > 0 ? TOPK * 8 : 1];
	TopCache tc{ nullptr, 0xFFFFFFFFu };
	if( TOPK > 0 ) tc = stageTop<( TOPK > 0 ? TOPK : 1 )>( v, s_top );
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	const int x = blockIdx.x * ( WARPS_X * 8 ) + ( warp % WARPS_X ) * 8 + ( lane & 7u );
	const int y = y0 + blockIdx.y * ( WARPS_Y * 4 ) + ( warp / WARPS_X ) * 4 + ( lane >> 3 );
	if( x >= w || y >= y1 ) return;

	V3 ro, rd;
	cameraShoot( cam, x, y, 0.5f, 0.5f, w, h, &ro, &rd );
	float t = HAKO_MAX_FLOAT;
	int nMajor = -1;
	uint32_t vIndex = 0xFFFFFFFFu;
	const bool hit = octreeTraverseImpl<( TOPK > 0 )>( v, ro, rd, false, &t, &nMajor, &vIndex, tc );

	const uint64_t p = (uint64_t)( y - y0 ) * w + x;
	// streaming stores: the outputs are write-once and must not evict octree nodes from L2
	if( tOut ) __stcs( tOut + p, t );
	if( nMajorOut ) __stcs( nMajorOut + p, nMajor );
	if( vIndexOut ) __stcs( vIndexOut + p, vIndex );
	if( rgba )
	{
		uchar4 c = make_uchar4( 0, 0, 0, 255 );
		if( hit )
		{
			if( showVertexColor )
			{
				const hako_voxel_attrib a = v.attrs[vIndex];
				c = make_uchar4( a.color[0], a.color[1], a.color[2], a.color[3] );
			}
			else
			{
				// voxKernel.cu:470-476
				V3 n = getHitN( nMajor, rd );
				V3 col = ( n + V3{ 1.0f, 1.0f, 1.0f } ) * 0.5f;
				c = make_uchar4( (uint8_t)( 255 * col.x + 0.5f ), (uint8_t)( 255 * col.y + 0.5f ), (uint8_t)( 255 * col.z + 0.5f ), 255 );
			}
		}
		__stcs( rgba + p, c );
	}
}
int launchRenderPrimary( hako_ctx* ctx, cudaStream_t s, const OctreeView& v, const hako_camera& cam, int w, int h, int y0, int y1,
						 int showVertexColor, uint8_t* rgba, float* t, int32_t* nMajor, uint32_t* vIndex )
{
	if( w <= 0 || y1 <= y0 ) return 0;
	// block shape: 2x4 warps = 16x16 pixels; 128/64/32-thread blocks measured within 1 % of this (the kernel is issue-bound, not occupancy-bound)
	dim3 grid( divUp( w, 16 ), divUp( y1 - y0, 16 ) );
	// HAKO_SMEM_TOP = 64 / 512: stage the top of the tree in shared memory (north_star's design point; measured, not the default:
	// DESIGN.md section 7a)
	static const int topEnv = getenv( "HAKO_SMEM_TOP" ) ? atoi( getenv( "HAKO_SMEM_TOP" ) ) : 0;
	auto k = renderPrimaryKernel<2, 4, 0>;
	if( topEnv == 64 ) k = renderPrimaryKernel<2, 4, 64>;
	if( topEnv == 512 ) k = renderPrimaryKernel<2, 4, 512>;
	HAKO_LAUNCH( ctx, k, grid, 256, 0, s, v, cam, w, h, y0, y1, showVertexColor, (uchar4*)rgba, t, nMajor, vIndex );
	return 0;
}
} // namespace hako
