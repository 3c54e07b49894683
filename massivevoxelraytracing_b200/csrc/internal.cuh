// internal.cuh -- host-side objects behind the opaque handles of include/hako.h, and the launch plumbing.
#pragma once

#include <atomic>
#include <functional>
#include <vector>

#include "common.cuh"

namespace hako
{
struct Stager;
}

struct hako_ctx
{
	int device = 0;
	int smCount = 148;
	size_t l2Bytes = 0;
	cudaMemPool_t pool = nullptr;
	std::atomic<uint64_t> launches{ 0 };
	hako_build_stats stats{};
	bool statsPending = false; // the stage times of the last build are still in the events ev[0..5]
	// voxelizer: voxels of the LARGEST 32-triangle warp in the last build of an equally large mesh (sizes the next build's record slabs)
	uint32_t lastVoxTriangles = 0;
	int lastVoxGrid = 0;
	double lastVoxPerWarp = 0.0;
	uint64_t voxFreeBytes = 0; // free device memory as last queried by the voxelizer (cudaMemGetInfo costs 0.1 - 3 ms and more: asked once per mesh, not per build)
	cudaEvent_t ev[8] = {};
	// pinned staging word(s) for the few device->host counter reads a build needs
	uint64_t* pinned = nullptr; // 64 x u64
	// device->host pipeline of the host-buffer entry points: band k is copied on copyStream while band k+1 is traced
	cudaStream_t copyStream = nullptr;
	cudaStream_t bandStream[2] = {}; // band kernels alternate between these so that one band's tail overlaps the next band
	cudaEvent_t bandEv[17] = {};
	// host->device staging of pageable caller memory (h2d.cu): HAKO_H2D_WORKERS threads, each with two pinned chunks and a stream
	hako::Stager* stager = nullptr;
};

// sorted dumped voxels of one shard (hako_voxelize_sorted)
struct hako_pairs
{
	hako_ctx* ctx = nullptr;
	uint64_t* keys = nullptr;
	uint64_t* values = nullptr;
	uint64_t n = 0;
};

struct hako_octree
{
	hako_ctx* ctx = nullptr;
	hako::TraceNode* nodes = nullptr;
	uint8_t* masks = nullptr;
	hako_voxel_attrib* attrs = nullptr;
	uint64_t* mortons = nullptr; // sorted unique keys (kept for export / parity), may be null for imported trees
	uint32_t numberOfNodes = 0;
	uint32_t numberOfVoxels = 0;
	hako::V3 lower{}, upper{};
	float dps = 0.0f;
	float emissionScale = 7.5f; // IntersectorOctreeGPU.hpp:273
	uint32_t hasEmission = 0;
	int gridRes = 0;
	// 1: plain 32-bit child indices, the child's mask comes from `masks` (trees of >= 2^24 - 1 nodes, which the reference's
	// index | mask << 24 child pointers cannot express); 0: the reference layout
	int wide = 0;

	hako::OctreeView view() const
	{
		hako::OctreeView v;
		v.nodes = nodes;
		v.masks = masks;
		v.attrs = attrs;
		v.numberOfNodes = numberOfNodes;
		v.numberOfVoxels = numberOfVoxels;
		v.lower = lower;
		v.upper = upper;
		v.emissionScale = emissionScale;
		v.hasEmission = hasEmission;
		v.wide = (uint32_t)wide;
		return v;
	}
};

namespace hako
{
// stream-ordered allocation from the context's pool (replaces the ~12 oroMalloc/oroFree per build,
// IntersectorOctreeGPU.hpp:94-228; the pool keeps freed blocks so steady-state builds do not hit the driver)
template <class T>
inline int dalloc( hako_ctx* ctx, T** p, uint64_t count, cudaStream_t s )
{
	*p = nullptr;
	uint64_t bytes = count * sizeof( T );
	if( bytes == 0 ) bytes = 16;
	HAKO_CUDA( cudaMallocFromPoolAsync( (void**)p, bytes, ctx->pool, s ) );
	return 0;
}
template <class T>
inline void dfree( T* p, cudaStream_t s )
{
	if( p ) cudaFreeAsync( (void*)p, s );
}

// every kernel of the library goes through this so hako_ctx_launch_count() is a true count
#define HAKO_LAUNCH( ctx, kernel, grid, block, smem, stream, ... )              \
	do                                                                          \
	{                                                                           \
		kernel<<<( grid ), ( block ), ( smem ), ( stream )>>>( __VA_ARGS__ );   \
		( ctx )->launches.fetch_add( 1, std::memory_order_relaxed );            \
		cudaError_t le__ = cudaGetLastError();                                  \
		if( le__ != cudaSuccess ) return hako::cudaFail( le__, #kernel, __FILE__, __LINE__ ); \
	} while( 0 )

// the same inside a `do { ... } while( 0 )` body (or a loop guarded by r == 0) whose owner releases resources afterwards: a launch
// error sets r and leaves the body instead of returning past the clean-up.  One statement, safe as an unbraced if / else arm.
#define HAKO_LAUNCH_R( ctx, r, kernel, grid, block, smem, stream, ... )                             \
	if( ( r = [&]() -> int {                                                                        \
			kernel<<<( grid ), ( block ), ( smem ), ( stream )>>>( __VA_ARGS__ );                   \
			( ctx )->launches.fetch_add( 1, std::memory_order_relaxed );                            \
			const cudaError_t le__ = cudaGetLastError();                                            \
			return le__ == cudaSuccess ? 0 : hako::cudaFail( le__, #kernel, __FILE__, __LINE__ );   \
		}() ) != 0 )                                                                                \
		break;                                                                                      \
	else                                                                                            \
		( (void)0 )

// ---- stage entry points (defined in the per-stage .cu files; all stream-ordered, device pointers) ----
int traceImportNodes( hako_ctx* ctx, cudaStream_t s, const hako_node68* nodes68_d, uint32_t n, TraceNode* out, uint8_t* masks );
int traceExportNodes( hako_ctx* ctx, cudaStream_t s, const TraceNode* nodes, const uint8_t* masks, uint32_t n, hako_node68* out68_d );
int launchIntersectBatch( hako_ctx* ctx, cudaStream_t s, const OctreeView& v, const float* ro, const float* rd, uint64_t n, int shadow,
						  float* t, int32_t* nMajor, uint32_t* vIndex );
int launchRenderPrimary( hako_ctx* ctx, cudaStream_t s, const OctreeView& v, const hako_camera& cam, int w, int h, int y0, int y1,
						 int showVertexColor, uint8_t* rgba, float* t, int32_t* nMajor, uint32_t* vIndex );

// h2d.cu: upload `n` host segments; pageable sources go through worker threads that memcpy into pinned chunks while the
// previous chunk is on the wire (a plain cudaMemcpyAsync from pageable memory runs at ~11 GB/s on this box).  Stream-ordered on s.
struct H2DSegment
{
	void* dst;
	const void* src;
	size_t bytes;
};
int uploadSegments( hako_ctx* ctx, cudaStream_t s, const H2DSegment* segs, int n );
int uploadBegin( hako_ctx* ctx, cudaStream_t s, const H2DSegment* segs, int n, bool orderBehindStream, int groupSegs = 0,
				 const std::function<void( int )>& onGroup = std::function<void( int )>() ); // copies in flight; s does not wait yet
int uploadFinish( hako_ctx* ctx, cudaStream_t s );								  // s waits for every copy begun so far
int uploadMark( hako_ctx* ctx, cudaEvent_t ev );									  // ev is recorded behind every copy begun so far
void destroyStager( hako_ctx* ctx );

// voxelize.cu
// beforeEmit (may be empty): called once, after the counting pass, before the first kernel that reads col / emi is launched
int voxelizeDevice( hako_ctx* ctx, cudaStream_t s, const float* pos, const float* col, const float* emi, uint32_t nTriangles,
					V3 origin, float dps, int gridRes, int sixSeparating, uint64_t** keysOut, uint64_t** valuesOut, uint64_t* D,
					const std::function<int( int )>& beforeEmit = std::function<int( int )>(), int emitSlices = 1 );
// sort.cu
int sortPairsDevice( hako_ctx* ctx, cudaStream_t s, uint64_t* keys, uint64_t* values, uint64_t* keysTmp, uint64_t* valuesTmp,
					 uint64_t n, int beginBit, int endBit, int* resultInTmp );
// unique.cu
int uniqueMergeLaunch( hako_ctx* ctx, cudaStream_t s, const uint64_t* keys, const uint64_t* values, uint64_t n,
					   uint64_t* keysOut, hako_voxel_attrib* attrsOut, unsigned long long* counters_d ); // stream-ordered, no host wait
int countRunsDevice( hako_ctx* ctx, cudaStream_t s, const uint64_t* keys, uint64_t n, uint64_t* U, uint64_t* levelBoundaries = nullptr ); // distinct keys + per-level boundary counts of a sorted array (synchronises)
int uniqueMergeDevice( hako_ctx* ctx, cudaStream_t s, const uint64_t* keys, const uint64_t* values, uint64_t n,
					   uint64_t* keysOut, hako_voxel_attrib* attrsOut, uint64_t* U, uint32_t* hasEmission );
// octree.cu
// levelBoundaries: null, or the 22 per-level boundary counts of the key list (countRunsDevice) -- then the stage needs no host read of its own
int buildOctreeDevice( hako_ctx* ctx, cudaStream_t s, const uint64_t* mortons, uint64_t n, int gridRes, int dag,
					   TraceNode** nodesOut, uint8_t** masksOut, uint32_t* nNodes, uint32_t* launchesOut, int* wideOut, const uint64_t* levelBoundaries = nullptr );
} // namespace hako
