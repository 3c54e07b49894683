/*
 * hako.h -- C ABI of libhako_b200.so: the B200-native (sm_100a, ahead-of-time CUDA) replacement for the
 * GPU half of Ushio/MassiveVoxelRayTracing ("Hako").  Citations are relative to the reference tree.
 *
 * What this replaces: the reference has no FFI layer; its "operator API" is three C++ types whose GPU
 * halves are reached by launch-by-name on an nvrtc module (hipUtil.hpp:112-186):
 *   IntersectorOctree      IntersectorOctree.hpp:214-266     (CPU build / intersect)
 *   IntersectorOctreeGPU   IntersectorOctreeGPU.hpp:21-275   (GPU build, device intersect)
 *   PathTracer             PathTracer.hpp:14-170             (setup / updateScene / step / resolve)
 * Each entry point below names the reference interface it stands in for.
 *
 * Conventions
 *   - every function returns 0 on success, a negative HAKO_E_* code otherwise; hako_last_error() returns
 *     a thread-local message.  (The reference has no return codes: __debugbreak()/abort()/assert,
 *     hipUtil.hpp:18-22, IntersectorOctreeGPU.hpp:48-51.)
 *   - `stream` is a cudaStream_t passed as void* (stand-in for oroStream); NULL = the legacy default stream.
 *     Calls are stream-ordered.  Functions whose pointer arguments are HOST buffers return after the
 *     results have landed in them; the *_device variants take DEVICE pointers and only enqueue work.
 *   - handles are opaque and owned by the library; one hako_ctx per device, no hidden statics
 *     (unlike `static thrs::RadixSort`, IntersectorOctreeGPU.hpp:58).
 *   - a hako_ctx (and the octrees / path tracers made from it) may be used by ONE thread at a time: it owns staging buffers, build
 *     statistics and internal streams that calls share.  Use one context per thread (several contexts per device are fine), as
 *     tests/cpp/shim_demo.cpp does with one thread per GPU.
 *   - there is NO CPU fallback: every call fails with HAKO_E_CUDA when no sm_100 device is usable.
 */
#ifndef HAKO_H
#define HAKO_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HAKO_OK 0
#define HAKO_E_INVALID -1 /* bad argument (e.g. gridRes not a power of two: IntersectorOctreeGPU.hpp:48-51 abort()) */
#define HAKO_E_CUDA -2	  /* CUDA runtime error / no device */
#define HAKO_E_LIMIT -3	  /* 32-bit format ceiling exceeded (voxel / node counters), or a reference-layout export of a wide-index tree */
#define HAKO_E_STATE -4	  /* call order (e.g. step before resize) */

typedef struct hako_ctx hako_ctx;
typedef struct hako_octree hako_octree;
typedef struct hako_pt hako_pt;
typedef struct hako_pairs hako_pairs; /* sorted dumped voxels of one triangle shard, resident on the device */
typedef struct hako_mesh hako_mesh;
typedef struct hako_seq hako_seq;
typedef struct hako_dist hako_dist; /* one NCCL communicator rank bound to a hako_ctx */
typedef void* hako_stream; /* cudaStream_t */

/* == OctreeNode, voxCommon.hpp:133-138 (68 bytes, align 4).  children[i] = index | mask<<24 after embedMasks
 * (voxCommon.hpp:183-195); absent and leaf-level children are 0xFFFFFFFF. */
typedef struct hako_node68
{
	uint8_t mask;
	uint8_t _pad[3];
	uint32_t children[8];
	uint32_t nVoxelsPSum[8];
} hako_node68;

/* == VoxelAttirb, voxCommon.hpp:121-125 */
typedef struct hako_voxel_attrib
{
	uint8_t color[4];
	uint8_t emission[4];
} hako_voxel_attrib;

/* == CameraPinhole members, renderCommon.hpp:77-83 (60 bytes) */
typedef struct hako_camera
{
	float o[3];
	float front[3];
	float up[3];
	float right[3];
	float tanHthetaY;
	float lensR;
	float focus;
} hako_camera;

/* == the POD members of IntersectorOctreeGPU, IntersectorOctreeGPU.hpp:265-274 */
typedef struct hako_octree_info
{
	uint32_t numberOfNodes;
	uint32_t numberOfVoxels;
	float lower[3];
	float upper[3];
	float dps;
	float emissionScale;
	uint32_t hasEmission;
	int32_t gridRes;
	uint64_t deviceBytes; /* nodes + attributes (+ kept Morton keys) resident in HBM */
	/* 1: the tree has >= 2^24 - 1 nodes, more than the reference's child pointers (index | mask << 24, voxCommon.hpp:183-195) can
	 * address; it is kept with plain 32-bit child indices (the traversal reads child masks from a side array) and cannot be exported
	 * in the 68-byte reference layout.  0: reference layout. */
	uint32_t wideIndex;
	uint32_t reserved;
} hako_octree_info;

/* per-stage record of the last hako_octree_build* call on a context (voxRTGPU.cpp:169-173 prints one total) */
typedef struct hako_build_stats
{
	uint64_t nTriangles;
	uint64_t dumpedVoxels; /* D: voxels emitted before sort/unique */
	uint64_t uniqueVoxels; /* U */
	uint32_t numberOfNodes;
	uint32_t sortPasses;
	/* device times between stream events.  hako_octree_build uploads the positions first (ms_h2d); the colour / emission arrays
	 * follow by triangle range behind the voxelizer's counting pass and the emission of earlier ranges, so the exposed
	 * remainder of that upload is inside ms_voxelize. */
	float ms_h2d, ms_voxelize, ms_sort, ms_unique, ms_octree, ms_total;
	uint32_t hostSyncs;
	uint32_t kernelLaunches;
} hako_build_stats;

const char* hako_last_error( void );
int hako_version( void );

/* ---- context (replaces oroInitialize/oroCtxCreate/Shader nvrtc compile, voxRTGPU.cpp:25-68) ---- */
int hako_ctx_create( int device, hako_ctx** out );
void hako_ctx_destroy( hako_ctx* ctx );
int hako_ctx_synchronize( hako_ctx* ctx, hako_stream stream );
int hako_ctx_build_stats( const hako_ctx* ctx, hako_build_stats* out );
/* number of kernel launches issued through this context since creation (bench.py's gpu_launches claim) */
uint64_t hako_ctx_launch_count( const hako_ctx* ctx );
/* map `bytes` of device memory into the context's stream-ordered pool now (kept there), so that the first build does not grow the pool
 * on the caller's critical path; optional -- the pool grows on demand otherwise */
int hako_ctx_reserve( hako_ctx* ctx, hako_stream stream, uint64_t bytes );
/* measurement: the L2 read roof of the device (traversal is served by L2 while the node array fits it).  Sweeps a
 * buffer of setBytes (0 = a quarter of the L2) `reps` times; *l2Bytes = cudaDevAttrL2CacheSize, *readGBs = best of three. */
int hako_ctx_l2_probe( hako_ctx* ctx, hako_stream stream, uint64_t setBytes, int reps, uint64_t* l2Bytes, float* readGBs );

/* ---- build: IntersectorOctreeGPU::build, IntersectorOctreeGPU.hpp:40-241 ----
 * pos/col/emi: 9 floats per triangle (v0,v1,v2), as trianglesFlattened (voxUtil.hpp:8-65) produces.
 * HOST pointers; the H2D copy is part of the call, as in the reference (:68-70).  On return the host arrays are free to reuse
 * and every count is final; the last device copy of the node array may still be in flight on `stream` (the reference returns
 * with its embedMasks launch in flight, :230-238): use the octree on the same stream, or synchronise the stream first. */
int hako_octree_build( hako_ctx* ctx, hako_stream stream, const float* pos, const float* col, const float* emi,
					   uint32_t nTriangles, const float origin[3], float dps, int gridRes, hako_octree** out );
/* same with DEVICE-resident vertex data (per-frame rebuilds skip the H2D) */
int hako_octree_build_device( hako_ctx* ctx, hako_stream stream, const float* pos_d, const float* col_d, const float* emi_d,
							  uint32_t nTriangles, const float origin[3], float dps, int gridRes, hako_octree** out );
/* IntersectorOctree::buildDAGReference / build, IntersectorOctree.hpp:221-235: sorted unique Morton list
 * (HOST) -> octree.  attrs may be NULL.  dag = 1: DAG (reference numbering), 0: naive (no dedupe). */
int hako_octree_build_from_morton( hako_ctx* ctx, hako_stream stream, const uint64_t* mortons, const hako_voxel_attrib* attrs,
								   uint64_t n, const float origin[3], float dps, int gridRes, int dag, hako_octree** out );
/* upload a tree built elsewhere (canonical 68-byte layout, masks embedded); parity tooling */
int hako_octree_import( hako_ctx* ctx, hako_stream stream, const hako_node68* nodes, uint32_t nNodes,
						const hako_voxel_attrib* attrs, uint32_t nVoxels, const float lower[3], const float upper[3], float dps,
						hako_octree** out );
int hako_octree_info_get( const hako_octree* o, hako_octree_info* out );
/* download in the reference layout.  Any of nodes / mortons / attrs may be NULL.  HOST pointers. */
int hako_octree_export( const hako_octree* o, hako_stream stream, hako_node68* nodes, uint64_t* mortons, hako_voxel_attrib* attrs );
/* IntersectorOctreeGPU::cleanUp, IntersectorOctreeGPU.hpp:26-38 */
void hako_octree_destroy( hako_octree* o );
/* the same without the device-wide wait: the arrays are released behind the work already queued on `stream` (frame loops) */
void hako_octree_destroy_async( hako_octree* o, hako_stream stream );
/* device addresses of the resident arrays (for NCCL broadcast by the host layer); sizes in bytes */
int hako_octree_device_arrays( const hako_octree* o, void** nodes_d, uint64_t* nodeBytes, void** masks_d, uint64_t* maskBytes,
							   void** attrs_d, uint64_t* attrBytes );
/* allocate an empty replica of given sizes to receive a broadcast; fill via hako_octree_device_arrays */
int hako_octree_create_replica( hako_ctx* ctx, const hako_octree_info* info, hako_octree** out );

/* ---- a build sharded over GPUs (SURVEY.md 8e "sharding the build by Morton range"; no counterpart in the reference, whose
 * IntersectorOctreeGPU::build runs on one device).  The same stages as hako_octree_build, cut where the exchanges go:
 *   1. hako_voxelize_sorted       a rank's shard of the triangles (HOST arrays, scene frame) -> sorted (key, packed attribute)
 *                                 pairs on the device; the caller exchanges them by Morton range (NCCL all-to-all);
 *   2. hako_sort_unique_device    the pairs a rank received for its range (DEVICE, any order) -> sorted unique keys and
 *                                 integer-mean attributes (== the `unique` kernel) in place at the front of the same arrays;
 *   3. hako_octree_build_from_morton_device   the gathered unique voxels (DEVICE) -> octree, on every rank.
 * Results are the bytes hako_octree_build produces on one GPU (tests/test_gpu_build.py, tests/test_dist_gpu.py). */
int hako_voxelize_sorted( hako_ctx* ctx, hako_stream stream, const float* pos, const float* col, const float* emi, uint32_t nTriangles,
						  const float origin[3], float dps, int gridRes, hako_pairs** out );
int hako_pairs_info( const hako_pairs* p, uint64_t* n, void** keys_d, void** values_d );
void hako_pairs_destroy( hako_pairs* p );
int hako_sort_unique_device( hako_ctx* ctx, hako_stream stream, uint64_t* keys_d, uint64_t* values_d, uint64_t n, int gridRes, uint64_t* U, uint32_t* hasEmission );
int hako_octree_build_from_morton_device( hako_ctx* ctx, hako_stream stream, const uint64_t* mortons_d, const hako_voxel_attrib* attrs_d,
										  uint64_t n, const float origin[3], float dps, int gridRes, int dag, int hasEmission, hako_octree** out );
/* stage 3 without the staging copy: the octree object first -- its resident key / attribute arrays, sized for n voxels, are
 * returned so that the gathering collective can write into them -- then the nodes, built from those arrays in place */
int hako_octree_create_for_voxels( hako_ctx* ctx, hako_stream stream, uint64_t n, const float origin[3], float dps, int gridRes, hako_octree** out,
								   void** mortons_d, void** attrs_d );
int hako_octree_build_nodes( hako_octree* octree, hako_stream stream, int dag, int hasEmission );

/* ---- build stages, exposed for stage-level parity and the build sweep (BASELINE config 4) ---- */
/* voxCount + voxelize, voxKernel.cu:58-166.  HOST in, HOST out.  Call with keys == NULL to get D only.
 * Emission order is unspecified (the reference's is an atomicAdd race); compare as a multiset. */
int hako_voxelize( hako_ctx* ctx, hako_stream stream, const float* pos, const float* col, const float* emi, uint32_t nTriangles,
				   const float origin[3], float dps, int gridRes, int sixSeparating,
				   uint64_t* keys, hako_voxel_attrib* attrs, uint64_t capacity, uint64_t* D );
/* thrs::RadixSort::sortPairs, tinyhipradixsort.hpp:849-944: stable LSD sort of (u64 key, u64 value) on bits
 * [beginBit, endBit).  DEVICE pointers, sorted in place.  values_d may be NULL (sortKeys). */
int hako_sort_pairs_device( hako_ctx* ctx, hako_stream stream, uint64_t* keys_d, uint64_t* values_d, uint64_t n, int beginBit, int endBit );
/* HOST convenience wrapper of the above */
int hako_sort_pairs( hako_ctx* ctx, hako_stream stream, uint64_t* keys, uint64_t* values, uint64_t n, int beginBit, int endBit );
/* `unique` kernel, voxKernel.cu:170-243 == mergeVoxels, voxRT.cpp:14-51: sorted (key, attrib) -> first of each run,
 * integer-mean attributes.  HOST in/out (in place), *U receives the count. */
int hako_unique_merge( hako_ctx* ctx, hako_stream stream, uint64_t* keys, hako_voxel_attrib* attrs, uint64_t n, uint64_t* U, uint32_t* hasEmission );
/* The same merge for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * NO COUNTERPART IN THE REFERENCE.  Calls no cudaMalloc / cudaFree / pool allocation and does
 * not wait for the stream: it enqueues two memsets and the merge kernel.  DEVICE pointers throughout: sorted keys_d / attrs_d (n entries,
 * 16-byte aligned) -> keysOut_d / attrsOut_d (room for n entries) and counts_d[2] = { number of runs U, hasEmission }, all written
 * stream-ordered.  The kernel's scratch (scan descriptors) is a static device array of 64.3 KB that every GPU loading the library holds
 * until exit, which bounds n: more than HAKO_UNIQUE_NOALLOC_MAX_ITEMS entries -> HAKO_E_LIMIT.  All calls on a device share that array:
 * calls that overlap on the GPU (different streams without a dependency) race; calls on one stream, or ordered by events, do not.
 * The first call on a device resolves the array's address (cudaGetSymbolAddress); make it once outside any graph capture. */
#define HAKO_UNIQUE_NOALLOC_MAX_ITEMS ( 1ull << 24 )
int hako_unique_merge_noalloc( hako_ctx* ctx, hako_stream stream, const uint64_t* keys_d, const hako_voxel_attrib* attrs_d, uint64_t n,
							   uint64_t* keysOut_d, hako_voxel_attrib* attrsOut_d, uint64_t* counts_d );

/* ---- intersect: IntersectorOctreeGPU::intersect (device, :243-251) / IntersectorOctree::intersect (:248-257) ----
 * ro/rd: 3 floats per ray (rd un-normalised).  Caller presets t = FLT_MAX; a miss leaves t / nMajor / vIndex
 * untouched (voxRT.cpp:313-325).  HOST pointers. */
int hako_intersect_batch( const hako_octree* o, hako_stream stream, const float* ro, const float* rd, uint64_t n, int isShadowRay,
						  float* t, int32_t* nMajor, uint32_t* vIndex );
int hako_intersect_batch_device( const hako_octree* o, hako_stream stream, const float* ro_d, const float* rd_d, uint64_t n, int isShadowRay,
								 float* t_d, int32_t* nMajor_d, uint32_t* vIndex_d );

/* ---- primary rays: `render` kernel, voxKernel.cu:437-483 / renderLine, voxRT.cpp:307-347 ----
 * Pixel-centre rays of rows [y0, y1) of a w x h image (the row range is how image tiles are sharded over GPUs).
 * Outputs are (y1-y0) x w, row-major; any may be NULL.  rgba: hit -> (hitN+1)/2 or voxel colour, miss -> (0,0,0,255).
 * t / nMajor / vIndex: miss -> FLT_MAX / -1 / 0xFFFFFFFF. */
int hako_render_primary( const hako_octree* o, hako_stream stream, const hako_camera* cam, int w, int h, int y0, int y1,
						 int showVertexColor, uint8_t* rgba, float* t, int32_t* nMajor, uint32_t* vIndex );
int hako_render_primary_device( const hako_octree* o, hako_stream stream, const hako_camera* cam, int w, int h, int y0, int y1,
								int showVertexColor, uint8_t* rgba_d, float* t_d, int32_t* nMajor_d, uint32_t* vIndex_d );
/* A run of nViews whole frames in one call -- the orbit loop of voxRTGPU.cpp:150-208 without the window: cams[i] -> frame i of
 * every non-NULL output (HOST pointers, nViews x h x w, frame-major).  Frame i+1 is traced while frame i is copied out, so the
 * call costs about max( trace, D2H ) per frame instead of their sum; results equal nViews hako_render_primary calls. */
int hako_render_primary_views( const hako_octree* o, hako_stream stream, const hako_camera* cams, int nViews, int w, int h,
							   int showVertexColor, uint8_t* rgba, float* t, int32_t* nMajor, uint32_t* vIndex );

/* ---- camera (host-side math; no device work) ----
 * CameraPinhole::initFromPerspective( viewMatrix, projMatrix, focus, lensR ), renderCommon.hpp:21-35.
 * view / proj: column-major 4x4 as glm stores them (glm::value_ptr). */
int hako_camera_from_matrices( const float view[16], const float proj[16], float focus, float lensR, hako_camera* out );
/* pr::GetCameraMatrix (libs/prlib/src/pr.cpp:1099-1191) == glm::lookAt + glm::perspectiveFov, then initFromPerspective:
 * what voxRT.cpp:300-305 / PathTracer::step (PathTracer.hpp:150-160) do with a pr::Camera3D. */
int hako_camera_lookat( const float eye[3], const float center[3], const float up[3], float fovy, int width, int height,
						float zNear, float zFar, float focus, float lensR, hako_camera* out );

/* ---- path tracer: PathTracer, PathTracer.hpp:14-170 ---- */
/* setup (:43-69): PMJ table generation (pmjSampler.hpp:114-134) + upload.  No stack pool (DynamicAllocatorGPU is replaced). */
int hako_pt_create( hako_ctx* ctx, hako_stream stream, hako_pt** out );
void hako_pt_destroy( hako_pt* pt ); /* cleanUp, :70-79 */
int hako_pt_resize( hako_pt* pt, hako_stream stream, int w, int h ); /* resizeFrameBufferIfNeeded, :81-97 */
int hako_pt_clear( hako_pt* pt, hako_stream stream );				 /* clearFrameBuffer, :98-103 */
/* loadHDRI (:104-116) minus the file decode: RGBA32F pixels (HOST), optional separate primary image (may be NULL). */
int hako_pt_load_hdri( hako_pt* pt, hako_stream stream, const float* rgba, int w, int h, const float* rgbaPrimary, int wP, int hP );
/* step (:150-169): one call = 16 spp; `iteration` is m_steps (spp = iteration*16 + localSpp, voxKernel.cu:642).
 * Rows [y0,y1) only (tile sharding); maxDepth: reference = 8 (voxKernel.cu:691). */
int hako_pt_step( hako_pt* pt, const hako_octree* o, hako_stream stream, const hako_camera* cam, int iteration, int y0, int y1, int maxDepth );
int hako_pt_resolve( hako_pt* pt, hako_stream stream );										/* resolve, :118-127 */
int hako_pt_read_rgba8( hako_pt* pt, hako_stream stream, uint8_t* rgba );					/* toImageAsync, :128-137 (HOST) */
int hako_pt_read_rgba8_async( hako_pt* pt, hako_stream stream, uint8_t* rgba );			/* the same without the host wait: the caller synchronises the stream (voxPTGPU.cpp:188-189) */
int hako_pt_read_accum( hako_pt* pt, hako_stream stream, float* rgbw );						/* float4 accumulators (HOST) */
int hako_pt_accum_device( hako_pt* pt, void** accum_d, uint64_t* bytes );					/* for ncclReduce by the host layer */
int hako_pt_ray_count( hako_pt* pt, hako_stream stream, uint64_t* raysTraced, int reset ); /* traversal calls since reset */
/* node visits of the secondary rays (trace waves of the wavefront path tracer) since reset: with hako_pt_ray_count the measure of how long
   those rays are; 0 in the persistent-kernel modes (HAKO_PT_WAVEFRONT=0/1) */
int hako_pt_trace_steps( hako_pt* pt, hako_stream stream, uint64_t* steps, int reset );
/* pieces of the sampler / environment exposed for parity tests */
int hako_pt_pmj_table( hako_pt* pt, hako_stream stream, float* samples /* [128][4096][2], HOST */ );
int hako_pt_hdri_sats( hako_pt* pt, hako_stream stream, uint32_t* sats /* [7][w*h], HOST */ );

int hako_pt_rgba8_device( hako_pt* pt, void** rgba8_d, int* w, int* h );				/* m_frameBufferU8 (PathTracer.hpp:93), for the frame loop */
hako_ctx* hako_pt_context( hako_pt* pt );

/* ---- input stage: voxUtil.hpp:8-77, voxRT.cpp:188-190 ---- */
/* triangle meshes from Wavefront OBJ (what prlib loads through tiny_obj_loader, libs/prlib/src/prg.cpp:1231-1241; "v x y z r g b"
 * vertex colours are read as Cd) or Stanford PLY (ascii / binary, the xyzrgb dragon's native format).  Polygons are fanned. */
int hako_mesh_load( const char* path, hako_mesh** out );
void hako_mesh_destroy( hako_mesh* m );
int hako_mesh_info( const hako_mesh* m, uint64_t* nVertices, uint64_t* nTriangles, int* hasColors );
/* trianglesFlattened (voxUtil.hpp:8-65): 9 floats per triangle into pos / col / emi (col, emi may be NULL); colour = Cd or 1, emission = 0 */
int hako_mesh_flatten( const hako_mesh* m, float* pos, float* col, float* emi );
/* getBoundingBox (voxUtil.hpp:66-77) as a device reduction over DEVICE-resident vertices (3 floats each); HOST results */
int hako_bounds_device( hako_ctx* ctx, hako_stream stream, const float* pos_d, uint64_t nVertices, float lower[3], float upper[3] );
/* voxRT.cpp:188-190: origin = bbox lower, dps = max( bbox size ) / (float)gridRes */
int hako_scene_params( const float lower[3], const float upper[3], int gridRes, float origin[3], float* dps );

/* ---- images ---- */
/* Radiance .hdr -> RGBA32F exactly as stbi_loadf( file, &w, &h, 0, 4 ) (libs/prlib/src/prp.cpp:353-366): the decode step of
 * PathTracer::loadHDRI (PathTracer.hpp:104-116).  *rgba is malloc'ed; release with hako_image_free. */
int hako_image_load_hdr( const char* path, float** rgba, int* w, int* h );
void hako_image_free( float* rgba );
/* Image2DRGBA8::saveAsPngUncompressed (libs/prlib/src/prp.cpp:215-223): RGBA8 PNG with stored deflate blocks */
int hako_image_save_png( const char* path, const uint8_t* rgba, int w, int h );

/* ---- frame loop: RTCamp.cpp:122-196, headless ----
 * Per frame: updateScene -> clearFrameBuffer -> `iterations` steps -> resolve -> D2D into one of nBuffers pooled frame buffers ->
 * a pool thread downloads it and writes the PNG while the next frame builds and renders.  HOST vertex pointers. */
int hako_seq_create( hako_pt* pt, int nBuffers, int nWriters, hako_seq** out );
int hako_seq_render_frame( hako_seq* q, const float* pos, const float* col, const float* emi, uint32_t nTriangles, const float origin[3], float dps,
						   int gridRes, const hako_camera* cam, int iterations, int maxDepth, const char* pngPath, float* updateMs, float* renderMs );
int hako_seq_wait( hako_seq* q ); /* taskGroup.waitForAllElementsToFinish, RTCamp.cpp:196 */
void hako_seq_destroy( hako_seq* q );

/* ---- multi-GPU: one hako_ctx per GPU, NCCL over NVLink / NVSwitch.  No counterpart in the reference (its only
 * distribution is usecase2_submission/run.py:13-27).  A C++ host written like voxPTGPU.cpp:88-193 / RTCamp.cpp:138-196 drives it
 * with one process per GPU under any launcher, or one thread per GPU in one process; libnccl.so.2 is loaded at hako_dist_init.
 *   rank 0:  hako_dist_unique_id( id )  -> ship the 128 bytes to the other ranks  -> every rank: hako_dist_init( ctx, id, rank, world )
 *   octree:  build on one rank, hako_dist_broadcast_octree          (or hako_dist_build_sharded: every rank builds 1/G of it)
 *   primary rays: rows / views per rank, no collective (hako_render_primary's y0, y1)
 *   path tracing: rank r runs hako_pt_step for iterations r, r + G, ... then ONE hako_dist_reduce_accum; resolve on the root. */
#define HAKO_DIST_ID_BYTES 128
int hako_dist_unique_id( uint8_t id[HAKO_DIST_ID_BYTES] ); /* ncclGetUniqueId */
int hako_dist_init( hako_ctx* ctx, const uint8_t id[HAKO_DIST_ID_BYTES], int rank, int world, hako_dist** out ); /* collective; warms the connections up */
void hako_dist_destroy( hako_dist* d );
int hako_dist_info( const hako_dist* d, int* rank, int* world, int* ncclVersion );
int hako_dist_barrier( hako_dist* d, hako_stream stream );					 /* returns when every rank's stream has reached it */
int hako_dist_max_f32( hako_dist* d, hako_stream stream, float* value );	 /* max over ranks of a host float (device-timed ms) */
/* Replicate rank root's octree (IntersectorOctreeGPU's two device buffers, IntersectorOctreeGPU.hpp:265-266, + the mask side array).
 * src: the octree on root, ignored elsewhere.  *out: src on root, elsewhere a new replica (hako_octree_destroy it).  A 64-byte
 * header broadcast and one host wait size the receivers' arrays; nodes + masks + attributes then travel as ONE NCCL group on
 * `stream` and the call returns with it enqueued. */
int hako_dist_broadcast_octree( hako_dist* d, hako_stream stream, int root, hako_octree* src, hako_octree** out );
/* ncclReduce( sum ) of the float4 accumulators of pt (PathTracer::m_frameBufferF32, PathTracer.hpp:24) onto root, in place. */
int hako_dist_reduce_accum( hako_dist* d, hako_stream stream, hako_pt* pt, int root );
int hako_dist_reduce_f32( hako_dist* d, hako_stream stream, float* buf_d, uint64_t count, int root ); /* any DEVICE float buffer */
int hako_dist_broadcast_bytes( hako_dist* d, hako_stream stream, void* buf_d, uint64_t bytes, int root );
/* IntersectorOctreeGPU::build (IntersectorOctreeGPU.hpp:40-241) sharded over the ranks: pos / col / emi = THIS rank's triangles
 * (HOST, scene frame).  Every rank ends with the octree of the whole scene, byte-identical to hako_octree_build on one GPU.
 * timesMs: NULL or 6 floats (voxelize+sort, splitters, all-to-all, sort+unique, gather, nodes; host clock, synchronised). */
int hako_dist_build_sharded( hako_dist* d, hako_stream stream, const float* pos, const float* col, const float* emi, uint32_t nTriangles,
							 const float origin[3], float dps, int gridRes, hako_octree** out, float* timesMs );

/* ---- debug export: saveVoxelsAsMesh, voxMesh.cpp:111-219 ----
 * binary little-endian PLY: 8 corners per voxel, one quad per face without a neighbouring voxel.  HOST pointers. */
int hako_voxel_mesh_save_ply( const char* path, const uint64_t* sortedMortons, uint64_t n, const float origin[3], float dps );

#ifdef __cplusplus
}
#endif
#endif
