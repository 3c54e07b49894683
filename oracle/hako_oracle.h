/*
 * hako_oracle.h -- C ABI of the CPU oracle (TEST INFRASTRUCTURE ONLY).
 *
 * The oracle is a CPU restatement of the reference's own CPU path
 * (voxRT.cpp + voxelization.hpp + morton.hpp + IntersectorOctree.hpp +
 * voxCommon.hpp + renderCommon.hpp + pmjSampler.hpp).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load it.  The product (libhako_b200.so) never links or calls it.
 *
 * The same ABI is implemented twice:
 *   oracle/hako_oracle.cpp      -> liboracle.so       (our restatement, "port")
 *   oracle/ref_driver.cpp       -> _ref/libhako_ref.so (the reference's own headers,
 *                                  compiled where they lie; only in the build container)
 * so the restatement can be pinned bit-for-bit against the real reference.
 */
#ifndef HAKO_ORACLE_H
#define HAKO_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* 68-byte canonical node == reference OctreeNode (voxCommon.hpp:133-138). */
typedef struct OrcNode68
{
	uint8_t mask;
	uint8_t _pad[3];
	uint32_t children[8];
	uint32_t nVoxelsPSum[8];
} OrcNode68;

/* == reference CameraPinhole members (renderCommon.hpp:77-83), 60 bytes. */
typedef struct OrcCamera
{
	float o[3];
	float front[3];
	float up[3];
	float right[3];
	float tanHthetaY;
	float lensR;
	float focus;
} OrcCamera;

typedef struct OrcOctree OrcOctree;

const char* orc_impl_name( void ); /* "port" or "reference" */

/* morton.hpp:70-77 / :101-115 */
uint64_t orc_morton_encode( uint32_t x, uint32_t y, uint32_t z );
uint64_t orc_morton_encode_naive( uint32_t x, uint32_t y, uint32_t z );
void orc_morton_decode( uint64_t m, uint32_t* x, uint32_t* y, uint32_t* z );
int orc_sort_bits_morton( uint32_t gridRes ); /* voxCommon.hpp:113-118 */

/* voxRT.cpp:198-240 voxel loop.  Writes at most `capacity` entries, returns the
 * total number of dumped voxels D.  colors/emissions are uchar4 arrays. */
uint64_t orc_voxelize( const float* pos, const float* col, const float* emi, uint32_t nTriangles,
					   const float origin[3], float dps, int gridRes, int sixSeparating,
					   uint64_t* keys, uint8_t* colors, uint8_t* emissions, uint64_t capacity );

/* voxRT.cpp:14-51 mergeVoxels, in place; returns U. */
uint64_t orc_merge_voxels( uint64_t* keys, uint8_t* colors, uint8_t* emissions, uint64_t n );

/* IntersectorOctree.hpp:221-246 */
OrcOctree* orc_octree_build_dag( const uint64_t* mortons, uint64_t n, const float origin[3], float dps, int gridRes );
OrcOctree* orc_octree_build_naive( const uint64_t* mortons, uint64_t n, const float origin[3], float dps, int gridRes );
void orc_octree_free( OrcOctree* o );
uint32_t orc_octree_num_nodes( const OrcOctree* o );
void orc_octree_bounds( const OrcOctree* o, float lower[3], float upper[3] );
void orc_octree_copy_nodes( const OrcOctree* o, OrcNode68* out );

/* IntersectorOctree.hpp:248-257 -> voxCommon.hpp:231-423.  t must be preset by the caller
 * (FLT_MAX); a miss leaves t / nMajor / vIndex untouched.
 * stats (optional, 4 x uint64): loop visits, descents, pushes, max stack depth. */
void orc_intersect_batch( const OrcOctree* o, const float* ro, const float* rd, uint64_t n, int isShadowRay,
						  float* t, int* nMajor, uint32_t* vIndex, uint64_t* stats );

/* glm::lookAt + glm::perspectiveFov + CameraPinhole::initFromPerspective
 * (libs/prlib/src/pr.cpp:1104-1112,1154-1156; renderCommon.hpp:21-35). */
void orc_camera_lookat( const float eye[3], const float center[3], const float up[3], float fovy,
						int width, int height, float zNear, float zFar, float focus, float lensR, OrcCamera* out );
/* renderCommon.hpp:37-49 / :50-75 */
void orc_camera_shoot( const OrcCamera* cam, int x, int y, float ox, float oy, int w, int h, float ro[3], float rd[3] );
void orc_camera_shoot_thin_lens( const OrcCamera* cam, int x, int y, float ox, float oy, int w, int h, float u0, float u1, float ro[3], float rd[3] );

/* voxRT.cpp:307-347 render loop with pixel-centre rays; rows [y0,y1); miss => t=FLT_MAX, vIndex=0xFFFFFFFF, nMajor=-1.
 * nThreads > 1 splits rows over std::thread (the reference's ParallelFor is serial on Linux). */
void orc_render_primary( const OrcOctree* o, const OrcCamera* cam, int w, int h, int y0, int y1, int nThreads,
						 float* t, int* nMajor, uint32_t* vIndex, uint64_t* stats );

/* diagnostic (tools/ray_length_probe.py): traversal loop visits of every pixel-centre primary ray of a w x h frame */
void orc_render_primary_visits( const OrcOctree* o, const OrcCamera* cam, int w, int h, int nThreads, uint32_t* visits );

/* FNV-1a-64 over a byte buffer (used for the golden checksums in SURVEY.md section 8c). */
uint64_t orc_fnv1a64( const void* p, uint64_t bytes );

/* ---- path-tracing half (hako_oracle_pt.cpp / ref_driver_pt.cpp) ----
 * Pieces the reference can run on the host are implemented by BOTH libraries (pinning); orc_hdri_build_sats,
 * orc_hdri_sample_nearest, orc_pt_render and orc_pt_resolve restate device-only code (voxKernel.cu:485-795,
 * renderCommon.hpp:353-366) and exist only in the port ("parity unpinned" for those). */
uint32_t orc_murmur_combine( uint32_t seed, const uint32_t* values, int n );			  /* voxCommon.hpp:45-109 */
void orc_pcg32_stream( uint64_t seed, uint64_t stream, int n, uint32_t* out );			  /* renderCommon.hpp:86-111 */
uint32_t orc_nested_uniform_scramble( uint32_t x, uint32_t seed );						  /* pmjSampler.hpp:62-90 */
float orc_scramble_f32( float x, uint32_t seed );										  /* pmjSampler.hpp:92-102 */
void orc_pmj_table( float* out /* [128][4096][2] */ );									  /* pmjSampler.hpp:114-134 */
void orc_pmj_sample2d( const float* table, uint32_t sampleIdx, uint32_t dimension, uint32_t stream, float* xy ); /* :158-173 */
void orc_hdri_importance_sample( const float* rgba, const uint32_t* sats, int w, int h, float scale, const float* N, int axisAligned,
								 float u0, float u1, float u2, float u3, float* dir, float* L, float* srPDF, uint32_t* XY ); /* renderCommon.hpp:367-435 */
void orc_sample_lambertian( float a, float b, const float* Ng, float* out );			  /* renderCommon.hpp:119-151 */
/* port only */
void orc_hdri_build_sats( const float* rgba, int w, int h, uint32_t* sats /* [7][w*h] */ ); /* renderCommon.hpp:214-314 + voxKernel.cu:485-608 */
void orc_hdri_sample_nearest( const float* rgba, int w, int h, float scale, const float* d, float* out );
void orc_pt_render( const OrcOctree* octree, const uint8_t* attrs, uint32_t hasEmission, const OrcCamera* cam, int w, int h, int y0, int y1,
					int iteration, const float* hdriRgba, const uint32_t* sats, int hw, int hh, float hdriScale,
					const float* primaryRgba, int wP, int hP, const float* pmjTable, int maxDepth, float* accum, uint64_t* stats, int nThreads );
void orc_pt_resolve( const float* accum, uint64_t n, uint8_t* rgba );

#ifdef __cplusplus
}
#endif
#endif
