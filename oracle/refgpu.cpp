/*
 * refgpu.cpp -- driver-API launcher for the REFERENCE's own GPU kernels (TEST INFRASTRUCTURE ONLY).
 *
 * oracle/Makefile stages /root/reference/voxKernel.cu with its headers (the same 2-line scope patch of voxCommon.hpp that the
 * CPU build needs) and compiles it ahead of time with `nvcc -arch=sm_100a -cubin` into oracle/_ref/voxKernel.sm_100a*.cubin.
 * Nothing of the reference is copied into the repository: this file only fills the by-value kernel arguments the reference's
 * host code would fill (ShaderArgument::add memcpy, hipUtil.hpp:79-103) and launches the kernels by name as
 * Shader::launch does (hipUtil.hpp:161-180):
 *
 *   render               voxKernel.cu:437-483   launched as voxRTGPU.cpp:191-203
 *   HDRIstoreImportance  voxKernel.cu:485-524 \
 *   buildSATh / buildSATv          :557-598    > launched as HDRI::load, renderCommon.hpp:214-314
 *   buildSAT2u32                   :600-608   /
 *   renderPT             voxKernel.cu:610-777   launched as PathTracer::step, PathTracer.hpp:150-169
 *   renderResolve        voxKernel.cu:779-795   launched as PathTracer::resolve, PathTracer.hpp:130-137
 *
 * It is the only available run of the code that exists solely on the device in the reference (renderPT, the SAT build,
 * HDRI::sampleNearest renderCommon.hpp:353-366), so it is what pins those rows; and it is the same-box GPU baseline
 * ("the bar the new traversal kernel must beat").
 *
 * libcuda is dlopen'ed so that this library loads (for symbol checks) on a machine without a driver.  The primary context of
 * the device is used, so the launcher coexists with libhako_b200.so (CUDA runtime) in one process.
 */
#include <cuda.h>
#include <dlfcn.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

#include "hako_oracle.h"

namespace
{
thread_local std::string g_err;

struct Api
{
	void* lib = nullptr;
#define F( name ) decltype( &::name ) name = nullptr;
	F( cuInit )
	F( cuDeviceGet )
	F( cuDevicePrimaryCtxRetain )
	F( cuDevicePrimaryCtxRelease )
	F( cuCtxSetCurrent )
	F( cuCtxSynchronize )
	F( cuModuleLoad )
	F( cuModuleUnload )
	F( cuModuleGetFunction )
	F( cuLaunchKernel )
	F( cuMemAlloc )
	F( cuMemFree )
	F( cuMemcpyHtoD )
	F( cuMemcpyDtoH )
	F( cuMemcpyDtoD )
	F( cuMemsetD32 )
	F( cuEventCreate )
	F( cuEventDestroy )
	F( cuEventRecord )
	F( cuEventSynchronize )
	F( cuEventElapsedTime )
	F( cuGetErrorString )
	F( cuFuncGetParamInfo )
	F( cuFuncGetAttribute )
#undef F
};
Api g;

bool loadApi()
{
	if( g.lib ) return true;
	void* h = dlopen( "libcuda.so.1", RTLD_NOW | RTLD_GLOBAL );
	if( !h )
	{
		g_err = std::string( "libcuda.so.1: " ) + dlerror();
		return false;
	}
	bool ok = true;
#define S2( x ) #x
#define S( x ) S2( x )
	/* cuda.h maps the plain names to their versioned symbols by macro (cuMemAlloc -> cuMemAlloc_v2 ...); S() expands them */
#define L( name )                                                    \
	g.name = (decltype( g.name ))dlsym( h, S( name ) );              \
	if( !g.name )                                                    \
	{                                                                \
		g_err = std::string( "missing driver symbol " ) + S( name ); \
		ok = false;                                                  \
	}
	L( cuInit )
	L( cuDeviceGet )
	L( cuDevicePrimaryCtxRetain )
	L( cuDevicePrimaryCtxRelease )
	L( cuCtxSetCurrent )
	L( cuCtxSynchronize )
	L( cuModuleLoad )
	L( cuModuleUnload )
	L( cuModuleGetFunction )
	L( cuLaunchKernel )
	L( cuMemAlloc )
	L( cuMemFree )
	L( cuMemcpyHtoD )
	L( cuMemcpyDtoH )
	L( cuMemcpyDtoD )
	L( cuMemsetD32 )
	L( cuEventCreate )
	L( cuEventDestroy )
	L( cuEventRecord )
	L( cuEventSynchronize )
	L( cuEventElapsedTime )
	L( cuGetErrorString )
	L( cuFuncGetParamInfo )
	L( cuFuncGetAttribute )
#undef L
	if( !ok ) return false;
	g.lib = h;
	return true;
}

int fail( CUresult r, const char* what )
{
	const char* s = nullptr;
	if( g.cuGetErrorString ) g.cuGetErrorString( r, &s );
	char buf[256];
	snprintf( buf, sizeof( buf ), "%s: CUDA driver error %d (%s)", what, (int)r, s ? s : "?" );
	g_err = buf;
	return -2;
}
#define CK( call )                                          \
	do                                                      \
	{                                                       \
		CUresult r_ = ( call );                             \
		if( r_ != CUDA_SUCCESS ) return fail( r_, #call ); \
	} while( 0 )

/* ---- by-value kernel arguments, laid out as nvcc lays out the reference's structs (sizes checked against the cubin) ---- */
struct f3
{
	float x, y, z;
};
struct i2
{
	int x, y;
};
/* == IntersectorOctreeGPU data members, IntersectorOctreeGPU.hpp:265-274 */
struct ArgIntersector
{
	CUdeviceptr vAttributeBuffer;
	CUdeviceptr nodeBuffer;
	uint32_t numberOfNodes;
	uint32_t numberOfVoxels;
	f3 lower;
	f3 upper;
	float dps;
	float emissionScale;
	uint32_t hasEmission;
	uint32_t _tail;
};
static_assert( sizeof( ArgIntersector ) == 64, "IntersectorOctreeGPU is 64 bytes" );
/* == DynamicAllocatorGPU<StackElement> data members, IntersectorOctreeGPU.hpp:356-360 */
struct ArgAllocator
{
	CUdeviceptr memory;
	CUdeviceptr locks;
	int numberOfBlock;
	int blockSize;
	int nElementPerThread;
	int _tail;
};
static_assert( sizeof( ArgAllocator ) == 32, "DynamicAllocatorGPU is 32 bytes" );
/* == HDRI data members, renderCommon.hpp:472-480 */
struct ArgHdri
{
	CUdeviceptr pixels;
	CUdeviceptr pixelsPrimary;
	CUdeviceptr sat;
	CUdeviceptr sats[6];
	int width, height, widthPrimary, heightPrimary;
	float scale;
	int _tail;
};
static_assert( sizeof( ArgHdri ) == 96, "HDRI is 96 bytes" );
static_assert( sizeof( OrcCamera ) == 60, "CameraPinhole is 60 bytes" );

} // namespace

struct RefGpu
{
	CUdevice dev = 0;
	CUcontext ctx = nullptr;
	CUmodule mod = nullptr;
	CUfunction render = nullptr, storeImportance = nullptr, satH = nullptr, satV = nullptr, sat2u32 = nullptr, renderPT = nullptr, resolve = nullptr;
	CUevent e0 = nullptr, e1 = nullptr;
	ArgIntersector isect{};
	ArgAllocator alloc{};
	ArgHdri hdri{};
	CUdeviceptr pmj = 0, counter = 0, fbF32 = 0, fbU8 = 0, frame = 0;
	size_t frameBytes = 0;
	int w = 0, h = 0;
};

namespace
{
int checkParams( CUfunction f, const char* name, const std::vector<size_t>& sizes )
{
	for( size_t i = 0; i < sizes.size(); i++ )
	{
		size_t off = 0, sz = 0;
		CUresult r = g.cuFuncGetParamInfo( f, i, &off, &sz );
		if( r != CUDA_SUCCESS ) return fail( r, name );
		if( sz != sizes[i] )
		{
			char buf[200];
			snprintf( buf, sizeof( buf ), "kernel %s: parameter %zu is %zu bytes in the cubin, launcher passes %zu", name, i, sz, sizes[i] );
			g_err = buf;
			return -1;
		}
	}
	size_t off = 0, sz = 0;
	if( g.cuFuncGetParamInfo( f, sizes.size(), &off, &sz ) == CUDA_SUCCESS )
	{
		g_err = std::string( "kernel " ) + name + " has more parameters than the launcher passes";
		return -1;
	}
	return 0;
}
void freeIf( CUdeviceptr& p )
{
	if( p ) g.cuMemFree( p );
	p = 0;
}
} // namespace

extern "C" {

const char* refgpu_last_error( void ) { return g_err.c_str(); }

void refgpu_close( RefGpu* r )
{
	if( !r ) return;
	if( r->ctx )
	{
		g.cuCtxSetCurrent( r->ctx );
		g.cuCtxSynchronize();
		freeIf( r->isect.vAttributeBuffer );
		freeIf( r->isect.nodeBuffer );
		freeIf( r->alloc.memory );
		freeIf( r->alloc.locks );
		freeIf( r->hdri.pixels );
		freeIf( r->hdri.pixelsPrimary );
		freeIf( r->hdri.sat );
		for( auto& s : r->hdri.sats ) freeIf( s );
		freeIf( r->pmj );
		freeIf( r->counter );
		freeIf( r->fbF32 );
		freeIf( r->fbU8 );
		freeIf( r->frame );
		if( r->e0 ) g.cuEventDestroy( r->e0 );
		if( r->e1 ) g.cuEventDestroy( r->e1 );
		if( r->mod ) g.cuModuleUnload( r->mod );
		g.cuDevicePrimaryCtxRelease( r->dev );
	}
	delete r;
}

/* Shader::Shader + DynamicAllocatorGPU::setup( 16 * 256, RENDER_NUMBER_OF_THREAD, 32 ) (PathTracer.hpp:43-69, voxRTGPU.cpp:104-106) */
int refgpu_open( const char* cubinPath, int device, RefGpu** out )
{
	*out = nullptr;
	if( !loadApi() ) return -2;
	CK( g.cuInit( 0 ) );
	RefGpu* r = new RefGpu;
	CUresult e;
#define CKR( call )                  \
	if( ( e = ( call ) ) != CUDA_SUCCESS ) \
	{                                \
		fail( e, #call );            \
		refgpu_close( r );           \
		return -2;                   \
	}
	CKR( g.cuDeviceGet( &r->dev, device ) );
	CKR( g.cuDevicePrimaryCtxRetain( &r->ctx, r->dev ) );
	CKR( g.cuCtxSetCurrent( r->ctx ) );
	CKR( g.cuModuleLoad( &r->mod, cubinPath ) );
	CKR( g.cuModuleGetFunction( &r->render, r->mod, "render" ) );
	CKR( g.cuModuleGetFunction( &r->storeImportance, r->mod, "HDRIstoreImportance" ) );
	CKR( g.cuModuleGetFunction( &r->satH, r->mod, "buildSATh" ) );
	CKR( g.cuModuleGetFunction( &r->satV, r->mod, "buildSATv" ) );
	CKR( g.cuModuleGetFunction( &r->sat2u32, r->mod, "buildSAT2u32" ) );
	CKR( g.cuModuleGetFunction( &r->renderPT, r->mod, "renderPT" ) );
	CKR( g.cuModuleGetFunction( &r->resolve, r->mod, "renderResolve" ) );
	CKR( g.cuEventCreate( &r->e0, CU_EVENT_DEFAULT ) );
	CKR( g.cuEventCreate( &r->e1, CU_EVENT_DEFAULT ) );
	int bad = 0;
	bad |= checkParams( r->render, "render", { 8, 8, 8, 32, 60, 64, 4 } );
	bad |= checkParams( r->storeImportance, "HDRIstoreImportance", { 8, 8, 8, 4, 12 } );
	bad |= checkParams( r->satH, "buildSATh", { 8, 8 } );
	bad |= checkParams( r->satV, "buildSATv", { 8, 8 } );
	bad |= checkParams( r->sat2u32, "buildSAT2u32", { 8, 8, 4 } );
	bad |= checkParams( r->renderPT, "renderPT", { 4, 8, 8, 60, 64, 32, 96, 8 } );
	bad |= checkParams( r->resolve, "renderResolve", { 8, 8, 4 } );
	if( bad )
	{
		refgpu_close( r );
		return -1;
	}
	/* stack pool: 4096 blocks x 256 threads x 32 StackElement (32 B) = 1 GiB, PathTracer.hpp:46 */
	r->alloc.numberOfBlock = 16 * 256;
	r->alloc.blockSize = 256;
	r->alloc.nElementPerThread = 32;
	CKR( g.cuMemAlloc( &r->alloc.memory, (size_t)32 * r->alloc.nElementPerThread * r->alloc.blockSize * r->alloc.numberOfBlock ) );
	CKR( g.cuMemAlloc( &r->alloc.locks, sizeof( uint32_t ) * r->alloc.numberOfBlock ) );
	CKR( g.cuMemsetD32( r->alloc.locks, 0, r->alloc.numberOfBlock ) );
	CKR( g.cuMemAlloc( &r->counter, 4 ) );
	r->isect.emissionScale = 7.5f; /* IntersectorOctreeGPU.hpp:273 */
	r->hdri.scale = 1.75f;		   /* renderCommon.hpp:480 */
#undef CKR
	*out = r;
	return 0;
}

/* the POD an IntersectorOctreeGPU::build leaves behind (IntersectorOctreeGPU.hpp:75-77,136,209-210): nodes in the reference's
 * 68-byte layout with masks embedded, root last; attrs = VoxelAttirb[nVoxels] */
int refgpu_set_octree( RefGpu* r, const OrcNode68* nodes, uint32_t nNodes, const uint8_t* attrs, uint32_t nVoxels, const float lower[3],
					   const float upper[3], float dps, uint32_t hasEmission )
{
	CK( g.cuCtxSetCurrent( r->ctx ) );
	freeIf( r->isect.nodeBuffer );
	freeIf( r->isect.vAttributeBuffer );
	CK( g.cuMemAlloc( &r->isect.nodeBuffer, (size_t)68 * ( nNodes ? nNodes : 1 ) ) );
	CK( g.cuMemAlloc( &r->isect.vAttributeBuffer, (size_t)8 * ( nVoxels ? nVoxels : 1 ) ) );
	if( nNodes ) CK( g.cuMemcpyHtoD( r->isect.nodeBuffer, nodes, (size_t)68 * nNodes ) );
	if( nVoxels ) CK( g.cuMemcpyHtoD( r->isect.vAttributeBuffer, attrs, (size_t)8 * nVoxels ) );
	r->isect.numberOfNodes = nNodes;
	r->isect.numberOfVoxels = nVoxels;
	r->isect.lower = { lower[0], lower[1], lower[2] };
	r->isect.upper = { upper[0], upper[1], upper[2] };
	r->isect.dps = dps;
	r->isect.hasEmission = hasEmission;
	return 0;
}

/* replace the attribute array only (a test encodes the voxel rank in the colour bytes to read hit ids out of `render`) */
int refgpu_set_attrs( RefGpu* r, const uint8_t* attrs, uint32_t nVoxels )
{
	if( nVoxels != r->isect.numberOfVoxels )
	{
		g_err = "refgpu_set_attrs: voxel count differs from the octree's";
		return -1;
	}
	CK( g.cuCtxSetCurrent( r->ctx ) );
	if( nVoxels ) CK( g.cuMemcpyHtoD( r->isect.vAttributeBuffer, attrs, (size_t)8 * nVoxels ) );
	return 0;
}

/* voxRTGPU.cpp:176-207: `reps` launches of `render`, each timed with events; the frame of the last one is copied out */
int refgpu_render( RefGpu* r, const OrcCamera* cam, int w, int h, int showVertexColor, uint8_t* rgba, int reps, float* msBest, float* msMean )
{
	CK( g.cuCtxSetCurrent( r->ctx ) );
	size_t bytes = (size_t)w * h * 4;
	if( r->frameBytes < bytes )
	{
		freeIf( r->frame );
		CK( g.cuMemAlloc( &r->frame, bytes ) );
		r->frameBytes = bytes;
	}
	i2 res = { w, h };
	int show = showVertexColor;
	OrcCamera c = *cam;
	void* args[] = { &r->frame, &res, &r->counter, &r->alloc, &c, &r->isect, &show };
	unsigned grid = ( (unsigned)( w * h ) + 255u ) / 256u;
	float best = 1e30f, sum = 0.0f;
	for( int i = 0; i < ( reps > 0 ? reps : 1 ); i++ )
	{
		CK( g.cuMemsetD32( r->counter, 0, 1 ) );
		CK( g.cuEventRecord( r->e0, 0 ) );
		CK( g.cuLaunchKernel( r->render, grid, 1, 1, 256, 1, 1, 0, 0, args, nullptr ) );
		CK( g.cuEventRecord( r->e1, 0 ) );
		CK( g.cuEventSynchronize( r->e1 ) );
		float ms = 0;
		CK( g.cuEventElapsedTime( &ms, r->e0, r->e1 ) );
		best = ms < best ? ms : best;
		sum += ms;
	}
	if( msBest ) *msBest = best;
	if( msMean ) *msMean = sum / (float)( reps > 0 ? reps : 1 );
	if( rgba ) CK( g.cuMemcpyDtoH( rgba, r->frame, bytes ) );
	return 0;
}

/* HDRI::load + HDRI::loadPrimary, renderCommon.hpp:214-326 (SAT_BLOCK_SIZE 512); *ms = device time of the 7 table builds */
int refgpu_load_hdri( RefGpu* r, const float* rgba, int w, int h, const float* rgbaPrimary, int wP, int hP, float* ms )
{
	CK( g.cuCtxSetCurrent( r->ctx ) );
	freeIf( r->hdri.pixels );
	freeIf( r->hdri.pixelsPrimary );
	freeIf( r->hdri.sat );
	for( auto& s : r->hdri.sats ) freeIf( s );
	r->hdri.width = w;
	r->hdri.height = h;
	r->hdri.widthPrimary = r->hdri.heightPrimary = 0;
	size_t n = (size_t)w * h;
	CK( g.cuMemAlloc( &r->hdri.pixels, n * 16 ) );
	CK( g.cuMemcpyHtoD( r->hdri.pixels, rgba, n * 16 ) );
	CUdeviceptr f64 = 0;
	CK( g.cuMemAlloc( &f64, n * 8 ) );
	CK( g.cuEventRecord( r->e0, 0 ) );
	const f3 axes[7] = { { 0, 0, 0 }, { 1, 0, 0 }, { -1, 0, 0 }, { 0, 1, 0 }, { 0, -1, 0 }, { 0, 0, 1 }, { 0, 0, -1 } };
	for( int k = 0; k < 7; k++ )
	{
		CUdeviceptr* dst = k == 0 ? &r->hdri.sat : &r->hdri.sats[k - 1];
		CK( g.cuMemAlloc( dst, n * 4 ) );
		i2 res = { w, h };
		int cosWeighted = k == 0 ? 0 : 1;
		f3 axis = axes[k];
		void* a0[] = { &r->hdri.pixels, &res, &f64, &cosWeighted, &axis };
		CK( g.cuLaunchKernel( r->storeImportance, ( w + 7 ) / 8, ( h + 7 ) / 8, 1, 8, 8, 1, 0, 0, a0, nullptr ) );
		void* a1[] = { &res, &f64 };
		CK( g.cuLaunchKernel( r->satH, h, 1, 1, 512, 1, 1, 0, 0, a1, nullptr ) );
		CK( g.cuLaunchKernel( r->satV, w, 1, 1, 512, 1, 1, 0, 0, a1, nullptr ) );
		int cnt = (int)n;
		void* a2[] = { dst, &f64, &cnt };
		CK( g.cuLaunchKernel( r->sat2u32, (unsigned)( ( n + 63 ) / 64 ), 1, 1, 64, 1, 1, 0, 0, a2, nullptr ) );
	}
	CK( g.cuEventRecord( r->e1, 0 ) );
	CK( g.cuEventSynchronize( r->e1 ) );
	if( ms ) CK( g.cuEventElapsedTime( ms, r->e0, r->e1 ) );
	g.cuMemFree( f64 );
	if( rgbaPrimary )
	{
		size_t nP = (size_t)wP * hP;
		CK( g.cuMemAlloc( &r->hdri.pixelsPrimary, nP * 16 ) );
		CK( g.cuMemcpyHtoD( r->hdri.pixelsPrimary, rgbaPrimary, nP * 16 ) );
		r->hdri.widthPrimary = wP;
		r->hdri.heightPrimary = hP;
	}
	return 0;
}

/* the 7 tables (global, +x, -x, +y, -y, +z, -z), w*h uint32 each */
int refgpu_get_sats( RefGpu* r, uint32_t* out )
{
	CK( g.cuCtxSetCurrent( r->ctx ) );
	size_t n = (size_t)r->hdri.width * r->hdri.height;
	if( !r->hdri.sat )
	{
		g_err = "refgpu_get_sats: no HDRI loaded";
		return -4;
	}
	for( int k = 0; k < 7; k++ ) CK( g.cuMemcpyDtoH( out + n * k, k == 0 ? r->hdri.sat : r->hdri.sats[k - 1], n * 4 ) );
	return 0;
}

/* PMJSampler::setup( true ) (pmjSampler.hpp:114-134) minus the table generation: the caller passes the table
 * (orc_pmj_table of the "reference" oracle runs the reference's own host code for it) */
int refgpu_set_pmj( RefGpu* r, const float* table )
{
	CK( g.cuCtxSetCurrent( r->ctx ) );
	size_t bytes = sizeof( float ) * 2 * 4096 * 128;
	if( !r->pmj ) CK( g.cuMemAlloc( &r->pmj, bytes ) );
	CK( g.cuMemcpyHtoD( r->pmj, table, bytes ) );
	return 0;
}

/* resizeFrameBufferIfNeeded + clearFrameBuffer, PathTracer.hpp:81-103.  renderPT writes frameBuffer[blockIdx * 256 + tid] without
 * a bounds check (voxKernel.cu:770-774), so the allocation is rounded up to whole blocks. */
int refgpu_pt_resize( RefGpu* r, int w, int h )
{
	CK( g.cuCtxSetCurrent( r->ctx ) );
	size_t n = ( ( (size_t)w * h + 255 ) / 256 ) * 256;
	freeIf( r->fbF32 );
	freeIf( r->fbU8 );
	CK( g.cuMemAlloc( &r->fbF32, n * 16 ) );
	CK( g.cuMemAlloc( &r->fbU8, n * 4 ) );
	CK( g.cuMemsetD32( r->fbF32, 0, n * 4 ) );
	r->w = w;
	r->h = h;
	return 0;
}
int refgpu_pt_clear( RefGpu* r )
{
	CK( g.cuCtxSetCurrent( r->ctx ) );
	size_t n = ( ( (size_t)r->w * r->h + 255 ) / 256 ) * 256;
	CK( g.cuMemsetD32( r->fbF32, 0, n * 4 ) );
	return 0;
}

/* PathTracer::step, PathTracer.hpp:158-168: one launch = 16 spp; `iteration` is m_steps.  hdriScale <= 0 disables the
 * environment (HDRI::isEnabled, renderCommon.hpp:467-470). */
int refgpu_pt_step( RefGpu* r, const OrcCamera* cam, int iteration, float* ms )
{
	CK( g.cuCtxSetCurrent( r->ctx ) );
	if( !r->fbF32 || !r->pmj || !r->isect.nodeBuffer || !r->hdri.sat )
	{
		g_err = "refgpu_pt_step: octree / HDRI / PMJ table / frame buffer not set";
		return -4;
	}
	i2 res = { r->w, r->h };
	OrcCamera c = *cam;
	int it = iteration;
	void* args[] = { &it, &r->fbF32, &res, &c, &r->isect, &r->alloc, &r->hdri, &r->pmj };
	unsigned grid = ( (unsigned)( r->w * r->h ) + 255u ) / 256u;
	CK( g.cuEventRecord( r->e0, 0 ) );
	CK( g.cuLaunchKernel( r->renderPT, grid, 1, 1, 256, 1, 1, 0, 0, args, nullptr ) );
	CK( g.cuEventRecord( r->e1, 0 ) );
	CK( g.cuEventSynchronize( r->e1 ) );
	if( ms ) CK( g.cuEventElapsedTime( ms, r->e0, r->e1 ) );
	return 0;
}

int refgpu_pt_read_accum( RefGpu* r, float* rgbw )
{
	CK( g.cuCtxSetCurrent( r->ctx ) );
	CK( g.cuMemcpyDtoH( rgbw, r->fbF32, (size_t)r->w * r->h * 16 ) );
	return 0;
}
/* overwrite the accumulators (to run renderResolve on somebody else's sums) */
int refgpu_pt_write_accum( RefGpu* r, const float* rgbw )
{
	CK( g.cuCtxSetCurrent( r->ctx ) );
	CK( g.cuMemcpyHtoD( r->fbF32, rgbw, (size_t)r->w * r->h * 16 ) );
	return 0;
}

/* PathTracer::toImageAsync, PathTracer.hpp:118-129 */
int refgpu_pt_resolve( RefGpu* r, uint8_t* rgba )
{
	CK( g.cuCtxSetCurrent( r->ctx ) );
	int n = r->w * r->h;
	void* args[] = { &r->fbU8, &r->fbF32, &n };
	CK( g.cuLaunchKernel( r->resolve, ( n + 127 ) / 128, 1, 1, 128, 1, 1, 0, 0, args, nullptr ) );
	CK( g.cuMemcpyDtoH( rgba, r->fbU8, (size_t)n * 4 ) );
	return 0;
}

void refgpu_set_scales( RefGpu* r, float emissionScale, float hdriScale )
{
	r->isect.emissionScale = emissionScale;
	r->hdri.scale = hdriScale;
}

} /* extern "C" */
