// stubs.cu -- TEMPORARY: entry points whose stage has not landed yet fail loudly (no fallback).
#include "internal.cuh"
#define NOT_YET( name ) { hako::setError( name ": not implemented yet" ); return HAKO_E_STATE; }
extern "C" {
int hako_pt_create( hako_ctx*, hako_stream, hako_pt** ) NOT_YET( "hako_pt_create" )
void hako_pt_destroy( hako_pt* ) {}
int hako_pt_resize( hako_pt*, hako_stream, int, int ) NOT_YET( "hako_pt" )
int hako_pt_clear( hako_pt*, hako_stream ) NOT_YET( "hako_pt" )
int hako_pt_load_hdri( hako_pt*, hako_stream, const float*, int, int, const float*, int, int ) NOT_YET( "hako_pt" )
int hako_pt_step( hako_pt*, const hako_octree*, hako_stream, const hako_camera*, int, int, int, int ) NOT_YET( "hako_pt" )
int hako_pt_resolve( hako_pt*, hako_stream ) NOT_YET( "hako_pt" )
int hako_pt_read_rgba8( hako_pt*, hako_stream, uint8_t* ) NOT_YET( "hako_pt" )
int hako_pt_read_accum( hako_pt*, hako_stream, float* ) NOT_YET( "hako_pt" )
int hako_pt_accum_device( hako_pt*, void**, uint64_t* ) NOT_YET( "hako_pt" )
int hako_pt_ray_count( hako_pt*, hako_stream, uint64_t*, int ) NOT_YET( "hako_pt" )
int hako_pt_pmj_table( hako_pt*, hako_stream, float* ) NOT_YET( "hako_pt" )
int hako_pt_hdri_sats( hako_pt*, hako_stream, uint32_t* ) NOT_YET( "hako_pt" )
}
