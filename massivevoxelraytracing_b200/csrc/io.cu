// io.cu -- host-side file formats around the hot path: no device code in this file.
//
//   hako_image_load_hdr      Radiance RGBE -> RGBA32F, what PathTracer::loadHDRI gets from stbi_loadf( file, .., 4 )
//                            (PathTracer.hpp:104-116, libs/prlib/src/prp.cpp:353-366; stb_image.h stbi__hdr_load / stbi__hdr_convert)
//   hako_image_save_png      Image2DRGBA8::saveAsPngUncompressed (libs/prlib/src/prp.cpp:215-223, libattopng): RGBA8, stored deflate blocks
//   hako_mesh_load           triangle soup from Wavefront OBJ (what prlib reads through tiny_obj_loader, libs/prlib/src/prg.cpp:1231-1241)
//                            or Stanford PLY (the xyzrgb dragon's native format; the reference consumed an Alembic conversion)
//   hako_mesh_flatten        trianglesFlattened (voxUtil.hpp:8-65): 3 vertices per face, colour = Cd or 1, emission = 0
//   hako_voxel_mesh_save_ply saveVoxelsAsMesh (voxMesh.cpp:111-219): 8 corners per voxel + one quad per exposed face
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "internal.cuh"

using namespace hako;

#define IO_ARG( cond, msg )              \
	do                                   \
	{                                    \
		if( !( cond ) )                  \
		{                                \
			hako::setError( "%s", msg ); \
			return HAKO_E_INVALID;       \
		}                                \
	} while( 0 )

struct hako_mesh
{
	std::vector<float> positions; // xyz per vertex
	std::vector<float> colors;	  // rgb per vertex in [0,1], empty when the file has none
	std::vector<uint32_t> indices; // 3 per triangle
};

namespace
{
bool readFile( const char* path, std::vector<uint8_t>& out )
{
	FILE* fp = fopen( path, "rb" );
	if( !fp ) return false;
	fseek( fp, 0, SEEK_END );
	const long n = ftell( fp );
	fseek( fp, 0, SEEK_SET );
	out.resize( n > 0 ? (size_t)n : 0 );
	const size_t got = out.empty() ? 0 : fread( out.data(), 1, out.size(), fp );
	fclose( fp );
	return got == out.size();
}

// one text line starting at `pos` (without the terminator); false at the end of the buffer
bool nextLine( const std::vector<uint8_t>& buf, size_t& pos, std::string& line )
{
	if( pos >= buf.size() ) return false;
	size_t e = pos;
	while( e < buf.size() && buf[e] != '\n' ) e++;
	line.assign( (const char*)buf.data() + pos, e - pos );
	if( !line.empty() && line.back() == '\r' ) line.pop_back();
	pos = e + 1;
	return true;
}

// ---- CRC-32 / Adler-32 for the PNG container ----
uint32_t crc32( const uint8_t* p, size_t n, uint32_t crc )
{
	static uint32_t table[256];
	static bool ready = false;
	if( !ready )
	{
		for( uint32_t i = 0; i < 256; i++ )
		{
			uint32_t c = i;
			for( int k = 0; k < 8; k++ ) c = ( c & 1u ) ? 0xEDB88320u ^ ( c >> 1 ) : c >> 1;
			table[i] = c;
		}
		ready = true;
	}
	crc = ~crc;
	for( size_t i = 0; i < n; i++ ) crc = table[( crc ^ p[i] ) & 0xFFu] ^ ( crc >> 8 );
	return ~crc;
}
void put32be( std::vector<uint8_t>& v, uint32_t x )
{
	v.push_back( (uint8_t)( x >> 24 ) ), v.push_back( (uint8_t)( x >> 16 ) ), v.push_back( (uint8_t)( x >> 8 ) ), v.push_back( (uint8_t)x );
}
void pngChunk( std::vector<uint8_t>& out, const char type[4], const std::vector<uint8_t>& data )
{
	put32be( out, (uint32_t)data.size() );
	const size_t start = out.size();
	out.insert( out.end(), type, type + 4 );
	out.insert( out.end(), data.begin(), data.end() );
	put32be( out, crc32( out.data() + start, out.size() - start, 0 ) );
}
} // namespace

extern "C" {

// ---------------------------------------------------------------- Radiance .hdr
int hako_image_load_hdr( const char* path, float** rgba, int* w, int* h )
{
	IO_ARG( path && rgba && w && h, "hako_image_load_hdr: null argument" );
	*rgba = nullptr;
	*w = *h = 0;
	std::vector<uint8_t> buf;
	if( !readFile( path, buf ) )
	{
		setError( "hako_image_load_hdr: cannot read %s", path );
		return HAKO_E_INVALID;
	}
	size_t pos = 0;
	std::string line;
	if( !nextLine( buf, pos, line ) || ( line != "#?RADIANCE" && line != "#?RGBE" ) )
	{
		setError( "hako_image_load_hdr: %s is not a Radiance file", path );
		return HAKO_E_INVALID;
	}
	bool formatOk = false;
	while( nextLine( buf, pos, line ) )
	{
		if( line.empty() ) break;
		if( line == "FORMAT=32-bit_rle_rgbe" ) formatOk = true;
	}
	int width = 0, height = 0;
	if( !formatOk || !nextLine( buf, pos, line ) || sscanf( line.c_str(), "-Y %d +X %d", &height, &width ) != 2 || width <= 0 || height <= 0 )
	{
		setError( "hako_image_load_hdr: unsupported header in %s (need FORMAT=32-bit_rle_rgbe and a \"-Y h +X w\" resolution line)", path );
		return HAKO_E_INVALID;
	}
	float* out = (float*)malloc( (size_t)width * height * 4 * sizeof( float ) );
	if( !out )
	{
		setError( "hako_image_load_hdr: out of memory" );
		return HAKO_E_INVALID;
	}
	// stbi__hdr_convert: mantissa * 2^(e - 136), alpha = 1; e == 0 is black
	auto convert = []( float* dst, const uint8_t* rgbe ) {
		if( rgbe[3] != 0 )
		{
			const float f = ldexpf( 1.0f, (int)rgbe[3] - ( 128 + 8 ) );
			dst[0] = rgbe[0] * f, dst[1] = rgbe[1] * f, dst[2] = rgbe[2] * f;
		}
		else
			dst[0] = dst[1] = dst[2] = 0.0f;
		dst[3] = 1.0f;
	};
	const uint8_t* p = buf.data() + pos;
	const uint8_t* end = buf.data() + buf.size();
	bool ok = true;
	std::vector<uint8_t> scan( (size_t)width * 4 );
	const bool maybeRle = width >= 8 && width < 32768;
	bool flat = !maybeRle;
	if( maybeRle && end - p >= 4 && !( p[0] == 2 && p[1] == 2 && ( p[2] & 0x80 ) == 0 ) ) flat = true; // stb: the first scanline decides
	if( flat )
	{
		ok = (size_t)( end - p ) >= (size_t)width * height * 4;
		for( size_t i = 0; ok && i < (size_t)width * height; i++ ) convert( out + i * 4, p + i * 4 );
	}
	else
	{
		for( int y = 0; ok && y < height; y++ )
		{
			if( end - p < 4 || p[0] != 2 || p[1] != 2 || ( ( (int)p[2] << 8 ) | p[3] ) != width )
			{
				ok = false;
				break;
			}
			p += 4;
			for( int c = 0; ok && c < 4; c++ )
			{
				int x = 0;
				while( x < width )
				{
					if( p >= end )
					{
						ok = false;
						break;
					}
					int count = *p++;
					if( count > 128 )
					{
						count -= 128; // a run
						if( p >= end || count > width - x )
						{
							ok = false;
							break;
						}
						const uint8_t v = *p++;
						for( int k = 0; k < count; k++ ) scan[( x++ ) * 4 + c] = v;
					}
					else
					{
						if( count == 0 || count > width - x || end - p < count )
						{
							ok = false;
							break;
						}
						for( int k = 0; k < count; k++ ) scan[( x++ ) * 4 + c] = *p++;
					}
				}
			}
			for( int x = 0; ok && x < width; x++ ) convert( out + ( (size_t)y * width + x ) * 4, scan.data() + (size_t)x * 4 );
		}
	}
	if( !ok )
	{
		free( out );
		setError( "hako_image_load_hdr: corrupt pixel data in %s", path );
		return HAKO_E_INVALID;
	}
	*rgba = out;
	*w = width;
	*h = height;
	return HAKO_OK;
}

void hako_image_free( float* rgba ) { free( rgba ); }

// ---------------------------------------------------------------- PNG (stored deflate blocks)
int hako_image_save_png( const char* path, const uint8_t* rgba, int w, int h )
{
	IO_ARG( path && rgba && w > 0 && h > 0, "hako_image_save_png: bad argument" );
	std::vector<uint8_t> raw;
	raw.reserve( (size_t)h * ( (size_t)w * 4 + 1 ) );
	for( int y = 0; y < h; y++ )
	{
		raw.push_back( 0 ); // filter: none
		raw.insert( raw.end(), rgba + (size_t)y * w * 4, rgba + (size_t)( y + 1 ) * w * 4 );
	}
	std::vector<uint8_t> z;
	z.reserve( raw.size() + raw.size() / 65535 * 5 + 16 );
	z.push_back( 0x78 ), z.push_back( 0x01 );
	uint32_t a = 1, b = 0;
	for( size_t off = 0; off < raw.size(); )
	{
		const size_t n = std::min<size_t>( 65535, raw.size() - off );
		z.push_back( off + n == raw.size() ? 1 : 0 );
		z.push_back( (uint8_t)( n & 0xFF ) ), z.push_back( (uint8_t)( n >> 8 ) );
		z.push_back( (uint8_t)( ~n & 0xFF ) ), z.push_back( (uint8_t)( ( ~n >> 8 ) & 0xFF ) );
		z.insert( z.end(), raw.begin() + off, raw.begin() + off + n );
		for( size_t i = 0; i < n; )
		{
			// Adler-32 with deferred modulo (5552 bytes keep the sums below 2^32)
			const size_t m = std::min<size_t>( 5552, n - i );
			for( size_t k = 0; k < m; k++ )
			{
				a += raw[off + i + k];
				b += a;
			}
			a %= 65521u, b %= 65521u;
			i += m;
		}
		off += n;
	}
	put32be( z, ( b << 16 ) | a );
	std::vector<uint8_t> out = { 0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A };
	std::vector<uint8_t> ihdr;
	put32be( ihdr, (uint32_t)w ), put32be( ihdr, (uint32_t)h );
	ihdr.push_back( 8 ), ihdr.push_back( 6 ), ihdr.push_back( 0 ), ihdr.push_back( 0 ), ihdr.push_back( 0 ); // 8-bit RGBA
	pngChunk( out, "IHDR", ihdr );
	pngChunk( out, "IDAT", z );
	pngChunk( out, "IEND", std::vector<uint8_t>() );
	FILE* fp = fopen( path, "wb" );
	if( !fp )
	{
		setError( "hako_image_save_png: cannot open %s", path );
		return HAKO_E_INVALID;
	}
	const bool ok = fwrite( out.data(), 1, out.size(), fp ) == out.size();
	fclose( fp );
	if( !ok )
	{
		setError( "hako_image_save_png: short write to %s", path );
		return HAKO_E_INVALID;
	}
	return HAKO_OK;
}

// ---------------------------------------------------------------- meshes
static int loadObj( const std::vector<uint8_t>& buf, hako_mesh* m )
{
	size_t pos = 0;
	std::string line;
	std::vector<long> face;
	while( nextLine( buf, pos, line ) )
	{
		const char* s = line.c_str();
		while( *s == ' ' || *s == '\t' ) s++;
		if( s[0] == 'v' && ( s[1] == ' ' || s[1] == '\t' ) )
		{
			// "v x y z [r g b]": the colour extension is what tiny_obj_loader exposes as attrib.colors
			float v[6];
			const int n = sscanf( s + 1, "%f %f %f %f %f %f", v, v + 1, v + 2, v + 3, v + 4, v + 5 );
			if( n < 3 )
			{
				setError( "hako_mesh_load: bad vertex line \"%s\"", line.c_str() );
				return HAKO_E_INVALID;
			}
			m->positions.insert( m->positions.end(), v, v + 3 );
			if( n == 6 )
			{
				if( m->colors.size() != m->positions.size() - 3 ) m->colors.resize( m->positions.size() - 3, 1.0f );
				m->colors.insert( m->colors.end(), v + 3, v + 6 );
			}
		}
		else if( s[0] == 'f' && ( s[1] == ' ' || s[1] == '\t' ) )
		{
			face.clear();
			const char* p = s + 1;
			for( ;; )
			{
				while( *p == ' ' || *p == '\t' ) p++;
				if( !*p ) break;
				char* e = nullptr;
				long idx = strtol( p, &e, 10 );
				if( e == p )
				{
					setError( "hako_mesh_load: bad face line \"%s\"", line.c_str() );
					return HAKO_E_INVALID;
				}
				const long nv = (long)( m->positions.size() / 3 );
				idx = idx > 0 ? idx - 1 : nv + idx; // negative = relative to the vertices read so far
				if( idx < 0 || idx >= nv )
				{
					setError( "hako_mesh_load: face index out of range in \"%s\"", line.c_str() );
					return HAKO_E_INVALID;
				}
				face.push_back( idx );
				p = e;
				while( *p && *p != ' ' && *p != '\t' ) p++; // skip /vt/vn
			}
			// triangle fan (the reference asserts triangles, voxUtil.hpp:39; tiny_obj_loader triangulates polygons the same way)
			for( size_t k = 2; k < face.size(); k++ )
			{
				m->indices.push_back( (uint32_t)face[0] );
				m->indices.push_back( (uint32_t)face[k - 1] );
				m->indices.push_back( (uint32_t)face[k] );
			}
		}
	}
	if( !m->colors.empty() && m->colors.size() != m->positions.size() ) m->colors.resize( m->positions.size(), 1.0f );
	return HAKO_OK;
}

namespace
{
struct PlyProp
{
	std::string name;
	int type; // bytes of a scalar; negative: float type of that size
	bool isList;
	int countType, itemType;
};
int plyTypeSize( const std::string& t )
{
	if( t == "char" || t == "int8" || t == "uchar" || t == "uint8" ) return 1;
	if( t == "short" || t == "int16" || t == "ushort" || t == "uint16" ) return 2;
	if( t == "int" || t == "int32" || t == "uint" || t == "uint32" ) return 4;
	if( t == "float" || t == "float32" ) return -4;
	if( t == "double" || t == "float64" ) return -8;
	return 0;
}
bool plyTypeSigned( const std::string& t ) { return t == "char" || t == "int8" || t == "short" || t == "int16" || t == "int" || t == "int32"; }
double plyReadScalar( const uint8_t* p, int type, bool isSigned, bool swap )
{
	uint8_t b[8];
	const int n = type < 0 ? -type : type;
	for( int i = 0; i < n; i++ ) b[i] = swap ? p[n - 1 - i] : p[i];
	if( type == -4 )
	{
		float f;
		memcpy( &f, b, 4 );
		return f;
	}
	if( type == -8 )
	{
		double d;
		memcpy( &d, b, 8 );
		return d;
	}
	uint64_t u = 0;
	memcpy( &u, b, n );
	if( isSigned )
	{
		if( n == 1 ) return (int8_t)u;
		if( n == 2 ) return (int16_t)u;
		return (int32_t)u;
	}
	return (double)u;
}
} // namespace

static int loadPly( const std::vector<uint8_t>& buf, hako_mesh* m )
{
	size_t pos = 0;
	std::string line;
	nextLine( buf, pos, line ); // "ply"
	int format = -1;			// 0 ascii, 1 little endian, 2 big endian
	struct Elem
	{
		std::string name;
		size_t count;
		std::vector<PlyProp> props;
		std::vector<bool> propSigned, countSigned, itemSigned;
	};
	std::vector<Elem> elems;
	bool headerDone = false;
	while( nextLine( buf, pos, line ) )
	{
		char a[64] = "", b[64] = "", c[64] = "", d[64] = "", e[64] = "";
		const int n = sscanf( line.c_str(), "%63s %63s %63s %63s %63s", a, b, c, d, e );
		if( n >= 1 && !strcmp( a, "end_header" ) )
		{
			headerDone = true;
			break;
		}
		if( n >= 2 && !strcmp( a, "format" ) ) format = !strcmp( b, "ascii" ) ? 0 : ( !strcmp( b, "binary_little_endian" ) ? 1 : ( !strcmp( b, "binary_big_endian" ) ? 2 : -1 ) );
		else if( n >= 3 && !strcmp( a, "element" ) )
		{
			Elem el;
			el.name = b;
			el.count = (size_t)strtoull( c, nullptr, 10 );
			elems.push_back( el );
		}
		else if( n >= 3 && !strcmp( a, "property" ) && !elems.empty() )
		{
			Elem& el = elems.back();
			PlyProp p;
			if( !strcmp( b, "list" ) && n >= 5 )
			{
				p.isList = true, p.countType = plyTypeSize( c ), p.itemType = plyTypeSize( d ), p.type = 0, p.name = e;
				el.countSigned.push_back( plyTypeSigned( c ) ), el.itemSigned.push_back( plyTypeSigned( d ) ), el.propSigned.push_back( false );
				if( !p.countType || !p.itemType ) format = -1;
			}
			else
			{
				p.isList = false, p.type = plyTypeSize( b ), p.countType = p.itemType = 0, p.name = c;
				el.countSigned.push_back( false ), el.itemSigned.push_back( false ), el.propSigned.push_back( plyTypeSigned( b ) );
				if( !p.type ) format = -1;
			}
			el.props.push_back( p );
		}
	}
	if( !headerDone || format < 0 )
	{
		setError( "hako_mesh_load: unsupported PLY header" );
		return HAKO_E_INVALID;
	}
	const bool swap = format == 2;
	const uint8_t* p = buf.data() + pos;
	const uint8_t* end = buf.data() + buf.size();
	// ascii: tokenise the remainder once
	std::vector<double> tokens;
	size_t tok = 0;
	if( format == 0 )
	{
		const char* s = (const char*)p;
		const char* se = (const char*)end;
		std::string body( s, se );
		char* q = &body[0];
		for( ;; )
		{
			char* e2 = nullptr;
			const double v = strtod( q, &e2 );
			if( e2 == q ) break;
			tokens.push_back( v );
			q = e2;
		}
	}
	auto fail = []() {
		setError( "hako_mesh_load: truncated PLY body" );
		return HAKO_E_INVALID;
	};
	std::vector<double> item;
	std::vector<double> rgb( 3 );
	for( const Elem& el : elems )
	{
		const bool isVertex = el.name == "vertex", isFace = el.name == "face";
		int ix = -1, iy = -1, iz = -1, ir = -1, ig = -1, ib = -1;
		bool colourIsByte = true;
		for( size_t k = 0; k < el.props.size(); k++ )
		{
			const std::string& nm = el.props[k].name;
			if( nm == "x" ) ix = (int)k;
			if( nm == "y" ) iy = (int)k;
			if( nm == "z" ) iz = (int)k;
			if( nm == "red" || nm == "diffuse_red" ) ir = (int)k, colourIsByte = el.props[k].type == 1;
			if( nm == "green" || nm == "diffuse_green" ) ig = (int)k;
			if( nm == "blue" || nm == "diffuse_blue" ) ib = (int)k;
		}
		if( isVertex && ( ix < 0 || iy < 0 || iz < 0 ) )
		{
			setError( "hako_mesh_load: PLY vertex element without x / y / z" );
			return HAKO_E_INVALID;
		}
		const bool hasColour = isVertex && ir >= 0 && ig >= 0 && ib >= 0;
		std::vector<double> vals( el.props.size() );
		for( size_t i = 0; i < el.count; i++ )
		{
			for( size_t k = 0; k < el.props.size(); k++ )
			{
				const PlyProp& pr = el.props[k];
				if( !pr.isList )
				{
					if( format == 0 )
					{
						if( tok >= tokens.size() ) return fail();
						vals[k] = tokens[tok++];
					}
					else
					{
						const int n = pr.type < 0 ? -pr.type : pr.type;
						if( end - p < n ) return fail();
						vals[k] = plyReadScalar( p, pr.type, el.propSigned[k], swap );
						p += n;
					}
					continue;
				}
				size_t cnt;
				if( format == 0 )
				{
					if( tok >= tokens.size() ) return fail();
					cnt = (size_t)tokens[tok++];
				}
				else
				{
					const int n = pr.countType < 0 ? -pr.countType : pr.countType;
					if( end - p < n ) return fail();
					cnt = (size_t)plyReadScalar( p, pr.countType, el.countSigned[k], swap );
					p += n;
				}
				item.resize( cnt );
				for( size_t j = 0; j < cnt; j++ )
				{
					if( format == 0 )
					{
						if( tok >= tokens.size() ) return fail();
						item[j] = tokens[tok++];
					}
					else
					{
						const int n = pr.itemType < 0 ? -pr.itemType : pr.itemType;
						if( end - p < n ) return fail();
						item[j] = plyReadScalar( p, pr.itemType, el.itemSigned[k], swap );
						p += n;
					}
				}
				if( isFace && ( pr.name == "vertex_indices" || pr.name == "vertex_index" ) )
				{
					const size_t nv = m->positions.size() / 3;
					for( size_t j = 0; j < cnt; j++ )
						if( item[j] < 0 || (size_t)item[j] >= nv )
						{
							setError( "hako_mesh_load: PLY face index out of range" );
							return HAKO_E_INVALID;
						}
					for( size_t j = 2; j < cnt; j++ )
					{
						m->indices.push_back( (uint32_t)item[0] );
						m->indices.push_back( (uint32_t)item[j - 1] );
						m->indices.push_back( (uint32_t)item[j] );
					}
				}
			}
			if( isVertex )
			{
				m->positions.push_back( (float)vals[ix] ), m->positions.push_back( (float)vals[iy] ), m->positions.push_back( (float)vals[iz] );
				if( hasColour )
				{
					const float sc = colourIsByte ? 1.0f / 255.0f : 1.0f;
					m->colors.push_back( (float)vals[ir] * sc ), m->colors.push_back( (float)vals[ig] * sc ), m->colors.push_back( (float)vals[ib] * sc );
				}
			}
		}
	}
	return HAKO_OK;
}

int hako_mesh_load( const char* path, hako_mesh** out )
{
	IO_ARG( path && out, "hako_mesh_load: null argument" );
	*out = nullptr;
	std::vector<uint8_t> buf;
	if( !readFile( path, buf ) )
	{
		setError( "hako_mesh_load: cannot read %s", path );
		return HAKO_E_INVALID;
	}
	hako_mesh* m = new hako_mesh();
	const bool ply = buf.size() >= 3 && !memcmp( buf.data(), "ply", 3 );
	const int r = ply ? loadPly( buf, m ) : loadObj( buf, m );
	if( r )
	{
		delete m;
		return r;
	}
	*out = m;
	return HAKO_OK;
}
void hako_mesh_destroy( hako_mesh* m ) { delete m; }
int hako_mesh_info( const hako_mesh* m, uint64_t* nVertices, uint64_t* nTriangles, int* hasColors )
{
	IO_ARG( m, "hako_mesh_info: null mesh" );
	if( nVertices ) *nVertices = m->positions.size() / 3;
	if( nTriangles ) *nTriangles = m->indices.size() / 3;
	if( hasColors ) *hasColors = m->colors.empty() ? 0 : 1;
	return HAKO_OK;
}
// trianglesFlattened, voxUtil.hpp:8-65 (identity localToWorld): colour = the vertex colour ("Cd") or 1, emission = 0
int hako_mesh_flatten( const hako_mesh* m, float* pos, float* col, float* emi )
{
	IO_ARG( m && pos, "hako_mesh_flatten: null argument" );
	const size_t n = m->indices.size();
	for( size_t i = 0; i < n; i++ )
	{
		const uint32_t v = m->indices[i];
		memcpy( pos + i * 3, m->positions.data() + (size_t)v * 3, 12 );
		if( col )
		{
			if( m->colors.empty() )
				col[i * 3] = col[i * 3 + 1] = col[i * 3 + 2] = 1.0f;
			else
				memcpy( col + i * 3, m->colors.data() + (size_t)v * 3, 12 );
		}
		if( emi ) emi[i * 3] = emi[i * 3 + 1] = emi[i * 3 + 2] = 0.0f;
	}
	return HAKO_OK;
}

// ---------------------------------------------------------------- saveVoxelsAsMesh, voxMesh.cpp:111-219
int hako_voxel_mesh_save_ply( const char* path, const uint64_t* sortedMortons, uint64_t n, const float origin[3], float dps )
{
	IO_ARG( path && origin && ( n == 0 || sortedMortons ), "hako_voxel_mesh_save_ply: null argument" );
	IO_ARG( n * 8 < 0xFFFFFFFFull, "hako_voxel_mesh_save_ply: more than 2^32 / 8 voxels (the reference indexes corners with uint32)" );
	auto compact = []( uint64_t x ) { // inverse of splitBy3 (morton.hpp:117-134 decodes with PEXT)
		x &= 0x1249249249249249ull;
		x = ( x ^ ( x >> 2 ) ) & 0x10c30c30c30c30c3ull;
		x = ( x ^ ( x >> 4 ) ) & 0x100f00f00f00f00full;
		x = ( x ^ ( x >> 8 ) ) & 0x1f0000ff0000ffull;
		x = ( x ^ ( x >> 16 ) ) & 0x1f00000000ffffull;
		x = ( x ^ ( x >> 32 ) ) & 0x1fffffull;
		return (uint32_t)x;
	};
	auto present = [&]( uint32_t x, uint32_t y, uint32_t z ) { return std::binary_search( sortedMortons, sortedMortons + n, mortonEncode( x, y, z ) ); };
	std::vector<float> points;
	points.reserve( n * 24 );
	std::vector<uint8_t> faces;
	uint32_t nface = 0;
	for( uint64_t i = 0; i < n; i++ )
	{
		const uint64_t mc = sortedMortons[i];
		const uint32_t x = compact( mc ), y = compact( mc >> 1 ), z = compact( mc >> 2 );
		const float px = origin[0] + x * dps, py = origin[1] + y * dps, pz = origin[2] + z * dps;
		const float c[8][3] = { { px, py, pz }, { px + dps, py, pz }, { px + dps, py, pz + dps }, { px, py, pz + dps },
								{ px, py + dps, pz }, { px + dps, py + dps, pz }, { px + dps, py + dps, pz + dps }, { px, py + dps, pz + dps } };
		points.insert( points.end(), &c[0][0], &c[0][0] + 24 );
		const uint32_t b = (uint32_t)i * 8;
		auto quad = [&]( uint32_t i0, uint32_t i1, uint32_t i2, uint32_t i3 ) {
			const uint32_t q[4] = { b + i0, b + i1, b + i2, b + i3 };
			faces.push_back( 4 );
			const uint8_t* qb = (const uint8_t*)q;
			faces.insert( faces.end(), qb, qb + 16 );
			nface++;
		};
		if( y == 0 || !present( x, y - 1, z ) ) quad( 3, 2, 1, 0 );
		if( !present( x, y + 1, z ) ) quad( 4, 5, 6, 7 );
		if( z == 0 || !present( x, y, z - 1 ) ) quad( 0, 1, 5, 4 );
		if( !present( x + 1, y, z ) ) quad( 1, 2, 6, 5 );
		if( !present( x, y, z + 1 ) ) quad( 2, 3, 7, 6 );
		if( x == 0 || !present( x - 1, y, z ) ) quad( 3, 0, 4, 7 );
	}
	FILE* fp = fopen( path, "wb" );
	if( !fp )
	{
		setError( "hako_voxel_mesh_save_ply: cannot open %s", path );
		return HAKO_E_INVALID;
	}
	fprintf( fp, "ply\nformat binary_little_endian 1.0\nelement vertex %llu\nproperty float x\nproperty float y\nproperty float z\n", (unsigned long long)( points.size() / 3 ) );
	fprintf( fp, "element face %u\nproperty list uchar uint vertex_indices\nend_header\n", nface );
	bool ok = points.empty() || fwrite( points.data(), 4, points.size(), fp ) == points.size();
	ok = ok && ( faces.empty() || fwrite( faces.data(), 1, faces.size(), fp ) == faces.size() );
	fclose( fp );
	if( !ok )
	{
		setError( "hako_voxel_mesh_save_ply: short write to %s", path );
		return HAKO_E_INVALID;
	}
	return HAKO_OK;
}
}
