// oracle/shim/shim.cpp — TEST INFRASTRUCTURE, not product code.
// Implementation side of the headless pr.hpp stand-in: stb-based JPEG decode exactly as prlib does
// it (prp.cpp:179-193), the u8 -> float conversion (prp.cpp:796-807), the frame counter that ends
// the reference's while-loop, and the image dump behind ITexture::upload (tex0 is the 2nd texture
// main.cpp creates, main.cpp:253,309; its uploads happen once per iteration, main.cpp:794).
#include "pr.hpp"

#define STB_IMAGE_IMPLEMENTATION
#include "stb_image.h"

#include <chrono>

namespace pr
{
static std::string g_dataDir = ".";
static int g_frames = 0;
static int g_textures = 0;

static const char* env_or( const char* k, const char* d )
{
	const char* v = getenv( k );
	return ( v && *v ) ? v : d;
}

void SetDataDir( std::string ) { g_dataDir = env_or( "PR_DATA", "." ); }
std::string ExecutableDir() { return "."; }
void Initialize( Config ) {}
void CleanUp() {}
double GetElapsedTime()
{
	using namespace std::chrono;
	return duration<double>( steady_clock::now().time_since_epoch() ).count();
}
double GetFrameRate() { return 0.0; }

// PR_PRIMS=file: every PrimVertex of the frames listed in PR_PRIMS_FRAMES (comma separated, default "0") is appended as
// { int32 frame; float x, y, z; uint8 r, g, b, 0 } (20 bytes)
void PrimVertex( glm::vec3 p, glm::u8vec3 c )
{
	static const char* path = getenv( "PR_PRIMS" );
	if( !path )
		return;
	static FILE* fp = fopen( path, "wb" );
	static std::string frames = std::string( "," ) + env_or( "PR_PRIMS_FRAMES", "0" ) + ",";
	const int frame = g_frames - 1; // NextFrame() has already counted the frame being drawn
	if( !fp || frames.find( "," + std::to_string( frame ) + "," ) == std::string::npos )
		return;
	struct
	{
		int32_t frame;
		float x, y, z;
		uint8_t r, g, b, pad;
	} rec = { frame, p.x, p.y, p.z, c.x, c.y, c.z, 0 };
	fwrite( &rec, sizeof( rec ), 1, fp );
	fflush( fp );
}

bool NextFrame()
{
	static int iters = atoi( env_or( "PR_ITERS", "20" ) );
	return g_frames++ >= iters;
}

Result Image2DRGBA8::load( const char* filename )
{
	std::string name = env_or( "PR_IMAGE", filename );
	std::string full = g_dataDir + "/" + name;
	stbi_uc* pixels = stbi_load( full.c_str(), &_width, &_height, 0, 4 );
	if( pixels == nullptr )
	{
		fprintf( stderr, "shim: cannot load %s\n", full.c_str() );
		exit( 2 );
	}
	_values.resize( (size_t)_width * _height );
	memcpy( _values.data(), pixels, (size_t)_width * _height * 4 );
	stbi_image_free( pixels );
	if( const char* t = getenv( "PR_TARGET" ) )
	{
		FILE* fp = fopen( t, "wb" );
		int32_t wh[2] = { _width, _height };
		fwrite( wh, sizeof( wh ), 1, fp );
		fwrite( _values.data(), 4, (size_t)_width * _height, fp );
		fclose( fp );
	}
	return Result::Sucess;
}

Image2DRGBA32 Image2DRGBA8_to_Image2DRGBA32( const Image2DRGBA8& src )
{
	Image2DRGBA32 image;
	image.allocate( src.width(), src.height() );
	for( int j = 0; j < image.height(); ++j )
		for( int i = 0; i < image.width(); ++i )
			image( i, j ) = glm::vec4( src( i, j ) ) / glm::vec4( 255.0f );
	return image;
}

ITexture* CreateTexture()
{
	ITexture* t = new ITexture();
	t->_id = g_textures++;
	return t;
}

void ITexture::upload( const Image2DRGBA32& image )
{
	_w = image.width();
	_h = image.height();
	const char* dir = getenv( "PR_DUMP" );
	if( dir && _id == 1 ) // tex0: the rendered image, one upload per iteration
	{
		char path[1024];
		snprintf( path, sizeof( path ), "%s/itr%05d.rgba32f", dir, _uploads );
		FILE* fp = fopen( path, "wb" );
		if( fp )
		{
			fwrite( image.data(), sizeof( glm::vec4 ), (size_t)_w * _h, fp );
			fclose( fp );
		}
	}
	_uploads++;
}
} // namespace pr
