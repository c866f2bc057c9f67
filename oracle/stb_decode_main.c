/* oracle/stb_decode_main.c — TEST INFRASTRUCTURE. Decodes one image with the reference's own vendored stb_image.h
 * (compiled from where it lies, -I$(REF)/libs/prlib/libs/stb) exactly like prlib's Image2DRGBA8::load does
 * (prp.cpp:185: stbi_load( path, &w, &h, 0, 4 )) and writes int32 w, int32 h, then w*h*4 bytes RGBA to argv[2].
 * Used by tests/golden/make_jpeg_golden.py to pin the library's own JPEG decoder (gs2d_jpeg.cpp). */
#define STB_IMAGE_IMPLEMENTATION
#include "stb_image.h"
#include <stdio.h>

int main( int argc, char** argv )
{
	if( argc < 3 )
		return 2;
	int w = 0, h = 0, n = 0;
	unsigned char* px = stbi_load( argv[1], &w, &h, &n, 4 );
	if( !px )
	{
		fprintf( stderr, "stb: %s\n", stbi_failure_reason() );
		return 1;
	}
	FILE* f = fopen( argv[2], "wb" );
	if( !f )
		return 3;
	fwrite( &w, 4, 1, f );
	fwrite( &h, 4, 1, f );
	fwrite( px, 4, (size_t)w * h, f );
	fclose( f );
	stbi_image_free( px );
	return 0;
}
