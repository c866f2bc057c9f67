// gs2d_io.cpp — host code either side of the hot path: checkpoint files (f3), PPM in/out as the
// headless stand-in for the reference's GL texture (f1), overlay geometry as data (f4). Plain C++ on top of the public
// C ABI (include/gs2d.h); nothing here touches the device directly.
//
// Checkpoint file layout (little endian, no padding):
//   char     magic[8]   = "GS2DCKPT"
//   uint32   version    = 1
//   uint32   n_images, n_splats, W, H
//   uint32   optimize_opacity                         (main.cpp:317)
//   per image:
//     float    beta1t, beta2t                         (main.cpp:274-275)
//     int32    iterations                             (main.cpp:278)
//     gs2d_splat      splats[n_splats]     36 B each  (== std::vector<Splat>,     main.cpp:272)
//     gs2d_splat_adam adams[n_splats]      72 B each  (== std::vector<SplatAdam>, main.cpp:276)
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/gs2d.h"

namespace
{
struct File
{
	FILE* f = nullptr;
	File( const char* path, const char* mode ) { f = path ? fopen( path, mode ) : nullptr; }
	~File()
	{
		if( f )
			fclose( f );
	}
	bool put( const void* p, size_t n ) { return fwrite( p, 1, n, f ) == n; }
	bool get( void* p, size_t n ) { return fread( p, 1, n, f ) == n; }
};
const char CKPT_MAGIC[8] = { 'G', 'S', '2', 'D', 'C', 'K', 'P', 'T' };
} // namespace

extern "C" int gs2d_save_checkpoint( gs2d_ctx* ctx, const char* path )
{
	int W, H, n, n_images;
	if( !ctx || !path || gs2d_get_dims( ctx, &W, &H, &n, &n_images ) != GS2D_OK )
		return GS2D_ERR_INVALID;
	// write to a temporary and rename: a crash mid-write never leaves a truncated checkpoint under the final name
	const std::string tmp = std::string( path ) + ".tmp";
	{
		File fp( tmp.c_str(), "wb" );
		if( !fp.f )
			return GS2D_ERR_IO;
		int opt = 0;
		if( gs2d_get_optimize_opacity( ctx, &opt ) != GS2D_OK )
			return GS2D_ERR_INVALID;
		const uint32_t hdr[6] = { 1u, (uint32_t)n_images, (uint32_t)n, (uint32_t)W, (uint32_t)H, (uint32_t)opt };
		if( !fp.put( CKPT_MAGIC, 8 ) || !fp.put( hdr, sizeof( hdr ) ) )
			return GS2D_ERR_IO;
		std::vector<gs2d_splat> s( (size_t)n );
		std::vector<gs2d_splat_adam> a( (size_t)n );
		for( int img = 0; img < n_images; img++ )
		{
			float b1 = 1.0f, b2 = 1.0f;
			int it = 0;
			int rc = gs2d_download_state( ctx, img, s.data(), a.data(), &b1, &b2, &it );
			if( rc != GS2D_OK && rc != GS2D_ERR_NONFINITE ) // a diverged state can still be saved for inspection
				return rc;
			const int32_t it32 = it;
			if( !fp.put( &b1, 4 ) || !fp.put( &b2, 4 ) || !fp.put( &it32, 4 ) ||
			    !fp.put( s.data(), sizeof( gs2d_splat ) * s.size() ) ||
			    !fp.put( a.data(), sizeof( gs2d_splat_adam ) * a.size() ) )
				return GS2D_ERR_IO;
		}
		if( fflush( fp.f ) != 0 )
			return GS2D_ERR_IO;
	}
	if( rename( tmp.c_str(), path ) != 0 )
		return GS2D_ERR_IO;
	return GS2D_OK;
}

extern "C" int gs2d_load_checkpoint( gs2d_ctx* ctx, const char* path )
{
	int W, H, n, n_images;
	if( !ctx || !path || gs2d_get_dims( ctx, &W, &H, &n, &n_images ) != GS2D_OK )
		return GS2D_ERR_INVALID;
	File fp( path, "rb" );
	if( !fp.f )
		return GS2D_ERR_IO;
	char magic[8];
	uint32_t hdr[6];
	if( !fp.get( magic, 8 ) || !fp.get( hdr, sizeof( hdr ) ) )
		return GS2D_ERR_FORMAT;
	if( memcmp( magic, CKPT_MAGIC, 8 ) != 0 || hdr[0] != 1u )
		return GS2D_ERR_FORMAT;
	// the Gaussians live in pixel coordinates of the image they were fitted to: size and counts must match
	if( hdr[1] != (uint32_t)n_images || hdr[2] != (uint32_t)n || hdr[3] != (uint32_t)W || hdr[4] != (uint32_t)H )
		return GS2D_ERR_FORMAT;
	// read everything first, so a truncated file leaves the context untouched
	std::vector<gs2d_splat> s( (size_t)n * n_images );
	std::vector<gs2d_splat_adam> a( (size_t)n * n_images );
	std::vector<float> b1( n_images ), b2( n_images );
	std::vector<int32_t> it( n_images );
	for( int img = 0; img < n_images; img++ )
	{
		if( !fp.get( &b1[img], 4 ) || !fp.get( &b2[img], 4 ) || !fp.get( &it[img], 4 ) ||
		    !fp.get( s.data() + (size_t)img * n, sizeof( gs2d_splat ) * n ) ||
		    !fp.get( a.data() + (size_t)img * n, sizeof( gs2d_splat_adam ) * n ) )
			return GS2D_ERR_FORMAT;
	}
	for( int img = 0; img < n_images; img++ )
	{
		int rc = gs2d_upload_state( ctx, img, s.data() + (size_t)img * n, a.data() + (size_t)img * n, b1[img], b2[img],
		                            it[img] );
		if( rc != GS2D_OK )
			return rc;
	}
	return gs2d_set_optimize_opacity( ctx, hdr[5] ? 1 : 0 );
}

// ------------------------------------------------------------------------------------------------------- PPM ---
extern "C" int gs2d_image_write_ppm( const char* path, const float* rgba, int W, int H )
{
	if( !path || !rgba || W <= 0 || H <= 0 )
		return GS2D_ERR_INVALID;
	File fp( path, "wb" );
	if( !fp.f )
		return GS2D_ERR_IO;
	fprintf( fp.f, "P6\n%d %d\n255\n", W, H );
	std::vector<uint8_t> row( (size_t)W * 3 );
	for( int y = 0; y < H; y++ )
	{
		for( int x = 0; x < W; x++ )
			for( int ch = 0; ch < 3; ch++ )
			{
				float v = rgba[( (size_t)y * W + x ) * 4 + ch];
				v = v < 0.0f ? 0.0f : ( v > 1.0f ? 1.0f : v ); // NaN -> 1 by the second comparison's else branch: fine
				row[(size_t)x * 3 + ch] = (uint8_t)lrintf( v * 255.0f );
			}
		if( !fp.put( row.data(), row.size() ) )
			return GS2D_ERR_IO;
	}
	return GS2D_OK;
}

static bool ppm_token( const uint8_t* d, size_t n, size_t& i, int& out )
{
	for( ;; )
	{
		while( i < n && ( d[i] == ' ' || d[i] == '\t' || d[i] == '\n' || d[i] == '\r' ) )
			i++;
		if( i < n && d[i] == '#' )
		{
			while( i < n && d[i] != '\n' )
				i++;
			continue;
		}
		break;
	}
	if( i >= n || d[i] < '0' || d[i] > '9' )
		return false;
	long v = 0;
	while( i < n && d[i] >= '0' && d[i] <= '9' && v < 100000000 )
		v = v * 10 + ( d[i++] - '0' );
	out = (int)v;
	return true;
}

static int decode_ppm( const uint8_t* d, size_t n, uint8_t** rgba, int* W, int* H )
{
	size_t i = 2;
	int w, h, maxv;
	if( !ppm_token( d, n, i, w ) || !ppm_token( d, n, i, h ) || !ppm_token( d, n, i, maxv ) )
		return GS2D_ERR_FORMAT;
	if( w <= 0 || h <= 0 || w > 65535 || h > 65535 || maxv != 255 || i >= n )
		return GS2D_ERR_FORMAT;
	i++; // the single whitespace byte after maxval
	if( n - i < (size_t)w * h * 3 )
		return GS2D_ERR_FORMAT;
	uint8_t* out = (uint8_t*)malloc( (size_t)w * h * 4 );
	if( !out )
		return GS2D_ERR_OOM;
	for( size_t p = 0; p < (size_t)w * h; p++ )
	{
		out[p * 4 + 0] = d[i + p * 3 + 0];
		out[p * 4 + 1] = d[i + p * 3 + 1];
		out[p * 4 + 2] = d[i + p * 3 + 2];
		out[p * 4 + 3] = 255;
	}
	*rgba = out;
	*W = w;
	*H = h;
	return GS2D_OK;
}

extern "C" int gs2d_image_load_rgba8( const char* path, uint8_t** rgba, int* W, int* H )
{
	if( !path || !rgba || !W || !H )
		return GS2D_ERR_INVALID;
	File fp( path, "rb" );
	if( !fp.f )
		return GS2D_ERR_IO;
	std::vector<uint8_t> bytes;
	uint8_t buf[65536];
	size_t got;
	while( ( got = fread( buf, 1, sizeof( buf ), fp.f ) ) > 0 )
		bytes.insert( bytes.end(), buf, buf + got );
	if( bytes.size() < 4 )
		return GS2D_ERR_FORMAT;
	if( bytes[0] == 'P' && bytes[1] == '6' )
		return decode_ppm( bytes.data(), bytes.size(), rgba, W, H );
	return gs2d_image_decode_jpeg_rgba8( bytes.data(), bytes.size(), rgba, W, H );
}

extern "C" void gs2d_image_free( uint8_t* rgba ) { free( rgba ); }

// --------------------------------------------------------------------------------------------------- overlay ---
// Same float operations, in the same order, as main.cpp:187-234 and :421-465 (sinf/cosf from libm as there).
extern "C" int gs2d_overlay_geometry( const gs2d_splat* splats, int n, int nvtx, gs2d_overlay* out, float* ellipse )
{
	if( !splats || !out || n < 0 || nvtx < 1 )
		return GS2D_ERR_INVALID;
	const float step = 3.14159265358979323846264338327950288f * 2.0f / (float)nvtx; // glm::pi<float>() * 2.0f / nvtx
	const float sinD = sinf( step ), cosD = cosf( step );
	for( int i = 0; i < n; i++ )
	{
		const gs2d_splat& s = splats[i];
		// cov_of, main.cpp:206-221
		const float c = cosf( s.rot ), sn = sinf( s.rot );
		const float l0 = s.sx * s.sx, l1 = s.sy * s.sy;
		const float s11 = l0 * c * c + l1 * sn * sn;
		const float s12 = ( l0 - l1 ) * sn * c;
		const float s22 = l0 + l1 - s11;
		// eignValues, main.cpp:187-196
		const float mean = ( s11 + s22 ) * 0.5f;
		const float det = s11 * s22 - s12 * s12;
		float dd = mean * mean - det;
		dd = sqrtf( dd < 0.0f ? 0.0f : dd );
		const float lambda0 = mean + dd, lambda1 = mean - dd;
		// eigen_vectors_of_cov, main.cpp:223-234
		const float eps = 1e-15f;
		float ex = s11 < s22 ? s12 + eps : lambda0 - s22;
		float ey = s11 < s22 ? lambda0 - s11 : s12 + eps;
		const float inv_len = 1.0f / sqrtf( ex * ex + ey * ey ); // glm::normalize = v * inversesqrt( dot( v, v ) )
		ex *= inv_len;
		ey *= inv_len;
		const float r0 = sqrtf( lambda0 ), r1 = sqrtf( lambda1 );
		gs2d_overlay& o = out[i];
		o.center[0] = s.pos[0];
		o.center[1] = s.pos[1];
		o.axis0[0] = ex * r0;
		o.axis0[1] = ey * r0;
		o.axis1[0] = -ey * r1;
		o.axis1[1] = ex * r1;
		// inv_cov = adj / det (main.cpp:432-436); the box half-sizes of :464-465 multiply det back in
		const float inv00 = s22 / det, inv11 = s11 / det;
		o.half_extent[0] = sqrtf( inv11 * det );
		o.half_extent[1] = sqrtf( inv00 * det );
		if( ellipse )
		{
			float sinT = 0.0f, cosT = 1.0f; // CircleGenerator( step ), prp.hpp:276-313
			float* e = ellipse + (size_t)i * ( nvtx + 1 ) * 2;
			for( int k = 0; k <= nvtx; k++ )
			{
				e[2 * k + 0] = s.pos[0] + o.axis0[0] * sinT + o.axis1[0] * cosT;
				e[2 * k + 1] = s.pos[1] + o.axis0[1] * sinT + o.axis1[1] * cosT;
				const float ns = sinT * cosD + cosT * sinD;
				const float nc = cosT * cosD - sinT * sinD;
				sinT = ns;
				cosT = nc;
			}
		}
	}
	return GS2D_OK;
}
