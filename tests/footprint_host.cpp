// tests/footprint_host.cpp — TEST INFRASTRUCTURE: the product's footprint arithmetic compiled for the HOST.
//
// 2dgaussiansplatting_b200/csrc/gs2d_footprint.cuh holds the device functions behind every coverage decision of the
// CUDA path: splat_geometry (cov_of / inverse / row range, main.cpp:206-221, :432-436, :489-491), row_mask (the exact
// scan-line footprint of one tile row, main.cpp:171-185, :492-514) and tile_span (the conservative tile columns the
// binning emits). This file includes that very header with the CUDA rounding intrinsics mapped to plain IEEE float
// operations (compile with -ffp-contract=off, like the reference build) and exports two C functions for
// tests/test_footprint_host.py, which checks them against the CPU oracle without a GPU.
#include "host_shim.h"

#include "../2dgaussiansplatting_b200/csrc/gs2d_footprint.cuh"

// cosf / sinf of libm, like the oracle and the reference (the device rounds a double sincos once; the two agree
// except in the last bit of ~1 % of the arguments, which is not what these tests are about)
static SplatGeom geometry( float py, float sx, float sy, float rot ) { return splat_geometry( py, sx, sy, cosf( rot ), sinf( rot ) ); }

extern "C" {

// Coverage of ONE Gaussian over a W x H image, decided tile row by tile row with row_mask exactly as k_forward's
// solve_batch does. out: H * W bytes (1 = covered). Returns the number of covered pixels.
long long fp_coverage( int W, int H, float px, float py, float sx, float sy, float rot, uint8_t* out )
{
	const SplatGeom ge = geometry( py, sx, sy, rot );
	memset( out, 0, (size_t)W * H );
	long long n = 0;
	const int tiles_x = ( W + GS2D_TILE - 1 ) / GS2D_TILE, tiles_y = ( H + GS2D_TILE - 1 ) / GS2D_TILE;
	for( int ty = 0; ty < tiles_y; ty++ )
		for( int tx = 0; tx < tiles_x; tx++ )
			for( int r = 0; r < GS2D_TILE; r++ )
			{
				const int y = ty * GS2D_TILE + r;
				const unsigned m = row_mask( px, py, ge.a, ge.b, ge.d, ge.begY, ge.endY, y, tx * GS2D_TILE, W, H );
				for( int x = 0; x < GS2D_TILE; x++ )
					if( ( m >> x ) & 1u )
					{
						out[(size_t)y * W + tx * GS2D_TILE + x] = 1; // a bit outside the image would write out of bounds: caught by the caller's guard row
						n++;
					}
			}
	return n;
}

// The tiles the binning would emit for this Gaussian: the conservative rectangle (mirrors k_preprocess,
// gs2d_binning.cu) narrowed per tile row by tile_span. tiles: tiles_y * tiles_x bytes (1 = emitted).
// counts[0] = emitted pairs, counts[1] = pairs of the rectangle alone. Returns 0.
int fp_binned_tiles( int W, int H, float px, float py, float sx, float sy, float rot, int cull, uint8_t* tiles,
                     long long counts[2] )
{
	const SplatGeom ge = geometry( py, sx, sy, rot );
	const int tiles_x = ( W + GS2D_TILE - 1 ) / GS2D_TILE, tiles_y = ( H + GS2D_TILE - 1 ) / GS2D_TILE;
	memset( tiles, 0, (size_t)tiles_x * tiles_y );
	counts[0] = counts[1] = 0;
	// rectangle: k_preprocess
	int rtx0 = 1, rtx1 = 0, ty0 = 1, ty1 = 0;
	const int y0 = max( ge.begY, 0 ), y1 = min( ge.endY, H - 1 );
	const float hx = half_width_x( ge );
	const float m = 2.0f + hx * 1.0e-4f;
	const float fx0 = floorf( px - hx - m ), fx1 = floorf( px + hx + m );
	const int x0 = ( fx0 == fx0 ) ? (int)fmaxf( fx0, 0.0f ) : 0;
	int x1 = ( fx1 == fx1 ) ? (int)fminf( fx1, (float)( W - 1 ) ) : W - 1;
	if( fx1 < 0.0f )
		x1 = -1;
	if( y0 <= y1 && x0 <= x1 )
	{
		ty0 = max( y0 / GS2D_TILE, 0 );
		ty1 = min( y1 / GS2D_TILE, tiles_y - 1 );
		rtx0 = x0 / GS2D_TILE;
		rtx1 = x1 / GS2D_TILE;
	}
	if( !( ty0 <= ty1 && rtx0 <= rtx1 ) )
		return 0;
	SpanConsts k;
	k.px = px;
	k.py = py;
	k.a = ge.a;
	k.b = ge.b;
	k.d = ge.d;
	k.inv_a = 1.0f / ge.a;
	k.vx_max = hx;
	k.vy_at = -ge.b * hx / ge.d;
	k.delta = span_delta( px, hx, cull );
	k.begY = ge.begY;
	k.endY = ge.endY;
	for( int ty = ty0; ty <= ty1; ty++ )
	{
		counts[1] += rtx1 - rtx0 + 1;
		int lo, hi;
		if( !tile_span( k, ty, H, rtx0, rtx1, lo, hi ) )
			continue;
		for( int tx = lo; tx <= hi; tx++ )
			tiles[(size_t)ty * tiles_x + tx] = 1;
		counts[0] += hi - lo + 1;
	}
	return 0;
}
}
