// tests/update_host.cpp — TEST INFRASTRUCTURE: the product's state arithmetic compiled for the HOST.
//
// 2dgaussiansplatting_b200/csrc/gs2d_update.cuh holds the per-scalar device functions of the state kernels: adam_one
// (Adam::optimize, main.cpp:144-156), clamp_glm and init_splat / pcg3d (main.cpp:17-24, :288-304). This file includes
// that very header (intrinsics mapped by host_shim.h, compile with -ffp-contract=off) and wraps it in the loops of
// k_advance / k_adam / k_init_reference (gs2d_adam.cu), over the reference's AoS types, for
// tests/test_update_host.py, which compares the results bit for bit with the CPU oracle without a GPU.
#include "host_shim.h"

#include "../2dgaussiansplatting_b200/csrc/gs2d_footprint.cuh" // the contraction-free x* helpers
#include "../2dgaussiansplatting_b200/csrc/gs2d_update.cuh"
#include "../include/gs2d.h"

extern "C" {

// k_init_reference: splats = n * 9 floats in the Splat field order
void up_init( int W, int H, int n, float* splats )
{
	for( int i = 0; i < n; i++ )
		init_splat( (unsigned)i, W, H, splats + (size_t)i * 9 );
}

// k_advance (beta powers first, main.cpp:718-719) + k_adam (update, clamps main.cpp:740-749, finite check :752-785).
// splats / grads: n * 9 floats; adam: n * 18 floats ({m, v} per scalar, SplatAdam order). Returns the non-finite flag.
int up_adam( int W, int H, int n, float* splats, float* adam, const float* grads, float* beta1t, float* beta2t,
             const gs2d_params* hp )
{
	*beta1t = xmul( *beta1t, hp->beta1 );
	*beta2t = xmul( *beta2t, hp->beta2 );
	const float om_b1t = xsub( 1.0f, *beta1t ), om_b2t = xsub( 1.0f, *beta2t );
	const float wmax = xsub( (float)W, 1.0f ), hmax = xsub( (float)H, 1.0f );
	int bad = 0;
	for( int i = 0; i < n; i++ )
		for( int pl = 0; pl < 9; pl++ )
		{
			float x = splats[(size_t)i * 9 + pl];
			if( pl != 8 || hp->optimize_opacity )
				x = adam_one( x, grads[(size_t)i * 9 + pl], adam[(size_t)i * 18 + 2 * pl], adam[(size_t)i * 18 + 2 * pl + 1],
				              hp->lr, hp->beta1, hp->beta2, hp->eps, om_b1t, om_b2t );
			if( pl == 0 )
				x = clamp_glm( x, 0.0f, wmax );
			else if( pl == 1 )
				x = clamp_glm( x, 0.0f, hmax );
			else if( pl == 2 || pl == 3 )
				x = clamp_glm( x, hp->s_min, hp->s_max );
			else if( pl == 5 || pl == 6 || pl == 7 )
				x = clamp_glm( x, 0.0f, 1.0f );
			else if( pl == 8 )
				x = clamp_glm( x, hp->o_min, hp->o_max );
			splats[(size_t)i * 9 + pl] = x;
			if( pl != 1 && pl != 8 && !std::isfinite( x ) )
				bad = 1;
		}
	return bad;
}

// The two band tests of the multi-GPU partition (gs2d_footprint.cuh), for the simulation of the sparse parameter
// broadcast in tests/test_update_host.py. n splats; rejects[r * n + i] = band_rejects on rank r's band
// [r * band_px, (r+1) * band_px); masks[i] = reach_mask.
void up_band_tests( int n, const float* py, const float* sx, const float* sy, int band_px, int n_ranks, float slack,
                    uint8_t* rejects, uint32_t* masks )
{
	for( int i = 0; i < n; i++ )
	{
		for( int r = 0; r < n_ranks; r++ )
			rejects[(size_t)r * n + i] = band_rejects( py[i], sx[i], sy[i], r * band_px, ( r + 1 ) * band_px ) ? 1 : 0;
		masks[i] = reach_mask( py[i], sx[i], sy[i], band_px, n_ranks, slack );
	}
}
}
