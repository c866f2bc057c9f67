// gs2d_update.cuh — the per-scalar arithmetic of the state kernels (gs2d_adam.cu), free of anything CUDA-specific but
// the rounding intrinsics: Adam::optimize (main.cpp:144-156), glm::clamp, pcg3d (main.cpp:17-24) and the reference's
// init() formula (main.cpp:288-304).
//
// Kept in its own header, like gs2d_footprint.cuh, so that tests/test_update_host.py can compile these very functions
// for the HOST and check them bit for bit against the CPU oracle without a GPU. Needs the x* helpers of
// gs2d_footprint.cuh and the plane order PL_* (value = field order of the reference's Splat, main.cpp:85-93).
#pragma once

// glm::clamp = min( max( x, lo ), hi ) with comparison-based min/max (func_common.inl:17-30): NaN survives.
__device__ __forceinline__ float clamp_glm( float x, float lo, float hi )
{
	float t = ( x < lo ) ? lo : x;
	return ( hi < t ) ? hi : t;
}


// Adam::optimize, main.cpp:144-156 — reference operand order, IEEE division and square root, no contraction.
__device__ __forceinline__ float adam_one( float value, float gr, float& m_io, float& v_io, float lr, float beta1,
                                           float beta2, float eps, float om_b1t, float om_b2t )
{
	const float m = xadd( xmul( beta1, m_io ), xmul( xsub( 1.0f, beta1 ), gr ) );
	const float v = xadd( xmul( beta2, v_io ), xmul( xmul( xsub( 1.0f, beta2 ), gr ), gr ) );
	m_io = m;
	v_io = v;
	const float m_hat = xdiv( m, om_b1t );
	const float v_hat = xdiv( v, om_b2t );
	return xsub( value, xdiv( xmul( lr, m_hat ), xadd( xsqrt( v_hat ), eps ) ) );
}


__device__ __forceinline__ void pcg3d( unsigned& x, unsigned& y, unsigned& z ) // main.cpp:17-24
{
	x = x * 1664525u + 1013904223u;
	y = y * 1664525u + 1013904223u;
	z = z * 1664525u + 1013904223u;
	x += y * z;
	y += z * x;
	z += x * y;
	x ^= x >> 16u;
	y ^= y >> 16u;
	z ^= z >> 16u;
	x += y * z;
	y += z * x;
	z += x * y;
}
__device__ __forceinline__ float mix_glm( float x, float y, float a ) // x*(1-a) + y*a, func_common.inl:123-133
{
	return xadd( xmul( x, xsub( 1.0f, a ) ), xmul( y, a ) );
}


// init(), main.cpp:288-304, for splat i of an image: v[plane], plane order == Splat field order.
__device__ __forceinline__ void init_splat( unsigned i, int W, int H, float v[9] )
{
	unsigned ax = i, ay = 0u, az = 0xFFFFFFFFu, bx = i, by = 1u, bz = 0xFFFFFFFFu;
	pcg3d( ax, ay, az );
	pcg3d( bx, by, bz );
	const float denom = 4294967296.0f; // (float)0xFFFFFFFF, main.cpp:290-291
	const float r0x = xdiv( __uint2float_rn( ax ), denom ), r0y = xdiv( __uint2float_rn( ay ), denom );
	const float r1x = xdiv( __uint2float_rn( bx ), denom ), r1y = xdiv( __uint2float_rn( by ), denom ),
	            r1z = xdiv( __uint2float_rn( bz ), denom );
	v[0] = mix_glm( r0x, xsub( (float)W, 1.0f ), r0x );   // pos.x
	v[1] = mix_glm( r0y, xsub( (float)H, 1.0f ), r0y );   // pos.y
	v[2] = mix_glm( 6.0f, 10.0f, r1x );                   // sx
	v[3] = mix_glm( 6.0f, 10.0f, r1y );                   // sy
	v[4] = xmul( 3.14159265358979323846264338327950288f, r1z ); // rot
	v[5] = 0.5f;                                          // color
	v[6] = 0.5f;
	v[7] = 0.5f;
	v[8] = 1.0f;                                          // opacity
}
