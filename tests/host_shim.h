// tests/host_shim.h — TEST INFRASTRUCTURE: lets the product's arithmetic headers (csrc/gs2d_footprint.cuh,
// csrc/gs2d_update.cuh) compile for the host. The CUDA rounding intrinsics become plain IEEE float operations, which is
// what they are on the device; the translation unit must be compiled with -ffp-contract=off so that the host compiler
// does not fuse what the _rn intrinsics keep apart.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>

#define __device__
#define __forceinline__ inline
#define GS2D_TILE 16
static inline float __fmul_rn( float a, float b ) { return a * b; }
static inline float __fadd_rn( float a, float b ) { return a + b; }
static inline float __fsub_rn( float a, float b ) { return a - b; }
static inline float __fdiv_rn( float a, float b ) { return a / b; }
static inline float __fsqrt_rn( float a ) { return sqrtf( a ); }
static inline float __fmaf_rn( float a, float b, float c ) { return fmaf( a, b, c ); }
static inline int __float2int_rz( float f ) { return (int)f; } // trunc_x86 has excluded NaN / out of range before
static inline float __int_as_float( int i )
{
	float f;
	memcpy( &f, &i, 4 );
	return f;
}
static inline float __uint2float_rn( unsigned u ) { return (float)u; } // round to nearest even, like cvt.rn.f32.u32
using std::max;
using std::min;

