// gs2d_static.cu — gs2d_mse_static: the loss of main.cpp:796-805 over caller-owned DEVICE buffers, for callers that
// may not allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc). It has no counterpart in the
// reference. Nothing here touches a gs2d_ctx or any other entry
// point's path.
//
// The per-tile partial sums that gs2d_forward keeps in a cudaMalloc'd buffer (Buffers::loss_partial) live here in a
// file-scope __device__ array of GS2D_MSE_STATIC_MAX_TILES doubles (512 KiB), which every GPU that loads the library
// holds until the process exits. Both kernels repeat the summation ORDER of the product path (the forward epilogue,
// gs2d_forward.cu, and k_finalize_loss, gs2d_adam.cu), so the result equals gs2d_last_mse bit for bit.
//
// The array is shared by every call on a device and keeps the last call's contents. Nothing has to start at zero:
// k_tile_sq_error overwrites partial[0, n_tiles) and k_sum_partials reads exactly that range. Two calls that overlap
// on the GPU (different streams, or one stream per host thread) race on it: serialise them.
#include "gs2d_internal.cuh"

#include <mutex>
#include <stdarg.h>
#include <stdio.h>

#define STATIC_MAX_DEVICES 64
#define SQ_WARPS 8
#define SUM_THREADS 1024

static __device__ __align__( 256 ) double g_static_partial[GS2D_MSE_STATIC_MAX_TILES];

// One warp per 16x16 tile; lane -> pixels and the order of the sums as in the forward epilogue (gs2d_forward.cu):
// lane pixel (lane & 7, lane >> 3), slot j at (+8 * (j & 1), +4 * (j >> 1)), then a butterfly over the warp.
__global__ void __launch_bounds__( SQ_WARPS * 32 ) k_tile_sq_error( const float4* __restrict__ image,
                                                                 const float4* __restrict__ target, int W, int H,
                                                                 int tiles_x, int n_tiles, double* __restrict__ partial )
{
	const int lane = threadIdx.x & 31;
	const int tile_id = blockIdx.x * SQ_WARPS + ( threadIdx.x >> 5 );
	if( tile_id >= n_tiles )
		return;
	const int tile_x = tile_id % tiles_x, tile_y = tile_id / tiles_x;
	const int px_x = tile_x * GS2D_TILE + ( lane & 7 ), px_y = tile_y * GS2D_TILE + ( lane >> 3 );
	double e = 0.0;
#pragma unroll
	for( int j = 0; j < 8; j++ )
	{
		const int x = px_x + 8 * ( j & 1 ), y = px_y + 4 * ( j >> 1 );
		if( x < W && y < H )
		{
			const size_t pix = (size_t)y * W + x;
			const float4 c = image[pix], ref = target[pix];
			const float dr = xmul( xsub( c.x, ref.x ), 255.0f ), dg = xmul( xsub( c.y, ref.y ), 255.0f ),
			            db = xmul( xsub( c.z, ref.z ), 255.0f );
			e += (double)xadd( xadd( xmul( dr, dr ), xmul( dg, dg ) ), xmul( db, db ) );
		}
	}
#pragma unroll
	for( int o = 16; o > 0; o >>= 1 )
		e += __shfl_xor_sync( 0xffffffffu, e, o );
	if( lane == 0 )
		partial[tile_id] = e;
}

// One block: the fixed-order sum of k_finalize_loss (gs2d_adam.cu) and the division of main.cpp:805.
__global__ void __launch_bounds__( SUM_THREADS ) k_sum_partials( const double* __restrict__ partial, int n_tiles, int W,
                                                                int H, double* __restrict__ mse )
{
	__shared__ double red[SUM_THREADS];
	double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
	int t = threadIdx.x;
	for( ; t + 3 * SUM_THREADS < n_tiles; t += 4 * SUM_THREADS )
	{
		s0 += partial[t];
		s1 += partial[t + SUM_THREADS];
		s2 += partial[t + 2 * SUM_THREADS];
		s3 += partial[t + 3 * SUM_THREADS];
	}
	for( ; t < n_tiles; t += SUM_THREADS )
		s0 += partial[t];
	red[threadIdx.x] = ( s0 + s1 ) + ( s2 + s3 );
	__syncthreads();
	for( int o = SUM_THREADS / 2; o > 0; o >>= 1 )
	{
		if( threadIdx.x < o )
			red[threadIdx.x] += red[threadIdx.x + o];
		__syncthreads();
	}
	if( threadIdx.x == 0 )
		mse[0] = red[0] / (double)( H * W * 3 );
}

// In host code g_static_partial is a stand-in for the symbol, not a device address: the address the kernels are given
// comes from cudaGetSymbolAddress, asked once per device with that device current.
static std::mutex g_static_lock;
static double* g_static_ptr[STATIC_MAX_DEVICES];

static int say( char* err, size_t cap, int code, const char* fmt, ... )
{
	if( err && cap )
	{
		va_list ap;
		va_start( ap, fmt );
		vsnprintf( err, cap, fmt, ap );
		va_end( ap );
	}
	return code;
}

extern "C" int gs2d_mse_static( int device, const float* image_dev, const float* target_dev, int W, int H,
                                double* mse_dev, void* cuda_stream, char* err, size_t err_cap )
{
	if( err && err_cap )
		err[0] = 0;
	if( !image_dev || !target_dev || !mse_dev || W <= 0 || H <= 0 )
		return say( err, err_cap, GS2D_ERR_INVALID, "gs2d_mse_static: null pointer or empty image (W %d, H %d)", W, H );
	if( ( ( (uintptr_t)image_dev | (uintptr_t)target_dev ) & 15 ) || ( (uintptr_t)mse_dev & 7 ) )
		return say( err, err_cap, GS2D_ERR_INVALID, "gs2d_mse_static: image / target must be 16-byte aligned, mse 8-byte" );
	if( device < 0 || device >= STATIC_MAX_DEVICES )
		return say( err, err_cap, GS2D_ERR_INVALID, "gs2d_mse_static: device index %d (limit %d)", device,
		            STATIC_MAX_DEVICES - 1 );
	const long long tiles_x = ( W + GS2D_TILE - 1 ) / GS2D_TILE, tiles_y = ( H + GS2D_TILE - 1 ) / GS2D_TILE;
	const long long n_tiles = tiles_x * tiles_y;
	if( n_tiles > GS2D_MSE_STATIC_MAX_TILES )
		return say( err, err_cap, GS2D_ERR_INVALID,
		            "gs2d_mse_static: image too large: %d x %d is %lld 16x16 tiles, the static scratch holds %d", W, H,
		            n_tiles, GS2D_MSE_STATIC_MAX_TILES );
	cudaError_t e = cudaSetDevice( device );
	if( e != cudaSuccess )
		return say( err, err_cap, GS2D_ERR_NO_DEVICE, "gs2d_mse_static: cudaSetDevice: %s (device %d)",
		            cudaGetErrorString( e ), device );
	double* partial;
	{
		std::lock_guard<std::mutex> hold( g_static_lock );
		if( !g_static_ptr[device] )
		{
			void* p = nullptr;
			e = cudaGetSymbolAddress( &p, g_static_partial );
			if( e != cudaSuccess || !p )
				return say( err, err_cap, GS2D_ERR_CUDA, "gs2d_mse_static: cudaGetSymbolAddress: %s (device %d)",
				            cudaGetErrorString( e ), device );
			g_static_ptr[device] = (double*)p;
		}
		partial = g_static_ptr[device];
	}
	cudaStream_t s = (cudaStream_t)cuda_stream;
	const int blocks = (int)( ( n_tiles + SQ_WARPS - 1 ) / SQ_WARPS );
	k_tile_sq_error<<<blocks, SQ_WARPS * 32, 0, s>>>( reinterpret_cast<const float4*>( image_dev ),
	                                                  reinterpret_cast<const float4*>( target_dev ), W, H, (int)tiles_x,
	                                                  (int)n_tiles, partial );
	k_sum_partials<<<1, SUM_THREADS, 0, s>>>( partial, (int)n_tiles, W, H, mse_dev );
	e = cudaGetLastError();
	if( e != cudaSuccess )
		return say( err, err_cap, GS2D_ERR_CUDA, "gs2d_mse_static: launch: %s (device %d)",
		            cudaGetErrorString( e ), device );
	return GS2D_OK;
}
