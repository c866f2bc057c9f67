// gs2d_microbench.cu — the two roofline denominators the blend/Adam kernels are reported against, measured on the
// device at hand: FP32 FMA issue rate of the CUDA cores and HBM copy bandwidth.
#include "gs2d_internal.cuh"

// 8 independent FMA chains per thread, 4096 FMAs each: pure FFMA issue, no memory.
__global__ void __launch_bounds__( 256 ) k_fma_peak( float* out, float a, float b, int iters )
{
	float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
	for( int i = 0; i < iters; i++ )
	{
#pragma unroll
		for( int k = 0; k < 16; k++ )
		{
			x0 = fmaf( x0, a, b );
			x1 = fmaf( x1, a, b );
			x2 = fmaf( x2, a, b );
			x3 = fmaf( x3, a, b );
			x4 = fmaf( x4, a, b );
			x5 = fmaf( x5, a, b );
			x6 = fmaf( x6, a, b );
			x7 = fmaf( x7, a, b );
		}
	}
	float s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
	if( s == 12345.678f )
		out[0] = s;
}

__global__ void __launch_bounds__( 256 ) k_copy4( const float4* __restrict__ src, float4* __restrict__ dst, size_t n )
{
	size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
	const size_t stride = (size_t)gridDim.x * 256;
	for( ; i < n; i += stride )
		dst[i] = src[i];
}

extern "C" int gs2d_microbench_fp32( int device, double* tflops )
{
	if( !tflops || cudaSetDevice( device ) != cudaSuccess )
		return GS2D_ERR_NO_DEVICE;
	cudaDeviceProp prop;
	cudaGetDeviceProperties( &prop, device );
	float* out = nullptr;
	if( cudaMalloc( &out, 4 ) != cudaSuccess )
		return GS2D_ERR_OOM;
	cudaEvent_t e0, e1;
	cudaEventCreate( &e0 );
	cudaEventCreate( &e1 );
	const int blocks = prop.multiProcessorCount * 8, iters = 256;
	double best = 0.0;
	for( int rep = 0; rep < 5; rep++ )
	{
		cudaEventRecord( e0 );
		k_fma_peak<<<blocks, 256>>>( out, 0.999f, 0.001f, iters );
		cudaEventRecord( e1 );
		cudaEventSynchronize( e1 );
		float ms = 0;
		cudaEventElapsedTime( &ms, e0, e1 );
		double flops = 2.0 * 8 * 16 * (double)iters * 256.0 * blocks;
		best = fmax( best, flops / ( ms * 1e-3 ) / 1e12 );
	}
	cudaEventDestroy( e0 );
	cudaEventDestroy( e1 );
	cudaFree( out );
	*tflops = best;
	return cudaGetLastError() == cudaSuccess ? GS2D_OK : GS2D_ERR_CUDA;
}

extern "C" int gs2d_microbench_hbm( int device, double* gbs )
{
	if( !gbs || cudaSetDevice( device ) != cudaSuccess )
		return GS2D_ERR_NO_DEVICE;
	cudaDeviceProp prop;
	cudaGetDeviceProperties( &prop, device );
	const size_t n = ( 1ull << 30 ) / sizeof( float4 ); // 1 GiB each way: far larger than the 126 MB L2
	float4 *a = nullptr, *b = nullptr;
	if( cudaMalloc( &a, n * sizeof( float4 ) ) != cudaSuccess || cudaMalloc( &b, n * sizeof( float4 ) ) != cudaSuccess )
	{
		cudaFree( a );
		return GS2D_ERR_OOM;
	}
	cudaMemset( a, 1, n * sizeof( float4 ) );
	cudaEvent_t e0, e1;
	cudaEventCreate( &e0 );
	cudaEventCreate( &e1 );
	double best = 0.0;
	for( int rep = 0; rep < 5; rep++ )
	{
		cudaEventRecord( e0 );
		k_copy4<<<prop.multiProcessorCount * 16, 256>>>( a, b, n );
		cudaEventRecord( e1 );
		cudaEventSynchronize( e1 );
		float ms = 0;
		cudaEventElapsedTime( &ms, e0, e1 );
		best = fmax( best, 2.0 * n * sizeof( float4 ) / ( ms * 1e-3 ) / 1e9 );
	}
	cudaEventDestroy( e0 );
	cudaEventDestroy( e1 );
	cudaFree( a );
	cudaFree( b );
	*gbs = best;
	return cudaGetLastError() == cudaSuccess ? GS2D_OK : GS2D_ERR_CUDA;
}
