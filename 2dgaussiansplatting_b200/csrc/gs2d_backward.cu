// gs2d_backward.cu — analytic gradients of the blend (the reference's backward loop, main.cpp:552-712).
//
// Same tiling, staging and exact footprint as k_forward. The reference's backward walks the splats in FORWARD order,
// rebuilds the prefix colour/transmittance (image1) and obtains the colour from behind by subtraction,
// S = final - prefix (main.cpp:623-627); alpha is not clamped below 1, so the 3DGS-style reverse sweep (divide T by
// 1-alpha) is unsafe here and the forward-order sweep is what preserves the reference's numbers.
// The prefix chain (P, T) repeats the forward kernel's contraction-free arithmetic, so the set of live pixels and
// the point where a tile terminates are identical to the forward pass.
//
// Gradient accumulation: per (splat, warp) the 8 (9 with opacity) per-pixel terms are reduced over the 32 pixels
// of the warp with a transposed butterfly (9 shuffles for 8 values instead of 40), summed over the tile's warps in
// shared memory, scaled by the hoisted per-splat factors and added to the SoA gradient planes with ONE atomic per
// (tile, splat, parameter).
#include "gs2d_internal.cuh"

#define BWD_NV_MAX 9

struct BwdSmem
{
	SplatRec rec[2][GS2D_BATCH];                 // 4 KB
	unsigned short mask[GS2D_BATCH][GS2D_TILE];  // 1 KB
	unsigned idx[2][GS2D_BATCH];                 // global splat index of every staged entry
	unsigned visited[GS2D_BATCH];                // bit w set: warp w contributed to this entry
	float part[GS2D_BATCH][BWD_NV_MAX][8];       // per-warp partial sums, 9 KB
};

// Sum 8 values over the 32 lanes with 9 shuffles. On return, lane L holds in the result the total of value
// index ((L>>4)&1)*4 + ((L>>3)&1)*2 + ((L>>2)&1)  (all four lanes with the same bits 4..2 hold the same total).
__device__ __forceinline__ float warp_reduce8( float v[8], int lane )
{
	const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
	float r4[4];
#pragma unroll
	for( int i = 0; i < 4; i++ )
	{
		const float keep = b4 ? v[i + 4] : v[i];
		const float send = b4 ? v[i] : v[i + 4];
		r4[i] = keep + __shfl_xor_sync( 0xffffffffu, send, 16 );
	}
	float r2[2];
#pragma unroll
	for( int i = 0; i < 2; i++ )
	{
		const float keep = b3 ? r4[i + 2] : r4[i];
		const float send = b3 ? r4[i] : r4[i + 2];
		r2[i] = keep + __shfl_xor_sync( 0xffffffffu, send, 8 );
	}
	const float keep = b2 ? r2[1] : r2[0];
	const float send = b2 ? r2[0] : r2[1];
	float r = keep + __shfl_xor_sync( 0xffffffffu, send, 4 );
	r += __shfl_xor_sync( 0xffffffffu, r, 2 );
	r += __shfl_xor_sync( 0xffffffffu, r, 1 );
	return r;
}

template <bool OPACITY>
__global__ void __launch_bounds__( 256 ) k_backward( Geometry g, const SplatRec* __restrict__ rec,
                                                     const unsigned* __restrict__ vals,
                                                     const uint2* __restrict__ tile_range,
                                                     const float4* __restrict__ target,
                                                     const float4* __restrict__ image, float* __restrict__ grads,
                                                     float min_throughput, const DevState* __restrict__ st )
{
	constexpr int NV = OPACITY ? 9 : 8;
	if( st->need_capacity )
		return;
	__shared__ BwdSmem sm;

	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const int tile_x = blockIdx.x, tile_y = blockIdx.y + g.band_ty0, img = blockIdx.z;
	const unsigned tile_id = (unsigned)( ( img * g.tiles_y + tile_y ) * g.tiles_x + tile_x );
	const int tile_x0 = tile_x * GS2D_TILE, tile_y0 = tile_y * GS2D_TILE;
	const int x = tile_x0 + ( lane & 15 ), y = tile_y0 + warp * 2 + ( lane >> 4 );
	const bool inside = x < g.W && y < g.H;
	const float fpx = (float)x + 0.5f, fpy = (float)y + 0.5f;

	const uint2 range = tile_range[tile_id];
	const unsigned start = range.x, end = range.y;
	if( start >= end )
		return;

	// per-pixel constants: final colour (image0) and dL/dC = final - reference (main.cpp:613-616)
	float Fr = 0.0f, Fg = 0.0f, Fb = 0.0f, dLr = 0.0f, dLg = 0.0f, dLb = 0.0f;
	if( inside )
	{
		const size_t pix = ( (size_t)img * g.H + y ) * g.W + x;
		const float4 fin = image[pix];
		const float4 ref = target[pix];
		Fr = fin.x;
		Fg = fin.y;
		Fb = fin.z;
		dLr = xsub( fin.x, ref.x );
		dLg = xsub( fin.y, ref.y );
		dLb = xsub( fin.z, ref.z );
	}
	float T = 1.0f, Pr = 0.0f, Pg = 0.0f, Pb = 0.0f; // image1, main.cpp:549
	bool alive = inside;

	// gather role: 4 threads per record (64 bytes), threads 0..127
	const int gk = tid >> 2, gq = tid & 3;
	const bool gatherer = tid < GS2D_BATCH * 4;
	unsigned idx_nxt = 0;
	if( gatherer && start + gk < end )
	{
		unsigned idx0 = vals[start + gk];
		cp_async16( reinterpret_cast<float4*>( &sm.rec[0][gk] ) + gq, reinterpret_cast<const float4*>( rec + idx0 ) + gq );
		if( gq == 0 )
			sm.idx[0][gk] = idx0;
	}
	cp_async_commit();
	if( gatherer && start + GS2D_BATCH + gk < end )
		idx_nxt = vals[start + GS2D_BATCH + gk];

	unsigned livemask = __ballot_sync( 0xffffffffu, alive );
	const unsigned* words = reinterpret_cast<const unsigned*>( &sm.mask[0][0] ) + warp;

	int stage = 0;
	for( unsigned base = start; base < end; base += GS2D_BATCH, stage ^= 1 )
	{
		const int nb = min( (unsigned)GS2D_BATCH, end - base );
		const unsigned nbase = base + GS2D_BATCH;
		if( nbase < end )
		{
			if( gatherer && nbase + gk < end )
			{
				cp_async16( reinterpret_cast<float4*>( &sm.rec[stage ^ 1][gk] ) + gq,
				            reinterpret_cast<const float4*>( rec + idx_nxt ) + gq );
				if( gq == 0 )
					sm.idx[stage ^ 1][gk] = idx_nxt;
			}
			cp_async_commit();
			if( gatherer && nbase + GS2D_BATCH + gk < end )
				idx_nxt = vals[nbase + GS2D_BATCH + gk];
			cp_async_wait<1>();
		}
		else
			cp_async_wait<0>();
		if( tid < GS2D_BATCH )
			sm.visited[tid] = 0u;
		__syncthreads();

		{
			const int r = tid & 15;
#pragma unroll
			for( int h = 0; h < GS2D_BATCH / 16; h++ )
			{
				const int k = ( tid >> 4 ) + h * 16;
				unsigned m = 0;
				if( k < nb )
				{
					const SplatRec& s = sm.rec[stage][k];
					m = row_mask( s.px, s.py, s.a, s.b, s.d, s.begY, s.endY, tile_y0 + r, tile_x0, g.W, g.H );
				}
				sm.mask[k][r] = (unsigned short)m;
			}
		}
		__syncthreads();

		if( livemask )
		{
#pragma unroll 1
			for( int k = 0; k < nb; k++ )
			{
				const unsigned word = words[k * ( GS2D_TILE / 2 )];
				if( ( word & livemask ) == 0u )
					continue;
				float v[8] = { 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f };
				float vop = 0.0f;
				if( ( ( word >> lane ) & 1u ) && alive )
				{
					const SplatRec& s = sm.rec[stage][k];
					const float4 q0 = *reinterpret_cast<const float4*>( &s.px );
					const float4 q1 = *reinterpret_cast<const float4*>( &s.d );
					const float op = s.op;
					const float cT = s.cosT, sT = s.sinT;
					const float vx = xsub( fpx, q0.x ), vy = xsub( fpy, q0.y );
					const float rx = xadd( xmul( q0.z, vx ), xmul( q0.w, vy ) );
					const float ry = xadd( xmul( q0.w, vx ), xmul( q1.x, vy ) );
					const float d2 = xadd( xmul( vx, rx ), xmul( vy, ry ) );
					const float G = gauss_of_d2( d2 );       // main.cpp:610
					const float alpha = xmul( G, op );       // main.cpp:611
					const float wgt = alpha * T;             // dC_dc, main.cpp:618
					v[0] = dLr * wgt;                        // main.cpp:619
					v[1] = dLg * wgt;
					v[2] = dLb * wgt;
					// prefix colour, main.cpp:623-625 (same rounding as the forward pass)
					const float tcr = xmul( T, q1.y ), tcg = xmul( T, q1.z ), tcb = xmul( T, q1.w );
					Pr = xadd( Pr, xmul( tcr, alpha ) );
					Pg = xadd( Pg, xmul( tcg, alpha ) );
					Pb = xadd( Pb, xmul( tcb, alpha ) );
					// S = colour from behind, main.cpp:627;  dC/dalpha = c*T - S/(1 - alpha + 1e-15), main.cpp:628
					const float oma = xsub( 1.0f, alpha );
					const float inv = __frcp_rn( xadd( oma, 1.0e-15f ) );
					const float dar = tcr - xsub( Fr, Pr ) * inv;
					const float dag = tcg - xsub( Fg, Pg ) * inv;
					const float dab = tcb - xsub( Fb, Pb ) * inv;
					const float gsum = dLr * dar + dLg * dag + dLb * dab; // dL_dalpha_rgb, main.cpp:629-630
					const float ga = gsum * alpha;
					v[3] = ga * rx; // 0.5*alpha*(2a vx + (b+c) vy) == alpha*(a vx + b vy), main.cpp:639, :654
					v[4] = ga * ry; // main.cpp:640, :655
					// rotated frame: u = cos*vx + sin*vy, w = cos*vy - sin*vx  =>  the three quadratic forms of
					// main.cpp:657-662 and :680-683 are u*u, w*w and u*w
					const float u = cT * vx + sT * vy;
					const float w = cT * vy - sT * vx;
					const float gu = ga * u;
					v[5] = gu * u;        // * 1/sx^3 at flush time, main.cpp:657-659, :677
					v[6] = ( ga * w ) * w; // * 1/sy^3, main.cpp:660-662, :678
					v[7] = gu * w;        // * (sx^2-sy^2)/(sx^2 sy^2), main.cpp:680-685
					if( OPACITY )
						vop = gsum * G;   // main.cpp:703-704
					T = xmul( T, oma );   // main.cpp:707
					alive = !( T < min_throughput );
				}
				const float r8 = warp_reduce8( v, lane );
				if( ( lane & 3 ) == 0 )
					sm.part[k][lane >> 2][warp] = r8;
				if( OPACITY )
				{
#pragma unroll
					for( int o = 16; o > 0; o >>= 1 )
						vop += __shfl_xor_sync( 0xffffffffu, vop, o );
					if( lane == 0 )
						sm.part[k][8][warp] = vop;
				}
				if( lane == 0 )
					atomicOr( &sm.visited[k], 1u << warp );
				livemask = __ballot_sync( 0xffffffffu, alive );
				if( livemask == 0u )
					break;
			}
		}
		const int any_alive = __syncthreads_or( livemask != 0u );

		// flush: one atomic per (entry, parameter)
		for( int e = tid; e < nb * NV; e += 256 )
		{
			const int k = e / NV, vi = e - k * NV;
			const unsigned vis = sm.visited[k];
			if( vis == 0u )
				continue;
			float sum = 0.0f;
#pragma unroll
			for( int w = 0; w < 8; w++ )
				if( vis & ( 1u << w ) )
					sum += sm.part[k][vi][w];
			const SplatRec& s = sm.rec[stage][k];
			int plane;
			switch( vi )
			{
			case 0: plane = PL_CR; break;
			case 1: plane = PL_CG; break;
			case 2: plane = PL_CB; break;
			case 3: plane = PL_PX; break;
			case 4: plane = PL_PY; break;
			case 5: plane = PL_SX; sum *= s.isx3; break;
			case 6: plane = PL_SY; sum *= s.isy3; break;
			case 7: plane = PL_ROT; sum *= s.rotc; break;
			default: plane = PL_OP; break;
			}
			atomicAdd( &grads[(size_t)plane * g.n_pad + sm.idx[stage][k]], sum );
		}
		if( !any_alive )
			break;
		__syncthreads(); // part/visited/mask are rewritten by the next round
	}
	cp_async_wait<0>();
}

int launch_backward( const Geometry& g, const Buffers& b, int sorted_buf, bool with_opacity, cudaStream_t s )
{
	const int rows = g.band_ty1 - g.band_ty0;
	if( rows <= 0 || g.tiles_x <= 0 )
		return 0;
	dim3 grid( g.tiles_x, rows, g.n_images );
	if( with_opacity )
		k_backward<true><<<grid, 256, 0, s>>>( g, b.rec, b.vals[sorted_buf], b.tile_range, b.target, b.image, b.grads,
		                                       g.min_throughput, b.state );
	else
		k_backward<false><<<grid, 256, 0, s>>>( g, b.rec, b.vals[sorted_buf], b.tile_range, b.target, b.image, b.grads,
		                                        g.min_throughput, b.state );
	return 1;
}
