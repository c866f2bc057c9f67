// gs2d_backward.cu — analytic gradients of the blend (the reference's backward loop, main.cpp:552-712).
//
// Same tiling, staging, pixel-to-thread map and exact footprint as k_forward (gs2d_blend.cuh). The reference's
// backward walks the splats in FORWARD order, rebuilds the prefix colour/transmittance (image1) and obtains the
// colour from behind by subtraction, S = final - prefix (main.cpp:623-627); alpha is not clamped below 1, so the
// 3DGS-style reverse sweep (divide T by 1-alpha) is unsafe here and the forward-order sweep is what preserves the
// reference's numbers. The prefix chain (P, T) repeats the forward kernel's contraction-free
// arithmetic, so the set of live pixels and the point where a tile terminates are identical to the forward pass.
//
// Gradient accumulation: every thread first sums the 8 (9 with opacity) per-pixel terms over its own P pixels in
// registers, then the warp reduces them over its 32 lanes with a transposed butterfly (9 shuffles for 8 values
// instead of 40) ONCE per (splat, warp), warps of a tile are summed in shared memory, the hoisted per-splat factors
// are applied and ONE atomic per (tile, splat, parameter) goes to the SoA gradient planes.
#include "gs2d_blend.cuh"

#define BWD_NV_MAX 9
#ifndef BWD_CONST_SMEM
#define BWD_CONST_SMEM 1   // keep the per-pixel constants (final colour, dL/dC) in shared memory instead of registers
#endif
#ifndef BWD_MINB
#define BWD_MINB 16   // P = 8: cap at 128 registers -> 16 resident warps per SM (measured best)
#endif

// Sum 8 values over the 32 lanes with 9 shuffles. On return, lane L holds the total of value index
// ((L>>4)&1)*4 + ((L>>3)&1)*2 + ((L>>2)&1)  (all four lanes with the same bits 4..2 hold the same total).
__device__ __forceinline__ float warp_reduce8( const float v[8], int lane )
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

template <int NW>
struct BwdExtra
{
	unsigned idx[2][GS2D_BATCH];               // global splat index of every staged entry
	unsigned visited[GS2D_BATCH];              // bit w set: warp w contributed to this entry
	float part[GS2D_BATCH][BWD_NV_MAX][NW];    // per-warp partial sums
};

template <int P, bool OPACITY>
__global__ void __launch_bounds__( BlendCfg<P>::NT, P == 8 ? BWD_MINB : 1 ) k_backward( Geometry g, const SplatRec* __restrict__ rec,
                                                                 const unsigned* __restrict__ vals,
                                                                 const uint2* __restrict__ tile_range,
                                                                 const float4* __restrict__ target,
                                                                 const float4* __restrict__ image,
                                                                 float* __restrict__ grads,
                                                                 const DevState* __restrict__ st )
{
	constexpr int NW = BlendCfg<P>::NW, NT = BlendCfg<P>::NT;
	constexpr int NV = OPACITY ? 9 : 8;
	if( st->need_capacity )
		return;
	__shared__ StageSmem sm;
	__shared__ BwdExtra<NW> sx;

	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const int tile_x = blockIdx.x, tile_y = blockIdx.y + g.band_ty0, img = blockIdx.z;
	const unsigned tile_id = (unsigned)( ( img * g.tiles_y + tile_y ) * g.tiles_x + tile_x );
	const int tile_x0 = tile_x * GS2D_TILE, tile_y0 = tile_y * GS2D_TILE;
	const int x = tile_x0 + ( lane & 15 ), y0 = tile_y0 + warp * 2 * P + ( lane >> 4 );
	const float fpx = (float)x + 0.5f;
	const float min_T = g.min_throughput;

	const uint2 range = tile_range[tile_id];
	const unsigned start = range.x, end = range.y;
	if( start >= end )
		return;

	// per-pixel constants: final colour (image0) and dL/dC = final - reference (main.cpp:613-616);
	// running prefix (image1, main.cpp:549): colour 0, throughput 1 (0 = dead for pixels outside the image)
#if BWD_CONST_SMEM
	__shared__ float cst[6][P][NT]; // [component][slot][thread]: conflict-free, 6 KB per tile
#define FR( j ) cst[0][j][tid]
#define FG( j ) cst[1][j][tid]
#define FB( j ) cst[2][j][tid]
#define DLR( j ) cst[3][j][tid]
#define DLG( j ) cst[4][j][tid]
#define DLB( j ) cst[5][j][tid]
#else
	float Fr[P], Fg[P], Fb[P], dLr[P], dLg[P], dLb[P];
#define FR( j ) Fr[j]
#define FG( j ) Fg[j]
#define FB( j ) Fb[j]
#define DLR( j ) dLr[j]
#define DLG( j ) dLg[j]
#define DLB( j ) dLb[j]
#endif
	float T[P], Pr[P], Pg[P], Pb[P];
	unsigned livemask[P];
#pragma unroll
	for( int j = 0; j < P; j++ )
	{
		const int y = y0 + 2 * j;
		FR( j ) = FG( j ) = FB( j ) = DLR( j ) = DLG( j ) = DLB( j ) = 0.0f;
		T[j] = 0.0f;
		if( x < g.W && y < g.H )
		{
			const size_t pix = ( (size_t)img * g.H + y ) * g.W + x;
			const float4 fin = image[pix];
			const float4 ref = target[pix];
			FR( j ) = fin.x;
			FG( j ) = fin.y;
			FB( j ) = fin.z;
			DLR( j ) = xsub( fin.x, ref.x );
			DLG( j ) = xsub( fin.y, ref.y );
			DLB( j ) = xsub( fin.z, ref.z );
			T[j] = 1.0f;
		}
		Pr[j] = Pg[j] = Pb[j] = 0.0f;
		livemask[j] = __ballot_sync( 0xffffffffu, !( T[j] < min_T ) );
	}

	// gather role: 4 quads (64 bytes) per record
	constexpr int GI = ( GS2D_BATCH * 4 + NT - 1 ) / NT;
	unsigned idx_nxt[GI];
#pragma unroll
	for( int i = 0; i < GI; i++ )
	{
		const int t = tid + i * NT, gk = t >> 2, gq = t & 3;
		idx_nxt[i] = 0;
		if( t < GS2D_BATCH * 4 && start + gk < end )
		{
			const unsigned idx0 = vals[start + gk];
			cp_async16( reinterpret_cast<float4*>( &sm.rec[0][gk] ) + gq,
			            reinterpret_cast<const float4*>( rec + idx0 ) + gq );
			if( gq == 0 )
				sx.idx[0][gk] = idx0;
		}
	}
	cp_async_commit();
#pragma unroll
	for( int i = 0; i < GI; i++ )
	{
		const int t = tid + i * NT, gk = t >> 2;
		if( t < GS2D_BATCH * 4 && start + GS2D_BATCH + gk < end )
			idx_nxt[i] = vals[start + GS2D_BATCH + gk];
	}

	int stage = 0;
	for( unsigned base = start; base < end; base += GS2D_BATCH, stage ^= 1 )
	{
		const int nb = min( (unsigned)GS2D_BATCH, end - base );
		const unsigned nbase = base + GS2D_BATCH;
		if( nbase < end )
		{
#pragma unroll
			for( int i = 0; i < GI; i++ )
			{
				const int t = tid + i * NT, gk = t >> 2, gq = t & 3;
				if( t < GS2D_BATCH * 4 && nbase + gk < end )
				{
					cp_async16( reinterpret_cast<float4*>( &sm.rec[stage ^ 1][gk] ) + gq,
					            reinterpret_cast<const float4*>( rec + idx_nxt[i] ) + gq );
					if( gq == 0 )
						sx.idx[stage ^ 1][gk] = idx_nxt[i];
				}
			}
			cp_async_commit();
#pragma unroll
			for( int i = 0; i < GI; i++ )
			{
				const int t = tid + i * NT, gk = t >> 2;
				if( t < GS2D_BATCH * 4 && nbase + GS2D_BATCH + gk < end )
					idx_nxt[i] = vals[nbase + GS2D_BATCH + gk];
			}
			cp_async_wait<1>();
		}
		else
			cp_async_wait<0>();
		if( tid < GS2D_BATCH )
			sx.visited[tid] = 0u;
		cta_sync<NT>();
		solve_masks<NT>( sm, stage, nb, tid, tile_x0, tile_y0, g.W, g.H );
		cta_sync<NT>();

		unsigned anylive = 0;
#pragma unroll
		for( int j = 0; j < P; j++ )
			anylive |= livemask[j];
		if( anylive )
		{
#pragma unroll 1
			for( int k = 0; k < nb; k++ )
			{
				unsigned wd[P];
				load_words<P>( sm, k, warp, wd );
				unsigned hit = 0;
#pragma unroll
				for( int j = 0; j < P; j++ )
					hit |= wd[j] & livemask[j];
				if( hit == 0u )
					continue;
				const SplatRec& s = sm.rec[stage][k];
				const float4 q0 = *reinterpret_cast<const float4*>( &s.px ); // px py a b
				const float4 q1 = *reinterpret_cast<const float4*>( &s.d );  // d cr cg cb
				const float op = s.op;
				const float cT = s.cosT, sT = s.sinT;
				const float vx = xsub( fpx, q0.x );
				const float avx = xmul( q0.z, vx ), bvx = xmul( q0.w, vx );
				const float cvx = cT * vx, svx = sT * vx;
				float v[8] = { 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f };
				float vop = 0.0f;
#pragma unroll
				for( int j = 0; j < P; j++ )
				{
					if( ( wd[j] & livemask[j] ) == 0u )
						continue;
					if( ( ( wd[j] >> lane ) & 1u ) && !( T[j] < min_T ) )
					{
						const float vy = xsub( (float)( y0 + 2 * j ) + 0.5f, q0.y );
						const float rx = xadd( avx, xmul( q0.w, vy ) );
						const float ry = xadd( bvx, xmul( q1.x, vy ) );
						const float d2 = xadd( xmul( vx, rx ), xmul( vy, ry ) );
						const float G = gauss_of_d2( d2 );       // main.cpp:610
						const float alpha = xmul( G, op );       // main.cpp:611
						const float wgt = alpha * T[j];          // dC_dc, main.cpp:618
						const float dlr = DLR( j ), dlg = DLG( j ), dlb = DLB( j );
						v[0] = fmaf( dlr, wgt, v[0] );           // main.cpp:619
						v[1] = fmaf( dlg, wgt, v[1] );
						v[2] = fmaf( dlb, wgt, v[2] );
						// prefix colour, main.cpp:623-625 (same rounding as the forward pass)
						const float tcr = xmul( T[j], q1.y ), tcg = xmul( T[j], q1.z ), tcb = xmul( T[j], q1.w );
						Pr[j] = xadd( Pr[j], xmul( tcr, alpha ) );
						Pg[j] = xadd( Pg[j], xmul( tcg, alpha ) );
						Pb[j] = xadd( Pb[j], xmul( tcb, alpha ) );
						// S = colour from behind, main.cpp:627;  dC/dalpha = c*T - S/(1 - alpha + 1e-15), main.cpp:628
						const float oma = xsub( 1.0f, alpha );
						const float inv = __fdividef( 1.0f, xadd( oma, 1.0e-15f ) ); // MUFU.RCP: gradients are not bit-exact anyway
						const float dar = tcr - xsub( FR( j ), Pr[j] ) * inv;
						const float dag = tcg - xsub( FG( j ), Pg[j] ) * inv;
						const float dab = tcb - xsub( FB( j ), Pb[j] ) * inv;
						const float gsum = dlr * dar + dlg * dag + dlb * dab; // dL_dalpha_rgb, main.cpp:629-630
						const float ga = gsum * alpha;
						v[3] = fmaf( ga, rx, v[3] ); // 0.5*alpha*(2a vx + (b+c) vy) == alpha*(a vx + b vy), main.cpp:639, :654
						v[4] = fmaf( ga, ry, v[4] ); // main.cpp:640, :655
						// rotated frame: u = cos*vx + sin*vy, w = cos*vy - sin*vx  =>  the three quadratic forms of
						// main.cpp:657-662 and :680-683 are u*u, w*w and u*w
						const float u = fmaf( sT, vy, cvx );
						const float w = fmaf( cT, vy, -svx );
						const float gu = ga * u;
						v[5] = fmaf( gu, u, v[5] );      // * 1/sx^3 at flush time, main.cpp:657-659, :677
						v[6] = fmaf( ga * w, w, v[6] );  // * 1/sy^3, main.cpp:660-662, :678
						v[7] = fmaf( gu, w, v[7] );      // * (sx^2-sy^2)/(sx^2 sy^2), main.cpp:680-685
						if( OPACITY )
							vop = fmaf( gsum, G, vop ); // main.cpp:703-704
						T[j] = xmul( T[j], oma );       // main.cpp:707
					}
					livemask[j] = __ballot_sync( 0xffffffffu, !( T[j] < min_T ) );
				}
				const float r8 = warp_reduce8( v, lane );
				if( ( lane & 3 ) == 0 )
					sx.part[k][lane >> 2][warp] = r8;
				if( OPACITY )
				{
#pragma unroll
					for( int o = 16; o > 0; o >>= 1 )
						vop += __shfl_xor_sync( 0xffffffffu, vop, o );
					if( lane == 0 )
						sx.part[k][8][warp] = vop;
				}
				if( lane == 0 )
				{
					if( NW > 1 )
						atomicOr( &sx.visited[k], 1u << warp );
					else
						sx.visited[k] = 1u;
				}
			}
			anylive = 0;
#pragma unroll
			for( int j = 0; j < P; j++ )
				anylive |= livemask[j];
		}
		int any_alive;
		if( NT > 32 )
			any_alive = __syncthreads_or( anylive != 0u );
		else
		{
			any_alive = anylive != 0u;
			__syncwarp();
		}

		// flush: one atomic per (entry, parameter)
		for( int e = tid; e < nb * NV; e += NT )
		{
			const int k = e / NV, vi = e - k * NV;
			const unsigned vis = sx.visited[k];
			if( vis == 0u )
				continue;
			float sum = 0.0f;
#pragma unroll
			for( int w = 0; w < NW; w++ )
				if( vis & ( 1u << w ) )
					sum += sx.part[k][vi][w];
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
			atomicAdd( &grads[(size_t)plane * g.n_pad + sx.idx[stage][k]], sum );
		}
		if( !any_alive )
			break;
		cta_sync<NT>(); // part/visited/mask are rewritten by the next round
	}
	cp_async_wait<0>();
}

int launch_backward( const Geometry& g, const Buffers& b, int sorted_buf, bool with_opacity, cudaStream_t s )
{
	const int rows = g.band_ty1 - g.band_ty0;
	if( rows <= 0 || g.tiles_x <= 0 )
		return 0;
	dim3 grid( g.tiles_x, rows, g.n_images );
#define LAUNCH_BWD( P )                                                                                               \
	do                                                                                                                \
	{                                                                                                                 \
		if( with_opacity )                                                                                            \
			k_backward<P, true><<<grid, BlendCfg<P>::NT, 0, s>>>( g, b.rec, b.vals[sorted_buf], b.tile_range,         \
			                                                     b.target, b.image, b.grads, b.state );              \
		else                                                                                                          \
			k_backward<P, false><<<grid, BlendCfg<P>::NT, 0, s>>>( g, b.rec, b.vals[sorted_buf], b.tile_range,        \
			                                                      b.target, b.image, b.grads, b.state );             \
	} while( 0 )
	switch( gs2d_pixels_per_thread( 1 ) )
	{
	case 1: LAUNCH_BWD( 1 ); break;
	case 2: LAUNCH_BWD( 2 ); break;
	case 4: LAUNCH_BWD( 4 ); break;
	default: LAUNCH_BWD( 8 ); break;
	}
#undef LAUNCH_BWD
	return 1;
}
