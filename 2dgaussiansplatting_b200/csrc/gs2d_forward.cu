// gs2d_forward.cu — per-tile front-to-back alpha blend (the reference's forward splat loop, main.cpp:419-538,
// turned inside out: one CTA per 16x16 tile, the tile's splat list walked in index order == the reference's order).
//
// Per round, GS2D_BATCH list entries are gathered into shared memory with cp.async (double-buffered, the indices of
// the round after next already in flight), the exact scan-line footprint of every (splat, tile row) is solved ONCE
// (row_mask: main.cpp:489-509 incl. the float quadratic and the x86 truncation) and stored as a 16-bit column mask,
// and then every warp walks the batch over its 2P rows (see gs2d_blend.cuh for the pixel-to-thread map): one 32-bit
// word per (splat, 16x2 strip) says which pixels are covered, strips without a covered live pixel cost one AND.
// A pixel is live while T >= MIN_THROUGHPUT (main.cpp:520); T only decreases, so a dead pixel never revives and a
// tile stops as soon as all its pixels are dead.
//
// Arithmetic: the transmittance/colour chain is evaluated with the reference's operand order and WITHOUT FMA
// contraction, so the image is bit-identical to the CPU reference whenever cosf/sinf agree (see k_preprocess).
#include "gs2d_blend.cuh"

#ifndef FWD_MINB
#define FWD_MINB 32   // P = 8 (one warp per tile): cap at 64 registers -> 32 resident warps per SM (measured best)
#endif

template <int P>
__global__ void __launch_bounds__( BlendCfg<P>::NT, P == 8 ? FWD_MINB : 1 ) k_forward( Geometry g, const SplatRec* __restrict__ rec,
                                                                const unsigned* __restrict__ vals,
                                                                const uint2* __restrict__ tile_range,
                                                                const float4* __restrict__ target,
                                                                float4* __restrict__ image,
                                                                double* __restrict__ loss_partial,
                                                                unsigned long long* __restrict__ live_slots,
                                                                unsigned long long* __restrict__ staged_slots,
                                                                const DevState* __restrict__ st )
{
	constexpr int NW = BlendCfg<P>::NW, NT = BlendCfg<P>::NT;
	if( st->need_capacity )
		return;
	__shared__ StageSmem sm;
	__shared__ double red[NW];
	__shared__ unsigned red_live[NW];

	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const int tile_x = blockIdx.x, tile_y = blockIdx.y + g.band_ty0, img = blockIdx.z;
	const unsigned tile_id = (unsigned)( ( img * g.tiles_y + tile_y ) * g.tiles_x + tile_x );
	const int tile_x0 = tile_x * GS2D_TILE, tile_y0 = tile_y * GS2D_TILE;
	const int x = tile_x0 + ( lane & 15 ), y0 = tile_y0 + warp * 2 * P + ( lane >> 4 ); // slot j: y = y0 + 2j
	const float fpx = (float)x + 0.5f; // glm::vec2 p = { x + 0.5f, y + 0.5f }, main.cpp:523
	const float min_T = g.min_throughput;

	const uint2 range = tile_range[tile_id];
	const unsigned start = range.x, end = range.y;

	// main.cpp:414: colour 0, throughput 1. Pixels outside the image start dead (T = 0).
	float T[P], Cr[P], Cg[P], Cb[P];
	unsigned livemask[P];
#pragma unroll
	for( int j = 0; j < P; j++ )
	{
		T[j] = ( x < g.W && y0 + 2 * j < g.H ) ? 1.0f : 0.0f;
		Cr[j] = Cg[j] = Cb[j] = 0.0f;
		livemask[j] = __ballot_sync( 0xffffffffu, !( T[j] < min_T ) );
	}
	unsigned nlive = 0, nstaged = 0;

	// gather role: 3 quads (48 bytes) per record
	constexpr int GI = ( GS2D_BATCH * 3 + NT - 1 ) / NT;
	unsigned idx_nxt[GI];
#pragma unroll
	for( int i = 0; i < GI; i++ )
	{
		const int t = tid + i * NT, gk = t / 3, gq = t - gk * 3;
		idx_nxt[i] = 0;
		if( t < GS2D_BATCH * 3 && start + gk < end )
		{
			const unsigned idx0 = vals[start + gk];
			cp_async16( reinterpret_cast<float4*>( &sm.rec[0][gk] ) + gq,
			            reinterpret_cast<const float4*>( rec + idx0 ) + gq );
		}
	}
	cp_async_commit();
#pragma unroll
	for( int i = 0; i < GI; i++ )
	{
		const int t = tid + i * NT, gk = t / 3;
		if( t < GS2D_BATCH * 3 && start + GS2D_BATCH + gk < end )
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
				const int t = tid + i * NT, gk = t / 3, gq = t - gk * 3;
				if( t < GS2D_BATCH * 3 && nbase + gk < end )
					cp_async16( reinterpret_cast<float4*>( &sm.rec[stage ^ 1][gk] ) + gq,
					            reinterpret_cast<const float4*>( rec + idx_nxt[i] ) + gq );
			}
			cp_async_commit();
#pragma unroll
			for( int i = 0; i < GI; i++ )
			{
				const int t = tid + i * NT, gk = t / 3;
				if( t < GS2D_BATCH * 3 && nbase + GS2D_BATCH + gk < end )
					idx_nxt[i] = vals[nbase + GS2D_BATCH + gk];
			}
			cp_async_wait<1>();
		}
		else
			cp_async_wait<0>();
		cta_sync<NT>();
		solve_masks<NT>( sm, stage, nb, tid, tile_x0, tile_y0, g.W, g.H );
		cta_sync<NT>();
		nstaged += nb;

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
				const float vx = xsub( fpx, q0.x );
				const float avx = xmul( q0.z, vx ), bvx = xmul( q0.w, vx );
#pragma unroll
				for( int j = 0; j < P; j++ )
				{
					if( ( wd[j] & livemask[j] ) == 0u )
						continue;
					nlive += __popc( wd[j] & livemask[j] ); // warp-uniform count of the live pairs blended below
					if( ( ( wd[j] >> lane ) & 1u ) && !( T[j] < min_T ) ) // covered (main.cpp:492-514) and live (:520)
					{
						const float vy = xsub( (float)( y0 + 2 * j ) + 0.5f, q0.y );
						// d2 = dot( v, inv_cov * v ), main.cpp:526 (inv_cov[0][1] == inv_cov[1][0] == b)
						const float rx = xadd( avx, xmul( q0.w, vy ) );
						const float ry = xadd( bvx, xmul( q1.x, vy ) );
						const float d2 = xadd( xmul( vx, rx ), xmul( vy, ry ) );
						const float alpha = xmul( gauss_of_d2( d2 ), op );          // main.cpp:527
						Cr[j] = xadd( Cr[j], xmul( xmul( T[j], q1.y ), alpha ) );   // main.cpp:529-531
						Cg[j] = xadd( Cg[j], xmul( xmul( T[j], q1.z ), alpha ) );
						Cb[j] = xadd( Cb[j], xmul( xmul( T[j], q1.w ), alpha ) );
						T[j] = xmul( T[j], xsub( 1.0f, alpha ) );                   // main.cpp:533
					}
					livemask[j] = __ballot_sync( 0xffffffffu, !( T[j] < min_T ) );
				}
			}
			anylive = 0;
#pragma unroll
			for( int j = 0; j < P; j++ )
				anylive |= livemask[j];
		}
		if( NT > 32 )
		{
			if( !__syncthreads_or( anylive != 0u ) )
				break;
		}
		else
		{
			if( anylive == 0u )
				break;
			__syncwarp();
		}
	}
	cp_async_wait<0>();

	// epilogue: image0 with w reset to 1 (main.cpp:543-546) and this tile's share of the MSE sum (main.cpp:796-805)
	double e = 0.0;
#pragma unroll
	for( int j = 0; j < P; j++ )
	{
		const int y = y0 + 2 * j;
		if( x < g.W && y < g.H )
		{
			const size_t pix = ( (size_t)img * g.H + y ) * g.W + x;
			const float4 ref = target[pix];
			image[pix] = make_float4( Cr[j], Cg[j], Cb[j], 1.0f );
			const float dr = xmul( xsub( Cr[j], ref.x ), 255.0f ), dg = xmul( xsub( Cg[j], ref.y ), 255.0f ),
			            db = xmul( xsub( Cb[j], ref.z ), 255.0f );
			e += (double)xadd( xadd( xmul( dr, dr ), xmul( dg, dg ) ), xmul( db, db ) );
		}
	}
#pragma unroll
	for( int o = 16; o > 0; o >>= 1 )
	{
		e += __shfl_xor_sync( 0xffffffffu, e, o );
	}
	if( lane == 0 )
	{
		red[warp] = e;
		red_live[warp] = nlive;
	}
	cta_sync<NT>();
	if( tid == 0 )
	{
		double sum = 0.0;
		unsigned long long lv = 0;
#pragma unroll
		for( int w = 0; w < NW; w++ )
		{
			sum += red[w];
			lv += red_live[w];
		}
		loss_partial[tile_id] = sum;
		atomicAdd( &live_slots[tile_id & ( GS2D_LIVE_SLOTS - 1 )], lv );
		atomicAdd( &staged_slots[tile_id & ( GS2D_LIVE_SLOTS - 1 )], (unsigned long long)nstaged );
	}
}

int launch_forward( const Geometry& g, const Buffers& b, int sorted_buf, cudaStream_t s )
{
	cudaMemsetAsync( b.live_slots, 0, sizeof( unsigned long long ) * GS2D_LIVE_SLOTS * 2, s );
	const int rows = g.band_ty1 - g.band_ty0;
	if( rows <= 0 || g.tiles_x <= 0 )
		return 0;
	dim3 grid( g.tiles_x, rows, g.n_images );
#define LAUNCH_FWD( P )                                                                                               \
	k_forward<P><<<grid, BlendCfg<P>::NT, 0, s>>>( g, b.rec, b.vals[sorted_buf], b.tile_range, b.target, b.image,     \
	                                               b.loss_partial, b.live_slots, b.staged_slots, b.state )
	switch( gs2d_pixels_per_thread( 0 ) )
	{
	case 1: LAUNCH_FWD( 1 ); break;
	case 2: LAUNCH_FWD( 2 ); break;
	case 4: LAUNCH_FWD( 4 ); break;
	default: LAUNCH_FWD( 8 ); break;
	}
#undef LAUNCH_FWD
	return 1;
}
