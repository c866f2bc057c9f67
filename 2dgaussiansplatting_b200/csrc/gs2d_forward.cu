// gs2d_forward.cu — per-tile front-to-back alpha blend (the reference's forward splat loop, main.cpp:419-538,
// turned inside out: one CTA per 16x16 tile, one pixel per thread, the tile's splat list walked in index order).
//
// Per round, GS2D_BATCH list entries are gathered into shared memory with cp.async (double-buffered, the indices of
// the round after next already in flight), the exact scan-line footprint of every (splat, tile row) is solved ONCE
// by one thread (row_mask: main.cpp:489-509 incl. the float quadratic and the x86 truncation) and stored as a 16-bit
// column mask, and then every warp (2 tile rows) walks the batch: one 32-bit word per (splat, warp) says which of
// its 32 pixels are covered; warps skip splats that touch none of their still-live pixels without a vote.
// A pixel is live while T >= MIN_THROUGHPUT (main.cpp:520); T only decreases, so a dead pixel never revives and a
// tile stops as soon as all its pixels are dead (block-wide vote per round).
//
// Arithmetic: the transmittance/colour chain is evaluated with the reference's operand order and WITHOUT FMA
// contraction, so the image is bit-identical to the CPU reference whenever cosf/sinf agree (see k_preprocess).
#include "gs2d_internal.cuh"

struct FwdSmem
{
	SplatRec rec[2][GS2D_BATCH];                 // 4 KB
	unsigned short mask[GS2D_BATCH][GS2D_TILE];  // 1 KB, read back as 8 words per splat
};

__global__ void __launch_bounds__( 256 ) k_forward( Geometry g, const SplatRec* __restrict__ rec,
                                                    const unsigned* __restrict__ vals,
                                                    const uint2* __restrict__ tile_range,
                                                    const float4* __restrict__ target, float4* __restrict__ image,
                                                    double* __restrict__ loss_partial,
                                                    unsigned long long* __restrict__ live_slots,
                                                    unsigned long long* __restrict__ staged_slots,
                                                    float min_throughput, const DevState* __restrict__ st )
{
	if( st->need_capacity )
		return;
	__shared__ FwdSmem sm;
	__shared__ double red[8];

	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const int tile_x = blockIdx.x, tile_y = blockIdx.y + g.band_ty0, img = blockIdx.z;
	const unsigned tile_id = (unsigned)( ( img * g.tiles_y + tile_y ) * g.tiles_x + tile_x );
	const int tile_x0 = tile_x * GS2D_TILE, tile_y0 = tile_y * GS2D_TILE;
	const int x = tile_x0 + ( lane & 15 ), y = tile_y0 + warp * 2 + ( lane >> 4 );
	const bool inside = x < g.W && y < g.H;
	const float fpx = (float)x + 0.5f, fpy = (float)y + 0.5f; // glm::vec2 p = { x + 0.5f, y + 0.5f }, main.cpp:523

	const uint2 range = tile_range[tile_id];
	const unsigned start = range.x, end = range.y;

	float T = 1.0f, Cr = 0.0f, Cg = 0.0f, Cb = 0.0f; // main.cpp:414
	bool alive = inside;
	unsigned nlive = 0, nstaged = 0;

	// gather role: threads 0..95 move quad (tid % 3) of record (tid / 3)
	const int gk = tid / 3, gq = tid - gk * 3;
	const bool gatherer = tid < GS2D_BATCH * 3;
	unsigned idx_nxt = 0;
	if( gatherer && start + gk < end )
	{
		unsigned idx0 = vals[start + gk];
		cp_async16( reinterpret_cast<float4*>( &sm.rec[0][gk] ) + gq, reinterpret_cast<const float4*>( rec + idx0 ) + gq );
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
				cp_async16( reinterpret_cast<float4*>( &sm.rec[stage ^ 1][gk] ) + gq,
				            reinterpret_cast<const float4*>( rec + idx_nxt ) + gq );
			cp_async_commit();
			if( gatherer && nbase + GS2D_BATCH + gk < end )
				idx_nxt = vals[nbase + GS2D_BATCH + gk];
			cp_async_wait<1>();
		}
		else
			cp_async_wait<0>();
		__syncthreads();

		// exact scan-line footprint: thread -> (splat k, tile row r), two splats per thread
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
		nstaged += nb;

		if( livemask )
		{
#pragma unroll 1
			for( int k = 0; k < nb; k++ )
			{
				const unsigned word = words[k * ( GS2D_TILE / 2 )];
				if( ( word & livemask ) == 0u )
					continue;
				if( ( ( word >> lane ) & 1u ) && alive )
				{
					const SplatRec& s = sm.rec[stage][k];
					const float4 q0 = *reinterpret_cast<const float4*>( &s.px );
					const float4 q1 = *reinterpret_cast<const float4*>( &s.d );
					const float op = s.op;
					const float vx = xsub( fpx, q0.x ), vy = xsub( fpy, q0.y );
					// d2 = dot( v, inv_cov * v ), main.cpp:526 (inv_cov[0][1] == inv_cov[1][0] == b)
					const float rx = xadd( xmul( q0.z, vx ), xmul( q0.w, vy ) );
					const float ry = xadd( xmul( q0.w, vx ), xmul( q1.x, vy ) );
					const float d2 = xadd( xmul( vx, rx ), xmul( vy, ry ) );
					const float alpha = xmul( gauss_of_d2( d2 ), op ); // main.cpp:527
					Cr = xadd( Cr, xmul( xmul( T, q1.y ), alpha ) );    // main.cpp:529-531
					Cg = xadd( Cg, xmul( xmul( T, q1.z ), alpha ) );
					Cb = xadd( Cb, xmul( xmul( T, q1.w ), alpha ) );
					T = xmul( T, xsub( 1.0f, alpha ) );                 // main.cpp:533
					nlive++;
					alive = !( T < min_throughput );                   // main.cpp:520
				}
				livemask = __ballot_sync( 0xffffffffu, alive );
				if( livemask == 0u )
					break;
			}
		}
		if( !__syncthreads_or( livemask != 0u ) )
			break;
	}
	cp_async_wait<0>();

	// epilogue: image0 with w reset to 1 (main.cpp:543-546) and this tile's share of the MSE sum (main.cpp:796-805)
	double e = 0.0;
	if( inside )
	{
		const size_t pix = ( (size_t)img * g.H + y ) * g.W + x;
		const float4 ref = target[pix];
		image[pix] = make_float4( Cr, Cg, Cb, 1.0f );
		const float dr = xmul( xsub( Cr, ref.x ), 255.0f ), dg = xmul( xsub( Cg, ref.y ), 255.0f ),
		            db = xmul( xsub( Cb, ref.z ), 255.0f );
		e = (double)xadd( xadd( xmul( dr, dr ), xmul( dg, dg ) ), xmul( db, db ) );
	}
#pragma unroll
	for( int o = 16; o > 0; o >>= 1 )
	{
		e += __shfl_xor_sync( 0xffffffffu, e, o );
		nlive += __shfl_xor_sync( 0xffffffffu, nlive, o );
	}
	__shared__ unsigned red_live[8];
	if( lane == 0 )
	{
		red[warp] = e;
		red_live[warp] = nlive;
	}
	__syncthreads();
	if( tid == 0 )
	{
		double sum = 0.0;
		unsigned long long lv = 0;
#pragma unroll
		for( int w = 0; w < 8; w++ )
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
	k_forward<<<grid, 256, 0, s>>>( g, b.rec, b.vals[sorted_buf], b.tile_range, b.target, b.image, b.loss_partial,
	                                b.live_slots, b.staged_slots, g.min_throughput, b.state );
	return 1;
}
