// gs2d_forward.cu — per-tile front-to-back alpha blend (the reference's forward splat loop, main.cpp:419-538,
// turned inside out: one warp per 16x16 tile, the tile's splat list walked in index order == the reference's order).
//
// Per round, GS2D_BATCH list entries are gathered into shared memory with cp.async (double-buffered) and the exact
// scan-line footprint of every (splat, tile row) - solved once per (splat, image row) by k_row_table with the
// reference's float quadratic and x86 truncation (main.cpp:489-509) - is clipped to the tile and stored as coverage
// words of the eight 8x4 pixel blocks of the tile (gs2d_blend.cuh).
// The warp then walks the entries that touch a live pixel of the tile: a block is visited only when
// (coverage & live) != 0, and the visit is branch-free: uncovered or dead lanes blend with alpha = 0, which leaves
// their colour and throughput bit-for-bit unchanged (x + (+0) == x, T * 1 == T).
// A pixel is live while T >= MIN_THROUGHPUT (main.cpp:520); T only decreases, so a dead pixel never revives and a
// tile stops as soon as all its pixels are dead.
//
// Arithmetic, two modes (template parameter EXACT, tuning knob GS2D_EXACT):
//   exact   : the transmittance/colour chain is evaluated with the reference's operand order and WITHOUT FMA
//             contraction, so the image is bit-identical to the CPU reference whenever cosf/sinf agree (k_preprocess);
//             28 issue slots per block visit;
//   relaxed : (default) what the north star allows - "within a stated FP32 tolerance, blend order preserved":
//             d2 and exp_approx's 1 + x/8 collapse into two FMAs on per-entry column terms (the record carries the
//             inverse covariance pre-scaled by -opacity^(1/8)/16), the cut-off t < 1.8e-5 -> 0 (main.cpp:60) becomes
//             the saturation of that FMA (t in (0, 1.8e-5) gives alpha < 1e-38 instead of 0), colour accumulates as
//             C += (T alpha) c with FMAs and T -= T alpha: 17 issue slots per block visit. The covered-pixel set
//             (footprint), the blend order and the T < 1/256 skip rule are the reference's in both modes.
#include "gs2d_blend.cuh"

// MINB = resident warps (tiles) per SM the register allocation is capped for: 32 -> 64 registers, 24 -> 80, 20 -> 96.
// COUNT: also count the live (pixel, splat) pairs (statistics for the roofline; one more instruction per block visit).
// developer statistic (GS2D_VISIT_HIST=1, counting variant only): histogram of the number of covered-and-live lanes
// per 8x4 block visit, and of the live pixels left in the tile at each visit
__device__ unsigned long long g_visit_hist[2][33];

// PDL (default, GS2D_PDL=0 turns it off; gs2d_blend.cuh): the backward grid is launched with programmatic stream serialization and fills the
// SM slots the forward pass's last wave frees; a tile's forward warp publishes tile_done[tile] when its image pixels,
// coverage words and loss partial are written, and the tile's backward warp waits for exactly that.
template <int MINB, bool COUNT, bool PDL, bool EXACT>
__global__ void __launch_bounds__( 32, MINB ) k_forward( Geometry g, const SplatRec* __restrict__ rec,
                                                             const unsigned* __restrict__ vals,
                                                             const uint2* __restrict__ tile_range,
                                                             const float4* __restrict__ target,
                                                             float4* __restrict__ image,
                                                             unsigned* __restrict__ cover,
                                                             double* __restrict__ loss_partial,
                                                             unsigned long long* __restrict__ live_slots,
                                                             unsigned long long* __restrict__ staged_slots,
                                                             const DevState* __restrict__ st, int visit_hist,
                                                             unsigned* __restrict__ tile_done,
                                                             const unsigned* __restrict__ rowtab,
                                                             const RowRec* __restrict__ rowrec,
                                                             unsigned* __restrict__ round_vis )
{
	if( PDL )
		pdl_launch_dependents(); // the dependent grid may start once every forward warp has got this far (or exited)
	if( st->need_capacity )
		return;
	__shared__ TileSmem<1> sm;

	const int lane = threadIdx.x;
	const int tile_x = blockIdx.x, tile_y = blockIdx.y + g.band_ty0, img = blockIdx.z;
	const unsigned tile_id = (unsigned)( ( img * g.tiles_y + tile_y ) * g.tiles_x + tile_x );
	const int tile_x0 = tile_x * GS2D_TILE, tile_y0 = tile_y * GS2D_TILE;
	const int px_x = tile_x0 + ( lane & 7 ), px_y = tile_y0 + ( lane >> 3 ); // slot j: x + 8*(j&1), y + 4*(j>>1)
	// glm::vec2 p = { x + 0.5f, y + 0.5f }, main.cpp:523 (small integers + 0.5 are exact, so is adding 4 or 8)
	float fpx0 = (float)px_x + 0.5f, fpx1 = fpx0 + 8.0f;
	asm volatile( "" : "+f"( fpx0 ), "+f"( fpx1 ) ); // kept in registers (else rebuilt from the integer per entry)
	const float fpy0 = (float)px_y + 0.5f;
	// the four row coordinates of a lane, kept in registers (opaque to the compiler, which would otherwise rebuild
	// them from the integer at every block visit: I2FP + 2 FADD)
	float fpyr[4] = { fpy0, fpy0 + 4.0f, fpy0 + 8.0f, fpy0 + 12.0f };
#pragma unroll
	for( int r = 0; r < 4; r++ )
		asm volatile( "" : "+f"( fpyr[r] ) );
	const float min_T = g.min_throughput;
	unsigned lanebit = 1u << lane;
	asm volatile( "" : "+r"( lanebit ) ); // kept in a register (else rebuilt with a shift at every block visit)

	const uint2 range = tile_range[tile_id];
	const unsigned start = range.x, end = range.y;

	// main.cpp:414: colour 0, throughput 1. Pixels outside the image start dead (T = 0).
	float T[8], Cr[8], Cg[8], Cb[8];
	unsigned live[8];
#pragma unroll
	for( int j = 0; j < 8; j++ )
	{
		T[j] = ( px_x + 8 * ( j & 1 ) < g.W && px_y + 4 * ( j >> 1 ) < g.H ) ? 1.0f : 0.0f;
		Cr[j] = Cg[j] = Cb[j] = 0.0f;
		live[j] = __ballot_sync( 0xffffffffu, !( T[j] < min_T ) );
	}
	unsigned nlive = 0, nstaged = 0;

	// lane k gathers record k of the batch: 3 quads (48 bytes)
	{
		const bool v = start + lane < end;
		const unsigned idx0 = v ? vals[start + lane] : 0u;
		gather_record<3>( sm, 0, lane, rec, idx0, v );
		cp_async_commit();
	}
	unsigned idx_nxt = ( start + GS2D_BATCH + lane < end ) ? vals[start + GS2D_BATCH + lane] : 0u;

	int stage = 0;
	for( unsigned base = start; base < end; base += GS2D_BATCH, stage ^= 1 )
	{
		const int nb = min( (unsigned)GS2D_BATCH, end - base );
		const unsigned nbase = base + GS2D_BATCH;
		if( nbase < end )
		{
			gather_record<3>( sm, stage ^ 1, lane, rec, idx_nxt, nbase + lane < end );
			cp_async_commit();
			idx_nxt = ( nbase + GS2D_BATCH + lane < end ) ? vals[nbase + GS2D_BATCH + lane] : 0u;
			cp_async_wait<1>();
		}
		else
			cp_async_wait<0>();
		__syncwarp();
		build_words( sm, stage, nb, lane, tile_x0, tile_y, rowtab, rowrec, g.W, g.H );
		__syncwarp();
		nstaged += nb;
		// the coverage words go to HBM for the backward pass (32 bytes per staged entry, coalesced)
		uint4 cwa, cwb;
		touching_entries( sm, 0, nb, lane, cwa, cwb );
		if( lane < nb )
		{
			uint4* dst = reinterpret_cast<uint4*>( cover + ( (size_t)base + lane ) * 8 );
			__stcg( dst, cwa );
			__stcg( dst + 1, cwb );
		}
		// entries worth walking: those that cover a pixel that is still live now (lane k tests entry k). Pixels only
		// die, so an entry dropped here could not have contributed; one that loses its last live pixel while the
		// batch is walked fails its eight block tests below.
		unsigned ent = __ballot_sync( 0xffffffffu, ( ( cwa.x & live[0] ) | ( cwa.y & live[1] ) | ( cwa.z & live[2] ) |
		                                             ( cwa.w & live[3] ) | ( cwb.x & live[4] ) | ( cwb.y & live[5] ) |
		                                             ( cwb.z & live[6] ) | ( cwb.w & live[7] ) ) != 0u );

		// the entries of this round that blended into at least one pixel (bit = list position in the round): the backward
		// pass walks exactly these instead of testing every entry against its live masks again (round_visited)
		unsigned vis_bits = 0u;
#pragma unroll 1
		while( ent )
		{
			const int k = __ffs( ent ) - 1;
			ent &= ent - 1u;
			const unsigned kbit = 1u << k;
			const uint4 wa = *reinterpret_cast<const uint4*>( &sm.words[0][k][0] );
			const uint4 wb = *reinterpret_cast<const uint4*>( &sm.words[0][k][4] );
			const unsigned wd[8] = { wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w };
			const SplatRec& s = sm.rec[stage][k];
			const float4 q0 = *reinterpret_cast<const float4*>( &s.px ); // px py B D
			const float4 q1 = *reinterpret_cast<const float4*>( &s.A );  // A cr cg cb
			const float op = s.op;
			const float vx0 = xsub( fpx0, q0.x ), vx1 = xsub( fpx1, q0.x );
			if( EXACT )
			{
				const float avx0 = xmul( q1.x, vx0 ), avx1 = xmul( q1.x, vx1 );
				const float bvx0 = xmul( q0.z, vx0 ), bvx1 = xmul( q0.z, vx1 );
				const f32x2 bd = pack2( q0.z, q0.w ), cgb = pack2( q1.z, q1.w ); // packed multiplies: see xmul2
#pragma unroll
				for( int j = 0; j < 8; j++ )
				{
					const unsigned wl = wd[j] & live[j];
					if( wl == 0u ) // warp-uniform: no covered live pixel in this 8x4 block
						continue;
					if( COUNT )
						nlive += count_visit( wl, lanebit, live, visit_hist, lane, g_visit_hist );
					vis_bits |= kbit;
					const float vx = ( j & 1 ) ? vx1 : vx0, avx = ( j & 1 ) ? avx1 : avx0, bvx = ( j & 1 ) ? bvx1 : bvx0;
					const float vy = xsub( fpyr[j >> 1], q0.y );
					// d2 = dot( v, inv_cov * v ), main.cpp:526 (inv_cov[0][1] == inv_cov[1][0] == b)
					float bvy, dvy, vxrx, vyry;
					unpack2( xmul2( bd, pack2( vy, vy ) ), bvy, dvy );
					const float rx = xadd( avx, bvy );
					const float ry = xadd( bvx, dvy );
					unpack2( xmul2( pack2( vx, vy ), pack2( rx, ry ) ), vxrx, vyry );
					const float d2 = xadd( vxrx, vyry );
					// exp_approx( -0.5f * d2 ), main.cpp:49-83 (see gauss_of_d2), and alpha, main.cpp:527
					const float t = __fmaf_rn( d2, -0.0625f, 1.0f );
					const float t2 = xmul( t, t ), t4 = xmul( t2, t2 ), t8 = xmul( t4, t4 );
					// alpha = ( cov && !( t < threshold ) ) ? t8 * op : 0 with the two conditions folded into one predicate
					const float alpha = select_alpha( wl & lanebit, t, xmul( t8, op ) );
					float tga, tba;
					unpack2( xmul2( xmul2( pack2( T[j], T[j] ), cgb ), pack2( alpha, alpha ) ), tga, tba );
					Cr[j] = xadd( Cr[j], xmul( xmul( T[j], q1.y ), alpha ) );   // main.cpp:529-531
					Cg[j] = xadd( Cg[j], tga );
					Cb[j] = xadd( Cb[j], tba );
					T[j] = xmul( T[j], xsub( 1.0f, alpha ) );                   // main.cpp:533
					live[j] = __ballot_sync( 0xffffffffu, !( T[j] < min_T ) );
				}
			}
			else
			{
				// s (1 - d2/16) = ( s + A vx^2 ) + vy ( B vx + D vy ), A B D pre-scaled (SplatRec); per entry and block
				// column: B vx and s + A vx^2
				const float bvx0 = __fmul_rn( q0.z, vx0 ), bvx1 = __fmul_rn( q0.z, vx1 );
				const float axx0 = __fmaf_rn( __fmul_rn( q1.x, vx0 ), vx0, op );
				const float axx1 = __fmaf_rn( __fmul_rn( q1.x, vx1 ), vx1, op );
#pragma unroll
				for( int j = 0; j < 8; j++ )
				{
					const unsigned wl = wd[j] & live[j];
					if( wl == 0u ) // warp-uniform: no covered live pixel in this 8x4 block
						continue;
					if( COUNT )
						nlive += count_visit( wl, lanebit, live, visit_hist, lane, g_visit_hist );
					vis_bits |= kbit;
					const float vy = xsub( fpyr[j >> 1], q0.y );
					const float alpha = relaxed_alpha( wl & lanebit, vy, q0.w, ( j & 1 ) ? bvx1 : bvx0, ( j & 1 ) ? axx1 : axx0 );
					const float ta = __fmul_rn( T[j], alpha );
					Cr[j] = __fmaf_rn( ta, q1.y, Cr[j] );                       // main.cpp:529-531
					Cg[j] = __fmaf_rn( ta, q1.z, Cg[j] );
					Cb[j] = __fmaf_rn( ta, q1.w, Cb[j] );
					T[j] = __fsub_rn( T[j], ta );                               // main.cpp:533
					live[j] = __ballot_sync( 0xffffffffu, !( T[j] < min_T ) );
				}
			}
		}
		if( lane == 0 )
			__stcg( round_vis + round_vis_index( start, tile_id, base ), vis_bits );
		unsigned anylive = 0;
#pragma unroll
		for( int j = 0; j < 8; j++ )
			anylive |= live[j];
		if( anylive == 0u )
			break;
		__syncwarp();
	}
	cp_async_wait<0>();

	// epilogue: image0 with w reset to 1 (main.cpp:543-546) and this tile's share of the MSE sum (main.cpp:796-805)
	double e = 0.0;
#pragma unroll
	for( int j = 0; j < 8; j++ )
	{
		const int x = px_x + 8 * ( j & 1 ), y = px_y + 4 * ( j >> 1 );
		if( x < g.W && y < g.H )
		{
			const size_t pix = ( (size_t)img * g.H + y ) * g.W + x;
			const float4 ref = load_target( g, target, pix );
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
		nlive += __shfl_xor_sync( 0xffffffffu, nlive, o );
	}
	if( lane == 0 )
	{
		loss_partial[tile_id] = e;
		atomicAdd( &live_slots[tile_id & ( GS2D_LIVE_SLOTS - 1 )], (unsigned long long)nlive );
		atomicAdd( &staged_slots[tile_id & ( GS2D_LIVE_SLOTS - 1 )], (unsigned long long)nstaged );
	}
	if( PDL )
		pdl_publish_tile( tile_done + tile_id, lane );
}

// both PDL blend kernels ask for the same shared-memory carveout, so that warps of the two grids can share an SM
void prepare_forward_pdl()
{
#define CARVE( COUNT, EXACT )                                                                                         \
	cudaFuncSetAttribute( k_forward<GS2D_FWD_MINB_DEFAULT, COUNT, true, EXACT>,                                      \
	                      cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared )
	CARVE( false, false );
	CARVE( true, false );
	CARVE( false, true );
	CARVE( true, true );
#undef CARVE
}

template <int MINB, bool COUNT, bool PDL, bool EXACT>
static void launch_forward_as( dim3 grid, const Geometry& g, const Buffers& b, int sorted_buf, int vh, cudaStream_t s )
{
	k_forward<MINB, COUNT, PDL, EXACT><<<grid, 32, 0, s>>>( g, b.rec, b.vals[sorted_buf], b.tile_range,
	                                                       g.target_u8 ? reinterpret_cast<const float4*>( b.target8 ) : b.target, b.image,
	                                                       b.cover, b.loss_partial, b.live_slots, b.staged_slots, b.state,
	                                                       COUNT ? vh : 0, PDL ? b.tile_done : nullptr, b.rowtab, b.rowrec,
	                                                       b.cover + 8 * (size_t)b.capacity );
}
template <int MINB, bool PDL>
static void launch_forward_minb( dim3 grid, const Geometry& g, const Buffers& b, int sorted_buf, bool count_live,
                                 bool exact, int vh, cudaStream_t s )
{
	if( count_live )
	{
		if( exact )
			launch_forward_as<MINB, true, PDL, true>( grid, g, b, sorted_buf, vh, s );
		else
			launch_forward_as<MINB, true, PDL, false>( grid, g, b, sorted_buf, vh, s );
	}
	else
	{
		if( exact )
			launch_forward_as<MINB, false, PDL, true>( grid, g, b, sorted_buf, vh, s );
		else
			launch_forward_as<MINB, false, PDL, false>( grid, g, b, sorted_buf, vh, s );
	}
}

int launch_forward( const Geometry& g, const Buffers& b, int sorted_buf, bool count_live, cudaStream_t s, bool pdl )
{
	cudaMemsetAsync( b.live_slots, 0, sizeof( unsigned long long ) * GS2D_LIVE_SLOTS * 2, s );
	const int rows = g.band_ty1 - g.band_ty0;
	if( rows <= 0 || g.tiles_x <= 0 )
		return 0;
	dim3 grid( g.tiles_x, rows, g.n_images );
	const int vh = gs2d_tune( "GS2D_VISIT_HIST", 0 );
	const bool exact = gs2d_exact_mode();
	if( pdl ) // only the default register cap (32 warps per SM) has PDL twins: gs2d_pdl_usable()
	{
		cudaMemsetAsync( b.tile_done, 0, sizeof( unsigned ) * (size_t)g.n_tiles_total, s );
		launch_forward_minb<GS2D_FWD_MINB_DEFAULT, true>( grid, g, b, sorted_buf, count_live, exact, vh, s );
		return 1;
	}
	switch( gs2d_tune( "GS2D_FWD_MINB", GS2D_FWD_MINB_DEFAULT ) )
	{
	case 24: launch_forward_minb<24, false>( grid, g, b, sorted_buf, count_live, exact, vh, s ); break;
	default: launch_forward_minb<32, false>( grid, g, b, sorted_buf, count_live, exact, vh, s ); break;
	}
	return 1;
}

// developer statistic, see g_visit_hist
extern "C" int gs2d_debug_visit_hist( unsigned long long out[66], int reset )
{
	if( !out )
		return GS2D_ERR_INVALID;
	if( cudaMemcpyFromSymbol( out, g_visit_hist, sizeof( unsigned long long ) * 66 ) != cudaSuccess )
		return GS2D_ERR_CUDA;
	if( reset )
	{
		unsigned long long z[66] = { 0 };
		cudaMemcpyToSymbol( g_visit_hist, z, sizeof( z ) );
	}
	return GS2D_OK;
}
