// gs2d_backward.cu — analytic gradients of the blend (the reference's backward loop, main.cpp:552-712).
//
// Same tiling, staging, pixel-to-lane map and exact footprint as k_forward (gs2d_blend.cuh). The reference's
// backward walks the splats in FORWARD order, rebuilds the prefix colour/transmittance (image1) and obtains the
// colour from behind by subtraction, S = final - prefix (main.cpp:623-627); alpha is not clamped below 1, so the
// 3DGS-style reverse sweep (divide T by 1-alpha) is unsafe here and the forward-order sweep is what preserves the
// reference's numbers. The transmittance chain T repeats the forward kernel's contraction-free
// arithmetic, so the set of live pixels and the point where a tile terminates are identical to the forward pass.
//
// What a live pair costs (per pixel state: T, Q and the constant dL/dC — 5 registers instead of 10):
//   The reference needs S only inside g = sum_c dL/dC_c * (c_c*T - S_c/(1-alpha+1e-15)) (main.cpp:628-630), i.e. only
//   the scalar Q = sum_c dL/dC_c * S_c. With D = T * sum_c dL/dC_c * c_c the prefix update P += T*c*alpha
//   (main.cpp:623-625) becomes Q -= alpha*D and g = D - Q/(1-alpha+1e-15). When the pixel dies at this splat, nothing
//   behind it contributes, the reference's S is exactly 0 (prefix == final bit for bit), and Q is snapped to 0.
//   The five geometric derivatives (main.cpp:635-685) are linear in the moments of w = g*alpha over the pixels:
//   sum w, sum w*vy, sum w*vy^2 per lane and block column (vx is constant per lane and column), combined into
//   sum w*{vx, vy, vx^2, vx*vy, vy^2} once per (entry, lane): 4 instructions per pair instead of 12.
// Per entry the warp reduces 8 values (3 colour sums + 5 moments) over its lanes with a transposed butterfly
// (9 shuffles instead of 40); at the end of a round lane k turns the moments of entry k into the reference's
// d{pos, sx, sy, rot} with the hoisted per-splat factors and issues ONE red.global.add.f32 per (tile, splat, parameter).
//
// Arithmetic modes (template parameter EXACT, tuning knob GS2D_EXACT; gs2d_forward.cu): in both, alpha and the
// transmittance chain are evaluated exactly as the forward kernel of the same mode evaluates them, so the two passes
// walk the same live set. Relaxed mode reads the pre-scaled record (SplatRec), needs two FMAs for alpha's argument
// and gets d/d opacity from the moments (sum g*G = sum g*alpha / opacity) instead of a sum of its own.
#include "gs2d_blend.cuh"


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

// MINB = resident warps (tiles) per SM the register allocation is capped for: 20 -> 96 registers, 16 -> 128.
// COUNT: also count the live (pixel, splat) pairs into live_slots (same set as the forward pass).
// lane k gathers the 8 coverage words the forward pass stored for list position base + k
__device__ __forceinline__ void gather_cover( TileSmem<2>& sm, int stage, int lane, const unsigned* __restrict__ cover,
                                              unsigned base, bool valid )
{
	if( valid )
	{
		const uint4* src = reinterpret_cast<const uint4*>( cover + ( (size_t)base + lane ) * 8 );
		uint4* dst = reinterpret_cast<uint4*>( &sm.words[stage][lane][0] );
		cp_async16( dst, src );
		cp_async16( dst + 1, src + 1 );
	}
}

// Totals of the n (<= STAGE_SLOTS) entries parked in the staging area: lane (half = lane >> 4, slot = (lane >> 3) & 1,
// value = lane & 7) adds 16 of the 32 lane-partials of one value of one entry, one shuffle joins the halves. A slot is
// padded by 8 floats and the halves are 16 lanes apart, so the 32 lanes of a load hit 32 different banks
// ((8 slot + 8 l + value) mod 32 over half * 128) with plain immediate offsets. The totals go to
// words_of_stage[k][value], k = the entry's list position.
#define STAGE_SLOTS 2
#define STAGE_STRIDE ( 32 * 8 + 8 )
__device__ __forceinline__ void reduce_group( const float ( *stg )[STAGE_STRIDE], unsigned* words_of_stage,
                                              unsigned grp_k, unsigned n, int lane )
{
	__syncwarp();
	const unsigned half = (unsigned)lane >> 4, slot = ( (unsigned)lane >> 3 ) & 1u, val = (unsigned)lane & 7u;
	const float* src = &stg[slot][half * 128u + val];
	float s0 = src[0], s1 = src[8], s2 = src[16], s3 = src[24]; // (not 0 + x: the compiler keeps that add, -0 + 0 is +0)
#pragma unroll
	for( int l = 4; l < 16; l += 4 )
	{
		s0 += src[l * 8];
		s1 += src[l * 8 + 8];
		s2 += src[l * 8 + 16];
		s3 += src[l * 8 + 24];
	}
	float tot = ( s0 + s1 ) + ( s2 + s3 );
	tot += __shfl_xor_sync( 0xffffffffu, tot, 16 );
	if( half == 0u && slot < n )
	{
		const unsigned k = ( grp_k >> ( 8u * slot ) ) & 0xffu;
		words_of_stage[k * 8u + val] = __float_as_uint( tot );
	}
	__syncwarp();
}

// PDL: launched behind k_forward<.., true> with programmatic stream serialization; waits for its own tile's forward
// pass instead of the whole grid (gs2d_blend.cuh).
template <int MINB, bool OPACITY, bool COUNT, bool PDL, bool EXACT>
__global__ void __launch_bounds__( 32, MINB ) k_backward( Geometry g, const SplatRec* __restrict__ rec,
                                                              const unsigned* __restrict__ vals,
                                                              const uint2* __restrict__ tile_range,
                                                              const float4* __restrict__ target,
                                                              const float4* __restrict__ image,
                                                              const unsigned* __restrict__ cover,
                                                              float* __restrict__ grads,
                                                              unsigned long long* __restrict__ live_slots,
                                                              const DevState* __restrict__ st,
                                                              const unsigned* tile_done,
                                                              const unsigned* __restrict__ round_vis )
{
	if( st->need_capacity )
	{
		if( PDL )
			pdl_grid_wait();
		return;
	}
	__shared__ TileSmem<2> sm;
	// Cross-lane sums. Every lane parks its 8 partial sums of a visited entry here (two 128-bit stores); after two
	// entries the warp adds them up transposed (reduce_group: 16 conflict-free loads + adds and one shuffle for two
	// entries, where a shuffle butterfly costs ~45 instructions per entry).
	// The totals of entry k overwrite its own coverage words sm.words[stage][k][0..7], which are dead once the entry
	// has been walked - so no separate array is needed for them.
	__shared__ __align__( 16 ) float stage4[STAGE_SLOTS][STAGE_STRIDE]; // [slot][lane * 8 + value], padded
	__shared__ float vop_tot[OPACITY ? GS2D_BATCH : 1]; // per-entry total of the opacity gradient

	const int lane = threadIdx.x;
	const int tile_x = blockIdx.x, tile_y = blockIdx.y + g.band_ty0, img = blockIdx.z;
	const unsigned tile_id = (unsigned)( ( img * g.tiles_y + tile_y ) * g.tiles_x + tile_x );
	const int tile_x0 = tile_x * GS2D_TILE, tile_y0 = tile_y * GS2D_TILE;
	const int px_x = tile_x0 + ( lane & 7 ), px_y = tile_y0 + ( lane >> 3 );
	float fpx0 = (float)px_x + 0.5f, fpx1 = fpx0 + 8.0f;
	asm volatile( "" : "+f"( fpx0 ), "+f"( fpx1 ) ); // kept in registers (else rebuilt from the integer per entry)
	const float fpy0 = (float)px_y + 0.5f;
	// row coordinates and lane bit kept in registers (see k_forward)
	float fpyr[4] = { fpy0, fpy0 + 4.0f, fpy0 + 8.0f, fpy0 + 12.0f };
#pragma unroll
	for( int r = 0; r < 4; r++ )
		asm volatile( "" : "+f"( fpyr[r] ) );
	const float min_T = g.min_throughput;
	unsigned lanebit = 1u << lane;
	asm volatile( "" : "+r"( lanebit ) );

	const uint2 range = tile_range[tile_id];
	const unsigned start = range.x, end = range.y;
	if( start >= end )
	{
		if( PDL )
			pdl_grid_wait(); // an empty tile's forward warp still writes its pixels and loss partial
		return;
	}
	if( PDL )
	{
		// tile_range and the lists were complete before k_forward started; the tile's pixels and coverage words are
		// complete once its forward warp has published the flag
		if( !pdl_wait_tile( tile_done + tile_id, lane ) )
		{
			if( lane == 0 )
				atomicExch( const_cast<int*>( &st->peer_fault ), 2 );
			pdl_grid_wait();
			return;
		}
	}

	// per-pixel constants dL/dC = final - reference (main.cpp:613-616) and the running state: throughput T of the
	// prefix (image1.w, main.cpp:549) and Q = sum_c dL/dC_c * (final_c - prefix_c); prefix colour starts at 0.
	float T[8], Q[8], dLr[8], dLg[8], dLb[8];
	unsigned live[8];
#pragma unroll
	for( int j = 0; j < 8; j++ )
	{
		const int x = px_x + 8 * ( j & 1 ), y = px_y + 4 * ( j >> 1 );
		T[j] = 0.0f;
		Q[j] = dLr[j] = dLg[j] = dLb[j] = 0.0f;
		if( x < g.W && y < g.H )
		{
			const size_t pix = ( (size_t)img * g.H + y ) * g.W + x;
			const float4 fin = PDL ? __ldcg( image + pix ) : image[pix]; // PDL: written by a grid that is still running
			const float4 ref = load_target( g, target, pix );
			dLr[j] = xsub( fin.x, ref.x );
			dLg[j] = xsub( fin.y, ref.y );
			dLb[j] = xsub( fin.z, ref.z );
			Q[j] = dLr[j] * fin.x + dLg[j] * fin.y + dLb[j] * fin.z;
			T[j] = 1.0f;
		}
		live[j] = __ballot_sync( 0xffffffffu, !( T[j] < min_T ) );
	}

	unsigned nlive = 0;

	// lane k gathers record k of the batch: 4 quads (64 bytes); idx_cur stays in a register for the flush
	unsigned idx_cur = 0;
	{
		const bool v = start + lane < end;
		idx_cur = v ? vals[start + lane] : 0u;
		gather_record<4>( sm, 0, lane, rec, idx_cur, v );
		gather_cover( sm, 0, lane, cover, start, v );
		cp_async_commit();
	}
	unsigned idx_nxt = ( start + GS2D_BATCH + lane < end ) ? vals[start + GS2D_BATCH + lane] : 0u;
	// The entries of a round that blended into a pixel in the forward pass (k_forward left one mask per round): the
	// two passes evaluate the same alpha and transmittance chain bit for bit, so these are exactly the entries that
	// would pass a test against this pass's live masks - 20 instructions per popped entry that are not spent here.
	// The mask of the next round is requested a round ahead (through L2: the forward grid may still be running).
	unsigned rvi = ( start >> 5 ) + tile_id; // round_vis_index( start, tile_id, base ), advanced round by round
	unsigned ent_nxt = __ldcg( round_vis + rvi );

	int stage = 0;
	for( unsigned base = start; base < end; base += GS2D_BATCH, stage ^= 1 )
	{
		const int nb = min( (unsigned)GS2D_BATCH, end - base );
		const unsigned nbase = base + GS2D_BATCH;
		const unsigned idx_this = idx_cur;
		unsigned ent = ent_nxt;
		rvi++;
		if( nbase < end )
			ent_nxt = __ldcg( round_vis + rvi ); // (a round the forward pass never reached: unwritten, and never used here either)
		if( nbase < end )
		{
			gather_record<4>( sm, stage ^ 1, lane, rec, idx_nxt, nbase + lane < end );
			gather_cover( sm, stage ^ 1, lane, cover, nbase, nbase + lane < end );
			cp_async_commit();
			idx_cur = idx_nxt;
			idx_nxt = ( nbase + GS2D_BATCH + lane < end ) ? vals[nbase + GS2D_BATCH + lane] : 0u;
			cp_async_wait<1>();
		}
		else
			cp_async_wait<0>();
		__syncwarp();
		const unsigned visited = ent;
		unsigned grp_n = 0u, grp_k = 0u; // visited entries parked in stage4, their list positions (8 bits each)
#pragma unroll 1
		while( ent )
		{
			const int k = __ffs( ent ) - 1;
			ent &= ent - 1u;
			const uint4 wa = *reinterpret_cast<const uint4*>( &sm.words[stage][k][0] );
			const uint4 wb = *reinterpret_cast<const uint4*>( &sm.words[stage][k][4] );
			const unsigned wd[8] = { wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w };
			const SplatRec& s = sm.rec[stage][k];
			const float4 q0 = *reinterpret_cast<const float4*>( &s.px ); // px py B D
			const float4 q1 = *reinterpret_cast<const float4*>( &s.A );  // A cr cg cb
			const float op = s.op;
			const float vx0 = xsub( fpx0, q0.x ), vx1 = xsub( fpx1, q0.x );
			// per entry and block column: exact  a vx, b vx;  relaxed  s + A vx^2, B vx  (relaxed_alpha)
			const float avx0 = EXACT ? xmul( q1.x, vx0 ) : __fmaf_rn( __fmul_rn( q1.x, vx0 ), vx0, op );
			const float avx1 = EXACT ? xmul( q1.x, vx1 ) : __fmaf_rn( __fmul_rn( q1.x, vx1 ), vx1, op );
			const float bvx0 = xmul( q0.z, vx0 ), bvx1 = xmul( q0.z, vx1 );
			float gc0 = 0.f, gc1 = 0.f, gc2 = 0.f;          // sum dL/dC * alpha * T            (main.cpp:618-619)
			float A0[2] = { 0.f, 0.f }, A1[2] = { 0.f, 0.f }; // sum w, sum w*vy per block column   (w = g * alpha)
			float A2 = 0.f, vop = 0.f;                      // sum w*vy^2; sum g*G               (main.cpp:703-704)
#pragma unroll
			for( int j = 0; j < 8; j++ )
			{
				const unsigned wl = wd[j] & live[j];
				if( wl == 0u )
					continue;
				const bool cov = ( wl & lanebit ) != 0u;
				if( COUNT )
					nlive += cov ? 1u : 0u;
				const float vx = ( j & 1 ) ? vx1 : vx0, avx = ( j & 1 ) ? avx1 : avx0, bvx = ( j & 1 ) ? bvx1 : bvx0;
				const float vy = xsub( fpyr[j >> 1], q0.y );
				float alpha, G = 0.0f, wgt, Tn;
				if( EXACT )
				{
					// scalar multiplies here: the packed forms of k_forward need aligned register pairs, which at this
					// kernel's 80-register cap spill (measured: lane bit reloaded from local memory at every visit)
					const float rx = xadd( avx, xmul( q0.z, vy ) );
					const float ry = xadd( bvx, xmul( q0.w, vy ) );
					const float d2 = xadd( xmul( vx, rx ), xmul( vy, ry ) );
					const float t = __fmaf_rn( d2, -0.0625f, 1.0f );
					const float t2 = xmul( t, t ), t4 = xmul( t2, t2 ), t8 = xmul( t4, t4 );
					// alpha exactly as k_forward forms it (main.cpp:610-611): select AFTER the multiply, so that a NaN / inf
					// opacity cannot leak into uncovered or dead lanes (0 * NaN) and make the two passes' live sets differ
					alpha = select_alpha( wl & lanebit, t, xmul( t8, op ) );
					if( OPACITY )
						G = select_alpha( wl & lanebit, t, t8 );
					wgt = alpha * T[j];                      // dC_dc, main.cpp:618
					Tn = xmul( T[j], xsub( 1.0f, alpha ) );  // main.cpp:707
				}
				else
				{
					alpha = relaxed_alpha( wl & lanebit, vy, q0.w, bvx, avx );
					wgt = __fmul_rn( T[j], alpha );          // == k_forward's T * alpha
				}
				// every use of the old T in front of the new one (else the compiler keeps both and moves one at the end)
				const float D = T[j] * fmaf( dLb[j], q1.w, fmaf( dLg[j], q1.z, dLr[j] * q1.y ) );
				if( !EXACT )
					Tn = __fsub_rn( T[j], wgt );             // == k_forward's transmittance chain
				gc0 = fmaf( dLr[j], wgt, gc0 );
				gc1 = fmaf( dLg[j], wgt, gc1 );
				gc2 = fmaf( dLb[j], wgt, gc2 );
				const float oma = xsub( 1.0f, alpha );
				const bool dead = Tn < min_T;
				float Qn = fmaf( -alpha, D, Q[j] );
				Qn = dead ? 0.0f : Qn;                       // nothing behind a dead pixel: the reference's S is exactly 0
				// 1 / ( 1 - alpha + 1e-15 ), main.cpp:628: one MUFU.RCP (the argument is >= 1e-15, never subnormal, so
				// the range fix-up __fdividef wraps around the instruction is not needed)
				float inv;
				asm( "rcp.approx.ftz.f32 %0, %1;" : "=f"( inv ) : "f"( oma + 1.0e-15f ) );
				const float gsum = fmaf( -inv, Qn, D );      // dL_dalpha_rgb, main.cpp:629-630
				const float ga = gsum * alpha;
				const float gav = ga * vy;
				A0[j & 1] += ga;
				A1[j & 1] += gav;
				A2 = fmaf( gav, vy, A2 );
				if( OPACITY && EXACT )
					vop = fmaf( gsum, G, vop );
				T[j] = Tn;
				Q[j] = Qn;
				live[j] = __ballot_sync( 0xffffffffu, !dead );
			}
			if( OPACITY && !EXACT )
				vop = A0[0] + A0[1]; // sum g*alpha = opacity * sum g*G: divided by the opacity in the flush
			// per-lane moments of w: vx is constant per lane and block column
			float v[8];
			v[0] = gc0;
			v[1] = gc1;
			v[2] = gc2;
			const float w0 = vx0 * A0[0], w1 = vx1 * A0[1];
			v[3] = w0 + w1;                                  // sum w*vx
			v[4] = A1[0] + A1[1];                            // sum w*vy
			v[5] = fmaf( vx0, w0, vx1 * w1 );                // sum w*vx^2
			v[6] = fmaf( vx0, A1[0], vx1 * A1[1] );          // sum w*vx*vy
			v[7] = A2;                                       // sum w*vy^2
			{
				float4* dst = reinterpret_cast<float4*>( &stage4[grp_n][lane * 8] );
				dst[0] = make_float4( v[0], v[1], v[2], v[3] );
				dst[1] = make_float4( v[4], v[5], v[6], v[7] );
			}
			grp_k |= (unsigned)k << ( 8u * grp_n );
			if( OPACITY )
			{
#pragma unroll
				for( int o = 16; o > 0; o >>= 1 )
					vop += __shfl_xor_sync( 0xffffffffu, vop, o );
				if( lane == 0 )
					vop_tot[k] = vop;
			}
			if( ++grp_n == (unsigned)STAGE_SLOTS )
			{
				reduce_group( stage4, &sm.words[stage][0][0], grp_k, (unsigned)STAGE_SLOTS, lane );
				grp_n = 0u;
				grp_k = 0u;
			}
		}
		if( grp_n )
			reduce_group( stage4, &sm.words[stage][0][0], grp_k, grp_n, lane );
		__syncwarp();

		// flush: lane k owns entry k. d{pos,sx,sy,rot} from the moments (main.cpp:635-685), one atomic per parameter.
		if( ( visited >> lane ) & 1u )
		{
			const float4 m0 = *reinterpret_cast<const float4*>( &sm.words[stage][lane][0] ); // dcr dcg dcb Mx
			const float4 m1 = *reinterpret_cast<const float4*>( &sm.words[stage][lane][4] ); // My Mxx Mxy Myy
			const SplatRec& s = sm.rec[stage][lane];
			// inv_cov: as stored (exact) or recovered from the pre-scaled record (relaxed: A = -s a/16, B = -s b/8,
			// D = -s d/16, s = opacity^(1/8) > 0 for every splat that contributed anything)
			float a = s.A, b = s.B, d = s.D, inv_op = 0.0f;
			if( !EXACT )
			{
				const float is = s.op > 0.0f ? __frcp_rn( s.op ) : 0.0f;
				a *= -16.0f * is;
				b *= -8.0f * is;
				d *= -16.0f * is;
				const float is2 = is * is, is4 = is2 * is2;
				inv_op = is4 * is4; // 1 / opacity
			}
			const float c = s.cosT, sn = s.sinT;
			const float c2 = c * c, s2 = sn * sn, sc = sn * c;
			const float Mx = m0.w, My = m1.x, Mxx = m1.y, Mxy = m1.z, Myy = m1.w;
			float* gp = grads + idx_this;
			const size_t np = (size_t)g.n_pad;
			atomicAdd( gp + PL_CR * np, m0.x );
			atomicAdd( gp + PL_CG * np, m0.y );
			atomicAdd( gp + PL_CB * np, m0.z );
			// 0.5*alpha*(2a vx + (b+c) vy) == alpha*(a vx + b vy), main.cpp:639-640, :654-655
			atomicAdd( gp + PL_PX * np, fmaf( a, Mx, b * My ) );
			atomicAdd( gp + PL_PY * np, fmaf( b, Mx, d * My ) );
			// main.cpp:657-662, :677-678: alpha/s^3 * ( R^T v )_i^2 ;  :680-685: the rotation term
			const float tsc = 2.0f * sc * Mxy;
			atomicAdd( gp + PL_SX * np, s.isx3 * ( fmaf( c2, Mxx, fmaf( s2, Myy, tsc ) ) ) );
			atomicAdd( gp + PL_SY * np, s.isy3 * ( fmaf( s2, Mxx, fmaf( c2, Myy, -tsc ) ) ) );
			atomicAdd( gp + PL_ROT * np, s.rotc * ( fmaf( c2 - s2, Mxy, -sc * ( Mxx - Myy ) ) ) );
			if( OPACITY )
				atomicAdd( gp + PL_OP * np, EXACT ? vop_tot[lane] : vop_tot[lane] * inv_op );
		}

		unsigned anylive = 0;
#pragma unroll
		for( int j = 0; j < 8; j++ )
			anylive |= live[j];
		if( anylive == 0u )
			break;
		__syncwarp(); // part / words / rec[stage] are rewritten by the next rounds
	}
	cp_async_wait<0>();
	if( COUNT )
	{
#pragma unroll
		for( int o = 16; o > 0; o >>= 1 )
			nlive += __shfl_xor_sync( 0xffffffffu, nlive, o );
		if( lane == 0 )
			atomicAdd( &live_slots[tile_id & ( GS2D_LIVE_SLOTS - 1 )], (unsigned long long)nlive );
	}
	if( PDL )
		pdl_grid_wait();
}

template <int MINB, bool OPACITY, bool COUNT, bool PDL, bool EXACT>
static void launch_backward_as( dim3 grid, const Geometry& g, const Buffers& b, int sorted_buf,
                                unsigned long long* slots, cudaStream_t s )
{
	auto kernel = k_backward<MINB, OPACITY, COUNT, PDL, EXACT>;
	cudaLaunchAttribute attr[1];
	attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
	attr[0].val.programmaticStreamSerializationAllowed = 1;
	cudaLaunchConfig_t cfg = {};
	cfg.gridDim = grid;
	cfg.blockDim = dim3( 32 );
	cfg.dynamicSmemBytes = 0;
	cfg.stream = s;
	cfg.attrs = attr;
	cfg.numAttrs = PDL ? 1 : 0;
	cudaLaunchKernelEx( &cfg, kernel, g, (const SplatRec*)b.rec, (const unsigned*)b.vals[sorted_buf],
	                    (const uint2*)b.tile_range,
	                    g.target_u8 ? reinterpret_cast<const float4*>( b.target8 ) : (const float4*)b.target, (const float4*)b.image,
	                    (const unsigned*)b.cover, b.grads, slots, (const DevState*)b.state,
	                    (const unsigned*)( PDL ? b.tile_done : nullptr ),
	                    (const unsigned*)( b.cover + 8 * (size_t)b.capacity ) );
}
template <int MINB, bool PDL, bool COUNT>
static void launch_backward_minb( dim3 grid, const Geometry& g, const Buffers& b, int sorted_buf, bool with_opacity,
                                  bool exact, unsigned long long* slots, cudaStream_t s )
{
	if( with_opacity )
	{
		if( exact )
			launch_backward_as<MINB, true, COUNT, PDL, true>( grid, g, b, sorted_buf, slots, s );
		else
			launch_backward_as<MINB, true, COUNT, PDL, false>( grid, g, b, sorted_buf, slots, s );
	}
	else
	{
		if( exact )
			launch_backward_as<MINB, false, COUNT, PDL, true>( grid, g, b, sorted_buf, slots, s );
		else
			launch_backward_as<MINB, false, COUNT, PDL, false>( grid, g, b, sorted_buf, slots, s );
	}
}

void prepare_backward_pdl()
{
#define CARVE( OP, EXACT )                                                                                            \
	cudaFuncSetAttribute( k_backward<GS2D_BWD_MINB_DEFAULT, OP, false, true, EXACT>,                                 \
	                      cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared )
	CARVE( false, false );
	CARVE( true, false );
	CARVE( false, true );
	CARVE( true, true );
#undef CARVE
}

// PDL twins exist for the default register caps
bool gs2d_pdl_usable()
{
	return gs2d_tune( "GS2D_PDL", 1 ) != 0 &&
	       gs2d_tune( "GS2D_FWD_MINB", GS2D_FWD_MINB_DEFAULT ) == GS2D_FWD_MINB_DEFAULT &&
	       gs2d_tune( "GS2D_BWD_MINB", GS2D_BWD_MINB_DEFAULT ) == GS2D_BWD_MINB_DEFAULT;
}

int launch_backward( const Geometry& g, const Buffers& b, int sorted_buf, bool with_opacity, bool count_live,
                     cudaStream_t s, bool pdl )
{
	const int rows = g.band_ty1 - g.band_ty0;
	if( rows <= 0 || g.tiles_x <= 0 )
		return 0;
	dim3 grid( g.tiles_x, rows, g.n_images );
	// the backward's own live-pair count goes to the third slot group (stats[6]); it must equal the forward's
	unsigned long long* slots = b.live_slots + 2 * GS2D_LIVE_SLOTS;
	if( count_live )
		cudaMemsetAsync( slots, 0, sizeof( unsigned long long ) * GS2D_LIVE_SLOTS, s );
	const bool exact = gs2d_exact_mode();
	if( pdl ) // the counting variant is only launched stand-alone (gs2d_backward)
	{
		launch_backward_minb<GS2D_BWD_MINB_DEFAULT, true, false>( grid, g, b, sorted_buf, with_opacity, exact, slots, s );
		return 1;
	}
	const int minb = gs2d_tune( "GS2D_BWD_MINB", GS2D_BWD_MINB_DEFAULT );
	if( count_live )
	{
		if( minb == 20 )
			launch_backward_minb<20, false, true>( grid, g, b, sorted_buf, with_opacity, exact, slots, s );
		else
			launch_backward_minb<24, false, true>( grid, g, b, sorted_buf, with_opacity, exact, slots, s );
	}
	else
	{
		if( minb == 20 )
			launch_backward_minb<20, false, false>( grid, g, b, sorted_buf, with_opacity, exact, slots, s );
		else
			launch_backward_minb<24, false, false>( grid, g, b, sorted_buf, with_opacity, exact, slots, s );
	}
	return 1;
}
