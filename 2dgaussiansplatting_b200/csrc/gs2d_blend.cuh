// gs2d_blend.cuh — pieces shared by the forward and backward blend kernels (gs2d_forward.cu, gs2d_backward.cu).
//
// Work decomposition (both kernels): ONE WARP per 16x16 tile, 8 pixels per lane, no block barrier anywhere.
// Lane (lx = lane & 7, ly = lane >> 3) owns, for slot j = 0..7 (bx = j & 1, by = j >> 1), the pixel
//     x = tile_x0 + 8*bx + lx,   y = tile_y0 + 4*by + ly,
// so the 32 lanes of slot j are one compact 8x4 block of the tile and the coverage of that block by a list entry is
// ONE 32-bit word (bit = lane). Compact blocks instead of 16x2 strips: liveness and coverage are spatially coherent,
// so fewer blocks are visited for the same live pairs (238 vs 276 per tile at the 4096^2 / 1 M workload) and more
// lanes are useful per visit.
//
// Per round 32 list entries are staged: lane k gathers record k with 16-byte cp.async (double-buffered; the indices
// of the round after next are already in flight), then the exact scan-line footprint (main.cpp:489-509) of every
// (entry, tile row) is solved once and scattered as bytes into the per-entry block words.
#pragma once
#include "gs2d_internal.cuh"
#include "gs2d_bins.cuh" // round_vis_index

// ( covered != 0 && !( t < exp_approx's cut-off ) ) ? value : 0 in three instructions (LOP3 with predicate output,
// FSETP.GEU.AND on that predicate, FSEL): main.cpp:60 (`t < 0.000018...` -> 0) and the footprint / liveness test.
__device__ __forceinline__ float select_alpha( unsigned covered, float t, float value )
{
	float r;
	asm( "{\n\t.reg .pred pc, pt;\n\t"
	     "setp.ne.u32 pc, %1, 0;\n\t"
	     "setp.geu.and.f32 pt, %2, 0f379837F1, pc;\n\t" // !( t < 1.8145861e-5f ); NaN -> true like the reference's `<`
	     "selp.f32 %0, %3, 0f00000000, pt;\n\t}"
	     : "=f"( r )
	     : "r"( covered ), "f"( t ), "f"( value ) );
	return r;
}

// ---- relaxed arithmetic (GS2D_EXACT=0, the default) ------------------------------------------------------------------
// Everything below is written with explicit _rn intrinsics, not operators: the forward and the backward kernel must
// evaluate the SAME alpha and transmittance chain bit for bit (the backward re-walks the forward's live set), which
// must not depend on what the compiler chooses to contract in either of them.
// fma with the result clamped to [0, 1] (FFMA.SAT; NaN -> 0): s * (1 - d2/16) is at most s <= 1, and everything
// below exp_approx's cut-off (main.cpp:60) that matters is negative.
__device__ __forceinline__ float fma_sat( float a, float b, float c )
{
	float r;
	asm( "fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"( r ) : "f"( a ), "f"( b ), "f"( c ) );
	return r;
}
// covered != 0 ? value : 0 (LOP3 with predicate output + FSEL)
__device__ __forceinline__ float select_covered( unsigned covered, float value )
{
	float r;
	asm( "{\n\t.reg .pred pc;\n\t"
	     "setp.ne.u32 pc, %1, 0;\n\t"
	     "selp.f32 %0, %2, 0f00000000, pc;\n\t}"
	     : "=f"( r )
	     : "r"( covered ), "f"( value ) );
	return r;
}
// alpha of one pixel in relaxed arithmetic: t = clamp( axx + vy * ( bvx + D * vy ), 0, 1 ) with the per-entry column
// terms bvx = B vx, axx = s + A vx^2 (SplatRec: A, B, D pre-scaled by -s/16, -s/8, -s/16, s = opacity^(1/8));
// alpha = t^8 = opacity * (1 - d2/16)^8 = main.cpp:527 up to rounding. 0 for uncovered or dead lanes.
__device__ __forceinline__ float relaxed_alpha( unsigned covered, float vy, float D, float bvx, float axx )
{
	const float t = fma_sat( vy, __fmaf_rn( D, vy, bvx ), axx );
	const float t2 = __fmul_rn( t, t ), t4 = __fmul_rn( t2, t2 );
	return select_covered( covered, __fmul_rn( t4, t4 ) );
}

// statistics of the counting variants: 1 if this lane's pixel is covered and live (GS2D_VISIT_HIST: also the
// developer histograms of useful lanes per block visit and of live pixels left in the tile)
__device__ __forceinline__ unsigned count_visit( unsigned wl, unsigned lanebit, const unsigned live[8], int visit_hist,
                                                 int lane, unsigned long long ( *hist )[33] )
{
	if( visit_hist && lane == 0 )
	{
		atomicAdd( &hist[0][__popc( wl )], 1ull );
		unsigned tl = 0;
#pragma unroll
		for( int jj = 0; jj < 8; jj++ )
			tl += __popc( live[jj] );
		atomicAdd( &hist[1][min( 32u, ( tl + 7u ) / 8u )], 1ull );
	}
	return ( wl & lanebit ) != 0u ? 1u : 0u;
}

// Two floats in an aligned register pair, and the packed multiply mul.rn.f32x2 (SASS FMUL2): each half is the IEEE
// product, exactly __fmul_rn, but the pair costs ONE issue slot (the blend kernels are issue-bound).
// Only multiplies are packed: ptxas contracts a packed multiply feeding a packed add into FFMA2 (one rounding, not
// the reference's two), whereas a scalar add.rn.f32 of one half is never contracted. A scalar operand is written
// pack2( s, s ); ptxas encodes it as a broadcast (R.F32), no move.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2( float lo, float hi )
{
	f32x2 r;
	asm( "mov.b64 %0, {%1, %2};" : "=l"( r ) : "f"( lo ), "f"( hi ) );
	return r;
}
__device__ __forceinline__ void unpack2( f32x2 p, float& lo, float& hi )
{
	asm( "mov.b64 {%0, %1}, %2;" : "=f"( lo ), "=f"( hi ) : "l"( p ) );
}
__device__ __forceinline__ f32x2 xmul2( f32x2 a, f32x2 b )
{
	f32x2 r;
	asm( "mul.rn.f32x2 %0, %1, %2;" : "=l"( r ) : "l"( a ), "l"( b ) );
	return r;
}

// ---- forward -> backward overlap through programmatic dependent launch (default; GS2D_PDL=0 turns it off) -----------------------
// The blend grids end in a tail: one warp per tile, so the last resident tiles keep their SMs mostly idle for one
// tile's run time (~50 us; twice per iteration, a quarter of a rank's blend time at N = 8). The backward pass of a
// tile needs the forward pass of THAT tile only (its final pixels, coverage words), so in this mode
//   * every forward warp issues griddepcontrol.launch_dependents first thing: the backward grid becomes eligible
//     once the LAST forward warp has started, i.e. all forward warps already hold their SM slots (no deadlock:
//     a waiting backward warp can never keep a forward warp from being scheduled);
//   * the backward grid is launched with the programmatic-stream-serialization attribute and does not START with
//     griddepcontrol.wait (that would wait for the whole forward grid); instead the warp of tile t spins on
//     tile_done[t], which the forward warp of t stores with release semantics after its last global write. Every
//     backward warp ENDS with griddepcontrol.wait (pdl_grid_wait) to keep the stream's transitive ordering.
// Backward warps are dispatched in tile order, so the tiles they ask for finished long ago; the wait is a single
// acquire load except for the last wave. What the backward reads of the forward's output goes through L2
// (cp.async.cg, ld.global.cg): L1 is not coherent with stores of a concurrently running grid.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile( "griddepcontrol.launch_dependents;" ::: "memory" ); }
// Executed by every backward warp as its LAST instruction (also on the early-exit paths: empty tile, capacity
// overflow): the backward grid then completes only after the whole forward grid has completed and flushed, so
// whatever follows the backward grid on the stream (k_finalize_loss reads loss_partial / live_slots of ALL forward
// warps, empty tiles included) is transitively ordered behind the forward grid, as PTX requires of a dependent of a
// grid that uses launch_dependents. Free in practice: the forward grid is in its tail when the first backward warp
// starts, and a backward tile runs longer than that tail.
__device__ __forceinline__ void pdl_grid_wait() { asm volatile( "griddepcontrol.wait;" ::: "memory" ); }
__device__ __forceinline__ void pdl_publish_tile( unsigned* flag, int lane )
{
	__threadfence(); // this lane's pixel / coverage / loss stores are performed at gpu scope
	__syncwarp();
	if( lane == 0 )
		asm volatile( "st.release.gpu.global.u32 [%0], %1;" ::"l"( flag ), "r"( 1u ) : "memory" );
}
// false: the flag did not arrive within ~2 s (cannot happen with the launch order above; never hang the device)
__device__ __forceinline__ bool pdl_wait_tile( const unsigned* flag, int lane )
{
	unsigned ok = 1u;
	if( lane == 0 )
	{
		unsigned v = 0u;
#pragma unroll 1
		for( unsigned spin = 0; spin < ( 1u << 22 ); spin++ )
		{
			asm volatile( "ld.acquire.gpu.global.u32 %0, [%1];" : "=r"( v ) : "l"( flag ) : "memory" );
			if( v != 0u )
				break;
			__nanosleep( 500 );
		}
		ok = v;
	}
	ok = __shfl_sync( 0xffffffffu, ok, 0 ); // also orders the other lanes' loads after lane 0's acquire
	return ok != 0u;
}

// WSTAGES = 1: the coverage words are solved in place (forward); 2: they are gathered like the records (backward)
template <int WSTAGES>
struct TileSmem
{
	SplatRec rec[2][GS2D_BATCH];              // 4 KB, double-buffered records of the staged list entries
	unsigned words[WSTAGES][GS2D_BATCH][8];   // 1 KB each, exact coverage: word j of entry k = 8x4 block j (bit = lane)
};

// lane k gathers QUADS float4 of record idx into sm.rec[stage][k]
template <int QUADS, int WSTAGES>
__device__ __forceinline__ void gather_record( TileSmem<WSTAGES>& sm, int stage, int lane, const SplatRec* __restrict__ rec,
                                               unsigned idx, bool valid )
{
	if( valid )
	{
		const float4* src = reinterpret_cast<const float4*>( rec + idx );
		float4* dst = reinterpret_cast<float4*>( &sm.rec[stage][lane] );
#pragma unroll
		for( int q = 0; q < QUADS; q++ )
			cp_async16( dst + q, src + q );
	}
	else
		sm.rec[stage][lane].idx = 0u; // a slot beyond the batch: build_words looks it up like any other (splat 0 exists)
}

// Exact coverage words of the nb staged entries of `stage`: 16 rows x nb entries.
// First lane k turns the row range of entry k (SplatRec::flags) into the 16-bit mask of the tile's rows that lie
// inside it (bit 16: the splat is in the row table) and leaves it in the staged record. Then lane (h = lane >> 4,
// r = lane & 15) handles row r of entries 2i + h: it fetches the packed span of (splat, image row) from the row table
// (k_row_table solved it once for all the tiles the row crosses; the 16 rows of an entry are 64 contiguous bytes),
// clips it to the tile's 16 columns and scatters the mask as two bytes into the block words of the entry. Four
// entries per lane are in flight at a time.
// Splats that are not in the row table (taller than GS2D_ROWTAB_STRIDE rows, or GS2D_ROWTAB=0) are solved here, in a
// second loop that a warp enters only if its batch holds one: the reference's quadratic per (entry, tile row) from
// the unscaled RowRec (row_mask, gs2d_footprint.cuh) - same arithmetic as k_row_table, same words.
#define ROWMASK_IN_TABLE 0x10000u
__device__ __forceinline__ void build_words( TileSmem<1>& sm, int stage, int nb, int lane, int tile_x0, int tile_y,
                                             const unsigned* __restrict__ rowtab, const RowRec* __restrict__ rowrec,
                                             int W, int H )
{
	const int h = lane >> 4, r = lane & 15;
	const int tile_y0 = tile_y * GS2D_TILE;
	bool mine_big = false;
	{
		// a slot beyond the batch gets "in the table, no row": looked up like any other (gather_record gave it index 0),
		// masked to nothing
		unsigned rowmask = ROWMASK_IN_TABLE;
		if( lane < nb )
		{
			const unsigned f = sm.rec[stage][lane].flags;
			const int ry0 = (int)( ( f >> SPLAT_ROW0_SHIFT ) & ( ( 1u << SPLAT_ROW0_BITS ) - 1u ) );
			const int nrows = (int)( f >> SPLAT_NROWS_SHIFT ) + 1;
			const int lo = min( max( ry0 - tile_y0, 0 ), GS2D_TILE ), hi = min( max( ry0 + nrows - tile_y0, 0 ), GS2D_TILE );
			mine_big = ( f & SPLAT_IN_TABLE ) == 0u;
			rowmask = mine_big ? 0u : ( ( ( 0xffffffffu << lo ) & ~( 0xffffffffu << hi ) ) | ROWMASK_IN_TABLE );
		}
		sm.rec[stage][lane].flags = rowmask;
	}
	__syncwarp();
	const unsigned ring = (unsigned)( ( tile_y0 + r ) & ( GS2D_ROWTAB_STRIDE - 1 ) );
	const unsigned rbit = 1u << r;
	// byte (r & 3) of word 2*(r >> 2) of entry h; the right half of the row is 4 bytes on, the next entry pair 64
	unsigned char* wb0 = reinterpret_cast<unsigned char*>( &sm.words[0][h][0] ) + ( r >> 2 ) * 8 + ( r & 3 );
	{
		// every staged slot is processed and every entry is looked up, in the table or not (the table is allocated for
		// every splat and was cleared once): no branch in the loop; what is not in the table, or not a row of the
		// splat, is masked out (entries that are not in the table are then solved by the second loop)
		unsigned char* wb = wb0;
		const SplatRec* s = &sm.rec[stage][h];
		const unsigned* tab = rowtab + ring;
#pragma unroll 1
		for( int k0 = h; k0 < nb; k0 += 8, s += 8, wb += 256 )
		{
			unsigned span[4];
			bool ok[4];
#pragma unroll
			for( int u = 0; u < 4; u++ )
			{
				const uint2 f = *reinterpret_cast<const uint2*>( &s[2 * u].idx ); // idx, row mask
				span[u] = __ldg( tab + (size_t)f.x * GS2D_ROWTAB_STRIDE );
				ok[u] = ( f.y & rbit ) != 0u;
			}
#pragma unroll
			for( int u = 0; u < 4; u++ )
			{
				const unsigned m = ok[u] ? mask_of_span( span[u], tile_x0 ) : 0u;
				wb[64 * u] = (unsigned char)m;
				wb[64 * u + 4] = (unsigned char)( m >> 8 );
			}
		}
	}
	if( __any_sync( 0xffffffffu, mine_big ) )
	{
		const int y = tile_y0 + r;
		unsigned char* wb = wb0;
		const SplatRec* s = &sm.rec[stage][h];
#pragma unroll 1
		for( int k = h; k < nb; k += 2, s += 2, wb += 64 )
		{
			if( s->flags & ROWMASK_IN_TABLE )
				continue;
			const RowRec* rr = rowrec + s->idx;
			const float4 q0 = __ldg( reinterpret_cast<const float4*>( rr ) );     // px py a b
			const float4 q1 = __ldg( reinterpret_cast<const float4*>( rr ) + 1 ); // d begY endY rows
			const unsigned m = row_mask( q0.x, q0.y, q0.z, q0.w, q1.x, __float_as_int( q1.y ), __float_as_int( q1.z ), y,
			                             tile_x0, W, H );
			wb[0] = (unsigned char)( m & 0xffu );
			wb[4] = (unsigned char)( m >> 8 );
		}
	}
}

// Mask of the staged entries that cover at least one pixel of the tile (the others are never looked at again);
// lane k owns entry k. wa/wb return the lane's own 8 coverage words.
template <int WSTAGES>
__device__ __forceinline__ unsigned touching_entries( const TileSmem<WSTAGES>& sm, int wstage, int nb, int lane,
                                                      uint4& wa, uint4& wb )
{
	wa = make_uint4( 0u, 0u, 0u, 0u );
	wb = wa;
	if( lane < nb )
	{
		wa = *reinterpret_cast<const uint4*>( &sm.words[wstage][lane][0] );
		wb = *reinterpret_cast<const uint4*>( &sm.words[wstage][lane][4] );
	}
	const unsigned any = wa.x | wa.y | wa.z | wa.w | wb.x | wb.y | wb.z | wb.w;
	return __ballot_sync( 0xffffffffu, any != 0u );
}
