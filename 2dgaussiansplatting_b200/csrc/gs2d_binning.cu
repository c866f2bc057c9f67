// gs2d_binning.cu — per-Gaussian preprocessing and tile binning.
//
// Replaces, for a pixel-parallel GPU pass, what the reference does implicitly by iterating splats in index order
// and scan-converting each one (main.cpp:419-538): here every 16x16 tile gets the list of splats whose footprint
// may touch it, ordered by splat index (== the reference's blend order).
//
// The (tile | order) keys are sorted MSD-first without a sorting pass over the keys:
//   bucket = (tile, index slice): the splat indices of an image are cut into n_slices equal slices (<= 65535 splats
//   each), so a bucket holds the few splats of one slice that touch one tile (3.3 on average at 4096^2 / 1 M).
//   k_preprocess   : covariance, inverse, row range (cov_of main.cpp:206-221, :432-436, :489-491) -> 64-byte record;
//                    then, warp-cooperatively over (splat, tile row) work items:
//                    the tiles the 3-sigma ellipse can touch (per tile row: the x-extent of the ellipse inside the
//                    band of that row's pixel centres, see tile_span) are emitted as unordered (bucket, splat, rank)
//                    triples packed into one 64-bit word, rank = arrival order in the bucket. The bucket counters are 16 bits wide and the
//                    counters of 4 consecutive tiles share one 64-bit word, so ONE atomicAdd serves up to 4 tiles
//                    of a tile row (the L2 atomic rate, not HBM, bounds this kernel).
//   k_scan_*       : exclusive scan of the bucket counts in (tile, slice) order = bucket offsets; capacity check
//   k_place        : one thread per triple: list[offset[bucket] + rank] = splat (no atomics)
//   k_rank         : one thread per bucket orders its few entries by splat index (rank by counting) and the first /
//                    last bucket of a tile write the tile's [start,end)
// Splat indices are unique inside a bucket, so the sorted lists - and everything downstream - are deterministic
// although the arrival order inside a bucket is not.
#include "gs2d_internal.cuh"
#include "gs2d_bins.cuh"

#include <algorithm>

#define PRE_THREADS 256

// ------------------------------------------------------------------------------------------------ preprocess ---
// Per-splat part: record + span constants + conservative tile rectangle (tile units; empty if rtx0 > rtx1).
struct SplatBin
{
	SpanConsts k;
	int rtx0, rtx1, ty0, ty1;
	int tile_base; // first tile id of the splat's image
	int slice;     // index slice of the splat
	int local;     // index of the splat inside its slice
};

#define ITEMS_PER_CHUNK 256 // (splat, tile row) work items a warp handles per reservation of the pair list

// One emitted (tile, splat) pair is ONE 64-bit word: bucket = tile * n_slices + slice in the low half, then the
// splat's index inside its slice (16 bits; slice_size <= 65535) and the arrival order inside the bucket (16 bits).
// The image is implied by the tile and the slice by the bucket, so the global splat index is rebuilt from them.
struct PairOut
{
	unsigned long long* triple;  // bucket | local << 32 | rank << 48
	unsigned long long* cnt64;   // [n_slices][nt_pad / 4] words of four 16-bit bucket counters (consecutive tiles)
	// slot binning (GS2D_SLOTS, the default): the first GS2D_BUCKET_SLOTS arrivals of a bucket store their 16-bit index
	// inside the slice straight into the bucket's own 16-byte slot row, [n_slices][nt_pad][GS2D_BUCKET_SLOTS]; only later
	// arrivals (rare: 3.3 entries per bucket on average) become triples, appended at *n_over, and go through k_place
	unsigned short* slots;
	unsigned long long* n_over;
	// group rows (GS2D_SLOTS=2, the default; gs2d_bins.cuh): ONE 32-bit atomic and ONE 32-bit store per (splat, group of
	// four consecutive tiles) instead of one 16-bit store per (splat, tile): the kernel is bound by the number of L2
	// requests its scattered stores and atomics make (14 M + 4.7 M at 4096^2 / 1 M; now 4.7 M + 4.7 M)
	unsigned* grp_cnt;           // [nt_pad / 4][n_slices] arrivals of a (group, slice), also those that overflowed
	unsigned* grp_rows;          // [nt_pad / 4][n_slices][GS2D_GROUP_SLOTS] keys (the memory of `slots`)
	unsigned long long* ovf64;   // [nt_pad / 4][n_slices] four 16-bit counters: overflow arrivals per tile of the group
	// A rank that owns a band of tile rows: the splats whose rows k_row_table has to solve (1 in 7 at 8 ranks) are
	// appended here, so that its warps run with full lanes (nullptr: every splat is looked at, the single-GPU case)
	unsigned* rt_list;
	unsigned long long* rt_count;
	// ... and the splats that passed the band test at all (k_band_filter): k_preprocess<.., LISTED> walks this list
	// with full warps instead of finding 1 splat in 7 (8 ranks) in every warp
	const unsigned* band_list;
	const unsigned long long* band_count;
};

// Emit the pairs of tiles id_lo..id_hi (one tile row of one splat) at pos.., one 64-bit atomic per 4-tile group.
template <int SLOTS>
__device__ __forceinline__ void emit_row( const Geometry& g, const PairOut& o, unsigned id_lo, unsigned id_hi,
                                          int slice, unsigned local, unsigned long long pos,
                                          unsigned long long capacity )
{
	unsigned long long* cnt = o.cnt64 + (size_t)slice * ( g.nt_pad >> 2 );
	unsigned short* slot_row = o.slots + (size_t)slice * g.nt_pad * GS2D_BUCKET_SLOTS;
	for( unsigned grp = id_lo >> 2; grp <= ( id_hi >> 2 ); grp++ )
	{
		const unsigned f0 = max( id_lo, grp * 4u ) - grp * 4u, f1 = min( id_hi, grp * 4u + 3u ) - grp * 4u;
		// ones in the 16-bit fields f0..f1
		const unsigned long long add = ( 0x0001000100010001ull >> ( 16u * ( 3u - ( f1 - f0 ) ) ) ) << ( 16u * f0 );
		const unsigned long long old = atomicAdd( cnt + grp, add );
		for( unsigned f = f0; f <= f1; f++, pos++ )
		{
			const unsigned tile = grp * 4u + f;
			const unsigned bucket = tile * (unsigned)g.n_slices + (unsigned)slice;
			const unsigned long long rank = ( old >> ( 16u * f ) ) & 0xffffull;
			const unsigned long long t = (unsigned long long)bucket | ( (unsigned long long)local << 32 ) | ( rank << 48 );
			if( !SLOTS )
				__stcs( o.triple + pos, t );
			else if( rank < (unsigned long long)GS2D_BUCKET_SLOTS )
				slot_row[(size_t)tile * GS2D_BUCKET_SLOTS + (unsigned)rank] = (unsigned short)local;
			else
			{
				const unsigned long long at = atomicAdd( o.n_over, 1ull );
				if( at < capacity ) // cannot fail: the chunk's pairs were reserved against the capacity
					__stcs( o.triple + at, t );
			}
		}
	}
}

// Group rows (GS2D_SLOTS=2): the arrival of (splat, tiles f0..f1 of group grp) that took position `old` of the row of
// its (group, slice). Beyond the row: per-tile overflow ranks from one 64-bit atomic, one triple per tile.
__device__ __forceinline__ void emit_group( const Geometry& g, const PairOut& o, unsigned grp, unsigned f0, unsigned f1,
                                            unsigned old, int slice, unsigned local, unsigned long long capacity,
                                            int probe = 0 )
{
	const size_t pair = (size_t)grp * (unsigned)g.n_slices + (unsigned)slice;
	if( old < (unsigned)GS2D_GROUP_SLOTS )
	{
		if( !( probe & 1 ) )
			o.grp_rows[pair * GS2D_GROUP_SLOTS + old] = group_key( local, f0, f1 - f0 );
		return;
	}
	const unsigned long long oldv = atomicAdd( o.ovf64 + pair, field_ones( f0, f1 ) );
	for( unsigned f = f0; f <= f1; f++ )
	{
		const unsigned bucket = ( grp * 4u + f ) * (unsigned)g.n_slices + (unsigned)slice;
		const unsigned long long rank = ( oldv >> ( 16u * f ) ) & 0xffffull;
		const unsigned long long at = atomicAdd( o.n_over, 1ull );
		if( at < capacity ) // cannot fail: the chunk's pairs were reserved against the capacity
			__stcs( o.triple + at, (unsigned long long)bucket | ( (unsigned long long)local << 32 ) | ( rank << 48 ) );
	}
}
// one tile row of one splat, tiles id_lo..id_hi. Two groups per step: both atomics are in flight before the first
// store waits for its position (the emission is bound by the latency of the returning atomics per lane).
// probe (developer timing knob GS2D_K1_PROBE, results WRONG when set): 1 = no key stores, 2 = no atomics
__device__ __forceinline__ void emit_row_groups( const Geometry& g, const PairOut& o, unsigned id_lo, unsigned id_hi,
                                                 int slice, unsigned local, unsigned long long capacity, int probe )
{
	const unsigned ns = (unsigned)g.n_slices, g1 = id_hi >> 2;
	for( unsigned grp = id_lo >> 2; grp <= g1; grp += 2u )
	{
		const bool two = grp < g1;
		unsigned* cnt = o.grp_cnt + (size_t)grp * ns + (unsigned)slice;
		const unsigned old_a = ( probe & 2 ) ? ( local & 7u ) : atomicAdd( cnt, 1u );
		unsigned old_b = ( probe & 2 ) ? ( local & 7u ) : 0u;
		if( two && !( probe & 2 ) )
			old_b = atomicAdd( cnt + ns, 1u );
		const unsigned fa0 = max( id_lo, grp * 4u ) - grp * 4u, fa1 = min( id_hi, grp * 4u + 3u ) - grp * 4u;
		emit_group( g, o, grp, fa0, fa1, old_a, slice, local, capacity, probe );
		if( two ) // the row continues into the next group: it starts at that group's first tile
			emit_group( g, o, grp + 1u, 0u, min( id_hi, grp * 4u + 7u ) - ( grp * 4u + 4u ), old_b, slice, local, capacity,
			            probe );
	}
}

// relaxed: write the record in the scaled form the relaxed-arithmetic blend kernels read (SplatRec, gs2d_internal.cuh)
// rowtab: 1 = splats of at most GS2D_ROWTAB_STRIDE rows get their rows solved by k_row_table (else every
// splat is left to the blend kernels' in-kernel solve: GS2D_ROWTAB=0, the A/B reference)
// LISTED: the splats are those of out.band_list (they passed k_band_filter's band test, which has also written the
// touch bitmap and the reach masks); else every splat is looked at and tested here.
template <int SLOTS, bool LISTED>
__global__ void __launch_bounds__( PRE_THREADS, 6 ) k_preprocess( Geometry g, const float* __restrict__ params,
                                                               SplatRec* __restrict__ rec, RowRec* __restrict__ rowrec,
                                                               PairOut out, unsigned long long capacity, int cull,
                                                               int relaxed, int rowtab, unsigned* __restrict__ touch,
                                                               Buffers::Reach reach, DevState* __restrict__ st )
{
	int gi = blockIdx.x * PRE_THREADS + threadIdx.x;
	if( LISTED )
	{
		const unsigned long long n_listed = *out.band_count;
		if( (unsigned long long)blockIdx.x * PRE_THREADS >= n_listed )
			return;
		gi = (unsigned long long)gi < n_listed ? (int)out.band_list[gi] : g.n_total;
	}
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	SplatBin sb;
	sb.rtx0 = 1;
	sb.rtx1 = 0;
	sb.ty0 = 1;
	sb.ty1 = 0;
	sb.tile_base = 0;
	sb.slice = 0;
	sb.local = 0;
	sb.k.begY = 1;
	sb.k.endY = 0;
	bool in_band = false; // passed the band test: the only splats that can receive a gradient on this rank
	bool want_rt = false; // has rows for k_row_table to solve
	unsigned reaches = 0; // sparse exchange: the ranks this splat can reach (computed for the quads this rank owns)
	if( gi < g.n_total )
	{
		const float px = params[PL_PX * g.n_pad + gi];
		const float py = params[PL_PY * g.n_pad + gi];
		const float sx = params[PL_SX * g.n_pad + gi];
		const float sy = params[PL_SY * g.n_pad + gi];
		const float rot = params[PL_ROT * g.n_pad + gi];
		if( !LISTED && reach.now && ( gi >> 2 ) >= reach.q0 && ( gi >> 2 ) < reach.q1 )
			reaches = reach_mask( py, sx, sy, reach.band_px, reach.n_ranks, reach.slack );

		// a rank that owns a band of tile rows skips the splats that cannot reach it (band_rejects, gs2d_footprint.cuh)
		// before the double sincos, the five IEEE divisions and the records, which only this rank's blend would read
		const bool off_band = LISTED ? false : band_rejects( py, sx, sy, g.band_ty0 * GS2D_TILE, g.band_ty1 * GS2D_TILE );
		in_band = !off_band;
		if( off_band )
			rowrec[gi].rows = ROWREC_NONE; // k_row_table looks at every splat
		else
		{
			// cosf/sinf of the reference are glibc's (< 1 ulp, almost always correctly rounded). Evaluate in double and
			// round once: agrees with it bit for bit except in rare double-rounding cases.
			double sd, cd;
			sincos( (double)rot, &sd, &cd );
			const float c = (float)cd, s = (float)sd;

			// cov_of, its inverse and the row range (main.cpp:206-221, :432-436, :489-491): gs2d_footprint.cuh
			const SplatGeom ge = splat_geometry( py, sx, sy, c, s );
			const float det = ge.det, a = ge.a, b = ge.b, d = ge.d;
			const int begY = ge.begY, endY = ge.endY;

			const float op = params[PL_OP * g.n_pad + gi];
			SplatRec r;
			r.px = px;
			r.py = py;
			if( relaxed )
			{
				// s = opacity^(1/8) (three correctly rounded square roots), at most 1: alpha = ( s * (1 - d2/16) )^8 with
				// the product clamped to [0, 1] by the blend's saturating FMA. The scalings by 1/8, 1/16 are exact.
				const float sop = fminf( xsqrt( xsqrt( xsqrt( op ) ) ), 1.0f );
				r.A = xmul( sop, xmul( a, -0.0625f ) );
				r.B = xmul( sop, xmul( b, -0.125f ) );
				r.D = xmul( sop, xmul( d, -0.0625f ) );
				r.op = sop;
			}
			else
			{
				r.A = a;
				r.B = b;
				r.D = d;
				r.op = op;
			}
			r.cr = params[PL_CR * g.n_pad + gi];
			r.cg = params[PL_CG * g.n_pad + gi];
			r.cb = params[PL_CB * g.n_pad + gi];
			r.idx = (unsigned)gi;
			// hoisted backward invariants, main.cpp:658, :661, :682 (same operand order)
			r.rotc = xdiv( xsub( xmul( sx, sx ), xmul( sy, sy ) ), xmul( xmul( xmul( sx, sx ), sy ), sy ) );
			r.cosT = c;
			r.sinT = s;
			r.isx3 = xdiv( 1.0f, xmul( xmul( sx, sx ), sx ) );
			r.isy3 = xdiv( 1.0f, xmul( xmul( sy, sy ), sy ) );
			// rows of the splat inside the image; the row table holds splats of at most GS2D_ROWTAB_STRIDE rows (whole
			// image: the flag must not depend on the band)
			const int ry0 = max( begY, 0 ), ry1 = min( endY, g.H - 1 );
			const bool in_table = rowtab && ry0 <= ry1 && ry1 - ry0 < GS2D_ROWTAB_STRIDE;
			r.flags = in_table ? ( SPLAT_IN_TABLE | ( (unsigned)ry0 << SPLAT_ROW0_SHIFT ) |
			                       ( (unsigned)( ry1 - ry0 ) << SPLAT_NROWS_SHIFT ) )
			                   : 0u;
			float4* dst = reinterpret_cast<float4*>( rec + gi );
			dst[0] = make_float4( r.px, r.py, r.B, r.D );
			dst[1] = make_float4( r.A, r.cr, r.cg, r.cb );
			dst[2] = make_float4( __uint_as_float( r.idx ), __uint_as_float( r.flags ), r.op, r.rotc );
			dst[3] = make_float4( r.cosT, r.sinT, r.isx3, r.isy3 );
			// what k_row_table solves: the rows inside the image AND this rank's band
			unsigned rr_rows = ROWREC_NONE;
			if( in_table )
			{
				const int b0 = max( ry0, g.band_ty0 * GS2D_TILE ), b1 = min( ry1, g.band_ty1 * GS2D_TILE - 1 );
				if( b0 <= b1 )
					rr_rows = (unsigned)b0 | ( (unsigned)( b1 - b0 ) << 20 );
			}
			want_rt = rr_rows != ROWREC_NONE;
			float4* rdst = reinterpret_cast<float4*>( rowrec + gi );
			rdst[0] = make_float4( px, py, a, b );
			rdst[1] = make_float4( d, __int_as_float( begY ), __int_as_float( endY ), __uint_as_float( rr_rows ) );

			// Conservative tile rectangle. Rows: exactly the reference's [begY, endY]. Columns: every chord of the 3-sigma
			// ellipse lies within pos.x +- 3*sqrt(cov[0][0]) = 3*sqrt(d*det); widened by 2 px + 1e-4 relative for the
			// truncation rule and the float error of the quadratic solve. Only a superset is needed here: the exact
			// per-pixel set is decided by row_mask() in the forward blend.
			const int H = g.H, W = g.W;
			const int y0 = max( begY, 0 ), y1 = min( endY, H - 1 );
			const float hx = half_width_x( ge );
			const float m = 2.0f + hx * 1.0e-4f;
			const float fx0 = floorf( px - hx - m ), fx1 = floorf( px + hx + m );
			const int x0 = ( fx0 == fx0 ) ? (int)fmaxf( fx0, 0.0f ) : 0;
			int x1 = ( fx1 == fx1 ) ? (int)fminf( fx1, (float)( W - 1 ) ) : W - 1;
			if( fx1 < 0.0f )
				x1 = -1;
			if( y0 <= y1 && x0 <= x1 )
			{
				sb.ty0 = max( y0 / GS2D_TILE, g.band_ty0 );
				sb.ty1 = min( y1 / GS2D_TILE, g.band_ty1 - 1 );
				sb.rtx0 = x0 / GS2D_TILE;
				sb.rtx1 = x1 / GS2D_TILE;
			}
			const int img = gi / g.n_splats;
			sb.tile_base = img * g.tiles_y * g.tiles_x;
			sb.slice = ( gi - img * g.n_splats ) / g.slice_size;
			sb.local = ( gi - img * g.n_splats ) - sb.slice * g.slice_size;
			sb.k.px = px;
			sb.k.py = py;
			sb.k.a = a;
			sb.k.b = b;
			sb.k.d = d;
			sb.k.inv_a = 1.0f / a;
			sb.k.vx_max = hx; // 3 * sqrt( cov[0][0] ) == 3 * sqrt( d / (a d - b^2) )
			sb.k.vy_at = -b * hx / d;
			sb.k.delta = span_delta( px, hx, cull & 1 );
			sb.k.begY = begY;
			sb.k.endY = endY;
		}
	}
	if( !LISTED && reach.now )
	{
		// OR over the four splats of the quad; the first lane of the quad stores the byte
		unsigned m = reaches;
		m |= __shfl_xor_sync( 0xffffffffu, m, 1 );
		m |= __shfl_xor_sync( 0xffffffffu, m, 2 );
		const int q = gi >> 2;
		if( ( lane & 3 ) == 0 && q >= reach.q0 && q < reach.q1 )
			reach.now[q] = (unsigned char)m;
	}
	if( out.rt_list )
	{
		const unsigned bal = __ballot_sync( 0xffffffffu, want_rt );
		if( bal )
		{
			unsigned long long at = 0;
			if( lane == 0 )
				at = atomicAdd( out.rt_count, (unsigned long long)__popc( bal ) );
			at = __shfl_sync( 0xffffffffu, at, 0 );
			if( want_rt )
				out.rt_list[at + __popc( bal & ( ( 1u << lane ) - 1u ) )] = (unsigned)gi;
		}
	}
	if( !LISTED && touch )
	{
		// sparse gradient exchange (PeerPtrs::touch): one bit per quad of 4 consecutive splats, the 8 quads of a warp
		// are 8 aligned bits of one word
		const unsigned bal = __ballot_sync( 0xffffffffu, in_band );
		if( lane == 0 && bal )
		{
			unsigned bits = 0;
#pragma unroll
			for( int qd = 0; qd < 8; qd++ )
				bits |= ( ( bal >> ( 4 * qd ) ) & 0xfu ) ? ( 1u << qd ) : 0u;
			const unsigned q = (unsigned)gi >> 2; // first quad of the warp: a multiple of 8
			atomicOr( touch + ( q >> 5 ), bits << ( q & 31u ) );
		}
	}
	const bool has = sb.ty0 <= sb.ty1 && sb.rtx0 <= sb.rtx1;

	// ---- emission, warp-cooperative -----------------------------------------------------------------------------
	// A work item is one (splat, tile row). The items of the warp's 32 splats are numbered consecutively and dealt to
	// the lanes 32 at a time, so a warp whose splats have 1, 9 or no rows at all (a rank whose band most splats miss)
	// still runs with full lanes, and a splat at the 1024 clamp (hundreds of tile rows) is shared by the whole warp.
	// Items are handled in chunks of ITEMS_PER_CHUNK: spans first (stashed in shared memory), ONE reservation of the
	// pair list per chunk, then the emission.
	__shared__ SplatBin s_bin[PRE_THREADS];
	__shared__ unsigned s_rows[PRE_THREADS];          // inclusive scan of the row counts inside the warp
	__shared__ unsigned s_span[PRE_THREADS / 32][ITEMS_PER_CHUNK]; // lo | hi << 12 | owner lane << 24, ~0 = no tile
	SplatBin* wbin = s_bin + warp * 32;
	unsigned* wrows = s_rows + warp * 32;
	unsigned* wspan = s_span[warp];
	wbin[lane] = sb;
	unsigned rows_incl = has ? (unsigned)( sb.ty1 - sb.ty0 + 1 ) : 0u;
#pragma unroll
	for( int o = 1; o < 32; o <<= 1 )
	{
		const unsigned v = __shfl_up_sync( 0xffffffffu, rows_incl, o );
		if( lane >= o )
			rows_incl += v;
	}
	wrows[lane] = rows_incl;
	const unsigned R = __shfl_sync( 0xffffffffu, rows_incl, 31 );
	__syncwarp();

	if constexpr( SLOTS == 2 )
	{
		// Group rows need no place in the pair list, so there is nothing to reserve and nothing to stash: span and
		// emission in one pass. (The number of pairs - DevState::n_keys, the capacity check - comes out of the scan of
		// the bucket totals, k_scan_sums; the per-chunk atomicAdd on that one counter was what a block waited for.)
		for( unsigned item = lane; item < R; item += 32 )
		{
			unsigned owner = 0;
#pragma unroll
			for( int step = 16; step > 0; step >>= 1 )
				if( wrows[owner + step - 1] <= item )
					owner += step;
			const SplatBin& o = wbin[owner];
			const unsigned first = owner ? wrows[owner - 1] : 0u;
			const int ty = o.ty0 + (int)( item - first );
			int lo, hi;
			if( tile_span( o.k, ty, g.H, o.rtx0, o.rtx1, lo, hi ) )
			{
				const unsigned row = (unsigned)( o.tile_base + ty * g.tiles_x );
				emit_row_groups( g, out, row + (unsigned)lo, row + (unsigned)hi, o.slice, (unsigned)o.local, capacity, cull >> 8 );
			}
		}
		return;
	}
	for( unsigned chunk = 0; chunk < R; chunk += ITEMS_PER_CHUNK )
	{
		const unsigned n_items = min( (unsigned)ITEMS_PER_CHUNK, R - chunk );
		// pass A: spans of the chunk's items
		unsigned cnt = 0;
		for( unsigned it = lane; it < n_items; it += 32 )
		{
			const unsigned item = chunk + it;
			// owner = first lane whose inclusive row count exceeds the item number (binary search, 5 steps)
			unsigned owner = 0;
#pragma unroll
			for( int step = 16; step > 0; step >>= 1 )
				if( wrows[owner + step - 1] <= item )
					owner += step;
			const SplatBin& o = wbin[owner];
			const unsigned first = owner ? wrows[owner - 1] : 0u;
			const int ty = o.ty0 + (int)( item - first );
			int lo, hi;
			unsigned span = 0xffffffffu;
			if( tile_span( o.k, ty, g.H, o.rtx0, o.rtx1, lo, hi ) )
			{
				span = (unsigned)lo | ( (unsigned)hi << 12 ) | ( owner << 24 ); // tiles_x <= 4096 (gs2d_create)
				cnt += (unsigned)( hi - lo + 1 );
			}
			wspan[it] = span;
		}
		// one reservation per chunk
		unsigned incl = cnt;
#pragma unroll
		for( int o = 1; o < 32; o <<= 1 )
		{
			const unsigned v = __shfl_up_sync( 0xffffffffu, incl, o );
			if( lane >= o )
				incl += v;
		}
		const unsigned total = __shfl_sync( 0xffffffffu, incl, 31 );
		unsigned long long base = 0;
		if( lane == 0 && total )
			base = atomicAdd( &st->n_keys, (unsigned long long)total );
		base = __shfl_sync( 0xffffffffu, base, 0 );
		__syncwarp();
		// pass B: emit (skipped when the list is full: the scan flags the overflow and the host redoes the iteration)
		if( total && base + total <= capacity )
		{
			unsigned long long pos = base + incl - cnt;
			for( unsigned it = lane; it < n_items; it += 32 )
			{
				const unsigned span = wspan[it];
				if( span == 0xffffffffu )
					continue;
				const unsigned item = chunk + it, owner = span >> 24;
				const SplatBin& o = wbin[owner];
				const unsigned first = owner ? wrows[owner - 1] : 0u;
				const int ty = o.ty0 + (int)( item - first );
				const unsigned lo = span & 0xfffu, hi = ( span >> 12 ) & 0xfffu;
				const unsigned row = (unsigned)( o.tile_base + ty * g.tiles_x );
				emit_row<SLOTS>( g, out, row + lo, row + hi, o.slice, (unsigned)o.local, pos, capacity );
				pos += hi - lo + 1u;
			}
		}
		__syncwarp();
	}
}

// ---------------------------------------------------------------------------------------------- band filter ---
// A rank that owns a band of tile rows (N ranks): the band test of every splat, streamed - three parameters per
// splat - with what hangs on it: the reach masks of the quads this rank owns and the touch bitmap of the sparse
// exchange (PeerPtrs), and the list of the splats that passed, which k_preprocess<.., LISTED> then walks with full
// warps. In k_preprocess itself 1 splat in 7 passes at 8 ranks, so that nearly every warp ran the double-precision
// sincos, the divisions and the record stores for four or five lanes.
#define FILTER_THREADS 1024
__global__ void __launch_bounds__( FILTER_THREADS ) k_band_filter( Geometry g, const float* __restrict__ params,
                                                                   RowRec* __restrict__ rowrec, int mark_rowrec,
                                                                   unsigned* __restrict__ touch, Buffers::Reach reach,
                                                                   unsigned* __restrict__ band_list,
                                                                   unsigned long long* __restrict__ band_count )
{
	const int gi = blockIdx.x * FILTER_THREADS + threadIdx.x;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	bool in_band = false;
	unsigned reaches = 0;
	if( gi < g.n_total )
	{
		const float py = params[PL_PY * g.n_pad + gi];
		const float sx = params[PL_SX * g.n_pad + gi];
		const float sy = params[PL_SY * g.n_pad + gi];
		if( reach.now && ( gi >> 2 ) >= reach.q0 && ( gi >> 2 ) < reach.q1 )
			reaches = reach_mask( py, sx, sy, reach.band_px, reach.n_ranks, reach.slack );
		in_band = !band_rejects( py, sx, sy, g.band_ty0 * GS2D_TILE, g.band_ty1 * GS2D_TILE );
		if( !in_band && mark_rowrec )
			rowrec[gi].rows = ROWREC_NONE; // k_row_table without its list looks at every splat
	}
	if( reach.now )
	{
		unsigned m = reaches;
		m |= __shfl_xor_sync( 0xffffffffu, m, 1 );
		m |= __shfl_xor_sync( 0xffffffffu, m, 2 );
		const int q = gi >> 2;
		if( ( lane & 3 ) == 0 && q >= reach.q0 && q < reach.q1 )
			reach.now[q] = (unsigned char)m;
	}
	const unsigned bal = __ballot_sync( 0xffffffffu, in_band );
	if( touch && lane == 0 && bal )
	{
		// one bit per quad of 4 consecutive splats, the 8 quads of a warp are 8 aligned bits of one word
		unsigned bits = 0;
#pragma unroll
		for( int qd = 0; qd < 8; qd++ )
			bits |= ( ( bal >> ( 4 * qd ) ) & 0xfu ) ? ( 1u << qd ) : 0u;
		const unsigned q = (unsigned)gi >> 2; // first quad of the warp: a multiple of 8
		atomicOr( touch + ( q >> 5 ), bits << ( q & 31u ) );
	}
	// one reservation of the list per block (a reservation per warp is 31 000 returning atomics on one address)
	__shared__ unsigned wcount[FILTER_THREADS / 32];
	__shared__ unsigned long long block_at;
	if( lane == 0 )
		wcount[warp] = (unsigned)__popc( bal );
	__syncthreads();
	if( warp == 0 )
	{
		const unsigned mine = wcount[lane];
		unsigned incl = mine;
#pragma unroll
		for( int o = 1; o < 32; o <<= 1 )
		{
			const unsigned v = __shfl_up_sync( 0xffffffffu, incl, o );
			if( lane >= o )
				incl += v;
		}
		wcount[lane] = incl - mine;
		if( lane == 31 )
			block_at = incl ? atomicAdd( band_count, (unsigned long long)incl ) : 0ull;
	}
	__syncthreads();
	if( in_band )
		band_list[block_at + wcount[warp] + __popc( bal & ( ( 1u << lane ) - 1u ) )] = (unsigned)gi;
}

// ---------------------------------------------------------------------------------------------- row table ---
// The exact scan-line footprint (solve_quadratic per image row, main.cpp:171-185, :492-509: two IEEE divisions and a
// square root) used to be solved by the forward blend for every (staged list entry, tile row): all 16 rows of every
// entry of every tile a splat touches (~4 tiles per tile row), by warps that spent a quarter of their issue slots on
// it. Here every (splat, image row) is solved ONCE, by a dense grid: 8 lanes per splat walk its rows 8 at a time
// (36..60 rows at the reference's initial scales: 5..8 steps, lanes 82 % busy), 4 splats per warp. Only the rows
// inside [begY, endY] (and inside the image and this rank's band) are solved and written; the blend masks the other
// rows of a tile out itself (SplatRec::flags). Same arithmetic as the in-kernel solve (row_span_dense == row_span bit
// for bit, tests/test_gpu_parity.py::test_row_table_equals_the_in_kernel_solve), so the coverage words - and the
// image - are identical to it. Measured at 4096^2 / 1 M (profiles/r2_*): 148 M warp-instructions, 0.16 ms at 87 %
// issue-active, against 0.39 ms of blend time saved.
// Tried and dropped: the same loop fused into k_preprocess behind (or, for every other warp, in front of) its
// emission phase, on the reading that the emission leaves 62 % of the issue slots idle - the fused kernel took
// exactly the sum of the two (0.213 + 0.154 = 0.367 ms at 60 % issue-active, either order): with 10 resident warps
// per scheduler the solve's dependent chain (sqrt, two divisions) finds no more free slots than it does alone. Two
// kernels on two streams do not overlap either (the first kernel's blocks hold the SMs).
#define ROWTAB_LANES 8
// rt_list / rt_count: the splats to solve (k_preprocess appended them; a rank's band) - a grid-stride loop over the
// list; nullptr: every splat, one group of lanes each.
__global__ void __launch_bounds__( 256 ) k_row_table( Geometry g, const RowRec* __restrict__ rowrec,
                                                      unsigned* __restrict__ rowtab,
                                                      const unsigned* __restrict__ rt_list,
                                                      const unsigned long long* __restrict__ rt_count )
{
	const unsigned n = rt_list ? (unsigned)*rt_count : (unsigned)g.n_total;
	for( unsigned t = blockIdx.x * 256u + threadIdx.x; t < n * ROWTAB_LANES; t += gridDim.x * 256u )
	{
		const unsigned sub = t % ROWTAB_LANES;
		const unsigned gi = rt_list ? rt_list[t / ROWTAB_LANES] : t / ROWTAB_LANES;
		const float4* rr = reinterpret_cast<const float4*>( rowrec + gi );
		const float4 q1 = __ldg( rr + 1 ); // d begY endY rows
		const unsigned rows = (unsigned)__float_as_int( q1.w );
		if( rows == ROWREC_NONE )
			continue;
		const float4 q0 = __ldg( rr );     // px py a b
		// [y0, y1] lies inside [begY, endY] and inside the image (k_preprocess), so the row tests of row_span are void
		const int y0 = (int)( rows & 0xfffffu ), y1 = y0 + (int)( rows >> 20 );
		const float px = q0.x, py = q0.y, a = q0.z, b = q0.w, d = q1.x;
		const float a4 = xmul( 4.0f, a ), b2 = xmul( 2.0f, b );
		const bool a_ok = far_from_the_edges( a );
		const float ra = rcp_refined( a );
		unsigned* tab = rowtab + (size_t)gi * GS2D_ROWTAB_STRIDE;
#pragma unroll 1
		for( int y = y0 + (int)sub; y <= y1; y += ROWTAB_LANES )
		{
			int begX = 0, endX = 0;
			const bool any = row_span_dense( px, py, a, b, d, a4, b2, ra, a_ok, y, begX, endX );
			tab[y & ( GS2D_ROWTAB_STRIDE - 1 )] = pack_row_span( any, begX, endX, g.W );
		}
	}
}

// ------------------------------------------------------------------------------- exclusive scan of the counts ---
#define SCAN_THREADS 256
#define SCAN_PER_BLOCK 1024
// scan index i = tile * n_slices + slice; the 16-bit counter of bucket i lives at cnt16[slice * nt_pad + tile]
// (group rows: k_group_count has written the totals in scan order, tot16[i])
__device__ __forceinline__ unsigned bucket_count( const Geometry& g, const unsigned short* __restrict__ cnt16,
                                                  const unsigned short* __restrict__ tot16, unsigned i )
{
	if( tot16 )
		return tot16[i];
	const unsigned tile = i / (unsigned)g.n_slices, slice = i - tile * (unsigned)g.n_slices;
	return cnt16[(size_t)slice * g.nt_pad + tile];
}

// offset[i] = block-local exclusive scan of the bucket counts; sums[block] = block total
__global__ void __launch_bounds__( SCAN_THREADS ) k_scan_local( Geometry g, const unsigned short* __restrict__ cnt16,
                                                                const unsigned short* __restrict__ tot16,
                                                                unsigned* __restrict__ offset, unsigned base,
                                                                unsigned n, unsigned n_off, unsigned* __restrict__ sums )
{
	// scans the bucket window [base, n): the buckets of this rank's band of tile rows; offsets are written for
	// [base, n_off), n_off = n or n + 1 (group rows: the end of the last bucket is looked up as the start of the next)
	__shared__ unsigned wtot[SCAN_THREADS / 32];
	const unsigned i0 = base + blockIdx.x * SCAN_PER_BLOCK + threadIdx.x * 4;
	unsigned v[4];
#pragma unroll
	for( int k = 0; k < 4; k++ )
		v[k] = ( i0 + k < n ) ? bucket_count( g, cnt16, tot16, i0 + k ) : 0u;
	unsigned tsum = v[0] + v[1] + v[2] + v[3], incl = tsum;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
	for( int o = 1; o < 32; o <<= 1 )
	{
		unsigned t = __shfl_up_sync( 0xffffffffu, incl, o );
		if( lane >= o )
			incl += t;
	}
	if( lane == 31 )
		wtot[warp] = incl;
	__syncthreads();
	unsigned wb = 0, tot = 0;
#pragma unroll
	for( int w = 0; w < SCAN_THREADS / 32; w++ )
	{
		unsigned t = wtot[w];
		if( w < warp )
			wb += t;
		tot += t;
	}
	unsigned run = wb + incl - tsum;
#pragma unroll
	for( int k = 0; k < 4; k++ )
	{
		if( i0 + k < n_off )
			offset[i0 + k] = run;
		run += v[k];
	}
	if( threadIdx.x == 0 )
		sums[blockIdx.x] = tot;
}
// one block: exclusive scan of the block totals; the capacity check of the iteration
// total_is_keys (group rows): k_preprocess has not counted the pairs; their number is the grand total of the scan
__global__ void __launch_bounds__( 1024 ) k_scan_sums( unsigned* __restrict__ sums, unsigned nb,
                                                       unsigned long long capacity, DevState* __restrict__ st,
                                                       int total_is_keys )
{
	__shared__ unsigned wtot[32];
	__shared__ unsigned carry_s;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	if( threadIdx.x == 0 )
	{
		carry_s = 0;
		// n_keys was accumulated by k_preprocess (every pair reserved, written or not)
		st->need_capacity = ( st->n_keys > capacity ) ? st->n_keys : 0ull;
	}
	__syncthreads();
	for( unsigned base = 0; base < nb; base += 1024 )
	{
		unsigned i = base + threadIdx.x;
		unsigned v = ( i < nb ) ? sums[i] : 0u, incl = v;
#pragma unroll
		for( int o = 1; o < 32; o <<= 1 )
		{
			unsigned t = __shfl_up_sync( 0xffffffffu, incl, o );
			if( lane >= o )
				incl += t;
		}
		if( lane == 31 )
			wtot[warp] = incl;
		__syncthreads();
		unsigned wb = 0, tot = 0;
		for( int w = 0; w < 32; w++ )
		{
			unsigned t = wtot[w];
			if( w < warp )
				wb += t;
			tot += t;
		}
		unsigned carry = carry_s;
		if( i < nb )
			sums[i] = carry + wb + incl - v;
		__syncthreads();
		if( threadIdx.x == 0 )
			carry_s = carry + tot;
		__syncthreads();
	}
	if( total_is_keys && threadIdx.x == 0 )
	{
		st->n_keys = carry_s;
		st->need_capacity = ( carry_s > capacity ) ? carry_s : 0ull;
	}
}
// The block-local offsets are not patched in a third pass: the consumers add the scanned block total themselves
// (16 KB of sums, cache resident).
__device__ __forceinline__ unsigned bucket_offset( const unsigned* __restrict__ offset,
                                                   const unsigned* __restrict__ sums, unsigned bucket, unsigned base )
{
	return __ldg( offset + bucket ) + __ldg( sums + ( ( bucket - base ) / SCAN_PER_BLOCK ) );
}

// ------------------------------------------------------------------------------------------- place and order ---
__device__ __forceinline__ unsigned splat_of_triple( const Geometry& g, unsigned bucket, unsigned local,
                                                    unsigned tiles_per_image )
{
	const unsigned tile = bucket / (unsigned)g.n_slices, slice = bucket - tile * (unsigned)g.n_slices;
	return ( tile / tiles_per_image ) * (unsigned)g.n_splats + slice * (unsigned)g.slice_size + local;
}

__global__ void __launch_bounds__( 256 ) k_place( Geometry g, const unsigned long long* __restrict__ triple,
                                                  const unsigned* __restrict__ offset,
                                                  const unsigned* __restrict__ sums, unsigned base,
                                                  unsigned* __restrict__ list, const DevState* __restrict__ st,
                                                  const unsigned long long* __restrict__ n_triples, int from_end )
{
	if( st->need_capacity )
		return;
	if( from_end )
	{
		// group rows: the triples are the overflow arrivals of their buckets, ranked among themselves; they fill the
		// bucket from its end (k_rank_groups puts the entries of the row in front of them)
		const unsigned long long n = *n_triples;
		const unsigned tpi = (unsigned)( g.tiles_x * g.tiles_y );
		for( unsigned long long i = (unsigned long long)blockIdx.x * 256 + threadIdx.x; i < n;
		     i += (unsigned long long)gridDim.x * 256 )
		{
			const unsigned long long t = triple[i];
			list[bucket_offset( offset, sums, (unsigned)t + 1u, base ) - 1u - (unsigned)( t >> 48 )] =
			    splat_of_triple( g, (unsigned)t, (unsigned)( t >> 32 ) & 0xffffu, tpi );
		}
		return;
	}
	const unsigned long long n = *n_triples; // all pairs (st->n_keys) or, with slot binning, the overflow arrivals
	const unsigned long long stride = (unsigned long long)gridDim.x * 256;
	const unsigned tpi = (unsigned)( g.tiles_x * g.tiles_y );
	unsigned long long i = (unsigned long long)blockIdx.x * 256 + threadIdx.x;
	// four independent gather chains in flight per thread (the kernel is bound by the latency of its scattered accesses)
	for( ; i + 3 * stride < n; i += 4 * stride )
	{
		unsigned long long t[4];
		unsigned of[4];
#pragma unroll
		for( int k = 0; k < 4; k++ )
			t[k] = __ldcs( triple + i + k * stride );
#pragma unroll
		for( int k = 0; k < 4; k++ )
			of[k] = bucket_offset( offset, sums, (unsigned)t[k], base );
#pragma unroll
		for( int k = 0; k < 4; k++ )
			list[of[k] + (unsigned)( t[k] >> 48 )] =
			    splat_of_triple( g, (unsigned)t[k], (unsigned)( t[k] >> 32 ) & 0xffffu, tpi );
	}
	for( ; i < n; i += stride )
	{
		const unsigned long long t = triple[i];
		list[bucket_offset( offset, sums, (unsigned)t, base ) + (unsigned)( t >> 48 )] =
		    splat_of_triple( g, (unsigned)t, (unsigned)( t >> 32 ) & 0xffffu, tpi );
	}
}

// One thread per bucket: order its entries by splat index (rank by counting: the indices are unique), list -> out.
// SLOTS: the first GS2D_BUCKET_SLOTS entries come from the bucket's slot row (16-bit indices inside the slice, one
// 16-byte load), later arrivals from list[start + GS2D_BUCKET_SLOTS ..) where k_place put them.
template <bool SLOTS>
__global__ void __launch_bounds__( 256 ) k_rank( Geometry g, const unsigned short* __restrict__ cnt16,
                                                 const unsigned* __restrict__ offset,
                                                 const unsigned* __restrict__ sums,
                                                 unsigned* __restrict__ list, unsigned* __restrict__ out,
                                                 uint2* __restrict__ tile_range, unsigned base, unsigned n_buckets,
                                                 const DevState* __restrict__ st,
                                                 const unsigned short* __restrict__ slots )
{
	if( st->need_capacity )
		return;
	const unsigned i = base + blockIdx.x * 256 + threadIdx.x;
	if( i >= n_buckets )
		return;
	const unsigned tile = i / (unsigned)g.n_slices, slice = i - tile * (unsigned)g.n_slices;
	// all four loads of a bucket (counter, offset, block sum, slot row) are independent: requested together, one
	// memory round trip instead of two (the kernel waits on L2, profiles/r2_ncu.txt: 9 long-scoreboard stalls per issue)
	uint4 w = make_uint4( 0u, 0u, 0u, 0u );
	if( SLOTS )
		w = __ldcs( reinterpret_cast<const uint4*>( slots + ( (size_t)slice * g.nt_pad + tile ) * GS2D_BUCKET_SLOTS ) );
	const unsigned n = cnt16[(size_t)slice * g.nt_pad + tile];
	const unsigned start = bucket_offset( offset, sums, i, base );
	if( slice == 0 )
		tile_range[tile].x = start;
	if( slice == (unsigned)g.n_slices - 1 )
		tile_range[tile].y = start + n;
	unsigned* p = list + start;
	unsigned* q = out + start;
	unsigned sl[GS2D_BUCKET_SLOTS];
	if( SLOTS )
	{
		static_assert( GS2D_BUCKET_SLOTS == 8, "k_rank reads a slot row as one uint4 and sorts 8 registers" );
		if( n == 0u )
			return;
		// all entries of a bucket share the image and the slice: global index = first + index inside the slice
		const unsigned first = ( tile / (unsigned)( g.tiles_x * g.tiles_y ) ) * (unsigned)g.n_splats + slice * (unsigned)g.slice_size;
		sl[0] = first + ( w.x & 0xffffu );
		sl[1] = first + ( w.x >> 16 );
		sl[2] = first + ( w.y & 0xffffu );
		sl[3] = first + ( w.y >> 16 );
		sl[4] = first + ( w.z & 0xffffu );
		sl[5] = first + ( w.z >> 16 );
		sl[6] = first + ( w.w & 0xffffu );
		sl[7] = first + ( w.w >> 16 );
		if( n > 8u ) // rare: join the slot entries with the overflow arrivals in the unordered list, then rank below
		{
#pragma unroll
			for( int a = 0; a < 8; a++ )
				p[a] = sl[a];
		}
	}
	if( n <= 8u )
	{
		// the common case (3.3 entries on average): a fixed 19-comparator network over 8 registers padded with
		// +inf - no data-dependent control flow, so the lanes of a warp (buckets of all sizes) stay converged
		unsigned e[8];
#pragma unroll
		for( int a = 0; a < 8; a++ )
			e[a] = (unsigned)a < n ? ( SLOTS ? sl[a] : p[a] ) : 0xffffffffu;
#define CE( x, y )                                                                                                    \
	{                                                                                                                 \
		const unsigned lo_ = min( e[x], e[y] ), hi_ = max( e[x], e[y] );                                              \
		e[x] = lo_;                                                                                                   \
		e[y] = hi_;                                                                                                   \
	}
		CE( 0, 1 ) CE( 2, 3 ) CE( 4, 5 ) CE( 6, 7 )
		CE( 0, 2 ) CE( 1, 3 ) CE( 4, 6 ) CE( 5, 7 )
		CE( 1, 2 ) CE( 5, 6 ) CE( 0, 4 ) CE( 3, 7 )
		CE( 1, 5 ) CE( 2, 6 )
		CE( 1, 4 ) CE( 3, 6 )
		CE( 2, 4 ) CE( 3, 5 )
		CE( 3, 4 )
#undef CE
#pragma unroll
		for( int a = 0; a < 8; a++ )
			if( (unsigned)a < n )
				q[a] = e[a];
	}
	else
		for( unsigned a = 0; a < n; a++ )
		{
			const unsigned e = p[a];
			unsigned r = 0;
			for( unsigned b = 0; b < n; b++ )
				r += p[b] < e ? 1u : 0u;
			q[r] = e;
		}
}

// ------------------------------------------------------------------------------------- group rows (GS2D_SLOTS=2) ---
// One thread per (group of four consecutive tiles, index slice), consecutive threads = consecutive slices of a group:
// the rows (64 bytes), the arrival counters and the overflow words of a warp are contiguous.
// store into the block's staging array through its 32-bit shared-memory address (a generic pointer makes the
// compiler rebuild the shared window's base in front of every predicated store)
struct PutShared
{
	enum { SCALE = 4 }; // positions are byte addresses
	__device__ __forceinline__ void operator()( unsigned at, unsigned v ) const
	{
		asm volatile( "st.shared.u32 [%0], %1;" ::"r"( at ), "r"( v ) : "memory" );
	}
};
#define RANK_STAGE 8192 // list entries a block of k_rank_groups can stage in shared memory (32 KB)
struct GroupWindow
{
	unsigned grp_lo, n_pairs;    // pairs [grp_lo * n_slices, .. + n_pairs)
	unsigned tile_lo, tile_hi;   // tiles of this rank's band: what the scan covers
};
// the keys of one row: only the 32-byte sectors that hold entries are fetched
__device__ __forceinline__ void load_row( unsigned* w, const unsigned* __restrict__ rows, size_t pair, unsigned c )
{
	const uint4* r = reinterpret_cast<const uint4*>( rows ) + pair * ( GS2D_GROUP_SLOTS / 4 );
	uint4 q[4] = { make_uint4( 0u, 0u, 0u, 0u ), make_uint4( 0u, 0u, 0u, 0u ), make_uint4( 0u, 0u, 0u, 0u ),
	               make_uint4( 0u, 0u, 0u, 0u ) };
	if( c > 0u )
	{
		q[0] = __ldcs( r );
		q[1] = __ldcs( r + 1 );
	}
	if( c > 8u )
	{
		q[2] = __ldcs( r + 2 );
		q[3] = __ldcs( r + 3 );
	}
#pragma unroll
	for( int i = 0; i < 4; i++ )
	{
		w[4 * i] = q[i].x;
		w[4 * i + 1] = q[i].y;
		w[4 * i + 2] = q[i].z;
		w[4 * i + 3] = q[i].w;
	}
}

// entries per (tile, slice) bucket in scan order: what the exclusive scan turns into the list offsets
__global__ void __launch_bounds__( 256 ) k_group_count( Geometry g, const unsigned* __restrict__ grp_cnt,
                                                        const unsigned* __restrict__ rows,
                                                        const unsigned long long* __restrict__ ovf64,
                                                        unsigned short* __restrict__ tot16, GroupWindow win )
{
	const unsigned p = blockIdx.x * 256u + threadIdx.x;
	const bool valid = p < win.n_pairs; // no early exit: the warp votes below
	const unsigned ns = (unsigned)g.n_slices;
	const size_t pair = (size_t)win.grp_lo * ns + ( valid ? p : 0u );
	const unsigned grp = (unsigned)( pair / ns ), slice = (unsigned)( pair - (size_t)grp * ns );
	const unsigned c_raw = valid ? grp_cnt[pair] : 0u, c = min( c_raw, (unsigned)GS2D_GROUP_SLOTS );
	unsigned w[GS2D_GROUP_SLOTS], n[4];
	load_row( w, rows, pair, c );
	group_counts( w, c, __reduce_max_sync( 0xffffffffu, c ), n );
	if( c_raw > (unsigned)GS2D_GROUP_SLOTS )
	{
		const unsigned long long ov = ovf64[pair];
#pragma unroll
		for( unsigned f = 0; f < 4u; f++ )
			n[f] += (unsigned)( ov >> ( 16u * f ) ) & 0xffffu;
	}
#pragma unroll
	for( unsigned f = 0; f < 4u; f++ )
	{
		const unsigned tile = grp * 4u + f;
		if( valid && tile >= win.tile_lo && tile < win.tile_hi )
			tot16[(size_t)tile * ns + slice] = (unsigned short)n[f];
	}
}

// k_group_count + k_scan_local + k_scan_sums in ONE kernel (n_slices <= 256: a block of 256 (group, slice) pairs is
// 256 / n_slices whole groups = 1024 consecutive buckets of the scan order). Every thread counts the entries of the
// four buckets of its pair straight from the group row; the counts change places through shared memory (pair order ->
// scan order), the block scans them, and the block totals are chained by a decoupled look-back (each block publishes
// its total, then its inclusive prefix, in one 64-bit status word; a block adds up the totals of its predecessors
// until it meets an inclusive prefix). Blocks take their place in the chain from a ticket counter, so a block's
// predecessors have always started. Same outputs as the three kernels: block-local offsets + sums[block]; the last
// block writes the number of pairs and the capacity verdict.
#define SCAN_FLAG_AGG ( 1ull << 62 )
#define SCAN_FLAG_INC ( 2ull << 62 )
#define SCAN_VALUE_MASK ( ( 1ull << 62 ) - 1ull )
__global__ void __launch_bounds__( 256 ) k_group_scan( Geometry g, const unsigned* __restrict__ grp_cnt,
                                                       const unsigned* __restrict__ rows,
                                                       const unsigned long long* __restrict__ ovf64,
                                                       unsigned* __restrict__ offset, unsigned* __restrict__ sums,
                                                       unsigned long long* __restrict__ ticket,
                                                       unsigned long long* __restrict__ status, GroupWindow win,
                                                       unsigned n_blocks, unsigned n_off,
                                                       unsigned long long capacity, DevState* __restrict__ st )
{
	__shared__ unsigned short cnt_s[SCAN_PER_BLOCK];
	__shared__ unsigned wtot[8];
	__shared__ unsigned blk_s;
	if( threadIdx.x == 0 )
		blk_s = (unsigned)atomicAdd( ticket, 1ull );
	__syncthreads();
	const unsigned blk = blk_s;
	const unsigned ns = (unsigned)g.n_slices;
	const unsigned p = blk * 256u + threadIdx.x;
	const bool valid = p < win.n_pairs;
	const size_t pair = (size_t)win.grp_lo * ns + ( valid ? p : 0u );
	const unsigned grp = (unsigned)( pair / ns ), slice = (unsigned)( pair - (size_t)grp * ns );
	{
		const unsigned c_raw = valid ? grp_cnt[pair] : 0u, c = min( c_raw, (unsigned)GS2D_GROUP_SLOTS );
		unsigned w[GS2D_GROUP_SLOTS], n[4];
		load_row( w, rows, pair, c );
		group_counts( w, c, __reduce_max_sync( 0xffffffffu, c ), n );
		if( c_raw > (unsigned)GS2D_GROUP_SLOTS )
		{
			const unsigned long long ov = ovf64[pair];
#pragma unroll
			for( unsigned f = 0; f < 4u; f++ )
				n[f] += (unsigned)( ov >> ( 16u * f ) ) & 0xffffu;
		}
		// pair order -> scan order: bucket (tile, slice) of the block is item (tile - first tile of the block) * ns + slice
		// (from the thread index: a thread beyond the window's pairs still clears its four items)
		const unsigned g_local = threadIdx.x / ns, s_local = threadIdx.x - g_local * ns;
#pragma unroll
		for( unsigned f = 0; f < 4u; f++ )
		{
			const unsigned tile = grp * 4u + f;
			const bool in = valid && tile >= win.tile_lo && tile < win.tile_hi;
			cnt_s[( g_local * 4u + f ) * ns + s_local] = (unsigned short)( in ? n[f] : 0u );
		}
	}
	__syncthreads();
	// block-local exclusive scan of the 1024 items, four per thread (as k_scan_local)
	const unsigned base = win.grp_lo * 4u * ns; // first bucket of block 0
	const unsigned i0 = base + blk * SCAN_PER_BLOCK + threadIdx.x * 4u;
	unsigned v[4];
#pragma unroll
	for( int k = 0; k < 4; k++ )
		v[k] = cnt_s[threadIdx.x * 4u + k];
	const unsigned tsum = v[0] + v[1] + v[2] + v[3];
	unsigned incl = tsum;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
	for( int o = 1; o < 32; o <<= 1 )
	{
		const unsigned t = __shfl_up_sync( 0xffffffffu, incl, o );
		if( lane >= o )
			incl += t;
	}
	if( lane == 31 )
		wtot[warp] = incl;
	__syncthreads();
	unsigned wb = 0, tot = 0;
#pragma unroll
	for( int w = 0; w < 8; w++ )
	{
		const unsigned t = wtot[w];
		if( w < warp )
			wb += t;
		tot += t;
	}
	unsigned run = wb + incl - tsum;
#pragma unroll
	for( int k = 0; k < 4; k++ )
	{
		if( i0 + k < n_off )
			offset[i0 + k] = run;
		run += v[k];
	}
	// chain of the block totals: the first warp looks back 32 predecessors at a time
	if( warp == 0 )
	{
		if( lane == 0 && blk > 0u )
			asm volatile( "st.release.gpu.global.u64 [%0], %1;" ::"l"( status + blk ), "l"( SCAN_FLAG_AGG | (unsigned long long)tot ) : "memory" );
		unsigned long long prefix = 0;
		for( int j = (int)blk - 1; j >= 0; j -= 32 )
		{
			const int idx = j - lane;
			unsigned long long sv = SCAN_FLAG_INC; // in front of block 0: an inclusive prefix of 0
			if( idx >= 0 )
			{
				unsigned spin = 0;
				do
				{
					asm volatile( "ld.acquire.gpu.global.u64 %0, [%1];" : "=l"( sv ) : "l"( status + idx ) : "memory" );
					if( sv == 0ull )
					{
						__nanosleep( 64 ); // leave the issue slots and the L2 to the blocks that are still counting
						if( ++spin > ( 1u << 22 ) )
						{
							st->peer_fault = 3; // cannot happen (ticket order); never hang the device
							sv = SCAN_FLAG_INC;
						}
					}
				} while( sv == 0ull );
			}
			// the nearest predecessor with an inclusive prefix ends the walk; the totals of the blocks in front of it count
			const unsigned inc = __ballot_sync( 0xffffffffu, ( sv & SCAN_FLAG_INC ) != 0ull );
			const unsigned upto = inc ? (unsigned)__ffs( inc ) - 1u : 31u;
			const unsigned mine = (unsigned)lane <= upto ? (unsigned)( sv & 0xffffffffull ) : 0u; // totals fit 32 bits (offsets are)
			prefix += __reduce_add_sync( 0xffffffffu, mine );
			if( inc )
				break;
		}
		if( lane == 0 )
		{
			asm volatile( "st.release.gpu.global.u64 [%0], %1;" ::"l"( status + blk ), "l"( SCAN_FLAG_INC | ( prefix + tot ) ) : "memory" );
			sums[blk] = (unsigned)prefix;
			if( blk == n_blocks - 1u )
			{
				const unsigned long long total = prefix + tot;
				st->n_keys = total;
				st->need_capacity = ( total > capacity ) ? total : 0ull;
			}
		}
	}
}

// The lists of the four tiles of a group, ordered by splat index: sort the row's keys once, deal them to the tiles
// they cover (gs2d_bins.cuh). A bucket with overflow arrivals (k_place put them at its end) is joined in the
// unordered list and ranked by counting, like the sort-free path does for every long bucket.
__global__ void __launch_bounds__( 256 ) k_rank_groups( Geometry g, const unsigned* __restrict__ grp_cnt,
                                                        const unsigned* __restrict__ rows,
                                                        const unsigned long long* __restrict__ ovf64,
                                                        const unsigned* __restrict__ offset,
                                                        const unsigned* __restrict__ sums, unsigned* __restrict__ list,
                                                        unsigned* __restrict__ out, uint2* __restrict__ tile_range,
                                                        unsigned base, GroupWindow win, const DevState* __restrict__ st )
{
	if( st->need_capacity )
		return;
	const unsigned p = blockIdx.x * 256u + threadIdx.x;
	const bool valid = p < win.n_pairs; // no early exit: the warp votes and the block synchronises below
	const unsigned ns = (unsigned)g.n_slices;
	// A block that holds whole groups (all slices of 256 / n_slices groups) writes ONE contiguous piece of the lists:
	// the entries are dealt into shared memory and leave in full 128-byte lines. Dealt straight to global memory, a
	// warp's store of "entry i of tile f" touches ~13 sectors with a few bytes each, and the kernel ran at the L2
	// request rate (77 us for 56 MB of lists) like the emission it replaced.
	__shared__ unsigned sm_out[RANK_STAGE];
	unsigned blk_base = 0u, blk_n = 0u;
	bool staged = false;
	if( ns <= 256u )
	{
		const unsigned gpb = 256u / ns, grp_a = win.grp_lo + blockIdx.x * gpb;
		const unsigned t_a = max( grp_a * 4u, win.tile_lo ), t_b = min( ( grp_a + gpb ) * 4u, win.tile_hi );
		if( t_a < t_b )
		{
			blk_base = bucket_offset( offset, sums, t_a * ns, base );
			blk_n = bucket_offset( offset, sums, t_b * ns, base ) - blk_base;
			staged = blk_n <= (unsigned)RANK_STAGE;
		}
	}
	const size_t pair = (size_t)win.grp_lo * ns + ( valid ? p : 0u );
	const unsigned grp = (unsigned)( pair / ns ), slice = (unsigned)( pair - (size_t)grp * ns );
	const unsigned c_raw = valid ? grp_cnt[pair] : 0u, c = min( c_raw, (unsigned)GS2D_GROUP_SLOTS );
	const unsigned tpi = (unsigned)( g.tiles_x * g.tiles_y );
	unsigned start[4], first[4];
	bool in[4];
#pragma unroll
	for( unsigned f = 0; f < 4u; f++ )
	{
		const unsigned tile = grp * 4u + f;
		in[f] = valid && tile >= win.tile_lo && tile < win.tile_hi;
		start[f] = in[f] ? bucket_offset( offset, sums, tile * ns + slice, base ) : 0u;
		first[f] = ( tile / tpi ) * (unsigned)g.n_splats + slice * (unsigned)g.slice_size;
	}
	unsigned w[GS2D_GROUP_SLOTS], e[GS2D_GROUP_SLOTS];
	load_row( w, rows, pair, c );
	const unsigned long long ov = c_raw > (unsigned)GS2D_GROUP_SLOTS ? ovf64[pair] : 0ull;
	row_keys( e, w, c );
	const unsigned c_max = __reduce_max_sync( 0xffffffffu, c );
	sort8_keys( e );
	if( c_max > 8u )
	{
		sort8_keys( e + 8 );
		merge16_keys( e );
	}
	unsigned pos[4] = { 0u, 0u, 0u, 0u };
	if( staged )
	{
		unsigned off[4];
#pragma unroll
		for( unsigned f = 0; f < 4u; f++ )
			off[f] = start[f] - blk_base;
		unsigned sm_addr = (unsigned)__cvta_generic_to_shared( sm_out );
		asm volatile( "" : "+r"( sm_addr ) ); // kept in a register (else rebuilt from the CTA id in front of every store)
#pragma unroll
		for( unsigned f = 0; f < 4u; f++ )
			off[f] = sm_addr + off[f] * 4u;
		group_deal( e, c, c_max, PutShared{}, off, in, first, pos );
	}
	else
		group_deal( e, c, c_max, PutArray{ out }, start, in, first, pos );
#pragma unroll
	for( unsigned f = 0; f < 4u; f++ )
	{
		if( !in[f] )
			continue;
		const unsigned tile = grp * 4u + f, novf = (unsigned)( ov >> ( 16u * f ) ) & 0xffffu, n = pos[f] + novf;
		if( slice == 0u )
			tile_range[tile].x = start[f];
		if( slice == ns - 1u )
			tile_range[tile].y = start[f] + n;
		if( novf )
		{
			// rare: the entries of the row join the bucket's overflow arrivals (k_place put those at its end) in the
			// unordered list; then rank by counting, like the sort-free path does for every long bucket
			unsigned* q = staged ? sm_out + ( start[f] - blk_base ) : out + start[f];
			unsigned* pl = list + start[f];
			for( unsigned a = 0; a < pos[f]; a++ )
				pl[a] = q[a];
			for( unsigned a = 0; a < n; a++ )
			{
				const unsigned ea = pl[a];
				unsigned r = 0;
				for( unsigned b = 0; b < n; b++ )
					r += pl[b] < ea ? 1u : 0u;
				q[r] = ea;
			}
		}
	}
	if( staged )
	{
		__syncthreads();
		for( unsigned i = threadIdx.x; i < blk_n; i += 256u )
			out[blk_base + i] = sm_out[i];
	}
}

// ---------------------------------------------------------------------------------------------------- launch ---
bool gs2d_exact_mode() { return gs2d_tune( "GS2D_EXACT", 0 ) != 0; }

int launch_preprocess_and_bin( const Geometry& g, const Buffers& b, int* sorted_buf, cudaStream_t s,
                               cudaEvent_t after_emit, bool use_slots, const SideStream* side )
{
	int launches = 0;
	*sorted_buf = 0; // the ordered lists end up in vals[0]
	cudaMemsetAsync( &b.state->n_keys, 0, 2 * sizeof( unsigned long long ), s );
	if( g.n_total == 0 )
	{
		cudaMemsetAsync( b.tile_range, 0, sizeof( uint2 ) * (size_t)g.n_tiles_total, s );
		if( after_emit )
			cudaEventRecord( after_emit, s );
		return 0;
	}
	const unsigned nbk = (unsigned)g.n_tiles_total * (unsigned)g.n_slices;
	// 0: every pair a triple; 1: slot rows per (tile, slice); 2: group rows (gs2d_bins.cuh), the default
	const int mode = use_slots ? gs2d_tune( "GS2D_SLOTS", 2 ) : 0;
	const bool slots = mode != 0;
	const size_t n_cnt16 = (size_t)g.nt_pad * g.n_slices, n_pairs_all = (size_t)( g.nt_pad / 4 ) * g.n_slices;
	PairOut po;
	po.triple = b.triples;
	po.cnt64 = reinterpret_cast<unsigned long long*>( b.bucket_cnt );
	po.slots = b.bucket_slots;
	// the overflow counter sits behind the bucket counters and is cleared with them; behind it the arrival counters
	// of the group rows (cleared too when they are in use) and the totals in scan order (every one is written)
	po.n_over = reinterpret_cast<unsigned long long*>( b.bucket_cnt + n_cnt16 );
	// n_over, rt_count, band_count, then the ticket counter and the status words of k_group_scan's look-back
	unsigned long long* scan_ticket = po.n_over + 3;
	const size_t n_status = (size_t)g.n_tiles_total * g.n_slices / SCAN_PER_BLOCK + 4;
	po.grp_cnt = reinterpret_cast<unsigned*>( scan_ticket + 1 + n_status );
	po.grp_rows = reinterpret_cast<unsigned*>( b.bucket_slots );
	po.ovf64 = po.cnt64;
	unsigned short* tot16 = reinterpret_cast<unsigned short*>( po.grp_cnt + n_pairs_all );
	// a rank with a band (not the whole image): k_row_table works from the list of the splats that have rows in it
	const bool banded = g.band_ty0 > 0 || g.band_ty1 < g.tiles_y;
	po.rt_count = po.n_over + 1;
	po.rt_list = banded && b.rt_list && gs2d_tune( "GS2D_RT_LIST", 1 ) ? b.rt_list : nullptr;
	// (worth it when at most a third of the Gaussians pass: with two ranks more than half do, and the filter pass costs
	// what the fuller warps save - measured 0.105 vs 0.102 ms for one rank's band of 2; GS2D_BAND_LIST=2 forces it)
	const int band_list = gs2d_tune( "GS2D_BAND_LIST", 1 );
	const bool listed = banded && b.band_list && band_list &&
	                    ( band_list > 1 || ( g.band_ty1 - g.band_ty0 ) * 3 <= g.tiles_y );
	po.band_list = b.band_list;
	po.band_count = po.n_over + 2;
	cudaMemsetAsync( b.bucket_cnt, 0,
	                 sizeof( unsigned short ) * n_cnt16 + ( 4 + n_status ) * sizeof( unsigned long long ) +
	                     ( mode == 2 ? sizeof( unsigned ) * n_pairs_all : 0 ),
	                 s );
	if( b.touch )
		cudaMemsetAsync( b.touch, 0, sizeof( unsigned ) * ( (size_t)g.n_pad / 128 + 1 ), s );
	const int nb = ( g.n_total + PRE_THREADS - 1 ) / PRE_THREADS;
	const int relaxed = gs2d_exact_mode() ? 0 : 1, rowtab = gs2d_tune( "GS2D_ROWTAB", 1 ) ? 1 : 0;
	const int cull = gs2d_tune( "GS2D_CULL", 1 ) | ( gs2d_tune( "GS2D_K1_PROBE", 0 ) << 8 );
	if( listed )
	{
		k_band_filter<<<( g.n_total + FILTER_THREADS - 1 ) / FILTER_THREADS, FILTER_THREADS, 0, s>>>(
		    g, b.params, b.rowrec, po.rt_list ? 0 : 1, b.touch, b.reach, b.band_list, po.n_over + 2 );
		launches += 1;
	}
#define K1( MODE, LISTED )                                                                                            \
	k_preprocess<MODE, LISTED><<<nb, PRE_THREADS, 0, s>>>( g, b.params, b.rec, b.rowrec, po, b.capacity, cull, relaxed,  \
	                                                       rowtab, b.touch, b.reach, b.state )
	if( mode == 2 )
	{
		if( listed )
			K1( 2, true );
		else
			K1( 2, false );
	}
	else if( slots )
	{
		if( listed )
			K1( 1, true );
		else
			K1( 1, false );
	}
	else
	{
		if( listed )
			K1( 0, true );
		else
			K1( 0, false );
	}
#undef K1
	launches += 1;
	if( after_emit )
		cudaEventRecord( after_emit, s );
	// The row table needs the RowRecs only, the scan / place / rank kernels the bucket counters only: the two run side
	// by side (k_row_table is issue-bound, the binning kernels wait on L2 / HBM most of the time).
	bool forked = false;
	if( rowtab )
	{
		const unsigned rt_blocks = (unsigned)( ( (size_t)g.n_total * ROWTAB_LANES + 255 ) / 256 );
		cudaStream_t rs = s;
		if( side && side->stream )
		{
			cudaEventRecord( side->ev_fork, s );
			cudaStreamWaitEvent( side->stream, side->ev_fork, 0 );
			rs = side->stream;
			forked = true;
		}
		if( po.rt_list ) // persistent grid over the list
			k_row_table<<<std::min( rt_blocks, 148u * 8u ), 256, 0, rs>>>( g, b.rowrec, b.rowtab, po.rt_list, po.rt_count );
		else
			k_row_table<<<rt_blocks, 256, 0, rs>>>( g, b.rowrec, b.rowtab, nullptr, nullptr );
		launches += 1;
		if( forked )
			cudaEventRecord( side->ev_join, rs );
	}
	auto join = [&]() {
		if( forked )
			cudaStreamWaitEvent( s, side->ev_join, 0 );
	};
	// only the buckets of this rank's band of tile rows hold entries (one contiguous window for a single image)
	unsigned lo = 0, hi = nbk;
	if( g.n_images == 1 )
	{
		lo = (unsigned)g.band_ty0 * (unsigned)g.tiles_x * (unsigned)g.n_slices;
		hi = (unsigned)g.band_ty1 * (unsigned)g.tiles_x * (unsigned)g.n_slices;
	}
	if( hi <= lo )
	{
		join();
		return launches;
	}
	if( mode == 2 )
	{
		GroupWindow win;
		win.tile_lo = lo / (unsigned)g.n_slices;
		win.tile_hi = hi / (unsigned)g.n_slices;
		win.grp_lo = win.tile_lo >> 2;
		win.n_pairs = ( ( ( win.tile_hi + 3u ) >> 2 ) - win.grp_lo ) * (unsigned)g.n_slices;
		const unsigned pair_nb = ( win.n_pairs + 255u ) / 256u;
		unsigned scan_base = lo;
		if( g.n_slices <= 256 && gs2d_tune( "GS2D_SCAN_FUSED", 1 ) )
		{
			// one kernel: counts from the group rows, block scan, block totals chained by look-back. Its blocks are whole
			// groups, so its first bucket is the first bucket of the window's first GROUP; one more scan item than
			// buckets (the end of the window's last bucket)
			scan_base = win.grp_lo * 4u * (unsigned)g.n_slices;
			const unsigned scan_nb = ( hi - scan_base + 1u + SCAN_PER_BLOCK - 1 ) / SCAN_PER_BLOCK;
			k_group_scan<<<scan_nb, 256, 0, s>>>( g, po.grp_cnt, po.grp_rows, po.ovf64, b.bucket_off, b.scan_sums, scan_ticket,
			                                      scan_ticket + 1, win, scan_nb, hi + 1u, b.capacity, b.state );
			launches += 1;
		}
		else
		{
			const unsigned scan_nb = ( hi - lo + 1u + SCAN_PER_BLOCK - 1 ) / SCAN_PER_BLOCK;
			k_group_count<<<pair_nb, 256, 0, s>>>( g, po.grp_cnt, po.grp_rows, po.ovf64, tot16, win );
			k_scan_local<<<scan_nb, SCAN_THREADS, 0, s>>>( g, b.bucket_cnt, tot16, b.bucket_off, lo, hi, hi + 1u, b.scan_sums );
			k_scan_sums<<<1, 1024, 0, s>>>( b.scan_sums, scan_nb, b.capacity, b.state, 1 );
			launches += 3;
		}
		k_place<<<148, 256, 0, s>>>( g, b.triples, b.bucket_off, b.scan_sums, scan_base, b.vals[1], b.state, po.n_over, 1 );
		k_rank_groups<<<pair_nb, 256, 0, s>>>( g, po.grp_cnt, po.grp_rows, po.ovf64, b.bucket_off, b.scan_sums, b.vals[1],
		                                       b.vals[0], b.tile_range, scan_base, win, b.state );
		launches += 2;
		join();
		return launches;
	}
	const unsigned scan_nb = ( hi - lo + SCAN_PER_BLOCK - 1 ) / SCAN_PER_BLOCK;
	k_scan_local<<<scan_nb, SCAN_THREADS, 0, s>>>( g, b.bucket_cnt, nullptr, b.bucket_off, lo, hi, hi, b.scan_sums );
	k_scan_sums<<<1, 1024, 0, s>>>( b.scan_sums, scan_nb, b.capacity, b.state, 0 );
	if( slots )
	{
		// few triples, if any: a small grid (the kernel is a grid-stride loop over *n_over)
		k_place<<<148, 256, 0, s>>>( g, b.triples, b.bucket_off, b.scan_sums, lo, b.vals[1], b.state, po.n_over, 0 );
		k_rank<true><<<( hi - lo + 255 ) / 256, 256, 0, s>>>( g, b.bucket_cnt, b.bucket_off, b.scan_sums, b.vals[1],
		                                                     b.vals[0], b.tile_range, lo, hi, b.state, b.bucket_slots );
	}
	else
	{
		k_place<<<148 * 16, 256, 0, s>>>( g, b.triples, b.bucket_off, b.scan_sums, lo, b.vals[1], b.state,
		                                  &b.state->n_keys, 0 );
		k_rank<false><<<( hi - lo + 255 ) / 256, 256, 0, s>>>( g, b.bucket_cnt, b.bucket_off, b.scan_sums, b.vals[1],
		                                                      b.vals[0], b.tile_range, lo, hi, b.state, nullptr );
	}
	launches += 4;
	join();
	return launches;
}
