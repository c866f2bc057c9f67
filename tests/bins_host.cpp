// tests/bins_host.cpp — TEST INFRASTRUCTURE: the product's group-row binning arithmetic compiled for the HOST.
//
// 2dgaussiansplatting_b200/csrc/gs2d_bins.cuh holds the device functions behind the default binning path
// (GS2D_SLOTS=2): the key an arrival stores, the comparator networks that sort a row and the deal of the sorted keys
// to the lists of the four tiles of a group. This file includes that very header and walks it through the sequence
// of the kernels (k_preprocess emission -> k_group_count -> scan -> k_place -> k_rank_groups, gs2d_binning.cu) in
// plain sequential C++: arrivals come in the order the caller gives (the device's order is arbitrary), one "thread"
// per (group, slice). tests/test_bins_host.py checks the lists against a brute-force enumeration.
#include "host_shim.h"

#include "../2dgaussiansplatting_b200/csrc/gs2d_bins.cuh"

#include <vector>

extern "C" {

// arrivals: n_arr x 4 ints (tile_lo, tile_hi, slice, local): one tile row of one splat, tiles tile_lo..tile_hi
// (consecutive ids, may cross group boundaries). slice_size: splat = slice * slice_size + local.
// Window [tile_lo_w, tile_hi_w) as for a rank's band. Outputs: out_list (capacity cap), tile_range (2 per tile).
// Returns the total number of list entries, or -1 if cap is too small.
// fused != 0: offsets the way k_group_scan lays them out (blocks of 256 pairs = 1024 consecutive buckets from the first
// bucket of the window's first GROUP, block-local offsets + sums[block]); else the plain scan of the three kernels.
long long bins_build( int n_tiles, int n_slices, int slice_size, int tile_lo_w, int tile_hi_w, const int* arrivals,
                      long long n_arr, unsigned* out_list, long long cap, unsigned* tile_range, long long* n_overflow,
                      int fused )
{
	const unsigned ns = (unsigned)n_slices;
	const int nt_pad = ( n_tiles + 3 ) & ~3;
	const size_t n_pairs = (size_t)( nt_pad / 4 ) * ns;
	std::vector<unsigned> grp_cnt( n_pairs, 0u ), rows( n_pairs * GS2D_GROUP_SLOTS, 0xdeadbeefu ); // rows are never cleared
	std::vector<unsigned long long> ovf64( n_pairs, 0ull );
	struct Triple { unsigned bucket, local, rank; };
	std::vector<Triple> triples;

	// ---- emission (emit_row_groups / emit_group)
	for( long long i = 0; i < n_arr; i++ )
	{
		const unsigned id_lo = (unsigned)arrivals[4 * i], id_hi = (unsigned)arrivals[4 * i + 1];
		const unsigned slice = (unsigned)arrivals[4 * i + 2], local = (unsigned)arrivals[4 * i + 3];
		for( unsigned grp = id_lo >> 2; grp <= ( id_hi >> 2 ); grp++ )
		{
			const unsigned f0 = max( id_lo, grp * 4u ) - grp * 4u, f1 = min( id_hi, grp * 4u + 3u ) - grp * 4u;
			const size_t pair = (size_t)grp * ns + slice;
			const unsigned old = grp_cnt[pair]++;
			if( old < (unsigned)GS2D_GROUP_SLOTS )
				rows[pair * GS2D_GROUP_SLOTS + old] = group_key( local, f0, f1 - f0 );
			else
			{
				const unsigned long long oldv = ovf64[pair];
				ovf64[pair] += field_ones( f0, f1 );
				for( unsigned f = f0; f <= f1; f++ )
					triples.push_back( { ( grp * 4u + f ) * ns + slice, local, (unsigned)( oldv >> ( 16u * f ) ) & 0xffffu } );
			}
		}
	}
	*n_overflow = (long long)triples.size();

	// ---- k_group_count: totals in scan order
	const unsigned lo = (unsigned)tile_lo_w * ns, hi = (unsigned)tile_hi_w * ns;
	std::vector<unsigned short> tot16( (size_t)n_tiles * ns + 1, 0xabcd );
	const unsigned grp_lo = (unsigned)tile_lo_w >> 2, grp_hi = ( (unsigned)tile_hi_w + 3u ) >> 2;
	for( size_t pair = (size_t)grp_lo * ns; pair < (size_t)grp_hi * ns; pair++ )
	{
		const unsigned grp = (unsigned)( pair / ns ), slice = (unsigned)( pair % ns );
		const unsigned c_raw = grp_cnt[pair], c = min( c_raw, (unsigned)GS2D_GROUP_SLOTS );
		unsigned n[4];
		group_counts( &rows[pair * GS2D_GROUP_SLOTS], c, c, n );
		if( c_raw > (unsigned)GS2D_GROUP_SLOTS )
			for( unsigned f = 0; f < 4u; f++ )
				n[f] += (unsigned)( ovf64[pair] >> ( 16u * f ) ) & 0xffffu;
		for( unsigned f = 0; f < 4u; f++ )
		{
			const unsigned tile = grp * 4u + f;
			if( tile >= (unsigned)tile_lo_w && tile < (unsigned)tile_hi_w )
				tot16[(size_t)tile * ns + slice] = (unsigned short)n[f];
		}
	}
	// ---- exclusive scan over [lo, hi], one item more than buckets
	std::vector<unsigned> offset( (size_t)( n_tiles + 4 ) * ns + 1026, 0u );
	unsigned run = 0;
	if( !fused )
	{
		for( unsigned i = lo; i <= hi; i++ )
		{
			offset[i] = run;
			if( i < hi )
				run += tot16[i];
		}
	}
	else
	{
		// k_group_scan: thread t of block blk is pair blk * 256 + t of the window; its four counts go to items
		// ( t / ns * 4 + f ) * ns + t % ns of the block; items are scanned in order, offsets stay block-local
		const unsigned base = grp_lo * 4u * ns, n_pairs_w = ( grp_hi - grp_lo ) * ns, n_off = hi + 1u;
		const unsigned n_blocks = ( hi - base + 1u + 1023u ) / 1024u;
		std::vector<unsigned> local( offset.size(), 0u ), sums( n_blocks + 1, 0u );
		for( unsigned blk = 0; blk < n_blocks; blk++ )
		{
			unsigned short cnt_s[1024];
			for( unsigned t = 0; t < 256u; t++ )
			{
				const unsigned p = blk * 256u + t;
				const bool valid = p < n_pairs_w;
				const size_t pair = (size_t)grp_lo * ns + ( valid ? p : 0u );
				const unsigned grp = (unsigned)( pair / ns );
				const unsigned c_raw = valid ? grp_cnt[pair] : 0u, c = min( c_raw, (unsigned)GS2D_GROUP_SLOTS );
				unsigned n[4];
				group_counts( &rows[pair * GS2D_GROUP_SLOTS], c, c, n );
				if( c_raw > (unsigned)GS2D_GROUP_SLOTS )
					for( unsigned f = 0; f < 4u; f++ )
						n[f] += (unsigned)( ovf64[pair] >> ( 16u * f ) ) & 0xffffu;
				const unsigned g_local = t / ns, s_local = t - g_local * ns;
				for( unsigned f = 0; f < 4u; f++ )
				{
					const unsigned tile = grp * 4u + f;
					const bool in = valid && tile >= (unsigned)tile_lo_w && tile < (unsigned)tile_hi_w;
					cnt_s[( g_local * 4u + f ) * ns + s_local] = (unsigned short)( in ? n[f] : 0u );
				}
			}
			unsigned r = 0;
			for( unsigned i = 0; i < 1024u; i++ )
			{
				if( base + blk * 1024u + i < n_off )
					local[base + blk * 1024u + i] = r;
				r += cnt_s[i];
			}
			sums[blk] = run;
			run += r;
		}
		// what the consumers compute: bucket_offset( offset, sums, bucket, base )
		for( unsigned i = lo; i <= hi; i++ )
			offset[i] = local[i] + sums[( i - base ) / 1024u];
	}
	if( (long long)run > cap )
		return -1;
	std::vector<unsigned> list( (size_t)run + 1, 0xfeedfaceu );
	// ---- k_place( from_end )
	for( const Triple& t : triples )
		list[offset[t.bucket + 1] - 1u - t.rank] = ( t.bucket % ns ) * (unsigned)slice_size + t.local;
	// ---- k_rank_groups
	for( size_t pair = (size_t)grp_lo * ns; pair < (size_t)grp_hi * ns; pair++ )
	{
		const unsigned grp = (unsigned)( pair / ns ), slice = (unsigned)( pair % ns );
		const unsigned c_raw = grp_cnt[pair], c = min( c_raw, (unsigned)GS2D_GROUP_SLOTS );
		unsigned e[GS2D_GROUP_SLOTS];
		row_keys( e, &rows[pair * GS2D_GROUP_SLOTS], c );
		sort8_keys( e );
		if( c > 8u ) // the device takes this branch per warp (__any_sync): entries 8.. are all GS2D_KEY_NONE otherwise
		{
			sort8_keys( e + 8 );
			merge16_keys( e );
		}
		const unsigned long long ov = c_raw > (unsigned)GS2D_GROUP_SLOTS ? ovf64[pair] : 0ull;
		unsigned first[4], start[4], novf[4], pos[4] = { 0u, 0u, 0u, 0u };
		bool in[4];
		for( unsigned f = 0; f < 4u; f++ )
		{
			const unsigned tile = grp * 4u + f;
			in[f] = tile >= (unsigned)tile_lo_w && tile < (unsigned)tile_hi_w;
			start[f] = in[f] ? offset[tile * ns + slice] : 0u;
			first[f] = slice * (unsigned)slice_size;
			novf[f] = (unsigned)( ov >> ( 16u * f ) ) & 0xffffu;
		}
		group_deal( e, c, c, PutArray{ out_list }, start, in, first, pos );
		for( unsigned f = 0; f < 4u; f++ )
		{
			if( !in[f] )
				continue;
			const unsigned tile = grp * 4u + f, n = pos[f] + novf[f];
			if( slice == 0u )
				tile_range[2 * tile] = start[f];
			if( slice == ns - 1u )
				tile_range[2 * tile + 1] = start[f] + n;
			if( novf[f] )
			{
				// the entries of the row join the overflow arrivals in the unordered list, then everything is ranked
				for( unsigned a = 0; a < pos[f]; a++ )
					list[start[f] + a] = out_list[start[f] + a];
				const unsigned* pl = list.data() + start[f];
				for( unsigned a = 0; a < n; a++ )
				{
					unsigned r = 0;
					for( unsigned b = 0; b < n; b++ )
						r += pl[b] < pl[a] ? 1u : 0u;
					out_list[start[f] + r] = pl[a];
				}
			}
		}
	}
	return (long long)run;
}

// where the forward pass leaves the mask of round `round` of a tile whose list starts at `start` (round_vis_index)
long long bins_round_vis_index( unsigned start, unsigned tile, unsigned round ) { return (long long)round_vis_index( start, tile, start + 32u * round ); }

// the comparator networks on their own: sorts 16 keys in place (two sorted halves merged)
void bins_sort16( unsigned* e )
{
	sort8_keys( e );
	sort8_keys( e + 8 );
	merge16_keys( e );
}
}
