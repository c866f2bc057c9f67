// gs2d_bins.cuh — group rows: how the binning (gs2d_binning.cu, GS2D_SLOTS=2, the default) records the (tile, splat)
// pairs of a splat's tile row with ONE 32-bit store per group of four consecutive tiles, and how the per-tile lists
// are rebuilt from them in splat-index order (== the reference's blend order, main.cpp:419).
//
// A tile row of a splat spans a few consecutive tiles; the tiles it covers inside one aligned group of four are the
// fields f0 .. f0 + ext of that group. k_preprocess takes the arrival's position in the row of its (group, index
// slice) from one 32-bit atomicAdd and stores one key there:
//     key = index inside the slice << 4 | f0 << 2 | ext          (0xffffffff: no entry)
// A row holds GS2D_GROUP_SLOTS keys (64 bytes); later arrivals - rare: 4.5 arrivals per row on average at 4096^2 /
// 1 M - overflow into the (bucket, index, rank) triples of the sort-free path.
// One thread per (group, slice) then sorts the row's keys with a fixed comparator network (the index is in the high
// bits: sorted by splat index) and deals every key to the lists of the tiles it covers: the subsequence of a sorted
// sequence is sorted, so each tile's entries leave in blend order with no per-tile sort.
//
// Free of anything CUDA-specific, so that tests/test_bins_host.py can compile these very functions for the host and
// check them against a brute-force enumeration without a GPU.
#pragma once

#define GS2D_GROUP_SLOTS 16
#define GS2D_KEY_NONE 0xffffffffu

__device__ __forceinline__ unsigned group_key( unsigned local, unsigned f0, unsigned ext )
{
	return ( local << 4 ) | ( f0 << 2 ) | ext;
}
// does the arrival cover field f of its group? (f - f0 wraps to a huge number for f < f0)
__device__ __forceinline__ bool key_covers( unsigned key, unsigned f ) { return ( f - ( ( key >> 2 ) & 3u ) ) <= ( key & 3u ); }
__device__ __forceinline__ unsigned key_local( unsigned key ) { return key >> 4; }

// ones in the 16-bit fields f0 .. f1 of a 64-bit word of four counters
__device__ __forceinline__ unsigned long long field_ones( unsigned f0, unsigned f1 )
{
	return ( 0x0001000100010001ull >> ( 16u * ( 3u - ( f1 - f0 ) ) ) ) << ( 16u * f0 );
}

#define GS2D_CE( x, y )                                                                                               \
	{                                                                                                                 \
		const unsigned lo_ = min( e[x], e[y] ), hi_ = max( e[x], e[y] );                                              \
		e[x] = lo_;                                                                                                   \
		e[y] = hi_;                                                                                                   \
	}
// ascending, fixed 19-comparator network (no data-dependent control flow)
__device__ __forceinline__ void sort8_keys( unsigned* e )
{
	GS2D_CE( 0, 1 ) GS2D_CE( 2, 3 ) GS2D_CE( 4, 5 ) GS2D_CE( 6, 7 )
	GS2D_CE( 0, 2 ) GS2D_CE( 1, 3 ) GS2D_CE( 4, 6 ) GS2D_CE( 5, 7 )
	GS2D_CE( 1, 2 ) GS2D_CE( 5, 6 ) GS2D_CE( 0, 4 ) GS2D_CE( 3, 7 )
	GS2D_CE( 1, 5 ) GS2D_CE( 2, 6 )
	GS2D_CE( 1, 4 ) GS2D_CE( 3, 6 )
	GS2D_CE( 2, 4 ) GS2D_CE( 3, 5 )
	GS2D_CE( 3, 4 )
}
// Batcher's odd-even merge of the sorted halves e[0..8) and e[8..16): 25 comparators
__device__ __forceinline__ void merge16_keys( unsigned* e )
{
	GS2D_CE( 0, 8 ) GS2D_CE( 4, 12 ) GS2D_CE( 4, 8 ) GS2D_CE( 2, 10 ) GS2D_CE( 6, 14 ) GS2D_CE( 6, 10 )
	GS2D_CE( 2, 4 ) GS2D_CE( 6, 8 ) GS2D_CE( 10, 12 ) GS2D_CE( 1, 9 ) GS2D_CE( 5, 13 ) GS2D_CE( 5, 9 )
	GS2D_CE( 3, 11 ) GS2D_CE( 7, 15 ) GS2D_CE( 7, 11 ) GS2D_CE( 3, 5 ) GS2D_CE( 7, 9 ) GS2D_CE( 11, 13 )
	GS2D_CE( 1, 2 ) GS2D_CE( 3, 4 ) GS2D_CE( 5, 6 ) GS2D_CE( 7, 8 ) GS2D_CE( 9, 10 ) GS2D_CE( 11, 12 ) GS2D_CE( 13, 14 )
}

// The row of one (group, slice): key i of the four 128-bit words, GS2D_KEY_NONE beyond the c entries that were stored.
__device__ __forceinline__ void row_keys( unsigned* e, const unsigned* w, unsigned c )
{
#pragma unroll
	for( int i = 0; i < GS2D_GROUP_SLOTS; i++ )
		e[i] = (unsigned)i < c ? w[i] : GS2D_KEY_NONE;
}

// entries of the four tiles of the group among the row's c keys (any order); c_max >= c bounds the work in steps of
// four keys (the device passes the maximum over the warp, so the skip is a uniform branch).
// The 4-bit mask of the fields a key covers is spread to one byte per field (bit k -> bit 8k: the four partial
// products of m * 0x00204081 never meet in a bit position, so there are no carries) and added: four counters per add.
__device__ __forceinline__ void group_counts( const unsigned* e, unsigned c, unsigned c_max, unsigned* n )
{
	unsigned acc = 0u;
#pragma unroll
	for( int i0 = 0; i0 < GS2D_GROUP_SLOTS; i0 += 4 )
	{
		if( (unsigned)i0 >= c_max )
			break;
#pragma unroll
		for( int i = i0; i < i0 + 4; i++ )
		{
			const unsigned key = e[i];
			const unsigned m = ( ( 2u << ( key & 3u ) ) - 1u ) << ( ( key >> 2 ) & 3u ); // fields f0 .. f0 + ext (<= 3)
			acc += (unsigned)i < c ? ( ( m * 0x00204081u ) & 0x01010101u ) : 0u;
		}
	}
	n[0] = acc & 0xffu;
	n[1] = ( acc >> 8 ) & 0xffu;
	n[2] = ( acc >> 16 ) & 0xffu;
	n[3] = acc >> 24;
}

// Deal the sorted keys to the lists of the four tiles: base[off[f] + pos[f]++] = first[f] + index inside the slice for
// every tile f the key covers and that is inside the window (in[f]). pos[] must come in as 0 and returns the number
// of entries written per tile. c_max: as above.
// The four running positions live in the bytes of one word, like group_counts' counters (a row deals at most 16
// entries to a tile): one add per key advances all four, and byte f is the position of the key in tile f's list.
// put( off[f] + Put::SCALE * position, value ) stores one list entry (the device's shared-memory store works on byte
// addresses, SCALE = 4; the global-memory store and the host's on array indices).
template <class Put>
__device__ __forceinline__ void group_deal( const unsigned* e, unsigned c, unsigned c_max, const Put& put,
                                            const unsigned* off, const bool* in, const unsigned* first, unsigned* pos )
{
	// byte f of acc = Put::SCALE * entries dealt to tile f so far (at most 16 * 4)
	unsigned acc = 0u;
	const unsigned window = ( in[0] ? 1u : 0u ) | ( in[1] ? 2u : 0u ) | ( in[2] ? 4u : 0u ) | ( in[3] ? 8u : 0u );
#pragma unroll
	for( int i0 = 0; i0 < GS2D_GROUP_SLOTS; i0 += 4 )
	{
		if( (unsigned)i0 >= c_max )
			break;
#pragma unroll
		for( int i = i0; i < i0 + 4; i++ )
		{
			const unsigned key = e[i], local = key_local( key );
			// fields f0 .. f0 + ext (<= 3) inside the window; nothing for a slot beyond the row's entries
			const unsigned m = (unsigned)i < c ? ( ( ( ( 2u << ( key & 3u ) ) - 1u ) << ( ( key >> 2 ) & 3u ) ) & window ) : 0u;
#pragma unroll
			for( unsigned f = 0; f < 4u; f++ )
				if( m & ( 1u << f ) )
					put( off[f] + ( ( acc >> ( 8u * f ) ) & 0xffu ), first[f] + local );
			acc += ( ( m * 0x00204081u ) & 0x01010101u ) * (unsigned)Put::SCALE;
		}
	}
	pos[0] = ( acc & 0xffu ) / (unsigned)Put::SCALE;
	pos[1] = ( ( acc >> 8 ) & 0xffu ) / (unsigned)Put::SCALE;
	pos[2] = ( ( acc >> 16 ) & 0xffu ) / (unsigned)Put::SCALE;
	pos[3] = ( acc >> 24 ) / (unsigned)Put::SCALE;
}
// the plain store
struct PutArray
{
	enum { SCALE = 1 };
	unsigned* base;
	__device__ __forceinline__ void operator()( unsigned at, unsigned v ) const { base[at] = v; }
};

// ---- rounds of a tile's list (forward -> backward blend, gs2d_blend.cuh) -------------------------------------------
// Where the forward pass leaves the 32-bit mask of a round's entries that blended into a pixel (k_forward -> k_backward):
// one word per round of a tile. Rounds of tile t number (start_t >> 5) + t + round: floor((s + n)/32) >=
// floor(s/32) + ceil(n/32) - 1, so the words of consecutive tiles never meet; capacity/32 + n_tiles + 1 words in all,
// kept behind the coverage words (Buffers::cover). Here, not in gs2d_blend.cuh, so that the host test can see it.
__device__ __forceinline__ size_t round_vis_index( unsigned start, unsigned tile_id, unsigned base )
{
	return (size_t)( start >> 5 ) + tile_id + ( ( base - start ) >> 5 );
}

