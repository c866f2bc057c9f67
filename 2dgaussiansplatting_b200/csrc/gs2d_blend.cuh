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

struct TileSmem
{
	SplatRec rec[2][GS2D_BATCH];        // 4 KB, double-buffered records of the staged list entries
	unsigned words[GS2D_BATCH][8];      // 1 KB, exact coverage: word j of entry k = 8x4 block j (bit = lane)
};

// lane k gathers QUADS float4 of record idx into sm.rec[stage][k]
template <int QUADS>
__device__ __forceinline__ void gather_record( TileSmem& sm, int stage, int lane, const SplatRec* __restrict__ rec,
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
}

// Exact footprint of the nb staged entries of `stage`: 16 rows x nb entries = up to 512 solves, 16 per lane.
// Lane (h = lane >> 4, r = lane & 15) solves row r of entries 2i + h. Returns the mask of entries that cover at
// least one pixel of the tile (entries outside it are never looked at again).
__device__ __forceinline__ unsigned solve_batch( TileSmem& sm, int stage, int nb, int lane, int tile_x0, int tile_y0,
                                                 int W, int H )
{
	const int h = lane >> 4, r = lane & 15;
	const int y = tile_y0 + r;
	const int boff = ( r >> 2 ) * 8 + ( r & 3 ); // byte (r & 3) of word 2*(r >> 2); the right half is 4 bytes on
	unsigned ent = 0;
#pragma unroll 2
	for( int i = 0; 2 * i < nb; i++ )
	{
		const int k = 2 * i + h;
		unsigned m = 0;
		if( k < nb )
		{
			const SplatRec& s = sm.rec[stage][k];
			m = row_mask( s.px, s.py, s.a, s.b, s.d, s.begY, s.endY, y, tile_x0, W, H );
			unsigned char* wb = reinterpret_cast<unsigned char*>( &sm.words[k][0] ) + boff;
			wb[0] = (unsigned char)( m & 0xffu );
			wb[4] = (unsigned char)( m >> 8 );
		}
		const unsigned bal = __ballot_sync( 0xffffffffu, m != 0u );
		ent |= ( ( bal & 0xffffu ) ? 1u : 0u ) << ( 2 * i );
		ent |= ( ( bal >> 16 ) ? 2u : 0u ) << ( 2 * i );
	}
	return ent;
}
