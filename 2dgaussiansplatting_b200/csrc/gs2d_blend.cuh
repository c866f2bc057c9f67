// gs2d_blend.cuh — pieces shared by the forward and backward blend kernels (gs2d_forward.cu, gs2d_backward.cu).
//
// Work decomposition (both kernels): one CTA per 16x16 tile; every thread owns P pixels of one column
// ("register tiling over pixels"), a warp therefore owns 2P tile rows and a tile needs 8/P warps (P = 8: one warp per
// tile and no block barrier at all). Slot j of lane (lx = lane & 15, hy = lane >> 4) of warp w is pixel
//     x = tile_x0 + lx,   y = tile_y0 + w*2P + 2j + hy,
// so the 32 lanes of slot j are exactly the 16x2 strip whose coverage is ONE 32-bit word of the per-entry mask table
// (mask word i = rows 2i, 2i+1). Per list entry a warp pays the record load, the loop control and (backward) the
// shuffle reduction once for 32*P pixels instead of once per 32 pixels.
#pragma once
#include "gs2d_internal.cuh"

template <int P>
struct BlendCfg
{
	static constexpr int NW = 8 / P;       // warps per tile
	static constexpr int NT = 32 * NW;     // threads per CTA
};

struct StageSmem
{
	SplatRec rec[2][GS2D_BATCH];                 // 4 KB, double-buffered records of the staged list entries
	unsigned short mask[GS2D_BATCH][GS2D_TILE];  // 1 KB, exact scan-line coverage: 16 columns x 16 rows per entry
};

template <int NT>
__device__ __forceinline__ void cta_sync()
{
	if( NT > 32 )
		__syncthreads();
	else
		__syncwarp();
}

// Gather QUADS float4 of every record of a batch with cp.async. idx[] are this thread's prefetched list values.
template <int NT, int QUADS>
struct Gather
{
	static constexpr int ITEMS = ( GS2D_BATCH * QUADS + NT - 1 ) / NT; // cp.async per thread per batch
};

// Exact footprint of the batch in `stage`: thread -> (entry k, tile row r) pairs, 512 solves per batch.
template <int NT>
__device__ __forceinline__ void solve_masks( StageSmem& sm, int stage, int nb, int tid, int tile_x0, int tile_y0,
                                             int W, int H )
{
#pragma unroll 2
	for( int t = tid; t < GS2D_BATCH * GS2D_TILE; t += NT )
	{
		const int k = t >> 4, r = t & 15;
		unsigned m = 0;
		if( k < nb )
		{
			const SplatRec& s = sm.rec[stage][k];
			m = row_mask( s.px, s.py, s.a, s.b, s.d, s.begY, s.endY, tile_y0 + r, tile_x0, W, H );
		}
		sm.mask[k][r] = (unsigned short)m;
	}
}

// P consecutive mask words of entry k for warp w (rows w*2P .. w*2P + 2P - 1), vector loads where possible.
template <int P>
__device__ __forceinline__ void load_words( const StageSmem& sm, int k, int warp, unsigned wd[P] )
{
	const unsigned* base = reinterpret_cast<const unsigned*>( &sm.mask[k][0] ) + warp * P;
	if( P == 8 )
	{
		const uint4 a = *reinterpret_cast<const uint4*>( base );
		const uint4 b = *reinterpret_cast<const uint4*>( base + 4 );
		wd[0] = a.x; wd[1 % P] = a.y; wd[2 % P] = a.z; wd[3 % P] = a.w;
		wd[4 % P] = b.x; wd[5 % P] = b.y; wd[6 % P] = b.z; wd[7 % P] = b.w;
	}
	else if( P == 4 )
	{
		const uint4 a = *reinterpret_cast<const uint4*>( base );
		wd[0] = a.x; wd[1 % P] = a.y; wd[2 % P] = a.z; wd[3 % P] = a.w;
	}
	else if( P == 2 )
	{
		const uint2 a = *reinterpret_cast<const uint2*>( base );
		wd[0] = a.x; wd[1 % P] = a.y;
	}
	else
		wd[0] = base[0];
}
