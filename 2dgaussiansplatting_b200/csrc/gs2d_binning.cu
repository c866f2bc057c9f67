// gs2d_binning.cu — per-Gaussian preprocessing, tile binning and the on-device radix sort.
//
// Replaces, for a pixel-parallel GPU pass, what the reference does implicitly by iterating splats in index order
// and scan-converting each one (main.cpp:419-538): here every 16x16 tile gets the list of splats whose footprint
// may touch it, ordered by splat index (== the reference's blend order).
//
//   k_preprocess   : covariance, inverse, row range (cov_of main.cpp:206-221, :432-436, :489-491), conservative tile
//                    rectangle, per-splat tile count + block-local exclusive scan of the counts
//   k_scan_top     : scan of the block sums, total key count, capacity check
//   k_emit         : duplicated (tile | splat) keys, written in ascending splat order
//   k_radix_*      : stable LSD radix passes over the TILE bits only (8 bits per pass). Because the emission is
//                    already ordered by splat index and every pass is stable, the result is ordered by
//                    (tile, splat index) after ceil(log2(tiles)/8) passes instead of the ~5 a full 64-bit key needs.
//   k_tile_ranges  : [start,end) of every tile's list
#include "gs2d_internal.cuh"

#define RADIX_BITS 8
#define RADIX_BINS 256
#define RADIX_THREADS 256
#define RADIX_ITEMS 16
#define RADIX_CHUNK ( RADIX_THREADS * RADIX_ITEMS )

#define PRE_THREADS 256

// ------------------------------------------------------------------------------------------------ preprocess ---
__global__ void __launch_bounds__( PRE_THREADS ) k_preprocess( Geometry g, const float* __restrict__ params,
                                                               SplatRec* __restrict__ rec, ushort4* __restrict__ rect,
                                                               unsigned* __restrict__ cnt_local,
                                                               unsigned* __restrict__ block_sum )
{
	const int gi = blockIdx.x * PRE_THREADS + threadIdx.x;
	unsigned cnt = 0;
	if( gi < g.n_total )
	{
		const float px = params[PL_PX * g.n_pad + gi];
		const float py = params[PL_PY * g.n_pad + gi];
		const float sx = params[PL_SX * g.n_pad + gi];
		const float sy = params[PL_SY * g.n_pad + gi];
		const float rot = params[PL_ROT * g.n_pad + gi];

		// cosf/sinf of the reference are glibc's (< 1 ulp, almost always correctly rounded). Evaluate in double and
		// round once: agrees with it bit for bit except in rare double-rounding cases.
		double sd, cd;
		sincos( (double)rot, &sd, &cd );
		const float c = (float)cd, s = (float)sd;

		// cov_of, main.cpp:214-220
		const float l0 = xmul( sx, sx );
		const float l1 = xmul( sy, sy );
		const float s11 = xadd( xmul( xmul( l0, c ), c ), xmul( xmul( l1, s ), s ) );
		const float s12 = xmul( xmul( xsub( l0, l1 ), s ), c );
		const float s22 = xsub( xadd( l0, l1 ), s11 );
		const float det = xsub( xmul( s11, s22 ), xmul( s12, s12 ) ); // main.cpp:191, :560
		// inv_cov = mat2( cov[1][1], -cov[0][1], -cov[1][0], cov[0][0] ) / det, main.cpp:432-436
		const float a = xdiv( s22, det );
		const float b = xdiv( -s12, det );
		const float d = xdiv( s11, det );
		// row range, main.cpp:489-491
		const float h = xmul( xsqrt( xmul( a, det ) ), 3.0f );
		const int begY = trunc_x86( xsub( py, h ) );
		const int endY = trunc_x86( xadd( py, h ) );

		SplatRec r;
		r.px = px;
		r.py = py;
		r.a = a;
		r.b = b;
		r.d = d;
		r.cr = params[PL_CR * g.n_pad + gi];
		r.cg = params[PL_CG * g.n_pad + gi];
		r.cb = params[PL_CB * g.n_pad + gi];
		r.op = params[PL_OP * g.n_pad + gi];
		r.begY = begY;
		r.endY = endY;
		// hoisted backward invariants, main.cpp:658, :661, :682 (same operand order)
		r.rotc = xdiv( xsub( xmul( sx, sx ), xmul( sy, sy ) ), xmul( xmul( xmul( sx, sx ), sy ), sy ) );
		r.cosT = c;
		r.sinT = s;
		r.isx3 = xdiv( 1.0f, xmul( xmul( sx, sx ), sx ) );
		r.isy3 = xdiv( 1.0f, xmul( xmul( sy, sy ), sy ) );
		float4* dst = reinterpret_cast<float4*>( rec + gi );
		const float4* src = reinterpret_cast<const float4*>( &r );
		dst[0] = src[0];
		dst[1] = src[1];
		dst[2] = src[2];
		dst[3] = src[3];

		// Conservative tile rectangle. Rows: exactly the reference's [begY, endY]. Columns: every chord of the 3-sigma
		// ellipse lies within pos.x +- 3*sqrt(cov[0][0]) = 3*sqrt(d*det); widened by 2 px + 1e-4 relative for the
		// truncation rule and the float error of the quadratic solve. Only a superset is needed here: the exact
		// per-pixel set is decided by row_mask() in the blend kernels.
		const int H = g.H, W = g.W;
		int y0 = max( begY, 0 ), y1 = min( endY, H - 1 );
		float hx = xmul( xsqrt( xmul( d, det ) ), 3.0f );
		float m = 2.0f + hx * 1.0e-4f;
		float fx0 = floorf( px - hx - m ), fx1 = floorf( px + hx + m );
		int x0 = ( fx0 == fx0 ) ? (int)fmaxf( fx0, 0.0f ) : 0;
		int x1 = ( fx1 == fx1 ) ? (int)fminf( fx1, (float)( W - 1 ) ) : W - 1;
		if( fx1 < 0.0f )
			x1 = -1;
		ushort4 rc = make_ushort4( 1, 1, 0, 0 ); // empty rectangle (width 0)
		if( y0 <= y1 && x0 <= x1 )
		{
			int ty0 = max( y0 / GS2D_TILE, g.band_ty0 ), ty1 = min( y1 / GS2D_TILE, g.band_ty1 - 1 );
			int tx0 = x0 / GS2D_TILE, tx1 = x1 / GS2D_TILE;
			if( ty0 <= ty1 )
			{
				cnt = (unsigned)( ( ty1 - ty0 + 1 ) * ( tx1 - tx0 + 1 ) );
				rc = make_ushort4( (unsigned short)tx0, (unsigned short)ty0, (unsigned short)tx1, (unsigned short)ty1 );
			}
		}
		rect[gi] = rc;
	}

	// block-local exclusive scan of cnt (warp shuffles, then the 8 warp totals)
	__shared__ unsigned warp_tot[PRE_THREADS / 32];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	unsigned incl = cnt;
#pragma unroll
	for( int o = 1; o < 32; o <<= 1 )
	{
		unsigned v = __shfl_up_sync( 0xffffffffu, incl, o );
		if( lane >= o )
			incl += v;
	}
	if( lane == 31 )
		warp_tot[warp] = incl;
	__syncthreads();
	unsigned wbase = 0, total = 0;
#pragma unroll
	for( int w = 0; w < PRE_THREADS / 32; w++ )
	{
		unsigned t = warp_tot[w];
		if( w < warp )
			wbase += t;
		total += t;
	}
	if( gi < g.n_total )
		cnt_local[gi] = wbase + incl - cnt; // exclusive, block-local
	if( threadIdx.x == 0 )
		block_sum[blockIdx.x] = total;
}

// One block: exclusive scan of the block sums (u64 accumulation), total, capacity check.
__global__ void __launch_bounds__( 1024 ) k_scan_top( const unsigned* __restrict__ block_sum, int nb,
                                                      unsigned long long* __restrict__ block_base,
                                                      unsigned long long capacity, DevState* st )
{
	__shared__ unsigned long long wtot[32];
	__shared__ unsigned long long carry_s;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	if( threadIdx.x == 0 )
		carry_s = 0;
	__syncthreads();
	for( int base = 0; base < nb; base += 1024 )
	{
		int i = base + threadIdx.x;
		unsigned long long v = ( i < nb ) ? block_sum[i] : 0ull, incl = v;
#pragma unroll
		for( int o = 1; o < 32; o <<= 1 )
		{
			unsigned long long t = __shfl_up_sync( 0xffffffffu, incl, o );
			if( lane >= o )
				incl += t;
		}
		if( lane == 31 )
			wtot[warp] = incl;
		__syncthreads();
		unsigned long long wb = 0, tot = 0;
		for( int w = 0; w < 32; w++ )
		{
			unsigned long long t = wtot[w];
			if( w < warp )
				wb += t;
			tot += t;
		}
		unsigned long long carry = carry_s;
		if( i < nb )
			block_base[i] = carry + wb + incl - v;
		__syncthreads();
		if( threadIdx.x == 0 )
			carry_s = carry + tot;
		__syncthreads();
	}
	if( threadIdx.x == 0 )
	{
		unsigned long long total = carry_s;
		st->n_keys = total;
		st->need_capacity = ( total > capacity ) ? total : 0ull;
	}
}

// One warp per 32 splats; for each of them the 32 lanes write its tile keys side by side (coalesced).
__global__ void __launch_bounds__( 256 ) k_emit( Geometry g, const ushort4* __restrict__ rect,
                                                 const unsigned* __restrict__ cnt_local,
                                                 const unsigned long long* __restrict__ block_base,
                                                 unsigned* __restrict__ keys, unsigned* __restrict__ vals,
                                                 const DevState* __restrict__ st )
{
	if( st->need_capacity )
		return;
	const int lane = threadIdx.x & 31;
	const int gi = ( blockIdx.x * 256 + threadIdx.x );
	ushort4 rc = make_ushort4( 1, 1, 0, 0 );
	unsigned long long off = 0;
	if( gi < g.n_total )
	{
		rc = rect[gi];
		off = block_base[gi / PRE_THREADS] + cnt_local[gi];
	}
	const int w = (int)rc.z - (int)rc.x + 1, h = (int)rc.w - (int)rc.y + 1; // empty rect is stored as (1,1,0,0): w = 0
	const unsigned my_cnt = ( w > 0 && h > 0 ) ? (unsigned)( w * h ) : 0u;
	const int img_tile_base = ( gi < g.n_total ) ? ( gi / g.n_splats ) * g.tiles_y : 0;
#pragma unroll 1
	for( int j = 0; j < 32; j++ )
	{
		unsigned c = __shfl_sync( 0xffffffffu, my_cnt, j );
		if( c == 0 )
			continue;
		unsigned long long o = __shfl_sync( 0xffffffffu, off, j );
		int jw = __shfl_sync( 0xffffffffu, w, j );
		int jx0 = __shfl_sync( 0xffffffffu, (int)rc.x, j );
		int jy0 = __shfl_sync( 0xffffffffu, (int)rc.y, j );
		int jbase = __shfl_sync( 0xffffffffu, img_tile_base, j );
		int jg = __shfl_sync( 0xffffffffu, gi, j );
		for( unsigned e = lane; e < c; e += 32 )
		{
			int ty = (int)( e / (unsigned)jw ), tx = (int)( e - (unsigned)ty * (unsigned)jw );
			unsigned tile = (unsigned)( ( jbase + jy0 + ty ) * g.tiles_x + jx0 + tx );
			keys[o + e] = tile;
			vals[o + e] = (unsigned)jg;
		}
	}
}

// ------------------------------------------------------------------------------------------------ radix sort ---
__global__ void __launch_bounds__( RADIX_THREADS ) k_radix_hist( const unsigned* __restrict__ keys, int shift,
                                                                 unsigned* __restrict__ hist, unsigned max_chunks,
                                                                 const DevState* __restrict__ st )
{
	__shared__ unsigned h[RADIX_BINS];
	const unsigned long long n = st->need_capacity ? 0ull : st->n_keys;
	const unsigned long long base = (unsigned long long)blockIdx.x * RADIX_CHUNK;
	h[threadIdx.x] = 0;
	__syncthreads();
	if( base < n )
	{
#pragma unroll
		for( int i = 0; i < RADIX_ITEMS; i++ )
		{
			unsigned long long idx = base + (unsigned)( i * RADIX_THREADS ) + threadIdx.x;
			if( idx < n )
				atomicAdd( &h[( keys[idx] >> shift ) & ( RADIX_BINS - 1 )], 1u );
		}
	}
	__syncthreads();
	hist[threadIdx.x * max_chunks + blockIdx.x] = h[threadIdx.x];
}

// generic in-place exclusive scan of a u32 array (three launches)
#define SCAN_THREADS 256
#define SCAN_PER_BLOCK 1024
__global__ void __launch_bounds__( SCAN_THREADS ) k_scan_local( unsigned* __restrict__ a, unsigned n,
                                                                unsigned* __restrict__ sums )
{
	__shared__ unsigned wtot[SCAN_THREADS / 32];
	const unsigned i0 = blockIdx.x * SCAN_PER_BLOCK + threadIdx.x * 4;
	unsigned v[4];
#pragma unroll
	for( int k = 0; k < 4; k++ )
		v[k] = ( i0 + k < n ) ? a[i0 + k] : 0u;
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
		if( i0 + k < n )
			a[i0 + k] = run;
		run += v[k];
	}
	if( threadIdx.x == 0 )
		sums[blockIdx.x] = tot;
}
__global__ void __launch_bounds__( 1024 ) k_scan_sums( unsigned* __restrict__ sums, unsigned nb )
{
	__shared__ unsigned wtot[32];
	__shared__ unsigned carry_s;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	if( threadIdx.x == 0 )
		carry_s = 0;
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
}
__global__ void __launch_bounds__( SCAN_THREADS ) k_scan_add( unsigned* __restrict__ a, unsigned n,
                                                              const unsigned* __restrict__ sums )
{
	const unsigned add = sums[blockIdx.x];
	const unsigned i0 = blockIdx.x * SCAN_PER_BLOCK + threadIdx.x * 4;
#pragma unroll
	for( int k = 0; k < 4; k++ )
		if( i0 + k < n )
			a[i0 + k] += add;
}

// Stable scatter of one chunk. Warp w owns keys [base + 512 w, base + 512 (w+1)) and walks them in 16 rounds of 32
// consecutive keys; match.any gives every key its rank among equal digits of the round, a per-warp digit counter
// carries the rank across rounds, an exclusive scan over the 8 warps (+ the scanned global histogram) the rest.
__global__ void __launch_bounds__( RADIX_THREADS ) k_radix_scatter(
    const unsigned* __restrict__ keys_in, const unsigned* __restrict__ vals_in, unsigned* __restrict__ keys_out,
    unsigned* __restrict__ vals_out, const unsigned* __restrict__ hist_scanned, unsigned max_chunks, int shift,
    const DevState* __restrict__ st )
{
	__shared__ unsigned wcnt[RADIX_THREADS / 32][RADIX_BINS];
	const unsigned long long n = st->need_capacity ? 0ull : st->n_keys;
	const unsigned long long base = (unsigned long long)blockIdx.x * RADIX_CHUNK;
	if( base >= n )
		return;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
	for( int w = 0; w < RADIX_THREADS / 32; w++ )
		wcnt[w][threadIdx.x] = 0;
	__syncthreads();

	const unsigned long long seg = base + (unsigned)( warp * ( RADIX_CHUNK / ( RADIX_THREADS / 32 ) ) );
	unsigned key[RADIX_ITEMS];
	unsigned short pos[RADIX_ITEMS];
	const unsigned lt = ( 1u << lane ) - 1u;
#pragma unroll
	for( int r = 0; r < RADIX_ITEMS; r++ )
	{
		unsigned long long idx = seg + (unsigned)( r * 32 ) + lane;
		bool valid = idx < n;
		unsigned k = valid ? keys_in[idx] : 0xffffffffu;
		unsigned dg = valid ? ( ( k >> shift ) & ( RADIX_BINS - 1 ) ) : RADIX_BINS;
		unsigned peers = __match_any_sync( 0xffffffffu, dg );
		unsigned pre = valid ? wcnt[warp][dg] : 0u;
		__syncwarp();
		if( valid && ( peers & lt ) == 0u )
			wcnt[warp][dg] = pre + __popc( peers );
		__syncwarp();
		key[r] = k;
		pos[r] = (unsigned short)( pre + __popc( peers & lt ) );
	}
	__syncthreads();
	{
		unsigned run = hist_scanned[threadIdx.x * max_chunks + blockIdx.x];
#pragma unroll
		for( int w = 0; w < RADIX_THREADS / 32; w++ )
		{
			unsigned c = wcnt[w][threadIdx.x];
			wcnt[w][threadIdx.x] = run;
			run += c;
		}
	}
	__syncthreads();
#pragma unroll
	for( int r = 0; r < RADIX_ITEMS; r++ )
	{
		unsigned long long idx = seg + (unsigned)( r * 32 ) + lane;
		if( idx < n )
		{
			unsigned dg = ( key[r] >> shift ) & ( RADIX_BINS - 1 );
			unsigned dst = wcnt[warp][dg] + pos[r];
			keys_out[dst] = key[r];
			vals_out[dst] = vals_in[idx];
		}
	}
}

__global__ void __launch_bounds__( 256 ) k_tile_ranges( const unsigned* __restrict__ keys,
                                                        uint2* __restrict__ tile_range,
                                                        const DevState* __restrict__ st )
{
	const unsigned long long n = st->need_capacity ? 0ull : st->n_keys;
	const unsigned long long i = (unsigned long long)blockIdx.x * 256 + threadIdx.x;
	if( i >= n )
		return;
	const unsigned k = keys[i];
	if( i == 0 || keys[i - 1] != k )
		tile_range[k].x = (unsigned)i;
	if( i + 1 == n || keys[i + 1] != k )
		tile_range[k].y = (unsigned)( i + 1 );
}

// ---------------------------------------------------------------------------------------------------- launch ---
int launch_preprocess_and_bin( const Geometry& g, const Buffers& b, int* sorted_buf, cudaStream_t s,
                               cudaEvent_t after_emit )
{
	int launches = 0;
	if( g.n_total == 0 )
	{
		cudaMemsetAsync( b.tile_range, 0, sizeof( uint2 ) * (size_t)g.n_tiles_total, s );
		cudaMemsetAsync( &b.state->n_keys, 0, 2 * sizeof( unsigned long long ), s );
		*sorted_buf = 0;
		if( after_emit )
			cudaEventRecord( after_emit, s );
		return 0;
	}
	const int nb = ( g.n_total + PRE_THREADS - 1 ) / PRE_THREADS;
	unsigned* block_sum = reinterpret_cast<unsigned*>( b.scan_tmp );         // [nb] u32
	unsigned long long* block_base = b.scan_tmp + ( ( nb + 1 ) / 2 + 1 );      // [nb] u64
	k_preprocess<<<nb, PRE_THREADS, 0, s>>>( g, b.params, b.rec, b.rect, b.cnt, block_sum );
	k_scan_top<<<1, 1024, 0, s>>>( block_sum, nb, block_base, b.capacity, b.state );
	k_emit<<<nb, 256, 0, s>>>( g, b.rect, b.cnt, block_base, b.keys[0], b.vals[0], b.state );
	launches += 3;
	if( after_emit )
		cudaEventRecord( after_emit, s );

	int bits = 0;
	while( ( 1ll << bits ) < (long long)g.n_tiles_total )
		bits++;
	const int passes = ( bits + RADIX_BITS - 1 ) / RADIX_BITS;
	const unsigned chunks = b.max_chunks;
	const unsigned hist_n = RADIX_BINS * chunks;
	const unsigned scan_nb = ( hist_n + SCAN_PER_BLOCK - 1 ) / SCAN_PER_BLOCK;
	unsigned* scan_sums = b.radix_hist + hist_n;
	int cur = 0;
	for( int p = 0; p < passes; p++ )
	{
		const int shift = p * RADIX_BITS;
		k_radix_hist<<<chunks, RADIX_THREADS, 0, s>>>( b.keys[cur], shift, b.radix_hist, chunks, b.state );
		k_scan_local<<<scan_nb, SCAN_THREADS, 0, s>>>( b.radix_hist, hist_n, scan_sums );
		k_scan_sums<<<1, 1024, 0, s>>>( scan_sums, scan_nb );
		k_scan_add<<<scan_nb, SCAN_THREADS, 0, s>>>( b.radix_hist, hist_n, scan_sums );
		k_radix_scatter<<<chunks, RADIX_THREADS, 0, s>>>( b.keys[cur], b.vals[cur], b.keys[cur ^ 1], b.vals[cur ^ 1],
		                                                  b.radix_hist, chunks, shift, b.state );
		launches += 5;
		cur ^= 1;
	}
	cudaMemsetAsync( b.tile_range, 0, sizeof( uint2 ) * (size_t)g.n_tiles_total, s );
	const unsigned long long cap_blocks = ( b.capacity + 255 ) / 256;
	k_tile_ranges<<<(unsigned)cap_blocks, 256, 0, s>>>( b.keys[cur], b.tile_range, b.state );
	launches += 1;
	*sorted_buf = cur;
	return launches;
}
