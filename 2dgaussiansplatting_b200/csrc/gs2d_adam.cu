// gs2d_adam.cu — loss finalisation, Adam + clamps (main.cpp:714-785), reference init (main.cpp:280-305) and the
// AoS <-> SoA conversions at the ABI.
#include "gs2d_internal.cuh"
#include "gs2d_update.cuh"

// ------------------------------------------------------------------------------------------------------ loss ---
// One block per image: deterministic sum of the per-tile partials of this rank's band (double).
#define LOSS_THREADS 1024
__global__ void __launch_bounds__( LOSS_THREADS ) k_finalize_loss( Geometry g, const double* __restrict__ loss_partial,
                                                          double* __restrict__ loss,
                                                          const unsigned long long* __restrict__ slots,
                                                          unsigned long long* __restrict__ stats,
                                                          const DevState* __restrict__ st )
{
	__shared__ double red[LOSS_THREADS];
	const int img = blockIdx.x;
	const int t0 = ( img * g.tiles_y + g.band_ty0 ) * g.tiles_x;
	const int t1 = ( img * g.tiles_y + g.band_ty1 ) * g.tiles_x;
	// four independent partial sums per thread (loads in flight), combined in a fixed order: deterministic
	double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
	int t = t0 + threadIdx.x;
	for( ; t + 3 * LOSS_THREADS < t1; t += 4 * LOSS_THREADS )
	{
		s0 += loss_partial[t];
		s1 += loss_partial[t + LOSS_THREADS];
		s2 += loss_partial[t + 2 * LOSS_THREADS];
		s3 += loss_partial[t + 3 * LOSS_THREADS];
	}
	for( ; t < t1; t += LOSS_THREADS )
		s0 += loss_partial[t];
	red[threadIdx.x] = ( s0 + s1 ) + ( s2 + s3 );
	__syncthreads();
	for( int o = LOSS_THREADS / 2; o > 0; o >>= 1 )
	{
		if( threadIdx.x < o )
			red[threadIdx.x] += red[threadIdx.x + o];
		__syncthreads();
	}
	if( threadIdx.x == 0 )
		loss[img] = red[0];
	// overflow vote: summed over ranks together with the loss, so that every rank skips the same iterations
	if( img == 0 && threadIdx.x == 0 )
		loss[g.n_images] = st->need_capacity ? 1.0 : 0.0;
	// stats: [0] live pairs, [1] staged entries of this forward; [2],[3] their running sums; [4] sum of keys; [5] forwards
	if( img == 0 && threadIdx.x < 2 )
	{
		unsigned long long a = 0;
		for( int i = 0; i < GS2D_LIVE_SLOTS; i++ )
			a += slots[threadIdx.x * GS2D_LIVE_SLOTS + i];
		stats[threadIdx.x] = a;
		stats[2 + threadIdx.x] += a;
	}
	if( img == 0 && threadIdx.x == 2 && !st->need_capacity )
	{
		stats[4] += st->n_keys;
		stats[5] += 1ull;
	}
}

// One thread block: publish the mse values (main.cpp:805-807), advance beta powers (main.cpp:718-719) and the
// iteration counter (main.cpp:809). Runs after the (all-reduced) loss is final and before k_adam.
__global__ void k_advance( Geometry g, const double* __restrict__ loss, double* __restrict__ mse_ring, int mse_cap,
                           float beta1, float beta2, DevState* st )
{
	if( loss[g.n_images] != 0.0 )
		return; // some rank ran out of key capacity: the iteration is void everywhere (redone by the host)
	if( st->nonfinite )
	{
		// the previous iteration's update produced a non-finite parameter: the reference has aborted (main.cpp:752-785);
		// nothing advances any more (DevState::frozen)
		if( threadIdx.x == 0 )
			st->frozen = 1;
		return;
	}
	const int slot = st->mse_slot;
	for( int i = threadIdx.x; i < g.n_images; i += blockDim.x )
		if( slot < mse_cap )
			mse_ring[(size_t)slot * g.n_images + i] = loss[i] / (double)( g.H * g.W * 3 );
	__syncthreads();
	if( threadIdx.x == 0 )
	{
		st->mse_slot = slot + 1;
		st->beta1t = xmul( st->beta1t, beta1 );
		st->beta2t = xmul( st->beta2t, beta2 );
		st->iteration += 1;
	}
}

// ------------------------------------------------------------------------------------------------------ Adam ---
// adam_one, clamp_glm: gs2d_update.cuh
// One thread per (plane, 4 consecutive splats) - blockIdx.y is the plane, the nine scalars of a splat are independent
// in Adam (main.cpp:723-738) - 128-bit accesses on every SoA plane: reads p, g, m, v, writes p, m, v and zeroes g for the
// next iteration (28 + 4 bytes per scalar; HBM-bound).
__global__ void __launch_bounds__( 256 ) k_adam( Geometry g, float* __restrict__ params, float* __restrict__ grads,
                                                 float* __restrict__ adam_m, float* __restrict__ adam_v,
                                                 gs2d_params hp, const double* __restrict__ loss, DevState* st )
{
	if( loss[g.n_images] != 0.0 || st->frozen )
		return;
	const int i4 = ( blockIdx.x * 256 + threadIdx.x ) * 4;
	if( i4 >= g.n_total )
		return;
	const int pl = blockIdx.y;
	const size_t o = (size_t)pl * g.n_pad + i4;
	const bool optimised = pl != PL_OP || hp.optimize_opacity;
	const float4 p = __ldcs( reinterpret_cast<const float4*>( params + o ) );
	float pv[4] = { p.x, p.y, p.z, p.w };
	if( optimised )
	{
		const float4 gr = __ldcs( reinterpret_cast<const float4*>( grads + o ) );
		const float4 m = __ldcs( reinterpret_cast<const float4*>( adam_m + o ) );
		const float4 v = __ldcs( reinterpret_cast<const float4*>( adam_v + o ) );
		const float om_b1t = xsub( 1.0f, st->beta1t ), om_b2t = xsub( 1.0f, st->beta2t );
		float gv[4] = { gr.x, gr.y, gr.z, gr.w }, mv[4] = { m.x, m.y, m.z, m.w }, vv[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
		for( int k = 0; k < 4; k++ )
			pv[k] = adam_one( pv[k], gv[k], mv[k], vv[k], hp.lr, hp.beta1, hp.beta2, hp.eps, om_b1t, om_b2t );
		*reinterpret_cast<float4*>( adam_m + o ) = make_float4( mv[0], mv[1], mv[2], mv[3] );
		*reinterpret_cast<float4*>( adam_v + o ) = make_float4( vv[0], vv[1], vv[2], vv[3] );
	}
	// constraints, main.cpp:740-749 (the opacity clamp applies even when opacity is not optimised)
	const float wmax = xsub( (float)g.W, 1.0f ), hmax = xsub( (float)g.H, 1.0f );
	bool bad = false;
#pragma unroll
	for( int k = 0; k < 4; k++ )
	{
		float x = pv[k];
		if( pl == PL_PX )
			x = clamp_glm( x, 0.0f, wmax );
		else if( pl == PL_PY )
			x = clamp_glm( x, 0.0f, hmax );
		else if( pl == PL_SX || pl == PL_SY )
			x = clamp_glm( x, hp.s_min, hp.s_max );
		else if( pl == PL_CR || pl == PL_CG || pl == PL_CB )
			x = clamp_glm( x, 0.0f, 1.0f );
		else if( pl == PL_OP )
			x = clamp_glm( x, hp.o_min, hp.o_max );
		pv[k] = x;
		// finite check, main.cpp:752-785: colour, sx, sy, rot, pos.x (pos.y and opacity are not checked)
		if( pl != PL_PY && pl != PL_OP && i4 + k < g.n_total && !isfinite( x ) )
			bad = true;
	}
	*reinterpret_cast<float4*>( params + o ) = make_float4( pv[0], pv[1], pv[2], pv[3] );
	*reinterpret_cast<float4*>( grads + o ) = make_float4( 0.f, 0.f, 0.f, 0.f );
	if( bad )
		atomicOr( &st->nonfinite, 1 );
}

// ---------------------------------------------------------------------------------- Adam over peer memory ------
// 128-bit load that is never served from a non-coherent cache: the peers' gradient planes were written by their
// backward kernels (red.global.add in the home GPU's L2) and are read here over NVLink.
__device__ __forceinline__ float4 ld_sys4( const float* p )
{
	float4 v;
	asm volatile( "ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];"
	              : "=f"( v.x ), "=f"( v.y ), "=f"( v.z ), "=f"( v.w )
	              : "l"( p ) );
	return v;
}

// ---- flag barriers over the peers' mailboxes ---------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys( unsigned long long* p, unsigned long long v )
{
	asm volatile( "st.release.sys.global.u64 [%0], %1;" ::"l"( p ), "l"( v ) : "memory" );
}
__device__ __forceinline__ unsigned long long ld_acquire_sys( const unsigned long long* p )
{
	unsigned long long v;
	asm volatile( "ld.acquire.sys.global.u64 %0, [%1];" : "=l"( v ) : "l"( p ) : "memory" );
	return v;
}
// spin until rank r's sequence number has reached seq; gives up after ~4 s (a peer died / diverged) and reports it
__device__ __forceinline__ void wait_seq( const unsigned long long* flag, unsigned long long seq, DevState* st )
{
	const long long t0 = clock64();
	while( ld_acquire_sys( flag ) < seq )
	{
		if( clock64() - t0 > 8000000000ll )
		{
			st->peer_fault = 1;
			break;
		}
		__nanosleep( 100 );
	}
}
__device__ __forceinline__ double* mailbox_loss( unsigned char* mb, int r, int L )
{
	return reinterpret_cast<double*>( mb + sizeof( PeerMailbox ) ) + (size_t)r * L;
}

// Barrier 1: "my backward pass is complete" + all-reduce of the loss. Thread t ships loss[t] to every rank's mailbox,
// thread r publishes the sequence number to rank r and waits for rank r's; then the partial sums are added in rank
// order, so every rank computes the identical total.
// The sequence number of a barrier lives on the device (seq_ctr, one counter per barrier kind, bumped by the kernel
// itself): every rank runs the identical sequence of barrier kernels, so the counters agree without the host passing
// them in - which is what lets a multi-rank iteration be replayed as a CUDA graph (no per-launch arguments).
__global__ void __launch_bounds__( 256 ) k_peer_loss_exchange( Geometry g, PeerPtrs pp, double* __restrict__ loss,
                                                               unsigned long long* __restrict__ seq_ctr, DevState* st )
{
	const int L = g.n_images + 1, me = pp.rank, n = pp.n_ranks;
	__shared__ unsigned long long seq_s;
	if( threadIdx.x == 0 )
		seq_s = ++( *seq_ctr );
	__syncthreads();
	const unsigned long long seq = seq_s;
	for( int t = threadIdx.x; t < L; t += 256 )
	{
		const double v = loss[t];
		for( int r = 0; r < n; r++ )
			*( (volatile double*)mailbox_loss( pp.mbox[r], me, L ) + t ) = v;
	}
	__threadfence_system();
	__syncthreads();
	if( threadIdx.x < n )
	{
		PeerMailbox* theirs = reinterpret_cast<PeerMailbox*>( pp.mbox[threadIdx.x] );
		st_release_sys( &theirs->seq1[me], seq );
		const PeerMailbox* mine = reinterpret_cast<const PeerMailbox*>( pp.mbox[me] );
		wait_seq( &mine->seq1[threadIdx.x], seq, st );
	}
	__syncthreads();
	for( int t = threadIdx.x; t < L; t += 256 )
	{
		double sum = 0.0;
		for( int r = 0; r < n; r++ )
			sum += *( (volatile double*)mailbox_loss( pp.mbox[me], r, L ) + t );
		loss[t] = sum;
	}
}

// Barrier 2: "my stores into your parameter planes are complete" + OR of the non-finite flags.
__global__ void __launch_bounds__( 32 ) k_peer_barrier2( PeerPtrs pp, unsigned long long* __restrict__ seq_ctr,
                                                         DevState* st )
{
	const int me = pp.rank, n = pp.n_ranks, r = threadIdx.x;
	unsigned long long seq = 0;
	if( r == 0 )
		seq = ++( *seq_ctr );
	seq = __shfl_sync( 0xffffffffu, seq, 0 );
	int any = 0;
	if( r < n )
	{
		PeerMailbox* theirs = reinterpret_cast<PeerMailbox*>( pp.mbox[r] );
		*( (volatile int*)&theirs->nf[me] ) = st->nonfinite;
		__threadfence_system();
		st_release_sys( &theirs->seq2[me], seq );
		const PeerMailbox* mine = reinterpret_cast<const PeerMailbox*>( pp.mbox[me] );
		wait_seq( &mine->seq2[r], seq, st );
		any = *( (volatile const int*)&mine->nf[r] );
	}
	any = __any_sync( 0xffffffffu, any != 0 );
	if( r == 0 && any )
		st->nonfinite = 1;
}

// Reduce-scatter + Adam + all-gather in ONE kernel (the exchange step of the band-sharded fit): one
// thread per (plane, quad) of the slice this rank owns (blockIdx.y = plane: the planes are independent scalars, and
// nine times more threads keep nine times more NVLink loads in flight). NR = number of ranks (compile time, so that
// the peer pointers live in registers and all NR loads are issued before the first add). Gradients are summed in
// rank order (deterministic), and every splat is updated by exactly one rank, so the replicated parameters stay
// bit-identical on all ranks.
// the same through L2 only (ld.global.cg): what a remote line needs - it lives in its home GPU's L2 - without the
// system-scope ordering of ld.volatile (probe 16: is the strong load what makes the kernel slow?)
__device__ __forceinline__ float4 ld_cg4( const float* p ) { return __ldcg( reinterpret_cast<const float4*>( p ) ); }
__device__ __forceinline__ unsigned ld_sys_u32( const unsigned* p )
{
	unsigned v;
	asm volatile( "ld.volatile.global.u32 %0, [%1];" : "=r"( v ) : "l"( p ) );
	return v;
}
template <int NR>
__global__ void __launch_bounds__( 256 ) k_adam_peer( Geometry g, PeerPtrs pp, float* __restrict__ adam_m,
                                                      float* __restrict__ adam_v, gs2d_params hp,
                                                      const double* __restrict__ loss, DevState* st, int q0, int q1 )
{
	if( loss[g.n_images] != 0.0 || st->frozen )
		return;
	const int q = q0 + blockIdx.x * blockDim.x + threadIdx.x;
	if( q >= q1 )
		return;
	const int i4 = q * 4;
	const int pl = blockIdx.y;
	const size_t o = (size_t)pl * g.n_pad + i4;
	const bool optimised = pl != PL_OP || hp.optimize_opacity;
	// Everything this thread reads from its OWN memory is requested first, so that it is in flight while the two
	// dependent rounds of NVLink loads below (touch words, then gradient quads) make their way: with the loads in
	// program order behind them the kernel was a chain of four memory latencies per wave (measured: 40 us at 8 ranks
	// with every peer access switched off, where the arithmetic and the HBM traffic account for 8).
	const float* own = pp.params[pp.rank];
	const float4 p = __ldcs( reinterpret_cast<const float4*>( own + o ) );
	float4 m = make_float4( 0.f, 0.f, 0.f, 0.f ), v = m;
	if( optimised )
	{
		m = __ldcs( reinterpret_cast<const float4*>( adam_m + o ) );
		v = __ldcs( reinterpret_cast<const float4*>( adam_v + o ) );
	}
	unsigned reach_last = 0, reach_cur = 0;
	const int par = st->iteration & 1; // k_advance has run: the parity of the iteration being finished
	const size_t nq = (size_t)g.n_pad / 4;
	if( pp.sparse )
	{
		reach_cur = pp.reach_now[q];
		reach_last = pp.reach_prev[(size_t)par * nq + q];
	}
	const float om_b1t = xsub( 1.0f, st->beta1t ), om_b2t = xsub( 1.0f, st->beta2t );
	const float wmax = xsub( (float)g.W, 1.0f ), hmax = xsub( (float)g.H, 1.0f );
	// which ranks may hold a non-zero gradient for this quad (their band test passed one of its Gaussians)
	unsigned tmask = ( 1u << NR ) - 1u;
	if( pp.sparse && pp.sum_all && !( pp.probe & 4 ) )
	{
		unsigned w[NR];
#pragma unroll
		for( int r = 0; r < NR; r++ )
			w[r] = ld_sys_u32( pp.touch[r] + ( q >> 5 ) ); // one word per 32 consecutive quads: a broadcast per warp
		tmask = 0;
#pragma unroll
		for( int r = 0; r < NR; r++ )
			tmask |= ( ( w[r] >> ( q & 31 ) ) & 1u ) << r;
	}
	float gv[4] = { 0.f, 0.f, 0.f, 0.f };
	if( optimised )
	{
		if( pp.sum_all )
		{
			float4 part[NR];
#pragma unroll
			for( int r = 0; r < NR; r++ )
				part[r] = ( ( ( tmask >> r ) & 1u ) && !( ( pp.probe & 1 ) && r != pp.rank ) )
				              ? ( ( pp.probe & 16 ) ? ld_cg4( pp.grads[r] + o ) : ld_sys4( pp.grads[r] + o ) )
				              : make_float4( 0.f, 0.f, 0.f, 0.f ); // in flight together
#pragma unroll
			for( int r = 0; r < NR; r++ )
			{
				gv[0] = xadd( gv[0], part[r].x );
				gv[1] = xadd( gv[1], part[r].y );
				gv[2] = xadd( gv[2], part[r].z );
				gv[3] = xadd( gv[3], part[r].w );
			}
		}
		else
		{
			const float4 t = ld_sys4( pp.grads[pp.rank] + o );
			gv[0] = t.x, gv[1] = t.y, gv[2] = t.z, gv[3] = t.w;
		}
	}
	float pv[4] = { p.x, p.y, p.z, p.w };
	if( optimised )
	{
		float mv[4] = { m.x, m.y, m.z, m.w }, vv[4] = { v.x, v.y, v.z, v.w };
		if( !( pp.probe & 8 ) )
		{
#pragma unroll
			for( int k = 0; k < 4; k++ )
				pv[k] = adam_one( pv[k], gv[k], mv[k], vv[k], hp.lr, hp.beta1, hp.beta2, hp.eps, om_b1t, om_b2t );
		}
		__stcs( reinterpret_cast<float4*>( adam_m + o ), make_float4( mv[0], mv[1], mv[2], mv[3] ) );
		__stcs( reinterpret_cast<float4*>( adam_v + o ), make_float4( vv[0], vv[1], vv[2], vv[3] ) );
	}
	bool bad = false;
#pragma unroll
	for( int k = 0; k < 4; k++ )
	{
		float x = pv[k];
		if( pl == PL_PX )
			x = clamp_glm( x, 0.0f, wmax );
		else if( pl == PL_PY )
			x = clamp_glm( x, 0.0f, hmax );
		else if( pl == PL_SX || pl == PL_SY )
			x = clamp_glm( x, hp.s_min, hp.s_max );
		else if( pl == PL_CR || pl == PL_CG || pl == PL_CB )
			x = clamp_glm( x, 0.0f, 1.0f );
		else if( pl == PL_OP )
			x = clamp_glm( x, hp.o_min, hp.o_max );
		pv[k] = x;
		if( pl != PL_PY && pl != PL_OP && i4 + k < g.n_total && !isfinite( x ) )
			bad = true;
	}
	const float4 out = make_float4( pv[0], pv[1], pv[2], pv[3] );
	// a parameter that cannot have changed (opacity while it is not optimised and already inside its clamp) is not sent
	const bool changed = optimised || out.x != p.x || out.y != p.y || out.z != p.z || out.w != p.w;
	// whom to send it to: everyone (dense), or the ranks a Gaussian of the quad reaches now or reached last time
	unsigned send = ( 1u << NR ) - 1u;
	if( pp.sparse )
	{
		send = reach_cur | reach_last | ( 1u << pp.rank );
		if( pl == 0 )
			pp.reach_prev[(size_t)( par ^ 1 ) * nq + q] = (unsigned char)reach_cur;
	}
	if( changed )
	{
#pragma unroll
		for( int r = 0; r < NR; r++ )
			if( ( ( send >> r ) & 1u ) && !( ( pp.probe & 2 ) && r != pp.rank ) )
				*reinterpret_cast<float4*>( pp.params[r] + o ) = out;
	}
	if( pp.traffic )
	{
		// NVLink quads of this warp: gradient quads read from peers, parameter quads written to peers
		const unsigned peers = ( ( 1u << NR ) - 1u ) & ~( 1u << pp.rank );
		unsigned nr = ( optimised && pp.sum_all ) ? __popc( tmask & peers ) : 0u, nw = changed ? __popc( send & peers ) : 0u;
		const unsigned act = __activemask(); // the last warp of the slice may be partial
		nr = __reduce_add_sync( act, nr );
		nw = __reduce_add_sync( act, nw );
		if( ( threadIdx.x & 31 ) == __ffs( act ) - 1 )
		{
			atomicAdd( pp.traffic, (unsigned long long)nr );
			atomicAdd( pp.traffic + 1, (unsigned long long)nw );
		}
	}
	if( bad )
		atomicOr( &st->nonfinite, 1 );
	// No fence here (a system-scope fence per thread costs a third of the kernel): the stores into the peers' planes
	// are complete when the grid is, and the barrier kernel that follows on the stream issues a cumulative
	// __threadfence_system() before it publishes this rank's sequence number.
}

// ------------------------------------------------------------------------------------------------------ init ---
// pcg3d, mix_glm, init_splat: gs2d_update.cuh
__global__ void __launch_bounds__( 256 ) k_init_reference( Geometry g, float* __restrict__ params )
{
	const int gi = blockIdx.x * 256 + threadIdx.x;
	if( gi >= g.n_total )
		return;
	float v[GS2D_NPLANES];
	init_splat( (unsigned)( gi % g.n_splats ), g.W, g.H, v );
#pragma unroll
	for( int pl = 0; pl < GS2D_NPLANES; pl++ )
		params[pl * g.n_pad + gi] = v[pl];
}

// Image2DRGBA8_to_Image2DRGBA32: vec4( u8 ) / vec4( 255.0f ), prp.cpp:796-807
__global__ void __launch_bounds__( 256 ) k_u8_to_target( const uchar4* __restrict__ src, float4* __restrict__ dst,
                                                         size_t n )
{
	const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
	if( i >= n )
		return;
	const uchar4 c = src[i];
	dst[i] = make_float4( xdiv( (float)c.x, 255.0f ), xdiv( (float)c.y, 255.0f ), xdiv( (float)c.z, 255.0f ),
	                      xdiv( (float)c.w, 255.0f ) );
}

// ------------------------------------------------------------------------------------------- AoS <-> SoA -------
// Splat field f (0..8) <-> plane f; SplatAdam stores {m,v} pairs in the same field order (main.cpp:158-166).
__global__ void __launch_bounds__( 256 ) k_aos_to_soa( Geometry g, int image, const float* __restrict__ splats,
                                                       const float* __restrict__ adam, float* __restrict__ params,
                                                       float* __restrict__ adam_m, float* __restrict__ adam_v )
{
	const int i = blockIdx.x * 256 + threadIdx.x;
	if( i >= g.n_splats )
		return;
	const size_t gi = (size_t)image * g.n_splats + i;
#pragma unroll
	for( int f = 0; f < GS2D_NPLANES; f++ )
	{
		params[(size_t)f * g.n_pad + gi] = splats[(size_t)i * 9 + f];
		adam_m[(size_t)f * g.n_pad + gi] = adam ? adam[(size_t)i * 18 + 2 * f] : 0.0f;
		adam_v[(size_t)f * g.n_pad + gi] = adam ? adam[(size_t)i * 18 + 2 * f + 1] : 0.0f;
	}
}
__global__ void __launch_bounds__( 256 ) k_soa_to_aos( Geometry g, int image, float* __restrict__ splats,
                                                       float* __restrict__ adam, float* __restrict__ gsplats,
                                                       const float* __restrict__ params,
                                                       const float* __restrict__ adam_m,
                                                       const float* __restrict__ adam_v,
                                                       const float* __restrict__ grads )
{
	const int i = blockIdx.x * 256 + threadIdx.x;
	if( i >= g.n_splats )
		return;
	const size_t gi = (size_t)image * g.n_splats + i;
#pragma unroll
	for( int f = 0; f < GS2D_NPLANES; f++ )
	{
		if( splats )
			splats[(size_t)i * 9 + f] = params[(size_t)f * g.n_pad + gi];
		if( adam )
		{
			adam[(size_t)i * 18 + 2 * f] = adam_m[(size_t)f * g.n_pad + gi];
			adam[(size_t)i * 18 + 2 * f + 1] = adam_v[(size_t)f * g.n_pad + gi];
		}
		if( gsplats )
			gsplats[(size_t)i * 9 + f] = grads[(size_t)f * g.n_pad + gi];
	}
}

// ---------------------------------------------------------------------------------------------------- launch ---
int launch_init_reference( const Geometry& g, const Buffers& b, cudaStream_t s )
{
	const size_t plane_bytes = sizeof( float ) * (size_t)g.n_pad * GS2D_NPLANES;
	cudaMemsetAsync( b.params, 0, plane_bytes, s );
	cudaMemsetAsync( b.grads, 0, plane_bytes, s );
	cudaMemsetAsync( b.adam_m, 0, plane_bytes, s );
	cudaMemsetAsync( b.adam_v, 0, plane_bytes, s );
	if( g.n_total == 0 )
		return 0;
	k_init_reference<<<( g.n_total + 255 ) / 256, 256, 0, s>>>( g, b.params );
	return 1;
}
__global__ void k_u8_lut( float* __restrict__ lut ) { lut[threadIdx.x] = xdiv( (float)threadIdx.x, 255.0f ); }
int launch_u8_lut( float* lut, cudaStream_t s )
{
	k_u8_lut<<<1, 256, 0, s>>>( lut );
	return 1;
}
int launch_u8_to_target( const uint8_t* src_dev, float4* dst, size_t n_px, cudaStream_t s )
{
	k_u8_to_target<<<(unsigned)( ( n_px + 255 ) / 256 ), 256, 0, s>>>( reinterpret_cast<const uchar4*>( src_dev ), dst,
	                                                                   n_px );
	return 1;
}
// gs2d_mse_ticket: the loop-carried scalars and the loss of the newest iteration, stored straight into a slot of
// mapped page-locked host memory (no copy engine involved: a small D2H copy can queue behind a large target upload).
__global__ void k_publish( const DevState* __restrict__ st, const double* __restrict__ loss, int n_images,
                           unsigned* __restrict__ slot_state, double* __restrict__ slot_loss )
{
	const unsigned* src = reinterpret_cast<const unsigned*>( st );
	for( int i = threadIdx.x; i < (int)( sizeof( DevState ) / sizeof( unsigned ) ); i += blockDim.x )
		slot_state[i] = src[i];
	for( int i = threadIdx.x; i < n_images; i += blockDim.x )
		slot_loss[i] = loss[i];
	__threadfence_system();
}
int launch_publish( const Buffers& b, int n_images, unsigned char* slot, cudaStream_t s )
{
	static_assert( sizeof( DevState ) % sizeof( double ) == 0, "the loss follows the state in a ticket slot" );
	k_publish<<<1, 64, 0, s>>>( b.state, b.loss, n_images, reinterpret_cast<unsigned*>( slot ),
	                            reinterpret_cast<double*>( slot + sizeof( DevState ) ) );
	return 1;
}
int launch_finalize_loss( const Geometry& g, const Buffers& b, cudaStream_t s )
{
	k_finalize_loss<<<g.n_images, LOSS_THREADS, 0, s>>>( g, b.loss_partial, b.loss, b.live_slots, b.stats, b.state );
	return 1;
}
int launch_advance( const Geometry& g, const Buffers& b, const gs2d_params& p, cudaStream_t s )
{
	k_advance<<<1, 256, 0, s>>>( g, b.loss, b.mse_ring, b.mse_cap, p.beta1, p.beta2, b.state );
	return 1;
}
int launch_adam( const Geometry& g, const Buffers& b, const gs2d_params& p, cudaStream_t s )
{
	if( g.n_total == 0 )
		return 0;
	const int n4 = ( g.n_total + 3 ) / 4;
	k_adam<<<dim3( ( n4 + 255 ) / 256, GS2D_NPLANES ), 256, 0, s>>>( g, b.params, b.grads, b.adam_m, b.adam_v, p, b.loss,
	                                                                 b.state );
	return 1;
}
int launch_adam_peer( const Geometry& g, const Buffers& b, const gs2d_params& p, const PeerPtrs& pp, int q0, int q1,
                      cudaStream_t s )
{
	if( q1 <= q0 )
		return 0;
	const int bs = ( pp.probe & 32 ) ? 256 : 128;
	const dim3 grid( ( q1 - q0 + bs - 1 ) / bs, GS2D_NPLANES );
#define PEER_CASE( NR )                                                                                               \
	case NR: k_adam_peer<NR><<<grid, bs, 0, s>>>( g, pp, b.adam_m, b.adam_v, p, b.loss, b.state, q0, q1 ); break;
	switch( pp.n_ranks )
	{
		PEER_CASE( 1 ) PEER_CASE( 2 ) PEER_CASE( 3 ) PEER_CASE( 4 ) PEER_CASE( 5 ) PEER_CASE( 6 ) PEER_CASE( 7 ) PEER_CASE( 8 )
	default: return 0;
	}
#undef PEER_CASE
	return 1;
}
int launch_peer_loss_exchange( const Geometry& g, const Buffers& b, const PeerPtrs& pp, unsigned long long* seq_ctr,
                               cudaStream_t s )
{
	k_peer_loss_exchange<<<1, 256, 0, s>>>( g, pp, b.loss, seq_ctr, b.state );
	return 1;
}
int launch_peer_barrier2( const Geometry& g, const Buffers& b, const PeerPtrs& pp, unsigned long long* seq_ctr,
                          cudaStream_t s )
{
	k_peer_barrier2<<<1, 32, 0, s>>>( pp, seq_ctr, b.state );
	return 1;
}
int launch_aos_to_soa( const Geometry& g, const Buffers& b, int image, const gs2d_splat* splats_dev,
                       const gs2d_splat_adam* adam_dev, cudaStream_t s )
{
	if( g.n_splats == 0 )
		return 0;
	k_aos_to_soa<<<( g.n_splats + 255 ) / 256, 256, 0, s>>>( g, image, reinterpret_cast<const float*>( splats_dev ),
	                                                         reinterpret_cast<const float*>( adam_dev ), b.params,
	                                                         b.adam_m, b.adam_v );
	return 1;
}
int launch_soa_to_aos( const Geometry& g, const Buffers& b, int image, gs2d_splat* splats_dev,
                       gs2d_splat_adam* adam_dev, gs2d_splat* grads_dev, cudaStream_t s )
{
	if( g.n_splats == 0 )
		return 0;
	k_soa_to_aos<<<( g.n_splats + 255 ) / 256, 256, 0, s>>>( g, image, reinterpret_cast<float*>( splats_dev ),
	                                                         reinterpret_cast<float*>( adam_dev ),
	                                                         reinterpret_cast<float*>( grads_dev ), b.params, b.adam_m,
	                                                         b.adam_v, b.grads );
	return 1;
}
