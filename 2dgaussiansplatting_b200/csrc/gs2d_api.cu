// gs2d_api.cu — host side of the C ABI (include/gs2d.h): context, device memory, the per-iteration launch sequence,
// state upload/download in the reference's AoS layouts, NCCL (dlopen'ed) for band-sharded multi-GPU fits.
//
// There is no CPU implementation of anything here: without an sm_100 device gs2d_create fails.
#include "gs2d_internal.cuh"

#include <dlfcn.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <string>
#include <vector>

// ---- minimal NCCL surface, resolved at run time (only needed when gs2d_comm_init is called) --------------------
typedef struct ncclComm* ncclComm_t;
typedef struct
{
	char internal[128];
} ncclUniqueId_;
struct NcclApi
{
	void* handle = nullptr;
	int ( *GetUniqueId )( ncclUniqueId_* ) = nullptr;
	int ( *CommInitRank )( ncclComm_t*, int, ncclUniqueId_, int ) = nullptr;
	int ( *CommDestroy )( ncclComm_t ) = nullptr;
	int ( *AllReduce )( const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t ) = nullptr;
	int ( *AllGather )( const void*, void*, size_t, int, ncclComm_t, cudaStream_t ) = nullptr;
	int ( *GroupStart )() = nullptr;
	int ( *GroupEnd )() = nullptr;
	const char* ( *GetErrorString )( int ) = nullptr;
};
static NcclApi g_nccl;
static bool nccl_load()
{
	if( g_nccl.handle )
		return true;
	const char* names[] = { "libnccl.so.2", "libnccl.so" };
	void* h = nullptr;
	for( const char* n : names )
		if( ( h = dlopen( n, RTLD_NOW | RTLD_GLOBAL ) ) )
			break;
	if( !h )
		return false;
	g_nccl.GetUniqueId = (int ( * )( ncclUniqueId_* ))dlsym( h, "ncclGetUniqueId" );
	g_nccl.CommInitRank = (int ( * )( ncclComm_t*, int, ncclUniqueId_, int ))dlsym( h, "ncclCommInitRank" );
	g_nccl.CommDestroy = (int ( * )( ncclComm_t ))dlsym( h, "ncclCommDestroy" );
	g_nccl.AllReduce =
	    (int ( * )( const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t ))dlsym( h, "ncclAllReduce" );
	g_nccl.AllGather =
	    (int ( * )( const void*, void*, size_t, int, ncclComm_t, cudaStream_t ))dlsym( h, "ncclAllGather" );
	g_nccl.GroupStart = (int ( * )())dlsym( h, "ncclGroupStart" );
	g_nccl.GroupEnd = (int ( * )())dlsym( h, "ncclGroupEnd" );
	g_nccl.GetErrorString = (const char* ( * )( int ))dlsym( h, "ncclGetErrorString" );
	if( !g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllReduce || !g_nccl.CommDestroy )
		return false;
	g_nccl.handle = h;
	return true;
}
enum
{
	NCCL_UINT8 = 1,
	NCCL_INT32 = 2,
	NCCL_FLOAT32 = 7,
	NCCL_FLOAT64 = 8,
	NCCL_SUM = 0,
	NCCL_MAX = 2,
	NCCL_MIN = 3
};

static std::map<std::string, int>& tune_cache()
{
	static std::map<std::string, int> cache;
	return cache;
}
int gs2d_tune( const char* env_name, int dflt )
{
	auto& cache = tune_cache();
	auto it = cache.find( env_name );
	if( it != cache.end() )
		return it->second;
	const char* e = getenv( env_name );
	const int v = e ? atoi( e ) : dflt;
	cache[env_name] = v;
	return v;
}
static int g_tune_epoch = 0; // captured iteration graphs are rebuilt when a tuning knob changes
extern "C" int gs2d_set_tuning( const char* name, int value )
{
	if( !name )
		return GS2D_ERR_INVALID;
	g_tune_epoch++;
	tune_cache()[name] = value;
	return GS2D_OK;
}

// ---- context ----------------------------------------------------------------------------------------------------
struct gs2d_ctx
{
	int device = 0;
	cudaStream_t stream = nullptr;
	Geometry g{};
	Buffers b{};
	gs2d_params hp{};
	int sorted_buf = 0;
	bool have_forward = false;
	uint64_t launches = 0;
	char err[512] = { 0 };

	void* pinned = nullptr; // host staging
	size_t pinned_bytes = 0;
	void* dstage = nullptr; // device staging (AoS <-> SoA, u8 targets)
	size_t dstage_bytes = 0;
	DevState* h_state = nullptr; // pinned mirror
	// Slot binning (gs2d_binning.cu) is built for buckets of a few entries; when Gaussians have grown so large that most
	// arrivals overflow their bucket's slot row, every overflow costs an atomic on ONE counter and the triple path
	// (one reservation per warp chunk) is the faster one. Both build identical lists, so the context simply switches
	// at its synchronisation points: off when more than a quarter of the last iteration's pairs overflowed, on again
	// (to look) 256 iterations later.
	bool slot_binning = true;
	long long slot_retry_at = 0;
	unsigned long long* h_n_over = nullptr; // pinned: overflow arrivals of the last iteration

	// asynchronous target upload (gs2d_set_target_rgba8_async): H2D on its own stream into one of two device staging
	// buffers; the u8 -> float conversion is enqueued on the main stream by the next forward (apply_pending_target)
	// second stream of an iteration: the row table kernel runs beside the scan / place / rank kernels (SideStream)
	SideStream side{};
	cudaStream_t copy_stream = nullptr;
	uint8_t* ustage[2] = { nullptr, nullptr };
	size_t ustage_bytes = 0;
	cudaEvent_t copy_done[2] = { nullptr, nullptr }, convert_done[2] = { nullptr, nullptr };
	int ustage_next = 0;
	int pending_buf = -1, pending_image = 0; // an uploaded target that has not been taken over yet
	// One-image contexts read an asynchronously uploaded RGBA8 target where the copy engine put it: Buffers::target8
	// points at staging buffer `adopted_buf` (no device-to-device copy on the compute stream); target8_own is the
	// buffer the blocking uploads and image batches use. convert_done[k] then means "nothing enqueued reads buffer k".
	uchar4* target8_own = nullptr;
	int adopted_buf = -1;
	std::vector<char> img_fmt;               // per image: 0 no target yet (zeros), 1 float32, 2 RGBA8 (target_store_u8)
	size_t pending_px0 = 0, pending_npx = 0;

	// One iteration captured as a CUDA graph (single GPU, profiling off): 13 dependent nodes replayed per iteration
	// instead of 13 launches - what the small configurations (268x213, 1024 Gaussians: 0.16 ms per iteration) are
	// bound by. Rebuilt whenever anything baked into the nodes changes (GraphKey).
	struct GraphKey
	{
		Geometry g;
		gs2d_params hp;
		void* ptr[8];
		void* target8;
		unsigned long long capacity;
		int mse_cap, count_live, tune_epoch, pad;
	};
	// two slots: a caller that streams targets alternates between the two staging buffers (GraphKey::target8)
	struct GraphSlot
	{
		cudaGraphExec_t exec = nullptr;
		GraphKey key{};
		int nodes = 0;
		uint64_t used = 0;
	};
	GraphSlot graphs[2];
	uint64_t graph_clock = 0;

	long long expected_iteration = 0; // what the device counter will read once everything enqueued has run

	// pipelined loss read-back (gs2d_mse_ticket / gs2d_mse_wait): a ring of mapped page-locked slots; a one-block
	// kernel enqueued behind an iteration stores { DevState, loss[n_images] } into one, an event follows it
	static const int N_TICKETS = 8;
	struct Ticket
	{
		cudaEvent_t ev = nullptr;
		uint64_t id = 0;
		long long expected = 0; // iteration counter the copy must show if nothing was skipped
	};
	Ticket tickets[N_TICKETS];
	uint64_t next_ticket = 1;
	unsigned char* ticket_pin = nullptr;
	size_t ticket_stride = 0;
	cudaEvent_t ev[10]{};
	float last_ms = 0.0f;

	ncclComm_t comm = nullptr;
	int rank = 0, n_ranks = 1;
	// peer-memory update (setup_peer_update): the other ranks' gradient / parameter / moment planes mapped through
	// CUDA IPC; this rank owns the quads [own_q0, own_q1) of every plane (Adam moments are valid only there)
	bool peer_update = false;
	bool self_peer = false; // developer knob GS2D_PEER_SELF=1: the peer-memory update with ONE rank (profiling under ncu)
	PeerPtrs peers{};
	float* peer_m[GS2D_MAX_RANKS] = {};
	float* peer_v[GS2D_MAX_RANKS] = {};
	std::vector<void*> ipc_opened;
	int own_q0 = 0, own_q1 = 0;
	unsigned char* mbox = nullptr;          // this rank's mailbox (PeerMailbox), mapped by every peer
	unsigned long long* peer_seq = nullptr; // device: [2] sequence numbers of barrier 1 / barrier 2, bumped by the kernels
	unsigned* touch_mem = nullptr;          // device: this rank's touch bitmap (Buffers::touch points here while the sparse exchange is on)
	unsigned char* reach_prev = nullptr;    // device: PeerPtrs::reach_prev
	unsigned char* reach_now = nullptr;     // device: PeerPtrs::reach_now
	unsigned long long* peer_traffic = nullptr; // device: PeerPtrs::traffic
	long long traffic_iter0 = 0;            // expected_iteration when the traffic counters were last reset
	bool peer_flags = true;                 // barriers over the mailboxes (else tiny NCCL all-reduces): GS2D_PEER_FLAGS
	// developer timing of the pieces of the peer-memory update (GS2D_PEER_DEBUG=1): events around barrier 1, the
	// fused kernel, barrier 2 and the gradient clear of every iteration; averages are printed by gs2d_destroy
	std::vector<cudaEvent_t> dbg_ev;

	// per-iteration CUDA-event profiling (gs2d_set_profiling): 7 events per enqueued iteration
	struct ProfRec
	{
		cudaEvent_t ev[7];
	};
	std::vector<ProfRec> prof_pool;
	size_t prof_used = 0;
	bool profiling = false;
	bool count_live = false;             // gs2d_set_stats: count live pairs / staged entries in the forward blend of gs2d_step
	unsigned long long bwd_live = 0;     // live pairs counted by the last gs2d_backward
	double prof_sum[6] = { 0, 0, 0, 0, 0, 0 };
	long long prof_iters = 0;
};

#define CK( call )                                                                                                    \
	do                                                                                                                \
	{                                                                                                                 \
		cudaError_t e_ = ( call );                                                                                    \
		if( e_ != cudaSuccess )                                                                                       \
		{                                                                                                             \
			snprintf( c->err, sizeof( c->err ), "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString( e_ ) ); \
			return ( e_ == cudaErrorMemoryAllocation ) ? GS2D_ERR_OOM : GS2D_ERR_CUDA;                                \
		}                                                                                                             \
	} while( 0 )

static int check_launch( gs2d_ctx* c )
{
	cudaError_t e = cudaGetLastError();
	if( e != cudaSuccess )
	{
		snprintf( c->err, sizeof( c->err ), "kernel launch failed: %s", cudaGetErrorString( e ) );
		return GS2D_ERR_CUDA;
	}
	return GS2D_OK;
}

extern "C" void gs2d_default_params( gs2d_params* p )
{
	p->lr = 0.05f;
	p->beta1 = 0.9f;
	p->beta2 = 0.99f;
	p->eps = 1.0e-15f;
	p->bounds = 3.0f;
	p->min_throughput = 1.0f / 256.0f;
	p->s_min = 1.0f;
	p->s_max = 1024.0f;
	p->o_min = 0.1f;
	p->o_max = 1.0f;
	p->optimize_opacity = 0;
}
extern "C" int gs2d_version( void ) { return GS2D_VERSION; }
extern "C" const char* gs2d_strerror( int code )
{
	switch( code )
	{
	case GS2D_OK: return "ok";
	case GS2D_ERR_INVALID: return "invalid argument";
	case GS2D_ERR_NO_DEVICE: return "no sm_100 CUDA device available (there is no CPU fallback)";
	case GS2D_ERR_CUDA: return "CUDA error";
	case GS2D_ERR_NONFINITE: return "non-finite splat parameter (the reference would abort here, main.cpp:752-785)";
	case GS2D_ERR_OOM: return "out of memory";
	case GS2D_ERR_NCCL: return "NCCL unavailable or failed";
	case GS2D_ERR_STATE: return "call sequence error";
	case GS2D_ERR_IO: return "file could not be opened, read or written";
	case GS2D_ERR_FORMAT: return "unsupported or mismatching file contents";
	default: return "unknown error";
	}
}
extern "C" const char* gs2d_last_error( const gs2d_ctx* c ) { return c ? c->err : ""; }

static void owned_quads( const gs2d_ctx* c, int r, int* q0, int* q1 )
{
	const int n4 = c->g.n_pad / 4, per = ( n4 + c->n_ranks - 1 ) / c->n_ranks;
	*q0 = std::min( r * per, n4 );
	*q1 = std::min( ( r + 1 ) * per, n4 );
}

static int nccl_fail( gs2d_ctx* c, const char* what, int e )
{
	snprintf( c->err, sizeof( c->err ), "%s failed: %s", what, g_nccl.GetErrorString ? g_nccl.GetErrorString( e ) : "?" );
	return GS2D_ERR_NCCL;
}

static int ensure_staging( gs2d_ctx* c, size_t bytes )
{
	if( bytes > c->pinned_bytes )
	{
		if( c->pinned )
			cudaFreeHost( c->pinned );
		c->pinned = nullptr;
		c->pinned_bytes = 0;
		CK( cudaMallocHost( &c->pinned, bytes ) );
		c->pinned_bytes = bytes;
	}
	if( bytes > c->dstage_bytes )
	{
		if( c->dstage )
			cudaFree( c->dstage );
		c->dstage = nullptr;
		c->dstage_bytes = 0;
		CK( cudaMalloc( &c->dstage, bytes ) );
		c->dstage_bytes = bytes;
	}
	return GS2D_OK;
}

static int alloc_keys( gs2d_ctx* c, unsigned long long capacity )
{
	Buffers& b = c->b;
	for( int i = 0; i < 2; i++ )
	{
		if( b.vals[i] )
			cudaFree( b.vals[i] );
		b.vals[i] = nullptr;
	}
	if( b.triples )
		cudaFree( b.triples );
	b.triples = nullptr;
	if( b.cover )
		cudaFree( b.cover );
	b.cover = nullptr;
	capacity = std::min<unsigned long long>( std::max<unsigned long long>( capacity, 1ull << 16 ), 0xFFFF0000ull );
	CK( cudaMalloc( &b.triples, sizeof( unsigned long long ) * capacity ) );
	CK( cudaMalloc( &b.vals[0], sizeof( unsigned ) * capacity ) );
	CK( cudaMalloc( &b.vals[1], sizeof( unsigned ) * capacity ) );
	// coverage words, and behind them one word per round of a tile (round_vis_index, gs2d_blend.cuh)
	CK( cudaMalloc( &b.cover, sizeof( unsigned ) * ( 8 * capacity + capacity / 32 + (size_t)c->g.n_tiles_total + 8 ) ) );
	b.capacity = capacity;
	return GS2D_OK;
}

static void set_band( gs2d_ctx* c, int rank, int n_ranks )
{
	const int per = ( c->g.tiles_y + n_ranks - 1 ) / n_ranks;
	c->g.band_ty0 = std::min( rank * per, c->g.tiles_y );
	c->g.band_ty1 = std::min( ( rank + 1 ) * per, c->g.tiles_y );
	c->rank = rank;
	c->n_ranks = n_ranks;
}

extern "C" void gs2d_destroy( gs2d_ctx* c )
{
	if( !c )
		return;
	cudaSetDevice( c->device );
	if( c->stream )
		cudaStreamSynchronize( c->stream );
	if( !c->dbg_ev.empty() )
	{
		double sum[4] = { 0, 0, 0, 0 };
		size_t n = c->dbg_ev.size() / 5, skip = n > 20 ? 10 : 0;
		for( size_t i = skip; i < n; i++ )
			for( int k = 0; k < 4; k++ )
			{
				float ms = 0;
				cudaEventElapsedTime( &ms, c->dbg_ev[i * 5 + k], c->dbg_ev[i * 5 + k + 1] );
				sum[k] += ms;
			}
		fprintf( stderr, "[gs2d rank %d] peer update over %zu iterations: barrier1 %.1f us, advance+fused kernel %.1f us, "
		                 "barrier2 %.1f us, clear %.1f us\n",
		         c->rank, n - skip, 1e3 * sum[0] / ( n - skip ), 1e3 * sum[1] / ( n - skip ), 1e3 * sum[2] / ( n - skip ),
		         1e3 * sum[3] / ( n - skip ) );
		for( cudaEvent_t e : c->dbg_ev )
			cudaEventDestroy( e );
	}
	for( gs2d_ctx::GraphSlot& gs : c->graphs )
		if( gs.exec )
			cudaGraphExecDestroy( gs.exec );
	for( auto& t : c->tickets )
		if( t.ev )
			cudaEventDestroy( t.ev );
	if( c->ticket_pin )
		cudaFreeHost( c->ticket_pin );
	for( void* p : c->ipc_opened )
		cudaIpcCloseMemHandle( p );
	if( c->comm && g_nccl.CommDestroy )
		g_nccl.CommDestroy( c->comm );
	Buffers& b = c->b;
	if( c->side.stream )
	{
		cudaStreamSynchronize( c->side.stream );
		cudaStreamDestroy( c->side.stream );
	}
	if( c->side.ev_fork )
		cudaEventDestroy( c->side.ev_fork );
	if( c->side.ev_join )
		cudaEventDestroy( c->side.ev_join );
	void* ptrs[] = { b.params,     b.grads,  b.adam_m, b.adam_v,       b.rec,        b.bucket_cnt, b.bucket_off, b.triples,
	                 b.rowrec,     b.rowtab,     b.rt_list, b.band_list, c->target8_own, b.u8_lut,
	                 b.scan_sums,  b.vals[0],    b.vals[1],    b.cover, b.tile_range,
	                 b.target,     b.image,  b.loss_partial, b.live_slots, b.stats, b.loss,  b.mse_ring,
	                 b.state,      c->dstage, c->mbox,   b.tile_done, b.bucket_slots, c->peer_seq, c->touch_mem, c->reach_prev, c->reach_now, c->peer_traffic };
	for( void* p : ptrs )
		if( p )
			cudaFree( p );
	for( int k = 0; k < 2; k++ )
	{
		if( c->ustage[k] )
			cudaFree( c->ustage[k] );
		if( c->copy_done[k] )
			cudaEventDestroy( c->copy_done[k] );
		if( c->convert_done[k] )
			cudaEventDestroy( c->convert_done[k] );
	}
	if( c->copy_stream )
	{
		cudaStreamSynchronize( c->copy_stream );
		cudaStreamDestroy( c->copy_stream );
	}
	if( c->pinned )
		cudaFreeHost( c->pinned );
	if( c->h_state )
		cudaFreeHost( c->h_state );
	if( c->h_n_over )
		cudaFreeHost( c->h_n_over );
	for( auto& e : c->ev )
		if( e )
			cudaEventDestroy( e );
	for( auto& r : c->prof_pool )
		for( auto& e : r.ev )
			cudaEventDestroy( e );
	if( c->stream )
		cudaStreamDestroy( c->stream );
	delete c;
}

extern "C" int gs2d_create( gs2d_ctx** out, int W, int H, int n_splats, int n_images, int device,
                            const gs2d_params* params )
{
	if( !out )
		return GS2D_ERR_INVALID;
	*out = nullptr;
	// tile columns are 12-bit fields in k_preprocess; row spans are packed into 16 bits per end (pack_row_span)
	if( W <= 0 || H <= 0 || n_splats < 0 || n_images <= 0 || W > 65535 || H > 65535 * GS2D_TILE )
		return GS2D_ERR_INVALID;
	if( (long long)n_splats * n_images > 0x7FFF0000ll || n_images > 65535 )
		return GS2D_ERR_INVALID;
	int n_dev = 0;
	if( cudaGetDeviceCount( &n_dev ) != cudaSuccess || device < 0 || device >= n_dev )
		return GS2D_ERR_NO_DEVICE;
	cudaDeviceProp prop;
	if( cudaGetDeviceProperties( &prop, device ) != cudaSuccess || prop.major != 10 )
		return GS2D_ERR_NO_DEVICE; // the kernels are compiled for sm_100a only
	gs2d_ctx* c = new gs2d_ctx();
	c->device = device;
	if( params )
		c->hp = *params;
	else
		gs2d_default_params( &c->hp );
	if( c->hp.bounds != 3.0f )
	{
		delete c;
		return GS2D_ERR_INVALID; // the footprint solve hard-codes SPLAT_BOUNDS = 3 (main.cpp:7)
	}
	auto fail = [&]( int code ) {
		gs2d_destroy( c );
		return code;
	};
#define CKC( call )                                                                                                   \
	do                                                                                                                \
	{                                                                                                                 \
		cudaError_t e_ = ( call );                                                                                    \
		if( e_ != cudaSuccess )                                                                                       \
		{                                                                                                             \
			fprintf( stderr, "gs2d_create: %s -> %s\n", #call, cudaGetErrorString( e_ ) );                           \
			return fail( e_ == cudaErrorMemoryAllocation ? GS2D_ERR_OOM : GS2D_ERR_CUDA );                            \
		}                                                                                                             \
	} while( 0 )
	CKC( cudaSetDevice( device ) );
	{
		// The side stream (k_row_table: pure arithmetic, 31 000 short blocks) gets the LOWEST priority and the main
		// stream the highest: the binning kernels that run beside it (scan / place / rank: few instructions, long
		// waits on L2) then take the SM slots they need as soon as row-table blocks retire, and the row table fills
		// what is left - with equal priorities the kernel launched first simply occupies every SM.
		int prio_lo = 0, prio_hi = 0;
		cudaDeviceGetStreamPriorityRange( &prio_lo, &prio_hi );
		CKC( cudaStreamCreateWithPriority( &c->stream, cudaStreamNonBlocking, prio_hi ) );
		CKC( cudaStreamCreateWithPriority( &c->side.stream, cudaStreamNonBlocking, prio_lo ) );
	}
	CKC( cudaEventCreateWithFlags( &c->side.ev_fork, cudaEventDisableTiming ) );
	CKC( cudaEventCreateWithFlags( &c->side.ev_join, cudaEventDisableTiming ) );
	prepare_forward_pdl();
	prepare_backward_pdl();
	cudaGetLastError(); // a refused carveout hint is not an error of this call
	for( auto& e : c->ev )
		CKC( cudaEventCreate( &e ) );

	Geometry& g = c->g;
	g.W = W;
	g.H = H;
	g.n_splats = n_splats;
	g.n_images = n_images;
	g.n_total = n_splats * n_images;
	g.n_pad = ( ( g.n_total + 3 ) / 4 ) * 4 + 4;
	g.tiles_x = ( W + GS2D_TILE - 1 ) / GS2D_TILE;
	g.tiles_y = ( H + GS2D_TILE - 1 ) / GS2D_TILE;
	g.n_tiles_total = n_images * g.tiles_x * g.tiles_y;
	g.nt_pad = ( g.n_tiles_total + 3 ) & ~3;
	// index slices: buckets of ~3 entries at the reference density (~14 tiles per splat), 16-bit bucket counters
	g.n_slices = 1;
	while( g.n_slices < 64 && (long long)g.n_slices * 4 * g.tiles_x * g.tiles_y < (long long)n_splats * 14 )
		g.n_slices *= 2;
	while( ( n_splats + g.n_slices - 1 ) / g.n_slices > 65535 )
		g.n_slices *= 2;
	g.slice_size = std::max( 1, ( n_splats + g.n_slices - 1 ) / g.n_slices );
	if( (long long)g.n_tiles_total * g.n_slices > 0x7FFF0000ll )
		return fail( GS2D_ERR_INVALID );
	g.min_throughput = c->hp.min_throughput;
	set_band( c, 0, 1 );

	Buffers& b = c->b;
	const size_t plane_bytes = sizeof( float ) * (size_t)g.n_pad * GS2D_NPLANES;
	const size_t n_px = (size_t)n_images * W * H;
	const int nb = ( g.n_total + 255 ) / 256;
	CKC( cudaMalloc( &b.params, plane_bytes ) );
	CKC( cudaMalloc( &b.grads, plane_bytes ) );
	CKC( cudaMalloc( &b.adam_m, plane_bytes ) );
	CKC( cudaMalloc( &b.adam_v, plane_bytes ) );
	CKC( cudaMalloc( &b.rec, sizeof( SplatRec ) * (size_t)std::max( g.n_total, 1 ) ) );
	CKC( cudaMalloc( &b.rowrec, sizeof( RowRec ) * (size_t)std::max( g.n_total, 1 ) ) );
	// row table: 512 bytes per splat; cleared once - the blend looks every staged entry up, also those whose rows are
	// not in the table (their words are overwritten by the in-kernel solve, build_words)
	CKC( cudaMalloc( &b.rowtab, sizeof( unsigned ) * GS2D_ROWTAB_STRIDE * (size_t)std::max( g.n_total, 1 ) ) );
	CKC( cudaMemset( b.rowtab, 0, sizeof( unsigned ) * GS2D_ROWTAB_STRIDE * (size_t)std::max( g.n_total, 1 ) ) );
	CKC( cudaMemset( b.rowrec, 0xff, sizeof( RowRec ) * (size_t)std::max( g.n_total, 1 ) ) );
	CKC( cudaMalloc( &b.rt_list, sizeof( unsigned ) * (size_t)std::max( g.n_total, 1 ) ) );
	CKC( cudaMalloc( &b.band_list, sizeof( unsigned ) * (size_t)std::max( g.n_total, 1 ) ) );
	const size_t n_buckets = (size_t)g.n_tiles_total * g.n_slices;
	// [16-bit counters per (tile, slice) | 8-byte overflow counter of the slot binning, 8-byte lengths of rt_list and band_list, ticket and status words of the scan's look-back | group rows: 32-bit arrival
	// counters per (group of 4 tiles, slice) | 16-bit totals per bucket in scan order] - the first two (three with group
	// rows) are cleared by one memset per iteration (launch_preprocess_and_bin)
	CKC( cudaMalloc( &b.bucket_cnt, sizeof( unsigned short ) * (size_t)g.nt_pad * g.n_slices +
	                                    ( 4 + n_buckets / 1024 + 4 ) * sizeof( unsigned long long ) +
	                                    sizeof( unsigned ) * (size_t)( g.nt_pad / 4 ) * g.n_slices +
	                                    sizeof( unsigned short ) * ( n_buckets + 8 ) ) );
	// slot rows: 16 bytes per (tile, slice) = group rows: 64 bytes per (group of 4 tiles, slice)
	static_assert( sizeof( unsigned short ) * GS2D_BUCKET_SLOTS * 4 == sizeof( unsigned ) * 16, "slot rows and group rows share their memory" );
	CKC( cudaMalloc( &b.bucket_slots, sizeof( unsigned short ) * GS2D_BUCKET_SLOTS * (size_t)g.nt_pad * g.n_slices ) );
	// cleared once: k_rank loads a bucket's whole 16-byte slot row, of which only the first `count` slots were written
	CKC( cudaMemset( b.bucket_slots, 0, sizeof( unsigned short ) * GS2D_BUCKET_SLOTS * (size_t)g.nt_pad * g.n_slices ) );
	CKC( cudaMalloc( &b.bucket_off, sizeof( unsigned ) * ( n_buckets + 4 ) ) );
	CKC( cudaMalloc( &b.scan_sums, sizeof( unsigned ) * ( n_buckets / 1024 + 16 ) ) );
	CKC( cudaMalloc( &b.tile_range, sizeof( uint2 ) * (size_t)g.n_tiles_total ) );
	CKC( cudaMalloc( &b.target, sizeof( float4 ) * n_px ) );
	CKC( cudaMalloc( &b.u8_lut, sizeof( float ) * 256 ) );
	c->launches += launch_u8_lut( b.u8_lut, c->stream );
	g.u8_lut = b.u8_lut;
	g.target_u8 = 0;
	c->img_fmt.assign( (size_t)n_images, 0 );
	CKC( cudaMalloc( &b.image, sizeof( float4 ) * n_px ) );
	CKC( cudaMalloc( &b.loss_partial, sizeof( double ) * (size_t)g.n_tiles_total ) );
	CKC( cudaMalloc( &b.tile_done, sizeof( unsigned ) * (size_t)g.n_tiles_total ) );
	CKC( cudaMalloc( &b.live_slots, sizeof( unsigned long long ) * GS2D_LIVE_SLOTS * 3 ) );
	b.staged_slots = b.live_slots + GS2D_LIVE_SLOTS;
	CKC( cudaMalloc( &b.stats, sizeof( unsigned long long ) * 8 ) );
	CKC( cudaMalloc( &b.loss, sizeof( double ) * ( (size_t)n_images + 1 ) ) );
	b.mse_cap = 64;
	CKC( cudaMalloc( &b.mse_ring, sizeof( double ) * (size_t)b.mse_cap * n_images ) );
	CKC( cudaMalloc( &b.state, sizeof( DevState ) ) );
	CKC( cudaMallocHost( &c->h_state, sizeof( DevState ) ) );
	CKC( cudaMallocHost( &c->h_n_over, sizeof( unsigned long long ) ) );
	*c->h_n_over = 0;
	CKC( cudaMemset( b.params, 0, plane_bytes ) );
	CKC( cudaMemset( b.grads, 0, plane_bytes ) );
	CKC( cudaMemset( b.adam_m, 0, plane_bytes ) );
	CKC( cudaMemset( b.adam_v, 0, plane_bytes ) );
	CKC( cudaMemset( b.target, 0, sizeof( float4 ) * n_px ) );
	CKC( cudaMemset( b.image, 0, sizeof( float4 ) * n_px ) );
	CKC( cudaMemset( b.loss_partial, 0, sizeof( double ) * (size_t)g.n_tiles_total ) );
	CKC( cudaMemset( b.live_slots, 0, sizeof( unsigned long long ) * GS2D_LIVE_SLOTS * 3 ) );
	CKC( cudaMemset( b.stats, 0, sizeof( unsigned long long ) * 8 ) );
	CKC( cudaMemset( b.loss, 0, sizeof( double ) * ( (size_t)n_images + 1 ) ) );
	DevState s0{};
	s0.beta1t = 1.0f;
	s0.beta2t = 1.0f;
	CKC( cudaMemcpy( b.state, &s0, sizeof( s0 ), cudaMemcpyHostToDevice ) );
	{
		// ~17 tiles per splat at the reference's initial scales; start with twice that
		int rc = alloc_keys( c, 40ull * (unsigned long long)std::max( g.n_total, 1 ) );
		if( rc != GS2D_OK )
			return fail( rc );
	}
#undef CKC
	*out = c;
	return GS2D_OK;
}

// ---- targets, state ---------------------------------------------------------------------------------------------
// Host pointer already page-locked (cudaMallocHost / cudaHostRegister / torch pin_memory)? Then DMA straight from it.
static const void* h2d_source( gs2d_ctx* c, const void* user, size_t bytes )
{
	cudaPointerAttributes attr;
	if( cudaPointerGetAttributes( &attr, user ) == cudaSuccess && attr.type == cudaMemoryTypeHost )
		return user;
	cudaGetLastError();
	memcpy( c->pinned, user, bytes );
	return c->pinned;
}

static int apply_pending_target( gs2d_ctx* c );
static int release_adopted( gs2d_ctx* c );

extern "C" int gs2d_set_target_rgba32f( gs2d_ctx* c, int image, const float* rgba )
{
	if( !c || !rgba || image < 0 || image >= c->g.n_images )
		return GS2D_ERR_INVALID;
	CK( cudaSetDevice( c->device ) );
	const size_t n_px = (size_t)c->g.W * c->g.H, bytes = n_px * sizeof( float4 );
	int rc = apply_pending_target( c );
	if( rc )
		return rc;
	rc = ensure_staging( c, bytes );
	if( rc )
		return rc;
	if( c->g.target_u8 )
	{
		// the context has been reading RGBA8 targets: from now on it reads float32 ones, the 8-bit images converted
		// once (u8 / 255.0f, prp.cpp:796-807)
		for( int i = 0; i < c->g.n_images; i++ )
			if( c->img_fmt[(size_t)i] == 2 )
			{
				c->launches += launch_u8_to_target( reinterpret_cast<const uint8_t*>( c->b.target8 + (size_t)i * n_px ),
				                                    c->b.target + (size_t)i * n_px, n_px, c->stream );
				c->img_fmt[(size_t)i] = 1;
			}
		c->g.target_u8 = 0;
		rc = release_adopted( c );
		if( rc )
			return rc;
	}
	CK( cudaMemcpyAsync( c->b.target + (size_t)image * n_px, h2d_source( c, rgba, bytes ), bytes, cudaMemcpyHostToDevice,
	                     c->stream ) );
	CK( cudaStreamSynchronize( c->stream ) );
	c->img_fmt[(size_t)image] = 1;
	return check_launch( c );
}

// An RGBA8 target (rows [px0, px0 + npx) of `image`, already on the device at dev_src) becomes the image's target.
// As long as no image of the context has a float32 target, the 8-bit pixels are kept as they are and the blend kernels
// convert on the fly (Geometry::target_u8): one device-to-device copy of 4 bytes per pixel instead of a conversion
// pass that reads 4 and writes 16. Otherwise they are converted into the float32 target as before.
// adopt >= 0: dev_src is staging buffer `adopt`, holding the rows at their place in a whole-image buffer (ustage[adopt] +
// px0): a one-image context then reads the target there instead of copying it (gs2d_ctx::adopted_buf).
static int release_adopted( gs2d_ctx* c )
{
	if( c->adopted_buf >= 0 )
	{
		// everything enqueued so far may read the buffer; the copy stream waits for this before it overwrites it
		CK( cudaEventRecord( c->convert_done[c->adopted_buf], c->stream ) );
		c->adopted_buf = -1;
	}
	c->b.target8 = c->target8_own;
	return GS2D_OK;
}

static int target_store_u8( gs2d_ctx* c, int image, const uint8_t* dev_src, size_t px0, size_t npx, int adopt = -1 )
{
	const size_t n_px = (size_t)c->g.W * c->g.H;
	bool any_float = false;
	for( char f : c->img_fmt )
		any_float = any_float || f == 1;
	if( any_float )
	{
		if( npx )
			c->launches += launch_u8_to_target( dev_src, c->b.target + (size_t)image * n_px + px0, npx, c->stream );
		c->img_fmt[(size_t)image] = 1;
		return check_launch( c );
	}
	if( adopt >= 0 && c->g.n_images == 1 && gs2d_tune( "GS2D_ADOPT_TARGET", 1 ) != 0 )
	{
		int rc = release_adopted( c ); // the buffer read so far is free once what has been enqueued has run
		if( rc )
			return rc;
		c->adopted_buf = adopt;
		c->b.target8 = reinterpret_cast<uchar4*>( c->ustage[adopt] );
		c->img_fmt[(size_t)image] = 2;
		c->g.target_u8 = 1;
		return GS2D_OK;
	}
	int rc = release_adopted( c );
	if( rc )
		return rc;
	if( !c->target8_own )
	{
		const size_t bytes = sizeof( uchar4 ) * n_px * (size_t)c->g.n_images;
		CK( cudaMalloc( &c->target8_own, bytes ) );
		CK( cudaMemsetAsync( c->target8_own, 0, bytes, c->stream ) ); // images without a target: zeros, as in the float32 buffer
	}
	c->b.target8 = c->target8_own;
	if( npx )
		CK( cudaMemcpyAsync( c->b.target8 + (size_t)image * n_px + px0, dev_src, npx * 4, cudaMemcpyDeviceToDevice, c->stream ) );
	c->img_fmt[(size_t)image] = 2;
	c->g.target_u8 = 1;
	return GS2D_OK;
}

// Pixel range [px0, px0 + npx) of an image this context needs: all of it, or - once a communicator shards the image
// into tile-row bands (gs2d_comm_init) - only the rows of this rank's band; the blend kernels read nothing else.
static void target_rows( const gs2d_ctx* c, size_t* px0, size_t* npx )
{
	int y0 = 0, y1 = c->g.H;
	if( c->comm && c->n_ranks > 1 )
	{
		y0 = std::min( c->g.band_ty0 * GS2D_TILE, c->g.H );
		y1 = std::min( c->g.band_ty1 * GS2D_TILE, c->g.H );
	}
	*px0 = (size_t)y0 * c->g.W;
	*npx = (size_t)( y1 - y0 ) * c->g.W;
}

// take an asynchronously uploaded target over on the main stream (before the next forward)
static int apply_pending_target( gs2d_ctx* c )
{
	if( c->pending_buf < 0 )
		return GS2D_OK;
	const int k = c->pending_buf;
	c->pending_buf = -1;
	CK( cudaStreamWaitEvent( c->stream, c->copy_done[k], 0 ) );
	int rc = target_store_u8( c, c->pending_image, c->ustage[k] + c->pending_px0 * 4, c->pending_px0, c->pending_npx, k );
	if( rc )
		return rc;
	if( c->adopted_buf != k ) // copied or converted: the staging buffer is free once that has run
		CK( cudaEventRecord( c->convert_done[k], c->stream ) );
	return check_launch( c );
}

extern "C" int gs2d_set_target_rgba8( gs2d_ctx* c, int image, const uint8_t* rgba )
{
	if( !c || !rgba || image < 0 || image >= c->g.n_images )
		return GS2D_ERR_INVALID;
	CK( cudaSetDevice( c->device ) );
	int rc = apply_pending_target( c ); // keep the order of uploads
	if( rc )
		return rc;
	size_t px0, npx;
	target_rows( c, &px0, &npx );
	const size_t bytes = npx * 4;
	rc = ensure_staging( c, std::max<size_t>( bytes, 16 ) );
	if( rc )
		return rc;
	if( npx )
	{
		CK( cudaMemcpyAsync( c->dstage, h2d_source( c, rgba + px0 * 4, bytes ), bytes, cudaMemcpyHostToDevice, c->stream ) );
	}
	rc = target_store_u8( c, image, (const uint8_t*)c->dstage, px0, npx );
	if( rc )
		return rc;
	CK( cudaStreamSynchronize( c->stream ) );
	return check_launch( c );
}

extern "C" int gs2d_set_target_rgba8_async( gs2d_ctx* c, int image, const uint8_t* rgba )
{
	if( !c || !rgba || image < 0 || image >= c->g.n_images )
		return GS2D_ERR_INVALID;
	CK( cudaSetDevice( c->device ) );
	cudaPointerAttributes attr;
	if( cudaPointerGetAttributes( &attr, rgba ) != cudaSuccess || attr.type != cudaMemoryTypeHost )
	{
		cudaGetLastError();
		return GS2D_ERR_INVALID; // an asynchronous copy needs page-locked memory; use gs2d_set_target_rgba8 otherwise
	}
	int rc = apply_pending_target( c ); // at most one unconverted upload: an older one takes effect first
	if( rc )
		return rc;
	if( !c->copy_stream )
	{
		CK( cudaStreamCreateWithFlags( &c->copy_stream, cudaStreamNonBlocking ) );
		for( int k = 0; k < 2; k++ )
		{
			CK( cudaEventCreateWithFlags( &c->copy_done[k], cudaEventDisableTiming ) );
			CK( cudaEventCreateWithFlags( &c->convert_done[k], cudaEventDisableTiming ) );
		}
	}
	size_t px0, npx;
	target_rows( c, &px0, &npx );
	const size_t bytes = npx * 4, full = (size_t)c->g.W * c->g.H * 4;
	if( c->ustage_bytes < full )
	{
		CK( cudaStreamSynchronize( c->copy_stream ) );
		CK( cudaStreamSynchronize( c->stream ) );
		for( int k = 0; k < 2; k++ )
		{
			if( c->ustage[k] )
				cudaFree( c->ustage[k] );
			c->ustage[k] = nullptr;
			CK( cudaMalloc( &c->ustage[k], full ) );
			CK( cudaMemsetAsync( c->ustage[k], 0, full, c->stream ) ); // rows outside a band: zeros, like target8_own
			CK( cudaEventRecord( c->convert_done[k], c->stream ) );
		}
		c->ustage_bytes = full;
	}
	const int k = c->ustage_next;
	c->ustage_next ^= 1;
	// the buffer is free once the conversion that last read it has run
	CK( cudaStreamWaitEvent( c->copy_stream, c->convert_done[k], 0 ) );
	if( bytes )
		CK( cudaMemcpyAsync( c->ustage[k] + px0 * 4, rgba + px0 * 4, bytes, cudaMemcpyHostToDevice, c->copy_stream ) );
	CK( cudaEventRecord( c->copy_done[k], c->copy_stream ) );
	c->pending_buf = k;
	c->pending_image = image;
	c->pending_px0 = px0;
	c->pending_npx = npx;
	return GS2D_OK;
}

static int write_state( gs2d_ctx* c, float b1t, float b2t, int iteration )
{
	DevState s{};
	s.beta1t = b1t;
	s.beta2t = b2t;
	s.iteration = iteration;
	*c->h_state = s;
	CK( cudaMemcpyAsync( c->b.state, c->h_state, sizeof( DevState ), cudaMemcpyHostToDevice, c->stream ) );
	// every rank holds the complete new state: the next exchange sends every quad to everyone once (sparse exchange)
	if( c->reach_prev )
		CK( cudaMemsetAsync( c->reach_prev, 0xff, 2 * ( (size_t)c->g.n_pad / 4 ), c->stream ) );
	CK( cudaStreamSynchronize( c->stream ) );
	c->expected_iteration = iteration;
	c->have_forward = false;
	return GS2D_OK;
}

extern "C" int gs2d_init_reference( gs2d_ctx* c )
{
	if( !c )
		return GS2D_ERR_INVALID;
	CK( cudaSetDevice( c->device ) );
	c->launches += launch_init_reference( c->g, c->b, c->stream );
	int rc = check_launch( c );
	if( rc )
		return rc;
	return write_state( c, 1.0f, 1.0f, 0 );
}

extern "C" int gs2d_upload_state( gs2d_ctx* c, int image, const gs2d_splat* splats, const gs2d_splat_adam* adam,
                                  float beta1t, float beta2t, int iteration )
{
	if( !c || ( !splats && c->g.n_splats > 0 ) || image < 0 || image >= c->g.n_images )
		return GS2D_ERR_INVALID;
	CK( cudaSetDevice( c->device ) );
	const size_t n = (size_t)c->g.n_splats;
	const size_t sb = n * sizeof( gs2d_splat ), ab = n * sizeof( gs2d_splat_adam );
	int rc = ensure_staging( c, sb + ab + 64 );
	if( rc )
		return rc;
	if( n )
	{
		memcpy( c->pinned, splats, sb );
		if( adam )
			memcpy( (char*)c->pinned + sb, adam, ab );
		CK( cudaMemcpyAsync( c->dstage, c->pinned, sb + ( adam ? ab : 0 ), cudaMemcpyHostToDevice, c->stream ) );
		c->launches += launch_aos_to_soa( c->g, c->b, image, (const gs2d_splat*)c->dstage,
		                                  adam ? (const gs2d_splat_adam*)( (char*)c->dstage + sb ) : nullptr, c->stream );
		// gradients of a freshly uploaded state are zero (dSplats is value-initialised every iteration, main.cpp:550)
		for( int f = 0; f < GS2D_NPLANES; f++ )
			CK( cudaMemsetAsync( c->b.grads + (size_t)f * c->g.n_pad + (size_t)image * n, 0, sizeof( float ) * n,
			                     c->stream ) );
		rc = check_launch( c );
		if( rc )
			return rc;
	}
	return write_state( c, beta1t, beta2t, iteration );
}

static int fetch_state( gs2d_ctx* c )
{
	CK( cudaMemcpyAsync( c->h_state, c->b.state, sizeof( DevState ), cudaMemcpyDeviceToHost, c->stream ) );
	// the overflow counter of the slot binning sits behind the bucket counters (launch_preprocess_and_bin)
	CK( cudaMemcpyAsync( c->h_n_over, c->b.bucket_cnt + (size_t)c->g.nt_pad * c->g.n_slices, sizeof( unsigned long long ),
	                     cudaMemcpyDeviceToHost, c->stream ) );
	CK( cudaStreamSynchronize( c->stream ) );
	return GS2D_OK;
}

// see gs2d_ctx::slot_binning; called with a fresh h_state / h_n_over and nothing in flight
static void adapt_binning( gs2d_ctx* c )
{
	const DevState& s = *c->h_state;
	const bool was = c->slot_binning;
	if( c->slot_binning )
	{
		if( s.need_capacity == 0 && s.n_keys > 0 && *c->h_n_over * 4ull > s.n_keys )
		{
			c->slot_binning = false;
			c->slot_retry_at = c->expected_iteration + 256;
		}
	}
	else if( c->expected_iteration >= c->slot_retry_at )
		c->slot_binning = true;
	if( was != c->slot_binning && gs2d_tune( "GS2D_DEBUG_BINNING", 0 ) )
		fprintf( stderr, "[gs2d] slot binning %s at iteration %lld (%llu of %llu pairs overflowed their slot row)\n",
		         c->slot_binning ? "on" : "off", c->expected_iteration, *c->h_n_over, s.n_keys );
}

static int download_aos( gs2d_ctx* c, int image, gs2d_splat* splats, gs2d_splat_adam* adam, gs2d_splat* grads )
{
	const size_t n = (size_t)c->g.n_splats;
	if( n == 0 )
		return GS2D_OK;
	const size_t sb = n * sizeof( gs2d_splat ), ab = n * sizeof( gs2d_splat_adam );
	int rc = ensure_staging( c, 2 * sb + ab + 64 );
	if( rc )
		return rc;
	char* d = (char*)c->dstage;
	c->launches += launch_soa_to_aos( c->g, c->b, image, splats ? (gs2d_splat*)d : nullptr,
	                                  adam ? (gs2d_splat_adam*)( d + sb ) : nullptr,
	                                  grads ? (gs2d_splat*)( d + sb + ab ) : nullptr, c->stream );
	rc = check_launch( c );
	if( rc )
		return rc;
	CK( cudaMemcpyAsync( c->pinned, c->dstage, 2 * sb + ab, cudaMemcpyDeviceToHost, c->stream ) );
	CK( cudaStreamSynchronize( c->stream ) );
	char* h = (char*)c->pinned;
	if( splats )
		memcpy( splats, h, sb );
	if( adam )
		memcpy( adam, h + sb, ab );
	if( grads )
		memcpy( grads, h + sb + ab, sb );
	return GS2D_OK;
}

static int resolve( gs2d_ctx* c ); // sync + capacity recovery, below

extern "C" int gs2d_download_state( gs2d_ctx* c, int image, gs2d_splat* splats, gs2d_splat_adam* adam, float* beta1t,
                                    float* beta2t, int* iteration )
{
	if( !c || image < 0 || image >= c->g.n_images )
		return GS2D_ERR_INVALID;
	CK( cudaSetDevice( c->device ) );
	int rc = resolve( c );
	if( rc && rc != GS2D_ERR_NONFINITE )
		return rc;
	if( ( adam || splats ) && c->comm && c->peer_update )
	{
		// the Adam moments of a slice live on the rank that owns it, and with the sparse exchange so does the only
		// complete copy of its parameters (this rank's copy is exact for the Gaussians that reach its band, stale for
		// the others): fetch the other slices over NVLink. The peers are quiescent: they cannot start their next update
		// before this rank joins the next iteration's barrier.
		for( int r = 0; r < c->n_ranks; r++ )
		{
			int q0, q1;
			owned_quads( c, r, &q0, &q1 );
			if( r == c->rank || q1 <= q0 )
				continue;
			for( int pl = 0; pl < GS2D_NPLANES; pl++ )
			{
				const size_t o = (size_t)pl * c->g.n_pad + (size_t)q0 * 4, bytes = sizeof( float ) * 4 * (size_t)( q1 - q0 );
				if( adam )
				{
					CK( cudaMemcpyAsync( c->b.adam_m + o, c->peer_m[r] + o, bytes, cudaMemcpyDeviceToDevice, c->stream ) );
					CK( cudaMemcpyAsync( c->b.adam_v + o, c->peer_v[r] + o, bytes, cudaMemcpyDeviceToDevice, c->stream ) );
				}
				if( splats && c->peers.sparse && !gs2d_tune( "GS2D_DOWNLOAD_LOCAL", 0 ) ) // (developer knob: this rank's own copy)
					CK( cudaMemcpyAsync( c->b.params + o, c->peers.params[r] + o, bytes, cudaMemcpyDeviceToDevice, c->stream ) );
			}
		}
	}
	int rc2 = ( splats || adam ) ? download_aos( c, image, splats, adam, nullptr ) : GS2D_OK;
	if( rc2 )
		return rc2;
	if( beta1t )
		*beta1t = c->h_state->beta1t;
	if( beta2t )
		*beta2t = c->h_state->beta2t;
	if( iteration )
		*iteration = c->h_state->iteration;
	return rc;
}

extern "C" int gs2d_set_optimize_opacity( gs2d_ctx* c, int enabled )
{
	if( !c )
		return GS2D_ERR_INVALID;
	c->hp.optimize_opacity = enabled ? 1 : 0;
	return GS2D_OK;
}

extern "C" int gs2d_get_optimize_opacity( const gs2d_ctx* c, int* enabled )
{
	if( !c || !enabled )
		return GS2D_ERR_INVALID;
	*enabled = c->hp.optimize_opacity;
	return GS2D_OK;
}

extern "C" int gs2d_get_dims( const gs2d_ctx* c, int* W, int* H, int* n_splats, int* n_images )
{
	if( !c )
		return GS2D_ERR_INVALID;
	if( W )
		*W = c->g.W;
	if( H )
		*H = c->g.H;
	if( n_splats )
		*n_splats = c->g.n_splats;
	if( n_images )
		*n_images = c->g.n_images;
	return GS2D_OK;
}

// ---- the iteration ----------------------------------------------------------------------------------------------
static int enqueue_forward( gs2d_ctx* c )
{
	int rcp = apply_pending_target( c );
	if( rcp )
		return rcp;
	c->launches += launch_preprocess_and_bin( c->g, c->b, &c->sorted_buf, c->stream, nullptr, c->slot_binning, &c->side );
	c->launches += launch_forward( c->g, c->b, c->sorted_buf, true, c->stream );
	c->launches += launch_finalize_loss( c->g, c->b, c->stream );
	c->have_forward = true;
	return check_launch( c );
}

static int enqueue_allreduce( gs2d_ctx* c )
{
	if( !c->comm )
		return GS2D_OK;
	// gradients of all 9 planes (padding included: zeros) and the per-image loss sums (+ the overflow vote)
	int e1 = g_nccl.AllReduce( c->b.grads, c->b.grads, (size_t)c->g.n_pad * GS2D_NPLANES, NCCL_FLOAT32, NCCL_SUM, c->comm,
	                           c->stream );
	int e2 = g_nccl.AllReduce( c->b.loss, c->b.loss, (size_t)c->g.n_images + 1, NCCL_FLOAT64, NCCL_SUM, c->comm,
	                           c->stream );
	if( e1 || e2 )
	{
		snprintf( c->err, sizeof( c->err ), "ncclAllReduce failed: %s",
		          g_nccl.GetErrorString ? g_nccl.GetErrorString( e1 ? e1 : e2 ) : "?" );
		return GS2D_ERR_NCCL;
	}
	return GS2D_OK;
}

// The exchange + update of one iteration (main.cpp:714-750 across ranks), after the backward pass.
//   one GPU            : k_advance, k_adam
//   NCCL path          : all-reduce of the 9 gradient planes and of the loss, then the same two kernels on every rank
//   peer-memory path   : all-reduce of the loss only (it doubles as the "all gradients are final" barrier), then ONE
//                        kernel that sums the peers' gradients for the slice this rank owns, runs Adam on it and stores
//                        the new parameters into every rank's planes; a one-int all-reduce (the non-finite flag) is
//                        the "all parameters have arrived" barrier; each rank then clears its own gradient planes.
// ev_mid (optional) is recorded between the collective and the update (profiling groups 4 and 5).
static int enqueue_update( gs2d_ctx* c, cudaEvent_t ev_mid, bool grads_already_reduced )
{
	cudaStream_t s = c->stream;
	const bool peer_path = ( c->comm || c->self_peer ) && c->peer_update;
	const bool dbg = peer_path && gs2d_tune( "GS2D_PEER_DEBUG", 0 ) && c->dbg_ev.size() < 5 * 4096;
	auto mark = [&]() {
		if( dbg )
		{
			cudaEvent_t e;
			cudaEventCreate( &e );
			cudaEventRecord( e, s );
			c->dbg_ev.push_back( e );
		}
	};
	if( peer_path )
	{
		mark();
		int e = 0;
		if( !grads_already_reduced ) // else gs2d_backward has reduced gradients and loss already
		{
			if( c->peer_flags )
				c->launches += launch_peer_loss_exchange( c->g, c->b, c->peers, c->peer_seq, s );
			else
				e = g_nccl.AllReduce( c->b.loss, c->b.loss, (size_t)c->g.n_images + 1, NCCL_FLOAT64, NCCL_SUM, c->comm, s );
		}
		if( e )
			return nccl_fail( c, "ncclAllReduce(loss)", e );
		if( ev_mid )
			cudaEventRecord( ev_mid, s );
		mark();
		c->launches += launch_advance( c->g, c->b, c->hp, s );
		PeerPtrs pp = c->peers;
		pp.sum_all = grads_already_reduced ? 0 : 1;
		c->launches += launch_adam_peer( c->g, c->b, c->hp, pp, c->own_q0, c->own_q1, s );
		mark();
		if( c->peer_flags )
			c->launches += launch_peer_barrier2( c->g, c->b, c->peers, c->peer_seq + 1, s );
		else
			e = g_nccl.AllReduce( &c->b.state->nonfinite, &c->b.state->nonfinite, 1, NCCL_INT32, NCCL_MAX, c->comm, s );
		if( e )
			return nccl_fail( c, "ncclAllReduce(flag)", e );
		mark();
		CK( cudaMemsetAsync( c->b.grads, 0, sizeof( float ) * (size_t)c->g.n_pad * GS2D_NPLANES, s ) );
		mark();
		return GS2D_OK;
	}
	if( !grads_already_reduced )
	{
		int rc = enqueue_allreduce( c );
		if( rc )
			return rc;
	}
	if( ev_mid )
		cudaEventRecord( ev_mid, s );
	c->launches += launch_advance( c->g, c->b, c->hp, s );
	c->launches += launch_adam( c->g, c->b, c->hp, s );
	return GS2D_OK;
}

static int enqueue_iteration( gs2d_ctx* c )
{
	cudaStream_t s = c->stream;
	int rcp = apply_pending_target( c );
	if( rcp )
		return rcp;
	gs2d_ctx::ProfRec* pr = nullptr;
	if( c->profiling )
	{
		if( c->prof_used == c->prof_pool.size() )
		{
			gs2d_ctx::ProfRec r;
			for( auto& e : r.ev )
				CK( cudaEventCreate( &e ) );
			c->prof_pool.push_back( r );
		}
		pr = &c->prof_pool[c->prof_used++];
	}
#define PROF( i )                                                                                                     \
	if( pr )                                                                                                          \
	cudaEventRecord( pr->ev[i], s )
	PROF( 0 );
	c->launches += launch_preprocess_and_bin( c->g, c->b, &c->sorted_buf, s, pr ? pr->ev[1] : nullptr, c->slot_binning,
	                                          gs2d_tune( "GS2D_SIDE_STREAM", 1 ) ? &c->side : nullptr );
	PROF( 2 );
	if( gs2d_pdl_usable() )
	{
		// forward -> backward overlap (gs2d_blend.cuh): the two grids must be adjacent on the stream, so the loss is
		// finalised behind the backward pass and profiling group 2 (forward) is empty, group 3 holds both passes
		PROF( 3 );
		c->launches += launch_forward( c->g, c->b, c->sorted_buf, c->count_live, s, true );
		c->launches += launch_backward( c->g, c->b, c->sorted_buf, c->hp.optimize_opacity != 0, false, s, true );
		c->launches += launch_finalize_loss( c->g, c->b, s );
	}
	else
	{
		c->launches += launch_forward( c->g, c->b, c->sorted_buf, c->count_live, s );
		c->launches += launch_finalize_loss( c->g, c->b, s );
		PROF( 3 );
		c->launches += launch_backward( c->g, c->b, c->sorted_buf, c->hp.optimize_opacity != 0, false, s );
	}
	PROF( 4 );
	int rc = enqueue_update( c, pr ? pr->ev[5] : nullptr, false );
	if( rc )
		return rc;
	PROF( 6 );
#undef PROF
	c->expected_iteration++;
	c->have_forward = false; // parameters changed: image/grads no longer correspond to them
	return check_launch( c );
}

// n iterations: replays of the captured graph where possible, plain launches otherwise
static int enqueue_iterations( gs2d_ctx* c, int n )
{
	// A multi-rank iteration is captured too when its exchange is made of this library's own kernels only (peer-memory
	// path with flag barriers: their sequence numbers live on the device); the NCCL path keeps plain launches.
	const bool comm_graph = !c->comm || ( c->peer_update && c->peer_flags && !gs2d_tune( "GS2D_PEER_DEBUG", 0 ) &&
	                                      gs2d_tune( "GS2D_COMM_GRAPH", 1 ) != 0 );
	const bool use_graph = n > 0 && !c->profiling && comm_graph && gs2d_tune( "GS2D_GRAPH", 1 ) != 0;
	if( use_graph )
	{
		int rc = apply_pending_target( c ); // stays outside the graph
		if( rc )
			return rc;
		gs2d_ctx::GraphKey key{};
		memset( &key, 0, sizeof( key ) );
		key.g = c->g;
		key.hp = c->hp;
		key.ptr[0] = c->b.triples;
		key.ptr[1] = c->b.vals[0];
		key.ptr[2] = c->b.vals[1];
		key.ptr[3] = c->b.cover;
		key.ptr[4] = c->b.mse_ring;
		key.ptr[5] = c->b.bucket_cnt;
		key.ptr[6] = c->comm;
		key.ptr[7] = c->peer_update ? (void*)c->peers.mbox[c->rank] : nullptr;
		key.target8 = c->b.target8;
		key.capacity = c->b.capacity;
		key.mse_cap = c->b.mse_cap;
		key.count_live = c->count_live ? 1 : 0;
		key.tune_epoch = g_tune_epoch;
		key.pad = c->slot_binning ? 1 : 0;
		gs2d_ctx::GraphSlot* slot = nullptr;
		for( gs2d_ctx::GraphSlot& gs : c->graphs )
			if( gs.exec && memcmp( &key, &gs.key, sizeof( key ) ) == 0 )
				slot = &gs;
		if( !slot )
		{
			// replace the slot used longest ago (or an empty one)
			slot = &c->graphs[0];
			for( gs2d_ctx::GraphSlot& gs : c->graphs )
				if( !gs.exec || ( slot->exec && gs.used < slot->used ) )
					slot = &gs;
			if( slot->exec )
				cudaGraphExecDestroy( slot->exec );
			slot->exec = nullptr;
			const uint64_t launches0 = c->launches;
			const long long expected0 = c->expected_iteration;
			cudaGraph_t graph = nullptr;
			CK( cudaStreamBeginCapture( c->stream, cudaStreamCaptureModeThreadLocal ) );
			rc = enqueue_iteration( c );
			cudaError_t e = cudaStreamEndCapture( c->stream, &graph );
			slot->nodes = (int)( c->launches - launches0 );
			c->launches = launches0;
			c->expected_iteration = expected0;
			if( rc == GS2D_OK && e == cudaSuccess && graph )
				e = cudaGraphInstantiate( &slot->exec, graph, 0 );
			if( graph )
				cudaGraphDestroy( graph );
			if( rc != GS2D_OK || e != cudaSuccess )
			{
				cudaGetLastError();
				slot->exec = nullptr; // capture unavailable: plain launches below
			}
			else
				slot->key = key;
		}
		if( slot->exec )
		{
			slot->used = ++c->graph_clock;
			for( int i = 0; i < n; i++ )
			{
				CK( cudaGraphLaunch( slot->exec, c->stream ) );
				c->launches += (uint64_t)slot->nodes;
				c->expected_iteration++;
			}
			c->have_forward = false;
			return GS2D_OK;
		}
	}
	for( int i = 0; i < n; i++ )
	{
		int rc = enqueue_iteration( c );
		if( rc )
			return rc;
	}
	return GS2D_OK;
}

// everything enqueued has completed: fold the recorded events into the per-group sums
static void collect_profile( gs2d_ctx* c )
{
	for( size_t i = 0; i < c->prof_used; i++ )
	{
		const gs2d_ctx::ProfRec& r = c->prof_pool[i];
		for( int k = 0; k < 6; k++ )
		{
			float ms = 0.0f;
			if( cudaEventElapsedTime( &ms, r.ev[k], r.ev[k + 1] ) == cudaSuccess )
				c->prof_sum[k] += ms;
		}
		c->prof_iters++;
	}
	cudaGetLastError();
	c->prof_used = 0;
}

// Wait for everything enqueued; if an iteration was skipped because the key buffers were too small (on this or on
// any other rank), grow them and run the missing iterations again.
static int resolve( gs2d_ctx* c )
{
	for( int attempt = 0; attempt < 8; attempt++ )
	{
		int rc = fetch_state( c );
		if( rc )
			return rc;
		const DevState& s = *c->h_state;
		if( s.peer_fault )
		{
			if( s.peer_fault == 2 )
			{
				snprintf( c->err, sizeof( c->err ), "a forward -> backward tile flag did not arrive within its time-out (GS2D_PDL=0 serialises the two blend grids)" );
				return GS2D_ERR_CUDA;
			}
			snprintf( c->err, sizeof( c->err ), "a peer-memory barrier timed out (a rank died or the ranks' call sequences diverged)" );
			return GS2D_ERR_NCCL;
		}
		if( s.nonfinite )
		{
			// the device froze at the failing iteration (DevState::frozen): nothing enqueued behind it advanced
			c->expected_iteration = s.iteration;
			adapt_binning( c );
			collect_profile( c );
			return GS2D_ERR_NONFINITE;
		}
		const long long missing = c->expected_iteration - (long long)s.iteration;
		if( s.need_capacity == 0 && missing <= 0 )
		{
			adapt_binning( c );
			collect_profile( c );
			return s.nonfinite ? GS2D_ERR_NONFINITE : GS2D_OK;
		}
		if( s.need_capacity )
		{
			rc = alloc_keys( c, s.need_capacity + s.need_capacity / 2 );
			if( rc )
				return rc;
			CK( cudaMemsetAsync( &c->b.state->need_capacity, 0, sizeof( unsigned long long ), c->stream ) );
		}
		if( missing > 0 )
		{
			// the skipped iterations may have left partial gradients behind (ranks that did not overflow)
			CK( cudaMemsetAsync( c->b.grads, 0, sizeof( float ) * (size_t)c->g.n_pad * GS2D_NPLANES, c->stream ) );
			c->expected_iteration = s.iteration;
			for( long long i = 0; i < missing; i++ )
				if( ( rc = enqueue_iteration( c ) ) )
					return rc;
		}
		else if( c->have_forward )
		{
			// a stand-alone gs2d_forward overflowed: redo it
			if( ( rc = enqueue_forward( c ) ) )
				return rc;
		}
	}
	snprintf( c->err, sizeof( c->err ), "key capacity did not converge" );
	return GS2D_ERR_OOM;
}

extern "C" int gs2d_step( gs2d_ctx* c, int n_iters, double* mse )
{
	if( !c || n_iters < 0 )
		return GS2D_ERR_INVALID;
	CK( cudaSetDevice( c->device ) );
	Buffers& b = c->b;
	if( mse && n_iters > b.mse_cap )
	{
		CK( cudaStreamSynchronize( c->stream ) );
		cudaFree( b.mse_ring );
		b.mse_ring = nullptr;
		b.mse_cap = n_iters;
		CK( cudaMalloc( &b.mse_ring, sizeof( double ) * (size_t)b.mse_cap * c->g.n_images ) );
	}
	CK( cudaMemsetAsync( &b.state->mse_slot, 0, sizeof( int ), c->stream ) );
	CK( cudaEventRecord( c->ev[0], c->stream ) );
	{
		int rc = enqueue_iterations( c, n_iters );
		if( rc )
			return rc;
	}
	CK( cudaEventRecord( c->ev[1], c->stream ) );
	if( !mse )
		return GS2D_OK;
	int rc = resolve( c );
	if( rc && rc != GS2D_ERR_NONFINITE )
		return rc;
	cudaEventElapsedTime( &c->last_ms, c->ev[0], c->ev[1] );
	const size_t bytes = sizeof( double ) * (size_t)n_iters * c->g.n_images;
	if( bytes )
	{
		int rc2 = ensure_staging( c, bytes );
		if( rc2 )
			return rc2;
		CK( cudaMemcpyAsync( c->pinned, b.mse_ring, bytes, cudaMemcpyDeviceToHost, c->stream ) );
		CK( cudaStreamSynchronize( c->stream ) );
		memcpy( mse, c->pinned, bytes );
	}
	return rc;
}

extern "C" int gs2d_forward( gs2d_ctx* c )
{
	if( !c )
		return GS2D_ERR_INVALID;
	CK( cudaSetDevice( c->device ) );
	int rc = enqueue_forward( c );
	if( rc )
		return rc;
	rc = resolve( c );
	return rc == GS2D_ERR_NONFINITE ? GS2D_OK : rc;
}

extern "C" int gs2d_backward( gs2d_ctx* c )
{
	if( !c )
		return GS2D_ERR_INVALID;
	if( !c->have_forward )
		return GS2D_ERR_STATE;
	CK( cudaSetDevice( c->device ) );
	CK( cudaMemsetAsync( c->b.grads, 0, sizeof( float ) * (size_t)c->g.n_pad * GS2D_NPLANES, c->stream ) );
	c->launches += launch_backward( c->g, c->b, c->sorted_buf, true, true, c->stream );
	int rc = check_launch( c );
	if( rc )
		return rc;
	rc = enqueue_allreduce( c );
	if( rc )
		return rc;
	unsigned long long slots[GS2D_LIVE_SLOTS];
	CK( cudaMemcpyAsync( slots, c->b.live_slots + 2 * GS2D_LIVE_SLOTS, sizeof( slots ), cudaMemcpyDeviceToHost,
	                     c->stream ) );
	CK( cudaStreamSynchronize( c->stream ) );
	c->bwd_live = 0;
	for( unsigned long long v : slots )
		c->bwd_live += v;
	return GS2D_OK;
}

extern "C" int gs2d_adam( gs2d_ctx* c )
{
	if( !c )
		return GS2D_ERR_INVALID;
	CK( cudaSetDevice( c->device ) );
	int rcu = enqueue_update( c, nullptr, true ); // gs2d_backward has already summed gradients and loss over ranks
	if( rcu )
		return rcu;
	c->expected_iteration++;
	c->have_forward = false;
	int rc = check_launch( c );
	if( rc )
		return rc;
	rc = fetch_state( c );
	if( rc )
		return rc;
	return c->h_state->nonfinite ? GS2D_ERR_NONFINITE : GS2D_OK;
}

extern "C" int gs2d_download_image_rgba32f( gs2d_ctx* c, int image, float* rgba )
{
	if( !c || !rgba || image < 0 || image >= c->g.n_images )
		return GS2D_ERR_INVALID;
	CK( cudaSetDevice( c->device ) );
	const size_t n_px = (size_t)c->g.W * c->g.H, bytes = n_px * sizeof( float4 );
	int rc = ensure_staging( c, bytes );
	if( rc )
		return rc;
	CK( cudaMemcpyAsync( c->pinned, c->b.image + (size_t)image * n_px, bytes, cudaMemcpyDeviceToHost, c->stream ) );
	CK( cudaStreamSynchronize( c->stream ) );
	memcpy( rgba, c->pinned, bytes );
	return GS2D_OK;
}

extern "C" int gs2d_download_grads( gs2d_ctx* c, int image, gs2d_splat* dsplats )
{
	if( !c || ( !dsplats && c->g.n_splats > 0 ) || image < 0 || image >= c->g.n_images )
		return GS2D_ERR_INVALID;
	CK( cudaSetDevice( c->device ) );
	return download_aos( c, image, nullptr, nullptr, dsplats );
}

extern "C" int gs2d_last_mse( gs2d_ctx* c, double* mse )
{
	if( !c || !mse )
		return GS2D_ERR_INVALID;
	CK( cudaSetDevice( c->device ) );
	int rc = resolve( c ); // everything enqueued has run (re-running iterations skipped for lack of list capacity)
	if( rc && rc != GS2D_ERR_NONFINITE )
		return rc;
	rc = ensure_staging( c, sizeof( double ) * ( (size_t)c->g.n_images + 1 ) );
	if( rc )
		return rc;
	CK( cudaMemcpyAsync( c->pinned, c->b.loss, sizeof( double ) * c->g.n_images, cudaMemcpyDeviceToHost, c->stream ) );
	CK( cudaStreamSynchronize( c->stream ) );
	const double* src = (const double*)c->pinned;
	for( int i = 0; i < c->g.n_images; i++ )
		mse[i] = src[i] / (double)( c->g.H * c->g.W * 3 ); // main.cpp:805
	return GS2D_OK;
}

// Pipelined read-back: the caller enqueues iteration i+1 before it waits for the loss of iteration i, so the device
// never idles across the host round trip. The copies ride the compute stream right behind the iteration they belong
// to; waiting is on the ticket's own event, not on the stream.
extern "C" int gs2d_mse_ticket( gs2d_ctx* c, uint64_t* ticket )
{
	if( !c || !ticket )
		return GS2D_ERR_INVALID;
	CK( cudaSetDevice( c->device ) );
	if( !c->ticket_pin )
	{
		c->ticket_stride = ( sizeof( DevState ) + sizeof( double ) * (size_t)c->g.n_images + 63 ) / 64 * 64;
		CK( cudaHostAlloc( (void**)&c->ticket_pin, c->ticket_stride * gs2d_ctx::N_TICKETS, cudaHostAllocMapped ) );
		for( auto& t : c->tickets )
			CK( cudaEventCreateWithFlags( &t.ev, cudaEventDisableTiming ) );
	}
	const uint64_t id = c->next_ticket++;
	gs2d_ctx::Ticket& t = c->tickets[id % gs2d_ctx::N_TICKETS];
	unsigned char* slot = c->ticket_pin + ( id % gs2d_ctx::N_TICKETS ) * c->ticket_stride;
	c->launches += launch_publish( c->b, c->g.n_images, slot, c->stream ); // stores into mapped host memory
	CK( cudaEventRecord( t.ev, c->stream ) );
	t.id = id;
	t.expected = c->expected_iteration;
	*ticket = id;
	return GS2D_OK;
}

extern "C" int gs2d_mse_wait( gs2d_ctx* c, uint64_t ticket, double* mse )
{
	if( !c || !mse || ticket == 0 )
		return GS2D_ERR_INVALID;
	gs2d_ctx::Ticket& t = c->tickets[ticket % gs2d_ctx::N_TICKETS];
	if( t.id != ticket )
	{
		snprintf( c->err, sizeof( c->err ), "mse ticket %llu is unknown or older than the %d newest", (unsigned long long)ticket,
		          gs2d_ctx::N_TICKETS );
		return GS2D_ERR_INVALID;
	}
	CK( cudaSetDevice( c->device ) );
	CK( cudaEventSynchronize( t.ev ) );
	const unsigned char* slot = c->ticket_pin + ( ticket % gs2d_ctx::N_TICKETS ) * c->ticket_stride;
	DevState s;
	memcpy( &s, slot, sizeof( s ) );
	if( s.peer_fault || s.need_capacity || (long long)s.iteration < t.expected )
		return gs2d_last_mse( c, mse ); // an iteration was skipped (list capacity): recover, report the newest loss
	const double* loss = reinterpret_cast<const double*>( slot + sizeof( DevState ) );
	for( int i = 0; i < c->g.n_images; i++ )
		mse[i] = loss[i] / (double)( c->g.H * c->g.W * 3 ); // main.cpp:805
	return s.nonfinite ? GS2D_ERR_NONFINITE : GS2D_OK;
}

extern "C" int gs2d_get_stats( gs2d_ctx* c, gs2d_stats* out )
{
	if( !c || !out )
		return GS2D_ERR_INVALID;
	CK( cudaSetDevice( c->device ) );
	int rc = fetch_state( c );
	if( rc )
		return rc;
	unsigned long long st[8] = { 0 };
	CK( cudaMemcpy( st, c->b.stats, sizeof( st ), cudaMemcpyDeviceToHost ) );
	out->keys = c->h_state->n_keys;
	out->live_pairs = st[0];
	out->staged = st[1];
	out->capacity = c->b.capacity;
	out->live_pairs_sum = st[2];
	out->staged_sum = st[3];
	out->keys_sum = st[4];
	out->n_forwards = st[5];
	out->live_pairs_backward = c->bwd_live;
	return GS2D_OK;
}

extern "C" int gs2d_sync( gs2d_ctx* c )
{
	if( !c )
		return GS2D_ERR_INVALID;
	CK( cudaSetDevice( c->device ) );
	int rc = resolve( c );
	if( rc == GS2D_OK || rc == GS2D_ERR_NONFINITE )
		cudaEventElapsedTime( &c->last_ms, c->ev[0], c->ev[1] );
	return rc;
}

extern "C" int gs2d_last_step_ms( gs2d_ctx* c, float* ms )
{
	if( !c || !ms )
		return GS2D_ERR_INVALID;
	*ms = c->last_ms;
	return GS2D_OK;
}

extern "C" int gs2d_launch_count( gs2d_ctx* c, uint64_t* n )
{
	if( !c || !n )
		return GS2D_ERR_INVALID;
	*n = c->launches;
	return GS2D_OK;
}

extern "C" int gs2d_set_profiling( gs2d_ctx* c, int enable )
{
	if( !c )
		return GS2D_ERR_INVALID;
	CK( cudaSetDevice( c->device ) );
	int rc = resolve( c );
	if( rc && rc != GS2D_ERR_NONFINITE )
		return rc;
	c->profiling = enable != 0;
	if( enable )
		CK( cudaMemset( c->b.stats + 2, 0, sizeof( unsigned long long ) * 6 ) );
	for( double& v : c->prof_sum )
		v = 0.0;
	c->prof_iters = 0;
	return GS2D_OK;
}

extern "C" int gs2d_set_stats( gs2d_ctx* c, int enable )
{
	if( !c )
		return GS2D_ERR_INVALID;
	c->count_live = enable != 0;
	return GS2D_OK;
}

extern "C" int gs2d_get_profile( gs2d_ctx* c, double sums_ms[6], long long* n_iters )
{
	if( !c || !sums_ms || !n_iters )
		return GS2D_ERR_INVALID;
	CK( cudaSetDevice( c->device ) );
	int rc = resolve( c );
	if( rc && rc != GS2D_ERR_NONFINITE )
		return rc;
	for( int k = 0; k < 6; k++ )
		sums_ms[k] = c->prof_sum[k];
	*n_iters = c->prof_iters;
	return GS2D_OK;
}

extern "C" int gs2d_profile_step( gs2d_ctx* c, float out_ms[6] )
{
	if( !c || !out_ms )
		return GS2D_ERR_INVALID;
	const bool was = c->profiling;
	int rc = gs2d_set_profiling( c, 1 );
	if( rc )
		return rc;
	rc = enqueue_iteration( c );
	if( rc )
		return rc;
	double sums[6];
	long long n = 0;
	rc = gs2d_get_profile( c, sums, &n );
	for( int k = 0; k < 6; k++ )
		out_ms[k] = n ? (float)( sums[k] / (double)n ) : 0.0f;
	c->profiling = was;
	return rc;
}

// ---- multi-GPU --------------------------------------------------------------------------------------------------
extern "C" int gs2d_set_band( gs2d_ctx* c, int rank, int n_ranks )
{
	if( !c || n_ranks <= 0 || rank < 0 || rank >= n_ranks )
		return GS2D_ERR_INVALID;
	CK( cudaSetDevice( c->device ) );
	CK( cudaStreamSynchronize( c->stream ) );
	set_band( c, rank, n_ranks );
	CK( cudaMemset( c->b.loss_partial, 0, sizeof( double ) * (size_t)c->g.n_tiles_total ) );
	c->have_forward = false;
	return GS2D_OK;
}

extern "C" int gs2d_comm_unique_id( uint8_t id[128] )
{
	if( !id || !nccl_load() )
		return GS2D_ERR_NCCL;
	ncclUniqueId_ u;
	if( g_nccl.GetUniqueId( &u ) )
		return GS2D_ERR_NCCL;
	memcpy( id, u.internal, 128 );
	return GS2D_OK;
}

// Map the other ranks' gradient, parameter and Adam-moment planes into this process (CUDA IPC; NVLink peer access)
// for the fused reduce + Adam + broadcast kernel. Collective: every rank calls it from gs2d_comm_init. If any rank
// cannot map any peer (no P2P, IPC unavailable in this container, GS2D_PEER_ADAM=0) ALL ranks keep the NCCL path.
static int setup_peer_update( gs2d_ctx* c )
{
	c->peer_update = false;
	const int n = c->n_ranks;
	if( n > GS2D_MAX_RANKS || !g_nccl.AllGather )
		return GS2D_OK;
	cudaStream_t s = c->stream;
	const int NH = 6;
	struct Blob
	{
		cudaIpcMemHandle_t h[6];
		int ok;
		int pad[3];
	};
	Blob mine{};
	if( !c->mbox )
	{
		CK( cudaMalloc( &c->mbox, peer_mailbox_bytes( c->g.n_images ) ) );
		CK( cudaMemset( c->mbox, 0, peer_mailbox_bytes( c->g.n_images ) ) );
		CK( cudaMalloc( &c->peer_seq, 2 * sizeof( unsigned long long ) ) );
		CK( cudaMemset( c->peer_seq, 0, 2 * sizeof( unsigned long long ) ) );
		const size_t touch_bytes = sizeof( unsigned ) * ( (size_t)c->g.n_pad / 128 + 1 ), nq = (size_t)c->g.n_pad / 4;
		CK( cudaMalloc( &c->touch_mem, touch_bytes ) );
		CK( cudaMemset( c->touch_mem, 0, touch_bytes ) );
		CK( cudaMalloc( &c->reach_prev, 2 * nq ) );
		CK( cudaMemset( c->reach_prev, 0xff, 2 * nq ) );
		CK( cudaMalloc( &c->reach_now, nq ) );
		CK( cudaMemset( c->reach_now, 0xff, nq ) );
		CK( cudaMalloc( &c->peer_traffic, 2 * sizeof( unsigned long long ) ) );
		CK( cudaMemset( c->peer_traffic, 0, 2 * sizeof( unsigned long long ) ) );
	}
	void* local[6] = { c->b.grads, c->b.params, c->b.adam_m, c->b.adam_v, c->mbox, c->touch_mem };
	mine.ok = gs2d_tune( "GS2D_PEER_ADAM", 1 ) ? 1 : 0;
	c->peer_flags = gs2d_tune( "GS2D_PEER_FLAGS", 1 ) != 0;
	for( int k = 0; k < NH && mine.ok; k++ )
		if( cudaIpcGetMemHandle( &mine.h[k], local[k] ) != cudaSuccess )
			mine.ok = 0;
	cudaGetLastError();
	Blob* dev = nullptr;
	CK( cudaMalloc( &dev, sizeof( Blob ) * ( (size_t)n + 1 ) ) );
	CK( cudaMemcpyAsync( dev + n, &mine, sizeof( Blob ), cudaMemcpyHostToDevice, s ) );
	int e = g_nccl.AllGather( dev + n, dev, sizeof( Blob ), NCCL_UINT8, c->comm, s );
	std::vector<Blob> all( (size_t)n );
	if( !e )
	{
		CK( cudaMemcpyAsync( all.data(), dev, sizeof( Blob ) * (size_t)n, cudaMemcpyDeviceToHost, s ) );
		CK( cudaStreamSynchronize( s ) );
	}
	int ok = e ? 0 : 1;
	for( int r = 0; r < n && ok; r++ )
		ok = all[r].ok;
	PeerPtrs pp{};
	pp.n_ranks = n;
	pp.rank = c->rank;
	pp.sum_all = 1;
	for( int r = 0; r < n && ok; r++ )
	{
		void* ptr[6] = { local[0], local[1], local[2], local[3], local[4], local[5] };
		if( r != c->rank )
			for( int k = 0; k < NH && ok; k++ )
			{
				if( cudaIpcOpenMemHandle( &ptr[k], all[r].h[k], cudaIpcMemLazyEnablePeerAccess ) != cudaSuccess )
				{
					ok = 0;
					cudaGetLastError();
				}
				else
					c->ipc_opened.push_back( ptr[k] );
			}
		pp.grads[r] = (const float*)ptr[0];
		pp.params[r] = (float*)ptr[1];
		c->peer_m[r] = (float*)ptr[2];
		c->peer_v[r] = (float*)ptr[3];
		pp.mbox[r] = (unsigned char*)ptr[4];
		pp.touch[r] = (const unsigned*)ptr[5];
	}
	// sparse exchange from three ranks on: with two, 94 % of the quads are touched by either rank and more than half of
	// the Gaussians reach either band, so the bitmaps cost more than they save (849 vs 841 it/s); GS2D_PEER_SPARSE=1/0
	// forces it (every rank derives the same default from n)
	const int sparse_knob = gs2d_tune( "GS2D_PEER_SPARSE", -1 );
	pp.sparse = sparse_knob < 0 ? ( n > 2 ? 1 : 0 ) : ( sparse_knob ? 1 : 0 );
	pp.probe = gs2d_tune( "GS2D_PEER_PROBE", 0 );
	pp.reach_prev = c->reach_prev;
	pp.reach_now = c->reach_now;
	pp.traffic = gs2d_tune( "GS2D_PEER_TRAFFIC", 1 ) ? c->peer_traffic : nullptr;
	pp.band_tile_rows = ( c->g.tiles_y + n - 1 ) / n; // set_band
	// agree: the peer path is used only if every rank mapped every peer
	int* flag = reinterpret_cast<int*>( dev );
	CK( cudaMemcpyAsync( flag, &ok, sizeof( int ), cudaMemcpyHostToDevice, s ) );
	e = g_nccl.AllReduce( flag, flag, 1, NCCL_INT32, NCCL_MIN, c->comm, s );
	int agreed = 0;
	if( !e )
	{
		CK( cudaMemcpyAsync( &agreed, flag, sizeof( int ), cudaMemcpyDeviceToHost, s ) );
		CK( cudaStreamSynchronize( s ) );
	}
	cudaFree( dev );
	if( e )
		return nccl_fail( c, "peer setup collective", e );
	c->b.touch = ( agreed && pp.sparse ) ? c->touch_mem : nullptr; // dense exchange or NCCL path: k_preprocess sets no bits
	if( agreed )
	{
		c->peers = pp;
		owned_quads( c, c->rank, &c->own_q0, &c->own_q1 );
		c->peer_update = true;
		if( pp.sparse )
		{
			// k_preprocess computes the reach masks of the quads this rank owns (PeerPtrs::reach_now)
			c->b.reach.now = c->reach_now;
			c->b.reach.q0 = c->own_q0;
			c->b.reach.q1 = c->own_q1;
			c->b.reach.n_ranks = n;
			c->b.reach.band_px = pp.band_tile_rows * GS2D_TILE;
			c->b.reach.slack = 1.0f + 3.0f * c->hp.lr * 1.001f + c->hp.lr;
		}
	}
	else
	{
		for( void* p : c->ipc_opened )
			cudaIpcCloseMemHandle( p );
		c->ipc_opened.clear();
	}
	return GS2D_OK;
}

extern "C" int gs2d_comm_mode( const gs2d_ctx* c, int* n_ranks, int* peer_memory )
{
	if( !c )
		return GS2D_ERR_INVALID;
	if( n_ranks )
		*n_ranks = c->comm ? c->n_ranks : 1;
	if( peer_memory )
		*peer_memory = ( c->comm && c->peer_update ) ? 1 : 0;
	return GS2D_OK;
}

extern "C" int gs2d_comm_traffic( gs2d_ctx* c, uint64_t* bytes_read, uint64_t* bytes_written, int* sparse, int reset )
{
	if( !c )
		return GS2D_ERR_INVALID;
	unsigned long long t[2] = { 0, 0 };
	if( c->comm && c->peer_update && c->peer_traffic )
	{
		CK( cudaSetDevice( c->device ) );
		CK( cudaStreamSynchronize( c->stream ) );
		CK( cudaMemcpy( t, c->peer_traffic, sizeof( t ), cudaMemcpyDeviceToHost ) );
		if( reset )
			CK( cudaMemset( c->peer_traffic, 0, sizeof( t ) ) );
	}
	if( bytes_read )
		*bytes_read = 16ull * t[0];
	if( bytes_written )
		*bytes_written = 16ull * t[1];
	if( sparse )
		*sparse = ( c->comm && c->peer_update && c->peers.sparse ) ? 1 : 0;
	return GS2D_OK;
}

extern "C" int gs2d_comm_init( gs2d_ctx* c, int rank, int n_ranks, const uint8_t id[128] )
{
	if( !c || !id || n_ranks <= 0 || rank < 0 || rank >= n_ranks )
		return GS2D_ERR_INVALID;
	if( !nccl_load() )
	{
		snprintf( c->err, sizeof( c->err ), "libnccl.so.2 not found" );
		return GS2D_ERR_NCCL;
	}
	CK( cudaSetDevice( c->device ) );
	int rc = gs2d_set_band( c, rank, n_ranks );
	if( rc )
		return rc;
	if( n_ranks == 1 )
	{
		if( gs2d_tune( "GS2D_PEER_SELF", 0 ) )
		{
			// developer knob: run the peer-memory update path with this rank as its only peer, so that its kernels can be
			// profiled on one GPU (ncu cannot replay a kernel that waits for another process)
			if( !c->mbox )
			{
				CK( cudaMalloc( &c->mbox, peer_mailbox_bytes( c->g.n_images ) ) );
				CK( cudaMemset( c->mbox, 0, peer_mailbox_bytes( c->g.n_images ) ) );
				CK( cudaMalloc( &c->peer_seq, 2 * sizeof( unsigned long long ) ) );
				CK( cudaMemset( c->peer_seq, 0, 2 * sizeof( unsigned long long ) ) );
				const size_t touch_bytes = sizeof( unsigned ) * ( (size_t)c->g.n_pad / 128 + 1 ), nq = (size_t)c->g.n_pad / 4;
				CK( cudaMalloc( &c->touch_mem, touch_bytes ) );
				CK( cudaMemset( c->touch_mem, 0, touch_bytes ) );
				CK( cudaMalloc( &c->reach_prev, 2 * nq ) );
				CK( cudaMemset( c->reach_prev, 0xff, 2 * nq ) );
				CK( cudaMalloc( &c->reach_now, nq ) );
				CK( cudaMemset( c->reach_now, 0xff, nq ) );
				CK( cudaMalloc( &c->peer_traffic, 2 * sizeof( unsigned long long ) ) );
				CK( cudaMemset( c->peer_traffic, 0, 2 * sizeof( unsigned long long ) ) );
			}
			PeerPtrs pp{};
			pp.n_ranks = 1;
			pp.rank = 0;
			pp.sum_all = 1;
			pp.grads[0] = c->b.grads;
			pp.params[0] = c->b.params;
			pp.mbox[0] = c->mbox;
			pp.touch[0] = c->touch_mem;
			pp.sparse = gs2d_tune( "GS2D_PEER_SPARSE", -1 ) ? 1 : 0; // (unset: sparse, what this mode is there to profile)
			pp.probe = gs2d_tune( "GS2D_PEER_PROBE", 0 );
			pp.reach_prev = c->reach_prev;
			pp.reach_now = c->reach_now; // stays 0xff: everything reaches the only rank
			pp.traffic = c->peer_traffic;
			pp.band_tile_rows = c->g.tiles_y;
			c->peer_m[0] = c->b.adam_m;
			c->peer_v[0] = c->b.adam_v;
			c->b.touch = pp.sparse ? c->touch_mem : nullptr;
			c->peers = pp;
			// GS2D_PEER_SELF=k > 1: own only 1/k of the quads, like one rank of k (the update of the others is skipped)
			const int k = std::max( 1, gs2d_tune( "GS2D_PEER_SELF", 0 ) );
			const int n4 = c->g.n_pad / 4;
			c->own_q0 = 0;
			c->own_q1 = ( n4 + k - 1 ) / k;
			c->peer_update = true;
			c->self_peer = true;
			c->peer_flags = true;
		}
		return GS2D_OK;
	}
	ncclUniqueId_ u;
	memcpy( u.internal, id, 128 );
	int e = g_nccl.CommInitRank( &c->comm, n_ranks, u, rank );
	if( e )
	{
		snprintf( c->err, sizeof( c->err ), "ncclCommInitRank failed: %s",
		          g_nccl.GetErrorString ? g_nccl.GetErrorString( e ) : "?" );
		c->comm = nullptr;
		return GS2D_ERR_NCCL;
	}
	return setup_peer_update( c );
}
