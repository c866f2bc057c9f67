// gs2d_internal.cuh — shared declarations of the sm_100a kernels behind include/gs2d.h.
//
// Device data layout (all in HBM, owned by gs2d_ctx):
//   params / grads / adam_m / adam_v : 9 SoA planes of n_total_pad floats each, plane order == field order of the
//       reference's Splat (main.cpp:85-93): pos.x pos.y sx sy rot color.r color.g color.b opacity.
//       n_total = n_images * n_splats; global splat index g = image * n_splats + i.
//   rec      : one 64-byte record per splat, rewritten every iteration by k_preprocess (SplatRec below)
//   bucket_cnt / bucket_slots / bucket_off / triples / vals: the sort-free binning of gs2d_binning.cu (see Buffers)
//   tile_range: [start,end) of every tile's list, ordered by splat index
//   target / image : RGBA float32, row-major, n_images * H * W float4 (== pr::Image2DRGBA32, prp.cpp:751-756)
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "../../include/gs2d.h"

#define GS2D_TILE 16          // tile edge in pixels; one warp per tile, 8 pixels per lane (gs2d_blend.cuh)
#define GS2D_BATCH 32         // list entries staged into shared memory per round
#define GS2D_NPLANES 9
#define GS2D_BUCKET_SLOTS 8   // GS2D_SLOTS=1: direct slots per (tile, index slice) bucket, 16-bit entries (gs2d_binning.cu);
                              // the default binning keeps 16 32-bit keys per (group of 4 tiles, slice): GS2D_GROUP_SLOTS, gs2d_bins.cuh
#define GS2D_LIVE_SLOTS 64    // spread of the live-pair counter atomics
#define GS2D_FWD_MINB_DEFAULT 32 // resident warps per SM the forward blend is register-capped for (env GS2D_FWD_MINB)
#define GS2D_BWD_MINB_DEFAULT 24 // same for the backward blend (env GS2D_BWD_MINB)

enum
{
	PL_PX = 0,
	PL_PY,
	PL_SX,
	PL_SY,
	PL_ROT,
	PL_CR,
	PL_CG,
	PL_CB,
	PL_OP
};

// 64-byte per-splat record (4 x float4), rewritten every iteration by k_preprocess. q0..q2 are what the forward pass
// reads; q3 holds the per-splat invariants hoisted out of the backward pass (cos, sin, sx^-3, sy^-3,
// (sx^2-sy^2)/(sx^2 sy^2)).
// Two flavours of (A, B, D, op), chosen by the arithmetic mode of the context (GS2D_EXACT, gs2d_blend.cuh):
//   exact   : A, B, D = inv_cov[0][0], inv_cov[1][0], inv_cov[1][1] (a, b, d) and op = opacity: the blend kernels
//             evaluate d2, exp_approx and alpha operation for operation like main.cpp:526-533, never contracted;
//   relaxed : with s = opacity^(1/8):  A = -s*a/16, B = -s*b/8, D = -s*d/16, op = s, so that
//             s * (1 - d2/16) = (s + A vx^2) + vy (B vx + D vy)   [two FMAs per pixel]  and  alpha = (that)^8.
struct __align__( 16 ) SplatRec
{
	float px, py, B, D;        // q0: position; (B, D) and (cg, cb) sit in aligned register pairs (xmul2)
	float A, cr, cg, cb;       // q1
	unsigned idx;              // q2: global splat index (row table / RowRec lookup while the coverage words are built)
	unsigned flags;            //     bit 0: the splat's rows are in the row table (SPLAT_IN_TABLE); then bits 1..20 its
	                           //     first row inside the image, bits 21..27 its number of rows - 1 (<= 127).
	                           //     (idx, flags) are read as one aligned 64-bit word by build_words
	float op;                  //     opacity (exact) / opacity^(1/8) (relaxed)
	float rotc;                //     (sx^2 - sy^2) / (sx^2 sy^2)
	float cosT, sinT, isx3, isy3; // q3
};
static_assert( sizeof( SplatRec ) == 64 && offsetof( SplatRec, idx ) == 32, "SplatRec: 64 bytes, (idx, flags) 8-byte aligned" );
#define SPLAT_IN_TABLE 1u
#define SPLAT_ROW0_SHIFT 1
#define SPLAT_ROW0_BITS 20
#define SPLAT_NROWS_SHIFT 21

// What the exact scan-line footprint needs (main.cpp:489-509), always unscaled; written by k_preprocess for every
// splat, read by k_row_table (every splat of the band) and by the blend's in-kernel fallback (splats too tall for the
// row table only).
struct __align__( 16 ) RowRec
{
	float px, py, a, b;
	float d;
	int begY, endY;   // row range, main.cpp:490-491 (x86 truncation semantics)
	unsigned rows;    // first row | number of rows - 1 << 20 of the splat inside the image AND this rank's band: what
	                  // k_row_table solves; ROWREC_NONE: nothing, or too tall for the table (more than GS2D_ROWTAB_STRIDE rows)
};
static_assert( sizeof( RowRec ) == 32, "RowRec must be 32 bytes" );
#define ROWREC_NONE 0xffffffffu
// Row table: per splat a ring of GS2D_ROWTAB_STRIDE packed row spans (pack_row_span), image row y at ring position
// y % GS2D_ROWTAB_STRIDE: 512 bytes per splat; the 16 rows of a tile row are 64 contiguous bytes. Only the rows inside
// [begY, endY] are written; the blend masks the others out with the row range carried in SplatRec::flags.
#define GS2D_ROWTAB_STRIDE 128

// Loop-carried scalars + flags, one per context, lives in device memory so that gs2d_step can be fully async.
struct DevState
{
	float beta1t, beta2t;      // main.cpp:274-275
	int iteration;             // main.cpp:278
	int nonfinite;             // set by k_adam where the reference would abort() (main.cpp:752-785)
	unsigned long long n_keys; // total list entries of the current iteration (written by the scan)
	unsigned long long need_capacity; // != 0: the key buffers were too small; the iteration was skipped
	int mse_slot;              // next slot of the mse ring written by k_finalize
	int peer_fault;            // a peer-memory barrier timed out (a rank died or the call sequences diverged)
	// The reference abort()s in the iteration whose Adam step produced a non-finite parameter (main.cpp:752-785), after
	// the whole update loop and before that iteration's mse line. Mirror: k_advance of the NEXT iteration sees
	// `nonfinite`, sets `frozen` and returns; from then on k_advance / k_adam are no-ops, so `iteration` stays at the
	// failing iteration's index + 1 and the parameters are the ones the reference held when it aborted.
	int frozen;
	int pad_;
};
static_assert( sizeof( DevState ) == 48, "DevState is copied as 12 words (k_publish) and mirrored by the host" );

struct Geometry
{
	int W, H;
	int n_splats;   // per image
	int n_images;
	int n_total;    // n_images * n_splats
	int n_pad;      // plane stride (multiple of 4)
	int tiles_x, tiles_y;    // per image
	int band_ty0, band_ty1;  // this rank's tile rows [ty0, ty1)
	int n_tiles_total;       // n_images * tiles_y * tiles_x  (key space)
	int nt_pad;              // n_tiles_total rounded up to a multiple of 4
	int n_slices, slice_size; // index slices of an image's splats: slice = i / slice_size, slice_size <= 65535
	float min_throughput;    // MIN_THROUGHPUT, main.cpp:8
	// Target format (not geometry, but it travels to the same kernels and into the same graph key): 1 = the blend
	// kernels read the target as RGBA8 (Buffers::target8, the reference's ingest format, prp.cpp:179-193) and convert
	// with u8_lut[c] == (float)c / 255.0f (prp.cpp:796-807) - 4 bytes per pixel instead of 16, and no conversion pass
	// per uploaded target; 0 = RGBA float32 (Buffers::target, == imageRef).
	int target_u8;
	const float* u8_lut;     // [256]
};

#include "gs2d_footprint.cuh" // contraction-free helpers, exp_approx, exact row footprint, tile-row spans

// ---------------------------------------------------------------------------------------------------------------
// cp.async (LDGSTS) 16-byte gather into shared memory.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16( void* smem_dst, const void* gmem_src )
{
	unsigned s = (unsigned)__cvta_generic_to_shared( smem_dst );
	asm volatile( "cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"( s ), "l"( gmem_src ) );
}
__device__ __forceinline__ void cp_async_commit() { asm volatile( "cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait()
{
	asm volatile( "cp.async.wait_group %0;\n" ::"n"( N ) );
}

// ---------------------------------------------------------------------------------------------------------------
// Launch wrappers (each returns the number of kernels it launched; errors are picked up by cudaGetLastError).
// ---------------------------------------------------------------------------------------------------------------
struct Buffers
{
	float* params;  // [9][n_pad]
	float* grads;   // [9][n_pad]
	float* adam_m;  // [9][n_pad]
	float* adam_v;  // [9][n_pad]
	SplatRec* rec;  // [n_total]
	RowRec* rowrec; // [n_total]
	unsigned* rowtab; // [n_total][GS2D_ROWTAB_STRIDE] packed row spans (k_row_table -> k_forward)
	unsigned* rt_list; // [n_total] a banded rank: the splats k_row_table has to solve this iteration (k_preprocess appends)
	unsigned* band_list; // [n_total] a banded rank: the splats that passed the band test (k_band_filter appends)
	unsigned short* bucket_cnt; // [n_slices][nt_pad] entries per (tile, index slice) bucket, 4 tiles per 64-bit word; behind
	                            // them the counters of the group rows, the lists' lengths and the scan's scratch (gs2d_create)
	unsigned short* bucket_slots; // group rows [nt_pad / 4][n_slices][16] 32-bit keys (default), or slot rows
	                              // [n_slices][nt_pad][GS2D_BUCKET_SLOTS] of 16-bit indices (GS2D_SLOTS=1): same size
	unsigned* bucket_off;       // [n_tiles_total * n_slices] exclusive scan of the counts in (tile, slice) order
	unsigned* scan_sums;        // scratch of the scan
	unsigned long long* triples; // [capacity] unordered (bucket, splat-in-slice, arrival rank) words, gs2d_binning.cu
	unsigned* vals[2];          // vals[1]: the lists bucketed but unordered; vals[0]: ordered by splat index (final)
	unsigned* cover;          // [capacity][8] exact coverage words of the staged list entries (forward -> backward), then
	                          // [capacity / 32 + n_tiles + 8] round masks (round_vis_index)
	unsigned long long capacity;
	uint2* tile_range;        // [n_tiles_total]
	float4* target;           // [n_images][H][W]
	uchar4* target8;          // [n_images][H][W] RGBA8 targets (allocated at the first 8-bit upload; Geometry::target_u8)
	float* u8_lut;            // [256] c / 255.0f
	float4* image;            // [n_images][H][W]
	double* loss_partial;     // [n_tiles_total]
	unsigned* tile_done;      // [n_tiles_total] forward -> backward tile flags of the PDL overlap (gs2d_blend.cuh)
	unsigned* touch;          // [n_pad / 128 + 1] bitmap of the quads that passed this rank's band test (PeerPtrs::touch); nullptr on one GPU
	// sparse parameter broadcast (PeerPtrs::reach_now): k_preprocess computes, for the quads this rank OWNS
	// [reach_q0, reach_q1), the mask of the ranks whose band a Gaussian of the quad can reach
	struct Reach
	{
		unsigned char* now; // [n_pad / 4]; nullptr: not needed
		int q0, q1, n_ranks, band_px;
		float slack;
	} reach;
	unsigned long long* live_slots;   // [GS2D_LIVE_SLOTS]
	unsigned long long* staged_slots; // [GS2D_LIVE_SLOTS], contiguous after live_slots
	unsigned long long* stats;        // [2] live pairs, staged entries of the last forward (summed slots)
	double* loss;             // [n_images]  sum over this rank's band of the squared 8-bit error
	double* mse_ring;         // [mse_cap][n_images]
	int mse_cap;
	DevState* state;
};

int launch_init_reference( const Geometry& g, const Buffers& b, cudaStream_t s );
int launch_u8_to_target( const uint8_t* src_dev, float4* dst, size_t n_px, cudaStream_t s );
int launch_u8_lut( float* lut, cudaStream_t s );
// the reference pixel of the loss / of dL/dC (imageRef, main.cpp:616, :800) in either target format
__device__ __forceinline__ float4 load_target( const Geometry& g, const float4* __restrict__ target, size_t pix )
{
	if( g.target_u8 )
	{
		const uchar4 c = reinterpret_cast<const uchar4*>( target )[pix];
		return make_float4( __ldg( g.u8_lut + c.x ), __ldg( g.u8_lut + c.y ), __ldg( g.u8_lut + c.z ), 1.0f );
	}
	return target[pix];
}
// use_slots: group rows (GS2D_SLOTS=2, default) / slot rows (=1) unless the context has switched to the triple path for
// now (gs2d_ctx::slot_binning: most arrivals overflow their rows) or GS2D_SLOTS=0
// side: a second stream the row table kernel runs on, concurrently with the scan / place / rank kernels (fork and
// join through ev_fork / ev_join; also inside a stream capture); nullptr: everything on s
struct SideStream
{
	cudaStream_t stream;
	cudaEvent_t ev_fork, ev_join;
};
int launch_preprocess_and_bin( const Geometry& g, const Buffers& b, int* sorted_buf, cudaStream_t s,
                               cudaEvent_t after_emit, bool use_slots, const SideStream* side );
// arithmetic mode of the blend kernels: GS2D_EXACT (default 0 = relaxed: FMA contraction + folded opacity)
bool gs2d_exact_mode();
// developer tuning knob: integer environment variable read once per name, else the default
int gs2d_tune( const char* env_name, int dflt );
// pdl: forward/backward overlap through programmatic dependent launch (gs2d_blend.cuh); must be set for both or none,
// backward directly behind forward on the stream, default register caps (gs2d_pdl_usable)
int launch_forward( const Geometry& g, const Buffers& b, int sorted_buf, bool count_live, cudaStream_t s,
                    bool pdl = false );
int launch_backward( const Geometry& g, const Buffers& b, int sorted_buf, bool with_opacity, bool count_live,
                     cudaStream_t s, bool pdl = false );
bool gs2d_pdl_usable();
void prepare_forward_pdl();  // per-device kernel attributes of the PDL twins (called once by gs2d_create)
void prepare_backward_pdl();
int launch_finalize_loss( const Geometry& g, const Buffers& b, cudaStream_t s );
int launch_publish( const Buffers& b, int n_images, unsigned char* slot_mapped_host, cudaStream_t s );
int launch_advance( const Geometry& g, const Buffers& b, const gs2d_params& p, cudaStream_t s );
int launch_adam( const Geometry& g, const Buffers& b, const gs2d_params& p, cudaStream_t s );
// Fused gradient reduction + Adam + parameter broadcast over peer memory (NVLink / NVSwitch, CUDA IPC mappings):
// this rank owns the quads [q0, q1) of every plane; it sums the gradients of all ranks for them straight out of the
// peers' HBM, runs Adam on them (its m, v are the only valid ones for that slice) and stores the new parameters into
// every rank's parameter planes.
#define GS2D_MAX_RANKS 8
struct PeerPtrs
{
	const float* grads[GS2D_MAX_RANKS];
	float* params[GS2D_MAX_RANKS];
	int n_ranks, rank;
	int sum_all; // 1: sum the gradients of all ranks; 0: this rank's gradients are already the total
	// Sparse exchange (GS2D_PEER_SPARSE; the default from three ranks on, dense at two). A Gaussian reaches 1.15 of 8 bands on average, so most of what the
	// dense exchange moves is zeros (gradients) or parameters the receiver never looks at:
	//   touch[r] : rank r's bitmap of the quads (4 consecutive Gaussians) that passed its band test in k_preprocess this
	//              iteration, i.e. that CAN hold a non-zero gradient on r; the owner of a quad reads the gradient quads
	//              of the ranks whose bit is set and takes zeros for the others;
	//   reach    : the owner sends a quad's new parameters only to the ranks whose band a Gaussian of the quad can
	//              reach (rows py +- (3 max(sx, sy) + margin), a superset of the receiver's own band test after one
	//              Adam step) - and once more to a rank it has just stopped reaching, so that the receiver's stale copy
	//              is one its band test rejects. A rank's copy of the parameters is therefore exact for every Gaussian
	//              its kernels look at; gs2d_download_state gathers the rest from the owners.
	const unsigned* touch[GS2D_MAX_RANKS];
	unsigned char* reach_prev; // [2][n_pad / 4] double-buffered by iteration parity: ranks reached at the last exchange
	const unsigned char* reach_now; // [n_pad / 4] ranks a quad's Gaussians reach, computed by k_preprocess of this iteration
	                           // (from the parameters BEFORE the Adam step; the slack of the test covers the step)
	unsigned long long* traffic; // [2] 16-byte quads read from / written to peers over NVLink (running sums)
	int sparse;
	int band_tile_rows;        // tile rows per rank (set_band)
	int probe;                 // developer timing knob GS2D_PEER_PROBE (results are WRONG when set): 1 no remote gradient
	                           // loads, 2 no remote parameter stores, 4 no touch-bitmap loads, 8 no Adam arithmetic
	// mailboxes for the flag barriers (one per rank, mapped everywhere): see PeerMailbox
	unsigned char* mbox[GS2D_MAX_RANKS];
};
// Layout of a rank's mailbox; slot r of every array is written only by rank r (so no atomics are needed):
//   seq1[r], seq2[r]  : sequence numbers of the last barrier-1 / barrier-2 rank r has arrived at (release stores)
//   nf[r]             : rank r's non-finite flag, payload of barrier 2
//   loss[r][0..L)     : rank r's band sums of the loss + its overflow vote, payload of barrier 1 (L = n_images + 1)
struct PeerMailbox
{
	unsigned long long seq1[GS2D_MAX_RANKS];
	unsigned long long seq2[GS2D_MAX_RANKS];
	int nf[GS2D_MAX_RANKS];
	// double loss[GS2D_MAX_RANKS][L] follows
};
static inline size_t peer_mailbox_bytes( int n_images )
{
	return sizeof( PeerMailbox ) + sizeof( double ) * GS2D_MAX_RANKS * ( (size_t)n_images + 1 );
}
// Barrier 1 + all-reduce of the loss over the mailboxes (replaces ncclAllReduce of n_images + 1 doubles), and
// barrier 2 + OR of the non-finite flags. One small block each; seq_ctr: this rank's device-resident counter of the
// barrier kind, bumped by the kernel (every rank launches the same sequence of barriers).
int launch_peer_loss_exchange( const Geometry& g, const Buffers& b, const PeerPtrs& pp, unsigned long long* seq_ctr,
                               cudaStream_t s );
int launch_peer_barrier2( const Geometry& g, const Buffers& b, const PeerPtrs& pp, unsigned long long* seq_ctr,
                          cudaStream_t s );
int launch_adam_peer( const Geometry& g, const Buffers& b, const gs2d_params& p, const PeerPtrs& pp, int q0, int q1,
                      cudaStream_t s );
int launch_aos_to_soa( const Geometry& g, const Buffers& b, int image, const gs2d_splat* splats_dev,
                       const gs2d_splat_adam* adam_dev, cudaStream_t s );
int launch_soa_to_aos( const Geometry& g, const Buffers& b, int image, gs2d_splat* splats_dev,
                       gs2d_splat_adam* adam_dev, gs2d_splat* grads_dev, cudaStream_t s );
