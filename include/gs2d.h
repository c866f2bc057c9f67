/* include/gs2d.h — C ABI of the B200-native 2D Gaussian-splatting image-fitting engine.
 *
 * The reference (Ushio/2dGaussianSplatting) has no plugin/FFI boundary: its optimisation step is the body of
 * `while( pr::NextFrame() == false )` in main() (main.cpp:334, hot part main.cpp:414-809) operating on locals.
 * This header therefore puts the boundary at the DATA CONTRACT of that loop body: the AoS types below are
 * byte-identical to the reference's `Splat` (main.cpp:85-93) and `SplatAdam` (main.cpp:158-166); the target and
 * the rendered image are row-major RGBA float32 exactly like `pr::Image2DRGBA32` (prp.cpp:347-352, 751-756).
 * A patched main.cpp-style driver calls gs2d_step() in place of lines 414-809.
 *
 * Conventions
 *   - every function returns GS2D_OK (0) or a negative GS2D_ERR_* code; nothing aborts or throws across the ABI
 *   - the caller owns every host pointer; the library copies (through its own pinned staging)
 *   - a context is driven by ONE host thread (mirrors the reference); distinct contexts are independent
 *   - there is NO CPU fallback: gs2d_create fails with GS2D_ERR_NO_DEVICE unless an sm_100 device is present
 *   - "image" arguments index independent fitting problems (BASELINE config 5); pass 0 for a single image
 */
#ifndef GS2D_H
#define GS2D_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GS2D_VERSION 100

enum
{
	GS2D_OK = 0,
	GS2D_ERR_INVALID = -1,   /* bad argument */
	GS2D_ERR_NO_DEVICE = -2, /* no sm_100 CUDA device / driver */
	GS2D_ERR_CUDA = -3,      /* a CUDA call failed; gs2d_last_error() has the text */
	GS2D_ERR_NONFINITE = -4, /* non-aborting mirror of main.cpp:752-785 (colour, sx, sy, rot or pos.x not finite) */
	GS2D_ERR_OOM = -5,       /* device or pinned allocation failed */
	GS2D_ERR_NCCL = -6,      /* NCCL missing or a collective failed */
	GS2D_ERR_STATE = -7,     /* call sequence error (e.g. backward before forward) */
	GS2D_ERR_IO = -8,        /* file missing / unreadable / unwritable */
	GS2D_ERR_FORMAT = -9     /* not a file this library understands, or it does not match the context */
};

/* == struct Splat, 36 bytes (main.cpp:85-93); also the layout of dSplats (main.cpp:550) */
typedef struct gs2d_splat
{
	float pos[2];
	float sx, sy;
	float rot;
	float color[3];
	float opacity;
} gs2d_splat;

/* == struct Adam (main.cpp:139-143) */
typedef struct gs2d_moment
{
	float m, v;
} gs2d_moment;

/* == struct SplatAdam, 72 bytes (main.cpp:158-166) */
typedef struct gs2d_splat_adam
{
	gs2d_moment pos[2];
	gs2d_moment sx, sy;
	gs2d_moment rot;
	gs2d_moment color[3];
	gs2d_moment opacity;
} gs2d_splat_adam;

/* Every literal of the reference's loop body, defaulting to the reference's values (gs2d_default_params). */
typedef struct gs2d_params
{
	float lr;               /* 0.05   main.cpp:715 */
	float beta1, beta2;     /* .9 .99 main.cpp:136-137 */
	float eps;              /* 1e-15  main.cpp:154 */
	float bounds;           /* 3.0    main.cpp:7   (only 3.0 is supported by the CUDA path) */
	float min_throughput;   /* 1/256  main.cpp:8 */
	float s_min, s_max;     /* 1,1024 main.cpp:744-745 */
	float o_min, o_max;     /* .1, 1  main.cpp:749 */
	int optimize_opacity;   /* 0      main.cpp:317 */
} gs2d_params;

/* Per-iteration work counters (device-side counts, for the roofline). */
typedef struct gs2d_stats
{
	uint64_t keys;          /* (tile, splat) list entries binned this iteration */
	uint64_t live_pairs;    /* (pixel, splat) pairs blended with T >= min_throughput, forward pass */
	uint64_t staged;        /* list entries actually staged by the forward kernel before tile termination */
	uint64_t capacity;      /* current capacity of the key buffers (entries) */
	uint64_t live_pairs_sum; /* running sums over all forwards since create / the last gs2d_set_profiling */
	uint64_t staged_sum;
	uint64_t keys_sum;
	uint64_t n_forwards;
	uint64_t live_pairs_backward; /* live pairs counted by the last gs2d_backward (same set => same count) */
} gs2d_stats;

typedef struct gs2d_ctx gs2d_ctx;

void gs2d_default_params( gs2d_params* p );
int gs2d_version( void );
const char* gs2d_strerror( int code );
const char* gs2d_last_error( const gs2d_ctx* ctx ); /* text of the last CUDA/NCCL failure ("" if none) */

/* Create a context on CUDA device `device` for `n_images` independent W x H problems with n_splats Gaussians each.
 * params == NULL selects gs2d_default_params. */
int gs2d_create( gs2d_ctx** out, int W, int H, int n_splats, int n_images, int device, const gs2d_params* params );
void gs2d_destroy( gs2d_ctx* ctx );

/* Target = imageRef (main.cpp:254-258): H*W*4 floats, row-major RGBA, A ignored. */
int gs2d_set_target_rgba32f( gs2d_ctx* ctx, int image, const float* rgba );
/* Target from 8-bit RGBA as decoded by stb (prp.cpp:185): converted on device as u8/255.0f (prp.cpp:796-807). */
int gs2d_set_target_rgba8( gs2d_ctx* ctx, int image, const uint8_t* rgba );
/* The same without blocking: rgba must be page-locked host memory (else GS2D_ERR_INVALID) and stay untouched until the
 * next gs2d_step / gs2d_forward / gs2d_sync has been issued. The copy runs on its own stream into one of two staging
 * buffers, so the upload of the NEXT target overlaps the iterations still running on the current one; the new target
 * takes effect at the next forward pass. Once gs2d_comm_init has sharded the image, both calls upload only the rows
 * of this rank's band (the other rows of the buffer are not read). */
int gs2d_set_target_rgba8_async( gs2d_ctx* ctx, int image, const uint8_t* rgba );

/* init() of the reference (main.cpp:280-305) evaluated on the device for every image; resets Adam moments,
 * beta powers and the iteration counter ("Restart" button, main.cpp:828-831). */
int gs2d_init_reference( gs2d_ctx* ctx );

/* Checkpoint format == the reference's loop-carried state (main.cpp:271-278). adam may be NULL (zeros) on upload
 * and NULL (skipped) on download; scalars may be NULL on download. */
int gs2d_upload_state( gs2d_ctx* ctx, int image, const gs2d_splat* splats, const gs2d_splat_adam* adam, float beta1t,
                       float beta2t, int iteration );
int gs2d_download_state( gs2d_ctx* ctx, int image, gs2d_splat* splats, gs2d_splat_adam* adam, float* beta1t,
                         float* beta2t, int* iteration );

int gs2d_set_optimize_opacity( gs2d_ctx* ctx, int enabled ); /* the ImGui checkbox, main.cpp:825 */
int gs2d_get_optimize_opacity( const gs2d_ctx* ctx, int* enabled );

/* n_iters full iterations (binning, forward, backward, [gradient exchange], Adam), main.cpp:414-809.
 * mse: NULL => enqueue only (asynchronous); else n_iters*n_images doubles, mse[it*n_images+img] = the value the
 * reference prints at main.cpp:807 (render made with the pre-update parameters), and the call blocks. */
int gs2d_step( gs2d_ctx* ctx, int n_iters, double* mse );

/* The three phases separately, for parity tests. forward = bin + sort + blend + loss (main.cpp:414-546, 796-805);
 * backward = main.cpp:548-712 (always includes d/d opacity); adam = main.cpp:714-785. */
int gs2d_forward( gs2d_ctx* ctx );
int gs2d_backward( gs2d_ctx* ctx );
int gs2d_adam( gs2d_ctx* ctx );

int gs2d_download_image_rgba32f( gs2d_ctx* ctx, int image, float* rgba ); /* image0 with w == 1 (main.cpp:788-792) */
int gs2d_download_grads( gs2d_ctx* ctx, int image, gs2d_splat* dsplats );  /* == dSplats (main.cpp:550) */
int gs2d_last_mse( gs2d_ctx* ctx, double* mse /* n_images */ );            /* loss of the last forward */
int gs2d_get_stats( gs2d_ctx* ctx, gs2d_stats* out );                      /* counters of the last forward */

/* Pipelined loss read-back for callers that stream targets (the reference prints the mse of EVERY frame,
 * main.cpp:807; reading it with gs2d_last_mse drains the device once per iteration). gs2d_mse_ticket enqueues, behind
 * everything enqueued so far, a device->host copy of the loss of the newest enqueued iteration and returns at once;
 * gs2d_mse_wait blocks until THAT copy has landed (iterations enqueued after the ticket keep running) and returns
 * the same value gs2d_last_mse would have returned at the ticket's position in the stream. The 8 newest tickets
 * can be waited for; an older one returns GS2D_ERR_INVALID. If an iteration had to be re-run because the key
 * buffers grew, the wait falls back to gs2d_last_mse (newest loss). */
int gs2d_mse_ticket( gs2d_ctx* ctx, uint64_t* ticket );
int gs2d_mse_wait( gs2d_ctx* ctx, uint64_t ticket, double* mse /* n_images */ );

/* Tile-row band sharding across ranks (BASELINE config 3/4): this context renders and back-propagates only tile
 * rows [rank*ceil(R/n), ...) of every image; gradients and the loss are summed over ranks before Adam.
 * gs2d_comm_unique_id fills a 128-byte NCCL id on rank 0; the caller ships it to the other ranks (any transport)
 * and every rank calls gs2d_comm_init. One process per GPU. */
int gs2d_comm_unique_id( uint8_t id[128] );
int gs2d_comm_init( gs2d_ctx* ctx, int rank, int n_ranks, const uint8_t id[128] );
/* How the ranks exchange gradients. peer_memory = 1: gs2d_comm_init mapped every peer's gradient / parameter / moment
 * planes through CUDA IPC (NVLink peer access) and the exchange is ONE kernel per rank that sums the peers' gradients
 * for the slice of Gaussians it owns, runs Adam on it and stores the new parameters into every rank's planes (the
 * Adam moments are then sharded; gs2d_download_state gathers them). peer_memory = 0: NCCL all-reduce of the nine
 * gradient planes followed by the identical Adam on every rank (chosen automatically when a peer cannot be mapped,
 * or with the environment variable GS2D_PEER_ADAM=0). */
int gs2d_comm_mode( const gs2d_ctx* ctx, int* n_ranks, int* peer_memory );
/* NVLink payload of the peer-memory exchange since the context was created or the counters were last reset: bytes of
 * gradient quads this rank read out of its peers' memory and bytes of parameter quads it wrote into it (counted by the
 * exchange kernel itself; 0 on one GPU and on the NCCL path). *sparse = 1 when the exchange skips what a rank cannot
 * have touched / will not look at (the default from three ranks on; environment variable GS2D_PEER_SPARSE=0 / 1: every
 * quad to and from every rank / sparse also at two ranks). Waits for the stream. Any pointer may be NULL. */
int gs2d_comm_traffic( gs2d_ctx* ctx, uint64_t* bytes_read, uint64_t* bytes_written, int* sparse, int reset );
/* Band sharding WITHOUT a communicator (host-side tests on one GPU: render bands separately and add them up). */
int gs2d_set_band( gs2d_ctx* ctx, int rank, int n_ranks );

/* Device time of the last completed gs2d_step call, milliseconds, CUDA events on the context's stream. */
int gs2d_last_step_ms( gs2d_ctx* ctx, float* ms );
/* Per-kernel-group device times of one extra profiled iteration (does advance the fit by one iteration):
 * out[0]=preprocess+scan+emit, [1]=sort, [2]=forward, [3]=backward, [4]=allreduce, [5]=adam+finalise. */
int gs2d_profile_step( gs2d_ctx* ctx, float out_ms[6] );
/* Continuous profiling: while enabled, every iteration enqueued by gs2d_step is bracketed by CUDA events on the
 * context's stream; gs2d_get_profile waits for the stream and returns the summed device time of each kernel group
 * (same order as gs2d_profile_step) and the number of iterations measured. gs2d_set_profiling resets the sums.
 * With the forward -> backward overlap on (the default, see GS2D_PDL below) the two blend grids run concurrently and
 * only the pair can be bracketed: group 2 (forward) then holds two memsets and group 3 (backward) both passes. */
int gs2d_set_profiling( gs2d_ctx* ctx, int enable );
int gs2d_get_profile( gs2d_ctx* ctx, double sums_ms[6], long long* n_iters );
/* Developer knobs (process-wide; same names as the environment variables they override): "GS2D_CULL" 0/1 = bin
 * with the conservative tile rectangle only / with the per-tile-row ellipse spans (default; same images, fewer keys),
 * "GS2D_FWD_MINB" 32/24/20 and "GS2D_BWD_MINB" 24/20/16 = register caps of the blend kernels,
 * "GS2D_SLOTS" 2/1/0 = group rows (default: one key per (Gaussian, group of 4 tiles), sorted once per group) / slot
 * rows per (tile, index slice) / every (tile, Gaussian) pair through the place pass (same lists from all three),
 * "GS2D_SCAN_FUSED" 1/0 = bucket totals + scan + chained block totals in one kernel (default) / three kernels,
 * "GS2D_RT_LIST", "GS2D_BAND_LIST" 1/0 = a rank with a band walks lists of the Gaussians that reach it (default) / every
 * Gaussian in every kernel (same results), "GS2D_K1_PROBE" = timing knock-outs of the binning kernel (results WRONG),
 * "GS2D_PDL" 1/0 = backward grid launched inside the forward grid's tail (default; needs the default register
 * caps) / the two grids serialised (same results), "GS2D_GRAPH" 1/0 = iterations replayed as a CUDA graph. */
int gs2d_set_tuning( const char* name, int value );
/* Work counters on/off (default off): while off, the forward blend launched by gs2d_step does not count live pairs
 * (one instruction per 8x4 block visit); gs2d_forward / gs2d_backward always count. */
int gs2d_set_stats( gs2d_ctx* ctx, int enable );
/* Number of kernel launches the library has issued on this context so far. */
int gs2d_launch_count( gs2d_ctx* ctx, uint64_t* n );

int gs2d_sync( gs2d_ctx* ctx );

/* Dimensions the context was created with (any pointer may be NULL). */
int gs2d_get_dims( const gs2d_ctx* ctx, int* W, int* H, int* n_splats, int* n_images );

/* ---- the steps either side of the path (host code, gs2d_io.cpp / gs2d_jpeg.cpp) -------------- */

/* f3 — checkpoint = the reference's loop-carried state {splats, splatAdams, beta1t, beta2t, iterations}
 * (main.cpp:271-278) of every image of the context, plus optimizeOpacity (main.cpp:317). The reference itself saves
 * nothing; the file layout is documented in gs2d_io.cpp. Loading requires equal n_splats/n_images. */
int gs2d_save_checkpoint( gs2d_ctx* ctx, const char* path );
int gs2d_load_checkpoint( gs2d_ctx* ctx, const char* path );

/* f2 — target ingest (prp.cpp:179-193): decode a JPEG (baseline or progressive Huffman, 8 bit, 1 or 3 components,
 * any sampling factors) to RGBA8 with the same integer arithmetic as stb_image's decoder, which the reference uses
 * through prlib (libs/prlib/libs/stb/stb_image.h): bit-identical pixels, A = 255. Binary PPM (P6, maxval 255) is
 * accepted too. *rgba is malloc'ed; release it with gs2d_image_free. */
int gs2d_image_load_rgba8( const char* path, uint8_t** rgba, int* W, int* H );
int gs2d_image_decode_jpeg_rgba8( const uint8_t* bytes, size_t n_bytes, uint8_t** rgba, int* W, int* H );
void gs2d_image_free( uint8_t* rgba );
/* Headless stand-in for tex0->upload( image0 ) (main.cpp:794): round( clamp( c, 0, 1 ) * 255 ) as binary PPM. */
int gs2d_image_write_ppm( const char* path, const float* rgba32f, int W, int H );

/* f4 — the overlay main() draws per splat (main.cpp:421-476) as data: the two eigen-axes scaled by sqrt(lambda)
 * (eignValues :187-196, eigen_vectors_of_cov :223-234), the nvtx-segment ellipse polyline (:453-461, prlib's
 * CircleGenerator recurrence prp.hpp:276-313) and the exact half-extent of the 1-sigma bounding box (:464-465).
 * y is image-down here (the reference flips it for GL). ellipse: n * (nvtx + 1) * 2 floats or NULL. */
typedef struct gs2d_overlay
{
	float center[2];
	float axis0[2], axis1[2]; /* eigenvector * sqrt(eigenvalue), axis0 = larger */
	float half_extent[2];     /* sqrt( inv_cov[1][1] * det ), sqrt( inv_cov[0][0] * det ) */
} gs2d_overlay;
int gs2d_overlay_geometry( const gs2d_splat* splats, int n, int nvtx, gs2d_overlay* out, float* ellipse );

/* Microbenchmarks that give the roofline denominators on the device at hand:
 * FP32 FMA issue rate (TFLOP/s, FMA = 2 flop) and HBM copy bandwidth (GB/s, read+write bytes). */
int gs2d_microbench_fp32( int device, double* tflops );
int gs2d_microbench_hbm( int device, double* gbs );

/* The loss of main.cpp:796-805, mean over 3*W*H of (255*(image - target))^2, on DEVICE buffers the caller owns, without
 * allocating anything. This entry point has no counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it). It calls no cudaMalloc / cudaFree / allocating library routine and needs no gs2d_ctx.
 *   image_dev, target_dev  row-major RGBA float32 (16-byte aligned) of W x H pixels on `device`; alpha is ignored
 *   mse_dev                one double on `device`, written by the second of the two kernels enqueued on cuda_stream
 *   cuda_stream            a cudaStream_t (NULL = the default stream); the call returns once both kernels are enqueued
 *   err, err_cap           NULL/0, or a buffer that receives the text of a failure ("" on success)
 * The per-tile partial sums live in a file-scope static __device__ array of GS2D_MSE_STATIC_MAX_TILES doubles
 * (512 KiB, held by every GPU that loads the library until the process exits): images of more 16x16 tiles than that
 * (e.g. beyond 4096 x 4096) are rejected with GS2D_ERR_INVALID. All calls on a device share the array: calls that
 * overlap on the GPU (two streams) race - serialise them. Nothing in it has to be cleared between calls (each call
 * overwrites exactly the range it then reads). The summation order is the one of gs2d_forward + gs2d_last_mse, so the
 * value is the same bit for bit. The call makes `device` current; the first call per device looks the array's
 * address up (cudaGetSymbolAddress) - make that one outside a stream capture. */
#define GS2D_MSE_STATIC_MAX_TILES 65536
int gs2d_mse_static( int device, const float* image_dev, const float* target_dev, int W, int H, double* mse_dev,
                     void* cuda_stream, char* err, size_t err_cap );

#ifdef __cplusplus
}
#endif
#endif /* GS2D_H */
