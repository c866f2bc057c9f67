/* oracle/gs2d_oracle.h — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * "Oracle level 1": a function-ised, plain-C restatement of the reference's optimisation step
 * (/root/reference/main.cpp:414-809) with the image size, Gaussian count, target and state passed in.
 * It exists because oracle level 0 (the unmodified main.cpp behind oracle/shim, built into oracle/_ref/ref_main)
 * hard-wires one JPEG name and 1024 splats (main.cpp:257, :271) and keeps dSplats / Adam state in locals.
 * Parity status: PINNED — tests/test_oracle.py proves this file bit-identical (rendered image every iteration and
 * the printed mse stream) to level 0 on both reference JPEGs; tests/golden/ holds the level-0 outputs.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this library.
 * Must be compiled with -ffp-contract=off (the reference build has no FMA contraction).
 */
#ifndef GS2D_ORACLE_H
#define GS2D_ORACLE_H

#include "../include/gs2d.h"

#ifdef __cplusplus
extern "C" {
#endif

/* pcg3d, main.cpp:17-24 */
void oracle_pcg3d( uint32_t v[3] );
/* init(), main.cpp:280-305 (Adam state is simply zeros, beta powers 1) */
void oracle_init_splats( int W, int H, int n, gs2d_splat* splats );

/* Forward, main.cpp:414-546. image: W*H*4 floats out, w reset to 1. counts (may be NULL): [0] visited pairs
 * (inside the scan-line footprint and the image), [1] live pairs (T >= min_throughput). */
void oracle_forward( int W, int H, int n, const gs2d_splat* splats, const gs2d_params* p, float* image,
                     uint64_t counts[2] );
/* Backward, main.cpp:548-712. image = output of oracle_forward; scratch = W*H*4 floats (image1); dsplats out. */
void oracle_backward( int W, int H, int n, const gs2d_splat* splats, const gs2d_params* p, const float* image,
                      const float* target, float* scratch, gs2d_splat* dsplats );
/* Adam + clamps + finite check, main.cpp:714-785. Returns 0, or 1 where the reference would abort(). */
int oracle_adam( int W, int H, int n, gs2d_splat* splats, gs2d_splat_adam* adam, const gs2d_splat* dsplats,
                 float* beta1t, float* beta2t, const gs2d_params* p );
/* MSE, main.cpp:796-805. */
double oracle_mse( int W, int H, const float* image, const float* target );

/* One whole loop body. image and scratch: W*H*4 floats each. Returns the mse the reference prints (main.cpp:807);
 * *nonfinite (may be NULL) receives oracle_adam's return. dsplats may be NULL. */
double oracle_step( int W, int H, int n, gs2d_splat* splats, gs2d_splat_adam* adam, float* beta1t, float* beta2t,
                    const float* target, const gs2d_params* p, float* image, float* scratch, gs2d_splat* dsplats,
                    int* nonfinite, uint64_t counts[2] );

#ifdef __cplusplus
}
#endif
#endif
