// gs2d_fit.cpp — headless fit driver: the reference's main() (main.cpp:236-334, :794-831) with the
// window removed and the loop body (main.cpp:414-809) replaced by gs2d_step(). C++ host code over the C ABI only.
//
// What main() does, and where it is here:
//   Image2DRGBA8::load( "squirrel_cls_mini.jpg" ) + Image2DRGBA8_to_Image2DRGBA32   -> gs2d_image_load_rgba8 +
//                                                                                     gs2d_set_target_rgba8
//   NSplat = 1024, init()                         (main.cpp:271-307)                -> --splats, gs2d_init_reference
//   one iteration per frame, forever              (main.cpp:334)                    -> --iters
//   printf( "%d itr, mse %.4f\n", iterations, mse ) (main.cpp:807)                  -> the same line on stdout
//   "Optimize opacity" checkbox                   (main.cpp:825)                    -> --opacity-from ITER
//   "Restart" button -> init()                    (main.cpp:828-831)                -> --restart-at ITER
//   tex0->upload( image0 )                        (main.cpp:794)                    -> --out image.ppm (last render)
//   abort() on a non-finite parameter             (main.cpp:752-785)                -> exit code 3 after the message
//   (nothing is saved)                                              -> --save / --load checkpoint
// The stdout of `gs2d_fit squirrel_cls_mini.jpg --iters K` can be diffed against the unmodified reference's stdout
// (tests/golden/c1_mse.txt); tests/test_gpu_parity.py does exactly that.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/gs2d.h"

static void usage()
{
	fprintf( stderr,
	         "usage: gs2d_fit <target.jpg|target.ppm> [--splats N=1024] [--iters K=300] [--device D=0]\n"
	         "                [--opacity-from ITER] [--restart-at ITER] [--load ckpt] [--save ckpt]\n"
	         "                [--out render.ppm] [--overlay overlay.csv] [--print-every M=1] [--exact]\n"
	         "  --exact: the blend kernels keep the reference's operand order without FMA contraction (image bit-identical\n"
	         "           to the CPU reference; default: relaxed arithmetic, results within FP32 round-off of it)\n" );
}

#define CHECK( call )                                                                                                 \
	do                                                                                                                \
	{                                                                                                                 \
		int rc_ = ( call );                                                                                           \
		if( rc_ != GS2D_OK )                                                                                          \
		{                                                                                                             \
			fprintf( stderr, "gs2d_fit: %s -> %s (%s)\n", #call, gs2d_strerror( rc_ ), ctx ? gs2d_last_error( ctx ) : "" ); \
			if( ctx )                                                                                                 \
				gs2d_destroy( ctx );                                                                                  \
			return rc_ == GS2D_ERR_NONFINITE ? 3 : 1;                                                                 \
		}                                                                                                             \
	} while( 0 )

int main( int argc, char** argv )
{
	const char* target = nullptr;
	const char *load = nullptr, *save = nullptr, *out = nullptr, *overlay = nullptr;
	int n_splats = 1024, iters = 300, device = 0, opacity_from = -1, restart_at = -1, print_every = 1;
	for( int i = 1; i < argc; i++ )
	{
		const std::string a = argv[i];
		auto val = [&]() -> const char* { return i + 1 < argc ? argv[++i] : nullptr; };
		const char* v = nullptr;
		if( a == "--splats" && ( v = val() ) )
			n_splats = atoi( v );
		else if( a == "--iters" && ( v = val() ) )
			iters = atoi( v );
		else if( a == "--device" && ( v = val() ) )
			device = atoi( v );
		else if( a == "--opacity-from" && ( v = val() ) )
			opacity_from = atoi( v );
		else if( a == "--restart-at" && ( v = val() ) )
			restart_at = atoi( v );
		else if( a == "--print-every" && ( v = val() ) )
			print_every = std::max( 1, atoi( v ) );
		else if( a == "--load" && ( v = val() ) )
			load = v;
		else if( a == "--save" && ( v = val() ) )
			save = v;
		else if( a == "--out" && ( v = val() ) )
			out = v;
		else if( a == "--overlay" && ( v = val() ) )
			overlay = v;
		else if( a == "--exact" )
			gs2d_set_tuning( "GS2D_EXACT", 1 );
		else if( a[0] != '-' && !target )
			target = argv[i];
		else
		{
			usage();
			return 2;
		}
	}
	if( !target || n_splats < 0 || iters < 0 )
	{
		usage();
		return 2;
	}

	gs2d_ctx* ctx = nullptr;
	uint8_t* rgba8 = nullptr;
	int W = 0, H = 0;
	CHECK( gs2d_image_load_rgba8( target, &rgba8, &W, &H ) );
	int rc = gs2d_create( &ctx, W, H, n_splats, 1, device, nullptr );
	if( rc != GS2D_OK )
	{
		fprintf( stderr, "gs2d_fit: gs2d_create -> %s\n", gs2d_strerror( rc ) );
		gs2d_image_free( rgba8 );
		return 1;
	}
	rc = gs2d_set_target_rgba8( ctx, 0, rgba8 );
	gs2d_image_free( rgba8 );
	CHECK( rc );
	CHECK( gs2d_init_reference( ctx ) );
	if( load )
		CHECK( gs2d_load_checkpoint( ctx, load ) );

	// `iterations` continues from the checkpoint (main.cpp:278 is part of the saved state)
	int iterations = 0;
	CHECK( gs2d_download_state( ctx, 0, nullptr, nullptr, nullptr, nullptr, &iterations ) );

	// The loop. UI events of the reference happen between frames; here they split the run into chunks of gs2d_step.
	std::vector<double> mse;
	int done = 0;
	while( done < iters )
	{
		// events take effect before the iteration whose counter equals their argument
		if( opacity_from >= 0 && iterations >= opacity_from )
			CHECK( gs2d_set_optimize_opacity( ctx, 1 ) );
		if( restart_at >= 0 && iterations == restart_at && done > 0 )
		{
			CHECK( gs2d_init_reference( ctx ) ); // the "Restart" button: init() (main.cpp:828-831)
			iterations = 0;
			restart_at = -1;
		}
		int chunk = std::min( iters - done, 256 );
		if( opacity_from > iterations )
			chunk = std::min( chunk, opacity_from - iterations );
		if( restart_at > iterations )
			chunk = std::min( chunk, restart_at - iterations );
		mse.resize( (size_t)chunk );
		rc = gs2d_step( ctx, chunk, mse.data() );
		if( rc != GS2D_OK && rc != GS2D_ERR_NONFINITE )
			CHECK( rc );
		int valid = chunk;
		if( rc == GS2D_ERR_NONFINITE )
		{
			// The device froze in the iteration whose Adam step produced the non-finite parameter; its counter reads that
			// iteration's index + 1. The reference abort()s there BEFORE printing that iteration's line (main.cpp:752-785
			// precedes :796-807), so only the lines of the iterations before it are printed.
			int it_dev = iterations;
			gs2d_download_state( ctx, 0, nullptr, nullptr, nullptr, nullptr, &it_dev );
			valid = std::max( 0, std::min( chunk, it_dev - 1 - iterations ) );
		}
		for( int k = 0; k < valid; k++, iterations++ )
			if( iterations % print_every == 0 )
				printf( "%d itr, mse %.4f\n", iterations, mse[(size_t)k] ); // main.cpp:807
		fflush( stdout );
		done += chunk;
		if( rc == GS2D_ERR_NONFINITE )
		{
			fprintf( stderr, "gs2d_fit: non-finite splat parameter in iteration %d - the reference aborts here (main.cpp:752-785)\n", iterations );
			if( save )
				gs2d_save_checkpoint( ctx, save );
			gs2d_destroy( ctx );
			return 3;
		}
	}

	if( out )
	{
		// image0 as uploaded to tex0 at the top of the next frame: the render of the LAST iteration's forward pass
		std::vector<float> img( (size_t)W * H * 4 );
		if( iters == 0 )
			CHECK( gs2d_forward( ctx ) );
		CHECK( gs2d_download_image_rgba32f( ctx, 0, img.data() ) );
		CHECK( gs2d_image_write_ppm( out, img.data(), W, H ) );
	}
	if( overlay )
	{
		std::vector<gs2d_splat> s( (size_t)n_splats );
		rc = gs2d_download_state( ctx, 0, s.data(), nullptr, nullptr, nullptr, nullptr );
		if( rc != GS2D_OK && rc != GS2D_ERR_NONFINITE )
			CHECK( rc );
		std::vector<gs2d_overlay> o( (size_t)n_splats );
		CHECK( gs2d_overlay_geometry( s.data(), n_splats, 16, o.data(), nullptr ) );
		FILE* f = fopen( overlay, "w" );
		if( !f )
			CHECK( GS2D_ERR_IO );
		fprintf( f, "cx,cy,axis0x,axis0y,axis1x,axis1y,half_w,half_h,r,g,b,opacity\n" );
		for( int i = 0; i < n_splats; i++ )
			fprintf( f, "%.9g,%.9g,%.9g,%.9g,%.9g,%.9g,%.9g,%.9g,%.9g,%.9g,%.9g,%.9g\n", o[i].center[0], o[i].center[1],
			         o[i].axis0[0], o[i].axis0[1], o[i].axis1[0], o[i].axis1[1], o[i].half_extent[0], o[i].half_extent[1],
			         s[i].color[0], s[i].color[1], s[i].color[2], s[i].opacity );
		fclose( f );
	}
	if( save )
		CHECK( gs2d_save_checkpoint( ctx, save ) );
	gs2d_destroy( ctx );
	return 0;
}
