#!/usr/bin/env python
"""oracle/patched/make_patched.py - TEST INFRASTRUCTURE: the drop-in boundary exercised from the reference's own main().

Applies the patch of INTEGRATION.md to a COPY of /root/reference/main.cpp (never committed: the copy is written to a
temporary directory and only the linked binary is kept, as oracle/_ref/patched_main): the loop body main.cpp:414-809
(forward, backward, Adam, clamps, finite check, MSE) is cut out and replaced by gs2d_step(); init() (main.cpp:280-305)
keeps filling `splats` on the host exactly as the reference does and then hands them over with gs2d_upload_state(),
so that the "Restart" button (main.cpp:828-831) still works. Everything else of main() is the reference's text,
compiled against the same headless shim as oracle level 0 and linked with -lgs2d.

The binary prints the reference's "%d itr, mse %.4f" lines; tests/test_gpu_driver.py diffs them with level 0's
(tests/golden/c1_mse.txt) on the GPU box. Needs /root/reference, so it is built in the build container only
(`make -C oracle patched`, run by __graft_entry__.build()).
"""
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE = os.path.dirname(HERE)
ROOT = os.path.dirname(ORACLE)
REF = os.environ.get("REF", "/root/reference")

CREATE = r'''
	// ---- gs2d drop-in (INTEGRATION.md): the context owns the device state; created once the target is known ----
	gs2d_ctx* gs = nullptr;
	if( gs2d_create( &gs, imageRef.width(), imageRef.height(), NSplat, 1, 0, nullptr ) != GS2D_OK )
	{
		fprintf( stderr, "patched main: gs2d_create failed (no sm_100 device?)\n" );
		return 1;
	}
	static_assert( sizeof( Splat ) == sizeof( gs2d_splat ) && sizeof( SplatAdam ) == sizeof( gs2d_splat_adam ), "AoS layouts" );
	gs2d_set_target_rgba32f( gs, 0, (const float*)imageRef.data() ); // Image2DRGBA32 is row-major vec4 (prp.cpp:751-756)
'''
UPLOAD = r'''
		// gs2d drop-in: the state init() has just written becomes the device state (Adam moments zero, beta powers 1)
		gs2d_upload_state( gs, 0, (const gs2d_splat*)splats.data(), nullptr, beta1t, beta2t, iterations );
'''
STEP = r'''
		// ---- gs2d drop-in: replaces main.cpp:414-809 (forward, backward, Adam + clamps, finite check, MSE) ----------
		gs2d_set_optimize_opacity( gs, optimizeOpacity );                 // the checkbox, consumed at main.cpp:735
		double mse = 0.0;
		{
			const int rc = gs2d_step( gs, 1, &mse );                      // one loop body; mse = the value of main.cpp:805
			if( rc == GS2D_ERR_NONFINITE )
				abort();                                                  // main.cpp:752-785
			if( rc != GS2D_OK )
			{
				fprintf( stderr, "patched main: gs2d_step -> %s (%s)\n", gs2d_strerror( rc ), gs2d_last_error( gs ) );
				return 1;
			}
		}
		gs2d_download_image_rgba32f( gs, 0, (float*)image0.data() );      // only because tex0->upload wants host memory
		// the loop-carried locals (main.cpp:271-278), for the overlay / a debugger / whoever reads them next frame
		gs2d_download_state( gs, 0, (gs2d_splat*)splats.data(), (gs2d_splat_adam*)splatAdams.data(), &beta1t, &beta2t, nullptr );
		tex0->upload(image0);                                             // main.cpp:794
		printf( "%d itr, mse %.4f\n", iterations, mse );                  // main.cpp:807
		iterations++;                                                     // main.cpp:809
'''


def patch(lines):
    """lines: main.cpp as a list (index 0 = line 1). Returns the patched text. Every anchor is asserted."""
    L = lambda n: lines[n - 1]
    assert "std::fill( image0.data()" in L(414), L(414)
    assert "iterations++;" in L(809), L(809)
    assert "tex0->upload(image0);" in L(794) and "printf(" in L(807)
    assert "int iterations = 0;" in L(278), L(278)
    assert L(305).strip() == "};" and "splats[i] = s;" in L(303), (L(303), L(305))
    assert L(307).strip() == "init();"
    first_include = next(i for i, l in enumerate(lines) if l.startswith("#include"))
    out = []
    for i, l in enumerate(lines, start=1):
        if i - 1 == first_include:
            out.append('#include "gs2d.h" // gs2d drop-in\n')
        if 414 <= i <= 809:
            if i == 414:
                out.append(STEP)
            continue
        if i == 305:  # end of the init lambda: hand the fresh state to the device
            out.append(UPLOAD)
        out.append(l)
        if i == 278:
            out.append(CREATE)
    return "".join(out)


def main():
    src = os.path.join(REF, "main.cpp")
    out_bin = os.path.join(ORACLE, "_ref", "patched_main")
    if not os.path.exists(src):
        print(f"oracle: {REF} absent - keeping prebuilt _ref/patched_main")
        return 0
    lib_dir = os.path.join(ROOT, "2dgaussiansplatting_b200")
    if not os.path.exists(os.path.join(lib_dir, "libgs2d.so")):
        print("oracle: libgs2d.so not built yet - build it first (make -C 2dgaussiansplatting_b200/csrc)")
        return 1
    os.makedirs(os.path.dirname(out_bin), exist_ok=True)
    with tempfile.TemporaryDirectory() as d:
        patched = os.path.join(d, "main_patched.cpp")
        with open(src) as f:
            text = patch(f.readlines())
        with open(patched, "w") as f:
            f.write(text)
        cmd = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-w",
               "-I" + os.path.join(ORACLE, "shim"), "-I" + os.path.join(REF, "libs/prlib/src"),
               "-I" + os.path.join(REF, "libs/prlib/libs/stb"), "-I" + os.path.join(ROOT, "include"),
               patched, os.path.join(ORACLE, "shim", "shim.cpp"), "-o", out_bin,
               "-L" + lib_dir, "-lgs2d", "-Wl,-rpath,$ORIGIN/../../2dgaussiansplatting_b200", "-lm"]
        subprocess.run(cmd, check=True)
    print("oracle: built", out_bin)
    return 0


if __name__ == "__main__":
    sys.exit(main())
