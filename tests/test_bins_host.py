"""CPU tests of the product's group-row binning (no GPU): 2dgaussiansplatting_b200/csrc/gs2d_bins.cuh - the very header
gs2d_binning.cu includes for its default path (GS2D_SLOTS=2) - is compiled for the host by tests/bins_host.cpp, which
walks it through the kernels' sequence (emission in arbitrary arrival order -> totals -> scan, plain and in the block
layout of the fused kernel -> overflow placement ->
one sort per (group of 4 tiles, index slice) and the deal to the tiles' lists). Every tile's list must hold exactly the
splats whose tile rows cover it, in splat-index order (== the reference's blend order, main.cpp:419), whatever the
arrival order, for tile rows that straddle groups, groups that straddle tile rows (tiles_x not a multiple of 4), rows
that overflow their 16 slots, and a rank's window of tile rows. The comparator networks are checked exhaustively with
the 0-1 principle. The device side of the same path: tests/test_gpu_parity.py (lists equal to the sort-free path's)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def bins(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("bins") / "libbins_host.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", os.path.join(HERE, "bins_host.cpp"), "-o", so],
                   check=True)
    L = C.CDLL(so)
    L.bins_build.argtypes = [C.c_int] * 5 + [C.c_void_p, C.c_longlong, C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p, C.c_int]
    L.bins_build.restype = C.c_longlong
    L.bins_sort16.argtypes = [C.c_void_p]
    L.bins_round_vis_index.argtypes = [C.c_uint, C.c_uint, C.c_uint]
    L.bins_round_vis_index.restype = C.c_longlong
    return L


def test_comparator_networks_sort_every_0_1_input(bins):
    # 0-1 principle: a comparator network that sorts all 2^16 binary inputs sorts every input
    for v in range(1 << 16):
        e = np.array([(v >> i) & 1 for i in range(16)], np.uint32)
        bins.bins_sort16(e.ctypes.data)
        assert np.all(e[:-1] <= e[1:]), v


def make_case(rng, tiles_x, tiles_y, n_splats, n_slices, max_w, max_h):
    """Every splat covers a few tile rows, in each a contiguous span of tiles (like tile_span's output)."""
    slice_size = (n_splats + n_slices - 1) // n_slices
    arrivals, truth = [], [[] for _ in range(tiles_x * tiles_y)]
    for sidx in range(n_splats):
        y0 = rng.integers(0, tiles_y)
        h = rng.integers(1, max_h + 1)
        cx = rng.integers(0, tiles_x)
        for ty in range(y0, min(y0 + h, tiles_y)):
            w = rng.integers(1, max_w + 1)
            lo = max(0, cx - w // 2 + rng.integers(-1, 2))
            hi = min(tiles_x - 1, lo + w - 1)
            if lo > hi:
                continue
            arrivals.append((ty * tiles_x + lo, ty * tiles_x + hi, sidx // slice_size, sidx % slice_size))
            for tx in range(lo, hi + 1):
                truth[ty * tiles_x + tx].append(sidx)
    return np.array(arrivals, np.int32).reshape(-1, 4), truth, slice_size


@pytest.mark.parametrize("tiles_x,tiles_y,n_splats,n_slices,max_w,max_h", [
    (16, 12, 3000, 8, 5, 4),     # the usual shape: spans of 1..5 tiles crossing group boundaries
    (5, 9, 800, 4, 5, 3),        # tiles_x not a multiple of 4: groups straddle tile rows (one splat, two arrivals per row)
    (3, 7, 500, 2, 3, 7),        # tiles_x < 4: a group holds parts of two or three tile rows
    (1, 5, 100, 1, 1, 5),        # a single tile column, one slice
    (8, 4, 4000, 2, 8, 4),       # dense: most rows overflow their 16 slots (overflow arrivals ranked per tile)
    (32, 8, 2000, 64, 6, 3),     # more slices than entries per bucket
])
def test_lists_are_complete_and_in_index_order(bins, tiles_x, tiles_y, n_splats, n_slices, max_w, max_h):
    rng = np.random.default_rng(tiles_x * 1000 + n_splats)
    arr, truth, slice_size = make_case(rng, tiles_x, tiles_y, n_splats, n_slices, max_w, max_h)
    n_tiles = tiles_x * tiles_y
    for trial, (row_lo, row_hi) in enumerate([(0, tiles_y), (0, tiles_y), (tiles_y // 3, max(tiles_y // 3 + 1, 2 * tiles_y // 3))]):
        a = arr[rng.permutation(len(arr))]  # arrival order is arbitrary on the device
        # a rank emits only the tile rows of its band
        rows_of = a[:, 0] // tiles_x
        a = np.ascontiguousarray(a[(rows_of >= row_lo) & (rows_of < row_hi)])
        cap = int(sum(len(t) for t in truth)) + 8
        out = np.full(cap, 0xFFFFFFFF, np.uint32)
        tr = np.full((n_tiles, 2), 0xFFFFFFFF, np.uint32)
        n_over = C.c_longlong(0)
        # trial 0: offsets by the plain scan (k_group_count + k_scan_*), then by k_group_scan's block layout
        total = bins.bins_build(n_tiles, n_slices, slice_size, row_lo * tiles_x, row_hi * tiles_x, a.ctypes.data, len(a),
                                out.ctypes.data, cap, tr.ctypes.data, C.byref(n_over), int(trial > 0))
        want_total = sum(len(truth[t]) for t in range(row_lo * tiles_x, row_hi * tiles_x))
        assert total == want_total
        if (tiles_x, n_splats) == (8, 4000):
            assert n_over.value > 0.3 * total  # the overflow path is what this case is about
        pos = 0
        for t in range(row_lo * tiles_x, row_hi * tiles_x):
            s, e = int(tr[t, 0]), int(tr[t, 1])
            assert s == pos and e - s == len(truth[t]), (t, s, e, pos, len(truth[t]))  # lists are contiguous, in tile order
            assert list(out[s:e]) == sorted(truth[t]), t
            pos = e


def test_round_mask_words_of_all_tiles_are_distinct_and_fit(bins):
    """k_forward leaves one 32-bit mask per round (32 list entries) of a tile for k_backward at round_vis_index. The
    tiles' lists are contiguous and of any length (0, 1, 31, 32, 33, ...): no two (tile, round) pairs may share a
    word, and every index must lie below capacity / 32 + n_tiles + 1 (what gs2d_api.cu allocates behind the coverage
    words)."""
    rng = np.random.default_rng(3)
    for trial in range(20):
        n_tiles = int(rng.integers(1, 400))
        lens = rng.choice([0, 0, 1, 2, 31, 32, 33, 63, 64, 65, 200], n_tiles) if trial % 2 else rng.integers(0, 300, n_tiles)
        starts = np.concatenate([[0], np.cumsum(lens)[:-1]]) + int(rng.integers(0, 1000)) * (trial % 3 == 0)
        total = int(starts[-1] + lens[-1])
        seen = set()
        for t in range(n_tiles):
            for r in range((int(lens[t]) + 31) // 32):
                i = bins.bins_round_vis_index(int(starts[t]), t, r)
                assert i not in seen, (trial, t, r)
                seen.add(i)
                assert 0 <= i < total // 32 + n_tiles + 1, (trial, t, r, i, total)
