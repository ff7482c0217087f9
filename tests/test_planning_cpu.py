"""Host-side planning logic of the native library (no device needed): TMA pixel boxes and the measured
(N tile, cluster) rules (profiles/pool_fusion.md).  Also proves that the in-tree .so builds and loads on a CPU box."""
import ctypes as C

import pytest

from ddl_b200.ops import native


def _lib():
    try:
        return native.lib()
    except native.NativeError as e:          # no nvcc and no prebuilt library
        pytest.skip("native library unavailable: %s" % e)


def box(N, H, W, max_rows=128, row_mult=1, even=False):
    out = (C.c_int * 3)()
    assert _lib().ddl_plan_choose_box(N, H, W, max_rows, row_mult, int(even), out) == 0
    return tuple(out)


def staged(pix, nout, kiters):
    out = (C.c_int * 2)()
    assert _lib().ddl_plan_pick_staged(pix, nout, kiters, out) == 0
    return tuple(out)


def test_library_loads_on_cpu_and_reports_version():
    assert _lib().ddl_version() >= 100


@pytest.mark.parametrize("N,H,W", [(100, 14, 14), (100, 7, 7), (100, 4, 4), (37, 14, 14), (513, 4, 4), (10, 28, 28)])
def test_pixel_boxes_fit_a_tile_and_cover_the_tensor(N, H, W):
    for even in (False, True):
        bw, bh, bn = box(N, H, W, even=even)
        assert 1 <= bw * bh * bn <= 128
        assert bw == W or W % bw == 0                       # full rows or exact divisors of the width
        if even:                                            # 2x2 pooling windows never straddle two tiles
            assert bw == W or bw % 2 == 0
            assert bh >= H or bh % 2 == 0


def test_wgrad_boxes_are_multiples_of_eight_rows():
    for (N, H, W, cap) in [(100, 14, 14, 48), (100, 7, 7, 56), (100, 4, 4, 40), (100, 1, 1, 128)]:
        bw, bh, bn = box(N, H, W, max_rows=cap, row_mult=8)
        assert (bw * bh * bn) % 8 == 0 and bw * bh * bn <= cap


def test_staged_tiling_rule_reproduces_the_measured_optima():
    # (pixel tiles, output channels, K blocks) of the six fused conv kernels of the reference model at batch 100
    assert staged(175, 64, 25) == (64, 1)        # conv2 fprop + pool
    assert staged(48, 128, 50) == (128, 4)       # conv3 fprop + pool
    assert staged(13, 256, 100) == (128, 8)      # conv4 fprop + pool
    assert staged(168, 32, 50) == (32, 1)        # conv2 dgrad + unpool
    assert staged(42, 64, 100) == (64, 4)        # conv3 dgrad + unpool
    assert staged(13, 128, 200) == (64, 8)       # conv4 dgrad + unpool


def test_staged_tiling_large_batches_use_the_widest_tile_unsplit():
    assert staged(1250, 256, 100) == (256, 1)
    assert staged(2000, 64, 25) == (64, 1)
    nt, cl = staged(168, 192, 150)               # widened model: no pair fits one wave
    assert 192 % nt == 0 and cl == 1


def test_fc_tiling_spreads_the_weight_stream_over_a_wave():
    out = (C.c_int * 2)()
    L = _lib()
    L.ddl_plan_pick_fc(1024, 32, out)
    assert tuple(out) == (32, 4)                 # fc1: 32 N tiles x cluster 4 = 128 CTAs
    L.ddl_plan_pick_fc(512, 32, out)
    assert tuple(out) == (32, 8)                 # fc2: 16 N tiles x cluster 8
    L.ddl_plan_pick_fc(6144, 192, out)
    assert tuple(out) == (32, 1)                 # widened fc1: enough N tiles, no split


def test_fc3_fused_kernel_shape_limits():
    L = _lib()
    assert L.ddl_fc3_fused_fits(100, 512, 10) == 1
    assert L.ddl_fc3_fused_fits(256, 2048, 16) in (0, 1)
    assert L.ddl_fc3_fused_fits(257, 512, 10) == 0      # one row per warp: at most 32 x 8 rows
    assert L.ddl_fc3_fused_fits(100, 3072, 10) == 0     # k-slice of a CTA is at most 256 columns
    assert L.ddl_fc3_fused_fits(100, 510, 10) == 0      # K % 4


def test_patch_resident_conv_selection_rule():
    """csrc/tc_patch.cu host side: the patch kernel needs the padded patch [H+4][W+4][cin] plus a weight ring in 227 KB of
    shared memory and channel counts in multiples of 32; it is PREFERRED only where >= 70% of its M rows are real pixels
    (pad columns are M rows): the reference model's conv2 (14x14) yes, conv3 (7x7: 38%) and conv4 (4x4) no
    (profiles/patch_conv_r2.md)."""
    L = _lib()
    assert L.ddl_patch_conv_fits(14, 14, 32, 64) == 1 and L.ddl_patch_conv_preferred(14, 14, 32, 64) == 1     # conv2
    assert L.ddl_patch_conv_fits(7, 7, 64, 128) == 1 and L.ddl_patch_conv_preferred(7, 7, 64, 128) == 0       # conv3
    assert L.ddl_patch_conv_preferred(4, 4, 128, 256) == 0                                                    # conv4
    assert L.ddl_patch_conv_fits(14, 14, 33, 64) == 0                          # channels not a multiple of 32
    assert L.ddl_patch_conv_fits(14, 14, 192, 384) == 0                        # widened x6 conv2: the patch does not fit
    assert L.ddl_patch_conv_preferred(14, 14, 64, 128) == 1                    # widened x2 conv2 still does
    for H, cin, cout in [(14, 32, 64), (7, 64, 128), (28, 32, 32), (10, 32, 96)]:
        if L.ddl_patch_conv_preferred(H, H, cin, cout):
            rpt = 128 // (H + 4)
            tiles = -(-H // rpt)
            assert H * H / (tiles * 128.0) >= 0.7


def test_committed_sass_evidence_belongs_to_the_sources_in_the_tree():
    """profiles/sass_evidence.md and profiles/sass/*.txt are generated from the built library (tools/sass_extract.py, also
    run by ``python -m ddl_b200.ops.build -v``); the digest they quote must be the digest of the sources as they are now."""
    import os, re
    from ddl_b200.ops import build
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(root, "profiles", "sass_evidence.md")).read()
    m = re.search(r"Library source digest: `([0-9a-f]{64})`", text)
    assert m, "no digest in profiles/sass_evidence.md"
    assert m.group(1) == build._digest(), "profiles/sass* are stale: run python tools/sass_extract.py"
    for name in ("tc_act_gemm_kernel", "tc_wgrad_kernel", "tc_patch_conv_kernel"):
        listing = open(os.path.join(root, "profiles", "sass", name + ".txt")).read()
        assert "UTCHMMA" in listing and "UTMALDG" in listing and m.group(1) in listing
    # the whole-library counts of the table are what cuobjdump says about the shipped .so right now
    import shutil, subprocess
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if os.path.exists(exe) and not build.needs_build():
        sass = subprocess.run([exe, "-sass", build.LIB], capture_output=True, text=True).stdout
        row = re.search(r"\| \*\*whole library\*\* \| \| (\d+) \| (\d+) \| (\d+) \| (\d+) \|", text)
        assert row, "no whole-library row in profiles/sass_evidence.md"
        lines = sass.splitlines()
        count = lambda k: sum(1 for ln in lines if re.search(r"(?<![A-Z])" + k, ln))
        assert [int(x) for x in row.groups()] == [count("UTCHMMA"), count("UTMALDG"), count("UTMASTG"), count("LDTM")]
        assert count("HMMA") == 0                            # no legacy mma.sync anywhere


def test_baseline_tables_quote_the_committed_bench_lines():
    """BASELINE.md section 6 is regenerated from the JSON lines under profiles/r2_*: every line's samples/s value must
    appear in the document (thousands separators), so a table row cannot outlive the measurement it came from."""
    import glob
    import json
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(root, "BASELINE.md")).read()
    files = sorted(glob.glob(os.path.join(root, "profiles", "r2_8gpu", "*.json")) +
                   glob.glob(os.path.join(root, "profiles", "r2_scale", "*.json")))
    assert len(files) >= 15
    seen = 0
    for f in files:
        for line in open(f):
            line = line.strip()
            if not (line.startswith("{") and '"metric"' in line):
                continue
            d = json.loads(line)
            if os.path.basename(f) == "async_dedicated_1ps_1w_n2.json":
                continue                                  # 2-GPU plumbing check of the dedicated layout, not a table row
            v = "{:,}".format(int(round(d["value"])))
            assert v in text, "%s: %s samples/s is not quoted in BASELINE.md" % (os.path.relpath(f, root), v)
            if isinstance(d.get("e2e"), dict) and d.get("impl", "ours") == "ours" and "r2_8gpu" in f:
                e = "{:,}".format(int(round(d["e2e"]["value"])))
                assert e in text, "%s: e2e %s is not quoted in BASELINE.md" % (os.path.relpath(f, root), e)
            seen += 1
    assert seen >= 15
