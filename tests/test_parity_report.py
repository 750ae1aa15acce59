"""tools/parity_report.py on the CPU: the report's bookkeeping (offender listing, strip selection, the reference-undefined anti-alias pixels)."""
import importlib.util
import os
import numpy as np
from conftest import ROOT

_spec = importlib.util.spec_from_file_location("parity_report", os.path.join(ROOT, "tools", "parity_report.py"))
PR = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(PR)


def test_undefined_pixels_are_the_h_pixels_of_survey_a1():
    """with anti-aliasing, column x = W-1 for y >= 1 (3 of 9 samples each) and pixel (W-2, H-1) read past the end of pixel_loc."""
    assert PR.undefined_pixels(320, 180, False) == []
    px = PR.undefined_pixels(320, 180, True)
    assert len(px) == 180 and (319, 1) in px and (319, 179) in px and (318, 179) in px and (319, 0) not in px


def test_columns_of_interleaved_strips():
    assert np.array_equal(PR.columns_of(40, 1, 0), np.arange(40))
    cols = PR.columns_of(100, 4, 1)                                   # strips of 16: s = 1, 5 -> columns 16..31, 80..95
    assert np.array_equal(cols, np.concatenate([np.arange(16, 32), np.arange(80, 96)]))
    from b200rt.sharding import columns_of_rank
    for world in (2, 4, 8):
        for r in range(world):
            assert np.array_equal(PR.columns_of(7680, world, r), columns_of_rank(7680, r, world, 16))


def test_compare_lists_every_offender_and_applies_the_tolerance():
    rng = np.random.default_rng(0)
    W, H = 64, 48
    ref = rng.integers(0, 256, (3, W, H), dtype=np.uint8)
    got = ref.copy()
    got[0, 5, 7] = np.clip(int(ref[0, 5, 7]) + 1, 0, 255)             # 1 level: differing, not an offender (unless clipped to 0 difference)
    got[1, 10, 11] = (int(ref[1, 10, 11]) + 40) % 256                 # far off: offender
    got[2, 63, 20] = (int(ref[2, 63, 20]) + 9) % 256                  # offender in the reference-undefined column
    cols = np.arange(W)
    rep = PR.compare(got, ref, cols, W, H, True)
    xs = sorted((o[0], o[1]) for o in rep["offenders"])
    assert xs == [(10, 11), (63, 20)] and rep["pixels_over_1_level"] == 2
    assert rep["pixels_compared"] == W * H and rep["within_tolerance"]            # 2 <= max(2, 0.1 % of 3072)
    o = [o for o in rep["offenders"] if (o[0], o[1]) == (10, 11)][0]
    assert o[2] == [int(v) for v in ref[:, 10, 11]] and o[3] == [int(v) for v in got[:, 10, 11]]
    und = rep["reference_undefined"]
    assert und["count"] == H and [63, 20] in [d[:2] for d in und["differing"]]
    # a restricted column set only sees its own pixels
    rep2 = PR.compare(got, ref, PR.columns_of(W, 2, 0), W, H, True)               # strips 0, 2: columns 0..15, 32..47
    assert [(o[0], o[1]) for o in rep2["offenders"]] == [(10, 11)] and rep2["reference_undefined"]["count"] == 0
    many = ref.copy(); many[:, :8, :] ^= 0x80
    assert not PR.compare(many, ref, cols, W, H, False)["within_tolerance"]
