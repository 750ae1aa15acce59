import os
import sys
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "python-ray-tracer_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def pytest_sessionfinish(session, exitstatus):
    """Checked build of the library (tools/gpu_check.sh): report the index-assertion failures of the whole session."""
    if os.environ.get("B200RT_REPORT_CHECKS"):
        from b200rt import _lib
        if _lib._lib is not None:
            f = int(_lib._lib.b200rt_debug_check_failures())
            print("\n[b200rt checked build] index-assertion failures in this session:", "not a checked build" if f == 2**64 - 1 else f)
            if f not in (0, 2**64 - 1):
                session.exitstatus = 3


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return {k: z[k] for k in z.files}


def golden_regions(g):
    """[(x0,x1,y0,y1, expected uint8 (3,x1-x0,y1-y0))]"""
    if "result" in g:
        return [(0, int(g["W"]), 0, int(g["H"]), g["result"])]
    return [(int(r[0]), int(r[1]), int(r[2]), int(r[3]), g[f"result_{k}"]) for k, r in enumerate(g["regions"])]


def pixel_loc_of(g):
    """Rebuild the float64 (3,W,H) pixel grid from the stored closed form (checked bit-exact at generation time)."""
    W, H = int(g["W"]), int(g["H"])
    f, ar, sy, sz = [float(v) for v in g["grid"]]
    out = np.empty((3, W, H))
    out[0] = f
    out[1] = (np.arange(W) * sy + ar)[:, None]
    out[2] = (np.arange(H) * sz + 1.0)[None, :]
    return out


def parity_report(got, ref):
    """SURVEY §8(d) parity report for uint8 (3,...) arrays: max |d|, fraction of pixels with any channel |d|>1."""
    d = np.abs(got.astype(np.int16) - ref.astype(np.int16))
    px = d.max(axis=0)
    bad = np.argwhere(px > 1)
    return dict(max=int(d.max()) if d.size else 0, n=int(px.size), n_diff=int((px > 0).sum()), n_bad=int(len(bad)),
                frac_bad=float(len(bad)) / max(1, px.size), bad=[tuple(int(v) for v in b) for b in bad[:50]])


FRAME_GOLDENS = ["c1_aa0", "c1_aa1", "rand_a", "rand_b", "rand_c", "edge_s0", "edge_l0", "edge_p0", "edge_d0", "c4_small"]
CROP_GOLDENS = ["c2_aa0_crops", "c2_aa1_crops", "c3_crops", "c4_crops", "c5_crops_k0", "c5_crops_k23"]


def have_golden(name):
    return os.path.exists(os.path.join(GOLDEN, name + ".npz"))


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle_f64
    oracle_f64.build()
    return oracle_f64
