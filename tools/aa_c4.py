import sys, os, time
sys.path.insert(0, "python-ray-tracer_b200"); sys.path.insert(0, ".")
import numpy as np
import b200rt
from b200rt.workloads import workload
w = workload("c4")
cam = w["cameras"][0]
for aa in (True,):
    t0 = time.time()
    img, st = b200rt.render_frame(w["scene"], cam, w["amb"], w["lamb"], w["refl"], w["refl_depth"], aa, device=0, stats=True)
    print("AA", aa, img.shape, "rays", st["rays"], "ms_total", round(st["ms_total"], 2), "wall", round(time.time() - t0, 2), "sum", int(img.sum()))
    t0 = time.time()
    img2 = b200rt.render_frame(w["scene"], cam, w["amb"], w["lamb"], w["refl"], w["refl_depth"], aa, device=0)
    print("second frame wall", round(time.time() - t0, 3), "equal", bool(np.array_equal(img, img2)))
