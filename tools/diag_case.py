"""Debug helper (GPU box): re-creates one case of tests/test_gpu_parity.py::test_randomised_scenes_vs_oracle and prints
the hypotheses whose flags / models differ from the oracle."""
import math, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "graph-cut-ransac_b200"), os.path.join(ROOT, "tests")]
import numpy as np, gcr_capi, gcr_scenes as gs
from oracle_lib import Oracle
target = int(sys.argv[1])
port = Oracle("port")
rng = np.random.Generator(np.random.PCG64(20261104))
for case in range(target + 1):
    ns, no = int(rng.integers(2, 700)), int(rng.integers(2, 400))
    ratio = float(rng.choice([0.05, 0.2, 0.5, 0.9]))
    lvl = float(rng.choice([0.25, 1.0, 3.0, 8.0, 25.0]))
    sc = gs.make_scene(ns, no, ratio, 5000 + case, thr_scale=0.05 * lvl, thr_orient=math.radians(lvl))
    scale, orient = sc.scale.copy(), sc.orient.copy()
    if case % 3 == 0 and ns > 12 and no > 12:
        scale[3] = scale[2]; scale[5, 2] = 0.0; scale[6, 2] = -1.5; scale[7, :2] = [3e5, -2e5]; scale[8, 2] = 1e-11; scale[9, 2] = 1e11
        orient[4] = orient[1]; orient[6, 2] = orient[5, 2]; orient[8, 2] = orient[5, 2] + math.pi
        orient[9, 2] = 9.5 * math.pi; orient[10, 2] = -2.75 * math.pi
    H = int(rng.integers(1, 1500))
    smp = gs.draw_samples(H, ns, no, 900 + case)
    if case % 3 == 0 and ns > 12 and no > 12 and H > 64:
        smp[:64, :2] = rng.integers(0, 11, (64, 2)); smp[:64, 2:] = rng.integers(0, 11, (64, 2))
        same = smp[:64, 0] == smp[:64, 1]; smp[:64, 1][same] = (smp[:64, 1][same] + 1) % 11
        same = smp[:64, 2] == smp[:64, 3]; smp[:64, 3][same] = (smp[:64, 3][same] + 1) % 11
ora = port.replay(scale, orient, smp, sc.thr, nthreads=0)
e = gcr_capi.Engine(0); e.set_features(scale, orient)
e.run(H, sc.thr, samples=smp, flags=gcr_capi.RUN_KEEP_PER_HYP); dev = e.fetch()
np.set_printoptions(precision=17, linewidth=200)
for h in np.nonzero((dev["flags"] & 7) != (ora["flags"] & 7))[0]:
    print("hyp", h, "sample", smp[h], "dev flags", dev["flags"][h], "oracle flags", ora["flags"][h])
    print(" scale rows", scale[smp[h, 0]], scale[smp[h, 1]]); print(" orient rows", orient[smp[h, 2]], orient[smp[h, 3]])
    print(" dev model", dev["models"][h], "oracle model", ora["models"][h])
