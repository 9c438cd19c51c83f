"""SURVEY.md 8(f) row f4 through the C ABI with HOST inputs (copies inside the timed call; pyramids device-resident, as a tracker
keeps them), next to the oracle port on ONE host thread (these are per-frame, sequential calls in the reference: Tracking thread /
LocalMapping thread).  Each GPU result is checked against the oracle before timing.  One JSON line."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb-ygz-slam_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as O  # noqa: E402
import synth  # noqa: E402
import ygzorb_ctypes as Y  # noqa: E402
from test_direct_oracle import synth_observations  # noqa: E402


def timed(fn, reps):
    fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    return (time.perf_counter() - t0) / reps


def main():
    mt = Y.Matcher(0)
    out = {}
    # ---- SparseImgAlign::run: 1000 features, levels 7..1
    c = synth.sparse_align_case(1, n=1000)
    args = {k: v for k, v in c.items() if k != "t_true"}
    ref, cur = Y.Pyramid(c["ref_levels"]), Y.Pyramid(c["cur_levels"])
    sa = Y.SparseAlign(mt)

    def gpu_align():
        sa.begin(ref, cur, c["kp_xy"], c["valid"], c["world_pos"], c["ref_pose"], c["focal"], c["cur_intr"], c["inv_scale"])
        return sa.run(7, 1, c["t_cur_ref"])
    T, nt, stats = gpu_align()
    eT, en, estats = O.sparse_align_run(max_level=7, min_level=1, **args)
    assert nt == en and np.abs(T - eT).max() < 2e-4
    g = timed(gpu_align, 50)
    cpu = timed(lambda: O.sparse_align_run(max_level=7, min_level=1, **args), 5)
    out["SparseImgAlign::run (1000 features, 7 levels)"] = {"gpu_ms": g * 1e3, "cpu_ms_1_thread": cpu * 1e3, "iterations": float(stats[0]),
                                                            "tracked": int(nt), "launches_per_run": 1}
    sa.close(); ref.close(); cur.close()
    # ---- FindDirectProjection: 2000 (MapPoint, keyframe) pairs
    d = synth.direct_projection_case(2, n=2000)
    pyrs = [Y.Pyramid(k["levels"]) for k in d["kfs"]]
    curp = Y.Pyramid(d["cur_levels"])
    kfs = [dict(k, pyramid=p) for k, p in zip(d["kfs"], pyrs)]
    dargs = {k: v for k, v in d.items() if k not in ("kfs", "cur_levels")}
    g_px, g_lv, g_ok = mt.find_direct_projection(kfs, curp, **dargs)
    e_px, e_lv, e_ok = O.find_direct_projection(**d)
    assert g_px.tobytes() == e_px.tobytes() and np.array_equal(g_ok, e_ok)
    g = timed(lambda: mt.find_direct_projection(kfs, curp, **dargs), 50)
    cpu = timed(lambda: O.find_direct_projection(**d), 5)
    out["FindDirectProjection (2000 pairs)"] = {"gpu_ms": g * 1e3, "cpu_ms_1_thread": cpu * 1e3, "converged": float(e_ok.mean())}
    for p in pyrs + [curp]:
        p.close()
    # ---- ComputeDistinctiveDescriptors: 20000 MapPoints, 3..40 observations each
    desc, start = synth_observations(20000, 5)
    gi, gm, gd = mt.distinctive_descriptors(desc, start)
    ei, em, ed = O.distinctive_descriptors(desc, start)
    assert np.array_equal(gi, ei) and np.array_equal(gm, em)
    g = timed(lambda: mt.distinctive_descriptors(desc, start), 20)
    cpu = timed(lambda: O.distinctive_descriptors(desc, start), 3)
    out["ComputeDistinctiveDescriptors (20000 MapPoints, %d observations)" % len(desc)] = {"gpu_ms": g * 1e3, "cpu_ms_1_thread": cpu * 1e3}
    mt.close()
    print(json.dumps({"bench": "f4: direct-tracking pyramid consumers + descriptor post-processing, host inputs through the C ABI",
                      "results": out}))


if __name__ == "__main__":
    main()
