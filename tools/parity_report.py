"""Parity report on the benchmarked configurations (tests/parity_util.py): one JSON line per (config, frame).

  python tools/parity_report.py [C1 C2 C3 C5] > profiles/r02_parity.jsonl

Runs the fused CUDA wavefront with the diagnostic path record on, the compiled reference kernels (oracle/_ref) on
the host cores, and classifies every pixel outside 1e-4 by the first bounce at which the two records differ.
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import mcrt_b200  # noqa: E402,F401
import parity_util  # noqa: E402


def bdpt_c4():
    import numpy as np
    from mcrt_b200 import scene_library
    from test_bdpt_gpu import gpu_frames, ref_frames, rel_err
    scene = scene_library.cornell_box(True)
    w = h = 1024
    t = time.time()
    got, cnt = gpu_frames(scene, w, h, 5, 2)
    want, _ = ref_frames(scene, w, h, 5, 2)
    for f in range(2):
        e = rel_err(got[f][:, :3], want[f][:, :3])
        print(json.dumps({"config": "C4", "frame": f, "pixels": w * h, "frac_within": float((e <= 1e-4).mean()),
                          "mean_got": got[f][:, :3].mean(0).tolist(), "mean_want": want[f][:, :3].mean(0).tolist()}), flush=True)
    a = 0.5 * (got[0][:, :3] + got[1][:, :3])
    b = 0.5 * (want[0][:, :3] + want[1][:, :3])
    print(json.dumps({"config": "C4", "rmse_2spp": float(np.sqrt(np.mean((a - b) ** 2))),
                      "half_rms_frame0_vs_frame1_oracle": float(np.sqrt(np.mean((want[0][:, :3] - want[1][:, :3]) ** 2))) / 2.0,
                      "seconds_config": round(time.time() - t, 1)}), flush=True)


def main():
    names = sys.argv[1:] or ["C1", "C2", "C3", "C5", "C4"]
    for n in names:
        t = time.time()
        if n == "C4":
            bdpt_c4()
            continue
        for rep in parity_util.run_config(n):
            rep["seconds_config"] = round(time.time() - t, 1)
            print(json.dumps(rep), flush=True)


if __name__ == "__main__":
    main()
