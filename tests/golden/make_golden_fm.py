#!/usr/bin/env python
"""Generates tests/golden/feature_manager.npz — golden vectors of the reference's FeatureManager
(/root/reference/src/frontend/feature_manager.cpp), produced by the reference's OWN compiled code
(oracle/_ref/libvsref_fm.so: that file whole and unmodified, see oracle/Makefile) driven through the synthetic estimator
loop of oracle/fm_scenario.py.  The reference holds no vectors for this path; these travel with the repo so that the
Python restatement (and through it the device feature bank) stays pinned on machines where /root/reference and the compiled
library are absent.

Per scenario (seed, speed): the key-frame decision and last_track_num_ of every frame, and the final list (ids, start
frames, observation counts, depths, solve flags, observations).  Eigen::JacobiSVD is LAPACK behind the stand-in, so the
depths carry that caveat.

Run from the repo root where /root/reference exists:  python tests/golden/make_golden_fm.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import fm_scenario, ref_fm_binding  # noqa: E402

SCENARIOS = [(1, 0.06), (2, 0.004), (4, 0.15)]
N_FRAMES = 45


def main():
    out = {"scenarios": np.array(SCENARIOS), "n_frames": np.array(N_FRAMES)}
    for seed, speed in SCENARIOS:
        fm = ref_fm_binding.CompiledRefFeatureManager(10.0, 460.0, 5.0)
        tracked = []
        keys = fm_scenario.drive(fm, seed, N_FRAMES, speed=speed,
                                 on_step=lambda op, m, r: tracked.append(m.last_track_num) if op[0] == "add" else None)
        ids, start, n_obs, depth, flag, obs = fm.export()
        p = f"s{seed}_"
        out.update({p + "keys": np.array(keys, dtype=np.uint8), p + "tracked": np.array(tracked, dtype=np.int32), p + "ids": ids,
                    p + "start": start, p + "n_obs": n_obs, p + "depth": depth, p + "flag": flag, p + "obs": obs})
        print(f"seed {seed} speed {speed}: {sum(keys)} key frames of {N_FRAMES}, {len(ids)} features in the final list")
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "feature_manager.npz"), **out)


if __name__ == "__main__":
    main()
