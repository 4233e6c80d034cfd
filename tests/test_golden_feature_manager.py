"""oracle/feature_manager.py against the committed golden vectors of the reference's compiled FeatureManager
(tests/golden/feature_manager.npz, generator tests/golden/make_golden_fm.py): runs everywhere, /root/reference not needed."""
import os

import numpy as np

from oracle import fm_scenario
from oracle.feature_manager import FeatureManager

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "feature_manager.npz"))


def test_restatement_reproduces_the_golden_vectors():
    n_frames = int(G["n_frames"])
    for seed, speed in G["scenarios"]:
        seed = int(seed)
        fm = FeatureManager(10.0, 460.0, 5.0)
        tracked = []
        keys = fm_scenario.drive(fm, seed, n_frames, speed=float(speed),
                                 on_step=lambda op, m, r: tracked.append(m.last_track_num) if op[0] == "add" else None)
        p = f"s{seed}_"
        assert np.array_equal(np.array(keys, dtype=np.uint8), G[p + "keys"]), f"seed {seed}: key-frame decisions"
        assert np.array_equal(np.array(tracked, dtype=np.int32), G[p + "tracked"]), f"seed {seed}: last_track_num_"
        ids, start, n_obs, depth, flag, obs = fm.export()
        for name, a in (("ids", ids), ("start", start), ("n_obs", n_obs), ("depth", depth), ("flag", flag), ("obs", obs)):
            b = G[p + name]
            if name == "depth":      # through numpy.linalg.svd on both sides: another LAPACK build may differ in the last bits
                assert a.shape == b.shape and np.allclose(a, b, rtol=1e-9, atol=0), f"seed {seed}: depth"
            else:
                assert a.shape == b.shape and np.array_equal(a.view(np.uint8), b.view(np.uint8)), f"seed {seed}: {name}"
    assert len(G["s1_ids"]) > 50
