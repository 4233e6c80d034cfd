"""The Python restatement of frontend::FeatureManager (oracle/feature_manager.py) against the REFERENCE's own compiled
feature_manager.cpp (oracle/_ref/libvsref_fm.so): bit-equal state after every operation of a synthetic estimator loop
(reference: /root/reference/src/frontend/feature_manager.cpp:27-73,142-294)."""
import numpy as np
import pytest

from oracle import fm_scenario, ref_fm_binding
from oracle.feature_manager import WINDOW_SIZE, FeatureManager

pytestmark = pytest.mark.skipif(not ref_fm_binding.available(), reason="oracle/_ref/libvsref_fm.so not built (needs /root/reference)")


def _same_state(a, b, where):
    ea, eb = a.export(), b.export()
    names = ["ids", "start_frame", "n_obs", "depth", "solve_flag", "obs"]
    for n, x, y in zip(names, ea, eb):
        assert x.shape == y.shape, f"{where}: {n} shape {x.shape} vs {y.shape}"
        assert np.array_equal(x.view(np.uint8), y.view(np.uint8)), f"{where}: {n} differs"


def _lockstep(seed, n_frames, **kw):
    py = FeatureManager(10.0, 460.0, 5.0)
    ref = ref_fm_binding.CompiledRefFeatureManager(10.0, 460.0, 5.0)
    gen = fm_scenario.steps(seed, n_frames, **kw)
    op = next(gen, None)
    n_ops = {}
    keys = []
    while op is not None:
        n_ops[op[0]] = n_ops.get(op[0], 0) + 1
        send = None
        if op[0] == "add":
            k1 = py.add_feature_and_check_parallax(op[1], op[2])
            k2 = ref.add_feature_and_check_parallax(op[1], op[2])
            assert k1 == k2 and py.last_track_num == ref.last_track_num, f"add at frame_count {op[1]}"
            keys.append(k1)
            send = k1
        elif op[0] == "triangulate":
            py.triangulate_across_all_views(op[1], op[2], op[3])
            ref.triangulate_across_all_views(op[1], op[2], op[3])
        elif op[0] == "depth_vector":
            d1, d2 = py.get_depth_vector(), ref.get_depth_vector()
            assert d1 == d2
            assert py.get_feature_count() == ref.get_feature_count() == len(d1)
            send = d1
        elif op[0] == "set_depth":
            py.set_depth(op[1]); ref.set_depth(op[1])
        elif op[0] == "remove_failures":
            py.remove_failures(); ref.remove_failures()
        elif op[0] == "remove_back":
            py.remove_back(); ref.remove_back()
        elif op[0] == "remove_back_shift_depth":
            py.remove_back_shift_depth(*op[1:]); ref.remove_back_shift_depth(*op[1:])
        elif op[0] == "remove_front":
            py.remove_front(op[1]); ref.remove_front(op[1])
        _same_state(py, ref, f"seed {seed} after {op[0]} #{n_ops[op[0]]}")
        c1, c2 = py.get_corresponding(WINDOW_SIZE - 3, WINDOW_SIZE), ref.get_corresponding(WINDOW_SIZE - 3, WINDOW_SIZE)
        assert c1 == c2
        try:
            op = gen.send(send)
        except StopIteration:
            op = None
    return keys, n_ops


@pytest.mark.parametrize("seed,speed", [(1, 0.06), (2, 0.004), (3, 0.02), (4, 0.3)])
def test_restatement_equals_compiled_reference(seed, speed):
    keys, n_ops = _lockstep(seed, 60, speed=speed)
    # the loop went through every operation of the manager
    assert n_ops["add"] == 60 and n_ops.get("triangulate", 0) > 30
    assert n_ops.get("remove_back_shift_depth", 0) + n_ops.get("remove_back", 0) + n_ops.get("remove_front", 0) == 60 - WINDOW_SIZE
    assert n_ops.get("remove_failures", 0) > 5


def test_scenarios_cover_both_decisions():
    slow, _ = _lockstep(2, 60, speed=0.004)
    fast, _ = _lockstep(4, 60, speed=0.3)
    assert not all(slow[WINDOW_SIZE:]), "low parallax must give frames that are not key frames (removeFront path)"
    assert all(fast), "high parallax: every frame a key frame"


def test_few_tracked_features_is_a_key_frame():
    # fm:52-55: fewer than 20 tracked features => key frame whatever the parallax
    keys, _ = _lockstep(5, 30, speed=0.0005, max_feat=15)
    assert all(keys)


def test_clear_state():
    py = FeatureManager(); ref = ref_fm_binding.CompiledRefFeatureManager()
    img = {3: [0.1, 0.2, 1, 10, 20, 0, 0], 9: [0.3, -0.2, 1, 30, 40, 0, 0]}
    for m in (py, ref):
        m.add_feature_and_check_parallax(0, img)
        m.clear_state()
        m.add_feature_and_check_parallax(0, img)
    _same_state(py, ref, "after clear")
    assert len(py.bank) == 2


def _reuse_ops():
    """A feature erased by removeFailures while the tracker still publishes its id comes back as a NEW list entry at the
    back (the list is then no longer ordered by id); an empty frame and a single-feature frame ride along."""
    def img(ids, k):
        return {i: [0.01 * i + 0.002 * k, -0.005 * i, 1.0, 100.0 + i, 50.0 + k, 0.1, -0.1] for i in ids}
    ops = [("add", 0, img(range(1, 31), 0)), ("add", 1, img(range(1, 31), 1)), ("add", 2, img(range(1, 31), 2))]
    ops.append(("set_depth_by_id", {5: -2.0, 17: -0.5}))       # inverse depths: negative => solve_flag 2
    ops.append(("remove_failures",))
    ops.append(("add", 3, img(list(range(1, 31)) + [40, 33], 3)))    # 5 and 17 return, 33 and 40 are new: order 5, 17, 33, 40
    ops.append(("add", 4, {}))
    ops.append(("add", 5, img([33], 5)))
    ops.append(("remove_back",))      # (removeFront is only defined for frame_count == WINDOW_SIZE: estimator.cpp:311)
    ops.append(("remove_back",))
    return ops


def _apply_reuse(fm, op):
    if op[0] == "add":
        return fm.add_feature_and_check_parallax(op[1], op[2])
    if op[0] == "set_depth_by_id":
        ids, start, n_obs, depth, flag, obs = fm.export()
        x = [op[1].get(int(i), 0.25) for i, s, n in zip(ids, start, n_obs) if n >= 2 and s < WINDOW_SIZE - 2]
        return fm.set_depth(x)
    return getattr(fm, op[0])(*op[1:])


def test_erased_feature_returns_at_the_back_of_the_list():
    py = FeatureManager(); ref = ref_fm_binding.CompiledRefFeatureManager()
    for k, op in enumerate(_reuse_ops()):
        r1, r2 = _apply_reuse(py, op), _apply_reuse(ref, op)
        assert r1 == r2, (k, op[0])
        _same_state(py, ref, f"reuse op {k} {op[0]}")
    ids = [it.feature_id for it in py.bank]
    assert ids != sorted(ids) and ids.index(5) > ids.index(30) and ids.index(17) > ids.index(30)
    assert ids.index(5) < ids.index(17) < ids.index(33) < ids.index(40)
