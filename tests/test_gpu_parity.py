"""-m gpu: the CUDA step vs the oracle through the C-ABI, bit-exact unless a tolerance is stated."""
import pytest
import parity
from box2d_b200 import scenes

pytestmark = pytest.mark.gpu


def test_small_pyramid_bit_exact():
    res = parity.run_pair(lambda w: scenes.add_pyramid(w, 10), steps=20, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_mixed_drop_bit_exact():
    res = parity.run_pair(lambda w: scenes.mixed_drop(w), steps=60, verbose=True)
    assert res["max_rel_err"] == 0.0
    assert res["toi"] == res["toi_oracle"]      # fast bodies take the no-hit continuous path on both sides


def test_large_pyramid_10_steps():
    res = parity.run_pair(lambda w: scenes.large_pyramid(w, 40), steps=10, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_joint_grid_small():
    res = parity.run_pair(lambda w: scenes.joint_grid(w, 12), steps=30, verbose=True)
    assert res["max_rel_err"] == 0.0


def test_tumbler_small():
    res = parity.run_pair(lambda w: scenes.tumbler(w, n=600, half=6.0), steps=40, verbose=True)
    assert res["max_rel_err"] == 0.0


@pytest.mark.parametrize("name", ["pyramid10", "mixed_drop", "joint_grid8", "tumbler300"])
def test_cuda_matches_committed_golden(name):
    """The CUDA path against tests/golden/*.npz (bit-exact pair sets, ids, colors, feature ids, events, body states)."""
    from box2d_b200 import api
    from test_oracle_golden import check_against_golden
    w = api.World(api.CudaBackend(), scenes.world_def())
    try:
        check_against_golden(w, name)
    finally:
        w.Destroy()
