"""rrl_checkpoint_save / rrl_checkpoint_load: a run interrupted between epochs and resumed
in a NEW learner from the checkpoint file continues bit-identically to the uninterrupted run; rrl_set_opt_state (which
the loader uses) is covered on the way.  Damaged or mismatched files are refused and leave the learner untouched."""
import numpy as np
import pytest

from conftest import load_golden
from test_gpu_learner import _ingest_epoch, _mk

pytestmark = pytest.mark.gpu


def test_resume_matches_the_uninterrupted_run_bit_for_bit(tmp_path):
    from relayrl_b200 import RelayRLB200Error
    g = load_golden("epoch_vf_h128")
    path = str(tmp_path / "learner.ckpt")
    A_, *_ = _mk(g)                       # uninterrupted: epochs 0, 1, then epoch 0's data again
    for ep in (0, 1, 0):
        _ingest_epoch(A_, g, ep)
        A_.train_epoch()
    B_, *_ = _mk(g)                       # interrupted after epoch 0
    _ingest_epoch(B_, g, 0)
    B_.train_epoch()
    B_.checkpoint_save(path, tag=1)
    mid = (B_.get_params(0), B_.get_params(1))
    B_.close()
    C_, *_ = _mk(g)                       # a fresh process would do exactly this
    C_.set_params(0, np.zeros_like(mid[0]))
    assert C_.checkpoint_load(path) == 1
    assert np.array_equal(C_.get_params(0), mid[0]) and np.array_equal(C_.get_params(1), mid[1])
    m, v, step = C_.get_opt_state(1)
    assert step == int(g["iters"]) and np.abs(m).max() > 0 and v.min() >= 0
    for ep in (1, 0):
        _ingest_epoch(C_, g, ep)
        C_.train_epoch()
    for which in (0, 1):
        assert np.array_equal(C_.get_params(which), A_.get_params(which))
        for x, y in zip(C_.get_opt_state(which), A_.get_opt_state(which)):
            assert np.array_equal(x, y)
    # refused: a truncated file, a file for another shape -- and the learner keeps its state
    before = C_.get_params(1)
    raw = open(path, "rb").read()
    bad = str(tmp_path / "bad.ckpt")
    open(bad, "wb").write(raw[:-100])
    with pytest.raises(RelayRLB200Error) as e:
        C_.checkpoint_load(bad)
    assert e.value.status == -6
    open(bad, "wb").write(b"NOTACKPT" + raw[8:])
    with pytest.raises(RelayRLB200Error):
        C_.checkpoint_load(bad)
    g64 = load_golden("epoch_vf_h64")
    D_, *_ = _mk(g64)
    with pytest.raises(RelayRLB200Error) as e:
        D_.checkpoint_load(path)
    assert e.value.status == -1
    assert np.array_equal(C_.get_params(1), before)
    for L in (A_, C_, D_):
        L.close()
