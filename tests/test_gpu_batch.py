"""PathBatch uploads: the action block of a large pinned fp32 batch travels on a copy stream behind the observations
(mjrl_b200/batch.py); consumers must see exactly the host data, on whichever stream they run."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_action_block_uploaded_behind_the_observations_is_complete_on_first_use():
    import torch
    from mjrl_b200 import batch as B
    rng = np.random.default_rng(3)
    N, n, m = 400_000, 11, 3
    obs = torch.from_numpy(rng.standard_normal((N, n)).astype(np.float32)).pin_memory()
    act = torch.from_numpy(rng.standard_normal((N, m)).astype(np.float32)).pin_memory()
    rew = torch.from_numpy(rng.standard_normal(N)).pin_memory()
    off = np.arange(0, N + 1, 1000, dtype=np.int64)
    term = np.zeros(len(off) - 1, dtype=np.uint8)
    assert B.OVERLAP_ACT
    for trial in range(3):
        b = B.PathBatch(obs, act, rew, off, term)
        assert b._act_event is not None                      # in flight on the copy stream
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):                        # a consumer on another stream than the one that built the batch
            s = b.act.double().sum()
        assert b._act_event is None
        torch.cuda.synchronize()
        assert abs(float(s) - float(act.double().sum())) < 1e-6
        assert torch.equal(b.act.cpu(), act)
        assert torch.equal(b.obs.cpu(), obs)
    keep, B.OVERLAP_ACT = B.OVERLAP_ACT, False
    try:
        b2 = B.PathBatch(obs, act, rew, off, term)
        assert b2._act_event is None and torch.equal(b2.act.cpu(), act)
    finally:
        B.OVERLAP_ACT = keep
