"""The RAFT harness (oracle/raft_harness.py + frozen reference sub-networks) against the
reference's own RAFT.forward, frozen in tests/golden/raft_e2e_small.npz by
oracle/make_ref_modules.py.  CPU; skipped where the frozen modules were not built."""
import numpy as np
import pytest
import torch

from oracle import raft_harness as Hn

from conftest import load_golden


@pytest.mark.skipif(not Hn.available(), reason="oracle/_ref/*.pt not built (needs the reference checkout)")
def test_harness_reproduces_reference_forward():
    g = load_golden("raft_e2e_small")
    nets = Hn.load(torch.device("cpu"))
    low, up = Hn.raft_forward(nets, torch.from_numpy(g["image1"]), torch.from_numpy(g["image2"]),
                              iters=int(g["iters"]), corr_cls=Hn.TorchCorrBlock)
    # same modules, same weights, same torch calls: agreement to fp32 round-off
    assert np.abs(low.numpy() - g["flow_low"]).max() < 1e-3
    assert np.abs(up.numpy() - g["flow_up"]).max() < 1e-2
    epe = np.sqrt(((up.numpy() - g["flow_up"]) ** 2).sum(1)).mean()
    assert epe < 1e-3, epe


@pytest.mark.skipif(not Hn.available(small=True), reason="oracle/_ref/raft_small_*.pt not built (needs the reference checkout)")
def test_small_harness_reproduces_reference_forward():
    """RAFT-small (D = 128, radius 3, no mask -> upflow8): the harness around the traced small sub-networks against
    the reference's own RAFT(args.small=True).forward frozen in tests/golden/raft_small_e2e.npz."""
    g = load_golden("raft_small_e2e")
    nets = Hn.load(torch.device("cpu"), small=True)
    low, up = Hn.raft_small_forward(nets, torch.from_numpy(g["image1"]), torch.from_numpy(g["image2"]),
                                    iters=int(g["iters"]))
    assert np.abs(low.numpy() - g["flow_low"]).max() < 1e-3
    epe = np.sqrt(((up.numpy() - g["flow_up"]) ** 2).sum(1)).mean()
    assert epe < 1e-3, epe
