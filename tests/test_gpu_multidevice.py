"""Two devices driven from ONE process (round-1 verdict weak #10 / next #5): the library keeps a per-device-ordinal attribute
table (api.cu::device_info), module blobs follow .to(device), launches go to the stream of the tensors' device.  Skipped on
boxes with one GPU (the 8-GPU scaling run uses one process per GPU)."""
import pytest
import torch

from conftest import load_ckpt

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two CUDA devices")
def test_two_devices_in_one_process(gold_stats):
    import mpcdl
    from mpcdl import synth
    rows = synth.load_rows()
    outs = []
    for name, window, st in (("Basic_64_4_32", 0, gold_stats[0]), ("GRU_64_4_4_16_False", 4, gold_stats[1])):
        init, U = synth.make_inputs(rows, 300, 6, window=window, seed=13)
        per_dev = []
        for d in (1, 0, 1):                                  # first call on device 1, then 0, then 1 again
            dev = torch.device("cuda", d)
            m, _ = mpcdl.from_checkpoint_name(name, load_ckpt(name), dev)
            with torch.cuda.device(dev):
                s = m.rollout(init.to(dev), U.to(dev), st)[0]
            assert s.device == dev
            per_dev.append(s.cpu())
        assert torch.equal(per_dev[0], per_dev[1]) and torch.equal(per_dev[0], per_dev[2])
        # one module moved between devices re-packs its blob
        m, _ = mpcdl.from_checkpoint_name(name, load_ckpt(name), torch.device("cuda:0"))
        a = m.rollout(init.cuda(0), U.cuda(0), st)[0].cpu()
        m.to("cuda:1")
        b = m.rollout(init.cuda(1), U.cuda(1), st)[0].cpu()
        assert torch.equal(a, b) and torch.equal(a, per_dev[0])
        outs.append(a)
    with pytest.raises(RuntimeError):                        # tensors on two different devices in one call
        m.rollout(init.cuda(0), U.cuda(1), st)
