"""affb200.prefetch.DevicePrefetcher: batches arrive intact and in order while the compute stream is busy."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_prefetched_batches_arrive_in_order():
    from affb200.prefetch import DevicePrefetcher
    dev = torch.device("cuda", 0)
    host = [[torch.full((1 << 20,), float(i)).pin_memory(), torch.arange(8).add(i).pin_memory()] for i in range(5)]
    pf = DevicePrefetcher(dev)
    busy = torch.randn(2048, 2048, device=dev)
    sums = []
    pf.submit(host[0])
    for i in range(5):
        if i + 1 < 5:
            pf.submit(host[i + 1])
        x, y = pf.next()
        for _ in range(4):
            busy = busy @ busy * 1e-3            # keep the compute stream occupied while the next copy runs
        sums.append((x.sum() / x.numel(), y.clone()))
        pf.release()
    torch.cuda.synchronize()
    for i, (s, y) in enumerate(sums):
        assert float(s) == float(i)
        assert torch.equal(y.cpu(), torch.arange(8) + i)


def test_submit_refuses_a_third_batch_in_flight():
    from affb200.prefetch import DevicePrefetcher
    pf = DevicePrefetcher(torch.device("cuda", 0))
    h = [torch.zeros(4).pin_memory()]
    pf.submit(h)
    pf.submit(h)
    with pytest.raises(RuntimeError, match="in flight"):
        pf.submit(h)
