"""train_step.InputPrefetcher: batches come out in order and intact while copies and consumers overlap."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_input_prefetcher_order_and_reuse():
    from maavss_b200 import train_step
    host = [(torch.full((1 << 20,), float(i)).pin_memory(), torch.arange(257, dtype=torch.float32).add(i).pin_memory()) for i in range(7)]
    feed = train_step.InputPrefetcher(host[0], "cuda")
    feed.put(host[0])
    sums = []
    for i in range(7):
        if i + 1 < 7:
            feed.put(host[i + 1])
        a, b = feed.get()
        # a long-running consumer so that the next copy really has to wait for release()
        acc = a
        for _ in range(20):
            acc = acc * 1.0 + 0.0
        sums.append((acc.sum(), b.sum()))
        feed.release()
    for i, (sa, sb) in enumerate(sums):
        assert float(sa) == float(i) * (1 << 20)
        assert float(sb) == float(sum(range(257)) + 257 * i)
    with pytest.raises(AssertionError):
        feed.put(host[0]); feed.put(host[1]); feed.put(host[2])
