"""HostBatchStream (double-buffered H2D staging in front of the train step): order, content, buffer reuse."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_host_batch_stream_delivers_batches_in_order_and_reuses_buffers():
    import gpflow_b200 as G
    host = [(torch.full((1000, 2), float(i)).pin_memory(), torch.full((1000, 2), -float(i)).pin_memory()) for i in range(7)]
    stream = G.HostBatchStream(iter(host), "cuda")
    ptrs, sums = set(), []
    for k, (a, b) in enumerate(stream):
        ptrs.add(a.data_ptr())
        # consumer work on the main stream that must not be overtaken by the next copy into the same slot
        s = (a + b).abs().sum() + a.sum()
        torch.cuda._sleep(2_000_000)
        sums.append((s, k))
    assert len(sums) == 7 and len(ptrs) == 2
    for s, k in sums:
        assert float(s) == 2000.0 * k
    assert stream.bytes_copied == 7 * 2 * 1000 * 2 * 4
    assert list(G.HostBatchStream(iter([]), "cuda")) == []
