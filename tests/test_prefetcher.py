"""DevicePrefetcher (the copy of train_eval_utils.py:30-31 of the reference, moved to a side stream): every
yielded batch holds its own host batch's bytes while the consumer's work on it runs, whatever the
order of shapes, with the two resident buffer sets reused in turn."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _batches(shapes, pin):
    g = torch.Generator().manual_seed(11)
    out = []
    for i, (h, w) in enumerate(shapes):
        x = torch.rand(2, 3, h, w, generator=g) + i
        gt = torch.rand(2, 1, h // 8, w // 8, generator=g) + i
        out.append((x.pin_memory(), gt.pin_memory()) if pin else (x, gt))
    return out


@pytest.mark.parametrize("pin", [True, False])
@pytest.mark.parametrize("shapes", [[(512, 768)] * 6, [(64, 64), (512, 768), (512, 768), (96, 128), (64, 64), (512, 768)],
                                    [(256, 256)], []])
def test_every_batch_arrives_intact_while_slots_are_reused(dev, shapes, pin):
    from can_b200.utils.train_eval_utils import DevicePrefetcher
    host = _batches(shapes, pin)
    sums, n = [], 0
    spin = torch.empty(64 << 20, device=dev)
    for (xd, gd), (xh, gh) in zip(DevicePrefetcher(host, dev), host):
        assert xd.device == dev and xd.shape == xh.shape and gd.shape == gh.shape
        # slow consumer on the current stream: the next copies must not overtake these reads
        for _ in range(4):
            spin.add_(1.0)
        sums.append((xd.double().sum() + gd.double().sum(), float(xh.double().sum() + gh.double().sum())))
        assert torch.equal(xd.cpu(), xh) and torch.equal(gd.cpu(), gh)
        n += 1
    assert n == len(host)
    for got, want in sums:
        assert float(got) == pytest.approx(want, rel=1e-12)


def test_deferred_reads_see_the_right_batch(dev):
    """A consumer that only enqueues work (no host sync inside the loop) still reads each batch before its slot
    is overwritten: the copy stream waits for the step that used the slot."""
    from can_b200.utils.train_eval_utils import DevicePrefetcher
    host = _batches([(512, 768)] * 8, True)
    spin = torch.empty(64 << 20, device=dev)
    results = []
    for xd, gd in DevicePrefetcher(host, dev):
        for _ in range(6):
            spin.add_(1.0)
        results.append(xd.double().sum() + gd.double().sum())   # device scalars, read after the loop
    torch.cuda.synchronize()
    for got, (xh, gh) in zip(results, host):
        assert float(got) == pytest.approx(float(xh.double().sum() + gh.double().sum()), rel=1e-12)


def test_steady_state_allocates_nothing(dev):
    from can_b200.utils.train_eval_utils import DevicePrefetcher
    host = _batches([(256, 384)] * 6, True)
    ptrs = set()
    for xd, gd in DevicePrefetcher(host, dev):
        ptrs.add((xd.data_ptr(), gd.data_ptr()))
    assert len(ptrs) == 2
