"""Host logic of DevicePrefetcher without a GPU: the CUDA stream / event objects are replaced by recorders, the
"device" is the CPU.  Checked: every batch arrives, the two buffer sets alternate, a slot is rewritten only after the
event recorded behind its consumer, and leaving the loop early still records that event."""
import contextlib

import pytest
import torch


class _Event:
    count = 0

    def __init__(self):
        _Event.count += 1
        self.name = f"ev{_Event.count}"
        self.records = 0

    def record(self, stream=None):
        self.records += 1
        LOG.append(("record", self.name, getattr(stream, "name", None)))


class _Stream:
    def __init__(self, name):
        self.name = name

    def wait_event(self, ev):
        LOG.append(("wait", self.name, ev.name))


LOG = []


@pytest.fixture()
def prefetcher(monkeypatch):
    from can_b200.utils import train_eval_utils as t
    LOG.clear()
    cur = _Stream("compute")
    monkeypatch.setattr(t.torch.cuda, "Event", _Event)
    monkeypatch.setattr(t.torch.cuda, "Stream", lambda device=None: _Stream("copy"))
    monkeypatch.setattr(t.torch.cuda, "stream", lambda s: contextlib.nullcontext())
    monkeypatch.setattr(t.torch.cuda, "current_stream", lambda d=None: cur)
    monkeypatch.setattr(t.torch.cuda, "current_device", lambda: 0)
    monkeypatch.setattr(torch.Tensor, "record_stream", lambda self, s: None, raising=False)
    monkeypatch.setattr(t.DevicePrefetcher, "_streams", {})
    return t.DevicePrefetcher


def _host(n, shape=(2, 3)):
    return [(torch.full(shape, float(i)), torch.full((1,), float(i))) for i in range(n)]


def test_batches_arrive_in_order_from_two_alternating_buffer_sets(prefetcher):
    pf = prefetcher(_host(5), torch.device("cpu"))
    seen, ptrs = [], []
    for x, g in pf:
        seen.append((float(x[0, 0]), float(g[0])))
        ptrs.append(x.data_ptr())
    assert seen == [(float(i), float(i)) for i in range(5)]
    assert ptrs[0] == ptrs[2] == ptrs[4] and ptrs[1] == ptrs[3] and ptrs[0] != ptrs[1]
    assert len(list(prefetcher([], torch.device("cpu")))) == 0


def test_a_slot_is_rewritten_only_behind_its_consumer(prefetcher):
    pf = prefetcher(_host(4), torch.device("cpu"))
    for _ in pf:
        LOG.append(("step",))
    freed = [e.name for e in pf._freed]
    copied = [e.name for e in pf._copied]
    # batch 2 goes to slot 0: the copy stream must wait for the event recorded after step 0, and that wait must come
    # after step 0 was enqueued and before the copy is marked done a second time
    i_step0 = LOG.index(("step",))
    i_rec_freed0 = LOG.index(("record", freed[0], "compute"))
    i_wait_freed0 = LOG.index(("wait", "copy", freed[0]))
    second_copy_slot0 = [i for i, e in enumerate(LOG) if e == ("record", copied[0], "copy")][1]
    assert i_step0 < i_rec_freed0 < i_wait_freed0 < second_copy_slot0
    # the compute stream waits for each batch's copy before the consumer sees it
    assert [e for e in LOG if e[0] == "wait" and e[1] == "compute"] == [("wait", "compute", copied[i & 1]) for i in range(4)]


def test_leaving_the_loop_early_still_releases_the_slot(prefetcher):
    pf = prefetcher(_host(4), torch.device("cpu"))
    for _ in pf:
        break
    assert pf._freed[0] is not None and pf._freed[0].records == 1
    assert [float(x[0, 0]) for x, _ in pf] == [0.0, 1.0, 2.0, 3.0]   # a second pass over the same object works


def test_buffers_follow_shape_changes(prefetcher):
    host = _host(2, (2, 3)) + _host(2, (4, 5)) + _host(1, (2, 3))
    out = [tuple(x.shape) for x, _ in prefetcher(host, torch.device("cpu"))]
    assert out == [(2, 3), (2, 3), (4, 5), (4, 5), (2, 3)]
