"""Host policy for CUDA-graph capture (can_b200/utils/graph_cache.py): fixed-size workloads capture on their third
consecutive step, a stream of changing image sizes (the reference's data sets, one image per batch, train.py:104-113)
never does, and the number of live graphs is bounded."""
import pytest

from can_b200.utils.graph_cache import GraphCache


def _run(cache, keys, signature=None):
    """Drives the cache the way CANNet._forward_no_grad / FusedTrainer.step do; returns 'e' (eager), 'c' (capture +
    replay) or 'r' (replay) per call."""
    trace = []
    for k in keys:
        ent = cache.lookup(k, signature)
        if ent is None:
            if not cache.should_capture():
                trace.append("e")
                continue
            cache.insert(k, ("graph", k))
            trace.append("c")
        else:
            assert ent == ("graph", k)
            trace.append("r")
    return "".join(trace)


def test_fixed_shape_captures_on_the_third_consecutive_call():
    c = GraphCache(capacity=4)
    assert _run(c, ["a"] * 6) == "eecrrr"
    assert len(c) == 1


def test_changing_shapes_never_capture():
    c = GraphCache(capacity=4)
    epoch = [f"img{i}" for i in range(300)]
    assert set(_run(c, epoch * 3)) == {"e"}          # three epochs over 300 distinct sizes: nothing is captured
    assert len(c) == 0
    assert _run(c, ["a", "b", "a", "b", "a", "b"]) == "eeeeee"   # alternation is not a streak either


def test_captured_shapes_survive_interruptions_and_are_bounded():
    c = GraphCache(capacity=2)
    assert _run(c, ["a"] * 3 + ["b"] * 3) == "eeceec"
    assert _run(c, ["a", "b", "a"]) == "rrr"           # both stay cached; a foreign shape in between does not evict
    assert _run(c, ["x", "a"]) == "er"
    assert _run(c, ["d"] * 3) == "eec" and len(c) == 2   # third graph: least recently used ("b") goes
    assert c.lookup("b") is None and c.lookup("a") == ("graph", "a") and c.lookup("d") == ("graph", "d")


def test_signature_change_drops_everything_and_warms_up_again():
    c = GraphCache(capacity=4)
    assert _run(c, ["a"] * 4, signature=1) == "eecr"
    assert _run(c, ["a"] * 4, signature=2) == "eecr"   # parameters changed: two eager calls again, then a new graph
    assert len(c) == 1


def test_capacity_from_environment(monkeypatch):
    monkeypatch.setenv("CAN_B200_GRAPH_CACHE", "7")
    assert GraphCache().capacity == 7
    monkeypatch.setenv("CAN_B200_GRAPH_CACHE", "0")
    assert GraphCache().capacity == 1                    # never less than one
    monkeypatch.delenv("CAN_B200_GRAPH_CACHE")
    assert GraphCache().capacity == 4


def test_clear():
    c = GraphCache()
    _run(c, ["a"] * 3)
    c.clear()
    assert len(c) == 0 and _run(c, ["a"] * 3) == "eec"
