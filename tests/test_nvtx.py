"""NVTX ranges (SURVEY.md section 5 asks the new repo for them; the reference has no tracing): off by default and then
free, balanced push / pop when on (checked without a GPU by swapping the two torch.cuda.nvtx calls)."""
import importlib
import os
import sys

import torch


def _load(on):
    os.environ["CAN_B200_NVTX"] = "1" if on else "0"
    try:
        sys.modules.pop("can_b200.utils.nvtx", None)
        return importlib.import_module("can_b200.utils.nvtx")
    finally:
        os.environ.pop("CAN_B200_NVTX", None)


def test_off_by_default_and_free():
    m = _load(False)
    assert not m.ENABLED
    assert m.push("x") is None and m.pop() is None
    with m.span("can.step"):
        pass


def test_ranges_balance_when_on(monkeypatch):
    seen = []
    monkeypatch.setattr(torch.cuda.nvtx, "range_push", lambda name: seen.append(("push", name)))
    monkeypatch.setattr(torch.cuda.nvtx, "range_pop", lambda: seen.append(("pop", None)))
    m = _load(True)
    assert m.ENABLED
    try:
        with m.span("can.step"):
            m.push("can.fwd.frontend")
            m.pop()
            raise RuntimeError("inside")
    except RuntimeError:
        pass
    assert seen == [("push", "can.step"), ("push", "can.fwd.frontend"), ("pop", None), ("pop", None)]
    _load(False)


def test_engine_sections_push_and_pop_in_pairs():
    """Every nvtx.push in the engine / trainer has its pop (a generator that yields between them closes the range first)."""
    import inspect
    import can_b200.engine as eng
    import can_b200.utils.train_eval_utils as tu
    for mod in (eng, tu):
        src = inspect.getsource(mod)
        assert src.count("nvtx.push(") == src.count("nvtx.pop()") and src.count("nvtx.push(") > 0


def test_cold_read_pass_ignores_small_and_cpu_buffers(monkeypatch):
    """ops._read_if_fresh (the one-off read pass over freshly allocated activation buffers >= 512 MB)
    must be free for everything else: CPU tensors and small buffers return before any CUDA call."""
    from can_b200 import ops
    called = []
    monkeypatch.setattr(torch.cuda, "memory_reserved", lambda *a, **k: called.append(1) or 0)
    ops._read_if_fresh(torch.empty(16))
    ops._read_if_fresh(torch.empty(1 << 20, dtype=torch.bfloat16))
    assert not called
    assert ops._COLD_READ_MIN == 512 << 20
