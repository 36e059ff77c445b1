"""can_b200 -- B200-native (sm_100a) CANNet training / inference step.

Drop-in for the hot path of zgzhengSEU/CAN-distributed-pytorch: ``can_b200.model.CANNet``
mirrors ``model/CANNet.py``; ``can_b200.utils`` mirrors ``utils/train_eval_utils.py`` and
``utils/distributed_utils.py``.  All arithmetic runs in ``libcan_b200.so``.
"""
from ._lib import CanB200Error, LIB_PATH, exported_symbols, lib  # noqa: F401

__all__ = ["CanB200Error", "LIB_PATH", "exported_symbols", "lib"]
