from .CANNet import CANNet, make_layers  # noqa: F401

__all__ = ["CANNet", "make_layers"]
