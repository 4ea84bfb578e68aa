"""pointdsc_b200: B200 (sm_100a) implementation of PointDSC's spatial-consistency hot path
behind the reference's ``PointDSC.forward(data)`` contract."""
from .model import PointDSC, NonLocalNet, NonLocalBlock  # noqa: F401
from . import _cabi as cabi  # noqa: F401

__all__ = ["PointDSC", "NonLocalNet", "NonLocalBlock", "cabi"]
