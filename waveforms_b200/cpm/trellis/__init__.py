from .encoder import TrellisEncoder
from .model import (
    Branch,
    FiniteStateMachine,
    SimpleTrellis2,
    SimpleTrellis4,
    SOQPSKTrellis4x2,
    SOQPSKTrellis4x2DiffEncoded,
    SOQPSKTrellis8x1,
    Trellis,
)

__all__ = ["Branch", "Trellis", "FiniteStateMachine", "TrellisEncoder", "SOQPSKTrellis8x1", "SOQPSKTrellis4x2",
           "SOQPSKTrellis4x2DiffEncoded", "SimpleTrellis2", "SimpleTrellis4"]
