from .glfsr import GLFSR
from .pn import PNSequence

__all__ = ["GLFSR", "PNSequence"]
