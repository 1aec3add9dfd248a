from .algorithm import SOQPSKTrellisDetector

__all__ = ["SOQPSKTrellisDetector"]
