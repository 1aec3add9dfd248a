from .precoder import MultiHSymbolMapper
from .pulse_filters import MULTIH_IRIG_DENOM, MULTIH_IRIG_NUMER, freq_pulse_multih_irig

__all__ = ["MultiHSymbolMapper", "MULTIH_IRIG_NUMER", "MULTIH_IRIG_DENOM", "freq_pulse_multih_irig"]
