"""pymasc_b200 -- B200-native (sm_100a) replacement for PyMaSC's bit-array cross-correlation path.

Public surface (mirrors PyMaSC/core/bitarray/mscc.pyx and PyMaSC/result.py):
    CCBitArrayCalculator, ReadUnsortedError, NCCResult, MSCCResult, BothChromResult,
    NCCGenomeWideResult, MSCCGenomeWideResult, BothGenomeWideResult
"""
from .exceptions import DeviceError, PositionOutOfRange, ReadUnsortedError  # noqa: F401
from .result import (  # noqa: F401
    BothChromResult, BothGenomeWideResult, EmptyMSCCResult, EmptyNCCResult,
    MSCCGenomeWideResult, MSCCResult, NCCGenomeWideResult, NCCResult,
)
from .calculator import CCBitArrayCalculator, DeviceEngine  # noqa: F401
from .multifile import MultiFileCalculation  # noqa: F401

__version__ = "0.1.0"
