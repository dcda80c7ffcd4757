"""Exceptions of the calculator boundary.

``ReadUnsortedError`` must be the reference's own class when PyMaSC is importable, because
PyMaSC/pymasc.py:249-250 catches it by identity; stand-alone it is an ``IndexError``
subclass of the same name (PyMaSC/core/exceptions.py:4).
"""
try:
    from PyMaSC.core.exceptions import ReadUnsortedError  # type: ignore  # noqa: F401
except Exception:
    class ReadUnsortedError(IndexError):
        """Input reads are not sorted by chromosome and position."""


class DeviceError(RuntimeError):
    """The CUDA shim reported an error (code and message from pmb_last_error)."""


class PositionOutOfRange(IndexError):
    """A read end or mappability interval lies beyond the chromosome's bit array
    (chrom_len + read_len + max_shift + 100 bits).  The reference aborts the process on
    a BitArray assert here (external/BitArray/bit_array.c:670, :825); we raise instead."""
