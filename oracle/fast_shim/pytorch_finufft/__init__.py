"""TEST INFRASTRUCTURE ONLY -- `pytorch_finufft` module name for the UNMODIFIED reference in oracle/_ref, backed by the
C/OpenMP ES-kernel NUFFT of oracle/cnufft.c (FINUFFT itself is not installable in this image).  Used by bench.py's CPU
arm only; the parity fixtures use the exact NUDFT of oracle/pytorch_finufft instead."""
from . import functional  # noqa: F401
