"""kfac_b200: B200-native K-FAC / KAISA gradient preconditioning (drop-in for kfac_pytorch's KFACPreconditioner).

The Python layer is host logic only; all arithmetic runs in hand-written sm_100a kernels behind the C ABI of
include/kfac_b200.h (libkfac_b200.so).  There is no CPU or eager-PyTorch fallback.
"""
from __future__ import annotations

from kfac_b200 import assignment
from kfac_b200 import base_preconditioner
from kfac_b200 import distributed
from kfac_b200 import enums
from kfac_b200 import layers
from kfac_b200 import preconditioner
from kfac_b200.preconditioner import KFACPreconditioner

__version__ = '0.1.0'
__all__ = ['KFACPreconditioner', 'assignment', 'base_preconditioner', 'distributed', 'enums', 'layers',
           'preconditioner']
