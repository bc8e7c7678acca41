"""B200-native loss head of KoDohwan/MIL-NCE_HowTo100M.

Drop-ins for the reference's ``loss.MILNCELoss``, ``soft_dtw_cuda.SoftDTW`` and
``utils.AllGather``; all arithmetic is in libb200head.so (csrc/, sm_100a only).
"""
from . import _lib
from .loss import GraphedMILNCEStep, MILNCELoss
from .soft_dtw_cuda import SoftDTW
from .composite import CDTW, SDTW_3, SDTW_CIDM, SDTW_negative
from .utils import AllGather, GatheredMILNCELoss, GraphedGatheredMILNCEStep

__all__ = ["MILNCELoss", "GraphedMILNCEStep", "SoftDTW", "AllGather", "GatheredMILNCELoss",
           "GraphedGatheredMILNCEStep", "CDTW", "SDTW_3", "SDTW_CIDM", "SDTW_negative", "_lib"]
