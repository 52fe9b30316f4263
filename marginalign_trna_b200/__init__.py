"""marginalign_trna_b200: B200-native anchor-banded pair-HMM forward-backward (posterior match probabilities and
EM expectations) behind the cPecanRealign call sites of marginAlign / marginCaller / cPecanEm.

(The project is named marginalign-trna_b200; the package directory carries the importable spelling.)
"""
from . import build, caller, capi, chain_sam, em, hmm, realign_sam, seqio, shard, synth  # noqa: F401
from .hmm import Hmm  # noqa: F401
from .pairhmm import (AlignmentParameters, DevicePlan, PairHmmEngine, PairHmmModel, Plan, PlanCache,  # noqa: F401
                      PosteriorPairs, RealignedCigars, jitter_table)
from .synth import AlignmentBatch  # noqa: F401

__version__ = "0.1.0"
