"""EM training loop around the GPU E-step: the in-process replacement of
submodules/cPecan/cPecanEm.py:107-242 (expectationMaximisation / expectationMaximisation2 / calculateMaximisation /
expectationMaximisationTrials), keeping its option names and semantics.

What changes: `calculateExpectations` (cPecanEm.py:178-180) spawned one `cPecanRealign --outputExpectations` per split
file and the M-step summed `%f` text files; here the E-step is one batched GPU call per iteration
(get_expectations_using_anchors), summed across ranks by one all-reduce, and the M-step runs identically on every
rank.  What stays: model initialisation order, sampling of alignments by aligned length, normalise / tieEmissions /
keep-old-emissions rules, running likelihoods, model file written every iteration (the reference's checkpoint),
random-restart trials with best-likelihood selection.
"""
import inspect
import random
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

from .hmm import Hmm
from .pairhmm import AlignmentParameters, PairHmmModel, PlanCache


@dataclass
class EmOptions:
    """Option group of cPecanEm.py:361-402 with marginAlign's overrides (src/margin/marginAlign.py:38-54)."""
    modelType: str = "fiveStateAsymmetric"
    inputModel: Optional[str] = None
    iterations: int = 100
    trials: int = 3
    randomStart: bool = True
    setJukesCantorStartingEmissions: Optional[float] = None
    trainEmissions: bool = True
    tieEmissions: bool = False
    maxAlignmentLengthPerJob: int = 700000
    maxAlignmentLengthToSample: int = 50000000
    outputTrialHmms: bool = False
    # optionsToRealign of marginAlign.py:41: --diagonalExpansion=10 --splitMatrixBiggerThanThis=300
    diagonalExpansion: int = 10
    splitMatrixBiggerThanThis: int = 300
    seed: Optional[int] = None
    py2_compat_files: bool = False  # write models with Python-2 str(float) rounding (12 significant digits)
    log: List[str] = field(default_factory=list)

    def params(self):
        return AlignmentParameters(diagonal_expansion=self.diagonalExpansion,
                                   split_matrix_bigger_than=self.splitMatrixBiggerThanThis)


def initial_hmm(options, rng=random):
    """cPecanEm.py:109-123."""
    if options.inputModel is not None:
        hmm = Hmm.loadHmm(options.inputModel)
        hmm.normalise()
    else:
        hmm = Hmm(options.modelType)
        if options.randomStart:
            hmm.randomise(rng)
        else:
            hmm.equalise()
    if options.setJukesCantorStartingEmissions is not None:
        hmm.setEmissionsToJukesCantor(float(options.setJukesCantorStartingEmissions))
    return hmm


def sample_alignments(batch, options, rng=random):
    """cPecanEm.py:128-158: alignments are grouped, in input order, into chunks that close once their summed aligned
    length (mean of the two spans) exceeds maxAlignmentLengthPerJob; the chunks are shuffled and taken until
    maxAlignmentLengthToSample bases are covered.  Returns the selected alignment indices (ascending)."""
    lx = (batch.off_x[1:] - batch.off_x[:-1]).astype(np.float64)
    ly = (batch.off_y[1:] - batch.off_y[:-1]).astype(np.float64)
    length = (lx + ly) / 2.0
    chunks, cur, acc = [], [], 0.0
    for i in range(batch.n):
        cur.append(i)
        acc += length[i]
        if acc > options.maxAlignmentLengthPerJob:
            chunks.append((cur, acc))
            cur, acc = [], 0.0
    if cur:
        chunks.append((cur, acc))
    rng.shuffle(chunks)
    picked, total = [], 0.0
    for idx, l in chunks:
        total += l
        picked.extend(idx)
        if total >= options.maxAlignmentLengthToSample:
            break
    return np.array(sorted(picked), dtype=np.int64)


def maximisation(expectations: Hmm, previous: Hmm, options):
    """calculateMaximisation (cPecanEm.py:182-202): normalise; tie or keep emissions."""
    hmm = expectations
    hmm.normalise()
    if options.trainEmissions:
        if options.tieEmissions:
            hmm.tieEmissions()
    else:
        hmm.emissions = previous.emissions.copy()
    return hmm


def expectation_maximisation(runner, batch, output_model, options, rng=random):
    """One EM run.  `runner` provides get_expectations_using_anchors(model, batch, params) -> (T, E, likelihood):
    a PairHmmEngine (one GPU) or a shard.ShardedPairHmm (all ranks; every rank must call this with the same batch
    and the same rng seed).  Returns the final Hmm; writes `output_model` every iteration when a path is given."""
    hmm = initial_hmm(options, rng)
    if output_model:
        hmm.write(output_model, py2_compat=options.py2_compat_files)
    idx = sample_alignments(batch, options, rng)
    work = batch if len(idx) == batch.n else batch.subset(idx)
    params = options.params()
    running = []
    # the band plans do not depend on the model: runners that can keep them in HBM plan `work` once (pairhmm.PlanCache)
    extra = {}
    if "plan_cache" in inspect.signature(runner.get_expectations_using_anchors).parameters:
        extra["plan_cache"] = PlanCache()
    for iteration in range(options.iterations):
        model = PairHmmModel.from_hmm(hmm)
        T, E, lik = runner.get_expectations_using_anchors(model, work, params, **extra)
        new = Hmm(hmm.modelType)
        new.addExpectations(T, E, lik)
        hmm = maximisation(new, hmm, options)
        running.append(hmm.likelihood)
        options.log.append("On %i iteration got likelihood: %s for model-type: %s" % (iteration, hmm.likelihood,
                                                                                      hmm.modelType))
        if output_model:
            hmm.write(output_model, py2_compat=options.py2_compat_files)
    hmm.runningLikelihoods = running
    if output_model:
        hmm.write(output_model, py2_compat=options.py2_compat_files)
    return hmm


def expectation_maximisation_trials(runner, batch, output_model, options):
    """expectationMaximisationTrials (cPecanEm.py:217-242): with a fixed start there is one run; with random starts
    `trials` independent runs and the highest final likelihood wins."""
    rng = random.Random(options.seed) if options.seed is not None else random
    if options.inputModel is not None or not options.randomStart:
        return expectation_maximisation(runner, batch, output_model, options, rng)
    hmms = []
    for i in range(options.trials):
        trial_path = (output_model + "_%i" % i) if (output_model and options.outputTrialHmms) else None
        hmms.append(expectation_maximisation(runner, batch, trial_path, options, rng))
    best = max(hmms, key=lambda h: h.likelihood)
    if output_model:
        best.write(output_model, py2_compat=options.py2_compat_files)
    return best
