"""TEST INFRASTRUCTURE ONLY: CPU restatement of marginCaller's use of the posterior pairs.

Follows src/margin/marginCallerLib.py line by line -- posteriorProbabilityCalculationTargetFn (:39-68),
calcBasePosteriorProbs (:81-92), loadHmmSubstitutionMatrix (:94-100), variantCallSamFileTargetFn (:171-209) -- with the
posterior pairs coming from the oracle's getAlignedPairsUsingAnchors restatement and passing through the same text
round trip as in the reference ("%f" of p_int / 1e7, cPecanRealign.c:346-351, parsed back with float()).  Python 3
dicts iterate in insertion order where Python 2's did not; only the order of floating-point sums over the four
bases depends on it (last-bit differences).
"""
import math
from functools import reduce
from itertools import product

BASES = "ACGT"


def collate(pairs_per_read, reads, ref_names, ref_starts, read_starts=None):
    """:39-68.  pairs_per_read[i]: rows (p_int, x, y) of read i relative to its sub-sequences."""
    expectations = {}
    for i, pairs in enumerate(pairs_per_read):
        query = reads[i]
        for p_int, x, y in pairs:
            text = "%d\t%d\t%f" % (ref_starts[i] + x, y, float(p_int) / 10000000)  # writePosteriorProbs
            ref_position, query_position, posterior = map(float, text.split())
            assert 0.0 <= posterior <= 1.01
            key = (ref_names[i], int(ref_position))
            if key not in expectations:
                expectations[key] = dict(zip(BASES, [0.0] * len(BASES)))
            base = query[int(query_position)].upper()
            if base in BASES:
                expectations[key][base] += posterior
    return expectations


def substitution_matrix(emissions):
    """:94-100."""
    m = list(emissions[:len(BASES) ** 2])
    m = [m[i] / sum(m[4 * (i // 4):4 * (1 + i // 4)]) for i in range(len(m))]
    return dict(zip(product(BASES, BASES), m))


def null_matrix():
    return dict(zip(product(BASES, BASES), [1.0] * len(BASES) ** 2))


def calc_base_posterior_probs(observations, ref_base, evo, err):
    """:81-92."""
    log_probs = [math.log(evo[(ref_base.upper(), missing)]) +
                 reduce(lambda x, y: x + y, [math.log(err[(missing, obs)]) * observations[obs] for obs in BASES])
                 for missing in BASES]
    total = reduce(lambda x, y: x + math.log(1 + math.exp(y - x)), log_probs)
    return dict(zip(BASES, [math.exp(lp - total) for lp in log_probs]))


def variant_calls(expectations, references, err, threshold=0.3, evo=None):
    """:189-209.  Returns (calls, posteriors per key)."""
    evo = evo or null_matrix()
    calls, post = [], {}
    for (name, pos), e in expectations.items():
        ref_base = references[name][pos]
        total = sum(e.values())
        if not total > 0 or ref_base.upper() not in BASES:
            continue
        probs = calc_base_posterior_probs({b: float(e[b]) / total for b in BASES}, ref_base, evo, err)
        post[(name, pos)] = probs
        for b in BASES:
            if b != ref_base.upper() and probs[b] >= threshold:
                calls.append((name, pos, b, probs[b]))
    return calls, post
