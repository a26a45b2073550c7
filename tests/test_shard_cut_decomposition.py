"""Ground for exact sharded serving (SURVEY.md 8e exactness caveat), checked on the CPU oracle's own
operator* / operator+ (IndexSequence.cs:218-322).

Claim: every gap-run an operator node forms is a sub-chain of U = the union of ALL leaf lists of the request, chained with
gaps <= R_max (the largest |R| of the request), because a node's operands are subsets of U and its absR <= R_max.  Hence the
result of a request decomposes at the chain breaks of U: cutting the coordinate axis at a shard boundary X changes results
only inside the one chain of U that bridges X (if any).  Shard-local evaluation is exact outside that window, and evaluating
the leaf slices of the window alone reproduces the whole-corpus result inside it."""
import numpy as np
import pytest

from oracle_py import OracleIndex


def evaluate(o, expr, leaves):
    """expr: nested tuples ('*' | '+', left, right) over leaf names; leaves: name -> (sorted uint64 values, R)."""
    if isinstance(expr, str):
        v, R = leaves[expr]
        return o.encode(v), R
    op, a, b = expr
    wa, Ra = evaluate(o, a, leaves)
    wb, Rb = evaluate(o, b, leaves)
    return o.op(op, wa, Ra, wb, Rb)


def values(o, expr, leaves):
    return o.decode(evaluate(o, expr, leaves)[0])


def restrict(leaves, lo, hi):
    return {k: (v[(v >= lo) & (v < hi)], R) for k, (v, R) in leaves.items()}


EXPRS = [("*", "a", "b"), ("*", ("*", "a", "b"), "c"), ("*", "a", ("+", "b", "c")), ("+", ("*", "a", "b"), ("*", "b", "c")),
         ("*", ("*", "a", "c"), ("*", "b", "c"))]


@pytest.mark.parametrize("seed", range(6))
def test_results_decompose_at_chain_breaks(seed):
    rng = np.random.default_rng(1000 + seed)
    o = OracleIndex()
    span = 200_000
    checked_windows = differing = 0
    for trial in range(12):
        dens = rng.choice([40, 150, 600, 3000], size=3)                     # mean gap of each leaf: dense lists chain far
        ordered = bool(rng.integers(0, 2))
        leaves = {}
        for name, d in zip("abc", dens):
            v = np.unique(rng.integers(0, span, size=max(3, span // int(d))).astype(np.uint64))
            R = int(rng.choice([6, 44, 120, 262]))
            leaves[name] = (v, -R if ordered and name != "c" else R)
        rmax = max(abs(R) for _, R in leaves.values())
        U = np.unique(np.concatenate([v for v, _ in leaves.values()]))
        X = int(rng.integers(span // 4, 3 * span // 4))                     # the shard cut: first coordinate of the right shard
        # window = the chain of U (gaps <= rmax) that has elements on both sides of X
        li = int(np.searchsorted(U, X)) - 1                                 # last element < X
        lo = hi = None
        if 0 <= li < len(U) - 1 and int(U[li + 1]) - int(U[li]) <= rmax:
            i = li
            while i > 0 and int(U[i]) - int(U[i - 1]) <= rmax:
                i -= 1
            j = li + 1
            while j + 1 < len(U) and int(U[j + 1]) - int(U[j]) <= rmax:
                j += 1
            lo, hi = int(U[i]), int(U[j]) + 1
        for expr in EXPRS:
            whole = values(o, expr, leaves)
            left = values(o, expr, restrict(leaves, 0, X))
            right = values(o, expr, restrict(leaves, X, span + 1))
            shards = np.concatenate([left, right])
            if lo is None:
                assert np.array_equal(shards, whole), (trial, expr)         # no bridging chain: shard-local results are exact
                continue
            outside = lambda v: v[(v < lo) | (v >= hi)]                     # noqa: E731
            assert np.array_equal(outside(shards), outside(whole)), (trial, expr)
            window = values(o, expr, restrict(leaves, lo, hi))
            assert np.array_equal(window, whole[(whole >= lo) & (whole < hi)]), (trial, expr)
            checked_windows += 1
            differing += int(not np.array_equal(shards, whole))
    assert checked_windows > 0
    if seed == 0:
        assert differing > 0   # the caveat is real: shard-local evaluation does differ inside a bridging window
