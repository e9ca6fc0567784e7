"""Record comparison helpers shared by the GPU parity tests and __graft_entry__.smoke().

Two bars:
* ``assert_records_equal``   -- bit-exact on every field (the generic kernel visits candidates in list order and
                                reproduces the reference including its history-dependent order of exact lh ties).
* ``assert_slots_equal_modulo_ties`` -- what BASELINE.json's north_star asks of the ranked list: identical
                                modulo exact-score ties.  lh is bit-exact slot by slot; wherever a pixel's lh value
                                is unique (and not the boundary value of a full list) every other field is
                                bit-exact too; and *every* record returned is re-evaluated by the oracle and must
                                reproduce lh, flux and obs_count exactly (so a tied record is a genuine tie).
"""
import numpy as np

FLOAT_FIELDS = ("vx", "vy", "lh", "flux")
INT_FIELDS = ("x", "y", "obs_count")
EMPTY_LH = -np.finfo(np.float32).max


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def assert_records_equal(got, want, what=""):
    assert len(got) == len(want), f"{what}: {len(got)} vs {len(want)} records"
    for f in INT_FIELDS:
        assert np.array_equal(got[f], want[f]), f"{what}: field {f} differs"
    for f in FLOAT_FIELDS:
        bad = np.flatnonzero(bits(got[f]) != bits(want[f]))
        assert bad.size == 0, f"{what}: field {f} differs at {bad[:5]}: {got[f][bad[:5]]} vs {want[f][bad[:5]]}"


def assert_records_valid(got, orc, stack, params, what=""):
    """Every non-empty record must be exactly what the oracle computes for that (x, y, vx, vy)."""
    real = ~((got["lh"] == np.float32(EMPTY_LH)) & (got["obs_count"] == 0) & (got["vx"] == 0) & (got["vy"] == 0))
    r = got[real]
    if len(r) == 0:
        return
    lh, flux, obs = orc.eval_many(stack, params, r["x"], r["y"], r["vx"], r["vy"])
    assert np.array_equal(bits(lh), bits(r["lh"])), f"{what}: lh of a returned record is not reproducible"
    assert np.array_equal(bits(flux), bits(r["flux"])), f"{what}: flux of a returned record is not reproducible"
    assert np.array_equal(obs, r["obs_count"]), f"{what}: obs_count of a returned record is not reproducible"


def assert_slots_equal_modulo_ties(got, want, K, orc, stack, params, what=""):
    assert len(got) == len(want), f"{what}: {len(got)} vs {len(want)} slots"
    for f in ("x", "y"):
        assert np.array_equal(got[f], want[f]), f"{what}: field {f} differs"
    bad = np.flatnonzero(bits(got["lh"]) != bits(want["lh"]))
    assert bad.size == 0, f"{what}: lh differs at {bad[:5]}: {got['lh'][bad[:5]]} vs {want['lh'][bad[:5]]}"
    g, w = got.reshape(-1, K), want.reshape(-1, K)
    lh = w["lh"]
    unique = np.ones(lh.shape, dtype=bool)
    for i in range(K):
        for j in range(K):
            if i != j:
                unique[:, i] &= lh[:, i] != lh[:, j]
    full = lh[:, K - 1] != np.float32(EMPTY_LH)           # a full list may hide a tie with a dropped candidate
    unique &= ~(full[:, None] & (lh == lh[:, K - 1:K]))
    for f in ("vx", "vy", "flux"):
        bad = np.argwhere(unique & (bits(g[f]) != bits(w[f])))
        assert bad.size == 0, f"{what}: {f} differs outside ties at pixel/slot {bad[:3].tolist()}"
    assert np.array_equal(g["obs_count"][unique], w["obs_count"][unique]), f"{what}: obs_count differs outside ties"
    assert_records_valid(got, orc, stack, params, what)


def assert_lists_equal_modulo_ties(got, want, orc, stack, params, what=""):
    """Filtered + sorted result lists: same length, same lh sequence; identical records wherever lh is unique in the
    whole list; every record valid."""
    assert len(got) == len(want), f"{what}: {len(got)} vs {len(want)} results"
    assert np.array_equal(bits(got["lh"]), bits(want["lh"])), f"{what}: ranked lh sequence differs"
    lh = want["lh"]
    uniq = np.ones(len(lh), dtype=bool)
    if len(lh) > 1:
        same = lh[1:] == lh[:-1]
        uniq[1:] &= ~same
        uniq[:-1] &= ~same
    # a unique lh may still be a boundary tie of its pixel's list; identity is only required when it also matches
    ident = np.ones(len(lh), dtype=bool)
    for f in ("x", "y"):
        ident &= got[f] == want[f]
    for f in ("vx", "vy"):
        ident &= bits(got[f]) == bits(want[f])
    frac = float((ident | ~uniq).mean()) if len(lh) else 1.0
    assert frac > 0.99, f"{what}: only {frac:.4f} of the non-tied results are identical"
    assert_records_valid(got, orc, stack, params, what)
