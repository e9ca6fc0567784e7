"""Generate tests/golden/stamps_case.npz with the REFERENCE's own stamp / coadd functions.

    python tests/golden/make_golden_stamps.py        (build container: needs /root/reference and torch on the CPU)

core/stamp_utils.py and trajectory_utils.py are loaded from where they lie, unmodified.  Two of their imports do not
exist here and are stubbed: ``numba`` (its ``jit`` decorator becomes the identity, so the same Python source runs
uncompiled) and ``kbmod.core.image_stack_py`` / ``kbmod.search`` (class names only used for isinstance checks and
type hints on code paths not exercised).  The loop over results is the one of filters/stamp_filters.py:110-149
(append_coadds, nightly=False), which itself needs astropy tables and is therefore repeated here line for line.
"""
import importlib.util
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
REF = os.environ.get("KBMOD_REFERENCE", "/root/reference")


def load_reference_stamp_utils():
    numba = types.ModuleType("numba")
    numba.jit = lambda *a, **k: (lambda f: f)
    numba.typed = types.SimpleNamespace(List=list)
    sys.modules["numba"] = numba
    for name in ("kbmod", "kbmod.core", "kbmod.core.image_stack_py", "kbmod.search"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["kbmod.core.image_stack_py"].ImageStackPy = type("ImageStackPy", (), {})
    spec = importlib.util.spec_from_file_location("ref_stamp_utils", os.path.join(REF, "src", "kbmod", "core", "stamp_utils.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def reference_positions(times, x0, vx):
    # trajectory_utils.py:64-73 (predict_pixel_locations, centered=True, as_int=True), verbatim arithmetic
    pos = vx[:, np.newaxis] * times[np.newaxis, :] + x0[:, np.newaxis]
    pos = pos + 0.5
    return pos.astype(int)


def main():
    from cases import make_case
    su = load_reference_stamp_utils()
    case = make_case(seed=53, T=14, H=40, W=48, n_cand=30, mask_frac=0.08)
    sci, var, times = case["sci"], case["var"], np.asarray(case["times"], dtype=np.float64)
    var[3][5:9, 6:10] = 0.0                       # zero variance: ignored by the weighted coadd
    rng = np.random.default_rng(8)
    R, radius = 24, 3
    x = rng.integers(-3, case["W"] + 3, R).astype(np.float64)
    y = rng.integers(-3, case["H"] + 3, R).astype(np.float64)
    vx = rng.uniform(-6, 6, R).astype(np.float32).astype(np.float64)
    vy = rng.uniform(-6, 6, R).astype(np.float32).astype(np.float64)
    x[0], y[0], vx[0], vy[0] = -50, -50, 0.0, 0.0   # never on the image
    obs_valid = rng.random((R, len(times))) < 0.8
    obs_valid[1, :] = False                          # nothing selected -> all coadds zero
    obs_valid[2, :] = True
    xvals, yvals = reference_positions(times, x, vx), reference_positions(times, y, vy)
    out = {"x": x, "y": y, "vx": vx, "vy": vy, "obs_valid": obs_valid, "radius": np.int64(radius),
           "sci": np.stack(sci), "var": np.stack(var), "times": times}
    import warnings
    # T x H x W arrays: for a *list* of images the reference returns [] when nothing is selected and its own
    # coadd_mean then fails on the shape; the array form returns an empty (0, S, S) stack and all-zero coadds.
    sci, var = np.stack(sci), np.stack(var)
    for tag, use_valid in (("all", False), ("valid", True)):
        co = {k: np.zeros((R, 2 * radius + 1, 2 * radius + 1), dtype=np.float32) for k in ("sum", "mean", "median", "weighted")}
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            for idx in range(R):
                inc = obs_valid[idx] if use_valid else np.full(len(times), True)
                s = np.asanyarray(su.extract_stamp_stack(sci, xvals[idx, :], yvals[idx, :], radius, to_include=inc))
                v = np.asanyarray(su.extract_stamp_stack(var, xvals[idx, :], yvals[idx, :], radius, to_include=inc))
                co["mean"][idx][:, :] = su.coadd_mean(s)
                co["median"][idx][:, :] = su.coadd_median(s, device="cpu")
                co["sum"][idx][:, :] = su.coadd_sum(s)
                co["weighted"][idx][:, :] = su.coadd_weighted(s, v)
        for k, a in co.items():
            out[f"{tag}_{k}"] = a
    out["all_stamps_r2"] = np.asanyarray(su.extract_stamp_stack(sci, xvals[2], yvals[2], radius)).astype(np.float32)
    out["curve_values"] = np.asarray(su.extract_curve_values(sci, xvals, yvals), dtype=np.float64)     # R x T
    out["curve_values_one"] = np.asarray(su.extract_curve_values(sci, xvals[2], yvals[2]), dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, "stamps_case.npz"), **out)
    print("wrote stamps_case.npz")


if __name__ == "__main__":
    main()
