"""Generate tests/golden/sigmag_filter_case.npz with the REFERENCE's own sigma-G clipping class.

Run in the build container (needs /root/reference and torch on the CPU):
    python tests/golden/make_golden_filters.py
The reference module (src/kbmod/filters/sigma_g_filter.py) is loaded from where it lies, unmodified; the only thing
it imports from its package is ``kbmod.search.DebugTimer`` (a logging helper), which is stubbed because the kbmod
package itself cannot be imported here (astropy etc. are not installed).  The likelihood-curve and
likelihood-update expectations are the numpy expressions of results.py:596-606 and :630-645 evaluated verbatim.
"""
import importlib.util
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("KBMOD_REFERENCE", "/root/reference")


def load_reference_filter():
    sys.modules.setdefault("kbmod", types.ModuleType("kbmod"))
    stub = types.ModuleType("kbmod.search")

    class DebugTimer:
        def __init__(self, *a, **k):
            pass

        def stop(self):
            pass
    stub.DebugTimer = DebugTimer
    sys.modules["kbmod.search"] = stub
    spec = importlib.util.spec_from_file_location(
        "ref_sigma_g_filter", os.path.join(REF, "src", "kbmod", "filters", "sigma_g_filter.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def make_curves(rng, R, T):
    """psi/phi curves as get_all_psi_phi_curves returns them (invalid -> 0.0), plus a few hostile rows."""
    phi = (0.2 + 0.1 * rng.random((R, T))).astype(np.float32)
    psi = (rng.normal(0.0, 1.0, (R, T)) * np.sqrt(phi) + 1.5 * phi).astype(np.float32)
    out = rng.random((R, T)) < 0.03
    psi[out] += (rng.normal(0, 25, (R, T)).astype(np.float32) * np.sqrt(phi))[out]
    bad = rng.random((R, T)) < 0.1
    psi[bad] = 0.0
    phi[bad] = 0.0
    if R >= 6:
        psi[0], phi[0] = 0.0, 0.0                      # no data at all
        psi[1] = phi[1] * np.float32(3.0)              # (nearly) constant curve -> delta floor
        psi[2, :] = np.float32(2.0); phi[2, :] = np.float32(4.0)   # exactly constant
        psi[3] = -np.abs(psi[3])                       # all negative (clip_negative clips everything)
        psi[4, ::3] = np.nan                           # non-finite psi
        phi[5, 1::4] = np.inf                          # non-finite phi
    return psi, phi


def results_lh_curves(psi, phi, obs_valid=None, mask_value=np.nan):   # results.py:596-606
    valid = (phi != 0) & np.isfinite(psi) & np.isfinite(phi)
    if obs_valid is not None:
        valid = valid & obs_valid
    lh_matrix = np.full(psi.shape, mask_value, dtype=np.float32)
    lh_matrix[valid] = psi[valid] / np.sqrt(phi[valid])
    return lh_matrix


def results_update_likelihood(psi, phi, obs_valid):                   # results.py:630-645
    num_rows = psi.shape[0]
    phi_sum = (phi * obs_valid).sum(axis=1)
    psi_sum = (psi * obs_valid).sum(axis=1)
    num_obs = obs_valid.sum(axis=1)
    non_zero = phi_sum != 0
    lh = np.zeros((num_rows))
    lh[non_zero] = psi_sum[non_zero] / np.sqrt(phi_sum[non_zero])
    flux = np.zeros((num_rows))
    flux[non_zero] = psi_sum[non_zero] / phi_sum[non_zero]
    return lh, flux, num_obs


def main():
    ref = load_reference_filter()
    rng = np.random.default_rng(20261107)
    out = {}
    cases = [(40, 20, 25, 75, 2, False), (40, 20, 25, 75, 2, True), (25, 137, 10, 90, 3, False),
             (12, 1, 25, 75, 2, False), (30, 400, 15, 60, 1.5, True), (9, 7, 0.5, 99.5, 2, False)]
    out["cases"] = np.array(cases, dtype=np.float64)
    with np.errstate(all="ignore"):
        for i, (R, T, lo, hi, ns, neg) in enumerate(cases):
            psi, phi = make_curves(rng, R, T)
            prior = rng.random((R, T)) < 0.95                 # an earlier filter's obs_valid
            lh = results_lh_curves(psi, phi, prior)
            clipper = ref.SigmaGClipping(lo, hi, ns, neg)
            valid = clipper.compute_clipped_sigma_g_matrix(lh)
            ulh, uflux, uobs = results_update_likelihood(psi, phi, valid)
            out[f"c{i}_psi"], out[f"c{i}_phi"], out[f"c{i}_prior"] = psi, phi, prior
            out[f"c{i}_lh_curves"], out[f"c{i}_valid"] = lh, valid
            out[f"c{i}_coeff"] = np.float64(clipper.coeff)
            out[f"c{i}_lh"], out[f"c{i}_flux"], out[f"c{i}_obs"] = ulh, uflux, uobs.astype(np.int64)
        # single curves through compute_clipped_sigma_g (numpy float64 percentiles): rows of the first cases, as float64
        singles = []
        for i in (0, 2, 4):
            R, T, lo, hi, ns, neg = cases[i]
            lh = np.nan_to_num(out[f"c{i}_lh_curves"].astype(np.float64), nan=0.0)
            for neg_flag in (False, True):
                clipper = ref.SigmaGClipping(lo, hi, ns, neg_flag)
                for r in range(min(int(R), 8)):
                    singles.append((i, r, neg_flag, np.asarray(clipper.compute_clipped_sigma_g(lh[r]), dtype=np.int64)))
        out["single_meta"] = np.array([(a, b, int(c)) for a, b, c, _ in singles], dtype=np.int64)
        for k, (_, _, _, idx) in enumerate(singles):
            out[f"single_{k}"] = idx
    np.savez_compressed(os.path.join(HERE, "sigmag_filter_case.npz"), **out)
    print("wrote sigmag_filter_case.npz:", {k: v.shape for k, v in out.items() if k.endswith("_valid")})


if __name__ == "__main__":
    main()
