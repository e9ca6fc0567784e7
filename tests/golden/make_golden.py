"""Generate the golden fixtures in this directory from the REFERENCE ITSELF (oracle/_ref, built from
/root/reference by oracle/build_ref.sh).  Run here, in the container that has /root/reference; the outputs
(small .npz files) are committed so the GPU box -- which has no /root/reference -- can check against them.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from cases import SEARCH_CONFIGS, make_case  # noqa: E402
from oracle import ref  # noqa: E402

S, Hn = ref.load()


def ref_stack(case, nb):
    ppa = S.PsiPhiArray()
    S.fill_psi_phi_array_from_image_arrays(ppa, nb, [a.copy() for a in case["sci"]], [a.copy() for a in case["var"]],
                                           list(case["psfs"]), list(case["times"]))
    return ppa


def main():
    case = make_case()
    out = {}
    cands = [S.Trajectory(vx=float(a), vy=float(b)) for a, b in zip(case["vx"], case["vy"])]
    # P2 / P1: convolution and psi/phi images
    out["conv_sci0"] = np.asarray(S.convolve_image_cpu(case["sci"][0].copy(), case["psfs"][0]))
    out["psi0"] = np.asarray(S.generate_psi(case["sci"][0], case["var"][0], case["psfs"][0]))
    out["phi0"] = np.asarray(S.generate_phi(case["var"][0], case["psfs"][0]))
    for nb in (-1, 2, 1):
        ppa = ref_stack(case, nb)
        tag = {-1: "f32", 2: "u16", 1: "u8"}[nb]
        out[f"stack_{tag}"] = np.asarray(Hn.psi_phi_bytes(ppa))
        out[f"meta_{tag}"] = np.array([ppa.psi_min_val, ppa.psi_max_val, ppa.psi_scale, ppa.phi_min_val,
                                       ppa.phi_max_val, ppa.phi_scale], dtype=np.float32)
        for ci, (mo, ml, sg, b, K) in enumerate(SEARCH_CONFIGS):
            b = b or (0, case["W"], 0, case["H"])
            p = Hn.make_params(mo, ml, sg is not None, *(sg or (0.25, 0.75, -1.0)), nb, b[0], b[1], b[2], b[3], K)
            r = Hn.search_slots(ppa, p, cands)
            rec = np.zeros(len(r[0]), dtype=[("vx", "<f4"), ("vy", "<f4"), ("lh", "<f4"), ("flux", "<f4"),
                                             ("x", "<i4"), ("y", "<i4"), ("obs_count", "<i4")])
            for name, arr in zip(("x", "y", "vx", "vy", "lh", "flux", "obs_count"), r):
                rec[name] = np.asarray(arr)
            out[f"slots_{tag}_{ci}"] = rec
        if nb == -1:
            # single trajectories through the host-compiled evaluateTrajectory, sigma-G on
            p = Hn.make_params(3, 0.0, True, 0.25, 0.75, 0.7413, nb, 0, case["W"], 0, case["H"], 8)
            ev = []
            for (ox, oy, ovx, ovy) in case["objects"]:
                t = Hn.eval_host(ppa, p, S.Trajectory(x=ox, y=oy, vx=ovx, vy=ovy))
                ev.append((t.x, t.y, t.vx, t.vy, t.lh, t.flux, t.obs_count))
            out["eval_sigmag"] = np.array(ev, dtype=np.float64)
            # psi/phi curves (stack_search.cpp:14-31) through the reference StackSearch
            ss = S.StackSearch([a.copy() for a in case["sci"]], [a.copy() for a in case["var"]], list(case["psfs"]),
                               list(case["times"]), -1)
            trjs = [S.Trajectory(x=ox, y=oy, vx=ovx, vy=ovy) for (ox, oy, ovx, ovy) in case["objects"]]
            trjs.append(S.Trajectory(x=3, y=3, vx=0.0, vy=0.0))
            trjs.append(S.Trajectory(x=-4, y=2, vx=2.5, vy=0.25))
            out["curves"] = np.asarray(ss.get_all_psi_phi_curves(trjs))
            out["curves_trj"] = np.array([(t.x, t.y, t.vx, t.vy) for t in trjs], dtype=np.float64)
            # the reference's own CPU search_all (stack_search.cpp:221-284, on_gpu=False), K=4, no filtering
            ss.set_min_obs(0)
            ss.set_min_lh(-1.0)
            ss.set_results_per_pixel(4)
            ss.search_all(cands, False)
            res = ss.get_all_results()
            out["cpu_search_all"] = np.array([(t.x, t.y, t.vx, t.vy, t.lh, t.flux, t.obs_count) for t in res],
                                             dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, "small_case.npz"), **out)
    print("wrote", os.path.join(HERE, "small_case.npz"), {k: v.shape for k, v in out.items() if k.startswith("stack")})


if __name__ == "__main__":
    main()
