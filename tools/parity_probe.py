"""How far one device iteration is from one oracle iteration, quantity by quantity (lock-step on config-3 blends), and how
the free-running trajectories separate.  Prints JSON.   python tools/parity_probe.py [--blends 8] [--iters 40]"""
import argparse
import copy
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
from scarlet_b200 import _native as nat  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--blends", type=int, default=8)
    ap.add_argument("--iters", type=int, default=40)
    ap.add_argument("--first", type=int, default=300)
    ap.add_argument("--converge", type=int, default=0, help="also: fit(200, 1e-2) of this many blends as ONE batch vs the oracle")
    a = ap.parse_args()
    nat.require_device()
    conv = []
    if a.converge:
        blends = [helpers.synthetic_blend(a.first + s)[3] for s in range(a.converge)]
        packs = helpers.pack(blends)
        bt = nat.Batch(**packs)
        bt.fit(200, 1e-2)
        out = bt.download()
        bt.close()
        off = packs["comp_off"]
        for i, bl in enumerate(blends):
            bl.fit(200, 1e-2)
            sl = slice(off[i], off[i + 1])
            got = np.einsum("kc,kyx->cyx", out["sed"][sl].astype(np.float64), out["morph"][sl].astype(np.float64))
            want = bl.get_model().astype(np.float64)
            n = min(int(out["n_iter"][i]), bl.it)
            dl = np.abs(out["logL"][i, :n] / np.array(bl.mse[:n], dtype=np.float64) - 1)
            md = np.abs(np.array([c.morph for c in bl.components]) - out["morph"][sl]).reshape(off[i + 1] - off[i], -1).max(axis=1)
            sd = np.abs(np.array([c.sed for c in bl.components]) - out["sed"][sl]).max(axis=1) / \
                (np.abs(np.array([c.sed for c in bl.components])).max() + 1e-30)
            conv.append(dict(seed=a.first + i, it_dev=int(out["n_iter"][i]), it_ref=bl.it, logL_max=float(dl.max()),
                             logL_last=float(dl[-1]), model=float(np.abs(got - want).max() / np.abs(want).max()),
                             morph_max=float(md.max()), sed_max=float(sd.max()),
                             centers_equal=bool((out["center"][sl] == np.array([s_.pixel_center for s_ in bl.sources])).all())))
    worst = dict(logL=0.0, L_sed=0.0, L_morph=0.0, sed=0.0, morph=0.0)
    per_iter = []
    free = []
    for s in range(a.first, a.first + a.blends):
        blend = helpers.synthetic_blend(s)[3]
        fr = copy.deepcopy(blend)
        # free-running device trajectory vs free-running oracle
        bt = nat.Batch(**helpers.pack([blend]))
        bt.fit(a.iters, 0.0)
        out = bt.download()
        bt.close()
        for _ in range(a.iters):
            fr.step(0.0)
        dl = np.abs(out["logL"][0, :a.iters] / np.array(fr.mse, dtype=np.float64) - 1)
        got = np.einsum("kc,kyx->cyx", out["sed"].astype(np.float64), out["morph"].astype(np.float64))
        want = fr.get_model().astype(np.float64)
        centers_equal = bool((out["center"] == np.array([s_.pixel_center for s_ in fr.sources])).all())
        free.append(dict(seed=s, logL_dev=[float(dl[i]) for i in (0, 1, 2, 4, 9, 19, a.iters - 1) if i < a.iters],
                         model_dev=float(np.abs(got - want).max() / np.abs(want).max()), centers_equal=centers_equal))
        # lock-step: one iteration from the oracle's state
        for it in range(a.iters):
            bt = nat.Batch(**helpers.pack([blend]))
            bt.fit(1, 0.0)
            o = bt.download()
            bt.close()
            blend.step(0.0)
            sed = np.array([c.sed for c in blend.components])
            morph = np.array([c.morph for c in blend.components])
            e = dict(logL=abs(o["logL"][0, 0] / blend.mse[-1] - 1), L_sed=abs(o["lipschitz"][0, 0] / blend.L_sed - 1),
                     L_morph=abs(o["lipschitz"][0, 1] / blend.L_morph - 1),
                     sed=float(np.abs(o["sed"] - sed).max() / np.abs(sed).max()), morph=float(np.abs(o["morph"] - morph).max()))
            for k in worst:
                worst[k] = max(worst[k], float(e[k]))
            if s == a.first:
                per_iter.append({k: float(v) for k, v in e.items()})
    print(json.dumps(dict(worst_single_iteration=worst, first_blend_per_iteration=per_iter[:12], free_running=free,
                          converged_batch=conv)))


if __name__ == "__main__":
    main()
