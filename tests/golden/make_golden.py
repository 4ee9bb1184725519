#!/usr/bin/env python
"""Generate tests/golden/golden.npz by running the REFERENCE defenses themselves.

Run in the build container only (needs /root/reference, which the GPU box does not
have):  PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

The reference package is loaded read-only under the alias ``ref_defenses`` so it
cannot collide with this repo's own ``src.defenses`` mirror.  Every case stores the
flattened fp32 input matrix X[n, d], the per-tensor segment sizes the update lists
were cut into, the defense config and everything the reference produced (aggregate,
selected/rejected lists, scores, norms, trust scores ...).  The reference ships no
golden vectors of its own (its tests assert only qualitative facts), so these
fixtures are what pins the oracle and, through it, the CUDA path.
"""
import importlib.util
import json
import os
import sys

import numpy as np
import torch

REF = "/root/reference/src/defenses"
REF_ATTACKS = "/root/reference/src/attacks"
HERE = os.path.dirname(os.path.abspath(__file__))


def load_reference():
    spec = importlib.util.spec_from_file_location(
        "ref_defenses", os.path.join(REF, "__init__.py"), submodule_search_locations=[REF])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["ref_defenses"] = mod
    spec.loader.exec_module(mod)
    return mod


def load_reference_attacks():
    """model_poisoning.py alone (the package __init__ pulls in the dataset attacks)."""
    import types
    pkg = types.ModuleType("ref_attacks")
    pkg.__path__ = [REF_ATTACKS]
    sys.modules["ref_attacks"] = pkg
    for name in ("base_attack", "model_poisoning"):
        spec = importlib.util.spec_from_file_location(f"ref_attacks.{name}", os.path.join(REF_ATTACKS, f"{name}.py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[f"ref_attacks.{name}"] = mod
        spec.loader.exec_module(mod)
    return sys.modules["ref_attacks.model_poisoning"]


def make_rows(rng, n, d, n_byz=0, kind="replace", theta_sigma=0.05, delta_sigma=1e-3):
    """Rows as run_experiments.py:238 produces them: full weights theta + delta_i."""
    theta = rng.normal(0, theta_sigma, d)
    X = theta[None, :] + rng.normal(0, delta_sigma, (n, d))
    for b in range(n_byz):
        if kind == "replace":          # model-replacement style, cf. model_poisoning.py:65-80
            X[b] = theta + 10 * (X[b] - theta)
        elif kind == "scale":          # scaling attack, model_poisoning.py:236
            X[b] = 100 * X[b]
        elif kind == "flip":           # IPM-like sign flip, model_poisoning.py:274-276
            X[b] = theta - (X[b] - theta)
    return X.astype(np.float32)


def to_lists(X, seg):
    """[n, d] matrix -> list[n][P] of tensors (segment p reshaped 2-D when even)."""
    out = []
    offs = np.concatenate([[0], np.cumsum(seg)])
    for i in range(X.shape[0]):
        row = []
        for p, s in enumerate(seg):
            t = torch.from_numpy(X[i, offs[p]:offs[p + 1]].copy())
            if s % 2 == 0 and s > 2:
                t = t.reshape(2, s // 2)
            row.append(t)
        out.append(row)
    return out


def flat(lst):
    return torch.cat([t.flatten().float() for t in lst]).numpy()


def main():
    ref = load_reference()
    rng = np.random.RandomState(20261104)
    cases, arrays = {}, {}

    def put(name, cfg, X, seg, **outs):
        cases[name] = dict(cfg=cfg, seg=[int(s) for s in seg],
                           fields=sorted(outs.keys()))
        arrays[f"{name}/X"] = X
        for k, v in outs.items():
            arrays[f"{name}/{k}"] = np.asarray(v)

    ones = lambda n: [100] * n  # noqa: E731

    # ---------------- median ----------------
    edge = np.array([0.0, -0.0, 1.0, -1.0, 2.5, np.inf, -np.inf, 1e-38, -1e-38, 3e38], dtype=np.float32)
    for n, d, seg in [(2, 64, [64]), (3, 70, [30, 40]), (5, 257, [257]), (10, 1000, [288, 32, 680]),
                      (16, 300, [300]), (33, 129, [129]), (100, 200, [120, 80]), (128, 96, [96]),
                      (130, 64, [64])]:
        X = make_rows(rng, n, d, n_byz=n // 5)
        X[:, : d // 4] = rng.choice(edge, size=(n, d // 4))          # ties, +-0, +-inf, denormals
        X[:, d // 4] = 0.0                                             # all-equal column
        X[::2, d // 4 + 1] = -0.0                                      # mixed-sign zeros only
        X[1::2, d // 4 + 1] = 0.0
        d_ = ref.MedianDefense({})
        out = flat(d_.aggregate(to_lists(X, seg), ones(n)))
        put(f"median_n{n}", {}, X, seg, out=out)
    # NaN columns (payloads preserved): any NaN -> first NaN
    n, d = 7, 96
    X = make_rows(rng, n, d)
    Xi = X.view(np.uint32)
    for j in range(0, d, 3):
        r = rng.randint(0, n)
        Xi[r, j] = rng.choice([0x7FC00000, 0xFFC00001, 0x7F800123, 0xFFFFFFFF])
        if j % 2 == 0:
            Xi[(r + 2) % n, j] = 0x7FC00ABC
    out = flat(ref.MedianDefense({}).aggregate(to_lists(X, [d]), ones(n)))
    put("median_nan", {}, X, [d], out=out)

    # ---------------- trimmed mean ----------------
    for n, ratio, d, seg, byz in [(5, 0.2, 130, [100, 30], 1), (10, 0.1, 500, [500], 2),
                                  (20, 0.1, 1000, [512, 488], 4), (20, 0.2, 1000, [1000], 4),
                                  (100, 0.29, 120, [120], 10), (33, 0.3, 99, [99], 5),
                                  (2, 0.5, 40, [40], 0), (3, 0.4, 41, [41], 0), (4, 0.6, 33, [33], 1)]:
        X = make_rows(rng, n, d, n_byz=byz, kind="scale")
        if n >= 10:
            X[:, :8] = rng.choice(edge, size=(n, 8))
            X.view(np.uint32)[1, 9] = 0x7FC00000      # one NaN: sorts last, trimmed off
        df = ref.TrimmedMeanDefense({"trim_ratio": ratio})
        out = flat(df.aggregate(to_lists(X, seg), ones(n)))
        put(f"trimmed_n{n}_r{ratio}", {"trim_ratio": ratio}, X, seg, out=out,
            k=df.num_trimmed_per_end)

    # ---------------- Krum / Multi-Krum ----------------
    for name, n, f, mk, d, seg, byz, dup in [("krum_single", 7, 2, 1, 400, [300, 100], 2, False),
                                             ("krum_c1like", 10, 2, 5, 2000, [288, 32, 1680], 2, True),
                                             ("krum_multi3", 9, 3, 3, 513, [513], 3, False),
                                             ("krum_n20", 20, 4, 10, 800, [800], 4, True)]:
        X = make_rows(rng, n, d, n_byz=byz, kind="replace")
        if dup:
            X[1] = X[0]                       # two colluders with identical rows (tie in scores)
        df = ref.KrumDefense({"num_malicious": f, "multi_k": mk})
        res = df.aggregate(to_lists(X, seg), ones(n))
        put(name, {"num_malicious": f, "multi_k": mk}, X, seg, out=flat(res),
            selected=df.selected_clients, rejected=df.rejected_clients,
            scores=np.array(df.client_scores, dtype=np.float64),
            D=df._compute_distances(to_lists(X, seg)))

    # ---------------- FedAvg ----------------
    n, d, seg = 6, 700, [200, 500]
    X = make_rows(rng, n, d)
    ne = [10, 200, 3000, 7, 64, 16777217]
    out = flat(ref.NoDefense({}).aggregate(to_lists(X, seg), ne))
    put("fedavg", {}, X, seg, out=out, num_examples=ne)

    # ---------------- FLTrust (synthetic g0 through an override of the training step) ----------------
    class FLTrustFixed(ref.FLTrustDefense):
        def __init__(self, cfg, g0):
            super().__init__(cfg)
            self._g0 = g0

        def _compute_server_gradient(self, global_params):     # fltrust.py:93-150 is model training
            self.server_gradient = self._g0
            return self._g0

    for name, n, d, seg, mode in [("fltrust_mixed", 12, 900, [400, 500], "mixed"),
                                  ("fltrust_allneg", 4, 128, [128], "allneg"),
                                  ("fltrust_zero_row", 5, 200, [200], "zero_row")]:
        g = rng.normal(0, 1e-3, d).astype(np.float32)
        X = (g[None, :] * rng.uniform(0.5, 2.0, (n, 1)) + rng.normal(0, 1e-3, (n, d))).astype(np.float32)
        if mode == "mixed":
            X[0] = -X[0]
            X[1] = 100 * X[1]
            X[2] = rng.normal(0, 1e-3, d)
        elif mode == "allneg":
            X = -np.abs(X) * np.sign(g)[None, :]
        elif mode == "zero_row":
            X[3] = 0.0
        g0 = to_lists(g[None, :], seg)[0]
        df = FLTrustFixed({}, g0)
        res = df.aggregate(to_lists(X, seg), ones(n), global_params=g0)
        put(name, {}, X, seg, out=flat(res), g0=g, trust=np.array(df.trust_scores, dtype=np.float64),
            detected=np.array(df.detect_malicious(None, None), dtype=np.int64))

    # ---------------- clipping family ----------------
    n, d, seg = 8, 640, [600, 40]
    X = make_rows(rng, n, d, n_byz=2, kind="scale", theta_sigma=0.01)
    ne = [50, 60, 70, 80, 90, 100, 110, 120]
    med_norm = float(np.median(np.linalg.norm(X.astype(np.float64), axis=1)))
    for ctype, clip in [("l2", med_norm), ("linf", float(np.median(np.abs(X).max(axis=1))))]:
        df = ref.GradientClippingDefense({"clip_norm": clip, "clip_type": ctype})
        res = df.aggregate(to_lists(X, seg), ne)
        put(f"gradclip_{ctype}", {"clip_norm": clip, "clip_type": ctype}, X, seg, out=flat(res),
            norms=np.array(df.original_norms), clipped_count=df.clipped_count, num_examples=ne)
    df = ref.DPSGDDefense({"clip_norm": med_norm, "noise_multiplier": 0.0})
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        res = df.aggregate(to_lists(X, seg), ne)         # sigma = 0: the randn term is exactly 0
    put("dpsgd_nonoise", {"clip_norm": med_norm, "noise_multiplier": 0.0}, X, seg, out=flat(res),
        num_examples=ne)
    for name, lo, hi in [("normbound_some", 0.0, med_norm * 1.5), ("normbound_all_rejected", 1e6, 2e6),
                         ("normbound_min", med_norm * 0.999, 1e9)]:
        df = ref.NormBoundingDefense({"max_norm": hi, "min_norm": lo})
        res = df.aggregate(to_lists(X, seg), ne)
        put(name, {"max_norm": hi, "min_norm": lo}, X, seg, out=flat(res),
            rejected=np.array(df.rejected_clients, dtype=np.int64), num_examples=ne)

    # ---------------- geometric median ----------------
    n, d, seg = 9, 300, [300]
    X = make_rows(rng, n, d, n_byz=2, kind="scale")
    df = ref.GeometricMedianDefense({"max_iters": 50})
    res = df.aggregate(to_lists(X, seg), ones(n))
    put("geomed", {"max_iters": 50}, X, seg, out=flat(res), num_iters=df.num_iters)

    # ======== round 2 additions (appended: the cases above keep their random streams) ========
    # ---------------- Krum with non-finite clients (NaN / Inf rows are ranked last, krum.py:95,126,174) --------
    import warnings as _w
    U32 = np.uint32
    for name, n, f, mk, d, seg, spec in [
            ("krum_nan_n9", 9, 2, 3, 400, [300, 100], {4: "nan"}),
            ("krum_inf_n9", 9, 2, 3, 400, [400], {2: "+inf@7", 6: "-inf@9"}),
            ("krum_infpair_n12", 12, 3, 4, 500, [256, 244], {1: "+inf@5", 8: "+inf@5", 10: "nan"}),
            ("krum_nan_n40", 40, 5, 10, 300, [300], {0: "nan", 17: "+inf@3"}),     # row 0 is always a centre row
            ("krum_nan_n70", 70, 10, 35, 260, [260], {33: "nan"})]:
        X = make_rows(rng, n, d, n_byz=2, kind="replace")
        for r, what in spec.items():
            if what == "nan":
                X.view(U32)[r, 11] = 0x7FC00000
            else:
                X[r, int(what.split("@")[1])] = np.inf if what[0] == "+" else -np.inf
        df = ref.KrumDefense({"num_malicious": f, "multi_k": mk})
        with _w.catch_warnings(), np.errstate(all="ignore"):
            _w.simplefilter("ignore")
            res = df.aggregate(to_lists(X, seg), ones(n))
            D = df._compute_distances(to_lists(X, seg))
        put(name, {"num_malicious": f, "multi_k": mk}, X, seg, out=flat(res), selected=df.selected_clients,
            rejected=df.rejected_clients, scores=np.array(df.client_scores, dtype=np.float64), D=D)

    # ---------------- geometric median, several tensors, more iterations ----------------
    n, d, seg = 15, 1000, [512, 100, 388]
    X = make_rows(rng, n, d, n_byz=3, kind="replace")
    df = ref.GeometricMedianDefense({"max_iters": 100, "tolerance": 1e-6})
    res = df.aggregate(to_lists(X, seg), ones(n))
    put("geomed_p3", {"max_iters": 100, "tolerance": 1e-6}, X, seg, out=flat(res), num_iters=df.num_iters)

    # ---------------- model-poisoning attacks (src/attacks/model_poisoning.py) ----------------
    att = load_reference_attacks()
    n, d, seg = 6, 900, [400, 37, 463]
    X = make_rows(rng, n, d)
    X[0, :6] = [0.0, -0.0, np.inf, -np.inf, 1e-40, -3e38]
    X.view(U32)[0, 6] = 0x7FC00000
    glob = rng.normal(0, 0.05, d).astype(np.float32)
    direction = rng.normal(0, 1.0, d).astype(np.float32)
    G = to_lists(glob[None, :], seg)[0]
    L = to_lists(X, seg)
    with _w.catch_warnings(), np.errstate(all="ignore"):
        _w.simplefilter("ignore")
        put("attack_replace", {"scale_factor": 10.0, "num_malicious": 2}, X, seg, glob=glob, num_clients=20,
            out=flat(att.ModelReplacementAttack({"scale_factor": 10.0, "num_malicious": 2}).poison_update(L[0], G, 20)))
        put("attack_replace_boost", {"boost_factor": 3.7}, X, seg, glob=glob, num_clients=20,
            out=flat(att.ModelReplacementAttack({"boost_factor": 3.7}).poison_update(L[0], G, 20)))
        put("attack_scaling", {"scale": 100.0}, X, seg, glob=glob,
            out=flat(att.ScalingAttack({"scale": 100.0}).poison_update(L[0], G, 20)))
        put("attack_ipm_negate", {"epsilon": 0.1}, X, seg, glob=glob,
            out=flat(att.InnerProductManipulationAttack({"epsilon": 0.1}).poison_update(L[0], G, 20)))
        put("attack_ipm_sign", {"epsilon": 0.1}, X, seg, glob=glob,
            out=flat(att.InnerProductManipulationAttack({"epsilon": 0.1}).poison_update(L[1], G, 20, benign_mean=L[0])))
        ak = att.AdaptiveKrumAttack({"perturbation_scale": 0.05, "target_direction": to_lists(direction[None, :], seg)[0]})
        res = ak.poison_update(L[1], G, 20, benign_updates=L[1:])
        put("attack_adaptive_krum", {"perturbation_scale": 0.05}, X, seg, glob=glob, direction=direction,
            out=flat(res), center=flat(ak.estimated_benign_center))

    np.savez_compressed(os.path.join(HERE, "golden.npz"), **arrays)
    with open(os.path.join(HERE, "golden.json"), "w") as fh:
        json.dump(dict(torch=torch.__version__, numpy=np.__version__, cases=cases), fh, indent=1, sort_keys=True)
    print(f"wrote {len(cases)} cases, {sum(a.nbytes for a in arrays.values()) / 1e6:.2f} MB raw")


if __name__ == "__main__":
    main()
