"""GPU parity, part 4: the rows SURVEY.md section 8f adds around the aggregation path --
model-poisoning attacks as device row transforms (8f-3), the round epilogue that writes the
aggregate straight into the global model's parameter tensors (8f-4), and the Weiszfeld loop of
the geometric median with the host out of the iteration.  Fixtures come from the reference's own
classes (tests/golden/make_golden.py)."""
import numpy as np
import pytest
import torch

from conftest import bits, golden_names, same_bits_nan_aware
from test_gpu_golden import flat, lists

pytestmark = pytest.mark.gpu

DEVICES = ["cuda", "cpu"]


def one(c, key, device):
    """A [d] fixture array as the P tensors of one client."""
    return lists(dict(X=np.asarray(c[key])[None, :], seg=c["seg"]), device)[0]


# ------------------------------------------------------------------------------------------
# attacks (src/attacks/model_poisoning.py:48-80, 125-186, 229-236, 266-284)
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("device", DEVICES)
def test_attacks_bit_exact(golden, device):
    from src.attacks import (AdaptiveKrumAttack, InnerProductManipulationAttack, ModelReplacementAttack,
                             ScalingAttack, get_attack)
    c = golden.case("attack_replace")
    L, G = lists(c, device), one(c, "glob", device)
    a = ModelReplacementAttack(dict(c["cfg"]))
    out = a.poison_update(L[0], G, int(c["num_clients"]))
    assert all(o.device == t.device and o.shape == t.shape for o, t in zip(out, L[0]))
    assert same_bits_nan_aware(flat(out), c["out"])
    c = golden.case("attack_replace_boost")
    out = ModelReplacementAttack(dict(c["cfg"])).poison_update(L[0], G, 20)
    assert same_bits_nan_aware(flat(out), c["out"])
    c = golden.case("attack_scaling")
    out = ScalingAttack(dict(c["cfg"])).poison_update(L[0], G, 20)
    assert same_bits_nan_aware(flat(out), c["out"])
    c = golden.case("attack_ipm_negate")
    out = InnerProductManipulationAttack(dict(c["cfg"])).poison_update(L[0], G, 20)
    assert same_bits_nan_aware(flat(out), c["out"])
    c = golden.case("attack_ipm_sign")
    out = InnerProductManipulationAttack(dict(c["cfg"])).poison_update(L[1], G, 20, benign_mean=L[0])
    assert same_bits_nan_aware(flat(out), c["out"])
    c = golden.case("attack_adaptive_krum")
    ak = AdaptiveKrumAttack({"perturbation_scale": c["cfg"]["perturbation_scale"],
                             "target_direction": one(c, "direction", device)})
    out = ak.poison_update(L[1], G, 20, benign_updates=L[1:])
    assert same_bits_nan_aware(flat(ak.estimated_benign_center), c["center"])
    assert same_bits_nan_aware(flat(out), c["out"])
    assert isinstance(get_attack("scaling", {}), ScalingAttack)
    with pytest.raises(ValueError):
        get_attack("no_such_attack", {})
    assert set(a.get_metrics()) == {"attack_type", "scale_factor", "num_malicious"}


def test_attacks_in_place_on_resident_rows(golden):
    """out=local_update: the Byzantine rows of a resident update matrix are rewritten where they sit
    and the defense then reads them (attack -> defense without leaving the GPU)."""
    from src.attacks import ModelReplacementAttack, ScalingAttack
    from src.defenses import KrumDefense
    c = golden.case("attack_replace")
    L, G = lists(c, "cuda"), one(c, "glob", "cuda")
    ptrs = [t.data_ptr() for t in L[0]]
    res = ModelReplacementAttack(dict(c["cfg"])).poison_update(L[0], G, int(c["num_clients"]), out=L[0])
    assert [t.data_ptr() for t in res] == ptrs
    assert same_bits_nan_aware(flat(L[0]), c["out"])
    # a scaled honest row is rejected by Krum on the same resident matrix
    c = golden.case("krum_n20")
    ups = lists(c, "cuda")
    ScalingAttack({"scale": 100.0}).poison_update(ups[7], None, 20, out=ups[7])
    d = KrumDefense({"num_malicious": 4, "multi_k": 10})
    d.aggregate(ups, [1] * len(ups))
    assert 7 in d.rejected_clients and d.rejected_clients[-1] == 7


# ------------------------------------------------------------------------------------------
# round epilogue: 'output_params' (experiments/run_experiments.py:255-259)
# ------------------------------------------------------------------------------------------
def _model_params(upd):
    return [torch.full_like(t, 7.0, dtype=torch.float32) for t in upd[0]]


def _epilogue_pair(cls, cfg, upd, num_examples, **kw):
    plain = cls(dict(cfg)).aggregate(upd, num_examples, **kw)
    params = _model_params(upd)
    d = cls(dict(cfg, output_params=params))
    res = d.aggregate(upd, num_examples, **kw)
    return plain, params, res, d


@pytest.mark.parametrize("which", ["median", "trimmed", "multikrum", "krum1", "fedavg", "gradclip", "normbound", "geomed"])
def test_output_params_epilogue(golden, which):
    """With 'output_params' the producing kernel writes the model's own tensors; the values are those
    of the plain call bit for bit and aggregate() returns the caller's tensors, so the simulator's
    per-tensor ``param.copy_(new_val)`` loop becomes a self-copy."""
    from src.defenses import (GeometricMedianDefense, GradientClippingDefense, KrumDefense, MedianDefense, NoDefense,
                              NormBoundingDefense, TrimmedMeanDefense)
    table = {
        "median": (MedianDefense, "median_n10", {}),
        "trimmed": (TrimmedMeanDefense, "trimmed_n20_r0.1", None),
        "multikrum": (KrumDefense, "krum_c1like", None),
        "krum1": (KrumDefense, "krum_single", None),
        "fedavg": (NoDefense, "fedavg", {}),
        "gradclip": (GradientClippingDefense, "gradclip_l2", None),
        "normbound": (NormBoundingDefense, "normbound_some", None),
        "geomed": (GeometricMedianDefense, "geomed_p3", None),
    }
    cls, name, cfg = table[which]
    c = golden.case(name)
    cfg = dict(c["cfg"]) if cfg is None else cfg
    upd = lists(c, "cuda")
    ne = c["num_examples"].tolist() if "num_examples" in c["fields"] else [100] * len(upd)
    plain, params, res, d = _epilogue_pair(cls, cfg, upd, ne)
    if which == "krum1":
        assert res is upd[d.selected_clients[0]]            # the reference's return value is kept
    else:
        assert all(r is p for r, p in zip(res, params))
    assert np.array_equal(bits(flat(params)), bits(flat(plain)))
    with pytest.raises(ValueError):
        cls(dict(cfg, output_params=params[:-1] if len(params) > 1 else params + params)).aggregate(upd, ne)
    with pytest.raises(ValueError):
        cls(dict(cfg, output_params=params)).aggregate(lists(c, "cpu"), ne)


def test_output_params_fltrust(golden):
    from test_gpu_golden import _fltrust
    from src.defenses import FLTrustDefense

    class Fixed(FLTrustDefense):
        def _compute_server_gradient(self, global_params):
            self.server_gradient = global_params
            return global_params

    for name in ("fltrust_mixed", "fltrust_allneg"):
        c = golden.case(name)
        _, plain, upd = _fltrust(c, "cuda")
        params = _model_params(upd)
        g0 = one(c, "g0", "cuda")
        res = Fixed({"output_params": params}).aggregate(upd, [100] * len(upd), global_params=g0)
        assert all(r is p for r, p in zip(res, params))
        assert np.array_equal(bits(flat(params)), bits(flat(plain))), name


# ------------------------------------------------------------------------------------------
# geometric median: device-side Weiszfeld loop (trimmed_mean.py:240-262)
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("name", golden_names("geomed"))
def test_geometric_median_iterations_and_value(golden, name, device):
    """Same number of Weiszfeld iterations as the reference and the result within 1e-5 (north_star's
    tolerance for FP32 reductions): per coordinate against |ref| + the cohort's spread, and norm-wise."""
    from src.defenses import GeometricMedianDefense
    c = golden.case(name)
    upd = lists(c, device)
    d = GeometricMedianDefense(dict(c["cfg"]))
    res = d.aggregate(upd, [100] * len(upd))
    assert d.num_iters == int(c["num_iters"])
    got, ref = flat(res).astype(np.float64), c["out"].astype(np.float64)
    assert np.linalg.norm(got - ref) <= 1e-5 * np.linalg.norm(ref)
    spread = np.abs(c["X"] - c["out"][None, :]).mean(axis=0)
    assert np.all(np.abs(got - ref) <= 1e-5 * (np.abs(ref) + spread))
    assert d.get_metrics()["num_iters"] == d.num_iters
