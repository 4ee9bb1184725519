"""The oracle (oracle/) against the fixtures the REFERENCE produced (tests/golden/).

This is what pins the oracle: every restated function must reproduce the
reference's own outputs on the same inputs -- bit-for-bit wherever the reference's
arithmetic is order-deterministic, and to 1e-6 where only an FP64 restatement is
compared.  CPU only.
"""
import numpy as np
import pytest
import torch

from conftest import bits, golden_names
from oracle import oracle64, ref_port


def T(x):
    return torch.from_numpy(np.ascontiguousarray(x))


@pytest.mark.parametrize("name", golden_names("median"))
def test_median_port_and_independent_restatement(golden, name):
    c = golden.case(name)
    assert np.array_equal(bits(ref_port.median(T(c["X"])).numpy()), bits(c["out"]))
    assert np.array_equal(bits(oracle64.median_stable(c["X"])), bits(c["out"]))


@pytest.mark.parametrize("name", golden_names("trimmed"))
def test_trimmed_mean(golden, name):
    c = golden.case(name)
    out, k, fell_back = ref_port.trimmed_mean(T(c["X"]), c["cfg"]["trim_ratio"])
    assert k == int(c["k"])
    got, ref = out.numpy(), c["out"]
    assert np.array_equal(np.isnan(got), np.isnan(ref))
    assert np.array_equal(bits(got)[~np.isnan(ref)], bits(ref)[~np.isnan(ref)])
    n = c["X"].shape[0]
    if fell_back:
        assert np.array_equal(bits(oracle64.median_stable(c["X"])), bits(ref))
    else:
        with np.errstate(invalid="ignore"):
            t64 = oracle64.trimmed_mean64(c["X"], k)
        fin = np.isfinite(ref) & np.isfinite(t64)
        scale = oracle64.trimmed_scale(c["X"], k)
        assert np.all(np.abs(t64[fin] - ref[fin]) <= 1e-6 * (np.abs(ref[fin]) + scale[fin]))
        assert n - 2 * k >= 1


def test_trim_count_double_rounding():
    assert ref_port.trim_count(100, 0.29) == 28      # int(100*0.29) == 28, not 29
    assert ref_port.trim_count(5, 0.1) == 1


@pytest.mark.parametrize("name", golden_names("krum"))
def test_krum(golden, name):
    c = golden.case(name)
    r = ref_port.krum(T(c["X"]), c["cfg"]["num_malicious"], c["cfg"]["multi_k"])
    assert np.array_equal(r["D"], c["D"], equal_nan=True)       # NaN / inf rows: the krum_nan_* / krum_inf* fixtures
    assert np.array_equal(np.array(r["scores"]), c["scores"], equal_nan=True)
    assert r["selected"] == c["selected"].tolist()
    assert r["rejected"] == c["rejected"].tolist()
    assert np.array_equal(bits(r["out"].numpy()), bits(c["out"]))
    # FP64 truth: the reference's fp32 norms are accurate at this d
    with np.errstate(all="ignore"):
        D64 = oracle64.distances64(c["X"])
        assert np.allclose(D64, c["D"], rtol=2e-6, atol=0, equal_nan=True)
        s64 = oracle64.krum_scores64(D64, c["cfg"]["num_malicious"])
        assert np.allclose(s64, c["scores"], rtol=2e-6, equal_nan=True)
    # Multi-Krum mean == strict left-to-right fp32 sum in score order
    if c["cfg"]["multi_k"] > 1:
        seq = oracle64.combine_seq32(c["X"], None, None, c["cfg"]["multi_k"], rows=c["selected"].tolist())
        assert np.array_equal(bits(seq), bits(c["out"]))


def test_krum_too_few_clients():
    with pytest.raises(ValueError):
        ref_port.krum(torch.zeros(3, 8), 2, 1)


def test_fedavg(golden):
    c = golden.case("fedavg")
    ne = c["num_examples"].tolist()
    assert np.array_equal(bits(ref_port.fedavg(T(c["X"]), ne).numpy()), bits(c["out"]))
    w = np.array(ne, dtype=np.float64).astype(np.float32)
    seq = oracle64.combine_seq32(c["X"], w, None, np.float32(np.float64(sum(ne))))
    assert np.array_equal(bits(seq), bits(c["out"]))


@pytest.mark.parametrize("name", golden_names("fltrust"))
def test_fltrust(golden, name):
    c = golden.case(name)
    r = ref_port.fltrust(T(c["X"]), T(c["g0"]))
    assert np.array_equal(np.array(r["trust"]), c["trust"])
    assert np.array_equal(bits(r["out"].numpy()), bits(c["out"]))
    # closed form in FP64: ts_i = relu(cos(x_i, g0))
    ss, dot = oracle64.row_sumsq64(c["X"]), oracle64.row_dot64(c["X"], c["g0"])
    gg = float(np.dot(c["g0"].astype(np.float64), c["g0"].astype(np.float64)))
    with np.errstate(invalid="ignore", divide="ignore"):
        cos = np.where(ss > 0, dot / np.sqrt(ss * gg), 0.0)
    assert np.allclose(np.maximum(cos, 0), c["trust"], rtol=0, atol=2e-6)


@pytest.mark.parametrize("ctype", ["l2", "linf"])
def test_gradient_clipping(golden, ctype):
    c = golden.case(f"gradclip_{ctype}")
    r = ref_port.gradient_clipping(T(c["X"]), c["num_examples"].tolist(), c["cfg"]["clip_norm"], ctype)
    assert r["norms"] == c["norms"].tolist()
    assert r["clipped_count"] == int(c["clipped_count"])
    assert np.array_equal(bits(r["out"].numpy()), bits(c["out"]))


def test_dpsgd_pre_noise(golden):
    c = golden.case("dpsgd_nonoise")
    r = ref_port.dp_sgd_clip_mean(T(c["X"]), c["num_examples"].tolist(), c["cfg"]["clip_norm"])
    # sigma = 0 adds randn*0 = +-0 to every element: values equal, sign of zero aside
    assert np.array_equal(r["out"].numpy(), c["out"])


@pytest.mark.parametrize("name", golden_names("normbound"))
def test_norm_bounding(golden, name):
    c = golden.case(name)
    r = ref_port.norm_bounding(T(c["X"]), c["num_examples"].tolist(), c["cfg"]["max_norm"], c["cfg"]["min_norm"])
    assert r["rejected"] == c["rejected"].tolist()
    assert np.array_equal(bits(r["out"].numpy()), bits(c["out"]))


@pytest.mark.parametrize("name", golden_names("geomed"))
def test_geometric_median(golden, name):
    c = golden.case(name)
    r = ref_port.geometric_median(T(c["X"]), c["cfg"]["max_iters"], c["cfg"].get("tolerance", 1e-5))
    assert r["num_iters"] == int(c["num_iters"])
    assert np.array_equal(bits(r["out"].numpy()), bits(c["out"]))


def test_attacks(golden):
    """The reference's model-poisoning attacks (src/attacks/model_poisoning.py) restated: bit-exact
    against fixtures produced by its own classes, incl. +-0 / Inf / NaN / denormal elements."""
    c = golden.case("attack_replace")
    X, g = T(c["X"]), T(c["glob"])
    with np.errstate(all="ignore"):
        out = ref_port.attack_model_replacement(X[0], g, int(c["num_clients"]), 10.0, 2)
        assert np.array_equal(bits(out.numpy()), bits(c["out"]))
        c = golden.case("attack_replace_boost")
        out = ref_port.attack_model_replacement(X[0], g, 20, boost_factor=3.7)
        assert np.array_equal(bits(out.numpy()), bits(c["out"]))
        assert np.array_equal(bits(ref_port.attack_scaling(X[0], 100.0).numpy()), bits(golden.case("attack_scaling")["out"]))
        assert np.array_equal(bits(ref_port.attack_ipm(X[0], 0.1).numpy()), bits(golden.case("attack_ipm_negate")["out"]))
        assert np.array_equal(bits(ref_port.attack_ipm(X[1], 0.1, benign_mean=X[0]).numpy()),
                              bits(golden.case("attack_ipm_sign")["out"]))
        c = golden.case("attack_adaptive_krum")
        out, center = ref_port.attack_adaptive_krum(X[1:], T(c["direction"]), 0.05)
        assert np.array_equal(bits(center.numpy()), bits(c["center"]))
        assert np.array_equal(bits(out.numpy()), bits(c["out"]))


def test_median_stable_matches_torch_on_random_ties():
    rng = np.random.RandomState(7)
    vals = np.array([0.0, -0.0, 1.0, -1.0, 2.0, np.inf, -np.inf], dtype=np.float32)
    for n in list(range(1, 40)) + [63, 64, 65, 100, 127, 128, 129, 200]:
        X = rng.choice(vals, size=(n, 64))
        assert np.array_equal(bits(oracle64.median_stable(X)), bits(torch.median(T(X), dim=0)[0].numpy())), n
