"""The assertions of the reference's own tests/test_defenses.py (17 tests, CPU tensors in,
structural / qualitative checks) restated against the drop-in classes.  The reference file
itself cannot travel to the GPU box; in the build container it can also be run unmodified
with PYTHONPATH=multimodal-fl-security_b200 on a machine that has a GPU."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def mock_updates(num_clients=5, shapes=((10, 10), (10,))):
    return [[torch.randn(s) for s in shapes] for _ in range(num_clients)]


def test_krum_single_multi_and_outlier():
    from src.defenses.krum import KrumDefense, MultiKrumDefense
    ups = mock_updates(5)
    d = KrumDefense({'num_malicious': 1, 'multi_k': 1})
    res = d.aggregate(ups, [100] * 5)
    assert len(res) == len(ups[0]) and len(d.selected_clients) == 1
    d = KrumDefense({'num_malicious': 1, 'multi_k': 2})
    res = d.aggregate(ups, [100] * 5)
    assert len(res) == len(ups[0]) and len(d.selected_clients) == 2
    base = torch.zeros(10, 10)
    ups = [[base + torch.randn_like(base) * 0.01] for _ in range(4)] + [[torch.randn(10, 10) * 100]]
    d = KrumDefense({'num_malicious': 1, 'multi_k': 1})
    d.aggregate(ups, [100] * 5)
    assert 4 not in d.selected_clients
    with pytest.raises(ValueError):
        KrumDefense({'num_malicious': 2}).aggregate(mock_updates(3), [100] * 3)
    cfg = {'num_malicious': 1}
    assert MultiKrumDefense(cfg).multi_k == 3 and cfg['multi_k'] == 3


def test_trimmed_mean_and_median():
    from src.defenses.trimmed_mean import TrimmedMeanDefense, MedianDefense, GeometricMedianDefense
    d = TrimmedMeanDefense({'trim_ratio': 0.1})
    assert d.trim_ratio == 0.1
    ups = mock_updates(5)
    assert len(TrimmedMeanDefense({'trim_ratio': 0.2}).aggregate(ups, [100] * 5)) == len(ups[0])
    base = torch.zeros(10)
    ups = [[base + 0.1], [base + 0.2], [base], [base + 100], [base - 100]]
    res = TrimmedMeanDefense({'trim_ratio': 0.2}).aggregate(ups, [100] * 5)
    assert res[0].abs().mean() < 1.0
    ups = [[torch.tensor([1.0, 1.0])], [torch.tensor([1.1, 0.9])], [torch.tensor([0.9, 1.1])],
           [torch.tensor([1.0, 1.0])], [torch.tensor([100.0, 100.0])]]
    res = MedianDefense({}).aggregate(ups, [100] * 5)
    assert torch.allclose(res[0], torch.tensor([1.0, 1.0]), atol=0.2)
    assert len(MedianDefense().aggregate(mock_updates(5), [100] * 5)) == 2
    g = GeometricMedianDefense({'max_iters': 10})
    g.aggregate(mock_updates(5), [100] * 5)
    assert g.num_iters > 0


def test_dp_and_clipping():
    from src.defenses.differential_privacy import DPSGDDefense, GradientClippingDefense
    d = DPSGDDefense({'clip_norm': 1.0, 'noise_multiplier': 0.1})
    clipped, norm = d.clip_gradient([torch.randn(10, 10) * 10])
    assert torch.norm(torch.cat([c.flatten() for c in clipped])) <= 1.0 + 1e-6
    upd = [torch.zeros(10, 10)]
    assert not torch.allclose(d.add_noise(upd, 5)[0], upd[0])
    res = d.aggregate(mock_updates(5), [100] * 5)
    assert len(res) == 2 and d.rounds_completed == 1
    before = d.get_privacy_spent()
    d.aggregate(mock_updates(5), [100] * 5)
    assert d.get_privacy_spent() > before
    ups = mock_updates(5)
    ups[0] = [u * 100 for u in ups[0]]
    g = GradientClippingDefense({'clip_norm': 1.0})
    assert len(g.aggregate(ups, [100] * 5)) == 2 and g.clipped_count >= 1
