"""Detector decisions vs the unmodified reference module + golden values."""
import json

import numpy as np
import pytest
import torch

from trustworthy_dl import AttackDetector, AttackType
from trustworthy_dl.ops.reference_impl import STAT_NAMES, tensor_statistics


def test_twelve_statistics_golden():
    s = dict(zip(STAT_NAMES, tensor_statistics(torch.tensor([1., 2., 4., 8., 16.])).tolist()))
    assert s["mean"] == pytest.approx(6.2) and s["std"] == pytest.approx(5.4553, abs=1e-4)
    assert s["median"] == 4 and s["skewness"] == pytest.approx(0.889, abs=1e-3)
    assert s["kurtosis"] == pytest.approx(-0.6741, abs=1e-4)
    assert s["percentile_25"] == 2.0 and s["percentile_75"] == 8.0
    assert s["norm_l1"] == 31 and s["norm_l2"] == pytest.approx(18.4662, abs=1e-4) and s["norm_inf"] == 16


def test_statistics_match_reference(ref_detector):
    g = torch.Generator().manual_seed(3)
    x = torch.randn(7, 33, 5, generator=g) * 2.5 + 0.3
    ref = ref_detector.AttackDetector()._calculate_tensor_statistics(x.numpy().flatten())
    ours = dict(zip(STAT_NAMES, tensor_statistics(x).tolist()))
    for k in STAT_NAMES:
        assert ours[k] == pytest.approx(ref[k], rel=1e-5, abs=1e-6), k


@pytest.mark.parametrize("poison", ["x1.5", "x2", "x3", "x10", "x100", "flip", "noise"])
def test_decisions_match_reference(ref_detector, poison):
    """30 benign steps then one poisoned step; decisions must be identical at every step."""
    torch.manual_seed(0)
    ours, ref = AttackDetector(), ref_detector.AttackDetector()
    for step in range(30):
        x = torch.randn(4, 64, 128)
        a, b = ours.detect_output_anomaly(x, 0, step), ref.detect_output_anomaly(x, 0, step)
        assert a == b and a is False
    x = torch.randn(4, 64, 128)
    if poison.startswith("x"):
        x = x * float(poison[1:])
    elif poison == "flip":
        x = -x
    else:
        x = x + 0.5 * torch.randn_like(x)
    a, b = ours.detect_output_anomaly(x, 0, 30), ref.detect_output_anomaly(x, 0, 30)
    assert a == b
    assert a == (poison != "flip")                                  # sign flip is invisible to these stats
    if a and poison.startswith("x"):
        assert ours.last_result[0].attack_type is AttackType.GRADIENT_POISONING
        assert ours.last_result[0].confidence > 0.5


@pytest.mark.parametrize("poison", ["x1.5", "x2", "x3", "x5", "x10", "x100", "flip", "noise", "shift", "x0.2"])
def test_injected_poisoning_same_verdict_with_and_without_evidence_rule(ref_detector, poison):
    """The trainer's default (``detector_require_evidence=True``, an extension) must not change any verdict on
    injected poisoning: every scenario gets the verdict of the unmodified reference rule, step by step."""
    torch.manual_seed(0)
    strict, plain, ref = AttackDetector(require_evidence=True), AttackDetector(require_evidence=False), ref_detector.AttackDetector()
    for step in range(30):
        x = torch.randn(4, 64, 128)
        assert strict.detect_output_anomaly(x, 0, step) == plain.detect_output_anomaly(x, 0, step) \
            == ref.detect_output_anomaly(x, 0, step)
    x = torch.randn(4, 64, 128)
    if poison.startswith("x"):
        x = x * float(poison[1:])
    elif poison == "flip":
        x = -x
    elif poison == "shift":
        x = x + 0.75
    else:
        x = x + 0.5 * torch.randn_like(x)
    a, b, c = (d.detect_output_anomaly(x, 0, 30) for d in (strict, plain, ref))
    assert a == b == c
    if a:
        assert strict.last_result[0].attack_type == plain.last_result[0].attack_type


def test_first_nine_calls_never_flag_and_none_is_attack():
    d = AttackDetector()
    assert d.detect_output_anomaly(None, 0, 0) is True
    for i in range(9):
        assert d.detect_output_anomaly(torch.randn(100) * (100 if i == 8 else 1), 1, i) is False


def test_gradient_poisoning_mixed_shapes_no_crash():
    d = AttackDetector()
    torch.manual_seed(1)
    for s in range(20):
        assert d.detect_gradient_poisoning([torch.randn(8, 8), torch.randn(8)], 0, s) is False
    assert d.detect_gradient_poisoning([torch.randn(8, 8) * 50, torch.randn(8) * 50], 0, 20) is True


def test_byzantine_and_backdoor(ref_detector):
    torch.manual_seed(0)
    base = torch.randn(256)
    outs = {i: base + 0.01 * torch.randn(256) for i in range(4)}
    outs[2] = -outs[2]
    ours, ref = AttackDetector(), ref_detector.AttackDetector()
    assert ours.detect_byzantine_behavior(outs, 0) == ref.detect_byzantine_behavior(outs, 0) == [0, 1, 2, 3]  # Q16
    assert AttackDetector(byzantine_mode="median").detect_byzantine_behavior(outs, 0) == [2]
    assert ours.detect_byzantine_behavior({0: base, 1: base}, 0) == []
    x = torch.randn(8, 10)
    assert ours.detect_backdoor_attack(x, x, 0) is False
    assert ours.detect_backdoor_attack(10 * x, -10 * x, 0) is True


def test_ml_models_and_exports(tmp_path):
    torch.manual_seed(0)
    d = AttackDetector()
    for s in range(60):
        d.detect_output_anomaly(torch.randn(2000), 0, s)
    d.update_detection_models()
    assert 0 in d.anomaly_detectors and 0 in d.clustering_models
    assert d.detect_with_ml_models(d._calculate_tensor_statistics(torch.randn(2000) * 10), 0) is True
    p = tmp_path / "det.json"
    d.export_detection_data(str(p))
    data = json.loads(p.read_text())
    assert set(data) == {"detection_stats", "baselines", "history_lengths"} and data["history_lengths"]["0"] == 60
    st = d.get_detection_statistics()
    assert st["nodes_monitored"] == 1 and st["average_history_length"] == 60
    d2 = AttackDetector()
    d2.load_state_dict(d.state_dict())
    x = torch.randn(2000) * 3
    assert d2.detect_output_anomaly(x, 0, 61) == d.detect_output_anomaly(x, 0, 61) is True
    d.reset_node_history(0)
    assert len(d.output_history[0]) == 0
    d.cleanup()


def test_history_ring_is_bounded():
    d = AttackDetector(history_size=16)
    for s in range(40):
        d.detect_output_anomaly(torch.randn(64), 0, s)
    assert len(d.output_history[0]) == 16
    assert d.output_history[0].steps_view().tolist() == list(range(24, 40))


def test_reference_private_helpers_exist_and_agree():
    """The reference's private decision helpers (attack_detector.py:241-363) are available under their names."""
    import torch
    from trustworthy_dl.security.attack_detection import AttackDetector
    d = AttackDetector()
    torch.manual_seed(0)
    for i in range(12):
        d.detect_output_anomaly(torch.randn(2000), 0, i)
    d._update_output_baseline(0)
    base = d.output_baselines[0]
    assert set(base["mean"].keys()) >= {"mean", "std", "min", "max", "percentile_5", "percentile_95"}
    calm = d._detect_statistical_anomaly(d._calculate_tensor_statistics(torch.randn(2000)), base, 0)
    assert not calm.is_attack
    loud = d._detect_statistical_anomaly(d._calculate_tensor_statistics(torch.randn(2000) * 40), base, 0)
    assert loud.is_attack and loud.attack_type == d._classify_attack_type(loud.evidence)
    assert d._detect_gradient_anomaly(d._calculate_tensor_statistics(torch.randn(2000) * 40), base).is_attack
    assert d._classify_attack_type({}) is None
