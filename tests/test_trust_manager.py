"""Trust layer vs the unmodified reference module (oracle) + golden values of SURVEY Appendix A."""
import json

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from trustworthy_dl import NodeStatus, TrustManager
from trustworthy_dl.core.trust_manager import LogicalClock, STATUS_CODES


def test_healthy_trajectory_golden():
    tm = TrustManager(4)
    vals = [tm.update_trust_score(0, 0.0, 1.0) for _ in range(5)]
    assert np.allclose(vals, [0.99, 0.981, 0.9729, 0.96561, 0.959049], atol=1e-6)
    assert tm.get_node_status(0) is NodeStatus.TRUSTED
    for _ in range(400):
        tm.update_trust_score(0, 0.0, 1.0)
    assert abs(tm.get_trust_score(0) - 0.9) < 1e-6


def test_compromise_and_recovery_golden():
    tm = TrustManager(2)
    tm.mark_compromised(1, "gradient_poisoning")
    assert tm.get_trust_score(1) == pytest.approx(0.1)
    assert tm.get_node_status(1) is NodeStatus.COMPROMISED
    assert tm.attack_history[1][0]["previous_trust"] == pytest.approx(1.0)     # Q10 fixed
    statuses = []
    for _ in range(14):
        tm.update_trust_score(1, 0.0, 1.0)
        statuses.append(tm.get_node_status(1))
    assert statuses[0] is NodeStatus.COMPROMISED and statuses[1] is NodeStatus.COMPROMISED
    assert statuses[2] is NodeStatus.SUSPICIOUS
    assert statuses[12] is NodeStatus.SUSPICIOUS and statuses[13] is NodeStatus.TRUSTED


def test_always_bad_never_compromised_by_scoring():
    tm = TrustManager(1)
    st_ = [(tm.update_trust_score(0, 1.0, 0.0), tm.get_node_status(0)) for _ in range(200)]
    assert st_[4][1] is NodeStatus.TRUSTED and st_[5][1] is NodeStatus.SUSPICIOUS
    assert st_[5][0] == pytest.approx(0.672, abs=1e-3)
    assert all(s is not NodeStatus.COMPROMISED for _, s in st_)


def test_matches_reference_step_by_step(ref_trust, monkeypatch):
    """Drive both implementations with the same random behaviour under a frozen clock."""
    fake = {"t": 1000.0}
    monkeypatch.setattr(ref_trust.time, "time", lambda: fake["t"])
    rng = np.random.default_rng(0)
    ours = TrustManager(6, trust_threshold=0.7, clock=lambda: fake["t"])
    ref = ref_trust.TrustManager(6, trust_threshold=0.7)
    for step in range(300):
        fake["t"] += float(rng.uniform(0, 0.5))
        node = int(rng.integers(0, 6))
        if rng.random() < 0.03:
            ours.mark_compromised(node, "x"); ref.mark_compromised(node, "x")
        # initiate_recovery is exercised separately: recovery_rate is live here (Q9), dead in the oracle
        dev, gc = float(rng.uniform(0, 1.5)), float(rng.uniform(0, 1))
        extra = {"error_rate": float(rng.uniform(0, 0.3))} if rng.random() < 0.3 else {}
        ours.update_trust_score(node, dev, gc, **extra)
        ref.update_trust_score(node, dev, gc, **extra)
        for n in range(6):
            assert ours.get_trust_score(n) == pytest.approx(ref.get_trust_score(n), abs=1e-12)
            assert ours.get_node_status(n).value == ref.get_node_status(n).value
    assert ours.calculate_system_trust() == pytest.approx(ref.calculate_system_trust())
    so, sr = ours.get_trust_statistics(), ref.get_trust_statistics()
    for k in ("mean_trust", "std_trust", "min_trust", "max_trust", "system_trust"):
        assert so[k] == pytest.approx(sr[k])
    assert so["node_status_counts"] == sr["node_status_counts"]
    assert ours.select_best_nodes(3) == ref.select_best_nodes(3)
    assert ours.get_recommendations() == ref.get_recommendations()
    for n in range(6):
        assert ours.predict_node_reliability(n) == pytest.approx(float(ref.predict_node_reliability(n)))
    ours.adaptive_threshold_adjustment(); ref.adaptive_threshold_adjustment()
    assert ours.trust_threshold == pytest.approx(ref.trust_threshold)


def test_queries_and_selection():
    tm = TrustManager(4)
    for i, v in enumerate([0.95, 0.2, 0.6, 0.85]):
        tm._value[i] = v
        tm._update_node_status(i, v)
    assert tm.select_best_nodes(2) == [0, 3]
    assert tm.get_compromised_nodes() == [1] and tm.get_suspicious_nodes() == [2]
    assert tm.get_trust_score(99) == 0.0 and tm.get_node_status(99) is NodeStatus.OFFLINE
    assert tm.calculate_system_trust() == pytest.approx(sum(v * v for v in [0.95, 0.2, 0.6, 0.85]) / 2.6)
    w = tm.aggregation_weights()
    assert w[1] == 0.0 and w.sum() == pytest.approx(1.0)


def test_adaptive_threshold_cases():
    for mean, expect in ((0.65, 0.7), (0.95, 0.8), (0.30, 0.3)):
        tm = TrustManager(2)
        tm._value[:] = mean
        assert tm.adaptive_threshold_adjustment() == pytest.approx(expect)


def test_readme_constructor_and_recovery_rate():
    tm = TrustManager(initial_trust=0.9, decay_rate=0.1, recovery_rate=0.05)     # README.md:71-75
    assert tm.num_nodes == 0
    tm.resize(3)
    assert tm.trust_scores[2].value == 0.9 and tm.trust_scores[2].decay_rate == 0.1
    slow, fast = TrustManager(1), TrustManager(1, recovery_rate=0.05)
    for t in (slow, fast):
        t.mark_compromised(0)
        for _ in range(5):
            t.update_trust_score(0, 0.0, 1.0)
    assert fast.get_trust_score(0) > slow.get_trust_score(0)
    tm2 = TrustManager(1)
    tm2.mark_compromised(0); tm2.initiate_recovery(0)
    assert tm2.get_node_status(0) is NodeStatus.RECOVERING and tm2.trust_scores[0].recovery_rate == 0.02


def test_logical_clock_decay_is_deterministic():
    clk = LogicalClock()
    tm = TrustManager(1, clock=clk)
    clk.advance(10.0)
    v = tm.update_trust_score(0, 0.0, 1.0)
    assert v == pytest.approx(0.9 * np.exp(-0.1) + 0.09)


def test_export_and_state_roundtrip(tmp_path):
    tm = TrustManager(3)
    tm.update_trust_score(0, 0.2, 0.9); tm.mark_compromised(2, "byz")
    p = tmp_path / "trust.json"
    tm.export_trust_data(str(p))
    d = json.loads(p.read_text())
    assert set(d) == {"trust_scores", "node_status", "trust_history", "attack_history", "statistics"}
    assert d["node_status"]["2"] == "compromised"
    tm2 = TrustManager(3)
    tm2.load_state_dict(tm.state_dict())
    assert np.allclose(tm2.trust_vector(), tm.trust_vector()) and (tm2.status_vector() == tm.status_vector()).all()
    assert tm2.update_trust_score(1, 0.1, 1.0) == pytest.approx(tm.update_trust_score(1, 0.1, 1.0))


@settings(max_examples=200, deadline=None)
@given(st.floats(0, 1), st.sampled_from(list(STATUS_CODES.values())), st.floats(0.31, 0.95))
def test_fsm_properties(trust, current, threshold):
    tm = TrustManager(1, trust_threshold=threshold)
    new = tm.next_status_code(current, trust)
    if trust < 0.3:
        assert new == STATUS_CODES[NodeStatus.COMPROMISED]
    elif trust < threshold:
        assert new == STATUS_CODES[NodeStatus.SUSPICIOUS]
    else:
        assert new in (STATUS_CODES[NodeStatus.TRUSTED], STATUS_CODES[NodeStatus.RECOVERING])


@settings(max_examples=100, deadline=None)
@given(st.lists(st.tuples(st.floats(0, 3), st.floats(0, 1)), min_size=1, max_size=60))
def test_trust_stays_clipped(seq):
    tm = TrustManager(1)
    for dev, gc in seq:
        v = tm.update_trust_score(0, dev, gc)
        assert 0.0 <= v <= 1.0


def test_heartbeat_monitor_marks_lagging_and_silent_ranks_offline():
    from trustworthy_dl.core import HeartbeatMonitor, NodeStatus, TrustManager
    tm = TrustManager(num_nodes=4)
    hb = HeartbeatMonitor(4, max_step_lag=2, timeout_s=30.0, on_offline=tm.mark_offline)
    assert hb.observe([1, 1, 1, 1], [10.0, 10.0, 10.1, 10.0]) == []
    # rank 2 stops advancing: 2 steps behind is tolerated, 3 is not
    assert hb.observe([3, 3, 1, 3], [12.0, 12.0, 10.1, 12.0]) == []
    assert hb.observe([4, 4, float("nan"), 4], [13.0, 13.0, float("nan"), 13.0]) == [2]
    assert tm.get_node_status(2) == NodeStatus.OFFLINE
    # rank 0 keeps the step count but its clock is a minute old
    assert hb.observe([5, 5, 5, 5], [14.0 - 61.0, 14.0, 14.0, 14.0]) == [0]
    assert [(e["node_id"], e["event"]) for e in hb.events] == [(2, "offline"), (0, "offline"), (2, "online")]
    hb2 = HeartbeatMonitor(4)
    hb2.load_state_dict(hb.state_dict())
    assert hb2.offline.tolist() == hb.offline.tolist()
