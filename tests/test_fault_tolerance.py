"""Heartbeats / dashboard and node replacement with replica recovery (stubs in the reference:
src/system/heartbeat_info.*, dashboard.*, executor.cc:112-138, ManageNode::REPLACE)."""
import numpy as np

from parameter_server_b200.runtime import host


def test_heartbeat_reports_reach_the_dashboard():
    table, dead = host.run_heartbeat_scenario(interval_s=1, run_ms=2600, dead_timeout_s=60)
    for node in ("S0", "W0", "W1"):
        assert node in table, table
    assert dead == 0
    # with a zero timeout every node that has reported is overdue
    _, dead0 = host.run_heartbeat_scenario(interval_s=1, run_ms=2600, dead_timeout_s=0)
    assert dead0 >= 3


def test_dead_server_is_replaced_and_recovers_from_replica():
    n = 40
    vals, (lo, hi) = host.run_replace_scenario(n)
    k = np.arange(n)
    on_s1 = (k >= lo) & (k < hi)
    assert on_s1.any() and (~on_s1).any()
    # surviving shard: initial value + the push; recovered shard: what the replica saw pushed
    np.testing.assert_allclose(vals[~on_s1], 0.1 * k[~on_s1] + 1.0, rtol=1e-6)
    np.testing.assert_allclose(vals[on_s1], 1.0, rtol=1e-6)
