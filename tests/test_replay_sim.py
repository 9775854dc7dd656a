"""Unit tests of the sink model behind BASELINE configs[4] (scripts/replay_sim.py): token-bucket
arithmetic, worker pool, retry back-off — with the reference's defaults
(.helm/values.yaml:141-161; services/supervisor.go:41-47,107-117)."""
import importlib.util
import os

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("replay_sim", os.path.join(ROOT, "scripts", "replay_sim.py"))
sim = importlib.util.module_from_spec(spec)
spec.loader.exec_module(sim)


def test_token_bucket_burst_then_rate():
    b = sim.TokenBucket(rate=10, burst=100)
    times = [b.take(0.0) for _ in range(103)]
    assert times[:100] == [0.0] * 100  # the burst goes at once
    assert times[100] == pytest.approx(0.1) and times[101] == pytest.approx(0.2) and times[102] == pytest.approx(0.3)
    # an idle second refills ten tokens, never more than the burst
    b = sim.TokenBucket(rate=10, burst=5)
    for _ in range(5):
        b.take(0.0)
    assert b.take(0.0) == pytest.approx(0.1)
    assert [b.take(10.0) for _ in range(6)][:5] == [10.0] * 5
    assert b.take(10.0) == pytest.approx(10.2)


def test_backoff_doubles_up_to_the_cap():
    d = [sim.backoff_delay(a, 0.1, 1.0) for a in range(1, 7)]
    assert d == pytest.approx([0.1, 0.2, 0.4, 0.8, 1.0, 1.0])


def test_reference_defaults_are_rate_bound():
    # 1000 elements, 10/s, burst 100, 2 workers, 1 ms each: 100 at once, then one every 100 ms
    span, calls, dropped, _ = sim.simulate([False] * 1000, rate=10, burst=100, workers=2, latency_s=1e-3)
    assert calls == 1000 and dropped == 0
    assert span == pytest.approx(90.0 + 1e-3, rel=1e-3)


def test_worker_bound_when_the_limit_is_lifted():
    span, calls, _, _ = sim.simulate([False] * 1000, rate=float("inf"), burst=1, workers=2, latency_s=0.01)
    assert calls == 1000 and span == pytest.approx(5.0, rel=1e-6)  # 2 workers x 100 elements/s each


def test_failing_elements_are_retried_with_backoff_then_dropped():
    span, calls, dropped, first = sim.simulate([True], rate=float("inf"), burst=1, workers=1, latency_s=0.0, base_delay=0.1,
                                        max_delay=1.0, max_retries=3)
    assert calls == 4 and dropped == 1  # the first try and three retries
    assert span == pytest.approx(0.1 + 0.2 + 0.4) and first == 0.0


def test_stream_comes_from_the_pipeline():
    st = sim.pipeline_stream(200_000)
    assert st["job_rows"] > 80_000 and 0.5 < st["upserts"] / st["job_rows"] < 0.7
    assert st["elements"] >= st["upserts"] and st["err_rows"] > 0 and st["encoded_bytes"] > 0
    assert len(st["fails"]) == st["elements"] and int(st["fails"].sum()) == st["err_rows"]
