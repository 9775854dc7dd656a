"""The reference's TestSupervisor (services/supervisor_test.go:542-580), replayed against the C++ host
mirror of the service interface (nexus-supervisor_b200/host) on top of the CUDA library.  The scenarios
come from tests/golden/supervisor_test_scenarios.json (extracted from the reference test)."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "nexus-supervisor_b200", "lib", "supervisor_test")
STAGE_CONST = {"LifecycleStageSchedulingFailed": "SCHEDULING_FAILED", "LifecycleStageDeadlineExceeded":
               "DEADLINE_EXCEEDED", "LifecycleStageRunning": "RUNNING", "LifecycleStageFailed": "FAILED",
               "LifecycleStageCancelled": "CANCELLED"}


def _scenarios_tsv(tmp_path):
    golden = json.load(open(os.path.join(ROOT, "tests", "golden", "supervisor_test_scenarios.json")))
    path = tmp_path / "scenarios.tsv"
    with open(path, "w") as f:
        for s in golden["scenarios"]:
            f.write("\t".join([s["kind"], s["reason"], s["involved_name"], s["request_id"], s["algorithm"],
                               s["seeded_stage"], STAGE_CONST[s["expected_stage_const"]],
                               ",".join(s["objects_present"]), s["message"]]) + "\n")
    return str(path)


def _build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "nexus-supervisor_b200", "host")])


@pytest.mark.gpu
def test_reference_scenarios_through_cpp_host_mirror(tmp_path):
    _build()
    out = subprocess.run([BIN, _scenarios_tsv(tmp_path)], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "ok: TestSupervisor scenarios and error paths pass" in out.stdout


def test_cpp_host_mirror_has_no_cpu_fallback(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    _build()
    out = subprocess.run([BIN, _scenarios_tsv(tmp_path)], capture_output=True, text=True, timeout=60)
    assert out.returncode != 0 and "no CUDA device" in out.stdout
