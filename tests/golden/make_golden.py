#!/usr/bin/env python3
"""Extracts the reference's own known-answer scenarios into a JSON fixture.

Reads (in the build container only — /root/reference does not exist on the GPU box)
  /root/reference/services/supervisor_test.go   the 8 scenarios of TestSupervisor (:46-580)
  /root/reference/test-resources/checkpoints.cql the seeded checkpoint rows (:31-101)
and writes tests/golden/supervisor_test_scenarios.json: for every event the test
feeds, its involved-object kind, reason, message, the request id, the seeded
lifecycle_stage and the lifecycle stage constant the test asserts afterwards.
Nothing is typed by hand; rerun to regenerate:  python tests/golden/make_golden.py
"""
import json
import os
import re
import sys

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def seeded_stages(cql):
    rows = {}
    for m in re.finditer(r"VALUES \('([^']+)',\s*'([^']+)',\s*'([^']+)'", cql):
        algorithm, rid, stage = m.groups()
        rows[rid] = {"algorithm": algorithm, "lifecycle_stage": stage}
    return rows


def split_funcs(src):
    funcs = {}
    for m in re.finditer(r"^func (\w+)\((.*?)\)[^\n]*\{\n(.*?)^\}\n", src, re.S | re.M):
        funcs[m.group(1)] = {"params": m.group(2), "body": m.group(3),
                             "line": src[:m.start()].count("\n") + 1}
    return funcs


def events_of(body):
    """Event literals: the InvolvedObject block followed by Reason/Message."""
    out = []
    pat = re.compile(
        r"InvolvedObject: corev1\.ObjectReference\{\s*Kind:\s*\"(\w+)\",\s*Name:\s*([^\n]+?),\s*"
        r"Namespace:\s*\"(\w+)\",\s*\},\s*Reason:\s*\"(\w+)\",\s*Message:\s*\"([^\"]*)\"", re.S)
    for m in pat.finditer(body):
        out.append({"kind": m.group(1), "name_expr": m.group(2).strip(), "namespace": m.group(3),
                    "reason": m.group(4), "message": m.group(5)})
    return out


def objects_of(body):
    kinds = re.findall(
        r"TypeMeta: metav1\.TypeMeta\{\s*APIVersion:\s*\"[^\"]+\",\s*Kind:\s*\"(\w+)\"", body)
    return sorted(set(k for k in kinds if k != "Event"))


def main():
    test = open(os.path.join(REF, "services/supervisor_test.go")).read()
    cql = open(os.path.join(REF, "test-resources/checkpoints.cql")).read()
    seeds = seeded_stages(cql)
    funcs = split_funcs(test)
    main_body = funcs["TestSupervisor"]["body"]

    getters = re.findall(r"append\(k8sObjects, (\w+)\(([^)]*)\)\.\.\.\)", main_body)
    validators = re.findall(r"^\s*(validate\w+)\(f, (.*?), t\)", main_body, re.M)
    cfg = dict(re.findall(
        r"(RateLimitElementsPerSecond|RateLimitElementsBurst|Workers):\s*(\d+)", main_body))

    expected = {}  # request id -> asserted stage constant
    for vname, args in validators:
        ids = re.findall(r"\"([^\"]+)\"", args)
        consts = re.findall(r"LifecycleStage != models\.(\w+)", funcs[vname]["body"])
        # validateDeadlinedJobObjects checks result1 (first id) then result2 (second id)
        for rid, const in zip(ids, consts):
            expected[rid] = const

    scenarios = []
    for gname, args in getters:
        ids = re.findall(r"\"([^\"]+)\"", args)
        params = [p.strip().split()[0] for p in funcs[gname]["params"].split(",")]
        binding = dict(zip(params, ids))
        for ev in events_of(funcs[gname]["body"]):
            expr = ev.pop("name_expr")
            m = re.match(r'fmt\.Sprintf\("%s(-\w+)", (\w+)\)', expr)
            if m:
                rid = binding[m.group(2)]
                involved = rid + m.group(1)
            else:
                rid = binding[expr]
                involved = rid
            ev.update({
                "scenario": gname,
                "test_line": funcs[gname]["line"],
                "involved_name": involved,
                "request_id": rid,
                "objects_present": objects_of(funcs[gname]["body"]),
                "algorithm": seeds[rid]["algorithm"],
                "seeded_stage": seeds[rid]["lifecycle_stage"],
                "expected_stage_const": expected[rid],
            })
            scenarios.append(ev)

    out = {
        "source": ["services/supervisor_test.go", "test-resources/checkpoints.cql"],
        "processing_config": {k: int(v) for k, v in cfg.items()},
        "scenarios": scenarios,
    }
    path = os.path.join(HERE, "supervisor_test_scenarios.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1)
        f.write("\n")
    print(f"wrote {len(scenarios)} scenarios to {path}")


if __name__ == "__main__":
    sys.exit(main())
