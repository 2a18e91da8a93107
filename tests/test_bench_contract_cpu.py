"""bench.py's line contract, checked without a GPU: the reference arm (`--impl reference`: the oracle port on
the host cores) prints ONE JSON line with the keys the driver reads, the same `config` keys as the product arm
builds through `base_config`, and an `e2e` that moves no bytes; under a multi-rank launch only rank 0 prints."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(extra_env=None, args=()):
    env = dict(os.environ)
    env.pop("RANK", None); env.pop("WORLD_SIZE", None); env.pop("LOCAL_RANK", None)
    env.update(extra_env or {})
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1", *args],
                       cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    return [l for l in p.stdout.splitlines() if l.startswith("{")]


@pytest.fixture(scope="module")
def line():
    lines = _run()
    assert len(lines) == 1
    return json.loads(lines[0])


def test_reference_arm_line_has_the_contract_keys(line):
    import bench
    assert line["impl"] == "reference"
    assert line["metric"] == bench.METRIC and line["unit"] == bench.UNIT
    assert line["higher_is_better"] is True and line["n_gpus"] == 1 and line["steps"] == 1
    assert line["warmup"] >= 3                     # the timing rules ask for at least three warm-up steps
    assert line["value"] > 0 and line["ms_per_step"] == pytest.approx(1e3 / line["value"], rel=1e-6)
    assert line["vs_baseline"] is None and line["data"] == "synthetic" and line["dtype"] == "f32"
    cb = line["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == line["value"] and "fit iterations" in cb["sample"]
    e = line["e2e"]
    assert e["value"] == line["value"] and e["unit"] == line["unit"]
    assert e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0


def test_every_arm_builds_config_with_the_same_keys(line):
    import bench
    want = set(bench.base_config(0, "p", "e", "x", "l"))
    assert set(line["config"]) == want
    assert line["config"]["workload"] == bench.WORKLOAD and line["config"]["workload"].startswith("C2")
    assert (line["config"]["H"], line["config"]["W"], line["config"]["gaussians"]) == (1080, 1920, 50_000)
    assert line["config"]["intersections"] > 100_000        # the model as initialised (same seed as the product arm's rank 0)


def test_other_ranks_of_a_multi_rank_launch_print_nothing():
    assert _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}, ("--gpus", "2")) == []
