"""bench.py contract pieces that run without a GPU: the reference arm prints exactly one JSON
line with the agreed keys (the unmodified reference is a CPU program)."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


@pytest.mark.skipif(not (ROOT / "baseline" / "_ref" / "simple_distributed.py").exists()
                    and not Path("/root/reference/simple_distributed.py").exists(),
                    reason="reference not available")
def test_reference_arm_prints_one_json_line():
    if not (ROOT / "baseline" / "_ref" / "simple_distributed.py").exists():
        subprocess.run([sys.executable, str(ROOT / "baseline" / "install_reference.py")], check=True,
                       stdout=subprocess.DEVNULL)
    env = dict(os.environ, MASTER_PORT="29733")
    env.pop("RANK", None); env.pop("WORLD_SIZE", None)
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--steps", "6",
                          "--warmup", "2"], capture_output=True, text=True, timeout=600, env=env, cwd=str(ROOT))
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, out.stdout + out.stderr[-2000:]
    d = json.loads(lines[0])
    assert d["impl"] == "reference"
    if "unavailable" in d:
        return
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
                "config", "e2e", "dtype", "data"):
        assert key in d, key
    assert d["steps"] == 6 and d["warmup"] == 2 and d["value"] > 0
    assert "LeNet" in d["config"]["model"] and d["config"]["global_batch"] == 60


def test_bench_layouts_cover_named_configs():
    sys.path.insert(0, str(ROOT))
    import bench
    class A:  # minimal stand-in for argparse output
        microbatches = 0; schedule = ""; mlp_microbatches = 8
    assert bench.layout_for("lenet", 1, A)[:3] == (1, 1, 1)
    assert bench.layout_for("lenet", 8, A)[:3] == (1, 8, 1)
    assert bench.layout_for("lenet_pp", 8, A)[:4] == (2, 4, 1, 8)        # config 4: 2 stages x 4 replicas
    assert bench.layout_for("lenet_pp", 2, A)[:5] == (2, 1, 1, 8, "1f1b")  # config 2
    assert bench.layout_for("mlp", 8, A)[:3] == (8, 1, 1)                # config 3
    assert bench.layout_for("mlp_tp", 8, A)[:3] == (4, 1, 2)             # config 5
    assert bench.layout_for("mlp", 8, A)[3] == 8                         # the named line keeps 8 micro-batches at pp8
    assert bench.layout_for("mlp_mb16", 8, A)[3] == 16 and bench.layout_for("mlp_mb32", 8, A)[3] == 32
    assert bench.layout_for("mlp_tp_fp8", 4, A)[:3] == (2, 1, 2)


def test_default_run_measures_every_named_config_that_fits():
    """Config 2 / 4 (`lenet_pp`), 3 (`mlp`), 5 (`mlp_tp`, `mlp_tp_fp8`) + fp8 and DP variants by default at
    N >= 2; one GPU cannot hold a pipeline, a DP group or a TP pair."""
    sys.path.insert(0, str(ROOT))
    import bench
    assert bench.default_extras(1) == ["mlp", "mlp_fp8"]
    for n in (2, 4, 8):
        ex = bench.default_extras(n)
        for name in ("lenet_pp", "mlp", "mlp_fp8", "mlp_dp", "mlp_tp", "mlp_tp_fp8"):
            assert name in ex, (n, name)
    assert "mlp_mb16" in bench.default_extras(8) and "mlp_mb32" in bench.default_extras(8)
    assert "mlp_mb16" not in bench.default_extras(4)
