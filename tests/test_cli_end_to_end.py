"""The reference's own two-terminal recipe (readme:19; ref:138-186) with this repo's script:
one process per rank, the five reference flags, loopback interface -- on CPUs."""
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def test_reference_command_line_two_ranks(free_port, tmp_path):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="1")
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    base = [sys.executable, str(ROOT / "simple_distributed.py"), "--world_size=2", "--interface=lo",
            "--master_addr=127.0.0.1", f"--master_port={free_port}",
            # not in the reference: keep the run short and write checkpoints
            "--epochs=1", "--train_samples=600", "--test_samples=120", f"--ckpt_dir={tmp_path}"]
    procs = [subprocess.Popen(base + [f"--rank={r}"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                              env=env, cwd=str(tmp_path)) for r in (0, 1)]
    outs = []
    try:
        for p in procs:
            outs.append(p.communicate(timeout=300))
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
    assert [p.returncode for p in procs] == [0, 0], outs[0][1][-1500:] + outs[1][1][-1500:]
    # the rank that owns the loss (last stage) prints the reference's two log formats; the other is silent
    printed = outs[1][0]
    assert "Train Epoch: 1 [0/600 (0%)]\tLoss: " in printed
    assert "\nTest set: Average loss: " in printed and "Accuracy: " in printed
    assert outs[0][0].strip() == ""
    assert (tmp_path / "stage0.pt").exists() and (tmp_path / "stage1.pt").exists()


def test_local_launcher_starts_all_ranks_and_propagates_failure(tmp_path):
    """`python -m ...launch --nproc 2 -- <flags>`: both ranks run to completion; a bad flag on the
    ranks makes the launcher return non-zero instead of hanging."""
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="1", PYTHONPATH=str(ROOT))
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    base = [sys.executable, "-m", "simple_distributed_machine_learning_b200.launch", "--nproc", "2", "--"]
    ok = subprocess.run(base + ["--epochs=1", "--train_samples=240", "--test_samples=60", "--device=cpu",
                                f"--ckpt_dir={tmp_path}"], env=env, cwd=str(tmp_path), capture_output=True,
                        text=True, timeout=300)
    assert ok.returncode == 0, ok.stderr[-1500:]
    assert "Test set: Average loss:" in ok.stdout
    assert (tmp_path / "stage0.pt").exists() and (tmp_path / "stage1.pt").exists()
    bad = subprocess.run(base + ["--schedule=lockstep", "--microbatches=4"], env=env, cwd=str(tmp_path),
                         capture_output=True, text=True, timeout=120)
    assert bad.returncode != 0
