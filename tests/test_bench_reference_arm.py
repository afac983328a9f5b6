"""bench.py --impl reference (the CPU arm: oracle/train_ref.py under gloo DDP, world_size 2) must run on its own and
as rank 0 of a torchrun launch -- the driver starts it exactly like the GPU arm at N = 2, 4, 8."""
import json
import os
import socket
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return int(s.getsockname()[1])


def _line(stdout: str) -> dict:
    # ONE JSON line and nothing else on stdout: bench.py keeps a private copy of the real stdout for its line and points
    # file descriptor 1 at stderr, so that what libraries print (NCCL's banner on the GPU boxes) cannot land in front of it
    rows = [ln for ln in stdout.splitlines() if ln.strip()]
    assert len(rows) == 1 and rows[0].startswith("{"), stdout
    return json.loads(rows[0])


def test_reference_arm_runs_under_torchrun_world2():
    # torchrun exports TORCHELASTIC_USE_AGENT_STORE, RANK, WORLD_SIZE, MASTER_*: the arm's own gloo workers must not
    # inherit them (they rendezvous over their own tcp:// address)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), "bench.py", "--impl", "reference", "--gpus", "2", "--steps",
           "2", "--warmup", "1", "--rows-per-step", "4096"]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr[-2000:]
    line = _line(r.stdout)
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["steps"] == 2
    assert line["value"] > 0 and line["cpu_baseline"]["kind"] == "port"
    assert line["e2e"]["h2d_bytes_per_step"] == 0


def test_reference_arm_uses_the_gpu_arms_rows_per_step():
    r = subprocess.run([sys.executable, "bench.py", "--impl", "reference", "--steps", "2", "--warmup", "1",
                        "--rows-per-step", "2048"], cwd=ROOT, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr[-2000:]
    line = _line(r.stdout)
    assert line["config"]["rows_per_step_per_gpu"] == 2048          # same configuration as the GPU arm's line
    assert "1024 rows/rank x 2 ranks" in line["cpu_baseline"]["sample"]
