"""jobs/launch_ddp.py: the one-node form of the DAG task that starts the ranks (dags/2_pytorch_training.py:49-78) --
env contract per rank, staggered start, success only if every rank exits 0."""
import json
import os
import subprocess
import sys
import tempfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LAUNCH = os.path.join(ROOT, "jobs", "launch_ddp.py")


def test_launcher_env_contract_stagger_and_exit_codes():
    tmp = tempfile.mkdtemp()
    probe = ("import os, time, json, sys; r = os.environ['NODE_RANK']; "
             f"json.dump({{k: os.environ.get(k) for k in ('NODE_RANK', 'LOCAL_RANK', 'WORLD_SIZE', 'MASTER_ADDR', 'MASTER_PORT')}} | {{'t': time.time()}}, "
             f"open(os.path.join({tmp!r}, 'rank' + r + '.json'), 'w')); sys.exit(int(os.environ.get('FAIL_RANK', '-1')) == int(r))")
    r = subprocess.run([sys.executable, LAUNCH, "--ranks", "3", "--stagger", "0.5", "--", sys.executable, "-c", probe],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "✓ Training Success" in r.stdout, r.stdout + r.stderr
    recs = [json.load(open(os.path.join(tmp, f"rank{i}.json"))) for i in range(3)]
    for i, rec in enumerate(recs):
        assert rec["NODE_RANK"] == str(i) and rec["LOCAL_RANK"] == str(i) and rec["WORLD_SIZE"] == "3"
        assert rec["MASTER_ADDR"] == "127.0.0.1"
    assert len({rec["MASTER_PORT"] for rec in recs}) == 1
    assert recs[1]["t"] - recs[0]["t"] >= 0.4 and recs[2]["t"] - recs[1]["t"] >= 0.4       # rank 0 first, then the others
    env = dict(os.environ, FAIL_RANK="1")
    r = subprocess.run([sys.executable, LAUNCH, "--ranks", "2", "--stagger", "0", "--", sys.executable, "-c", probe],
                       capture_output=True, text=True, timeout=120, env=env)
    assert r.returncode == 1 and "✗ Training Failed" in r.stdout and "rank 1: 1" in r.stdout


@pytest.mark.gpu
def test_launcher_runs_the_cli_on_two_gpus_with_a_staggered_start():
    """WORLD_SIZE=2 on two GPUs through the launcher, rank 1 started two seconds after rank 0: both exit 0, rank 0 writes
    the checkpoints, and the many-steps-per-launch path with the in-kernel DDP exchange is the one that ran."""
    import numpy as np
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from oracle import preprocess_ref, synth

    tmp = tempfile.mkdtemp()
    t = synth.synth_table(11, 4000, 5)
    pp = preprocess_ref.preprocess_table(t["raw"], t["labels"])
    preprocess_ref.write_parquet_dir(os.path.join(tmp, "processed"), t["names"], pp["norm"], pp["labels32"], n_parts=2)
    env = dict(os.environ, WX_DATA_DIR=os.path.join(tmp, "processed"), WX_MODEL_DIR=os.path.join(tmp, "models"),
               WX_MAX_EPOCHS="2", WX_METRICS_DIR=os.path.join(tmp, "metrics"), WX_DISABLE_MLFLOW="1")
    env.pop("WORLD_SIZE", None)
    r = subprocess.run([sys.executable, LAUNCH, "--ranks", "2", "--stagger", "2"], env=env, capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "✓ Training Success" in r.stdout and "Training finished." in r.stdout
    files = sorted(os.listdir(os.path.join(tmp, "models")))
    assert "last.ckpt" in files and any(f.startswith("weather-best-epoch=") for f in files)
    ck = torch.load(os.path.join(tmp, "models", "last.ckpt"), weights_only=False)
    # 4000 rows -> 3200 train rows -> 1600 per rank -> 400 optimizer steps per epoch, 2 epochs
    assert ck["global_step"] == 800 and np.isfinite(ck["state_dict"]["net.0.weight"].numpy()).all()
