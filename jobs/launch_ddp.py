#!/usr/bin/env python3
"""One-node stand-in for the DAG task that starts the training ranks (dags/2_pytorch_training.py:49-78): the
reference runs `python3 /workspace/jobs/train_lightning_ddp.py` once per container -- rank 0 first, the next rank five
seconds later -- with MASTER_ADDR / MASTER_PORT / NODE_RANK / WORLD_SIZE in each container's environment
(docker-compose.yml:120-125,140-145), waits for all of them and succeeds only if every rank exits 0.

Here the "nodes" are the GPUs of one box: rank i gets NODE_RANK=i, LOCAL_RANK=i (its GPU), the same WORLD_SIZE,
MASTER_ADDR=127.0.0.1 and one MASTER_PORT.  The staggered start is kept (the rendezvous has to tolerate it).

    python3 jobs/launch_ddp.py [--ranks N] [--stagger SECONDS] [--timeout SECONDS] [--] [command ...]

Defaults: N = WORLD_SIZE or the number of visible GPUs, stagger 5 s (the DAG's `sleep 5`), timeout 3 h (the DAG's
execution_timeout), command = this repository's jobs/train_lightning_ddp.py.
"""
import argparse
import os
import socket
import subprocess
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))


def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return int(s.getsockname()[1])


def visible_gpus() -> int:
    cvd = os.environ.get("CUDA_VISIBLE_DEVICES")
    if cvd is not None:
        return len([d for d in cvd.split(",") if d.strip()])
    try:
        out = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True, timeout=30).stdout
        return sum(1 for ln in out.splitlines() if ln.startswith("GPU "))
    except Exception:
        return 0


def rank_env(rank: int, world: int, port: int, base=None) -> dict:
    env = dict(os.environ if base is None else base)
    env.update({"NODE_RANK": str(rank), "RANK": str(rank), "LOCAL_RANK": str(rank), "WORLD_SIZE": str(world),
                "MASTER_ADDR": env.get("MASTER_ADDR", "127.0.0.1"), "MASTER_PORT": str(port)})
    return env


def launch(cmd, ranks: int, stagger: float, timeout: float, port=None) -> int:
    port = port or int(os.environ.get("MASTER_PORT", 0)) or free_port()
    print("🚀 Starting Distributed Training...", flush=True)
    procs = []
    for r in range(ranks):
        if r > 0 and stagger > 0:
            time.sleep(stagger)                                   # `sleep 5` between master and worker (:57)
        procs.append(subprocess.Popen(cmd, env=rank_env(r, ranks, port)))
    deadline = time.time() + timeout
    codes = []
    for p in procs:                                               # `wait $MASTER_PID; wait $WORKER_PID` (:63-66)
        try:
            codes.append(p.wait(timeout=max(1.0, deadline - time.time())))
        except subprocess.TimeoutExpired:
            codes.append(None)
    if any(c is None for c in codes):
        for p in procs:
            if p.poll() is None:
                p.kill()
        print(f"✗ Training Failed (timeout after {timeout:.0f} s; exit codes {codes})", flush=True)
        return 1
    if all(c == 0 for c in codes):
        print("✓ Training Success", flush=True)
        return 0
    print("✗ Training Failed (" + ", ".join(f"rank {i}: {c}" for i, c in enumerate(codes)) + ")", flush=True)
    return 1


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--ranks", type=int, default=0)
    ap.add_argument("--stagger", type=float, default=5.0)
    ap.add_argument("--timeout", type=float, default=3 * 3600.0)
    ap.add_argument("command", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    ranks = a.ranks or int(os.environ.get("WORLD_SIZE", 0)) or visible_gpus()
    if ranks < 1:
        print("launch_ddp: no GPUs visible and no --ranks / WORLD_SIZE given", file=sys.stderr)
        return 2
    cmd = [c for c in a.command if c != "--"] or [sys.executable, os.path.join(HERE, "train_lightning_ddp.py")]
    return launch(cmd, ranks, a.stagger, a.timeout)


if __name__ == "__main__":
    sys.exit(main())
