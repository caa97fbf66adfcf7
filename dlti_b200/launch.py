"""Single-node launcher: one process per GPU.

    python -m dlti_b200.launch --num_gpus=4 training/train_deepspeed_zero3.py --deepspeed configs/...

Mirrors what `deepspeed --num_gpus=N script ...` does in the reference notebook
(training/train.ipynb:137,231,249-253,822-824): sets MASTER_ADDR=127.0.0.1, MASTER_PORT (29500),
RANK / LOCAL_RANK / WORLD_SIZE, passes --local_rank=i, and tears every rank down as soon as one
exits non-zero (fail fast; §5.3). torchrun works too — the scripts only read the env.
"""
from __future__ import annotations

import argparse
import os
import signal
import subprocess
import sys
import time


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="dlti_b200.launch")
    ap.add_argument("--num_gpus", "--nproc_per_node", dest="n", type=int, default=1)
    ap.add_argument("--master_port", type=int, default=29500)
    ap.add_argument("--master_addr", type=str, default="127.0.0.1")
    ap.add_argument("--no_local_rank", action="store_true", help="do not append --local_rank=i")
    ap.add_argument("script")
    ap.add_argument("script_args", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    procs = []
    for r in range(a.n):
        env = dict(os.environ, MASTER_ADDR=a.master_addr, MASTER_PORT=str(a.master_port), RANK=str(r),
                   LOCAL_RANK=str(r), WORLD_SIZE=str(a.n))
        cmd = [sys.executable, "-u", a.script, *a.script_args]
        if not a.no_local_rank:
            cmd.append(f"--local_rank={r}")
        procs.append(subprocess.Popen(cmd, env=env, start_new_session=True))
    rc = 0
    try:
        alive = set(range(a.n))
        while alive:
            for r in list(alive):
                code = procs[r].poll()
                if code is None:
                    continue
                alive.discard(r)
                if code != 0:
                    rc = code
                    print(f"[launch] rank {r} exited with code {code}; killing the other ranks", file=sys.stderr)
                    for o in alive:
                        try:
                            os.killpg(procs[o].pid, signal.SIGTERM)
                        except ProcessLookupError:
                            pass
                    alive.clear()
            time.sleep(0.2)
    except KeyboardInterrupt:
        for p in procs:
            try:
                os.killpg(p.pid, signal.SIGTERM)
            except ProcessLookupError:
                pass
        rc = 130
    for p in procs:
        try:
            p.wait(timeout=30)
        except subprocess.TimeoutExpired:
            os.killpg(p.pid, signal.SIGKILL)
    return rc


if __name__ == "__main__":
    sys.exit(main())
