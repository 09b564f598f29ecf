import json
import os
import socket
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def run_tiny(nproc=1, timeout=600, env_extra=None, **kw):
    """Launch tests/helpers/run_tiny.py (torchrun when nproc > 1) and return its RESULT dict."""
    script = os.path.join(ROOT, "tests", "helpers", "run_tiny.py")
    flags = []
    for k, v in kw.items():
        if v is True:
            flags.append(f"--{k}")
        elif v is not None and v is not False:
            flags.append(f"--{k}={v}")        # "=" form: values may themselves start with "--"
    for attempt in range(3):
        env = dict(os.environ)
        env["MASTER_ADDR"] = "127.0.0.1"
        env["MASTER_PORT"] = str(free_port())
        env.update(env_extra or {})
        if nproc == 1:
            env.update(RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
            cmd = [sys.executable, script] + flags
        else:
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
                   "--master-addr", "127.0.0.1", "--master-port", env["MASTER_PORT"], script] + flags
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env, cwd=ROOT)
        # the probed port can be taken again before the rendezvous binds it: try another one
        if r.returncode != 0 and "EADDRINUSE" in r.stderr and "RESULT " not in r.stdout:
            continue
        break
    lines = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
    assert r.returncode == 0 and lines, f"run_tiny failed (rc={r.returncode}):\n{r.stdout[-2000:]}\n{r.stderr[-4000:]}"
    return json.loads(lines[-1][7:])
