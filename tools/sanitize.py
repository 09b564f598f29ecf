"""compute-sanitizer over the kernel self-checks (the reference ships no
race / memory checking of any kind).

    python tools/sanitize.py                       # memcheck on the small checks of every kernel
    python tools/sanitize.py --tool racecheck attn_fwd_one_tile gemm_nt_pair_small
    python tools/sanitize.py --tool synccheck --all

ONE ``compute-sanitizer --tool T --error-exitcode 3`` invocation wraps a child (this script with
``--inner``) that runs the selected checks back to back in one process; the summary is written to
``gpurun_out/sanitize_<tool>.md``. racecheck only sees shared-memory hazards between threads; the
cross-GPU protocols are covered by the bounded waits plus ``--check_dp_consistency_interval`` and
``TPWorkspace.check()``. Runs ~50-100x slower than native: keep to the small shapes (default set).
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# one small shape per kernel family; tcgen05/TMA kernels included
DEFAULT = ["rmsnorm", "layernorm", "rope", "glu", "softmax", "xent", "adam",
           "gemm_nt_pair_small", "gemm_tn_pair_acc", "gemm_nt_f32", "attn_fwd_one_tile", "attn_fwd_gen1",
           "attn_bwd_small"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tool", default="memcheck", choices=["memcheck", "racecheck", "synccheck", "initcheck"])
    ap.add_argument("--all", action="store_true", help="every registered self-check")
    ap.add_argument("--timeout", type=int, default=600)
    ap.add_argument("--inner", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("names", nargs="*")
    a = ap.parse_args()
    sys.path.insert(0, ROOT)
    from epfl_megatron_b200.ops import selfcheck
    names = a.names or (list(selfcheck.CHECKS) if a.all else [n for n in DEFAULT if n in selfcheck.CHECKS])
    if a.inner:
        import torch
        for n in names:
            t0 = time.time()
            try:
                selfcheck.CHECKS[n]()
                torch.cuda.synchronize()
                print(f"INNER {n} ok {time.time() - t0:.0f}s", flush=True)
            except Exception as e:  # noqa: BLE001
                print(f"INNER {n} FAIL {type(e).__name__}: {str(e)[:200]}", flush=True)
        return 0
    cmd = ["compute-sanitizer", "--tool", a.tool, "--error-exitcode", "3", "--print-limit", "10",
           sys.executable, os.path.abspath(__file__), "--inner", "--tool", a.tool] + names
    t0 = time.time()
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=a.timeout, cwd=ROOT)
        out, rc = r.stdout + r.stderr, r.returncode
    except subprocess.TimeoutExpired as e:
        out, rc = (e.stdout or b"").decode(errors="replace") if isinstance(e.stdout, bytes) else (e.stdout or ""), "timeout"
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"sanitize_{a.tool}.raw.log"), "w") as f:
        f.write(" ".join(cmd) + "\n" + out)
    inner = {l.split()[1]: " ".join(l.split()[2:]) for l in out.splitlines() if l.startswith("INNER ")}
    summary = [l.strip() for l in out.splitlines() if "ERROR SUMMARY" in l]
    reports = [l.strip()[:200] for l in out.splitlines() if l.startswith("=========") and
               any(k in l for k in ("Invalid", "Race", "hazard", "Uninitialized", "Barrier error"))][:20]
    rows = [(n, inner.get(n, "not reached"), "", "") for n in names]
    bad = int(rc != 0 or any("ok" not in inner.get(n, "") for n in names))
    print("\n".join(f"{n}: {inner.get(n, 'not reached')}" for n in names))
    print(f"compute-sanitizer rc={rc} {summary[-1] if summary else ''} ({time.time() - t0:.0f}s)")
    rows.append(("(whole run)", f"rc={rc}", f"{time.time() - t0:.0f}s", (summary[-1] if summary else "") + " " + " / ".join(reports)))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"sanitize_{a.tool}.md"), "w") as f:
        f.write(f"# compute-sanitizer --tool {a.tool} over ops.selfcheck\n\n| check | result | time | summary |\n|---|---|---|---|\n")
        for row in rows:
            f.write("| " + " | ".join(row) + " |\n")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
