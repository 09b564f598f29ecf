import sys, json
sys.path.insert(0, '.')
from tests.helpers.launch import run_tiny
base = run_tiny(steps=4, layers=4, num_microbatches=4)
print("base           ", base["losses"], flush=True)
for name, kw in [("vpp auto", {}), ("vpp kernels off", {"kernels": "off"}), ("vpp gemm cublas", {"gemm": "cublas"}),
                 ("pp2 1f1b auto", {"vpp_layers": None})]:
    k = dict(nproc=2, pp=2, vpp_layers=1, layers=4, steps=4, num_microbatches=4)
    k.update(kw)
    try:
        r = run_tiny(**k)
        print(name.ljust(16), r["losses"], flush=True)
    except AssertionError as e:
        print(name, "FAILED", str(e)[-600:], flush=True)
