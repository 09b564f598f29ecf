"""Installs the UNMODIFIED reference under baseline/_ref (git-ignored).

1. ``pip install --no-index --no-build-isolation --target baseline/_ref <copy of /root/reference>``.
   The reference's setup.py packages ``find_packages(include=("megatron.core"))`` -- a plain
   string, so setuptools matches no package and the wheel contains metadata only.
2. Therefore the source tree itself (``megatron/``, ``finetune.py``) is copied byte for
   byte next to that metadata. Nothing is patched; API gaps of this image (apex missing,
   torch 2.11 removals) are bridged from OUTSIDE by baseline/shims and baseline/run_reference.py.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.environ.get("REFERENCE_SRC", "/root/reference")
DST = os.path.join(HERE, "_ref")


def main():
    if not os.path.isdir(REF_SRC):
        print(f"reference source {REF_SRC} not found", file=sys.stderr)
        return 1
    tmp = "/tmp/_refcopy_install"
    shutil.rmtree(tmp, ignore_errors=True)
    shutil.copytree(REF_SRC, tmp)
    os.makedirs(DST, exist_ok=True)
    r = subprocess.run([sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation",
                        "--find-links", "/opt/wheelhouse", "--no-deps", "--upgrade", "--target", DST, tmp],
                       capture_output=True, text=True)
    print(r.stdout[-400:], r.stderr[-400:])
    for name in ("megatron", "finetune.py", "LICENSE"):
        src, dst = os.path.join(REF_SRC, name), os.path.join(DST, name)
        if os.path.isdir(src):
            shutil.rmtree(dst, ignore_errors=True)
            shutil.copytree(src, dst, ignore=shutil.ignore_patterns("__pycache__", "*.so", "*.o"))
        else:
            shutil.copy2(src, dst)
    print("reference installed under", DST)
    return 0


if __name__ == "__main__":
    sys.exit(main())
