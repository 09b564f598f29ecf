"""Run autopep8 over the python sources (tools/linter.py:7-36 surface). autopep8 is optional;
without it the tool lists the files it would have formatted."""
import os
import os.path as osp
import sys


def format_file(path, apply=True):
    try:
        import autopep8
    except ImportError:
        print("autopep8 not installed; would format", path)
        return
    with open(path) as f:
        src = f.read()
    out = autopep8.fix_code(src, options={"max_line_length": 120, "aggressive": 1})
    if out != src and apply:
        with open(path, "w") as f:
            f.write(out)
        print("formatted", path)


def main(roots):
    for root in roots:
        for d, _, files in os.walk(root):
            if any(part.startswith(".") or part in ("_ref", "_lib") for part in d.split(os.sep)):
                continue
            for name in files:
                if name.endswith(".py"):
                    format_file(osp.join(d, name))


if __name__ == "__main__":
    here = osp.abspath(osp.join(osp.dirname(__file__), osp.pardir))
    main(sys.argv[1:] or [here])
