"""Finds names that are read but never bound anywhere they could come from (the class of bug that
only shows when a rarely used branch finally runs). No third-party linter is needed:

    python tools/check_names.py [paths...]        # default: the whole repository

Scope rules are the ones of the language, approximated conservatively: a name is fine if it is
bound in the function (or an enclosing function), at module level, or is a builtin; class bodies
are visible only to code directly in the class body. Star imports and ``exec`` silence a module.
"""
import ast
import builtins
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SKIP_DIRS = {".git", "baseline", "gpurun_out", "__pycache__", "_lib", "build", ".pytest_cache"}
BUILTINS = set(dir(builtins)) | {"__file__", "__name__", "__doc__", "__path__", "__spec__", "__builtins__", "__class__"}


def _bound_names(node):
    """Names bound directly in this scope (not in nested functions / classes / comprehensions)."""
    names = set()

    def targets(t):
        if isinstance(t, ast.Name):
            names.add(t.id)
        elif isinstance(t, (ast.Tuple, ast.List)):
            for e in t.elts:
                targets(e)
        elif isinstance(t, ast.Starred):
            targets(t.value)

    def visit(n):
        for child in ast.iter_child_nodes(n):
            if isinstance(child, (ast.FunctionDef, ast.AsyncFunctionDef, ast.ClassDef)):
                names.add(child.name)
                for d in child.decorator_list:
                    visit(d)
                continue
            if isinstance(child, ast.Lambda):
                continue
            if isinstance(child, (ast.ListComp, ast.SetComp, ast.DictComp, ast.GeneratorExp)):
                # walrus targets inside comprehensions bind in the enclosing scope
                for sub in ast.walk(child):
                    if isinstance(sub, ast.NamedExpr):
                        targets(sub.target)
                continue
            if isinstance(child, (ast.Import, ast.ImportFrom)):
                for a in child.names:
                    names.add((a.asname or a.name).split(".")[0])
            elif isinstance(child, (ast.Assign,)):
                for t in child.targets:
                    targets(t)
            elif isinstance(child, (ast.AugAssign, ast.AnnAssign)):
                targets(child.target)
            elif isinstance(child, (ast.For, ast.AsyncFor)):
                targets(child.target)
            elif isinstance(child, (ast.With, ast.AsyncWith)):
                for item in child.items:
                    if item.optional_vars is not None:
                        targets(item.optional_vars)
            elif isinstance(child, ast.ExceptHandler) and child.name:
                names.add(child.name)
            elif isinstance(child, ast.NamedExpr):
                targets(child.target)
            elif isinstance(child, (ast.Global, ast.Nonlocal)):
                names.update(child.names)
            elif isinstance(child, ast.MatchAs) and child.name:
                names.add(child.name)
            elif isinstance(child, ast.MatchStar) and child.name:
                names.add(child.name)
            visit(child)

    visit(node)
    return names


def _args(fn):
    a = fn.args
    out = {x.arg for x in a.posonlyargs + a.args + a.kwonlyargs}
    if a.vararg:
        out.add(a.vararg.arg)
    if a.kwarg:
        out.add(a.kwarg.arg)
    return out


def check_file(path):
    src = open(path, encoding="utf-8").read()
    try:
        tree = ast.parse(src, path)
    except SyntaxError as e:
        return [(path, e.lineno or 0, f"syntax error: {e.msg}")]
    for n in ast.walk(tree):
        if isinstance(n, ast.ImportFrom) and any(a.name == "*" for a in n.names):
            return []
        if isinstance(n, ast.Call) and isinstance(n.func, ast.Name) and n.func.id in ("exec", "eval"):
            return []
    module_names = _bound_names(tree) | BUILTINS
    problems = []

    def walk(node, visible, in_class_body):
        """visible: names usable from this scope's code."""
        for child in ast.iter_child_nodes(node):
            if isinstance(child, (ast.FunctionDef, ast.AsyncFunctionDef, ast.Lambda)):
                if not isinstance(child, ast.Lambda):
                    for d in child.decorator_list:
                        walk_expr(d, visible)
                    for default in child.args.defaults + [d for d in child.args.kw_defaults if d is not None]:
                        walk_expr(default, visible)
                inner = visible if not in_class_body else visible - in_class_body
                local = _args(child) | (_bound_names(child) if not isinstance(child, ast.Lambda) else set())
                walk(child, inner | local, None)
            elif isinstance(child, ast.ClassDef):
                for b in child.bases + [k.value for k in child.keywords] + child.decorator_list:
                    walk_expr(b, visible)
                body = _bound_names(child)
                # names bound only in the class body are hidden from the methods defined in it
                walk(child, visible | body, body - visible)
            elif isinstance(child, (ast.ListComp, ast.SetComp, ast.DictComp, ast.GeneratorExp)):
                comp_names = set()
                for g in child.generators:
                    for sub in ast.walk(g.target):
                        if isinstance(sub, ast.Name):
                            comp_names.add(sub.id)
                walk(child, visible | comp_names, in_class_body)
            elif isinstance(child, ast.Name):
                if isinstance(child.ctx, ast.Load) and child.id not in visible:
                    problems.append((path, child.lineno, f"undefined name '{child.id}'"))
            else:
                walk(child, visible, in_class_body)

    def walk_expr(expr, visible):
        holder = ast.Expr(value=expr)
        walk(holder, visible, None)

    walk(tree, module_names, None)
    return problems


def main(argv):
    paths = argv or [ROOT]
    files = []
    for p in paths:
        if os.path.isfile(p):
            files.append(p)
            continue
        for d, dirs, fs in os.walk(p):
            dirs[:] = [x for x in dirs if x not in SKIP_DIRS]
            files += [os.path.join(d, f) for f in fs if f.endswith(".py")]
    bad = []
    for f in sorted(files):
        bad += check_file(f)
    for path, line, msg in bad:
        print(f"{os.path.relpath(path, ROOT)}:{line}: {msg}")
    print(f"{len(files)} files, {len(bad)} problem(s)")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
