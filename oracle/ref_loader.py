"""Imports the UNMODIFIED reference solvers from /root/reference (build container) or from the copy that
`__graft_entry__.build()` stages into the git-ignored oracle/_ref/ (travels to the GPU box with the working tree).

TEST / BENCH INFRASTRUCTURE: used by oracle/make_golden.py, by tests that pin the C oracle against the live
reference, and by bench.py's CPU legs (`cpu_baseline`, `--impl reference`), which time the reference's own solvers.
Nothing in the product path imports this.

The reference package (`/root/reference/fmap2pointmap_solvers/__init__.py:1-8`) imports `gin`, which is not
installed here; a no-op `gin.configurable` shim is injected so the reference code itself stays untouched.
"""
import importlib
import os
import sys
import types

_HERE = os.path.dirname(os.path.abspath(__file__))
STAGED_ROOT = os.path.join(_HERE, "_ref")
REFERENCE_FILES = ("fmap2pointmap_solvers/__init__.py", "fmap2pointmap_solvers/naive.py",
                   "fmap2pointmap_solvers/spacial_filtering.py")


def _pick_root():
    env = os.environ.get("F2P_REFERENCE_ROOT")
    if env:
        return env
    if os.path.isfile("/root/reference/fmap2pointmap_solvers/naive.py"):
        return "/root/reference"
    return STAGED_ROOT


REFERENCE_ROOT = _pick_root()


def reference_root() -> str:
    return REFERENCE_ROOT


def stage_reference(src_root="/root/reference") -> bool:
    """Copies the reference's solver package (3 files, 117 lines) byte for byte into oracle/_ref/ so that the GPU box —
    which has no /root/reference — can time the reference's own code.  oracle/_ref/ is git-ignored build output, never
    committed.  Returns False when the source tree is absent (the GPU box: the staged copy is used as it is)."""
    import hashlib
    import json
    import shutil
    if not os.path.isfile(os.path.join(src_root, REFERENCE_FILES[1])):
        return False
    manifest = {}
    for rel in REFERENCE_FILES:
        dst = os.path.join(STAGED_ROOT, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(src_root, rel), dst)
        manifest[rel] = hashlib.sha256(open(dst, "rb").read()).hexdigest()
    with open(os.path.join(STAGED_ROOT, "MANIFEST.json"), "w") as f:
        json.dump({"source": src_root, "sha256": manifest}, f, indent=1)
    return True


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "fmap2pointmap_solvers", "naive.py"))


def _gin_shim():
    gin = types.ModuleType("gin")

    def configurable(*args, **kwargs):
        if args and callable(args[0]) and not kwargs:
            return args[0]
        return lambda fn: fn

    gin.configurable = configurable
    gin.REQUIRED = object()
    return gin


def load_reference():
    """Returns the reference `fmap2pointmap_solvers` module object (not the drop-in of the same name)."""
    if not reference_available():
        raise RuntimeError(f"reference not found under {REFERENCE_ROOT}")
    saved_gin = sys.modules.get("gin")
    saved_pkg = {k: v for k, v in sys.modules.items() if k == "fmap2pointmap_solvers" or k.startswith("fmap2pointmap_solvers.")}
    for k in saved_pkg:
        del sys.modules[k]
    if saved_gin is None:
        sys.modules["gin"] = _gin_shim()
    sys.path.insert(0, REFERENCE_ROOT)
    try:
        mod = importlib.import_module("fmap2pointmap_solvers")
        sub = {k: v for k, v in sys.modules.items() if k.startswith("fmap2pointmap_solvers")}
    finally:
        sys.path.remove(REFERENCE_ROOT)
        for k in list(sys.modules):
            if k == "fmap2pointmap_solvers" or k.startswith("fmap2pointmap_solvers."):
                del sys.modules[k]
        sys.modules.update(saved_pkg)
        if saved_gin is None:
            del sys.modules["gin"]
    # keep the reference modules alive under private names so pickling/inspection still works
    for k, v in sub.items():
        sys.modules["_f2p_reference." + k] = v
    assert os.path.realpath(mod.__file__).startswith(os.path.realpath(REFERENCE_ROOT)), mod.__file__
    return mod


def load_reference_utils():
    """The reference's `utils/utils.py` module object (compute_inlier_ratio, C_from_sparse_P, ...), unmodified."""
    import importlib.util
    path = os.path.join(REFERENCE_ROOT, "utils", "utils.py")
    if not os.path.isfile(path):
        raise RuntimeError(f"reference utils not found under {REFERENCE_ROOT}")
    saved_gin = sys.modules.get("gin")
    if saved_gin is None:
        sys.modules["gin"] = _gin_shim()
    try:
        spec = importlib.util.spec_from_file_location("_f2p_reference.utils_utils", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        if saved_gin is None:
            del sys.modules["gin"]
    return mod


def load_reference_method(rel_path, class_name, func_name):
    """Compiles ONE method of a reference class from its unmodified source text, without importing the module (whose
    top-level imports — open3d, robust_laplacian ... — are not installed here).  Used for `base_object_dataset.find_positives`
    (dataset/object.py:281-288), which only needs NumPy."""
    import ast
    import numpy as np
    path = os.path.join(REFERENCE_ROOT, rel_path)
    src = open(path).read()
    tree = ast.parse(src)
    for node in ast.walk(tree):
        if isinstance(node, ast.ClassDef) and node.name == class_name:
            for item in node.body:
                if isinstance(item, ast.FunctionDef) and item.name == func_name:
                    code = ast.get_source_segment(src, item)
                    import textwrap
                    ns = {"np": np}
                    exec(compile(textwrap.dedent(code), path, "exec"), ns)
                    return ns[func_name]
    raise RuntimeError(f"{class_name}.{func_name} not found in {path}")
