"""TEST / BASELINE INFRASTRUCTURE -- builds the reference's own env modules into oracle/_ref/.

    python oracle/make_ref.py            (build container only: needs /root/reference)

The reference is Python; its three env modules (othello.py, simple_policies.py, util.py) are COMPILED -- from the
sources where they lie under /root/reference, nothing is copied -- to marshalled CPython code objects in git-ignored
oracle/_ref/ (`othello.code`, ...; `*.pyc` files do not survive the snapshot to the GPU box), which
`load_modules()` executes into fresh modules.  oracle/_ref/ is not gpurun-ignored, so
the build travels to the GPU box like the CUDA library does, and `bench.py --impl reference` / the `cpu_baseline`
legs time the REFERENCE'S OWN code there (`cpu_baseline.kind: "reference"`); when oracle/_ref/ is absent they fall
back to the port oracle/rules_py.py and say so.  The GPU box runs the same image, hence the same CPython magic
number (recorded in oracle/_ref/BUILD_INFO.json and checked at import).
Never imported by the product package.
"""
import importlib.util
import json
import os
import marshal
import types
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("OTHELLO_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(HERE, "_ref")
MODULES = ("othello", "simple_policies", "util")


def build(force=False):
    """-> OUT, or None when the reference tree is not present (GPU box: the prebuilt files are used as they are)."""
    if not os.path.isfile(os.path.join(REF_ROOT, "othello.py")):
        return None
    os.makedirs(OUT, exist_ok=True)
    for m in MODULES:
        src, dst = os.path.join(REF_ROOT, m + ".py"), os.path.join(OUT, m + ".code")
        if force or not os.path.isfile(dst) or os.path.getmtime(dst) < os.path.getmtime(src):
            with open(src, "rb") as f:
                code = compile(f.read(), "reference/" + m + ".py", "exec", dont_inherit=True, optimize=0)
            with open(dst, "wb") as f:
                marshal.dump(code, f)
    with open(os.path.join(OUT, "BUILD_INFO.json"), "w") as f:
        json.dump({"modules": list(MODULES), "python": sys.version.split()[0],
                   "magic": importlib.util.MAGIC_NUMBER.hex(), "built_from": REF_ROOT}, f)
    return OUT


def available():
    info = os.path.join(OUT, "BUILD_INFO.json")
    if not os.path.isfile(info):
        return False
    try:
        return json.load(open(info)).get("magic") == importlib.util.MAGIC_NUMBER.hex() and all(
            os.path.isfile(os.path.join(OUT, m + ".code")) for m in MODULES)
    except (OSError, ValueError):
        return False


def load_modules():
    """Executes the built code objects into modules named like the reference's (their own `import othello` etc.
    resolve through sys.modules) -> (othello, simple_policies, util).  The caller installs the gym / pyglet stubs."""
    mods = []
    for m in MODULES:                                  # dependency order: othello <- simple_policies <- util
        if m in sys.modules and getattr(sys.modules[m], "__oracle_ref__", False):
            mods.append(sys.modules[m])
            continue
        with open(os.path.join(OUT, m + ".code"), "rb") as f:
            code = marshal.load(f)
        mod = types.ModuleType(m)
        mod.__file__ = os.path.join(OUT, m + ".code")
        mod.__oracle_ref__ = True
        sys.modules[m] = mod
        exec(code, mod.__dict__)
        mods.append(mod)
    return tuple(mods)


if __name__ == "__main__":
    print(build(force="--force" in sys.argv) or "reference tree not present: nothing built")
