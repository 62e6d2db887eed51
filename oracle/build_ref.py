"""Recipe for oracle/_ref: the UNMODIFIED reference package, staged so that it travels to the GPU box.

TEST / BENCH INFRASTRUCTURE ONLY (see oracle/__init__.py).  The reference (godhj93/bayesian-torch) is pure
Python: "building" it is archiving its package directory, as it lies under /root/reference, into ONE importable
artefact oracle/_ref/bayesian_torch_ref.zip (zipimport; git-ignored, NOT gpurun-ignored -- it ships with the
snapshot like a built .so) together with a two-line `termcolor` stand-in (conv_variational.py:55 imports it,
the image does not have it).  No reference source enters the repository's history and nothing
is edited.

    python oracle/build_ref.py            # stage (needs /root/reference; a no-op elsewhere)
    python oracle/build_ref.py --check    # report what import_reference() would load

Who uses it: bench.py `--impl reference` and its `cpu_baseline` / `eager_b200` legs (the reference's own eager code,
timed on the box's host cores and on the box's GPU), and oracle/make_golden.py.  Never the product path.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys
import types
import zipfile

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/bayesian_torch"
DST = os.path.join(HERE, "_ref")
ZIP = os.path.join(DST, "bayesian_torch_ref.zip")
SKIP_DIRS = {"examples", "scripts", "__pycache__"}   # drivers and shell scripts: not on the path


def _tree_hash(root: str) -> str:
    h = hashlib.sha256()
    for d, dirs, files in sorted(os.walk(root)):
        dirs[:] = sorted(x for x in dirs if x not in SKIP_DIRS)
        for f in sorted(files):
            if f.endswith(".py"):
                p = os.path.join(d, f)
                h.update(os.path.relpath(p, root).encode())
                with open(p, "rb") as fh:
                    h.update(fh.read())
    return h.hexdigest()


def build(verbose: bool = False) -> str | None:
    """Stage the reference under oracle/_ref.  Returns the archive path, or None when neither /root/reference
    nor a staged archive exists (on the GPU box the snapshot already carries oracle/_ref)."""
    if not os.path.isdir(SRC):
        return ZIP if os.path.exists(ZIP) else None
    want = _tree_hash(SRC)
    info_path = os.path.join(DST, "BUILD_INFO.json")
    if os.path.exists(info_path) and os.path.exists(ZIP):
        try:
            if json.load(open(info_path)).get("sha256") == want:
                return ZIP
        except (OSError, ValueError):
            pass
    shutil.rmtree(DST, ignore_errors=True)
    os.makedirs(DST)
    with zipfile.ZipFile(ZIP, "w", zipfile.ZIP_DEFLATED) as z:
        for d, dirs, files in os.walk(SRC):
            dirs[:] = sorted(x for x in dirs if x not in SKIP_DIRS)
            for f in sorted(files):
                if f.endswith(".py"):
                    p = os.path.join(d, f)
                    z.write(p, os.path.join("bayesian_torch", os.path.relpath(p, SRC)))
        z.writestr("termcolor.py", '"""Stand-in for the `termcolor` package the reference imports for one coloured print."""\n'
                                   "def colored(s, *a, **k):\n    return s\n")
    with open(info_path, "w") as f:
        json.dump({"source": SRC, "sha256": want, "skipped": sorted(SKIP_DIRS),
                   "note": "archive of the unmodified reference package (build artefact, git-ignored)"}, f, indent=1)
    if verbose:
        print(f"staged {SRC} -> {ZIP} ({want[:16]})")
    return ZIP


def import_reference():
    """(layers, dnn_to_bnn module, util module, where) of the reference, from the staged archive when it exists,
    else from /root/reference; raises ImportError when neither exists."""
    root = None
    if os.path.exists(ZIP):
        root = ZIP
    elif os.path.isdir(SRC):
        root = os.path.dirname(SRC)
        stub = types.ModuleType("termcolor")
        stub.colored = lambda s, *a, **k: s
        sys.modules.setdefault("termcolor", stub)
    if root is None:
        raise ImportError("the reference is not staged: run `python oracle/build_ref.py` where /root/reference exists")
    if root not in sys.path:
        sys.path.insert(0, root)
    import bayesian_torch.layers as L  # noqa: E402
    from bayesian_torch.models import dnn_to_bnn as D  # noqa: E402
    from bayesian_torch.utils import util as U  # noqa: E402
    return L, D, U, root


if __name__ == "__main__":
    if "--check" in sys.argv:
        L, D, U, root = import_reference()
        print("reference importable from", root, "->", L.__file__)
    else:
        print(build(verbose=True))
