"""Install the UNMODIFIED reference package into ``baseline/_ref`` (git-ignored, travels to the GPU
box with the snapshot).  TEST / BENCH INFRASTRUCTURE -- see oracle/__init__.py.

    python -m pip install --no-index --no-build-isolation --no-deps --target baseline/_ref <copy>

run on a copy of ``/root/reference`` under /tmp (the tree is read-only and setup.py writes
``vllm_hymeta/_version.py``).  The reference's setup.py imports ``setuptools_scm`` at the top
(setup.py:7), which this image does not have; a three-line stand-in whose ``get_version`` raises
``LookupError`` is put on PYTHONPATH for the install only, which is the path setup.py:10-13 already
handles (VERSION = "0.0.1").  No reference file is edited and nothing is copied into tracked paths.

Used by (and only by): ``tests/test_gpu_live_libs.py`` (the reference's own Triton decode kernel and
TorchScript merge as live checkers on the B200 box) and ``bench.py --impl reference`` (the
reference's hf_7B_model torch modules timed on the host cores).  Everything that uses it skips or
falls back to the oracle port when ``baseline/_ref`` is absent.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SRC = "/root/reference"
TARGET = os.path.join(ROOT, "baseline", "_ref")


def installed() -> bool:
    return os.path.isfile(os.path.join(TARGET, "vllm_hymeta", "models", "my_fused_recurrent.py"))


def install(force: bool = False) -> bool:
    """Returns True when baseline/_ref holds the reference afterwards."""
    if installed() and not force:
        return True
    if not os.path.isdir(REF_SRC):
        return False
    tmp = tempfile.mkdtemp(prefix="hymeta_ref_")
    try:
        src = os.path.join(tmp, "reference")
        shutil.copytree(REF_SRC, src, ignore=shutil.ignore_patterns(".git"))
        stub = os.path.join(tmp, "stub", "setuptools_scm")
        os.makedirs(stub)
        with open(os.path.join(stub, "__init__.py"), "w") as f:
            f.write("def get_version(*a, **k):\n"
                    "    raise LookupError('setuptools_scm is not installed in this image')\n")
        env = dict(os.environ)
        env["PYTHONPATH"] = os.path.dirname(stub) + os.pathsep + env.get("PYTHONPATH", "")
        if os.path.isdir(TARGET):
            shutil.rmtree(TARGET)
        cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--no-deps",
               "--find-links", "/opt/wheelhouse", "--target", TARGET, src]
        r = subprocess.run(cmd, env=env, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write("reference install failed:\n" + r.stdout[-2000:] + r.stderr[-2000:])
            return False
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return installed()


def load_ref_file(relpath: str, name: str):
    """Import ONE file of the installed reference by path (its package ``__init__`` imports vLLM
    0.10.0 symbols that the installed vLLM no longer has -- SURVEY F4)."""
    import importlib.util
    path = os.path.join(TARGET, relpath)
    if not os.path.isfile(path):
        return None
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    ok = install(force="--force" in sys.argv)
    print("baseline/_ref", "installed" if ok else "NOT installed")
    sys.exit(0 if ok else 1)
