"""Install the UNMODIFIED reference into ``baseline/_ref`` (git-ignored, travels with gpurun).

``ensure_installed()`` is called by ``__graft_entry__.build()`` (runs where ``/root/reference``
exists) and lazily by ``bench.py --impl reference``.  Route, in order:

1. the documented command
   ``pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target baseline/_ref /root/reference``
   -- fails on this reference ("neither setup.py nor pyproject.toml"), recorded in
   ``baseline/_ref/INSTALL.json``;
2. the same command with ``--no-deps`` from a copy under /tmp: ``Server/dtds`` plus the two analysis
   scripts, byte for byte, next to a 3-line ``setup.py`` (the reference ships no packaging metadata);
3. a plain byte-identical copy of the same files if pip itself is unusable.

Nothing in the copied tree is edited; ``verify()`` re-checks every file's SHA-256 against
``/root/reference`` whenever that tree is present.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DST = os.path.join(HERE, "_ref")
REF_SRC = os.environ.get("FEDTGAN_REFERENCE_SRC", "/root/reference")
WHEELHOUSE = "/opt/wheelhouse"
SCRIPTS = ("similarity_analysis.py", "utility_analysis.py")


def installed() -> bool:
    return os.path.isfile(os.path.join(REF_DST, "dtds", "distributed.py"))


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as fh:
        for blk in iter(lambda: fh.read(1 << 20), b""):
            h.update(blk)
    return h.hexdigest()


def _py_files(root):
    out = []
    for d, _, files in os.walk(root):
        if "__pycache__" in d:
            continue
        for f in files:
            if f.endswith(".py"):
                out.append(os.path.relpath(os.path.join(d, f), root))
    return sorted(out)


def verify() -> dict:
    """Compare every installed reference .py with its source (when the source tree is present)."""
    src_pkg = os.path.join(REF_SRC, "Server", "dtds")
    if not os.path.isdir(src_pkg):
        return {"checked": False, "reason": f"{REF_SRC} not present on this machine"}
    bad = []
    n = 0
    for rel in _py_files(src_pkg):
        dst = os.path.join(REF_DST, "dtds", rel)
        n += 1
        if not os.path.isfile(dst) or _sha(dst) != _sha(os.path.join(src_pkg, rel)):
            bad.append(rel)
    return {"checked": True, "files": n, "modified_or_missing": bad}


def _stage_copy(tmp):
    """Byte-identical copy of Server/dtds + analysis scripts with minimal packaging metadata."""
    srv = os.path.join(REF_SRC, "Server")
    shutil.copytree(os.path.join(srv, "dtds"), os.path.join(tmp, "dtds"),
                    ignore=shutil.ignore_patterns("__pycache__", "saved_models"))
    # the reference has no __init__.py files: make the directories importable as regular packages so that
    # setuptools picks them up (namespace-package import works without them as well)
    for d, _, _ in os.walk(os.path.join(tmp, "dtds")):
        init = os.path.join(d, "__init__.py")
        if not os.path.exists(init):
            open(init, "w").close()
    for s in SCRIPTS:
        shutil.copy2(os.path.join(srv, s), os.path.join(tmp, s))
    with open(os.path.join(tmp, "setup.py"), "w") as fh:
        fh.write("from setuptools import setup, find_packages\n"
                 "setup(name='fed_tgan_reference', version='0', packages=find_packages(),\n"
                 "      py_modules=['similarity_analysis', 'utility_analysis'])\n")


def ensure_installed(verbose: bool = False) -> bool:
    """Make ``baseline/_ref/dtds`` exist.  Returns True when the reference is importable from there."""
    if installed():
        return True
    if not os.path.isdir(os.path.join(REF_SRC, "Server", "dtds")):
        return False
    log = {"source": REF_SRC, "attempts": []}
    pip = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links", WHEELHOUSE,
           "--target", REF_DST]

    def run(cmd, cwd=None):
        try:
            r = subprocess.run(cmd, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
            log["attempts"].append({"cmd": " ".join(cmd), "rc": r.returncode, "tail": r.stdout[-400:]})
            if verbose:
                print(r.stdout[-800:])
            return r.returncode == 0
        except Exception as exc:  # pip missing / timeout
            log["attempts"].append({"cmd": " ".join(cmd), "rc": None, "tail": repr(exc)[:300]})
            return False

    os.makedirs(REF_DST, exist_ok=True)
    ok = run(pip + [REF_SRC]) and installed()
    if not ok:
        tmp = tempfile.mkdtemp(prefix="ref_pkg_")
        try:
            _stage_copy(tmp)
            ok = run(pip + ["--no-deps", tmp]) and installed()
            if not ok:
                # pip unusable: plain copy of the same staged tree
                shutil.rmtree(os.path.join(REF_DST, "dtds"), ignore_errors=True)
                shutil.copytree(os.path.join(tmp, "dtds"), os.path.join(REF_DST, "dtds"))
                for s in SCRIPTS:
                    shutil.copy2(os.path.join(tmp, s), os.path.join(REF_DST, s))
                log["attempts"].append({"cmd": "copytree (byte-identical)", "rc": 0, "tail": ""})
                ok = installed()
        finally:
            shutil.rmtree(tmp, ignore_errors=True)
    log["installed"] = ok
    log["verify"] = verify() if ok else None
    with open(os.path.join(REF_DST, "INSTALL.json"), "w") as fh:
        json.dump(log, fh, indent=1)
    if ok and log["verify"]["checked"] and log["verify"]["modified_or_missing"]:
        raise RuntimeError(f"reference install differs from {REF_SRC}: {log['verify']['modified_or_missing']}")
    return ok


if __name__ == "__main__":
    ok = ensure_installed(verbose=True)
    print(json.dumps({"installed": ok, "verify": verify() if ok else None}))
    sys.exit(0 if ok else 1)
