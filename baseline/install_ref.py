"""Put the UNMODIFIED reference package under ``baseline/_ref`` (git-ignored, shipped to GPU boxes).

Order of attempts, first success wins:

1. ``pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target
   baseline/_ref <copy of /root/reference under /tmp>`` (the source tree is read-only, so pip
   builds from a copy), then the same with ``--no-deps``.
2. If the build backend named in the reference's ``pyproject.toml`` (``poetry-core``) is not
   installable offline, a byte-for-byte copy of the pure-Python package directory
   ``/root/reference/nanofed`` → ``baseline/_ref/nanofed``.  For a pure-Python project this is
   exactly the set of files a wheel install would have produced.

In every case a manifest (``baseline/_ref/MANIFEST.sha256.json``) records the sha256 of every
file of the installed package, the sha256 of the same file in the source tree, and how the
install was done.  ``verify()`` recomputes the hashes; ``baseline/reference_arm.py`` refuses to
run a package whose files differ from the manifest and prints ``ref_source`` / ``ref_sha``.

No reference file is ever edited.
"""

from __future__ import annotations

import hashlib
import json
import os
import shutil
import subprocess
import sys
import tempfile
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF_DIR = HERE / "_ref"
MANIFEST = REF_DIR / "MANIFEST.sha256.json"
SOURCE = Path(os.environ.get("NANOFED_REFERENCE_SRC", "/root/reference"))
WHEELHOUSE = Path(os.environ.get("NANOFED_WHEELHOUSE", "/opt/wheelhouse"))


def _sha(path: Path) -> str:
    h = hashlib.sha256()
    with open(path, "rb") as fh:
        for block in iter(lambda: fh.read(1 << 20), b""):
            h.update(block)
    return h.hexdigest()


def _package_files(pkg: Path) -> list[Path]:
    return sorted(p for p in pkg.rglob("*")
                  if p.is_file() and "__pycache__" not in p.parts and p.suffix != ".pyc")


def _tree_sha(files: dict[str, str]) -> str:
    h = hashlib.sha256()
    for name in sorted(files):
        h.update(name.encode())
        h.update(files[name].encode())
    return h.hexdigest()


def _pip_install(extra: list[str]) -> tuple[bool, str]:
    with tempfile.TemporaryDirectory(prefix="nanofed_ref_src_") as tmp:
        src = Path(tmp) / "reference"
        shutil.copytree(SOURCE, src, symlinks=True,
                        ignore=shutil.ignore_patterns(".git", "__pycache__", "*.pyc"))
        cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation",
               "--find-links", str(WHEELHOUSE), "--target", str(REF_DIR), *extra, str(src)]
        try:
            done = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
        except Exception as exc:  # pip missing, timeout, …
            return False, repr(exc)[:300]
        tail = (done.stderr or done.stdout).strip().splitlines()[-1:] or [""]
        return done.returncode == 0 and (REF_DIR / "nanofed" / "__init__.py").exists(), tail[0][:300]


def _write_manifest(how: str, notes: list[str]) -> dict:
    installed = {str(p.relative_to(REF_DIR)): _sha(p) for p in _package_files(REF_DIR / "nanofed")}
    source = {}
    if (SOURCE / "nanofed").is_dir():
        source = {str(p.relative_to(SOURCE)): _sha(p) for p in _package_files(SOURCE / "nanofed")}
    manifest = {
        "how": how, "notes": notes, "source": str(SOURCE),
        "identical_to_source": bool(source) and installed == source,
        "tree_sha256": _tree_sha(installed), "files": installed,
    }
    MANIFEST.write_text(json.dumps(manifest, indent=1, sort_keys=True))
    return manifest


def verify() -> dict:
    """Recompute the hashes of ``baseline/_ref/nanofed``; raise if anything differs."""
    if not MANIFEST.exists():
        raise FileNotFoundError(f"{MANIFEST} missing: run `python -m baseline.install_ref`")
    manifest = json.loads(MANIFEST.read_text())
    now = {str(p.relative_to(REF_DIR)): _sha(p) for p in _package_files(REF_DIR / "nanofed")}
    if now != manifest["files"]:
        changed = sorted(set(now.items()) ^ set(manifest["files"].items()))[:3]
        raise RuntimeError(f"baseline/_ref/nanofed differs from its manifest: {changed}")
    return manifest


def install(force: bool = False) -> dict:
    """Idempotent: returns the manifest of a verified install, installing first if needed."""
    if not force and MANIFEST.exists():
        try:
            return verify()
        except Exception:
            pass
    if not (SOURCE / "nanofed" / "__init__.py").exists():
        raise FileNotFoundError(f"reference source tree not found at {SOURCE}")
    if REF_DIR.exists():
        shutil.rmtree(REF_DIR)
    REF_DIR.mkdir(parents=True)
    notes: list[str] = []
    for extra in ([], ["--no-deps"]):
        ok, msg = _pip_install(extra)
        label = "pip install --target" + (" --no-deps" if extra else "")
        if ok:
            return _write_manifest(label, notes)
        notes.append(f"{label} failed: {msg}")
        shutil.rmtree(REF_DIR, ignore_errors=True)
        REF_DIR.mkdir(parents=True)
    shutil.copytree(SOURCE / "nanofed", REF_DIR / "nanofed",
                    ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    return _write_manifest("verbatim copy of the pure-Python package directory", notes)


def main() -> None:
    m = install(force="--force" in sys.argv)
    print(json.dumps({k: m[k] for k in ("how", "notes", "identical_to_source", "tree_sha256")},
                     indent=1))


if __name__ == "__main__":
    main()
