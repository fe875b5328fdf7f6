#!/usr/bin/env python
"""Prepare a release: bump the version, refresh the changelog, stage the files.

    python scripts/prepare_release.py 0.2.0 [--dry-run]

Steps: (1) check the version string and that the tree is clean, (2) write the version into
``pyproject.toml`` and ``nanofed_b200/__init__.py``, (3) run ``scripts/changelog.py`` since the
last tag, (4) ``git add`` the touched files.  Tagging and pushing are left to ``scripts/release.sh``.
(The reference ships a script of the same name; this one was written for this repository's layout.)
"""
import argparse
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
VERSION_RE = re.compile(r"^\d+\.\d+\.\d+([abrc]+\d+)?$")


def sh(*cmd, check=True):
    return subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, check=check).stdout.strip()


def replace_once(path: Path, pattern: str, repl: str, dry: bool) -> None:
    text = path.read_text()
    new, n = re.subn(pattern, repl, text, count=1, flags=re.M)
    if n != 1:
        sys.exit(f"{path}: version field not found")
    if not dry:
        path.write_text(new)
    print(f"{'would update' if dry else 'updated'} {path.relative_to(ROOT)}")


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("version")
    ap.add_argument("--dry-run", action="store_true")
    ap.add_argument("--allow-dirty", action="store_true")
    a = ap.parse_args()
    if not VERSION_RE.match(a.version):
        sys.exit(f"'{a.version}' is not a MAJOR.MINOR.PATCH version")
    if sh("git", "status", "--porcelain") and not a.allow_dirty:
        sys.exit("working tree is not clean (commit or pass --allow-dirty)")
    replace_once(ROOT / "pyproject.toml", r'^version\s*=\s*"[^"]+"', f'version = "{a.version}"',
                 a.dry_run)
    replace_once(ROOT / "nanofed_b200" / "__init__.py", r'^__version__\s*=\s*"[^"]+"',
                 f'__version__ = "{a.version}"', a.dry_run)
    last_tag = sh("git", "describe", "--tags", "--abbrev=0", check=False)
    cmd = [sys.executable, "scripts/changelog.py", "--version", a.version, "--output", "CHANGELOG.md"]
    if last_tag:
        cmd += ["--since", last_tag]
    if a.dry_run:
        print("would run:", " ".join(cmd))
        return
    subprocess.run(cmd, cwd=ROOT, check=True)
    sh("git", "add", "pyproject.toml", "nanofed_b200/__init__.py", "CHANGELOG.md")
    print(f"staged release {a.version}; review with `git diff --cached`, then scripts/release.sh")


if __name__ == "__main__":
    main()
