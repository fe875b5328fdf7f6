#!/usr/bin/env python
"""Generate a changelog section from conventional-commit style git history.

    python scripts/changelog.py [--since TAG] [--version X.Y.Z] [--output CHANGELOG.md]

Commits are grouped by their prefix (`feat:`, `fix:`, `perf:`, `docs:`, `test:`, other);
the section is printed or prepended to the output file.  (Counterpart of the reference's
release tooling, `scripts/changelog.py` there; written independently.)
"""
import argparse
import datetime as dt
import re
import subprocess
from collections import OrderedDict
from pathlib import Path

GROUPS = OrderedDict([("feat", "Features"), ("fix", "Fixes"), ("perf", "Performance"),
                      ("docs", "Documentation"), ("test", "Tests"), ("other", "Other changes")])


def commits(since):
    rng = f"{since}..HEAD" if since else "HEAD"
    out = subprocess.run(["git", "log", "--pretty=%h%x09%s", rng], capture_output=True, text=True,
                         check=True).stdout
    return [line.split("\t", 1) for line in out.splitlines() if "\t" in line]


def classify(subject):
    m = re.match(r"^(\w+)(\([^)]*\))?!?:\s*(.+)$", subject)
    if m and m.group(1).lower() in GROUPS:
        return m.group(1).lower(), m.group(3)
    low = subject.lower()
    for key in ("fix", "perf", "doc", "test"):
        if low.startswith(key):
            return {"doc": "docs"}.get(key, key), subject
    return "other", subject


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--since")
    ap.add_argument("--version", default="Unreleased")
    ap.add_argument("--output")
    a = ap.parse_args()
    buckets = {k: [] for k in GROUPS}
    for sha, subject in commits(a.since):
        kind, text = classify(subject)
        buckets[kind].append(f"- {text} ({sha})")
    lines = [f"## {a.version} — {dt.date.today().isoformat()}", ""]
    for key, title in GROUPS.items():
        if buckets[key]:
            lines += [f"### {title}", *buckets[key], ""]
    section = "\n".join(lines)
    if a.output:
        path = Path(a.output)
        old = path.read_text() if path.exists() else "# Changelog\n\n"
        head, _, rest = old.partition("\n\n")
        path.write_text(f"{head}\n\n{section}\n{rest}")
    else:
        print(section)


if __name__ == "__main__":
    main()
