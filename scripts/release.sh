#!/usr/bin/env bash
# Cut a release: build + CPU tests, prepare_release.py, commit, tag vX.Y.Z.
#   scripts/release.sh 0.2.0 [--push]
# The tag triggers .github/workflows/release.yml (sdist/wheel with the sm_100a library built in).
set -euo pipefail
version="${1:?usage: scripts/release.sh X.Y.Z [--push]}"
push="${2:-}"
cd "$(dirname "$0")/.."
branch="$(git rev-parse --abbrev-ref HEAD)"
if [ "$branch" != "main" ]; then echo "releases are cut from main (on $branch)"; exit 1; fi
python -m nanofed_b200.build
python -m pytest tests -q -m "not gpu"
python scripts/prepare_release.py "$version"
git commit -m "release: v$version"
git tag -a "v$version" -m "nanofed_b200 $version"
if [ "$push" = "--push" ]; then
  git push origin main "v$version"
else
  echo "tagged v$version locally; push with: git push origin main v$version"
fi
