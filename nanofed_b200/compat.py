"""Import-name compatibility: make ``import nanofed…`` resolve to this package.

``install_alias()`` registers a meta-path finder that maps ``nanofed`` and every
``nanofed.x.y`` to ``nanofed_b200`` / ``nanofed_b200.x.y`` (the SAME module
objects, so ``isinstance`` and ``mock.patch`` targets agree).  With it, code —
and the reference's own test-suite — written against ``nanofed`` runs unchanged
on this framework::

    import nanofed_b200.compat; nanofed_b200.compat.install_alias()
    from nanofed import Coordinator, HTTPServer          # → nanofed_b200

It refuses to shadow a real ``nanofed`` installation unless ``force=True``.
"""

from __future__ import annotations

import importlib
import importlib.abc
import importlib.machinery
import importlib.util
import sys

_ALIAS = "nanofed"
_TARGET = "nanofed_b200"


class _AliasLoader(importlib.abc.Loader):
    def __init__(self, real_name: str) -> None:
        self._real = real_name

    def create_module(self, spec):  # noqa: D401
        return importlib.import_module(self._real)

    def exec_module(self, module) -> None:
        pass


class _AliasFinder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path=None, target=None):
        if fullname != _ALIAS and not fullname.startswith(_ALIAS + "."):
            return None
        real = _TARGET + fullname[len(_ALIAS):]
        try:
            real_spec = importlib.util.find_spec(real)
        except (ImportError, ValueError):
            return None
        if real_spec is None:
            return None
        spec = importlib.machinery.ModuleSpec(
            fullname, _AliasLoader(real), is_package=real_spec.submodule_search_locations is not None
        )
        return spec


_installed: _AliasFinder | None = None


def install_alias(force: bool = False) -> None:
    global _installed
    if _installed is not None:
        return
    if not force and _ALIAS not in sys.modules:
        try:
            existing = importlib.util.find_spec(_ALIAS)
        except (ImportError, ValueError):
            existing = None
        if existing is not None:
            raise RuntimeError(
                "a real 'nanofed' package is importable; pass force=True to shadow it"
            )
    _installed = _AliasFinder()
    sys.meta_path.insert(0, _installed)


def uninstall_alias() -> None:
    global _installed
    if _installed is not None:
        sys.meta_path.remove(_installed)
        _installed = None
    for name in [m for m in sys.modules if m == _ALIAS or m.startswith(_ALIAS + ".")]:
        del sys.modules[name]
