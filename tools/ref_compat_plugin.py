"""pytest plugin: run the REFERENCE's test-suite against nanofed_b200.

    python -m pytest /root/reference/tests -p tools.ref_compat_plugin \
        --rootdir /tmp/refcompat -c /dev/null -q -p no:cacheprovider

It aliases ``nanofed`` → ``nanofed_b200`` (``nanofed_b200.compat``) and runs
``@pytest.mark.asyncio`` coroutine tests with ``asyncio.run`` (pytest-asyncio is
not installed on this image).  Nothing from the reference is copied into this
repository; the tests are executed from where they live.
"""

import asyncio
import inspect
import sys
from pathlib import Path

import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import nanofed_b200.compat as _compat  # noqa: E402

_compat.install_alias(force=True)


def pytest_configure(config):
    config.addinivalue_line("markers", "asyncio: run coroutine test with asyncio.run")


@pytest.hookimpl(tryfirst=True)
def pytest_pyfunc_call(pyfuncitem):
    fn = pyfuncitem.obj
    if inspect.iscoroutinefunction(fn):
        names = pyfuncitem._fixtureinfo.argnames
        kwargs = {k: pyfuncitem.funcargs[k] for k in names}
        asyncio.run(fn(**kwargs))
        return True
    return None
