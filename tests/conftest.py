import importlib.util
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu)")


def load_pkg():
    """Import depth-conversions_b200/ (hyphenated directory) as a module."""
    name = "depth_conversions_b200"
    if name in sys.modules:
        return sys.modules[name]
    path = os.path.join(ROOT, "depth-conversions_b200", "__init__.py")
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="session")
def pkg():
    return load_pkg()


@pytest.fixture(scope="session")
def oracle():
    from oracle_lib import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def gpu_ctx(pkg):
    """A context on cuda:0. Fails loudly (no skip, no fallback) without one."""
    ctx = pkg.Context(0)
    yield ctx
    ctx.close()
