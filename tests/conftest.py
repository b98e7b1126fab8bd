import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


# RT_B200_TEST_LIB=<path>: run the tests against another build of the library (the checked build of
# tools/build_variant.sh).  Read here, by the test harness; the library itself never reads the environment.
if os.environ.get("RT_B200_TEST_LIB"):
    from rust_tracer_b200 import _lib as _product_lib
    _product_lib.LIB_PATH = os.path.join(ROOT, os.environ["RT_B200_TEST_LIB"])


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # build the product library and the oracle if they are not there yet (seconds)
    import __graft_entry__ as entry
    lib = os.path.join(ROOT, "rust_tracer_b200", "librt_b200.so")
    orc = os.path.join(ROOT, "oracle", "liborc.so")
    if not (os.path.exists(lib) and os.path.exists(orc)):
        entry.build()


def _has_gpu():
    try:
        import rust_tracer_b200 as rt
        return rt.lib().rt_device_count() > 0
    except Exception:
        return False


HAS_GPU = _has_gpu()


def pytest_collection_modifyitems(config, items):
    if HAS_GPU:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def default_desc():
    import rust_tracer_b200 as rt
    return rt.SceneDesc.builtin("default")
