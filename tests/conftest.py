import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")
    config.addinivalue_line("markers", "ref: needs oracle/_ref/libprt_ref.so (built where /root/reference exists)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    return oracle_lib.load_oracle()


@pytest.fixture(scope="session")
def ref():
    import oracle_lib
    lib = oracle_lib.load_ref()
    if lib is None:
        pytest.skip("oracle/_ref/libprt_ref.so not built (needs /root/reference; run `make -C oracle ref`)")
    return lib


def pytest_sessionfinish(session, exitstatus):
    """PRT_CHECKED=1 python -m pytest tests -m gpu runs every GPU test through the bounds-checked library; no check may have fired
    in this process by the end (the provoked overflow of test_checked_build.py happens in a subprocess of its own)."""
    if os.environ.get("PRT_CHECKED") != "1":
        return
    try:
        import torch
        if not torch.cuda.is_available():
            return
        from parallelraytracing_b200.capi import Context
        with Context(0) as c:
            flags = c.check_flags()
    except Exception as e:                                            # no library / no device: nothing to report
        print("\n[checked build] flags not read:", e)
        return
    print("\n[checked build] check flags at session end: 0x%x" % flags)
    if flags != 0x40000000:
        session.exitstatus = 1
