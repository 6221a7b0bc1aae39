import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def orc():
    from oracle import oracle as O
    return O.Oracle()


@pytest.fixture(scope="session")
def reflib():
    """The reference's own kernels compiled in place (oracle/_ref). Present in the build container and, as
    prebuilt .so files, on the GPU box."""
    from oracle import oracle as O
    try:
        return O.RefLib()
    except (FileNotFoundError, OSError) as e:  # pragma: no cover
        pytest.skip(f"oracle/_ref not available: {e}")


@pytest.fixture(scope="session")
def ctx():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import vlut_b200  # noqa: F401
    from vlut_b200 import api
    c = api.Context(0)
    yield c
    c.close()


@pytest.fixture
def knobs():
    """Set VLUT_* environment knobs for one test.  The library resolves its knobs once (never on the launch path), so the
    fixture makes it re-read them after every change and again after restoring the environment."""
    import os
    import vlut_b200  # noqa: F401
    from vlut_b200 import api
    saved = {}

    def set_(**kv):
        for k, v in kv.items():
            if k not in saved:
                saved[k] = os.environ.get(k)
            os.environ[k] = str(v)
        api.reload_env()

    yield set_
    for k, v in saved.items():
        if v is None:
            os.environ.pop(k, None)
        else:
            os.environ[k] = v
    api.reload_env()
