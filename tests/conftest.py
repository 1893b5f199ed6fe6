import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, HERE):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_ref
    if not oracle_ref.available():
        pytest.skip("oracle/_ref/libteqp_ref_oracle.so not built (make -C oracle ref)")
    return oracle_ref


@pytest.fixture(scope="session")
def emul():
    import emul_lib
    if not emul_lib.available():
        pytest.skip("tests/emul/libtq_emul.so not built (make -C tests/emul)")
    return emul_lib
