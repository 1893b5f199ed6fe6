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


def pytest_terminal_summary(terminalreporter, exitstatus, config):
    """per-output report of the parity rule: how many points needed the absolute floor, and the largest error in units of the scale S"""
    import json
    import parity
    if not parity.REPORT:
        return
    tr = terminalreporter
    tr.write_line("")
    tr.write_line("parity report: %d checks; outputs that needed the absolute floor at values that are NOT small against their family" % len(parity.REPORT))
    tr.write_line("  (fraction among |ref| >= %.0e S | limit | fraction over all points | max |d|/S):" % parity.SIGNIFICANT)
    agg = {}
    for name, n, ff, in_s, lim, fs in parity.REPORT:
        a = agg.setdefault(name, [0, 0.0, 0.0, lim, 0.0])
        a[0] += n; a[1] = max(a[1], ff); a[2] = max(a[2], in_s); a[3] = min(a[3], lim); a[4] = max(a[4], fs)
    shown = 0
    for name, (n, ff, in_s, lim, fs) in sorted(agg.items(), key=lambda kv: (-kv[1][4], -kv[1][1])):
        if (fs > 0 or shown < 12) and shown < 40:
            tr.write_line("  %-48s %7.4f | %5.2f | %7.4f | %9.2e" % (name[:48], fs, lim, ff, in_s)); shown += 1
    worst = max(parity.REPORT, key=lambda r: r[3])
    tr.write_line("  largest error in units of S: %.2e (%s); FTOL = %.1e" % (worst[3], worst[0], parity.FTOL))
    try:
        out = os.path.join(ROOT, "gpurun_out")
        os.makedirs(out, exist_ok=True)
        json.dump([dict(name=k, points=v[0], floor_frac=v[1], max_err_over_S=v[2], limit=v[3], floor_frac_significant=v[4]) for k, v in agg.items()],
                  open(os.path.join(out, "parity_report.json"), "w"), indent=0)
    except Exception:
        pass
