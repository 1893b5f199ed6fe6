"""get_Arxy_batch_noalloc (include/teqpb200.h): the no-allocation variant of get_Arxy_batch whose covering table lives in a static device
array instead of a cudaMallocAsync block.  Checked through the C ABI against get_Arxy_batch itself, on two calls in a row with different
inputs (the array keeps the last call's contents), plus its error paths.  Collected last on purpose
(file name)."""
import ctypes as C

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu

PAIRS = [(0, 0), (0, 1), (0, 2), (1, 0), (1, 1), (2, 0), (0, 3), (1, 2)]
RTOL = 1e-12   # the parity rule's relative tolerance (tests/parity.py); both entry points run the same sweeps, so they agree to the bit


def _dp(t):
    return C.cast(C.c_void_p(t.data_ptr()), C.POINTER(C.c_double))


def _noalloc(lib, model, pairs, T, rho, x, out, flags=1):
    NT = (C.c_int * len(pairs))(*[p[0] for p in pairs])
    ND = (C.c_int * len(pairs))(*[p[1] for p in pairs])
    err = C.create_string_buffer(2000)
    rc = lib.get_Arxy_batch_noalloc(model._uuid, C.c_int(len(pairs)), NT, ND, C.c_longlong(T.numel()), _dp(T), _dp(rho), _dp(x),
                                    C.c_int(x.shape[0]), _dp(out), C.c_int(flags), err, C.c_int(2000))
    return rc, err.value.decode(errors="replace")


def _points(name, n, seed):
    import torch
    spec, (Tlo, Thi), (rlo, rhi), z0 = cases.SPECS[name]
    rng = np.random.default_rng(seed)
    T = Tlo + (Thi - Tlo) * rng.random(n)
    rho = np.exp(np.log(rlo) + (np.log(rhi) - np.log(rlo)) * rng.random(n))
    x = np.asarray(z0)[:, None] * np.exp(0.2 * (2 * rng.random((len(z0), n)) - 1))
    x /= x.sum(axis=0)
    return spec, tuple(torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (T, rho, x))


@pytest.mark.parametrize("name", ["PR2k", "GERG10"])
def test_noalloc_matches_get_Arxy_batch_twice(name):
    import torch
    import teqp_b200
    from teqp_b200.model import make_model
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    lib = teqp_b200.lib()
    lib.teqpb200_noalloc_max_doubles.restype = C.c_longlong
    assert lib.teqpb200_noalloc_max_doubles() == 4 << 20
    err = C.create_string_buffer(2000)
    assert lib.teqpb200_noalloc_setup(err, C.c_int(2000)) == 0, err.value
    model = None
    for call, (n, seed) in enumerate([(20000, 11), (7001, 12)]):   # two calls in a row, different sizes and inputs
        spec, (T, rho, x) = _points(name, n, seed)
        if model is None:
            model = make_model(spec)
        ref = model.get_Arxy_batch(PAIRS, T, rho, x)              # existing entry point (also generates the per-model kernels)
        out = torch.full((len(PAIRS), n), float("nan"), dtype=torch.float64, device="cuda")
        rc, msg = _noalloc(lib, model, PAIRS, T, rho, x, out)
        assert rc == 0, msg
        torch.cuda.synchronize()
        a, b = out.cpu().numpy(), ref.cpu().numpy()
        assert np.isfinite(b).all()
        np.testing.assert_allclose(a, b, rtol=RTOL, atol=0.0, err_msg="call %d" % call)


def test_noalloc_rejects_what_it_cannot_serve():
    import torch
    import teqp_b200
    from teqp_b200.model import make_model
    lib = teqp_b200.lib()
    # (1) more scratch than the static array holds: 9 block rows x 600 000 points > 4 Mi doubles
    n = 600000
    spec, (T, rho, x) = _points("PR2k", n, 5)
    model = make_model(spec)
    out = torch.empty((2, n), dtype=torch.float64, device="cuda")
    rc, msg = _noalloc(lib, model, [(2, 2), (1, 1)], T, rho, x, out)
    assert rc != 0 and "exceed the static scratch array" in msg, msg
    # (2) host pointers
    rc, msg = _noalloc(lib, model, [(0, 1)], T, rho, x, out, flags=0)
    assert rc != 0 and "device pointers only" in msg, msg
    # (3) a generated-kernel family whose shape has never been used: an error, not a compile
    spec, (T, rho, x) = _points("GERG2004_3", 1000, 6)
    fresh = make_model(spec)
    out = torch.empty((1, 1000), dtype=torch.float64, device="cuda")
    rc, msg = _noalloc(lib, fresh, [(0, 5)], T, rho, x, out)
    assert rc != 0 and "has not been generated yet" in msg, msg
    torch.cuda.synchronize()
