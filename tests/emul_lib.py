"""ctypes front-end of tests/emul/libtq_emul.so (host replay of the kernels' per-point functions; TEST HARNESS ONLY)."""
import ctypes as C
import json
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "emul", "libtq_emul.so")
_dp = C.POINTER(C.c_double)
_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(LIB_PATH)
    return _lib


def _p(a):
    return a.ctypes.data_as(_dp) if a is not None else None


def ar_table(spec, NT, ND, T, rho, x, mode=0):
    T = np.ascontiguousarray(T, dtype=np.float64); rho = np.ascontiguousarray(rho, dtype=np.float64)
    x = np.ascontiguousarray(x, dtype=np.float64)
    n = T.size
    out = np.empty(((NT + 1) * (ND + 1), n))
    err = C.create_string_buffer(1000)
    rc = lib().emul_ar_table(json.dumps(spec).encode(), C.c_int(mode), C.c_int(NT), C.c_int(ND), C.c_longlong(n), _p(T), _p(rho), _p(x),
                             C.c_int(x.shape[0]), C.c_int(0), _p(out), err, C.c_int(1000))
    if rc:
        raise RuntimeError(err.value.decode())
    return out.reshape(NT + 1, ND + 1, n)


def ATrhoXi_table(spec, T, rho, x):
    T = np.ascontiguousarray(T, dtype=np.float64); rho = np.ascontiguousarray(rho, dtype=np.float64)
    x = np.ascontiguousarray(x, dtype=np.float64)
    n = T.size; nc = x.shape[0]
    out = np.empty((nc * 36, n))
    err = C.create_string_buffer(1000)
    rc = lib().emul_ATrhoXi_table(json.dumps(spec).encode(), C.c_longlong(n), _p(T), _p(rho), _p(x), C.c_int(nc), _p(out), err, C.c_int(1000))
    if rc:
        raise RuntimeError(err.value.decode())
    return out.reshape(nc, 3, 3, 4, n)


def cross_table(spec, idx, T, rho, x):
    """idx = (i, j) -> [2][2][3][3][n]; idx = (i, j, k) -> [2][2][2][2][2][n]"""
    T = np.ascontiguousarray(T, dtype=np.float64); rho = np.ascontiguousarray(rho, dtype=np.float64); x = np.ascontiguousarray(x, dtype=np.float64)
    n = T.size; nc = x.shape[0]
    which = len(idx)
    out = np.empty((36 if which == 2 else 32, n))
    err = C.create_string_buffer(1000)
    i, j = idx[0], idx[1]; k = idx[2] if which == 3 else -1
    rc = lib().emul_cross_table(json.dumps(spec).encode(), C.c_int(which), C.c_int(i), C.c_int(j), C.c_int(k), C.c_longlong(n), _p(T), _p(rho), _p(x),
                                C.c_int(nc), _p(out), err, C.c_int(1000))
    if rc:
        raise RuntimeError(err.value.decode())
    return out.reshape((2, 2, 3, 3, n) if which == 2 else (2, 2, 2, 2, 2, n))


ISOPROPS = ["pr", "splus", "dpdT", "d2PsirdTdrhoi", "dpdrhovec", "pmv", "chempotVLE", "dchempotdT", "dlnphi_dT", "dlnphi_drho", "dlnphi_dv", "dlnphi_dx"]


def isoprops(spec, T, rhovec):
    """dict of the 12 derived isochoric properties (IsoPropsOut)"""
    T = np.ascontiguousarray(T, dtype=np.float64); rv = np.ascontiguousarray(rhovec, dtype=np.float64)
    n = T.size; nc = rv.shape[0]
    rows = [1, 1, 1] + [nc] * 8 + [nc * nc]
    outs = [np.empty((r, n)) for r in rows]
    arr = (C.POINTER(C.c_double) * 12)(*[_p(o) for o in outs])
    err = C.create_string_buffer(1000)
    rc = lib().emul_isoprops(json.dumps(spec).encode(), C.c_longlong(n), _p(T), _p(rv), C.c_int(nc), arr, err, C.c_int(1000))
    if rc:
        raise RuntimeError(err.value.decode())
    return {k: (o[0] if r == 1 else o) for k, o, r in zip(ISOPROPS, outs, rows)}


def isochoric(spec, T, rhovec, hess=True):
    T = np.ascontiguousarray(T, dtype=np.float64); rv = np.ascontiguousarray(rhovec, dtype=np.float64)
    n = T.size; nc = rv.shape[0]
    f = np.empty(n); g = np.empty((nc, n)); H = np.empty((nc * nc, n)) if hess else None
    lnphi = np.empty((nc, n)); phi = np.empty((nc, n))
    err = C.create_string_buffer(1000)
    rc = lib().emul_isochoric(json.dumps(spec).encode(), C.c_longlong(n), _p(T), _p(rv), C.c_int(nc), _p(f), _p(g), _p(H), _p(lnphi), _p(phi),
                              err, C.c_int(1000))
    if rc:
        raise RuntimeError(err.value.decode())
    return f, g, H, lnphi, phi


# ---- generated per-model kernels (csrc/tq_gerg_codegen.cu): the SAME translation unit NVRTC compiles for the GPU, compiled here with
# g++ and run as a host loop, so the CPU tier checks the generated mathematics against the oracle ----------------------------------
_JIT_CACHE = {}


def virial_table(spec, NT, ND, T, x):
    """[NT+1][ND+1][n]: (1/T)^n d^(n+m) alphar / d(1/T)^n drho^m at rho = 0 (unit-seeded density jet)"""
    T = np.ascontiguousarray(T, dtype=np.float64); x = np.ascontiguousarray(x, dtype=np.float64)
    n = T.size; bc = x.ndim == 1
    out = np.empty(((NT + 1) * (ND + 1), n))
    err = C.create_string_buffer(1000)
    rc = lib().emul_extras(json.dumps(spec).encode(), C.c_int(0), C.c_int(NT), C.c_int(ND), C.c_longlong(n), _p(T), _p(x), None, C.c_int(x.shape[0]),
                           C.c_int(1 if bc else 0), _p(out), err, C.c_int(1000))
    if rc:
        raise RuntimeError(err.value.decode())
    return out.reshape(NT + 1, ND + 1, n)


def Ataudelta_table(spec, tau, delta, x):
    tau = np.ascontiguousarray(tau, dtype=np.float64); delta = np.ascontiguousarray(delta, dtype=np.float64); x = np.ascontiguousarray(x, dtype=np.float64)
    n = tau.size; nc = x.shape[0]
    out = np.empty((nc * 36, n))
    err = C.create_string_buffer(1000)
    rc = lib().emul_Ataudelta(json.dumps(spec).encode(), C.c_int(1), C.c_int(0), C.c_int(0), C.c_int(0), C.c_longlong(n), _p(tau), _p(delta), _p(x), C.c_int(nc),
                              _p(out), err, C.c_int(1000))
    if rc:
        raise RuntimeError(err.value.decode())
    return out.reshape(nc, 3, 3, 4, n)


def dmBnvirdTm(spec, Nvir, NT, T, x):
    """the combination libteqpb200's virial_call forms from the table (csrc/tq_capi.cu)"""
    import math
    tab = virial_table(spec, NT, Nvir - 1, T, x)
    D = tab[:, Nvir - 1]
    f = math.factorial(Nvir - 2)
    T = np.asarray(T, dtype=np.float64)
    return (D[0] if NT == 0 else -D[1] / T if NT == 1 else (D[2] + 2 * D[1]) / T**2) / f


def sigma_derivs(spec, T, rv, v):
    T = np.ascontiguousarray(T, dtype=np.float64); rv = np.ascontiguousarray(rv, dtype=np.float64); v = np.ascontiguousarray(v, dtype=np.float64)
    n = T.size
    out = np.empty((5, n))
    err = C.create_string_buffer(1000)
    rc = lib().emul_extras(json.dumps(spec).encode(), C.c_int(1), C.c_int(0), C.c_int(0), C.c_longlong(n), _p(T), _p(rv), _p(v), C.c_int(rv.shape[0]),
                           C.c_int(0), _p(out), err, C.c_int(1000))
    if rc:
        raise RuntimeError(err.value.decode())
    return out


def jit_ar_table(spec, NT, ND, T, rho, x):
    import hashlib
    import subprocess
    import tempfile
    import teqp_b200
    key = (json.dumps(spec, sort_keys=True), NT, ND)
    if key not in _JIT_CACHE:
        m = teqp_b200.make_model(spec)
        L = teqp_b200.lib()

        def fetch(which):
            n = C.c_ulonglong(0); err = C.create_string_buffer(2000)
            rc = L.teqpb200_generated_kernel(m._uuid, C.c_int(NT), C.c_int(ND), C.c_int(which), None, C.byref(n), err, C.c_int(2000))
            if rc:
                raise RuntimeError(err.value.decode())
            buf = C.create_string_buffer(n.value)
            rc = L.teqpb200_generated_kernel(m._uuid, C.c_int(NT), C.c_int(ND), C.c_int(which), buf, C.byref(n), err, C.c_int(2000))
            if rc:
                raise RuntimeError(err.value.decode())
            return buf.raw[:n.value]
        src = fetch(0).rstrip(b"\0")
        consts = np.frombuffer(fetch(1), dtype=np.float64).copy()
        d = os.path.join(tempfile.gettempdir(), "tqj_emul")
        os.makedirs(d, exist_ok=True)
        h = hashlib.sha1(src).hexdigest()[:16]
        so = os.path.join(d, h + ".so")
        if not os.path.exists(so):
            cpp = os.path.join(d, h + ".cpp")
            open(cpp, "wb").write(src)
            subprocess.run(["g++", "-std=c++17", "-O1", "-mfma", "-fPIC", "-shared", "-w", cpp, "-o", so + ".tmp"], check=True)
            os.replace(so + ".tmp", so)
        _JIT_CACHE[key] = (C.CDLL(so), consts)
    lib_, consts = _JIT_CACHE[key]
    T = np.ascontiguousarray(T, dtype=np.float64); rho = np.ascontiguousarray(rho, dtype=np.float64); x = np.ascontiguousarray(x, dtype=np.float64)
    n = T.size
    bc = x.ndim == 1
    out = np.empty(((NT + 1) * (ND + 1), n))
    lib_.tqj_host(_p(consts), C.c_longlong(n), _p(T), _p(rho), _p(x), C.c_longlong(1 if bc else n), C.c_int(1 if bc else 0), _p(out))
    return out.reshape(NT + 1, ND + 1, n)


def pure_vle_update(A, rho2, done):
    n = done.size
    A = np.ascontiguousarray(A, dtype=np.float64)
    assert A.shape == (3, 2 * n) and rho2.shape == (2 * n,)
    lib().emul_pure_vle_update(C.c_longlong(n), _p(A), _p(rho2), _p(done))


def mix_vle_update(R, T, xspec, f, g, H, tols, last_iter, rv2, code):
    n = code.size
    f, g, H, xspec, T = (np.ascontiguousarray(a, dtype=np.float64) for a in (f, g, H, xspec, T))
    assert f.shape == (2 * n,) and g.shape == (2, 2 * n) and H.shape == (4, 2 * n) and rv2.shape == (2, 2 * n) and xspec.shape == (2, n)
    lib().emul_mix_vle_update(C.c_longlong(n), C.c_double(R), _p(T), _p(xspec), _p(f), _p(g), _p(H), *(C.c_double(t) for t in tols), C.c_int(int(last_iter)),
                              _p(rv2), _p(code))


def mix_vle_tp_update(R, T, p, f, g, H, rv2):
    n = T.size
    f, g, H, T, p = (np.ascontiguousarray(a, dtype=np.float64) for a in (f, g, H, T, p))
    assert f.shape == (2 * n,) and g.shape == (2, 2 * n) and H.shape == (4, 2 * n) and rv2.shape == (2, 2 * n)
    lib().emul_mix_vle_tp_update(C.c_longlong(n), C.c_double(R), _p(T), _p(p), _p(f), _p(g), _p(H), _p(rv2))
