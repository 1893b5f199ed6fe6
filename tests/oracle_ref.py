"""ctypes front-end of the parity oracle (oracle/_ref/libteqp_ref_oracle.so, built by oracle/Makefile from the reference's own
sources plus the clean-room AD shim).  TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the CPU legs
of bench.py -- never by the teqp_b200 package."""
import ctypes as C
import json
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "..", "oracle", "_ref", "libteqp_ref_oracle.so")
_lib = None
_dp = C.POINTER(C.c_double)


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(LIB_PATH)
        _lib.ref_build_model.restype = C.c_void_p
        _lib.ref_build_model.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        _lib.ref_free_model.argtypes = [C.c_void_p]
        _lib.ref_hardware_threads.restype = C.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(_dp)


def _soa(x, npts):
    x = np.ascontiguousarray(x, dtype=np.float64)
    assert x.ndim == 2 and x.shape[1] == npts, x.shape
    return x


class RefModel:
    """One model of the reference, evaluated point by point on host threads."""

    def __init__(self, spec):
        self.spec = spec if isinstance(spec, str) else json.dumps(spec)
        err = C.create_string_buffer(2000)
        self.h = lib().ref_build_model(self.spec.encode(), err, 2000)
        if not self.h:
            raise RuntimeError("oracle: " + err.value.decode())
        self.nthreads = os.cpu_count() or 1

    def __del__(self):
        if getattr(self, "h", None):
            lib().ref_free_model(C.c_void_p(self.h))
            self.h = None

    def _call(self, fn, *args):
        err = C.create_string_buffer(2000)
        rc = fn(C.c_void_p(self.h), *args, C.c_int(self.nthreads), err, C.c_int(2000))
        if rc != 0:
            raise RuntimeError("oracle rc=%d: %s" % (rc, err.value.decode()))

    def get_R(self, z):
        z = np.ascontiguousarray(z, dtype=np.float64)
        out = C.c_double()
        err = C.create_string_buffer(2000)
        rc = lib().ref_get_R(C.c_void_p(self.h), _p(z), C.c_int(len(z)), C.byref(out), err, C.c_int(2000))
        if rc != 0:
            raise RuntimeError(err.value.decode())
        return out.value

    def Ar0n(self, nmax, T, rho, x):
        T = np.ascontiguousarray(T, dtype=np.float64); rho = np.ascontiguousarray(rho, dtype=np.float64)
        n = T.size; x = _soa(x, n)
        out = np.empty((nmax + 1, n))
        self._call(lib().ref_get_Ar0n, C.c_int(nmax), C.c_longlong(n), _p(T), _p(rho), _p(x), C.c_int(x.shape[0]), _p(out))
        return out

    def Arn0(self, nmax, T, rho, x):
        T = np.ascontiguousarray(T, dtype=np.float64); rho = np.ascontiguousarray(rho, dtype=np.float64)
        n = T.size; x = _soa(x, n)
        out = np.empty((nmax + 1, n))
        self._call(lib().ref_get_Arn0, C.c_int(nmax), C.c_longlong(n), _p(T), _p(rho), _p(x), C.c_int(x.shape[0]), _p(out))
        return out

    def ArNM(self, NT, ND, T, rho, x):
        """compile-time dispatched get_Ar{NT}{ND} (NT 0..2, ND 0..4)"""
        T = np.ascontiguousarray(T, dtype=np.float64); rho = np.ascontiguousarray(rho, dtype=np.float64)
        n = T.size; x = _soa(x, n)
        out = np.empty(n)
        self._call(lib().ref_get_ArNM, C.c_int(NT), C.c_int(ND), C.c_longlong(n), _p(T), _p(rho), _p(x), C.c_int(x.shape[0]), _p(out))
        return out

    def Arxy(self, NT, ND, T, rho, x):
        T = np.ascontiguousarray(T, dtype=np.float64); rho = np.ascontiguousarray(rho, dtype=np.float64)
        n = T.size; x = _soa(x, n)
        out = np.empty(n)
        self._call(lib().ref_get_Arxy, C.c_int(NT), C.c_int(ND), C.c_longlong(n), _p(T), _p(rho), _p(x), C.c_int(x.shape[0]), _p(out))
        return out

    def ATrhoXi(self, NT, ND, i, NXi, T, rho, x):
        T = np.ascontiguousarray(T, dtype=np.float64); rho = np.ascontiguousarray(rho, dtype=np.float64)
        n = T.size; x = _soa(x, n)
        out = np.empty(n)
        self._call(lib().ref_get_ATrhoXi, C.c_int(NT), C.c_int(ND), C.c_int(i), C.c_int(NXi), C.c_longlong(n), _p(T), _p(rho), _p(x),
                   C.c_int(x.shape[0]), _p(out))
        return out

    def ATrhoXiXj(self, NT, ND, i, NXi, j, NXj, T, rho, x):
        T = np.ascontiguousarray(T, dtype=np.float64); rho = np.ascontiguousarray(rho, dtype=np.float64)
        n = T.size; x = _soa(x, n)
        out = np.empty(n)
        self._call(lib().ref_get_ATrhoXiXj, C.c_int(NT), C.c_int(ND), C.c_int(i), C.c_int(NXi), C.c_int(j), C.c_int(NXj), C.c_longlong(n),
                   _p(T), _p(rho), _p(x), C.c_int(x.shape[0]), _p(out))
        return out

    def ATrhoXiXjXk(self, NT, ND, i, NXi, j, NXj, k, NXk, T, rho, x):
        T = np.ascontiguousarray(T, dtype=np.float64); rho = np.ascontiguousarray(rho, dtype=np.float64)
        n = T.size; x = _soa(x, n)
        out = np.empty(n)
        self._call(lib().ref_get_ATrhoXiXjXk, C.c_int(NT), C.c_int(ND), C.c_int(i), C.c_int(NXi), C.c_int(j), C.c_int(NXj), C.c_int(k), C.c_int(NXk),
                   C.c_longlong(n), _p(T), _p(rho), _p(x), C.c_int(x.shape[0]), _p(out))
        return out

    def isochoric_scalar(self, what, T, rhovec):
        """what: 0 get_pr, 1 get_splus, 2 get_dpdT_constrhovec"""
        T = np.ascontiguousarray(T, dtype=np.float64)
        n = T.size; rv = _soa(rhovec, n)
        out = np.empty(n)
        self._call(lib().ref_isochoric_scalar, C.c_int(what), C.c_longlong(n), _p(T), _p(rv), C.c_int(rv.shape[0]), _p(out))
        return out

    def isochoric_array(self, what, T, rhovec):
        T = np.ascontiguousarray(T, dtype=np.float64)
        n = T.size; rv = _soa(rhovec, n)
        out = np.empty((rv.shape[0], n))
        self._call(lib().ref_isochoric_array, C.c_int(what), C.c_longlong(n), _p(T), _p(rv), C.c_int(rv.shape[0]), _p(out))
        return out

    def fgradHessian(self, T, rhovec):
        T = np.ascontiguousarray(T, dtype=np.float64)
        n = T.size; rv = _soa(rhovec, n); nc = rv.shape[0]
        f = np.empty(n); g = np.empty((nc, n)); H = np.empty((nc * nc, n))
        self._call(lib().ref_Psir_fgradHessian, C.c_int(0), C.c_longlong(n), _p(T), _p(rv), C.c_int(nc), _p(f), _p(g), _p(H))
        return f, g, H

    def lnphi_family(self, what, T, rhovec):
        T = np.ascontiguousarray(T, dtype=np.float64)
        n = T.size; rv = _soa(rhovec, n); nc = rv.shape[0]
        out = np.empty((nc * nc if what == 3 else nc, n))
        self._call(lib().ref_lnphi_family, C.c_int(what), C.c_longlong(n), _p(T), _p(rv), C.c_int(nc), _p(out))
        return out
