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

    def Ataudelta(self, which, Ntau, Ndelta, idx, tau, delta, x):
        """which 1: get_AtaudeltaXi idx = (i, NXi); 2: XiXj idx = (i, NXi, j, NXj); 3: XiXjXk idx = (i, NXi, j, NXj, k, NXk)"""
        tau = np.ascontiguousarray(tau, dtype=np.float64); delta = np.ascontiguousarray(delta, dtype=np.float64)
        n = tau.size; x = _soa(x, n)
        out = np.empty(n)
        ii = (C.c_int * 6)(*(list(idx) + [0] * (6 - len(idx))))
        self._call(lib().ref_get_Ataudelta, C.c_int(which), C.c_int(Ntau), C.c_int(Ndelta), ii, C.c_longlong(n), _p(tau), _p(delta), _p(x), C.c_int(x.shape[0]), _p(out))
        return out

    def dmBnvirdTm(self, Nvir, NT, T, x):
        T = np.ascontiguousarray(T, dtype=np.float64)
        n = T.size
        x = np.ascontiguousarray(x, dtype=np.float64)
        if x.ndim == 1:
            x = np.ascontiguousarray(np.tile(x[:, None], (1, n)))
        out = np.empty(n)
        self._call(lib().ref_get_dmBnvirdTm, C.c_int(Nvir), C.c_int(NT), C.c_longlong(n), _p(T), _p(x), C.c_int(x.shape[0]), _p(out))
        return out

    def sigma_derivs(self, T, rhovec, v):
        T = np.ascontiguousarray(T, dtype=np.float64)
        n = T.size; rv = _soa(rhovec, n); vv = _soa(v, n)
        out = np.empty((5, n))
        self._call(lib().ref_Psir_sigma_derivs, C.c_longlong(n), _p(T), _p(rv), _p(vv), C.c_int(rv.shape[0]), _p(out))
        return out

    def pure_VLE_T(self, T, rhoL0, rhoV0, maxiter, z):
        """teqp::pure_VLE_T + dpsatdT_pure per isotherm: (rhoL, rhoV, psat, dpsatdT)"""
        T, rhoL0, rhoV0, z = (np.ascontiguousarray(a, dtype=np.float64) for a in (T, rhoL0, rhoV0, z))
        n = T.size
        rL, rV, ps, dp = (np.empty(n) for _ in range(4))
        self._call(lib().ref_pure_VLE_T, C.c_longlong(n), _p(T), _p(rhoL0), _p(rhoV0), C.c_int(maxiter), _p(z), C.c_int(z.size), _p(rL), _p(rV), _p(ps), _p(dp))
        return rL, rV, ps, dp

    def mix_VLE_Tx(self, T, rvL0, rvV0, xspec, atol=1e-10, reltol=1e-10, axtol=1e-10, relxtol=1e-10, maxiter=10):
        """oracle port of teqp::mix_VLE_Tx (VLE.hpp:76-192) on the reference's AbstractModel: (rhovecL [2][n], rhovecV [2][n], return_code [n])"""
        T = np.ascontiguousarray(T, dtype=np.float64); n = T.size
        a, b, x = _soa(rvL0, n), _soa(rvV0, n), _soa(xspec, n)
        oL, oV, oc = np.empty((2, n)), np.empty((2, n)), np.empty(n)
        self._call(lib().ref_mix_VLE_Tx, C.c_longlong(n), _p(T), _p(a), _p(b), _p(x), C.c_double(atol), C.c_double(reltol), C.c_double(axtol),
                   C.c_double(relxtol), C.c_int(maxiter), _p(oL), _p(oV), _p(oc))
        return oL, oV, oc

    def mix_VLE_Tp(self, T, p, rvL0, rvV0, maxiter=10):
        """oracle port of teqp::mix_VLE_Tp (VLE.hpp:196-380, Newton branch): (rhovecL [2][n], rhovecV [2][n])"""
        T = np.ascontiguousarray(T, dtype=np.float64); p = np.ascontiguousarray(p, dtype=np.float64); n = T.size
        a, b = _soa(rvL0, n), _soa(rvV0, n)
        oL, oV = np.empty((2, n)), np.empty((2, n))
        self._call(lib().ref_mix_VLE_Tp, C.c_longlong(n), _p(T), _p(p), _p(a), _p(b), C.c_int(maxiter), _p(oL), _p(oV))
        return oL, oV

    def deriv_mat2(self, T, rho, x):
        T = np.ascontiguousarray(T, dtype=np.float64); rho = np.ascontiguousarray(rho, dtype=np.float64)
        n = T.size; x = _soa(x, n)
        out = np.empty((6, n))
        self._call(lib().ref_deriv_mat2, C.c_longlong(n), _p(T), _p(rho), _p(x), C.c_int(x.shape[0]), _p(out))
        return out

    def reducing(self, which, z):
        z = np.ascontiguousarray(z, dtype=np.float64)
        out = C.c_double()
        err = C.create_string_buffer(2000)
        rc = lib().ref_get_reducing(C.c_void_p(self.h), C.c_int(which), _p(z), C.c_int(len(z)), C.byref(out), err, C.c_int(2000))
        if rc != 0:
            raise RuntimeError(err.value.decode())
        return out.value

    def lnphi_family(self, what, T, rhovec):
        T = np.ascontiguousarray(T, dtype=np.float64)
        n = T.size; rv = _soa(rhovec, n); nc = rv.shape[0]
        out = np.empty((nc * nc if what == 3 else nc, n))
        self._call(lib().ref_lnphi_family, C.c_int(what), C.c_longlong(n), _p(T), _p(rv), C.c_int(nc), _p(out))
        return out


def _two_sum_split(v):
    """double -> (hi, lo) pair; here lo = 0 (helper for truth_alphar_dd callers that perturb in quad)"""
    v = np.ascontiguousarray(v, dtype=np.float64)
    return v, np.zeros_like(v)


class TruthModel:
    """The reference's model templates in __float128 (oracle/ref_truth.cpp): the adjudicator between libteqpb200 and RefModel."""

    def __init__(self, spec):
        self.spec = spec if isinstance(spec, str) else json.dumps(spec)
        L = lib()
        L.truth_build_model.restype = C.c_void_p
        L.truth_build_model.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        L.truth_free_model.argtypes = [C.c_void_p]
        err = C.create_string_buffer(2000)
        self.h = L.truth_build_model(self.spec.encode(), err, 2000)
        if not self.h:
            raise RuntimeError("truth oracle: " + err.value.decode())
        self.nthreads = os.cpu_count() or 1

    def __del__(self):
        if getattr(self, "h", None):
            lib().truth_free_model(C.c_void_p(self.h))
            self.h = None

    _call = RefModel._call

    def ar_table(self, NT, ND, T, rho, x):
        """Lambda_nm, n <= NT, m <= ND -> [NT+1][ND+1][npts]; shapes (0,6), (4,0), (2,4)"""
        T = np.ascontiguousarray(T, dtype=np.float64); rho = np.ascontiguousarray(rho, dtype=np.float64)
        n = T.size; x = _soa(x, n)
        out = np.empty(((NT + 1) * (ND + 1), n))
        self._call(lib().truth_ar_table, C.c_int(NT), C.c_int(ND), C.c_longlong(n), _p(T), _p(rho), _p(x), C.c_int(x.shape[0]), _p(out))
        return out.reshape(NT + 1, ND + 1, n)

    def Ar0n_refdriver(self, nmax, T, rho, x):
        T = np.ascontiguousarray(T, dtype=np.float64); rho = np.ascontiguousarray(rho, dtype=np.float64)
        n = T.size; x = _soa(x, n)
        out = np.empty((nmax + 1, n))
        self._call(lib().truth_Ar0n_refdriver, C.c_int(nmax), C.c_longlong(n), _p(T), _p(rho), _p(x), C.c_int(x.shape[0]), _p(out))
        return out

    def ATrhoXi_table(self, T, rho, x):
        """[ncomp][3][3][4][npts], same layout as Model.get_ATrhoXi_table_batch"""
        T = np.ascontiguousarray(T, dtype=np.float64); rho = np.ascontiguousarray(rho, dtype=np.float64)
        n = T.size; x = _soa(x, n); nc = x.shape[0]
        out = np.empty((nc * 36, n))
        self._call(lib().truth_ATrhoXi_table, C.c_longlong(n), _p(T), _p(rho), _p(x), C.c_int(nc), _p(out))
        return out.reshape(nc, 3, 3, 4, n)

    def fgradHessian(self, T, rhovec, want_H=True):
        T = np.ascontiguousarray(T, dtype=np.float64)
        n = T.size; rv = _soa(rhovec, n); nc = rv.shape[0]
        f = np.empty(n); g = np.empty((nc, n)); H = np.empty((nc * nc, n)) if want_H else None
        self._call(lib().truth_Psir_fgradHessian, C.c_longlong(n), _p(T), _p(rv), C.c_int(nc), _p(f), _p(g), _p(H) if want_H else None)
        return f, g, H

    def alphar_dd(self, T, rho, x):
        """alphar at (T, rho, x) given as (hi, lo) double-double pairs; returns (hi, lo) of the quad result"""
        (Th, Tl), (rh, rl), (xh, xl) = T, rho, x
        Th = np.ascontiguousarray(Th, dtype=np.float64); n = Th.size
        arrs = [np.ascontiguousarray(a, dtype=np.float64) for a in (Th, Tl, rh, rl)]
        xh = _soa(xh, n); xl = _soa(xl, n)
        oh = np.empty(n); ol = np.empty(n)
        self._call(lib().truth_alphar_dd, C.c_longlong(n), _p(arrs[0]), _p(arrs[1]), _p(arrs[2]), _p(arrs[3]), _p(xh), _p(xl), C.c_int(xh.shape[0]),
                   _p(oh), _p(ol))
        return oh, ol
