"""AbstractModel mirror over the C ABI.  Method names follow teqp::cppinterface::AbstractModel
(/root/reference/include/teqp/cpp/teqpcpp.hpp:94-191): get_Ar00..get_Ar24, get_Ar0{1..6}n, get_Ar{1..4}0n, get_Arxy,
get_ATrhoXi, get_R, get_fugacity_coefficients, build_Psir_gradient_autodiff, build_Psir_Hessian_autodiff,
build_Psir_fgradHessian_autodiff -- scalar calls are batches of one state point on the GPU.  The *_batch methods are the new
array entry points."""
import ctypes as C
import json

import numpy as np

from . import BROADCAST_X, DEVICE_PTRS, TeqpError, _Arg, _check, _dp, _ip, _is_torch, lib


def _torch_stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class Model:
    def __init__(self, spec):
        self.spec = spec if isinstance(spec, str) else json.dumps(spec)
        self._uuid = C.c_longlong(-1)
        err = C.create_string_buffer(2000)
        _check(lib().build_model(self.spec.encode(), C.byref(self._uuid), err, C.c_int(2000)), err)
        n = C.c_int(0)
        _check(lib().teqpb200_get_ncomp(self._uuid, C.byref(n), err, C.c_int(2000)), err)
        self.ncomp = n.value

    def __del__(self):
        try:
            if self._uuid.value >= 0:
                err = C.create_string_buffer(200)
                lib().free_model(self._uuid, err, C.c_int(200))
                self._uuid = C.c_longlong(-1)
        except Exception:
            pass

    # ------------------------------------------------------------------ helpers
    def _prep(self, T, rho, x):
        dev = _is_torch(T)
        flags = DEVICE_PTRS if dev else 0
        n = int(T.numel() if dev else np.size(T))
        xs = tuple(x.shape)
        if len(xs) == 1:
            flags |= BROADCAST_X
            nc = xs[0]
        else:
            nc = xs[0]
            assert xs[1] == n, "molefrac must be [Ncomp][npts] (or [Ncomp] to broadcast)"
        return dev, flags, n, nc, _Arg(T, dev), _Arg(rho, dev), _Arg(x, dev)

    def _out(self, shape, dev, like, out=None):
        if out is not None:
            return out
        if dev:
            import torch
            return torch.empty(shape, dtype=torch.float64, device=like.device)
        return np.empty(shape)

    def _enter(self, dev):
        if dev:
            lib().teqpb200_set_stream(_torch_stream_ptr())

    # ------------------------------------------------------------------ batched entry points
    def get_Ar0n_batch(self, nmax, T, rho, x, out=None):
        dev, flags, n, nc, aT, ar, ax = self._prep(T, rho, x)
        out = self._out((nmax + 1, n), dev, T, out)
        ao = _Arg(out, dev)
        err = C.create_string_buffer(2000)
        self._enter(dev)
        _check(lib().get_Ar0n_batch(self._uuid, C.c_int(nmax), C.c_longlong(n), aT.ptr, ar.ptr, ax.ptr, C.c_int(nc), ao.ptr,
                                    C.c_int(flags), err, C.c_int(2000)), err)
        return ao.obj

    def get_Arn0_batch(self, nmax, T, rho, x, out=None):
        dev, flags, n, nc, aT, ar, ax = self._prep(T, rho, x)
        out = self._out((nmax + 1, n), dev, T, out)
        ao = _Arg(out, dev)
        err = C.create_string_buffer(2000)
        self._enter(dev)
        _check(lib().get_Arn0_batch(self._uuid, C.c_int(nmax), C.c_longlong(n), aT.ptr, ar.ptr, ax.ptr, C.c_int(nc), ao.ptr,
                                    C.c_int(flags), err, C.c_int(2000)), err)
        return ao.obj

    def get_Arxy_batch(self, pairs, T, rho, x, out=None):
        """pairs: list of (NT, ND); returns [len(pairs)][npts]"""
        dev, flags, n, nc, aT, ar, ax = self._prep(T, rho, x)
        out = self._out((len(pairs), n), dev, T, out)
        ao = _Arg(out, dev)
        NT = (C.c_int * len(pairs))(*[p[0] for p in pairs])
        ND = (C.c_int * len(pairs))(*[p[1] for p in pairs])
        err = C.create_string_buffer(2000)
        self._enter(dev)
        _check(lib().get_Arxy_batch(self._uuid, C.c_int(len(pairs)), NT, ND, C.c_longlong(n), aT.ptr, ar.ptr, ax.ptr, C.c_int(nc), ao.ptr,
                                    C.c_int(flags), err, C.c_int(2000)), err)
        return ao.obj

    def get_ATrhoXi_table_batch(self, T, rho, x, out=None):
        """returns [Ncomp][3][3][4][npts]: index (i, NT, ND, NXi)"""
        dev, flags, n, nc, aT, ar, ax = self._prep(T, rho, x)
        out = self._out((nc, 3, 3, 4, n), dev, T, out)
        ao = _Arg(out, dev)
        err = C.create_string_buffer(2000)
        self._enter(dev)
        _check(lib().get_ATrhoXi_table_batch(self._uuid, C.c_longlong(n), aT.ptr, ar.ptr, ax.ptr, C.c_int(nc), ao.ptr, C.c_int(flags),
                                             err, C.c_int(2000)), err)
        return ao.obj

    def get_ATrhoXiXj_table_batch(self, i, j, T, rho, x, out=None):
        """returns [2][2][3][3][npts]: index (NT, ND, NXi, NXj)"""
        dev, flags, n, nc, aT, ar, ax = self._prep(T, rho, x)
        out = self._out((2, 2, 3, 3, n), dev, T, out)
        ao = _Arg(out, dev)
        err = C.create_string_buffer(2000)
        self._enter(dev)
        _check(lib().get_ATrhoXiXj_table_batch(self._uuid, C.c_int(i), C.c_int(j), C.c_longlong(n), aT.ptr, ar.ptr, ax.ptr, C.c_int(nc), ao.ptr,
                                               C.c_int(flags), err, C.c_int(2000)), err)
        return ao.obj

    def get_ATrhoXiXjXk_table_batch(self, i, j, k, T, rho, x, out=None):
        """returns [2][2][2][2][2][npts]: index (NT, ND, NXi, NXj, NXk)"""
        dev, flags, n, nc, aT, ar, ax = self._prep(T, rho, x)
        out = self._out((2, 2, 2, 2, 2, n), dev, T, out)
        ao = _Arg(out, dev)
        err = C.create_string_buffer(2000)
        self._enter(dev)
        _check(lib().get_ATrhoXiXjXk_table_batch(self._uuid, C.c_int(i), C.c_int(j), C.c_int(k), C.c_longlong(n), aT.ptr, ar.ptr, ax.ptr,
                                                 C.c_int(nc), ao.ptr, C.c_int(flags), err, C.c_int(2000)), err)
        return ao.obj

    def _iso(self, fn, T, rhovec, nrows_list, out=None):
        dev = _is_torch(T)
        flags = DEVICE_PTRS if dev else 0
        n = int(T.numel() if dev else np.size(T))
        nc = rhovec.shape[0]
        aT, av = _Arg(T, dev), _Arg(rhovec, dev)
        if out is not None and not isinstance(out, (tuple, list)):
            out = (out,)
        outs = [_Arg(self._out((r, n), dev, T, None if out is None else out[q]), dev) for q, r in enumerate(nrows_list)]
        err = C.create_string_buffer(2000)
        self._enter(dev)
        _check(fn(self._uuid, C.c_longlong(n), aT.ptr, av.ptr, C.c_int(nc), *[o.ptr for o in outs], C.c_int(flags), err, C.c_int(2000)), err)
        return [o.obj for o in outs]

    ISOPROPS = ["pr", "splus", "dpdT", "d2PsirdTdrhoi", "dpdrhovec", "pmv", "chempotVLE", "dchempotdT", "dlnphi_dT", "dlnphi_drho", "dlnphi_dv", "dlnphi_dx"]

    def isochoric_properties_batch(self, T, rhovec, which=None, out=None):
        """dict of derived isochoric properties (names: Model.ISOPROPS; teqpcpp.hpp:52-75, 128-160), one launch"""
        which = list(self.ISOPROPS) if which is None else list(which)
        dev = _is_torch(T)
        flags = DEVICE_PTRS if dev else 0
        n = int(T.numel() if dev else np.size(T))
        nc = rhovec.shape[0]
        aT, av = _Arg(T, dev), _Arg(rhovec, dev)
        rows = dict(zip(self.ISOPROPS, [1, 1, 1] + [nc] * 8 + [nc * nc]))
        outs = {k: _Arg(self._out((rows[k], n), dev, T, None if out is None else out.get(k)), dev) for k in which}
        null = C.POINTER(C.c_double)()
        ptrs = [outs[k].ptr if k in outs else null for k in self.ISOPROPS]
        err = C.create_string_buffer(2000)
        self._enter(dev)
        _check(lib().isochoric_properties_batch(self._uuid, C.c_longlong(n), aT.ptr, av.ptr, C.c_int(nc), *ptrs, C.c_int(flags), err, C.c_int(2000)), err)
        return {k: (outs[k].obj[0] if rows[k] == 1 else outs[k].obj) for k in which}

    def fgradHessian_and_properties_batch(self, T, rhovec, which=None, out=None, out_props=None):
        """build_Psir_fgradHessian_batch and isochoric_properties_batch in one call (one kernel launch for the GERG / multifluid models):
        returns (Psir[npts], grad[Ncomp][npts], H[Ncomp][Ncomp][npts], {name: array}); out = (f[1][npts], g, H[Ncomp*Ncomp][npts]) and out_props reuse buffers"""
        which = list(self.ISOPROPS) if which is None else list(which)
        dev = _is_torch(T)
        flags = DEVICE_PTRS if dev else 0
        n = int(T.numel() if dev else np.size(T))
        nc = rhovec.shape[0]
        aT, av = _Arg(T, dev), _Arg(rhovec, dev)
        fgh = [_Arg(self._out((r, n), dev, T, None if out is None else out[q]), dev) for q, r in enumerate([1, nc, nc * nc])]
        rows = dict(zip(self.ISOPROPS, [1, 1, 1] + [nc] * 8 + [nc * nc]))
        outs = {k: _Arg(self._out((rows[k], n), dev, T, None if out_props is None else out_props.get(k)), dev) for k in which}
        null = C.POINTER(C.c_double)()
        ptrs = [a.ptr for a in fgh] + [outs[k].ptr if k in outs else null for k in self.ISOPROPS]
        err = C.create_string_buffer(2000)
        self._enter(dev)
        _check(lib().build_Psir_fgradHessian_and_properties_batch(self._uuid, C.c_longlong(n), aT.ptr, av.ptr, C.c_int(nc), *ptrs, C.c_int(flags), err, C.c_int(2000)), err)
        H = fgh[2].obj
        return fgh[0].obj[0], fgh[1].obj, H.reshape((nc, nc) + tuple(H.shape[1:])), {k: (outs[k].obj[0] if rows[k] == 1 else outs[k].obj) for k in which}

    def build_Psir_fgradHessian_batch(self, T, rhovec, out=None):
        """returns (Psir[npts], grad[Ncomp][npts], H[Ncomp][Ncomp][npts]); out = (f[1][npts], g[Ncomp][npts], H[Ncomp*Ncomp][npts]) reuses buffers"""
        nc = rhovec.shape[0]
        f, g, H = self._iso(lib().build_Psir_fgradHessian_batch, T, rhovec, [1, nc, nc * nc], out)
        return f[0], g, H.reshape((nc, nc) + tuple(H.shape[1:]))

    def get_ln_fugacity_coefficients_batch(self, T, rhovec, out=None):
        return self._iso(lib().get_ln_fugacity_coefficients_batch, T, rhovec, [rhovec.shape[0]], out)[0]

    def get_fugacity_coefficients_batch(self, T, rhovec, out=None):
        return self._iso(lib().get_fugacity_coefficients_batch, T, rhovec, [rhovec.shape[0]], out)[0]

    def get_AtaudeltaXi_table_batch(self, tau, delta, x):
        """[Ncomp][3][3][4][npts]: index (i, Ntau, Ndelta, NXi); multifluid models only (derivs.hpp:341-351)"""
        tau = np.ascontiguousarray(tau, dtype=np.float64); delta = np.ascontiguousarray(delta, dtype=np.float64); x = np.ascontiguousarray(x, dtype=np.float64)
        n = tau.size; nc = x.shape[0]
        flags = BROADCAST_X if x.ndim == 1 else 0
        out = np.empty((nc, 3, 3, 4, n))
        err = C.create_string_buffer(2000)
        _check(lib().get_AtaudeltaXi_table_batch(self._uuid, C.c_longlong(n), tau.ctypes.data_as(_dp), delta.ctypes.data_as(_dp), x.ctypes.data_as(_dp), C.c_int(nc),
                                                 out.ctypes.data_as(_dp), C.c_int(flags), err, C.c_int(2000)), err)
        return out

    def _taudelta_scalar(self, fn, tau, Ntau, delta, Ndelta, molefrac, *idx):
        z = np.ascontiguousarray(molefrac, dtype=np.float64)
        out = C.c_double()
        err = C.create_string_buffer(2000)
        _check(fn(self._uuid, C.c_double(tau), C.c_int(Ntau), C.c_double(delta), C.c_int(Ndelta), z.ctypes.data_as(_dp), C.c_int(z.size),
                  *[C.c_int(v) for v in idx], C.byref(out), err, C.c_int(2000)), err)
        return out.value

    def get_AtaudeltaXi(self, tau, Ntau, delta, Ndelta, molefrac, i, NXi):
        return self._taudelta_scalar(lib().get_AtaudeltaXi, tau, Ntau, delta, Ndelta, molefrac, i, NXi)

    def get_AtaudeltaXiXj(self, tau, Ntau, delta, Ndelta, molefrac, i, NXi, j, NXj):
        return self._taudelta_scalar(lib().get_AtaudeltaXiXj, tau, Ntau, delta, Ndelta, molefrac, i, NXi, j, NXj)

    def get_AtaudeltaXiXjXk(self, tau, Ntau, delta, Ndelta, molefrac, i, NXi, j, NXj, k, NXk):
        return self._taudelta_scalar(lib().get_AtaudeltaXiXjXk, tau, Ntau, delta, Ndelta, molefrac, i, NXi, j, NXj, k, NXk)

    def get_dmBnvirdTm_batch(self, Nvir, NT, T, x, out=None):
        """d^NT B_Nvir / dT^NT for an array of temperatures (VirialDerivatives::get_dmBnvirdTm, derivs.hpp:780-841); x: [Ncomp][npts] or [Ncomp]"""
        dev = _is_torch(T)
        flags = DEVICE_PTRS if dev else 0
        n = int(T.numel() if dev else np.size(T))
        if len(tuple(x.shape)) == 1:
            flags |= BROADCAST_X
        aT, ax = _Arg(T, dev), _Arg(x, dev)
        ao = _Arg(self._out((n,), dev, T, out), dev)
        err = C.create_string_buffer(2000)
        self._enter(dev)
        _check(lib().get_dmBnvirdTm_batch(self._uuid, C.c_int(Nvir), C.c_int(NT), C.c_longlong(n), aT.ptr, ax.ptr, C.c_int(x.shape[0]), ao.ptr,
                                          C.c_int(flags), err, C.c_int(2000)), err)
        return ao.obj

    def get_Psir_sigma_derivs_batch(self, T, rhovec, v, out=None):
        """[5][npts]: d^k Psir(T, rhovec + sigma v)/dsigma^k, k = 0..4 (IsochoricDerivatives::get_Psir_sigma_derivs, derivs.hpp:1496-1509)"""
        dev = _is_torch(T)
        flags = DEVICE_PTRS if dev else 0
        n = int(T.numel() if dev else np.size(T))
        aT, ar, av = _Arg(T, dev), _Arg(rhovec, dev), _Arg(v, dev)
        ao = _Arg(self._out((5, n), dev, T, out), dev)
        err = C.create_string_buffer(2000)
        self._enter(dev)
        _check(lib().get_Psir_sigma_derivs_batch(self._uuid, C.c_longlong(n), aT.ptr, ar.ptr, av.ptr, C.c_int(rhovec.shape[0]), ao.ptr, C.c_int(flags),
                                                 err, C.c_int(2000)), err)
        return ao.obj

    def pure_VLE_T_batch(self, T, rhoL0, rhoV0, maxiter=10, molefrac=None, want_extra=True):
        """teqp::pure_VLE_T (algorithms/VLE_pure.hpp:108-143) for many isotherms at once: returns (rhoL, rhoV, psat, dpsatdT), the last two None
        unless want_extra (dpsatdT as teqp::dpsatdT_pure, :158-171).  molefrac: ONE composition for all points (default [1.0])."""
        dev = _is_torch(T)
        flags = (DEVICE_PTRS if dev else 0) | BROADCAST_X
        n = int(T.numel() if dev else np.size(T))
        if molefrac is None:
            molefrac = np.ones(1)
        if dev and not _is_torch(molefrac):
            import torch
            molefrac = torch.as_tensor(np.ascontiguousarray(molefrac, dtype=np.float64), device=T.device)
        aT, aL, aV, ax = _Arg(T, dev), _Arg(rhoL0, dev), _Arg(rhoV0, dev), _Arg(molefrac, dev)
        nc = int(ax.obj.numel() if dev else ax.obj.size)
        outs = [_Arg(self._out((n,), dev, T), dev) for _ in range(4 if want_extra else 2)]
        null = C.cast(None, _dp)
        err = C.create_string_buffer(2000)
        self._enter(dev)
        _check(lib().pure_VLE_T_batch(self._uuid, C.c_longlong(n), aT.ptr, aL.ptr, aV.ptr, C.c_int(maxiter), ax.ptr, C.c_int(nc), outs[0].ptr, outs[1].ptr,
                                      outs[2].ptr if want_extra else null, outs[3].ptr if want_extra else null, C.c_int(flags), err, C.c_int(2000)), err)
        return outs[0].obj, outs[1].obj, (outs[2].obj if want_extra else None), (outs[3].obj if want_extra else None)

    def mix_VLE_Tx_batch(self, T, rhovecL0, rhovecV0, xspec, atol=1e-10, reltol=1e-10, axtol=1e-10, relxtol=1e-10, maxiter=10):
        """teqp::mix_VLE_Tx (algorithms/VLE.hpp:76-192) for many binary-mixture problems at once; arrays [2][npts]; returns (rhovecL, rhovecV, return_code)"""
        dev = _is_torch(T)
        flags = DEVICE_PTRS if dev else 0
        n = int(T.numel() if dev else np.size(T))
        aT, aL, aV, ax = _Arg(T, dev), _Arg(rhovecL0, dev), _Arg(rhovecV0, dev), _Arg(xspec, dev)
        oL, oV, oc = _Arg(self._out((2, n), dev, T), dev), _Arg(self._out((2, n), dev, T), dev), _Arg(self._out((n,), dev, T), dev)
        err = C.create_string_buffer(2000)
        self._enter(dev)
        _check(lib().mix_VLE_Tx_batch(self._uuid, C.c_longlong(n), aT.ptr, aL.ptr, aV.ptr, ax.ptr, C.c_double(atol), C.c_double(reltol), C.c_double(axtol),
                                      C.c_double(relxtol), C.c_int(maxiter), oL.ptr, oV.ptr, oc.ptr, C.c_int(flags), err, C.c_int(2000)), err)
        return oL.obj, oV.obj, oc.obj

    def mix_VLE_Tp_batch(self, T, p, rhovecL0, rhovecV0, maxiter=10):
        """teqp::mix_VLE_Tp (algorithms/VLE.hpp:196-380) for many binary problems at once; rhovec arrays [2][npts]; returns (rhovecL, rhovecV)"""
        dev = _is_torch(T)
        flags = DEVICE_PTRS if dev else 0
        n = int(T.numel() if dev else np.size(T))
        aT, ap, aL, aV = _Arg(T, dev), _Arg(p, dev), _Arg(rhovecL0, dev), _Arg(rhovecV0, dev)
        oL, oV = _Arg(self._out((2, n), dev, T), dev), _Arg(self._out((2, n), dev, T), dev)
        err = C.create_string_buffer(2000)
        self._enter(dev)
        _check(lib().mix_VLE_Tp_batch(self._uuid, C.c_longlong(n), aT.ptr, ap.ptr, aL.ptr, aV.ptr, C.c_int(maxiter), oL.ptr, oV.ptr, C.c_int(flags), err, C.c_int(2000)), err)
        return oL.obj, oV.obj

    def point_status_batch(self, T, rho, x, out=None):
        """int32 [npts] bit set per point (teqpb200.h: 1 bad input, 2 non-finite output, 4 mole fractions do not sum to one) for the inputs of a
        batched call and, if given, the [rows][npts] (or [..., npts]) output array it wrote -- the per-point counterpart of the exceptions the
        reference throws per call"""
        dev, flags, n, nc, aT, ar, ax = self._prep(T, rho, x)
        nrows = 0
        optr = None
        if out is not None:
            ao = _Arg(out.reshape(-1, n), dev)
            nrows = int(ao.obj.shape[0])
            optr = ao.ptr
        if dev:
            import torch
            st = torch.empty(n, dtype=torch.int32, device=T.device)
            sp = C.cast(C.c_void_p(st.data_ptr()), _ip)
        else:
            st = np.empty(n, dtype=np.int32)
            sp = st.ctypes.data_as(_ip)
        err = C.create_string_buffer(2000)
        self._enter(dev)
        _check(lib().teqpb200_point_status_batch(self._uuid, C.c_longlong(n), aT.ptr, ar.ptr, ax.ptr, C.c_int(nc), C.c_int(nrows), optr, sp,
                                                 C.c_int(flags), err, C.c_int(2000)), err)
        return st

    def get_reducing_temperature(self, molefrac):
        return self._reducing(0, molefrac)

    def get_reducing_density(self, molefrac):
        return self._reducing(1, molefrac)

    def _reducing(self, which, molefrac):
        z = np.ascontiguousarray(molefrac, dtype=np.float64)
        out = C.c_double()
        err = C.create_string_buffer(2000)
        _check(lib().teqpb200_get_reducing(self._uuid, C.c_int(which), z.ctypes.data_as(_dp), C.c_int(z.size), C.byref(out), err, C.c_int(2000)), err)
        return out.value

    def get_dmBnvirdTm(self, Nvir, NT, T, molefrac):
        z = np.ascontiguousarray(molefrac, dtype=np.float64)
        out = C.c_double()
        err = C.create_string_buffer(2000)
        _check(lib().get_dmBnvirdTm(self._uuid, C.c_int(Nvir), C.c_int(NT), C.c_double(T), z.ctypes.data_as(_dp), C.c_int(z.size), C.byref(out), err,
                                    C.c_int(2000)), err)
        return out.value

    def get_B2vir(self, T, molefrac):
        return self.get_dmBnvirdTm(2, 0, T, molefrac)

    def get_Psir_sigma_derivs(self, T, rhovec, v):
        r = np.ascontiguousarray(rhovec, dtype=np.float64).reshape(-1, 1); d = np.ascontiguousarray(v, dtype=np.float64).reshape(-1, 1)
        return self.get_Psir_sigma_derivs_batch(np.array([float(T)]), r, d)[:, 0]

    def get_deriv_mat2(self, T, rho, molefrac):
        z = np.ascontiguousarray(molefrac, dtype=np.float64)
        pairs = [(i, j) for i in range(3) for j in range(3)]
        return self.get_Arxy_batch(pairs, np.array([float(T)]), np.array([float(rho)]), z.reshape(-1, 1))[:, 0].reshape(3, 3)

    # ------------------------------------------------------------------ reference scalar surface
    def get_R(self, molefrac):
        z = np.ascontiguousarray(molefrac, dtype=np.float64)
        out = C.c_double()
        err = C.create_string_buffer(2000)
        _check(lib().teqpb200_get_R(self._uuid, z.ctypes.data_as(_dp), C.c_int(z.size), C.byref(out), err, C.c_int(2000)), err)
        return out.value

    def get_Arxy(self, NT, ND, T, rho, molefrac):
        z = np.ascontiguousarray(molefrac, dtype=np.float64)
        out = C.c_double()
        err = C.create_string_buffer(2000)
        _check(lib().get_Arxy(self._uuid, C.c_int(NT), C.c_int(ND), C.c_double(T), C.c_double(rho), z.ctypes.data_as(_dp), C.c_int(z.size),
                              C.byref(out), err, C.c_int(2000)), err)
        return out.value

    def get_ATrhoXi(self, T, NT, rho, ND, molefrac, i, NXi):
        z = np.ascontiguousarray(molefrac, dtype=np.float64)
        out = C.c_double()
        err = C.create_string_buffer(2000)
        _check(lib().get_ATrhoXi(self._uuid, C.c_double(T), C.c_int(NT), C.c_double(rho), C.c_int(ND), z.ctypes.data_as(_dp), C.c_int(z.size),
                                 C.c_int(i), C.c_int(NXi), C.byref(out), err, C.c_int(2000)), err)
        return out.value

    def get_ATrhoXiXj(self, T, NT, rho, ND, molefrac, i, NXi, j, NXj):
        z = np.ascontiguousarray(molefrac, dtype=np.float64)
        out = C.c_double()
        err = C.create_string_buffer(2000)
        _check(lib().get_ATrhoXiXj(self._uuid, C.c_double(T), C.c_int(NT), C.c_double(rho), C.c_int(ND), z.ctypes.data_as(_dp), C.c_int(z.size),
                                   C.c_int(i), C.c_int(NXi), C.c_int(j), C.c_int(NXj), C.byref(out), err, C.c_int(2000)), err)
        return out.value

    def get_ATrhoXiXjXk(self, T, NT, rho, ND, molefrac, i, NXi, j, NXj, k, NXk):
        z = np.ascontiguousarray(molefrac, dtype=np.float64)
        out = C.c_double()
        err = C.create_string_buffer(2000)
        _check(lib().get_ATrhoXiXjXk(self._uuid, C.c_double(T), C.c_int(NT), C.c_double(rho), C.c_int(ND), z.ctypes.data_as(_dp), C.c_int(z.size),
                                     C.c_int(i), C.c_int(NXi), C.c_int(j), C.c_int(NXj), C.c_int(k), C.c_int(NXk), C.byref(out), err,
                                     C.c_int(2000)), err)
        return out.value

    def _one(self, T, rho, molefrac):
        z = np.ascontiguousarray(molefrac, dtype=np.float64).reshape(-1, 1)
        return np.array([float(T)]), np.array([float(rho)]), z

    def get_fugacity_coefficients(self, T, rhovec):
        rv = np.ascontiguousarray(rhovec, dtype=np.float64).reshape(-1, 1)
        return self.get_fugacity_coefficients_batch(np.array([float(T)]), rv)[:, 0]

    def build_Psir_gradient_autodiff(self, T, rhovec):
        rv = np.ascontiguousarray(rhovec, dtype=np.float64).reshape(-1, 1)
        return self.build_Psir_fgradHessian_batch(np.array([float(T)]), rv)[1][:, 0]

    def build_Psir_Hessian_autodiff(self, T, rhovec):
        rv = np.ascontiguousarray(rhovec, dtype=np.float64).reshape(-1, 1)
        return self.build_Psir_fgradHessian_batch(np.array([float(T)]), rv)[2][:, :, 0]

    def build_Psir_fgradHessian_autodiff(self, T, rhovec):
        rv = np.ascontiguousarray(rhovec, dtype=np.float64).reshape(-1, 1)
        f, g, H = self.build_Psir_fgradHessian_batch(np.array([float(T)]), rv)
        return float(f[0]), g[:, 0], H[:, :, 0]


def _add_scalar_methods():
    # get_Ar{i}{j}(T, rho, z) -> double (teqpcpp.hpp:22-37,107-109); get_Ar0{n}n / get_Ar{n}0n -> array (:39-51,111-117)
    def mk_ij(i, j):
        def f(self, T, rho, molefrac):
            Tn, rn, z = self._one(T, rho, molefrac)
            return float(self.get_Arxy_batch([(i, j)], Tn, rn, z)[0, 0])
        f.__name__ = "get_Ar%d%d" % (i, j)
        return f
    for i in range(3):
        for j in range(5):
            setattr(Model, "get_Ar%d%d" % (i, j), mk_ij(i, j))

    def mk_0n(n):
        def f(self, T, rho, molefrac):
            Tn, rn, z = self._one(T, rho, molefrac)
            return self.get_Ar0n_batch(n, Tn, rn, z)[:, 0].copy()
        return f

    def mk_n0(n):
        def f(self, T, rho, molefrac):
            Tn, rn, z = self._one(T, rho, molefrac)
            return self.get_Arn0_batch(n, Tn, rn, z)[:, 0].copy()
        return f
    for n in range(1, 7):
        setattr(Model, "get_Ar0%dn" % n, mk_0n(n))
    for n in range(1, 5):
        setattr(Model, "get_Ar%d0n" % n, mk_n0(n))


_add_scalar_methods()


def make_model(spec, validate=True):
    """teqp.make_model(json) equivalent (/root/reference/teqp/__init__.py:42-49); `validate` is accepted for signature parity."""
    return Model(spec)
