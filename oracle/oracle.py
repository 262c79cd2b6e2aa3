"""ctypes bindings for the CPU checkers (TEST INFRASTRUCTURE -- never imported by the product).

* ``Oracle``  -> oracle/liboracle.so        : plain-C restatement (tamcmc_oracle.c)
* ``RefLib``  -> oracle/_ref/libtamcmc_ref.so : the UNMODIFIED reference sources compiled against
  oracle/eigen_shim (present when built in the container that has /root/reference; it travels to
  the GPU box as a prebuilt, git-ignored file).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libtamcmc_ref.so")

_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_ip = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")


def build(force=False):
    """Compile liboracle.so (and oracle/_ref when /root/reference exists)."""
    if force or not os.path.exists(ORACLE_SO) or \
            os.path.getmtime(ORACLE_SO) < os.path.getmtime(os.path.join(HERE, "tamcmc_oracle.c")):
        subprocess.run(["make", "-C", HERE, os.path.join(HERE, "liboracle.so")], check=True,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if os.path.isdir("/root/reference/tamcmc/sources") and (force or not os.path.exists(REF_SO)):
        subprocess.run(["make", "-C", HERE, "ref"], check=True, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


class Oracle:
    def __init__(self):
        build()
        L = self.lib = C.CDLL(ORACLE_SO)
        L.to_model.restype = C.c_int
        L.to_model.argtypes = [C.c_int, _dp, _ip, _dp, C.c_long, _dp, C.POINTER(C.c_double)]
        L.to_call_likelihood.restype = C.c_longdouble
        L.to_call_likelihood.argtypes = [C.c_int, _dp, C.c_void_p, _dp, C.c_long, C.c_double, C.c_double]
        L.to_call_prior.restype = C.c_longdouble
        L.to_call_prior.argtypes = [C.c_int, _dp, _ip, _dp, _ip, _dp, C.POINTER(C.c_int)]
        L.to_generate_models.restype = C.c_int
        L.to_generate_models.argtypes = [C.c_int, C.c_int, C.c_int, _dp, C.c_int, C.c_int, _ip, _dp, _dp,
                                         C.c_void_p, C.c_long, _dp, _ip, _dp, C.c_double, _dp,
                                         _dp, _dp, _dp, C.c_void_p, C.c_void_p, C.c_void_p]
        L.to_amplitude_ratio.argtypes = [C.c_int, C.c_double, _dp]
        L.to_lin_interpol.restype = C.c_double
        L.to_lin_interpol.argtypes = [_dp, _dp, C.c_long, C.c_double]
        L.to_window.restype = C.c_int
        L.to_window.argtypes = [_dp, C.c_long, C.c_double, C.c_double, C.c_double, C.c_int, C.c_double,
                                C.c_double, C.POINTER(C.c_long), C.POINTER(C.c_long)]
        L.to_frstder_adaptive.argtypes = [_dp, C.c_int, _dp]
        L.to_scndder_adaptive.argtypes = [_dp, C.c_int, _dp]
        for name in ("to_logP_uniform", "to_logP_uniform_abs", "to_logP_uniform_cos", "to_logP_gaussian",
                     "to_logP_jeffrey", "to_logP_jeffrey_abs"):
            f = getattr(L, name)
            f.restype = C.c_longdouble
            f.argtypes = [C.c_longdouble] * 3
        for name in ("to_logP_uniform_gaussian", "to_logP_gaussian_uniform"):
            f = getattr(L, name)
            f.restype = C.c_longdouble
            f.argtypes = [C.c_longdouble] * 4
        L.to_logP_gaussian_uniform_gaussian.restype = C.c_longdouble
        L.to_logP_gaussian_uniform_gaussian.argtypes = [C.c_longdouble] * 5
        L.to_p1.restype = C.c_double
        L.to_p1.argtypes = [C.c_double] * 3
        L.to_cholesky.restype = C.c_int
        L.to_cholesky.argtypes = [_dp, C.c_int, _dp]
        L.to_update_proposal.argtypes = [_dp, C.c_double, C.c_int, C.c_double, C.c_double, C.c_double,
                                         C.c_double, _dp, _dp, C.POINTER(C.c_double)]
        L.to_mh_ratio.restype = C.c_double
        L.to_mh_ratio.argtypes = [C.c_double] * 3
        L.to_pt_ratio.restype = C.c_double
        L.to_pt_ratio.argtypes = [C.c_double] * 4

    # -- model / likelihood / prior --------------------------------------------------------
    def model(self, model_case, params, plength, x):
        x = _f64(x)
        out = np.empty_like(x)
        Lacc = C.c_double(0)
        st = self.lib.to_model(model_case, _f64(params), _i32(plength), x, x.size, out, C.byref(Lacc))
        return st, out, Lacc.value

    def likelihood(self, like_case, y, sigma_y, model, likelihood_params=1.0, T=1.0):
        y = _f64(y)
        sig = None if sigma_y is None else _f64(sigma_y).ctypes.data
        return float(self.lib.to_call_likelihood(like_case, y, sig, _f64(model), y.size, likelihood_params, T))

    def prior(self, prior_case, params, plength, priors, switch, extra):
        st = C.c_int(0)
        v = self.lib.to_call_prior(prior_case, _f64(params), _i32(plength), _f64(priors), _i32(switch),
                                   _f64(extra), C.byref(st))
        return float(v), st.value

    def generate_models(self, prob, params, Tcoefs, want_model=False):
        """Batched Model_def::generate_model over C chains (OpenMP over chains)."""
        params = _f64(params)
        Cn, Np = params.shape
        logL = np.empty(Cn); logPr = np.empty(Cn); logPo = np.empty(Cn)
        status = np.zeros(Cn, dtype=np.int32); Lout = np.zeros(Cn)
        model = np.empty((Cn, prob.x.size)) if want_model else None
        sig = None if prob.sigma_y is None else _f64(prob.sigma_y).ctypes.data
        self.lib.to_generate_models(prob.model_case, prob.like_case, prob.prior_case, params, Cn, Np,
                                    _i32(prob.plength), _f64(prob.x), _f64(prob.y), sig, prob.x.size,
                                    _f64(prob.priors), _i32(prob.priors_switch), _f64(prob.extra_priors),
                                    float(prob.likelihood_params), _f64(Tcoefs), logL, logPr, logPo,
                                    None if model is None else model.ctypes.data,
                                    status.ctypes.data, Lout.ctypes.data)
        return dict(logL=logL, logPrior=logPr, logPost=logPo, status=status, L=Lout, model=model)

    # -- helpers ---------------------------------------------------------------------------
    def amplitude_ratio(self, l, beta):
        out = np.empty(2 * l + 1)
        self.lib.to_amplitude_ratio(l, beta, out)
        return out

    def lin_interpol(self, x, y, xi):
        x = _f64(x)
        return self.lib.to_lin_interpol(x, _f64(y), x.size, xi)

    def window(self, x, fc, fs, gamma, l, step, c):
        x = _f64(x)
        a, b = C.c_long(0), C.c_long(0)
        st = self.lib.to_window(x, x.size, fc, fs, gamma, l, step, c, C.byref(a), C.byref(b))
        return st, a.value, b.value

    def frstder(self, y):
        y = _f64(y); out = np.empty_like(y); self.lib.to_frstder_adaptive(y, y.size, out); return out

    def scndder(self, y):
        y = _f64(y); out = np.empty_like(y); self.lib.to_scndder_adaptive(y, y.size, out); return out

    def cholesky(self, A):
        A = _f64(A); n = A.shape[0]; Lm = np.zeros_like(A)
        st = self.lib.to_cholesky(A, n, Lm)
        return st, Lm

    def update_proposal(self, vars_, acceptance, gamma, target, eps1, A1, mu, cov, sigma):
        mu = _f64(mu).copy(); cov = _f64(cov).copy(); s = C.c_double(sigma)
        self.lib.to_update_proposal(_f64(vars_), acceptance, mu.size, gamma, target, eps1, A1, mu, cov, C.byref(s))
        return mu, cov, s.value


class RefLib:
    """The unmodified reference sources (oracle/_ref).  ``available`` is False when absent."""

    def __init__(self):
        build()
        self.available = os.path.exists(REF_SO)
        if not self.available:
            return
        L = self.lib = C.CDLL(REF_SO)
        L.ref_model.restype = C.c_int
        L.ref_model.argtypes = [C.c_int, _dp, C.c_int, _ip, _dp, C.c_long, _dp]
        L.ref_likelihood.restype = C.c_longdouble
        L.ref_likelihood.argtypes = [C.c_int, _dp, C.c_void_p, _dp, C.c_long, C.c_double, C.c_double]
        L.ref_prior.restype = C.c_longdouble
        L.ref_prior.argtypes = [C.c_int, _dp, C.c_int, _ip, _dp, _ip, _dp]
        L.ref_generate_models.argtypes = [C.c_int, C.c_int, C.c_int, _dp, C.c_int, C.c_int, _ip, _dp, _dp,
                                          C.c_void_p, C.c_long, _dp, _ip, _dp, C.c_double, _dp, _dp, _dp, _dp]
        L.ref_amplitude_ratio.argtypes = [C.c_int, C.c_double, _dp]
        L.ref_lin_interpol.restype = C.c_double
        L.ref_lin_interpol.argtypes = [_dp, _dp, C.c_long, C.c_double]
        L.ref_frstder.argtypes = [_dp, C.c_int, _dp]
        L.ref_scndder.argtypes = [_dp, C.c_int, _dp]
        L.ref_logP.restype = C.c_longdouble
        L.ref_logP.argtypes = [C.c_int] + [C.c_longdouble] * 5

    def model(self, model_case, params, plength, x):
        x = _f64(x); params = _f64(params)
        out = np.empty_like(x)
        st = self.lib.ref_model(model_case, params, params.size, _i32(plength), x, x.size, out)
        return st, out

    def likelihood(self, like_case, y, sigma_y, model, likelihood_params=1.0, T=1.0):
        y = _f64(y)
        sig = None if sigma_y is None else _f64(sigma_y).ctypes.data
        return float(self.lib.ref_likelihood(like_case, y, sig, _f64(model), y.size, likelihood_params, T))

    def prior(self, prior_case, params, plength, priors, switch, extra):
        params = _f64(params)
        return float(self.lib.ref_prior(prior_case, params, params.size, _i32(plength), _f64(priors),
                                        _i32(switch), _f64(extra)))

    def generate_models(self, prob, params, Tcoefs):
        params = _f64(params)
        Cn, Np = params.shape
        logL = np.empty(Cn); logPr = np.empty(Cn); logPo = np.empty(Cn)
        sig = None if prob.sigma_y is None else _f64(prob.sigma_y).ctypes.data
        self.lib.ref_generate_models(prob.model_case, prob.like_case, prob.prior_case, params, Cn, Np,
                                     _i32(prob.plength), _f64(prob.x), _f64(prob.y), sig, prob.x.size,
                                     _f64(prob.priors), _i32(prob.priors_switch), _f64(prob.extra_priors),
                                     float(prob.likelihood_params), _f64(Tcoefs), logL, logPr, logPo)
        return dict(logL=logL, logPrior=logPr, logPost=logPo)

    def amplitude_ratio(self, l, beta):
        out = np.empty(2 * l + 1); self.lib.ref_amplitude_ratio(l, beta, out); return out

    def lin_interpol(self, x, y, xi):
        x = _f64(x); return self.lib.ref_lin_interpol(x, _f64(y), x.size, xi)

    def frstder(self, y):
        y = _f64(y); out = np.empty_like(y); self.lib.ref_frstder(y, y.size, out); return out

    def scndder(self, y):
        y = _f64(y); out = np.empty_like(y); self.lib.ref_scndder(y, y.size, out); return out

    def logP(self, kind, a, b, c, d, x):
        return float(self.lib.ref_logP(kind, a, b, c, d, x))
