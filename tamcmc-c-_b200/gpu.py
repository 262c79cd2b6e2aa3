"""ctypes binding of include/tamcmc_gpu.h and the host-side mirror of the reference interface.

``GpuModelDef`` mirrors ``Model_def`` (tamcmc/headers/model_def.h:58-83): public per-chain members
``params, vars, logLikelihood, logPrior, logPosterior, Pmove, moved`` and the methods
``generate_model / call_model / update_params_with_vars``; ``GpuMALA`` mirrors ``MALA``
(tamcmc/headers/MALA.h) with ``execute``.  All numerical work happens in libtamcmc_gpu.so.
"""
import ctypes as C
import os
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

from .build import LIB_PATH, build_library

ST_EMPTY_WINDOW, ST_NAN_WINDOW, ST_NAN_LOGL, ST_BAD_PRIOR, ST_CHOL_FAIL, ST_MODE_OVERFLOW = 1, 2, 4, 8, 16, 32
F_FIX_SWAP_BOOKKEEPING = 1
F_RESTORE_MEANS = 4

EXPORTS = ["tg_last_error", "tg_version", "tg_create", "tg_destroy", "tg_eval", "tg_eval_async", "tg_eval_wait", "tg_eval_noalloc", "tg_eval_upload", "tg_eval_resident",
           "tg_eval_fetch", "tg_last_work", "tg_get_state", "tg_set_state", "tg_run", "tg_sizes", "tg_step_begin",
           "tg_step_end", "tg_partials_device_ptr", "tg_shard_install_current", "tg_shard_unique_id", "tg_shard_attach", "tg_shard_detach", "tg_shard_ipc_export", "tg_shard_ipc_attach", "tg_shard_status", "tg_sync", "tg_kernel_time", "tg_set_tuning",
           "tg_measure_fp64_peak", "tg_work_profile", "tg_event_record", "tg_event_elapsed_ms",
           "tg_batch_create", "tg_batch_run", "tg_batch_destroy", "tg_batch_sizes", "tg_batch_kernel_time", "tg_batch_event_record",
           "tg_batch_prepass_time", "tg_prepass_time", "tg_farfield_profile", "tg_get_state_means", "tg_last_run_ms", "tg_batch_last_run_ms"]
# include/tamcmc_outputs.h (host code in the same library)
EXPORTS_OUT = ["tg_out_last_error", "tg_out_open", "tg_out_append", "tg_out_write_restore", "tg_out_write_restore_means", "tg_out_close", "tg_restore_read", "tg_restore_read_means"]


# include/tamcmc_frontend.h (host code in the same library)
EXPORTS_FE = ["tg_fe_last_error", "tg_fe_open", "tg_fe_close", "tg_fe_counts", "tg_fe_nslices", "tg_fe_setup", "tg_fe_problem",
              "tg_fe_run_info", "tg_fe_input_name", "tg_fe_prior_name", "tg_fe_dump", "tg_evidence", "tg_acceptance"]


class TgError(RuntimeError):
    pass


class _Problem(C.Structure):
    _fields_ = [
        ("Nx", C.c_long), ("x", C.c_void_p), ("y", C.c_void_p), ("sigma_y", C.c_void_p),
        ("Nx_global", C.c_long), ("shard_lo", C.c_long),
        ("x_first", C.c_double), ("x_second", C.c_double), ("x_last", C.c_double),
        ("model_case", C.c_int), ("likelihood_case", C.c_int), ("prior_case", C.c_int),
        ("likelihood_params", C.c_double), ("plength", C.c_int * 11), ("Np", C.c_int),
        ("params0", C.c_void_p), ("relax", C.c_void_p), ("priors", C.c_void_p), ("priors_switch", C.c_void_p),
        ("extra_priors", C.c_double * 4),
        ("Nchains", C.c_int), ("chains_per_ensemble", C.c_int), ("Tcoefs", C.c_void_p),
        ("c0", C.c_double), ("epsilon1", C.c_double), ("epsilon2", C.c_double), ("A1", C.c_double),
        ("target_acceptance", C.c_double),
        ("n_learn", C.c_int), ("Nt_learn", C.c_void_p), ("periods_learn", C.c_void_p), ("dN_mixing", C.c_long),
        ("init_err", C.c_void_p), ("seed", C.c_ulonglong), ("device", C.c_int), ("flags", C.c_int),
        ("philox_chain0", C.c_int), ("philox_ensemble0", C.c_int),
    ]


class _Records(C.Structure):
    _fields_ = [("vars", C.c_void_p), ("logL", C.c_void_p), ("logPrior", C.c_void_p), ("logPost", C.c_void_p),
                ("Pmove", C.c_void_p), ("moved", C.c_void_p), ("swap_attempt", C.c_void_p),
                ("swap_chain0", C.c_void_p), ("swap_Pswap", C.c_void_p), ("swap_done", C.c_void_p),
                ("status", C.c_void_p)]


class GpuLibrary:
    """Loaded libtamcmc_gpu.so.  Raises if the library cannot be loaded: there is no fallback."""

    def __init__(self, path=None):
        path = path or LIB_PATH
        if not os.path.exists(path):
            raise TgError("libtamcmc_gpu.so not built (%s); run tamcmc_b200.build_library()" % path)
        L = self.lib = C.CDLL(path)
        self.path = path
        L.tg_last_error.restype = C.c_char_p
        for name in EXPORTS[1:]:
            getattr(L, name).restype = C.c_int
        L.tg_create.argtypes = [C.POINTER(_Problem), C.POINTER(C.c_void_p)]
        L.tg_destroy.argtypes = [C.c_void_p]
        L.tg_eval.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p] + [C.c_void_p] * 5
        L.tg_eval_async.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.tg_eval_wait.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 4
        L.tg_eval_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p] + [C.c_void_p] * 4
        L.tg_eval_upload.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.tg_eval_resident.argtypes = [C.c_void_p, C.c_int]
        L.tg_eval_fetch.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 4
        L.tg_last_work.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.POINTER(C.c_int)]
        L.tg_get_state.argtypes = [C.c_void_p] + [C.c_void_p] * 7
        L.tg_set_state.argtypes = [C.c_void_p] + [C.c_void_p] * 4
        L.tg_run.argtypes = [C.c_void_p, C.c_long, C.c_long, C.POINTER(_Records)]
        L.tg_sizes.argtypes = [C.c_void_p] + [C.POINTER(C.c_int)] * 4 + [C.POINTER(C.c_long)]
        L.tg_step_begin.argtypes = [C.c_void_p, C.c_long]
        L.tg_step_end.argtypes = [C.c_void_p, C.c_long]
        L.tg_partials_device_ptr.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int)]
        L.tg_shard_install_current.argtypes = [C.c_void_p]
        L.tg_shard_unique_id.argtypes = [C.c_char_p]
        L.tg_shard_attach.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int]
        L.tg_shard_detach.argtypes = [C.c_void_p]
        L.tg_shard_ipc_export.argtypes = [C.c_void_p, C.c_int, C.c_char_p]
        L.tg_shard_ipc_attach.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int]
        L.tg_shard_status.argtypes = [C.c_void_p]
        L.tg_sync.argtypes = [C.c_void_p]
        L.tg_kernel_time.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_long), C.POINTER(C.c_double), C.POINTER(C.c_long)]
        L.tg_set_tuning.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
        L.tg_measure_fp64_peak.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.tg_event_record.argtypes = [C.c_void_p, C.c_int]
        L.tg_event_elapsed_ms.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.tg_work_profile.argtypes = [C.c_void_p, C.POINTER(C.c_long), C.POINTER(C.c_long), C.POINTER(C.c_int)]
        L.tg_batch_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.POINTER(C.c_void_p)]
        L.tg_batch_run.argtypes = [C.c_void_p, C.c_long, C.c_long, C.POINTER(C.POINTER(_Records))]
        L.tg_batch_destroy.argtypes = [C.c_void_p]
        L.tg_batch_sizes.argtypes = [C.c_void_p] + [C.POINTER(C.c_int)] * 3
        L.tg_batch_kernel_time.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_long), C.POINTER(C.c_double), C.POINTER(C.c_long)]
        L.tg_batch_event_record.argtypes = [C.c_void_p, C.c_int]
        L.tg_batch_prepass_time.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.tg_prepass_time.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.tg_farfield_profile.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_long)]
        L.tg_last_run_ms.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.tg_batch_last_run_ms.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.tg_get_state_means.argtypes = [C.c_void_p] + [C.c_void_p] * 4 + [C.POINTER(C.c_long), C.c_int]
        L.tg_out_last_error.restype = C.c_char_p
        for name in EXPORTS_OUT[1:]:
            getattr(L, name).restype = C.c_int
        L.tg_out_open.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        L.tg_out_append.argtypes = [C.c_void_p, C.POINTER(_Records), C.c_long, C.c_int]
        L.tg_out_write_restore.argtypes = [C.c_void_p, C.c_long] + [C.c_void_p] * 4
        L.tg_out_write_restore_means.argtypes = [C.c_void_p, C.c_long] + [C.c_void_p] * 8
        L.tg_out_close.argtypes = [C.c_void_p]
        L.tg_restore_read.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_long)] + [C.c_void_p] * 4
        L.tg_restore_read_means.argtypes = L.tg_restore_read.argtypes
        L.tg_fe_last_error.restype = C.c_char_p
        for name in ("tg_fe_open", "tg_fe_counts", "tg_fe_nslices", "tg_fe_setup", "tg_fe_problem", "tg_fe_run_info", "tg_evidence", "tg_acceptance"):
            getattr(L, name).restype = C.c_int
        L.tg_fe_close.restype = None
        L.tg_fe_input_name.restype = C.c_char_p
        L.tg_fe_prior_name.restype = C.c_char_p
        L.tg_fe_dump.restype = C.c_long
        L.tg_fe_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
        L.tg_fe_close.argtypes = [C.c_void_p]
        L.tg_fe_counts.argtypes = [C.c_void_p] + [C.POINTER(C.c_int)] * 4
        L.tg_fe_nslices.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int)]
        L.tg_fe_setup.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.tg_fe_problem.argtypes = [C.c_void_p, C.POINTER(_Problem)]
        L.tg_fe_run_info.argtypes = [C.c_void_p, C.c_void_p]
        L.tg_fe_input_name.argtypes = [C.c_void_p, C.c_int]
        L.tg_fe_prior_name.argtypes = [C.c_void_p, C.c_int]
        L.tg_fe_dump.argtypes = [C.c_void_p, C.c_char_p, C.c_long]
        L.tg_evidence.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_long, C.c_int, C.c_void_p, C.POINTER(C.c_double)]
        L.tg_acceptance.argtypes = [C.c_void_p, C.c_long, C.c_int, C.c_long, C.c_long, C.c_long, C.c_void_p, C.POINTER(C.c_double)]

    def check(self, rc):
        if rc != 0:
            raise TgError("libtamcmc_gpu error %d: %s" % (rc, self.lib.tg_last_error().decode()))

    def measure_fp64_peak(self, device=0):
        t, ms = C.c_double(0), C.c_double(0)
        self.check(self.lib.tg_measure_fp64_peak(device, C.byref(t), C.byref(ms)))
        return t.value, ms.value


_LIB = None


def shard_unique_id(lib=None):
    """128 bytes of ncclGetUniqueId (call on one rank, broadcast to the others, pass to GpuModelDef.shard_attach)."""
    lib = lib or load_library()
    buf = C.create_string_buffer(128)
    lib.check(lib.lib.tg_shard_unique_id(buf))
    return buf.raw


def load_library(build=True):
    global _LIB
    if _LIB is None:
        if build:
            build_library()
        _LIB = GpuLibrary()
    return _LIB


@dataclass
class MalaConfig:
    """!MALA section of Config/default/config_default.cfg (defaults as shipped)."""
    Nchains: int = 10
    lambda_temp: float = 1.7
    chains_per_ensemble: int = 0
    Tcoefs: Optional[np.ndarray] = None
    c0: float = 10.0
    epsilon1: float = 1e-12
    epsilon2: float = 1e-12
    A1: float = 1e14
    target_acceptance: float = 0.234
    Nt_learn: List[int] = field(default_factory=lambda: [500, 1500, 100000])
    periods_learn: List[int] = field(default_factory=lambda: [1, 1])
    dN_mixing: int = 1
    seed: int = 12345
    flags: int = 0
    philox_chain0: int = 0          # offsets of the chain / ensemble indices in the Philox counters (include/tamcmc_gpu.h)
    philox_ensemble0: int = 0

    def tcoefs(self):
        if self.Tcoefs is not None:
            return np.ascontiguousarray(self.Tcoefs, dtype=np.float64)
        ce = self.chains_per_ensemble or self.Nchains
        return np.power(self.lambda_temp, (np.arange(self.Nchains) % ce).astype(np.float64))   # MALA.cpp:103


def _p(a):
    return None if a is None else a.ctypes.data


class GpuModelDef:
    """Model_def-shaped adapter: owns one tg_handle (one GPU, one stream)."""

    def __init__(self, prob, mala: Optional[MalaConfig] = None, device=0, shard=None, lib=None):
        self.lib = lib or load_library()
        self.prob = prob
        self.mala = mala or MalaConfig()
        self._keep = []
        P = _Problem()
        f64 = lambda a: self._hold(np.ascontiguousarray(a, dtype=np.float64))
        i32 = lambda a: self._hold(np.ascontiguousarray(a, dtype=np.int32))
        i64 = lambda a: self._hold(np.ascontiguousarray(a, dtype=np.int64))
        x, y = f64(prob.x), f64(prob.y)
        P.Nx, P.x, P.y = x.size, _p(x), _p(y)
        P.sigma_y = _p(f64(prob.sigma_y)) if prob.sigma_y is not None else None
        if shard is not None:   # dict(Nx_global, shard_lo, x_first, x_second, x_last)
            P.Nx_global, P.shard_lo = shard["Nx_global"], shard["shard_lo"]
            P.x_first, P.x_second, P.x_last = shard["x_first"], shard["x_second"], shard["x_last"]
        P.model_case, P.likelihood_case, P.prior_case = prob.model_case, prob.like_case, prob.prior_case
        P.likelihood_params = float(prob.likelihood_params)
        P.plength = (C.c_int * 11)(*[int(v) for v in prob.plength])
        P.Np = prob.Np
        P.params0, P.relax = _p(f64(prob.params)), _p(i32(prob.relax))
        P.priors, P.priors_switch = _p(f64(prob.priors)), _p(i32(prob.priors_switch))
        P.extra_priors = (C.c_double * 4)(*[float(v) for v in prob.extra_priors])
        m = self.mala
        T = f64(m.tcoefs())
        P.Nchains, P.chains_per_ensemble, P.Tcoefs = m.Nchains, m.chains_per_ensemble, _p(T)
        P.c0, P.epsilon1, P.epsilon2, P.A1, P.target_acceptance = m.c0, m.epsilon1, m.epsilon2, m.A1, m.target_acceptance
        P.n_learn = len(m.periods_learn)
        P.Nt_learn, P.periods_learn = _p(i64(m.Nt_learn)), _p(i64(m.periods_learn))
        P.dN_mixing = m.dN_mixing
        P.init_err = _p(f64(prob.init_errors())) if prob.err_frac is not None else None
        P.seed, P.device, P.flags = m.seed, device, m.flags
        P.philox_chain0, P.philox_ensemble0 = m.philox_chain0, m.philox_ensemble0
        self.Tcoefs = T
        h = C.c_void_p()
        self.lib.check(self.lib.lib.tg_create(C.byref(P), C.byref(h)))
        self.h = h
        nv, np_, c, e, nx = C.c_int(), C.c_int(), C.c_int(), C.c_int(), C.c_long()
        self.lib.check(self.lib.lib.tg_sizes(h, C.byref(nv), C.byref(np_), C.byref(c), C.byref(e), C.byref(nx)))
        self.Nvars, self.Nparams, self.Nmodels, self.Nensembles, self.Nx = nv.value, np_.value, c.value, e.value, nx.value
        self.index_to_relax = prob.index_to_relax
        # Model_def public members (model_def.h:58-69)
        self.params = np.tile(np.asarray(prob.params, dtype=np.float64), (self.Nmodels, 1))
        self.vars = self.params[:, self.index_to_relax].copy()
        self.logLikelihood = np.zeros(self.Nmodels)
        self.logPrior = np.zeros(self.Nmodels)
        self.logPosterior = np.zeros(self.Nmodels)
        self.Pmove = np.zeros(self.Nmodels)
        self.moved = np.zeros(self.Nmodels, dtype=bool)
        self.status = np.zeros(self.Nmodels, dtype=np.int32)
        self.refresh_from_device()

    def _hold(self, a):
        self._keep.append(a)
        return a

    def close(self):
        if getattr(self, "h", None):
            self.lib.lib.tg_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- Model_def interface -------------------------------------------------------------
    def update_params_with_vars(self, m=None):
        """model_def.cpp:370-378"""
        if m is None:
            self.params[:, self.index_to_relax] = self.vars
        else:
            self.params[m, self.index_to_relax] = self.vars[m]

    def evaluate(self, params, T=None, want_model=False):
        """Batched generate_model for arbitrary parameter vectors [B, Np] (tg_eval)."""
        params = np.ascontiguousarray(np.atleast_2d(params), dtype=np.float64)
        B = params.shape[0]
        assert params.shape[1] == self.Nparams
        T = np.ones(B) if T is None else np.ascontiguousarray(T, dtype=np.float64)
        logL, logPr, logPo = np.empty(B), np.empty(B), np.empty(B)
        st = np.zeros(B, dtype=np.int32)
        model = np.empty((B, self.Nx)) if want_model else None
        self.lib.check(self.lib.lib.tg_eval(self.h, _p(params), B, _p(T), _p(logL), _p(logPr), _p(logPo), _p(st), _p(model)))
        return dict(logL=logL, logPrior=logPr, logPost=logPo, status=st, model=model)

    def evaluate_async(self, params, T, slot):
        """tg_eval_async: enqueue the evaluation of params [B, Np] through staging slot 0 / 1; collect with evaluate_wait."""
        params = np.ascontiguousarray(np.atleast_2d(params), dtype=np.float64)
        T = np.ascontiguousarray(T, dtype=np.float64)
        self.lib.check(self.lib.lib.tg_eval_async(self.h, _p(params), params.shape[0], _p(T), slot))
        return params.shape[0]

    def evaluate_wait(self, B, slot):
        logL, logPr, logPo = np.empty(B), np.empty(B), np.empty(B)
        st = np.zeros(B, dtype=np.int32)
        self.lib.check(self.lib.lib.tg_eval_wait(self.h, slot, _p(logL), _p(logPr), _p(logPo), _p(st)))
        return dict(logL=logL, logPrior=logPr, logPost=logPo, status=st)

    def evaluate_noalloc(self, params, T=None):
        """tg_eval_noalloc: the same evaluation through the library's static device scratch (no allocation in the call)."""
        params = np.ascontiguousarray(np.atleast_2d(params), dtype=np.float64)
        B = params.shape[0]
        assert params.shape[1] == self.Nparams
        T = np.ones(B) if T is None else np.ascontiguousarray(T, dtype=np.float64)
        logL, logPr, logPo = np.empty(B), np.empty(B), np.empty(B)
        st = np.zeros(B, dtype=np.int32)
        self.lib.check(self.lib.lib.tg_eval_noalloc(self.h, _p(params), B, _p(T), _p(logL), _p(logPr), _p(logPo), _p(st)))
        return dict(logL=logL, logPrior=logPr, logPost=logPo, status=st)

    def generate_model(self, m=None, Tcoefs=None):
        """model_def.cpp:358-367 for chain m (or all chains): fills logLikelihood/logPrior/logPosterior."""
        T = self.Tcoefs if Tcoefs is None else np.asarray(Tcoefs, dtype=np.float64)
        rows = np.arange(self.Nmodels) if m is None else np.atleast_1d(m)
        r = self.evaluate(self.params[rows], T[rows])
        self.logLikelihood[rows], self.logPrior[rows], self.logPosterior[rows] = r["logL"], r["logPrior"], r["logPost"]
        self.status[rows] = r["status"]
        return self.logPosterior[rows] if m is None else float(self.logPosterior[rows][0])

    def call_model(self, m=0):
        """model_def.cpp:210-289: the model spectrum of chain m (parity / getmodel path)."""
        return self.evaluate(self.params[m:m + 1], want_model=True)["model"][0]

    def call_model_explicit(self, params):
        """model_def.cpp:199-208"""
        return self.evaluate(params, want_model=True)["model"][0]

    def last_work(self, B):
        L = np.zeros(B)
        K = C.c_int(0)
        self.lib.check(self.lib.lib.tg_last_work(self.h, B, _p(L), C.byref(K)))
        return L, K.value

    # -- resident evaluation (no PCIe traffic inside the timed region) -----------------------
    def eval_upload(self, params, T):
        params = np.ascontiguousarray(params, dtype=np.float64)
        T = np.ascontiguousarray(T, dtype=np.float64)
        self.lib.check(self.lib.lib.tg_eval_upload(self.h, _p(params), params.shape[0], _p(T)))
        self.sync()

    def eval_resident(self, B):
        self.lib.check(self.lib.lib.tg_eval_resident(self.h, B))

    def eval_fetch(self, B):
        logL, logPr, logPo = np.empty(B), np.empty(B), np.empty(B)
        st = np.zeros(B, dtype=np.int32)
        self.lib.check(self.lib.lib.tg_eval_fetch(self.h, B, _p(logL), _p(logPr), _p(logPo), _p(st)))
        return dict(logL=logL, logPrior=logPr, logPost=logPo, status=st)

    # -- device state ------------------------------------------------------------------------
    def refresh_from_device(self):
        self.lib.check(self.lib.lib.tg_get_state(self.h, _p(self.vars), None, None, None, _p(self.logLikelihood),
                                                 _p(self.logPrior), _p(self.logPosterior)))
        self.update_params_with_vars()

    def get_proposal(self):
        Cn, Nv = self.Nmodels, self.Nvars
        mu, sigma, cov = np.empty((Cn, Nv)), np.empty(Cn), np.empty((Cn, Nv, Nv))
        self.lib.check(self.lib.lib.tg_get_state(self.h, None, _p(mu), _p(sigma), _p(cov), None, None, None))
        return mu, sigma, cov

    def get_state_means(self, reset=True):
        """buf_restore.*_mean of the reference (outputs.cpp:1758-1818): means of vars, mu, sigma, covarmat over the steps since the last
        reset, and their number.  Needs MalaConfig(flags=F_RESTORE_MEANS)."""
        Cn, Nv = self.Nmodels, self.Nvars
        v, mu, sigma, cov = np.empty((Cn, Nv)), np.empty((Cn, Nv)), np.empty(Cn), np.empty((Cn, Nv, Nv))
        n = C.c_long(0)
        self.lib.check(self.lib.lib.tg_get_state_means(self.h, _p(v), _p(mu), _p(sigma), _p(cov), C.byref(n), int(bool(reset))))
        return dict(vars=v, mu=mu, sigma=sigma, covarmat=cov, count=n.value)

    def set_state(self, vars_=None, mu=None, sigma=None, covarmat=None):
        f = lambda a: None if a is None else np.ascontiguousarray(a, dtype=np.float64)
        v, m_, s, c = f(vars_), f(mu), f(sigma), f(covarmat)
        self.lib.check(self.lib.lib.tg_set_state(self.h, _p(v), _p(m_), _p(s), _p(c)))
        self.refresh_from_device()

    def sync(self):
        self.lib.check(self.lib.lib.tg_sync(self.h))

    def shard_ipc_export(self, nranks: int) -> bytes:
        """Bin-sharded variant over peer memory: this rank's exchange buffer as a 64-byte CUDA IPC handle."""
        buf = C.create_string_buffer(64)
        self.lib.check(self.lib.lib.tg_shard_ipc_export(self.h, nranks, buf))
        return buf.raw

    def shard_ipc_attach(self, handles: bytes, nranks: int, rank: int):
        assert len(handles) == 64 * nranks
        self.lib.check(self.lib.lib.tg_shard_ipc_attach(self.h, handles, nranks, rank))
        self.refresh_from_device()

    def shard_status(self):
        self.lib.check(self.lib.lib.tg_shard_status(self.h))

    def shard_attach(self, unique_id: bytes, nranks: int, rank: int):
        """Bin-sharded variant inside the library: joins the NCCL communicator of the shard group (tg_shard_attach)."""
        assert len(unique_id) == 128
        self.lib.check(self.lib.lib.tg_shard_attach(self.h, unique_id, nranks, rank))
        self.refresh_from_device()

    def set_tuning(self, chains_per_block=0, light_subtiles=0, use_graph=-1, heavy_group=0):
        self.lib.check(self.lib.lib.tg_set_tuning(self.h, chains_per_block, heavy_group, light_subtiles, use_graph))

    def last_run_ms(self):
        """Device time of the steps of the last execute() on this handle (events recorded inside tg_run)."""
        ms = C.c_double(0)
        self.lib.check(self.lib.lib.tg_last_run_ms(self.h, C.byref(ms)))
        return ms.value

    def timer_start(self):
        self.lib.check(self.lib.lib.tg_event_record(self.h, 0))

    def timer_stop_ms(self):
        """Records the stop event and returns the device time since timer_start (synchronises)."""
        self.lib.check(self.lib.lib.tg_event_record(self.h, 1))
        ms = C.c_double(0)
        self.lib.check(self.lib.lib.tg_event_elapsed_ms(self.h, C.byref(ms)))
        return ms.value

    def work_profile(self):
        """(bins evaluated bin by bin, background-only moment tiles, work units) of the sampler's launches."""
        b, m, u = C.c_long(0), C.c_long(0), C.c_int(0)
        self.lib.check(self.lib.lib.tg_work_profile(self.h, C.byref(b), C.byref(m), C.byref(u)))
        return b.value, m.value, u.value

    def farfield_profile(self, B):
        """(micro-tiles prepared, marked, exact records, far components interpolated, micro-tiles holding the background, bins per micro-tile) of the last evaluate() batch."""
        out = (C.c_long * 6)()
        self.lib.check(self.lib.lib.tg_farfield_profile(self.h, B, out))
        return tuple(out)

    def kernel_time(self, reset=0):
        n, ms, tot = C.c_long(0), C.c_double(0), C.c_long(0)
        self.lib.check(self.lib.lib.tg_kernel_time(self.h, reset, C.byref(n), C.byref(ms), C.byref(tot)))
        return n.value, ms.value, tot.value

    def prepass_time(self):
        ms = C.c_double(0)
        self.lib.check(self.lib.lib.tg_prepass_time(self.h, C.byref(ms)))
        return ms.value


class GpuMALA:
    """MALA-shaped driver: ``execute`` runs the sampling loop on the device (MALA.cpp:540-739)."""

    def __init__(self, model_current: GpuModelDef):
        self.model = model_current
        self.Nchains = model_current.Nmodels
        self.Tcoefs = model_current.Tcoefs

    @staticmethod
    def _records(md, nsteps, record_vars=True):
        Cn, Nv, E = md.Nmodels, md.Nvars, md.Nensembles
        out = dict(
            vars=np.empty((nsteps, Cn, Nv)) if record_vars else None,
            logL=np.empty((nsteps, Cn)), logPrior=np.empty((nsteps, Cn)), logPost=np.empty((nsteps, Cn)),
            Pmove=np.empty((nsteps, Cn)), moved=np.empty((nsteps, Cn), dtype=np.uint8),
            swap_attempt=np.empty((nsteps, E), dtype=np.int32), swap_chain0=np.empty((nsteps, E), dtype=np.int32),
            swap_Pswap=np.empty((nsteps, E)), swap_done=np.empty((nsteps, E), dtype=np.int32),
            status=np.empty((nsteps, Cn), dtype=np.int32))
        rec = _Records(*[_p(out[k]) for k in ("vars", "logL", "logPrior", "logPost", "Pmove", "moved",
                                               "swap_attempt", "swap_chain0", "swap_Pswap", "swap_done", "status")])
        return out, rec

    def execute(self, i0, nsteps, record=True, record_vars=True):
        md = self.model
        out = None
        rec = None
        if record:
            out, rec = self._records(md, nsteps, record_vars)
        md.lib.check(md.lib.lib.tg_run(md.h, i0, nsteps, C.byref(rec) if rec is not None else None))
        md.refresh_from_device()
        return out

    def step_begin(self, i):
        self.model.lib.check(self.model.lib.lib.tg_step_begin(self.model.h, i))

    def step_end(self, i):
        self.model.lib.check(self.model.lib.lib.tg_step_end(self.model.h, i))

    def install_current(self):
        """Bin-sharded variant: install the current logL from the (all-reduced) partials buffer."""
        self.model.lib.check(self.model.lib.lib.tg_shard_install_current(self.model.h))
        self.model.refresh_from_device()

    def partials_ptr(self):
        p, n = C.c_void_p(), C.c_int()
        self.model.lib.check(self.model.lib.lib.tg_partials_device_ptr(self.model.h, C.byref(p), C.byref(n)))
        return p.value, n.value


class GpuBatch:
    """Launch group of several stars on one GPU (tg_batch_*): the reference runs stars as independent processes
    (scripts/slurm/job_array.sh:8,24, main.cpp:111-127); here every step of all of them is one proposal, one fused
    model+likelihood and one decision launch.  ``models`` are GpuModelDef objects on the same device; they stay usable."""

    def __init__(self, models):
        self.models = list(models)
        self.lib = self.models[0].lib
        arr = (C.c_void_p * len(self.models))(*[m.h for m in self.models])
        b = C.c_void_p()
        self.lib.check(self.lib.lib.tg_batch_create(arr, len(self.models), C.byref(b)))
        self.b = b

    def close(self):
        if getattr(self, "b", None):
            self.lib.lib.tg_batch_destroy(self.b)
            self.b = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sizes(self):
        n, c, u = C.c_int(), C.c_int(), C.c_int()
        self.lib.check(self.lib.lib.tg_batch_sizes(self.b, C.byref(n), C.byref(c), C.byref(u)))
        return n.value, c.value, u.value

    def execute(self, i0, nsteps, record=True, record_vars=True):
        """MALA::execute for every star of the group; returns the list of per-star record dicts (or None)."""
        outs, arr = None, None
        if record:
            pairs = [GpuMALA._records(m, nsteps, record_vars) for m in self.models]
            outs = [p[0] for p in pairs]
            self._recs = [p[1] for p in pairs]
            arr = (C.POINTER(_Records) * len(pairs))(*[C.pointer(r) for r in self._recs])
        self.lib.check(self.lib.lib.tg_batch_run(self.b, i0, nsteps, arr))
        for m in self.models:
            m.refresh_from_device()
        return outs

    def kernel_time(self, reset=0):
        n, ms, tot = C.c_long(0), C.c_double(0), C.c_long(0)
        self.lib.check(self.lib.lib.tg_batch_kernel_time(self.b, reset, C.byref(n), C.byref(ms), C.byref(tot)))
        return n.value, ms.value, tot.value

    def prepass_time(self):
        ms = C.c_double(0)
        self.lib.check(self.lib.lib.tg_batch_prepass_time(self.b, C.byref(ms)))
        return ms.value

    def last_run_ms(self):
        ms = C.c_double(0)
        self.lib.check(self.lib.lib.tg_batch_last_run_ms(self.b, C.byref(ms)))
        return ms.value

    def timer_start(self):
        self.lib.check(self.lib.lib.tg_batch_event_record(self.b, 0))

    def timer_stop_ms(self):
        self.lib.check(self.lib.lib.tg_batch_event_record(self.b, 1))
        ms = C.c_double(0)
        self.lib.check(self.lib.lib.tg_event_elapsed_ms(self.models[0].h, C.byref(ms)))
        return ms.value
