"""Host-side mirror of the reference's ``Config`` / ``Config_presets`` front end (include/tamcmc_frontend.h).
Marshalling only: the files are parsed by the host code of libtamcmc_gpu.so (csrc/tg_frontend.cpp)."""
import ctypes as C

import numpy as np

from .gpu import _Problem, _p, TgError, load_library


class _Run(C.Structure):
    _fields_ = [(k, C.c_char_p) for k in ("star_id", "phase", "model_fct_name", "likelihood_fct_name", "prior_fct_name", "dir_out", "restore_dir",
                                          "diags_dir", "output_root_name", "restore_file_in", "restore_file_out", "params_txt_fileout",
                                          "stat_txt_fileout", "parallel_tempering_txt_fileout", "acceptance_txt_fileout", "file_evidence_diags")] + \
               [("Nsamples", C.c_long), ("Nbuffer", C.c_long)] + \
               [(k, C.c_int) for k in ("erase_old_files", "do_restore_proposal", "do_restore_variables", "do_restore_last_index", "do_restore_proposal_mean",
                                       "get_params", "get_statcriteria", "get_parallel_tempering", "evidence_diags", "evidence_interpolation_factor",
                                       "Nslices", "Nvars", "Ncons")] + [("lambda_temp", C.c_double), ("file_format", C.c_int)]


class FrontEnd:
    """``Config`` + ``Config_presets`` of the reference for a working directory that holds ``Config/``."""

    def __init__(self, workdir, lib=None):
        self.lib = lib or load_library()
        self.h = C.c_void_p()
        self._check(self.lib.lib.tg_fe_open(str(workdir).encode(), C.byref(self.h)))

    def _check(self, rc):
        if rc != 0:
            raise TgError("tamcmc front end error %d: %s" % (rc, self.lib.lib.tg_fe_last_error().decode()))

    def counts(self):
        a, b, c, d = C.c_int(0), C.c_int(0), C.c_int(0), C.c_int(0)
        self._check(self.lib.lib.tg_fe_counts(self.h, C.byref(a), C.byref(b), C.byref(c), C.byref(d)))
        return dict(n_stars=a.value, n_phases=b.value, first_phase=c.value, last_phase=d.value)

    def nslices(self, star):
        n = C.c_int(0)
        self._check(self.lib.lib.tg_fe_nslices(self.h, int(star), C.byref(n)))
        return n.value

    def setup(self, star, slice_ind, phase):
        """apply_presets + Config::setup (0-based indices); returns the run description as a dict."""
        self._check(self.lib.lib.tg_fe_setup(self.h, int(star), int(slice_ind), int(phase)))
        r = _Run()
        self._check(self.lib.lib.tg_fe_run_info(self.h, C.byref(r)))
        return {k: (getattr(r, k).decode() if isinstance(getattr(r, k), bytes) else getattr(r, k)) for k, _ in _Run._fields_}

    def raw_problem(self):
        p = _Problem()
        self._check(self.lib.lib.tg_fe_problem(self.h, C.byref(p)))
        return p

    def problem(self):
        """The problem of the last setup as a ``workloads.Problem`` plus the matching ``MalaConfig`` (copies)."""
        from . import workloads
        from .gpu import MalaConfig
        p = self.raw_problem()
        arr = lambda ptr, n, t: np.ctypeslib.as_array(C.cast(ptr, C.POINTER(t)), shape=(n,)).copy()
        Np, Nx, Cn = p.Np, p.Nx, p.Nchains
        relax = arr(p.relax, Np, C.c_int)
        names = [self.lib.lib.tg_fe_input_name(self.h, i).decode() for i in range(Np)]
        nv = int(relax.sum())
        prob = workloads.Problem(name="front end", x=arr(p.x, Nx, C.c_double), y=arr(p.y, Nx, C.c_double), model_case=p.model_case, params=arr(p.params0, Np, C.c_double),
                                 plength=np.array(list(p.plength), dtype=np.int32), relax=relax.astype(np.int32),
                                 priors=arr(p.priors, 4 * Np, C.c_double).reshape(Np, 4), priors_switch=arr(p.priors_switch, Np, C.c_int).astype(np.int32),
                                 extra_priors=np.array(list(p.extra_priors)), prior_case=p.prior_case, like_case=p.likelihood_case,
                                 likelihood_params=p.likelihood_params, names=names)
        prob.sigma_y = arr(p.sigma_y, Nx, C.c_double)
        # initial errors A*x+B of the free variables, already evaluated by the front end: A = 0, B = error
        err = np.ones(Np)
        err[relax == 1] = arr(p.init_err, nv, C.c_double)
        prob.err_frac, prob.err_off = np.zeros(Np), err
        nl = p.n_learn
        cfg = MalaConfig(Nchains=Cn, Nt_learn=list(arr(p.Nt_learn, nl + 1, C.c_long)), periods_learn=list(arr(p.periods_learn, nl, C.c_long)), c0=p.c0,
                         epsilon1=p.epsilon1, epsilon2=p.epsilon2, A1=p.A1, target_acceptance=p.target_acceptance, dN_mixing=p.dN_mixing)
        cfg.Tcoefs = arr(p.Tcoefs, Cn, C.c_double)
        return prob, cfg

    def dump(self):
        n = self.lib.lib.tg_fe_dump(self.h, None, 0)
        if n < 0:
            raise TgError("tg_fe_dump: no setup yet")
        buf = C.create_string_buffer(n + 1)
        self.lib.lib.tg_fe_dump(self.h, buf, n + 1)
        return buf.value.decode()

    def close(self):
        if self.h:
            self.lib.lib.tg_fe_close(self.h)
            self.h = None


def evidence(Tcoefs, logL, interpolation_factor=1000, lib=None):
    """Diagnostics::evidence_calc (diagnostics.cpp:924-956): (L_beta[C], evidence) from the tempered logL records [nsamples][C]."""
    lib = lib or load_library()
    T = np.ascontiguousarray(Tcoefs, dtype=np.float64)
    L = np.ascontiguousarray(logL, dtype=np.float64)
    Lb, ev = np.empty(T.size), C.c_double(0)
    rc = lib.lib.tg_evidence(_p(T), T.size, _p(L), L.shape[0], int(interpolation_factor), _p(Lb), C.byref(ev))
    if rc != 0:
        raise TgError(lib.lib.tg_fe_last_error().decode())
    return Lb, ev.value


def acceptance(moved, Nbuffer=None, Ncopy=0, Nsamples_init=0, lib=None):
    """Outputs::reject_rate (outputs.cpp:1824-1843): (x-axis value, accepted moves / Nbuffer per chain) of one block of moves [nsteps][C]."""
    lib = lib or load_library()
    m = np.ascontiguousarray(moved, dtype=np.uint8)
    rate, x = np.empty(m.shape[1]), C.c_double(0)
    rc = lib.lib.tg_acceptance(_p(m), m.shape[0], m.shape[1], int(Nbuffer or m.shape[0]), int(Ncopy), int(Nsamples_init), _p(rate), C.byref(x))
    if rc != 0:
        raise TgError(lib.lib.tg_fe_last_error().decode())
    return x.value, rate
