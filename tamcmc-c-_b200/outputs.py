"""Host-side mirror of the reference's ``Outputs`` class for the records of the hot path
(tamcmc/sources/outputs.cpp; C interface: include/tamcmc_outputs.h).  Marshalling only: the files
are written by libtamcmc_gpu.so's host code (csrc/tg_outputs.cpp)."""
import ctypes as C
import os

import numpy as np

from .gpu import _Records, _p, TgError, load_library


class _OutDesc(C.Structure):
    _fields_ = [("params_root", C.c_char_p), ("stat_root", C.c_char_p), ("ptemp_root", C.c_char_p), ("restore_root", C.c_char_p),
                ("Nsamples", C.c_long), ("Nsamples_sofar", C.c_long), ("erase_old_files", C.c_int),
                ("Nchains", C.c_int), ("Nvars", C.c_int), ("Ncons", C.c_int), ("Np", C.c_int), ("n_plength", C.c_int),
                ("relax", C.c_void_p), ("plength", C.c_void_p), ("variable_names", C.POINTER(C.c_char_p)),
                ("constant_names", C.POINTER(C.c_char_p)), ("constant_values", C.c_void_p), ("Tcoefs", C.c_void_p), ("file_format", C.c_int)]


def _check(lib, rc):
    if rc != 0:
        raise TgError("tamcmc outputs error %d: %s" % (rc, lib.lib.tg_out_last_error().decode()))


class Outputs:
    """``Outputs`` of one phase of one star: ``<outdir>/<root>_params_chain-<m>.bin`` (+ ``.hdr``),
    ``<root>_stat_criteria.bin``, ``<root>_parallel_tempering.bin`` and the restore triplet
    ``<restore_dir>/<root>_restore_{1,2,3}.dat`` -- the names main.cpp/config_presets.cpp give them."""

    def __init__(self, prob, Tcoefs, outdir, root="run", restore_dir=None, Nsamples=0, Nsamples_sofar=0, erase_old_files=True, lib=None, file_format="binary"):
        self.lib = lib or load_library()
        os.makedirs(outdir, exist_ok=True)
        restore_dir = restore_dir or outdir
        os.makedirs(restore_dir, exist_ok=True)
        self.paths = dict(params=os.path.join(outdir, root + "_params"), stat=os.path.join(outdir, root + "_stat_criteria"),
                          ptemp=os.path.join(outdir, root + "_parallel_tempering"), restore=os.path.join(restore_dir, root + "_restore_"))
        idx = prob.index_to_relax
        cons = np.flatnonzero(np.asarray(prob.relax) == 0)
        names = list(prob.names) if prob.names is not None else ["p%d" % i for i in range(prob.params.size)]
        vn = [names[i].encode() for i in idx]
        cn = [names[i].encode() for i in cons]
        self._keep = dict(relax=np.ascontiguousarray(prob.relax, dtype=np.int32), plength=np.ascontiguousarray(prob.plength, dtype=np.int32),
                          cvals=np.ascontiguousarray(prob.params[cons], dtype=np.float64), T=np.ascontiguousarray(Tcoefs, dtype=np.float64),
                          vn=(C.c_char_p * len(vn))(*vn), cn=(C.c_char_p * max(len(cn), 1))(*cn))
        k = self._keep
        d = _OutDesc(self.paths["params"].encode(), self.paths["stat"].encode(), self.paths["ptemp"].encode(), self.paths["restore"].encode(),
                     int(Nsamples), int(Nsamples_sofar), 1 if erase_old_files else 0, len(k["T"]), len(vn), len(cn), int(prob.params.size),
                     int(k["plength"].size), _p(k["relax"]), _p(k["plength"]), k["vn"], k["cn"] if cn else None, _p(k["cvals"]), _p(k["T"]),
                     {"binary": 0, "text": 1, "debug": 2}[file_format])
        self.h = C.c_void_p()
        _check(self.lib, self.lib.lib.tg_out_open(C.byref(d), C.byref(self.h)))
        self.Nchains, self.Nvars = len(k["T"]), len(vn)

    def append(self, rec):
        """rec: the dict returned by ``GpuMALA.execute`` (Outputs::update_buffer_* + write_bin_*)."""
        n = rec["logL"].shape[0]
        E = rec["swap_attempt"].shape[1]
        keys = ("vars", "logL", "logPrior", "logPost", "Pmove", "moved", "swap_attempt", "swap_chain0", "swap_Pswap", "swap_done", "status")
        hold = [np.ascontiguousarray(rec[key]) if rec.get(key) is not None else None for key in keys]     # keeps any copy alive over the call
        r = _Records(*[_p(a) for a in hold])
        _check(self.lib, self.lib.lib.tg_out_append(self.h, C.byref(r), n, E))

    def write_restore(self, model_def, iteration):
        """Outputs::write_buffer_restore from the device state of a GpuModelDef."""
        mu, sigma, cov = model_def.get_proposal()
        v = np.ascontiguousarray(model_def.vars, dtype=np.float64)
        _check(self.lib, self.lib.lib.tg_out_write_restore(self.h, int(iteration), _p(v), _p(mu), _p(sigma), _p(cov)))

    def write_restore_arrays(self, iteration, vars_, mu, sigma, cov, means=None):
        """means: dict(vars, mu, sigma, covarmat) as GpuModelDef.get_state_means returns it -> the *_mean blocks."""
        f = lambda a: np.ascontiguousarray(a, dtype=np.float64)
        v, m, s, c = f(vars_), f(mu), f(sigma), f(cov)
        if means is None:
            _check(self.lib, self.lib.lib.tg_out_write_restore(self.h, int(iteration), _p(v), _p(m), _p(s), _p(c)))
        else:
            mv, mm, ms, mc = f(means["vars"]), f(means["mu"]), f(means["sigma"]), f(means["covarmat"])
            _check(self.lib, self.lib.lib.tg_out_write_restore_means(self.h, int(iteration), _p(v), _p(m), _p(s), _p(c), _p(mv), _p(mm), _p(ms), _p(mc)))

    def close(self):
        if self.h:
            self.lib.lib.tg_out_close(self.h)
            self.h = None


def read_restore(restore_root, do_restore_last_index=True, lib=None):
    """Config::read_restore_files (config.cpp:1321-1579): dict(iteration, vars, mu, sigma, covarmat)."""
    lib = lib or load_library()
    Cn, Nv, it = C.c_int(0), C.c_int(0), C.c_long(0)
    root = restore_root.encode()
    _check(lib, lib.lib.tg_restore_read(root, int(do_restore_last_index), C.byref(Cn), C.byref(Nv), C.byref(it), None, None, None, None))
    v, mu, sg, cov = np.empty((Cn.value, Nv.value)), np.empty((Cn.value, Nv.value)), np.empty(Cn.value), np.empty((Cn.value, Nv.value, Nv.value))
    _check(lib, lib.lib.tg_restore_read(root, int(do_restore_last_index), C.byref(Cn), C.byref(Nv), C.byref(it), _p(v), _p(mu), _p(sg), _p(cov)))
    return dict(iteration=it.value, vars=v, mu=mu, sigma=sg, covarmat=cov)


def read_restore_means(restore_root, do_restore_last_index=True, lib=None):
    """The *_mean blocks of the same files (MALA::restore_proposal with do_restore_proposal_mean = 1, MALA.cpp:206-232)."""
    lib = lib or load_library()
    Cn, Nv, it = C.c_int(0), C.c_int(0), C.c_long(0)
    root = restore_root.encode()
    _check(lib, lib.lib.tg_restore_read_means(root, int(do_restore_last_index), C.byref(Cn), C.byref(Nv), C.byref(it), None, None, None, None))
    v, mu, sg, cov = np.empty((Cn.value, Nv.value)), np.empty((Cn.value, Nv.value)), np.empty(Cn.value), np.empty((Cn.value, Nv.value, Nv.value))
    _check(lib, lib.lib.tg_restore_read_means(root, int(do_restore_last_index), C.byref(Cn), C.byref(Nv), C.byref(it), _p(v), _p(mu), _p(sg), _p(cov)))
    return dict(iteration=it.value, vars=v, mu=mu, sigma=sg, covarmat=cov)
