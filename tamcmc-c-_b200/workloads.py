"""Synthetic spectra and parameter vectors of SURVEY.md section 8(d) (host-side harness, numpy only).

A ``Problem`` is what the reference's ``Config::setup`` leaves in ``Config.modeling.inputs`` /
``Config.data.data`` (config.cpp:80-189): the flat parameter vector and its ``plength[11]``
layout (io_ms_global.cpp:996-1050, io_local.cpp:1059-1120), ``relax``, the 4 x Np prior matrix,
the prior switches (Config/default/primepriors_ctrl.list), ``extra_priors[4]`` and the
``Data{x, y, sigma_y}`` triple -- i.e. exactly what ``tg_create`` uploads.
"""
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

# Config/default/primepriors_ctrl.list
PRIOR_NONE, PRIOR_UNIFORM, PRIOR_GAUSSIAN, PRIOR_JEFFREYS = 0, 1, 2, 4
PRIOR_UG, PRIOR_GU, PRIOR_GUG, PRIOR_UNIFORM_ABS, PRIOR_JEFFREYS_ABS, PRIOR_AUTO = 5, 6, 7, 8, 10, 11

# Config/default/models_ctrl.list
MODEL_CASES = {
    "model_Test_Gaussian": 0, "model_Harvey_Gaussian": 1,
    "model_MS_Global_a1etaa3_HarveyLike": 2, "model_MS_Global_a1etaa3_HarveyLike_Classic": 3,
    "model_MS_Global_a1etaa3_Harvey1985": 4, "model_MS_Global_a1acta3_Harvey1985": 5,
    "model_MS_Global_a1l_etaa3_HarveyLike": 6, "model_MS_Global_a1n_etaa3_HarveyLike": 7,
    "model_MS_Global_a1nl_etaa3_HarveyLike": 8, "model_MS_Global_a1etaa3_AppWidth_HarveyLike_v1": 9,
    "model_MS_Global_a1etaa3_AppWidth_HarveyLike_v2": 10, "model_MS_local_basic": 11,
    "model_MS_Global_a1etaa3_HarveyLike_Classic_v2": 12, "model_MS_Global_a1etaa3_HarveyLike_Classic_v3": 13,
    "model_MS_local_Hnlm": 14,
}
PRIOR_CASES = {"priors_Test_Gaussian": 0, "priors_Harvey_Gaussian": 1, "io_MS_Global": 2, "io_local": 3}
LIKELIHOOD_CASES = {"chi(2,2p)": 0, "chi_square": 1}


@dataclass
class Problem:
    name: str
    model_case: int
    like_case: int
    prior_case: int
    plength: np.ndarray          # int32[11]
    params: np.ndarray           # float64[Np] initial vector
    relax: np.ndarray            # int32[Np]   1 = free
    priors: np.ndarray           # float64[Np,4]  == column-major 4 x Np of Input_Data.priors
    priors_switch: np.ndarray    # int32[Np]
    extra_priors: np.ndarray     # float64[4]
    x: np.ndarray
    y: np.ndarray
    sigma_y: Optional[np.ndarray] = None
    likelihood_params: float = 1.0
    names: List[str] = field(default_factory=list)
    err_frac: Optional[np.ndarray] = None    # errors_default.cfg: sigma_i = A*x_i + B  (MALA.cpp:253-258)
    err_off: Optional[np.ndarray] = None

    @property
    def Np(self):
        return int(self.params.size)

    @property
    def index_to_relax(self):
        return np.nonzero(self.relax == 1)[0].astype(np.int32)

    @property
    def Nv(self):
        return int((self.relax == 1).sum())

    def init_errors(self):
        """Initial proposal std-devs of the free variables, MALA::init_proposal (MALA.cpp:246-289)."""
        idx = self.index_to_relax
        return self.err_frac[idx] * self.params[idx] + self.err_off[idx]


class _Builder:
    """Appends named parameter blocks the way IO_models::add_param does (io_models.cpp)."""

    def __init__(self):
        self.v, self.relax, self.pri, self.sw, self.names, self.ea, self.eb = [], [], [], [], [], [], []

    def add(self, name, value, relax, kind=PRIOR_NONE, p=(0, 0, 0, 0), err=(0.0, 1.0)):
        p = list(p) + [-9999.0] * (4 - len(p))
        self.v.append(float(value)); self.relax.append(int(relax)); self.pri.append(p)
        self.sw.append(int(kind) if relax else PRIOR_NONE); self.names.append(name)
        self.ea.append(err[0]); self.eb.append(err[1])

    def arrays(self):
        return (np.array(self.v), np.array(self.relax, dtype=np.int32), np.array(self.pri, dtype=np.float64),
                np.array(self.sw, dtype=np.int32), list(self.names), np.array(self.ea), np.array(self.eb))


# errors_default.cfg (A, B) pairs
_ERR = dict(freq=(0.0, 0.07), height=(0.02, 0.01), width=(0.015, 0.005), v1=(0, 0.1), v2=(0, 0.05), v3=(0, 0.05),
            a1=(0.02, 0.0), a1cos=(0.1, 0.05), eta=(0.2, 1e-6), a3=(0.15, 0.005), asym=(0.02, 1.0), inc=(0.0, 5.0),
            hH=(0.03, 0.0), htc=(0.015, 0.0), hp=(0.015, 0.0), n0=(0.015, 0.0002))


def philox_uniform(n, seed):
    """Counter-based U(0,1) (numpy Philox4x64), open interval."""
    g = np.random.Generator(np.random.Philox(key=seed))
    u = g.random(n)
    return np.clip(u, 1e-300, 1.0)


def _chi2_spectrum(model, seed):
    """y_i = m_i * (-ln u_i): chi^2 2-dof noise around the limit spectrum."""
    return model * (-np.log(philox_uniform(model.size, seed)))


def s_local(Nx=10000, seed=12345, make_y=None):
    """S-local (cfg2): model_MS_local_basic, 3 modes l=0,1,2, white noise only."""
    x = 1000.0 + np.arange(Nx) * 0.01
    modes = [(0, 1030.0, 10.0, 1.0), (1, 1055.0, 15.0, 1.0), (2, 1024.0, 5.0, 1.0)]
    return _local_problem("S-local", x, modes, a1=1.0, inc=45.0, N0=1.0, trunc_c=20.0, seed=seed, make_y=make_y)


def _local_problem(name, x, modes, a1, inc, N0, trunc_c, seed, make_y, y=None, asym=0.0, do_amp=0):
    nfl = [sum(1 for m in modes if m[0] == l) for l in range(4)]
    b = _Builder()
    for l in range(4):
        for m in modes:
            if m[0] == l:
                b.add("Height_l", m[2], 1, PRIOR_JEFFREYS, (1.0, 10000.0), _ERR["height"])
    for l in range(4):
        for m in modes:
            if m[0] == l:
                b.add("Frequency_l", m[1], 1, PRIOR_UNIFORM, (m[1] - 3.0, m[1] + 3.0), _ERR["freq"])
    ci, si = np.sqrt(a1) * np.cos(np.deg2rad(inc)), np.sqrt(a1) * np.sin(np.deg2rad(inc))
    b.add("Splitting_a1", 0.0, 0)
    b.add("Asphericity_eta", 0.0, 0)
    b.add("Splitting_a3", 0.0, 0)
    b.add("sqrt(splitting_a1).cosi", ci, 1, PRIOR_UNIFORM, (0.0, 2.5), _ERR["a1cos"])
    b.add("sqrt(splitting_a1).sini", si, 1, PRIOR_UNIFORM, (0.0, 2.5), _ERR["a1cos"])
    b.add("Lorentzian_asymetry", asym, 1 if asym != 0 else 0, PRIOR_JEFFREYS_ABS, (5.0, 200.0), _ERR["asym"])
    for l in range(4):
        for m in modes:
            if m[0] == l:
                b.add("Width_l", m[3], 1, PRIOR_JEFFREYS, (x[1] - x[0], 20.0), _ERR["width"])
    b.add("White_Noise_N0", N0, 1, PRIOR_GAUSSIAN, (N0, 0.1 * N0), _ERR["n0"])
    b.add("Inclination", 0.0, 0)
    b.add("Truncation parameter", trunc_c, 0)
    b.add("Switch for fit of Amplitudes or Heights", float(do_amp), 0)
    v, relax, pri, sw, names, ea, eb = b.arrays()
    nf = sum(nfl)
    plength = np.array([nf, 0, nfl[0], nfl[1], nfl[2], nfl[3], 6, nf, 1, 1, 2], dtype=np.int32)
    assert plength.sum() == v.size
    prob = Problem(name=name, model_case=11, like_case=0, prior_case=3, plength=plength, params=v, relax=relax,
                   priors=pri, priors_switch=sw, extra_priors=np.array([0.0, 2.0, 0.2, 0.0]),
                   x=x, y=np.ones_like(x) if y is None else y, names=names, err_frac=ea, err_off=eb)
    if y is None and make_y is not None:
        prob.y = _chi2_spectrum(make_y(prob), seed)
    return prob


def s_global(Nx=1_000_000, model_case=3, seed=12345, make_y=None, step=0.0085, norders=14, lmax=3,
             numax=2700.0, dnu=120.186, nu_shift=0.0):
    """S-global (cfg3/4/5): 14 orders x l=0..3, a1/eta/a3 + Harvey-like noise, ~106 parameters.

    model_case 3 (``_Classic``: a1 and inclination direct) or 2 (sqrt(a1)cos i / sqrt(a1)sin i).
    """
    assert model_case in (2, 3)
    x = (np.arange(Nx) + 1.0) * step
    eps_dnu = 42.5632
    n0 = int(round((numax - eps_dnu) / dnu)) - norders // 2 + 1
    n = np.arange(n0, n0 + norders)
    f0 = n * dnu + eps_dnu + nu_shift
    f = [f0, f0 + dnu / 2.0 - 3.0, f0 - 8.0, f0 + dnu / 2.0 - 3.0 - 14.0]
    W = np.linspace(1.0, 3.0, norders)
    H = 1.3 * np.exp(-0.5 * ((f0 - numax) / (2.5 * dnu)) ** 2)
    V = (1.5, 0.53, 0.08)
    a1, inc = 1.4, 55.0
    # Asphericity_eta "Fix_Auto": io_ms_global.cpp:322,:743
    G, Dnl, Dnu_sun, R_sun, M_sun = 6.667e-8, 0.75, 135.1, 6.96342e5, 1.98855e30
    rho_sun = M_sun * 1e3 / (4 * np.pi * (R_sun * 1e5) ** 3 / 3)
    rho = (dnu / Dnu_sun) ** 2 * rho_sun
    eta = (4.0 / 3.0) * np.pi * Dnl * (a1 * 1e-6) ** 2 / (rho * G)
    b = _Builder()
    for k in range(norders):
        b.add("Height_l", H[k], 1, PRIOR_JEFFREYS, (1.0, 1000.0), _ERR["height"])
    for l in range(1, lmax + 1):
        b.add("Visibility_l%d" % l, V[l - 1], 1, PRIOR_GAUSSIAN, (V[l - 1], (0.15, 0.05, 0.02)[l - 1]), _ERR["v%d" % l])
    for l in range(lmax + 1):
        for k in range(norders):
            b.add("Frequency_l", f[l][k], 1, PRIOR_GUG, (f[l][k] - 4.0, f[l][k] + 4.0, 0.5, 0.5), _ERR["freq"])
    if model_case == 3:
        b.add("Splitting_a1", a1, 1, PRIOR_UNIFORM, (0.0, 5.0), _ERR["a1"])
    else:
        b.add("Splitting_a1", 0.0, 0)
    b.add("Asphericity_eta", eta, 0)
    b.add("Splitting_a3", 0.0, 0)
    if model_case == 3:
        b.add("Empty", 0.0, 0)
        b.add("Empty", 0.0, 0)
    else:
        b.add("sqrt(splitting_a1).cosi", np.sqrt(a1) * np.cos(np.deg2rad(inc)), 1, PRIOR_UNIFORM, (0.0, 2.5), _ERR["a1cos"])
        b.add("sqrt(splitting_a1).sini", np.sqrt(a1) * np.sin(np.deg2rad(inc)), 1, PRIOR_UNIFORM, (0.0, 2.5), _ERR["a1cos"])
    b.add("Lorentzian_asymetry", 0.0, 0)
    for k in range(norders):
        b.add("Width_l", W[k], 1, PRIOR_JEFFREYS, (step, dnu / 3.0), _ERR["width"])
    noise = [(0.0, 0.0, 1.0), (11.049588, 49.669854, 4.0), (0.93569041, 1.3516447, 2.0)]
    nerr = [(0, 0, 0), (0.30, 0.57, 0), (0.0032, 0.0039, 0)]
    for (h, tc, p), (eh, etc_, _) in zip(noise, nerr):
        free = 1 if tc != 0 else 0
        b.add("Harvey-Noise_H", h, free, PRIOR_GAUSSIAN, (h, max(eh, 1e-3)), _ERR["hH"])
        b.add("Harvey-Noise_tc", tc, free, PRIOR_GAUSSIAN, (tc, max(etc_, 1e-3)), _ERR["htc"])
        b.add("Harvey-Noise_p", p, 0)
    b.add("White_Noise_N0", 0.13392108, 1, PRIOR_GAUSSIAN, (0.13392108, 0.00034), _ERR["n0"])
    if model_case == 3:
        b.add("Inclination", inc, 1, PRIOR_UNIFORM, (0.0, 90.0), _ERR["inc"])
    else:
        b.add("Inclination", 0.0, 0)
    b.add("Truncation parameter", 20.0, 0)
    b.add("Switch for fit of Amplitudes or Heights", 0.0, 0)
    v, relax, pri, sw, names, ea, eb = b.arrays()
    plength = np.array([norders, lmax] + [norders if l <= lmax else 0 for l in range(4)] + [6, norders, 10, 1, 2],
                       dtype=np.int32)
    assert plength.sum() == v.size, (plength.sum(), v.size)
    prob = Problem(name="S-global", model_case=model_case, like_case=0, prior_case=2, plength=plength, params=v,
                   relax=relax, priors=pri, priors_switch=sw, extra_priors=np.array([1.0, 2.0, 0.2, 0.0]),
                   x=x, y=np.ones_like(x), names=names, err_frac=ea, err_off=eb)
    if make_y is not None:
        prob.y = _chi2_spectrum(make_y(prob), seed)
    return prob


def tcoefs(C, lam=1.7):
    """T_m = lambda^m (MALA.cpp:103)."""
    return np.power(lam, np.arange(C, dtype=np.float64))


def replicated_tcoefs(C, per_ensemble=8, lam=1.7):
    """cfg4: C chains = replicated tempered ensembles of ``per_ensemble`` temperatures each."""
    return np.power(lam, (np.arange(C) % per_ensemble).astype(np.float64))


def perturbed_params(prob, C, scale=0.3, seed=7):
    """C parameter vectors around prob.params: free variables moved by scale * initial error."""
    g = np.random.Generator(np.random.Philox(key=seed))
    P = np.tile(prob.params, (C, 1))
    idx = prob.index_to_relax
    err = prob.init_errors()
    P[:, idx] += scale * err[None, :] * g.standard_normal((C, idx.size))
    P[0] = prob.params
    return P


# ---------------------------------------------------------------------------------------------
# Small problems covering every live model case of Config/default/models_ctrl.list (parity tests)
# ---------------------------------------------------------------------------------------------
def small_global(model_case, Nx=6000, step=0.05, norders=4, lmax=3, asym=0.0, do_amp=0, trunc_c=20.0,
                 a3=0.0, x0=900.0, seed=3):
    """A reduced global fit (few orders) for model cases 2,3,6,7,8,9,10,12,13."""
    assert model_case in (2, 3, 6, 7, 8, 9, 10, 12, 13)
    g = np.random.Generator(np.random.Philox(key=seed + model_case))
    x = x0 + np.arange(Nx) * step
    dnu = 55.0
    f0 = x0 + 40.0 + dnu * np.arange(norders) + g.normal(0, 0.3, norders)
    f = [f0, f0 + dnu / 2 - 2.0 + g.normal(0, 0.2, norders), f0 - 5.0 + g.normal(0, 0.2, norders),
         f0 + dnu / 2 - 9.0 + g.normal(0, 0.2, norders)]
    W = np.linspace(0.6, 2.4, norders)            # straddles the Gamma=1 branch of the window logic
    H = 20.0 * np.exp(-0.5 * ((f0 - f0.mean()) / 80.0) ** 2)
    V = (1.5, 0.53, 0.08)
    a1, inc, eta = 1.3, 52.0, 3e-4
    b = _Builder()
    for k in range(norders):
        b.add("Height_l", H[k], 1, PRIOR_JEFFREYS, (1.0, 1000.0), _ERR["height"])
    for l in range(1, lmax + 1):
        b.add("Visibility_l%d" % l, V[l - 1], 1, PRIOR_GAUSSIAN, (V[l - 1], 0.1), _ERR["v%d" % l])
    for l in range(lmax + 1):
        for k in range(norders):
            b.add("Frequency_l", f[l][k], 1, PRIOR_GUG, (f[l][k] - 2.0, f[l][k] + 2.0, 0.3, 0.4), _ERR["freq"])
    # Snlm block: a1, eta, a3, sqrt(a1)cos i, sqrt(a1) sin i, asym [, extra a1 terms]
    direct = model_case in (3, 6, 7, 8, 12, 13)
    b.add("Splitting_a1", a1 if direct else 0.0, 1 if direct and model_case not in (7, 8) else 0, PRIOR_UNIFORM, (0.0, 5.0), _ERR["a1"])
    b.add("Asphericity_eta", eta, 1, PRIOR_UNIFORM_ABS, (0.0, 1e-2), _ERR["eta"])
    b.add("Splitting_a3", a3, 1 if a3 != 0 else 0, PRIOR_UNIFORM, (-0.2, 0.2), _ERR["a3"])
    if direct:
        b.add("Empty", 0.0, 0)
        b.add("Empty", 0.0, 0)
    else:
        b.add("sqrt(splitting_a1).cosi", np.sqrt(a1) * np.cos(np.deg2rad(inc)), 1, PRIOR_UNIFORM, (0.0, 2.5), _ERR["a1cos"])
        b.add("sqrt(splitting_a1).sini", np.sqrt(a1) * np.sin(np.deg2rad(inc)), 1, PRIOR_UNIFORM, (0.0, 2.5), _ERR["a1cos"])
    b.add("Lorentzian_asymetry", asym, 1 if asym != 0 else 0, PRIOR_JEFFREYS_ABS, (5.0, 200.0), _ERR["asym"])
    nsplit = 6
    if model_case == 6:
        b.add("Splitting_a1(l=2)", 0.8, 1, PRIOR_UNIFORM, (0.0, 5.0), _ERR["a1"]); nsplit = 7
    if model_case == 7:
        for k in range(norders):
            b.add("Splitting_a1(n)", 0.7 + 0.2 * k, 1, PRIOR_UNIFORM, (0.0, 5.0), _ERR["a1"])
        nsplit = 6 + norders
    if model_case == 8:
        for k in range(2 * norders):
            b.add("Splitting_a1(n,l)", 0.6 + 0.15 * k, 1, PRIOR_UNIFORM, (0.0, 5.0), _ERR["a1"])
        nsplit = 6 + 2 * norders
    if model_case == 9:      # numax_dip?, alpha, Gamma_alpha, Wdip, DeltaGamma_dip  (models.cpp:1408-1411)
        for nm, v in zip(("nudip", "alfa", "GammaAlfa", "Wdip", "DeltaGammadip"), (f0.mean() - 10.0, 3.0, 1.5, 400.0, 2.0)):
            b.add("Width:" + nm, v, 1, PRIOR_UNIFORM, (0.2 * v, 3.0 * v), _ERR["width"])
        nwidth = 5
    elif model_case == 10:   # numax, nudip, alpha, Gamma_alpha, Wdip, DeltaGamma_dip (models.cpp:1590-1594)
        for nm, v in zip(("numax", "nudip", "alfa", "GammaAlfa", "Wdip", "DeltaGammadip"),
                         (f0.mean(), f0.mean() - 10.0, 3.0, 1.5, 400.0, 2.0)):
            b.add("Width:" + nm, v, 1, PRIOR_UNIFORM, (0.2 * v, 3.0 * v), _ERR["width"])
        nwidth = 6
    else:
        for k in range(norders):
            b.add("Width_l", W[k], 1, PRIOR_JEFFREYS, (step, dnu / 3.0), _ERR["width"])
        nwidth = norders
    noise = [(0.0, 0.0, 1.0), (30.0, 9.0, 3.7), (8.0, 2.2, 2.0)]
    for (h, tc, p) in noise:
        free = 1 if tc != 0 else 0
        b.add("Harvey-Noise_H", h, free, PRIOR_GAUSSIAN, (h, 0.1 * max(h, 1)), _ERR["hH"])
        b.add("Harvey-Noise_tc", tc, free, PRIOR_GAUSSIAN, (tc, 0.1 * max(tc, 1)), _ERR["htc"])
        b.add("Harvey-Noise_p", p, free, PRIOR_UG, (1.0, 4.5, 0.2), _ERR["hp"])
    b.add("White_Noise_N0", 0.6, 1, PRIOR_GAUSSIAN, (0.6, 0.05), _ERR["n0"])
    if model_case == 12:     # 9 free height ratios: l=1 (m0, m1), l=2 (m0, m1, m2), l=3 (m0..m3)
        r = [0.4, 0.3, 0.3, 0.2, 0.15, 0.3, 0.15, 0.1, 0.1]
        for v in r:
            b.add("Hlm_ratio", v, 1, PRIOR_UNIFORM, (0.0, 1.0), (0.0, 0.05))
        ninc = 9
    elif model_case == 13:   # 2*Nfl1 + 3*Nfl2 + 4*Nfl3 heights
        ninc = (2 + 3 + 4) * norders
        for k in range(ninc):
            b.add("Hlm", 3.0 + 0.37 * (k % 11), 1, PRIOR_JEFFREYS, (1.0, 1000.0), _ERR["height"])
    else:
        b.add("Inclination", inc if direct else 0.0, 1 if direct else 0, PRIOR_UNIFORM, (0.0, 90.0), _ERR["inc"])
        ninc = 1
    b.add("Truncation parameter", trunc_c, 0)
    b.add("Switch for fit of Amplitudes or Heights", float(do_amp), 0)
    v, relax, pri, sw, names, ea, eb = b.arrays()
    plength = np.array([norders, lmax] + [norders if l <= lmax else 0 for l in range(4)] + [nsplit, nwidth, 10, ninc, 2],
                       dtype=np.int32)
    assert plength.sum() == v.size, (plength.sum(), v.size)
    extra = np.array([1.0, 2.0, 0.2, 1.0 if model_case == 12 else 0.0])
    return Problem(name="small-global-%d" % model_case, model_case=model_case, like_case=0, prior_case=2,
                   plength=plength, params=v, relax=relax, priors=pri, priors_switch=sw, extra_priors=extra,
                   x=x, y=np.ones_like(x), names=names, err_frac=ea, err_off=eb)


def small_local(model_case=11, Nx=1500, step=0.02, asym=0.0, do_amp=0, trunc_c=20.0, x0=95.0):
    """A local fit with l=0,1,2,3 modes for model cases 11 (basic) and 14 (Hnlm)."""
    x = x0 + np.arange(Nx) * step
    span = Nx * step
    modes = [(0, x0 + 0.30 * span, 900.0, 0.31), (0, x0 + 0.75 * span, 500.0, 1.2), (1, x0 + 0.52 * span, 700.0, 0.45),
             (2, x0 + 0.22 * span, 400.0, 0.33), (3, x0 + 0.62 * span, 90.0, 0.38)]
    if model_case == 11:
        return _local_problem("small-local-11", x, modes, a1=0.45, inc=63.0, N0=110.0, trunc_c=trunc_c, seed=5,
                              make_y=None, asym=asym, do_amp=do_amp)
    nfl = [sum(1 for m in modes if m[0] == l) for l in range(4)]
    b = _Builder()
    for l in range(4):
        for m in modes:
            if m[0] == l:
                for k in range(l + 1):
                    b.add("Height_lm", m[2] / (1.0 + k), 1, PRIOR_JEFFREYS, (1.0, 10000.0), _ERR["height"])
    for l in range(4):
        for m in modes:
            if m[0] == l:
                b.add("Frequency_l", m[1], 1, PRIOR_UNIFORM, (m[1] - 3.0, m[1] + 3.0), _ERR["freq"])
    b.add("Splitting_a1", 0.45, 1, PRIOR_UNIFORM, (0.0, 1.5), _ERR["a1"])
    b.add("Asphericity_eta", 0.0, 0)
    b.add("Splitting_a3", 0.01, 1, PRIOR_UNIFORM, (-0.05, 0.05), _ERR["a3"])
    b.add("Empty", 0.0, 0); b.add("Empty", 0.0, 0)
    b.add("Lorentzian_asymetry", asym, 1 if asym != 0 else 0, PRIOR_JEFFREYS_ABS, (5.0, 200.0), _ERR["asym"])
    for l in range(4):
        for m in modes:
            if m[0] == l:
                b.add("Width_l", m[3], 1, PRIOR_JEFFREYS, (step, 20.0), _ERR["width"])
    b.add("White_Noise_N0", 110.0, 1, PRIOR_GAUSSIAN, (110.0, 11.0), _ERR["n0"])
    b.add("Inclination", 0.0, 0)
    b.add("Truncation parameter", trunc_c, 0)
    b.add("Switch for fit of Amplitudes or Heights", float(do_amp), 0)
    v, relax, pri, sw, names, ea, eb = b.arrays()
    nh = nfl[0] + 2 * nfl[1] + 3 * nfl[2] + 4 * nfl[3]
    nf = sum(nfl)
    plength = np.array([nh, 0, nfl[0], nfl[1], nfl[2], nfl[3], 6, nf, 1, 1, 2], dtype=np.int32)
    assert plength.sum() == v.size
    return Problem(name="small-local-14", model_case=14, like_case=0, prior_case=3, plength=plength, params=v,
                   relax=relax, priors=pri, priors_switch=sw, extra_priors=np.array([0.0, 2.0, 0.2, 2.0]),
                   x=x, y=np.ones_like(x), names=names, err_frac=ea, err_off=eb)


def gauss_problem(model_case=0, Nx=800, like_case=1, seed=11):
    """model_Test_Gaussian (case 0) / model_Harvey_Gaussian (case 1)."""
    g = np.random.Generator(np.random.Philox(key=seed))
    x = np.linspace(50.0, 450.0, Nx)
    b = _Builder()
    b.add("Hmax", 8.0, 1, PRIOR_UNIFORM, (0.0, 50.0), (0.1, 0.0))
    b.add("Gsigma", 30.0, 1, PRIOR_UNIFORM, (1.0, 200.0), (0.01, 1.0))
    b.add("Numax", 240.0, 1, PRIOR_UNIFORM, (100.0, 400.0), (0.01, 2.0))
    if model_case == 0:
        b.add("White_Noise_N0", 1.5, 1, PRIOR_UNIFORM, (0.0, 10.0), (0.015, 0.002))
        plength = np.array([4] + [0] * 10, dtype=np.int32)
    else:
        b.add("Harvey-Noise_H", 12.0, 1, PRIOR_UNIFORM, (0.0, 100.0), _ERR["hH"])
        b.add("Harvey-Noise_tc", 9.0, 1, PRIOR_UNIFORM, (0.0, 100.0), _ERR["htc"])
        b.add("Harvey-Noise_p", 2.3, 1, PRIOR_UNIFORM, (1.0, 5.0), _ERR["hp"])
        b.add("White_Noise_N0", 1.5, 1, PRIOR_UNIFORM, (0.0, 10.0), _ERR["n0"])
        plength = np.array([7] + [0] * 10, dtype=np.int32)
    v, relax, pri, sw, names, ea, eb = b.arrays()
    m = v[0] * np.exp(-0.5 * (x - v[2]) ** 2 / v[1] ** 2) + (v[3] if model_case == 0 else 1.5 + 12.0 / (1 + (1e-3 * 9.0 * x) ** 2.3))
    sig = 0.25 * np.ones(Nx)
    y = m + sig * g.standard_normal(Nx) if like_case == 1 else m * g.exponential(1.0, Nx)
    return Problem(name="gauss-%d" % model_case, model_case=model_case, like_case=like_case, prior_case=model_case,
                   plength=plength, params=v, relax=relax, priors=pri, priors_switch=sw,
                   extra_priors=np.zeros(4), x=x, y=y, sigma_y=sig, names=names, err_frac=ea, err_off=eb)
