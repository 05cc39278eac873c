"""numpy restatement of the reference's sampler transitions (oracle side; TEST
INFRASTRUCTURE -- see oracle/__init__.py for who may import this).

Written as plain functions over one chain (1-D float64 ``q``), following the
reference's order of floating-point operations so that, with the same model
callables and the same draws, results are bit-identical to the reference
classes (checked by tests/test_oracle_vs_reference.py when /root/reference is
present).  Random draws come from a ``replay.Draws``-like object ``rng`` with
``normal(size=)`` and ``uniform(low, high, size=)``; pass ``np.random`` to get
the reference's own behaviour.

Citations are path:line under /root/reference.
"""
import numpy as np


class Counters:
    """[Hcount, Vgcount, leapcount] -- the reference's per-step work counters
    (src/pyhmc.py:12,21-23,31-33,36)."""

    def __init__(self):
        self.H = 0
        self.Vg = 0
        self.leap = 0

    def as_list(self):
        return [self.H, self.Vg, self.leap]


class Kinetics:
    """Diagonal-metric kinetic energy, src/pyhmc.py:14-19 (same at
    src/algorithms.py:85-93, src/mala.py:13-18, src/adaptivehmc.py:14-19)."""

    def __init__(self, invmetric_diag=None):
        self.invmetric_diag = 1.0 if invmetric_diag is None else invmetric_diag
        self.metricstd = self.invmetric_diag ** -0.5

    def KE(self, p):
        return 0.5 * (p ** 2 * self.invmetric_diag).sum()

    def KE_g(self, p):
        return p * self.invmetric_diag


def _V(model, q):
    # src/pyhmc.py:10
    return model.log_prob(q) * -1.0


def _V_g(model, q, cnt):
    # src/pyhmc.py:21-23
    cnt.Vg += 1
    return model.grad_log_prob(q) * -1.0


def _H(model, kin, q, p, cnt):
    # src/pyhmc.py:31-33
    cnt.H += 1
    return _V(model, q) + kin.KE(p)


def leapfrog(model, kin, q, p, N, step_size, cnt):
    """src/pyhmc.py:35-48 (= src/algorithms.py:133-147): half kick, N-1 x (drift, kick),
    drift, half kick -> N+1 gradient evaluations."""
    cnt.leap += 1
    p = p - 0.5 * step_size * _V_g(model, q, cnt)
    for _ in range(N - 1):
        q = q + step_size * kin.KE_g(p)
        p = p - step_size * _V_g(model, q, cnt)
    q = q + step_size * kin.KE_g(p)
    p = p - 0.5 * step_size * _V_g(model, q, cnt)
    return q, p


def _bad(prob):
    return bool(np.isnan(prob) or np.isinf(prob))


# ----------------------------------------------------------------------------- vanilla HMC

def hmc_step_pyhmc(model, q, N, step_size, rng, invmetric_diag=None):
    """pyhmc.HMC.step / PyHMC.hmc_step, src/pyhmc.py:64-69,133-138 with metropolis
    src/pyhmc.py:50-61.  Codes 1./0./2.; the uniform is drawn only when prob is
    finite and q moved; reject iff u > min(1, prob)."""
    kin, cnt = Kinetics(invmetric_diag), Counters()
    with np.errstate(all="ignore"):
        p = rng.normal(size=q.size).reshape(q.shape) * kin.metricstd
        q1, p1 = leapfrog(model, kin, q, p, N, step_size, cnt)
        H0 = _H(model, kin, q, p, cnt)
        H1 = _H(model, kin, q1, p1, cnt)
        prob = np.exp(H0 - H1)
        if _bad(prob) or (q - q1).sum() == 0:
            return q, p, 2.0, [H0, H1], cnt.as_list()
        if rng.uniform(0.0, 1.0, size=1) > min(1.0, prob):
            return q, p, 0.0, [H0, H1], cnt.as_list()
        return q1, p1, 1.0, [H0, H1], cnt.as_list()


def hmc_step_alg(model, q, nleap, step_size, rng, invmetric_diag=None):
    """algorithms.HMC.step, src/algorithms.py:192-198 with metropolis :175-189.
    Codes 1./0./-1; the uniform is ALWAYS drawn (before the NaN test, :184-185)."""
    kin, cnt = Kinetics(invmetric_diag), Counters()
    with np.errstate(all="ignore"):
        p = rng.normal(size=q.size).reshape(q.shape) * kin.metricstd
        q1, p1 = leapfrog(model, kin, q, p, nleap, step_size, cnt)
        H0 = _H(model, kin, q, p, cnt)
        H1 = _H(model, kin, q1, p1, cnt)
        prob = np.exp(H0 - H1)
        u = rng.uniform(0.0, 1.0, size=1)
        if _bad(prob) or (q - q1).sum() == 0:
            return q, p, -1, [H0, H1], cnt.as_list()
        if u > min(1.0, prob):
            return q, p, 0.0, [H0, H1], cnt.as_list()
        return q1, p1, 1.0, [H0, H1], cnt.as_list()


# ----------------------------------------------------------------------------- delayed rejection

def _dr_ghost(model, kin, mj, q, p, nleap, step_size, stepfactor, H0, cnt):
    """DRHMC.get_num, src/algorithms.py:474-497: product over the mj ghost
    proposals from (q, p) of (1 - alpha'_j); recursion depth <= mj."""
    avec = np.zeros(mj)
    for j in range(mj):
        fac = stepfactor ** j
        qj, pj = leapfrog(model, kin, q, p, int(nleap * fac), step_size / fac, cnt)
        Hj = _H(model, kin, qj, pj, cnt)
        pfac = np.exp(H0 - Hj)
        if (q - qj).sum() == 0:
            pfac = 0.0
        if j:
            den = np.prod(1 - avec[:j])
            num = _dr_ghost(model, kin, j, qj, -pj, nleap, step_size, stepfactor, Hj, cnt)
            prob = pfac * num / den
        else:
            prob = pfac
        if _bad(prob):
            return 0.0
        avec[j] = min(1.0, prob)
        if not np.prod(1 - avec):
            return np.prod(1 - avec)
    return np.prod(1 - avec)


def drhmc_step(model, q, nleap, step_size, nprops, stepfactor, rng, invmetric_diag=None):
    """algorithms.DRHMC.step, src/algorithms.py:500-530.  Stage j integrates
    int(nleap*f^j) steps of size eps/f^j from (q0, p0); one bare uniform() per
    executed stage; accept iff u < alpha_j; code j+1 on accept at stage j, 0 otherwise."""
    kin, cnt = Kinetics(invmetric_diag), Counters()
    avec = np.zeros(nprops)
    with np.errstate(all="ignore"):
        q0 = q
        p0 = rng.normal(size=q0.size).reshape(q0.shape) * kin.metricstd
        H0 = _H(model, kin, q0, p0, cnt)
        for j in range(nprops):
            fac = stepfactor ** j
            qj, pj = leapfrog(model, kin, q0, p0, int(nleap * fac), step_size / fac, cnt)
            Hj = _H(model, kin, qj, pj, cnt)
            pfac = np.exp(H0 - Hj)
            if (q0 - qj).sum() == 0:
                pfac = 0.0
            den = np.prod(1 - avec[:j])
            num = _dr_ghost(model, kin, j, qj, -pj, nleap, step_size, stepfactor, Hj, cnt)
            prob = pfac * num / den
            if _bad(prob):
                prob = 0.0
            avec[j] = min(1.0, prob)
            if rng.uniform() < avec[j]:
                return qj, pj, j + 1, avec, cnt.as_list()
    return q0, p0, 0, avec, cnt.as_list()


def dr_leapfrog_work(nleap, nprops, stepfactor):
    """Worst-case leapfrog steps per stage: L_j + G(j), G(m) = sum_{j<m}(L_j + G(j))
    (SURVEY.md section 3.2)."""
    L = [int(nleap * stepfactor ** j) for j in range(nprops)]
    G = [0] * (nprops + 1)
    for m in range(1, nprops + 1):
        G[m] = sum(L[j] + G[j] for j in range(m))
    return L, G


# ----------------------------------------------------------------------------- MALA

def mala_step(model, q, eps, rng, V_q=None, V_gq=None):
    """mala.MALA.step + metropolis, src/mala.py:45-60,24-42.  Returns
    (q2, code, V_q2, V_gq2, counts); codes 1./0./2.; conditional uniform."""
    cnt = Counters()
    with np.errstate(all="ignore"):
        z = rng.normal(size=q.size).reshape(q.shape)
        if V_gq is None:
            V_gq = _V_g(model, q, cnt)
        if V_q is None:
            V_q = _V(model, q)
        q1 = q - 0.5 * eps ** 2 * V_gq + eps * z
        V_gq1 = _V_g(model, q1, cnt)
        V_q1 = _V(model, q1)
        p0 = np.exp(V_q - V_q1)
        mu0 = q - 0.5 * eps ** 2 * V_gq
        mu1 = q1 - 0.5 * eps ** 2 * V_gq1
        logden = -0.5 * np.sum((q1 - mu0) ** 2) / eps ** 2
        lognum = -0.5 * np.sum((q - mu1) ** 2) / eps ** 2
        prob = p0 * np.exp(lognum - logden)
        if _bad(prob) or (q - q1).sum() == 0:
            return q, 2.0, V_q, V_gq, cnt.as_list()
        if rng.uniform(0.0, 1.0, size=1) > min(1.0, prob):
            return q, 0.0, V_q, V_gq, cnt.as_list()
        return q1, 1.0, V_q1, V_gq1, cnt.as_list()


# ----------------------------------------------------------------------------- dual averaging

class DualAveragingStepSize:
    """src/algorithms.py:34-75 (Nesterov dual averaging; mu = log(10 eps0))."""

    def __init__(self, initial_step_size, target_accept=0.65, gamma=0.05, t0=10.0, kappa=0.75, nadapt=0):
        self.initial_step_size = initial_step_size
        self.mu = np.log(10 * initial_step_size)
        self.target_accept = target_accept
        self.gamma = gamma
        self.t = t0
        self.kappa = kappa
        self.error_sum = 0
        self.log_averaged_step = 0
        self.nadapt = nadapt

    def update(self, p_accept):
        if np.isnan(p_accept) or np.isinf(p_accept):
            p_accept = 0.0
        if p_accept > 1:
            p_accept = 1.0
        self.error_sum += self.target_accept - p_accept
        log_step = self.mu - self.error_sum / (np.sqrt(self.t) * self.gamma)
        eta = self.t ** -self.kappa
        self.log_averaged_step = eta * log_step + (1 - eta) * self.log_averaged_step
        self.t += 1
        return np.exp(log_step), np.exp(self.log_averaged_step)


def hmc_adapt_stepsize(model, q, epsadapt, nleap, step_size, rng, invmetric_diag=None):
    """algorithms.HMC.adapt_stepsize, src/algorithms.py:209-230: epsadapt+1 HMC steps
    feeding the dual-averaging update with prob in [0,1] (0 if q did not change).
    Returns (q, fixed_step_size, trace of step sizes used)."""
    da = DualAveragingStepSize(step_size)
    used = []
    for i in range(epsadapt + 1):
        qprev = q.copy()
        used.append(step_size)
        q, p, acc, Hs, count = hmc_step_alg(model, q, nleap, step_size, rng, invmetric_diag)
        with np.errstate(all="ignore"):
            prob = 0 if (qprev == q).all() else np.exp(Hs[0] - Hs[1])
        if i < epsadapt:
            if np.isnan(prob):
                prob = 0.0
            if prob > 1:
                prob = 1.0
            step_size, _ = da.update(prob)
        else:
            _, step_size = da.update(prob)
    return q, step_size, used


def hmc_sample(model, q, nsamples, burnin, step_size, nleap, rng, skipburn=True, epsadapt=0, invmetric_diag=None):
    """algorithms.HMC.sample, src/algorithms.py:233-254: optional dual-averaging warm-up, then
    nsamples+burnin steps; ``accepts`` for every iteration, samples / Hs / counts only for
    i > burnin (strict) -> nsamples-1 rows.  Returns a dict of stacked arrays and the step size used."""
    if epsadapt:
        q, step_size, _ = hmc_adapt_stepsize(model, q, epsadapt, nleap, step_size, rng, invmetric_diag)
    samples, accepts, Hs, counts = [], [], [], []
    n = 0
    for i in range(nsamples + burnin):
        q, p, acc, H, count = hmc_step_alg(model, q, nleap, step_size, rng, invmetric_diag)
        n += 1
        accepts.append(acc)
        if skipburn & (i > burnin):
            samples.append(q)
            Hs.append(H)
            counts.append(count)
    return dict(samples=np.array(samples), accepts=np.array(accepts), Hs=np.array(Hs), counts=np.array(counts), i=n,
                step_size=step_size)


def drhmc_sample(model, q, nsamples, burnin, step_size, nleap, nprops, stepfactor, rng, p=None, invmetric_diag=None):
    """algorithms.DRHMC.sample, src/algorithms.py:534-555: one (unused) normal vector is drawn when p is
    None (:538); every iteration, burn-in included, is stored."""
    if p is None:
        p = rng.normal(0, 1, q.size).reshape(q.shape)
    samples, accepts, avecs, counts = [], [], [], []
    n = 0
    for i in range(nsamples + burnin):
        q, p, acc, avec, count = drhmc_step(model, q, nleap, step_size, nprops, stepfactor, rng, invmetric_diag)
        n += 1
        samples.append(q)
        accepts.append(acc)
        avecs.append(avec)
        counts.append(count)
    return dict(samples=np.array(samples), accepts=np.array(accepts), avecs=np.array(avecs), counts=np.array(counts), i=n)


# ----------------------------------------------------------------------------- AdHMC_eps0

def _eps0_trials(model, kin, q0, p0, smin, smax, ntry, nsteps, cnt):
    """The trial sweep shared by both uses of AdHMC_eps0.get_stepsize,
    src/adaptivehmc.py:111-124: ntry log-spaced step sizes, nsteps leapfrog steps each,
    weights exp(H0-H)*sqrt(eps), normalised."""
    H0 = _H(model, kin, q0, p0, cnt)
    Hs = np.zeros(ntry)
    steps = np.logspace(np.log10(smin), np.log10(smax), ntry)
    pwts = steps.copy() ** 0.5
    for i, ss in enumerate(steps):
        q1, p1 = leapfrog(model, kin, q0, p0, nsteps, ss, cnt)
        Hs[i] = _H(model, kin, q1, p1, cnt)
    pp = np.exp(H0 - Hs) * pwts
    pp[np.isnan(pp)] = 0
    pp[np.isinf(pp)] = 0
    pp /= pp.sum()
    return steps, pp, np.cumsum(pp)


def adhmc_eps0_step(model, q0, Nleap, rng, smin=0.01, smax=1.0, ntry=10, nsteps=1, invmetric_diag=None, Tint=0):
    """adaptivehmc.AdHMC_eps0.hmc_step, src/adaptivehmc.py:138-192 (Nleap steps, or with Tint given
    int(Tint/eps)+1 steps of the drawn step size, :178-179).
    Draw order: normal(d), uniform(low=cdf.min()), uniform(steps[i], steps[i+1]),
    conditional uniform(0,1,size=1).  Returns the reference's 6-tuple."""
    kin, cnt = Kinetics(invmetric_diag), Counters()
    with np.errstate(all="ignore"):
        p0 = rng.normal(size=q0.size).reshape(q0.shape) * kin.metricstd
        H0 = _H(model, kin, q0, p0, cnt)
        steps, pp, cdf = _eps0_trials(model, kin, q0, p0, smin, smax, ntry, nsteps, cnt)
        sx = rng.uniform(low=cdf.min())
        isx = np.where(sx > cdf)[0][-1]
        ss = rng.uniform(steps[isx], steps[isx + 1])
        pf_den = pp[isx + 1]
        q1, p1 = leapfrog(model, kin, q0, p0, Nleap if Tint == 0 else int(Tint / ss) + 1, ss, cnt)
        H1 = _H(model, kin, q1, p1, cnt)
        steps_b, pp_b, _ = _eps0_trials(model, kin, q1, -p1, smin, smax, ntry, nsteps, cnt)
        pb_num = pp_b[np.where(steps_b > ss)[0][0]]
        hastings = pb_num / pf_den
        prob = np.exp(H0 - H1) * hastings
        info = [[prob, prob / hastings, hastings], np.stack([pf_den, pb_num, ss]), cnt.as_list()]
        if _bad(prob) or (q0 - q1).sum() == 0:
            return (q0, p0, 2.0, *info)
        if rng.uniform(0.0, 1.0, size=1) > min(1.0, prob):
            return (q0, p0, 0.0, *info)
        return (q1, p1, 1.0, *info)


# ----------------------------------------------------------------------------- AdHMC / AdHMC_tint
# Step size adapted inside the trajectory: the trajectory is cut into segments; after each one the
# step size is multiplied by one of `ratios`, chosen with probability proportional to e^{-dH} of
# the candidate segment, and the probability of the reverse choice is collected for the Hastings
# factor.  The two classes differ only in the length of a segment: `nleap` steps (AdHMC) or
# max(1, int(tint/eps)) steps (AdHMC_tint).

def _adhmc_prob_of(model, kin, q, p, eps, smin, smax, ntry, nsteps, logspace, cnt):
    """get_stepsize(..., eps=eps), src/adaptivehmc.py:202-224: probability of the bin above eps."""
    H0 = _H(model, kin, q, p, cnt)
    Hs = np.zeros(ntry)
    steps = np.logspace(np.log10(smin), np.log10(smax), ntry) if logspace else np.linspace(smin, smax, ntry)
    pwts = steps.copy() ** 0.5
    for i, ss in enumerate(steps):
        q1, p1 = leapfrog(model, kin, q, p, nsteps, ss, cnt)
        Hs[i] = _H(model, kin, q1, p1, cnt)
    pp = np.exp(H0 - Hs) * pwts
    pp[np.isnan(pp)] = 0
    pp[np.isinf(pp)] = 0
    pp /= pp.sum()
    return steps, pp, np.cumsum(pp)


def adhmc_step(model, q0, rng, Nleap=100, nleap=10, ratios=(1 / np.sqrt(2), np.sqrt(2)), pwts0=(1.0, 1.0), smin=0.01,
               smax=1.0, ntry_eps0=10, nsteps_eps0=1, logeps=True, Tint=None, ntimes=None, invmetric_diag=None):
    """adaptivehmc.AdHMC.hmc_step (src/adaptivehmc.py:227-354) and, with Tint / ntimes given,
    adaptivehmc.AdHMC_tint.hmc_step (:394-519).  Draw order: normal(d); uniform(low=cdf.min());
    uniform(lo, hi); one np.random.choice(nr, p=pf) per later segment; conditional
    uniform(0, 1, size=1).  On the blow-up path (:296-301) the reference's np.NaN is spelled
    np.nan here (np.NaN no longer exists in numpy 2); values are the same."""
    kin, cnt = Kinetics(invmetric_diag), Counters()
    ratios = list(ratios)
    nr = len(ratios)
    tmode = Tint is not None
    with np.errstate(all="ignore"):
        p0 = rng.normal(size=q0.size).reshape(q0.shape) * kin.metricstd
        if tmode:
            tint = Tint / ntimes
            N = ntimes
            nsteps_of = lambda e: max(1, int(tint / e))
        else:
            N = int(Nleap // nleap)
            nsteps_of = lambda e: nleap
        eps, pf_den, pb_num = np.zeros(N), np.zeros(N), np.zeros(N)
        H0 = _H(model, kin, q0, p0, cnt)
        steps, pp, cdf = _adhmc_prob_of(model, kin, q0, p0, None, smin, smax, ntry_eps0, nsteps_eps0, logeps, cnt)
        sx = rng.uniform(low=cdf.min())
        isx = np.where(sx > cdf)[0][-1]
        ss = rng.uniform(steps[isx], steps[isx + 1])
        pf_den[0] = pp[isx + 1]
        eps[0] = ss
        q1, p1 = leapfrog(model, kin, q0, p0, nsteps_of(ss), ss, cnt)
        H1 = _H(model, kin, q1, p1, cnt)
        Hprev = H0
        for i in range(N - 1):
            pf, pb = np.zeros(nr), np.zeros(nr)
            qs, ps, Hs = [], [], []
            for j in range(nr):
                ss2 = ss * ratios[j]
                q, p = leapfrog(model, kin, q1, p1, nsteps_of(ss2), ss2, cnt)
                qs.append(q)
                ps.append(p)
                Hs.append(_H(model, kin, q, p, cnt))
                pH = np.exp(H1 - Hs[-1])
                if _bad(pH):
                    pH = 0
                pf[j] = pH
            pwts = np.ones(nr) * pwts0
            if smin > ss * ratios[0]:
                pwts[0] = 0
            if smax < ss * ratios[-1]:
                pwts[-1] = 0
            pf *= pwts
            pf /= pf.sum()
            if _bad(pf.sum()):
                return (q0, p0, 100 + i, [np.nan, np.nan, np.nan], np.stack([pf_den, pb_num, eps]), cnt.as_list())
            pind = rng.choice(nr, p=pf)
            ssnew = ss * ratios[pind]
            q2, p2, H2 = qs[pind], ps[pind], Hs[pind]
            pf_den[i + 1] = pf[pind]
            Hsb = []
            for j in range(nr):
                if np.allclose(ssnew * ratios[j], ss):
                    Hsb.append(Hprev)
                    pbind = j
                else:
                    ss2 = ssnew * ratios[j]
                    q, p = leapfrog(model, kin, q1, -p1, nsteps_of(ss2), ss2, cnt)
                    Hsb.append(_H(model, kin, q, p, cnt))
                pH = np.exp(H1 - Hsb[-1])
                if _bad(pH):
                    pH = 0
                pb[j] = pH
            pwts = np.ones(nr) * pwts0
            if smin > ssnew * ratios[0]:
                pwts[0] = 0
            if smax < ssnew * ratios[-1]:
                pwts[-1] = 0
            pb *= pwts
            pb /= pb.sum()
            pb_num[i] = pb[pbind]
            eps[i + 1] = ssnew
            ss = ssnew
            Hprev = H1
            q1, p1, H1 = q2, p2, H2
        if (ssnew > smin) and (ssnew < smax):
            steps_b, pp_b, _ = _adhmc_prob_of(model, kin, q2, -p2, ssnew, smin, smax, ntry_eps0, nsteps_eps0, logeps, cnt)
            pb_num[i + 1] = pp_b[np.where(steps_b > ssnew)[0][0]]
        hastings = np.prod(pb_num[:i + 2]) / np.prod(pf_den[:i + 2])
        prob = np.exp(H0 - H2) * hastings
        info = [[prob, prob / hastings, hastings], np.stack([pf_den, pb_num, eps]), cnt.as_list()]
        if _bad(prob) or (q0 - q1).sum() == 0:
            return (q0, p0, 2.0, *info)
        if rng.uniform(0.0, 1.0, size=1) > min(1.0, prob):
            return (q0, p0, 0.0, *info)
        return (q2, p2, 1.0, *info)


# ----------------------------------------------------------------------------- diagnostics

def get_ess_true(x, mean, std):
    """utils.get_ess_true, src/utils.py:8-27; x is [Nsamples, Nchains, Nparams]."""
    err = x.mean(axis=0) - mean
    rms = (err ** 2).mean(axis=0) ** 0.5
    ess = (std / rms) ** 2
    with np.errstate(all="ignore"):
        esschain = (std / err) ** 2
    return ess * x.shape[1], esschain * x.shape[1]


def get_rcc(x, c=5):
    """utils.get_rcc, src/utils.py:59-71: FFT autocorrelation along axis 0,
    tau = 1 + 2 cumsum(rho), first lag where idx >= c*tau (Sokal window)."""
    from scipy import signal

    tmp = x - x.mean(axis=0)
    ac = signal.fftconvolve(tmp, tmp[::-1], mode="full", axes=0)
    rcc = ac[ac.shape[0] // 2:]
    rcc = rcc / rcc[0]
    idx = np.arange(rcc.shape[0])
    taus = 2 * np.cumsum(rcc, axis=0) + 1
    window = np.argmin(idx.reshape(-1, 1, 1) < c * taus, axis=0)
    tc = np.take_along_axis(taus, window[None], axis=0)[0].astype(int)
    return rcc, tc


def autocorr_time_window(x, maxlag=500):
    """The numeric part of diagnostics.plot_autorcc, src/diagnostics.py:69-80:
    normalised autocorrelation per chain (np.correlate) and the integrated time
    1 + 2 sum_{l<m} rho_l at the first m (< maxlag) with m > 5*(1 + 2 sum_{l<m} rho_l).
    x is [Nsamples, Nchains] (one parameter)."""
    xp = (x - x.mean(axis=0)) / x.std(axis=0)
    rcc = np.array([np.correlate(xp[:, i], xp[:, i], mode="full") for i in range(xp.shape[1])])
    rcc = rcc[:, rcc.shape[1] // 2:].T
    rcc = rcc / rcc[0]
    tcc = []
    for j in range(x.shape[1]):
        m = 0
        for m in range(maxlag):
            if m > 5 * (1 + 2 * rcc[:m, j].sum()):
                break
        tcc.append(1 + 2 * rcc[:m, j].sum())
    return rcc, np.array(tcc)


def split_rhat(x):
    """Split-R-hat (Gelman et al. 2013) per parameter; x is [Nsamples, Nchains, Nparams].
    NOT in the reference's src/ (only arviz.rhat is called, scripts_old/saveranks/
    saveranksadapt.py:200, and arviz is absent) -> parity unpinned for this function."""
    T = x.shape[0] // 2
    halves = np.concatenate([x[:T], x[T:2 * T]], axis=1)
    m = halves.mean(axis=0)
    s2 = halves.var(axis=0, ddof=1)
    W = s2.mean(axis=0)
    B = T * m.var(axis=0, ddof=1)
    return np.sqrt(((T - 1) / T * W + B / T) / W)


# ----------------------------------------------------------------------------- PyHMC_multistep

def _ms_ghost(model, kin, m, q, p, N, ss, fsub, cnt):
    """PyHMC_multistep.get_num, src/pyhmc.py:538-561: as DRHMC.get_num but it evaluates the
    energy at its base itself (one more Hcount per call)."""
    avec = np.zeros(m)
    H0 = _H(model, kin, q, p, cnt)
    for j in range(m):
        fac = fsub ** j
        qj, pj = leapfrog(model, kin, q, p, int(N * fac), ss / fac, cnt)
        Hj = _H(model, kin, qj, pj, cnt)
        pfac = np.exp(H0 - Hj)
        if (q - qj).sum() == 0:
            pfac = 0.0
        if j:
            den = np.prod(1 - avec[:j])
            num = _ms_ghost(model, kin, j, qj, -pj, N, ss, fsub, cnt)
            prob = pfac * num / den
        else:
            prob = pfac
        if _bad(prob):
            return 0.0
        avec[j] = min(1.0, prob)
        if not np.prod(1 - avec):
            return np.prod(1 - avec)
    return np.prod(1 - avec)


def multistep_step(model, m, q0, N, ss, fsub, rng, invmetric_diag=None):
    """pyhmc.PyHMC_multistep.multi_step, src/pyhmc.py:564-598: first stage accepts iff
    u <= alpha_0; later stages iff u < alpha_j; code = accepting stage j, or -1."""
    kin, cnt = Kinetics(invmetric_diag), Counters()
    with np.errstate(all="ignore"):
        p0 = rng.normal(size=q0.size).reshape(q0.shape) * kin.metricstd
        avec = np.zeros(m)
        H0 = _H(model, kin, q0, p0, cnt)
        q1, p1 = leapfrog(model, kin, q0, p0, N, ss, cnt)
        H1 = _H(model, kin, q1, p1, cnt)
        prob = np.exp(H0 - H1)
        if _bad(prob) or (q0 - q1).sum() == 0:
            prob = 0.0
        avec[0] = min(1.0, prob)
        if rng.uniform() <= avec[0]:
            return q1, p1, 0, avec, cnt.as_list()
        for j in range(1, m):
            fac = fsub ** j
            qj, pj = leapfrog(model, kin, q0, p0, int(N * fac), ss / fac, cnt)
            Hj = _H(model, kin, qj, pj, cnt)
            pfac = np.exp(H0 - Hj)
            if (q0 - qj).sum() == 0:
                pfac = 0.0
            den = np.prod(1 - avec[:j])
            num = _ms_ghost(model, kin, j, qj, -pj, N, ss, fsub, cnt)
            prob = pfac * num / den
            if _bad(prob):
                prob = 0.0
            avec[j] = min(1.0, prob)
            if rng.uniform() < avec[j]:
                return qj, pj, j, avec, cnt.as_list()
    return q0, p0, -1, avec, cnt.as_list()


# ----------------------------------------------------------------------------- probabilistic delayed rejection
# Two rules for the probability of attempting the next stage after a rejection:
#   "metrop"   p_try_j = 1 - min(1, e^{H0-Hj})   (algorithms.Prob_DRHMC_Metrop, pyhmc.PyHMC_multistep_tries)
#   "hastings" p_try_j = 1 - alpha_j             (algorithms.Prob_DRHMC_Hastings, pyhmc.PyHMC_multistep_tries3)
# They differ from each other only in where ptries[j] is set, which the two functions keep.

def _prob_ghost(model, kin, rule, m, q, p, N, ss, fsub, cnt):
    """get_num of the probabilistic classes, src/algorithms.py:659-684 (metrop) and :744-771
    (hastings); src/pyhmc.py:280-305 and :409-436 are the same bodies."""
    avec = np.zeros(m)
    ptries = np.ones(m)
    H0 = _H(model, kin, q, p, cnt)
    for j in range(m):
        fac = fsub ** j
        qj, pj = leapfrog(model, kin, q, p, int(N * fac), ss / fac, cnt)
        Hj = _H(model, kin, qj, pj, cnt)
        pfac = np.exp(H0 - Hj)
        if rule == "metrop":
            ptries[j] = 1 if _bad(pfac) else 1 - min(1.0, pfac)      # before the no-move test
        if (q - qj).sum() == 0:
            pfac = 0.0
        if j:
            den = np.prod(1 - avec[:j]) * np.prod(ptries[:j])
            num = _prob_ghost(model, kin, rule, j, qj, -pj, N, ss, fsub, cnt)
            prob = pfac * num / den
        else:
            prob = pfac
        if _bad(prob):
            return 0.0
        avec[j] = min(1.0, prob)
        if rule == "hastings":
            ptries[j] = 1 if _bad(pfac) else 1 - avec[j]
        if not np.prod(1 - avec):
            return np.prod(1 - avec)
    return np.prod(1 - avec) * np.prod(ptries)


def prob_multistep_step(model, rule, m, q0, N, ss, fsub, rng, invmetric_diag=None):
    """multi_step of the probabilistic classes, src/algorithms.py:687-728 (metrop) and :774-818
    (hastings): before stage j >= 1 one more uniform decides whether the stage is attempted at
    all (code -2 when it is not); codes otherwise as PyHMC_multistep.  Returns the reference's
    six values (q, p, code, avec, counts, ptries)."""
    kin, cnt = Kinetics(invmetric_diag), Counters()
    with np.errstate(all="ignore"):
        p0 = rng.normal(size=q0.size).reshape(q0.shape) * kin.metricstd
        avec = np.zeros(m)
        ptries = np.ones(m)
        H0 = _H(model, kin, q0, p0, cnt)
        q1, p1 = leapfrog(model, kin, q0, p0, N, ss, cnt)
        H1 = _H(model, kin, q1, p1, cnt)
        pfac = np.exp(H0 - H1)
        prob = pfac
        if _bad(prob) or (q0 - q1).sum() == 0:
            prob = 0.0
        avec[0] = min(1.0, prob)
        if _bad(pfac):
            ptries[0] = 1
        else:
            ptries[0] = 1 - min(1.0, pfac) if rule == "metrop" else 1 - avec[0]
        if rng.uniform() <= avec[0]:
            return q1, p1, 0, avec, cnt.as_list(), ptries
        for j in range(1, m):
            if rng.uniform() > ptries[j - 1]:
                return q0, p0, -2, avec, cnt.as_list(), ptries
            fac = fsub ** j
            qj, pj = leapfrog(model, kin, q0, p0, int(N * fac), ss / fac, cnt)
            Hj = _H(model, kin, qj, pj, cnt)
            pfac = np.exp(H0 - Hj)
            if (q0 - qj).sum() == 0:
                pfac = 0.0
            den = np.prod(1 - avec[:j]) * np.prod(ptries[:j])
            num = _prob_ghost(model, kin, rule, j, qj, -pj, N, ss, fsub, cnt)
            if rule == "metrop":
                ptries[j] = 1 if _bad(pfac) else 1 - min(1.0, pfac)  # here after the no-move test
            prob = pfac * num / den
            if _bad(prob):
                prob = 0.0
            avec[j] = min(1.0, prob)
            if rule == "hastings":
                ptries[j] = 1 if _bad(pfac) else 1 - avec[j]
            if rng.uniform() < avec[j]:
                return qj, pj, j, avec, cnt.as_list(), ptries
    return q0, p0, -1, avec, cnt.as_list(), ptries


# ----------------------------------------------------------------------------- generalised HMC family
# These classes are built as Cls(log_prob, grad_log_prob, log_prob_and_grad): V_g then uses
# grad_log_prob (src/algorithms.py:100-106) and the last gradient of a trajectory comes from
# log_prob_and_grad through V_vandg (:109-115,163-164), which also yields the cached V_q1.

def _leapfrog_Vgq(model, kin, q, p, N, step_size, cnt, V_q=None, V_gq=None):
    """algorithms.HMC.leapfrog_Vgq, src/algorithms.py:150-172: the first half kick reuses a cached
    gradient; the last gradient evaluation also returns the potential."""
    cnt.leap += 1
    if V_gq is None:
        p = p - 0.5 * step_size * _V_g(model, q, cnt)
    else:
        p = p - 0.5 * step_size * V_gq
    for _ in range(N - 1):
        q = q + step_size * kin.KE_g(p)
        p = p - step_size * _V_g(model, q, cnt)
    q = q + step_size * kin.KE_g(p)
    cnt.Vg += 1
    v, v_g = model.log_prob_and_grad(q)
    V_q1, V_gq1 = v * -1.0, v_g * -1
    p = p - 0.5 * step_size * V_gq1
    return q, p, V_q1, V_gq1


def _metropolis_alg(model, kin, q0, p0, q1, p1, cnt, V_q0, V_q1, u):
    """algorithms.HMC.metropolis with cached potentials and a given uniform, src/algorithms.py:175-189."""
    if V_q0 is None:
        cnt.H += 1
        V_q0 = _V(model, q0)
    H0 = V_q0 + kin.KE(p0)
    if V_q1 is None:
        cnt.H += 1
        V_q1 = _V(model, q1)
    H1 = V_q1 + kin.KE(p1)
    prob = np.exp(H0 - H1)
    if _bad(prob) or (q0 - q1).sum() == 0:
        return q0, p0, -1, [H0, H1]
    if u > min(1.0, prob):
        return q0, p0, 0.0, [H0, H1]
    return q1, p1, 1.0, [H0, H1]


def ghmc_step(model, q, p, nleap, step_size, alpha, rng, V_q=None, V_gq=None, invmetric_diag=None):
    """algorithms.GHMC.step, src/algorithms.py:265-281: partial momentum refresh
    p <- alpha p + sqrt(1-alpha^2) pn, cached V / grad V, momentum flipped unless accepted.
    Returns (q, p, V_q, V_gq, accepted, Hs, counts)."""
    kin, cnt = Kinetics(invmetric_diag), Counters()
    with np.errstate(all="ignore"):
        pn = rng.normal(size=q.size).reshape(q.shape) * kin.metricstd
        p = alpha * p + np.sqrt(1 - alpha ** 2) * pn
        q1, p1, V_q1, V_gq1 = _leapfrog_Vgq(model, kin, q, p, nleap, step_size, cnt, V_q, V_gq)
        u = rng.uniform(0.0, 1.0, size=1)
        q, p, accepted, Hs = _metropolis_alg(model, kin, q, p, q1, p1, cnt, V_q, V_q1, u)
        if accepted == 1:
            V_q, V_gq = V_q1, V_gq1
        else:
            p = -p
    return q, p, V_q, V_gq, accepted, Hs, cnt.as_list()


def pghmc_step(model, q, p, u, nleap, step_size, alpha, delta, rng, V_q=None, V_gq=None, invmetric_diag=None):
    """algorithms.Persistent_GHMC.step, src/algorithms.py:314-333: GHMC whose Metropolis uniform
    is itself a persistent variable, u <- (u + 1 + delta) % 2 - 1, rescaled by 1/exp(H0-H1) on
    acceptance.  Returns (q, p, u, V_q, V_gq, accepted, Hs, counts)."""
    kin, cnt = Kinetics(invmetric_diag), Counters()
    with np.errstate(all="ignore"):
        pn = rng.normal(size=q.size).reshape(q.shape) * kin.metricstd
        p = alpha * p + np.sqrt(1 - alpha ** 2) * pn
        u = (u + 1 + delta) % 2 - 1
        q1, p1, V_q1, V_gq1 = _leapfrog_Vgq(model, kin, q, p, nleap, step_size, cnt, V_q, V_gq)
        q, p, accepted, Hs = _metropolis_alg(model, kin, q, p, q1, p1, cnt, V_q, V_q1, u)
        if accepted == 1:
            V_q, V_gq = V_q1, V_gq1
            u = u / np.exp(Hs[0] - Hs[1])
        else:
            p = -p
    return q, p, u, V_q, V_gq, accepted, Hs, cnt.as_list()


def ghmc_sample(model, q, nsamples, burnin, step_size, nleap, alpha, rng, p=None, invmetric_diag=None):
    """algorithms.GHMC.sample, src/algorithms.py:284-303: p starts as one normal vector when not given (:290), V and grad V
    start as None, every iteration (burn-in included) is stored."""
    if p is None:
        p = rng.normal(0, 1, q.size).reshape(q.shape)
    V_q, V_gq = None, None
    samples, accepts, Hs, counts = [], [], [], []
    for i in range(nsamples + burnin):
        q, p, V_q, V_gq, acc, H, count = ghmc_step(model, q, p, nleap, step_size, alpha, rng, V_q, V_gq, invmetric_diag)
        samples.append(q)
        accepts.append(acc)
        Hs.append(H)
        counts.append(count)
    return dict(samples=np.array(samples), accepts=np.array(accepts), Hs=np.array(Hs), counts=np.array(counts),
                i=nsamples + burnin, p=p, V_q=V_q, V_gq=V_gq)


def pghmc_sample(model, q, nsamples, burnin, step_size, nleap, alpha, delta, rng, p=None, invmetric_diag=None):
    """algorithms.Persistent_GHMC.sample, src/algorithms.py:336-357: as GHMC.sample, plus the persistent variable u, which
    starts as ONE scalar normal draw (:343, after the momentum draw) and is stored every iteration."""
    if p is None:
        p = rng.normal(0, 1, q.size).reshape(q.shape)
    u = rng.normal(0, 1)
    V_q, V_gq = None, None
    samples, accepts, Hs, counts, us = [], [], [], [], []
    for i in range(nsamples + burnin):
        q, p, u, V_q, V_gq, acc, H, count = pghmc_step(model, q, p, u, nleap, step_size, alpha, delta, rng, V_q, V_gq,
                                                       invmetric_diag)
        samples.append(q)
        accepts.append(acc)
        Hs.append(H)
        counts.append(count)
        us.append(u)
    return dict(samples=np.array(samples), accepts=np.array(accepts), Hs=np.array(Hs), counts=np.array(counts),
                u=np.array(us), i=nsamples + burnin, p=p)


def drghmc_step(model, q, p, nleap, step_size, alpha, nprops, stepfactor, rng, invmetric_diag=None):
    """algorithms.DRGHMC.step, src/algorithms.py:596-625: DRHMC.step started from the partially
    refreshed momentum p0 = alpha p + sqrt(1-alpha^2) pn; p0 is returned (not flipped) on rejection."""
    kin, cnt = Kinetics(invmetric_diag), Counters()
    avec = np.zeros(nprops)
    with np.errstate(all="ignore"):
        q0 = q
        pn = rng.normal(size=q.size).reshape(q.shape) * kin.metricstd
        p0 = alpha * p + np.sqrt(1 - alpha ** 2) * pn
        H0 = _H(model, kin, q0, p0, cnt)
        for j in range(nprops):
            fac = stepfactor ** j
            qj, pj = leapfrog(model, kin, q0, p0, int(nleap * fac), step_size / fac, cnt)
            Hj = _H(model, kin, qj, pj, cnt)
            pfac = np.exp(H0 - Hj)
            if (q0 - qj).sum() == 0:
                pfac = 0.0
            den = np.prod(1 - avec[:j])
            num = _dr_ghost(model, kin, j, qj, -pj, nleap, step_size, stepfactor, Hj, cnt)
            prob = pfac * num / den
            if _bad(prob):
                prob = 0.0
            avec[j] = min(1.0, prob)
            if rng.uniform() < avec[j]:
                return qj, pj, j + 1, avec, cnt.as_list()
    return q0, p0, 0, avec, cnt.as_list()
