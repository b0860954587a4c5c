"""Independent restatement of the deviate draws of the uncertainty loop, written from the text at the top of
canopy_b200/csrc/expr_tape.h (and include/canopy.h "Parameter uncertainty"), sharing no code with the library.

Test infrastructure only.  The reference declares the six deviates (src/mef/openpsa/expr/random_deviate.h:25-264) but has
no DoSample body, so -- like the sampling specs -- the draws are defined by this repo's written mapping:
uniform j of deviate d in trial t under `seed` is Philox4x32-10(ctr = (t_lo, t_hi, d, 0x500 + j // 2), key = seed) words
(2 (j & 1) + 1, 2 (j & 1)) -> ((w >> 11) + 0.5) / 2^53.
"""
import math

M0, M1, W0, W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
MASK = 0xFFFFFFFF


def philox(ctr, key):
    c0, c1, c2, c3 = ctr
    k0, k1 = key
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        c0, c1, c2, c3 = ((p1 >> 32) ^ c1 ^ k0) & MASK, p1 & MASK, ((p0 >> 32) ^ c3 ^ k1) & MASK, p0 & MASK
        k0, k1 = (k0 + W0) & MASK, (k1 + W1) & MASK
    return c0, c1, c2, c3


class Uniforms:
    def __init__(self, seed, trial, deviate):
        self.seed, self.trial, self.deviate, self.j = seed, trial, deviate, 0

    def __call__(self):
        j = self.j
        self.j += 1
        x = philox((self.trial & MASK, self.trial >> 32, self.deviate, 0x500 + (j >> 1)), (self.seed & MASK, self.seed >> 32))
        w = (x[2 * (j & 1) + 1] << 32) | x[2 * (j & 1)]
        return ((w >> 11) + 0.5) * 2.0 ** -53


def normal01(U):
    u0, u1 = U(), U()
    return math.sqrt(-2.0 * math.log(u0)) * math.cos(6.283185307179586 * u1)


def gamma1(U, k):
    kk = k + 1.0 if k < 1.0 else k
    d = kk - 1.0 / 3.0
    c = 1.0 / math.sqrt(9.0 * d)
    while True:
        z = normal01(U)
        t = 1.0 + c * z
        if t <= 0.0:
            continue
        v = t * t * t
        u = U()
        if math.log(u) < 0.5 * z * z + d - d * v + d * math.log(v):
            g = d * v
            break
    if k < 1.0:
        g *= U() ** (1.0 / k)
    return g


def inv_normal_cdf(p):
    """Phi^-1 by bisection on erfc (slow, independent of the library's rational approximation)."""
    lo, hi = -40.0, 40.0
    for _ in range(200):
        mid = (lo + hi) / 2
        if 0.5 * math.erfc(-mid / math.sqrt(2.0)) < p:
            lo = mid
        else:
            hi = mid
    return (lo + hi) / 2


def draw(kind, params, seed, trial, deviate):
    """kind: 'uniform' (a, b) | 'normal' (m, s) | 'lognormal' (mu, s) | 'lognormal_ef' (mean, ef, level) |
    'gamma' (k, theta) | 'beta' (a, b) | 'histogram' (bounds, weights)."""
    U = Uniforms(seed, trial, deviate)
    if kind == "uniform":
        return params[0] + (params[1] - params[0]) * U()
    if kind == "normal":
        return params[0] + params[1] * normal01(U)
    if kind == "lognormal":
        return math.exp(params[0] + params[1] * normal01(U))
    if kind == "lognormal_ef":
        s = math.log(params[1]) / inv_normal_cdf(params[2])
        mu = math.log(params[0]) - s * s / 2
        return math.exp(mu + s * normal01(U))
    if kind == "gamma":
        return gamma1(U, params[0]) * params[1]
    if kind == "beta":
        x = gamma1(U, params[0])
        y = gamma1(U, params[1])
        return x / (x + y)
    if kind == "histogram":
        bounds, weights = params
        target = U() * sum(weights)
        u1 = U()
        acc = 0.0
        for i, w in enumerate(weights):
            acc += w
            if target < acc or i == len(weights) - 1:
                return bounds[i] + (bounds[i + 1] - bounds[i]) * u1
    raise ValueError(kind)


def periodic_test5_ode(lam, mu, tau, theta, time, steps_per_hour=200):
    """Five-argument periodic-test by direct integration of W' = -lam W + mu R, R' = -mu R with R += F at every test
    (RK4); independent of the closed form in the library."""
    W, R = 1.0, 0.0
    t = 0.0
    next_test = theta

    def f(W, R):
        return -lam * W + mu * R, -mu * R

    while t < time - 1e-12:
        end = min(time, next_test)
        n = max(1, int(math.ceil((end - t) * steps_per_hour)))
        h = (end - t) / n
        for _ in range(n):
            k1 = f(W, R)
            k2 = f(W + h / 2 * k1[0], R + h / 2 * k1[1])
            k3 = f(W + h / 2 * k2[0], R + h / 2 * k2[1])
            k4 = f(W + h * k3[0], R + h * k3[1])
            W += h / 6 * (k1[0] + 2 * k2[0] + 2 * k3[0] + k4[0])
            R += h / 6 * (k1[1] + 2 * k2[1] + 2 * k3[1] + k4[1])
        t = end
        if t >= next_test - 1e-12 and t < time - 1e-12:
            R = 1.0 - W
            next_test += tau
    return 1.0 - W
