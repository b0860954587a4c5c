// Expression tape: the probability expressions of a model (reference src/mef/openpsa/expression.h, expr/*.h) compiled
// into a flat, topologically ordered array of nodes, so that Expression::value() (the mean) and Expression::Sample() /
// Reset() (expression.h:98-117: one draw per deviate per trial, shared by everything that refers to it) are one forward
// pass each, with no XML and no object graph.  This is what the parameter-uncertainty outer loop
// evaluates once per epistemic trial to obtain BasicEvent::p() of every event.
//
// The reference declares the six random deviates (expr/random_deviate.h:25-264) but ships no DoSample body and draws from
// one process-global std::mt19937 "not suitable for parallelized simulations" (random_deviate.h:20-22,43).  Here every
// draw is counter based: uniform j of deviate d in trial t is a pure function of (seed, t, d, j), so trials can be
// evaluated in any order, on any rank, and restarted:
//     x = Philox4x32-10(ctr = (t_lo, t_hi, d, 0x500 + (j >> 1)), key = (seed_lo, seed_hi));
//     u_j = ((x[2(j&1)+1] * 2^32 + x[2(j&1)]) >> 11) + 0.5) * 2^-53                     in (0, 1)
// Counter word 3 = 0x500 + ... is disjoint from every stream of the sample-level specs (0, 1, 0x100.., 0x200.., 0x300..,
// 0x400..), so the epistemic draws never collide with the aleatory ones of the same seed.
//   uniform(a, b)        a + (b - a) u_0
//   normal(m, s)         m + s z,  z = sqrt(-2 ln u_0) cos(2 pi u_1)                          (Box-Muller)
//   lognormal(mu, s)     exp(mu + s z);  lognormal(mean, EF, level): s = ln EF / Phi^-1(level), mu = ln mean - s^2 / 2
//   gamma(k, theta)      Marsaglia-Tsang (2000) on k' = k or k + 1, uniforms taken in order; for k < 1 times u^(1/k)
//   beta(a, b)           X / (X + Y), X ~ gamma(a, 1), Y ~ gamma(b, 1) from the same uniform sequence
//   histogram            u_0 picks the bin by cumulative weight, u_1 the point inside it
// tests/deviate_restatement.py restates this text in Python; tests compare the two and check closed-form moments.
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>

namespace canopy {

enum ExprOp : uint16_t {
    X_CONST, X_NEG, X_ADD, X_SUB, X_MUL, X_DIV, X_ABS, X_ACOS, X_ASIN, X_ATAN, X_COS, X_SIN, X_TAN, X_COSH, X_SINH, X_TANH,
    X_EXP, X_LOG, X_LOG10, X_SQRT, X_CEIL, X_FLOOR, X_MOD, X_POW, X_MIN, X_MAX, X_MEAN, X_NOT, X_AND, X_OR,
    X_EQ, X_DF, X_LT, X_GT, X_LEQ, X_GEQ, X_ITE, X_EXPONENTIAL, X_GLM, X_WEIBULL, X_PERIODIC4, X_PERIODIC5,
    X_UNIFORM, X_NORMAL, X_LOGNORMAL3, X_LOGNORMAL2, X_GAMMA, X_BETA, X_HISTOGRAM
};
inline bool is_deviate_op(uint16_t op) { return op >= X_UNIFORM; }

// ---- Philox4x32-10 on the host (same constants as kernels.cuh; Salmon et al. 2011) --------------------------------
inline void philox4x32_10_host(uint32_t c[4], uint32_t k0, uint32_t k1) {
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0, n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        c[1] = (uint32_t)p1; c[3] = (uint32_t)p0; c[0] = n0; c[2] = n2;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}

// The uniform sequence u_0, u_1, ... of one deviate in one trial.
struct UniformStream {
    uint64_t seed, trial;
    uint32_t deviate, next = 0;
    uint32_t x[4];
    UniformStream(uint64_t s, uint64_t t, uint32_t d) : seed(s), trial(t), deviate(d) {}
    double operator()() {
        const uint32_t j = next++;
        if ((j & 1u) == 0u) {
            x[0] = (uint32_t)trial; x[1] = (uint32_t)(trial >> 32); x[2] = deviate; x[3] = 0x500u + (j >> 1);
            philox4x32_10_host(x, (uint32_t)seed, (uint32_t)(seed >> 32));
        }
        const uint64_t w = ((uint64_t)x[2 * (j & 1u) + 1] << 32) | x[2 * (j & 1u)];
        return ((double)(w >> 11) + 0.5) * 0x1.0p-53;
    }
};

// Phi^-1 (Acklam's rational approximation, then one Halley step on erfc: full double precision).
inline double inv_normal_cdf(double p) {
    static const double a[] = {-3.969683028665376e+01, 2.209460984245205e+02, -2.759285104469687e+02, 1.383577518672690e+02, -3.066479806614716e+01, 2.506628277459239e+00};
    static const double b[] = {-5.447609879822406e+01, 1.615858368580409e+02, -1.556989798598866e+02, 6.680131188771972e+01, -1.328068155288572e+01};
    static const double c[] = {-7.784894002430293e-03, -3.223964580411365e-01, -2.400758277161838e+00, -2.549732539343734e+00, 4.374664141464968e+00, 2.938163982698783e+00};
    static const double d[] = {7.784695709041462e-03, 3.224671290700398e-01, 2.445134137142996e+00, 3.754408661907416e+00};
    double x;
    if (p < 0.02425) { const double q = std::sqrt(-2 * std::log(p)); x = (((((c[0] * q + c[1]) * q + c[2]) * q + c[3]) * q + c[4]) * q + c[5]) / ((((d[0] * q + d[1]) * q + d[2]) * q + d[3]) * q + 1); }
    else if (p > 1 - 0.02425) { const double q = std::sqrt(-2 * std::log(1 - p)); x = -(((((c[0] * q + c[1]) * q + c[2]) * q + c[3]) * q + c[4]) * q + c[5]) / ((((d[0] * q + d[1]) * q + d[2]) * q + d[3]) * q + 1); }
    else { const double q = p - 0.5, r = q * q; x = (((((a[0] * r + a[1]) * r + a[2]) * r + a[3]) * r + a[4]) * r + a[5]) * q / (((((b[0] * r + b[1]) * r + b[2]) * r + b[3]) * r + b[4]) * r + 1); }
    const double e = 0.5 * std::erfc(-x / 1.4142135623730951) - p;
    const double u = e * 2.5066282746310002 * std::exp(x * x / 2);
    return x - u / (1 + x * u / 2);
}

inline double draw_normal(UniformStream& U) {
    const double u0 = U(), u1 = U();
    return std::sqrt(-2.0 * std::log(u0)) * std::cos(6.283185307179586 * u1);
}
inline double draw_gamma(UniformStream& U, double k) {  // scale 1
    const double kk = k < 1.0 ? k + 1.0 : k, d = kk - 1.0 / 3.0, c = 1.0 / std::sqrt(9.0 * d);
    double g;
    for (;;) {
        const double z = draw_normal(U), t = 1.0 + c * z;
        if (t <= 0.0) continue;
        const double v = t * t * t, u = U();
        if (std::log(u) < 0.5 * z * z + d - d * v + d * std::log(v)) { g = d * v; break; }
    }
    if (k < 1.0) g *= std::pow(U(), 1.0 / k);
    return g;
}

// periodic-test (expr/exponential.h:105-245; the reference declares Compute without a body).  Four arguments: tests and
// repairs instantaneous and always successful, so the component is as good as new after every test at theta, theta + tau,
// ...: unavailability = 1 - exp(-lambda x time since the last test).  Five arguments: tests instantaneous and always
// successful, repair at rate mu starts when a test finds the failure.  States W (working), F (failed, not detected),
// R (in repair); between tests W' = -lambda W + mu R, F' = lambda W, R' = -mu R; at a test R += F, F = 0; value = F + R.
inline double periodic_test4(double lambda, double tau, double theta, double time) {
    if (time <= theta) return 1.0 - std::exp(-lambda * time);
    const double delta = time - theta;
    return 1.0 - std::exp(-lambda * (delta - std::floor(delta / tau) * tau));
}
inline double periodic_test5(double lambda, double mu, double tau, double theta, double time) {
    double W = 1.0, R = 0.0;  // F = 1 - W - R
    auto advance = [&](double t) {
        const double el = std::exp(-lambda * t), em = std::exp(-mu * t);
        const double flow = lambda == mu ? mu * t * el : mu / (mu - lambda) * (el - em);  // R -> W that has not failed again
        W = W * el + R * flow; R = R * em;
    };
    if (time <= theta) { advance(time); return 1.0 - W; }
    advance(theta); R = 1.0 - W;  // first test: every failure found
    double left = time - theta;
    const double n = std::floor(left / tau);
    for (double i = 0; i < n; i += 1.0) { advance(tau); R = 1.0 - W; }
    advance(left - n * tau);
    return 1.0 - W;
}

struct ExprNode { uint16_t op; uint32_t n_args, arg_off; double c; uint32_t deviate; };

struct ExprTape {
    std::vector<ExprNode> nodes;     // arguments always precede the node that uses them
    std::vector<uint32_t> args;
    uint32_t n_deviates = 0;

    uint32_t add(uint16_t op, const std::vector<uint32_t>& a, double c = 0.0) {
        ExprNode n{op, (uint32_t)a.size(), (uint32_t)args.size(), c, 0u};
        if (is_deviate_op(op)) n.deviate = n_deviates++;
        args.insert(args.end(), a.begin(), a.end());
        nodes.push_back(n);
        return (uint32_t)nodes.size() - 1;
    }
    uint32_t constant(double c) { return add(X_CONST, {}, c); }

    // One pass over the tape.  sample == false: Expression::value() (deviates give their mean); sample == true: the
    // `trial`-th joint draw of all deviates under `seed`.  v[i] = value of node i.
    void evaluate(bool sample, uint64_t seed, uint64_t trial, std::vector<double>& v) const {
        v.resize(nodes.size());
        for (size_t i = 0; i < nodes.size(); ++i) {
            const ExprNode& n = nodes[i];
            const uint32_t* a = args.data() + n.arg_off;
            auto A = [&](uint32_t j) { return v[a[j]]; };
            double r = 0.0;
            switch (n.op) {
                case X_CONST: r = n.c; break;
                case X_NEG: r = -A(0); break;
                case X_ADD: r = A(0); for (uint32_t j = 1; j < n.n_args; ++j) r += A(j); break;
                case X_SUB: r = A(0); for (uint32_t j = 1; j < n.n_args; ++j) r -= A(j); break;
                case X_MUL: r = A(0); for (uint32_t j = 1; j < n.n_args; ++j) r *= A(j); break;
                case X_DIV: r = A(0); for (uint32_t j = 1; j < n.n_args; ++j) r /= A(j); break;
                case X_ABS: r = std::fabs(A(0)); break;
                case X_ACOS: r = std::acos(A(0)); break;
                case X_ASIN: r = std::asin(A(0)); break;
                case X_ATAN: r = std::atan(A(0)); break;
                case X_COS: r = std::cos(A(0)); break;
                case X_SIN: r = std::sin(A(0)); break;
                case X_TAN: r = std::tan(A(0)); break;
                case X_COSH: r = std::cosh(A(0)); break;
                case X_SINH: r = std::sinh(A(0)); break;
                case X_TANH: r = std::tanh(A(0)); break;
                case X_EXP: r = std::exp(A(0)); break;
                case X_LOG: r = std::log(A(0)); break;
                case X_LOG10: r = std::log10(A(0)); break;
                case X_SQRT: r = std::sqrt(A(0)); break;
                case X_CEIL: r = std::ceil(A(0)); break;
                case X_FLOOR: r = std::floor(A(0)); break;
                case X_MOD: { const long long d = (long long)A(1); r = d ? (double)((long long)A(0) % d) : 0.0; break; }
                case X_POW: r = std::pow(A(0), A(1)); break;
                case X_MIN: r = A(0); for (uint32_t j = 1; j < n.n_args; ++j) r = std::fmin(r, A(j)); break;
                case X_MAX: r = A(0); for (uint32_t j = 1; j < n.n_args; ++j) r = std::fmax(r, A(j)); break;
                case X_MEAN: r = 0; for (uint32_t j = 0; j < n.n_args; ++j) r += A(j); r /= (double)n.n_args; break;
                case X_NOT: r = A(0) != 0.0 ? 0.0 : 1.0; break;
                case X_AND: r = 1.0; for (uint32_t j = 0; j < n.n_args; ++j) if (A(j) == 0.0) r = 0.0; break;
                case X_OR: r = 0.0; for (uint32_t j = 0; j < n.n_args; ++j) if (A(j) != 0.0) r = 1.0; break;
                case X_EQ: r = A(0) == A(1); break;
                case X_DF: r = A(0) != A(1); break;
                case X_LT: r = A(0) < A(1); break;
                case X_GT: r = A(0) > A(1); break;
                case X_LEQ: r = A(0) <= A(1); break;
                case X_GEQ: r = A(0) >= A(1); break;
                case X_ITE: r = A(0) != 0.0 ? A(1) : A(2); break;
                case X_EXPONENTIAL: r = 1.0 - std::exp(-A(0) * A(1)); break;
                case X_GLM: { const double g = A(0), l = A(1), m = A(2), t = A(3), s = l + m; r = s == 0 ? g : (l - (l - g * s) * std::exp(-s * t)) / s; break; }
                case X_WEIBULL: r = A(3) <= A(2) ? 0.0 : 1.0 - std::exp(-std::pow((A(3) - A(2)) / A(0), A(1))); break;
                case X_PERIODIC4: r = periodic_test4(A(0), A(1), A(2), A(3)); break;
                case X_PERIODIC5: r = periodic_test5(A(0), A(1), A(2), A(3), A(4)); break;
                default: r = sample ? draw(n, a, v, seed, trial) : mean(n, a, v); break;
            }
            v[i] = r;
        }
    }
    // Mean of a deviate (random_deviate.h: value() of each class).
    double mean(const ExprNode& n, const uint32_t* a, const std::vector<double>& v) const {
        switch (n.op) {
            case X_UNIFORM: return (v[a[0]] + v[a[1]]) / 2;
            case X_NORMAL: return v[a[0]];
            case X_LOGNORMAL3: return v[a[0]];
            case X_LOGNORMAL2: return std::exp(v[a[0]] + v[a[1]] * v[a[1]] / 2);
            case X_GAMMA: return v[a[0]] * v[a[1]];
            case X_BETA: return v[a[0]] / (v[a[0]] + v[a[1]]);
            default: {  // histogram: args = lower, then (upper, weight) per bin
                double prev = v[a[0]], sw = 0, swm = 0;
                for (uint32_t j = 1; j + 1 < n.n_args; j += 2) { const double b = v[a[j]], w = v[a[j + 1]]; sw += w; swm += w * (prev + b); prev = b; }
                return swm / (2 * sw);
            }
        }
    }
    double draw(const ExprNode& n, const uint32_t* a, const std::vector<double>& v, uint64_t seed, uint64_t trial) const {
        UniformStream U(seed, trial, n.deviate);
        switch (n.op) {
            case X_UNIFORM: return v[a[0]] + (v[a[1]] - v[a[0]]) * U();
            case X_NORMAL: return v[a[0]] + v[a[1]] * draw_normal(U);
            case X_LOGNORMAL3: {
                const double s = std::log(v[a[1]]) / inv_normal_cdf(v[a[2]]), mu = std::log(v[a[0]]) - s * s / 2;
                return std::exp(mu + s * draw_normal(U));
            }
            case X_LOGNORMAL2: return std::exp(v[a[0]] + v[a[1]] * draw_normal(U));
            case X_GAMMA: return draw_gamma(U, v[a[0]]) * v[a[1]];
            case X_BETA: { const double x = draw_gamma(U, v[a[0]]), y = draw_gamma(U, v[a[1]]); return x / (x + y); }
            default: {
                double sw = 0;
                for (uint32_t j = 2; j < n.n_args; j += 2) sw += v[a[j]];
                const double target = U() * sw, u1 = U();
                double prev = v[a[0]], acc = 0;
                for (uint32_t j = 1; j + 1 < n.n_args; j += 2) {
                    const double b = v[a[j]], w = v[a[j + 1]];
                    acc += w;
                    if (target < acc || j + 3 >= n.n_args) return prev + (b - prev) * u1;
                    prev = b;
                }
                return prev;
            }
        }
    }
};

}  // namespace canopy
