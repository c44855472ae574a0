/*
 * oracle/pg.c -- TEST INFRASTRUCTURE (see oracle/__init__.py).
 *
 * Plain-C restatement of the Polya-gamma PG(1, z) sampler the reference calls through
 * pypolyagamma (eigenmodel.py:8,826-828,868-871: ppg.pgdrawvpar(ppgs, ones, Mu.ravel(), out)).
 * pypolyagamma is NOT in /root/reference (un-vendored, version unpinned); this follows the
 * published algorithm: Polson, Scott & Windle (2013), Alg. 1 -- Devroye's exact sampler for
 * J*(1, z/2), truncation t = 0.64, as implemented in BayesLogit's PolyaGamma::draw_like_devroye.
 * Sample paths cannot match pypolyagamma's private RNG => parity for omega is distributional
 * ("parity unpinned"); everything downstream is tested with injected Omega.
 *
 * RNG: splitmix64-seeded xoshiro256** per call-range; Box-Muller normals.
 * Build:  gcc -O2 -fPIC -shared -fopenmp -o oracle/_build/libpgoracle.so oracle/pg.c -lm
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define PG_TRUNC 0.64
#define PG_PI 3.141592653589793238462643383279502884

typedef struct { uint64_t s[4]; } rng_t;

static uint64_t splitmix64(uint64_t *x) {
    uint64_t z = (*x += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
static void rng_seed(rng_t *r, uint64_t seed) {
    for (int i = 0; i < 4; ++i) r->s[i] = splitmix64(&seed);
}
static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
static uint64_t rng_next(rng_t *r) {
    uint64_t *s = r->s;
    uint64_t result = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
    return result;
}
static double rng_unif(rng_t *r) { return ((rng_next(r) >> 11) + 0.5) * (1.0 / 9007199254740992.0); }
static double rng_expon(rng_t *r) { return -log(rng_unif(r)); }
static double rng_norm(rng_t *r) {
    double u1 = rng_unif(r), u2 = rng_unif(r);
    return sqrt(-2.0 * log(u1)) * cos(2.0 * PG_PI * u2);
}

static double norm_cdf(double x) { return 0.5 * erfc(-x / sqrt(2.0)); }

/* a_n(x): coefficients of the alternating series for the J*(1,0) density */
static double a_coef(int n, double x) {
    double K = (n + 0.5) * PG_PI;
    if (x > PG_TRUNC) return K * exp(-0.5 * K * K * x);
    double expnt = -1.5 * (log(0.5 * PG_PI) + log(x)) + log(K) - 2.0 * (n + 0.5) * (n + 0.5) / x;
    return exp(expnt);
}

/* probability mass of the truncated-exponential (right) piece of the proposal */
static double mass_texpon(double Z) {
    double t = PG_TRUNC;
    double fz = 0.125 * PG_PI * PG_PI + 0.5 * Z * Z;
    double b = sqrt(1.0 / t) * (t * Z - 1);
    double a = -sqrt(1.0 / t) * (t * Z + 1);
    double x0 = log(fz) + fz * t;
    double xb = x0 - Z + log(norm_cdf(b));
    double xa = x0 + Z + log(norm_cdf(a));
    double qdivp = 4.0 / PG_PI * (exp(xb) + exp(xa));
    return 1.0 / (1.0 + qdivp);
}

/* inverse-Gaussian(mu = 1/Z, lambda = 1) truncated to (0, t) */
static double rtigauss(double Z, rng_t *r) {
    double t = PG_TRUNC, X = t + 1.0;
    Z = fabs(Z);
    if (1.0 / t > Z) {
        double alpha = 0.0;
        while (rng_unif(r) > alpha) {
            double E1 = rng_expon(r), E2 = rng_expon(r);
            while (E1 * E1 > 2.0 * E2 / t) { E1 = rng_expon(r); E2 = rng_expon(r); }
            X = 1.0 + E1 * t;
            X = t / (X * X);
            alpha = exp(-0.5 * Z * Z * X);
        }
    } else {
        double mu = 1.0 / Z;
        while (X > t) {
            double Y = rng_norm(r); Y *= Y;
            double half_mu = 0.5 * mu, mu_Y = mu * Y;
            X = mu + half_mu * mu_Y - half_mu * sqrt(4.0 * mu_Y + mu_Y * mu_Y);
            if (rng_unif(r) > mu / (mu + X)) X = mu * mu / X;
        }
    }
    return X;
}

static double pg1_draw(double z, rng_t *r) {
    double Z = fabs(z) * 0.5;
    double fz = 0.125 * PG_PI * PG_PI + 0.5 * Z * Z;
    double mass = mass_texpon(Z);
    for (;;) {
        double X;
        if (rng_unif(r) < mass) X = PG_TRUNC + rng_expon(r) / fz;
        else X = rtigauss(Z, r);
        double S = a_coef(0, X);
        double Y = rng_unif(r) * S;
        int n = 0;
        for (;;) {
            ++n;
            if (n & 1) { S -= a_coef(n, X); if (Y <= S) return 0.25 * X; }
            else       { S += a_coef(n, X); if (Y > S) break; }
        }
    }
}

/* out[i] ~ PG(1, z[i]); OpenMP over contiguous chunks, one RNG stream per chunk (seed, chunk) */
void pg1_draw_vec(const double *z, double *out, size_t n, uint64_t seed) {
    const size_t chunk = 4096;
    size_t nchunks = (n + chunk - 1) / chunk;
#pragma omp parallel for schedule(static)
    for (long c = 0; c < (long)nchunks; ++c) {
        rng_t r; rng_seed(&r, seed * 0x9E3779B97F4A7C15ULL + (uint64_t)c + 1);
        size_t lo = (size_t)c * chunk, hi = lo + chunk < n ? lo + chunk : n;
        for (size_t i = lo; i < hi; ++i) out[i] = pg1_draw(z[i], &r);
    }
}

/* fused variant used by the timed CPU baseline: z[n,m] = mu0 + sum_d F[n,d] lam[d] F[m,d] (diag = mu0),
 * eigenmodel.py:75-83 + 859-874 in one pass, rows [r0, r1) */
void pg1_draw_eigen(const double *F, const double *lam, double mu0, int N, int D,
                    int r0, int r1, double *out, uint64_t seed) {
#pragma omp parallel for schedule(static)
    for (int n = r0; n < r1; ++n) {
        rng_t r; rng_seed(&r, seed * 0x9E3779B97F4A7C15ULL + (uint64_t)n + 1);
        for (int m = 0; m < N; ++m) {
            double z = mu0;
            if (m != n) for (int d = 0; d < D; ++d) z += F[(size_t)n * D + d] * lam[d] * F[(size_t)m * D + d];
            out[(size_t)(n - r0) * N + m] = pg1_draw(z, &r);
        }
    }
}

int pg_oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
