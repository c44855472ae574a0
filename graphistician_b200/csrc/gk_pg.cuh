// gk_pg.cuh -- device Polya-gamma PG(1, z) sampler (K3) and the counter-based RNG it draws from.
//
// Replaces the pypolyagamma call of eigenmodel.py:859-874 (ppg.pgdrawvpar(ppgs, ones, Mu.ravel(), out)).
// Algorithm: Polson, Scott & Windle (2013) Alg. 1 -- Devroye's exact rejection sampler for J*(1, z/2) with
// truncation point t = 0.64: proposal = truncated inverse-Gaussian on (0, t] / exponential on (t, inf),
// accepted against the alternating series sum_n (-1)^n a_n(x).  omega = J* / 4.
// Arithmetic is fp32 (the draw is a random variate, not a conditional: parity is distributional).
#pragma once
#include <stdint.h>

#define GK_PG_T 0.64f
#define GK_PI_F 3.14159265358979f

// ---- Philox4x32-10, keyed by (seed_lo, seed_hi); counter = (c0, c1, c2, c3) ---------------------------
struct gk_philox {
    uint32_t key0, key1, c0, c1, c2, c3;
    uint32_t out[4];
    int have;
    __device__ __forceinline__ void init(uint64_t seed, uint32_t a, uint32_t b, uint32_t cc) {
        key0 = (uint32_t)seed; key1 = (uint32_t)(seed >> 32); c0 = 0; c1 = a; c2 = b; c3 = cc; have = 0;
    }
    __device__ __forceinline__ void refill() {
        uint32_t x0 = c0, x1 = c1, x2 = c2, x3 = c3, k0 = key0, k1 = key1;
#pragma unroll
        for (int r = 0; r < 10; ++r) {
            uint32_t hi0 = __umulhi(0xD2511F53u, x0), lo0 = 0xD2511F53u * x0;
            uint32_t hi1 = __umulhi(0xCD9E8D57u, x2), lo1 = 0xCD9E8D57u * x2;
            uint32_t y0 = hi1 ^ x1 ^ k0, y1 = lo1, y2 = hi0 ^ x3 ^ k1, y3 = lo0;
            x0 = y0; x1 = y1; x2 = y2; x3 = y3;
            k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
        }
        out[0] = x0; out[1] = x1; out[2] = x2; out[3] = x3;
        c0++; have = 4;
    }
    __device__ __forceinline__ uint32_t next() {
        if (have == 0) refill();
        // fixed consumption order out[0..3] without dynamic register indexing
        uint32_t v = have == 4 ? out[0] : have == 3 ? out[1] : have == 2 ? out[2] : out[3];
        have--;
        return v;
    }
    // uniform in (0, 1): never 0, never 1
    __device__ __forceinline__ float unif() { return ((float)(next() >> 8) + 0.5f) * (1.0f / 16777216.0f); }
    __device__ __forceinline__ float expon() { return -__logf(unif()); }
    __device__ __forceinline__ float normal() {
        float u1 = unif(), u2 = unif();
        return sqrtf(-2.0f * __logf(u1)) * __cosf(2.0f * GK_PI_F * u2);
    }
};

// log Phi(x), stable for very negative x (asymptotic series below -5)
__device__ __forceinline__ float gk_log_ndtr(float x) {
    if (x > -5.0f) return __logf(0.5f * erfcf(-x * 0.70710678f));
    float x2 = x * x, r = 1.0f / x2;
    return -0.5f * x2 - __logf(-x) - 0.91893853f + __logf(1.0f - r + 3.0f * r * r - 15.0f * r * r * r);
}

// a_n(x): coefficients of the alternating series of the J*(1,0) density (left form for x <= t, right form above)
__device__ __forceinline__ float gk_pg_acoef(int n, float x, float logx) {
    float K = ((float)n + 0.5f) * GK_PI_F;
    if (x > GK_PG_T) return K * __expf(-0.5f * K * K * x);
    float h = (float)n + 0.5f;
    return __expf(-1.5f * (0.45158271f + logx) + __logf(K) - 2.0f * h * h / x);   // 0.4515827 = log(pi/2)
}

// inverse-Gaussian(mu = 1/Z, lambda = 1) truncated to (0, t]
__device__ __forceinline__ float gk_pg_rtigauss(float Z, gk_philox &g) {
    const float t = GK_PG_T;
    float X = t + 1.0f;
    if (1.0f / t > Z) {
        float alpha = 0.0f;
        int guard = 0;
        while (g.unif() > alpha && guard++ < 1000) {
            float E1 = g.expon(), E2 = g.expon();
            int guard2 = 0;
            while (E1 * E1 > 2.0f * E2 / t && guard2++ < 1000) { E1 = g.expon(); E2 = g.expon(); }
            X = 1.0f + E1 * t;
            X = t / (X * X);
            alpha = __expf(-0.5f * Z * Z * X);
        }
    } else {
        float mu = 1.0f / Z;
        int guard = 0;
        while (X > t && guard++ < 1000) {
            float Y = g.normal(); Y *= Y;
            float half_mu = 0.5f * mu, mu_Y = mu * Y;
            X = mu + half_mu * mu_Y - half_mu * sqrtf(4.0f * mu_Y + mu_Y * mu_Y);
            if (g.unif() > mu / (mu + X)) X = mu * mu / X;
        }
    }
    return X;
}

// omega ~ PG(1, z)
__device__ __forceinline__ float gk_pg1_draw(float z, gk_philox &g) {
    const float t = GK_PG_T;
    float Z = 0.5f * fabsf(z);
    float fz = 0.125f * GK_PI_F * GK_PI_F + 0.5f * Z * Z;
    // probability of the exponential (right) piece of the proposal
    float rt = 1.25f;                                   // sqrt(1/t)
    float b = rt * (t * Z - 1.0f), a = -rt * (t * Z + 1.0f);
    float x0 = __logf(fz) + fz * t;
    float xb = x0 - Z + gk_log_ndtr(b), xa = x0 + Z + gk_log_ndtr(a);
    float qdivp = 1.27323954f * (__expf(fminf(xb, 80.f)) + __expf(fminf(xa, 80.f)));
    float mass = 1.0f / (1.0f + qdivp);
    for (int iter = 0; iter < 1000; ++iter) {
        float X;
        if (g.unif() < mass) X = t + g.expon() / fz;
        else X = gk_pg_rtigauss(Z, g);
        float logx = __logf(X);
        float S = gk_pg_acoef(0, X, logx);
        float Y = g.unif() * S;
        int n = 0;
        bool done = false, accept = false;
        while (!done && n < 200) {
            ++n;
            float an = gk_pg_acoef(n, X, logx);
            if (n & 1) { S -= an; if (Y <= S) { accept = true; done = true; } }
            else { S += an; if (Y > S) done = true; }
        }
        if (accept) return 0.25f * X;
    }
    return 0.25f * tanhf(Z) / fmaxf(Z, 1e-6f) * 0.5f;    // unreachable in practice: fall back to the mean
}

// ---- lane-queue form of the same sampler (K3 tile kernel) ---------------------------------------------
// A rejection sampler on a 32-wide warp idles every lane that has already accepted.  Here each lane owns a
// queue of entries and keeps proposing for ITS current entry until the queue is empty, so lanes only idle at
// the very end of the queue.  The alternating-series test is done in ratio form r_n = a_n / a_0 =
// (2n+1) exp(n (n+1) e), e = -pi^2 x / 2 (x > t) or -2 / x (x <= t): one exponential, no a_0.
struct gk_xoshiro {                      // xoshiro128++ (Blackman & Vigna), seeded from one Philox block
    uint32_t s0, s1, s2, s3;
    __device__ __forceinline__ void seed(uint64_t key, uint32_t a, uint32_t b, uint32_t c) {
        gk_philox ph; ph.init(key, a, b, c); ph.refill();
        s0 = ph.out[0]; s1 = ph.out[1]; s2 = ph.out[2]; s3 = ph.out[3];
        if ((s0 | s1 | s2 | s3) == 0u) s0 = 1u;
    }
    __device__ __forceinline__ uint32_t next() {
        const uint32_t r = __funnelshift_l(s0 + s3, s0 + s3, 7) + s0;
        const uint32_t t = s1 << 9;
        s2 ^= s0; s3 ^= s1; s1 ^= s2; s0 ^= s3; s2 ^= t; s3 = __funnelshift_l(s3, s3, 11);
        return r;
    }
    __device__ __forceinline__ float unif() { return ((float)(next() >> 8) + 0.5f) * (1.0f / 16777216.0f); }
    __device__ __forceinline__ float expon() { return -__logf(unif()); }
    __device__ __forceinline__ float normal() {
        float u1 = unif(), u2 = unif();
        return sqrtf(-2.0f * __logf(u1)) * __cosf(2.0f * GK_PI_F * u2);
    }
};

// probability of the exponential (right) piece of the proposal for J*(1, Z), Z = |z| / 2
__device__ __forceinline__ float gk_pg_mass(float Z) {
    const float t = GK_PG_T, rt = 1.25f;
    const float fz = 0.125f * GK_PI_F * GK_PI_F + 0.5f * Z * Z;
    const float b = rt * (t * Z - 1.0f), a = -rt * (t * Z + 1.0f);
    const float x0 = __logf(fz) + fz * t;
    const float xb = x0 - Z + gk_log_ndtr(b), xa = x0 + Z + gk_log_ndtr(a);
    const float qdivp = 1.27323954f * (__expf(fminf(xb, 80.f)) + __expf(fminf(xa, 80.f)));
    return 1.0f / (1.0f + qdivp);
}

// one proposal step for the lane's current draw.  in_left: the left (truncated inverse-Gaussian) piece was
// chosen and its inner rejection has not succeeded yet (it must be retried, NOT re-chosen).  Returns true and
// sets X (= J*, omega = X / 4) when a variate is accepted.
__device__ __forceinline__ bool gk_pg1_step(float Z, float mass, bool &in_left, gk_xoshiro &g, float &X) {
    const float t = GK_PG_T;
    bool have = false;
    float x = 0.f;
    if (!in_left) {
        if (g.unif() < mass) {
            const float fz = 0.125f * GK_PI_F * GK_PI_F + 0.5f * Z * Z;
            x = t + __fdividef(g.expon(), fz);
            have = true;
        } else in_left = true;
    }
    if (in_left) {
        if (Z < 1.0f / t) {
            const float E1 = g.expon(), E2 = g.expon();
            x = 1.0f + E1 * t;
            x = __fdividef(t, x * x);
            const bool ok = (E1 * E1 <= 2.0f * E2 / t) && (g.unif() <= __expf(-0.5f * Z * Z * x));
            if (ok) { have = true; in_left = false; }
        } else {
            const float mu = __fdividef(1.0f, Z);
            float Y = g.normal(); Y *= Y;
            const float half_mu = 0.5f * mu, mu_Y = mu * Y;
            x = mu + half_mu * mu_Y - half_mu * sqrtf(4.0f * mu_Y + mu_Y * mu_Y);
            if (g.unif() > __fdividef(mu, mu + x)) x = __fdividef(mu * mu, x);
            if (x <= t) { have = true; in_left = false; }
        }
    }
    if (!have) return false;
    const float e = x > t ? -0.5f * GK_PI_F * GK_PI_F * x : __fdividef(-2.0f, x);
    const float u = g.unif();
    float S = 1.0f - 3.0f * __expf(2.0f * e);
    bool accept = u <= S;
    if (!accept) {                                   // rare (< 0.6 %): continue the alternating series
        for (int n = 2; n < 24; ++n) {
            const float r = (float)(2 * n + 1) * __expf((float)(n * (n + 1)) * e);
            if (n & 1) { S -= r; if (u <= S) { accept = true; break; } }
            else { S += r; if (u > S) break; }
        }
    }
    if (accept) X = x;
    return accept;
}

// ---- piece-sorted form (K3 tile kernel, pass 2) ----------------------------------------------------------
// gk_pg1_step makes every lane walk through right piece / left piece (small Z) / left piece (inverse Gaussian) code
// whether or not it needs it (ncu: 11 of 32 lanes active per instruction).  The tile kernel instead decides the
// piece of every draw first (uniform work), finishes right-piece draws in place and files left-piece draws into two
// per-warp lists that are then drained densely, one code path at a time.  The alternating-series test rejects
// < 0.1 % of proposals; those restart through the generic step.
__device__ __forceinline__ bool gk_pg_series_accept(float e, gk_xoshiro &g) {
    const float u = g.unif();
    float S = 1.0f - 3.0f * __expf(2.0f * e);
    if (u <= S) return true;
    for (int n = 2; n < 24; ++n) {
        const float r = (float)(2 * n + 1) * __expf((float)(n * (n + 1)) * e);
        if (n & 1) { S -= r; if (u <= S) return true; }
        else { S += r; if (u > S) return false; }
    }
    return false;
}

__device__ __noinline__ float gk_pg1_generic(float Z, gk_xoshiro &g) {
    const float mass = gk_pg_mass(Z);
    bool in_left = false;
    float X = 0.f;
    while (!gk_pg1_step(Z, mass, in_left, g, X)) {}
    return X;
}

// right (exponential) piece: one proposal, then the series test
__device__ __forceinline__ float gk_pg1_right(float Z, gk_xoshiro &g) {
    const float fz = 0.125f * GK_PI_F * GK_PI_F + 0.5f * Z * Z;
    const float x = GK_PG_T + __fdividef(g.expon(), fz);
    if (gk_pg_series_accept(-0.5f * GK_PI_F * GK_PI_F * x, g)) return x;
    return gk_pg1_generic(Z, g);
}

// left piece, Z < 1/t: exponential-tail proposal for 1/sqrt(x) with inner rejection
__device__ __forceinline__ float gk_pg1_left_small(float Z, gk_xoshiro &g) {
    const float t = GK_PG_T;
    float x;
    for (;;) {
        const float E1 = g.expon(), E2 = g.expon();
        x = 1.0f + E1 * t;
        x = __fdividef(t, x * x);
        if (E1 * E1 <= 2.0f * E2 / t && g.unif() <= __expf(-0.5f * Z * Z * x)) break;
    }
    if (gk_pg_series_accept(__fdividef(-2.0f, x), g)) return x;
    return gk_pg1_generic(Z, g);
}

// left piece, Z >= 1/t: inverse Gaussian(1/Z, 1) truncated to x <= t by rejection
__device__ __forceinline__ float gk_pg1_left_ig(float Z, gk_xoshiro &g) {
    const float t = GK_PG_T, mu = __fdividef(1.0f, Z), half_mu = 0.5f * mu;
    float x;
    do {
        float Y = g.normal(); Y *= Y;
        const float mu_Y = mu * Y;
        x = mu + half_mu * mu_Y - half_mu * sqrtf(4.0f * mu_Y + mu_Y * mu_Y);
        if (g.unif() > __fdividef(mu, mu + x)) x = __fdividef(mu * mu, x);
    } while (x > t);
    if (gk_pg_series_accept(__fdividef(-2.0f, x), g)) return x;
    return gk_pg1_generic(Z, g);
}
