/*
 * fhe_oracle.c — CPU restatement of OpenFHE's native DCRTPoly add/mul/NTT path.
 * TEST INFRASTRUCTURE ONLY (see fhe_oracle.h).  Citations relative to /root/reference.
 */
#include "fhe_oracle.h"
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef unsigned __int128 u128;

/* lbcrypto::GetMSB (src/core/include/utils/inttypes / nbtheory): 64 - clz, 0 for 0 */
uint32_t orc_msb(uint64_t x) { return x ? 64u - (uint32_t)__builtin_clzll(x) : 0u; }

/* ubintnat.h:650-656: mu = floor(2^(2*msb(q)+3) / q) */
uint64_t orc_compute_mu(uint64_t q) {
    u128 t = (u128)1 << (2 * orc_msb(q) + 3);
    return (uint64_t)(t / (u128)q);
}

/* ubintnat.h:760-767 */
uint64_t orc_modadd_fast(uint64_t a, uint64_t b, uint64_t q) {
    uint64_t r = a + b;
    if (r >= q) r -= q;
    return r;
}

/* ubintnat.h:934-938 */
uint64_t orc_modsub_fast(uint64_t a, uint64_t b, uint64_t q) {
    if (a < b) return a + q - b;
    return a - b;
}

/* ubintnat.h:1398-1413 (generalized Barrett, alpha = n+3); MultD :1919-1932, RShiftD :1901-1903 */
uint64_t orc_modmul_barrett(uint64_t a, uint64_t b, uint64_t q, uint64_t mu) {
    u128 prod = (u128)a * b;
    int64_t n = (int64_t)orc_msb(q) - 2;
    uint64_t lo = (uint64_t)prod, hi = (uint64_t)(prod >> 64);
    uint64_t t = (lo >> n) | (hi << (64 - n));      /* RShiftD, truncated to one word */
    u128 t2 = (u128)t * mu;
    u128 rv = prod - (u128)q * (t2 >> (n + 7));
    uint64_t r = (uint64_t)rv;
    if (r >= q) r -= q;
    return r;
}

/* ubintnat.h:1447-1455 */
uint64_t orc_prep_mulconst(uint64_t w, uint64_t q) { return (uint64_t)(((u128)w << 64) / (u128)q); }

/* ubintnat.h:1491-1497 */
uint64_t orc_modmul_const(uint64_t a, uint64_t w, uint64_t q, uint64_t wp) {
    uint64_t qhat = (uint64_t)(((u128)a * wp) >> 64) + 1;
    int64_t y = (int64_t)(a * w - qhat * q);
    return (uint64_t)(y >= 0 ? y : y + (int64_t)q);
}

uint64_t orc_modexp(uint64_t b, uint64_t e, uint64_t q) {
    uint64_t r = 1 % q;
    b %= q;
    while (e) {
        if (e & 1) r = (uint64_t)(((u128)r * b) % q);
        b = (uint64_t)(((u128)b * b) % q);
        e >>= 1;
    }
    return r;
}

/* extended Euclid; q need not be prime (used for 2^-k) */
uint64_t orc_modinv(uint64_t a, uint64_t q) {
    __int128 t = 0, nt = 1, r = q, nr = a % q;
    while (nr != 0) {
        __int128 qt = r / nr, tmp;
        tmp = t - qt * nt; t = nt; nt = tmp;
        tmp = r - qt * nr; r = nr; nr = tmp;
    }
    if (t < 0) t += q;
    return (uint64_t)t;
}

/* Deterministic Miller-Rabin for 64-bit (bases 2..37): same verdict as the reference's
 * probabilistic MillerRabinPrimalityTest (nbtheory-impl.h) on true primes/composites. */
int orc_is_prime(uint64_t p) {
    static const uint64_t bases[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    if (p < 2) return 0;
    for (size_t i = 0; i < 12; i++) {
        if (p == bases[i]) return 1;
        if (p % bases[i] == 0) return 0;
    }
    uint64_t d = p - 1; uint32_t s = 0;
    while (!(d & 1)) { d >>= 1; s++; }
    for (size_t i = 0; i < 12; i++) {
        uint64_t x = orc_modexp(bases[i], d, p);
        if (x == 1 || x == p - 1) continue;
        int comp = 1;
        for (uint32_t r = 1; r < s; r++) {
            x = (uint64_t)(((u128)x * x) % p);
            if (x == p - 1) { comp = 0; break; }
        }
        if (comp) return 0;
    }
    return 1;
}

/* nbtheory-impl.h:334-359 */
uint64_t orc_first_prime(uint32_t nbits, uint64_t m) {
    uint64_t q = (uint64_t)1 << nbits;
    uint64_t r = q % m;
    uint64_t q2 = q + 1;
    if (r > 0) q2 += (m - r);
    while (!orc_is_prime(q2)) q2 += m;
    return q2;
}
/* nbtheory-impl.h:361-369 */
uint64_t orc_next_prime(uint64_t q, uint64_t m) {
    uint64_t c = q + m;
    while (!orc_is_prime(c)) c += m;
    return c;
}
/* nbtheory-impl.h:372-379 */
uint64_t orc_previous_prime(uint64_t q, uint64_t m) {
    uint64_t c = q - m;
    while (!orc_is_prime(c)) c -= m;
    return c;
}

/* nbtheory-impl.h:183-231.  The reference takes a random generator g, r = g^((q-1)/m), then
 * returns the MINIMUM over all primitive m-th roots (odd powers of r for power-of-two m).
 * The minimum does not depend on which primitive root one starts from, so any r with
 * r^(m/2) = -1 reproduces the reference's return value. */
uint64_t orc_root_of_unity(uint64_t m, uint64_t q) {
    if ((q - 1) % m != 0) return 0;
    uint64_t e = (q - 1) / m, r = 0;
    for (uint64_t g = 2; g < q; g++) {
        r = orc_modexp(g, e, q);
        if (orc_modexp(r, m / 2, q) == q - 1) break;
    }
    uint64_t r2 = (uint64_t)(((u128)r * r) % q);
    uint64_t x = r, min = r;
    for (uint64_t k = 1; k < m; k += 2) {       /* x = r^k over odd k */
        if (x < min && x != 1) min = x;
        x = (uint64_t)(((u128)x * r2) % q);
    }
    return min;
}

void orc_gen_dcrt_params(uint32_t nbits, uint32_t N, uint32_t T, uint64_t* q, uint64_t* psi) {
    uint64_t m = 2ull * N;
    uint64_t first = orc_first_prime(nbits, m);
    for (uint32_t i = 0; i < T; i++) {
        q[i] = orc_previous_prime(i ? q[i - 1] : first, m);
        psi[i] = orc_root_of_unity(m, q[i]);
    }
}

/* ---- one-tower vector ops ---- */
/* mubintvecnat.h:426-432 / mubintvecnat.cpp:245-264 */
void orc_vec_modadd(uint64_t* dst, const uint64_t* a, const uint64_t* b, size_t n, uint64_t q) {
    for (size_t i = 0; i < n; i++) dst[i] = orc_modadd_fast(a[i], b[i], q);
}
/* mubintvecnat.cpp:301-307 */
void orc_vec_modsub(uint64_t* dst, const uint64_t* a, const uint64_t* b, size_t n, uint64_t q) {
    for (size_t i = 0; i < n; i++) dst[i] = orc_modsub_fast(a[i], b[i], q);
}
/* mubintvecnat.h:501-513: mu once per vector, Barrett per element */
void orc_vec_modmul(uint64_t* dst, const uint64_t* a, const uint64_t* b, size_t n, uint64_t q) {
    uint64_t mu = orc_compute_mu(q);
    for (size_t i = 0; i < n; i++) dst[i] = orc_modmul_barrett(a[i], b[i], q, mu);
}
/* scalar forms: mubintvecnat.cpp:198-218 (ModAddEq scalar reduces b mod q first), :266-284, :310-332 */
void orc_vec_modadd_scalar(uint64_t* dst, const uint64_t* a, uint64_t s, size_t n, uint64_t q) {
    if (s >= q) s %= q;
    for (size_t i = 0; i < n; i++) dst[i] = orc_modadd_fast(a[i], s, q);
}
void orc_vec_modsub_scalar(uint64_t* dst, const uint64_t* a, uint64_t s, size_t n, uint64_t q) {
    if (s >= q) s %= q;
    for (size_t i = 0; i < n; i++) dst[i] = orc_modsub_fast(a[i], s, q);
}
void orc_vec_modmul_scalar(uint64_t* dst, const uint64_t* a, uint64_t s, size_t n, uint64_t q) {
    if (s >= q) s %= q;
    uint64_t mu = orc_compute_mu(q);
    for (size_t i = 0; i < n; i++) dst[i] = orc_modmul_barrett(a[i], s, q, mu);
}

/* ---- mod-ops of the PimData surface (pim/PimData.h:174-233 -> mod_kernel, pim/common.h:32-46).
 * CPU definitions: NativeVectorT::SwitchModulus / Mod / ModEq  mubintvecnat.cpp:111-195,
 * ModByTwo(Eq) :370-385, ModExp(Eq) :387-410, ModInverse(Eq) mubintvecnat.h:546-566,
 * MultiplyAndRound(Eq) :423-455, DivideAndRound(Eq) :457-487; scalars ubintnat.h:536-541,
 * 596-606, 1507-1530, 1565-1592.  kind is enum mod_kernel; the Eq and non-Eq forms compute the
 * same values.  Returns 0, or 1 when ModInverse meets an element without an inverse (the
 * reference throws math_error there) or a divisor is zero. */
static uint64_t modexp_ref(uint64_t v, uint64_t b, uint64_t m) {   /* ubintnat.h:1507-1530 */
    u128 t = v, pw = b, mm = m, r = 1;
    if (pw & 1) { r = r * t; if (r >= mm) r %= mm; }
    while (pw >>= 1) {
        t = t * t; if (t >= mm) t %= mm;
        if (pw & 1) { r = r * t; if (r >= mm) r %= mm; }
    }
    return (uint64_t)r;
}
static int modinverse_ref(uint64_t v, uint64_t m, uint64_t* out) {   /* ubintnat.h:1565-1592 */
    int64_t modulus = (int64_t)m, a = (int64_t)(v % m);
    if (a == 0) return 1;
    if (modulus == 1) { *out = 0; return 0; }
    int64_t y = 0, x = 1;
    while (a > 1) {
        int64_t t = modulus, qq = a / t;
        modulus = a % t; a = t; t = y;
        y = x - qq * y; x = t;
    }
    if (x < 0) x += (int64_t)m;
    *out = (uint64_t)x;
    return 0;
}
static uint64_t mulround_ref(uint64_t v, uint64_t p, uint64_t q) {    /* ubintnat.h:536-541 */
    volatile double ratio = (double)v / (double)q;       /* volatile: no fused multiply-add */
    volatile double prod = (double)p * ratio;
    return (uint64_t)(prod + 0.5);
}
static uint64_t divround_ref(uint64_t v, uint64_t q) {               /* ubintnat.h:596-606 */
    uint64_t ans = v / q, rem = v % q;
    return rem > (q >> 1) ? ans + 1 : ans;
}
int orc_vec_modop(int kind, uint64_t* dst, const uint64_t* src, size_t n, uint64_t mod, uint64_t p, uint64_t q) {
    uint64_t half = mod >> 1;
    int bad = 0;
    switch (kind) {
    case 1: case 2:                                   /* MOD_NORMAL, MODEQ: nm == 2 -> ModByTwo */
        if (p == 2) { for (size_t i = 0; i < n; i++) dst[i] = 1 & (src[i] ^ (uint64_t)(src[i] > half)); return 0; }
        /* fall through */
    case 0: {                                         /* SWITCHMODULUS */
        uint64_t om = mod, nm = p;
        if (nm == 0) return 1;
        if (nm > om) {
            uint64_t diff = nm - om;
            for (size_t i = 0; i < n; i++) { uint64_t v = src[i]; if (v > half) v += diff; dst[i] = v; }
        } else {
            uint64_t diff = nm - (om % nm);
            for (size_t i = 0; i < n; i++) {
                uint64_t v = src[i];
                if (v > half) v += diff;
                if (v >= nm) v %= nm;
                dst[i] = v;
            }
        }
        return 0;
    }
    case 3: case 4:                                   /* MODBYTWO(EQ) */
        for (size_t i = 0; i < n; i++) dst[i] = 1 & (src[i] ^ (uint64_t)(src[i] > half));
        return 0;
    case 5: case 6: {                                 /* MODEXP(EQ): exponent reduced mod the modulus */
        if (mod == 0) return 1;
        uint64_t b = p >= mod ? p % mod : p;
        for (size_t i = 0; i < n; i++) dst[i] = modexp_ref(src[i], b, mod);
        return 0;
    }
    case 7: case 8:                                   /* MODINVERSE(EQ) */
        if (mod == 0) return 1;
        for (size_t i = 0; i < n; i++) { uint64_t r = 0; bad |= modinverse_ref(src[i], mod, &r); dst[i] = r; }
        return bad;
    case 9: case 10:                                  /* MULTIPLYANDROUND(EQ) */
        if (q == 0 || mod == 0) return 1;
        for (size_t i = 0; i < n; i++) {
            uint64_t v = src[i];
            if (v > half) dst[i] = mod - mulround_ref(mod - v, p, q);
            else dst[i] = mulround_ref(v, p, q) % mod;
        }
        return 0;
    case 11: case 12:                                 /* DIVIDEANDROUND(EQ): divisor travels in p */
        if (p == 0) return 1;
        for (size_t i = 0; i < n; i++) {
            uint64_t v = src[i];
            if (v > half) dst[i] = mod - divround_ref(mod - v, p);
            else dst[i] = divround_ref(v, p);
        }
        return 0;
    }
    return 1;
}

/* lbcrypto::ReverseBits(i, msb) */
static uint32_t bitrev(uint32_t x, uint32_t bits) {
    uint32_t r = 0;
    for (uint32_t i = 0; i < bits; i++) { r = (r << 1) | (x & 1); x >>= 1; }
    return r;
}

/* transformnat-impl.h:708-763 */
void orc_precompute(uint64_t psi, uint32_t N, uint64_t q,
                    uint64_t* W, uint64_t* Wp, uint64_t* WI, uint64_t* WIp,
                    uint64_t* ninv, uint64_t* ninvp) {
    uint32_t msb = orc_msb((uint64_t)N - 1);      /* = log2 N */
    uint64_t mu = orc_compute_mu(q);
    uint64_t psiinv = orc_modinv(psi, q);
    uint64_t x = 1, xinv = 1;
    for (uint32_t i = 0; i < N; i++) {            /* :724-730 */
        uint32_t r = bitrev(i, msb);
        W[r] = x; WI[r] = xinv;
        x = orc_modmul_barrett(x, psi, q, mu);
        xinv = orc_modmul_barrett(xinv, psiinv, q, mu);
    }
    for (uint32_t i = 0; i < N; i++) {            /* :745-750 */
        Wp[i] = orc_prep_mulconst(W[i], q);
        WIp[i] = orc_prep_mulconst(WI[i], q);
    }
    *ninv = orc_modinv((uint64_t)1 << msb, q);    /* TableCOI[msb], :734-739 */
    *ninvp = orc_prep_mulconst(*ninv, q);         /* :752-756 */
}

/* transformnat-impl.h:300-354 (GCC branch): CT, natural in -> bit-reversed out */
void orc_ntt_fwd(uint64_t* x, uint32_t N, uint64_t q, const uint64_t* W, const uint64_t* Wp) {
    uint32_t n = N >> 1, t = n, logt = orc_msb(t);
    for (uint32_t m = 1; m < n; m <<= 1, t >>= 1, --logt) {
        for (uint32_t i = 0; i < m; ++i) {
            uint64_t w = W[i + m], wp = Wp[i + m];
            for (uint32_t j1 = i << logt, j2 = j1 + t; j1 < j2; ++j1) {
                uint64_t of = orc_modmul_const(x[j1 + t], w, q, wp);
                uint64_t lo = x[j1];
                uint64_t hi = lo + of;
                if (hi >= q) hi -= q;
                if (lo < of) lo += q;
                lo -= of;
                x[j1] = hi; x[j1 + t] = lo;
            }
        }
    }
    for (uint32_t i = 0; i < (n << 1); i += 2) {  /* peeled t = 1 stage, :332-353 */
        uint64_t w = W[(i >> 1) + n], wp = Wp[(i >> 1) + n];
        uint64_t of = orc_modmul_const(x[i + 1], w, q, wp);
        uint64_t lo = x[i];
        uint64_t hi = lo + of;
        if (hi >= q) hi -= q;
        if (lo < of) lo += q;
        lo -= of;
        x[i] = hi; x[i + 1] = lo;
    }
}

/* transformnat-impl.h:492-552 (GCC branch): GS, bit-reversed in -> natural out, N^-1 folded in stage 1 */
void orc_ntt_inv(uint64_t* x, uint32_t N, uint64_t q, const uint64_t* WI, const uint64_t* WIp,
                 uint64_t ninv, uint64_t ninvp) {
    uint32_t n = N;
    for (uint32_t i = 0; i < n; i += 2) {
        uint64_t w = WI[(i + n) >> 1], wp = WIp[(i + n) >> 1];
        uint64_t hi = x[i + 1], lo = x[i];
        uint64_t of = lo;
        if (of < hi) of += q;
        of -= hi;
        lo += hi;
        if (lo >= q) lo -= q;
        lo = orc_modmul_const(lo, ninv, q, ninvp);
        of = orc_modmul_const(of, w, q, wp);
        of = orc_modmul_const(of, ninv, q, ninvp);
        x[i] = lo; x[i + 1] = of;
    }
    for (uint32_t m = n >> 2, t = 2, logt = 2; m >= 1; m >>= 1, t <<= 1, ++logt) {
        for (uint32_t i = 0; i < m; ++i) {
            uint64_t w = WI[i + m], wp = WIp[i + m];
            for (uint32_t j1 = i << logt, j2 = j1 + t; j1 < j2; ++j1) {
                uint64_t hi = x[j1 + t], lo = x[j1];
                uint64_t of = lo;
                if (of < hi) of += q;
                of -= hi;
                lo += hi;
                if (lo >= q) lo -= q;
                of = orc_modmul_const(of, w, q, wp);
                x[j1] = lo; x[j1 + t] = of;
            }
        }
    }
}

/* ---- DCRTPoly level ---- */
struct orc_ctx {
    uint32_t N, T;
    uint64_t *q, *psi, *ninv, *ninvp;
    uint64_t *W, *Wp, *WI, *WIp;      /* [T][N] each */
};

orc_ctx* orc_ctx_create(uint32_t N, uint32_t T, const uint64_t* q, const uint64_t* psi) {
    orc_ctx* c = (orc_ctx*)calloc(1, sizeof(orc_ctx));
    c->N = N; c->T = T;
    c->q = (uint64_t*)malloc(8 * T); c->psi = (uint64_t*)malloc(8 * T);
    c->ninv = (uint64_t*)malloc(8 * T); c->ninvp = (uint64_t*)malloc(8 * T);
    size_t tn = (size_t)T * N;
    c->W = (uint64_t*)malloc(8 * tn); c->Wp = (uint64_t*)malloc(8 * tn);
    c->WI = (uint64_t*)malloc(8 * tn); c->WIp = (uint64_t*)malloc(8 * tn);
    memcpy(c->q, q, 8 * T); memcpy(c->psi, psi, 8 * T);
#pragma omp parallel for schedule(dynamic)
    for (uint32_t t = 0; t < T; t++)
        orc_precompute(psi[t], N, q[t], c->W + (size_t)t * N, c->Wp + (size_t)t * N,
                       c->WI + (size_t)t * N, c->WIp + (size_t)t * N, &c->ninv[t], &c->ninvp[t]);
    return c;
}

void orc_ctx_destroy(orc_ctx* c) {
    if (!c) return;
    free(c->q); free(c->psi); free(c->ninv); free(c->ninvp);
    free(c->W); free(c->Wp); free(c->WI); free(c->WIp);
    free(c);
}

/* ---- automorphism ---- */
static uint32_t ilog2_u32(uint32_t n) { uint32_t l = 0; while ((1u << l) < n) l++; return l; }

/* poly-impl.h:312-365: EVALUATION: result[rev(j)] = in[rev(((2j+1)k >> 1) & mask)];
 * COEFFICIENT: result[jk & mask] = bit logn of jk set ? q - in[j] : in[j]  (32-bit wrapping jk). */
int orc_automorphism(uint64_t* out, const uint64_t* in, uint32_t N, uint64_t q, uint32_t k, int format) {
    if (k % 2 == 0) return 1;                                    /* "Automorphism index not odd" */
    uint32_t logn = ilog2_u32(N), mask = (uint32_t)((1ull << logn) - 1);
    if (format == 0) {
        uint32_t jk = k;
        for (uint32_t j = 0; j < N; j++, jk += 2 * k)
            out[bitrev(j, logn)] = in[bitrev((jk >> 1) & mask, logn)];
    } else {
        uint32_t jk = 0;
        for (uint32_t j = 0; j < N; j++, jk += k)
            out[jk & mask] = ((jk >> logn) & 1) ? q - in[j] : in[j];
    }
    return 0;
}

/* nbtheory2.cpp:274-286 */
void orc_precompute_automap(uint32_t N, uint32_t k, uint32_t* map) {
    uint32_t logn = ilog2_u32(N), logm = logn + 1;
    for (uint32_t j = 0; j < N; j++) {
        uint32_t jt = (j << 1) + 1;
        uint32_t idx = ((jt * k) - (((jt * k) >> logm) << logm)) >> 1;
        map[bitrev(j, logn)] = bitrev(idx, logn);
    }
}

/* poly-impl.h:368-380 */
void orc_automorphism_map(uint64_t* out, const uint64_t* in, uint32_t N, const uint32_t* map) {
    for (uint32_t j = 0; j < N; j++) out[j] = in[map[j]];
}

/* dcrtpoly-impl.h:348-369: tower by tower */
int orc_dcrt_automorphism(const orc_ctx* c, uint64_t* out, const uint64_t* in, size_t B, uint32_t k, int format) {
    if (k % 2 == 0) return 1;
    long units = (long)(B * c->T);
#pragma omp parallel for schedule(static)
    for (long u = 0; u < units; u++)
        orc_automorphism(out + (size_t)u * c->N, in + (size_t)u * c->N, c->N, c->q[u % c->T], k, format);
    return 0;
}

/* ---- RNS basis conversion and scaling ---- */
static uint64_t mulmod_u128(uint64_t a, uint64_t b, uint64_t m) { return (uint64_t)(((u128)a * b) % m); }

void orc_switch_tables(const uint64_t* q, uint32_t sq, const uint64_t* p, uint32_t sp, uint64_t* hat_inv, uint64_t* hat_modp) {
    for (uint32_t i = 0; i < sq; i++) {
        uint64_t h = 1 % q[i];
        for (uint32_t k = 0; k < sq; k++) if (k != i) h = mulmod_u128(h, q[k] % q[i], q[i]);
        hat_inv[i] = orc_modinv(h, q[i]);
        for (uint32_t j = 0; j < sp; j++) {
            uint64_t g = 1 % p[j];
            for (uint32_t k = 0; k < sq; k++) if (k != i) g = mulmod_u128(g, q[k] % p[j], p[j]);
            hat_modp[(size_t)i * sp + j] = g;
        }
    }
}

/* Ql * [Ql^-1]_ql = 1 + m*ql, so floor(.../ql) = m, and m = -ql^-1 (mod q_i) because q_i divides Ql. */
void orc_rescale_tables(const uint64_t* q, uint32_t T, uint64_t* qlql, uint64_t* qlinv) {
    uint32_t l = T - 1;
    for (uint32_t i = 0; i < l; i++) {
        qlinv[i] = orc_modinv(q[l] % q[i], q[i]);
        qlql[i] = (q[i] - qlinv[i]) % q[i];
    }
}

static int add_carry_out(uint64_t a, uint64_t b, uint64_t* c) { *c = a + b; return *c < a; }

/* utils/utilities-int.h:60-110, statement by statement */
static uint64_t barrett128(u128 a, uint64_t modulus, u128 mu) {
    uint64_t a_lo = (uint64_t)a, a_hi = (uint64_t)(a >> 64), mu_lo = (uint64_t)mu, mu_hi = (uint64_t)(mu >> 64);
    uint64_t left_hi = (uint64_t)(((u128)a_lo * mu_lo) >> 64);
    u128 middle = (u128)a_lo * mu_hi;
    uint64_t middle_lo = (uint64_t)middle, middle_hi = (uint64_t)(middle >> 64), tmp1, tmp2;
    uint64_t carry = (uint64_t)add_carry_out(middle_lo, left_hi, &tmp1);
    tmp2 = middle_hi + carry;
    middle = (u128)a_hi * mu_lo;
    middle_lo = (uint64_t)middle; middle_hi = (uint64_t)(middle >> 64);
    carry = (uint64_t)((uint64_t)(middle_lo + tmp1) < middle_lo);           /* IsAdditionOverflow */
    left_hi = middle_hi + carry;
    tmp1 = a_hi * mu_hi + tmp2 + left_hi;
    uint64_t result = a_lo - tmp1 * modulus;
    while (result >= modulus) result -= modulus;
    return result;
}

void orc_approx_switch(uint64_t* out, const uint64_t* in, size_t B, uint32_t N, const uint64_t* q, uint32_t sq,
                       const uint64_t* p, uint32_t sp, const uint64_t* hat_inv, const uint64_t* hat_modp) {
    uint64_t* pre = (uint64_t*)malloc(sizeof(uint64_t) * sq);
    u128* mu = (u128*)malloc(sizeof(u128) * sp);
    for (uint32_t i = 0; i < sq; i++) pre[i] = orc_prep_mulconst(hat_inv[i], q[i]);
    for (uint32_t j = 0; j < sp; j++) mu[j] = (~(u128)0) / p[j];     /* floor(2^128 / p_j): p_j is not a power of two */
    long items = (long)(B * N);
#pragma omp parallel for schedule(static)
    for (long e = 0; e < items; e++) {
        size_t b = (size_t)e / N, ri = (size_t)e % N;
        u128 sum[64];
        for (uint32_t j0 = 0; j0 < sp; j0 += 64) {
            uint32_t jn = sp - j0 < 64 ? sp - j0 : 64;
            for (uint32_t j = 0; j < jn; j++) sum[j] = 0;
            for (uint32_t i = 0; i < sq; i++) {
                uint64_t xi = in[(b * sq + i) * N + ri];
                uint64_t y = orc_modmul_const(xi, hat_inv[i], q[i], pre[i]);
                for (uint32_t j = 0; j < jn; j++) sum[j] += (u128)y * hat_modp[(size_t)i * sp + j0 + j];
            }
            for (uint32_t j = 0; j < jn; j++) out[(b * sp + j0 + j) * N + ri] = barrett128(sum[j], p[j0 + j], mu[j0 + j]);
        }
    }
    free(pre); free(mu);
}

/* gather / scatter of a tower range between [B][Tsrc][N] and [B][cnt][N] */
static void towers_copy(uint64_t* dst, uint32_t Tdst, uint32_t d0, const uint64_t* src, uint32_t Tsrc, uint32_t s0,
                        uint32_t cnt, size_t B, uint32_t N) {
    for (size_t b = 0; b < B; b++)
        memcpy(dst + (b * Tdst + d0) * N, src + (b * Tsrc + s0) * N, (size_t)cnt * N * 8);
}
/* NativeVectorT::SwitchModulus on one value (mubintvecnat.cpp:112-135) */
static uint64_t switch_modulus1(uint64_t v, uint64_t om, uint64_t nm) {
    uint64_t half = om >> 1;
    if (nm > om) { if (v > half) v += nm - om; return v; }
    if (v > half) v += nm - (om % nm);
    if (v >= nm) v %= nm;
    return v;
}
/* NativeVectorT::ModMulEq(scalar) on one value (mubintvecnat.cpp:323-332) */
static uint64_t mulc1(uint64_t v, uint64_t c, uint64_t q) {
    static __thread uint64_t kc[2], kq[2], kp[2];          /* PrepModMulConst of the last two (c, q) pairs */
    static __thread int turn;
    if (c >= q) c %= q;
    int s = (kq[0] == q && kc[0] == c) ? 0 : (kq[1] == q && kc[1] == c) ? 1 : -1;
    if (s < 0) { s = turn; turn ^= 1; kc[s] = c; kq[s] = q; kp[s] = orc_prep_mulconst(c, q); }
    return orc_modmul_const(v, c, q, kp[s]);
}

void orc_approx_mod_up(const orc_ctx* Q, const orc_ctx* P, uint64_t* out, const uint64_t* in, size_t B, int format) {
    uint32_t N = Q->N, sq = Q->T, sp = P->T;
    size_t wq = B * sq * (size_t)N, wp = B * sp * (size_t)N;
    uint64_t* coef = (uint64_t*)malloc(wq * 8);
    uint64_t* partP = (uint64_t*)malloc(wp * 8);
    uint64_t* hi = (uint64_t*)malloc(sizeof(uint64_t) * sq);
    uint64_t* hm = (uint64_t*)malloc(sizeof(uint64_t) * sq * sp);
    memcpy(coef, in, wq * 8);
    if (format == 0) orc_dcrt_ntt(Q, 1, coef, B, 0);                     /* SetFormat(COEFFICIENT), :1096-1099 */
    orc_switch_tables(Q->q, sq, P->q, sp, hi, hm);
    orc_approx_switch(partP, coef, B, N, Q->q, sq, P->q, sp, hi, hm);    /* :1105-1106 */
    orc_dcrt_ntt(P, 0, partP, B, 0);                                     /* :1113-1116 */
    towers_copy(out, sq + sp, sq, partP, sp, 0, sp, B, N);
    if (format == 0) {
        towers_copy(out, sq + sp, 0, in, sq, 0, sq, B, N);               /* :1119-1123 the stored EVALUATION towers */
    } else {
        orc_dcrt_ntt(Q, 0, coef, B, 0);                                  /* :1124-1130 */
        towers_copy(out, sq + sp, 0, coef, sq, 0, sq, B, N);
    }
    free(coef); free(partP); free(hi); free(hm);
}

void orc_approx_mod_down(const orc_ctx* Q, const orc_ctx* P, uint64_t* out, const uint64_t* in, size_t B, uint64_t t) {
    uint32_t N = Q->N, sq = Q->T, sp = P->T;
    size_t wq = B * sq * (size_t)N, wp = B * sp * (size_t)N;
    uint64_t* partP = (uint64_t*)malloc(wp * 8);
    uint64_t* sw = (uint64_t*)malloc(wq * 8);
    uint64_t* hi = (uint64_t*)malloc(sizeof(uint64_t) * sp);
    uint64_t* hm = (uint64_t*)malloc(sizeof(uint64_t) * sq * sp);
    towers_copy(partP, sp, 0, in, sq + sp, sq, sp, B, N);                /* :1150-1151 */
    orc_dcrt_ntt(P, 1, partP, B, 0);
    if (t)                                                               /* :1152-1154, [t^-1]_{p_j} */
        for (size_t b = 0; b < B; b++)
            for (uint32_t j = 0; j < sp; j++) {
                uint64_t tinv = orc_modinv(t % P->q[j], P->q[j]);
                uint64_t* v = partP + (b * sp + j) * N;
                for (uint32_t n = 0; n < N; n++) v[n] = mulc1(v[n], tinv, P->q[j]);
            }
    orc_switch_tables(P->q, sp, Q->q, sq, hi, hm);
    orc_approx_switch(sw, partP, B, N, P->q, sp, Q->q, sq, hi, hm);      /* :1158-1159 */
    if (t)                                                               /* :1170-1171 */
        for (size_t b = 0; b < B; b++)
            for (uint32_t i = 0; i < sq; i++) {
                uint64_t* v = sw + (b * sq + i) * N;
                for (uint32_t n = 0; n < N; n++) v[n] = mulc1(v[n], t, Q->q[i]);
            }
    orc_dcrt_ntt(Q, 0, sw, B, 0);                                        /* :1172 */
    for (uint32_t i = 0; i < sq; i++) {                                  /* :1173, [P^-1]_{q_i} */
        uint64_t qi = Q->q[i], pm = 1 % qi;
        for (uint32_t j = 0; j < sp; j++) pm = mulmod_u128(pm, P->q[j] % qi, qi);
        uint64_t pinv = orc_modinv(pm, qi);
        for (size_t b = 0; b < B; b++) {
            const uint64_t* x = in + (b * (sq + sp) + i) * N;
            const uint64_t* s2 = sw + (b * sq + i) * N;
            uint64_t* o = out + (b * sq + i) * N;
            for (uint32_t n = 0; n < N; n++) o[n] = mulc1(orc_modsub_fast(x[n], s2[n], qi), pinv, qi);
        }
    }
    free(partP); free(sw); free(hi); free(hm);
}

void orc_rescale(const orc_ctx* full, const orc_ctx* rest, const orc_ctx* last, uint64_t* out, const uint64_t* in,
                 size_t B, int format, uint64_t t) {
    uint32_t N = full->N, T = full->T, l = T - 1;
    uint64_t ql = full->q[l];
    uint64_t* lastp = (uint64_t*)malloc(B * (size_t)N * 8);
    uint64_t* tmp = (uint64_t*)malloc(B * l * (size_t)N * 8);
    uint64_t* c = (uint64_t*)malloc(sizeof(uint64_t) * l);
    uint64_t* qlinv = (uint64_t*)malloc(sizeof(uint64_t) * l);
    orc_rescale_tables(full->q, T, c, qlinv);
    towers_copy(lastp, 1, 0, in, T, l, 1, B, N);
    if (format == 0) orc_dcrt_ntt(last, 1, lastp, B, 0);                 /* lastPoly.SetFormat(COEFFICIENT) */
    if (t) {                                                             /* ModReduce :798: delta *= negtInvModq */
        uint64_t negtinv = ql - orc_modinv(t % ql, ql);
        for (size_t k = 0; k < B * (size_t)N; k++) lastp[k] = mulc1(lastp[k], negtinv, ql);
    }
    for (size_t b = 0; b < B; b++)
        for (uint32_t i = 0; i < l; i++) {
            uint64_t qi = full->q[i];
            uint64_t* v = tmp + (b * l + i) * N;
            for (uint32_t n = 0; n < N; n++) {
                uint64_t x = switch_modulus1(lastp[b * N + n], ql, qi);  /* :757 / :806 */
                v[n] = t ? x : mulc1(x, c[i], qi);                       /* :758 */
            }
        }
    if (format == 0) orc_dcrt_ntt(rest, 0, tmp, B, 0);                   /* :759-760 / :807-808 */
    for (size_t b = 0; b < B; b++)
        for (uint32_t i = 0; i < l; i++) {
            uint64_t qi = full->q[i];
            const uint64_t* x = in + (b * T + i) * N;
            const uint64_t* v = tmp + (b * l + i) * N;
            uint64_t* o = out + (b * l + i) * N;
            for (uint32_t n = 0; n < N; n++)
                o[n] = t ? mulc1(orc_modadd_fast(x[n], mulc1(v[n], t, qi), qi), qlinv[i], qi)   /* :809-810 */
                         : orc_modadd_fast(mulc1(x[n], qlinv[i], qi), v[n], qi);                /* :761-762 */
        }
    if (format == 1 && !t) orc_dcrt_ntt(rest, 0, out, B, 0);             /* :763-764: towers are switched, m_format is not */
    free(lastp); free(tmp); free(c); free(qlinv);
}

/* ---- BFV (HPS) set ---- */
void orc_exact_tables(const uint64_t* q, uint32_t sq, const uint64_t* p, uint32_t sp, double* qinv, uint64_t* alpha_q_modp) {
    for (uint32_t i = 0; i < sq; i++) qinv[i] = 1. / (double)q[i];
    for (uint32_t j = 0; j < sp; j++) {
        uint64_t Qm = 1 % p[j];
        for (uint32_t k = 0; k < sq; k++) Qm = mulmod_u128(Qm, q[k] % p[j], p[j]);
        for (uint32_t a = 0; a <= sq; a++) alpha_q_modp[(size_t)a * sp + j] = mulmod_u128(Qm, a, p[j]);
    }
}

void orc_switch_crt_basis(uint64_t* out, const uint64_t* in, size_t B, uint32_t N, const uint64_t* q, uint32_t sq,
                          const uint64_t* p, uint32_t sp, const uint64_t* hat_inv, const uint64_t* hat_modp,
                          const double* qinv, const uint64_t* alpha_q_modp) {
    uint64_t* pre = (uint64_t*)malloc(sizeof(uint64_t) * sq);
    u128* mu = (u128*)malloc(sizeof(u128) * sp);
    for (uint32_t i = 0; i < sq; i++) pre[i] = orc_prep_mulconst(hat_inv[i], q[i]);
    for (uint32_t j = 0; j < sp; j++) mu[j] = (~(u128)0) / p[j];
    long items = (long)(B * N);
#pragma omp parallel for schedule(static)
    for (long e = 0; e < items; e++) {
        size_t b = (size_t)e / N, ri = (size_t)e % N;
        uint64_t* y = (uint64_t*)malloc(sizeof(uint64_t) * sq);
        volatile double nu = 0.5;                                  /* :1194, :1204: separate multiply and add */
        for (uint32_t i = 0; i < sq; i++) {
            y[i] = orc_modmul_const(in[(b * sq + i) * N + ri], hat_inv[i], q[i], pre[i]);
            volatile double prod = (double)y[i] * qinv[i];
            nu = nu + prod;
        }
        uint32_t alpha = (uint32_t)nu;                             /* :1208 */
        for (uint32_t j = 0; j < sp; j++) {
            u128 cur = 0;
            for (uint32_t i = 0; i < sq; i++) cur += (u128)y[i] * hat_modp[(size_t)i * sp + j];
            uint64_t v = barrett128(cur, p[j], mu[j]);
            out[(b * sp + j) * N + ri] = orc_modsub_fast(v, alpha_q_modp[(size_t)alpha * sp + j], p[j]);   /* :1226 */
        }
        free(y);
    }
    free(pre); free(mu);
}

void orc_expand_crt_basis(const orc_ctx* Q, const orc_ctx* P, uint64_t* out, const uint64_t* in, size_t B, int format,
                          int result_format) {
    uint32_t N = Q->N, sq = Q->T, sp = P->T;
    size_t wq = B * sq * (size_t)N, wp = B * sp * (size_t)N;
    uint64_t* coef = (uint64_t*)malloc(wq * 8);
    uint64_t* partP = (uint64_t*)malloc(wp * 8);
    uint64_t* hi = (uint64_t*)malloc(sizeof(uint64_t) * sq);
    uint64_t* hm = (uint64_t*)malloc(sizeof(uint64_t) * sq * sp);
    double* qinv = (double*)malloc(sizeof(double) * sq);
    uint64_t* aq = (uint64_t*)malloc(sizeof(uint64_t) * (sq + 1) * sp);
    memcpy(coef, in, wq * 8);
    if (format == 0) orc_dcrt_ntt(Q, 1, coef, B, 0);                       /* :1321-1324 */
    orc_switch_tables(Q->q, sq, P->q, sp, hi, hm);
    orc_exact_tables(Q->q, sq, P->q, sp, qinv, aq);
    orc_switch_crt_basis(partP, coef, B, N, Q->q, sq, P->q, sp, hi, hm, qinv, aq);
    if (result_format == 0) orc_dcrt_ntt(P, 0, partP, B, 0);               /* :1338-1341 SetFormat(resultFormat) */
    towers_copy(out, sq + sp, sq, partP, sp, 0, sp, B, N);
    if (result_format == 0) {
        if (format == 0) {
            towers_copy(out, sq + sp, 0, in, sq, 0, sq, B, N);             /* :1346-1349 */
        } else {
            orc_dcrt_ntt(Q, 0, coef, B, 0);                                /* :1350-1355 */
            towers_copy(out, sq + sp, 0, coef, sq, 0, sq, B, N);
        }
    } else {
        towers_copy(out, sq + sp, 0, coef, sq, 0, sq, B, N);               /* the Q towers stay in COEFFICIENT */
    }
    free(coef); free(partP); free(hi); free(hm); free(qinv); free(aq);
}

/* DCRTPolyImpl::ExpandCRTBasisReverseOrder (dcrtpoly-impl.h:1361-1410): the same values as ExpandCRTBasis with the
 * new towers first (temp = [partP | Q], :1381-1385). */
void orc_expand_crt_basis_reverse(const orc_ctx* Q, const orc_ctx* P, uint64_t* out, const uint64_t* in, size_t B, int format,
                                  int result_format) {
    uint32_t N = Q->N, sq = Q->T, sp = P->T;
    uint64_t* tmp = (uint64_t*)malloc(B * (size_t)(sq + sp) * N * 8);
    orc_expand_crt_basis(Q, P, tmp, in, B, format, result_format);
    towers_copy(out, sq + sp, 0, tmp, sq + sp, sq, sp, B, N);
    towers_copy(out, sq + sp, sp, tmp, sq + sp, 0, sq, B, N);
    free(tmp);
}

/* DCRTPolyImpl::FastExpandCRTBasisPloverQ (dcrtpoly-impl.h:1413-1468, HAVE_INT128 branch).  neg_pl_qhat_inv[sq] =
 * mPlQHatInvModq and qinv_modp[sq*spl] = qInvModp as bfvrns-cryptoparameters.cpp:501-523 builds them; the second step is
 * SwitchCRTBasis Pl -> Ql with the tables of :390-443 (orc_switch_tables / orc_exact_tables restate them).
 * in over Q [B][sq][N] -> out over QlPl [B][sql+spl][N]; no format change. */
void orc_fast_expand_pl_over_q(uint64_t* out, const uint64_t* in, size_t B, uint32_t N, const uint64_t* q, uint32_t sq,
                               const uint64_t* ql, uint32_t sql, const uint64_t* pl, uint32_t spl,
                               const uint64_t* neg_pl_qhat_inv, const uint64_t* qinv_modp) {
    size_t wpl = B * (size_t)spl * N, wql = B * (size_t)sql * N;
    uint64_t* partPl = (uint64_t*)malloc(wpl * 8);
    uint64_t* partQl = (uint64_t*)malloc(wql * 8);
    uint64_t* hi = (uint64_t*)malloc(sizeof(uint64_t) * spl);
    uint64_t* hm = (uint64_t*)malloc(sizeof(uint64_t) * spl * sql);
    double* pinv = (double*)malloc(sizeof(double) * spl);
    uint64_t* ap = (uint64_t*)malloc(sizeof(uint64_t) * (spl + 1) * sql);
    /* :1420-1444 is the loop of ApproxSwitchCRTBasis with these two tables in the places of QHatInvModq / QHatModp */
    orc_approx_switch(partPl, in, B, N, q, sq, pl, spl, neg_pl_qhat_inv, qinv_modp);
    orc_switch_tables(pl, spl, ql, sql, hi, hm);
    orc_exact_tables(pl, spl, ql, sql, pinv, ap);
    orc_switch_crt_basis(partQl, partPl, B, N, pl, spl, ql, sql, hi, hm, pinv, ap);          /* :1448-1450 */
    towers_copy(out, sql + spl, 0, partQl, sql, 0, sql, B, N);                                /* :1457-1465 */
    towers_copy(out, sql + spl, sql, partPl, spl, 0, spl, B, N);
    free(partPl); free(partQl); free(hi); free(hm); free(pinv); free(ap);
}

/* Tables of FastExpandCRTBasisPloverQ, bfvrns-cryptoparameters.cpp:501-523:
 *   neg[i] = q_i - [Pl * (Q/q_i)^-1]_{q_i}   (m_negRlQHatInvModq),   qinv_modp[i*spl+j] = [q_i^-1]_{p_j}   (m_qInvModr) */
void orc_pl_over_q_tables(const uint64_t* q, uint32_t sq, const uint64_t* pl, uint32_t spl, uint64_t* neg, uint64_t* qinv_modp) {
    for (uint32_t i = 0; i < sq; i++) {
        uint64_t qi = q[i], hat = 1 % qi, plm = 1 % qi;
        for (uint32_t k = 0; k < sq; k++)
            if (k != i) hat = mulmod_u128(hat, q[k] % qi, qi);
        for (uint32_t j = 0; j < spl; j++) plm = mulmod_u128(plm, pl[j] % qi, qi);
        neg[i] = qi - mulmod_u128(plm, orc_modinv(hat, qi), qi);
        for (uint32_t j = 0; j < spl; j++) qinv_modp[(size_t)i * spl + j] = orc_modinv(qi % pl[j], pl[j]);
    }
}

/* DCRTPolyImpl::ExpandCRTBasisQlHat (dcrtpoly-impl.h:1514-1534): in over the first sl towers of q [B][sl][N] -> out over
 * q [B][sq][N]: tower i < sl times c[i] (ModMulFastConstEq), zero towers above. */
void orc_expand_ql_hat(uint64_t* out, const uint64_t* in, size_t B, uint32_t N, const uint64_t* q, uint32_t sq, uint32_t sl,
                       const uint64_t* c) {
    for (size_t b = 0; b < B; b++)
        for (uint32_t i = 0; i < sq; i++) {
            uint64_t* o = out + (b * sq + i) * N;
            if (i >= sl) { memset(o, 0, (size_t)N * 8); continue; }
            const uint64_t* x = in + (b * sl + i) * N;
            uint64_t cp = orc_prep_mulconst(c[i], q[i]);
            for (uint32_t n = 0; n < N; n++) o[n] = orc_modmul_const(x[n], c[i], q[i], cp);
        }
}

/* DCRTPolyImpl::ScaleAndRoundPOverQ (dcrtpoly-impl.h:2052-2066): in over q plus one extra tower [B][sq+1][N] -> out over
 * q: ModSubEq of the extra tower (NativeIntegerT::ModSubEq reduces an operand >= q_i first, ubintnat.h:901-912), then
 * Times(pInvModq) (dcrtpoly-impl.h:618-627 -> a full modular multiply). */
void orc_scale_round_p_over_q(uint64_t* out, const uint64_t* in, size_t B, uint32_t N, const uint64_t* q, uint32_t sq,
                              const uint64_t* pinv_modq) {
    for (size_t b = 0; b < B; b++) {
        const uint64_t* last = in + (b * (sq + 1) + sq) * N;
        for (uint32_t i = 0; i < sq; i++) {
            const uint64_t* x = in + (b * (sq + 1) + i) * N;
            uint64_t* o = out + (b * sq + i) * N;
            uint64_t qi = q[i];
            for (uint32_t n = 0; n < N; n++) {
                uint64_t av = x[n] >= qi ? x[n] % qi : x[n], bv = last[n] >= qi ? last[n] % qi : last[n];
                uint64_t d = av >= bv ? av - bv : av + qi - bv;
                o[n] = mulmod_u128(d, pinv_modq[i], qi);
            }
        }
    }
}

void orc_scale_and_round(uint64_t* out, const uint64_t* in, size_t B, uint32_t N, uint32_t sI, const uint64_t* o, uint32_t sO,
                         int out_last, const double* frac, const uint64_t* modo) {
    uint32_t T = sI + sO, in0 = out_last ? 0 : sO, out0 = out_last ? sI : 0;
    u128* mu = (u128*)malloc(sizeof(u128) * sO);
    for (uint32_t j = 0; j < sO; j++) mu[j] = (~(u128)0) / o[j];
    long items = (long)(B * N);
#pragma omp parallel for schedule(static)
    for (long e = 0; e < items; e++) {
        size_t b = (size_t)e / N, ri = (size_t)e % N;
        volatile double nu = 0.5;                                              /* :1908-1913 */
        for (uint32_t i = 0; i < sI; i++) {
            volatile double prod = frac[i] * (double)in[(b * T + in0 + i) * N + ri];
            nu = nu + prod;
        }
        int small = nu <= 18446744073709551616.0;                              /* isConvertableToNativeInt, utilities.h:133-137 */
        u128 alpha = small ? (u128)(uint64_t)nu : (u128)nu;
        for (uint32_t j = 0; j < sO; j++) {
            u128 cur = 0;
            for (uint32_t i = 0; i < sI; i++) cur += (u128)in[(b * T + in0 + i) * N + ri] * modo[(size_t)j * (sI + 1) + i];
            cur += (u128)in[(b * T + out0 + j) * N + ri] * modo[(size_t)j * (sI + 1) + sI];
            uint64_t v = barrett128(cur, o[j], mu[j]);
            uint64_t a = small ? (uint64_t)(alpha % o[j]) : barrett128(alpha, o[j], mu[j]);   /* :1929-1931 / :1952 */
            out[(b * sO + j) * N + ri] = orc_modadd_fast(v, a, o[j]);
        }
    }
    free(mu);
}

/* ---- BFV decryption: DCRTPolyImpl::ScaleAndRound(t, ...) -> NativePoly (dcrtpoly-impl.h:1537-1816) --------------------
 * Tables as bfvrns-cryptoparameters.cpp:447-494 builds them (here with 128-bit integers: [(Q/q_i)^-1]_{q_i} * t < 2^124):
 *   X_i = [(Q/q_i)^-1]_{q_i} * t;  modt = floor(X_i/q_i) mod t;  frac = double(X_i mod q_i) / double(q_i);
 *   and, when qMSB + sizeQMSB >= 52, the same for X_i << (qMSB/2): bmodt, bfrac.
 * branch = 4 * (t power of two) + 2 * (split) + (products reduced modulo t), chosen exactly as :1546-1631 choose. */
int orc_decrypt_branch(const uint64_t* q, uint32_t T, uint64_t t) {
    uint32_t qMSB = orc_msb(q[0]), tMSB = orc_msb(t), sMSB = orc_msb(T);
    int pow2 = (t & (t - 1)) == 0, split = !(qMSB + sMSB < 52);
    uint32_t width = (split ? (qMSB >> 1) : qMSB) + tMSB + sMSB;
    int reduce = pow2 ? !(width < (split ? 62u : 63u)) : !(width < 52u);
    return pow2 * 4 + split * 2 + reduce;
}

void orc_decrypt_tables(const uint64_t* q, uint32_t T, uint64_t t, uint64_t* modt, uint64_t* bmodt, double* frac, double* bfrac) {
    uint32_t h = orc_msb(q[0]) >> 1;
    int split = (orc_decrypt_branch(q, T, t) >> 1) & 1;
    for (uint32_t i = 0; i < T; i++) {
        uint64_t qi = q[i], hat = 1 % qi;
        for (uint32_t k = 0; k < T; k++) if (k != i) hat = mulmod_u128(hat, q[k] % qi, qi);
        u128 X = (u128)orc_modinv(hat, qi) * t, D = X / qi;
        uint64_t R = (uint64_t)(X % qi);
        modt[i] = (uint64_t)(D % t);
        frac[i] = (double)(int64_t)R / (double)(int64_t)qi;
        bmodt[i] = 0; bfrac[i] = 0.0;
        if (split) {
            u128 RB = (u128)R << h;
            uint64_t two_h = (uint64_t)(((u128)1 << h) % t);
            bmodt[i] = (uint64_t)(((u128)(uint64_t)(D % t) * two_h + (RB / qi) % t) % t);
            bfrac[i] = (double)(int64_t)(uint64_t)(RB % qi) / (double)(int64_t)qi;
        }
    }
}

void orc_scale_and_round_native(uint64_t* out, const uint64_t* in, size_t B, uint32_t N, const uint64_t* q, uint32_t T, uint64_t t,
                                const uint64_t* modt, const uint64_t* bmodt, const double* frac, const double* bfrac) {
    const int br = orc_decrypt_branch(q, T, t), pow2 = br >> 2, split = (br >> 1) & 1, reduce = br & 1;
    const uint32_t h = orc_msb(q[0]) >> 1;
    uint64_t pre[256], bpre[256];
    for (uint32_t i = 0; i < T; i++) { pre[i] = orc_prep_mulconst(modt[i], t); bpre[i] = orc_prep_mulconst(bmodt[i], t); }
    const double td = (double)t, tInv = 1. / td;
    long items = (long)(B * N);
#pragma omp parallel for schedule(static)
    for (long e = 0; e < items; e++) {
        size_t b = (size_t)e / N, ri = (size_t)e % N;
        volatile double floatSum = pow2 ? 0.5 : 0.0;          /* volatile: every product and sum is rounded to double */
        uint64_t intSum = 0;
        for (uint32_t i = 0; i < T; i++) {
            uint64_t tmp = in[(b * T + i) * N + ri];
            if (!split) {
                volatile double prod = (double)tmp * frac[i];
                floatSum = floatSum + prod;
                intSum += reduce ? orc_modmul_const(tmp, modt[i], t, pre[i]) : tmp * modt[i];
            } else {
                uint64_t hi = tmp >> h, lo = tmp - (hi << h);
                volatile double p1 = (double)lo * frac[i];
                floatSum = floatSum + p1;
                volatile double p2 = (double)hi * bfrac[i];
                floatSum = floatSum + p2;
                intSum += reduce ? orc_modmul_const(lo, modt[i], t, pre[i]) : lo * modt[i];
                intSum += reduce ? orc_modmul_const(hi, bmodt[i], t, bpre[i]) : hi * bmodt[i];
            }
        }
        if (pow2) {
            intSum += (uint64_t)floatSum;
            out[e] = intSum & (t - 1);
        } else {
            floatSum = floatSum + (double)intSum;
            volatile double qd = floatSum * tInv;
            uint64_t quot = (uint64_t)qd;
            volatile double sub = td * (double)quot;
            floatSum = floatSum - sub;
            volatile double rr = floatSum + 0.5;
            out[e] = (uint64_t)rr;
        }
    }
}

/* ---- hybrid key switching (KeySwitchHYBRID, src/pke/lib/keyswitch/keyswitch-hybrid.cpp) ----------------------------
 * QlP = the ring over the sizeQl current towers of Q followed by the sizeP towers of P.  Digits of alpha towers
 * (the last may be shorter), numPart = ceil(sizeQl / alpha)  (:345-348, without the cap at GetNumberOfQPartitions:
 * the caller passes the alpha of its parameter set, for which the cap never binds at this level).
 * orc_keyswitch_precompute = EvalKeySwitchPrecomputeCore (:330-420): digits [numPart][B][sizeQl+sizeP][N], EVALUATION.
 *   per digit: its towers of c -> COEFFICIENT (:392-393) -> ApproxSwitchCRTBasis to the complement basis (the other
 *   towers of Ql, then P) with the tables of the (truncated) digit basis (:396-402; rns-cryptoparameters.cpp:289-340
 *   builds exactly orc_switch_tables of those moduli) -> EVALUATION (:404) -> complement towers and the digit's own
 *   stored towers put at their places (:406-418).
 * orc_keyswitch_inner = EvalFastKeySwitchCoreExt (:442-480): cTilda0 = sum_j digit_j * b_j, cTilda1 = sum_j digit_j * a_j
 *   tower by tower; the key polys are given over the same sizeQl+sizeP towers ([numPart][sizeQl+sizeP][N]: the caller
 *   picks towers i < sizeQl and sizeQ.. of its key, :463-476).
 * orc_keyswitch_hybrid = KeySwitchCore / EvalFastKeySwitchCore (:324-328, 422-440): the two ApproxModDown to Ql. */
static orc_ctx* sub_ctx(const orc_ctx* c, const uint32_t* idx, uint32_t cnt) {
    uint64_t q[256], psi[256];
    for (uint32_t k = 0; k < cnt; k++) { q[k] = c->q[idx[k]]; psi[k] = c->psi[idx[k]]; }
    return orc_ctx_create(c->N, cnt, q, psi);
}

uint32_t orc_keyswitch_digits(uint32_t sizeQl, uint32_t alpha) { return (sizeQl + alpha - 1) / alpha; }

void orc_keyswitch_precompute(const orc_ctx* QlP, uint32_t sizeQl, uint32_t alpha, uint64_t* digits, const uint64_t* c, size_t B) {
    const uint32_t N = QlP->N, T = QlP->T, np = orc_keyswitch_digits(sizeQl, alpha);
    for (uint32_t part = 0; part < np; part++) {
        const uint32_t s = alpha * part, e = s + alpha < sizeQl ? s + alpha : sizeQl, sp = e - s, sc = T - sp;
        uint32_t pidx[256], cidx[256];
        for (uint32_t k = 0; k < sp; k++) pidx[k] = s + k;
        for (uint32_t k = 0, i = 0; i < T; i++) if (i < s || i >= e) cidx[k++] = i;
        orc_ctx *part_ctx = sub_ctx(QlP, pidx, sp), *compl_ctx = sub_ctx(QlP, cidx, sc);
        uint64_t* coef = (uint64_t*)malloc(B * sp * (size_t)N * 8);
        uint64_t* comp = (uint64_t*)malloc(B * sc * (size_t)N * 8);
        uint64_t* hi = (uint64_t*)malloc(8 * sp);
        uint64_t* hm = (uint64_t*)malloc(8 * (size_t)sp * sc);
        towers_copy(coef, sp, 0, c, sizeQl, s, sp, B, N);
        orc_dcrt_ntt(part_ctx, 1, coef, B, 0);
        orc_switch_tables(part_ctx->q, sp, compl_ctx->q, sc, hi, hm);
        orc_approx_switch(comp, coef, B, N, part_ctx->q, sp, compl_ctx->q, sc, hi, hm);
        orc_dcrt_ntt(compl_ctx, 0, comp, B, 0);
        uint64_t* d = digits + (size_t)part * B * T * N;
        towers_copy(d, T, 0, comp, sc, 0, s, B, N);                       /* towers below the digit */
        towers_copy(d, T, s, c, sizeQl, s, sp, B, N);                     /* the digit's own towers, as stored */
        towers_copy(d, T, e, comp, sc, s, T - e, B, N);                   /* the rest of Ql, then P */
        free(coef); free(comp); free(hi); free(hm);
        orc_ctx_destroy(part_ctx); orc_ctx_destroy(compl_ctx);
    }
}

void orc_keyswitch_inner(const orc_ctx* QlP, uint32_t np, uint64_t* c0, uint64_t* c1, const uint64_t* digits,
                         const uint64_t* evk_b, const uint64_t* evk_a, size_t B) {
    const uint32_t N = QlP->N, T = QlP->T;
    long units = (long)(B * T);
#pragma omp parallel for schedule(static)
    for (long u = 0; u < units; u++) {
        const uint32_t t = (uint32_t)(u % T);
        const uint64_t q = QlP->q[t], mu = orc_compute_mu(q);
        uint64_t *o0 = c0 + (size_t)u * N, *o1 = c1 + (size_t)u * N;
        for (uint32_t n = 0; n < N; n++) o0[n] = o1[n] = 0;               /* DCRTPoly(paramsQlP, EVALUATION, true) */
        for (uint32_t j = 0; j < np; j++) {
            const uint64_t* d = digits + ((size_t)j * B * T + (size_t)u) * N;
            const uint64_t* kb = evk_b + ((size_t)j * T + t) * N;
            const uint64_t* ka = evk_a + ((size_t)j * T + t) * N;
            for (uint32_t n = 0; n < N; n++) {
                o0[n] = orc_modadd_fast(o0[n], orc_modmul_barrett(d[n], kb[n], q, mu), q);
                o1[n] = orc_modadd_fast(o1[n], orc_modmul_barrett(d[n], ka[n], q, mu), q);
            }
        }
    }
}

void orc_keyswitch_hybrid(const orc_ctx* QlP, uint32_t sizeQl, uint32_t alpha, uint64_t* ct0, uint64_t* ct1, const uint64_t* c,
                          const uint64_t* evk_b, const uint64_t* evk_a, size_t B, uint64_t t) {
    const uint32_t N = QlP->N, T = QlP->T, np = orc_keyswitch_digits(sizeQl, alpha), sizeP = T - sizeQl;
    uint64_t* digits = (uint64_t*)malloc((size_t)np * B * T * N * 8);
    uint64_t* c0 = (uint64_t*)malloc(B * T * (size_t)N * 8);
    uint64_t* c1 = (uint64_t*)malloc(B * T * (size_t)N * 8);
    orc_keyswitch_precompute(QlP, sizeQl, alpha, digits, c, B);
    orc_keyswitch_inner(QlP, np, c0, c1, digits, evk_b, evk_a, B);
    uint32_t qi[256], pi[256];
    for (uint32_t k = 0; k < sizeQl; k++) qi[k] = k;
    for (uint32_t k = 0; k < sizeP; k++) pi[k] = sizeQl + k;
    orc_ctx *Q = sub_ctx(QlP, qi, sizeQl), *P = sub_ctx(QlP, pi, sizeP);
    orc_approx_mod_down(Q, P, ct0, c0, B, t);
    orc_approx_mod_down(Q, P, ct1, c1, B, t);
    orc_ctx_destroy(Q); orc_ctx_destroy(P);
    free(digits); free(c0); free(c1);
}

int orc_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* dcrtpoly.h:142-148,164-200 / dcrtpoly-impl.h:398-416: per tower i, vector op with modulus q_i.
 * The reference runs `omp parallel for` over towers of ONE poly; here the loop is over all
 * (batch, tower) units so that nthreads > T can be used for the all-cores CPU figure. */
void orc_dcrt_elementwise(const orc_ctx* c, int op, uint64_t* dst, const uint64_t* a,
                          const uint64_t* b, size_t B, int nthreads) {
    long units = (long)(B * c->T);
    size_t N = c->N;
    if (nthreads <= 0) nthreads = orc_max_threads();
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (long u = 0; u < units; u++) {
        uint64_t q = c->q[u % c->T];
        size_t off = (size_t)u * N;
        if (op == 0) orc_vec_modadd(dst + off, a + off, b + off, N, q);
        else if (op == 1) orc_vec_modsub(dst + off, a + off, b + off, N, q);
        else orc_vec_modmul(dst + off, a + off, b + off, N, q);
    }
}

/* dcrtpoly-impl.h:2518-2524 -> poly-impl.h:412-432 -> transformnat-impl.h:575-602 / 637-666 */
void orc_dcrt_ntt(const orc_ctx* c, int inverse, uint64_t* x, size_t B, int nthreads) {
    long units = (long)(B * c->T);
    size_t N = c->N;
    if (nthreads <= 0) nthreads = orc_max_threads();
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (long u = 0; u < units; u++) {
        uint32_t t = (uint32_t)(u % c->T);
        uint64_t* p = x + (size_t)u * N;
        if (c->psi[t] == 0 || c->psi[t] == 1) continue;   /* transformnat-impl.h:580-582 */
        if (!inverse) orc_ntt_fwd(p, c->N, c->q[t], c->W + (size_t)t * N, c->Wp + (size_t)t * N);
        else orc_ntt_inv(p, c->N, c->q[t], c->WI + (size_t)t * N, c->WIp + (size_t)t * N,
                         c->ninv[t], c->ninvp[t]);
    }
}

/* ---- seeded generator (ours; the reference's Blake2 PRNG is not reproducible) ----
 * splitmix64 over a counter; rejection-sample the low bitlen(q-1) bits until < q. */
static inline uint64_t splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
uint64_t orc_uniform_at(uint64_t seed, uint64_t i, uint64_t q) {
    uint32_t bits = orc_msb(q - 1);
    uint64_t mask = bits >= 64 ? ~0ull : (((uint64_t)1 << bits) - 1);
    uint64_t s = splitmix64(seed ^ (i * 0xD1342543DE82EF95ull));
    for (;;) {
        s = splitmix64(s);
        uint64_t v = s & mask;
        if (v < q) return v;
    }
}
void orc_fill_uniform(uint64_t* x, size_t n, uint64_t q, uint64_t seed, uint64_t first_index) {
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; i++) x[i] = orc_uniform_at(seed, first_index + (uint64_t)i, q);
}

/* ---- BEHZ: the BFV multiplication technique of Bajard-Eynard-Hasan-Zucca ----------------------------------
 * DCRTPolyImpl::FastBaseConvqToBskMontgomery (dcrtpoly-impl.h:2069-2209), FastRNSFloorq (:2212-2271),
 * FastBaseConvSK (:2309-2420) and the decryption ScaleAndRound(moduliQ, t, tgamma, ...) (:2005-2050), with the
 * tables of CryptoParametersBFVRNS::PrecomputeCRTTables (src/pke/lib/scheme/bfvrns/bfvrns-cryptoparameters.cpp:
 * 662-885) restated in word arithmetic.  The ring is Q followed by Bsk = B u {msk}: nq towers of Q, nb = |B| + 1
 * towers of Bsk, the last one msk.  mtilde = 2^16 and gamma = 2^26 are the scheme's constants
 * (bfvrns-cryptoparameters.h: m_mtilde, m_gamma). */
#define BEHZ_TABLES 17
struct orc_behz {
    orc_ctx *Q, *Bsk;
    uint32_t N, nq, nb;
    uint64_t t;
    uint64_t* tab[BEHZ_TABLES];
    size_t len[BEHZ_TABLES];
};

static uint64_t prod_except(const uint64_t* f, uint32_t n, uint32_t skip, uint64_t m) {
    uint64_t r = 1 % m;
    for (uint32_t k = 0; k < n; k++) if (k != skip) r = mulmod_u128(r, f[k] % m, m);
    return r;
}
static uint64_t* behz_new(orc_behz* z, int which, size_t n) {
    z->tab[which] = (uint64_t*)calloc(n ? n : 1, 8);
    z->len[which] = n;
    return z->tab[which];
}

orc_behz* orc_behz_create(const orc_ctx* QBsk, uint32_t nq, uint64_t t) {
    if (!QBsk || nq == 0 || QBsk->T < nq + 2 || t < 2) return NULL;
    orc_behz* z = (orc_behz*)calloc(1, sizeof(orc_behz));
    const uint32_t nb = QBsk->T - nq, nB = nb - 1, NONE = 0xffffffffu;
    const uint64_t* q = QBsk->q;
    const uint64_t* bsk = QBsk->q + nq;
    const uint64_t msk = bsk[nB], mtilde = (uint64_t)1 << 16, gamma = (uint64_t)1 << 26;
    z->N = QBsk->N; z->nq = nq; z->nb = nb; z->t = t;
    z->Q = orc_ctx_create(QBsk->N, nq, q, QBsk->psi);
    z->Bsk = orc_ctx_create(QBsk->N, nb, bsk, QBsk->psi + nq);
    uint64_t* hatinv = (uint64_t*)malloc(8 * nq);                       /* [(Q/q_i)^-1]_{q_i} */
    for (uint32_t i = 0; i < nq; i++) hatinv[i] = orc_modinv(prod_except(q, nq, i, q[i]), q[i]);
    uint64_t* T0 = behz_new(z, 0, nq);                                  /* [mtilde*(Q/q_i)^-1]_{q_i}           :763-776 */
    uint64_t* T1 = behz_new(z, 1, (size_t)nq * nb);                     /* [Q/q_i]_{bsk_j}                     :737-750 */
    uint64_t* T2 = behz_new(z, 2, nq);                                  /* [Q/q_i]_{mtilde}                    :749 */
    uint64_t* T6 = behz_new(z, 6, nq);                                  /* [t*(Q/q_i)^-1]_{q_i}                :724-735 */
    uint64_t* T7 = behz_new(z, 7, (size_t)nq * nb);                     /* [q_i^-1]_{bsk_j}                    :752-758 */
    for (uint32_t i = 0; i < nq; i++) {
        T0[i] = mulmod_u128(hatinv[i], mtilde % q[i], q[i]);
        T6[i] = mulmod_u128(hatinv[i], t % q[i], q[i]);
        T2[i] = prod_except(q, nq, i, mtilde);
        for (uint32_t j = 0; j < nb; j++) {
            T1[(size_t)i * nb + j] = prod_except(q, nq, i, bsk[j]);
            T7[(size_t)i * nb + j] = orc_modinv(q[i] % bsk[j], bsk[j]);
        }
    }
    uint64_t* T3 = behz_new(z, 3, nb);                                  /* [Q]_{bsk_j}                         :783-791 */
    uint64_t* T5 = behz_new(z, 5, nb);                                  /* [mtilde^-1]_{bsk_j}                 :793-802 */
    uint64_t* T8 = behz_new(z, 8, nb);                                  /* [t*Q^-1]_{bsk_j}                    :804-813 */
    for (uint32_t j = 0; j < nb; j++) {
        T3[j] = prod_except(q, nq, NONE, bsk[j]);
        T5[j] = orc_modinv(mtilde % bsk[j], bsk[j]);
        T8[j] = mulmod_u128(orc_modinv(T3[j], bsk[j]), t % bsk[j], bsk[j]);
    }
    uint64_t* T4 = behz_new(z, 4, 1);                                   /* [-Q^-1]_{mtilde}                    :778-781 */
    T4[0] = mulmod_u128(mtilde - 1, orc_modinv(prod_except(q, nq, NONE, mtilde), mtilde), mtilde);
    uint64_t* T9 = behz_new(z, 9, nB);                                  /* [(B/b_i)^-1]_{b_i}                  :815-827 */
    uint64_t* T10 = behz_new(z, 10, (size_t)nB * nq);                   /* [B/b_i]_{q_j}                       :829-839 */
    uint64_t* T11 = behz_new(z, 11, nB);                                /* [B/b_i]_{msk}                       :841-846 */
    for (uint32_t i = 0; i < nB; i++) {
        T9[i] = orc_modinv(prod_except(bsk, nB, i, bsk[i]), bsk[i]);
        T11[i] = prod_except(bsk, nB, i, msk);
        for (uint32_t j = 0; j < nq; j++) T10[(size_t)i * nq + j] = prod_except(bsk, nB, i, q[j]);
    }
    uint64_t* T12 = behz_new(z, 12, 1);                                 /* [B^-1]_{msk}                        :848-850 */
    T12[0] = orc_modinv(prod_except(bsk, nB, NONE, msk), msk);
    uint64_t* T13 = behz_new(z, 13, nq);                                /* [B]_{q_i}                           :852-858 */
    for (uint32_t i = 0; i < nq; i++) T13[i] = prod_except(bsk, nB, NONE, q[i]);
    const uint64_t tgamma = t * gamma;                                  /* :862-864 */
    uint64_t* T14 = behz_new(z, 14, 1);
    T14[0] = tgamma;
    uint64_t* T15 = behz_new(z, 15, nq);                                /* [-q_i^-1]_{t*gamma}                 :866-876 */
    uint64_t* T16 = behz_new(z, 16, nq);                                /* [t*gamma*(Q/q_i)^-1]_{q_i}          :878-891 */
    for (uint32_t i = 0; i < nq; i++) {
        T15[i] = mulmod_u128(tgamma - 1, orc_modinv(q[i] % tgamma, tgamma), tgamma);
        T16[i] = mulmod_u128(mulmod_u128(hatinv[i], gamma % q[i], q[i]), t % q[i], q[i]);
    }
    free(hatinv);
    return z;
}

void orc_behz_destroy(orc_behz* z) {
    if (!z) return;
    for (int k = 0; k < BEHZ_TABLES; k++) free(z->tab[k]);
    orc_ctx_destroy(z->Q); orc_ctx_destroy(z->Bsk);
    free(z);
}

/* copies table `which` (numbering above) to out when out is not NULL; returns its length in words */
size_t orc_behz_table(const orc_behz* z, int which, uint64_t* out) {
    if (!z || which < 0 || which >= BEHZ_TABLES) return 0;
    if (out) memcpy(out, z->tab[which], 8 * z->len[which]);
    return z->len[which];
}

/* FastBaseConvqToBskMontgomery (dcrtpoly-impl.h:2069-2209): in over Q [B][nq][N] in `format` (0 EVALUATION,
 * 1 COEFFICIENT) -> out over QBsk [B][nq+nb][N], EVALUATION */
void orc_behz_q_to_bsk_montgomery(const orc_behz* z, uint64_t* out, const uint64_t* in, size_t B, int format) {
    const uint32_t N = z->N, nq = z->nq, nb = z->nb, T = nq + nb;
    const uint64_t* q = z->Q->q;
    const uint64_t* bsk = z->Bsk->q;
    const uint64_t mtilde = (uint64_t)1 << 16, mtilde_half = mtilde >> 1, mtilde_minus_1 = mtilde - 1;
    size_t wq = B * nq * (size_t)N, wb = B * nb * (size_t)N;
    uint64_t* coef = (uint64_t*)malloc(wq * 8);
    uint64_t* part = (uint64_t*)malloc(wb * 8);
    memcpy(coef, in, wq * 8);
    if (format == 0) orc_dcrt_ntt(z->Q, 1, coef, B, 0);                  /* :2087-2090 */
    /* step 0, mod Bsk (:2103-2127): the twist by [mtilde*(Q/q_i)^-1]_{q_i} and the 128-bit sums are
     * ApproxSwitchCRTBasis's own statements with these two tables */
    orc_approx_switch(part, coef, B, N, q, nq, bsk, nb, z->tab[0], z->tab[1]);
    uint64_t* pre0 = (uint64_t*)malloc(8 * nq);
    uint64_t *preQ = (uint64_t*)malloc(8 * nb), *preI = (uint64_t*)malloc(8 * nb);
    for (uint32_t i = 0; i < nq; i++) pre0[i] = orc_prep_mulconst(z->tab[0][i], q[i]);
    for (uint32_t j = 0; j < nb; j++) {
        preQ[j] = orc_prep_mulconst(z->tab[3][j], bsk[j]);
        preI[j] = orc_prep_mulconst(z->tab[5][j], bsk[j]);
    }
    long items = (long)(B * N);
#pragma omp parallel for schedule(static)
    for (long e = 0; e < items; e++) {
        size_t b = (size_t)e / N, k = (size_t)e % N;
        uint64_t r = 0;                                                  /* mod mtilde = 2^16 (:2148-2160) */
        for (uint32_t i = 0; i < nq; i++)
            r += orc_modmul_const(coef[(b * nq + i) * N + k], z->tab[0][i], q[i], pre0[i]) * z->tab[2][i];
        r &= mtilde_minus_1;
        r *= z->tab[4][0];                                               /* step 1 (:2168-2171) */
        r &= mtilde_minus_1;
        for (uint32_t j = 0; j < nb; j++) {                              /* :2173-2190 */
            uint64_t rm = r;
            if (r >= mtilde_half) rm += bsk[j] - mtilde;
            rm = orc_modmul_const(rm, z->tab[3][j], bsk[j], preQ[j]);
            uint64_t* c = part + (b * nb + j) * N + k;
            rm = orc_modadd_fast(rm, *c, bsk[j]);
            *c = orc_modmul_const(rm, z->tab[5][j], bsk[j], preI[j]);
        }
    }
    if (format == 0) {                                                   /* :2194-2197 the stored EVALUATION towers */
        towers_copy(out, T, 0, in, nq, 0, nq, B, N);
    } else {                                                             /* :2198-2202 */
        orc_dcrt_ntt(z->Q, 0, coef, B, 0);
        towers_copy(out, T, 0, coef, nq, 0, nq, B, N);
    }
    orc_dcrt_ntt(z->Bsk, 0, part, B, 0);                                 /* :2204-2206 */
    towers_copy(out, T, nq, part, nb, 0, nb, B, N);
    free(coef); free(part); free(pre0); free(preQ); free(preI);
}

/* FastRNSFloorq (dcrtpoly-impl.h:2212-2271), in place on x over QBsk [B][nq+nb][N]: the towers of Q are left
 * multiplied by [t*(Q/q_i)^-1]_{q_i} (:2233-2243), the towers of Bsk hold the result */
void orc_behz_fast_rns_floor_q(const orc_behz* z, uint64_t* x, size_t B) {
    const uint32_t N = z->N, nq = z->nq, nb = z->nb, T = nq + nb;
    const uint64_t* q = z->Q->q;
    const uint64_t* bsk = z->Bsk->q;
    size_t wq = B * nq * (size_t)N, wb = B * nb * (size_t)N;
    uint64_t* xq = (uint64_t*)calloc(wq, 8);
    uint64_t* conv = (uint64_t*)malloc(wb * 8);
    towers_copy(xq, nq, 0, x, T, 0, nq, B, N);
    /* :2245-2256 = ApproxSwitchCRTBasis's statements with [t*(Q/q_i)^-1]_{q_i} and [q_i^-1]_{bsk_j} */
    orc_approx_switch(conv, xq, B, N, q, nq, bsk, nb, z->tab[6], z->tab[7]);
    for (size_t b = 0; b < B; b++) {
        for (uint32_t i = 0; i < nq; i++) {
            uint64_t pre = orc_prep_mulconst(z->tab[6][i], q[i]);
            uint64_t* v = x + (b * T + i) * N;
            for (uint32_t k = 0; k < N; k++) v[k] = orc_modmul_const(v[k], z->tab[6][i], q[i], pre);
        }
        for (uint32_t j = 0; j < nb; j++) {                              /* :2260-2269 */
            uint64_t pre = orc_prep_mulconst(z->tab[8][j], bsk[j]);
            uint64_t* v = x + (b * T + nq + j) * N;
            const uint64_t* c = conv + (b * nb + j) * N;
            for (uint32_t k = 0; k < N; k++)
                v[k] = orc_modsub_fast(orc_modmul_const(v[k], z->tab[8][j], bsk[j], pre), c[k], bsk[j]);
        }
    }
    free(xq); free(conv);
}

/* FastBaseConvSK (dcrtpoly-impl.h:2309-2420): in over QBsk [B][nq+nb][N] (only the towers of Bsk are read) ->
 * out over Q [B][nq][N] */
void orc_behz_fast_base_conv_sk(const orc_behz* z, uint64_t* out, const uint64_t* in, size_t B) {
    const uint32_t N = z->N, nq = z->nq, nb = z->nb, nB = nb - 1, T = nq + nb;
    const uint64_t* q = z->Q->q;
    const uint64_t* bsk = z->Bsk->q;
    const uint64_t msk = bsk[nB], mskDivTwo = msk / 2;
    size_t wB = B * nB * (size_t)N;
    uint64_t* xb = (uint64_t*)calloc(wB, 8);
    uint64_t* alpha = (uint64_t*)malloc(B * (size_t)N * 8);
    towers_copy(xb, nB, 0, in, T, nq, nB, B, N);
    /* FastBaseconv(x, B, q) (:2327-2349) and FastBaseConv(x, B, msk) (:2351-2364): ApproxSwitchCRTBasis's statements
     * with [(B/b_i)^-1]_{b_i} and [B/b_i]_{q_j} / [B/b_i]_{msk} */
    orc_approx_switch(out, xb, B, N, bsk, nB, q, nq, z->tab[9], z->tab[10]);
    orc_approx_switch(alpha, xb, B, N, bsk, nB, &msk, 1, z->tab[9], z->tab[11]);
    const uint64_t binv = z->tab[12][0], binvp = orc_prep_mulconst(binv, msk);
    for (size_t b = 0; b < B; b++) {
        const uint64_t* xsk = in + (b * T + nq + nB) * N;
        uint64_t* a = alpha + b * N;
        for (uint32_t k = 0; k < N; k++)                                 /* :2366-2371 */
            a[k] = orc_modmul_const(orc_modsub_fast(a[k], xsk[k], msk), binv, msk, binvp);
        for (uint32_t i = 0; i < nq; i++) {                              /* :2373-2388 */
            uint64_t pre = orc_prep_mulconst(z->tab[13][i], q[i]);
            uint64_t* o = out + (b * nq + i) * N;
            for (uint32_t k = 0; k < N; k++) {
                uint64_t ab = a[k];
                if (ab > mskDivTwo) ab = orc_modsub_fast(ab, msk, q[i]);
                ab = orc_modmul_const(ab, z->tab[13][i], q[i], pre);
                o[k] = orc_modsub_fast(o[k], ab, q[i]);
            }
        }
    }
    free(xb); free(alpha);
}

/* ScaleAndRound(moduliQ, t, tgamma, tgammaQHatModq, .., negInvqModtgamma, ..) -> NativePoly
 * (dcrtpoly-impl.h:2005-2050): in over Q [B][nq][N], COEFFICIENT -> out [B][N] */
void orc_behz_scale_and_round(const orc_behz* z, uint64_t* out, const uint64_t* in, size_t B) {
    const uint32_t N = z->N, nq = z->nq;
    const uint64_t* q = z->Q->q;
    const uint64_t tgamma = z->tab[14][0], gammaMinus1 = (1 << 26) - 1;
    uint64_t *p1 = (uint64_t*)malloc(8 * nq), *p2 = (uint64_t*)malloc(8 * nq);
    for (uint32_t i = 0; i < nq; i++) {
        p1[i] = orc_prep_mulconst(z->tab[16][i], q[i]);
        p2[i] = orc_prep_mulconst(z->tab[15][i], tgamma);
    }
    long items = (long)(B * N);
#pragma omp parallel for schedule(static)
    for (long e = 0; e < items; e++) {
        size_t b = (size_t)e / N, k = (size_t)e % N;
        uint64_t s = 0;
        for (uint32_t i = 0; i < nq; i++) {
            uint64_t tmp = orc_modmul_const(in[(b * nq + i) * N + k], z->tab[16][i], q[i], p1[i]);
            tmp = orc_modmul_const(tmp, z->tab[15][i], tgamma, p2[i]);
            s = orc_modadd_fast(s, tmp, tgamma);
        }
        s += s & gammaMinus1;
        out[e] = s >> 26;
    }
    free(p1); free(p2);
}

/* DCRTPolyImpl::CRTDecompose(baseBits = 0) (dcrtpoly-impl.h:266-288; the digits of KeySwitchBV, keyswitch-bv.cpp:308-312):
 * in [B][T][N] in `format` (0 EVALUATION, 1 COEFFICIENT) -> out [T][B][T][N], EVALUATION: digit i keeps the stored tower i
 * and gets, in every other tower k, the COEFFICIENT tower i moved to q_k with NativeVector::SwitchModulus and transformed */
void orc_crt_decompose(const orc_ctx* c, uint64_t* out, const uint64_t* in, size_t B, int format) {
    const uint32_t N = c->N, T = c->T;
    size_t w = B * T * (size_t)N;
    uint64_t* coef = (uint64_t*)malloc(w * 8);
    uint64_t* eval = (uint64_t*)malloc(w * 8);
    memcpy(coef, in, w * 8);
    memcpy(eval, in, w * 8);
    if (format == 0) orc_dcrt_ntt(c, 1, coef, B, 0);                     /* :267-270 */
    else orc_dcrt_ntt(c, 0, eval, B, 0);
    for (uint32_t i = 0; i < T; i++) {
        uint64_t* digit = out + (size_t)i * w;
        for (size_t b = 0; b < B; b++)
            for (uint32_t k = 0; k < T; k++) {
                uint64_t* o = digit + (b * T + k) * N;
                if (k == i) {                                            /* :274 result(size, *eval) */
                    memcpy(o, eval + (b * T + i) * N, (size_t)N * 8);
                } else {                                                 /* :280-283 */
                    const uint64_t* x = coef + (b * T + i) * N;
                    for (uint32_t n = 0; n < N; n++) o[n] = switch_modulus1(x[n], c->q[i], c->q[k]);
                    if (!(c->psi[k] == 0 || c->psi[k] == 1))
                        orc_ntt_fwd(o, N, c->q[k], c->W + (size_t)k * N, c->Wp + (size_t)k * N);
                }
            }
    }
    free(coef); free(eval);
}
