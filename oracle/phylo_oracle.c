/*
 * phylo_oracle.c — CPU ORACLE (test infrastructure, NOT product code).
 *
 * A plain-C restatement of the reference's (sriram98v/phylo-rs, crate `phylo` 6.0.0)
 * Felsenstein-pruning log-likelihood and marginal-ASR path, in f64, written from the
 * reference's algorithm description.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library; the product path
 * (libphylo_b200.so) never links, imports or calls it.
 *
 * PARITY STATUS: "parity unpinned" by reference-blessed absolute values — the reference's own
 * tests hold NO numeric lnL / posterior golden values (src/tree/likelihood/integration_test.rs
 * only asserts structural invariants) and the reference cannot be compiled in this image (no
 * cargo/rustc).  The oracle IS pinned against every invariant the reference tests assert
 * (tests/test_oracle_reference_invariants.py), against brute-force enumeration over internal
 * states, analytic JC69, scipy.linalg.expm, and the KAT table in SURVEY.md §8c.
 *
 * Third-party arithmetic restated: nalgebra 0.33 (Cargo.toml:56, no lockfile)
 *   - symmetric_eigen (rate_matrix.rs:100): restated as cyclic Jacobi; P(t) is invariant to the
 *     eigenvector basis up to round-off.
 *   - small DMatrix products (rate_matrix.rs:125) and DMatrix*DVector (likelihood.rs:121,328):
 *     column-wise gemv/axpy accumulation order k = 0..n-1, separate multiply and add (Rust never
 *     contracts to FMA) — compile this file with -ffp-contract=off.
 *
 * Every function cites the reference file:line it follows (paths relative to /root/reference).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

/* Number of states: 4 = Nucleotide (src/alphabet.rs:26), the default build; 20 = AminoAcid (src/alphabet.rs:72),
 * built as libphylo_oracle_aa.so with -DNS=20.  Everything below the alphabet section is written over NS, as the
 * reference is generic over A: Alphabet (likelihood.rs:156-166, gtr.rs:36-40). */
#ifndef NS
#define NS 4
#endif

/* status codes mirror AsrError, src/error.rs:106-125 */
enum {
    ORC_OK = 0,
    ORC_ERR_MISSING_BRANCH_LENGTH = 1,
    ORC_ERR_ALPHABET_MISMATCH = 2,
    ORC_ERR_INVALID_ALIGNMENT = 3,
    ORC_ERR_EIGENDECOMP = 4,
    ORC_ERR_NUMERICAL_INSTABILITY = 5,
    ORC_ERR_INVALID_MODEL_PARAMETER = 6
};

/* ------------------------------------------------------------------------------------------
 * models/gamma.rs
 * ---------------------------------------------------------------------------------------- */

/* src/models/gamma.rs:100-126  Lanczos g=7 */
static double orc_ln_gamma(double x)
{
    static const double COEFFS[9] = {
        0.9999999999998099, 676.5203681218851, -1259.1392167224028,
        771.3234287776531, -176.6150291621406, 12.507343278686905,
        -0.13857109526572012, 9.984369578019572e-6, 1.5056327351493116e-7};
    const double G = 7.0;
    if (x < 0.5)
        return log(M_PI / sin(M_PI * x)) - orc_ln_gamma(1.0 - x);
    x = x - 1.0;
    double a = COEFFS[0];
    double t = x + G + 0.5;
    for (int i = 1; i < 9; ++i)
        a += COEFFS[i] / (x + (double)i);
    return 0.5 * log(2.0 * M_PI) + (x + 0.5) * log(t) - t + log(a);
}

static double clamp01(double v) { return v < 0.0 ? 0.0 : (v > 1.0 ? 1.0 : v); }

/* src/models/gamma.rs:129-176 regularised lower incomplete gamma P(shape, x) */
static double orc_incomplete_gamma(double x, double shape)
{
    if (x <= 0.0)
        return 0.0;
    double lg = orc_ln_gamma(shape);
    if (x < shape + 1.0) {
        double term = 1.0 / shape, sum = term, n = shape;
        for (int it = 0; it < 500; ++it) {
            n += 1.0;
            term *= x / n;
            sum += term;
            if (fabs(term) < fabs(sum) * 1e-15)
                break;
        }
        return clamp01(sum * exp(-x + shape * log(x) - lg));
    } else {
        double b = x + 1.0 - shape, c = 1e300, d = 1.0 / b, h = d;
        for (int i = 1; i < 500; ++i) {
            double an = -((double)i) * ((double)i - shape);
            b += 2.0;
            d = an * d + b;
            if (fabs(d) < 1e-300)
                d = 1e-300;
            c = b + an / c;
            if (fabs(c) < 1e-300)
                c = 1e-300;
            d = 1.0 / d;
            double del = d * c;
            h *= del;
            if (fabs(del - 1.0) < 1e-15)
                break;
        }
        double q = exp(-x + shape * log(x) - lg) * h;
        return clamp01(1.0 - q);
    }
}

/* src/models/gamma.rs:180-229 */
static double orc_inverse_normal_cdf(double p)
{
    static const double a[6] = {-3.969683028665376e+01, 2.209460984245205e+02,
                                -2.759285104469687e+02, 1.38357751867269e2,
                                -3.066479806614716e+01, 2.506628277459239e+00};
    static const double b[5] = {-5.447609879822406e+01, 1.615858368580409e+02,
                                -1.556989798598866e+02, 6.680131188771972e+01,
                                -1.328068155288572e+01};
    static const double c[6] = {-7.784894002430293e-03, -3.223964580411365e-01,
                                -2.400758277161838e+00, -2.549732539343734e+00,
                                4.374664141464968e+00,  2.938163982698783e+00};
    static const double d[4] = {7.784695709041462e-03, 3.224671290700398e-01,
                                2.445134137142996e+00, 3.754408661907416e+00};
    const double p_low = 0.02425, p_high = 1.0 - 0.02425;
    if (p < p_low) {
        double q = sqrt(-2.0 * log(p));
        return (((((c[0] * q + c[1]) * q + c[2]) * q + c[3]) * q + c[4]) * q + c[5]) /
               ((((d[0] * q + d[1]) * q + d[2]) * q + d[3]) * q + 1.0);
    } else if (p <= p_high) {
        double q = p - 0.5, r = q * q;
        return (((((a[0] * r + a[1]) * r + a[2]) * r + a[3]) * r + a[4]) * r + a[5]) * q /
               (((((b[0] * r + b[1]) * r + b[2]) * r + b[3]) * r + b[4]) * r + 1.0);
    } else {
        double q = sqrt(-2.0 * log(1.0 - p));
        return -(((((c[0] * q + c[1]) * q + c[2]) * q + c[3]) * q + c[4]) * q + c[5]) /
               ((((d[0] * q + d[1]) * q + d[2]) * q + d[3]) * q + 1.0);
    }
}

/* src/models/gamma.rs:59-97 */
static int orc_point_gamma(double prob, double shape, double rate, double *out)
{
    if (!(prob >= 0.0 && prob < 1.0))
        return ORC_ERR_INVALID_MODEL_PARAMETER;
    double lgs = orc_ln_gamma(shape);
    double base = 1.0 - 1.0 / (9.0 * shape) + orc_inverse_normal_cdf(prob) * sqrt(1.0 / (9.0 * shape));
    double ch = shape * (base * base * base);
    double x = ch > 0.0 ? ch / rate : 1e-10;
    for (int it = 0; it < 100; ++it) {
        double f = orc_incomplete_gamma(x * rate, shape) - prob;
        double ln_df = (shape - 1.0) * log(x * rate) - x * rate - lgs + log(rate);
        double df = exp(ln_df);
        if (!isfinite(df) || fabs(df) < 1e-300)
            break;
        double step = f / df;
        double x_new = x - step;
        if (!isfinite(x_new) || x_new <= 0.0)
            break;
        double xm = x > 1e-12 ? x : 1e-12;
        if (fabs(x_new - x) < 1e-12 * xm) {
            x = x_new;
            break;
        }
        x = x_new;
    }
    *out = x;
    return ORC_OK;
}

/* src/models/gamma.rs:15-55 */
int orc_discrete_gamma(double alpha, int k, double *rates)
{
    if (alpha <= 0.0 || k <= 0)
        return ORC_ERR_INVALID_MODEL_PARAMETER;
    if (k == 1) {
        rates[0] = 1.0;
        return ORC_OK;
    }
    double kf = (double)k;
    double *cdf = (double *)malloc(sizeof(double) * (size_t)(k + 1));
    cdf[0] = 0.0;
    for (int i = 1; i < k; ++i) {
        double b;
        int rc = orc_point_gamma((double)i / kf, alpha, alpha, &b);
        if (rc) {
            free(cdf);
            return rc;
        }
        cdf[i] = orc_incomplete_gamma(b * alpha, alpha + 1.0);
    }
    cdf[k] = 1.0;
    for (int i = 0; i < k; ++i)
        rates[i] = kf * (cdf[i + 1] - cdf[i]);
    free(cdf);
    return ORC_OK;
}

/* exported helpers so tests can pin the special functions against scipy */
double orc_test_ln_gamma(double x) { return orc_ln_gamma(x); }
double orc_test_incomplete_gamma(double x, double shape) { return orc_incomplete_gamma(x, shape); }
double orc_test_inverse_normal_cdf(double p) { return orc_inverse_normal_cdf(p); }

/* ------------------------------------------------------------------------------------------
 * models/rate_matrix.rs
 * ---------------------------------------------------------------------------------------- */

typedef struct {
    double pi[NS];
    double lambda[NS];
    double V[NS * NS]; /* column-major like nalgebra: V[i + NS*k] = i-th component of eigenvector k */
    double sqrt_pi[NS];
    double inv_sqrt_pi[NS];
} orc_rate_matrix;

/* cyclic Jacobi for a symmetric NSxNS matrix (stands in for nalgebra symmetric_eigen,
 * src/models/rate_matrix.rs:100).  A row-major, destroyed.  V column-major. */
static int jacobi_eigen(double A[NS][NS], double *lambda, double *V)
{
    double Vm[NS][NS];
    for (int i = 0; i < NS; ++i)
        for (int j = 0; j < NS; ++j)
            Vm[i][j] = (i == j) ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 100; ++sweep) {
        double off = 0.0;
        for (int i = 0; i < NS; ++i)
            for (int j = i + 1; j < NS; ++j)
                off += A[i][j] * A[i][j];
        if (off < 1e-300)
            break;
        for (int p = 0; p < NS; ++p)
            for (int q = p + 1; q < NS; ++q) {
                if (fabs(A[p][q]) < 1e-300)
                    continue;
                double theta = (A[q][q] - A[p][p]) / (2.0 * A[p][q]);
                double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < NS; ++k) {
                    double akp = A[k][p], akq = A[k][q];
                    A[k][p] = c * akp - s * akq;
                    A[k][q] = s * akp + c * akq;
                }
                for (int k = 0; k < NS; ++k) {
                    double apk = A[p][k], aqk = A[q][k];
                    A[p][k] = c * apk - s * aqk;
                    A[q][k] = s * apk + c * aqk;
                }
                for (int k = 0; k < NS; ++k) {
                    double vkp = Vm[k][p], vkq = Vm[k][q];
                    Vm[k][p] = c * vkp - s * vkq;
                    Vm[k][q] = s * vkp + c * vkq;
                }
            }
    }
    for (int k = 0; k < NS; ++k) {
        lambda[k] = A[k][k];
        for (int i = 0; i < NS; ++i)
            V[i + NS * k] = Vm[i][k];
    }
    return ORC_OK;
}

/* src/models/rate_matrix.rs:34-110.  W row-major (W[i*NS+j]). */
int orc_rate_matrix_new(const double *pi_in, const double *W, int normalize, orc_rate_matrix *out)
{
    double pi_sum = 0.0;
    for (int i = 0; i < NS; ++i)
        pi_sum += pi_in[i];
    if (pi_sum <= 0.0)
        return ORC_ERR_INVALID_MODEL_PARAMETER;
    double pi[NS];
    for (int i = 0; i < NS; ++i)
        pi[i] = pi_in[i] / pi_sum;
    double q[NS][NS];
    for (int i = 0; i < NS; ++i)
        for (int j = 0; j < NS; ++j)
            q[i][j] = (i != j) ? pi[j] * W[i * NS + j] : 0.0;
    for (int i = 0; i < NS; ++i) {
        double rs = 0.0;
        for (int j = 0; j < NS; ++j)
            rs += q[i][j];
        q[i][i] = -rs;
    }
    if (normalize) {
        double mu = 0.0;
        for (int i = 0; i < NS; ++i)
            mu -= pi[i] * q[i][i];
        if (mu <= 0.0)
            return ORC_ERR_NUMERICAL_INSTABILITY;
        for (int i = 0; i < NS; ++i)
            for (int j = 0; j < NS; ++j)
                q[i][j] /= mu;
    }
    double s[NS][NS];
    for (int i = 0; i < NS; ++i) {
        out->pi[i] = pi[i];
        out->sqrt_pi[i] = sqrt(pi[i]);
        out->inv_sqrt_pi[i] = 1.0 / out->sqrt_pi[i];
    }
    for (int i = 0; i < NS; ++i)
        for (int j = 0; j < NS; ++j)
            s[i][j] = out->sqrt_pi[i] * q[i][j] * out->inv_sqrt_pi[j];
    /* symmetric_eigen reads the lower triangle; symmetrise explicitly so Jacobi sees an exactly
     * symmetric input (S is symmetric up to round-off). */
    for (int i = 0; i < NS; ++i)
        for (int j = i + 1; j < NS; ++j)
            s[i][j] = s[j][i];
    return jacobi_eigen(s, out->lambda, out->V);
}

/* src/models/rate_matrix.rs:113-135.  P row-major. */
void orc_transition(const orc_rate_matrix *m, double t, double *P)
{
    if (t == 0.0) { /* :115-117 exact identity */
        for (int i = 0; i < NS; ++i)
            for (int j = 0; j < NS; ++j)
                P[i * NS + j] = (i == j) ? 1.0 : 0.0;
        return;
    }
    double e[NS];
    for (int k = 0; k < NS; ++k)
        e[k] = exp(m->lambda[k] * t);
    /* (V * E): column k scaled by e[k]; the zero terms of the dense product are exact zeros */
    double VE[NS * NS];
    for (int k = 0; k < NS; ++k)
        for (int i = 0; i < NS; ++i)
            VE[i + NS * k] = m->V[i + NS * k] * e[k];
    for (int i = 0; i < NS; ++i)
        for (int j = 0; j < NS; ++j) {
            double acc = VE[i + NS * 0] * m->V[j + NS * 0];
            for (int k = 1; k < NS; ++k)
                acc += VE[i + NS * k] * m->V[j + NS * k];
            P[i * NS + j] = m->inv_sqrt_pi[i] * acc * m->sqrt_pi[j];
        }
}

/* ------------------------------------------------------------------------------------------
 * alphabet.rs: Nucleotide::profile (src/alphabet.rs:44-64)
 * ---------------------------------------------------------------------------------------- */
static int __attribute__((unused)) nucleotide_profile(uint8_t c, double *p)
{
    const double t = 1.0 / 3.0;
    switch (c) {
    case 'A': p[0] = 1; p[1] = 0; p[2] = 0; p[3] = 0; return 1;
    case 'C': p[0] = 0; p[1] = 1; p[2] = 0; p[3] = 0; return 1;
    case 'G': p[0] = 0; p[1] = 0; p[2] = 1; p[3] = 0; return 1;
    case 'T':
    case 'U': p[0] = 0; p[1] = 0; p[2] = 0; p[3] = 1; return 1;
    case 'R': p[0] = .5; p[1] = 0; p[2] = .5; p[3] = 0; return 1;
    case 'Y': p[0] = 0; p[1] = .5; p[2] = 0; p[3] = .5; return 1;
    case 'S': p[0] = 0; p[1] = .5; p[2] = .5; p[3] = 0; return 1;
    case 'W': p[0] = .5; p[1] = 0; p[2] = 0; p[3] = .5; return 1;
    case 'K': p[0] = 0; p[1] = 0; p[2] = .5; p[3] = .5; return 1;
    case 'M': p[0] = .5; p[1] = .5; p[2] = 0; p[3] = 0; return 1;
    case 'B': p[0] = 0; p[1] = t; p[2] = t; p[3] = t; return 1;
    case 'D': p[0] = t; p[1] = 0; p[2] = t; p[3] = t; return 1;
    case 'H': p[0] = t; p[1] = t; p[2] = 0; p[3] = t; return 1;
    case 'V': p[0] = t; p[1] = t; p[2] = t; p[3] = 0; return 1;
    case 'N': p[0] = .25; p[1] = .25; p[2] = .25; p[3] = .25; return 1;
    case '-': p[0] = 1; p[1] = 1; p[2] = 1; p[3] = 1; return 1;
    default: return 0;
    }
}

/* AminoAcid::profile, src/alphabet.rs:76-116: CANONICAL = "ACDEFGHIKLMNPQRSTVWY" (index_of = position), B = D|N,
 * Z = E|Q, J = I|L at 0.5 each, X and the gap all ones. */
static int __attribute__((unused)) amino_acid_profile(uint8_t c, double *p)
{
    static const char canonical[] = "ACDEFGHIKLMNPQRSTVWY";
    for (int i = 0; i < 20; ++i)
        p[i] = 0.0;
    for (int i = 0; i < 20; ++i)
        if ((uint8_t)canonical[i] == c) {
            p[i] = 1.0;
            return 1;
        }
    switch (c) {
    case 'B': p[3] = 0.5; p[11] = 0.5; return 1;  /* written as in the reference: indices 3 and 11 */
    case 'Z': p[4] = 0.5; p[14] = 0.5; return 1;  /* 4 and 14 */
    case 'J': p[8] = 0.5; p[9] = 0.5; return 1;   /* 8 and 9 */
    case 'X':
    case '-':
        for (int i = 0; i < 20; ++i)
            p[i] = 1.0;
        return 1;
    default: return 0;
    }
}

#if NS == 4
#define alphabet_profile nucleotide_profile
int orc_nucleotide_profile(uint8_t c, double *p) { return nucleotide_profile(c, p); }
#elif NS == 20
#define alphabet_profile amino_acid_profile
int orc_amino_acid_profile(uint8_t c, double *p) { return amino_acid_profile(c, p); }
#else
#error "NS must be 4 or 20"
#endif
int orc_n_states(void) { return NS; }

/* Profile::scale, src/tree/likelihood/profile.rs:20-39 */
static void profile_scale(double *v, double *log_scale)
{
    double mx = -INFINITY;
    for (int i = 0; i < NS; ++i)
        mx = fmax(mx, v[i]);
    if (mx <= 0.0)
        return;
    for (int i = 0; i < NS; ++i)
        v[i] /= mx;
    *log_scale += log(mx);
}

/* log_sum_exp, src/tree/likelihood.rs:61-68 */
static double log_sum_exp(const double *xs, int n)
{
    double mx = -INFINITY;
    for (int i = 0; i < n; ++i)
        mx = fmax(mx, xs[i]);
    if (!isfinite(mx))
        return mx;
    double sum = 0.0;
    for (int i = 0; i < n; ++i)
        sum += exp(xs[i] - mx);
    return mx + log(sum);
}

/* ------------------------------------------------------------------------------------------
 * Flattened tree handed in by the Python restatement of the tree accessors (oracle/refhost.py):
 * dense ids 0..n_nodes-1; postorder per src/iter.rs:112-138; preorder per
 * src/iter/node_iter.rs:137-161; children in arena order; blen[v] = weight of edge parent->v
 * (f32 widened to f64, None -> 0.0: likelihood.rs:115-118); leaf_row[v] = row inside a pattern
 * for leaves (node without children, src/node/simple_rnode.rs:74-76), -1 = no row.
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    int n_nodes;
    const int *postorder;
    const int *preorder;
    const int *child_off; /* n_nodes+1 */
    const int *children;
    const int *parent; /* -1 for root */
    const double *blen;
    const int *leaf_row;
    int root;
} orc_tree;

typedef struct {
    const orc_tree *tree;
    const orc_rate_matrix *rm;
    int n_cat;
    const double *rates;
    const double *weights;
    int n_rows;
    int64_t n_patterns;
    const uint8_t *patterns; /* [pattern][row] ASCII bytes, CompressedColumns.patterns */
    const uint64_t *mult;    /* nullable => 1 */
    int ref_linear_search;   /* reproduce the O(n) leaf_id_map.position() scan, likelihood.rs:99 */
    const int *leaf_id_map;  /* row -> node id (for the linear search) */
    /* outputs */
    double *per_pattern; /* nullable */
    uint8_t *states;     /* nullable, [node][pattern] */
    double *posteriors;  /* nullable, [node][pattern][4] */
    /* work split */
    int64_t p_begin, p_end;
    int status;
} orc_job;

/* prune_pattern_category, src/tree/likelihood.rs:83-143.  prof: [n_nodes][NS], ls: [n_nodes].
 * P(t) is recomputed for every child exactly as the reference does (:119). */
static int prune_pattern_category(const orc_job *J, int cat, const uint8_t *pattern, double *prof,
                                  double *ls, double *cat_ll)
{
    const orc_tree *T = J->tree;
    double P[NS * NS];
    for (int k = 0; k < T->n_nodes; ++k) {
        int v = T->postorder[k];
        int nch = T->child_off[v + 1] - T->child_off[v];
        double *pv = prof + (size_t)v * NS;
        if (nch == 0) {
            int pos = -1;
            if (J->ref_linear_search) { /* :99 */
                for (int r = 0; r < J->n_rows; ++r)
                    if (J->leaf_id_map[r] == v) {
                        pos = r;
                        break;
                    }
            } else
                pos = T->leaf_row[v];
            if (pos < 0)
                return ORC_ERR_INVALID_ALIGNMENT; /* :100 */
            if (!alphabet_profile(pattern[pos], pv))
                return ORC_ERR_ALPHABET_MISMATCH; /* :106 */
            ls[v] = 0.0;
            profile_scale(pv, &ls[v]); /* :108 */
        } else {
            double vals[NS];
            for (int i = 0; i < NS; ++i)
                vals[i] = 1.0;
            double sum_ls = 0.0;
            for (int ci = T->child_off[v]; ci < T->child_off[v + 1]; ++ci) {
                int c = T->children[ci];
                const double *pc = prof + (size_t)c * NS;
                orc_transition(J->rm, J->rates[cat] * T->blen[c], P); /* gtr.rs:61-64 */
                double contrib[NS];
                for (int i = 0; i < NS; ++i) /* gemv, column order j = 0..NS-1 */
                    contrib[i] = P[i * NS + 0] * pc[0];
                for (int j = 1; j < NS; ++j)
                    for (int i = 0; i < NS; ++i)
                        contrib[i] += P[i * NS + j] * pc[j];
                for (int i = 0; i < NS; ++i)
                    vals[i] *= contrib[i];
                sum_ls += ls[c];
            }
            for (int i = 0; i < NS; ++i)
                pv[i] = vals[i];
            ls[v] = sum_ls;
            profile_scale(pv, &ls[v]);
        }
    }
    const double *pr = prof + (size_t)T->root * NS;
    double root_mass = 0.0;
    for (int i = 0; i < NS; ++i)
        root_mass += J->rm->pi[i] * pr[i];
    *cat_ll = log(J->weights[cat]) + log(root_mass) + ls[T->root]; /* :140 */
    return ORC_OK;
}

/* compute_log_likelihood pattern loop, src/tree/likelihood.rs:186-206 (writes per-pattern site_ll;
 * the caller does the sequential multiplicity-weighted sum) */
static void *lnl_ref_worker(void *arg)
{
    orc_job *J = (orc_job *)arg;
    const orc_tree *T = J->tree;
    double *prof = (double *)malloc(sizeof(double) * NS * (size_t)T->n_nodes);
    double *ls = (double *)malloc(sizeof(double) * (size_t)T->n_nodes);
    double *cat_lls = (double *)malloc(sizeof(double) * (size_t)J->n_cat);
    for (int64_t p = J->p_begin; p < J->p_end; ++p) {
        const uint8_t *pattern = J->patterns + (size_t)p * (size_t)J->n_rows;
        for (int c = 0; c < J->n_cat; ++c) {
            int rc = prune_pattern_category(J, c, pattern, prof, ls, &cat_lls[c]);
            if (rc) {
                J->status = rc;
                goto done;
            }
        }
        J->per_pattern[p] = log_sum_exp(cat_lls, J->n_cat);
    }
done:
    free(prof);
    free(ls);
    free(cat_lls);
    return NULL;
}

/* compute_marginal_asr pattern loop, src/tree/likelihood.rs:275-399 */
static void *asr_ref_worker(void *arg)
{
    orc_job *J = (orc_job *)arg;
    const orc_tree *T = J->tree;
    const int n = T->n_nodes, K = J->n_cat;
    double *prof = (double *)malloc(sizeof(double) * NS * (size_t)n);
    double *ls = (double *)malloc(sizeof(double) * (size_t)n);
    double *cat_lls = (double *)malloc(sizeof(double) * (size_t)K);
    double *mixw = (double *)malloc(sizeof(double) * (size_t)K);
    double *post = (double *)malloc(sizeof(double) * NS * (size_t)n * (size_t)K); /* [cat][node][NS] */
    double P[NS * NS];
    for (int64_t p = J->p_begin; p < J->p_end; ++p) {
        const uint8_t *pattern = J->patterns + (size_t)p * (size_t)J->n_rows;
        for (int c = 0; c < K; ++c) {
            int rc = prune_pattern_category(J, c, pattern, prof, ls, &cat_lls[c]); /* :286-297 */
            if (rc) {
                J->status = rc;
                goto done;
            }
            double *pc = post + (size_t)c * n * NS;
            for (int k = 0; k < n; ++k) { /* DOWN pass :303-344 */
                int v = T->preorder[k];
                double *pv = pc + (size_t)v * NS;
                if (v == T->root) {
                    double sum = 0.0;
                    for (int i = 0; i < NS; ++i) {
                        pv[i] = J->rm->pi[i] * prof[(size_t)v * NS + i];
                        sum += pv[i];
                    }
                    if (sum > 0.0)
                        for (int i = 0; i < NS; ++i)
                            pv[i] /= sum;
                } else {
                    int par = T->parent[v];
                    const double *pp = pc + (size_t)par * NS;
                    orc_transition(J->rm, J->rates[c] * T->blen[v], P);
                    double msg[NS]; /* P^T * post_p, gemv over the materialised transpose :327-328 */
                    for (int i = 0; i < NS; ++i)
                        msg[i] = P[0 * NS + i] * pp[0];
                    for (int j = 1; j < NS; ++j)
                        for (int i = 0; i < NS; ++i)
                            msg[i] += P[j * NS + i] * pp[j];
                    double sum = 0.0;
                    for (int i = 0; i < NS; ++i) {
                        pv[i] = prof[(size_t)v * NS + i] * msg[i];
                        sum += pv[i];
                    }
                    if (sum > 0.0)
                        for (int i = 0; i < NS; ++i)
                            pv[i] /= sum;
                }
            }
        }
        double site_ll = log_sum_exp(cat_lls, K); /* :349 */
        J->per_pattern[p] = site_ll;
        if (isfinite(site_ll)) /* :356-363 */
            for (int c = 0; c < K; ++c)
                mixw[c] = exp(cat_lls[c] - site_ll);
        else
            for (int c = 0; c < K; ++c)
                mixw[c] = 1.0 / (double)K;
        for (int v = 0; v < n; ++v) { /* :365-398 */
            double mixed[NS] = {0.0};
            for (int c = 0; c < K; ++c) {
                const double *pv = post + ((size_t)c * n + (size_t)v) * NS;
                for (int i = 0; i < NS; ++i)
                    mixed[i] += mixw[c] * pv[i];
            }
            double sum = 0.0;
            for (int i = 0; i < NS; ++i)
                sum += mixed[i];
            if (sum > 0.0)
                for (int i = 0; i < NS; ++i)
                    mixed[i] /= sum;
            int best = 0; /* Iterator::max_by => LAST maximal index :385-390 */
            for (int i = 1; i < NS; ++i)
                if (!(mixed[best] > mixed[i]))
                    best = i;
            if (J->states)
                J->states[(size_t)v * (size_t)J->n_patterns + (size_t)p] = (uint8_t)best;
            if (J->posteriors)
                for (int i = 0; i < NS; ++i)
                    J->posteriors[((size_t)v * (size_t)J->n_patterns + (size_t)p) * NS + i] = mixed[i];
        }
    }
done:
    free(prof);
    free(ls);
    free(cat_lls);
    free(mixw);
    free(post);
    return NULL;
}

/* compute_joint_asr pattern loop, src/tree/likelihood.rs:421-560 (Viterbi in log space, per category;
 * the category with the strictly greatest ln(weight) + root maximum wins, first on ties :525-527;
 * first maximal index for every inner argmax :496-503, :517-523; no category with a finite value =>
 * best_overall_states stays empty and every node keeps state 0 :458-459, :543-549).
 * cvals: [n_nodes][NS], ptrs: [n_nodes][NS] (pointer of edge parent(v)->v, stored at v). */
static void *joint_ref_worker(void *arg)
{
    orc_job *J = (orc_job *)arg;
    const orc_tree *T = J->tree;
    const int n = T->n_nodes, K = J->n_cat;
    double *cvals = (double *)malloc(sizeof(double) * NS * (size_t)n);
    int *ptrs = (int *)malloc(sizeof(int) * NS * (size_t)n);
    int *st = (int *)malloc(sizeof(int) * (size_t)n);
    int *best_states = (int *)malloc(sizeof(int) * (size_t)n);
    double P[NS * NS];
    for (int64_t p = J->p_begin; p < J->p_end; ++p) {
        const uint8_t *pattern = J->patterns + (size_t)p * (size_t)J->n_rows;
        double best_overall = -INFINITY;
        int have_best = 0;
        for (int c = 0; c < K; ++c) {
            for (int k = 0; k < n; ++k) {
                int v = T->postorder[k];
                double *cv = cvals + (size_t)v * NS;
                if (T->child_off[v + 1] == T->child_off[v]) {
                    int pos = T->leaf_row[v];
                    double prof[NS];
                    if (pos < 0) {
                        J->status = ORC_ERR_INVALID_ALIGNMENT; /* :469-473 */
                        goto done;
                    }
                    if (!alphabet_profile(pattern[pos], prof)) {
                        J->status = ORC_ERR_ALPHABET_MISMATCH; /* :475-476 */
                        goto done;
                    }
                    for (int i = 0; i < NS; ++i) /* :477-480 */
                        cv[i] = prof[i] > 0.0 ? log(prof[i]) : -INFINITY;
                } else {
                    for (int i = 0; i < NS; ++i)
                        cv[i] = 0.0;
                    for (int ci = T->child_off[v]; ci < T->child_off[v + 1]; ++ci) {
                        int ch = T->children[ci];
                        const double *cc = cvals + (size_t)ch * NS;
                        orc_transition(J->rm, J->rates[c] * T->blen[ch], P);
                        for (int i = 0; i < NS; ++i) { /* :494-507 */
                            double local_max = -INFINITY;
                            int local_arg = 0;
                            for (int j = 0; j < NS; ++j) {
                                double val = log(P[i * NS + j]) + cc[j];
                                if (val > local_max) {
                                    local_max = val;
                                    local_arg = j;
                                }
                            }
                            cv[i] += local_max;
                            ptrs[(size_t)ch * NS + i] = local_arg;
                        }
                    }
                }
            }
            const double *rc = cvals + (size_t)T->root * NS;
            int root_best = 0;
            double root_max = -INFINITY;
            for (int i = 0; i < NS; ++i) { /* :514-523 */
                double ll = log(J->rm->pi[i]) + rc[i];
                if (ll > root_max) {
                    root_max = ll;
                    root_best = i;
                }
            }
            double cat_ll = log(J->weights[c]) + root_max;
            if (cat_ll > best_overall) { /* :526-540 */
                best_overall = cat_ll;
                have_best = 1;
                st[T->root] = root_best;
                for (int k = 0; k < n; ++k) {
                    int v = T->preorder[k];
                    for (int ci = T->child_off[v]; ci < T->child_off[v + 1]; ++ci) {
                        int ch = T->children[ci];
                        st[ch] = ptrs[(size_t)ch * NS + st[v]];
                    }
                }
                memcpy(best_states, st, sizeof(int) * (size_t)n);
            }
        }
        J->per_pattern[p] = best_overall; /* :543 (the caller weights by multiplicity) */
        if (J->states)
            for (int v = 0; v < n; ++v)
                J->states[(size_t)v * (size_t)J->n_patterns + (size_t)p] = have_best ? (uint8_t)best_states[v] : 0;
    }
done:
    free(cvals);
    free(ptrs);
    free(st);
    free(best_states);
    return NULL;
}

/* ------------------------------------------------------------------------------------------
 * cpu_opt: same mathematics, P(t) cached per (edge, category), SoA blocks of OB patterns so the
 * inner loops vectorise, threads over pattern blocks.  The honest multi-core CPU competitor
 * (BASELINE.md §4 B2).  Divide-by-max scaling + ln per node kept (Profile::scale semantics).
 * ---------------------------------------------------------------------------------------- */
#define OB 8
typedef struct {
    orc_job job;
    const double *Pcache; /* [node][cat][NS*NS] row-major; identity for rate*t == 0 */
} orc_optjob;

static void *lnl_opt_worker(void *arg)
{
    orc_optjob *OJ = (orc_optjob *)arg;
    orc_job *J = &OJ->job;
    const orc_tree *T = J->tree;
    const int n = T->n_nodes, K = J->n_cat;
    /* clv[node][cat][state][OB], ls[node][cat][OB] */
    double *clv = (double *)malloc(sizeof(double) * (size_t)n * K * NS * OB);
    double *ls = (double *)malloc(sizeof(double) * (size_t)n * K * OB);
    double *cat_lls = (double *)malloc(sizeof(double) * (size_t)K);
    for (int64_t p0 = J->p_begin; p0 < J->p_end; p0 += OB) {
        int nb = (int)((J->p_end - p0) < OB ? (J->p_end - p0) : OB);
        for (int k = 0; k < n; ++k) {
            int v = T->postorder[k];
            int nch = T->child_off[v + 1] - T->child_off[v];
            if (nch == 0) {
                int row = T->leaf_row[v];
                if (row < 0) {
                    J->status = ORC_ERR_INVALID_ALIGNMENT;
                    goto done;
                }
                for (int b = 0; b < OB; ++b) {
                    double pv[NS], l = 0.0;
                    for (int i = 0; i < NS; ++i)
                        pv[i] = 1.0;
                    if (b < nb) {
                        if (!alphabet_profile(J->patterns[(size_t)(p0 + b) * J->n_rows + row], pv)) {
                            J->status = ORC_ERR_ALPHABET_MISMATCH;
                            goto done;
                        }
                        profile_scale(pv, &l);
                    }
                    for (int c = 0; c < K; ++c) {
                        for (int i = 0; i < NS; ++i)
                            clv[(((size_t)v * K + c) * NS + i) * OB + b] = pv[i];
                        ls[((size_t)v * K + c) * OB + b] = l;
                    }
                }
            } else {
                for (int c = 0; c < K; ++c) {
                    double *out = clv + ((size_t)v * K + c) * NS * OB;
                    double *lo = ls + ((size_t)v * K + c) * OB;
                    for (int i = 0; i < NS * OB; ++i)
                        out[i] = 1.0;
                    for (int b = 0; b < OB; ++b)
                        lo[b] = 0.0;
                    for (int ci = T->child_off[v]; ci < T->child_off[v + 1]; ++ci) {
                        int ch = T->children[ci];
                        const double *P = OJ->Pcache + ((size_t)ch * K + c) * (NS * NS);
                        const double *in = clv + ((size_t)ch * K + c) * NS * OB;
                        const double *li = ls + ((size_t)ch * K + c) * OB;
                        for (int i = 0; i < NS; ++i)
                            for (int b = 0; b < OB; ++b) {
                                double acc = P[i * NS + 0] * in[0 * OB + b];
                                for (int j = 1; j < NS; ++j)
                                    acc += P[i * NS + j] * in[j * OB + b];
                                out[i * OB + b] *= acc;
                            }
                        for (int b = 0; b < OB; ++b)
                            lo[b] += li[b];
                    }
                    for (int b = 0; b < OB; ++b) {
                        double mx = out[b];
                        for (int i = 1; i < NS; ++i)
                            mx = fmax(mx, out[i * OB + b]);
                        if (mx > 0.0) {
                            for (int i = 0; i < NS; ++i)
                                out[i * OB + b] /= mx;
                            lo[b] += log(mx);
                        }
                    }
                }
            }
        }
        for (int b = 0; b < nb; ++b) {
            for (int c = 0; c < K; ++c) {
                const double *r = clv + ((size_t)T->root * K + c) * NS * OB;
                double mass = 0.0;
                for (int i = 0; i < NS; ++i)
                    mass += J->rm->pi[i] * r[i * OB + b];
                cat_lls[c] = log(J->weights[c]) + log(mass) + ls[((size_t)T->root * K + c) * OB + b];
            }
            J->per_pattern[p0 + b] = log_sum_exp(cat_lls, K);
        }
    }
done:
    free(clv);
    free(ls);
    free(cat_lls);
    return NULL;
}

/* ------------------------------------------------------------------------------------------
 * Entry points
 * mode: 0 = ref_faithful (reference loop nest), 1 = ref_faithful + linear leaf search,
 *       2 = cpu_opt.
 * ---------------------------------------------------------------------------------------- */
static int run_jobs(void *(*fn)(void *), void *jobs, size_t job_size, int n_threads, int64_t n_patterns,
                    int align)
{
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)n_threads);
    int64_t chunk = (n_patterns + n_threads - 1) / n_threads;
    chunk = (chunk + align - 1) / align * align;
    int started = 0;
    for (int t = 0; t < n_threads; ++t) {
        orc_job *J = (orc_job *)((char *)jobs + job_size * (size_t)t);
        J->p_begin = (int64_t)t * chunk;
        J->p_end = J->p_begin + chunk < n_patterns ? J->p_begin + chunk : n_patterns;
        J->status = ORC_OK;
        if (J->p_begin >= J->p_end) {
            J->p_begin = J->p_end = 0;
            continue;
        }
        if (n_threads == 1)
            fn(J);
        else {
            pthread_create(&th[t], NULL, fn, J);
            started |= 1;
        }
    }
    int status = ORC_OK;
    for (int t = 0; t < n_threads; ++t) {
        orc_job *J = (orc_job *)((char *)jobs + job_size * (size_t)t);
        if (n_threads > 1 && J->p_begin < J->p_end)
            pthread_join(th[t], NULL);
        if (J->status && !status)
            status = J->status;
    }
    (void)started;
    free(th);
    return status;
}

int orc_log_likelihood(int n_nodes, const int *postorder, const int *preorder, const int *child_off,
                       const int *children, const int *parent, const double *blen,
                       const int *leaf_row, int root, const orc_rate_matrix *rm, int n_cat,
                       const double *rates, const double *weights, int n_rows, int64_t n_patterns,
                       const uint8_t *patterns, const uint64_t *mult, int mode, int n_threads,
                       double *total, double *per_pattern)
{
    orc_tree T = {n_nodes, postorder, preorder, child_off, children, parent, blen, leaf_row, root};
    if (n_threads < 1)
        n_threads = 1;
    int *leaf_id_map = (int *)malloc(sizeof(int) * (size_t)(n_rows > 0 ? n_rows : 1));
    for (int r = 0; r < n_rows; ++r)
        leaf_id_map[r] = -1;
    for (int v = 0; v < n_nodes; ++v)
        if (leaf_row[v] >= 0 && leaf_row[v] < n_rows)
            leaf_id_map[leaf_row[v]] = v;
    double *pp = per_pattern ? per_pattern : (double *)malloc(sizeof(double) * (size_t)(n_patterns > 0 ? n_patterns : 1));
    int status;
    if (mode == 2) {
        double *Pc = (double *)malloc(sizeof(double) * (NS * NS) * (size_t)n_nodes * (size_t)n_cat);
        for (int v = 0; v < n_nodes; ++v)
            for (int c = 0; c < n_cat; ++c)
                orc_transition(rm, rates[c] * blen[v], Pc + ((size_t)v * n_cat + c) * (NS * NS));
        orc_optjob *jobs = (orc_optjob *)calloc((size_t)n_threads, sizeof(orc_optjob));
        for (int t = 0; t < n_threads; ++t) {
            orc_job *J = &jobs[t].job;
            J->tree = &T; J->rm = rm; J->n_cat = n_cat; J->rates = rates; J->weights = weights;
            J->n_rows = n_rows; J->n_patterns = n_patterns; J->patterns = patterns; J->mult = mult;
            J->per_pattern = pp; J->leaf_id_map = leaf_id_map;
            jobs[t].Pcache = Pc;
        }
        status = run_jobs(lnl_opt_worker, jobs, sizeof(orc_optjob), n_threads, n_patterns, OB);
        free(jobs);
        free(Pc);
    } else {
        orc_job *jobs = (orc_job *)calloc((size_t)n_threads, sizeof(orc_job));
        for (int t = 0; t < n_threads; ++t) {
            orc_job *J = &jobs[t];
            J->tree = &T; J->rm = rm; J->n_cat = n_cat; J->rates = rates; J->weights = weights;
            J->n_rows = n_rows; J->n_patterns = n_patterns; J->patterns = patterns; J->mult = mult;
            J->per_pattern = pp; J->leaf_id_map = leaf_id_map; J->ref_linear_search = (mode == 1);
        }
        status = run_jobs(lnl_ref_worker, jobs, sizeof(orc_job), n_threads, n_patterns, 1);
        free(jobs);
    }
    if (!status) { /* sequential sum in pattern order, likelihood.rs:205 */
        double tot = 0.0;
        for (int64_t p = 0; p < n_patterns; ++p)
            tot += (mult ? (double)mult[p] : 1.0) * pp[p];
        *total = tot;
    }
    if (!per_pattern)
        free(pp);
    free(leaf_id_map);
    return status;
}

int orc_marginal_asr(int n_nodes, const int *postorder, const int *preorder, const int *child_off,
                     const int *children, const int *parent, const double *blen, const int *leaf_row,
                     int root, const orc_rate_matrix *rm, int n_cat, const double *rates,
                     const double *weights, int n_rows, int64_t n_patterns, const uint8_t *patterns,
                     const uint64_t *mult, int n_threads, uint8_t *states, double *posteriors,
                     double *total, double *per_pattern)
{
    orc_tree T = {n_nodes, postorder, preorder, child_off, children, parent, blen, leaf_row, root};
    if (n_threads < 1)
        n_threads = 1;
    double *pp = per_pattern ? per_pattern : (double *)malloc(sizeof(double) * (size_t)(n_patterns > 0 ? n_patterns : 1));
    orc_job *jobs = (orc_job *)calloc((size_t)n_threads, sizeof(orc_job));
    for (int t = 0; t < n_threads; ++t) {
        orc_job *J = &jobs[t];
        J->tree = &T; J->rm = rm; J->n_cat = n_cat; J->rates = rates; J->weights = weights;
        J->n_rows = n_rows; J->n_patterns = n_patterns; J->patterns = patterns; J->mult = mult;
        J->per_pattern = pp; J->states = states; J->posteriors = posteriors;
    }
    int status = run_jobs(asr_ref_worker, jobs, sizeof(orc_job), n_threads, n_patterns, 1);
    free(jobs);
    if (!status) {
        double tot = 0.0;
        for (int64_t p = 0; p < n_patterns; ++p)
            tot += (mult ? (double)mult[p] : 1.0) * pp[p];
        *total = tot;
    }
    if (!per_pattern)
        free(pp);
    return status;
}

int orc_joint_asr(int n_nodes, const int *postorder, const int *preorder, const int *child_off,
                  const int *children, const int *parent, const double *blen, const int *leaf_row,
                  int root, const orc_rate_matrix *rm, int n_cat, const double *rates,
                  const double *weights, int n_rows, int64_t n_patterns, const uint8_t *patterns,
                  const uint64_t *mult, int n_threads, uint8_t *states, double *total, double *per_pattern)
{
    orc_tree T = {n_nodes, postorder, preorder, child_off, children, parent, blen, leaf_row, root};
    if (n_threads < 1)
        n_threads = 1;
    double *pp = per_pattern ? per_pattern : (double *)malloc(sizeof(double) * (size_t)(n_patterns > 0 ? n_patterns : 1));
    orc_job *jobs = (orc_job *)calloc((size_t)n_threads, sizeof(orc_job));
    for (int t = 0; t < n_threads; ++t) {
        orc_job *J = &jobs[t];
        J->tree = &T; J->rm = rm; J->n_cat = n_cat; J->rates = rates; J->weights = weights;
        J->n_rows = n_rows; J->n_patterns = n_patterns; J->patterns = patterns; J->mult = mult;
        J->per_pattern = pp; J->states = states;
    }
    int status = run_jobs(joint_ref_worker, jobs, sizeof(orc_job), n_threads, n_patterns, 1);
    free(jobs);
    if (!status) { /* likelihood.rs:543, sequential in pattern order */
        double tot = 0.0;
        for (int64_t p = 0; p < n_patterns; ++p)
            tot += (mult ? (double)mult[p] : 1.0) * pp[p];
        *total = tot;
    }
    if (!per_pattern)
        free(pp);
    return status;
}

/* Brute-force likelihood by enumeration over all internal-node states (validation of the
 * restatement itself; exponential, only for trees with <= ~8 internal nodes).  Single category. */
double orc_bruteforce_site_likelihood(int n_nodes, const int *child_off, const int *children,
                                      const int *parent, const double *blen, const int *leaf_row,
                                      int root, const orc_rate_matrix *rm, double rate,
                                      const uint8_t *pattern)
{
    int internal[32], n_int = 0;
    (void)children;
    if (n_nodes <= 0 || n_nodes > 24)
        return NAN;
    for (int v = 0; v < n_nodes; ++v)
        if (child_off[v + 1] - child_off[v] > 0)
            internal[n_int++] = v;
    if (n_int > 12)
        return NAN;
    double *P = (double *)malloc(sizeof(double) * (NS * NS) * (size_t)n_nodes);
    for (int v = 0; v < n_nodes; ++v)
        orc_transition(rm, rate * blen[v], P + (size_t)v * (NS * NS));
    int *state = (int *)calloc((size_t)n_nodes, sizeof(int));
    long combos = 1;
    for (int i = 0; i < n_int; ++i)
        combos *= NS;
    double total = 0.0;
    for (long code = 0; code < combos; ++code) {
        long c = code;
        for (int i = 0; i < n_int; ++i) {
            state[internal[i]] = (int)(c % NS);
            c /= NS;
        }
        double pr = rm->pi[state[root]];
        for (int v = 0; v < n_nodes && pr != 0.0; ++v) {
            if (v == root)
                continue;
            int par = parent[v];
            const double *Pv = P + (size_t)v * (NS * NS);
            if (child_off[v + 1] - child_off[v] > 0) {
                pr *= Pv[state[par] * NS + state[v]];
            } else {
                double prof[NS];
                alphabet_profile(pattern[leaf_row[v]], prof);
                double s = 0.0;
                for (int j = 0; j < NS; ++j)
                    s += Pv[state[par] * NS + j] * prof[j];
                pr *= s;
            }
        }
        total += pr;
    }
    free(P);
    free(state);
    return total;
}
