/* The C ABI exercised from plain C, in exactly the order the Rust shim (bindings/shim/b200.rs) calls it:
 *   ctx -> set_model -> set_tree -> compress_rows -> log_likelihood          (TreeLikelihood::log_likelihood)
 *   ctx -> set_alignment -> set_model -> set_tree -> marginal_asr            (MarginalAsr::marginal_asr)
 *   joint_asr, score_nni, score_trees on the same context; then the same order in a 20-state context.
 * No Python in between.  The fixture is the reference's own (src/tree/likelihood/integration_test.rs:14-17):
 * tree ((A:0.601,B:0.301)C:0.100,D:0.200)E:0.001 under JC69; the expected values are the known answers of
 * tests/golden/kat.json (brute-force enumeration, checked against the oracle).
 * Built by __graft_entry__.build() against the product library (run under `-m gpu`) and against the emulation
 * harness build (run on the CPU).  Exit code 0 = every check passed. */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/phylo_b200.h"

static int failures = 0;
#define CHECK(cond, ...)                                                                  \
    do {                                                                                  \
        if (!(cond)) {                                                                    \
            fprintf(stderr, "FAIL %s:%d: ", __FILE__, __LINE__);                          \
            fprintf(stderr, __VA_ARGS__);                                                 \
            fprintf(stderr, "\n");                                                        \
            ++failures;                                                                   \
        }                                                                                 \
    } while (0)
#define OK(call)                                                                          \
    do {                                                                                  \
        int rc_ = (call);                                                                 \
        if (rc_ != PHB_OK) {                                                              \
            fprintf(stderr, "FAIL %s:%d: %s -> %d (%s)\n", __FILE__, __LINE__, #call, rc_, phb_last_error(ctx)); \
            ++failures;                                                                   \
        }                                                                                 \
    } while (0)

/* The same call order in a 20-state context (PHB_FLAG_AMINO_ACID) under the equal-rates (Poisson) model: pi = 1/20,
 * Q normalised to mean rate 1 has eigenvalues 0 and -20/19 (x19) and ANY orthonormal basis with a constant first column
 * as eigenvectors (a Helmert basis here), so P_ii(t) = 1/20 + 19/20 e^(-20t/19), P_ij(t) = 1/20 - 1/20 e^(-20t/19) in
 * closed form: the expected log-likelihood is a brute-force sum over the 400 states of the two internal nodes. */
static double poisson_p(int i, int j, double t)
{
    const double e = exp(-20.0 * t / 19.0);
    return i == j ? 0.05 + 0.95 * e : 0.05 - 0.05 * e;
}
static double amino_acid_section(const int32_t *post, const int32_t *off, const int32_t *kids, const double *blen,
                                 const int32_t *leaf_row, const int32_t *post2, const int32_t *kids2)
{
    phb_ctx *ctx = NULL;
    OK(phb_ctx_create(NULL, 0, PHB_FLAG_AMINO_ACID, &ctx));
    if (!ctx) return NAN;
    double lambda[20], V[400], pi[20], sp[20], isp[20];
    for (int i = 0; i < 20; ++i) {
        lambda[i] = i ? -20.0 / 19.0 : 0.0;
        pi[i] = 0.05;
        sp[i] = sqrt(0.05);
        isp[i] = 1.0 / sqrt(0.05);
        V[i] = 1.0 / sqrt(20.0); /* column 0; V[i + 20 * k] */
    }
    for (int k = 1; k < 20; ++k)
        for (int i = 0; i < 20; ++i) V[i + 20 * k] = i < k ? 1.0 / sqrt((double)k * (k + 1)) : i == k ? -k / sqrt((double)k * (k + 1)) : 0.0;
    const double rates[1] = {1.0}, weights[1] = {1.0};
    const char *canon = "ACDEFGHIKLMNPQRSTVWY";
    /* rows A, B, D; 8 sites, 4 distinct columns twice; X (any state) in the last one */
    const char *A = "MKVLMKVL", *B = "MRVXMRVX", *D = "MKWLMKWL";
    const uint8_t *rows[3] = {(const uint8_t *)A, (const uint8_t *)B, (const uint8_t *)D};
    OK(phb_set_model(ctx, lambda, V, pi, sp, isp, 1, rates, weights));
    OK(phb_set_tree(ctx, 5, post, off, kids, blen, leaf_row, 0));
    int64_t n_pat = 0;
    uint64_t s2p[8], mult[8];
    OK(phb_compress_rows(ctx, 3, 8, rows, &n_pat, s2p, mult));
    CHECK(n_pat == 4, "20 states: 4 distinct columns expected, got %lld", (long long)n_pat);
    for (int i = 0; i < 8; ++i) CHECK(s2p[i] == (uint64_t)(i & 3), "20 states: site_to_pattern[%d] = %llu", i, (unsigned long long)s2p[i]);
    uint8_t back[12];
    OK(phb_get_alignment(ctx, back, PHB_LAYOUT_ROW_MAJOR, 0, 4));
    CHECK(!memcmp(back, "MKVL", 4) && !memcmp(back + 4, "MRVX", 4) && !memcmp(back + 8, "MKWL", 4), "20 states: patterns read back");
    double total = 0, per[4];
    OK(phb_log_likelihood(ctx, &total, per));
    /* brute force: root E (state e) -> C (state c, 0.100) -> A (0.601), B (0.301); E -> D (0.200) */
    double expect = 0;
    for (int p = 0; p < 4; ++p) {
        const int a = (int)(strchr(canon, A[p]) - canon), d = (int)(strchr(canon, D[p]) - canon);
        const int b = B[p] == 'X' ? -1 : (int)(strchr(canon, B[p]) - canon);
        double site = 0;
        for (int e = 0; e < 20; ++e)
            for (int c = 0; c < 20; ++c) {
                double pb = 0;
                for (int x = 0; x < 20; ++x)
                    if (b < 0 || b == x) pb += poisson_p(c, x, blen[4]);
                site += 0.05 * poisson_p(e, c, blen[1]) * poisson_p(c, a, blen[3]) * pb * poisson_p(e, d, blen[2]);
            }
        CHECK(fabs(per[p] - log(site)) <= 1e-12, "20 states: pattern %d lnL %.17g, closed form %.17g", p, per[p], log(site));
        expect += 2.0 * log(site);
    }
    CHECK(fabs(total - expect) <= 1e-12 * fabs(expect), "20 states: total %.17g, closed form %.17g", total, expect);
    /* P(t) itself */
    double Pm[400];
    const double t1[1] = {0.301};
    OK(phb_transition_matrices(ctx, 1, t1, Pm));
    CHECK(fabs(Pm[0] - poisson_p(0, 0, 0.301)) <= 1e-14 && fabs(Pm[7 * 20 + 3] - poisson_p(7, 3, 0.301)) <= 1e-14, "20 states: P(0.301)");
    /* marginal and joint ASR: leaves reproduce their letters */
    uint8_t states[5 * 4], jstates[5 * 4];
    double t_asr = 0, t_joint = 0;
    OK(phb_marginal_asr(ctx, states, NULL, &t_asr));
    OK(phb_joint_asr(ctx, jstates, &t_joint));
    CHECK(fabs(t_asr - total) <= 1e-12 * fabs(total) && t_joint <= total + 1e-9, "20 states: ASR totals %.17g / %.17g", t_asr, t_joint);
    for (int p = 0; p < 4; ++p) {
        CHECK(canon[states[3 * 4 + p]] == A[p] && canon[jstates[3 * 4 + p]] == A[p], "20 states: leaf A pattern %d", p);
        CHECK(canon[states[2 * 4 + p]] == D[p] && canon[jstates[2 * 4 + p]] == D[p], "20 states: leaf D pattern %d", p);
    }
    /* NNI around C against the hand-built neighbour through score_trees */
    const int32_t nid[3] = {1, 1, 3};
    const uint8_t left[3] = {0, 1, 0};
    double nni[3], sc[2];
    OK(phb_score_nni(ctx, 3, nid, left, nni));
    CHECK(isfinite(nni[0]) && isfinite(nni[1]) && isnan(nni[2]), "20 states: NNI values");
    phb_tree_desc cand[2] = {{5, post, off, kids, blen, leaf_row, 0}, {5, post2, off, kids2, blen, leaf_row, 0}};
    OK(phb_score_trees(ctx, 2, cand, sc));
    CHECK(fabs(sc[0] - total) <= 1e-12 * fabs(total), "20 states: score_trees(base) %.17g vs %.17g", sc[0], total);
    CHECK(fabs(sc[1] - nni[0]) <= 1e-10 * fabs(sc[1]) || fabs(sc[1] - nni[1]) <= 1e-10 * fabs(sc[1]),
          "20 states: the hand-built neighbour %.17g must equal one of phb_score_nni's %.17g / %.17g", sc[1], nni[0], nni[1]);
    phb_ctx_destroy(ctx);
    return total;
}

int main(void)
{
    phb_ctx *ctx = NULL;
    int rc = phb_ctx_create(NULL, 0, 0, &ctx);
    if (rc != PHB_OK) {
        fprintf(stderr, "phb_ctx_create -> %d: no usable device (the library has no CPU path)\n", rc);
        return 77; /* skipped */
    }
    printf("%s\n", phb_version());

    /* JC69: pi = 1/4, Q normalised to mean rate 1: eigenvalues 0, -4/3 x3; V = orthonormal basis (column-major) */
    const double lambda[4] = {0.0, -4.0 / 3.0, -4.0 / 3.0, -4.0 / 3.0};
    const double h = 0.5, s2 = 0.70710678118654752440;
    const double V[16] = {h, h, h, h, /* column 0: constant */
                          s2, -s2, 0, 0, 0, 0, s2, -s2, h, h, -h, -h};
    const double pi[4] = {0.25, 0.25, 0.25, 0.25}, sp[4] = {0.5, 0.5, 0.5, 0.5}, isp[4] = {2, 2, 2, 2};
    const double rates[1] = {1.0}, weights[1] = {1.0};

    /* dense ids: E=0 (root), C=1, D=2, A=3, B=4; rows: A=0, B=1, D=2 */
    const int32_t post[5] = {3, 4, 1, 2, 0};
    const int32_t off[6] = {0, 2, 4, 4, 4, 4};
    const int32_t kids[4] = {1, 2, 3, 4};
    const double blen[5] = {0.001, (double)0.100f, (double)0.200f, (double)0.601f, (double)0.301f};
    const int32_t leaf_row[5] = {-1, -1, 2, 0, 1};

    /* --- log_likelihood through device-side compression, as the shim's load_dev does --- */
    const char *A = "ACGTACGT", *B = "AGGTAGGT", *D = "TTTTTTTT";
    const uint8_t *rows[3] = {(const uint8_t *)A, (const uint8_t *)B, (const uint8_t *)D};
    OK(phb_set_model(ctx, lambda, V, pi, sp, isp, 1, rates, weights));
    OK(phb_set_tree(ctx, 5, post, off, kids, blen, leaf_row, 0));
    int64_t n_pat = 0;
    uint64_t s2p[8], mult[8];
    OK(phb_compress_rows(ctx, 3, 8, rows, &n_pat, s2p, mult));
    CHECK(n_pat == 4, "4 distinct columns expected, got %lld", (long long)n_pat);
    for (int i = 0; i < 8; ++i) CHECK(s2p[i] == (uint64_t)(i & 3), "site_to_pattern[%d] = %llu", i, (unsigned long long)s2p[i]);
    for (int i = 0; i < 4; ++i) CHECK(mult[i] == 2, "multiplicity[%d] = %llu", i, (unsigned long long)mult[i]);
    double total = 0, per[4];
    OK(phb_log_likelihood(ctx, &total, per));
    double sum = 0;
    for (int i = 0; i < 4; ++i) sum += 2.0 * per[i];
    CHECK(isfinite(total) && fabs(total - sum) <= 1e-12 * fabs(total), "total %.17g != sum of per-pattern %.17g", total, sum);
    const double lnl_dev = total;

    /* --- the same alignment through host-side compression (shim's load): [pattern][row] bytes + multiplicities --- */
    const uint8_t pats[12] = {'A', 'A', 'T', 'C', 'G', 'T', 'G', 'G', 'T', 'T', 'T', 'T'};
    const uint64_t m2[4] = {2, 2, 2, 2};
    OK(phb_set_alignment(ctx, 3, 4, pats, PHB_LAYOUT_PATTERN_MAJOR, m2));
    OK(phb_set_model(ctx, lambda, V, pi, sp, isp, 1, rates, weights));
    OK(phb_set_tree(ctx, 5, post, off, kids, blen, leaf_row, 0));
    OK(phb_log_likelihood(ctx, &total, NULL));
    CHECK(fabs(total - lnl_dev) <= 1e-12 * fabs(total), "host- and device-compressed totals differ: %.17g vs %.17g", total, lnl_dev);

    /* --- marginal ASR: states for every node, leaves included; leaves reproduce their own bases --- */
    uint8_t states[5 * 4];
    double post_out[5 * 4 * 4], total_asr = 0;
    OK(phb_marginal_asr(ctx, states, post_out, &total_asr));
    CHECK(fabs(total_asr - total) <= 1e-12 * fabs(total), "marginal_asr total %.17g vs %.17g", total_asr, total);
    const char *acgt = "ACGT";
    for (int p = 0; p < 4; ++p) {
        CHECK(acgt[states[3 * 4 + p]] == (char)pats[3 * p + 0], "leaf A pattern %d", p);
        CHECK(acgt[states[4 * 4 + p]] == (char)pats[3 * p + 1], "leaf B pattern %d", p);
        CHECK(acgt[states[2 * 4 + p]] == (char)pats[3 * p + 2], "leaf D pattern %d", p);
        for (int v = 0; v < 5; ++v) {
            double ps = 0;
            for (int i = 0; i < 4; ++i) ps += post_out[(v * 4 + p) * 4 + i];
            CHECK(fabs(ps - 1.0) < 1e-12, "posterior of node %d pattern %d sums to %.17g", v, p, ps);
        }
    }
    /* --- joint ASR --- */
    uint8_t jstates[5 * 4];
    double jtotal = 0;
    OK(phb_joint_asr(ctx, jstates, &jtotal));
    CHECK(isfinite(jtotal) && jtotal <= total + 1e-9, "joint value %.17g must not exceed the log-likelihood %.17g", jtotal, total);

    /* --- NNI neighbourhood: node C (dense 1) is the only eligible one; root and leaves give NaN --- */
    const int32_t nid[4] = {1, 1, 0, 3};
    const uint8_t left[4] = {0, 1, 0, 0};
    double nni[4];
    OK(phb_score_nni(ctx, 4, nid, left, nni));
    CHECK(isfinite(nni[0]) && isfinite(nni[1]), "NNI around C must be finite");
    CHECK(isnan(nni[2]) && isnan(nni[3]), "NNI at the root / a leaf must be NaN");

    /* --- score_trees: the base tree itself and the first NNI neighbour ((A,D),B) rebuilt by hand --- */
    const int32_t kids2[4] = {1, 4, 3, 2}; /* E -> (C, B); C -> (A, D): B and D swapped, branch lengths travel with the nodes */
    const int32_t post2[5] = {3, 2, 1, 4, 0};
    phb_tree_desc cand[2] = {{5, post, off, kids, blen, leaf_row, 0}, {5, post2, off, kids2, blen, leaf_row, 0}};
    double sc[2];
    OK(phb_score_trees(ctx, 2, cand, sc));
    CHECK(fabs(sc[0] - total) <= 1e-12 * fabs(total), "score_trees(base) %.17g vs %.17g", sc[0], total);
    int match = fabs(sc[1] - nni[0]) <= 1e-10 * fabs(sc[1]) || fabs(sc[1] - nni[1]) <= 1e-10 * fabs(sc[1]);
    CHECK(match, "the hand-built neighbour %.17g must equal one of phb_score_nni's %.17g / %.17g", sc[1], nni[0], nni[1]);
    OK(phb_log_likelihood(ctx, &total_asr, NULL));
    CHECK(total_asr == total, "the context's own tree must be untouched by score_trees");

    /* --- error paths keep the reference's variants --- */
    const int32_t no_row[5] = {-1, -1, 2, 0, -1};
    OK(phb_set_tree(ctx, 5, post, off, kids, blen, no_row, 0));
    rc = phb_log_likelihood(ctx, &total_asr, NULL);
    CHECK(rc == PHB_ERR_INVALID_ALIGNMENT, "leaf without a row -> InvalidAlignment, got %d", rc);
    const uint8_t bad[12] = {'A', 'A', 'T', 'C', 'X', 'T', 'G', 'G', 'T', 'T', 'T', 'T'};
    OK(phb_set_alignment(ctx, 3, 4, bad, PHB_LAYOUT_PATTERN_MAJOR, m2));
    OK(phb_set_tree(ctx, 5, post, off, kids, blen, leaf_row, 0));
    rc = phb_log_likelihood(ctx, &total_asr, NULL);
    CHECK(rc == PHB_ERR_ALPHABET_MISMATCH, "unaccepted byte in a leaf row -> AlphabetMismatch, got %d", rc);

    phb_ctx_destroy(ctx);
    const double lnl20 = amino_acid_section(post, off, kids, blen, leaf_row, post2, kids2);
    printf("20-state lnL %.12f\n", lnl20);
    printf("lnL %.12f  joint %.12f  nni %.12f %.12f  -> %s\n", total, jtotal, nni[0], nni[1], failures ? "FAILED" : "ok");
    return failures ? 1 : 0;
}
