// phylo_models.cpp — alignment container, column compression and substitution-model construction of the
// C++ host mirror (phylo_host.hpp).  Host-side, once-per-call / once-per-model work that stays on the
// reference's side of the boundary; nothing here is on the per-pattern path.
#include "phylo_host.hpp"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <unordered_map>

namespace phylo {

// =====================================================================================================
// Alignment (src/alignment.rs:14-125)
// =====================================================================================================
namespace {
bool ascii_space(unsigned char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\v' || c == '\f'; }
} // namespace

const std::vector<std::uint8_t> *Alignment::get(const std::string &name) const
{
    for (const auto &kv : seqs)
        if (kv.first == name) return &kv.second;
    return nullptr;
}

void Alignment::insert(const std::string &name, std::vector<std::uint8_t> seq)
{
    for (auto &kv : seqs)
        if (kv.first == name) {
            kv.second = std::move(seq);
            return;
        }
    seqs.emplace_back(name, std::move(seq));
}

Alignment Alignment::from_fasta_bytes(std::string_view data)
{
    Alignment aln;
    std::string cur_id;
    std::vector<std::uint8_t> cur_seq;
    auto flush = [&]() {
        if (cur_id.empty()) return;
        if (aln.get(cur_id)) throw AsrError(ErrorKind::InvalidAlignment, "Duplicate taxon ID: " + cur_id);
        aln.seqs.emplace_back(cur_id, std::move(cur_seq));
        cur_seq.clear();
    };
    std::size_t pos = 0;
    while (pos <= data.size()) {
        std::size_t nl = data.find('\n', pos);
        if (nl == std::string_view::npos) nl = data.size();
        std::size_t b = pos, e = nl;
        while (b < e && ascii_space((unsigned char)data[b])) ++b;
        while (e > b && ascii_space((unsigned char)data[e - 1])) --e;
        pos = nl + 1;
        if (b == e) continue;
        if (data[b] == '>') {
            flush();
            std::size_t h = b + 1, he = h;
            while (he < e && !ascii_space((unsigned char)data[he])) ++he; // header = up to the first whitespace
            if (he == h) throw AsrError(ErrorKind::InvalidAlignment, "Empty taxon ID in FASTA header");
            cur_id.assign(data.substr(h, he - h));
            cur_seq.clear();
        } else {
            for (std::size_t k = b; k < e; ++k) {
                unsigned char c = (unsigned char)data[k];
                if (c >= 'a' && c <= 'z') c = (unsigned char)(c - 'a' + 'A'); // normalised to upper case
                cur_seq.push_back(c);
            }
        }
    }
    flush();
    if (aln.seqs.empty()) throw AsrError(ErrorKind::InvalidAlignment, "Empty FASTA file");
    aln.width = aln.seqs.front().second.size();
    for (const auto &kv : aln.seqs)
        if (kv.second.size() != aln.width)
            throw AsrError(ErrorKind::InvalidAlignment, "Ragged alignment: sequences have different lengths");
    return aln;
}

CompressedColumns Alignment::compress_columns() const
{
    // Distinct columns over ALL rows in first-appearance order (alignment.rs:93-125).  The column gather is
    // done through a transposed copy so that each key is contiguous.
    CompressedColumns out;
    const std::size_t rows = seqs.size();
    for (const auto &kv : seqs) out.leaf_order.push_back(kv.first);
    out.site_to_pattern.reserve(width);
    std::string cols(rows * width, '\0');
    for (std::size_t r = 0; r < rows; ++r) {
        const auto &s = seqs[r].second;
        for (std::size_t i = 0; i < width; ++i) cols[i * rows + r] = (char)s[i];
    }
    std::unordered_map<std::string_view, std::size_t> index;
    index.reserve(width);
    for (std::size_t i = 0; i < width; ++i) {
        const std::string_view key(cols.data() + i * rows, rows);
        auto it = index.find(key);
        std::size_t idx;
        if (it == index.end()) {
            idx = out.patterns.size();
            index.emplace(key, idx);
            out.patterns.emplace_back(key.begin(), key.end());
            out.multiplicity.push_back(0);
        } else
            idx = it->second;
        out.multiplicity[idx] += 1;
        out.site_to_pattern.push_back(idx);
    }
    return out;
}

// =====================================================================================================
// discrete gamma (src/models/gamma.rs:15-229): Yang 1994 mean-of-bin rates with the reference's own
// special functions, so that the rates handed to the device agree with the reference to round-off.
// =====================================================================================================
namespace {

double ln_gamma(double x)
{
    // Lanczos, g = 7, 9 coefficients (gamma.rs:100-126)
    static const double C[9] = {0.99999999999980993,  676.5203681218851,     -1259.1392167224028,
                                771.32342877765313,   -176.61502916214059,   12.507343278686905,
                                -0.13857109526572012, 9.9843695780195716e-6, 1.5056327351493116e-7};
    const double pi = 3.14159265358979323846;
    if (x < 0.5) return std::log(pi / std::sin(pi * x)) - ln_gamma(1.0 - x);
    x -= 1.0;
    double a = C[0];
    const double t = x + 7.0 + 0.5;
    for (int i = 1; i < 9; ++i) a += C[i] / (x + (double)i);
    return 0.5 * std::log(2.0 * pi) + (x + 0.5) * std::log(t) - t + std::log(a);
}

double clamp01(double v) { return v < 0.0 ? 0.0 : (v > 1.0 ? 1.0 : v); }

double incomplete_gamma(double x, double shape)
{
    // regularised lower incomplete gamma: series below shape + 1, continued fraction above (gamma.rs:129-176)
    if (x <= 0.0) return 0.0;
    const double lg = ln_gamma(shape);
    if (x < shape + 1.0) {
        double term = 1.0 / shape, sum = term, n = shape;
        for (int it = 0; it < 500; ++it) {
            n += 1.0;
            term *= x / n;
            sum += term;
            if (std::fabs(term) < std::fabs(sum) * 1e-15) break;
        }
        return clamp01(sum * std::exp(-x + shape * std::log(x) - lg));
    }
    double b = x + 1.0 - shape, c = 1e300, d = 1.0 / b, h = d;
    for (int i = 1; i < 500; ++i) {
        const double an = -(double)i * ((double)i - shape);
        b += 2.0;
        d = an * d + b;
        if (std::fabs(d) < 1e-300) d = 1e-300;
        c = b + an / c;
        if (std::fabs(c) < 1e-300) c = 1e-300;
        d = 1.0 / d;
        const double del = d * c;
        h *= del;
        if (std::fabs(del - 1.0) < 1e-15) break;
    }
    return clamp01(1.0 - std::exp(-x + shape * std::log(x) - lg) * h);
}

double inverse_normal_cdf(double p)
{
    // Acklam's rational approximation (gamma.rs:180-229); only seeds the Newton iteration below
    static const double a[6] = {-3.969683028665376e+01, 2.209460984245205e+02, -2.759285104469687e+02,
                                1.383577518672690e+02,  -3.066479806614716e+01, 2.506628277459239e+00};
    static const double b[5] = {-5.447609879822406e+01, 1.615858368580409e+02, -1.556989798598866e+02,
                                6.680131188771972e+01, -1.328068155288572e+01};
    static const double c[6] = {-7.784894002430293e-03, -3.223964580411365e-01, -2.400758277161838e+00,
                                -2.549732539343734e+00, 4.374664141464968e+00,  2.938163982698783e+00};
    static const double d[4] = {7.784695709041462e-03, 3.224671290700398e-01, 2.445134137142996e+00, 3.754408661907416e+00};
    const double lo = 0.02425, hi = 1.0 - lo;
    auto tail = [&](double q) {
        return (((((c[0] * q + c[1]) * q + c[2]) * q + c[3]) * q + c[4]) * q + c[5]) /
               ((((d[0] * q + d[1]) * q + d[2]) * q + d[3]) * q + 1.0);
    };
    if (p < lo) return tail(std::sqrt(-2.0 * std::log(p)));
    if (p <= hi) {
        const double q = p - 0.5, r = q * q;
        return (((((a[0] * r + a[1]) * r + a[2]) * r + a[3]) * r + a[4]) * r + a[5]) * q /
               (((((b[0] * r + b[1]) * r + b[2]) * r + b[3]) * r + b[4]) * r + 1.0);
    }
    return -tail(std::sqrt(-2.0 * std::log(1.0 - p)));
}

double point_gamma(double prob, double shape, double rate)
{
    // quantile of Gamma(shape, rate): Wilson-Hilferty start, <= 100 Newton steps (gamma.rs:59-97)
    if (!(prob >= 0.0 && prob < 1.0))
        throw AsrError(ErrorKind::InvalidModelParameter, "quantile probability must be in [0, 1)");
    const double lg = ln_gamma(shape);
    const double wh = 1.0 - 1.0 / (9.0 * shape) + inverse_normal_cdf(prob) * std::sqrt(1.0 / (9.0 * shape));
    const double ch = shape * (wh * wh * wh);
    double x = ch > 0.0 ? ch / rate : 1e-10;
    for (int it = 0; it < 100; ++it) {
        const double f = incomplete_gamma(x * rate, shape) - prob;
        const double df = std::exp((shape - 1.0) * std::log(x * rate) - x * rate - lg + std::log(rate));
        if (!std::isfinite(df) || std::fabs(df) < 1e-300) break;
        const double xn = x - f / df;
        if (!std::isfinite(xn) || xn <= 0.0) break;
        const bool done = std::fabs(xn - x) < 1e-12 * std::max(x, 1e-12);
        x = xn;
        if (done) break;
    }
    return x;
}

} // namespace

std::vector<double> discrete_gamma(double alpha, std::size_t k)
{
    if (!(alpha > 0.0)) throw AsrError(ErrorKind::InvalidModelParameter, "gamma shape parameter alpha must be positive");
    if (k == 0) throw AsrError(ErrorKind::InvalidModelParameter, "number of gamma categories must be at least 1");
    if (k == 1) return {1.0};
    const double kf = (double)k;
    std::vector<double> cdf{0.0};
    for (std::size_t i = 1; i < k; ++i) {
        const double bound = point_gamma((double)i / kf, alpha, alpha);
        cdf.push_back(incomplete_gamma(bound * alpha, alpha + 1.0));
    }
    cdf.push_back(1.0);
    std::vector<double> rates(k);
    for (std::size_t i = 0; i < k; ++i) rates[i] = kf * (cdf[i + 1] - cdf[i]);
    return rates;
}

// =====================================================================================================
// RateMatrix (src/models/rate_matrix.rs:34-110)
// =====================================================================================================
namespace {

// Cyclic Jacobi for a symmetric 4x4 (the reference calls nalgebra 0.33 `symmetric_eigen`; P(t) is
// invariant to the order and sign of the eigenvectors, so only round-off differs).
void jacobi4(double A[4][4], double lam[4], double V[4][4])
{
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) V[i][j] = i == j ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 64; ++sweep) {
        double off = 0.0;
        for (int i = 0; i < 4; ++i)
            for (int j = i + 1; j < 4; ++j) off += A[i][j] * A[i][j];
        if (off < 1e-300) break;
        for (int p = 0; p < 4; ++p)
            for (int q = p + 1; q < 4; ++q) {
                if (std::fabs(A[p][q]) < 1e-300) continue;
                const double theta = (A[q][q] - A[p][p]) / (2.0 * A[p][q]);
                const double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
                const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < 4; ++k) { // A <- A J
                    const double akp = A[k][p], akq = A[k][q];
                    A[k][p] = c * akp - s * akq;
                    A[k][q] = s * akp + c * akq;
                }
                for (int k = 0; k < 4; ++k) { // A <- J^T A
                    const double apk = A[p][k], aqk = A[q][k];
                    A[p][k] = c * apk - s * aqk;
                    A[q][k] = s * apk + c * aqk;
                }
                for (int k = 0; k < 4; ++k) {
                    const double vkp = V[k][p], vkq = V[k][q];
                    V[k][p] = c * vkp - s * vkq;
                    V[k][q] = s * vkp + c * vkq;
                }
            }
    }
    for (int i = 0; i < 4; ++i) lam[i] = A[i][i];
}

} // namespace

RateMatrix::RateMatrix(std::size_t n_states, std::vector<double> pi, const std::vector<double> &w, bool normalize)
{
    if (n_states != 4) throw AsrError(ErrorKind::AlphabetMismatch, "only the 4-state nucleotide alphabet is on this path");
    if (pi.size() != n_states) throw AsrError(ErrorKind::AlphabetMismatch, "pi length does not match alphabet states");
    if (w.size() != n_states * n_states)
        throw AsrError(ErrorKind::AlphabetMismatch, "W matrix dimensions do not match alphabet states");
    double sum = 0.0;
    for (double x : pi) sum += x;
    if (!(sum > 0.0)) throw AsrError(ErrorKind::InvalidModelParameter, "equilibrium frequencies must sum to a positive value");
    for (int i = 0; i < 4; ++i) pi_[i] = pi[i] / sum;
    double q[4][4];
    for (int i = 0; i < 4; ++i) {
        double row = 0.0;
        for (int j = 0; j < 4; ++j) {
            q[i][j] = i == j ? 0.0 : pi_[j] * w[i * 4 + j];
            row += q[i][j];
        }
        q[i][i] = -row;
    }
    if (normalize) {
        double mu = 0.0;
        for (int i = 0; i < 4; ++i) mu -= pi_[i] * q[i][i];
        if (!(mu > 0.0)) throw AsrError(ErrorKind::NumericalInstability, "numerical instability encountered during ASR scaling");
        for (auto &row : q)
            for (double &x : row) x /= mu;
    }
    for (int i = 0; i < 4; ++i) {
        sqrt_pi_[i] = std::sqrt(pi_[i]);
        inv_sqrt_pi_[i] = 1.0 / sqrt_pi_[i];
    }
    double s[4][4], V[4][4], lam[4];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) s[i][j] = sqrt_pi_[i] * q[i][j] * inv_sqrt_pi_[j];
    for (int i = 0; i < 4; ++i) // the symmetrised generator is symmetric up to round-off; use its symmetric part
        for (int j = i + 1; j < 4; ++j) s[i][j] = s[j][i] = 0.5 * (s[i][j] + s[j][i]);
    jacobi4(s, lam, V);
    for (int k = 0; k < 4; ++k) {
        eigenvalues_[k] = lam[k];
        for (int i = 0; i < 4; ++i) eigenvectors_[i + 4 * k] = V[i][k];
    }
}

// =====================================================================================================
// GtrModel (src/models/gtr.rs)
// =====================================================================================================
namespace {
void validate_positive(double v, const char *name)
{
    if (!(v > 0.0) || !std::isfinite(v))
        throw AsrError(ErrorKind::InvalidModelParameter, std::string(name) + " must be positive and finite");
}
// exchangeabilities in the order AC, AG, AT, CG, CT, GT (state order A, C, G, T)
std::vector<double> symmetric_w(double ac, double ag, double at, double cg, double ct, double gt)
{
    std::vector<double> w(16, 0.0);
    const int pr[6][2] = {{0, 1}, {0, 2}, {0, 3}, {1, 2}, {1, 3}, {2, 3}};
    const double r[6] = {ac, ag, at, cg, ct, gt};
    for (int k = 0; k < 6; ++k) w[pr[k][0] * 4 + pr[k][1]] = w[pr[k][1] * 4 + pr[k][0]] = r[k];
    return w;
}
std::vector<double> ti_tv_w(double kappa, double tv) { return symmetric_w(tv, kappa, tv, tv, kappa, tv); }
std::vector<double> vec4(const std::array<double, 4> &a) { return {a[0], a[1], a[2], a[3]}; }
const char *RATE_NAMES[6] = {"ac", "ag", "at", "cg", "ct", "gt"};
} // namespace

GtrModel::GtrModel(std::vector<double> pi, const std::vector<double> &w, bool normalize)
    : matrix_(Nucleotide::N_STATES, std::move(pi), w, normalize), categories_{{1.0, 1.0}}
{
}

GtrModel GtrModel::jukes_cantor() { return GtrModel(std::vector<double>(4, 0.25), std::vector<double>(16, 1.0), true); }

GtrModel GtrModel::k80(double kappa)
{
    validate_positive(kappa, "kappa");
    return GtrModel(std::vector<double>(4, 0.25), ti_tv_w(kappa, 1.0), true);
}

GtrModel GtrModel::f81(const std::array<double, 4> &pi) { return GtrModel(vec4(pi), std::vector<double>(16, 1.0), true); }

GtrModel GtrModel::hky85(const std::array<double, 4> &pi, double kappa)
{
    validate_positive(kappa, "kappa");
    return GtrModel(vec4(pi), ti_tv_w(kappa, 1.0), true);
}

GtrModel GtrModel::tn93(const std::array<double, 4> &pi, double kappa1, double kappa2)
{
    validate_positive(kappa1, "kappa1");
    validate_positive(kappa2, "kappa2");
    return GtrModel(vec4(pi), symmetric_w(1.0, kappa1, 1.0, 1.0, kappa2, 1.0), true);
}

GtrModel GtrModel::k81(double a, double b, double c)
{
    validate_positive(a, "a");
    validate_positive(b, "b");
    validate_positive(c, "c");
    return GtrModel(std::vector<double>(4, 0.25), symmetric_w(b, a, c, c, a, b), true);
}

GtrModel GtrModel::tim(const std::array<double, 4> &pi, double kappa1, double kappa2, double a, double b)
{
    validate_positive(kappa1, "kappa1");
    validate_positive(kappa2, "kappa2");
    validate_positive(a, "a");
    validate_positive(b, "b");
    return GtrModel(vec4(pi), symmetric_w(a, kappa1, b, b, kappa2, a), true);
}

GtrModel GtrModel::tvm(const std::array<double, 4> &pi, double kappa, double a, double b, double c)
{
    validate_positive(kappa, "kappa");
    validate_positive(a, "a");
    validate_positive(b, "b");
    validate_positive(c, "c");
    // the reference's fourth transversion rate is the geometric mean of the other three (gtr.rs:236-240)
    return GtrModel(vec4(pi), symmetric_w(a, kappa, b, c, kappa, std::cbrt(a * b * c)), true);
}

GtrModel GtrModel::sym(const std::array<double, 6> &r)
{
    for (int k = 0; k < 6; ++k) validate_positive(r[k], RATE_NAMES[k]);
    return GtrModel(std::vector<double>(4, 0.25), symmetric_w(r[0], r[1], r[2], r[3], r[4], r[5]), true);
}

GtrModel GtrModel::gtr(const std::array<double, 4> &pi, const std::array<double, 6> &r)
{
    for (int k = 0; k < 6; ++k) validate_positive(r[k], RATE_NAMES[k]);
    return GtrModel(vec4(pi), symmetric_w(r[0], r[1], r[2], r[3], r[4], r[5]), true);
}

double GtrModel::proportion_invariant() const
{
    double p = 0.0;
    for (const auto &c : categories_)
        if (c.rate == 0.0) p += c.weight;
    return p;
}

GtrModel GtrModel::with_gamma(double alpha, std::size_t k) const
{
    // gtr.rs:85-112: [invariant (if any)] + k equiprobable gamma categories sharing 1 - p_inv
    const double p_inv = proportion_invariant();
    const std::vector<double> rates = discrete_gamma(alpha, k);
    GtrModel out = *this;
    out.categories_.clear();
    if (p_inv > 0.0) out.categories_.push_back({0.0, p_inv});
    for (double r : rates) out.categories_.push_back({r, (1.0 - p_inv) / (double)k});
    return out;
}

GtrModel GtrModel::with_invariant(double p_inv) const
{
    // gtr.rs:117-165
    if (!(p_inv >= 0.0 && p_inv < 1.0))
        throw AsrError(ErrorKind::InvalidModelParameter, "proportion of invariant sites must be in [0, 1)");
    const double old_var = 1.0 - proportion_invariant(), new_var = 1.0 - p_inv;
    GtrModel out = *this;
    out.categories_.clear();
    if (p_inv > 0.0) out.categories_.push_back({0.0, p_inv});
    const double scale = old_var > 0.0 ? new_var / old_var : 0.0;
    for (const auto &c : categories_)
        if (c.rate != 0.0) out.categories_.push_back({c.rate, c.weight * scale});
    bool all_zero = true;
    for (const auto &c : out.categories_) all_zero = all_zero && c.rate == 0.0;
    if (all_zero) out.categories_.push_back({1.0, new_var});
    return out;
}

// =====================================================================================================
// Reconstruction
// =====================================================================================================
std::optional<std::string> Reconstruction::sequence_string(NodeID id) const
{
    auto it = sequences.find(id);
    if (it == sequences.end()) return std::nullopt;
    std::string s;
    s.reserve(it->second.size());
    for (std::size_t idx : it->second) s.push_back((char)Nucleotide::char_of(idx));
    return s;
}

} // namespace phylo
