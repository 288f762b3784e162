// model.cpp -- alignment, substitution-model eigen-systems and rate categories: the INPUTS of the likelihood
// path, produced on the host exactly once per model change (SURVEY.md section 8, rows a3 and the "model side" of
// 8c).  Nothing here is on the hot path.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <fstream>
#include <sstream>
#include <stdexcept>

#include "b200host.h"

namespace b200 {

void outError(const std::string &msg) {
    fprintf(stderr, "ERROR: %s\n", msg.c_str());
    throw std::runtime_error(msg);
}
void outWarning(const std::string &msg) { fprintf(stderr, "WARNING: %s\n", msg.c_str()); }

// ---------------------------------------------------------------------------------------------------------
// Alignment
// ---------------------------------------------------------------------------------------------------------
int Alignment::getSeqID(const std::string &name) const {
    for (size_t i = 0; i < seq_names.size(); i++)
        if (seq_names[i] == name) return (int)i;
    return -1;
}

// DNA: 0-3 ACGT, 4..17 ambiguity sets with bit mask code-3 over {A=1,C=2,G=4,T=8}, 18 unknown;
// protein: ARNDCQEGHILKMFPSTWYV, 20 B{N,D}, 21 Z{Q,E}, 22 J{I,L}, 23 unknown
int Alignment::encodeState(SeqType t, char ch) {
    char c = (char)toupper((unsigned char)ch);
    if (t == SEQ_DNA) {
        switch (c) {
            case 'A': return 0;
            case 'C': return 1;
            case 'G': return 2;
            case 'T': case 'U': return 3;
            case 'R': return 1 + 4 + 3;
            case 'Y': return 2 + 8 + 3;
            case 'W': return 1 + 8 + 3;
            case 'S': return 2 + 4 + 3;
            case 'M': return 1 + 2 + 3;
            case 'K': return 4 + 8 + 3;
            case 'B': return 2 + 4 + 8 + 3;
            case 'H': return 1 + 2 + 8 + 3;
            case 'D': return 1 + 4 + 8 + 3;
            case 'V': return 1 + 2 + 4 + 3;
            default: return 18;  // N ? - . ~ X O
        }
    }
    if (t == SEQ_PROTEIN) {
        static const char *aa = "ARNDCQEGHILKMFPSTWYV";
        const char *p = strchr(aa, c);
        if (p && c) return (int)(p - aa);
        if (c == 'B') return 20;
        if (c == 'Z') return 21;
        if (c == 'J') return 22;
        return 23;
    }
    if (c >= '0' && c <= '9') return c - '0';
    return -1;
}

void Alignment::computeConst() {
    // Pattern::const_char (alignment/alignment.cpp computeConst): a pattern is constant when the sets of states
    // compatible with every character intersect; const_char is that state (an ambiguity code if the intersection
    // still has several states), STATE_UNKNOWN for an all-gap column, and STATE_UNKNOWN+1 when not constant
    const size_t nptn = getNPattern(), nseq = getNSeq();
    const_char.assign(nptn, STATE_UNKNOWN + 1);
    std::vector<double> lk(num_states);
    for (size_t p = 0; p < nptn; p++) {
        std::vector<char> compat(num_states, 1);
        bool all_unknown = true;
        for (size_t t = 0; t < nseq; t++) {
            const int s = at(p, t);
            if (s == STATE_UNKNOWN) continue;
            all_unknown = false;
            ModelSubst::computeTipLikelihood(seq_type, num_states, STATE_UNKNOWN, s, lk.data());
            for (int x = 0; x < num_states; x++) compat[x] = compat[x] && lk[x] != 0.0;
        }
        if (all_unknown) { const_char[p] = STATE_UNKNOWN; continue; }
        int cnt = 0, first = -1, mask = 0;
        for (int x = 0; x < num_states; x++)
            if (compat[x]) { cnt++; if (first < 0) first = x; mask |= 1 << x; }
        if (cnt == 1) const_char[p] = first;
        else if (cnt > 1) {
            if (seq_type == SEQ_DNA) const_char[p] = mask + 3;
            else if (seq_type == SEQ_PROTEIN) {
                if (mask == (4 | 8)) const_char[p] = 20;
                else if (mask == (32 | 64)) const_char[p] = 21;
                else if (mask == (512 | 1024)) const_char[p] = 22;
            }
        }
    }
}

void Alignment::buildFromStates(const std::vector<std::string> &names, const uint8_t *codes, int64_t nsite_, SeqType t,
                                int nstates, bool compress) {
    seq_type = t;
    num_states = nstates;
    STATE_UNKNOWN = t == SEQ_DNA ? 18 : (t == SEQ_PROTEIN ? 23 : nstates);
    seq_names = names;
    nsite = nsite_;
    const size_t nseq = names.size();
    site_pattern.assign(nsite, 0);
    std::vector<int64_t> rep;  // representative site of each pattern
    std::vector<double> freq;
    if (compress) {
        std::map<std::string, int> seen;
        std::string key(nseq, '\0');
        for (int64_t s = 0; s < nsite; s++) {
            for (size_t q = 0; q < nseq; q++) key[q] = (char)codes[q * nsite + s];
            auto it = seen.find(key);
            if (it == seen.end()) {
                it = seen.emplace(key, (int)rep.size()).first;
                rep.push_back(s);
                freq.push_back(0.0);
            }
            freq[it->second] += 1.0;
            site_pattern[s] = it->second;
        }
    } else {
        rep.resize(nsite);
        freq.assign(nsite, 1.0);
        for (int64_t s = 0; s < nsite; s++) { rep[s] = s; site_pattern[s] = (int32_t)s; }
    }
    const size_t nptn = rep.size();
    states.resize(nseq * nptn);
    for (size_t q = 0; q < nseq; q++)
        for (size_t p = 0; p < nptn; p++) states[q * nptn + p] = codes[q * nsite + rep[p]];
    ptn_freq = freq;
    computeConst();
}

void Alignment::buildFromSequences(const std::vector<std::string> &names, const std::vector<std::string> &seqs, SeqType t,
                                   bool compress) {
    if (names.size() != seqs.size() || names.empty()) outError("alignment: names and sequences do not match");
    const size_t L = seqs[0].size();
    std::vector<uint8_t> codes(names.size() * L);
    for (size_t q = 0; q < names.size(); q++) {
        if (seqs[q].size() != L) outError("Sequence " + names[q] + " has a different length");
        for (size_t s = 0; s < L; s++) {
            int c = encodeState(t, seqs[q][s]);
            if (c < 0) outError(std::string("Invalid character ") + seqs[q][s] + " in sequence " + names[q]);
            codes[q * L + s] = (uint8_t)c;
        }
    }
    buildFromStates(names, codes.data(), (int64_t)L, t, t == SEQ_DNA ? 4 : 20, compress);
}

void Alignment::readPhylip(const std::string &path, SeqType t) {
    std::ifstream in(path.c_str());
    if (!in) outError("File not found " + path);
    size_t nseq = 0, L = 0;
    in >> nseq >> L;
    if (!in || nseq == 0 || L == 0) outError("Bad PHYLIP header in " + path);
    std::string line;
    std::getline(in, line);
    std::vector<std::string> names, seqs(nseq);
    size_t row = 0;
    bool first_block = true;
    while (std::getline(in, line)) {
        if (line.find_first_not_of(" \t\r\n") == std::string::npos) continue;
        std::string body = line;
        if (first_block) {
            std::istringstream ss(line);
            std::string nm;
            ss >> nm;
            names.push_back(nm);
            size_t pos = line.find(nm) + nm.size();
            body = line.substr(pos);
        }
        for (char c : body)
            if (!isspace((unsigned char)c)) seqs[row].push_back(c);
        row++;
        if (row == nseq) { row = 0; first_block = false; }
    }
    if (names.size() != nseq) outError("PHYLIP file ended before all sequence names were read");
    for (size_t q = 0; q < nseq; q++)
        if (seqs[q].size() != L) outError("Sequence " + names[q] + " has wrong length in " + path);
    buildFromSequences(names, seqs, t, true);
}

// ---------------------------------------------------------------------------------------------------------
// ModelSubst
// ---------------------------------------------------------------------------------------------------------
void ModelSubst::computeTipLikelihood(SeqType t, int n, int SU, int state, double *lk) {
    for (int x = 0; x < n; x++) lk[x] = 0.0;
    if (state < n) { lk[state] = 1.0; return; }
    if (state >= SU) { for (int x = 0; x < n; x++) lk[x] = 1.0; return; }
    if (t == SEQ_DNA) {
        const int mask = state - 3;
        for (int x = 0; x < 4; x++)
            if (mask & (1 << x)) lk[x] = 1.0;
        return;
    }
    if (t == SEQ_PROTEIN) {
        static const int amb[3][2] = {{2, 3}, {5, 6}, {9, 10}};  // B = N|D, Z = Q|E, J = I|L
        if (state - 20 < 3) { lk[amb[state - 20][0]] = 1.0; lk[amb[state - 20][1]] = 1.0; return; }
    }
    for (int x = 0; x < n; x++) lk[x] = 1.0;
}

void ModelSubst::getStateFrequency(double *f, int mixture) const {
    const int n = num_states;
    if (mixture >= 0) {
        for (int i = 0; i < n; i++) f[i] = state_freq[(size_t)mixture * n + i];
        return;
    }
    for (int i = 0; i < n; i++) {
        f[i] = 0.0;
        for (int m = 0; m < nmix; m++) f[i] += mixw[m] * state_freq[(size_t)m * n + i];
    }
}

void eigenSymmetric(int n, const double *a_in, double *w, double *v) {
    std::vector<double> a(a_in, a_in + (size_t)n * n);
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) v[i * n + j] = i == j ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 100; sweep++) {
        double off = 0.0;
        for (int i = 0; i < n; i++)
            for (int j = i + 1; j < n; j++) off += a[i * n + j] * a[i * n + j];
        if (off < 1e-300) break;
        for (int p = 0; p < n; p++)
            for (int q = p + 1; q < n; q++) {
                const double apq = a[p * n + q];
                if (fabs(apq) < 1e-300) continue;
                const double theta = (a[q * n + q] - a[p * n + p]) / (2.0 * apq);
                const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < n; k++) {
                    const double akp = a[k * n + p], akq = a[k * n + q];
                    a[k * n + p] = c * akp - s * akq;
                    a[k * n + q] = s * akp + c * akq;
                }
                for (int k = 0; k < n; k++) {
                    const double apk = a[p * n + k], aqk = a[q * n + k];
                    a[p * n + k] = c * apk - s * aqk;
                    a[q * n + k] = s * apk + c * aqk;
                }
                for (int k = 0; k < n; k++) {
                    const double vkp = v[k * n + p], vkq = v[k * n + q];
                    v[k * n + p] = c * vkp - s * vkq;
                    v[k * n + q] = s * vkp + c * vkq;
                }
            }
    }
    // ascending eigenvalues, eigenvectors permuted alongside
    std::vector<int> idx(n);
    for (int i = 0; i < n; i++) idx[i] = i;
    std::sort(idx.begin(), idx.end(), [&](int x, int y) { return a[x * n + x] < a[y * n + y]; });
    std::vector<double> vv(v, v + (size_t)n * n);
    for (int j = 0; j < n; j++) {
        w[j] = a[idx[j] * n + idx[j]];
        for (int k = 0; k < n; k++) v[k * n + j] = vv[k * n + idx[j]];
    }
}

ModelSubst ModelSubst::reversible(int n, const double *exch, const double *freqs, const std::string &nm) {
    ModelSubst m;
    m.num_states = n;
    m.nmix = 1;
    m.name = nm;
    std::vector<double> pi(freqs, freqs + n), Q((size_t)n * n, 0.0), S((size_t)n * n), w(n), V((size_t)n * n);
    double sum = 0.0;
    for (double f : pi) sum += f;
    for (double &f : pi) f /= sum;
    for (int x = 0; x < n; x++) {
        double row = 0.0;
        for (int y = 0; y < n; y++)
            if (x != y) { Q[x * n + y] = exch[x * n + y] * pi[y]; row += Q[x * n + y]; }
        Q[x * n + x] = -row;
    }
    double norm = 0.0;
    for (int x = 0; x < n; x++) norm -= pi[x] * Q[x * n + x];
    for (double &q : Q) q /= norm;
    // S = Pi^{1/2} Q Pi^{-1/2} is symmetric: U = Pi^{-1/2} V, U^-1 = V^T Pi^{1/2}
    std::vector<double> s(n);
    for (int x = 0; x < n; x++) s[x] = sqrt(pi[x]);
    for (int x = 0; x < n; x++)
        for (int y = 0; y < n; y++) S[x * n + y] = Q[x * n + y] * s[x] / s[y];
    for (int x = 0; x < n; x++)
        for (int y = x + 1; y < n; y++) S[x * n + y] = S[y * n + x] = 0.5 * (S[x * n + y] + S[y * n + x]);
    eigenSymmetric(n, S.data(), w.data(), V.data());
    m.eval = w;
    m.evec.resize((size_t)n * n);
    m.inv_evec.resize((size_t)n * n);
    for (int x = 0; x < n; x++)
        for (int i = 0; i < n; i++) {
            m.evec[x * n + i] = V[x * n + i] / s[x];
            m.inv_evec[i * n + x] = V[x * n + i] * s[x];
        }
    m.state_freq = pi;
    m.mixw.assign(1, 1.0);
    return m;
}

ModelSubst ModelSubst::GTR(const double rates[6], const double freqs[4]) {
    double R[16] = {0};
    int k = 0;
    for (int i = 0; i < 4; i++)
        for (int j = i + 1; j < 4; j++) R[i * 4 + j] = R[j * 4 + i] = rates[k++];
    return reversible(4, R, freqs, "GTR");
}

ModelSubst ModelSubst::JC() {
    const double r[6] = {1, 1, 1, 1, 1, 1}, f[4] = {0.25, 0.25, 0.25, 0.25};
    ModelSubst m = GTR(r, f);
    m.name = "JC";
    return m;
}

// ---- the protein models BASELINE configs 3 and 4 name, as data (host/protein_data.inc, generated by
// tools/extract_model_data.py from the published tables: model/modelprotein.cpp:287-310, model/modelmixture.cpp:917-977)
#include "protein_data.inc"

static void lgExchangeabilities(double *R) {  // [20][20] symmetric from the lower triangle, rows 1..19
    for (int i = 0, k = 0; i < 20; i++) {
        R[i * 20 + i] = 0.0;
        for (int j = 0; j < i; j++, k++) R[i * 20 + j] = R[j * 20 + i] = B200_LG_LOWER[k];
    }
}

// `-m LG`: LG exchangeabilities with LG's own frequencies (ModelProtein, model/modelprotein.cpp)
ModelSubst ModelSubst::LG() {
    double R[400];
    lgExchangeabilities(R);
    return reversible(20, R, B200_LG_FREQS, "LG");
}

// `-m LG+C60`: the POISSON of `model C60 = POISSON+G+FMIX{C60pi1:1:w1,...}` replaced by LG -- 60 classes sharing the LG
// exchangeabilities, class k with frequency profile C60pi_k, rate 1 and fixed weight w_k; every class is normalised to
// one expected substitution, and with weights summing to 1 the global rate is 1 (ModelMixture::initMixture,
// model/modelmixture.cpp:1411-1437)
ModelSubst ModelSubst::LGC60() {
    double R[400];
    lgExchangeabilities(R);
    std::vector<ModelSubst> classes;
    std::vector<double> w(B200_C60_WEIGHTS, B200_C60_WEIGHTS + 60);
    for (int k = 0; k < 60; k++) classes.push_back(reversible(20, R, B200_C60_PROFILES + 20 * k, "LG+FC60pi" + std::to_string(k + 1)));
    ModelSubst m = mixture(classes, w);
    m.name = "LG+C60";
    return m;
}

ModelSubst ModelSubst::mixture(const std::vector<ModelSubst> &classes, const std::vector<double> &weights) {
    if (classes.empty() || classes.size() != weights.size()) outError("mixture: classes and weights do not match");
    ModelSubst m;
    m.num_states = classes[0].num_states;
    m.nmix = (int)classes.size();
    m.name = "MIX";
    double sum = 0.0;
    for (double w : weights) sum += w;
    for (size_t c = 0; c < classes.size(); c++) {
        if (classes[c].num_states != m.num_states || classes[c].nmix != 1) outError("mixture: incompatible class");
        m.eval.insert(m.eval.end(), classes[c].eval.begin(), classes[c].eval.end());
        m.evec.insert(m.evec.end(), classes[c].evec.begin(), classes[c].evec.end());
        m.inv_evec.insert(m.inv_evec.end(), classes[c].inv_evec.begin(), classes[c].inv_evec.end());
        m.state_freq.insert(m.state_freq.end(), classes[c].state_freq.begin(), classes[c].state_freq.end());
        m.mixw.push_back(weights[c] / sum);
    }
    return m;
}

ModelSubst ModelSubst::fromEigen(int n, int nmix, const double *eval, const double *evec, const double *inv_evec,
                                 const double *freqs, const double *mixw) {
    ModelSubst m;
    m.num_states = n;
    m.nmix = nmix;
    m.name = "EIGEN";
    m.eval.assign(eval, eval + (size_t)nmix * n);
    m.evec.assign(evec, evec + (size_t)nmix * n * n);
    m.inv_evec.assign(inv_evec, inv_evec + (size_t)nmix * n * n);
    m.state_freq.assign(freqs, freqs + (size_t)nmix * n);
    m.mixw.assign(mixw, mixw + nmix);
    return m;
}

// ---------------------------------------------------------------------------------------------------------
// Rates
// ---------------------------------------------------------------------------------------------------------
double gammaIncP(double a, double x) {
    if (x <= 0.0) return 0.0;
    const double gln = lgamma(a);
    if (x < a + 1.0) {  // series
        double ap = a, sum = 1.0 / a, del = sum;
        for (int i = 0; i < 10000; i++) {
            ap += 1.0;
            del *= x / ap;
            sum += del;
            if (fabs(del) < fabs(sum) * 1e-17) break;
        }
        return sum * exp(-x + a * log(x) - gln);
    }
    // continued fraction (modified Lentz) for Q
    const double tiny = 1e-300;
    double b = x + 1.0 - a, c = 1.0 / tiny, d = 1.0 / b, h = d;
    for (int i = 1; i < 10000; i++) {
        const double an = -i * (i - a);
        b += 2.0;
        d = an * d + b;
        if (fabs(d) < tiny) d = tiny;
        c = b + an / c;
        if (fabs(c) < tiny) c = tiny;
        d = 1.0 / d;
        const double del = d * c;
        h *= del;
        if (fabs(del - 1.0) < 1e-16) break;
    }
    return 1.0 - exp(-x + a * log(x) - gln) * h;
}

double gammaIncPInv(double a, double p) {
    if (p <= 0.0) return 0.0;
    if (p >= 1.0) return INFINITY;
    // bracket, then bisection polished by Newton steps that stay inside the bracket
    double lo = 0.0, hi = a > 1.0 ? a : 1.0;
    while (gammaIncP(a, hi) < p) hi *= 2.0;
    double x = 0.5 * (lo + hi);
    const double gln = lgamma(a);
    for (int it = 0; it < 400; it++) {
        const double f = gammaIncP(a, x) - p;
        if (f > 0) hi = x; else lo = x;
        const double dens = exp(-x + (a - 1.0) * log(x) - gln);
        double xn = dens > 0 ? x - f / dens : 0.5 * (lo + hi);
        if (!(xn > lo && xn < hi)) xn = 0.5 * (lo + hi);
        if (fabs(xn - x) <= 1e-16 * fabs(x)) { x = xn; break; }
        x = xn;
    }
    return x;
}

RateHeterogeneity RateHeterogeneity::uniform() {
    RateHeterogeneity r;
    r.rates.assign(1, 1.0);
    r.props.assign(1, 1.0);
    return r;
}

RateHeterogeneity RateHeterogeneity::gamma(double alpha, int ncat, double p_invar) {
    RateHeterogeneity r;
    r.gamma_shape = alpha;
    r.p_invar = p_invar;
    r.rates.assign(ncat, 1.0);
    r.props.assign(ncat, (1.0 - p_invar) / ncat);
    if (ncat == 1) return r;
    // category means of Gamma(alpha, alpha): cut points at quantiles k/ncat, mean of a slice through the
    // incomplete gamma of shape alpha+1 (model/rategamma.cpp:120-170), then rescaled to mean 1 (:109-119)
    std::vector<double> upper(ncat);
    for (int k = 1; k < ncat; k++) upper[k - 1] = gammaIncP(alpha + 1.0, gammaIncPInv(alpha, (double)k / ncat));
    upper[ncat - 1] = 1.0;
    double prev = 0.0, mean = 0.0;
    for (int k = 0; k < ncat; k++) {
        r.rates[k] = (upper[k] - prev) * ncat;
        prev = upper[k];
        mean += r.rates[k];
    }
    mean /= ncat;
    for (double &x : r.rates) x /= mean;
    return r;
}

RateHeterogeneity RateHeterogeneity::fromCategories(const double *rates, const double *props, int ncat, double p_invar) {
    RateHeterogeneity r;
    r.rates.assign(rates, rates + ncat);
    r.props.assign(props, props + ncat);
    r.p_invar = p_invar;
    return r;
}

}  // namespace b200
