// ctc_oracle.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// CPU restatement (C++17, single precision, no FMA contraction, glibc libm) of the decode
// algorithms of nanoporetech/fast-ctc-decode v0.3.7.  It exists so that the CUDA path can be
// checked against something that follows the reference's arithmetic operation by operation.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.  The product (libfctc_b200.so) never links or calls it.
//
// The reference itself is Rust and cannot be built in this image (no cargo/rustc), so this
// restatement is pinned instead against every golden vector the reference's own Rust, Python
// and JS tests contain for the path (tests/golden/*.json, tests/test_oracle_golden.py).
// Items whose behaviour the reference leaves open are decided here and documented:
//   * score sort tie order: the reference uses slice::sort_unstable_by (Rust std pdqsort), whose
//     order among equal keys is unspecified; the oracle orders equal scores by ascending node
//     id (what an insertion sort of the node-sorted list yields).  "parity unpinned" for ties.
//   * process aborts (assert!/unwrap panics with panic="abort") become status ORACLE_PANIC.
//
// Build: g++ -O2 -std=c++17 -ffp-contract=off -fno-fast-math -fPIC -shared (see oracle/Makefile).
//
// Reference citations are file:line relative to /root/reference.

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

namespace {

enum Status : int {
    ST_OK = 0,
    ST_RAN_OUT_OF_BEAM = 1,      // src/lib.rs:38,46-48
    ST_INCOMPARABLE = 2,         // src/lib.rs:39,49-51
    ST_INVALID_ENVELOPE = 3,     // src/lib.rs:40,53
    ST_PANIC = 100,              // an assert!/unwrap in the reference would abort the process
};

constexpr int32_t ROOT_NODE = -1;  // src/tree.rs:88

// ---------------------------------------------------------------------------------------------
// Prefix trie ("SuffixTree", src/tree.rs:17-194) with its child table (src/vec2d.rs:5-41).
// Node ids are creation order; the root is id -1 and keeps its children in a separate row.
// ---------------------------------------------------------------------------------------------
template <class Payload>
class LabelTrie {
public:
    explicit LabelTrie(size_t n_labels) : width_(n_labels), root_row_(n_labels, -1) {}

    // src/tree.rs:105-111 — label of a node, -1 standing for None (the root has no label)
    int64_t label_of(int32_t node) const { return node >= 0 ? recs_[size_t(node)].label : -1; }
    int32_t parent_of(int32_t node) const { return recs_[size_t(node)].parent; }
    size_t size() const { return recs_.size(); }

    // src/tree.rs:147-161
    int32_t child(int32_t node, size_t label) const {
        int32_t idx = (node == ROOT_NODE) ? root_row_[label] : table_[size_t(node) * width_ + label];
        return idx >= 0 ? idx : -1;
    }

    // src/tree.rs:125-145 (returns -2 where the reference would assert)
    int32_t add(int32_t parent, size_t label, Payload data) {
        if (label >= width_ || recs_.size() >= size_t(std::numeric_limits<int32_t>::max())) return -2;
        int32_t id = int32_t(recs_.size());
        int32_t& slot = (parent == ROOT_NODE) ? root_row_[label] : table_[size_t(parent) * width_ + label];
        if (slot != -1) return -2;
        slot = id;
        recs_.push_back(Rec{int64_t(label), parent, std::move(data)});
        table_.insert(table_.end(), width_, -1);  // src/vec2d.rs:23-26
        return id;
    }

    // src/tree.rs:163-177
    Payload* payload(int32_t node) {
        return (node >= 0 && size_t(node) < recs_.size()) ? &recs_[size_t(node)].data : nullptr;
    }
    const Payload* payload(int32_t node) const {
        return (node >= 0 && size_t(node) < recs_.size()) ? &recs_[size_t(node)].data : nullptr;
    }

private:
    struct Rec {
        int64_t label;
        int32_t parent;
        Payload data;
    };
    size_t width_;
    std::vector<Rec> recs_;
    std::vector<int32_t> table_;
    std::vector<int32_t> root_row_;
};

// ---------------------------------------------------------------------------------------------
// phred (src/search.rs:31-36).  Returns the Unicode scalar value the reference turns into a
// char; *valid is cleared where char::from_u32(..).unwrap() would panic.
// ---------------------------------------------------------------------------------------------
uint32_t saturating_u32(float v) {  // Rust `as u32`: NaN -> 0, saturates at both ends
    if (!(v == v)) return 0u;
    if (v <= 0.0f) return 0u;
    if (v >= 4294967296.0f) return 0xFFFFFFFFu;
    return uint32_t(v);
}

uint32_t phred_code(float prob, float qscale, float qbias, bool* valid) {
    const float floor_p = 1e-4f;
    float miss = 1.0f - prob;
    float p = (miss < floor_p) ? floor_p : miss;
    float l = log10f(p);
    float q = -10.0f * l;
    q = q * qscale;
    q = q + qbias;
    uint32_t code = saturating_u32(roundf(q)) + 33u;  // release build: wrapping add
    if (valid) *valid = !(code > 0x10FFFFu || (code >= 0xD800u && code <= 0xDFFFu));
    return code;
}

struct Posterior2 {  // strided [T, C] view (src/lib.rs:198: zero-copy numpy view, any strides)
    const float* p;
    int64_t T, C, st, sc;
    float at(int64_t t, int64_t c) const { return p[t * st + c * sc]; }
};
struct Posterior3 {  // strided [T, S, C] view
    const float* p;
    int64_t T, S, C, st, ss, sc;
    float at(int64_t t, int64_t s, int64_t c) const { return p[t * st + s * ss + c * sc]; }
};

struct Decoded {
    std::vector<int32_t> labels;  // alphabet indices (blank = 0 never appears)
    std::vector<uint32_t> quals;  // phred char codes, one per label (only if qstring)
    std::vector<int64_t> path;
};

// ---------------------------------------------------------------------------------------------
// viterbi_search (src/search.rs:320-383) with find_max (src/search.rs:303-318)
// ---------------------------------------------------------------------------------------------
int viterbi(const Posterior2& m, bool qstring, float qscale, float qbias, bool collapse, Decoded* out) {
    if (m.C <= 0 || m.T <= 0) return ST_PANIC;  // src/search.rs:328-329
    int64_t prev = -1;  // None
    int64_t run_n = 0;
    float run_sum = 0.0f;
    bool ok_char = true;
    for (int64_t t = 0; t < m.T; ++t) {
        int64_t arg = 0;
        float top = m.at(t, 0);
        for (int64_t c = 1; c < m.C; ++c) {
            float v = m.at(t, c);
            if (v > top) { top = v; arg = c; }  // strict: ties keep the lower index, NaN never wins
        }
        if (arg != 0 && (!collapse || prev != arg)) {
            if (run_n > 0) {
                bool v; out->quals.push_back(phred_code(run_sum / float(run_n), qscale, qbias, &v));
                ok_char = ok_char && v;
                run_sum = 0.0f; run_n = 0;
            }
            out->labels.push_back(int32_t(arg));
            out->path.push_back(t);
        }
        if (arg != 0) { run_sum += top; run_n += 1; }
        prev = arg;
    }
    if (run_n > 0) {
        bool v; out->quals.push_back(phred_code(run_sum / float(run_n), qscale, qbias, &v));
        ok_char = ok_char && v;
    }
    if (!qstring) out->quals.clear();
    return ok_char ? ST_OK : ST_PANIC;
}

// ---------------------------------------------------------------------------------------------
// 1-D beam search (src/search.rs:159-301) and its crf twin (src/search.rs:38-157)
// ---------------------------------------------------------------------------------------------
struct Hyp {  // src/search.rs:7-28
    int32_t node;
    int64_t state;
    float label_prob;
    float gap_prob;
    float total() const { return label_prob + gap_prob; }
};

// Merge, rank, cut and renormalise one row's candidates: src/search.rs:244-282 (= :104-142).
int settle_row(std::vector<Hyp>& cand, size_t beam_size) {
    std::stable_sort(cand.begin(), cand.end(), [](const Hyp& a, const Hyp& b) { return a.node < b.node; });
    size_t w = 0;
    for (size_t i = 0; i < cand.size(); ++i) {
        if (w > 0 && cand[w - 1].node == cand[i].node) {
            cand[w - 1].label_prob += cand[i].label_prob;
            cand[w - 1].gap_prob += cand[i].gap_prob;
        } else {
            cand[w++] = cand[i];
        }
    }
    cand.resize(w);
    // sort_unstable_by(partial_cmp): a comparison sort of n >= 2 items looks at every item at
    // least once, so any NaN score is seen (IncomparableValues); with n < 2 nothing is compared.
    if (cand.size() >= 2)
        for (const Hyp& h : cand) { float s = h.total(); if (!(s == s)) return ST_INCOMPARABLE; }
    std::stable_sort(cand.begin(), cand.end(), [](const Hyp& a, const Hyp& b) { return a.total() > b.total(); });
    if (cand.size() > beam_size) cand.resize(beam_size);
    if (cand.empty()) return ST_RAN_OUT_OF_BEAM;
    float top = cand[0].total();
    for (Hyp& h : cand) { h.label_prob /= top; h.gap_prob /= top; }
    return ST_OK;
}

void trace_back(const LabelTrie<int64_t>& trie, int32_t node, Decoded* out) {  // src/search.rs:285-300
    for (int32_t n = node; n >= 0; n = trie.parent_of(n)) {
        out->labels.push_back(int32_t(trie.label_of(n)) + 1);
        out->path.push_back(*trie.payload(n));
    }
    std::reverse(out->labels.begin(), out->labels.end());
    std::reverse(out->path.begin(), out->path.end());
}

int beam1d(const Posterior2& m, size_t beam_size, float thr, bool collapse, Decoded* out, int64_t* n_nodes) {
    const size_t n_lab = size_t(m.C - 1);
    LabelTrie<int64_t> trie(n_lab);
    std::vector<Hyp> beam{Hyp{ROOT_NODE, 0, 0.0f, 1.0f}}, next;
    for (int64_t t = 0; t < m.T; ++t) {
        next.clear();
        const float p_blank = m.at(t, 0);
        for (const Hyp& h : beam) {
            const int64_t tip = trie.label_of(h.node);
            if (p_blank > thr) next.push_back(Hyp{h.node, h.state, 0.0f, (h.label_prob + h.gap_prob) * p_blank});
            for (size_t l = 0; l < n_lab; ++l) {
                const float p = m.at(t, int64_t(l) + 1);
                if (p < thr) continue;
                if (collapse && tip == int64_t(l)) {
                    next.push_back(Hyp{h.node, h.state, h.label_prob * p, 0.0f});
                    int32_t kid = trie.child(h.node, l);
                    if (kid < 0 && h.gap_prob > 0.0f) { kid = trie.add(h.node, l, t); if (kid < 0) return ST_PANIC; }
                    if (kid >= 0) next.push_back(Hyp{kid, h.state, h.gap_prob * p, 0.0f});
                } else {
                    int32_t kid = trie.child(h.node, l);
                    if (kid < 0) { kid = trie.add(h.node, l, t); if (kid < 0) return ST_PANIC; }
                    next.push_back(Hyp{kid, h.state, (h.label_prob + h.gap_prob) * p, 0.0f});
                }
            }
        }
        beam.swap(next);
        int st = settle_row(beam, beam_size);
        if (st != ST_OK) return st;
    }
    if (n_nodes) *n_nodes = int64_t(trie.size());
    trace_back(trie, beam[0].node, out);
    return ST_OK;
}

// ndarray-stats 0.5 QuantileExt::argmax / max: first maximum wins, NaN -> Err -> unwrap panics.
int first_argmax(const float* v, int64_t n, int64_t stride, int64_t* arg, float* val) {
    if (n <= 0) return ST_PANIC;
    int64_t a = 0; float best = v[0];
    if (!(best == best)) return ST_PANIC;  // indexed_iter compares element 0 with itself: NaN -> Err
    for (int64_t i = 1; i < n; ++i) {
        float x = v[i * stride];
        if (!(x == x) || !(best == best)) return ST_PANIC;
        if (x > best) { best = x; a = i; }
    }
    *arg = a; *val = best;
    return ST_OK;
}

int crf_beam1d(const Posterior3& m, const float* init, int64_t n_init, size_t beam_size, float thr, Decoded* out) {
    if (m.C <= 0 || m.T <= 0 || m.S <= 0) return ST_PANIC;  // src/search.rs:45-48
    const int64_t n_state = m.S, n_base = m.C - 1;
    LabelTrie<int64_t> trie{size_t(n_base)};
    int64_t s0; float top0;
    if (first_argmax(init, n_init, 1, &s0, &top0) != ST_OK) return ST_PANIC;
    std::vector<Hyp> beam{Hyp{ROOT_NODE, s0, top0, init[0]}}, next;  // src/search.rs:54-59
    for (int64_t t = 0; t < m.T; ++t) {
        next.clear();
        for (const Hyp& h : beam) {
            if (h.state < 0 || h.state >= n_state) return ST_PANIC;  // slice out of bounds
            const float p_blank = m.at(t, h.state, 0);
            if (p_blank > thr) next.push_back(Hyp{h.node, h.state, 0.0f, (h.label_prob + h.gap_prob) * p_blank});
            for (int64_t l = 0; l < n_base; ++l) {
                const float p = m.at(t, h.state, l + 1);
                if (p < thr) continue;
                int32_t kid = trie.child(h.node, size_t(l));
                if (kid < 0) { kid = trie.add(h.node, size_t(l), t); if (kid < 0) return ST_PANIC; }
                next.push_back(Hyp{kid, (h.state * n_base) % n_state + l, (h.label_prob + h.gap_prob) * p, 0.0f});
            }
        }
        beam.swap(next);
        int st = settle_row(beam, beam_size);
        if (st != ST_OK) return st;
    }
    trace_back(trie, beam[0].node, out);
    return ST_OK;
}

// crf_greedy_search (src/search.rs:385-423)
int crf_greedy(const Posterior3& m, const float* init, int64_t n_init, bool qstring, float qscale, float qbias,
               Decoded* out) {
    if (m.C <= 0 || m.T <= 0 || m.S <= 0) return ST_PANIC;
    const int32_t n_state = int32_t(m.S), n_base = int32_t(m.C) - 1;
    int64_t s0; float dummy;
    if (first_argmax(init, n_init, 1, &s0, &dummy) != ST_OK) return ST_PANIC;
    int32_t state = int32_t(s0);
    bool ok_char = true;
    for (int64_t t = 0; t < m.T; ++t) {
        if (state < 0 || state >= n_state) return ST_PANIC;
        int64_t arg; float top;
        if (first_argmax(m.p + t * m.st + int64_t(state) * m.ss, m.C, m.sc, &arg, &top) != ST_OK) return ST_PANIC;
        if (arg > 0) {
            out->path.push_back(t);
            out->labels.push_back(int32_t(arg));
            bool v; out->quals.push_back(phred_code(top, qscale, qbias, &v));
            ok_char = ok_char && v;
            state = (state * n_base) % n_state + (int32_t(arg) - 1);
        }
    }
    if (!qstring) out->quals.clear();
    return ok_char ? ST_OK : ST_PANIC;
}

// ---------------------------------------------------------------------------------------------
// Log-domain scalar (src/duplex.rs:7-80).  exp_mode 0 = exact expf (--no-default-features),
// 1 = the shipped default feature set, whose fastexp() is transliterated below.
// ---------------------------------------------------------------------------------------------
// src/fastexp.rs:29-62, operation by operation.  On a little-endian machine the f32 read back
// from the union is the low half of an i64 that was just shifted left by 52 bits, i.e. +0.0, so
// the product is (+/-)0.0 for every finite argument.
float fastexp_as_shipped(float a) {
    if (a > -500.0f) {
        float x = 1.442695041f * a;
        union { int64_t i; float f; } u;
        u.i = int64_t(x);  // callers only pass small - big <= 0, in range
        x -= float(u.i);
        float f2 = x, xt = x;
        u.i += 1023;
        u.i = int64_t(uint64_t(u.i) << 52);
        f2 *= 0.006935931f; xt += 4.831794110f; f2 += 0.019890581f; xt *= x;
        f2 *= x; f2 += 0.143440676f; f2 *= xt; f2 += 1.0f;
        return u.f * f2;
    }
    return 0.0f;
}

struct LogMath { int exp_mode; };
thread_local int g_exp_mode = 0;

struct Lg {
    float v;
    static Lg of(float x) { return Lg{logf(x)}; }
    static Lg zero() { return Lg{-std::numeric_limits<float>::infinity()}; }
    static Lg one() { return Lg{0.0f}; }
    Lg hi(Lg o) const { return v < o.v ? o : *this; }  // src/duplex.rs:33-39
};
inline float lg_exp(float a) { return g_exp_mode == 0 ? expf(a) : fastexp_as_shipped(a); }
inline float lg_add_ordered(float big, float small) {  // src/duplex.rs:45-52
    if (small == -std::numeric_limits<float>::infinity()) return big;
    return big + log1pf(lg_exp(small - big));
}
inline Lg operator+(Lg a, Lg b) {  // src/duplex.rs:57-61
    return (a.v <= b.v) ? Lg{lg_add_ordered(b.v, a.v)} : Lg{lg_add_ordered(a.v, b.v)};
}
inline Lg operator*(Lg a, Lg b) { return Lg{a.v + b.v}; }  // src/duplex.rs:71-73

struct Pair {  // src/duplex.rs:84-128
    Lg label, gap;
    static Pair zero() { return Pair{Lg::zero(), Lg::zero()}; }
    Lg total() const { return label + gap; }
};

struct Secondary {  // src/duplex.rs:152-210
    int64_t offset = 0;
    std::vector<Pair> probs;
    Lg max_prob = Lg::zero();

    Pair get(int64_t at) const {  // :167-179
        int64_t i = at - offset;
        if (i < 0 || size_t(i) >= probs.size()) return Pair::zero();
        return probs[size_t(i)];
    }
    void discard_until(int64_t keep_from) {  // :181-191
        if (keep_from > offset) {
            size_t first = size_t(keep_from - offset);
            if (first < probs.size()) probs.erase(probs.begin(), probs.begin() + std::ptrdiff_t(first));
            else probs.clear();
            offset = keep_from;
        }
    }
    void update_max(int64_t lo, int64_t hi) {  // :193-204
        int64_t len = int64_t(probs.size());
        // lo/hi may be the extreme isize values in the reference's own test; avoid overflow
        auto rel = [&](int64_t b) -> int64_t {
            __int128 d = __int128(b) - __int128(offset);
            if (d < 0) return 0;
            if (d > len) return len;
            return int64_t(d);
        };
        int64_t b = rel(lo);
        int64_t e = std::max(b, rel(hi));
        Lg mx = Lg::zero();
        for (int64_t i = b; i < e; ++i) mx = mx.hi(probs[size_t(i)].total());
        max_prob = mx;
    }
    int64_t end() const { return offset + int64_t(probs.size()); }  // :206-209
};

// One step of the network-2 forward recurrence shared by build_/extend_secondary_probs
// (src/duplex.rs:234-246 and :373-385; crf: :274-285, :323-334).
inline Pair secondary_step(Pair last, Pair parent_prev, Lg p_blank, Lg p_label, bool repeat) {
    Lg gap = last.total() * p_blank;
    Lg lab = repeat ? p_label * (last.label + parent_prev.gap) : p_label * (last.label + parent_prev.total());
    return Pair{lab, gap};
}

struct LogPosterior2 {  // network output mapped to log space once (src/duplex.rs:452-453)
    std::vector<float> v; int64_t T, C;
    Lg at(int64_t t, int64_t c) const { return Lg{v[size_t(t * C + c)]}; }
};
struct LogPosterior3 {
    std::vector<float> v; int64_t T, S, C;
    Lg at(int64_t t, int64_t s, int64_t c) const { return Lg{v[size_t((t * S + s) * C + c)]}; }
};

// src/duplex.rs:212-249
Secondary build_secondary(const LogPosterior2& n2, const Secondary& parent, int64_t label, bool repeat, int64_t lo,
                          int64_t hi) {
    Secondary s; s.offset = lo; s.probs.reserve(size_t(hi - lo));
    Pair last = Pair::zero();
    for (int64_t i = lo; i < hi; ++i) {
        last = secondary_step(last, parent.get(i - 1), n2.at(i, 0), n2.at(i, label + 1), repeat);
        s.probs.push_back(last);
        s.max_prob = s.max_prob.hi(last.total());
    }
    return s;
}

// src/duplex.rs:338-387
int64_t extend_secondary(Secondary& s, const LogPosterior2& n2, const Secondary& parent, int64_t label, bool repeat,
                         int64_t lo, int64_t hi) {
    if (lo > s.offset) {
        s.discard_until(lo - 1);
        if (s.probs.empty()) s.offset = lo;
        s.update_max(lo, hi);
    }
    int64_t cur = s.end();
    Pair last = s.probs.empty() ? Pair::zero() : s.probs.back();
    for (int64_t i = cur; i < hi; ++i) {
        last = secondary_step(last, parent.get(i - 1), n2.at(i, 0), n2.at(i, label + 1), repeat);
        s.probs.push_back(last);
        s.max_prob = s.max_prob.hi(last.total());
    }
    return hi > cur ? hi - cur : 0;  // recurrence cells evaluated (work counter for the benchmark)
}

// src/duplex.rs:251-288 and :290-336 (state-conditioned; no repeat case)
Secondary crf_build_secondary(const LogPosterior3& n2, const Secondary& parent, int64_t label, int64_t tstate,
                              int64_t lo, int64_t hi) {
    Secondary s; s.offset = lo; s.probs.reserve(size_t(hi - lo));
    Pair last = Pair::zero();
    for (int64_t i = lo; i < hi; ++i) {
        last = secondary_step(last, parent.get(i - 1), n2.at(i, tstate, 0), n2.at(i, tstate, label + 1), false);
        s.probs.push_back(last);
        s.max_prob = s.max_prob.hi(last.total());
    }
    return s;
}
void crf_extend_secondary(Secondary& s, const LogPosterior3& n2, const Secondary& parent, int64_t label,
                          int64_t tstate, int64_t lo, int64_t hi) {
    if (lo > s.offset) {
        s.discard_until(lo - 1);
        if (s.probs.empty()) s.offset = lo;
        s.update_max(lo, hi);
    }
    int64_t cur = s.end();
    Pair last = s.probs.empty() ? Pair::zero() : s.probs.back();
    for (int64_t i = cur; i < hi; ++i) {
        last = secondary_step(last, parent.get(i - 1), n2.at(i, tstate, 0), n2.at(i, tstate, label + 1), false);
        s.probs.push_back(last);
        s.max_prob = s.max_prob.hi(last.total());
    }
}

struct Hyp2 {  // src/duplex.rs:132-149
    int32_t node;
    int64_t state;
    Pair p1;
    Lg p2max;
    Lg total() const { return p1.total() * p2max; }
};

// merge / refresh / rank / cut: src/duplex.rs:597-635 (= :784-822)
int settle_row2(std::vector<Hyp2>& cand, const LabelTrie<Secondary>& trie, size_t beam_size) {
    std::stable_sort(cand.begin(), cand.end(), [](const Hyp2& a, const Hyp2& b) { return a.node < b.node; });
    size_t w = 0;
    for (size_t i = 0; i < cand.size(); ++i) {
        if (w > 0 && cand[w - 1].node == cand[i].node) {
            cand[w - 1].p1.label = cand[w - 1].p1.label + cand[i].p1.label;
            cand[w - 1].p1.gap = cand[w - 1].p1.gap + cand[i].p1.gap;
        } else {
            cand[w++] = cand[i];
        }
    }
    cand.resize(w);
    for (Hyp2& h : cand)
        if (const Secondary* s = trie.payload(h.node)) h.p2max = s->max_prob;
    std::vector<float> score(cand.size());
    for (size_t i = 0; i < cand.size(); ++i) score[i] = cand[i].total().v;
    if (cand.size() >= 2)
        for (float s : score) if (!(s == s)) return ST_INCOMPARABLE;
    std::vector<size_t> order(cand.size());
    for (size_t i = 0; i < order.size(); ++i) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return score[a] > score[b]; });
    std::vector<Hyp2> ranked; ranked.reserve(std::min(beam_size, cand.size()));
    for (size_t i = 0; i < order.size() && i < beam_size; ++i) ranked.push_back(cand[order[i]]);
    cand.swap(ranked);
    return cand.empty() ? ST_RAN_OUT_OF_BEAM : ST_OK;
}

// duplex::beam_search (src/duplex.rs:443-650)
int duplex2d(const Posterior2& a, const Posterior2& b, const uint64_t* env, int64_t env_st, int64_t env_sc,
             size_t beam_size, float thr_real, bool collapse, std::vector<int32_t>* labels, int64_t* cells) {
    LogPosterior2 n1{{}, a.T, a.C}, n2{{}, b.T, b.C};
    n1.v.resize(size_t(a.T * a.C)); n2.v.resize(size_t(b.T * b.C));
    for (int64_t t = 0; t < a.T; ++t) for (int64_t c = 0; c < a.C; ++c) n1.v[size_t(t * a.C + c)] = logf(a.at(t, c));
    for (int64_t t = 0; t < b.T; ++t) for (int64_t c = 0; c < b.C; ++c) n2.v[size_t(t * b.C + c)] = logf(b.at(t, c));
    const Lg thr = Lg::of(thr_real);
    const int64_t n_lab = a.C - 1;
    LabelTrie<Secondary> trie{size_t(n_lab)};
    std::vector<Hyp2> beam{Hyp2{ROOT_NODE, 0, Pair{Lg::zero(), Lg::one()}, Lg::one()}}, next;

    // root_probs (src/duplex.rs:389-409) over envelope[(0,1)] rows; indexing envelope[(0,1)] on an
    // empty envelope panics, and the slice panics if the bound exceeds network 2's length.
    if (a.T <= 0) return ST_PANIC;
    Secondary root; root.offset = -1; root.max_prob = Lg::one();
    {
        uint64_t ub0 = env[0 * env_st + 1 * env_sc];
        if (ub0 > uint64_t(b.T)) return ST_PANIC;
        Lg cur = Lg::one();
        root.probs.push_back(Pair{Lg::zero(), cur});
        for (uint64_t i = 0; i < ub0; ++i) { cur = cur * n2.at(int64_t(i), 0); root.probs.push_back(Pair{Lg::zero(), cur}); }
    }
    int64_t last_hi = 0;
    int64_t n_cells = 0;
    for (int64_t t = 0; t < a.T; ++t) {
        next.clear();
        uint64_t lo_u = env[t * env_st], hi_u = env[t * env_st + env_sc];
        int64_t lo = lo_u > uint64_t(std::numeric_limits<int64_t>::max()) ? std::numeric_limits<int64_t>::max() : int64_t(lo_u);
        int64_t hi = hi_u > uint64_t(b.T) ? b.T : int64_t(hi_u);
        if (lo >= hi || lo > last_hi) return ST_INVALID_ENVELOPE;
        if (hi > last_hi) {
            std::stable_sort(beam.begin(), beam.end(), [](const Hyp2& x, const Hyp2& y) { return x.node < y.node; });
            for (const Hyp2& h : beam) {
                if (h.node < 0) continue;
                int32_t par = trie.parent_of(h.node);
                int64_t lab = trie.label_of(h.node);
                const Secondary* pp = trie.payload(par);
                if (!pp) pp = &root;
                n_cells += extend_secondary(*trie.payload(h.node), n2, *pp, lab, trie.label_of(par) == lab, lo, hi);
            }
        }
        last_hi = hi;
        const Lg p_blank = n1.at(t, 0);
        for (const Hyp2& tip : beam) {
            const int64_t tip_label = trie.label_of(tip.node);
            if (p_blank.v > thr.v) {
                Hyp2 c = tip; c.p1 = Pair{Lg::zero(), tip.p1.total() * p_blank};
                next.push_back(c);
            }
            for (int64_t l = 0; l < n_lab; ++l) {
                const Lg p = n1.at(t, l + 1);
                if (p.v < thr.v) continue;
                const Secondary* src = trie.payload(tip.node);
                if (!src) src = &root;
                if (collapse && tip_label == l) {
                    Hyp2 c = tip; c.p1 = Pair{tip.p1.label * p, Lg::zero()};
                    next.push_back(c);
                    int32_t kid = trie.child(tip.node, size_t(l));
                    if (kid < 0 && tip.p1.gap.v > Lg::zero().v) {
                        Secondary s = build_secondary(n2, *src, l, true, lo, hi);
                        n_cells += hi - lo;
                        kid = trie.add(tip.node, size_t(l), std::move(s));
                        if (kid < 0) return ST_PANIC;
                    }
                    if (kid >= 0) {
                        Hyp2 k = tip; k.node = kid; k.p1 = Pair{tip.p1.gap * p, Lg::zero()};
                        next.push_back(k);
                    }
                } else {
                    int32_t kid = trie.child(tip.node, size_t(l));
                    if (kid < 0) {
                        Secondary s = build_secondary(n2, *src, l, false, lo, hi);
                        n_cells += hi - lo;
                        kid = trie.add(tip.node, size_t(l), std::move(s));
                        if (kid < 0) return ST_PANIC;
                    }
                    Hyp2 k = tip; k.node = kid; k.p1 = Pair{tip.p1.total() * p, Lg::zero()};
                    next.push_back(k);
                }
            }
        }
        beam.swap(next);
        int st = settle_row2(beam, trie, beam_size);
        if (st != ST_OK) return st;
    }
    for (int32_t n = beam[0].node; n >= 0; n = trie.parent_of(n)) labels->push_back(int32_t(trie.label_of(n)) + 1);
    std::reverse(labels->begin(), labels->end());
    if (cells) *cells = n_cells;
    return ST_OK;
}

// duplex::crf_beam_search (src/duplex.rs:652-834).  No golden vector exists for it in the
// reference: parity unpinned.
int crf_duplex2d(const Posterior3& a, const float* init1, int64_t n_init1, const Posterior3& b, const float* init2,
                 int64_t n_init2, const uint64_t* env, int64_t env_st, int64_t env_sc, size_t beam_size,
                 float thr_real, std::vector<int32_t>* labels) {
    LogPosterior3 n1{{}, a.T, a.S, a.C}, n2{{}, b.T, b.S, b.C};
    n1.v.resize(size_t(a.T * a.S * a.C)); n2.v.resize(size_t(b.T * b.S * b.C));
    for (int64_t t = 0; t < a.T; ++t) for (int64_t s = 0; s < a.S; ++s) for (int64_t c = 0; c < a.C; ++c)
        n1.v[size_t((t * a.S + s) * a.C + c)] = logf(a.at(t, s, c));
    for (int64_t t = 0; t < b.T; ++t) for (int64_t s = 0; s < b.S; ++s) for (int64_t c = 0; c < b.C; ++c)
        n2.v[size_t((t * b.S + s) * b.C + c)] = logf(b.at(t, s, c));
    const Lg thr = Lg::of(thr_real);
    const int64_t n_state = a.S, n_base = a.C - 1;
    LabelTrie<Secondary> trie{size_t(n_base)};
    int64_t s1, s2; float dummy;
    if (first_argmax(init1, n_init1, 1, &s1, &dummy) != ST_OK) return ST_PANIC;
    if (first_argmax(init2, n_init2, 1, &s2, &dummy) != ST_OK) return ST_PANIC;
    std::vector<Hyp2> beam{Hyp2{ROOT_NODE, s1, Pair{Lg::zero(), Lg::one()}, Lg::one()}}, next;
    if (a.T <= 0) return ST_PANIC;
    Secondary root; root.offset = -1; root.max_prob = Lg::one();
    {   // crf_root_probs, src/duplex.rs:411-441
        uint64_t ub0 = env[1 * env_sc];
        if (ub0 > uint64_t(b.T)) return ST_PANIC;
        Lg cur = Lg::one();
        root.probs.push_back(Pair{Lg::zero(), cur});
        int64_t st = s2;
        for (uint64_t i = 0; i < ub0; ++i) {
            if (st < 0 || st >= n_state) return ST_PANIC;
            cur = cur * n2.at(int64_t(i), st, 0);
            root.probs.push_back(Pair{Lg::zero(), cur});
            st = (st * n_base) % n_state;
        }
    }
    int64_t last_hi = 0;
    for (int64_t t = 0; t < a.T; ++t) {
        next.clear();
        uint64_t lo_u = env[t * env_st], hi_u = env[t * env_st + env_sc];
        int64_t lo = lo_u > uint64_t(std::numeric_limits<int64_t>::max()) ? std::numeric_limits<int64_t>::max() : int64_t(lo_u);
        int64_t hi = hi_u > uint64_t(b.T) ? b.T : int64_t(hi_u);
        if (lo >= hi || lo > last_hi) return ST_INVALID_ENVELOPE;
        if (hi > last_hi) {
            std::stable_sort(beam.begin(), beam.end(), [](const Hyp2& x, const Hyp2& y) { return x.node < y.node; });
            for (const Hyp2& h : beam) {
                if (h.node < 0) continue;
                if (h.state < 0 || h.state >= n_state) return ST_PANIC;
                int32_t par = trie.parent_of(h.node);
                const Secondary* pp = trie.payload(par);
                if (!pp) pp = &root;
                crf_extend_secondary(*trie.payload(h.node), n2, *pp, trie.label_of(h.node), h.state, lo, hi);
            }
        }
        last_hi = hi;
        for (const Hyp2& tip : beam) {
            if (tip.state < 0 || tip.state >= n_state) return ST_PANIC;
            const Lg p_blank = n1.at(t, tip.state, 0);
            if (p_blank.v > thr.v) {
                Hyp2 c = tip; c.p1 = Pair{Lg::zero(), tip.p1.total() * p_blank};
                next.push_back(c);
            }
            for (int64_t l = 0; l < n_base; ++l) {
                const Lg p = n1.at(t, tip.state, l + 1);
                if (p.v < thr.v) continue;
                int32_t kid = trie.child(tip.node, size_t(l));
                if (kid < 0) {
                    const Secondary* src = trie.payload(tip.node);
                    if (!src) src = &root;
                    Secondary s = crf_build_secondary(n2, *src, l, tip.state, lo, hi);
                    kid = trie.add(tip.node, size_t(l), std::move(s));
                    if (kid < 0) return ST_PANIC;
                }
                Hyp2 k = tip; k.node = kid; k.state = (tip.state * n_base) % n_state + l;
                k.p1 = Pair{tip.p1.total() * p, Lg::zero()};
                next.push_back(k);
            }
        }
        beam.swap(next);
        int st = settle_row2(beam, trie, beam_size);
        if (st != ST_OK) return st;
    }
    for (int32_t n = beam[0].node; n >= 0; n = trie.parent_of(n)) labels->push_back(int32_t(trie.label_of(n)) + 1);
    std::reverse(labels->begin(), labels->end());
    return ST_OK;
}

template <class F>
void parallel_reads(int64_t n, int n_threads, F&& body) {
    if (n_threads <= 0) n_threads = int(std::thread::hardware_concurrency());
    if (n_threads <= 0) n_threads = 1;
    if (int64_t(n_threads) > n) n_threads = int(std::max<int64_t>(n, 1));
    std::atomic<int64_t> next{0};
    auto worker = [&]() {
        for (;;) {
            int64_t i = next.fetch_add(1);
            if (i >= n) break;
            body(i);
        }
    };
    std::vector<std::thread> pool;
    for (int i = 1; i < n_threads; ++i) pool.emplace_back(worker);
    worker();
    for (auto& th : pool) th.join();
}

int64_t copy_out(const Decoded& d, int32_t* labels, uint32_t* quals, int64_t* path, int64_t cap) {
    int64_t n = int64_t(d.labels.size());
    int64_t m = std::min(n, cap);
    if (labels) std::copy_n(d.labels.begin(), m, labels);
    if (path) std::copy_n(d.path.begin(), m, path);
    if (quals && !d.quals.empty()) std::copy_n(d.quals.begin(), std::min<int64_t>(m, int64_t(d.quals.size())), quals);
    return n;
}

}  // namespace

// =============================================================================================
// C interface used by tests/ and bench.py (ctypes).  Outputs are alphabet indices; the caller
// maps them to label strings as src/search.rs:358,291 and src/lib.rs:144-146 do.
// Strides are in elements.  *_len receives the true length even when it exceeds `cap`.
// =============================================================================================
extern "C" {

int oracle_hardware_threads() { return int(std::thread::hardware_concurrency()); }

uint32_t oracle_phred(float prob, float qscale, float qbias, int* valid) {
    bool v; uint32_t c = phred_code(prob, qscale, qbias, &v);
    if (valid) *valid = v ? 1 : 0;
    return c;
}

// glibc's log10f over an array: what Rust's f32::log10 calls on linux-gnu (src/search.rs:34)
void oracle_log10f_n(const float* x, float* y, int64_t n) { for (int64_t i = 0; i < n; ++i) y[i] = log10f(x[i]); }

void oracle_logf_n(const float* x, float* y, int64_t n) { for (int64_t i = 0; i < n; ++i) y[i] = logf(x[i]); }
// the system libm's expf / log1pf: what Rust's f32::exp / f32::ln_1p call in LogSpace '+' (src/duplex.rs:45-52)
// The softmax recipe of the logits entry points (fast-ctc-decode_b200/csrc/input.cuh), with the system libm: m = first
// maximum, e_c = expf(z_c - m), S summed in index order, p_c = e_c / S, all float32.  Not part of the reference (its
// callers hand it posteriors); decoding p with the reference algorithms is what the GPU path must reproduce.
void oracle_softmax_rows(const float* z, float* p, int64_t n_rows, int64_t C) {
    for (int64_t r = 0; r < n_rows; ++r) {
        const float* zr = z + r * C;
        float* pr = p + r * C;
        float m = zr[0];
        for (int64_t c = 1; c < C; ++c)
            if (zr[c] > m) m = zr[c];
        float S = 0.0f;
        for (int64_t c = 0; c < C; ++c) {
            pr[c] = expf(zr[c] - m);
            S = (c == 0) ? pr[c] : S + pr[c];
        }
        for (int64_t c = 0; c < C; ++c) pr[c] = pr[c] / S;
    }
}
void oracle_expf_n(const float* x, float* y, int64_t n) { for (int64_t i = 0; i < n; ++i) y[i] = expf(x[i]); }
void oracle_log1pf_n(const float* x, float* y, int64_t n) { for (int64_t i = 0; i < n; ++i) y[i] = log1pf(x[i]); }

float oracle_fastexp_as_shipped(float a) { return fastexp_as_shipped(a); }
float oracle_logadd(float a, float b, int exp_mode) { g_exp_mode = exp_mode; return (Lg{a} + Lg{b}).v; }

int oracle_viterbi(const float* post, int64_t T, int64_t C, int64_t stride_t, int64_t stride_c, int qstring,
                   float qscale, float qbias, int collapse, int32_t* labels, uint32_t* quals, int64_t* path,
                   int64_t cap, int64_t* out_len) {
    Decoded d;
    int st = viterbi(Posterior2{post, T, C, stride_t, stride_c}, qstring != 0, qscale, qbias, collapse != 0, &d);
    *out_len = copy_out(d, labels, quals, path, cap);
    return st;
}

int oracle_beam_search(const float* post, int64_t T, int64_t C, int64_t stride_t, int64_t stride_c, int64_t beam_size,
                       float thr, int collapse, int32_t* labels, int64_t* path, int64_t cap, int64_t* out_len,
                       int64_t* n_nodes) {
    Decoded d;
    int st = beam1d(Posterior2{post, T, C, stride_t, stride_c}, size_t(beam_size), thr, collapse != 0, &d, n_nodes);
    *out_len = (st == ST_OK) ? copy_out(d, labels, nullptr, path, cap) : 0;
    return st;
}

int oracle_crf_greedy(const float* post, int64_t T, int64_t S, int64_t C, const float* init, int64_t n_init,
                      int qstring, float qscale, float qbias, int32_t* labels, uint32_t* quals, int64_t* path,
                      int64_t cap, int64_t* out_len) {
    Decoded d;
    int st = crf_greedy(Posterior3{post, T, S, C, S * C, C, 1}, init, n_init, qstring != 0, qscale, qbias, &d);
    *out_len = copy_out(d, labels, quals, path, cap);
    return st;
}

int oracle_crf_beam_search(const float* post, int64_t T, int64_t S, int64_t C, const float* init, int64_t n_init,
                           int64_t beam_size, float thr, int32_t* labels, int64_t* path, int64_t cap,
                           int64_t* out_len) {
    Decoded d;
    int st = crf_beam1d(Posterior3{post, T, S, C, S * C, C, 1}, init, n_init, size_t(beam_size), thr, &d);
    *out_len = (st == ST_OK) ? copy_out(d, labels, nullptr, path, cap) : 0;
    return st;
}

int oracle_beam_search_duplex(const float* post1, int64_t T1, const float* post2, int64_t T2, int64_t C,
                              const uint64_t* envelope /* [T1,2] or NULL */, int64_t beam_size, float thr,
                              int collapse, int exp_mode, int32_t* labels, int64_t cap, int64_t* out_len,
                              int64_t* cells) {
    g_exp_mode = exp_mode;
    std::vector<uint64_t> dflt;
    if (!envelope) {  // src/lib.rs:461-468
        dflt.resize(size_t(T1) * 2);
        for (int64_t t = 0; t < T1; ++t) { dflt[size_t(2 * t)] = 0; dflt[size_t(2 * t + 1)] = uint64_t(T2); }
        envelope = dflt.data();
    }
    std::vector<int32_t> seq;
    int st = duplex2d(Posterior2{post1, T1, C, C, 1}, Posterior2{post2, T2, C, C, 1}, envelope, 2, 1, size_t(beam_size),
                      thr, collapse != 0, &seq, cells);
    *out_len = (st == ST_OK) ? int64_t(seq.size()) : 0;
    if (st == ST_OK && labels) std::copy_n(seq.begin(), std::min<int64_t>(cap, int64_t(seq.size())), labels);
    return st;
}

int oracle_crf_beam_search_duplex(const float* post1, int64_t T1, const float* init1, int64_t n_init1,
                                  const float* post2, int64_t T2, const float* init2, int64_t n_init2, int64_t S,
                                  int64_t C, const uint64_t* envelope, int64_t beam_size, float thr, int exp_mode,
                                  int32_t* labels, int64_t cap, int64_t* out_len) {
    g_exp_mode = exp_mode;
    std::vector<uint64_t> dflt;
    if (!envelope) {
        dflt.resize(size_t(T1) * 2);
        for (int64_t t = 0; t < T1; ++t) { dflt[size_t(2 * t)] = 0; dflt[size_t(2 * t + 1)] = uint64_t(T2); }
        envelope = dflt.data();
    }
    std::vector<int32_t> seq;
    int st = crf_duplex2d(Posterior3{post1, T1, S, C, S * C, C, 1}, init1, n_init1,
                          Posterior3{post2, T2, S, C, S * C, C, 1}, init2, n_init2, envelope, 2, 1, size_t(beam_size),
                          thr, &seq);
    *out_len = (st == ST_OK) ? int64_t(seq.size()) : 0;
    if (st == ST_OK && labels) std::copy_n(seq.begin(), std::min<int64_t>(cap, int64_t(seq.size())), labels);
    return st;
}

// ---- batch drivers: one read per worker thread (the timed CPU baseline) --------
// Outputs use fixed slots of `cap` entries per read; lens[n] is the true length; status[n] the
// per-read status.  Any output pointer may be NULL (timing runs that discard results).
int oracle_viterbi_batch(const float* post, int64_t N, int64_t T, int64_t C, int qstring, float qscale, float qbias,
                         int collapse, int32_t* labels, uint32_t* quals, int64_t* path, int64_t cap, int64_t* lens,
                         int32_t* status, int n_threads) {
    parallel_reads(N, n_threads, [&](int64_t n) {
        Decoded d;
        int st = viterbi(Posterior2{post + n * T * C, T, C, C, 1}, qstring != 0, qscale, qbias, collapse != 0, &d);
        int64_t len = copy_out(d, labels ? labels + n * cap : nullptr, quals ? quals + n * cap : nullptr,
                               path ? path + n * cap : nullptr, cap);
        if (lens) lens[n] = len;
        if (status) status[n] = st;
    });
    return 0;
}

int oracle_beam_search_batch(const float* post, int64_t N, int64_t T, int64_t C, int64_t beam_size, float thr,
                             int collapse, int32_t* labels, int64_t* path, int64_t cap, int64_t* lens,
                             int32_t* status, int64_t* n_nodes, int n_threads) {
    parallel_reads(N, n_threads, [&](int64_t n) {
        Decoded d; int64_t nn = 0;
        int st = beam1d(Posterior2{post + n * T * C, T, C, C, 1}, size_t(beam_size), thr, collapse != 0, &d, &nn);
        int64_t len = 0;
        if (st == ST_OK)
            len = copy_out(d, labels ? labels + n * cap : nullptr, nullptr, path ? path + n * cap : nullptr, cap);
        if (lens) lens[n] = len;
        if (status) status[n] = st;
        if (n_nodes) n_nodes[n] = nn;
    });
    return 0;
}

int oracle_beam_search_duplex_batch(const float* post1, const float* post2, int64_t N, int64_t T1, int64_t T2,
                                    int64_t C, const uint64_t* envelope /* [N,T1,2] or NULL */, int64_t beam_size,
                                    float thr, int collapse, int exp_mode, int32_t* labels, int64_t cap, int64_t* lens,
                                    int32_t* status, int n_threads) {
    parallel_reads(N, n_threads, [&](int64_t n) {
        int64_t len = 0;
        int st = oracle_beam_search_duplex(post1 + n * T1 * C, T1, post2 + n * T2 * C, T2, C,
                                           envelope ? envelope + n * T1 * 2 : nullptr, beam_size, thr, collapse,
                                           exp_mode, labels ? labels + n * cap : nullptr, cap, &len, nullptr);
        if (lens) lens[n] = len;
        if (status) status[n] = st;
    });
    return 0;
}

// ---- hooks for the reference's unit tests of its support types --------------------------------
// SuffixTree assembly (src/tree.rs:201-269): build from (parent,label,data) triples and answer
// child / label / data / ancestor queries.
struct OracleTrie { LabelTrie<int64_t> t; explicit OracleTrie(size_t n) : t(n) {} };
void* oracle_trie_new(int64_t n_labels) { return new OracleTrie(size_t(n_labels)); }
void oracle_trie_free(void* h) { delete static_cast<OracleTrie*>(h); }
int32_t oracle_trie_add(void* h, int32_t parent, int64_t label, int64_t data) {
    return static_cast<OracleTrie*>(h)->t.add(parent, size_t(label), data);
}
int32_t oracle_trie_child(void* h, int32_t node, int64_t label) {
    return static_cast<OracleTrie*>(h)->t.child(node, size_t(label));
}
int64_t oracle_trie_label(void* h, int32_t node) { return static_cast<OracleTrie*>(h)->t.label_of(node); }
int oracle_trie_data(void* h, int32_t node, int64_t* data) {
    const int64_t* p = static_cast<OracleTrie*>(h)->t.payload(node);
    if (!p) return 0;
    *data = *p;
    return 1;
}
int64_t oracle_trie_ancestors(void* h, int32_t node, int64_t* labels, int64_t* data, int64_t cap) {
    auto& t = static_cast<OracleTrie*>(h)->t;
    int64_t k = 0;
    for (int32_t n = node; n >= 0; n = t.parent_of(n), ++k)
        if (k < cap) { labels[k] = t.label_of(n); data[k] = *t.payload(n); }
    return k;
}

// SecondaryProbs::get / update_max (src/duplex.rs:842-993).  Entries are given in linear
// probability as (label, gap) pairs, 0 meaning LogSpace::zero().
float oracle_secondary_get_gap(int64_t offset, const float* label_lin, const float* gap_lin, int64_t n, int64_t at) {
    Secondary s; s.offset = offset;
    for (int64_t i = 0; i < n; ++i) s.probs.push_back(Pair{Lg::of(label_lin[i]), Lg::of(gap_lin[i])});
    return s.get(at).gap.v;
}
float oracle_secondary_update_max(int64_t offset, const float* label_lin, const float* gap_lin, int64_t n, int64_t lo,
                                  int64_t hi, int exp_mode) {
    g_exp_mode = exp_mode;
    Secondary s; s.offset = offset; s.max_prob = Lg::one();
    for (int64_t i = 0; i < n; ++i) s.probs.push_back(Pair{Lg::of(label_lin[i]), Lg::of(gap_lin[i])});
    s.update_max(lo, hi);
    return s.max_prob.v;
}

}  // extern "C"
