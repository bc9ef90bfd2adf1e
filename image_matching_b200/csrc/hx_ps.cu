// hx_ps.cu -- Paterson-Stockmeyer evaluators and OpenFHEWrapper::chebyshevCompare
// (/root/reference/src/openFHE_wrapper.cpp:143-185) on top of the Engine ops.
// Host-side double arithmetic (coefficients, long divisions) follows OpenFHE v1.2.3's
// EvalChebyshevCoefficients / LongDivisionChebyshev / LongDivisionPoly / EvalChebyshevSeriesPS /
// EvalPolyPS in the FIXEDMANUAL branch: same formulas and association order, because every
// rounded constant feeds integer polynomial arithmetic.
#include <cmath>

#include <functional>

#include "hx_engine.h"

namespace hx {
namespace {

using Vec = std::vector<double>;

bool not_one(double v) {   // IsNotEqualOne: |v - 1| > 2^-20
    const double prec = std::ldexp(1.0, -20);
    return (1 - prec > v) || (1 + prec < v);
}
int degree_of(const Vec& c) {
    int d = (int)c.size() - 1;
    while (d > 0 && c[d] == 0) d--;
    return d;
}
Vec resized(const Vec& v, size_t n) {
    Vec r(v);
    r.resize(n, 0.0);
    return r;
}

// m (giant-step count) per polynomial degree, OpenFHE's PopulateParameterPS table.
int ps_m(int n) {
    static const int bounds[][2] = {{2, 1},   {11, 2},   {13, 3},   {17, 2},   {55, 3},   {59, 4},   {76, 3},
                                    {239, 4}, {247, 5},  {284, 4},  {991, 5},  {1021, 6}, {1078, 5}, {2204, 6}};
    for (auto& b : bounds)
        if (n <= b[0]) return b[1];
    throw std::runtime_error("polynomial degree above the Paterson-Stockmeyer table");
}
void degrees_ps(int n, int& k, int& m) {
    m = ps_m(n);
    k = n / ((1 << m) - 1) + 1;
}

struct Div { Vec q, r; };

Div long_div_cheb(const Vec& f, const Vec& g) {
    int n = degree_of(f), k = degree_of(g);
    if (n != (int)f.size() - 1 || k != (int)g.size() - 1) throw std::runtime_error("LongDivisionChebyshev: zero dominant coefficient");
    Div out;
    Vec r = f;
    if (n - k >= 0) {
        Vec q(n - k + 1, 0.0);
        while (n - k > 0) {
            q[n - k] = 2 * r.back();
            if (not_one(g[k])) q[n - k] /= g.back();
            Vec d(n + 1, 0.0);
            if (k == n - k) {
                d[0] = 2 * g[n - k];
                for (int i = 1; i < 2 * k + 1; i++) d[i] = g[std::abs(n - k - i)];
            } else if (k > n - k) {
                d[0] = 2 * g[n - k];
                for (int i = 1; i < k - (n - k) + 1; i++) d[i] = g[std::abs(n - k - i)] + g[n - k + i];
                for (int i = k - (n - k) + 1; i < n + 1; i++) d[i] = g[std::abs(i - n + k)];
            } else {
                d[n - k] = g[0];
                for (int i = n - 2 * k; i < n + 1; i++)
                    if (i != n - k) d[i] = g[std::abs(i - n + k)];
            }
            if (not_one(r.back()))
                for (auto& x : d) x *= r.back();
            if (not_one(g.back()))
                for (auto& x : d) x /= g.back();
            for (size_t i = 0; i < r.size(); i++) r[i] -= d[i];
            if (r.size() > 1) {
                n = degree_of(r);
                r.resize(n + 1);
            }
        }
        if (n == k) {
            q[0] = r.back();
            if (not_one(g.back())) q[0] /= g.back();
            Vec d = g;
            if (not_one(r.back()))
                for (auto& x : d) x *= r.back();
            if (not_one(g.back()))
                for (auto& x : d) x /= g.back();
            for (size_t i = 0; i < r.size(); i++) r[i] -= d[i];
            if (r.size() > 1) {
                n = degree_of(r);
                r.resize(n + 1);
            }
        }
        q[0] *= 2;   // [c0] rather than [c0/2] in the free slot
        out.q = q;
    } else {
        out.q = Vec(1, 0.0);
    }
    out.r = r;
    return out;
}

Div long_div_poly(const Vec& f, const Vec& g) {
    int n = degree_of(f), k = degree_of(g);
    if (n != (int)f.size() - 1 || k != (int)g.size() - 1) throw std::runtime_error("LongDivisionPoly: zero dominant coefficient");
    Div out;
    if (n - k < 0) {
        out.q = Vec(1, 0.0);
        out.r = f;
        return out;
    }
    Vec q(n - k + 1, 0.0), r = f;
    while (n - k >= 0) {
        Vec d(n - k, 0.0);
        d.insert(d.end(), g.begin(), g.end());
        q[n - k] = r.back();
        if (not_one(g[k])) q[n - k] /= g.back();
        for (auto& x : d) x *= q[n - k];
        for (size_t i = 0; i < r.size(); i++) r[i] -= d[i];
        if (r.size() > 1) {
            n = degree_of(r);
            r.resize(n + 1);
        } else {
            break;
        }
    }
    out.q = q;
    out.r = r;
    return out;
}

// Shared split of the PS recursion: coeffs = q * B_{k 2^(m-1)} + r ;  r - B_{k(2^(m-1)-1)} = c * q + s.
struct Split { Vec q, c, s2; };
Split ps_split(const Vec& coeffs, int k, int m, bool cheb) {
    int k2m2k = k * (1 << (m - 1)) - k;
    Vec tkm(k2m2k + k + 1, 0.0);
    tkm.back() = 1;
    Div qr = cheb ? long_div_cheb(coeffs, tkm) : long_div_poly(coeffs, tkm);
    Vec r2 = qr.r;
    if (k2m2k - degree_of(qr.r) <= 0) {
        r2[k2m2k] -= 1;
        r2.resize(degree_of(r2) + 1);
    } else {
        r2.resize(k2m2k + 1, 0.0);
        r2.back() = -1;
    }
    Div cs = cheb ? long_div_cheb(r2, qr.q) : long_div_poly(r2, qr.q);
    Split s;
    s.q = qr.q;
    s.c = cs.q;
    s.s2 = cs.r;
    s.s2.resize(k2m2k + 1, 0.0);
    s.s2.back() = 1;
    return s;
}

}  // namespace

// EvalLinearWSumMutable: sum_i w_i * ct_i, then one ModReduce.
Ct Engine::linear_wsum(const std::vector<Ct>& cts, const std::vector<double>& w) {
    Guard g(this);
    bool same = cts.size() <= (size_t)kMaxWsum;
    for (auto& c : cts) same = same && c->nl == cts[0]->nl && c->ncomp == cts[0]->ncomp && c->nb == cts[0]->nb && c->deg == cts[0]->deg;
    if (same) {   // one launch: sum_t w_t * ct_t (modular addition is associative: same bits as the add chain below)
        const Ct& f = cts[0];
        WsumArgs a;
        a.n = (int)cts.size();
        for (int t = 0; t < a.n; t++) {
            a.p[t] = cts[t]->d;
            ScalarTable s = scalar_table(const_for_mult(w[t], f->nl));
            for (int i = 0; i < f->nl; i++) { a.v[t][i] = s.v[i]; a.sh[t][i] = s.sh[i]; }
        }
        Ct acc = new_ct(f->ncomp, f->nl, f->level, f->deg + 1, f->nb);
        { Scope sc(this, PK_SCALAR, 1, 8.0 * N * f->nl * f->comps_total() * (a.n + 1)); launch_wsum(T, acc->d, a, f->nl, f->comps_total(), st); }
        return rescale(acc);
    }
    Ct acc = mult_const(cts[0], w[0]);
    for (size_t i = 1; i < cts.size(); i++) acc = add(acc, mult_const(cts[i], w[i]));
    return rescale(acc);
}

static Ct eval_c(Engine& e, Ct (Engine::*wsum)(const std::vector<Ct>&, const std::vector<double>&), const std::vector<Ct>& basis,
                 const std::vector<double>& c, bool half) {
    int dc = 0;
    for (int d = (int)c.size() - 1; d > 0; d--)
        if (c[d] != 0) { dc = d; break; }
    if (dc < 1) return nullptr;
    Ct cu;
    if (dc == 1) {
        cu = (c[1] != 1) ? e.rescale(e.mult_const(basis[0], c[1])) : basis[0];
    } else {
        std::vector<Ct> cts(basis.begin(), basis.begin() + dc);
        std::vector<double> w(c.begin() + 1, c.begin() + 1 + dc);
        cu = (e.*wsum)(cts, w);
    }
    return e.add_const(cu, half ? c[0] / 2 : c[0]);
}

// InnerEvalChebyshevPS / InnerEvalPolyPS, level-synchronous (see hx_engine.h).  A node computes
//   cu = c(basis) ; qu, su = child result or leaf polynomial ; out = Rescale((basis2[m-1] + cu) * qu) + su
// exactly as the upstream recursion does; only the order of independent operations differs.
std::vector<Ct> Engine::inner_ps_many(const std::vector<Vec>& roots, int k, int m0, std::vector<Ct>& B, std::vector<Ct>& B2, bool cheb) {
    struct Node { Split sp; int m, depth, qchild = -1, schild = -1; Ct out; };
    std::vector<Node> nodes;
    std::function<int(const Vec&, int, int)> build = [&](const Vec& c, int m, int depth) -> int {
        Node n;
        n.sp = ps_split(c, k, m, cheb);
        n.m = m; n.depth = depth;
        const int id = (int)nodes.size();
        nodes.push_back(n);
        if (degree_of(nodes[id].sp.q) > k) { int ch = build(nodes[id].sp.q, m - 1, depth + 1); nodes[id].qchild = ch; }
        if (degree_of(nodes[id].sp.s2) > k) { int ch = build(nodes[id].sp.s2, m - 1, depth + 1); nodes[id].schild = ch; }
        return id;
    };
    std::vector<int> root_ids;
    for (auto& r : roots) root_ids.push_back(build(r, m0, 0));
    int maxd = 0;
    for (auto& n : nodes) maxd = std::max(maxd, n.depth);
    auto leaf_q = [&](const Split& sp) -> Ct {
        Ct qu;
        if (cheb) {
            Vec qc = resized(sp.q, k);
            int dq = degree_of(qc);
            Ct sum = B[k - 1];
            for (int i = 0; i < std::log2(sp.q.back()); i++) sum = add(sum, sum);
            if (dq > 0) {
                std::vector<Ct> cts(B.begin(), B.begin() + dq);
                Vec w(sp.q.begin() + 1, sp.q.begin() + 1 + dq);
                qu = add(linear_wsum(cts, w), sum);
            } else {
                qu = sum;
            }
            return add_const(qu, sp.q.front() / 2);
        }
        Vec vc = resized(sp.q, k);
        int dv = degree_of(vc);
        if (dv > 0) {
            std::vector<Ct> cts(B.begin(), B.begin() + dv);
            Vec w(sp.q.begin() + 1, sp.q.begin() + 1 + dv);
            qu = add(linear_wsum(cts, w), B[k - 1]);
        } else {
            qu = B[k - 1];
        }
        return add_const(qu, sp.q.front());
    };
    auto leaf_s = [&](const Split& sp) -> Ct {
        Vec sc = resized(sp.s2, k);
        int ds = degree_of(sc);
        Ct su;
        if (ds > 0) {
            std::vector<Ct> cts(B.begin(), B.begin() + ds);
            Vec w(sp.s2.begin() + 1, sp.s2.begin() + 1 + ds);
            su = add(linear_wsum(cts, w), B[k - 1]);
        } else {
            su = B[k - 1];
        }
        su = add_const(su, cheb ? sp.s2.front() / 2 : sp.s2.front());
        return cheb ? level_reduce(su, 1) : su;
    };
    for (int d = maxd; d >= 0; d--) {
        std::vector<int> ids;
        for (int i = 0; i < (int)nodes.size(); i++)
            if (nodes[i].depth == d) ids.push_back(i);
        std::vector<Ct> res(ids.size()), qu(ids.size()), su(ids.size());
        for (size_t j = 0; j < ids.size(); j++) {
            Node& n = nodes[ids[j]];
            Ct cu = eval_c(*this, &Engine::linear_wsum, B, n.sp.c, cheb);
            qu[j] = n.qchild >= 0 ? nodes[n.qchild].out : leaf_q(n.sp);
            su[j] = n.schild >= 0 ? nodes[n.schild].out : leaf_s(n.sp);
            res[j] = cu ? add(B2[n.m - 1], cu) : add_const(B2[n.m - 1], cheb ? n.sp.c.front() / 2 : n.sp.c.front());
        }
        bool same = ids.size() > 1 && batch_blocks && !force_unfused && fused_ks_supported(logN);
        for (size_t j = 1; j < ids.size() && same; j++)
            same = res[j]->nl == res[0]->nl && res[j]->deg == res[0]->deg && res[j]->nb == res[0]->nb && res[j]->ncomp == res[0]->ncomp &&
                   qu[j]->nl == qu[0]->nl && qu[j]->deg == qu[0]->deg && qu[j]->nb == qu[0]->nb && qu[j]->ncomp == qu[0]->ncomp;
        if (same) {
            std::vector<Ct> prod = split(rescale(mult(concat(res), concat(qu))), (int)ids.size());
            for (size_t j = 0; j < ids.size(); j++) nodes[ids[j]].out = add(prod[j], su[j]);
        } else {
            for (size_t j = 0; j < ids.size(); j++) nodes[ids[j]].out = add(rescale(mult(res[j], qu[j])), su[j]);
        }
    }
    std::vector<Ct> out;
    for (int id : root_ids) out.push_back(nodes[id].out);
    return out;
}
Ct Engine::inner_cheb(const std::vector<double>& coeffs, int k, int m, std::vector<Ct>& T, std::vector<Ct>& T2) {
    return inner_ps_many({coeffs}, k, m, T, T2, true)[0];
}

// EvalChebyshevSeriesPS (cc->EvalChebyshevFunction, openFHE_wrapper.cpp:174).
// Rescale(f * a0 * b0) and Rescale(f * a1 * b1) (f = 2: the Chebyshev doubling) for two independent products whose operands sit
// at one level: ONE batched multiply / relinearise / rescale instead of two chains of ~14 dependent launches each. Per entry the
// operations and operands are those of the sequential code, hence the same bits.
std::pair<Ct, Ct> Engine::two_products(Ct a0, Ct b0, Ct a1, Ct b1, bool twice) {
    align(a0, b0);
    align(a1, b1);
    auto one = [&](const Ct& a, const Ct& b) {
        Ct p = mult(a, b);
        return rescale(twice ? add(p, p) : p);
    };
    auto shape_eq = [](const Ct& u, const Ct& v) { return u->nl == v->nl && u->deg == v->deg && u->nb == v->nb && u->ncomp == v->ncomp && u->level == v->level; };
    if (!(batch_blocks && !force_unfused && fused_ks_supported(logN) && shape_eq(a0, a1) && shape_eq(b0, b1) && a0->ncomp == 2 && b0->ncomp == 2))
        return {one(a0, b0), one(a1, b1)};
    Ct p = mult(concat({a0, a1}), concat({b0, b1}));
    std::vector<Ct> r = split(rescale(twice ? add(p, p) : p), 2);
    return {r[0], r[1]};
}

Ct Engine::chebyshev_series(const Ct& x, const std::vector<double>& coefficients, double a, double b) {
    Guard g(this);
    int n = degree_of(coefficients);
    if (n < 5) throw std::runtime_error("EvalChebyshevSeries: degree < 5 (linear evaluation) is not on the HyDia path");
    Vec f2 = coefficients;
    if (coefficients.back() == 0) f2.resize(n + 1);
    int k, m;
    degrees_ps(n, k, m);
    std::vector<Ct> T(k);
    if ((a - std::round(a) < 1e-10) && (b - std::round(b) < 1e-10) && std::round(a) == -1 && std::round(b) == 1) {
        T[0] = x;
    } else {
        double al = 2 / (b - a), be = 2 * a / (b - a);
        T[0] = add_const(rescale(mult_const(x, al)), -1.0 - be);
    }
    Ct y = T[0];
    for (int i = 2; i <= k; i++) {
        if (!(i & (i - 1)) || i % 2 == 0) {
            Ct sq = square(T[i / 2 - 1]);
            T[i - 1] = add_const(rescale(add(sq, sq)), -1.0);
        } else if (i + 1 <= k) {      // T_i = 2 T_(i-1)/2 T_(i+1)/2 - T_1 and T_(i+1) = 2 T_(i+1)/2 ^2 - 1 depend on earlier terms only
            auto pr = two_products(T[i / 2 - 1], T[i / 2], T[(i + 1) / 2 - 1], T[(i + 1) / 2 - 1], true);
            T[i - 1] = sub(pr.first, y);
            T[i] = add_const(pr.second, -1.0);
            i++;
        } else {
            Ct prod = mult(T[i / 2 - 1], T[i / 2]);
            T[i - 1] = sub(rescale(add(prod, prod)), y);
        }
    }
    for (int i = 1; i < k; i++) T[i - 1] = level_reduce(T[i - 1], T[k - 1]->level - T[i - 1]->level);
    std::vector<Ct> T2(m);
    T2[0] = T.back();
    Ct T2km1 = T2[0];
    if (m > 1) {
        Ct sq = square(T2[0]);
        T2[1] = add_const(rescale(add(sq, sq)), -1.0);
    }
    for (int i = 1; i < m; i++) {
        if (i + 1 < m) {              // T2[i+1] = 2 T2[i]^2 - 1 and the next T2km1 = 2 T2km1 T2[i] - T2[0] are independent
            auto pr = two_products(T2[i], T2[i], T2km1, T2[i], true);
            T2[i + 1] = add_const(pr.first, -1.0);
            T2km1 = sub(pr.second, T2[0]);
        } else {
            Ct prod = mult(T2km1, T2[i]);
            T2km1 = sub(rescale(add(prod, prod)), T2[0]);
        }
    }
    int k2m2k = k * (1 << (m - 1)) - k;
    f2.resize(2 * k2m2k + k + 1, 0.0);
    f2.back() = 1;
    Split sp = ps_split(f2, k, m, true);
    Ct cu = eval_c(*this, &Engine::linear_wsum, T, sp.c, true);
    if (cu) cu = level_reduce(cu, T2[m - 1]->level - cu->level);
    Ct qu, su;
    std::vector<Ct> both;
    if (degree_of(sp.q) > k && degree_of(sp.s2) > k) both = inner_ps_many({sp.q, sp.s2}, k, m - 1, T, T2, true);
    if (!both.empty()) {
        qu = both[0];
    } else if (degree_of(sp.q) > k) {
        qu = inner_cheb(sp.q, k, m - 1, T, T2);
    } else {
        Vec qc = resized(sp.q, k);
        int dq = degree_of(qc);
        if (dq > 0) {
            std::vector<Ct> cts(T.begin(), T.begin() + dq);
            Vec w(sp.q.begin() + 1, sp.q.begin() + 1 + dq);
            qu = add(linear_wsum(cts, w), add(T[k - 1], T[k - 1]));
        } else {
            qu = T[k - 1];
            for (int i = 1; i < sp.q.back(); i++) qu = add(qu, T[k - 1]);
        }
        qu = add_const(qu, sp.q.front() / 2);
    }
    if (!both.empty()) {
        su = both[1];
    } else if (degree_of(sp.s2) > k) {
        su = inner_cheb(sp.s2, k, m - 1, T, T2);
    } else {
        Vec sc = resized(sp.s2, k);
        int ds = degree_of(sc);
        if (ds > 0) {
            std::vector<Ct> cts(T.begin(), T.begin() + ds);
            Vec w(sp.s2.begin() + 1, sp.s2.begin() + 1 + ds);
            su = add(linear_wsum(cts, w), T[k - 1]);
        } else {
            su = T[k - 1];
        }
        su = add_const(su, sp.s2.front() / 2);
    }
    Ct res = cu ? add(T2[m - 1], cu) : add_const(T2[m - 1], sp.c.front() / 2);
    res = rescale(mult(res, qu));
    res = add(res, su);
    return sub(res, T2km1);
}

// InnerEvalPolyPS.
Ct Engine::inner_poly(const std::vector<double>& coeffs, int k, int m, std::vector<Ct>& pw, std::vector<Ct>& pw2) {
    return inner_ps_many({coeffs}, k, m, pw, pw2, false)[0];
}

// EvalPolyPS (cc->EvalPoly, openFHE_wrapper.cpp:179).
Ct Engine::eval_poly(const Ct& x, const std::vector<double>& coefficients) {
    Guard g(this);
    int n = degree_of(coefficients);
    if (n < 5) throw std::runtime_error("EvalPoly: degree < 5 (linear evaluation) is not on the HyDia path");
    Vec f2 = coefficients;
    if (coefficients.back() == 0) f2.resize(n + 1);
    int k, m;
    degrees_ps(n, k, m);
    std::vector<int> need(k, 0);
    for (int i = k; i > 0; i--) {
        need[i - 1] = 1;
        if (i & (i - 1)) {
            int p2 = 1 << (int)std::floor(std::log2(i));
            int rem = i % p2;
            need[rem - 1] = 1;
            while (rem & (rem - 1)) {
                p2 = 1 << (int)std::floor(std::log2(rem));
                rem = rem % p2;
                need[rem - 1] = 1;
            }
        }
    }
    std::vector<Ct> pw(k);
    pw[0] = x;
    for (int i = 2; i <= k; i++) {
        if (!(i & (i - 1))) {
            pw[i - 1] = rescale(square(pw[i / 2 - 1]));
        } else if (need[i - 1] && i + 1 <= k && !((i + 1) & i) && (i + 1) / 2 < i) {
            // x^i = x^p2 * x^rem and x^(i+1) = (x^((i+1)/2))^2 are independent of each other
            int p2 = 1 << (int)std::floor(std::log2(i));
            int rem = i % p2;
            pw[rem - 1] = level_reduce(pw[rem - 1], pw[p2 - 1]->level - pw[rem - 1]->level);
            auto pr = two_products(pw[p2 - 1], pw[rem - 1], pw[(i + 1) / 2 - 1], pw[(i + 1) / 2 - 1], false);
            pw[i - 1] = pr.first;
            pw[i] = pr.second;
            i++;
        } else if (need[i - 1]) {
            int p2 = 1 << (int)std::floor(std::log2(i));
            int rem = i % p2;
            pw[rem - 1] = level_reduce(pw[rem - 1], pw[p2 - 1]->level - pw[rem - 1]->level);
            pw[i - 1] = rescale(mult(pw[p2 - 1], pw[rem - 1]));
        }
    }
    for (int i = 1; i < k; i++)
        if (need[i - 1]) pw[i - 1] = level_reduce(pw[i - 1], pw[k - 1]->level - pw[i - 1]->level);
    std::vector<Ct> pw2(m);
    pw2[0] = pw.back();
    for (int i = 1; i < m; i++) pw2[i] = rescale(square(pw2[i - 1]));
    Ct p2km1 = pw2[0];
    for (int i = 1; i < m; i++) p2km1 = rescale(mult(p2km1, pw2[i]));
    int k2m2k = k * (1 << (m - 1)) - k;
    f2.resize(2 * k2m2k + k + 1, 0.0);
    f2.back() = 1;
    Ct res = inner_poly(f2, k, m, pw, pw2);
    return sub(res, p2km1);
}

// OpenFHEWrapper::chebyshevCompare (openFHE_wrapper.cpp:143-185).
Ct Engine::chebyshev_compare(const Ct& x, double delta, int sign_depth) {
    if (sign_depth < 7 || sign_depth > 15) return x;   // the reference prints an error and returns the input
    static const int kDepthToDegree[] = {-1, -1, -1, 5, 13, 27, 59, 119, 247, 495, 1007, 2031};
    static const Vec kF4 = {0.0, 315.0 / 128.0, 0.0, -420.0 / 128.0, 0.0, 378.0 / 128.0, 0.0, -180.0 / 128.0, 0.0, 35.0 / 128.0};
    int deg = kDepthToDegree[sign_depth - 4];
    // EvalChebyshevCoefficients(f, -1, 1, deg), f = step at delta
    int nc = deg + 1;
    double a = -1, b = 1;
    double bMinusA = 0.5 * (b - a), bPlusA = 0.5 * (b + a);
    double piByDeg = M_PI / (double)nc;
    Vec fp(nc), coeffs(nc, 0.0);
    for (int i = 0; i < nc; i++) {
        double pt = std::cos(piByDeg * (i + 0.5)) * bMinusA + bPlusA;
        fp[i] = (pt >= delta) ? 1.0 : -1.0;
    }
    double mf = 2.0 / (double)nc;
    for (int i = 0; i < nc; i++) {
        for (int j = 0; j < nc; j++) coeffs[i] += fp[j] * std::cos(piByDeg * i * (j + 0.5));
        coeffs[i] *= mf;
    }
    Ct r = chebyshev_series(x, coeffs, a, b);
    r = eval_poly(r, kF4);
    return add_const(r, 1.0);
}

}  // namespace hx
