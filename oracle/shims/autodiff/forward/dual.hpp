// Stand-in for autodiff/forward/dual.hpp (autodiff/autodiff @ cb2d6e8 is an empty submodule under
// /root/reference/cpp/external/autodiff). TEST INFRASTRUCTURE ONLY: lets the unmodified reference
// sources compile into oracle/_ref/. Written from scratch: a minimal forward-mode dual number with
// the API surface the reference uses (dual, dual2nd, val/grad, wrt/at, derivative(s), elementary
// functions). Not part of the product path.
#pragma once
#include <cmath>
#include <array>
#include <tuple>
#include <type_traits>
#include <ostream>

namespace autodiff {

template<class T> struct Dual;

namespace detail {
    template<class T> struct is_dual : std::false_type {};
    template<class T> struct is_dual<Dual<T>> : std::true_type {};
    template<class T> struct depth { static constexpr int value = 0; };
    template<class T> struct depth<Dual<T>> { static constexpr int value = 1 + depth<T>::value; };
    template<class A> constexpr bool is_arith = std::is_arithmetic<A>::value;
    inline double innermost(double x) { return x; }
    template<class T> double innermost(const Dual<T>& x) { return innermost(x.val); }
}

template<class T>
struct Dual {
    T val{};
    T grad{};
    Dual() = default;
    template<class A, std::enable_if_t<detail::is_arith<A>, int> = 0>
    Dual(A v) : val(static_cast<double>(v)), grad(0.0) {}
    // promote a lower-order dual (e.g. dual -> dual2nd): value part carries it, no seed
    template<class U = T, std::enable_if_t<detail::is_dual<U>::value, int> = 0>
    Dual(const U& v) : val(v), grad(0.0) {}
    Dual(const T& v, const T& g, int) : val(v), grad(g) {}
    explicit operator double() const { return detail::innermost(val); }
    template<class A, std::enable_if_t<detail::is_arith<A>, int> = 0>
    Dual& operator=(A v) { val = T(static_cast<double>(v)); grad = T(0.0); return *this; }
    Dual& operator+=(const Dual& o) { val = val + o.val; grad = grad + o.grad; return *this; }
    Dual& operator-=(const Dual& o) { val = val - o.val; grad = grad - o.grad; return *this; }
    Dual& operator*=(const Dual& o) { *this = (*this) * o; return *this; }
    Dual& operator/=(const Dual& o) { *this = (*this) / o; return *this; }
    template<class A, std::enable_if_t<detail::is_arith<A>, int> = 0> Dual& operator+=(A o) { val = val + T(static_cast<double>(o)); return *this; }
    template<class A, std::enable_if_t<detail::is_arith<A>, int> = 0> Dual& operator-=(A o) { val = val - T(static_cast<double>(o)); return *this; }
    template<class A, std::enable_if_t<detail::is_arith<A>, int> = 0> Dual& operator*=(A o) { val = val * T(static_cast<double>(o)); grad = grad * T(static_cast<double>(o)); return *this; }
    template<class A, std::enable_if_t<detail::is_arith<A>, int> = 0> Dual& operator/=(A o) { val = val / T(static_cast<double>(o)); grad = grad / T(static_cast<double>(o)); return *this; }
};

using dual = Dual<double>;
using dual1st = dual;
using dual2nd = Dual<dual>;

template<class T> Dual<T> mk(const T& v, const T& g) { return Dual<T>(v, g, 0); }

// ---- arithmetic: dual (op) dual
template<class T> Dual<T> operator+(const Dual<T>& a) { return a; }
template<class T> Dual<T> operator-(const Dual<T>& a) { return mk<T>(-a.val, -a.grad); }
template<class T> Dual<T> operator+(const Dual<T>& a, const Dual<T>& b) { return mk<T>(a.val + b.val, a.grad + b.grad); }
template<class T> Dual<T> operator-(const Dual<T>& a, const Dual<T>& b) { return mk<T>(a.val - b.val, a.grad - b.grad); }
template<class T> Dual<T> operator*(const Dual<T>& a, const Dual<T>& b) { return mk<T>(a.val * b.val, a.grad * b.val + a.val * b.grad); }
template<class T> Dual<T> operator/(const Dual<T>& a, const Dual<T>& b) {
    T q = a.val / b.val;
    return mk<T>(q, (a.grad - q * b.grad) / b.val);
}
// ---- mixed order (dual op dual2nd): the lower-order operand is promoted (no seed)
#define KG_AD_MIXED(op) \
    template<class T> Dual<Dual<T>> operator op(const Dual<T>& a, const Dual<Dual<T>>& b) { return Dual<Dual<T>>(a) op b; } \
    template<class T> Dual<Dual<T>> operator op(const Dual<Dual<T>>& a, const Dual<T>& b) { return a op Dual<Dual<T>>(b); }
KG_AD_MIXED(+) KG_AD_MIXED(-) KG_AD_MIXED(*) KG_AD_MIXED(/)
#undef KG_AD_MIXED
// ---- dual (op) arithmetic
#define KG_AD_ARITH template<class T, class A, std::enable_if_t<detail::is_arith<A>, int> = 0>
KG_AD_ARITH Dual<T> operator+(const Dual<T>& a, A b) { return mk<T>(a.val + T(static_cast<double>(b)), a.grad); }
KG_AD_ARITH Dual<T> operator+(A b, const Dual<T>& a) { return mk<T>(T(static_cast<double>(b)) + a.val, a.grad); }
KG_AD_ARITH Dual<T> operator-(const Dual<T>& a, A b) { return mk<T>(a.val - T(static_cast<double>(b)), a.grad); }
KG_AD_ARITH Dual<T> operator-(A b, const Dual<T>& a) { return mk<T>(T(static_cast<double>(b)) - a.val, -a.grad); }
KG_AD_ARITH Dual<T> operator*(const Dual<T>& a, A b) { return mk<T>(a.val * T(static_cast<double>(b)), a.grad * T(static_cast<double>(b))); }
KG_AD_ARITH Dual<T> operator*(A b, const Dual<T>& a) { return mk<T>(T(static_cast<double>(b)) * a.val, T(static_cast<double>(b)) * a.grad); }
KG_AD_ARITH Dual<T> operator/(const Dual<T>& a, A b) { return mk<T>(a.val / T(static_cast<double>(b)), a.grad / T(static_cast<double>(b))); }
KG_AD_ARITH Dual<T> operator/(A b, const Dual<T>& a) { return Dual<T>(b) / a; }
// ---- comparisons (on the innermost value)
#define KG_AD_CMP(op) \
    template<class T> bool operator op(const Dual<T>& a, const Dual<T>& b) { return detail::innermost(a) op detail::innermost(b); } \
    KG_AD_ARITH bool operator op(const Dual<T>& a, A b) { return detail::innermost(a) op static_cast<double>(b); } \
    KG_AD_ARITH bool operator op(A b, const Dual<T>& a) { return static_cast<double>(b) op detail::innermost(a); }
KG_AD_CMP(<) KG_AD_CMP(>) KG_AD_CMP(<=) KG_AD_CMP(>=) KG_AD_CMP(==) KG_AD_CMP(!=)
#undef KG_AD_CMP

// ---- elementary functions (chain rule; T-level math recurses through overloads / std)
using std::exp; using std::log; using std::sqrt; using std::pow; using std::tanh; using std::abs;
using std::sin; using std::cos; using std::cosh; using std::sinh; using std::erf;

template<class T> Dual<T> exp(const Dual<T>& x) { T e = exp(x.val); return mk<T>(e, e * x.grad); }
template<class T> Dual<T> log(const Dual<T>& x) { return mk<T>(log(x.val), x.grad / x.val); }
template<class T> Dual<T> sqrt(const Dual<T>& x) { T s = sqrt(x.val); return mk<T>(s, x.grad / (T(2.0) * s)); }
template<class T> Dual<T> sin(const Dual<T>& x) { return mk<T>(sin(x.val), cos(x.val) * x.grad); }
template<class T> Dual<T> cos(const Dual<T>& x) { return mk<T>(cos(x.val), -(sin(x.val) * x.grad)); }
template<class T> Dual<T> tanh(const Dual<T>& x) { T t = tanh(x.val); return mk<T>(t, (T(1.0) - t * t) * x.grad); }
template<class T> Dual<T> abs(const Dual<T>& x) { return (detail::innermost(x) < 0) ? -x : x; }
KG_AD_ARITH Dual<T> pow(const Dual<T>& x, A p_) {
    const double p = static_cast<double>(p_);
    return mk<T>(pow(x.val, p), T(p) * pow(x.val, p - 1.0) * x.grad);
}
template<class T> Dual<T> pow(const Dual<T>& x, const Dual<T>& p) {
    T v = pow(x.val, p.val);
    return mk<T>(v, v * (p.grad * log(x.val) + p.val * x.grad / x.val));
}
template<class T> Dual<Dual<T>> pow(const Dual<T>& x, const Dual<Dual<T>>& p) { return pow(Dual<Dual<T>>(x), p); }
template<class T> Dual<Dual<T>> pow(const Dual<Dual<T>>& x, const Dual<T>& p) { return pow(x, Dual<Dual<T>>(p)); }
KG_AD_ARITH Dual<T> pow(A b_, const Dual<T>& p) {
    const double b = static_cast<double>(b_);
    T v = pow(T(b), p.val);
    return mk<T>(v, v * T(std::log(b)) * p.grad);
}
#undef KG_AD_ARITH

template<class T> std::ostream& operator<<(std::ostream& os, const Dual<T>& x) { return os << detail::innermost(x); }

inline double val(double x) { return x; }
template<class T> double val(const Dual<T>& x) { return detail::innermost(x); }

// ---- wrt / at / derivative(s)
template<class... V> struct Wrt { std::tuple<V&...> vars; };
template<class... A> struct At { std::tuple<A&...> args; };
template<class... V> Wrt<V...> wrt(V&... v) { return Wrt<V...>{std::tuple<V&...>(v...)}; }
template<class... A> At<A...> at(A&... a) { return At<A...>{std::tuple<A&...>(a...)}; }

namespace detail {
    // seed the k-th order slot of x: order 1 -> x.grad(innermost val)=1 ; order 2 -> x.val.grad = 1
    inline void seed(dual& x, int /*order*/, double s) { x.grad = s; }
    inline void seed(dual2nd& x, int order, double s) {
        if (order == 1) x.grad.val = s; else x.val.grad = s;
    }
    inline std::array<double, 2> collect(const dual& u) { return {u.val, u.grad}; }
    inline std::array<double, 3> collect(const dual2nd& u) { return {u.val.val, u.grad.val, u.grad.grad}; }
}

// derivatives(f, wrt(x), at(x)) and derivatives(f, wrt(x, x), at(x)) -> {u, du/dx, d2u/dx2}
template<class F, class... V, class... A>
auto derivatives(const F& f, Wrt<V...> w, At<A...> a) {
    auto& x0 = std::get<0>(w.vars);
    auto& x1 = std::get<sizeof...(V) - 1>(w.vars);
    detail::seed(x0, 1, 1.0);
    detail::seed(x1, 2, 1.0);
    auto u = std::apply(f, a.args);
    detail::seed(x0, 1, 0.0);
    detail::seed(x1, 2, 0.0);
    return detail::collect(u);
}

template<class F, class... V, class... A>
double derivative(const F& f, Wrt<V...> w, At<A...> a) {
    return derivatives(f, w, a)[1];
}

} // namespace autodiff
