// oracle/stubs/Aboria.h -- STAND-IN for the Aboria particle library (test infrastructure only).
//
// The reference's calculators (irvingkirkwood/IrvingKirkwood.h, thermo/ThermoCalculator.h, ParticleSystem.cc, ...) are
// compiled UNMODIFIED from /root/reference against this header so that their loop nests, functors and observables run
// as written.  Real Aboria (pyglasstools/extern/Aboria @4ffdeb1) needs Boost, which this image does not have; this file
// supplies only the names those headers use:
//   ABORIA_VARIABLE, Aboria::vdouble3 / vbool3, Aboria::Particles<std::tuple<...>> (size, [], begin/end, get_query,
//   init_neighbour_search), Aboria::get<V>(particle | container), Aboria::id, Aboria::euclidean_search(...).dx().
// Semantics restated from Aboria's sources (what is restated is exactly what stays "unpinned"):
//   * init_neighbour_search -> enforce_domain (NeighbourSearchBase.h:208-237): positions on periodic axes are wrapped
//     into [low, high) by repeated +-(high - low); particles that are non-finite or outside a non-periodic axis die;
//   * euclidean_search (Search.h:152-159, 180-231, 438-465; detail/Distance.h norm2): for every image offset
//     k in {-1,0,1} per periodic axis, every live particle p with dx = p - (c + k (bmax - bmin)) and
//     ((0 + dx0^2) + dx1^2) + dx2^2 <= r^2 (inclusive); dx() returns that vector.
// The enumeration order here (image lattice z, y, x ascending; particles by index) is the order of the oracle's
// brute-force mode, so sums can be compared bit for bit; Aboria's own bucket order differs (summation order only).
#pragma once
#include <cmath>
#include <cstddef>
#include <tuple>
#include <vector>

#define ABORIA_VARIABLE(NAME, TYPE, STR) \
    struct NAME { typedef TYPE value_type; static const char* name() { return STR; } }

namespace Aboria {

template <class T, unsigned N>
struct Vector {
    T v[N];
    Vector() { for (unsigned i = 0; i < N; ++i) v[i] = T(); }
    Vector(T a, T b, T c) { static_assert(N == 3, "3-vector"); v[0] = a; v[1] = b; v[2] = c; }
    T& operator[](unsigned i) { return v[i]; }
    const T& operator[](unsigned i) const { return v[i]; }
    Vector& operator+=(const Vector& o) { for (unsigned i = 0; i < N; ++i) v[i] += o.v[i]; return *this; }
};
typedef Vector<double, 3> vdouble3;
typedef Vector<bool, 3> vbool3;

struct id { typedef std::size_t value_type; };
struct alive { typedef bool value_type; };
struct position_tag { typedef vdouble3 value_type; };

namespace detail {
template <class V, class... Vars> struct index_of;
template <class V, class... Rest> struct index_of<V, V, Rest...> { static const std::size_t value = 0; };
template <class V, class U, class... Rest> struct index_of<V, U, Rest...> { static const std::size_t value = 1 + index_of<V, Rest...>::value; };
}

template <class Tuple> class Particles;

template <class... Vars>
struct particle_ref {                         // a reference to one particle (Aboria hands out tuples of references)
    Particles<std::tuple<Vars...>>* P;
    std::size_t i;
};

template <class... Vars>
class Particles<std::tuple<Vars...>> {
public:
    typedef position_tag position;
    typedef Particles query_type;
    typedef particle_ref<Vars...> value_type;
    struct iterator {
        Particles* P;
        std::size_t i;
        value_type operator*() const { return value_type{ P, i }; }
        iterator& operator++() { ++i; return *this; }
        iterator operator+(std::ptrdiff_t d) const { return iterator{ P, (std::size_t)((std::ptrdiff_t)i + d) }; }
        bool operator!=(const iterator& o) const { return i != o.i; }
        bool operator==(const iterator& o) const { return i == o.i; }
    };

    explicit Particles(std::size_t n = 0) { resize(n); }
    std::size_t size() const { return pos.size(); }
    value_type operator[](std::size_t i) { return value_type{ this, i }; }
    iterator begin() { return iterator{ this, 0 }; }
    iterator end() { return iterator{ this, size() }; }
    const Particles& get_query() const { return *this; }

    void init_neighbour_search(const vdouble3& low, const vdouble3& high, const vbool3& periodic)
    {
        bmin = low; bmax = high; per = periodic;
        std::size_t w = 0;
        for (std::size_t p = 0; p < pos.size(); ++p) {
            vdouble3 r = pos[p];
            bool live = true;
            for (unsigned d = 0; d < 3; ++d) {
                if (!std::isfinite(r[d])) live = false;
                else if (per[d]) {
                    while (r[d] < low[d]) r[d] += (high[d] - low[d]);
                    while (r[d] >= high[d]) r[d] -= (high[d] - low[d]);
                } else if ((r[d] < low[d]) || (r[d] >= high[d])) live = false;
            }
            if (!live) continue;
            pos[w] = r;
            ids[w] = ids[p];
            move_vars(w, p, std::index_sequence_for<Vars...>{});
            ++w;
        }
        resize(w);
    }

    // after positions were edited: the domain is enforced again (Particles::update_positions)
    void update_positions() { init_neighbour_search(bmin, bmax, per); }

    // storage (structure of arrays)
    std::vector<vdouble3> pos;
    std::vector<std::size_t> ids;
    std::tuple<std::vector<typename Vars::value_type>...> vars;
    vdouble3 bmin, bmax;
    vbool3 per;

private:
    void resize(std::size_t n)
    {
        const std::size_t old = pos.size();
        pos.resize(n);
        ids.resize(n);
        for (std::size_t p = old; p < n; ++p) ids[p] = p;
        resize_vars(n, std::index_sequence_for<Vars...>{});
    }
    template <std::size_t... I> void resize_vars(std::size_t n, std::index_sequence<I...>)
    {
        int dummy[] = { (std::get<I>(vars).resize(n), 0)... };
        (void)dummy;
    }
    template <std::size_t... I> void move_vars(std::size_t w, std::size_t p, std::index_sequence<I...>)
    {
        int dummy[] = { (std::get<I>(vars)[w] = std::get<I>(vars)[p], 0)... };
        (void)dummy;
    }
};

// ---- get<Variable>(one particle) / get<Variable>(container)
template <class V, class... Vars>
struct getter {
    static typename V::value_type& of(Particles<std::tuple<Vars...>>& P, std::size_t i)
    {
        return std::get<detail::index_of<V, Vars...>::value>(P.vars)[i];
    }
    static std::vector<typename V::value_type>& all(Particles<std::tuple<Vars...>>& P)
    {
        return std::get<detail::index_of<V, Vars...>::value>(P.vars);
    }
};
template <class... Vars>
struct getter<position_tag, Vars...> {
    static vdouble3& of(Particles<std::tuple<Vars...>>& P, std::size_t i) { return P.pos[i]; }
    static std::vector<vdouble3>& all(Particles<std::tuple<Vars...>>& P) { return P.pos; }
};
template <class... Vars>
struct getter<id, Vars...> {
    static std::size_t& of(Particles<std::tuple<Vars...>>& P, std::size_t i) { return P.ids[i]; }
    static std::vector<std::size_t>& all(Particles<std::tuple<Vars...>>& P) { return P.ids; }
};

template <class V, class... Vars>
auto get(const particle_ref<Vars...>& ref) -> decltype(getter<V, Vars...>::of(*ref.P, ref.i))
{
    return getter<V, Vars...>::of(*ref.P, ref.i);
}
template <class V, class... Vars>
auto get(Particles<std::tuple<Vars...>>& P) -> decltype(getter<V, Vars...>::all(P))
{
    return getter<V, Vars...>::all(P);
}

// ---- euclidean_search: the hits are materialised when the iterator is built
template <class P>
class search_iterator {
public:
    search_iterator(const P& particles, const vdouble3& c, double radius) : m_p(const_cast<P*>(&particles)), m_at(0)
    {
        const double r2 = std::pow(radius, 2);
        const int k0[3] = { particles.per[0] ? -1 : 0, particles.per[1] ? -1 : 0, particles.per[2] ? -1 : 0 };
        const int k1[3] = { particles.per[0] ? 2 : 1, particles.per[1] ? 2 : 1, particles.per[2] ? 2 : 1 };
        for (int kz = k0[2]; kz < k1[2]; ++kz)
            for (int ky = k0[1]; ky < k1[1]; ++ky)
                for (int kx = k0[0]; kx < k1[0]; ++kx) {
                    const int k[3] = { kx, ky, kz };
                    vdouble3 cp;
                    for (unsigned d = 0; d < 3; ++d) cp[d] = c[d] + k[d] * (particles.bmax[d] - particles.bmin[d]);
                    for (std::size_t p = 0; p < particles.size(); ++p) {
                        vdouble3 dx;
                        for (unsigned d = 0; d < 3; ++d) dx[d] = particles.pos[p][d] - cp[d];
                        double accum = 0;
                        for (unsigned d = 0; d < 3; ++d) accum = accum + std::pow(dx[d], 2);   // detail/Distance.h norm2
                        if (!(accum > r2)) { m_idx.push_back(p); m_dx.push_back(dx); }
                    }
                }
    }
    bool operator!=(bool) const { return m_at < m_idx.size(); }      // `p != false`: more candidates
    search_iterator& operator++() { ++m_at; return *this; }
    typename P::value_type operator*() const { return typename P::value_type{ m_p, m_idx[m_at] }; }
    const vdouble3& dx() const { return m_dx[m_at]; }

private:
    P* m_p;
    std::size_t m_at;
    std::vector<std::size_t> m_idx;
    std::vector<vdouble3> m_dx;
};

template <class P>
search_iterator<P> euclidean_search(const P& query, const vdouble3& centre, double radius)
{
    return search_iterator<P>(query, centre, radius);
}

}  // namespace Aboria
