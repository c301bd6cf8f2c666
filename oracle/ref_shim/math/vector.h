/*
 * Stand-in for MVE's <math/vector.h> (nmoehrle/mve, un-vendored dependency of
 * the reference, README.md:24-25).  TEST INFRASTRUCTURE ONLY: lets the
 * reference's own libs/acc headers compile unmodified for oracle/_ref.
 *
 * Only the operations libs/acc/{bvh_tree,kd_tree,primitives}.h use are
 * provided, with MVE's published evaluation order:
 *   dot   = std::inner_product(v, v+N, o.v, T(0))  -> ((0 + a0*b0) + a1*b1) + a2*b2
 *   norm  = std::sqrt(square_norm())
 *   v / s = component-wise true division (no reciprocal multiply)
 *   cross = (a1*b2 - a2*b1, a2*b0 - a0*b2, a0*b1 - a1*b0)
 */
#ifndef ORACLE_REF_SHIM_MATH_VECTOR_H
#define ORACLE_REF_SHIM_MATH_VECTOR_H

#include <cmath>
#include <cstddef>
#include <initializer_list>

namespace math {

template <typename T, int N>
class Vector {
public:
    T v[N];

    Vector() {}
    explicit Vector(T const & fill) { for (int i = 0; i < N; ++i) v[i] = fill; }
    Vector(T const & x, T const & y, T const & z) { static_assert(N == 3, ""); v[0] = x; v[1] = y; v[2] = z; }
    Vector(T const & x, T const & y) { static_assert(N == 2, ""); v[0] = x; v[1] = y; }
    explicit Vector(T const * ptr) { for (int i = 0; i < N; ++i) v[i] = ptr[i]; }

    T & operator[](int i) { return v[i]; }
    T const & operator[](int i) const { return v[i]; }
    T * operator*() { return v; }
    T const * operator*() const { return v; }
    T * begin() { return v; }
    T const * begin() const { return v; }

    T dot(Vector const & o) const {
        T acc(0);
        for (int i = 0; i < N; ++i) acc = acc + v[i] * o.v[i];
        return acc;
    }
    T square_norm() const { return dot(*this); }
    T norm() const { return std::sqrt(square_norm()); }

    Vector<T, 3> cross(Vector const & o) const {
        static_assert(N == 3, "");
        return Vector<T, 3>(v[1] * o.v[2] - v[2] * o.v[1],
                            v[2] * o.v[0] - v[0] * o.v[2],
                            v[0] * o.v[1] - v[1] * o.v[0]);
    }

    Vector operator-() const { Vector r; for (int i = 0; i < N; ++i) r.v[i] = -v[i]; return r; }
    Vector operator+(Vector const & o) const { Vector r; for (int i = 0; i < N; ++i) r.v[i] = v[i] + o.v[i]; return r; }
    Vector operator-(Vector const & o) const { Vector r; for (int i = 0; i < N; ++i) r.v[i] = v[i] - o.v[i]; return r; }
    Vector operator*(T const & s) const { Vector r; for (int i = 0; i < N; ++i) r.v[i] = v[i] * s; return r; }
    Vector operator/(T const & s) const { Vector r; for (int i = 0; i < N; ++i) r.v[i] = v[i] / s; return r; }
    Vector & operator+=(Vector const & o) { for (int i = 0; i < N; ++i) v[i] += o.v[i]; return *this; }
    Vector & operator-=(Vector const & o) { for (int i = 0; i < N; ++i) v[i] -= o.v[i]; return *this; }
    Vector & operator*=(T const & s) { for (int i = 0; i < N; ++i) v[i] *= s; return *this; }
    Vector & operator/=(T const & s) { for (int i = 0; i < N; ++i) v[i] /= s; return *this; }

    /* used by libs/geom (sphere.h, volume.h): MVE's normalized() = *this / norm(), component-wise products / quotients,
     * converting constructor */
    Vector normalized() const { return *this / norm(); }
    Vector cw_mult(Vector const & o) const { Vector r; for (int i = 0; i < N; ++i) r.v[i] = v[i] * o.v[i]; return r; }
    Vector cw_div(Vector const & o) const { Vector r; for (int i = 0; i < N; ++i) r.v[i] = v[i] / o.v[i]; return r; }
    template <typename O> explicit Vector(Vector<O, N> const & o) { for (int i = 0; i < N; ++i) v[i] = T(o[i]); }
};

/* MVE prints a vector as its components separated by single spaces (math/vector.h, operator<<) */
template <typename T, int N, typename Stream>
inline Stream & operator<<(Stream & os, Vector<T, N> const & a) {
    for (int i = 0; i < N - 1; ++i) os << a.v[i] << " ";
    os << a.v[N - 1];
    return os;
}

template <typename T, int N>
inline Vector<T, N> operator*(T const & s, Vector<T, N> const & a) {
    Vector<T, N> r;
    for (int i = 0; i < N; ++i) r.v[i] = a.v[i] * s;   /* MVE: rhs * s */
    return r;
}

typedef Vector<float, 2> Vec2f;
typedef Vector<float, 3> Vec3f;
typedef Vector<float, 4> Vec4f;
typedef Vector<unsigned int, 3> Vec3ui;

}  // namespace math

#endif
