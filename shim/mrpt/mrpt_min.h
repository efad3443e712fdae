// mrpt_min.h — a minimal stand-in for the subset of MRPT 2.x that mola_metric_maps uses.
//
// MRPT is not installed in this environment. This header (written from the public MRPT
// API, no MRPT code) lets (a) the reference's own .cpp files compile unmodified into
// oracle/_ref, and (b) the C++ drop-in host classes in mola_b200/host compile and be
// tested here. Against a real MRPT installation none of this is used: drop shim/ from
// the include path. Arithmetic that matters for parity:
//   CPose3D::composePoint   ((R0*x + R1*y) + R2*z) + t in double, then cast to float
//   mrpt::round             lrint (ties to even)
//   TPoint3D::sqrNorm       (x*x + y*y) + z*z
//   CSimplePointsMap nn_*   brute force, exact, ties -> lowest index, radius test d2 < r2
#pragma once
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <iostream>
#include <limits>
#include <map>
#include <memory>
#include <optional>
#include <set>
#include <sstream>
#include <stdexcept>
#include <string>
#include <typeinfo>
#include <vector>

// ---------------------------------------------------------------- exceptions
#define MRPT_START
#define MRPT_END
#define MRPT_TRY_START
#define MRPT_TRY_END
#define THROW_EXCEPTION(msg) throw std::runtime_error(std::string(msg))
#define THROW_EXCEPTION_FMT(fmt, ...) throw std::runtime_error(mrpt::format(fmt, __VA_ARGS__))
#define ASSERT_(c)                                                  \
    do {                                                            \
        if (!(c)) throw std::runtime_error("Assert failed: " #c);   \
    } while (0)
#define ASSERTMSG_(c, m) ASSERT_(c)
#define ASSERT_EQUAL_(a, b) ASSERT_((a) == (b))
#define ASSERT_GT_(a, b) ASSERT_((a) > (b))
#define ASSERT_LT_(a, b) ASSERT_((a) < (b))
#define ASSERT_NEAR_(a, b, t) ASSERT_(std::abs((a) - (b)) <= (t))
#define ASSERTDEB_LT_(a, b)
#define ASSERTDEB_(c)
#define MRPT_THROW_UNKNOWN_SERIALIZATION_VERSION(v) throw std::runtime_error("unknown serialization version")
#define MRPT_UNUSED_PARAM(x) (void)(x)

namespace mrpt
{
inline std::string format(const char* fmt, ...)
{
    char    buf[2048];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    return std::string(buf);
}
template <typename T>
inline int round(const T v)
{
    return static_cast<int>(lrint(v));
}
template <typename T>
inline T square(const T x)
{
    return x * x;
}
template <typename T, typename K>
inline void keep_min(T& var, const K test_val)
{
    if (test_val < var) var = static_cast<T>(test_val);
}
template <typename T, typename K>
inline void keep_max(T& var, const K test_val)
{
    if (test_val > var) var = static_cast<T>(test_val);
}
template <typename T>
inline const T max3(const T& a, const T& b, const T& c)
{
    return std::max(a, std::max(b, c));
}
template <typename T>
inline const T min3(const T& a, const T& b, const T& c)
{
    return std::min(a, std::min(b, c));
}

// ---------------------------------------------------------------------- math
namespace math
{
template <typename T>
struct TPoint3D_
{
    T x{0}, y{0}, z{0};
    constexpr TPoint3D_() = default;
    constexpr TPoint3D_(T X, T Y, T Z) : x(X), y(Y), z(Z) {}
    template <typename U>
    explicit TPoint3D_(const TPoint3D_<U>& o) : x(static_cast<T>(o.x)), y(static_cast<T>(o.y)), z(static_cast<T>(o.z))
    {
    }
    template <typename U>
    TPoint3D_<U> cast() const
    {
        return TPoint3D_<U>(static_cast<U>(x), static_cast<U>(y), static_cast<U>(z));
    }
    T& operator[](size_t i) { return i == 0 ? x : (i == 1 ? y : z); }
    const T& operator[](size_t i) const { return i == 0 ? x : (i == 1 ? y : z); }
    TPoint3D_ operator+(const TPoint3D_& o) const { return {x + o.x, y + o.y, z + o.z}; }
    TPoint3D_ operator-(const TPoint3D_& o) const { return {x - o.x, y - o.y, z - o.z}; }
    TPoint3D_ operator*(T d) const { return {x * d, y * d, z * d}; }
    TPoint3D_& operator*=(T d)
    {
        x *= d, y *= d, z *= d;
        return *this;
    }
    TPoint3D_& operator+=(const TPoint3D_& o)
    {
        x += o.x, y += o.y, z += o.z;
        return *this;
    }
    bool operator==(const TPoint3D_& o) const { return x == o.x && y == o.y && z == o.z; }
    T sqrNorm() const { return x * x + y * y + z * z; }
    T norm() const { return std::sqrt(sqrNorm()); }
    T sqrDistanceTo(const TPoint3D_& p) const { return square(p.x - x) + square(p.y - y) + square(p.z - z); }
    std::string asString() const { return mrpt::format("[%f %f %f]", (double)x, (double)y, (double)z); }
};
template <typename T, typename S>
inline TPoint3D_<T> operator*(const S d, const TPoint3D_<T>& p)
{
    return {d * p.x, d * p.y, d * p.z};
}
template <typename T>
std::ostream& operator<<(std::ostream& o, const TPoint3D_<T>& p)
{
    return o << p.asString();
}
using TPoint3Df = TPoint3D_<float>;
using TPoint3D  = TPoint3D_<double>;
using TVector3D = TPoint3D_<double>;
using TVector3Df = TPoint3D_<float>;
template <typename T>
struct TPoint2D_
{
    T x{0}, y{0};
};
using TPoint2Df = TPoint2D_<float>;
using TPoint2D  = TPoint2D_<double>;

template <typename T>
struct TBoundingBox_
{
    TPoint3D_<T> min, max;
    TBoundingBox_() = default;
    TBoundingBox_(const TPoint3D_<T>& a, const TPoint3D_<T>& b) : min(a), max(b) {}
    static TBoundingBox_ PlusMinusInfinity()
    {
        const T i = std::numeric_limits<T>::max();
        return {{i, i, i}, {-i, -i, -i}};
    }
    void updateWithPoint(const TPoint3D_<T>& p)
    {
        keep_min(min.x, p.x), keep_min(min.y, p.y), keep_min(min.z, p.z);
        keep_max(max.x, p.x), keep_max(max.y, p.y), keep_max(max.z, p.z);
    }
    std::string asString() const { return min.asString() + "-" + max.asString(); }
};
using TBoundingBoxf = TBoundingBox_<float>;
using TBoundingBox  = TBoundingBox_<double>;

class CHistogram
{
   public:
    CHistogram(double mn, double mx, size_t nb) : min_(mn), max_(mx), bins_(nb, 0) {}
    void add(double x)
    {
        if (x < min_ || x > max_ || bins_.empty()) return;
        size_t b = static_cast<size_t>((x - min_) / (max_ - min_) * bins_.size());
        if (b >= bins_.size()) b = bins_.size() - 1;
        bins_[b] += 1;
        count_++;
    }
    void getHistogramNormalized(std::vector<double>& x, std::vector<double>& h) const
    {
        x.resize(bins_.size()), h.resize(bins_.size());
        for (size_t i = 0; i < bins_.size(); i++)
        {
            x[i] = min_ + (i + 0.5) * (max_ - min_) / bins_.size();
            h[i] = count_ ? bins_[i] / count_ : 0;
        }
    }

   private:
    double              min_, max_, count_ = 0;
    std::vector<double> bins_;
};
template <typename A, typename B>
inline void confidenceIntervalsFromHistogram(const A& x, const A& h, B& lo, B& hi, double ci)
{
    double acc = 0;
    lo = static_cast<B>(x.front()), hi = static_cast<B>(x.back());
    for (size_t i = 0; i < x.size(); i++)
    {
        acc += h[i];
        if (acc >= ci)
        {
            lo = static_cast<B>(x[i]);
            break;
        }
    }
    acc = 0;
    for (size_t i = x.size(); i-- > 0;)
    {
        acc += h[i];
        if (acc >= ci)
        {
            hi = static_cast<B>(x[i]);
            break;
        }
    }
}
}  // namespace math
}  // namespace mrpt
#include "mrpt_min_part2.h"
