// mrpt_min_part2.h — poses, serialization, config, rtti (see mrpt_min.h)
#pragma once

namespace mrpt
{
// --------------------------------------------------------------------- poses
namespace poses
{
class CPose3D
{
   public:
    CPose3D() { setIdentity(); }
    static CPose3D FromXYZYawPitchRoll(double x, double y, double z, double yaw, double pitch, double roll)
    {
        CPose3D      p;
        const double cy = std::cos(yaw), sy = std::sin(yaw), cp = std::cos(pitch), sp = std::sin(pitch), cr = std::cos(roll),
                     sr = std::sin(roll);
        const double R[9] = {cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr, sy * cp, sy * sp * sr + cy * cr,
                             sy * sp * cr - cy * sr, -sp,     cp * sr,                cp * cr};
        std::memcpy(p.R_, R, sizeof(R));
        p.t_[0] = x, p.t_[1] = y, p.t_[2] = z;
        return p;
    }
    static CPose3D FromRowMajor12(const double* m)
    {
        CPose3D p;
        std::memcpy(p.R_, m, 9 * sizeof(double));
        std::memcpy(p.t_, m + 9, 3 * sizeof(double));
        return p;
    }
    void setIdentity()
    {
        const double I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
        std::memcpy(R_, I, sizeof(I));
        t_[0] = t_[1] = t_[2] = 0;
    }
    void composePoint(double lx, double ly, double lz, double& gx, double& gy, double& gz) const
    {
        gx = R_[0] * lx + R_[1] * ly + R_[2] * lz + t_[0];
        gy = R_[3] * lx + R_[4] * ly + R_[5] * lz + t_[1];
        gz = R_[6] * lx + R_[7] * ly + R_[8] * lz + t_[2];
    }
    mrpt::math::TPoint3Df composePoint(const mrpt::math::TPoint3Df& l) const
    {
        double x, y, z;
        composePoint(l.x, l.y, l.z, x, y, z);
        return {static_cast<float>(x), static_cast<float>(y), static_cast<float>(z)};
    }
    mrpt::math::TPoint3D translation() const { return {t_[0], t_[1], t_[2]}; }
    CPose3D operator+(const CPose3D& b) const
    {
        CPose3D r;
        for (int i = 0; i < 3; i++)
        {
            for (int j = 0; j < 3; j++) r.R_[3 * i + j] = R_[3 * i] * b.R_[j] + R_[3 * i + 1] * b.R_[3 + j] + R_[3 * i + 2] * b.R_[6 + j];
            r.t_[i] = R_[3 * i] * b.t_[0] + R_[3 * i + 1] * b.t_[1] + R_[3 * i + 2] * b.t_[2] + t_[i];
        }
        return r;
    }
    const double* rot() const { return R_; }
    const double* trans() const { return t_; }

   private:
    double R_[9], t_[3];
};
class CPose2D
{
   public:
    CPose2D() = default;
    explicit CPose2D(const CPose3D& p) : x_(p.trans()[0]), y_(p.trans()[1]), phi_(std::atan2(p.rot()[3], p.rot()[0])) {}
    double x() const { return x_; }
    double y() const { return y_; }
    double phi() const { return phi_; }

   private:
    double x_ = 0, y_ = 0, phi_ = 0;
};
}  // namespace poses

// --------------------------------------------------------------------- image
namespace img
{
struct TColorf
{
    float R = 0, G = 0, B = 0, A = 1;
    TColorf() = default;
    TColorf(float r, float g, float b, float a = 1.f) : R(r), G(g), B(b), A(a) {}
};
enum TColormap : int8_t
{
    cmNONE = -1,
    cmGRAYSCALE = 0,
    cmJET,
    cmHOT
};
}  // namespace img

// ------------------------------------------------------------- serialization
namespace serialization
{
class CSerializable;
// Byte-stream archive over an in-memory buffer; fixed-width little-endian PODs like MRPT.
class CArchive
{
   public:
    std::vector<uint8_t> buf;
    size_t               rpos = 0;
    void                 write(const void* p, size_t n)
    {
        const uint8_t* b = static_cast<const uint8_t*>(p);
        buf.insert(buf.end(), b, b + n);
    }
    void read(void* p, size_t n)
    {
        if (rpos + n > buf.size()) throw std::runtime_error("CArchive: read past end");
        std::memcpy(p, buf.data() + rpos, n);
        rpos += n;
    }
    template <typename T, typename U>
    void WriteAs(const U& v)
    {
        const T t = static_cast<T>(v);
        write(&t, sizeof(T));
    }
    template <typename T>
    T ReadAs()
    {
        T t;
        read(&t, sizeof(T));
        return t;
    }
    template <typename STORED, typename CAST>
    void ReadAsAndCastTo(CAST& v)
    {
        v = static_cast<CAST>(ReadAs<STORED>());
    }
};
#define MRPT_MIN_POD_IO(T)                                       \
    inline CArchive& operator<<(CArchive& a, const T& v)         \
    {                                                            \
        a.write(&v, sizeof(T));                                  \
        return a;                                                \
    }                                                            \
    inline CArchive& operator>>(CArchive& a, T& v)               \
    {                                                            \
        a.read(&v, sizeof(T));                                   \
        return a;                                                \
    }
MRPT_MIN_POD_IO(bool)
MRPT_MIN_POD_IO(int8_t)
MRPT_MIN_POD_IO(uint8_t)
MRPT_MIN_POD_IO(int16_t)
MRPT_MIN_POD_IO(uint16_t)
MRPT_MIN_POD_IO(int32_t)
MRPT_MIN_POD_IO(uint32_t)
MRPT_MIN_POD_IO(int64_t)
MRPT_MIN_POD_IO(uint64_t)
MRPT_MIN_POD_IO(float)
MRPT_MIN_POD_IO(double)
#undef MRPT_MIN_POD_IO
inline CArchive& operator<<(CArchive& a, const mrpt::math::TPoint3Df& p) { return a << p.x << p.y << p.z; }
inline CArchive& operator>>(CArchive& a, mrpt::math::TPoint3Df& p) { return a >> p.x >> p.y >> p.z; }
inline CArchive& operator<<(CArchive& a, const mrpt::img::TColorf& c) { return a << c.R << c.G << c.B << c.A; }
inline CArchive& operator>>(CArchive& a, mrpt::img::TColorf& c) { return a >> c.R >> c.G >> c.B >> c.A; }

class CSerializable
{
   public:
    virtual ~CSerializable() = default;
    virtual uint8_t     serializeGetVersion() const                 = 0;
    virtual void        serializeTo(CArchive& out) const            = 0;
    virtual void        serializeFrom(CArchive& in, uint8_t version) = 0;
    virtual const char* className() const { return "CSerializable"; }
};
// MRPT object framing reduced to: class name (u8 length + chars), version byte, payload.
inline CArchive& operator<<(CArchive& a, const CSerializable& o)
{
    const std::string n = o.className();
    a.WriteAs<uint8_t>(n.size());
    a.write(n.data(), n.size());
    a << o.serializeGetVersion();
    o.serializeTo(a);
    return a;
}
inline CArchive& operator>>(CArchive& a, CSerializable& o)
{
    const uint8_t len = a.ReadAs<uint8_t>();
    std::string   n(len, ' ');
    a.read(n.data(), len);
    uint8_t v;
    a >> v;
    o.serializeFrom(a, v);
    return a;
}
}  // namespace serialization

// ---------------------------------------------------------------------- rtti
namespace rtti
{
struct TRuntimeClassId
{
    const char* className;
};
inline void registerClass(const TRuntimeClassId*) {}
}  // namespace rtti
#define CLASS_ID(T) (&T::runtimeClassId())
#define IS_CLASS(obj, T) (dynamic_cast<const T*>(&(obj)) != nullptr)
#define DEFINE_SERIALIZABLE(C, NS)                                                    \
   public:                                                                            \
    using Ptr      = std::shared_ptr<C>;                                              \
    using ConstPtr = std::shared_ptr<const C>;                                        \
    template <typename... Args>                                                       \
    static Ptr Create(Args&&... args)                                                 \
    {                                                                                 \
        return std::make_shared<C>(std::forward<Args>(args)...);                      \
    }                                                                                 \
    static const mrpt::rtti::TRuntimeClassId& runtimeClassId();                       \
    const char*                               className() const override;             \
    uint8_t                                   serializeGetVersion() const override;   \
    void serializeTo(mrpt::serialization::CArchive& out) const override;              \
    void serializeFrom(mrpt::serialization::CArchive& in, uint8_t version) override;  \
                                                                                      \
   private:
#define IMPLEMENTS_SERIALIZABLE(C, BASE, NS)                            \
    const mrpt::rtti::TRuntimeClassId& NS::C::runtimeClassId()          \
    {                                                                   \
        static const mrpt::rtti::TRuntimeClassId id{#NS "::" #C};       \
        return id;                                                      \
    }                                                                   \
    const char* NS::C::className() const { return #NS "::" #C; }
#define MRPT_INITIALIZER(f)              \
    static void f();                     \
    namespace                            \
    {                                    \
    struct f##_caller                    \
    {                                    \
        f##_caller() { f(); }            \
    } f##_instance;                      \
    }                                    \
    static void f()

// -------------------------------------------------------------------- config
namespace config
{
class CConfigFileBase
{
   public:
    virtual ~CConfigFileBase() = default;
    std::map<std::string, std::map<std::string, std::string>> data;
    bool sectionExists(const std::string& s) const { return data.count(s) != 0; }
    void write(const std::string& s, const std::string& k, const std::string& v) { data[s][k] = v; }
    bool has(const std::string& s, const std::string& k) const
    {
        auto it = data.find(s);
        return it != data.end() && it->second.count(k);
    }
    const std::string& raw(const std::string& s, const std::string& k) const { return data.at(s).at(k); }
    template <typename T>
    T read_as(const std::string& s, const std::string& k, const T& def, bool must = false) const
    {
        if (!has(s, k))
        {
            if (must) throw std::runtime_error("missing config key " + s + "/" + k);
            return def;
        }
        std::istringstream ss(raw(s, k));
        double             v;
        ss >> v;
        return static_cast<T>(v);
    }
    double   read_double(const std::string& s, const std::string& k, double d, bool m = false) const { return read_as<double>(s, k, d, m); }
    float    read_float(const std::string& s, const std::string& k, float d, bool m = false) const { return read_as<float>(s, k, d, m); }
    int      read_int(const std::string& s, const std::string& k, int d, bool m = false) const { return read_as<int>(s, k, d, m); }
    uint64_t read_uint64_t(const std::string& s, const std::string& k, uint64_t d, bool m = false) const { return read_as<uint64_t>(s, k, d, m); }
    bool     read_bool(const std::string& s, const std::string& k, bool d, bool m = false) const
    {
        if (!has(s, k)) return d;
        const std::string& v = raw(s, k);
        return v == "1" || v == "true" || v == "True" || v == "yes";
    }
    template <typename E>
    E read_enum(const std::string& s, const std::string& k, const E& d) const
    {
        return static_cast<E>(read_as<int>(s, k, static_cast<int>(d)));
    }
};
using CConfigFileMemory = CConfigFileBase;
class CLoadableOptions
{
   public:
    virtual ~CLoadableOptions() = default;
    virtual void loadFromConfigFile(const CConfigFileBase& source, const std::string& section) = 0;
    virtual void dumpToTextStream(std::ostream&) const {}
};
}  // namespace config
#define MRPT_LOAD_CONFIG_VAR(var, type, cfg, sect) var = (cfg).read_##type(sect, #var, var)
#define MRPT_LOAD_CONFIG_VAR_CS(var, type) MRPT_LOAD_CONFIG_VAR(var, type, c, s)
#define LOADABLEOPTS_DUMP_VAR(var, type) out << #var << " = " << static_cast<double>(var) << "\n";
}  // namespace mrpt
#include "mrpt_min_part3.h"
