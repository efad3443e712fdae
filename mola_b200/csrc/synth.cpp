// synth.cpp — deterministic synthetic LiDAR scans of a procedural street scene.
//
// Host-only helper (no CUDA, no reference code): the same buffers are fed to the
// CPU oracle and to the GPU library, so parity never depends on this file's
// arithmetic, only on it being deterministic within one run. Shapes follow
// HDL-64 (64 rings x 1875 az = 120,000 pts) and OS1-128
// (128 x 2048 = 262,144 pts), rays cast from a sensor 1.73 m above a ground
// plane between two rows of buildings with clutter boxes, range noise N(0, sigma).
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <thread>
#include <vector>

namespace
{
inline uint64_t splitmix(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
inline double u01(uint64_t h) { return (double)((h >> 11) + 1) * (1.0 / 9007199254740993.0); }

struct Box
{
    float lo[3], hi[3];
};

// Scene = ground z=0 plus, for every 10 m street segment s (along x): one
// building per side and two clutter boxes, all derived from hash(seed, s).
inline void segment_boxes(uint64_t seed, long s, Box* out /*4*/)
{
    const float    x0 = 10.0f * (float)s;
    const uint64_t h0 = splitmix(seed * 0x100000001B3ull + (uint64_t)(s + (1L << 40)));
    uint64_t       h  = h0;
    auto           r  = [&]() {
        h = splitmix(h);
        return (float)u01(h);
    };
    const float offL = 10.f + 4.f * r(), offR = 10.f + 4.f * r();
    const float hL = 6.f + 14.f * r(), hR = 6.f + 14.f * r();
    out[0] = {{x0, offL, 0.f}, {x0 + 10.f, offL + 15.f, hL}};
    out[1] = {{x0, -offR - 15.f, 0.f}, {x0 + 10.f, -offR, hR}};
    for (int k = 0; k < 2; k++)
    {
        const float cx = x0 + 10.f * r(), cy = -8.f + 16.f * r();
        const float sx = 0.5f + 1.5f * r(), sy = 0.5f + 1.5f * r(), sz = 0.5f + 2.0f * r();
        // keep the driving lane (|y| < 2) free
        const float yy = (std::fabs(cy) < 3.5f) ? (cy < 0 ? cy - 3.5f : cy + 3.5f) : cy;
        out[2 + k]     = {{cx - sx, yy - sy, 0.f}, {cx + sx, yy + sy, sz}};
    }
}

inline bool ray_box(const float o[3], const float d[3], const Box& b, float& tmin_out)
{
    float t0 = 0.f, t1 = 1e30f;
    for (int a = 0; a < 3; a++)
    {
        if (std::fabs(d[a]) < 1e-12f)
        {
            if (o[a] < b.lo[a] || o[a] > b.hi[a]) return false;
            continue;
        }
        const float inv = 1.0f / d[a];
        float       ta = (b.lo[a] - o[a]) * inv, tb = (b.hi[a] - o[a]) * inv;
        if (ta > tb) std::swap(ta, tb);
        if (ta > t0) t0 = ta;
        if (tb < t1) t1 = tb;
        if (t0 > t1) return false;
    }
    if (t0 <= 1e-4f) return false;  // origin inside a box: ignore it
    tmin_out = t0;
    return true;
}
}  // namespace

extern "C" {

// Casts rings*az_steps rays (ring-major order; further half-step-offset sweeps
// if some rays miss) until exactly n_out points are produced. Points are in the
// SENSOR frame (x forward); pose12_out receives the sensor pose in the map frame
// (row-major R then t) for a vehicle at (px,py) heading yaw, sensor height 1.73.
size_t mbs_lidar_scan(int rings, int az_steps, float elev_top_deg, float elev_bot_deg, double px,
                      double py, double yaw, uint64_t seed, uint64_t scan_id, float noise_sigma,
                      float max_range, size_t n_out, float* xs, float* ys, float* zs,
                      double* pose12_out)
{
    const double sh = 1.73;
    const double cyaw = std::cos(yaw), syaw = std::sin(yaw);
    if (pose12_out)
    {
        const double T[12] = {cyaw, -syaw, 0, syaw, cyaw, 0, 0, 0, 1, px, py, sh};
        for (int i = 0; i < 12; i++) pose12_out[i] = T[i];
    }
    // boxes of all segments within reach
    const long       s0 = (long)std::floor((px - max_range) / 10.0) - 1,
               s1 = (long)std::floor((px + max_range) / 10.0) + 1;
    std::vector<Box> boxes((size_t)(s1 - s0 + 1) * 4);
    for (long s = s0; s <= s1; s++) segment_boxes(seed, s, &boxes[(size_t)(s - s0) * 4]);

    const size_t         per_sweep = (size_t)rings * (size_t)az_steps;
    size_t               produced  = 0;
    std::vector<float>   bx(per_sweep), by(per_sweep), bz(per_sweep);
    std::vector<uint8_t> ok(per_sweep);
    const float          o[3] = {(float)px, (float)py, (float)sh};
    for (int sweep = 0; sweep < 8 && produced < n_out; sweep++)
    {
        const double az_phase = (sweep == 0) ? 0.0 : (double)sweep / 8.0;
        unsigned     nthr     = std::thread::hardware_concurrency();
        if (nthr == 0) nthr = 1;
        if (nthr > 32) nthr = 32;
        std::vector<std::thread> th;
        for (unsigned t = 0; t < nthr; t++)
            th.emplace_back([&, t] {
                for (size_t r = per_sweep * t / nthr; r < per_sweep * (t + 1) / nthr; r++)
                {
                    const int    ring = (int)(r / (size_t)az_steps), a = (int)(r % (size_t)az_steps);
                    const double el   = (elev_top_deg + (elev_bot_deg - elev_top_deg) * (rings > 1 ? (double)ring / (rings - 1) : 0.0)) *
                                      (M_PI / 180.0);
                    const double az = 2.0 * M_PI * ((double)a + az_phase) / (double)az_steps;
                    // direction in the sensor frame, then rotated by yaw into the map frame
                    const double lx = std::cos(el) * std::cos(az), ly = std::cos(el) * std::sin(az),
                                 lz = std::sin(el);
                    const float d[3] = {(float)(cyaw * lx - syaw * ly), (float)(syaw * lx + cyaw * ly), (float)lz};
                    float       tbest = 1e30f;
                    if (d[2] < -1e-6f) tbest = -o[2] / d[2];
                    for (const Box& b : boxes)
                    {
                        float tb;
                        if (ray_box(o, d, b, tb) && tb < tbest) tbest = tb;
                    }
                    ok[r] = 0;
                    if (tbest < max_range && tbest > 0.5f)
                    {
                        const uint64_t h  = splitmix(splitmix(seed ^ (scan_id * 0xD6E8FEB86659FD93ull)) + r + per_sweep * (uint64_t)sweep);
                        const double   u1 = u01(h), u2 = u01(splitmix(h));
                        const double   g  = std::sqrt(-2.0 * std::log(u1)) * std::cos(2.0 * M_PI * u2);
                        const double   rng = (double)tbest + (double)noise_sigma * g;
                        bx[r] = (float)(rng * lx), by[r] = (float)(rng * ly), bz[r] = (float)(rng * lz);
                        ok[r] = 1;
                    }
                }
            });
        for (auto& t : th) t.join();
        for (size_t r = 0; r < per_sweep && produced < n_out; r++)
            if (ok[r])
            {
                xs[produced] = bx[r], ys[produced] = by[r], zs[produced] = bz[r];
                produced++;
            }
    }
    return produced;
}

// Vehicle pose at arc-length s along the procedural trajectory: gentle curves
// inside the driving lane.
void mbs_trajectory_pose(double s, double* px, double* py, double* yaw)
{
    *px  = s;
    *py  = 1.5 * std::sin(s / 40.0);
    *yaw = std::atan(1.5 / 40.0 * std::cos(s / 40.0));
}
}
