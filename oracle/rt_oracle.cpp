// rt_oracle.cpp — CPU ORACLE for the `raytrace` path-tracing loop.  TEST INFRASTRUCTURE ONLY.
//
// A double-precision C++17 restatement of the reference's algorithm (UnaryPlus/raytrace,
// Haskell): the description tree is evaluated recursively, in the reference's order, with
// the reference's arithmetic.  Each function cites the reference file:line it follows.
// Nothing under raytrace_b200/ may link, import or call this file; only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs use it, and
// only as the checker / the timed CPU baseline.
//
// Pinning status:
//   * geometry, camera, integrator, materials, textures, Perlin lattice: restated line by line
//     from /root/reference/src and pinned against the reference's golden PNGs
//     (tests/test_oracle_golden.py: exact miss/light masks, block luminance, silhouette);
//   * RNG STREAM: PARITY UNPINNED.  `random` (SplitMix StdGen) is a third-party package that is
//     not in /root/reference (raytrace.cabal:47: random >= 1.3.0 && < 1.4); mkStdGen / split /
//     next are restated from the published SplitMix algorithm, the Word64→Double conversion is
//     a plain 53-bit one.  No reference test pins any stream value, so images are compared
//     statistically, never bit for bit; the Perlin gradient table (Noise.hs:88-92) inherits this.
//
// Build: see oracle/Makefile (g++ -O3 -march=x86-64-v3 -shared -fPIC).
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "../include/rt_b200.h"  // shared POD description structs (the interface vocabulary only)

namespace {

constexpr double INF = std::numeric_limits<double>::infinity();  // Core.hs:22-23
constexpr double PI = 3.14159265358979323846;

// ---------------------------------------------------------------------------------------------
// linear: V3 Double
// ---------------------------------------------------------------------------------------------
struct V3 {
  double x = 0, y = 0, z = 0;
};
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
inline V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }  // Num (V3 a)
inline V3 operator*(double s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }      // (*^)
inline V3 operator*(V3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }      // (^*)
inline V3 operator/(V3 a, double s) { return {a.x / s, a.y / s, a.z / s}; }      // (^/)
inline double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline double quadrance(V3 a) { return dot(a, a); }
inline double norm(V3 a) { return std::sqrt(quadrance(a)); }
inline V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline V3 from3(const double* p) { return {p[0], p[1], p[2]}; }
// linear's `normalize`: leaves the vector alone when |v|² is within 1e-12 of 0 or 1.
inline V3 normalize(V3 v) {
  double l = quadrance(v);
  if (std::fabs(l) <= 1e-12 || std::fabs(1 - l) <= 1e-12) return v;
  return v / std::sqrt(l);
}
inline double comp(V3 v, int d) { return d == 0 ? v.x : (d == 1 ? v.y : v.z); }

// ---------------------------------------------------------------------------------------------
// random: StdGen = SplitMix (restated from the published algorithm; PARITY UNPINNED)
// ---------------------------------------------------------------------------------------------
struct StdGen {
  uint64_t seed, gamma;
};
inline uint64_t mix64(uint64_t z) {
  z = (z ^ (z >> 33)) * 0xff51afd7ed558ccdULL;
  z = (z ^ (z >> 33)) * 0xc4ceb9fe1a85ec53ULL;
  return z ^ (z >> 33);
}
inline uint64_t mix64variant13(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
  return z ^ (z >> 31);
}
inline uint64_t mixGamma(uint64_t z) {
  uint64_t g = mix64variant13(z) | 1ULL;
  int n = __builtin_popcountll(g ^ (g >> 1));
  return n >= 24 ? g : g ^ 0xaaaaaaaaaaaaaaaaULL;
}
inline StdGen mkStdGen(uint64_t s) { return {mix64(s), mixGamma(s + 0x9e3779b97f4a7c15ULL)}; }
inline uint64_t nextWord64(StdGen& g) {
  g.seed += g.gamma;
  return mix64(g.seed);
}
// splitGen: returns the generator kept by the caller's element, advances g to the other half.
inline StdGen splitGen(StdGen& g) {
  uint64_t s1 = g.seed + g.gamma, s2 = s1 + g.gamma;
  StdGen left{s2, g.gamma};
  g = StdGen{mix64(s1), mixGamma(s2)};
  return left;
}
inline double randomDouble(StdGen& g) {  // `random :: Double`, uniform on [0,1)
  return double(nextWord64(g) >> 11) * (1.0 / 9007199254740992.0);
}
inline double randomR(StdGen& g, double lo, double hi) { return lo + (hi - lo) * randomDouble(g); }

// Core.hs:54-60
inline V3 randomUnitVector(StdGen& g) {
  for (;;) {
    V3 v;
    v.x = randomR(g, -1, 1);
    v.y = randomR(g, -1, 1);
    v.z = randomR(g, -1, 1);
    double q = quadrance(v);
    if (1e-8 <= q && q <= 1) return v / std::sqrt(q);
  }
}
// Core.hs:63-68
inline void randomInUnitDisk(StdGen& g, double& x, double& y) {
  for (;;) {
    x = randomR(g, -1, 1);
    y = randomR(g, -1, 1);
    if (x * x + y * y <= 1) return;
  }
}

// ---------------------------------------------------------------------------------------------
// Core.hs: Ray, Interval, Box, HitRecord
// ---------------------------------------------------------------------------------------------
struct Ray {
  V3 o, d;
};
struct Box {
  double lo[3], hi[3];
};
inline Box fromCorners(V3 a, V3 b) {  // Core.hs:112-113
  Box r;
  double pa[3] = {a.x, a.y, a.z}, pb[3] = {b.x, b.y, b.z};
  for (int k = 0; k < 3; k++) {
    if (pa[k] < pb[k]) { r.lo[k] = pa[k]; r.hi[k] = pb[k]; } else { r.lo[k] = pb[k]; r.hi[k] = pa[k]; }
  }
  return r;
}
inline Box boxJoin2(const Box& a, const Box& b) {  // Core.hs:116-117
  Box r;
  for (int k = 0; k < 3; k++) { r.lo[k] = std::min(a.lo[k], b.lo[k]); r.hi[k] = std::max(a.hi[k], b.hi[k]); }
  return r;
}
inline Box boxHull(const std::vector<V3>& pts) {  // Core.hs:120-125
  Box r;
  for (int k = 0; k < 3; k++) { r.lo[k] = INF; r.hi[k] = -INF; }
  for (auto& p : pts) {
    double c[3] = {p.x, p.y, p.z};
    for (int k = 0; k < 3; k++) { r.lo[k] = std::min(r.lo[k], c[k]); r.hi[k] = std::max(r.hi[k], c[k]); }
  }
  return r;
}
inline Box padBox(double pad, Box b) {  // Core.hs:135-136
  for (int k = 0; k < 3; k++) { b.lo[k] -= pad; b.hi[k] += pad; }
  return b;
}
inline Box shiftBox(V3 v, Box b) {  // Core.hs:139-140
  double s[3] = {v.x, v.y, v.z};
  for (int k = 0; k < 3; k++) { b.lo[k] += s[k]; b.hi[k] += s[k]; }
  return b;
}
inline int longestDim(const Box& b) {  // Core.hs:143-144 + argMax Core.hs:43-46
  double x = b.hi[0] - b.lo[0], y = b.hi[1] - b.lo[1], z = b.hi[2] - b.lo[2];
  if (x > y) return x > z ? 0 : 2;
  return y > z ? 1 : 2;
}
// Haskell's default Ord methods on Double: max x y = if x <= y then y else x; min x y = if x <= y then x else y
inline double hsMax(double x, double y) { return x <= y ? y : x; }
inline double hsMin(double x, double y) { return x <= y ? x : y; }
// Core.hs:95-106,147-152
inline bool isect(double& a, double& b, double c, double d) {
  double imin = hsMax(a, c), imax = hsMin(b, d);
  if (imin > imax) return false;
  a = imin; b = imax;
  return true;
}
inline void overlapsInterval(double lo, double hi, double x, double d, double& t0o, double& t1o) {
  double t0 = (lo - x) / d, t1 = (hi - x) / d;
  if (t0 < t1) { t0o = t0; t1o = t1; } else { t0o = t1; t1o = t0; }
}
inline bool overlapsBox(const Box& b, const Ray& r, double tmin, double tmax) {
  double a, c;
  overlapsInterval(b.lo[0], b.hi[0], r.o.x, r.d.x, a, c);
  if (!isect(tmin, tmax, a, c)) return false;
  overlapsInterval(b.lo[1], b.hi[1], r.o.y, r.d.y, a, c);
  if (!isect(tmin, tmax, a, c)) return false;
  overlapsInterval(b.lo[2], b.hi[2], r.o.z, r.d.z, a, c);
  return isect(tmin, tmax, a, c);
}
inline bool inInterval(double tmin, double tmax, double t) { return tmin < t && t < tmax; }  // Core.hs:83-84

struct Material;
struct HitRecord {  // Core.hs:155-162
  double t = 0;
  V3 p, n;
  bool front = false;
  // hr_uv is lazy in the reference (Geometry.hs:90): spheres keep the outward normal and
  // evaluate sphereUV only when a texture reads it.
  bool uv_is_sphere = false;
  V3 uv_src;  // sphere: outward normal; otherwise (u, v, -)
  uint32_t prim = 0xFFFFFFFFu;  // canonical primitive id (oracle bookkeeping, not in the reference)
};
inline void hitUV(const HitRecord& h, double& u, double& v) {
  if (h.uv_is_sphere) {  // sphereUV, Geometry.hs:100-104
    u = std::atan2(h.uv_src.x, h.uv_src.z) / (2 * PI) + 0.5;
    v = std::acos(-h.uv_src.y) / PI;
  } else {
    u = h.uv_src.x;
    v = h.uv_src.y;
  }
}

struct Counters {
  uint64_t samples = 0, rays = 0, box_tests = 0, sphere_tests = 0, planar_tests = 0, medium_queries = 0;
};
struct Ctx {
  StdGen rng;
  Counters cnt;
  bool skip_media = false;  // probe mode
};

// ---------------------------------------------------------------------------------------------
// Noise.hs
// ---------------------------------------------------------------------------------------------
const int kPermX[256] = {
    179, 60, 35, 16, 220, 94, 67, 236, 106, 112, 65, 166, 83, 101, 246, 140, 219, 186, 113, 88, 153, 70, 34, 63, 157, 210, 212, 188, 54, 74, 23, 161,
    28, 137, 126, 107, 183, 58, 134, 127, 211, 225, 17, 123, 150, 243, 160, 68, 75, 239, 173, 221, 89, 109, 61, 72, 159, 80, 154, 18, 214, 144, 197, 24,
    105, 32, 84, 226, 136, 29, 139, 97, 230, 167, 165, 238, 27, 14, 50, 193, 46, 253, 240, 111, 69, 196, 130, 102, 104, 118, 204, 12, 169, 202, 142, 25,
    245, 215, 149, 138, 185, 48, 223, 247, 47, 98, 143, 26, 87, 251, 103, 52, 234, 232, 218, 205, 92, 228, 162, 85, 122, 191, 242, 164, 129, 192, 255, 231,
    147, 91, 178, 213, 176, 36, 120, 155, 241, 222, 177, 20, 152, 141, 51, 171, 250, 95, 71, 119, 254, 172, 53, 146, 135, 124, 125, 163, 235, 99, 7, 22,
    100, 229, 93, 174, 3, 189, 116, 66, 217, 158, 237, 55, 151, 0, 148, 45, 86, 64, 216, 43, 252, 121, 200, 115, 39, 184, 82, 56, 9, 181, 62, 2,
    81, 209, 44, 79, 19, 110, 41, 10, 194, 15, 132, 224, 249, 96, 233, 117, 49, 203, 5, 37, 11, 59, 168, 114, 90, 131, 31, 145, 40, 206, 13, 187,
    133, 207, 4, 199, 170, 78, 30, 182, 248, 21, 6, 227, 57, 180, 73, 42, 128, 175, 108, 33, 244, 201, 198, 77, 195, 8, 38, 190, 76, 156, 208, 1};
const int kPermY[256] = {
    252, 123, 131, 151, 243, 143, 12, 247, 196, 179, 99, 0, 178, 109, 71, 160, 93, 205, 127, 38, 142, 117, 152, 124, 166, 95, 200, 121, 15, 17, 10, 190,
    116, 158, 173, 75, 248, 191, 197, 58, 70, 184, 226, 146, 6, 239, 165, 113, 218, 34, 83, 77, 74, 5, 176, 85, 112, 147, 59, 66, 14, 31, 2, 21,
    198, 108, 255, 11, 36, 156, 96, 73, 1, 189, 126, 50, 220, 16, 249, 23, 139, 135, 141, 92, 159, 4, 119, 174, 171, 253, 86, 227, 251, 172, 7, 149,
    207, 212, 224, 44, 187, 91, 175, 84, 28, 211, 180, 195, 52, 98, 244, 125, 138, 54, 210, 201, 209, 219, 133, 240, 9, 202, 199, 42, 51, 48, 104, 154,
    88, 53, 64, 105, 242, 63, 223, 222, 67, 238, 32, 134, 72, 62, 101, 150, 94, 161, 19, 236, 215, 97, 206, 22, 188, 230, 170, 18, 13, 162, 129, 90,
    246, 130, 35, 46, 43, 213, 29, 76, 241, 61, 30, 136, 235, 87, 114, 68, 183, 177, 250, 132, 3, 122, 110, 145, 81, 78, 232, 69, 60, 80, 216, 65,
    164, 237, 157, 203, 25, 221, 254, 181, 8, 182, 214, 24, 40, 102, 37, 106, 228, 39, 229, 163, 111, 186, 245, 55, 41, 225, 234, 49, 45, 231, 107, 103,
    168, 153, 47, 56, 118, 120, 137, 155, 148, 82, 33, 204, 79, 169, 144, 20, 27, 128, 192, 217, 100, 208, 115, 140, 233, 89, 193, 185, 167, 57, 26, 194};
const int kPermZ[256] = {
    153, 90, 163, 138, 20, 136, 79, 100, 93, 38, 185, 31, 193, 43, 161, 2, 30, 37, 87, 6, 127, 207, 96, 51, 27, 227, 203, 215, 155, 190, 106, 94,
    65, 183, 114, 71, 74, 219, 245, 39, 140, 216, 195, 191, 3, 214, 13, 23, 168, 179, 218, 133, 102, 53, 194, 124, 166, 108, 116, 246, 35, 109, 220, 121,
    205, 110, 83, 242, 252, 231, 25, 128, 61, 57, 187, 222, 189, 228, 148, 101, 239, 162, 48, 150, 55, 174, 178, 42, 200, 160, 58, 206, 11, 92, 204, 67,
    8, 113, 34, 172, 181, 177, 66, 16, 243, 159, 197, 135, 249, 241, 28, 147, 158, 139, 176, 88, 29, 149, 254, 226, 192, 201, 202, 186, 69, 45, 72, 199,
    107, 99, 209, 10, 230, 217, 247, 89, 50, 129, 15, 64, 248, 167, 180, 146, 210, 169, 4, 81, 97, 238, 5, 237, 125, 236, 95, 9, 104, 221, 32, 86,
    165, 134, 224, 182, 198, 234, 253, 59, 164, 52, 60, 208, 103, 22, 46, 188, 54, 12, 1, 112, 144, 137, 47, 156, 98, 78, 18, 82, 175, 225, 19, 184,
    244, 49, 120, 157, 130, 70, 80, 251, 212, 250, 119, 118, 196, 223, 105, 68, 84, 21, 145, 173, 56, 235, 91, 17, 126, 76, 131, 117, 152, 111, 33, 36,
    233, 141, 122, 85, 7, 123, 44, 62, 115, 26, 0, 40, 229, 211, 213, 63, 143, 77, 14, 24, 142, 154, 73, 132, 171, 41, 170, 240, 75, 151, 232, 255};

struct NoiseTables {
  int px[256], py[256], pz[256];
  V3 grad[256];
  NoiseTables() {
    std::memcpy(px, kPermX, sizeof px);
    std::memcpy(py, kPermY, sizeof py);
    std::memcpy(pz, kPermZ, sizeof pz);
    StdGen g = mkStdGen(666);  // Noise.hs:88-92
    for (int i = 0; i < 256; i++) grad[i] = randomUnitVector(g);
  }
};
inline double smoothstep(double x) { return x * x * (3 - 2 * x); }  // Noise.hs:15-16
// Noise.hs:21-39
double perlinNoise(const NoiseTables& T, V3 p) {
  double flx = std::floor(p.x), fly = std::floor(p.y), flz = std::floor(p.z);
  int64_t ix = (int64_t)flx, iy = (int64_t)fly, iz = (int64_t)flz;
  double fx = p.x - flx, fy = p.y - fly, fz = p.z - flz;
  double acc = 0;
  for (int i = 0; i < 2; i++)
    for (int j = 0; j < 2; j++)
      for (int k = 0; k < 2; k++) {
        double id = i, jd = j, kd = k;
        int h = T.px[(ix + i) & 255] ^ T.py[(iy + j) & 255] ^ T.pz[(iz + k) & 255];
        V3 rel{fx - id, fy - jd, fz - kd};
        double coef = smoothstep(id * fx + (1 - id) * (1 - fx)) * smoothstep(jd * fy + (1 - jd) * (1 - fy)) *
                      smoothstep(kd * fz + (1 - kd) * (1 - fz));
        acc += coef * dot(T.grad[h], rel);
      }
  return acc;
}
// Noise.hs:42-46
double fractalNoise(const NoiseTables& T, int depth, V3 p) {
  double acc = 0, coef = 1;
  for (int o = 0; o < depth; o++) {
    acc += coef * perlinNoise(T, p);
    coef /= 2;
    p = p * 2.0;
  }
  return acc;
}
inline double turbulence(const NoiseTables& T, int depth, V3 p) { return std::fabs(fractalNoise(T, depth, p)); }  // Noise.hs:49-50

// ---------------------------------------------------------------------------------------------
// Texture.hs
// ---------------------------------------------------------------------------------------------
inline int64_t floorMod(int64_t a, int64_t m) {  // Haskell `mod`
  int64_t r = a % m;
  return r < 0 ? r + m : r;
}
struct Texture {
  int kind = RT_TEX_CONSTANT;
  int i0 = 0, i1 = 0, w = 0, h = 0;
  double freq = 0;
  V3 c0, c1, shift, dir;
  std::vector<double> image;  // linear RGB doubles (readImage, Ray.hs:241-245)
  const NoiseTables* noise = nullptr;

  V3 eval(V3 p, const HitRecord& hit) const {
    switch (kind) {
      case RT_TEX_CONSTANT: return c0;  // Texture.hs:18-19
      case RT_TEX_CHECKER: {            // Texture.hs:45-53
        double u, v;
        hitUV(hit, u, v);
        int64_t i = (int64_t)std::floor(u * (double)i0), j = (int64_t)std::floor(v * (double)i1);
        return ((i + j) & 1) == 0 ? c0 : c1;
      }
      case RT_TEX_IMAGE: {              // Texture.hs:31-41
        double u, v;
        hitUV(hit, u, v);
        int64_t i = floorMod((int64_t)std::floor(u * (double)w), w);
        int64_t j = floorMod((int64_t)std::floor((1 - v) * (double)h), h);
        const double* t = &image[(size_t)(j * w + i) * 3];
        return {t[0], t[1], t[2]};
      }
      case RT_TEX_NOISE: {              // Texture.hs:56-67
        double scale = 0.5 / 0.8;
        double n = fractalNoise(*noise, i0, p * freq + shift) * scale + 0.5;
        return c0 + (c1 - c0) * n;
      }
      case RT_TEX_MARBLE: {             // Texture.hs:70-79
        double sinArg = freq * dot(dir, p);
        double n = 10 * turbulence(*noise, 7, (0.25 * freq) * p + shift);
        double g = 0.5 + 0.5 * std::sin(sinArg + n);
        return {g, g, g};
      }
    }
    return {};
  }
};

// ---------------------------------------------------------------------------------------------
// Material.hs
// ---------------------------------------------------------------------------------------------
enum ResKind { ABSORB, SCATTER, HEMISPHERE_F, SPHERE_F };
struct Material {
  int kind = RT_MAT_PITCH_BLACK;
  const Texture* tex = nullptr;
  double param = 0;
};
struct MatResult {
  ResKind kind = ABSORB;
  V3 attenuation, dir;  // Scatter
};
inline V3 reflect(V3 n, V3 v) { return v - (2 * dot(n, v)) * n; }  // Core.hs:49-51
// Material.hs:81-85
inline V3 refract(double iorRatio, double cosTheta, V3 n, V3 u) {
  V3 perp = iorRatio * (u + cosTheta * n);
  V3 para = -(std::sqrt(std::fabs(1 - quadrance(perp))) * n);
  return perp + para;
}
// material application: emitted colour + State action (Material.hs:41-129)
inline MatResult evalMaterial(const Material& m, V3 inDir, const HitRecord& h, StdGen& rng, V3& emitted) {
  MatResult r;
  emitted = {};
  switch (m.kind) {
    case RT_MAT_LIGHT: emitted = m.tex->eval(h.p, h); r.kind = ABSORB; break;      // :41-42
    case RT_MAT_PITCH_BLACK: r.kind = ABSORB; break;                               // :46-47
    case RT_MAT_LAMBERTIAN:                                                        // :51-53
    case RT_MAT_LOMMEL_SEELIGER: r.kind = HEMISPHERE_F; break;                     // :56-61
    case RT_MAT_MIRROR:                                                            // :64-67
      r.kind = SCATTER; r.attenuation = m.tex->eval(h.p, h); r.dir = reflect(h.n, inDir); break;
    case RT_MAT_METAL: {                                                           // :72-78
      V3 u = randomUnitVector(rng);
      V3 d2 = reflect(h.n, inDir) + m.param * u;
      if (dot(d2, h.n) > 0) { r.kind = SCATTER; r.attenuation = m.tex->eval(h.p, h); r.dir = normalize(d2); }
      else r.kind = ABSORB;
      break;
    }
    case RT_MAT_DIELECTRIC: {                                                      // :89-106
      double iorRatio = h.front ? 1 / m.param : m.param;
      double cosTheta = std::min(1.0, dot(h.n, -inDir));
      double sinTheta = std::sqrt(1 - cosTheta * cosTheta);
      bool cannotRefract = iorRatio * sinTheta > 1;
      double r0 = (1 - iorRatio) / (1 + iorRatio);
      double r0s = r0 * r0;
      double reflectance = r0s + (1 - r0s) * std::pow(1 - cosTheta, 5.0);
      double x = randomDouble(rng);
      r.kind = SCATTER; r.attenuation = {1, 1, 1};
      r.dir = (cannotRefract || x < reflectance) ? reflect(h.n, inDir) : refract(iorRatio, cosTheta, h.n, inDir);
      break;
    }
    case RT_MAT_TRANSPARENT:                                                       // :109-112
      r.kind = SCATTER; r.attenuation = m.tex->eval(h.p, h); r.dir = inDir; break;
    case RT_MAT_ISOTROPIC:                                                         // :116-118
    case RT_MAT_ANISOTROPIC: r.kind = SPHERE_F; break;                             // :124-129
  }
  return r;
}
// the function carried by HemisphereF / SphereF
inline V3 evalMatF(const Material& m, V3 inDir, const HitRecord& h, V3 outDir) {
  V3 c = m.tex->eval(h.p, h);
  switch (m.kind) {
    case RT_MAT_LOMMEL_SEELIGER: {
      double mu0 = -dot(inDir, h.n), mu1 = dot(outDir, h.n);
      return (0.25 / (mu0 + mu1)) * c;
    }
    case RT_MAT_ANISOTROPIC: {
      double g = m.param, mu = dot(inDir, outDir);
      double hg = (1 - g * g) / std::pow(1 + g * g - 2 * g * mu, 1.5);
      return hg * c;
    }
    default: return c;
  }
}

// ---------------------------------------------------------------------------------------------
// Geometry.hs — closures become virtual calls; evaluation order is the reference's
// ---------------------------------------------------------------------------------------------
struct Geom {
  Box bbox{};
  virtual ~Geom() = default;
  // Double -> Ray -> Interval -> m (Maybe (HitRecord, a))
  virtual bool hit(double time, const Ray& ray, double tmin, double tmax, Ctx& ctx, HitRecord& out,
                   const Material*& mat) const = 0;
  virtual void number(uint32_t& next) { (void)next; }          // canonical ids: DFS, first visit
  virtual void leafOrder(std::vector<uint32_t>& out) const { (void)out; }  // evaluation-order leaf ids
};
using GeomP = std::shared_ptr<Geom>;

struct Sphere final : Geom {  // Geometry.hs:58-94
  V3 center;
  double radius;
  uint32_t id = 0xFFFFFFFFu;
  Sphere(V3 c, double r) : center(c), radius(r) {
    V3 d{r, r, r};
    bbox = fromCorners(c - d, c + d);
  }
  bool hit(double, const Ray& ray, double tmin, double tmax, Ctx& ctx, HitRecord& out, const Material*&) const override {
    ctx.cnt.sphere_tests++;
    V3 oc = center - ray.o;
    double h = dot(ray.d, oc);
    double c = quadrance(oc) - radius * radius;
    double disc = h * h - c;
    if (!(disc >= 0)) return false;
    double sq = std::sqrt(disc);
    double root1 = h - sq, root2 = h + sq, t;
    if (inInterval(tmin, tmax, root1)) t = root1;
    else if (inInterval(tmin, tmax, root2)) t = root2;
    else return false;
    V3 p = ray.o + t * ray.d;
    V3 on = (p - center) / radius;
    bool front = dot(ray.d, on) <= 0;
    out.t = t; out.p = p; out.n = front ? on : -on; out.front = front;
    out.uv_is_sphere = true; out.uv_src = on; out.prim = id;
    return true;
  }
  void number(uint32_t& next) override { if (id == 0xFFFFFFFFu) id = next++; }
  void leafOrder(std::vector<uint32_t>& o) const override { o.push_back(id); }
};

struct PlaneShape final : Geom {  // Geometry.hs:109-144 with the two library shapes
  V3 q, u, v, normal, normalS;
  bool is_triangle;
  double uv0[2], uv1[2], uv2[2];
  uint32_t id = 0xFFFFFFFFu;
  PlaneShape(V3 q_, V3 u_, V3 v_, bool tri, const Box& hull) : q(q_), u(u_), v(v_), is_triangle(tri) {
    V3 cp = cross(u, v);
    double ncp = norm(cp);
    normal = cp / ncp;
    normalS = normal / ncp;
    bbox = padBox(0.0001, hull);
    uv0[0] = uv0[1] = uv1[1] = uv2[0] = 0; uv1[0] = uv2[1] = 1;
  }
  bool hit(double, const Ray& ray, double tmin, double tmax, Ctx& ctx, HitRecord& out, const Material*&) const override {
    ctx.cnt.planar_tests++;
    double denom = dot(normal, ray.d);
    if (!(std::fabs(denom) > 1e-8)) return false;
    double t = dot(normal, q - ray.o) / denom;
    if (!inInterval(tmin, tmax, t)) return false;
    V3 p = ray.o + t * ray.d;
    V3 pr = p - q;
    double a = dot(normalS, cross(pr, v));
    double b = dot(normalS, cross(u, pr));
    if (is_triangle) { if (!(a >= 0 && b >= 0 && a + b <= 1)) return false; }   // Geometry.hs:174
    else { if (!(0 <= a && a <= 1 && 0 <= b && b <= 1)) return false; }         // Geometry.hs:150
    bool front = denom < 0;
    out.t = t; out.p = p; out.n = front ? normal : -normal; out.front = front;
    out.uv_is_sphere = false;
    if (is_triangle) {  // Geometry.hs:175
      double c0 = 1 - a - b;
      out.uv_src = {c0 * uv0[0] + a * uv1[0] + b * uv2[0], c0 * uv0[1] + a * uv1[1] + b * uv2[1], 0};
    } else out.uv_src = {a, b, 0};
    out.prim = id;
    return true;
  }
  void number(uint32_t& next) override { if (id == 0xFFFFFFFFu) id = next++; }
  void leafOrder(std::vector<uint32_t>& o) const override { o.push_back(id); }
};
inline GeomP mkParallelogram(V3 q, V3 u, V3 v) {  // Geometry.hs:147-151
  return std::make_shared<PlaneShape>(q, u, v, false, boxHull({q, q + u, q + v, q + u + v}));
}
inline GeomP mkTriangle(V3 p0, V3 p1, V3 p2, const double* uv0, const double* uv1, const double* uv2) {  // :169-176
  auto t = std::make_shared<PlaneShape>(p0, p1 - p0, p2 - p0, true, boxHull({p0, p1, p2}));
  for (int k = 0; k < 2; k++) { t->uv0[k] = uv0[k]; t->uv1[k] = uv1[k]; t->uv2[k] = uv2[k]; }
  return t;
}

struct Group final : Geom {  // Geometry.hs:336-347
  std::vector<GeomP> obs;
  explicit Group(std::vector<GeomP> o) : obs(std::move(o)) {
    bbox = obs[0]->bbox;
    for (size_t i = 1; i < obs.size(); i++) bbox = boxJoin2(bbox, obs[i]->bbox);
  }
  bool hit(double time, const Ray& ray, double tmin, double tmax, Ctx& ctx, HitRecord& out, const Material*& mat) const override {
    bool any = false;
    HitRecord h;
    const Material* m = nullptr;
    for (auto& o : obs) {
      if (o->hit(time, ray, tmin, tmax, ctx, h, m)) { any = true; tmax = h.t; out = h; mat = m; }
    }
    return any;
  }
  void number(uint32_t& next) override { for (auto& o : obs) o->number(next); }
  void leafOrder(std::vector<uint32_t>& out) const override { for (auto& o : obs) o->leafOrder(out); }
};

struct BvhNode final : Geom {  // Geometry.hs:352-363
  GeomP l, r;
  BvhNode(GeomP l_, GeomP r_) : l(std::move(l_)), r(std::move(r_)) { bbox = boxJoin2(l->bbox, r->bbox); }
  bool hit(double time, const Ray& ray, double tmin, double tmax, Ctx& ctx, HitRecord& out, const Material*& mat) const override {
    ctx.cnt.box_tests++;
    if (!overlapsBox(bbox, ray, tmin, tmax)) return false;
    if (!l->hit(time, ray, tmin, tmax, ctx, out, mat)) return r->hit(time, ray, tmin, tmax, ctx, out, mat);
    HitRecord h2;
    const Material* m2 = nullptr;
    if (r->hit(time, ray, tmin, out.t, ctx, h2, m2)) { out = h2; mat = m2; }   // (<|> res): right wins only if it hit
    return true;
  }
  void number(uint32_t& next) override { l->number(next); r->number(next); }
  void leafOrder(std::vector<uint32_t>& out) const override { l->leafOrder(out); r->leafOrder(out); }
};
// Geometry.hs:369-377
GeomP bvhTree(std::vector<GeomP> obs) {
  if (obs.size() == 1) return obs[0];
  Box all = obs[0]->bbox;
  for (size_t i = 1; i < obs.size(); i++) all = boxJoin2(all, obs[i]->bbox);
  int d = longestDim(all);
  std::stable_sort(obs.begin(), obs.end(), [d](const GeomP& a, const GeomP& b) {  // sortOn is stable
    return (a->bbox.lo[d] + a->bbox.hi[d]) / 2 < (b->bbox.lo[d] + b->bbox.hi[d]) / 2;
  });
  size_t n = obs.size() / 2;
  std::vector<GeomP> left(obs.begin(), obs.begin() + n), right(obs.begin() + n, obs.end());
  return std::make_shared<BvhNode>(bvhTree(std::move(left)), bvhTree(std::move(right)));
}
// A bvhTree/mesh/cuboid description node numbers its children in LIST order (before sorting) but
// evaluates the built tree; this wrapper keeps both.
struct Built final : Geom {
  std::vector<GeomP> listOrder;
  GeomP tree;
  Built(std::vector<GeomP> lst, GeomP t) : listOrder(std::move(lst)), tree(std::move(t)) { bbox = tree->bbox; }
  bool hit(double time, const Ray& ray, double tmin, double tmax, Ctx& ctx, HitRecord& out, const Material*& mat) const override {
    return tree->hit(time, ray, tmin, tmax, ctx, out, mat);
  }
  void number(uint32_t& next) override { for (auto& o : listOrder) o->number(next); }
  void leafOrder(std::vector<uint32_t>& out) const override { tree->leafOrder(out); }
};

struct M34 {
  double a[3][4];
  V3 point(V3 p) const { return {a[0][0] * p.x + a[0][1] * p.y + a[0][2] * p.z + a[0][3], a[1][0] * p.x + a[1][1] * p.y + a[1][2] * p.z + a[1][3], a[2][0] * p.x + a[2][1] * p.y + a[2][2] * p.z + a[2][3]}; }
  V3 vector(V3 p) const { return {a[0][0] * p.x + a[0][1] * p.y + a[0][2] * p.z, a[1][0] * p.x + a[1][1] * p.y + a[1][2] * p.z, a[2][0] * p.x + a[2][1] * p.y + a[2][2] * p.z}; }
};
// inv44 (linear): general 4x4 inverse by cofactors
bool inv44(const double* m, double* inv) {
  inv[0] = m[5] * m[10] * m[15] - m[5] * m[11] * m[14] - m[9] * m[6] * m[15] + m[9] * m[7] * m[14] + m[13] * m[6] * m[11] - m[13] * m[7] * m[10];
  inv[4] = -m[4] * m[10] * m[15] + m[4] * m[11] * m[14] + m[8] * m[6] * m[15] - m[8] * m[7] * m[14] - m[12] * m[6] * m[11] + m[12] * m[7] * m[10];
  inv[8] = m[4] * m[9] * m[15] - m[4] * m[11] * m[13] - m[8] * m[5] * m[15] + m[8] * m[7] * m[13] + m[12] * m[5] * m[11] - m[12] * m[7] * m[9];
  inv[12] = -m[4] * m[9] * m[14] + m[4] * m[10] * m[13] + m[8] * m[5] * m[14] - m[8] * m[6] * m[13] - m[12] * m[5] * m[10] + m[12] * m[6] * m[9];
  inv[1] = -m[1] * m[10] * m[15] + m[1] * m[11] * m[14] + m[9] * m[2] * m[15] - m[9] * m[3] * m[14] - m[13] * m[2] * m[11] + m[13] * m[3] * m[10];
  inv[5] = m[0] * m[10] * m[15] - m[0] * m[11] * m[14] - m[8] * m[2] * m[15] + m[8] * m[3] * m[14] + m[12] * m[2] * m[11] - m[12] * m[3] * m[10];
  inv[9] = -m[0] * m[9] * m[15] + m[0] * m[11] * m[13] + m[8] * m[1] * m[15] - m[8] * m[3] * m[13] - m[12] * m[1] * m[11] + m[12] * m[3] * m[9];
  inv[13] = m[0] * m[9] * m[14] - m[0] * m[10] * m[13] - m[8] * m[1] * m[14] + m[8] * m[2] * m[13] + m[12] * m[1] * m[10] - m[12] * m[2] * m[9];
  inv[2] = m[1] * m[6] * m[15] - m[1] * m[7] * m[14] - m[5] * m[2] * m[15] + m[5] * m[3] * m[14] + m[13] * m[2] * m[7] - m[13] * m[3] * m[6];
  inv[6] = -m[0] * m[6] * m[15] + m[0] * m[7] * m[14] + m[4] * m[2] * m[15] - m[4] * m[3] * m[14] - m[12] * m[2] * m[7] + m[12] * m[3] * m[6];
  inv[10] = m[0] * m[5] * m[15] - m[0] * m[7] * m[13] - m[4] * m[1] * m[15] + m[4] * m[3] * m[13] + m[12] * m[1] * m[7] - m[12] * m[3] * m[5];
  inv[14] = -m[0] * m[5] * m[14] + m[0] * m[6] * m[13] + m[4] * m[1] * m[14] - m[4] * m[2] * m[13] - m[12] * m[1] * m[6] + m[12] * m[2] * m[5];
  inv[3] = -m[1] * m[6] * m[11] + m[1] * m[7] * m[10] + m[5] * m[2] * m[11] - m[5] * m[3] * m[10] - m[9] * m[2] * m[7] + m[9] * m[3] * m[6];
  inv[7] = m[0] * m[6] * m[11] - m[0] * m[7] * m[10] - m[4] * m[2] * m[11] + m[4] * m[3] * m[10] + m[8] * m[2] * m[7] - m[8] * m[3] * m[6];
  inv[11] = -m[0] * m[5] * m[11] + m[0] * m[7] * m[9] + m[4] * m[1] * m[11] - m[4] * m[3] * m[9] - m[8] * m[1] * m[7] + m[8] * m[3] * m[5];
  inv[15] = m[0] * m[5] * m[10] - m[0] * m[6] * m[9] - m[4] * m[1] * m[10] + m[4] * m[2] * m[9] + m[8] * m[1] * m[6] - m[8] * m[2] * m[5];
  double det = m[0] * inv[0] + m[1] * inv[4] + m[2] * inv[8] + m[3] * inv[12];
  if (det == 0) return false;
  det = 1.0 / det;
  for (int i = 0; i < 16; i++) inv[i] *= det;
  return true;
}

struct Transform final : Geom {  // Geometry.hs:382-391
  M34 m34, inv_m;
  GeomP child;
  Transform(const double* m, GeomP c) : child(std::move(c)) {
    double inv[16];
    inv44(m, inv);
    for (int r = 0; r < 3; r++) for (int k = 0; k < 4; k++) { m34.a[r][k] = m[r * 4 + k]; inv_m.a[r][k] = inv[r * 4 + k]; }
    std::vector<V3> corners;
    const Box& b = child->bbox;  // allCorners, Core.hs:128-132
    for (int i = 0; i < 2; i++) for (int j = 0; j < 2; j++) for (int k = 0; k < 2; k++)
      corners.push_back(m34.point({i ? b.hi[0] : b.lo[0], j ? b.hi[1] : b.lo[1], k ? b.hi[2] : b.lo[2]}));
    bbox = boxHull(corners);
  }
  bool hit(double time, const Ray& ray, double tmin, double tmax, Ctx& ctx, HitRecord& out, const Material*& mat) const override {
    Ray r2{inv_m.point(ray.o), inv_m.vector(ray.d)};
    if (!child->hit(time, r2, tmin, tmax, ctx, out, mat)) return false;
    out.p = m34.point(out.p);
    out.n = m34.vector(out.n);
    return true;
  }
  void number(uint32_t& next) override { child->number(next); }
  void leafOrder(std::vector<uint32_t>& out) const override { child->leafOrder(out); }
};

struct Moving final : Geom {  // Geometry.hs:449-456
  V3 v0, v1;
  GeomP child;
  Moving(V3 a, V3 b, GeomP c) : v0(a), v1(b), child(std::move(c)) { bbox = boxJoin2(shiftBox(v0, child->bbox), shiftBox(v1, child->bbox)); }
  bool hit(double time, const Ray& ray, double tmin, double tmax, Ctx& ctx, HitRecord& out, const Material*& mat) const override {
    V3 shift = (1 - time) * v0 + time * v1;
    Ray r2{ray.o - shift, ray.d};
    if (!child->hit(time, r2, tmin, tmax, ctx, out, mat)) return false;
    out.p = out.p + shift;
    return true;
  }
  void number(uint32_t& next) override { child->number(next); }
  void leafOrder(std::vector<uint32_t>& out) const override { child->leafOrder(out); }
};

struct ConstantMedium final : Geom {  // Geometry.hs:298-330
  double negInvDensity;
  GeomP boundary;
  ConstantMedium(double density, GeomP b) : negInvDensity(-(1 / density)), boundary(std::move(b)) { bbox = boundary->bbox; }
  bool hit(double time, const Ray& ray, double tmin, double tmax, Ctx& ctx, HitRecord& out, const Material*&) const override {
    if (ctx.skip_media) return false;
    ctx.cnt.medium_queries++;
    HitRecord h1, h2;
    const Material* dummy = nullptr;
    double t1, t2;
    if (!boundary->hit(time, ray, tmin, INF, ctx, h1, dummy)) return false;
    if (h1.front) {
      if (!(h1.t < tmax)) return false;
      if (!boundary->hit(time, ray, h1.t, INF, ctx, h2, dummy)) return false;
      t1 = h1.t; t2 = std::min(tmax, h2.t);
    } else { t1 = tmin; t2 = std::min(tmax, h1.t); }
    double rand = randomDouble(ctx.rng);
    double inDist = t2 - t1;
    double hitDist = negInvDensity * std::log(rand);
    if (!(hitDist < inDist)) return false;
    double t = t1 + hitDist;
    out.t = t; out.p = ray.o + t * ray.d; out.n = -ray.d; out.front = true;
    out.uv_is_sphere = false; out.uv_src = {0, 0, 0}; out.prim = 0xFFFFFFFEu;
    return true;
  }
  void number(uint32_t& next) override { boundary->number(next); }
};

struct WithMaterial final : Geom {  // `mat <$ g`, Geometry.hs:44-47
  const Material* m;
  GeomP child;
  WithMaterial(const Material* m_, GeomP c) : m(m_), child(std::move(c)) { bbox = child->bbox; }
  bool hit(double time, const Ray& ray, double tmin, double tmax, Ctx& ctx, HitRecord& out, const Material*& mat) const override {
    if (!child->hit(time, ray, tmin, tmax, ctx, out, mat)) return false;
    mat = m;
    return true;
  }
  void number(uint32_t& next) override { child->number(next); }
  void leafOrder(std::vector<uint32_t>& out) const override { child->leafOrder(out); }
};

// ---------------------------------------------------------------------------------------------
// Scene container
// ---------------------------------------------------------------------------------------------
struct Scene {
  std::vector<std::unique_ptr<Texture>> textures;
  std::vector<std::unique_ptr<Material>> materials;
  std::vector<GeomP> nodes;
  NoiseTables noise;
  GeomP root;
  uint32_t n_prims = 0;
  std::string error;
};

GeomP mkCuboid(V3 lo, V3 hi) {  // Geometry.hs:154-166
  double xmin = lo.x, xmax = hi.x, ymin = lo.y, ymax = hi.y, zmin = lo.z, zmax = hi.z;
  V3 dx{xmax - xmin, 0, 0}, dy{0, ymax - ymin, 0}, dz{0, 0, zmax - zmin};
  std::vector<GeomP> f = {
      mkParallelogram({xmin, ymin, zmax}, dx, dy),    // front
      mkParallelogram({xmax, ymin, zmin}, -dx, dy),   // back
      mkParallelogram({xmin, ymin, zmin}, dz, dy),    // left
      mkParallelogram({xmax, ymin, zmax}, -dz, dy),   // right
      mkParallelogram({xmin, ymax, zmax}, dx, -dz),   // top
      mkParallelogram({xmin, ymin, zmin}, dx, dz),    // bottom
  };
  return std::make_shared<Group>(std::move(f));
}

// ---------------------------------------------------------------------------------------------
// Ray.hs: camera + integrator
// ---------------------------------------------------------------------------------------------
struct RedirectTarget {  // Ray.hs:112-118
  V3 origin, U, V, crs;
  std::shared_ptr<PlaneShape> shape;
  double p;
};
struct Camera {
  int W, H, spp, maxDepth;
  V3 center, topLeft, pixelU, pixelV, defocusU, defocusV;
  int bg_kind;
  V3 bg0, bg1;
  std::vector<RedirectTarget> targets;
  std::vector<double> cumulative;
  double remProb;
};
int imageHeightOf(const RtCamera& c) { return (int)std::nearbyint((double)c.image_width / c.aspect_ratio); }  // Ray.hs:123 (round-half-even)

Camera makeCamera(const RtCamera& c) {  // Ray.hs:123-155
  Camera k;
  k.W = c.image_width;
  k.H = imageHeightOf(c);
  k.spp = c.samples_per_pixel;
  k.maxDepth = c.max_recursion_depth;
  V3 center = from3(c.center), lookAt = from3(c.look_at), up = from3(c.up);
  double viewportHeight = c.focus_dist * std::tan(c.vfov / 2) * 2;
  double viewportWidth = viewportHeight * (double)k.W / (double)k.H;
  V3 w = normalize(center - lookAt);
  V3 u = normalize(cross(up, w));
  V3 v = cross(w, u);
  V3 across = viewportWidth * u;
  V3 down = -(viewportHeight * v);
  k.center = center;
  k.topLeft = center - w * c.focus_dist - across / 2 - down / 2;
  k.pixelU = across / (double)k.W;
  k.pixelV = down / (double)k.H;
  double defocusRadius = c.focus_dist * std::tan(c.defocus_angle / 2);
  k.defocusU = u * defocusRadius;
  k.defocusV = v * defocusRadius;
  k.bg_kind = c.bg_kind; k.bg0 = from3(c.bg_c0); k.bg1 = from3(c.bg_c1);
  double acc = 0, sum = 0;
  for (int i = 0; i < c.n_redirect_targets; i++) {
    const RtRedirectTarget& t = c.redirect_targets[i];
    RedirectTarget r;
    r.origin = from3(t.q); r.U = from3(t.s0); r.V = from3(t.s1); r.crs = cross(r.U, r.V); r.p = t.p;
    r.shape = std::static_pointer_cast<PlaneShape>(mkParallelogram(r.origin, r.U, r.V));
    k.targets.push_back(r);
    acc += t.p;             // scanl1 (+) probs
    k.cumulative.push_back(acc);
    sum += t.p;
  }
  k.remProb = 1 - sum;
  return k;
}

struct Renderer {
  const Scene& sc;
  const Camera& cam;
  V3 background(const Ray& r) const {
    if (cam.bg_kind == RT_BG_CONSTANT) return cam.bg0;
    double a = 0.5 * (r.d.y + 1);  // sky / grayFade, test/Main.hs:19-28
    return (1 - a) * cam.bg0 + a * cam.bg1;
  }
  const RedirectTarget* getTarget(double r) const {  // Ray.hs:150-151: first threshold with r < cumulative
    for (size_t i = 0; i < cam.targets.size(); i++) if (!(r >= cam.cumulative[i])) return &cam.targets[i];
    return nullptr;
  }
  double mixturePdf(const HitRecord& hit, V3 dir, double pdf1, Ctx& ctx) const {  // Ray.hs:199-204
    double s = 0;
    for (auto& t : cam.targets) {
      HitRecord h;
      const Material* m = nullptr;
      double pd = 0;
      if (t.shape->hit(0, Ray{hit.p, dir}, 0, INF, ctx, h, m)) pd = h.t * h.t / std::fabs(dot(t.crs, dir));
      s += t.p * pd;
    }
    return cam.remProb * pdf1 + s;
  }
  // Ray.hs:174-224
  V3 rayColor(int depth, double time, const Ray& ray, Ctx& ctx) const {
    if (depth <= 0) return {};
    ctx.cnt.rays++;
    HitRecord hit;
    const Material* mat = nullptr;
    if (!sc.root->hit(time, ray, 0.0001, INF, ctx, hit, mat)) return background(ray);
    V3 emitted;
    MatResult res = evalMaterial(*mat, ray.d, hit, ctx.rng, emitted);
    switch (res.kind) {
      case ABSORB: return emitted;
      case SCATTER: return emitted + res.attenuation * rayColor(depth - 1, time, Ray{hit.p, res.dir}, ctx);
      case HEMISPHERE_F: {
        const RedirectTarget* choice = getTarget(randomDouble(ctx.rng));
        V3 dir;
        if (!choice) {
          V3 uu = randomUnitVector(ctx.rng);
          dir = normalize(hit.n + uu);
        } else {
          double i = randomDouble(ctx.rng), j = randomDouble(ctx.rng);
          V3 lightPt = choice->origin + i * choice->U + j * choice->V;
          dir = normalize(lightPt - hit.p);
        }
        double pdf1 = dot(dir, hit.n) / PI;
        if (pdf1 <= 0) return emitted;
        double pdf = mixturePdf(hit, dir, pdf1, ctx);
        V3 c = rayColor(depth - 1, time, Ray{hit.p, dir}, ctx);
        return emitted + evalMatF(*mat, ray.d, hit, dir) * c * (pdf1 / pdf);
      }
      case SPHERE_F: {
        const RedirectTarget* choice = getTarget(randomDouble(ctx.rng));
        V3 dir;
        if (!choice) dir = randomUnitVector(ctx.rng);
        else {
          double i = randomDouble(ctx.rng), j = randomDouble(ctx.rng);
          V3 lightPt = choice->origin + i * choice->U + j * choice->V;
          dir = normalize(lightPt - hit.p);
        }
        double pdf1 = 0.25 / PI;
        double pdf = mixturePdf(hit, dir, pdf1, ctx);
        V3 c = rayColor(depth - 1, time, Ray{hit.p, dir}, ctx);
        return emitted + evalMatF(*mat, ray.d, hit, dir) * c * (pdf1 / pdf);
      }
    }
    return {};
  }
  // Ray.hs:157-172
  Ray getRay(int i, int j, Ctx& ctx) const {
    double dx, dy;
    randomInUnitDisk(ctx.rng, dx, dy);
    V3 origin = cam.center + dx * cam.defocusU + dy * cam.defocusV;
    double x = randomDouble(ctx.rng), y = randomDouble(ctx.rng);
    V3 target = cam.topLeft + ((double)i + x) * cam.pixelU + ((double)j + y) * cam.pixelV;
    return Ray{origin, normalize(target - origin)};
  }
};

}  // namespace

// ---------------------------------------------------------------------------------------------
// C API (used through ctypes by oracle/harness.py)
// ---------------------------------------------------------------------------------------------
extern "C" {

struct OrcCounters {
  uint64_t samples, rays, box_tests, sphere_tests, planar_tests, medium_queries;
  double seconds;
  int32_t threads, _pad;
};

void* orc_scene_create() { return new Scene(); }
void orc_scene_destroy(void* s) { delete (Scene*)s; }
const char* orc_last_error(void* s) { return ((Scene*)s)->error.c_str(); }

int32_t orc_scene_set_noise_tables(void* sp, const int32_t* px, const int32_t* py, const int32_t* pz, const double* g) {
  Scene& s = *(Scene*)sp;
  for (int i = 0; i < 256; i++) {
    if (px) s.noise.px[i] = px[i];
    if (py) s.noise.py[i] = py[i];
    if (pz) s.noise.pz[i] = pz[i];
    if (g) s.noise.grad[i] = {g[3 * i], g[3 * i + 1], g[3 * i + 2]};
  }
  return 0;
}
int32_t orc_noise_tables(void* sp, int32_t* px, int32_t* py, int32_t* pz, double* g) {
  Scene& s = *(Scene*)sp;
  for (int i = 0; i < 256; i++) {
    px[i] = s.noise.px[i]; py[i] = s.noise.py[i]; pz[i] = s.noise.pz[i];
    g[3 * i] = s.noise.grad[i].x; g[3 * i + 1] = s.noise.grad[i].y; g[3 * i + 2] = s.noise.grad[i].z;
  }
  return 0;
}

int32_t orc_scene_add_texture(void* sp, const RtTextureDesc* d, uint32_t* id) {
  Scene& s = *(Scene*)sp;
  if (d->kind < RT_TEX_CONSTANT || d->kind > RT_TEX_MARBLE) { s.error = "oracle: user-defined / unknown texture"; return RT_ERR_USER_DEFINED; }
  auto t = std::make_unique<Texture>();
  t->kind = d->kind; t->i0 = d->i0; t->i1 = d->i1; t->w = d->width; t->h = d->height; t->freq = d->freq;
  t->c0 = from3(d->c0); t->c1 = from3(d->c1); t->shift = from3(d->shift); t->dir = from3(d->dir);
  t->noise = &s.noise;
  if (d->kind == RT_TEX_IMAGE) {
    size_t n = (size_t)d->width * d->height * 3;
    t->image.resize(n);
    for (size_t i = 0; i < n; i++) t->image[i] = d->image[i];
  }
  *id = (uint32_t)s.textures.size();
  s.textures.push_back(std::move(t));
  return 0;
}
int32_t orc_scene_add_material(void* sp, const RtMaterialDesc* d, uint32_t* id) {
  Scene& s = *(Scene*)sp;
  if (d->kind < RT_MAT_LIGHT || d->kind > RT_MAT_ANISOTROPIC) { s.error = "oracle: user-defined / unknown material"; return RT_ERR_USER_DEFINED; }
  if (d->texture >= s.textures.size()) { s.error = "oracle: bad texture id"; return RT_ERR_INVALID_ARG; }
  auto m = std::make_unique<Material>();
  m->kind = d->kind; m->tex = s.textures[d->texture].get(); m->param = d->param;
  *id = (uint32_t)s.materials.size();
  s.materials.push_back(std::move(m));
  return 0;
}
int32_t orc_scene_add_node(void* sp, const RtNodeDesc* d, uint32_t* id) {
  Scene& s = *(Scene*)sp;
  std::vector<GeomP> ch;
  for (uint32_t i = 0; i < d->n_children; i++) {
    if (d->children[i] >= s.nodes.size()) { s.error = "oracle: bad child id"; return RT_ERR_INVALID_ARG; }
    ch.push_back(s.nodes[d->children[i]]);
  }
  GeomP g;
  switch (d->kind) {
    case RT_NODE_SPHERE: g = std::make_shared<Sphere>(from3(d->v[0]), d->s[0]); break;
    case RT_NODE_PARALLELOGRAM: g = mkParallelogram(from3(d->v[0]), from3(d->v[1]), from3(d->v[2])); break;
    case RT_NODE_TRIANGLE: g = mkTriangle(from3(d->v[0]), from3(d->v[1]), from3(d->v[2]), d->uv[0], d->uv[1], d->uv[2]); break;
    case RT_NODE_CUBOID: g = mkCuboid(from3(d->v[0]), from3(d->v[1])); break;
    case RT_NODE_MESH: {  // triangleMesh, Geometry.hs:288-294
      if (d->n_faces == 0) { s.error = "bvhTree: empty list"; return RT_ERR_EMPTY_BVH; }
      std::vector<GeomP> tris;
      const double def0[2] = {0, 0}, def1[2] = {1, 0}, def2[2] = {0, 1};
      for (uint32_t f = 0; f < d->n_faces; f++) {
        const int32_t* fi = d->faces + 6 * f;
        const double* uv0 = fi[1] >= 0 ? d->uvs + 2 * fi[1] : def0;
        const double* uv1 = fi[3] >= 0 ? d->uvs + 2 * fi[3] : def1;
        const double* uv2 = fi[5] >= 0 ? d->uvs + 2 * fi[5] : def2;
        tris.push_back(mkTriangle(from3(d->verts + 3 * fi[0]), from3(d->verts + 3 * fi[2]), from3(d->verts + 3 * fi[4]), uv0, uv1, uv2));
      }
      auto lst = tris;
      g = std::make_shared<Built>(std::move(lst), bvhTree(std::move(tris)));
      break;
    }
    case RT_NODE_GROUP:
      if (ch.empty()) { s.error = "group: empty list"; return RT_ERR_INVALID_ARG; }
      g = std::make_shared<Group>(ch); break;
    case RT_NODE_BVH_NODE:
      if (ch.size() != 2) { s.error = "bvhNode needs 2 children"; return RT_ERR_INVALID_ARG; }
      g = std::make_shared<BvhNode>(ch[0], ch[1]); break;
    case RT_NODE_BVH_TREE: {
      if (ch.empty()) { s.error = "bvhTree: empty list"; return RT_ERR_EMPTY_BVH; }
      auto lst = ch;
      g = std::make_shared<Built>(std::move(lst), bvhTree(std::move(ch)));
      break;
    }
    case RT_NODE_TRANSFORM: g = std::make_shared<Transform>(d->m, ch.at(0)); break;
    case RT_NODE_MOVING: g = std::make_shared<Moving>(from3(d->v[0]), from3(d->v[1]), ch.at(0)); break;
    case RT_NODE_MEDIUM: g = std::make_shared<ConstantMedium>(d->s[0], ch.at(0)); break;
    case RT_NODE_WITH_MATERIAL:
      if (d->material >= s.materials.size()) { s.error = "oracle: bad material id"; return RT_ERR_INVALID_ARG; }
      g = std::make_shared<WithMaterial>(s.materials[d->material].get(), ch.at(0)); break;
    case RT_NODE_PURE: g = ch.at(0); break;  // pureGeometry, Geometry.hs:50-51
    default: s.error = "oracle: user-defined / unknown node"; return RT_ERR_USER_DEFINED;
  }
  *id = (uint32_t)s.nodes.size();
  s.nodes.push_back(std::move(g));
  return 0;
}
int32_t orc_scene_commit(void* sp, uint32_t root) {
  Scene& s = *(Scene*)sp;
  if (root >= s.nodes.size()) { s.error = "oracle: bad root"; return RT_ERR_INVALID_ARG; }
  s.root = s.nodes[root];
  uint32_t next = 0;
  s.root->number(next);
  s.n_prims = next;
  return 0;
}
uint32_t orc_scene_num_prims(void* sp) { return ((Scene*)sp)->n_prims; }
// canonical ids of the surface leaves in the reference's evaluation (DFS of the built trees) order
int32_t orc_scene_leaf_order(void* sp, uint32_t* out, uint32_t cap, uint32_t* n) {
  Scene& s = *(Scene*)sp;
  std::vector<uint32_t> o;
  s.root->leafOrder(o);
  *n = (uint32_t)o.size();
  for (uint32_t i = 0; i < o.size() && i < cap; i++) out[i] = o[i];
  return 0;
}
int32_t orc_image_height(const RtCamera* c) { return imageHeightOf(*c); }

// raytrace (Ray.hs:121-238).  rgb_mean: H*W*3 doubles = Σ/spp; rgb_sumsq optional: Σ x² per channel.
int32_t orc_render(void* sp, const RtCamera* rc, uint64_t seed, int32_t spp_override, int32_t n_threads,
                   double* rgb_mean, double* rgb_sumsq, OrcCounters* out) {
  Scene& s = *(Scene*)sp;
  if (!s.root) { s.error = "oracle: not committed"; return RT_ERR_NOT_COMMITTED; }
  Camera cam = makeCamera(*rc);
  if (spp_override > 0) cam.spp = spp_override;
  const int W = cam.W, H = cam.H;
  // seeds = randomArrayS seed (Sz (H :. W)) splitGen  (Ray.hs:235-236): sequential, row-major
  std::vector<StdGen> seeds((size_t)W * H);
  StdGen g = mkStdGen(seed);
  for (size_t k = 0; k < seeds.size(); k++) seeds[k] = splitGen(g);
  if (n_threads <= 0) n_threads = (int)std::max(1u, std::thread::hardware_concurrency());
  Renderer R{s, cam};
  std::atomic<int> nextRow{0};
  std::vector<Counters> cnts(n_threads);
  auto t0 = std::chrono::steady_clock::now();
  auto work = [&](int tid) {
    Counters local;
    for (;;) {
      int j = nextRow.fetch_add(1);
      if (j >= H) break;
      for (int i = 0; i < W; i++) {
        Ctx ctx;
        ctx.rng = seeds[(size_t)j * W + i];
        V3 sum, sq;
        for (int k = 0; k < cam.spp; k++) {  // pixelColor, Ray.hs:226-232
          double time = randomDouble(ctx.rng);
          Ray ray = R.getRay(i, j, ctx);
          V3 c = R.rayColor(cam.maxDepth, time, ray, ctx);
          sum = sum + c;
          sq = sq + c * c;
          ctx.cnt.samples++;
        }
        size_t o = ((size_t)j * W + i) * 3;
        rgb_mean[o] = sum.x / cam.spp; rgb_mean[o + 1] = sum.y / cam.spp; rgb_mean[o + 2] = sum.z / cam.spp;
        if (rgb_sumsq) { rgb_sumsq[o] = sq.x; rgb_sumsq[o + 1] = sq.y; rgb_sumsq[o + 2] = sq.z; }
        local.samples += ctx.cnt.samples; local.rays += ctx.cnt.rays; local.box_tests += ctx.cnt.box_tests;
        local.sphere_tests += ctx.cnt.sphere_tests; local.planar_tests += ctx.cnt.planar_tests;
        local.medium_queries += ctx.cnt.medium_queries;
      }
    }
    cnts[tid] = local;
  };
  std::vector<std::thread> th;
  for (int t = 0; t < n_threads; t++) th.emplace_back(work, t);
  for (auto& t : th) t.join();
  double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  if (out) {
    std::memset(out, 0, sizeof *out);
    for (auto& c : cnts) {
      out->samples += c.samples; out->rays += c.rays; out->box_tests += c.box_tests;
      out->sphere_tests += c.sphere_tests; out->planar_tests += c.planar_tests; out->medium_queries += c.medium_queries;
    }
    out->seconds = secs;
    out->threads = n_threads;
  }
  return 0;
}

// Deterministic probe: pixel-centre ray from cs_center (no lens sample), media skipped.
// (dx, dy) = sub-pixel position of the probe ray (0.5, 0.5 = pixel centre); offsets near the centre are used by
// the tests to find pixels whose answer is ill-conditioned (ray through an edge / silhouette).
int32_t orc_probe_primary(void* sp, const RtCamera* rc, double time, double dx, double dy, uint32_t* ids, double* ts) {
  Scene& s = *(Scene*)sp;
  if (!s.root) { s.error = "oracle: not committed"; return RT_ERR_NOT_COMMITTED; }
  Camera cam = makeCamera(*rc);
  for (int j = 0; j < cam.H; j++)
    for (int i = 0; i < cam.W; i++) {
      V3 target = cam.topLeft + ((double)i + dx) * cam.pixelU + ((double)j + dy) * cam.pixelV;
      Ray ray{cam.center, normalize(target - cam.center)};
      Ctx ctx;
      ctx.rng = mkStdGen(0);
      ctx.skip_media = true;
      HitRecord h;
      const Material* m = nullptr;
      size_t o = (size_t)j * cam.W + i;
      if (s.root->hit(time, ray, 0.0001, INF, ctx, h, m)) { ids[o] = h.prim; ts[o] = h.t; }
      else { ids[o] = 0xFFFFFFFFu; ts[o] = INF; }
    }
  return 0;
}

// Deterministic shading inputs of the pixel-centre primary hit: world normal facing the ray (3), uv (2), front flag,
// the material's texture colour at the hit (3; zero for texture-less materials) and the material kind.  10 doubles per
// pixel (zeros on a miss).  Pins hit records, sphereUV, the cuboid uv orientation and every texture (checker / image /
// Perlin) pointwise, without Monte Carlo noise.
int32_t orc_probe_surface(void* sp, const RtCamera* rc, double time, double* out) {
  Scene& s = *(Scene*)sp;
  if (!s.root) { s.error = "oracle: not committed"; return RT_ERR_NOT_COMMITTED; }
  Camera cam = makeCamera(*rc);
  for (int j = 0; j < cam.H; j++)
    for (int i = 0; i < cam.W; i++) {
      V3 target = cam.topLeft + ((double)i + 0.5) * cam.pixelU + ((double)j + 0.5) * cam.pixelV;
      Ray ray{cam.center, normalize(target - cam.center)};
      Ctx ctx;
      ctx.rng = mkStdGen(0);
      ctx.skip_media = true;
      HitRecord h;
      const Material* m = nullptr;
      double* o = out + 10 * ((size_t)j * cam.W + i);
      for (int k = 0; k < 10; k++) o[k] = 0;
      if (!s.root->hit(time, ray, 0.0001, INF, ctx, h, m) || !m) continue;
      o[0] = h.n.x; o[1] = h.n.y; o[2] = h.n.z;
      hitUV(h, o[3], o[4]);
      o[5] = h.front ? 1 : 0;
      if (m->tex) { V3 c = m->tex->eval(h.p, h); o[6] = c.x; o[7] = c.y; o[8] = c.z; }
      o[9] = m->kind;
    }
  return 0;
}

// unit-test hooks
double orc_perlin(void* sp, double x, double y, double z) { return perlinNoise(((Scene*)sp)->noise, {x, y, z}); }
double orc_fractal(void* sp, int32_t depth, double x, double y, double z) { return fractalNoise(((Scene*)sp)->noise, depth, {x, y, z}); }
int32_t orc_texture_eval(void* sp, uint32_t tex, const double* p, double u, double v, double* rgb) {
  Scene& s = *(Scene*)sp;
  if (tex >= s.textures.size()) return RT_ERR_INVALID_ARG;
  HitRecord h;
  h.uv_src = {u, v, 0};
  V3 c = s.textures[tex]->eval(from3(p), h);
  rgb[0] = c.x; rgb[1] = c.y; rgb[2] = c.z;
  return 0;
}
// one closest-hit query through the full tree with a fresh generator (tests of single primitives)
int32_t orc_hit(void* sp, const double* o, const double* d, double time, double tmin, double tmax,
                double* t, double* p, double* n, int32_t* front, double* uv, uint32_t* prim) {
  Scene& s = *(Scene*)sp;
  Ctx ctx;
  ctx.rng = mkStdGen(0);
  ctx.skip_media = true;
  HitRecord h;
  const Material* m = nullptr;
  if (!s.root->hit(time, Ray{from3(o), from3(d)}, tmin, tmax, ctx, h, m)) return 0;
  *t = h.t; p[0] = h.p.x; p[1] = h.p.y; p[2] = h.p.z; n[0] = h.n.x; n[1] = h.n.y; n[2] = h.n.z;
  *front = h.front; hitUV(h, uv[0], uv[1]); *prim = h.prim;
  return 1;
}
// SplitMix probes (documenting the restated stream)
uint64_t orc_splitmix_next(uint64_t seed_int, int32_t n) {
  StdGen g = mkStdGen(seed_int);
  uint64_t w = 0;
  for (int i = 0; i < n; i++) w = nextWord64(g);
  return w;
}

}  // extern "C"
