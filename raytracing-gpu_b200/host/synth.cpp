// Deterministic synthetic scenes built through the reference's scene API
// (empty_scene / add_object_to_scene / get_vertex / get_normal, gpu/headers/scene.h:40-61)
// and a .svati writer (format: SURVEY Appendix B).  The 10 M-triangle benchmark scene of
// BASELINE.json configs[4] is rtb_synth_terrain(grid = 2237): a tessellated height field
// (4 octaves of value noise on a 256x256 lattice seeded with splitmix64 + two sinusoids)
// over x in [-500,500], z in [0,1000], smooth vertex normals from central differences,
// split into 4x4 patches (16 objects), one ambient + one directional + one point light.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <charconv>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include "rt_host.h"

extern "C" int rtb_scene_reserve(struct scene *s, size_t triangles);   // scene.cpp

namespace {

uint64_t splitmix64(uint64_t &s) {
  uint64_t z = (s += 0x9e3779b97f4a7c15ull);
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  return z ^ (z >> 31);
}

struct Terrain {
  std::vector<float> lattice;   // 256 x 256 values in [-1, 1]
  explicit Terrain(uint64_t seed) : lattice(256 * 256) {
    uint64_t s = seed;
    for (float &v : lattice) v = (float)((double)(splitmix64(s) >> 11) * (1.0 / 9007199254740992.0) * 2.0 - 1.0);
  }
  double noise(double x, double y) const {
    const double fx = std::floor(x), fy = std::floor(y);
    const int ix = (int)fx & 255, iy = (int)fy & 255, jx = (ix + 1) & 255, jy = (iy + 1) & 255;
    double tx = x - fx, ty = y - fy;
    tx = tx * tx * (3 - 2 * tx);
    ty = ty * ty * (3 - 2 * ty);
    const double a = lattice[iy * 256 + ix], b = lattice[iy * 256 + jx], c = lattice[jy * 256 + ix], d = lattice[jy * 256 + jx];
    return (a + (b - a) * tx) + ((c + (d - c) * tx) - (a + (b - a) * tx)) * ty;
  }
  // height above the point (x, z), world units
  double height(double x, double z) const {
    const double u = (x + 500.0) / 1000.0, v = z / 1000.0;
    double h = 0, amp = 40, f = 8;
    for (int o = 0; o < 4; ++o) {
      h += amp * noise(u * f + 17.0 * o, v * f + 31.0 * o);
      amp *= 0.5;
      f *= 2;
    }
    return h + 15.0 * std::sin(x * 0.013) + 10.0 * std::cos(z * 0.021);
  }
};

template <class F>
void parallel_for(int n, F fn) {
  int nt = (int)std::thread::hardware_concurrency();
  if (nt < 1) nt = 1;
  if (nt > 32) nt = 32;
  if (nt > n) nt = n > 0 ? n : 1;
  std::vector<std::thread> pool;
  for (int t = 0; t < nt; ++t)
    pool.emplace_back([=]() {
      for (int i = t; i < n; i += nt) fn(i);
    });
  for (auto &th : pool) th.join();
}

}  // namespace

extern "C" int rtb_synth_terrain(struct scene *out, int grid, uint64_t seed, int width, int height) {
  if (!out || grid < 4 || grid > 11585) return -1;   // 2*grid^2 < 2^28
  *out = empty_scene();
  const int G = grid, V = G + 1;
  const Terrain T(seed);
  std::vector<float> hgt((size_t)V * V);
  std::vector<vector3> nrm((size_t)V * V);
  const double step = 1000.0 / G;
  parallel_for(V, [&](int iz) {
    for (int ix = 0; ix < V; ++ix) {
      const double x = -500.0 + ix * step, z = iz * step;
      hgt[(size_t)iz * V + ix] = (float)T.height(x, z);
      const double e = 0.5 * step;
      const double dx = (T.height(x + e, z) - T.height(x - e, z)) / (2 * e);
      const double dz = (T.height(x, z + e) - T.height(x, z - e)) / (2 * e);
      const double inv = 1.0 / std::sqrt(dx * dx + 1.0 + dz * dz);
      vector3 n;
      n.x = (float)(-dx * inv);
      n.y = (float)(inv);
      n.z = (float)(-dz * inv);
      nrm[(size_t)iz * V + ix] = n;
    }
  });
  // 4 x 4 patches -> 16 objects, appended in row-major patch order
  struct Patch { int x0, x1, z0, z1; struct object *obj; size_t first; };
  std::vector<Patch> patches;
  for (int pz = 0; pz < 4; ++pz)
    for (int px = 0; px < 4; ++px) {
      Patch p;
      p.x0 = G * px / 4; p.x1 = G * (px + 1) / 4;
      p.z0 = G * pz / 4; p.z1 = G * (pz + 1) / 4;
      p.obj = nullptr;
      p.first = 0;
      patches.push_back(p);
    }
  rtb_scene_reserve(out, 2ull * G * G);   // one allocation of the final size
  for (size_t k = 0; k < patches.size(); ++k) {
    Patch &p = patches[k];
    const uint32_t tris = (uint32_t)(2ull * (p.x1 - p.x0) * (p.z1 - p.z0));
    struct object *o = add_object_to_scene(out, tris);
    const int px = (int)k % 4, pz = (int)k / 4;
    o->ka.x = 0.5f; o->ka.y = 0.5f; o->ka.z = 0.5f;
    o->kd.x = 0.35f + 0.15f * px; o->kd.y = 0.8f - 0.1f * pz; o->kd.z = 0.3f + 0.05f * ((px + pz) % 4);
    o->ks.x = o->ks.y = o->ks.z = 0.3f;
    o->ns = 32.0f;
    o->nr = 0.0f;
  }
  // objects[] may have moved while it grew: take the pointers only now
  for (size_t k = 0; k < patches.size(); ++k) patches[k].obj = &out->objects[k];
  for (size_t k = 0; k < patches.size(); ++k) {
    const Patch p = patches[k];
    const struct triangles_layout tl = p.obj->triangles;
    const int w = p.x1 - p.x0;
    parallel_for(p.z1 - p.z0, [&, p, tl, w](int rz) {
      const int iz = p.z0 + rz;
      for (int ix = p.x0; ix < p.x1; ++ix) {
        const uint32_t t = (uint32_t)(2 * ((size_t)rz * w + (ix - p.x0)));
        auto put = [&](uint32_t tri, int c, int vx, int vz) {
          vector3 pnt;
          pnt.x = (float)(-500.0 + vx * step);
          pnt.y = hgt[(size_t)vz * V + vx];
          pnt.z = (float)(vz * step);
          get_vertex(tl, tri)[c] = pnt;
          get_normal(tl, tri)[c] = nrm[(size_t)vz * V + vx];
        };
        put(t, 0, ix, iz); put(t, 1, ix, iz + 1); put(t, 2, ix + 1, iz);
        put(t + 1, 0, ix + 1, iz); put(t + 1, 1, ix, iz + 1); put(t + 1, 2, ix + 1, iz + 1);
      }
    });
  }
  out->lights = static_cast<struct light *>(malloc(3 * sizeof(struct light)));
  out->light_count = 3;
  memset(out->lights, 0, 3 * sizeof(struct light));
  out->lights[0].type = AMBIENT; out->lights[0].r = out->lights[0].g = out->lights[0].b = 0.2f;
  out->lights[1].type = DIRECTIONAL; out->lights[1].r = out->lights[1].g = out->lights[1].b = 1.0f;
  out->lights[1].v.x = 0.5f; out->lights[1].v.y = -1.0f; out->lights[1].v.z = 0.5f;
  out->lights[2].type = POINT; out->lights[2].r = out->lights[2].g = out->lights[2].b = 1.0f;
  out->lights[2].v.x = 0.0f; out->lights[2].v.y = 300.0f; out->lights[2].v.z = 500.0f;
  out->camera.width = width; out->camera.height = height;
  out->camera.position.x = 0; out->camera.position.y = 220; out->camera.position.z = -150;
  out->camera.u.x = 1; out->camera.u.y = 0; out->camera.u.z = 0;
  out->camera.v.x = 0; out->camera.v.y = -0.8f; out->camera.v.z = -0.6f;
  out->camera.fov = 70;
  return 0;
}

// Scene -> .svati text.  %.9g round-trips binary32 exactly; vertices are written in the
// order that parse_object's LIFO reversal turns back into the memory order (App. B).
namespace {

// "<tag>x y z\n" for arr[n-1] .. arr[0] (the loader's LIFO reversal undone), shortest
// round-trip spelling of every float (std::to_chars), formatted on all threads in batches.
bool write_lines(FILE *f, const char *tag, const vector3 *arr, size_t n) {
  const size_t chunk = 1 << 16;
  const size_t n_chunks = (n + chunk - 1) / chunk;
  int nt = (int)std::thread::hardware_concurrency();
  if (nt < 1) nt = 1;
  if (nt > 64) nt = 64;
  std::vector<std::string> text((size_t)nt);
  const size_t tag_len = strlen(tag);
  for (size_t c0 = 0; c0 < n_chunks; c0 += (size_t)nt) {
    const size_t batch = std::min<size_t>((size_t)nt, n_chunks - c0);
    std::vector<std::thread> pool;
    for (size_t b = 0; b < batch; ++b)
      pool.emplace_back([&, b] {
        std::string &out = text[b];
        out.clear();
        const size_t lo = (c0 + b) * chunk, hi = std::min(n, lo + chunk);   // positions in output order
        out.reserve((hi - lo) * 40);
        char buf[64];
        for (size_t pos = lo; pos < hi; ++pos) {
          const vector3 &v = arr[n - 1 - pos];
          out.append(tag, tag_len);
          const float xyz[3] = {v.x, v.y, v.z};
          for (int a = 0; a < 3; ++a) {
            const std::to_chars_result r = std::to_chars(buf, buf + sizeof buf, xyz[a]);
            out.append(buf, (size_t)(r.ptr - buf));
            out.push_back(a == 2 ? '\n' : ' ');
          }
        }
      });
    for (std::thread &t : pool) t.join();
    for (size_t b = 0; b < batch; ++b)
      if (fwrite(text[b].data(), 1, text[b].size(), f) != text[b].size()) return false;
  }
  return true;
}

}  // namespace

extern "C" int rtb_write_svati(const struct scene *s, const char *path, int width, int height) {
  FILE *f = fopen(path, "w");
  if (!f) return -1;
  const struct camera &c = s->camera;
  fprintf(f, "camera %d %d %.9g %.9g %.9g %.9g %.9g %.9g %.9g %.9g %.9g %.9g\n\n", width, height, c.position.x,
          c.position.y, c.position.z, c.u.x, c.u.y, c.u.z, c.v.x, c.v.y, c.v.z, c.fov);
  for (size_t i = 0; i < s->light_count; ++i) {
    const struct light &l = s->lights[i];
    if (l.type == AMBIENT) fprintf(f, "a_light %.9g %.9g %.9g\n", l.r, l.g, l.b);
    else fprintf(f, "%s %.9g %.9g %.9g %.9g %.9g %.9g\n", l.type == DIRECTIONAL ? "d_light" : "p_light", l.r, l.g, l.b,
                 l.v.x, l.v.y, l.v.z);
  }
  for (size_t o = 0; o < s->object_count; ++o) {
    const struct object &ob = s->objects[o];
    const size_t nv = 3 * (size_t)ob.triangle_count;
    fprintf(f, "\nobject %zu\nNs %.9g\nKa %.9g %.9g %.9g\nKd %.9g %.9g %.9g\nKs %.9g %.9g %.9g\nNi %.9g\nNr %.9g\nd %.9g\n",
            nv, ob.ns, ob.ka.x, ob.ka.y, ob.ka.z, ob.kd.x, ob.kd.y, ob.kd.z, ob.ks.x, ob.ks.y, ob.ks.z, ob.ni, ob.nr, ob.d);
    if (!write_lines(f, "v ", ob.triangles.vertex, nv) || !write_lines(f, "vn ", ob.triangles.normal, nv)) {
      fclose(f);
      return -1;
    }
  }
  return fclose(f) == 0 ? 0 : -1;
}
