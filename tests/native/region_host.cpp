// Host build of csrc/region.cuh for the CPU tests (tests/test_region_cpu.py): the very code the
// CUDA kernels run for the cull, compiled by g++.  Test infrastructure, not product.
#include <stdint.h>
#include <string.h>
#include "region.cuh"

using namespace rtb::region;

extern "C" {

// cam: pos3 D03 u3 v3 Dmax Mo (14 doubles); tri: n x (v0 3, e1 3, e2 3) floats
// out per triangle: kind, px0, py0, px1, py1 (int32 x 5); rec: 12 floats
void rg_primary(const double *cam, int W, int H, const float *tri, int n, int32_t *out, float *rec) {
  CamFrame cf;
  memcpy(cf.pos, cam, 24); memcpy(cf.D0, cam + 3, 24); memcpy(cf.u, cam + 6, 24); memcpy(cf.v, cam + 9, 24);
  cf.Dmax = cam[12]; cf.Mo = cam[13];
  for (int i = 0; i < n; ++i) {
    PrimaryCull c;
    primary_cull(cf, W, H, tri + 9 * i, tri + 9 * i + 3, tri + 9 * i + 6, c);
    out[5 * i] = c.kind; out[5 * i + 1] = c.x0; out[5 * i + 2] = c.y0; out[5 * i + 3] = c.x1; out[5 * i + 4] = c.y1;
    memcpy(rec + 12 * i, c.rec.v, 48);
  }
}

// the edge functions themselves: lines[12 i ..] = cx[3] cy[3] c0[3] E[3]; ok[i] = 0: never accepted
void rg_primary_lines(const double *cam, const float *tri, int n, double *lines, int32_t *ok) {
  CamFrame cf;
  memcpy(cf.pos, cam, 24); memcpy(cf.D0, cam + 3, 24); memcpy(cf.u, cam + 6, 24); memcpy(cf.v, cam + 9, 24);
  cf.Dmax = cam[12]; cf.Mo = cam[13];
  for (int i = 0; i < n; ++i) {
    Lines L;
    memset(&L, 0, sizeof L);
    ok[i] = primary_lines(cf, tri + 9 * i, tri + 9 * i + 3, tri + 9 * i + 6, L) ? 1 : 0;
    memcpy(lines + 12 * i, &L, sizeof L);
  }
}

// frame: d3 ax3 ay3 g11 g12 g22 smax Pmax (14 doubles); dom: x0 x1 y0 y1
// out: kind + bounds (5 doubles); rec 12 floats (origin = centre of the bounds, rounded to float)
void rg_dir(const double *fr, const double *dom, const float *tri, int n, double *out, float *rec) {
  DirFrame f;
  memcpy(f.d, fr, 24); memcpy(f.ax, fr + 3, 24); memcpy(f.ay, fr + 6, 24);
  f.g11 = fr[9]; f.g12 = fr[10]; f.g22 = fr[11]; f.smax = fr[12]; f.Pmax = fr[13];
  for (int i = 0; i < n; ++i) {
    Lines L;
    Region R;
    memset(&R, 0, sizeof R);
    if (dir_lines(f, tri + 9 * i, tri + 9 * i + 3, tri + 9 * i + 6, L)) build(L, dom[0], dom[1], dom[2], dom[3], R);
    out[5 * i] = R.kind; out[5 * i + 1] = R.x0; out[5 * i + 2] = R.x1; out[5 * i + 3] = R.y0; out[5 * i + 4] = R.y1;
    Record r;
    memset(&r, 0, sizeof r);
    if (R.kind) {
      const double X0 = (double)(float)(0.5 * (R.x0 + R.x1)), Y0 = (double)(float)(0.5 * (R.y0 + R.y1));
      make_record(R, X0, Y0, (R.x1 - R.x0) + (R.y1 - R.y0), r);
    }
    memcpy(rec + 12 * i, r.v, 48);
  }
}

// frame: org3 hx3 hy3 hz3 smax dmax ratio (15 doubles)
void rg_point(const double *fr, const double *dom, const float *tri, int n, double *out, float *rec) {
  PointFrame f;
  memcpy(f.org, fr, 24); memcpy(f.hx, fr + 3, 24); memcpy(f.hy, fr + 6, 24); memcpy(f.hz, fr + 9, 24);
  f.smax = fr[12]; f.dmax = fr[13]; f.ratio = fr[14];
  for (int i = 0; i < n; ++i) {
    Lines L;
    Region R;
    memset(&R, 0, sizeof R);
    if (point_lines(f, tri + 9 * i, tri + 9 * i + 3, tri + 9 * i + 6, L)) build(L, dom[0], dom[1], dom[2], dom[3], R);
    out[5 * i] = R.kind; out[5 * i + 1] = R.x0; out[5 * i + 2] = R.x1; out[5 * i + 3] = R.y0; out[5 * i + 4] = R.y1;
    Record r;
    memset(&r, 0, sizeof r);
    if (R.kind) {
      const double X0 = (double)(float)(0.5 * (R.x0 + R.x1)), Y0 = (double)(float)(0.5 * (R.y0 + R.y1));
      make_record(R, X0, Y0, (R.x1 - R.x0) + (R.y1 - R.y0), r);
    }
    memcpy(rec + 12 * i, r.v, 48);
  }
}

// record_span for n records / rows: out a, b
void rg_span(const float *rec, const float *dx0, const int32_t *w, const float *dy, int n, int32_t *ab) {
  for (int i = 0; i < n; ++i) {
    const float *r = rec + 12 * i;
    int a, b;
    record_span(r[0], r[1], r[2], r[3], r[4], r[5], r[6], r[7], r[8], r[11], dx0[i], w[i], dy[i], a, b);
    ab[2 * i] = a; ab[2 * i + 1] = b;
  }
}

// record_band for n records / bands: out ok, lo, hi
void rg_band(const float *rec, const float *dy0, const float *dy1, const float *xext, int n, int32_t *ok, float *lohi) {
  for (int i = 0; i < n; ++i) {
    const float *r = rec + 12 * i;
    float lo, hi;
    ok[i] = record_band(r[0], r[1], r[2], r[3], r[4], r[5], r[6], r[7], r[8], r[11], dy0[i], dy1[i], xext[i], lo, hi) ? 1 : 0;
    lohi[2 * i] = lo; lohi[2 * i + 1] = hi;
  }
}

// row_range / box_touches of the primary region of one triangle (tile emission logic)
int rg_primary_rows(const double *cam, int W, int H, const float *tri, int tile, int32_t *rows, int max_rows) {
  CamFrame cf;
  memcpy(cf.pos, cam, 24); memcpy(cf.D0, cam + 3, 24); memcpy(cf.u, cam + 6, 24); memcpy(cf.v, cam + 9, 24);
  cf.Dmax = cam[12]; cf.Mo = cam[13];
  PrimaryCull c;
  primary_cull(cf, W, H, tri, tri + 3, tri + 6, c);
  if (!c.kind) return 0;
  int n = 0;
  for (int ty = c.y0 / tile; ty <= c.y1 / tile && n < max_rows; ++ty) {
    int ta, tb;
    if (!primary_row_tiles(c, W, H, tile, ty, ta, tb)) continue;
    rows[3 * n] = ty; rows[3 * n + 1] = ta; rows[3 * n + 2] = tb;
    ++n;
  }
  return n;
}

}
