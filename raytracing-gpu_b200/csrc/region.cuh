// Where can the ROUNDED Moeller-Trumbore test of the reference accept a ray?  (raster.cuh and
// lightgrid.cuh cull with the answer; plain C++ so that tests/ can compile it for the host.)
//
// gpu/hit.cu:20-37 accepts a ray when the COMPUTED u = f (s . h), v = f (d . q) satisfy u >= 0,
// v >= 0, u + v <= 1 with f = 1 / a, a = e1 . h, h = d x e2, s = o - v0, q = s x e1 (binary32, no
// FMA).  Write U = s . h, V = d . q, A = a for the exact values on the same inputs and W = A - U - V;
// then u = U / A, v = V / A and "inside" is "U, V, W all have the sign of A".  For a family of rays
// through one point (primary rays: the camera position; shadow rays of a point light: the light) or
// with one direction (directional light), U, V and W are LINEAR functions of the 2-D coordinates
// (x, y) that name the ray in that family -- the edge functions of homogeneous rasterisation:
//     F_k(x, y) = c0_k + cx_k x + cy_k y,   k = u, v, w.
// Rounding moves the computed numerators by at most E_k (callers derive E_k from the operation
// sequence, see raster.cuh / lightgrid.cuh), so every accepted ray lies in
//     S+ = { F_k >= -E_k for all k }   (computed a > 0)   or   S- = { F_k <= E_k for all k }   (a < 0),
// two convex regions.  Nothing here divides by a, projects a vertex or assumes the triangle is in
// front of anything: triangles seen edge-on, triangles that cross the plane through the centre of
// projection and degenerate triangles just give long, thin or half-open regions, clipped to the
// domain (the screen / the light's grid).  Typical triangle: S- is empty and S+ is the projected
// triangle grown by a fraction of a sample.
#pragma once
#include <math.h>

#ifdef __CUDACC__
#define RTB_RG __host__ __device__ __forceinline__
#else
#define RTB_RG inline
#endif

namespace rtb {
namespace region {

struct Lines {              // three edge functions and their allowances (E >= 0), inputs of build()
  double cx[3], cy[3], c0[3], E[3];
};

struct Region {
  int kind;                 // 0: no ray of the domain can be accepted, 1: one convex region, 2: S+ and S-
  double x0, x1, y0, y1;    // bounds of the region(s) inside the domain
  // unit outward normals of the (first) region: inside <=> n_k . (x, y) + c_k <= 0 for k = 0..2.
  // kind 2: the second region is  n_k . (x, y) + c_k >= -m  for k = 0..2.
  double nx[3], ny[3], c[3], m;
  double work;              // samples a scan of the region(s) looks at: polygon area + a rim along the outline
};

struct Poly { double x[8], y[8]; int n; };

// keep the part of convex polygon p where a x + b y + c >= 0 (Sutherland-Hodgman; <= 1 vertex more)
RTB_RG void clip(Poly &p, double a, double b, double c) {
  Poly q;
  q.n = 0;
  for (int i = 0; i < p.n; ++i) {
    const int j = i + 1 == p.n ? 0 : i + 1;
    const double fi = a * p.x[i] + b * p.y[i] + c, fj = a * p.x[j] + b * p.y[j] + c;
    if (fi >= 0.0) { q.x[q.n] = p.x[i]; q.y[q.n] = p.y[i]; ++q.n; }
    if ((fi >= 0.0) != (fj >= 0.0)) {
      const double t = fi / (fi - fj);
      q.x[q.n] = p.x[i] + t * (p.x[j] - p.x[i]);
      q.y[q.n] = p.y[i] + t * (p.y[j] - p.y[i]);
      ++q.n;
    }
  }
  p = q;
}

// area of the convex polygon plus one unit along its outline (bounded by the box's perimeter)
RTB_RG double poly_work(const Poly &p) {
  if (p.n < 1) return 0.0;
  double a2 = 0.0, x0 = p.x[0], x1 = p.x[0], y0 = p.y[0], y1 = p.y[0];
  for (int i = 0; i < p.n; ++i) {
    const int j = i + 1 == p.n ? 0 : i + 1;
    a2 += p.x[i] * p.y[j] - p.x[j] * p.y[i];
    x0 = fmin(x0, p.x[i]); x1 = fmax(x1, p.x[i]); y0 = fmin(y0, p.y[i]); y1 = fmax(y1, p.y[i]);
  }
  return 0.5 * fabs(a2) + 2.0 * ((x1 - x0) + (y1 - y0)) + 1.0;
}

RTB_RG void grow(const Poly &p, double &x0, double &x1, double &y0, double &y1) {
  for (int i = 0; i < p.n; ++i) {
    x0 = fmin(x0, p.x[i]); x1 = fmax(x1, p.x[i]);
    y0 = fmin(y0, p.y[i]); y1 = fmax(y1, p.y[i]);
  }
}

// sign = +1: S+ (F_k >= -E_k), -1: S- (F_k <= E_k), clipped to the domain
RTB_RG void side_polygon(const Lines &L, double sign, double dx0, double dx1, double dy0, double dy1, Poly &p) {
  p.n = 4;
  p.x[0] = dx0; p.y[0] = dy0; p.x[1] = dx1; p.y[1] = dy0; p.x[2] = dx1; p.y[2] = dy1; p.x[3] = dx0; p.y[3] = dy1;
  for (int k = 0; k < 3 && p.n; ++k) clip(p, sign * L.cx[k], sign * L.cy[k], sign * L.c0[k] + L.E[k]);
}

// Region of the domain [dx0, dx1] x [dy0, dy1] where some ray can be accepted.
RTB_RG void build(const Lines &L, double dx0, double dx1, double dy0, double dy1, Region &R) {
  const double BIG = 1.0e300;
  R.kind = 0;
  R.x0 = BIG; R.x1 = -BIG; R.y0 = BIG; R.y1 = -BIG;
  R.m = 0.0;
  R.work = 0.0;
  double len[3];
  for (int k = 0; k < 3; ++k) len[k] = sqrt(L.cx[k] * L.cx[k] + L.cy[k] * L.cy[k]);
  bool plus = false, minus = false, done = false;
  // Fast path: the three gradients positively span the plane (a bounded triangle).  Then S+ and S-
  // are both triangles (or empty) whose corners are the pairwise intersections of the offset lines;
  // when they lie inside the domain nothing has to be clipped.
  {
    const double x01 = L.cx[0] * L.cy[1] - L.cy[0] * L.cx[1], x12 = L.cx[1] * L.cy[2] - L.cy[1] * L.cx[2],
                 x20 = L.cx[2] * L.cy[0] - L.cy[2] * L.cx[0];
    // lambda_k = cross of the other two: sum lambda_k grad F_k = 0, so sum lambda_k F_k is the constant K
    const double l0 = x12, l1 = x20, l2 = x01;
    const double tiny = 1e-9;
    const bool pos = l0 > tiny * len[1] * len[2] && l1 > tiny * len[2] * len[0] && l2 > tiny * len[0] * len[1];
    const bool neg = l0 < -tiny * len[1] * len[2] && l1 < -tiny * len[2] * len[0] && l2 < -tiny * len[0] * len[1];
    if (pos || neg) {
      const double sg = pos ? 1.0 : -1.0;
      const double K = sg * (l0 * L.c0[0] + l1 * L.c0[1] + l2 * L.c0[2]);
      const double KE = sg * (l0 * L.E[0] + l1 * L.E[1] + l2 * L.E[2]);      // > 0
      // S+ = { F_k >= -E_k } is non-empty iff K >= -KE, S- = { F_k <= E_k } iff K <= KE (a hair of margin: a
      // part that is "just empty" is kept as a point-sized triangle)
      const bool has_p = K >= -1.000001 * KE, has_m = K <= 1.000001 * KE;
      double x0 = BIG, x1 = -BIG, y0 = BIG, y1 = -BIG, work = 0.0;
      for (int part = 0; part < 2; ++part) {
        if (!(part ? has_m : has_p)) continue;
        const double sp = part ? -1.0 : 1.0;          // the part is  sp F_k >= -E_k
        const double b0 = -sp * L.E[0] - L.c0[0], b1 = -sp * L.E[1] - L.c0[1], b2 = -sp * L.E[2] - L.c0[2];
        const double ax = (b0 * L.cy[1] - b1 * L.cy[0]) / x01, ay = (L.cx[0] * b1 - L.cx[1] * b0) / x01;
        const double bx = (b1 * L.cy[2] - b2 * L.cy[1]) / x12, by = (L.cx[1] * b2 - L.cx[2] * b1) / x12;
        const double cx = (b2 * L.cy[0] - b0 * L.cy[2]) / x20, cy = (L.cx[2] * b0 - L.cx[0] * b2) / x20;
        const double px0 = fmin(ax, fmin(bx, cx)), px1 = fmax(ax, fmax(bx, cx));
        const double py0 = fmin(ay, fmin(by, cy)), py1 = fmax(ay, fmax(by, cy));
        x0 = fmin(x0, px0); x1 = fmax(x1, px1); y0 = fmin(y0, py0); y1 = fmax(y1, py1);
        work += 0.5 * fabs((bx - ax) * (cy - ay) - (cx - ax) * (by - ay)) + 2.0 * ((px1 - px0) + (py1 - py0)) + 1.0;
      }
      if (has_p || has_m) {
        if (x1 < dx0 || x0 > dx1 || y1 < dy0 || y0 > dy1) return;             // off the domain (NaN falls through)
        if (x0 >= dx0 && x1 <= dx1 && y0 >= dy0 && y1 <= dy1) {
          R.x0 = x0; R.x1 = x1; R.y0 = y0; R.y1 = y1;
          R.work = work;
          plus = has_p; minus = has_m; done = true;
        }
      }
    }
  }
  if (!done) {
    Poly p;
    side_polygon(L, 1.0, dx0, dx1, dy0, dy1, p);
    plus = p.n > 0;
    grow(p, R.x0, R.x1, R.y0, R.y1);
    R.work += poly_work(p);
    side_polygon(L, -1.0, dx0, dx1, dy0, dy1, p);
    minus = p.n > 0;
    grow(p, R.x0, R.x1, R.y0, R.y1);
    R.work += poly_work(p);
  }
  if (!plus && !minus) return;
  R.kind = plus && minus ? 2 : 1;
  const double s = plus ? 1.0 : -1.0;                // orientation of the first region: s F_k >= -E_k
  for (int k = 0; k < 3; ++k) {
    if (len[k] > 1e-280) {
      R.nx[k] = -s * L.cx[k] / len[k]; R.ny[k] = -s * L.cy[k] / len[k];
      R.c[k] = (-s * L.c0[k] - L.E[k]) / len[k];
      R.m = fmax(R.m, 2.0 * L.E[k] / len[k]);
    } else {
      // a constant edge function: in or out for the whole domain
      R.nx[k] = 0.0; R.ny[k] = 0.0;
      const bool in_plus = L.c0[k] >= -L.E[k], in_minus = L.c0[k] <= L.E[k];
      R.c[k] = R.kind == 1 ? -1.0 : (in_plus && in_minus ? 0.0 : (in_plus ? -BIG : BIG));
    }
  }
  if (R.kind == 2) {
    // constant edges: c = 0 passes both tests (c <= 0, c >= -m), -BIG only S+, +BIG only S-
    for (int k = 0; k < 3; ++k)
      if (!(len[k] > 1e-280) && R.c[k] == 0.0) R.c[k] = -0.5 * R.m;
  }
}

// The x range of the region(s) inside the horizontal band ya <= y <= yb (a superset: every edge
// takes the y of the band that suits it best).  false: nothing of the region in this band.
RTB_RG bool row_range(const Region &R, double ya, double yb, double &xa, double &xb) {
  const double BIG = 1.0e300;
  double lo1 = -BIG, hi1 = BIG, lo2 = -BIG, hi2 = BIG;
  bool ok1 = true, ok2 = R.kind == 2;
  for (int k = 0; k < 3; ++k) {
    const double t0 = R.ny[k] * ya, t1 = R.ny[k] * yb;
    // first region: nx x <= -c - ny y for some y of the band
    const double r1 = -R.c[k] - fmin(t0, t1);
    if (R.nx[k] > 0.0) hi1 = fmin(hi1, r1 / R.nx[k]);
    else if (R.nx[k] < 0.0) lo1 = fmax(lo1, r1 / R.nx[k]);
    else if (r1 < 0.0) ok1 = false;
    // second region: nx x >= -m - c - ny y for some y of the band
    const double r2 = -R.m - R.c[k] - fmax(t0, t1);
    if (R.nx[k] > 0.0) lo2 = fmax(lo2, r2 / R.nx[k]);
    else if (R.nx[k] < 0.0) hi2 = fmin(hi2, r2 / R.nx[k]);
    else if (r2 > 0.0) ok2 = false;
  }
  ok1 = ok1 && lo1 <= hi1;
  ok2 = ok2 && lo2 <= hi2;
  if (!ok1 && !ok2) return false;
  xa = BIG; xb = -BIG;
  if (ok1) { xa = fmin(xa, lo1); xb = fmax(xb, hi1); }
  if (ok2) { xa = fmin(xa, lo2); xb = fmax(xb, hi2); }
  return true;
}

// Could the axis-aligned box [xa, xb] x [ya, yb] touch the region(s)?  (conservative: per-edge test)
RTB_RG bool box_touches(const Region &R, double xa, double xb, double ya, double yb) {
  bool in1 = true, in2 = R.kind == 2;
  for (int k = 0; k < 3; ++k) {
    const double lo = fmin(R.nx[k] * xa, R.nx[k] * xb) + fmin(R.ny[k] * ya, R.ny[k] * yb) + R.c[k];
    const double hi = fmax(R.nx[k] * xa, R.nx[k] * xb) + fmax(R.ny[k] * ya, R.ny[k] * yb) + R.c[k];
    in1 = in1 && lo <= 0.0;
    in2 = in2 && hi >= -R.m;
  }
  return in1 || in2;
}

// The record the per-sample / per-ray test reads: three float4.
//   r0 = (n0x, n0y, c0, n1x)  r1 = (n1y, c1, n2x, n2y)  r2 = (c2, X0, Y0, flag)
// with f_k = n_k . (x - X0, y - Y0) + c_k:  flag == 0: no test (everything inside the bounds passes);
// flag > 0: inside <=> f_k <= 0 for all k;  flag < 0: inside <=> all f_k <= 0 or all f_k >= flag.
// (X0, Y0) is a local origin chosen by the caller so that the float evaluation works on small
// numbers; ext = the largest |x - X0| + |y - Y0| the record will be evaluated at.  The constants
// are rounded outwards by the float evaluation's own error.
struct Record { float v[12]; };

RTB_RG void make_record(const Region &R, double X0, double Y0, double ext, Record &rec) {
  double slack_max = 0.0;
  float n[3][2], c[3];
  for (int k = 0; k < 3; ++k) {
    const double cl = R.c[k] + R.nx[k] * X0 + R.ny[k] * Y0;
    // float(n) is off by 2^-24 relative, three roundings in the evaluation: 4 * 2^-24 * (|n.dx| + |c|)
    const double slack = 3.0e-7 * (ext + fabs(cl)) + 1e-30;
    slack_max = fmax(slack_max, fmin(slack, 1e30));
    n[k][0] = (float)R.nx[k]; n[k][1] = (float)R.ny[k];
    double cs = cl - slack;
    if (cs > 1e30) cs = 1e30;
    if (cs < -1e30) cs = -1e30;
    c[k] = (float)cs;
    // (float) rounds to nearest: push once more so that the stored constant is not above cl - slack / 2
    if ((double)c[k] > cl - 0.5 * slack) c[k] = (float)(cs - 0.5 * slack - fabs(cs) * 1.2e-7);
  }
  float flag = 1.0f;
  if (R.kind == 2) {
    double m = (R.m + 2.0 * slack_max) * 1.000001 + 1e-30;
    if (m > 1e30) m = 1e30;
    flag = -(float)m;
    if ((double)(-flag) < m) flag = -(float)(m * 1.0000002);
  }
  rec.v[0] = n[0][0]; rec.v[1] = n[0][1]; rec.v[2] = c[0]; rec.v[3] = n[1][0];
  rec.v[4] = n[1][1]; rec.v[5] = c[1]; rec.v[6] = n[2][0]; rec.v[7] = n[2][1];
  rec.v[8] = c[2]; rec.v[9] = (float)X0; rec.v[10] = (float)Y0; rec.v[11] = flag;
}

// the test a sample / ray runs against a record; dx, dy relative to the record's origin
RTB_RG bool record_inside(float r0x, float r0y, float r0z, float r0w, float r1x, float r1y, float r1z, float r1w,
                          float r2x, float flag, float dx, float dy) {
  if (flag == 0.0f) return true;
  const float f0 = r0x * dx + r0y * dy + r0z, f1 = r0w * dx + r1x * dy + r1y, f2 = r1z * dx + r1w * dy + r2x;
  const bool in1 = f0 <= 0.0f && f1 <= 0.0f && f2 <= 0.0f;
  if (flag >= 0.0f) return in1;
  return in1 || (f0 >= flag && f1 >= flag && f2 >= flag);
}

// The samples of one row that can pass record_inside: offsets [a, b] inside an item row of w samples
// whose first sample is dx0 from the record's origin (a > b: none).  A superset by construction
// (the bound of each edge is widened by the rounding of the float evaluation and of the division);
// callers still run record_inside on the samples of the span.
RTB_RG void record_span(float r0x, float r0y, float r0z, float r0w, float r1x, float r1y, float r1z, float r1w,
                        float r2x, float flag, float dx0, int w, float dy, int &a, int &b) {
  a = 0; b = w - 1;
  if (flag == 0.0f) return;
  const float BIGF = 3.0e38f;
  const float nx[3] = {r0x, r0w, r1z}, ny[3] = {r0y, r1x, r1w}, c[3] = {r0z, r1y, r2x};
  const float xmax = fabsf(dx0) + (float)w;
  float lo1 = -BIGF, hi1 = BIGF, lo2 = -BIGF, hi2 = BIGF;
  bool ok1 = true, ok2 = flag < 0.0f;
  for (int k = 0; k < 3; ++k) {
    const float t = ny[k] * dy, bk = t + c[k];
    const float err = 6.0e-7f * (fabsf(nx[k]) * xmax + fabsf(t) + fabsf(c[k])) + 1e-30f;
    if (nx[k] != 0.0f) {
      const float inv = 1.0f / nx[k];
      const float v1 = -bk * inv, v2 = (flag - bk) * inv;
      const float t1 = err * fabsf(inv) + 4.0e-7f * fabsf(v1) + 0.01f, t2 = err * fabsf(inv) + 4.0e-7f * fabsf(v2) + 0.01f;
      if (nx[k] > 0.0f) { hi1 = fminf(hi1, v1 + t1); lo2 = fmaxf(lo2, v2 - t2); }
      else { lo1 = fmaxf(lo1, v1 - t1); hi2 = fminf(hi2, v2 + t2); }
    } else {
      if (bk > err) ok1 = false;
      if (bk < flag - err) ok2 = false;
    }
  }
  ok1 = ok1 && lo1 <= hi1;
  ok2 = ok2 && lo2 <= hi2;
  float lo = BIGF, hi = -BIGF;
  if (ok1) { lo = fminf(lo, lo1); hi = fmaxf(hi, hi1); }
  if (ok2) { lo = fminf(lo, lo2); hi = fmaxf(hi, hi2); }
  if (!(lo <= hi)) { a = 1; b = 0; return; }
  // offsets relative to the first sample of the row; the clamps keep the conversions in range
  const float fa = fminf(fmaxf(ceilf(lo - dx0), 0.0f), (float)w), fb = fmaxf(fminf(floorf(hi - dx0), (float)(w - 1)), -1.0f);
  a = (int)fa; b = (int)fb;
}

// The x range (relative to the record's origin) in which record_inside can pass for SOME dy of the
// band [dy0, dy1]: what the light grids use to list the cells of a region from the very record the
// rays are tested against.  A superset: every edge takes the dy of the band that suits it best, and
// the bounds are widened by the rounding of this evaluation and of record_inside's (xext = the
// largest |dx| of interest).  false: no point of the band passes.
RTB_RG bool record_band(float r0x, float r0y, float r0z, float r0w, float r1x, float r1y, float r1z, float r1w,
                        float r2x, float flag, float dy0, float dy1, float xext, float &lo, float &hi) {
  const float BIGF = 3.0e38f;
  lo = -BIGF; hi = BIGF;
  if (flag == 0.0f) return true;
  const float nx[3] = {r0x, r0w, r1z}, ny[3] = {r0y, r1x, r1w}, c[3] = {r0z, r1y, r2x};
  float lo1 = -BIGF, hi1 = BIGF, lo2 = -BIGF, hi2 = BIGF;
  bool ok1 = true, ok2 = flag < 0.0f;
  for (int k = 0; k < 3; ++k) {
    const float t0 = ny[k] * dy0, t1 = ny[k] * dy1;
    const float tmin = fminf(t0, t1), tmax = fmaxf(t0, t1);
    const float err = 1.0e-6f * (fabsf(nx[k]) * xext + fmaxf(fabsf(t0), fabsf(t1)) + fabsf(c[k])) + 1e-30f;
    // first part: nx dx <= -c - ny dy for some dy;   second part: nx dx >= flag - c - ny dy for some dy
    const float b1 = -c[k] - tmin + err, b2 = flag - c[k] - tmax - err;
    if (nx[k] != 0.0f) {
      const float inv = 1.0f / nx[k];
      const float v1 = b1 * inv, v2 = b2 * inv;
      const float w1 = 4.0e-7f * fabsf(v1) + 1e-30f, w2 = 4.0e-7f * fabsf(v2) + 1e-30f;
      if (nx[k] > 0.0f) { hi1 = fminf(hi1, v1 + w1); lo2 = fmaxf(lo2, v2 - w2); }
      else { lo1 = fmaxf(lo1, v1 - w1); hi2 = fminf(hi2, v2 + w2); }
    } else {
      if (b1 < 0.0f) ok1 = false;
      if (b2 > 0.0f) ok2 = false;
    }
  }
  ok1 = ok1 && lo1 <= hi1;
  ok2 = ok2 && lo2 <= hi2;
  if (!ok1 && !ok2) return false;
  lo = BIGF; hi = -BIGF;
  if (ok1) { lo = fminf(lo, lo1); hi = fmaxf(hi, hi1); }
  if (ok2) { lo = fminf(lo, lo2); hi = fmaxf(hi, hi2); }
  return true;
}

}  // namespace region
}  // namespace rtb

// ---------------------------------------------------------------------------------------------
// Edge functions and allowances for the three ray families.  eps = 2^-24 (binary32 unit roundoff).
// Rounding of the reference's operation sequence (first order; SAFETY covers the second):
//   h = fl(d x e2):   |dh| <= 2 sqrt2 eps |d||e2|      s = fl(o - v0):  |ds_k| <= eps |s_k|
//   U_c = fl(s . h):  |U_c - (o - v0) . (d x e2)| <= (1 + 2 sqrt2 + 3) eps |s||d||e2| < 7 eps |s||d||e2|
//   V_c = fl(d . fl(s x e1)): the same with e1;   a_c = fl(e1 . h): |a_c - e1 . (d x e2)| < 6 eps |e1||d||e2|
//   u_c + v_c <= 1 computed  =>  U_c + V_c <= a_c (1 + 4 eps)  (a_c > 0; mirrored for a_c < 0)
namespace rtb {
namespace region {

constexpr double EPS24 = 5.9604644775390625e-8;
constexpr double SAFETY = 1.25;
constexpr double A_CUTOFF = 1.0e-7 * (1.0 - 1e-6);     // |a| below EPSILON is rejected (gpu/hit.cu:24)

RTB_RG void cross3(const double a[3], const double b[3], double o[3]) {
  o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}
RTB_RG double dot3(const double a[3], const double b[3]) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
RTB_RG double len3(const double a[3]) { return sqrt(dot3(a, a)); }

// ---- primary rays (gpu/raytracer.cu:97-105; csrc/trace.cu primary_ray) ----
// Sample (x, y) = (px - W/2, py - H/2).  Ideal origin o* = C + u x + v y, ideal direction along
// D* = pos - o* = D0 - u x - v y (real arithmetic on the float camera vectors).  The computed origin
// is off by |do| <= 4 eps Mo, the computed direction is d_c = alpha (r + r o theta), r = pos - o_c,
// |theta_k| <= 2 eps, alpha = 1 / |r| up to rounding.  With p = pos - v0:
//   (o_c - v0) . (d_c x e2) = p . (d_c x e2) - r . (d_c x e2),   |r . (d_c x e2)| <= 2 eps |r||e2|
//   p . (d_c x e2) = alpha [ D* . (e2 x p) + (r o theta - do) . (e2 x p) ]
// so  U_c = alpha U*(x, y) + eta,  U* = D* . gu,  gu = e2 x p, and dividing by alpha (<= Dmax):
//   |eta| / alpha <= eps |e2| [ 9 Dmax^2 + |p| (9 Dmax + 4 Mo) ] = E_u.   V*: gv = p x e1, |e1|.
//   A* = D* . ga, ga = e2 x e1, |eta_a| / alpha <= eps |e1||e2| (8 Dmax + 4 Mo);  W* = A* - U* - V*.
struct CamFrame {
  double pos[3], D0[3], u[3], v[3];
  double Dmax;              // upper bound of |pos - o| over the frame's samples
  double Mo;                // | (|C_k| + |u_k| max|x| + |v_k| max|y|)_k |
};

// false: no primary ray can be accepted (|a| stays below the cut-off of gpu/hit.cu:24)
RTB_RG bool primary_lines(const CamFrame &cf, const float v0[3], const float e1f[3], const float e2f[3], Lines &L) {
  const double p[3] = {cf.pos[0] - (double)v0[0], cf.pos[1] - (double)v0[1], cf.pos[2] - (double)v0[2]};
  const double e1[3] = {e1f[0], e1f[1], e1f[2]}, e2[3] = {e2f[0], e2f[1], e2f[2]};
  double g[3][3];
  cross3(e2, p, g[0]);
  cross3(p, e1, g[1]);
  double ga[3];
  cross3(e2, e1, ga);
  for (int k = 0; k < 3; ++k) g[2][k] = ga[k] - g[0][k] - g[1][k];
  const double l1 = len3(e1), l2 = len3(e2), lp = len3(p), la = len3(ga);
  if (la * 1.000001 + 6.0 * EPS24 * l1 * l2 < A_CUTOFF) return false;
  for (int k = 0; k < 3; ++k) {
    L.c0[k] = dot3(g[k], cf.D0);
    L.cx[k] = -dot3(g[k], cf.u);
    L.cy[k] = -dot3(g[k], cf.v);
  }
  const double common = 9.0 * cf.Dmax * cf.Dmax + lp * (9.0 * cf.Dmax + 4.0 * cf.Mo);
  L.E[0] = SAFETY * EPS24 * l2 * common;
  L.E[1] = SAFETY * EPS24 * l1 * common;
  L.E[2] = L.E[0] + L.E[1] + SAFETY * EPS24 * (l1 * l2 * (8.0 * cf.Dmax + 4.0 * cf.Mo) + 4.0 * la * cf.Dmax);
  return true;
}

// ---- shadow rays of a directional light (gpu/light.cu:68-95): o = P (a hit point), d = -l.v ----
// U = (P - v0) . gu with gu = d x e2: linear in P, and only P's position across the light matters.
// The grid names a ray by x = fl(P . ax), y = fl(P . ay) (float frame ax, ay of lightgrid.cuh); gu is
// written as  alpha ax + beta ay + rho  (least squares; rho is what the float frame misses), so
//   | U - (alpha x + beta y - v0 . gu) | <= |rho| Pmax + 3.01 eps Pmax (|alpha| + |beta|).
struct DirFrame {
  double d[3], ax[3], ay[3];
  double g11, g12, g22;     // Gram matrix of (ax, ay)
  double smax;              // upper bound of |P - v0|: shadow-ray origins lie in the scene's box (lg_any checks)
  double Pmax;              // upper bound of |P|
};

RTB_RG void in_frame(const DirFrame &f, const double g[3], double &a, double &b, double &rho) {
  const double ga = dot3(g, f.ax), gb = dot3(g, f.ay), det = f.g11 * f.g22 - f.g12 * f.g12;
  a = (ga * f.g22 - gb * f.g12) / det;
  b = (gb * f.g11 - ga * f.g12) / det;
  const double r[3] = {g[0] - a * f.ax[0] - b * f.ay[0], g[1] - a * f.ax[1] - b * f.ay[1], g[2] - a * f.ax[2] - b * f.ay[2]};
  rho = len3(r);
}

RTB_RG bool dir_lines(const DirFrame &f, const float v0f[3], const float e1f[3], const float e2f[3], Lines &L) {
  const double v0[3] = {v0f[0], v0f[1], v0f[2]}, e1[3] = {e1f[0], e1f[1], e1f[2]}, e2[3] = {e2f[0], e2f[1], e2f[2]};
  double gu[3], gv[3];
  cross3(f.d, e2, gu);
  cross3(e1, f.d, gv);
  const double A = dot3(e1, gu);
  const double l1 = len3(e1), l2 = len3(e2), ld = len3(f.d);
  if (fabs(A) + 6.0 * EPS24 * l1 * l2 * ld < A_CUTOFF) return false;
  double a[2], b[2], rho[2];
  in_frame(f, gu, a[0], b[0], rho[0]);
  in_frame(f, gv, a[1], b[1], rho[1]);
  L.cx[0] = a[0]; L.cy[0] = b[0]; L.c0[0] = -dot3(v0, gu);
  L.cx[1] = a[1]; L.cy[1] = b[1]; L.c0[1] = -dot3(v0, gv);
  L.cx[2] = -a[0] - a[1]; L.cy[2] = -b[0] - b[1]; L.c0[2] = A - L.c0[0] - L.c0[1];
  const double xi = 3.01 * EPS24 * f.Pmax;
  L.E[0] = SAFETY * (7.0 * EPS24 * f.smax * ld * l2 + rho[0] * f.Pmax + xi * (fabs(a[0]) + fabs(b[0])));
  L.E[1] = SAFETY * (7.0 * EPS24 * f.smax * ld * l1 + rho[1] * f.Pmax + xi * (fabs(a[1]) + fabs(b[1])));
  L.E[2] = L.E[0] + L.E[1] + SAFETY * EPS24 * (6.0 * l1 * l2 * ld + 4.0 * fabs(A));
  return true;
}

// ---- shadow rays of a point light (gpu/light.cu:96-120): o = P, d = fl(l.v - P) ----
// With r = P - org, p = org - v0:  (P - v0) . (d_c x e2) = -r . gu + (terms below),  gu = e2 x p, and the
// grid names the ray by x = (r . ax) / (r . az), y = (r . ay) / (r . az): with the dual basis
// (hx, hy, hz) of the float frame (ax, ay, az),  r . gu = den (gu.hx x + gu.hy y + gu.hz),  den = r . az.
//   |U_c + den F_u(x, y)| <= eps |d||e2| (7 |s| + |d| + |p|)   (rounding of d_c included)
// divided by |den| >= |d| / RATIO inside the grid (|x|, |y| <= the grid's limit), plus the rounding of
// the ray's own (x, y): XI = 128 eps.  Which of S+ / S- a ray lands in depends on the sign of den;
// the union is taken anyway.
struct PointFrame {
  double org[3], hx[3], hy[3], hz[3];
  double smax, dmax;        // upper bounds of |P - v0| and |org - P|
  double ratio;             // upper bound of |r| / |r . az| for rays the grid answers
};

RTB_RG bool point_lines(const PointFrame &f, const float v0f[3], const float e1f[3], const float e2f[3], Lines &L) {
  const double p[3] = {f.org[0] - (double)v0f[0], f.org[1] - (double)v0f[1], f.org[2] - (double)v0f[2]};
  const double e1[3] = {e1f[0], e1f[1], e1f[2]}, e2[3] = {e2f[0], e2f[1], e2f[2]};
  double g[3][3], ga[3];
  cross3(e2, p, g[0]);
  cross3(p, e1, g[1]);
  cross3(e2, e1, ga);
  for (int k = 0; k < 3; ++k) g[2][k] = ga[k] - g[0][k] - g[1][k];
  const double l1 = len3(e1), l2 = len3(e2), lp = len3(p), la = len3(ga);
  if (f.dmax * (la + 7.0 * EPS24 * l1 * l2) * 1.000001 < A_CUTOFF) return false;
  for (int k = 0; k < 3; ++k) {
    L.cx[k] = dot3(g[k], f.hx);
    L.cy[k] = dot3(g[k], f.hy);
    L.c0[k] = dot3(g[k], f.hz);
  }
  const double XI = 128.0 * EPS24;
  const double common = f.ratio * EPS24 * (7.0 * f.smax + f.dmax + lp);
  L.E[0] = SAFETY * (l2 * common + XI * (fabs(L.cx[0]) + fabs(L.cy[0])));
  L.E[1] = SAFETY * (l1 * common + XI * (fabs(L.cx[1]) + fabs(L.cy[1])));
  L.E[2] = L.E[0] + L.E[1] + SAFETY * (f.ratio * EPS24 * (7.0 * l1 * l2 + 4.0 * la) + XI * (fabs(L.cx[2]) + fabs(L.cy[2])));
  return true;
}


// ---- what k_rast_setup needs per triangle: the sample rectangle, the record, the tiles of a row ----
struct PrimaryCull {
  int kind;                 // 0: no primary ray of this frame can be accepted; else Region::kind
  int x0, y0, x1, y1;       // sample rectangle (inclusive), inside the frame
  Region R;                 // in (x, y) = (px - W/2, py - H/2)
  Record rec;               // origin (x0, y0) in sample coordinates
};

RTB_RG void primary_cull(const CamFrame &cf, int W, int H, const float v0[3], const float e1[3], const float e2[3],
                         PrimaryCull &c) {
  c.kind = 0;
  Lines L;
  if (!primary_lines(cf, v0, e1, e2, L)) return;
  const double hx = (double)(W / 2), hy = (double)(H / 2);      // integer division as in primary_ray
  build(L, -hx - 0.5, (double)(W - 1) - hx + 0.5, -hy - 0.5, (double)(H - 1) - hy + 0.5, c.R);
  if (!c.R.kind) return;
  const double fx0 = fmax(ceil(c.R.x0 + hx - 1e-6), 0.0), fx1 = fmin(floor(c.R.x1 + hx + 1e-6), (double)(W - 1));
  const double fy0 = fmax(ceil(c.R.y0 + hy - 1e-6), 0.0), fy1 = fmin(floor(c.R.y1 + hy + 1e-6), (double)(H - 1));
  if (!(fx1 >= fx0 && fy1 >= fy0)) return;
  c.x0 = (int)fx0; c.x1 = (int)fx1; c.y0 = (int)fy0; c.y1 = (int)fy1;
  c.kind = c.R.kind;
  make_record(c.R, fx0 - hx, fy0 - hy, (fx1 - fx0) + (fy1 - fy0), c.rec);
  c.rec.v[9] = (float)c.x0; c.rec.v[10] = (float)c.y0;
}

// tiles [ta, tb] of tile row ty (tiles of `tile` samples) that the region can touch; false: none
RTB_RG bool primary_row_tiles(const PrimaryCull &c, int W, int H, int tile, int ty, int &ta, int &tb) {
  const double hx = (double)(W / 2), hy = (double)(H / 2);
  const int ya = ty * tile > c.y0 ? ty * tile : c.y0, yb = ty * tile + tile - 1 < c.y1 ? ty * tile + tile - 1 : c.y1;
  if (yb < ya) return false;
  double xa, xb;
  if (!row_range(c.R, (double)ya - hy, (double)yb - hy, xa, xb)) return false;
  const double fa = fmax(ceil(xa + hx - 1e-6), (double)c.x0), fb = fmin(floor(xb + hx + 1e-6), (double)c.x1);
  if (!(fb >= fa)) return false;
  ta = (int)fa / tile; tb = (int)fb / tile;
  return true;
}

}  // namespace region
}  // namespace rtb
