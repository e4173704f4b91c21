/*
 * oracle/raster_oracle.c -- TEST INFRASTRUCTURE ONLY (never on the product path).
 *
 * CPU restatement of what the reference's OpenGL renderer computes:
 *   /root/reference/lib/render_glumpy/render_py_multi.py:17-47    GLSL programs
 *   /root/reference/lib/render_glumpy/render_py_multi.py:90-147   draw, readback, linearise
 *   /root/reference/lib/render_glumpy/render_py_multi.py:152-178  projection / view matrices
 *   /root/reference/lib/render_glumpy/render_py_light_modelnet_multi.py:14-79,149-231 (Lambert)
 *
 * The GL renderer itself is unavailable on this box (no OpenGL/EGL/Vulkan
 * loader, no software renderer); glumpy is not vendored.  PARITY UNPINNED: the
 * reference ships no rendered golden images.  This file follows the fixed
 * function pipeline literally so that the CUDA rasteriser (which uses a
 * pinhole formulation) is checked against an independent derivation:
 *
 *   clip   = [pos, 1] . V_np . P_np        (glumpy uploads the numpy matrices as
 *                                           they are; GLSL reads them column-major,
 *                                           so P*V*M*v == v . M . V . P in numpy)
 *   clip against the near plane  z_c >= -w_c   (GL primitive clipping)
 *   ndc    = clip.xyz / clip.w
 *   window = ((ndc.x+1) W/2, (ndc.y+1) H/2, (ndc.z+1)/2), pixel centres at +0.5
 *   flipud (:133-134):  image row = H-1 - GL row
 *   coverage: vertices snapped to 1/256 px (GL sub-pixel precision), integer
 *             edge functions, top-left tie rule; no face culling (:90-94)
 *   depth test GL_LESS on the window depth (float32), clear depth 1
 *   varyings interpolated perspective-correctly (1/w)
 *   texture2D: NEAREST, CLAMP_TO_EDGE, image flipped on upload (:75-77)
 *   8-bit colour buffer, float readback n/255, then * 255 (:136-137)
 *   depth = 2 f n / (f + n - (f - n)(2 d - 1)) in float32, background -> 0 (:139-146)
 *
 * Unknowable without the GL driver: 24-bit depth quantisation, the driver's
 * exact tie rule and sub-pixel bit count.  See DESIGN.md "rasteriser parity".
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
  double c[4];     /* clip coordinates */
  double attr[8];  /* u, v, nx, ny, nz, px, py, pz (object-space normal / position) */
} OVert;

static void lerp_vert(const OVert *a, const OVert *b, double t, OVert *o) {
  for (int i = 0; i < 4; ++i) o->c[i] = a->c[i] + t * (b->c[i] - a->c[i]);
  for (int i = 0; i < 8; ++i) o->attr[i] = a->attr[i] + t * (b->attr[i] - a->attr[i]);
}

/* Sutherland-Hodgman against z_c + w_c >= 0 (near plane). Returns vertex count (0, 3 or 4). */
static int clip_near(const OVert *in, OVert *out) {
  int n = 0;
  for (int i = 0; i < 3; ++i) {
    const OVert *a = &in[i], *b = &in[(i + 1) % 3];
    const double da = a->c[2] + a->c[3], db = b->c[2] + b->c[3];
    if (da >= 0) out[n++] = *a;
    if ((da >= 0) != (db >= 0)) {
      /* always interpolate from the inside vertex so both triangles sharing the edge agree */
      if (da >= 0) lerp_vert(a, b, da / (da - db), &out[n++]);
      else lerp_vert(b, a, db / (db - da), &out[n++]);
    }
  }
  return n;
}

typedef struct {
  int W, H;
  float *zbuf;      /* window depth, float32 */
  int *tri;         /* winning triangle */
  float *rgb;       /* [H,W,3] in 0..1 steps of 1/255 stored as the 8-bit integer value */
} OTarget;

typedef struct {
  int mode;                  /* 0 textured, 1 textured * Lambert (ModelNet), 2 LINEMOD light */
  const unsigned char *tex;  /* [Ht,Wt,3] RGB, row 0 = top of the PNG */
  int Ht, Wt;
  double view[16];           /* GLSL view matrix (column-vector convention), row-major here */
  double light_pos[3], light_int[3];
  double amb, dif;           /* 1 - brightness_ratio, brightness_ratio */
} OShade;

static inline int64_t edge(int64_t ax, int64_t ay, int64_t bx, int64_t by, int64_t px, int64_t py) {
  return (bx - ax) * (py - ay) - (by - ay) * (px - ax);
}

static inline int tie_bias(int64_t dx, int64_t dy) {
  /* owns the edge if it is a "left" edge (dy > 0) or a "top" edge (dy == 0, dx < 0) */
  return (dy > 0 || (dy == 0 && dx < 0)) ? 0 : -1;
}

static void shade(const OShade *S, const double *attr, unsigned char *out3) {
  /* texture2D(u_texture, v_texcoord), NEAREST + CLAMP_TO_EDGE, flipped upload */
  int tx = (int)floor(attr[0] * S->Wt), ty = (int)floor(attr[1] * S->Ht);
  if (tx < 0) tx = 0; if (tx > S->Wt - 1) tx = S->Wt - 1;
  if (ty < 0) ty = 0; if (ty > S->Ht - 1) ty = S->Ht - 1;
  const unsigned char *t = S->tex + ((size_t)(S->Ht - 1 - ty) * S->Wt + tx) * 3;
  if (S->mode == 0) {
    out3[0] = t[0]; out3[1] = t[1]; out3[2] = t[2];
    return;
  }
  /* render_py_light_modelnet_multi.py:50-74 */
  const double *V = S->view;
  double n[3], p[3];
  for (int r = 0; r < 3; ++r) {
    /* normal matrix = inverse transpose of the view matrix = its rotation block (rigid) */
    n[r] = V[r * 4 + 0] * attr[2] + V[r * 4 + 1] * attr[3] + V[r * 4 + 2] * attr[4];
    p[r] = V[r * 4 + 0] * attr[5] + V[r * 4 + 1] * attr[6] + V[r * 4 + 2] * attr[7] + V[r * 4 + 3];
  }
  double s2l[3] = {S->light_pos[0] - p[0], S->light_pos[1] - p[1], S->light_pos[2] - p[2]};
  const double ln = sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
  const double ls = sqrt(s2l[0] * s2l[0] + s2l[1] * s2l[1] + s2l[2] * s2l[2]);
  double br = (n[0] * s2l[0] + n[1] * s2l[1] + n[2] * s2l[2]) / (ls * ln);
  br = fmax(fmin(br, 1.0), 0.0);
  for (int c = 0; c < 3; ++c) {
    /* ModelNet: t * ((a + b*br) * I)   render_py_light_modelnet_multi.py:74
     * LINEMOD : t * (a + b*br * I)     render_py_light.py:74, render_py_light_multi_program.py:76 */
    double v = (S->mode == 2) ? (t[c] / 255.0) * (S->amb + S->dif * br * S->light_int[c])
                              : (t[c] / 255.0) * ((S->amb + S->dif * br) * S->light_int[c]);
    v = fmax(fmin(v, 1.0), 0.0);
    out3[c] = (unsigned char)floor(v * 255.0 + 0.5); /* unorm8 conversion of the colour buffer */
  }
}

static void raster_tri(const OVert *v, int tri_id, const OShade *S, OTarget *T) {
  const int W = T->W, H = T->H;
  int64_t X[3], Y[3];
  double zw[3], iw[3];
  for (int i = 0; i < 3; ++i) {
    const double w = v[i].c[3];
    if (!(w > 0)) return; /* cannot happen after near clipping with n > 0 */
    const double xw = (v[i].c[0] / w + 1.0) * W * 0.5;
    const double yw = (v[i].c[1] / w + 1.0) * H * 0.5;
    zw[i] = (v[i].c[2] / w + 1.0) * 0.5;
    iw[i] = 1.0 / w;
    const double u = xw - 0.5, vv = (double)(H - 1) - (yw - 0.5);
    if (!(fabs(u) < 1e6 && fabs(vv) < 1e6)) return;
    X[i] = (int64_t)rint(u * 256.0);
    Y[i] = (int64_t)rint(vv * 256.0);
  }
  int64_t area = edge(X[0], Y[0], X[1], Y[1], X[2], Y[2]);
  if (area == 0) return;
  int a = 0, b = 1, c = 2;
  if (area < 0) { b = 2; c = 1; area = -area; }
  const int idx[3] = {a, b, c};
  int64_t minx = X[0], maxx = X[0], miny = Y[0], maxy = Y[0];
  for (int i = 1; i < 3; ++i) {
    if (X[i] < minx) minx = X[i]; if (X[i] > maxx) maxx = X[i];
    if (Y[i] < miny) miny = Y[i]; if (Y[i] > maxy) maxy = Y[i];
  }
  int x0 = (int)((minx + 255) >> 8), x1 = (int)(maxx >> 8);
  int y0 = (int)((miny + 255) >> 8), y1 = (int)(maxy >> 8);
  if (x0 < 0) x0 = 0; if (y0 < 0) y0 = 0;
  if (x1 > W - 1) x1 = W - 1; if (y1 > H - 1) y1 = H - 1;
  const int64_t ax = X[idx[0]], ay = Y[idx[0]], bx = X[idx[1]], by = Y[idx[1]], cx = X[idx[2]], cy = Y[idx[2]];
  const int b0 = tie_bias(cx - bx, cy - by), b1 = tie_bias(ax - cx, ay - cy), b2 = tie_bias(bx - ax, by - ay);
  for (int y = y0; y <= y1; ++y) {
    for (int x = x0; x <= x1; ++x) {
      const int64_t px = (int64_t)x << 8, py = (int64_t)y << 8;
      const int64_t w0 = edge(bx, by, cx, cy, px, py); /* weight of vertex a */
      const int64_t w1 = edge(cx, cy, ax, ay, px, py);
      const int64_t w2 = edge(ax, ay, bx, by, px, py);
      if (w0 + b0 < 0 || w1 + b1 < 0 || w2 + b2 < 0) continue;
      const double l0 = (double)w0 / (double)area, l1 = (double)w1 / (double)area,
                   l2 = (double)w2 / (double)area;
      const float d = (float)(l0 * zw[idx[0]] + l1 * zw[idx[1]] + l2 * zw[idx[2]]);
      if (!(d >= 0.0f && d <= 1.0f)) continue; /* depth clipping of the fragment (far plane) */
      const size_t p = (size_t)y * W + x;
      if (!(d < T->zbuf[p])) continue; /* GL_LESS */
      T->zbuf[p] = d;
      T->tri[p] = tri_id;
      /* perspective-correct varyings */
      const double q0 = l0 * iw[idx[0]], q1 = l1 * iw[idx[1]], q2 = l2 * iw[idx[2]];
      const double qs = q0 + q1 + q2;
      double attr[8];
      for (int k = 0; k < 8; ++k)
        attr[k] = (q0 * v[idx[0]].attr[k] + q1 * v[idx[1]].attr[k] + q2 * v[idx[2]].attr[k]) / qs;
      unsigned char rgb[3];
      shade(S, attr, rgb);
      T->rgb[p * 3 + 0] = rgb[0]; T->rgb[p * 3 + 1] = rgb[1]; T->rgb[p * 3 + 2] = rgb[2];
    }
  }
}

/*
 * verts [V,3], uv [V,2] (may be NULL), normals [V,3] (may be NULL), tris [T,3]
 * proj_np, view_np: the 4x4 numpy arrays exactly as the reference builds them
 *   (my_compute_calib_proj, render_py_multi.py:152-169; _get_view_mtx, :171-178 -- i.e.
 *   view_np is already transposed), row-major float64.
 * Outputs (image orientation, after the reference's flipud):
 *   bgr   [H,W,3] float32: mode 0: (n/255)*255 as float32 (:136-137); mode 1: the uint8 value
 *   depth [H,W]   float32 metres, 0 = background
 *   tri_id [H,W]  int32, -1 = background (may be NULL)
 */
void oracle_render(const double *verts, const double *uv, const double *normals, const int *tris,
                   int V, int T, const unsigned char *tex, int Ht, int Wt, const double *proj_np,
                   const double *view_np, int W, int H, double zNear, double zFar, int mode,
                   const double *light_pos, const double *light_int, double brightness_ratio,
                   float *bgr, float *depth, int *tri_id) {
  (void)V;
  OTarget Tg;
  Tg.W = W; Tg.H = H;
  Tg.zbuf = (float *)malloc(sizeof(float) * W * H);
  Tg.tri = (int *)malloc(sizeof(int) * W * H);
  Tg.rgb = (float *)calloc((size_t)W * H * 3, sizeof(float));
  for (int i = 0; i < W * H; ++i) { Tg.zbuf[i] = 1.0f; Tg.tri[i] = -1; }
  OShade S;
  memset(&S, 0, sizeof(S));
  S.mode = mode; S.tex = tex; S.Ht = Ht; S.Wt = Wt;
  for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) S.view[r * 4 + c] = view_np[c * 4 + r];
  if (mode >= 1) {
    for (int i = 0; i < 3; ++i) { S.light_pos[i] = light_pos[i]; S.light_int[i] = light_int[i]; }
    S.amb = 1.0 - brightness_ratio; S.dif = brightness_ratio;
  }
  /* MVP = view_np . proj_np (row-vector convention, u_model = identity :79) */
  double M[16];
  for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) {
    double s = 0;
    for (int k = 0; k < 4; ++k) s += view_np[r * 4 + k] * proj_np[k * 4 + c];
    M[r * 4 + c] = s;
  }
  for (int t = 0; t < T; ++t) {
    OVert in[3], out[4];
    for (int i = 0; i < 3; ++i) {
      const int vi = tris[t * 3 + i];
      const double *p = verts + (size_t)vi * 3;
      for (int c = 0; c < 4; ++c)
        in[i].c[c] = p[0] * M[0 * 4 + c] + p[1] * M[1 * 4 + c] + p[2] * M[2 * 4 + c] + M[3 * 4 + c];
      in[i].attr[0] = uv ? uv[(size_t)vi * 2 + 0] : 0.0;
      in[i].attr[1] = uv ? uv[(size_t)vi * 2 + 1] : 0.0;
      for (int k = 0; k < 3; ++k) {
        in[i].attr[2 + k] = normals ? normals[(size_t)vi * 3 + k] : 0.0;
        in[i].attr[5 + k] = p[k];
      }
    }
    const int n = clip_near(in, out);
    if (n >= 3) raster_tri(out, t, &S, &Tg);
    if (n == 4) {
      OVert q[3] = {out[0], out[2], out[3]};
      raster_tri(q, t, &S, &Tg);
    }
  }
  /* readback + post-processing, render_py_multi.py:133-146 (float32 arithmetic) */
  /* python-float scalars are evaluated in float64 and enter the float32 array ops as float32 */
  const float c0 = (float)(2.0 * zFar * zNear), c1 = (float)(zFar + zNear), c2 = (float)(zFar - zNear);
  for (int i = 0; i < W * H; ++i) {
    const float d = Tg.zbuf[i];
    float z = 0.0f;
    if (d != 1.0f) {
      const float t2 = 2.0f * d - 1.0f;
      z = c0 / (c1 - c2 * t2);
    }
    depth[i] = z;
    if (tri_id) tri_id[i] = Tg.tri[i];
    for (int c = 0; c < 3; ++c) {
      const float n8 = Tg.rgb[(size_t)i * 3 + c];
      float v = (mode == 0) ? (n8 / 255.0f) * 255.0f : n8;
      bgr[(size_t)i * 3 + (2 - c)] = v; /* [2,1,0] reorder to BGR */
    }
  }
  free(Tg.zbuf); free(Tg.tri); free(Tg.rgb);
}
