// Per-Gaussian / per-pair arithmetic of the splat render path, shared by every kernel.
//
// The functions are __host__ __device__ so that the same source also compiles with g++ into
// the test-only host harness (tests/host_harness), which lets the hand-derived VJPs be checked
// against the oracle without a GPU.  The harness is never linked into the product library.
//
// Boundary replaced: gsplat v0.1.11 kernels behind project_gaussians / spherical_harmonics /
// rasterize_gaussians as called at /root/reference/utils/gs_utils.py:78,82-95,96-109
// (semantics: SURVEY.md Appendix A).
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define SB_HD __host__ __device__ __forceinline__
#else
#define SB_HD inline
#endif

namespace sb {

constexpr float kAlphaMin = 1.0f / 255.0f;
constexpr float kAlphaMax = 0.999f;
constexpr float kTMin = 1e-4f;
constexpr float kRadiusMax = 1073741824.0f;  // 2^30, see oracle/ref_torch.py RADIUS_MAX
constexpr float kBlur = 0.3f;

// ------------------------------------------------------------------------------------------
// A.1 spherical harmonics basis.  Y must hold (deg+1)^2 floats.  (x,y,z) is normalised here.
// ------------------------------------------------------------------------------------------
SB_HD void sh_basis(int deg, float dx, float dy, float dz, float* Y) {
  const float n = sqrtf(dx * dx + dy * dy + dz * dz);
  const float x = dx / n, y = dy / n, z = dz / n;
  Y[0] = 0.28209479177387814f;
  if (deg < 1) return;
  Y[1] = -0.4886025119029199f * y;
  Y[2] = 0.4886025119029199f * z;
  Y[3] = -0.4886025119029199f * x;
  if (deg < 2) return;
  const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
  Y[4] = 1.0925484305920792f * xy;
  Y[5] = -1.0925484305920792f * yz;
  Y[6] = 0.31539156525252005f * (2.0f * zz - xx - yy);
  Y[7] = -1.0925484305920792f * xz;
  Y[8] = 0.5462742152960396f * (xx - yy);
  if (deg < 3) return;
  Y[9] = -0.5900435899266435f * y * (3.0f * xx - yy);
  Y[10] = 2.890611442640554f * xy * z;
  Y[11] = -0.4570457994644658f * y * (4.0f * zz - xx - yy);
  Y[12] = 0.3731763325901154f * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
  Y[13] = -0.4570457994644658f * x * (4.0f * zz - xx - yy);
  Y[14] = 1.445305721320277f * z * (xx - yy);
  Y[15] = -0.5900435899266435f * x * (xx - 3.0f * yy);
  if (deg < 4) return;
  Y[16] = 2.5033429417967046f * xy * (xx - yy);
  Y[17] = -1.7701307697799304f * yz * (3.0f * xx - yy);
  Y[18] = 0.9461746957575601f * xy * (7.0f * zz - 1.0f);
  Y[19] = -0.6690465435572892f * yz * (7.0f * zz - 3.0f);
  Y[20] = 0.10578554691520431f * (zz * (35.0f * zz - 30.0f) + 3.0f);
  Y[21] = -0.6690465435572892f * xz * (7.0f * zz - 3.0f);
  Y[22] = 0.47308734787878004f * (xx - yy) * (7.0f * zz - 1.0f);
  Y[23] = -1.7701307697799304f * xz * (xx - 3.0f * yy);
  Y[24] = 0.6258357354491761f * (xx * (xx - 3.0f * yy) - yy * (3.0f * xx - yy));
}

// ------------------------------------------------------------------------------------------
// A.2 / A.3 projection forward.
//
// ARITHMETIC CONTRACT: when this function is compiled with FMA contraction off
// (nvcc --fmad=false / g++ -ffp-contract=off) every operation below is one IEEE binary32
// rounding in the order written, which is the order oracle/ref_torch.py:project_gaussians
// uses; the results are then bit-identical to the oracle.
// ------------------------------------------------------------------------------------------
struct Cam {
  float w[9];   // world->camera rotation, row-major (viewmat[:3,:3])
  float p[3];   // viewmat[:3,3]
  float fx, fy, cx, cy;
};

struct ProjFwd {
  // outputs
  float px, py, depth, ca, cb, cc, comp;  // ca,cb,cc = conic
  float cov3d[6];
  int radius, tiles, min_x, max_x, min_y, max_y;
  bool front;   // passed the near-plane test (cov3d is meaningful)
  bool valid;   // contributes to the image
  // intermediates kept for the backward pass
  float tx, ty, tz, txc, tyc;
  bool clamp_x, clamp_y;
  float T0[3], T1[3];
  float a, b, c, det, det_orig;   // blurred cov2d entries and determinants
  float M[9];                     // R * diag(glob*scale)
  float R[9];
  float qn[4], qnorm;
};

SB_HD int sb_trunc_clamp(float v, int hi) {
  // (int) cast with saturation (NaN -> 0 like cvt.rzi), then clamp to [0, hi]
  int iv;
  if (!(v == v)) iv = 0;
  else if (v >= 2.0e9f) iv = 2000000000;
  else if (v <= -2.0e9f) iv = -2000000000;
  else iv = (int)v;
  return iv < 0 ? 0 : (iv > hi ? hi : iv);
}

SB_HD void tile_bbox(float px, float py, float radius, int tiles_x, int tiles_y, int bw,
                     int& min_x, int& max_x, int& min_y, int& max_y) {
  const float fbw = (float)bw;
  float tcx, tcy, tr;
  if ((bw & (bw - 1)) == 0) {
    // a power of two: multiplying by the (exact) reciprocal gives the bits of the division
    const float ibw = 1.0f / fbw;
    tcx = px * ibw; tcy = py * ibw; tr = radius * ibw;
  } else {
    tcx = px / fbw; tcy = py / fbw; tr = radius / fbw;
  }
  min_x = sb_trunc_clamp(tcx - tr, tiles_x);
  max_x = sb_trunc_clamp(tcx + tr + 1.0f, tiles_x);
  min_y = sb_trunc_clamp(tcy - tr, tiles_y);
  max_y = sb_trunc_clamp(tcy + tr + 1.0f, tiles_y);
}

SB_HD bool sb_finite(float v) { return (v - v) == 0.0f; }

// Wrapper-level quaternion normalisation, literally /root/reference/utils/gs_utils.py:46-51:
//   quats = q / norm(q);  mask = (norm(quats) - 1) < 1e-6;  quats[~mask] = (0, 0, 0, 1)
// (a zero, NaN or infinite quaternion fails the one-sided test and is replaced).  The projection then
// normalises once more, as gsplat.project_gaussians does with the wrapper's output.  Returns the norm
// of the raw quaternion, or 0 when the substitution happened (no gradient flows to such a quaternion).
SB_HD float wrapper_normalise_quat(float* q) {
  const float n = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  const float a = q[0] / n, b = q[1] / n, c = q[2] / n, d = q[3] / n;
  const float n1 = sqrtf(a * a + b * b + c * c + d * d);
  if (!(n1 - 1.0f < 1e-6f)) { q[0] = 0.0f; q[1] = 0.0f; q[2] = 0.0f; q[3] = 1.0f; return 0.0f; }
  q[0] = a; q[1] = b; q[2] = c; q[3] = d;
  return n;
}

// View-independent half of the projection: quaternion normalisation, rotation, M = R diag(glob * s),
// Sigma = M M^T.  One call per Gaussian; the fused multi-view kernels keep only `sg` across their view loop.
struct GaussCov {
  float sg[6];                    // Sigma: s00 s01 s02 s11 s12 s22  (= the cov3d output)
  float M[9], R[9], qn[4], qnorm;
};

SB_HD void gaussian_cov3d(const float* s, float glob, const float* q, GaussCov& c) {
  const float qn = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  const float qw = q[0] / qn, qx = q[1] / qn, qy = q[2] / qn, qz = q[3] / qn;
  c.qnorm = qn; c.qn[0] = qw; c.qn[1] = qx; c.qn[2] = qy; c.qn[3] = qz;
  float* R = c.R;
  R[0] = 1.0f - 2.0f * (qy * qy + qz * qz); R[1] = 2.0f * (qx * qy - qw * qz); R[2] = 2.0f * (qx * qz + qw * qy);
  R[3] = 2.0f * (qx * qy + qw * qz); R[4] = 1.0f - 2.0f * (qx * qx + qz * qz); R[5] = 2.0f * (qy * qz - qw * qx);
  R[6] = 2.0f * (qx * qz - qw * qy); R[7] = 2.0f * (qy * qz + qw * qx); R[8] = 1.0f - 2.0f * (qx * qx + qy * qy);
  const float s0 = glob * s[0], s1 = glob * s[1], s2 = glob * s[2];
  float* M = c.M;
  for (int i = 0; i < 3; ++i) { M[3 * i] = R[3 * i] * s0; M[3 * i + 1] = R[3 * i + 1] * s1; M[3 * i + 2] = R[3 * i + 2] * s2; }
#define SB_SIG(i, j) (M[3 * (i)] * M[3 * (j)] + M[3 * (i) + 1] * M[3 * (j) + 1] + M[3 * (i) + 2] * M[3 * (j) + 2])
  c.sg[0] = SB_SIG(0, 0); c.sg[1] = SB_SIG(0, 1); c.sg[2] = SB_SIG(0, 2);
  c.sg[3] = SB_SIG(1, 1); c.sg[4] = SB_SIG(1, 2); c.sg[5] = SB_SIG(2, 2);
#undef SB_SIG
}

// 1.3 * tan(fov / 2) limits of the EWA clamp (A.2): a property of the camera, not of the Gaussian.
SB_HD void cam_limits(const Cam& cam, int img_w, int img_h, float& lim_x, float& lim_y) {
  lim_x = 1.3f * ((0.5f * (float)img_w) / cam.fx);
  lim_y = 1.3f * ((0.5f * (float)img_h) / cam.fy);
}

// View-dependent half, given Sigma.  Does not touch o.cov3d / o.M / o.R / o.qn / o.qnorm.
// inv_bw: 1 / bw when that is exact (bw a power of two), else 0 (then the tile bbox divides by bw).
SB_HD void project_fwd_view(const float* m, const float* sg6, const Cam& cam, float lim_x, float lim_y,
                            int img_w, int img_h, int bw, float clip, ProjFwd& o) {
  o.valid = false; o.front = false; o.radius = 0; o.tiles = 0;
  o.px = o.py = o.depth = o.ca = o.cb = o.cc = o.comp = 0.0f;
  const float* w = cam.w;
  const float x = m[0], y = m[1], z = m[2];
  const float tx = w[0] * x + w[1] * y + w[2] * z + cam.p[0];
  const float ty = w[3] * x + w[4] * y + w[5] * z + cam.p[1];
  const float tz = w[6] * x + w[7] * y + w[8] * z + cam.p[2];
  o.tx = tx; o.ty = ty; o.tz = tz;
  if (!(tz > clip)) return;
  o.front = true;
  const float s00 = sg6[0], s01 = sg6[1], s02 = sg6[2], s11 = sg6[3], s12 = sg6[4], s22 = sg6[5];
  const float sg[9] = {s00, s01, s02, s01, s11, s12, s02, s12, s22};

  const float rz = 1.0f / tz;
  const float rz2 = rz * rz;
  const float ux = tx * rz, uy = ty * rz;
  o.clamp_x = (ux < -lim_x) || (ux > lim_x);
  o.clamp_y = (uy < -lim_y) || (uy > lim_y);
  const float txc = tz * fminf(lim_x, fmaxf(-lim_x, ux));
  const float tyc = tz * fminf(lim_y, fmaxf(-lim_y, uy));
  o.txc = txc; o.tyc = tyc;
  const float j00 = cam.fx * rz;
  const float j02 = -(cam.fx * txc) * rz2;
  const float j11 = cam.fy * rz;
  const float j12 = -(cam.fy * tyc) * rz2;
  float* T0 = o.T0; float* T1 = o.T1;
  for (int j = 0; j < 3; ++j) {
    T0[j] = j00 * w[j] + j02 * w[6 + j];
    T1[j] = j11 * w[3 + j] + j12 * w[6 + j];
  }
  float a0[3], a1[3];
  for (int j = 0; j < 3; ++j) {
    a0[j] = T0[0] * sg[j] + T0[1] * sg[3 + j] + T0[2] * sg[6 + j];
    a1[j] = T1[0] * sg[j] + T1[1] * sg[3 + j] + T1[2] * sg[6 + j];
  }
  const float c00 = a0[0] * T0[0] + a0[1] * T0[1] + a0[2] * T0[2];
  const float c01 = a0[0] * T1[0] + a0[1] * T1[1] + a0[2] * T1[2];
  const float c11 = a1[0] * T1[0] + a1[1] * T1[1] + a1[2] * T1[2];

  const float det_orig = c00 * c11 - c01 * c01;
  const float a = c00 + kBlur, b = c01, c = c11 + kBlur;
  const float det = a * c - b * b;
  o.a = a; o.b = b; o.c = c; o.det = det; o.det_orig = det_orig;
  if (det == 0.0f || !sb_finite(det)) return;
  const float ratio = det_orig / det;
  const float comp = ratio > 0.0f ? sqrtf(ratio) : 0.0f;
  const float inv_det = 1.0f / det;
  const float k0 = c * inv_det, k1 = -b * inv_det, k2 = a * inv_det;
  const float bmid = 0.5f * (a + c);
  const float disc = sqrtf(fmaxf(0.1f, bmid * bmid - det));
  const float v1 = bmid + disc, v2 = bmid - disc;
  const float radius = ceilf(3.0f * sqrtf(fmaxf(v1, v2)));
  const float rw = 1.0f / (tz + 1e-6f);
  const float px = cam.fx * (tx * rw) + cam.cx;
  const float py = cam.fy * (ty * rw) + cam.cy;
  if (!(sb_finite(radius) && radius <= kRadiusMax && sb_finite(px) && sb_finite(py))) return;
  const int tiles_x = (img_w + bw - 1) / bw, tiles_y = (img_h + bw - 1) / bw;
  tile_bbox(px, py, radius, tiles_x, tiles_y, bw, o.min_x, o.max_x, o.min_y, o.max_y);
  const int area = (o.max_x - o.min_x) * (o.max_y - o.min_y);
  if (area <= 0) return;
  o.valid = true;
  o.px = px; o.py = py; o.depth = tz; o.ca = k0; o.cb = k1; o.cc = k2; o.comp = comp;
  o.radius = (int)radius; o.tiles = area;
}

SB_HD void project_fwd_one(const float* m, const float* s, float glob, const float* q,
                           const Cam& cam, int img_w, int img_h, int bw, float clip, ProjFwd& o) {
  for (int i = 0; i < 6; ++i) o.cov3d[i] = 0.0f;
  // the near-plane test comes first in A.2: cov3d is only an output for Gaussians in front of it
  const float tz = cam.w[6] * m[0] + cam.w[7] * m[1] + cam.w[8] * m[2] + cam.p[2];
  GaussCov gc;
  float lim_x, lim_y;
  cam_limits(cam, img_w, img_h, lim_x, lim_y);
  if (tz > clip) {
    gaussian_cov3d(s, glob, q, gc);
    for (int i = 0; i < 6; ++i) o.cov3d[i] = gc.sg[i];
    for (int i = 0; i < 9; ++i) { o.M[i] = gc.M[i]; o.R[i] = gc.R[i]; }
    for (int i = 0; i < 4; ++i) o.qn[i] = gc.qn[i];
    o.qnorm = gc.qnorm;
  } else {
    for (int i = 0; i < 6; ++i) gc.sg[i] = 0.0f;
  }
  project_fwd_view(m, gc.sg, cam, lim_x, lim_y, img_w, img_h, bw, clip, o);
}

// ------------------------------------------------------------------------------------------
// A.7 projection backward (VJP of project_fwd_one for a Gaussian with valid == true, given
// the forward record `f`).  v_cov3d may be nullptr.  Outputs are *assigned*.
// v_vm (12 floats: d/d viewmat[:3,:4] row-major) is accumulated into if non-null.
// ------------------------------------------------------------------------------------------
// Part A, per (view, Gaussian): everything up to the gradient of Sigma.  Inputs are the forward record `f`
// of project_fwd_view and Sigma; outputs v_mean (assigned) and the symmetric 3x3 gradient G = dL/dSigma
// (assigned; its six distinct entries appear twice).  v_vm as in project_bwd_one.
SB_HD void project_bwd_view(const float* m, const float* sg6, const Cam& cam, const ProjFwd& f, float v_px,
                            float v_py, float v_depth, float v_k0, float v_k1, float v_k2, float v_comp,
                            float* v_mean, float* G, float* v_vm) {
  const float* w = cam.w;
  const float a = f.a, b = f.b, c = f.c, det = f.det;
  const float inv_det2 = 1.0f / (det * det);
  // conic = (c, -b, a) / det
  float v_a = (-c * c * v_k0 + b * c * v_k1 - b * b * v_k2) * inv_det2;
  float v_c = (-b * b * v_k0 + a * b * v_k1 - a * a * v_k2) * inv_det2;
  float v_b = (2.0f * b * c * v_k0 - (a * c + b * b) * v_k1 + 2.0f * a * b * v_k2) * inv_det2;
  // compensation = sqrt(det_orig / det)
  if (v_comp != 0.0f && f.det_orig / det > 0.0f) {
    const float comp = sqrtf(f.det_orig / det);
    const float v_r = v_comp / (2.0f * comp);
    const float a0 = a - kBlur, c0 = c - kBlur;
    v_a += v_r * (c0 * det - f.det_orig * c) * inv_det2;
    v_c += v_r * (a0 * det - f.det_orig * a) * inv_det2;
    v_b += v_r * (-2.0f * b * det + 2.0f * b * f.det_orig) * inv_det2;
  }
  // cov2d = T Sigma T^T ; symmetric gradient V2 = [[v_a, v_b/2],[v_b/2, v_c]]
  const float* T0 = f.T0; const float* T1 = f.T1;
  const float sg[9] = {sg6[0], sg6[1], sg6[2], sg6[1], sg6[3], sg6[4], sg6[2], sg6[4], sg6[5]};
  float st0[3], st1[3];  // Sigma * T0, Sigma * T1
  for (int i = 0; i < 3; ++i) {
    st0[i] = sg[3 * i] * T0[0] + sg[3 * i + 1] * T0[1] + sg[3 * i + 2] * T0[2];
    st1[i] = sg[3 * i] * T1[0] + sg[3 * i + 1] * T1[1] + sg[3 * i + 2] * T1[2];
  }
  float v_T0[3], v_T1[3];
  for (int i = 0; i < 3; ++i) {
    v_T0[i] = 2.0f * v_a * st0[i] + v_b * st1[i];
    v_T1[i] = v_b * st0[i] + 2.0f * v_c * st1[i];
  }
  // G = T^T V2 T (symmetric 3x3)
  const float hb = 0.5f * v_b;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j)
      G[3 * i + j] = v_a * T0[i] * T0[j] + hb * (T0[i] * T1[j] + T1[i] * T0[j]) + v_c * T1[i] * T1[j];
  // J chain
  const float tz = f.tz, rz = 1.0f / tz, rz2 = rz * rz;
  float v_j00 = 0.f, v_j02 = 0.f, v_j11 = 0.f, v_j12 = 0.f;
  for (int j = 0; j < 3; ++j) {
    v_j00 += v_T0[j] * w[j]; v_j02 += v_T0[j] * w[6 + j];
    v_j11 += v_T1[j] * w[3 + j]; v_j12 += v_T1[j] * w[6 + j];
  }
  float v_rz = cam.fx * v_j00 + cam.fy * v_j11 - 2.0f * cam.fx * f.txc * rz * v_j02
               - 2.0f * cam.fy * f.tyc * rz * v_j12;
  const float v_txc = -cam.fx * rz2 * v_j02;
  const float v_tyc = -cam.fy * rz2 * v_j12;
  float v_tx = 0.f, v_ty = 0.f, v_tz = 0.f;
  if (f.clamp_x) v_tz += (f.txc * rz) * v_txc; else v_tx += v_txc;
  if (f.clamp_y) v_tz += (f.tyc * rz) * v_tyc; else v_ty += v_tyc;
  v_tz += -rz2 * v_rz;
  // pixel centre and depth
  const float rw = 1.0f / (tz + 1e-6f);
  v_tx += cam.fx * rw * v_px;
  v_ty += cam.fy * rw * v_py;
  v_tz += -rw * rw * (cam.fx * f.tx * v_px + cam.fy * f.ty * v_py);
  v_tz += v_depth;
  v_mean[0] = w[0] * v_tx + w[3] * v_ty + w[6] * v_tz;
  v_mean[1] = w[1] * v_tx + w[4] * v_ty + w[7] * v_tz;
  v_mean[2] = w[2] * v_tx + w[5] * v_ty + w[8] * v_tz;
  if (v_vm) {
    const float j00 = cam.fx * rz, j11 = cam.fy * rz;
    const float j02 = -(cam.fx * f.txc) * rz2, j12 = -(cam.fy * f.tyc) * rz2;
    const float vt[3] = {v_tx, v_ty, v_tz};
    for (int i = 0; i < 3; ++i) {
      for (int j = 0; j < 3; ++j) v_vm[4 * i + j] += vt[i] * m[j];
      v_vm[4 * i + 3] += vt[i];
    }
    for (int j = 0; j < 3; ++j) {
      v_vm[j] += j00 * v_T0[j];
      v_vm[4 + j] += j11 * v_T1[j];
      v_vm[8 + j] += j02 * v_T0[j] + j12 * v_T1[j];
    }
  }
}

// Part B, per Gaussian: Sigma = M M^T with M = R(q / |q|) diag(glob * s)  ->  v_scale, v_quat from G = dL/dSigma.
// Linear in G, so the fused multi-view backward sums G over the views first and runs this once.
SB_HD void cov3d_chain_bwd(const float* s, float glob, const float* M, const float* R, const float* qn4,
                           float qnorm, const float* G, float* v_scale, float* v_quat) {
  float vM[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j)
      vM[3 * i + j] = 2.0f * (G[3 * i] * M[j] + G[3 * i + 1] * M[3 + j] + G[3 * i + 2] * M[6 + j]);
  const float sc[3] = {glob * s[0], glob * s[1], glob * s[2]};
  float vR[9];
  for (int j = 0; j < 3; ++j) {
    v_scale[j] = glob * (R[j] * vM[j] + R[3 + j] * vM[3 + j] + R[6 + j] * vM[6 + j]);
    for (int i = 0; i < 3; ++i) vR[3 * i + j] = vM[3 * i + j] * sc[j];
  }
  const float qw = qn4[0], qx = qn4[1], qy = qn4[2], qz = qn4[3];
  float vq[4];
  vq[0] = 2.0f * (-qz * vR[1] + qy * vR[2] + qz * vR[3] - qx * vR[5] - qy * vR[6] + qx * vR[7]);
  vq[1] = 2.0f * (qy * vR[1] + qz * vR[2] + qy * vR[3] - 2.0f * qx * vR[4] - qw * vR[5] + qz * vR[6]
                  + qw * vR[7] - 2.0f * qx * vR[8]);
  vq[2] = 2.0f * (-2.0f * qy * vR[0] + qx * vR[1] + qw * vR[2] + qx * vR[3] + qz * vR[5] - qw * vR[6]
                  + qz * vR[7] - 2.0f * qy * vR[8]);
  vq[3] = 2.0f * (-2.0f * qz * vR[0] - qw * vR[1] + qx * vR[2] + qw * vR[3] - 2.0f * qz * vR[4]
                  + qy * vR[5] + qx * vR[6] + qy * vR[7]);
  // through q / |q|
  const float dotp = qw * vq[0] + qx * vq[1] + qy * vq[2] + qz * vq[3];
  const float inv_n = 1.0f / qnorm;
  v_quat[0] = (vq[0] - qw * dotp) * inv_n;
  v_quat[1] = (vq[1] - qx * dotp) * inv_n;
  v_quat[2] = (vq[2] - qy * dotp) * inv_n;
  v_quat[3] = (vq[3] - qz * dotp) * inv_n;
}

SB_HD void project_bwd_one(const float* m, const float* s, float glob, const Cam& cam,
                           const ProjFwd& f, float v_px, float v_py, float v_depth, float v_k0,
                           float v_k1, float v_k2, float v_comp, const float* v_cov3d,
                           float* v_mean, float* v_scale, float* v_quat, float* v_vm) {
  float G[9];
  project_bwd_view(m, f.cov3d, cam, f, v_px, v_py, v_depth, v_k0, v_k1, v_k2, v_comp, v_mean, G, v_vm);
  if (v_cov3d) {
    // direct cov3d gradient
    G[0] += v_cov3d[0]; G[4] += v_cov3d[3]; G[8] += v_cov3d[5];
    G[1] += 0.5f * v_cov3d[1]; G[3] += 0.5f * v_cov3d[1];
    G[2] += 0.5f * v_cov3d[2]; G[6] += 0.5f * v_cov3d[2];
    G[5] += 0.5f * v_cov3d[4]; G[7] += 0.5f * v_cov3d[4];
  }
  cov3d_chain_bwd(s, glob, f.M, f.R, f.qn, f.qnorm, G, v_scale, v_quat);
}

// cov3d-only forward/backward used when the near-plane test passed but the Gaussian is not
// drawn (cov3d is still an API output with a gradient).
SB_HD void cov3d_bwd_one(const float* s, float glob, const ProjFwd& f, const float* v_cov3d,
                         float* v_scale, float* v_quat) {
  const float G[9] = {v_cov3d[0], 0.5f * v_cov3d[1], 0.5f * v_cov3d[2],
                      0.5f * v_cov3d[1], v_cov3d[3], 0.5f * v_cov3d[4],
                      0.5f * v_cov3d[2], 0.5f * v_cov3d[4], v_cov3d[5]};
  cov3d_chain_bwd(s, glob, f.M, f.R, f.qn, f.qnorm, G, v_scale, v_quat);
}

// ------------------------------------------------------------------------------------------
// A.5 / A.6 per (pixel, Gaussian) pair.
// ------------------------------------------------------------------------------------------
#if defined(__CUDA_ARCH__)
#define SB_EXP(x) __expf(x)
#else
#define SB_EXP(x) expf(x)
#endif

// Conservative half extents of {d : alpha(d) >= 1/255} around the centre; -3e38 = never visible,
// +3e38 = do not cull.
SB_HD void visible_extent(float ca, float cb, float cc, float opac, float& hx, float& hy) {
  hx = hy = -3.0e38f;
  const float r = opac * 255.0f;
  if (!(r >= 1.0f)) return;                       // alpha can never reach 1/255 (also NaN)
  const float tau = logf(r) * 1.0001f + 1e-4f;    // sigma <= tau (slightly inflated)
  const float det = ca * cc - cb * cb;
  if (!(det > 0.0f) || !(ca > 0.0f) || !(cc > 0.0f)) { hx = hy = 3.0e38f; return; }  // not an ellipse: no culling
  hx = sqrtf(2.0f * tau * cc / det) * 1.0001f + 1e-3f;
  hy = sqrtf(2.0f * tau * ca / det) * 1.0001f + 1e-3f;
  if (!sb_finite(hx) || !sb_finite(hy)) { hx = hy = 3.0e38f; }
}

}  // namespace sb
