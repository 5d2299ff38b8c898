// TEST-ONLY host build of the per-Gaussian arithmetic in splatformer_b200/csrc/sb_math.cuh.
// Compiled with g++ -ffp-contract=off by tests/host_harness/build.py; never part of the product
// library.  Lets the projection forward (bit-exact contract) and the hand-derived projection VJP
// be checked against the oracle on a machine without a GPU.
#include "../../splatformer_b200/csrc/sb_math.cuh"

extern "C" {

void hh_project_fwd(int n, const float* means, const float* scales, float glob, const float* quats,
                    const float* viewmat /*3x4 row-major*/, float fx, float fy, float cx, float cy,
                    int img_h, int img_w, int bw, float clip, float* xys, float* depths, int* radii,
                    float* conics, float* comp, int* tiles, float* cov3d) {
  sb::Cam cam;
  for (int i = 0; i < 3; ++i) { for (int j = 0; j < 3; ++j) cam.w[3 * i + j] = viewmat[4 * i + j]; cam.p[i] = viewmat[4 * i + 3]; }
  cam.fx = fx; cam.fy = fy; cam.cx = cx; cam.cy = cy;
  for (int g = 0; g < n; ++g) {
    sb::ProjFwd o;
    sb::project_fwd_one(means + 3 * g, scales + 3 * g, glob, quats + 4 * g, cam, img_w, img_h, bw, clip, o);
    xys[2 * g] = o.px; xys[2 * g + 1] = o.py; depths[g] = o.depth; radii[g] = o.radius;
    conics[3 * g] = o.ca; conics[3 * g + 1] = o.cb; conics[3 * g + 2] = o.cc; comp[g] = o.comp;
    tiles[g] = o.tiles;
    for (int k = 0; k < 6; ++k) cov3d[6 * g + k] = o.cov3d[k];
  }
}

void hh_project_bwd(int n, const float* means, const float* scales, float glob, const float* quats,
                    const float* viewmat, float fx, float fy, float cx, float cy, int img_h, int img_w,
                    int bw, float clip, const float* v_xys, const float* v_depths, const float* v_conics,
                    const float* v_comp, const float* v_cov3d, float* v_means, float* v_scales,
                    float* v_quats, float* v_viewmat /*12, accumulated*/) {
  sb::Cam cam;
  for (int i = 0; i < 3; ++i) { for (int j = 0; j < 3; ++j) cam.w[3 * i + j] = viewmat[4 * i + j]; cam.p[i] = viewmat[4 * i + 3]; }
  cam.fx = fx; cam.fy = fy; cam.cx = cx; cam.cy = cy;
  for (int g = 0; g < n; ++g) {
    sb::ProjFwd o;
    sb::project_fwd_one(means + 3 * g, scales + 3 * g, glob, quats + 4 * g, cam, img_w, img_h, bw, clip, o);
    float vm[3] = {0, 0, 0}, vs[3] = {0, 0, 0}, vq[4] = {0, 0, 0, 0};
    if (o.valid) {
      sb::project_bwd_one(means + 3 * g, scales + 3 * g, glob, cam, o, v_xys[2 * g], v_xys[2 * g + 1],
                          v_depths[g], v_conics[3 * g], v_conics[3 * g + 1], v_conics[3 * g + 2],
                          v_comp[g], v_cov3d + 6 * g, vm, vs, vq, v_viewmat);
    } else if (o.front) {
      sb::cov3d_bwd_one(scales + 3 * g, glob, o, v_cov3d + 6 * g, vs, vq);
    }
    for (int k = 0; k < 3; ++k) { v_means[3 * g + k] = vm[k]; v_scales[3 * g + k] = vs[k]; }
    for (int k = 0; k < 4; ++k) v_quats[4 * g + k] = vq[k];
  }
}

void hh_sh_basis(int n, int deg, const float* dirs, float* Y /*n x (deg+1)^2*/) {
  const int K = (deg + 1) * (deg + 1);
  for (int g = 0; g < n; ++g) sb::sh_basis(deg, dirs[3 * g], dirs[3 * g + 1], dirs[3 * g + 2], Y + K * g);
}

void hh_visible_extent(int n, const float* conics, const float* opac, float* hx, float* hy) {
  for (int g = 0; g < n; ++g) sb::visible_extent(conics[3 * g], conics[3 * g + 1], conics[3 * g + 2], opac[g], hx[g], hy[g]);
}
}
