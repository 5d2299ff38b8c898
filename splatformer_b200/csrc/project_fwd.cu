// Projection forward (reference call site /root/reference/utils/gs_utils.py:82-95).
//
// THIS TRANSLATION UNIT IS COMPILED WITH --fmad=false: the arithmetic contract in sb_math.cuh /
// oracle/ref_torch.py requires every float operation of the projection chain to be rounded
// separately, so that radii, tile counts, depth keys and pixel centres are bit-identical to the
// oracle.  The kernel is a pure stream (100 B per Gaussian per view), so losing FMA costs nothing.
#include "sh_stage.cuh"
#include "raster_common.cuh"

namespace sb {

constexpr int kProjBlock = 256;

__global__ void __launch_bounds__(kProjBlock) project_fwd_kernel(
    int n, const float* __restrict__ means, const float* __restrict__ scales, float glob,
    const float* __restrict__ quats, CamSrc cams, int img_h, int img_w, int bw, float clip,
    float* __restrict__ xys, float* __restrict__ depths, int* __restrict__ radii,
    float* __restrict__ conics, float* __restrict__ comp, int* __restrict__ tiles,
    float* __restrict__ cov3d) {
  const int g = blockIdx.x * kProjBlock + threadIdx.x;
  const int v = blockIdx.y;
  if (g >= n) return;
  const Cam cam = load_cam(cams, v);
  const float m[3] = {__ldg(means + 3 * g), __ldg(means + 3 * g + 1), __ldg(means + 3 * g + 2)};
  const float s[3] = {__ldg(scales + 3 * g), __ldg(scales + 3 * g + 1), __ldg(scales + 3 * g + 2)};
  const float4 q4 = __ldg(reinterpret_cast<const float4*>(quats) + g);
  const float q[4] = {q4.x, q4.y, q4.z, q4.w};
  ProjFwd o;
  project_fwd_one(m, s, glob, q, cam, img_w, img_h, bw, clip, o);
  const size_t i = (size_t)v * n + g;
  reinterpret_cast<float2*>(xys)[i] = make_float2(o.px, o.py);
  depths[i] = o.depth;
  radii[i] = o.radius;
  conics[3 * i] = o.ca; conics[3 * i + 1] = o.cb; conics[3 * i + 2] = o.cc;
  if (comp) comp[i] = o.comp;
  tiles[i] = o.tiles;
  if (cov3d) {
#pragma unroll
    for (int k = 0; k < 6; ++k) cov3d[6 * i + k] = o.cov3d[k];
  }
}

// ------------------------------------------------------------------------------------------------
// Fused multi-view variant (rows a1/a2: the view loop of rasterize_gaussians_to_multiimgs,
// /root/reference/utils/gs_utils.py:20-27, with the per-view SH colour of :67-79 folded in).
// Four threads per Gaussian keep its parameters and SH coefficients in registers and walk the V
// cameras (every fourth view each), writing the packed 48-byte raster record, depth, radius and tile
// count per view.  Coefficients are read from HBM once per scene instead of once per view.
//   DEG = -1 : color_src is (N,3) view-independent rgb (SplatFormer's degree-0 path, :65)
//   DEG >= 0 : colour = max(SH(dir) + 0.5, 0)  (:78-79); the coefficients are either one (N,K,3)
//              array (color_src, color_rest = NULL) or features_dc (N,3) + features_rest (N,K-1,3)
//              as SplatFormer stores them (:59-60), staged side by side without a concatenation.
// `flags` (sb_common.cuh) select raw inputs: log-scales, opacity logits, sigmoid colours, c2w cameras.
// ------------------------------------------------------------------------------------------------
constexpr int kViewsBlock = 128;
constexpr int kViewLanes = 4;                          // threads that share one Gaussian, each taking V/4 views
constexpr int kViewsRows = kViewsBlock / kViewLanes;   // Gaussians per block

// Shared-memory row stride of a Gaussian's SH coefficients: a multiple of four floats (16-byte loads) whose quarter
// is odd, so that the eight Gaussians of a warp (four lanes each, same address within a Gaussian) hit eight different
// 16-byte bank groups.
__host__ __device__ inline int fwd_sh_row(int k3) {
  const int r = (k3 + 3) & ~3;
  return (r & 7) ? r : r + 4;
}

#ifndef SB_PROJ_FWD_MINB
#define SB_PROJ_FWD_MINB 5        // 96 registers, no spills, with the SH coefficients read from shared memory: 0.191 ms (4: 115 regs 0.200; 6: 80 regs + spills 0.220)
#endif
#define SB_PROJ_FWD_BOUNDS __launch_bounds__(kViewsBlock, SB_PROJ_FWD_MINB)
template <int DEG>
__global__ void SB_PROJ_FWD_BOUNDS views_project_fwd_kernel(
    int num_views, int n, const float* __restrict__ means, const float* __restrict__ scales, float glob,
    const float* __restrict__ quats, const float* __restrict__ opacities,
    const float* __restrict__ color_src, const float* __restrict__ color_rest, int k3_store, int flags,
    CamSrc cams, const float* __restrict__ campos, int img_h, int img_w, int bw, float clip,
    float4* __restrict__ packed, float* __restrict__ depths, int* __restrict__ radii,
    int* __restrict__ tiles, int* __restrict__ tile_counts, unsigned long long* __restrict__ tile_keys,
    int tile_cap) {
  extern __shared__ float smem[];
  constexpr int KU = DEG < 0 ? 1 : (DEG + 1) * (DEG + 1);
  const int g0 = blockIdx.x * kViewsRows;
  const int nb = min(kViewsRows, n - g0);
  const int rowi = threadIdx.x / kViewLanes, sub = threadIdx.x % kViewLanes;
  // SH coefficients: they STAY in shared memory and are read, four at a time, where a view's colour is evaluated.
  // Holding the 48 floats of a degree-3 Gaussian in registers across the view loop was what pinned the kernel at 128
  // registers (four CTAs per SM, part of the array in local memory anyway) -- and this kernel is bound by latency,
  // not by issue slots.
  float c[3] = {0.f, 0.f, 0.f};           // DEG < 0: the colour itself
  if (DEG >= 0) {
    const int row = fwd_sh_row(k3_store);
    if (color_rest) {
      stage_rows_in(smem, color_src + (size_t)g0 * 3, nb, 3, kViewsBlock, row, 0);
      if (k3_store > 3) stage_rows_in(smem, color_rest + (size_t)g0 * (k3_store - 3), nb, k3_store - 3, kViewsBlock, row, 3);
    } else {
      stage_rows_in(smem, color_src + (size_t)g0 * k3_store, nb, k3_store, kViewsBlock, row, 0);
    }
    __syncthreads();
  }
  // cameras: built once per block (shared memory after the SH staging rows) when there are few enough views
  const bool cams_staged = num_views <= kMaxStagedCams;
  float* s_cams = smem + (DEG >= 0 ? kViewsRows * fwd_sh_row(k3_store) : 0);
  if (cams_staged) {
    stage_cams(cams, campos, num_views, img_w, img_h, s_cams);
    __syncthreads();
  }
  const int g = g0 + rowi;
  if (rowi >= nb) return;
  const float4* crow4 = reinterpret_cast<const float4*>(smem + rowi * fwd_sh_row(k3_store));
  if (DEG < 0) {
    c[0] = __ldg(color_src + 3 * g); c[1] = __ldg(color_src + 3 * g + 1); c[2] = __ldg(color_src + 3 * g + 2);
    if (flags & kFlagColorSigmoid) { c[0] = sigmoidf_(c[0]); c[1] = sigmoidf_(c[1]); c[2] = sigmoidf_(c[2]); }
  }
  const float m[3] = {__ldg(means + 3 * g), __ldg(means + 3 * g + 1), __ldg(means + 3 * g + 2)};
  float s[3] = {__ldg(scales + 3 * g), __ldg(scales + 3 * g + 1), __ldg(scales + 3 * g + 2)};
  if (flags & kFlagRawScales) { s[0] = expf(s[0]); s[1] = expf(s[1]); s[2] = expf(s[2]); }
  const float4 q4 = __ldg(reinterpret_cast<const float4*>(quats) + g);
  float q[4] = {q4.x, q4.y, q4.z, q4.w};
  if (flags & kFlagRawQuats) wrapper_normalise_quat(q);      // gs_utils.py:46-51
  float op = __ldg(opacities + g);
  if (flags & kFlagRawOpacities) op = sigmoidf_(op);
  // Sigma = (R S)(R S)^T does not depend on the view: once per Gaussian, only its six entries stay live
  float sg[6];
  {
    GaussCov gc;
    gaussian_cov3d(s, glob, q, gc);
#pragma unroll
    for (int i = 0; i < 6; ++i) sg[i] = gc.sg[i];
  }
  // blocks start at different views so that, at any moment, the per-tile count reductions of the
  // resident blocks spread over all V*T counters instead of the T counters of one view
  const int v0 = blockIdx.x % num_views;
  for (int vi = sub; vi < num_views; vi += kViewLanes) {
    const int v = vi + v0 < num_views ? vi + v0 : vi + v0 - num_views;
    Cam cam;
    float lim_x, lim_y;
    if (cams_staged) {
      cam = cam_from_smem(s_cams + v * kCamFloats, lim_x, lim_y);
    } else {
      cam = load_cam(cams, v);
      cam_limits(cam, img_w, img_h, lim_x, lim_y);
    }
    ProjFwd o;
    project_fwd_view(m, sg, cam, lim_x, lim_y, img_w, img_h, bw, clip, o);
    const size_t i = (size_t)v * n + g;
    radii[i] = o.radius;
    tiles[i] = o.tiles;
    depths[i] = o.depth;
    if (!o.valid) continue;     // packed record of a culled Gaussian is never read
    // binning for tile_sort.cu: claim a slot in the fixed-capacity segment of every covered tile and
    // drop the (depth bits, id) key there.  A tile that runs over keeps counting (the sort kernel
    // reports the overflow and the host falls back to the global sort).  In the common case (<= 4
    // tiles) all slots are claimed first and the keys are stored after the colour evaluation, so the
    // round trips of the atomics overlap each other and the SH arithmetic.
    size_t tt[4];
    int pos[4];
    const unsigned long long key = ((unsigned long long)(unsigned)__float_as_int(o.depth) << 32) | (unsigned)g;
    const bool few = o.tiles <= 4;
    if (tile_counts) {
      const int tiles_x = (img_w + bw - 1) / bw;
      const size_t t0 = (size_t)v * tiles_x * ((img_h + bw - 1) / bw);
      const int nx = o.max_x - o.min_x;
      if (few) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (k < o.tiles) {
            const int ky = (nx == 1) ? k : ((nx >= 4) ? 0 : (nx == 2 ? (k >> 1) : (k == 3)));
            const int kx = k - ky * nx;
            tt[k] = t0 + (size_t)(o.min_y + ky) * tiles_x + (o.min_x + kx);
            pos[k] = atomicAdd(tile_counts + tt[k], 1);
          }
        }
      } else {
        // many tiles: claim the slots eight at a time so that the atomics' round trips overlap instead
        // of queueing behind each other (a 1M-Gaussian room scene averages 47 tiles per Gaussian)
        // the bounding box is walked row by row with a running tile index (no integer division per tile)
        const int total = nx * (o.max_y - o.min_y);
        size_t cur = t0 + (size_t)o.min_y * tiles_x + o.min_x;
        int kx = 0;
        for (int b0 = 0; b0 < total; b0 += 8) {
          size_t tb[8];
          int pb[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            if (b0 + k < total) {
              tb[k] = cur;
              pb[k] = atomicAdd(tile_counts + cur, 1);
              if (++kx == nx) { kx = 0; cur += tiles_x - nx + 1; } else { ++cur; }
            }
          }
#pragma unroll
          for (int k = 0; k < 8; ++k)
            if (b0 + k < total && pb[k] < tile_cap) tile_keys[tb[k] * tile_cap + pb[k]] = key;
        }
      }
    }
    float r, gg, b;
    if (DEG >= 0) {
      float Y[KU], cp[3];
      if (cams_staged) { const float* d = s_cams + v * kCamFloats; cp[0] = d[18]; cp[1] = d[19]; cp[2] = d[20]; }
      else load_campos(cams, campos, v, cp);
      sh_basis(DEG, m[0] - cp[0], m[1] - cp[1], m[2] - cp[2], Y);
      r = 0.f; gg = 0.f; b = 0.f;
      // coefficient e = 3 k + channel; per channel the sum runs over ascending k, as it always did
#pragma unroll
      for (int q4 = 0; q4 < (KU * 3 + 3) / 4; ++q4) {
        const float4 cc = crow4[q4];
        const float vals[4] = {cc.x, cc.y, cc.z, cc.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int e = 4 * q4 + j;
          if (e < KU * 3) {
            // (explicit FMA: the colour is outside the bit-exact projection / tiling contract of this
            // --fmad=false unit, and gsplat.spherical_harmonics' own kernel, sh.cu, contracts too)
            if (e % 3 == 0) r = fmaf(Y[e / 3], vals[j], r);
            else if (e % 3 == 1) gg = fmaf(Y[e / 3], vals[j], gg);
            else b = fmaf(Y[e / 3], vals[j], b);
          }
        }
      }
      r = fmaxf(r + 0.5f, 0.f); gg = fmaxf(gg + 0.5f, 0.f); b = fmaxf(b + 0.5f, 0.f);
    } else {
      r = c[0]; gg = c[1]; b = c[2];
    }
    float4* p = packed + 3 * i;
    pack_record(o.px, o.py, o.ca, o.cb, o.cc, op, r, gg, b, p[0], p[1], p[2]);
    if (tile_counts && few) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (k < o.tiles && pos[k] < tile_cap) tile_keys[tt[k] * tile_cap + pos[k]] = key;
    }
  }
}

}  // namespace sb

extern "C" int sb_project_fwd(int num_views, int n, const float* means, const float* scales,
                              float glob_scale, const float* quats, const float* viewmats,
                              int viewmat_stride, const float* intrins, float fx, float fy, float cx,
                              float cy, int img_h, int img_w, int block_width, float clip_thresh,
                              float* xys, float* depths, int* radii, float* conics,
                              float* compensation, int* num_tiles_hit, float* cov3d,
                              sb_stream_t stream) {
  if (n < 0 || num_views < 1 || block_width < 2 || block_width > 16 || img_h < 1 || img_w < 1)
    return SB_ERR_BAD_ARG;
  if (n == 0) return 0;
  sb::CamSrc cams{viewmats, viewmat_stride, intrins, fx, fy, cx, cy, 0};
  dim3 grid(sb::ceil_div(n, sb::kProjBlock), num_views);
  sb::project_fwd_kernel<<<grid, sb::kProjBlock, 0, (cudaStream_t)stream>>>(
      n, means, scales, glob_scale, quats, cams, img_h, img_w, block_width, clip_thresh, xys, depths,
      radii, conics, compensation, num_tiles_hit, cov3d);
  SB_RETURN_LAST();
}

extern "C" int sb_views_project_fwd(int num_views, int n, const float* means, const float* scales,
                                    float glob_scale, const float* quats, const float* opacities,
                                    const float* color_src, const float* color_rest, int sh_degree,
                                    int flags, const float* viewmats, int viewmat_stride,
                                    const float* intrins, float fx, float fy, float cx, float cy,
                                    const float* campos, int img_h, int img_w, int block_width,
                                    float clip_thresh, float* packed, float* depths, int* radii,
                                    int* num_tiles_hit, int* tile_counts, uint64_t* tile_keys,
                                    int tile_capacity, sb_stream_t stream) {
  const bool c2w = flags & sb::kFlagC2wOpenGL;
  if (n < 0 || num_views < 1 || block_width < 2 || block_width > 16 || img_h < 1 || img_w < 1 ||
      sh_degree < -1 || sh_degree > 4 || (sh_degree >= 0 && !campos && !c2w) || (sh_degree < 0 && color_rest) ||
      ((flags & sb::kFlagColorSigmoid) && sh_degree >= 0) || (tile_counts && (!tile_keys || tile_capacity < 1)))
    return SB_ERR_BAD_ARG;
  if (n == 0) return 0;
  sb::CamSrc cams{viewmats, viewmat_stride, intrins, fx, fy, cx, cy, c2w ? 1 : 0,
                  (flags & sb::kFlagIntrinsShared) ? 1 : 0};
  const int k3 = sh_degree < 0 ? 3 : (sh_degree + 1) * (sh_degree + 1) * 3;
  const size_t smem = (sh_degree < 0 ? 0 : (size_t)sb::kViewsRows * sb::fwd_sh_row(k3) * sizeof(float)) +
                      (num_views <= sb::kMaxStagedCams ? (size_t)num_views * sb::kCamFloats * sizeof(float) : 0);
  dim3 grid(sb::ceil_div(n, sb::kViewsRows));
  cudaStream_t st = (cudaStream_t)stream;
  float4* pk = reinterpret_cast<float4*>(packed);
#define SB_CASE(D)                                                                                       \
  case D:                                                                                                \
    sb::views_project_fwd_kernel<D><<<grid, sb::kViewsBlock, smem, st>>>(                                \
        num_views, n, means, scales, glob_scale, quats, opacities, color_src, color_rest, k3, flags, cams, \
        campos, img_h, img_w, block_width, clip_thresh, pk, depths, radii, num_tiles_hit, tile_counts,    \
        reinterpret_cast<unsigned long long*>(tile_keys), tile_capacity);                                \
    break;
  switch (sh_degree) { SB_CASE(-1) SB_CASE(0) SB_CASE(1) SB_CASE(2) SB_CASE(3) SB_CASE(4) }
#undef SB_CASE
  SB_RETURN_LAST();
}
